"""The allocation-free fp32 matmul entry point against the ordinary one.

`lowbit_matmul_f32_noalloc_cuda` exists for hosts that cannot hand the library memory (an arena that
owns the whole device, a fixed memory budget, a thread that must stay off the allocator): its scratch
is a fixed array inside the library.  Through the C ABI it must give what `lowbit_matmul_cuda` gives
on the same inputs, call after call, because every call reuses the same array.

Collected last on purpose (file name): a run that stops at the first failure has then already run
every other test.
"""
import ctypes

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

MAX_ELEMS = 262144      # m * k limit stated in include/b200_lowbit.h


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def test_noalloc_matches_regular_entry_point_twice(native_lib, cuda_device):
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    ggml, dev = native_lib, cuda_device
    rng = np.random.default_rng(4242)
    N, K = 400, 1024                       # six 64-row groups + 16 tail rows
    qt = lbl.SYM_INT4
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float32)
    packed = torch.from_numpy(oracle.quantize(w, "sym_int4")).to(dev)
    tiled = lbl.ggml_q_format_convet_cpu2xpu(packed, N * K, qt, shape=(N, K))
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(dev)

    for call, M in enumerate((3, 1, 40)):  # calls in a row, different shapes and data; 40 rows = two passes over the array
        x = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(dev)
        want = lbl.lowbit_matmul(x, tiled, (N, K), qt, bias=bias, out_dtype=torch.float32)   # existing entry point
        got = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            ggml.lowbit_matmul_f32_noalloc_cuda(ctypes.c_void_p(tiled.data_ptr()), qt, N, K, ctypes.c_void_p(x.data_ptr()), M,
                                                ctypes.c_void_p(got.data_ptr()), ggml.DT_F32,
                                                ctypes.c_void_p(bias.data_ptr()), 0, _stream(dev))
        torch.cuda.synchronize()
        # same tolerance as the other matmul tests: summation-order bound 2e-6 * sum|x w|
        wd = np.abs(oracle.dequantize(packed.cpu().numpy(), (N, K), "sym_int4")).astype(np.float64)
        bound = 2e-6 * (np.abs(x.cpu().numpy()).astype(np.float64) @ wd.T) + 1e-30
        diff = np.abs(got.cpu().numpy().astype(np.float64) - want.cpu().numpy().astype(np.float64))
        assert np.isfinite(got.cpu().numpy()).all(), f"call {call}: output not written"
        assert (diff <= bound).all(), f"call {call}: max ratio {(diff / bound).max()}"
        # and against the oracle itself, so that the comparison is not two wrongs agreeing
        y64 = oracle.matmul_f64(packed.cpu().numpy(), (N, K), "sym_int4", x.cpu().numpy()) + bias.cpu().numpy().astype(np.float64)
        assert (np.abs(got.cpu().numpy() - y64) <= bound + 1e-6 * np.abs(y64)).all()


def test_noalloc_rejects_oversize(native_lib, cuda_device):
    ggml, dev = native_lib, cuda_device
    K = 4096
    M = MAX_ELEMS // K + 1
    dummy = torch.zeros(64, dtype=torch.uint8, device=dev)
    with pytest.raises(RuntimeError, match="exceeds the built-in scratch"):
        ggml.lowbit_matmul_f32_noalloc_cuda(ctypes.c_void_p(dummy.data_ptr()), 2, 64, K, ctypes.c_void_p(dummy.data_ptr()), M,
                                            ctypes.c_void_p(dummy.data_ptr()), ggml.DT_F32, None, 0, _stream(dev))


def test_noalloc_at_the_size_limit(native_lib, cuda_device):
    """m * k exactly at the limit (64 x 4096 = 262144 elements, two passes of 32 rows) is accepted and correct"""
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    ggml, dev = native_lib, cuda_device
    rng = np.random.default_rng(7)
    N, K = 128, 4096
    M = MAX_ELEMS // K
    qt = lbl.SYM_INT4
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float32)
    packed = torch.from_numpy(oracle.quantize(w, "sym_int4")).to(dev)
    tiled = lbl.ggml_q_format_convet_cpu2xpu(packed, N * K, qt, shape=(N, K))
    x = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(dev)
    got = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        ggml.lowbit_matmul_f32_noalloc_cuda(ctypes.c_void_p(tiled.data_ptr()), qt, N, K, ctypes.c_void_p(x.data_ptr()), M,
                                            ctypes.c_void_p(got.data_ptr()), ggml.DT_F32, None, 0, _stream(dev))
    torch.cuda.synchronize()
    wd = np.abs(oracle.dequantize(packed.cpu().numpy(), (N, K), "sym_int4")).astype(np.float64)
    bound = 2e-6 * (np.abs(x.cpu().numpy()).astype(np.float64) @ wd.T) + 1e-30
    y64 = oracle.matmul_f64(packed.cpu().numpy(), (N, K), "sym_int4", x.cpu().numpy())
    assert (np.abs(got.cpu().numpy() - y64) <= bound + 1e-6 * np.abs(y64)).all()
