"""C-ABI checks that need no GPU: the library builds, loads, exports every symbol the public
header declares, and fails LOUDLY (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_lowbit.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?(?:void|int|size_t|char\s*\*)\s*\*?\s*(\w+)\s*\(", text, flags=re.M)
    return sorted(set(names))


def test_header_symbols_exported(native_lib):
    lib = ctypes.CDLL(native_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 20, names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200_lowbit.h but not exported"
    for must in ("ggml_quantize_tensor", "ggml_type_size", "ggml_qk_size", "ggml_dequantize_q4_0", "ggml_dequantize",
                 "ggml_q_format_convet_cpu2xpu", "ggml_q_format_convet_xpu2cpu",
                 "ggml_compute_forward_mul_mat_q_fp32", "llama_init_backend", "lowbit_matmul_cuda"):
        assert must in names


def test_format_tables(native_lib):
    g = native_lib
    from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype as Q
    assert [g.ggml_qk_size(Q[n]) for n in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp8", "fp6", "woq_int4")] == [64] * 8
    assert [g.ggml_type_size(Q[n]) for n in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8", "fp6", "woq_int4")] == \
        [34, 36, 34, 66, 66, 50, 34]
    # out-of-scope qtypes report 0 (callers must not size buffers from them)
    assert g.ggml_qk_size(Q["sym_int8"]) == 0 and g.ggml_type_size(Q["q4_k"]) == 0
    import oracle
    for n in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"):
        assert g.ggml_type_size(Q[n]) == oracle.type_size(n) and g.ggml_qk_size(Q[n]) == oracle.qk_size(n)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(native_lib):
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    w = torch.randn(16, 64)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        lbl.ggml_convert_qtype(w, lbl.SYM_INT4, device="cpu")
    with pytest.raises(RuntimeError, match="no CUDA device"):
        lbl.ggml_convert_fp32(torch.zeros(16 * 34, dtype=torch.uint8), (16, 64), 16 * 64, lbl.SYM_INT4)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        lbl.ggml_matmul_src1_x_src0_t(torch.zeros(16 * 34, dtype=torch.uint8), torch.zeros(1, 64), (16, 64), lbl.SYM_INT4)
    # shape-only paths work without a device (load_low_bit on meta, low_bit_linear.py:150)
    assert lbl.ggml_convert_qtype(w, lbl.SYM_INT4, device="meta").shape == (16 * 34,)
    assert lbl.ggml_convert_qtype(w, lbl.NF4, convert_shape_only=True).dtype == torch.bfloat16


def test_python_surface(native_lib):
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear, FP4Params, SYM_INT4, reshape_lm_head_input
    lin = LowBitLinear(128, 32, SYM_INT4, bias=True)
    assert isinstance(lin.weight, FP4Params) and not lin.weight.quantized
    assert (lin.in_len, lin.out_len, lin.weight_shape, lin.qtype) == (128, 32, (32, 128), SYM_INT4)
    x = torch.zeros(2, 0, 128)
    assert lin(x).shape == (2, 0, 32)                         # empty-activation early-out
    assert reshape_lm_head_input(torch.zeros(2, 11, 8)).shape == (2, 1, 8)
    assert reshape_lm_head_input(torch.zeros(2, 10, 8)).shape == (2, 10, 8)
    with pytest.raises(RuntimeError):
        LowBitLinear(128, 32, 8)                               # sym_int8: out of scope, loud
    # meta -> meta sizes the packed buffer only
    m = FP4Params(torch.empty(32, 128, device="meta"), qtype=SYM_INT4).to("meta")
    assert m.quantized and m.data.shape == (32 * 2 * 34,) and tuple(m._shape) == (32, 128)


def test_argument_checks_of_the_newer_entry_points(native_lib):
    """bad arguments are refused before anything touches the device (so this runs without one): rotary embedding,
    FP8 KV cache / decode attention, the tensor-parallel collective"""
    lib = ctypes.CDLL(native_lib.LIB_PATH)
    lib.b200_lowbit_last_error_string.restype = ctypes.c_char_p
    st3 = (ctypes.c_int64 * 3)(1, 1, 1)
    st2 = (ctypes.c_int64 * 2)(0, 1)
    buf = ctypes.create_string_buffer(256)
    p = ctypes.cast(buf, ctypes.c_void_p)

    def refused(rc, needle):
        msg = lib.b200_lowbit_last_error_string().decode()
        lib.b200_lowbit_clear_error()
        return rc != 0 and needle in msg

    # rotary width larger than head_dim / odd width / neither inv_freq nor tables
    lib.lowbit_rotary_inplaced_cuda.restype = ctypes.c_int
    args = lambda R, inv: (p, st3, 2, None, None, 0, 1, 1, 1, 64, R, 0, inv, p if inv else None, st2 if inv else None,
                           None, None, 0, None, 0, None)
    assert refused(lib.lowbit_rotary_inplaced_cuda(*args(80, p)), "lowbit_rotary_inplaced_cuda")
    assert refused(lib.lowbit_rotary_inplaced_cuda(*args(31, p)), "lowbit_rotary_inplaced_cuda")
    assert refused(lib.lowbit_rotary_inplaced_cuda(*args(32, None)), "lowbit_rotary_inplaced_cuda")
    # decode attention: unsupported head size, too many new tokens, workspace
    lib.lowbit_sdp_fp8_supported.restype = ctypes.c_int
    assert lib.lowbit_sdp_fp8_supported(32, 8, 1, 128) == 1 and lib.lowbit_sdp_fp8_supported(32, 8, 33, 128) == 0
    assert lib.lowbit_sdp_fp8_supported(32, 8, 1, 80) == 0 and lib.lowbit_sdp_fp8_supported(32, 6, 1, 128) == 0
    lib.lowbit_sdp_fp8_workspace_size.restype = ctypes.c_size_t
    assert lib.lowbit_sdp_fp8_workspace_size(1, 32, 8, 1, 4096, 80) == 0
    # the collective: sizes and ranks
    lib.lowbit_comm_max_bytes.restype = ctypes.c_size_t
    limit = lib.lowbit_comm_max_bytes()
    assert limit == 256 * 1024
    out = ctypes.c_void_p()
    assert refused(lib.lowbit_comm_create(ctypes.c_size_t(0), ctypes.byref(out), buf), "lowbit_comm_create")
    assert refused(lib.lowbit_comm_create(ctypes.c_size_t(limit + 1), ctypes.byref(out), buf), "lowbit_comm_create")
    peers = (ctypes.c_void_p * 2)(p.value, p.value)
    lib.lowbit_allreduce_sum_cuda.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                              ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]
    assert refused(lib.lowbit_allreduce_sum_cuda(peers, 2, 2, 4096, p, p, 1024, 1, None), "lowbit_allreduce_sum_cuda")    # rank out of range
    assert refused(lib.lowbit_allreduce_sum_cuda(peers, 0, 9, 4096, p, p, 1024, 1, None), "lowbit_allreduce_sum_cuda")    # world > 8
    assert refused(lib.lowbit_allreduce_sum_cuda(peers, 0, 2, 4096, p, p, 1001, 1, None), "lowbit_allreduce_sum_cuda")    # 2002 bytes: not x16
    assert refused(lib.lowbit_allreduce_sum_cuda(peers, 0, 2, 4096, p, p, 4096, 1, None), "lowbit_allreduce_sum_cuda")    # larger than the communicator
    lib.lowbit_matmul_allreduce_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                                 ctypes.c_int64, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                                 ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                                 ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    assert refused(lib.lowbit_matmul_allreduce_cuda(p, 2, 4096, 4096, p, 1, 33, p, 1, None, 0, peers, 0, 2, 1 << 18, None, 0, None),
                   "lowbit_matmul_allreduce_cuda")                                                                         # M > 32
    assert refused(lib.lowbit_matmul_allreduce_cuda(p, 2, 4100, 4096, p, 1, 1, p, 1, None, 0, peers, 0, 2, 1 << 18, None, 0, None),
                   "lowbit_matmul_allreduce_cuda")                                                                         # rows not x64
