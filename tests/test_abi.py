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
