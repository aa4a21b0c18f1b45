"""Gates that decide which fused operator a decoder forward may call -- the B200 readings of the reference's
transformers/models/utils.py (MU below): ``mlp_fusion_check`` MU:289-302, ``use_quantize_kv_cache`` MU:77-99,
``use_fuse_moe`` MU:399-404, activation ids MU:30-31."""
import os
import warnings

import torch

from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype, B200_SUPPORTED_QTYPE_IDS

SILU = 0
GELU = 1


def mlp_fusion_check(x, qtype, training):
    """MU:289-302: one token, on the accelerator, a qtype the fused gate/up kernel serves, inference.
    The reference lists the qtypes its XPU kernel has (sym_int4, fp8_e5m2, fp4, iq2_xxs, fp6, woq_int4); the B200
    GEMV streams two matrices for every in-scope qtype (lowbit_mlp_cuda), and for up to 8 tokens."""
    if x.numel() // x.size(-1) > 8:
        return False
    if x.device.type != "cuda":
        return False
    if qtype not in B200_SUPPORTED_QTYPE_IDS:
        return False
    if training or x.requires_grad:
        return False
    return True


def use_quantize_kv_cache(linear: torch.nn.Module, x: torch.Tensor, num_heads: int = 1, num_kv_heads: int = 1) -> bool:
    """MU:77-99.  Environment switches first (same names); otherwise the reference decides by XPU model.  On B200
    the FP8 cache is the default whenever the weights are low-bit, the model computes in half precision and a fused
    attention kernel exists for it: decode attention is the largest per-token HBM stream after the weights."""
    if os.environ.get("BIGDL_QUANTIZE_KV_CACHE", None) is not None:
        warnings.warn("`BIGDL_QUANTIZE_KV_CACHE` is deprecated and will be removed in future releases. "
                      "Please use `IPEX_LLM_QUANTIZE_KV_CACHE` instead.")
        return os.environ["BIGDL_QUANTIZE_KV_CACHE"] == "1"
    if os.environ.get("IPEX_LLM_QUANTIZE_KV_CACHE", None) is not None:
        return os.environ["IPEX_LLM_QUANTIZE_KV_CACHE"] == "1"
    if os.environ.get("IPEX_LLM_LOW_MEM", None) is not None:
        return os.environ["IPEX_LLM_LOW_MEM"] == "1"
    weight = getattr(linear, "weight", None)
    if weight is None or not getattr(weight, "quantized", False):       # unquantized
        return False
    return x.device.type == "cuda" and x.dtype in (torch.float16, torch.bfloat16)


def use_fuse_moe(hidden_states: torch.Tensor, qtype: int) -> bool:
    """MU:399-404: one token on the accelerator and a qtype the fused expert kernel serves"""
    return (hidden_states.device.type == "cuda" and hidden_states.numel() == hidden_states.size(-1)
            and qtype in B200_SUPPORTED_QTYPE_IDS)
