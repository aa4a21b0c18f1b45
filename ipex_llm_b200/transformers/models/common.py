"""Rotary position embedding applied in place to the query / key heads of the merged qkv projection: the four
``xe_addons`` entry points the reference's decoders call (models/llama.py:151-165, models/chatglm2.py:251,
models/common.py:354-367), over ``lowbit_rotary_inplaced_cuda`` (csrc/rope.cu).

``half`` = rotate_half pairing (element i with i + R/2: Llama, Mistral, Qwen, ...), ``two`` = adjacent pairing
(2i with 2i + 1: GPT-J, ChatGLM).  R, the rotary width, is 2 * len(inv_freq) or the last dimension of the tables;
head elements beyond R keep their values.  The tensors are views of the qkv output ([B, H, S, D] with D
contiguous); nothing is copied.
"""
import torch

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.utils.common import invalidInputError
from ipex_llm_b200.transformers.low_bit_linear import _on_device


def _stream(device):
    return torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else torch.cuda.current_device())


def _check_qk(query_states, key_states):
    invalidInputError(query_states.dim() == 4 and query_states.stride(-1) == 1, "query_states must be [B, H, S, D] with D contiguous")
    if key_states is not None:
        invalidInputError(key_states.dim() == 4 and key_states.stride(-1) == 1 and key_states.dtype == query_states.dtype
                          and key_states.shape[0] == query_states.shape[0] and key_states.shape[2:] == query_states.shape[2:],
                          "key_states must match query_states in batch, tokens, head_dim and dtype")


def _by_freq(inv_freq, position_ids, query_states, key_states, two):
    _check_qk(query_states, key_states)
    inv_freq = inv_freq.to(device=query_states.device, dtype=torch.float32).contiguous()
    position_ids = position_ids.to(device=query_states.device, dtype=torch.int64)
    if position_ids.dim() == 1:
        position_ids = position_ids[None, :]
    invalidInputError(position_ids.shape[-1] == query_states.shape[2] and position_ids.shape[0] in (1, query_states.shape[0]),
                      "position_ids must be [B | 1, S]")
    rotary_dim = 2 * inv_freq.numel()
    invalidInputError(rotary_dim <= query_states.shape[-1], "rotary width exceeds head_dim")
    with torch.cuda.device(query_states.device):
        ggml.lowbit_rotary_inplaced(query_states, key_states, rotary_dim, two, inv_freq, position_ids, None, None, False,
                                    _stream(query_states.device))


def rotary_half_inplaced(inv_freq, position_ids, query_states, key_states):
    """xe_addons.rotary_half_inplaced"""
    _by_freq(inv_freq, position_ids, query_states, key_states, False)


def rotary_two_inplaced(inv_freq, position_ids, query_states, key_states):
    """xe_addons.rotary_two_inplaced"""
    _by_freq(inv_freq, position_ids, query_states, key_states, True)


def _tables(cos, sin, query_states):
    """[S, R] / [1 | B, S, R] / [1 | B, 1, S, R] -> [1 | B, S, R] with R contiguous"""
    out = []
    for t in (cos, sin):
        if t.dim() == 4:
            t = t.squeeze(1)
        if t.dim() == 2:
            t = t[None]
        if t.stride(-1) != 1:
            t = t.contiguous()
        out.append(t)
    cos, sin = out
    invalidInputError(cos.shape == sin.shape and cos.dtype == sin.dtype and cos.stride() == sin.stride()
                      and cos.shape[0] in (1, query_states.shape[0]) and cos.shape[1] == query_states.shape[2],
                      "cos / sin must be [B | 1, S, R]")
    return cos, sin


def rotary_half_with_cache_inplaced(query_states, key_states, cos, sin):
    """models/common.py:362-367: cos / sin in the rotate_half layout (entry i + R/2 repeats entry i)"""
    _check_qk(query_states, key_states)
    cos, sin = _tables(cos, sin, query_states)
    with torch.cuda.device(query_states.device):
        ggml.lowbit_rotary_inplaced(query_states, key_states, cos.shape[-1], False, None, None, cos, sin, False,
                                    _stream(query_states.device))


def rotary_two_with_cache_inplaced(query_states, key_states, cos, sin, half_layout):
    """models/common.py:354-359: adjacent pairing; the tables are either in the rotate_half layout (half_layout,
    entry i for pair i) or repeat every entry twice (entry 2i for pair i)"""
    _check_qk(query_states, key_states)
    cos, sin = _tables(cos, sin, query_states)
    with torch.cuda.device(query_states.device):
        ggml.lowbit_rotary_inplaced(query_states, key_states, cos.shape[-1], True, None, None, cos, sin, not half_layout,
                                    _stream(query_states.device))


# ------------------------------------------------------------------ callers of the low-bit linears
# (reference models/common.py: merge_linear / merge_qkv_base :23-53, fuse_mlp_base / mlp_silu_forward /
# mlp_gelu_forward :146-171, rms_norm_forward :187-205, scaled_dot_product_attention :250-306)
from typing import List                                                   # noqa: E402

from ipex_llm_b200.transformers.kv import scaled_dot_product_attention   # noqa: E402,F401  (re-export, reference location)


def merge_linear(linears: List[torch.nn.Linear]) -> torch.nn.Linear:
    """:23-37: one Linear whose weight is the row concatenation of the given ones (before quantisation; block
    records never straddle rows, so quantising the merged weight gives the concatenation of the parts' bytes)"""
    if not hasattr(linears[0], "weight"):
        return None
    new_weight = torch.cat([linear.weight.data for linear in linears], dim=0)
    has_bias = linears[0].bias is not None
    with torch.device("meta"):
        new_linear = torch.nn.Linear(1, 1, bias=has_bias)
    if has_bias:
        new_linear.bias = torch.nn.Parameter(torch.cat([linear.bias.data for linear in linears], dim=0), requires_grad=False)
    new_linear.weight = torch.nn.Parameter(new_weight, requires_grad=False)
    new_linear.in_features = new_weight.size(1)
    new_linear.out_features = new_weight.size(0)
    return new_linear


def merge_qkv_base(module: torch.nn.Module, attention_class):
    """:40-53: q_proj, k_proj, v_proj -> qkv_proj on every attention module of the given class (or class name)"""
    if (isinstance(attention_class, str) and module.__class__.__name__ == attention_class
            or not isinstance(attention_class, str) and isinstance(module, attention_class)):
        if not all(hasattr(module, n) for n in ("q_proj", "k_proj", "v_proj")):
            return
        qkv_proj = merge_linear([module.q_proj, module.k_proj, module.v_proj])
        if qkv_proj is not None:
            module.qkv_proj = qkv_proj
            del module.q_proj, module.k_proj, module.v_proj


PREFILL_MLP_MAX_SCRATCH = 256 << 20      # bytes of fp32 scratch (2 x tokens x intermediate x 4) the prefill MLP may ask for


def fuse_mlp_base(module: torch.nn.Module, act: int, x: torch.Tensor):
    """:146-161: gate and up projections + activation * multiply in one kernel at decode sizes
    (xe_linear.mlp_forward_xpu -> lowbit_mlp_cuda), then down_proj"""
    from ipex_llm_b200.transformers.models.utils import mlp_fusion_check
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear, mlp_forward
    gate, up = module.gate_proj, module.up_proj
    qtype = getattr(gate, "qtype", None)
    if (isinstance(gate, LowBitLinear) and isinstance(up, LowBitLinear) and gate.bias is None and up.bias is None
            and gate.mp_group is None and getattr(up, "qtype", None) == qtype and mlp_fusion_check(x, qtype, module.training)):
        x_2d = x.contiguous().view(-1, x.size(-1))
        hidden = mlp_forward(x_2d, gate.weight.data, up.weight.data, x_2d.size(0), x_2d.size(1), gate.out_len, act, qtype)
        return module.down_proj(hidden.view(x.shape[:-1] + (gate.out_len,)))
    if (isinstance(gate, LowBitLinear) and isinstance(up, LowBitLinear) and gate.bias is None and up.bias is None
            and gate.mp_group is None and getattr(up, "qtype", None) == qtype and x.device.type == "cuda"
            and x.dtype in (torch.float16, torch.bfloat16) and not module.training and not x.requires_grad
            and not gate.act_order and gate.weight.data.device.type == "cuda"
            and 8 * (x.numel() // x.size(-1)) * gate.out_len <= PREFILL_MLP_MAX_SCRATCH):
        # prefill-sized batches: the same entry point runs the two matmuls into fp32 scratch and one activation kernel --
        # three launches from one host call, where the module-by-module statement below is five ops and as many allocations
        # (the 512-token prefill of a 7B model is bound by the host's launch rate between its GEMMs)
        x_2d = x.contiguous().view(-1, x.size(-1))
        hidden = mlp_forward(x_2d, gate.weight.data, up.weight.data, x_2d.size(0), x_2d.size(1), gate.out_len, act, qtype)
        return module.down_proj(hidden.view(x.shape[:-1] + (gate.out_len,)))
    return module.down_proj(module.act_fn(module.gate_proj(x)) * module.up_proj(x))


def mlp_silu_forward(self, x: torch.Tensor):
    """:164-166"""
    from ipex_llm_b200.transformers.models.utils import SILU
    return fuse_mlp_base(self, SILU, x)


def mlp_gelu_forward(self, x: torch.Tensor):
    """:169-171"""
    from ipex_llm_b200.transformers.models.utils import GELU
    return fuse_mlp_base(self, GELU, x)


_RMS_DTYPES = (torch.float, torch.half, torch.bfloat16)


def rms_norm_forward(self, hidden_states: torch.Tensor):
    """:187-205: xe_addons.rms_norm -> lowbit_rms_norm_cuda on the GPU, the torch statement elsewhere"""
    weight = self.weight
    eps = getattr(self, "variance_epsilon", None)
    if eps is None:
        eps = getattr(self, "epsilon", None)
    if eps is None:
        eps = self.eps
    if hidden_states.device.type == "cuda" and hidden_states.dtype in _RMS_DTYPES and weight.dtype in _RMS_DTYPES:
        x_2d = hidden_states.reshape(-1, hidden_states.size(-1))
        if x_2d.stride(-1) != 1:
            x_2d = x_2d.contiguous()
        out_dtype = x_2d.dtype if weight.dtype == x_2d.dtype else torch.promote_types(weight.dtype, x_2d.dtype)
        out = torch.empty(x_2d.shape, dtype=out_dtype, device=x_2d.device)
        with _on_device(x_2d.device):
            ggml.lowbit_rms_norm(x_2d, weight, eps, out, _stream(x_2d.device))
        return out.reshape(hidden_states.shape)
    input_dtype = hidden_states.dtype
    variance = hidden_states.to(torch.float32).pow(2).mean(dim=-1, keepdim=True)
    hidden_states = hidden_states.to(torch.float32) * torch.rsqrt(variance + eps)
    return weight * hidden_states.to(input_dtype)
