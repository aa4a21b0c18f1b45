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


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


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
