"""FP8 key/value cache (reference transformers/kv.py:33-77 DynamicFp8Cache, models/utils.py:103-190) and the
attention entry point that reads it (models/common.py:250-306 scaled_dot_product_attention).

The cache keeps K and V as uint8 tensors of the fp16 cache's shape [batch, kv_heads, len, head_dim], grown in
blocks of FP8_KV_ALLOC_LENGTH positions through ``as_strided`` views of a larger storage -- the reference's own
scheme.  The byte format (e5m2, no scales) is defined by the kernels in csrc/kvcache.cu.
"""
import math
from typing import Any, Dict, Optional, Tuple

import torch

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.utils.common import invalidInputError

FP8_KV_ALLOC_LENGTH = 512


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def init_fp8_kv_cache(batch_size, num_heads, current_length, head_dim, device):
    """models/utils.py:103-116: storage for current_length + FP8_KV_ALLOC_LENGTH positions, views of length 0"""
    max_length = current_length + FP8_KV_ALLOC_LENGTH
    caches = []
    for _ in range(2):
        storage = torch.empty(batch_size, num_heads, max_length, head_dim, dtype=torch.uint8, device=device)
        caches.append(storage.as_strided((batch_size, num_heads, 0, head_dim), storage.stride(), storage_offset=0))
    return caches[0], caches[1]


def quantize_key_value(key, value, k_out, v_out):
    """xe_addons.quantize_key_value: fp K/V [B,H,S,D] -> e5m2 bytes written into the cache slices"""
    for src, dst in ((key, k_out), (value, v_out)):
        invalidInputError(src.shape == dst.shape and dst.dtype == torch.uint8, "cache slice and new states differ in shape")
        if src.stride(-1) != 1:
            src = src.contiguous()
        with torch.cuda.device(src.device):
            ggml.lowbit_kv_quantize(src, dst, _stream(src.device))


def append_fp8_kv_cache(k_cache, v_cache, key, value):
    """models/utils.py:119-139"""
    batch_size, num_heads, cur_length, head_dim = k_cache.shape
    new_length = cur_length + key.size(2)
    new_size = (batch_size, num_heads, new_length, head_dim)
    if k_cache.stride(1) < new_length * k_cache.size(3):
        new_k_cache, new_v_cache = init_fp8_kv_cache(batch_size, num_heads, new_length, head_dim, key.device)
        new_k_cache = new_k_cache.as_strided(new_size, new_k_cache.stride(), storage_offset=0)
        new_v_cache = new_v_cache.as_strided(new_size, new_v_cache.stride(), storage_offset=0)
        new_k_cache[:, :, :cur_length, :] = k_cache
        new_v_cache[:, :, :cur_length, :] = v_cache
    else:
        new_k_cache = k_cache.as_strided(new_size, k_cache.stride(), storage_offset=0)
        new_v_cache = v_cache.as_strided(new_size, v_cache.stride(), storage_offset=0)
    quantize_key_value(key, value, new_k_cache[:, :, cur_length:new_length, :], new_v_cache[:, :, cur_length:new_length, :])
    return new_k_cache, new_v_cache


def restore_fp8_kv_cache(k_cache, v_cache, dtype):
    """models/utils.py:181-188: xe_addons.dequantize_key_value"""
    key_states = torch.empty(k_cache.shape, device=k_cache.device, dtype=dtype)
    value_states = torch.empty(v_cache.shape, device=v_cache.device, dtype=dtype)
    with torch.cuda.device(k_cache.device):
        ggml.lowbit_kv_dequantize(k_cache, key_states, _stream(k_cache.device))
        ggml.lowbit_kv_dequantize(v_cache, value_states, _stream(v_cache.device))
    return key_states, value_states


class DynamicFp8Cache:
    """kv.py:33-77.  Same ``update`` contract as the reference class (returns the uint8 cache views of the layer).
    The reference subclasses transformers' DynamicCache; its internals changed across transformers versions
    (the reference carries work-arounds for 4.37 / 4.39 / 4.45), so this class keeps its own two lists and
    offers the methods decoders use: ``update``, ``get_seq_length``, ``__len__``, ``__getitem__``."""

    def __init__(self, num_hidden_layers: Optional[int] = None) -> None:
        self.key_cache = []
        self.value_cache = []
        self._seen_tokens = 0

    def update(self, key_states: torch.Tensor, value_states: torch.Tensor, layer_idx: int,
               cache_kwargs: Optional[Dict[str, Any]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        batch_size, num_heads, seq_len, head_dim = key_states.shape
        if layer_idx == 0:
            self._seen_tokens += seq_len
        if len(self.key_cache) <= layer_idx:
            k_cache, v_cache = init_fp8_kv_cache(batch_size, num_heads, seq_len, head_dim, device=key_states.device)
            k_cache, v_cache = append_fp8_kv_cache(k_cache, v_cache, key_states, value_states)
            self.key_cache.append(k_cache)
            self.value_cache.append(v_cache)
        else:
            k_cache, v_cache = append_fp8_kv_cache(self.key_cache[layer_idx], self.value_cache[layer_idx],
                                                   key_states, value_states)
            self.key_cache[layer_idx] = k_cache
            self.value_cache[layer_idx] = v_cache
        return self.key_cache[layer_idx], self.value_cache[layer_idx]

    def get_seq_length(self, layer_idx: int = 0) -> int:
        return 0 if len(self.key_cache) <= layer_idx else self.key_cache[layer_idx].shape[2]

    def __len__(self):
        return len(self.key_cache)

    def __getitem__(self, layer_idx):
        return self.key_cache[layer_idx], self.value_cache[layer_idx]


_sdp_workspaces = {}


def _workspace(device, need):
    ws = _sdp_workspaces.get(device)
    if ws is None or ws.numel() < need:
        ws = torch.empty(int(need * 1.5) + 256, dtype=torch.uint8, device=device)
        _sdp_workspaces[device] = ws
    return ws


def sdp_fp8(query, key, value, mask, scale, is_causal=False):
    """xe_addons.sdp_fp8(query, key, value, mask, scale): decode attention (new tokens <= 32) straight from the
    uint8 cache.  mask: None or additive [B, 1|H, S, L]."""
    B, Hq, S, D = query.shape
    Hkv, L = key.shape[1], key.shape[2]
    invalidInputError(key.dtype == torch.uint8 and value.dtype == torch.uint8, "sdp_fp8 reads the uint8 cache")
    invalidInputError(ggml.lowbit_sdp_fp8_supported(Hq, Hkv, S, D), f"unsupported attention shape (Hq {Hq}, Hkv {Hkv}, S {S}, D {D})")
    if query.stride(-1) != 1:
        query = query.contiguous()
    if mask is not None:
        if mask.stride(-1) != 1:
            mask = mask.contiguous()
        mask = mask.expand(B, Hq, S, L)            # broadcast dimensions get stride 0: nothing is copied
    out = torch.empty((B, Hq, S, D), dtype=query.dtype, device=query.device)
    ws = _workspace(query.device, ggml.lowbit_sdp_fp8_workspace_size(B, Hq, Hkv, S, L, D))
    with torch.cuda.device(query.device):
        ggml.lowbit_sdp_fp8(query, key, value, mask, is_causal, scale, out, ws, _stream(query.device))
    return out


def scaled_dot_product_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, mask: torch.Tensor = None,
                                 is_causal: bool = False, scale: float = None) -> torch.Tensor:
    """models/common.py:250-306.  uint8 key/value = the FP8 cache: new tokens <= 32 (decode) go to the fused kernel
    (sdp_fp8); longer queries (prefill) restore the cache to the query dtype and use torch's attention, as does a
    floating-point cache."""
    bsz, n_heads, seq_length, head_dim = query.shape
    _, n_kv_heads, kv_length, _ = key.shape
    scale = 1 / math.sqrt(head_dim) if scale is None else scale
    if key.dtype == torch.uint8:
        if seq_length <= 32 and ggml.lowbit_sdp_fp8_supported(n_heads, n_kv_heads, seq_length, head_dim):
            return sdp_fp8(query, key, value, mask, scale, is_causal and mask is None)
        key, value = restore_fp8_kv_cache(key, value, query.dtype)
    if n_heads != n_kv_heads:
        key = key.repeat_interleave(n_heads // n_kv_heads, dim=1)
        value = value.repeat_interleave(n_heads // n_kv_heads, dim=1)
    if mask is not None:
        mask = mask[..., :seq_length, :kv_length]
        return torch.nn.functional.scaled_dot_product_attention(query, key, value, mask, scale=scale)
    return torch.nn.functional.scaled_dot_product_attention(query, key, value, is_causal=is_causal and seq_length > 1, scale=scale)
