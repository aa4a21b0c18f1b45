"""FP8 key/value cache (reference transformers/kv.py:33-77 DynamicFp8Cache, models/utils.py:103-190) and the
attention entry point that reads it (models/common.py:250-306 scaled_dot_product_attention).

The cache keeps K and V as uint8 tensors of the fp16 cache's shape [batch, kv_heads, len, head_dim], grown in
blocks of FP8_KV_ALLOC_LENGTH positions through ``as_strided`` views of a larger storage -- the reference's own
scheme.  The byte format (e5m2, no scales) is defined by the kernels in csrc/kvcache.cu.
"""
import math
from typing import Any, Dict, Optional, Tuple

import os

import torch
from transformers.cache_utils import Cache, DynamicLayer

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.utils.common import invalidInputError

FP8_KV_ALLOC_LENGTH = 512
KV_ALLOC_BLOCK_LENGTH = int(os.environ.get("KV_CACHE_ALLOC_BLOCK_LENGTH", 256))


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


class Fp8KVLayer(DynamicLayer):
    """One layer of the FP8 cache: uint8 views [B, H, len, D] over a storage that is allocated FP8_KV_ALLOC_LENGTH
    positions ahead (init_fp8_kv_cache / append_fp8_kv_cache, the reference's scheme).  ``reserve`` = positions to
    allocate at once (a generate() call knows prompt + new tokens); ``storage`` = (k, v) uint8 [B, H, capacity, D]
    tensors owned by somebody else (the recorded decode step keeps its caches at fixed addresses)."""

    def __init__(self, reserve: int = 0, storage=None):
        super().__init__()
        self.reserve = int(reserve)
        self.storage = storage

    def lazy_initialization(self, key_states, value_states):
        self.dtype, self.device = key_states.dtype, key_states.device
        batch_size, num_heads, seq_len, head_dim = key_states.shape
        if self.storage is not None:
            ks, vs = self.storage
            invalidInputError(ks.shape[0] == batch_size and ks.shape[1] == num_heads and ks.shape[3] == head_dim,
                              "bound cache storage does not match the key states")
            self.keys = ks.as_strided((batch_size, num_heads, 0, head_dim), ks.stride(), storage_offset=0)
            self.values = vs.as_strided((batch_size, num_heads, 0, head_dim), vs.stride(), storage_offset=0)
        else:
            self.keys, self.values = init_fp8_kv_cache(batch_size, num_heads, max(seq_len, self.reserve - FP8_KV_ALLOC_LENGTH),
                                                       head_dim, key_states.device)
        self.is_initialized = True

    def update(self, key_states, value_states, *args, **kwargs):
        if not self.is_initialized:
            self.lazy_initialization(key_states, value_states)
        self.keys, self.values = append_fp8_kv_cache(self.keys, self.values, key_states, value_states)
        return self.keys, self.values

    def rope_append(self, qkv, n_heads, n_kv_heads, head_dim, cos, sin):
        """The first call of a sequence: rotary on the q / k heads of the merged projection's output ``qkv`` [B, S, (Hq + 2 Hkv) D]
        (in place) and the e5m2 cache write of its S tokens in ONE launch -- rotary_half_with_cache_inplaced + update()
        (= quantize_key_value twice) are three.  cos / sin: [S, D] tables of positions 0..S-1.  Returns the layer's
        (keys, values) views, or None when the layer already holds tokens (the separate ops then do the work)."""
        B, S = qkv.shape[0], qkv.shape[1]
        if self.is_initialized and self.keys.shape[2] != 0:
            return None
        if not self.is_initialized:
            k_view = qkv.view(B, S, n_heads + 2 * n_kv_heads, head_dim).transpose(1, 2)[:, n_heads:n_heads + n_kv_heads]
            self.lazy_initialization(k_view, k_view)
        if self.keys.stride(1) < S * head_dim:                    # storage shorter than the prompt (no reserve given)
            if self.storage is not None:
                return None
            self.keys, self.values = init_fp8_kv_cache(B, n_kv_heads, S, head_dim, qkv.device)
        cap = self.keys.stride(1) // head_dim
        k_full = self.keys.as_strided((B, n_kv_heads, cap, head_dim), self.keys.stride(), storage_offset=0)
        v_full = self.values.as_strided((B, n_kv_heads, cap, head_dim), self.values.stride(), storage_offset=0)
        with torch.cuda.device(qkv.device):
            ggml.lowbit_rope_kv_append(qkv, n_heads, n_kv_heads, head_dim, cos, sin, None, 0, k_full, v_full, _stream(qkv.device))
        self.set_length(S)
        return self.keys, self.values

    def set_length(self, length: int):
        """the views after ``length`` positions were written behind the host's back (the recorded decode step)"""
        b, h, _, d = self.keys.shape
        self.keys = self.keys.as_strided((b, h, length, d), self.keys.stride(), storage_offset=0)
        self.values = self.values.as_strided((b, h, length, d), self.values.stride(), storage_offset=0)


class _LayeredCache(Cache):
    """transformers' ``Cache`` (a list of per-layer objects) with the ``key_cache`` / ``value_cache`` lists the
    reference's cache classes expose"""
    LAYER = None

    def __init__(self, num_hidden_layers: Optional[int] = None, reserve: int = 0) -> None:
        super().__init__(layers=[])
        self.reserve = reserve

    def new_layer(self):
        return self.LAYER(self.reserve)

    def update(self, key_states, value_states, layer_idx, *args, **kwargs):
        while len(self.layers) <= layer_idx:
            self.layers.append(self.new_layer())
        return self.layers[layer_idx].update(key_states, value_states)

    @property
    def key_cache(self):
        return [l.keys for l in self.layers]

    @property
    def value_cache(self):
        return [l.values for l in self.layers]

    def __getitem__(self, layer_idx):
        return self.layers[layer_idx].keys, self.layers[layer_idx].values

    @classmethod
    def from_legacy_cache(cls, past_key_values=None, **kwargs):
        """an empty or missing cache -> a new one of this class; a filled one is re-stored layer by layer
        (the reference's from_legacy_cache, kv.py / transformers DynamicCache)"""
        cache = cls(**kwargs)
        if past_key_values is None:
            return cache
        layers = getattr(past_key_values, "layers", None)
        pairs = [(l.keys, l.values) for l in layers if getattr(l, "is_initialized", False) and l.get_seq_length() > 0] \
            if layers is not None else list(past_key_values)
        for i, (k, v) in enumerate(pairs):
            cache.update(k, v, i)
        return cache


class DynamicFp8Cache(_LayeredCache):
    """kv.py:33-77: K and V as e5m2 bytes; ``update`` returns the uint8 views of the layer."""
    LAYER = Fp8KVLayer

    def rope_append(self, layer_idx, qkv, n_heads, n_kv_heads, head_dim, cos, sin):
        """see Fp8KVLayer.rope_append"""
        while len(self.layers) <= layer_idx:
            self.layers.append(self.new_layer())
        return self.layers[layer_idx].rope_append(qkv, n_heads, n_kv_heads, head_dim, cos, sin)


class NormalKVLayer(DynamicLayer):
    """One layer of the floating-point cache, grown in blocks of KV_ALLOC_BLOCK_LENGTH positions instead of
    torch.cat per token (kv.py:80-140, models/utils.py init_kv_cache / extend_kv_cache / append_kv_cache)."""

    def __init__(self, reserve: int = 0):
        super().__init__()
        self.reserve = int(reserve)

    def lazy_initialization(self, key_states, value_states):
        self.dtype, self.device = key_states.dtype, key_states.device
        self.keys = self.values = None
        self.is_initialized = True

    def _alloc(self, like, length, capacity):
        b, h, _, d = like.shape
        storage = torch.empty(b, h, capacity, d, dtype=like.dtype, device=like.device)
        return storage.as_strided((b, h, length, d), storage.stride(), storage_offset=0)

    def update(self, key_states, value_states, *args, **kwargs):
        if not self.is_initialized:
            self.lazy_initialization(key_states, value_states)
        cur = 0 if self.keys is None else self.keys.size(2)
        new_len = cur + key_states.size(2)
        if self.keys is None or self.keys.stride(1) < new_len * self.keys.size(3):
            cap = max(new_len, self.reserve) + KV_ALLOC_BLOCK_LENGTH
            nk, nv = self._alloc(key_states, new_len, cap), self._alloc(value_states, new_len, cap)
            if cur:
                nk[:, :, :cur, :] = self.keys
                nv[:, :, :cur, :] = self.values
            self.keys, self.values = nk, nv
        else:
            b, h, _, d = self.keys.shape
            self.keys = self.keys.as_strided((b, h, new_len, d), self.keys.stride(), storage_offset=0)
            self.values = self.values.as_strided((b, h, new_len, d), self.values.stride(), storage_offset=0)
        self.keys[:, :, cur:new_len, :] = key_states
        self.values[:, :, cur:new_len, :] = value_states
        return self.keys, self.values

    def get_seq_length(self) -> int:
        return 0 if (not self.is_initialized or self.keys is None) else self.keys.shape[-2]


class DynamicNormalCache(_LayeredCache):
    """kv.py:80-140"""
    LAYER = NormalKVLayer


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


def sdp_fp8_prefill(query, key, value, mask, scale, is_causal):
    """xe_addons.sdp_fp8_causal / sdp_fp8_non_causal: attention of many new tokens straight from the uint8 cache"""
    B, Hq, S, D = query.shape
    if query.stride(-1) != 1:
        query = query.contiguous()
    if mask is not None:
        if mask.stride(-1) != 1:
            mask = mask.contiguous()
        mask = mask.expand(B, Hq, S, key.shape[2])
    # [B, Hq, S, D] as a view of a [B, S, Hq, D] buffer: the transpose(1, 2).contiguous() that follows attention in the model
    # forwards (models/llama.py:198) finds its layout already there
    out = torch.empty((B, S, Hq, D), dtype=query.dtype, device=query.device).transpose(1, 2)
    with torch.cuda.device(query.device):
        ggml.lowbit_sdp_fp8_prefill(query, key, value, mask, is_causal, scale, out, _stream(query.device))
    return out


def sdp_fp8_causal(query, key, value, mask, scale):
    """models/common.py:273"""
    return sdp_fp8_prefill(query, key, value, mask, scale, mask is None)


def sdp_fp8_non_causal(query, key, value, mask, scale):
    """models/common.py:284"""
    return sdp_fp8_prefill(query, key, value, mask, scale, False)


def _prefill_kernel_ok(query, key, value, mask):
    def aligned(t, unit):
        return t.data_ptr() % 16 == 0 and all(s % unit == 0 for s in t.stride()[:3])
    return (ggml.lowbit_sdp_fp8_prefill_supported(query.shape[1], key.shape[1], query.shape[3], query.dtype)
            and os.environ.get("IPEX_LLM_B200_SDP_PREFILL", "1") != "0"
            and aligned(key, 16) and aligned(value, 16) and key.stride(-1) == 1 and value.stride(-1) == 1
            and (query.stride(-1) != 1 or aligned(query, 8)) and key.shape[2] >= query.shape[2])


def _additive_mask(mask, dtype):
    """transformers >= 5 hands attention a boolean mask (True = attend); the kernels take an additive one"""
    if mask is not None and mask.dtype == torch.bool:
        return torch.zeros(mask.shape, dtype=dtype, device=mask.device).masked_fill_(~mask, float("-inf"))
    return mask


def scaled_dot_product_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, mask: torch.Tensor = None,
                                 is_causal: bool = False, scale: float = None) -> torch.Tensor:
    """models/common.py:250-306.  uint8 key/value = the FP8 cache: new tokens <= 32 (decode) go to the fused kernel
    (sdp_fp8); longer queries (prefill) to the many-token kernel (sdp_fp8_causal / sdp_fp8_non_causal; fp16 / bf16, head_dim
    128); what neither serves restores the cache to the query dtype and uses torch's attention, as does a floating-point
    cache."""
    bsz, n_heads, seq_length, head_dim = query.shape
    _, n_kv_heads, kv_length, _ = key.shape
    scale = 1 / math.sqrt(head_dim) if scale is None else scale
    if key.dtype == torch.uint8:
        if seq_length <= 32 and ggml.lowbit_sdp_fp8_supported(n_heads, n_kv_heads, seq_length, head_dim):
            mask = _additive_mask(mask, query.dtype if query.dtype != torch.bfloat16 else torch.float32)
            return sdp_fp8(query, key, value, mask, scale, is_causal and mask is None)
        if _prefill_kernel_ok(query, key, value, mask):
            mask = _additive_mask(mask, query.dtype)
            if mask is not None:
                mask = mask[..., :seq_length, :kv_length]
            return sdp_fp8_prefill(query, key, value, mask, scale, is_causal and mask is None)
        key, value = restore_fp8_kv_cache(key, value, query.dtype)      # other head sizes / fp32 queries: torch on the restored cache
    if n_heads != n_kv_heads:
        key = key.repeat_interleave(n_heads // n_kv_heads, dim=1)
        value = value.repeat_interleave(n_heads // n_kv_heads, dim=1)
    if mask is not None:
        mask = mask[..., :seq_length, :kv_length]
    kwargs = dict(attn_mask=mask, scale=scale) if mask is not None else dict(is_causal=is_causal and seq_length > 1, scale=scale)
    if query.dtype == torch.float32 and query.device.type == "cuda":
        # an fp32 model asks for fp32 arithmetic (the CPU reference's contract): the plain softmax(QK^T)V statement,
        # not a fused kernel that may feed the tensor cores reduced-precision operands
        from torch.nn.attention import SDPBackend, sdpa_kernel
        with sdpa_kernel(SDPBackend.MATH):
            return torch.nn.functional.scaled_dot_product_attention(query, key, value, **kwargs)
    return torch.nn.functional.scaled_dot_product_attention(query, key, value, **kwargs)
