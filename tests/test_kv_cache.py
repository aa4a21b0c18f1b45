"""FP8 key/value cache and decode attention (reference transformers/kv.py:33-77,
models/utils.py:103-190, models/common.py:250-306).

The byte format is build-defined (e5m2, round to nearest even): checked against torch.float8_e5m2.  The
attention kernel is checked against softmax(q k^T * scale + mask) v computed in float64 from the SAME
dequantised cache (so only the kernel's fp32 arithmetic and the output rounding are in the tolerance).
"""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kv(native_lib, cuda_device):
    from ipex_llm_b200.transformers import kv as kvmod
    return kvmod


def ref_attention(q, k_u8, v_u8, mask, scale, causal):
    """float64 attention from the e5m2 bytes; GQA by head repetition"""
    k = k_u8.view(torch.float8_e5m2).double()
    v = v_u8.view(torch.float8_e5m2).double()
    B, Hq, S, D = q.shape
    Hkv, L = k.shape[1], k.shape[2]
    k = k.repeat_interleave(Hq // Hkv, dim=1)
    v = v.repeat_interleave(Hq // Hkv, dim=1)
    s = q.double() @ k.transpose(-1, -2) * scale
    if mask is not None:
        s = s + mask.double()
    if causal:
        i = torch.arange(S, device=q.device)[:, None]
        j = torch.arange(L, device=q.device)[None, :]
        s = s.masked_fill(j > (L - S + i), float("-inf"))
    return torch.softmax(s, dim=-1) @ v


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_quantize_key_value_is_e5m2_rne(kv, cuda_device, dtype):
    gen = torch.Generator(device=cuda_device).manual_seed(1)
    key = (torch.randn(2, 3, 7, 128, device=cuda_device, generator=gen) * 3).to(dtype)
    val = (torch.randn(2, 3, 7, 128, device=cuda_device, generator=gen) * 0.01).to(dtype)
    key[0, 0, 0, :4] = torch.tensor([0.0, -0.0, 57344.0, -1e-8], device=cuda_device).to(dtype)
    k_cache, v_cache = kv.init_fp8_kv_cache(2, 3, 7, 128, cuda_device)
    k_cache, v_cache = kv.append_fp8_kv_cache(k_cache, v_cache, key.transpose(1, 2).contiguous().transpose(1, 2), val)
    assert k_cache.shape == (2, 3, 7, 128) and k_cache.dtype == torch.uint8 and k_cache.stride(1) == (7 + kv.FP8_KV_ALLOC_LENGTH) * 128
    assert torch.equal(k_cache, key.to(torch.float8_e5m2).view(torch.uint8))
    assert torch.equal(v_cache, val.to(torch.float8_e5m2).view(torch.uint8))
    k_back, v_back = kv.restore_fp8_kv_cache(k_cache, v_cache, dtype)
    assert torch.equal(k_back, k_cache.view(torch.float8_e5m2).to(dtype))
    assert torch.equal(v_back, v_cache.view(torch.float8_e5m2).to(dtype))
    # values beyond the e5m2 range saturate instead of becoming inf
    big = torch.full((1, 1, 1, 64), 1e6, device=cuda_device, dtype=torch.float32)
    kc, vc = kv.init_fp8_kv_cache(1, 1, 1, 64, cuda_device)
    kc, vc = kv.append_fp8_kv_cache(kc, vc, big, -big)
    assert float(kc.view(torch.float8_e5m2).float().max()) == 57344.0 and float(vc.view(torch.float8_e5m2).float().min()) == -57344.0


def test_dynamic_fp8_cache_grows_like_the_reference(kv, cuda_device):
    cache = kv.DynamicFp8Cache()
    gen = torch.Generator(device=cuda_device).manual_seed(2)
    chunks = [torch.randn(1, 4, n, 64, device=cuda_device, generator=gen).half() for n in (500, 1, 1, 520, 1)]
    for layer in range(2):
        for c in chunks[:1]:
            cache.update(c, c * 0.5, layer)
    storage_ptr = cache.key_cache[0].data_ptr()
    for c in chunks[1:3]:
        for layer in range(2):
            k, v = cache.update(c, c * 0.5, layer)
    assert cache.key_cache[0].data_ptr() == storage_ptr and k.shape[2] == 502       # appended in place
    for c in chunks[3:]:
        for layer in range(2):
            k, v = cache.update(c, c * 0.5, layer)
    assert k.shape == (1, 4, 1023, 64) and cache.get_seq_length() == 1023 and len(cache) == 2
    assert cache.key_cache[0].data_ptr() != storage_ptr        # 500 + 512 allocated positions used up: re-allocated, copied
    full = torch.cat(chunks, dim=2)
    assert torch.equal(cache.key_cache[1], full.to(torch.float8_e5m2).view(torch.uint8))
    assert torch.equal(cache.value_cache[1], (full * 0.5).to(torch.float8_e5m2).view(torch.uint8))


@pytest.mark.parametrize("Hq,Hkv,S,L,D,dtype,causal", [
    (32, 32, 1, 1000, 128, torch.float16, False),      # Llama-2-7B decode
    (32, 8, 1, 4097, 128, torch.float16, False),       # Llama-3-8B / Mistral GQA decode
    (32, 8, 5, 777, 128, torch.bfloat16, True),        # a few new tokens, causal
    (8, 2, 32, 300, 64, torch.float32, True),
    (6, 6, 3, 65, 96, torch.float16, True),
    (4, 1, 2, 129, 256, torch.float16, False),
    (8, 8, 1, 31, 192, torch.float16, False),
])
def test_sdp_fp8_matches_float64_attention(kv, cuda_device, Hq, Hkv, S, L, D, dtype, causal):
    gen = torch.Generator(device=cuda_device).manual_seed(L)
    B = 2
    q = torch.randn(B, Hq, S, D, device=cuda_device, generator=gen).to(dtype)
    cache = kv.DynamicFp8Cache()
    k_new = torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).to(dtype)
    v_new = torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).to(dtype)
    k_u8, v_u8 = cache.update(k_new, v_new, 0)
    scale = 1 / math.sqrt(D)
    out = kv.scaled_dot_product_attention(q, k_u8, v_u8, None, causal, scale)
    assert out.shape == (B, Hq, S, D) and out.dtype == dtype
    want = ref_attention(q, k_u8, v_u8, None, scale, causal)
    eps = {torch.float16: 1e-3, torch.bfloat16: 8e-3, torch.float32: 2e-5}[dtype]
    err = (out.double() - want).abs()
    assert float(err.max()) <= eps * float(want.abs().max()) + 1e-6, float(err.max())


def test_sdp_fp8_additive_mask_and_padding(kv, cuda_device):
    gen = torch.Generator(device=cuda_device).manual_seed(9)
    B, Hq, Hkv, S, L, D = 3, 16, 4, 1, 640, 128
    q = torch.randn(B, Hq, S, D, device=cuda_device, generator=gen).half()
    cache = kv.DynamicFp8Cache()
    k_u8, v_u8 = cache.update(torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).half(),
                              torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).half(), 0)
    mask = torch.zeros(B, 1, S, L, device=cuda_device, dtype=torch.float16)
    mask[0, :, :, :100] = float("-inf")            # left padding of sequence 0
    mask[1, :, :, 5:600] = -3.0                     # a finite bias
    mask[2, :, :, 1:] = float("-inf")               # a single visible key
    scale = 0.11
    out = kv.scaled_dot_product_attention(q, k_u8, v_u8, mask, False, scale)
    want = ref_attention(q, k_u8, v_u8, mask, scale, False)
    assert float((out.double() - want).abs().max()) <= 1e-3 * float(want.abs().max()) + 1e-6
    # per-head mask
    mask_h = torch.randn(B, Hq, S, L, device=cuda_device, generator=gen).half()
    out = kv.scaled_dot_product_attention(q, k_u8, v_u8, mask_h, False, scale)
    want = ref_attention(q, k_u8, v_u8, mask_h, scale, False)
    assert float((out.double() - want).abs().max()) <= 1e-3 * float(want.abs().max()) + 1e-6
    # deterministic
    again = kv.scaled_dot_product_attention(q, k_u8, v_u8, mask_h, False, scale)
    assert torch.equal(out, again)


def test_prefill_falls_back_to_restored_cache(kv, cuda_device):
    """more than 32 new tokens: the cache is restored to the query dtype and torch's attention runs"""
    gen = torch.Generator(device=cuda_device).manual_seed(11)
    B, H, S, D = 1, 8, 48, 64
    q = torch.randn(B, H, S, D, device=cuda_device, generator=gen).half()
    cache = kv.DynamicFp8Cache()
    k_u8, v_u8 = cache.update(torch.randn(B, H, S, D, device=cuda_device, generator=gen).half(),
                              torch.randn(B, H, S, D, device=cuda_device, generator=gen).half(), 0)
    out = kv.scaled_dot_product_attention(q, k_u8, v_u8, None, True)
    want = ref_attention(q, k_u8, v_u8, None, 1 / math.sqrt(D), True)
    assert float((out.double() - want).abs().max()) <= 2e-3 * float(want.abs().max())


def test_sdp_fp8_growing_scores_and_long_cache(kv, cuda_device):
    """keys whose scores keep growing along the cache (the running reference of the softmax is moved many times,
    by small and by large steps) and a cache long enough for several chunks per warp and several splits"""
    gen = torch.Generator(device=cuda_device).manual_seed(21)
    B, Hq, Hkv, S, L, D = 1, 8, 2, 1, 9000, 128
    q = torch.randn(B, Hq, S, D, device=cuda_device, generator=gen).half()
    ramp = torch.linspace(0.05, 3.0, L, device=cuda_device)[None, None, :, None]
    k = torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen) * ramp
    k[:, :, 5000] *= 4.0                               # one key far above the rest
    v = torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen)
    cache = kv.DynamicFp8Cache()
    k_u8, v_u8 = cache.update(k.half(), v.half(), 0)
    for scale in (1 / math.sqrt(D), 1.0):
        out = kv.scaled_dot_product_attention(q, k_u8, v_u8, None, False, scale)
        want = ref_attention(q, k_u8, v_u8, None, scale, False)
        assert float((out.double() - want).abs().max()) <= 1e-3 * float(want.abs().max()) + 1e-6


def test_sdp_fp8_strided_and_misaligned_caches(kv, cuda_device):
    """rows that are not densely stored (row stride 160 bytes: the staged copies take their general path) and a value
    cache that is only 4-byte aligned (falls back to the CUDA-core kernel): same results as the dense cache"""
    gen = torch.Generator(device=cuda_device).manual_seed(22)
    B, Hq, Hkv, S, L, D = 2, 8, 2, 2, 700, 128
    q = torch.randn(B, Hq, S, D, device=cuda_device, generator=gen).half()
    cache = kv.DynamicFp8Cache()
    k_u8, v_u8 = cache.update(torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).half(),
                              torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).half(), 0)
    scale = 1 / math.sqrt(D)
    dense = kv.sdp_fp8(q, k_u8, v_u8, None, scale, True)
    want = ref_attention(q, k_u8, v_u8, None, scale, True)
    assert float((dense.double() - want).abs().max()) <= 1e-3 * float(want.abs().max()) + 1e-6
    wide_k = torch.zeros(B, Hkv, L, 160, dtype=torch.uint8, device=cuda_device)
    wide_v = torch.zeros(B, Hkv, L, 160, dtype=torch.uint8, device=cuda_device)
    wide_k[..., :D] = k_u8
    wide_v[..., :D] = v_u8
    strided = kv.sdp_fp8(q, wide_k[..., :D], wide_v[..., :D], None, scale, True)
    assert torch.equal(strided, dense)                 # same kernel, same order of operations
    wide_v[..., 4:4 + D] = v_u8
    shifted = kv.sdp_fp8(q, wide_k[..., :D], wide_v[..., 4:4 + D], None, scale, True)
    assert float((shifted.double() - want).abs().max()) <= 1e-3 * float(want.abs().max()) + 1e-6
    # a query that is a strided view (heads and tokens transposed, as decoders produce it)
    q_t = q.transpose(1, 2).contiguous().transpose(1, 2)
    assert torch.equal(kv.sdp_fp8(q_t, k_u8, v_u8, None, scale, True), dense)


@pytest.mark.parametrize("B,Hq,Hkv,S,L,dtype,causal", [
    (1, 32, 32, 512, 512, torch.float16, True),        # Llama-2-7B prefill, 512-token prompt
    (2, 8, 2, 100, 300, torch.bfloat16, True),         # GQA, a chunked prefill behind 200 cached tokens, ragged tiles
    (1, 4, 4, 70, 70, torch.float16, False),           # non-causal
    (1, 4, 1, 33, 33, torch.float16, True),            # the smallest query count that leaves the decode kernel
    (1, 2, 2, 257, 1100, torch.float16, True),
])
def test_sdp_fp8_many_tokens_matches_float64(kv, cuda_device, B, Hq, Hkv, S, L, dtype, causal):
    """sdp_fp8_causal / sdp_fp8_non_causal (csrc/sdp_prefill.cuh): more than 32 new tokens over the e5m2 cache"""
    gen = torch.Generator(device=cuda_device).manual_seed(21)
    D = 128
    q = torch.randn(B, Hq, S, D, device=cuda_device, generator=gen).to(dtype)
    cache = kv.DynamicFp8Cache()
    k_u8, v_u8 = cache.update(torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).to(dtype),
                              torch.randn(B, Hkv, L, D, device=cuda_device, generator=gen).to(dtype), 0)
    scale = 1 / math.sqrt(D)
    out = kv.scaled_dot_product_attention(q, k_u8, v_u8, None, causal, scale)
    direct = kv.sdp_fp8_causal(q, k_u8, v_u8, None, scale) if causal else kv.sdp_fp8_non_causal(q, k_u8, v_u8, None, scale)
    assert torch.equal(out, direct)                                   # the dispatcher took the many-token kernel
    want = ref_attention(q, k_u8, v_u8, None, scale, causal)
    tol = (1e-3 if dtype == torch.float16 else 8e-3) * float(want.abs().max())
    assert float((out.double() - want).abs().max()) <= tol
    # an additive mask (left padding of sequence 0 + a finite bias), with and without the causal rule on top of it
    mask = torch.zeros(B, 1, S, L, device=cuda_device, dtype=dtype)
    mask[0, :, :, :7] = float("-inf")
    mask[-1, :, :, 9:20] = -2.5
    if causal:
        i = torch.arange(S, device=cuda_device)[:, None]
        j = torch.arange(L, device=cuda_device)[None, :]
        mask = mask.masked_fill(j > (L - S + i), float("-inf"))
    out = kv.scaled_dot_product_attention(q, k_u8, v_u8, mask, causal, scale)
    want = torch.nan_to_num(ref_attention(q, k_u8, v_u8, mask, scale, False), nan=0.0)     # a row with every key masked gives zeros
    assert float((out.double() - want).abs().max()) <= tol
    # a boolean mask (what transformers >= 5 passes) and a strided query view (the rotated slice of the qkv projection)
    qkv = torch.randn(B, S, (Hq + 2 * Hkv) * D, device=cuda_device, generator=gen).to(dtype)
    q_view = qkv.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)[:, :Hq]
    bmask = torch.isfinite(mask)
    out = kv.scaled_dot_product_attention(q_view, k_u8, v_u8, bmask, False, scale)
    want = torch.nan_to_num(ref_attention(q_view, k_u8, v_u8, torch.zeros_like(mask).masked_fill(~bmask, float("-inf")), scale, False), nan=0.0)
    assert float((out.double() - want).abs().max()) <= tol
