"""In-place rotary position embedding (second half; reference models/llama.py:151-165,
models/chatglm2.py:251, models/common.py:354-367).  Reference: the rotation evaluated in float64 from fp32
angles (position * inv_freq in fp32, as transformers computes them) or from the given tables; the kernel works
in fp32 and rounds once, so the tolerance is one rounding of the output dtype plus the fp32 sin / cos error.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

EPS = {torch.float32: 2e-6, torch.float16: 2.0 ** -11 * 1.01, torch.bfloat16: 2.0 ** -8 * 1.01}


@pytest.fixture(scope="module")
def rope(native_lib, cuda_device):
    from ipex_llm_b200.transformers.models import common
    return common


def rotate64(x, cos, sin, two, R):
    """x [B,H,S,D] float64, cos/sin [B|1,S,R/2] float64"""
    out = x.clone()
    c, s = cos[:, None], sin[:, None]
    if two:
        x0, x1 = x[..., 0:R:2], x[..., 1:R:2]
        out[..., 0:R:2] = x0 * c - x1 * s
        out[..., 1:R:2] = x1 * c + x0 * s
    else:
        x0, x1 = x[..., :R // 2], x[..., R // 2:R]
        out[..., :R // 2] = x0 * c - x1 * s
        out[..., R // 2:R] = x1 * c + x0 * s
    return out


def qkv_views(dev, dtype, B, S, Hq, Hkv, D, seed):
    """q / k as the decoders get them: strided views of the merged projection's output"""
    gen = torch.Generator(device=dev).manual_seed(seed)
    qkv = torch.randn(B, S, (Hq + 2 * Hkv) * D, device=dev, generator=gen).to(dtype)
    v4 = qkv.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)
    q, k, v = v4.split([Hq, Hkv, Hkv], dim=1)
    return qkv, q, k, v


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("two,R", [(False, 128), (True, 128), (True, 64), (False, 96)])
def test_rotary_inplaced_from_inv_freq(rope, cuda_device, dtype, two, R):
    dev = cuda_device
    B, S, Hq, Hkv, D = 2, 7, 8, 2, 128
    qkv, q, k, v = qkv_views(dev, dtype, B, S, Hq, Hkv, D, 1)
    before = qkv.clone()
    inv_freq = 1.0 / (10000.0 ** (torch.arange(0, R, 2, device=dev, dtype=torch.float32) / R))
    position_ids = torch.tensor([[0, 1, 2, 3, 4, 5, 6], [4090, 4091, 4092, 4093, 32000, 100000, 7]], device=dev)
    fn = rope.rotary_two_inplaced if two else rope.rotary_half_inplaced
    fn(inv_freq, position_ids, q, k)
    angle = (position_ids[:, :, None].float() * inv_freq[None, None, :]).double()      # fp32 product, as in transformers
    b4 = before.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2).double()
    want = rotate64(b4[:, :Hq + Hkv], torch.cos(angle), torch.sin(angle), two, R)
    got = qkv.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)
    mag = b4[:, :Hq + Hkv].abs().max()
    assert float((got[:, :Hq + Hkv].double() - want).abs().max()) <= EPS[dtype] * float(want.abs().max()) + 4e-7 * float(mag)
    assert torch.equal(got[:, Hq + Hkv:], before.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)[:, Hq + Hkv:])    # v untouched
    if R < D:
        assert torch.equal(got[..., R:], before.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)[..., R:])


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_rotary_with_cache_tables(rope, cuda_device, dtype):
    dev = cuda_device
    B, S, Hq, Hkv, D = 3, 5, 4, 4, 64
    gen = torch.Generator(device=dev).manual_seed(3)
    angle = torch.rand(B, S, D // 2, device=dev, generator=gen) * 6.28
    cos_h, sin_h = torch.cos(angle).to(dtype), torch.sin(angle).to(dtype)
    # rotate_half layout, as transformers >= 4.46 hands it over: [B, S, D] with the half repeated
    cos, sin = torch.cat([cos_h, cos_h], -1), torch.cat([sin_h, sin_h], -1)
    qkv, q, k, v = qkv_views(dev, dtype, B, S, Hq, Hkv, D, 4)
    before = qkv.clone().view(B, S, Hq + 2 * Hkv, D).transpose(1, 2).double()
    rope.rotary_half_with_cache_inplaced(q, k, cos, sin)
    want = rotate64(before[:, :Hq + Hkv], cos_h.double(), sin_h.double(), False, D)
    got = qkv.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)[:, :Hq + Hkv].double()
    assert float((got - want).abs().max()) <= EPS[dtype] * float(want.abs().max()) + 1e-7
    # adjacent pairing with both table layouts; tables broadcast over the batch ([1, S, R]) and q only
    for half_layout in (True, False):
        qkv, q, k, v = qkv_views(dev, dtype, B, S, Hq, Hkv, D, 5)
        before = qkv.clone().view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)
        c1, s1 = cos_h[:1], sin_h[:1]
        if half_layout:
            ct, stb = torch.cat([c1, c1], -1), torch.cat([s1, s1], -1)
        else:
            ct, stb = c1.repeat_interleave(2, -1), s1.repeat_interleave(2, -1)
        rope.rotary_two_with_cache_inplaced(q, None, ct, stb, half_layout)
        want = rotate64(before[:, :Hq].double(), c1.double(), s1.double(), True, D)
        got = qkv.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2)
        assert float((got[:, :Hq].double() - want).abs().max()) <= EPS[dtype] * float(want.abs().max()) + 1e-7
        assert torch.equal(got[:, Hq:], before[:, Hq:])                      # k and v untouched


def test_rotary_matches_transformers_llama_rope(rope, cuda_device):
    """against transformers' own apply_rotary_pos_emb in fp32 (the formula decoders would otherwise run)"""
    from transformers.models.llama.modeling_llama import apply_rotary_pos_emb
    dev = cuda_device
    B, S, H, D = 1, 9, 4, 128
    gen = torch.Generator(device=dev).manual_seed(6)
    q = torch.randn(B, H, S, D, device=dev, generator=gen)
    k = torch.randn(B, H, S, D, device=dev, generator=gen)
    inv_freq = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device=dev, dtype=torch.float32) / D))
    pos = torch.arange(100, 100 + S, device=dev)[None]
    freqs = pos[:, :, None].float() * inv_freq[None, None, :]
    emb = torch.cat([freqs, freqs], -1)
    q_ref, k_ref = apply_rotary_pos_emb(q, k, emb.cos(), emb.sin())
    rope.rotary_half_inplaced(inv_freq, pos, q, k)
    assert float((q - q_ref).abs().max()) <= 4e-6 and float((k - k_ref).abs().max()) <= 4e-6


def test_rotary_rejects_bad_arguments(rope, cuda_device):
    q = torch.zeros(1, 2, 3, 64, device=cuda_device, dtype=torch.float16)
    inv_freq = torch.ones(40, device=cuda_device)                             # rotary width 80 > head_dim
    with pytest.raises(RuntimeError):
        rope.rotary_half_inplaced(inv_freq, torch.zeros(1, 3, dtype=torch.int64, device=cuda_device), q, None)
    with pytest.raises(RuntimeError):
        rope.rotary_half_inplaced(inv_freq[:8], torch.zeros(1, 3, dtype=torch.int64, device=cuda_device), q.transpose(2, 3), None)
