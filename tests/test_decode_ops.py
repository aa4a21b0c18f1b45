"""The small operators of the recorded decode step (csrc/decode_ops.cu, lowbit_sdp_fp8_devlen_cuda), through the
C ABI, against the statements they restate: transformers' RMSNorm, the separate rotary + cache-quantise kernels,
the host-length decode attention, torch.argmax."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _st(dev):
    return torch.cuda.current_stream(dev).cuda_stream


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("rows,hidden", [(1, 4096), (5, 1000), (3, 8192)])
def test_rms_norm_matches_the_torch_statement(native_lib, cuda_device, dtype, rows, hidden):
    ggml, dev = native_lib, cuda_device
    gen = torch.Generator(device=dev).manual_seed(0)
    x = (torch.randn(rows, hidden, device=dev, generator=gen) * 3).to(dtype)
    w = (1 + 0.1 * torch.randn(hidden, device=dev, generator=gen)).to(dtype)
    eps = 1e-5
    out = torch.empty_like(x)
    ggml.lowbit_rms_norm(x, w, eps, out, _st(dev))
    # LlamaRMSNorm.forward in float64 for the statistics, the same two roundings
    xf = x.double()
    n = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(dtype)
    want = (w.double() * n.double()).to(dtype)
    diff = (out.double() - want.double()).abs()
    ulp = {torch.float32: 2.0 ** -22, torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}[dtype]
    assert float((diff / (want.double().abs() + 1e-6)).max()) <= 2.02 * ulp          # two roundings, each may move by one
    # strided input rows
    big = torch.zeros(rows, hidden + 64, device=dev, dtype=dtype)
    big[:, :hidden] = x
    out2 = torch.empty_like(x)
    ggml.lowbit_rms_norm(big[:, :hidden], w, eps, out2, _st(dev))
    assert torch.equal(out, out2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("B,S,Hq,Hkv,D", [(1, 1, 32, 32, 128), (2, 3, 8, 2, 128), (1, 5, 4, 4, 64)])
def test_rope_kv_append_equals_the_separate_kernels(native_lib, cuda_device, dtype, B, S, Hq, Hkv, D):
    """same bytes as rotary_half_with_cache_inplaced + quantize_key_value (the eager attention forward)"""
    from ipex_llm_b200.transformers import kv
    from ipex_llm_b200.transformers.models import common
    ggml, dev = native_lib, cuda_device
    gen = torch.Generator(device=dev).manual_seed(1)
    cap, pos0 = 64, 7
    tot = (Hq + 2 * Hkv) * D
    qkv = torch.randn(B, S, tot, device=dev, generator=gen).to(dtype)
    inv_freq = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device=dev, dtype=torch.float32) / D))
    ang = torch.arange(cap, device=dev, dtype=torch.float32)[:, None] * inv_freq[None, :]
    emb = torch.cat([ang, ang], dim=-1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)                         # [cap, D] tables, rotate_half layout
    # eager statement on a copy
    ref = qkv.clone()
    q, k, v = ref.view(B, S, Hq + 2 * Hkv, D).transpose(1, 2).split([Hq, Hkv, Hkv], dim=1)
    common.rotary_half_with_cache_inplaced(q, k, cos[None, pos0:pos0 + S], sin[None, pos0:pos0 + S])
    kc_ref = torch.zeros(B, Hkv, cap, D, dtype=torch.uint8, device=dev)
    vc_ref = torch.zeros_like(kc_ref)
    kv.quantize_key_value(k, v, kc_ref[:, :, pos0:pos0 + S], vc_ref[:, :, pos0:pos0 + S])
    # fused kernel, position read from device memory
    kc, vc = torch.zeros_like(kc_ref), torch.zeros_like(kc_ref)
    state = torch.tensor([pos0 - 2], dtype=torch.int32, device=dev)
    ggml.lowbit_rope_kv_append(qkv, Hq, Hkv, D, cos, sin, state, 2, kc, vc, _st(dev))
    assert torch.equal(qkv.view(B, S, -1, D)[:, :, :Hq], ref.view(B, S, -1, D)[:, :, :Hq])       # rotated queries
    assert torch.equal(kc, kc_ref) and torch.equal(vc, vc_ref)
    # positions at or beyond the capacity are not written
    kc2, vc2 = torch.zeros_like(kc_ref), torch.zeros_like(kc_ref)
    state2 = torch.tensor([cap - 1], dtype=torch.int32, device=dev)
    ggml.lowbit_rope_kv_append(qkv.clone(), Hq, Hkv, D, cos, sin, state2, 0, kc2, vc2, _st(dev))
    assert int(kc2[:, :, :cap - 1].count_nonzero()) == 0


@pytest.mark.parametrize("Hq,Hkv,D,dtype", [(32, 32, 128, torch.float16), (32, 8, 128, torch.bfloat16), (8, 8, 64, torch.float16),
                                             (6, 6, 96, torch.float16)])
def test_sdp_device_length_matches_host_length(native_lib, cuda_device, Hq, Hkv, D, dtype):
    from ipex_llm_b200.transformers import kv
    ggml, dev = native_lib, cuda_device
    gen = torch.Generator(device=dev).manual_seed(2)
    B, cap = 2, 768
    k = torch.randn(B, Hkv, cap, D, device=dev, generator=gen).to(torch.float8_e5m2).view(torch.uint8)
    v = torch.randn(B, Hkv, cap, D, device=dev, generator=gen).to(torch.float8_e5m2).view(torch.uint8)
    q = torch.randn(B, Hq, 1, D, device=dev, generator=gen).to(dtype)
    ws = torch.empty(ggml.lowbit_sdp_fp8_workspace_size(B, Hq, Hkv, 1, cap, D), dtype=torch.uint8, device=dev)
    scale = 1 / math.sqrt(D)
    state = torch.zeros(3, dtype=torch.int32, device=dev)
    for L in (1, 31, 200, 513, cap):
        state[0] = L - 1
        out = torch.full((B, Hq, 1, D), float("nan"), dtype=dtype, device=dev)
        ggml.lowbit_sdp_fp8_devlen(q, k, v, False, scale, state, 1, out, ws, _st(dev))
        want = kv.sdp_fp8(q, k[:, :, :L], v[:, :, :L], None, scale)
        kd = k[:, :, :L].view(torch.float8_e5m2).double().repeat_interleave(Hq // Hkv, 1)
        vd = v[:, :, :L].view(torch.float8_e5m2).double().repeat_interleave(Hq // Hkv, 1)
        exact = torch.softmax(q.double() @ kd.transpose(-1, -2) * scale, -1) @ vd
        tol = (1e-3 if dtype == torch.float16 else 8e-3) * float(exact.abs().max())
        assert float((out.double() - exact).abs().max()) <= tol, L
        assert float((out.double() - want.double()).abs().max()) <= tol, L


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_greedy_advance(native_lib, cuda_device, dtype):
    ggml, dev = native_lib, cuda_device
    gen = torch.Generator(device=dev).manual_seed(3)
    B, V, cap = 3, 32000, 16
    state = torch.tensor([40, 0, 0, 0, 0, 0, 0, 0], dtype=torch.int32, device=dev)
    tokens = torch.full((B, cap), -1, dtype=torch.int64, device=dev)
    next_ids = torch.zeros(B, dtype=torch.int64, device=dev)
    picks = []
    for step in range(3):
        logits = torch.randn(B, V, device=dev, generator=gen).to(dtype)
        logits[1, 777] = logits[1].max()                   # a tie: the lowest index wins
        logits[1, 31999] = logits[1, 777]
        ggml.lowbit_greedy_advance(logits, next_ids, tokens, state, _st(dev))
        want = torch.stack([torch.nonzero(logits[b] == logits[b].max())[0, 0] for b in range(B)])
        assert torch.equal(next_ids, want)
        picks.append(want)
    assert state.tolist()[:3] == [43, 3, 0]
    assert torch.equal(tokens[:, :3], torch.stack(picks, dim=1)) and int((tokens[:, 3:] != -1).sum()) == 0


def test_greedy_advance_min_new_tokens_rule(native_lib, cuda_device):
    """ids in the suppress list cannot be picked while fewer than ``suppress_below`` tokens were generated"""
    ggml, dev = native_lib, cuda_device
    V = 1000
    logits = torch.zeros(1, V, device=dev)
    logits[0, 7] = 5.0
    logits[0, 3] = 4.0
    state = torch.tensor([10, 0, 0, 2, 1, 0, 0, 0], dtype=torch.int32, device=dev)      # suppress id 7 for the first 2 tokens
    sup = torch.tensor([7, -1, -1, -1, -1, -1, -1, -1], dtype=torch.int64, device=dev)
    tokens = torch.zeros(1, 8, dtype=torch.int64, device=dev)
    nxt = torch.zeros(1, dtype=torch.int64, device=dev)
    for _ in range(3):
        ggml.lowbit_greedy_advance(logits, nxt, tokens, state, _st(dev), sup)
    assert tokens[0, :3].tolist() == [3, 3, 7]


@pytest.mark.parametrize("B,Hq,Hkv,dtype", [(1, 32, 32, torch.float16), (2, 32, 8, torch.bfloat16), (1, 64, 8, torch.float16),
                                             (3, 8, 1, torch.float16), (1, 40, 40, torch.float16)])
def test_decode_attention_one_launch_equals_the_separate_kernels(native_lib, cuda_device, B, Hq, Hkv, dtype):
    """lowbit_decode_attention_cuda (rotary + cache write + attention + in-cluster merge in one launch) against
    lowbit_rope_kv_append_cuda + lowbit_sdp_fp8_devlen_cuda and against float64 attention on the same cache bytes"""
    ggml, dev = native_lib, cuda_device
    gen = torch.Generator(device=dev).manual_seed(5)
    D, cap = 128, 768
    tot = (Hq + 2 * Hkv) * D
    inv_freq = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device=dev, dtype=torch.float32) / D))
    ang = torch.arange(cap, device=dev, dtype=torch.float32)[:, None] * inv_freq[None, :]
    emb = torch.cat([ang, ang], dim=-1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    k0 = torch.randn(B, Hkv, cap, D, device=dev, generator=gen).to(torch.float8_e5m2).view(torch.uint8)
    v0 = torch.randn(B, Hkv, cap, D, device=dev, generator=gen).to(torch.float8_e5m2).view(torch.uint8)
    ws = torch.empty(max(256, ggml.lowbit_decode_attention_workspace_size(B, Hq, Hkv, cap, D)), dtype=torch.uint8, device=dev)
    scale = 1 / math.sqrt(D)
    launches = set()
    for P in (0, 5, 31, 32, 95, 200, 511, 700, cap - 1):
        qkv = torch.randn(B, 1, tot, device=dev, generator=gen).to(dtype)
        state = torch.tensor([P, 0, 0, 0, 0, 0, 0, 0], dtype=torch.int32, device=dev)
        # separate kernels
        qa, ka, va = qkv.clone(), k0.clone(), v0.clone()
        ggml.lowbit_rope_kv_append(qa, Hq, Hkv, D, cos, sin, state, 0, ka, va, _st(dev))
        q_view = qa.view(B, 1, Hq + 2 * Hkv, D).transpose(1, 2)[:, :Hq]
        want = torch.empty(B, Hq, 1, D, dtype=dtype, device=dev)
        ggml.lowbit_sdp_fp8_devlen(q_view, ka, va, False, scale, state, 1, want, ws, _st(dev))
        # one call
        qb, kb, vb = qkv.clone(), k0.clone(), v0.clone()
        got = torch.full((B, Hq, 1, D), float("nan"), dtype=dtype, device=dev)
        ggml.lowbit_decode_attention(qb, Hq, Hkv, D, cos, sin, state, kb, vb, scale, got, ws, _st(dev), cache_complete=bool(P % 2))
        launches.add(ggml.last_launch_count())
        assert torch.equal(kb, ka) and torch.equal(vb, va), P                     # the new row, and nothing else, was written
        L = P + 1
        kd = ka[:, :, :L].view(torch.float8_e5m2).double().repeat_interleave(Hq // Hkv, 1)
        vd = va[:, :, :L].view(torch.float8_e5m2).double().repeat_interleave(Hq // Hkv, 1)
        exact = torch.softmax(q_view.double() @ kd.transpose(-1, -2) * scale, -1) @ vd
        tol = (1e-3 if dtype == torch.float16 else 8e-3) * float(exact.abs().max())
        assert float((got.double() - exact).abs().max()) <= tol, P
        assert float((got.double() - want.double()).abs().max()) <= tol, P
    assert launches == {1}, launches                     # every case here fits the one-launch kernel
