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


# ------------------------------------------------------------------ RMSNorm inside the consuming GEMV
def _tiled_weight(dev, rows, k, qtype, seed):
    from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    gen = torch.Generator(device=dev).manual_seed(seed)
    w = torch.randn(rows, k, device=dev, generator=gen) * 0.05
    q = ggml_tensor_qtype[qtype]
    packed = lbl.ggml_convert_qtype(w, q, device=dev)
    return lbl.ggml_q_format_convet_cpu2xpu(packed, rows * k, q, shape=(rows, k)), q


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "fp8_e5m2", "nf4"])
@pytest.mark.parametrize("M,N,K", [(1, 12288, 4096), (1, 192, 1024), (3, 4096, 4096), (8, 1024, 2048), (5, 256, 16384)])
def test_fused_rms_norm_gemv_equals_the_two_launches(native_lib, cuda_device, dtype, qtype, M, N, K):
    """lowbit_fuse_rms_norm + lowbit_matmul_cuda on the raw rows == lowbit_rms_norm_cuda, then lowbit_matmul_cuda: the
    prologue reproduces the stand-alone kernel's summation order, so the outputs carry the same bits"""
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    ggml, dev = native_lib, cuda_device
    if not ggml.lowbit_fuse_rms_norm_supported(lbl.ggml_tensor_qtype[qtype], N, K, M, dtype):
        pytest.skip("format without a tensor-core codec")
    w, q = _tiled_weight(dev, N, K, qtype, 11)
    gen = torch.Generator(device=dev).manual_seed(12)
    x = (torch.randn(M, K, device=dev, generator=gen) * 2.5).to(dtype)
    nw = (1 + 0.2 * torch.randn(K, device=dev, generator=gen)).to(dtype)
    eps = 1e-5
    xn = torch.empty_like(x)
    ggml.lowbit_rms_norm(x, nw, eps, xn, _st(dev))
    for out_dtype in (dtype, torch.float32):
        want = lbl.lowbit_matmul(xn, w, (N, K), q, out_dtype=out_dtype)
        got = lbl.lowbit_matmul(x, w, (N, K), q, out_dtype=out_dtype, norm=(nw, eps))
        assert ggml.last_launch_count() == 1
        assert torch.equal(got, want), float((got.float() - want.float()).abs().max())
    # the request is one-shot: the next call is an ordinary matmul of the raw rows
    plain = lbl.lowbit_matmul(x, w, (N, K), q)
    assert torch.equal(plain, lbl.lowbit_matmul(x, w, (N, K), q))
    assert not torch.equal(plain.float(), want)
    # bias and accumulate go through the same epilogue
    bias = torch.randn(N, device=dev, generator=gen).to(dtype)
    base = torch.randn(M, N, device=dev, generator=gen).to(dtype)
    a, b = base.clone(), base.clone()
    lbl.lowbit_matmul(xn, w, (N, K), q, out=a, bias=bias, accumulate=True)
    lbl.lowbit_matmul(x, w, (N, K), q, out=b, bias=bias, accumulate=True, norm=(nw, eps))
    assert torch.equal(a, b)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("M", [1, 2, 7, 8])
def test_fused_rms_norm_mlp_equals_the_two_launches(native_lib, cuda_device, dtype, M):
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    ggml, dev = native_lib, cuda_device
    K, N = 4096, 11008
    wg, q = _tiled_weight(dev, N, K, "sym_int4", 21)
    wu, _ = _tiled_weight(dev, N, K, "sym_int4", 22)
    gen = torch.Generator(device=dev).manual_seed(23)
    x = (torch.randn(M, K, device=dev, generator=gen) * 0.7).to(dtype)
    nw = (1 + 0.2 * torch.randn(K, device=dev, generator=gen)).to(dtype)
    xn = torch.empty_like(x)
    ggml.lowbit_rms_norm(x, nw, 1e-6, xn, _st(dev))
    want = torch.empty(M, N, device=dev, dtype=dtype)
    got = torch.empty_like(want)
    lbl.mlp_forward(xn, wg, wu, M, K, N, 0, q, out=want)
    lbl.mlp_forward(x, wg, wu, M, K, N, 0, q, out=got, norm=(nw, 1e-6))
    assert ggml.last_launch_count() == 1
    assert torch.equal(got, want), float((got.float() - want.float()).abs().max())


def test_fused_rms_norm_is_refused_where_it_cannot_run(native_lib, cuda_device):
    """M > 8, fp32 rows or an output length off the 64-row grid: an argument error, and the request is dropped"""
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    ggml, dev = native_lib, cuda_device
    K = 1024
    nw16 = torch.ones(K, device=dev, dtype=torch.float16)
    for M, N, dt in [(9, 256, torch.float16), (40, 256, torch.float16), (2, 256, torch.float32), (2, 208, torch.float16)]:
        w, q = _tiled_weight(dev, N, K, "sym_int4", 31)
        x = torch.randn(M, K, device=dev).to(dt)
        assert not ggml.lowbit_fuse_rms_norm_supported(q, N, K, M, dt)
        with pytest.raises(Exception):
            lbl.lowbit_matmul(x, w, (N, K), q, norm=(nw16.to(dt), 1e-5))
        ref = lbl.lowbit_matmul(x, w, (N, K), q)              # nothing is left pending
        assert torch.equal(ref, lbl.lowbit_matmul(x, w, (N, K), q))
    # the two-matmul MLP route (M > 8) refuses it as well
    wg, q = _tiled_weight(dev, 256, K, "sym_int4", 32)
    x = torch.randn(12, K, device=dev).half()
    with pytest.raises(Exception):
        lbl.mlp_forward(x, wg, wg, 12, K, 256, 0, q, norm=(nw16, 1e-5))
