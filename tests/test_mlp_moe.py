"""GPU parity of the fused decode ops built on the dequant-GEMV:

  * mlp_forward  = xe_linear.mlp_forward_xpu  (models/common.py:146-161)
  * moe_forward  = xe_linear.moe_forward_vec   (models/common.py:341-354 fast path)

Bar: each inner matmul within the summation-order bound of test_gpu_parity (2e-6 * sum|x w|),
propagated through act(g) * u with the activation's Lipschitz constant (<= 1.13 for SiLU / GELU).
"""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lbl(native_lib, cuda_device):
    from ipex_llm_b200.transformers import low_bit_linear
    return low_bit_linear


def qid(name):
    from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
    return ggml_tensor_qtype[name]


def to_tiled(lbl, dev, packed, shape, qtype):
    p = torch.from_numpy(packed)
    if qtype == "nf4":
        p = p.view(torch.bfloat16)
    return lbl.ggml_q_format_convet_cpu2xpu(p.to(dev), shape[0] * shape[1], qid(qtype), shape=shape)


def act64(v, act):
    if act == 0:
        return v / (1.0 + np.exp(-v))
    from scipy.special import erf
    return 0.5 * v * (1.0 + erf(v / np.sqrt(2.0)))


def bound64(packed, shape, qtype, x, xdt=None):
    wd = np.abs(oracle.dequantize(packed, shape, qtype)).astype(np.float64)
    rel = 2e-6
    if qtype == "nf4":      # table values rounded to fp16 (fp16 activations) / split into two bf16 (others)
        rel = 2.6e-4 if xdt == torch.float16 else 1e-5
    return rel * (np.abs(x).astype(np.float64) @ wd.T) + 1e-30


def mlp_expect(pg, pu, shape, qtype, x, act, xdt=None):
    g = oracle.matmul_f64(pg, shape, qtype, x)
    u = oracle.matmul_f64(pu, shape, qtype, x)
    bg, bu = bound64(pg, shape, qtype, x, xdt), bound64(pu, shape, qtype, x, xdt)
    a = act64(g, act)
    want = a * u
    # |d(act(g) u)| <= L |dg| |u| + |act(g)| |du| + |dg||du| L, plus fp32 rounding of act and the product
    tol = 1.13 * bg * (np.abs(u) + bu) + np.abs(a) * bu + 4e-7 * np.abs(want) + 4e-7 * np.abs(u) * (np.abs(g) + 1) + 1e-30
    return want, tol


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "fp8_e4m3", "fp8_e5m2", "fp6", "nf4"])
@pytest.mark.parametrize("xdt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,act", [(1, 0), (4, 1), (8, 0), (12, 0), (16, 1)])
def test_mlp_fused_matches_oracle(lbl, cuda_device, qtype, xdt, M, act):
    rng = np.random.default_rng(4100 + M)
    N, K = 192, 1024
    wg = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    wu = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    pg, pu = oracle.quantize(wg, qtype), oracle.quantize(wu, qtype)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(xdt)
    x = x_t.float().numpy()
    out = torch.empty((M, N), dtype=torch.float32, device=cuda_device)
    lbl.mlp_forward(x_t.to(cuda_device), to_tiled(lbl, cuda_device, pg, (N, K), qtype),
                    to_tiled(lbl, cuda_device, pu, (N, K), qtype), M, K, N, act, qid(qtype), out=out)
    # one launch up to 8 rows, and up to 16 in half precision (two n-tiles)
    if xdt != torch.float32:                           # (fp32 rows add the launch that splits them into bf16 planes)
        assert lbl.ggml.last_launch_count() == 1
    want, tol = mlp_expect(pg, pu, (N, K), qtype, x, act, xdt)
    got = out.cpu().numpy().astype(np.float64)
    assert (np.abs(got - want) <= tol).all(), float((np.abs(got - want) / tol).max())


@pytest.mark.parametrize("shape,M", [((11008, 4096), 1), ((14336, 4096), 2), ((1000, 704), 3), ((256, 2048), 12),
                                     ((256, 2048), 48), ((64, 64), 1), ((11008, 4096), 24), ((1280, 640), 33), ((512, 1024), 100),
                                     ((1280, 640), 200), ((11008, 4096), 512), ((2816, 1024), 300)])
def test_mlp_shapes_and_fallbacks(lbl, cuda_device, shape, M):
    """Llama-2-7B / Llama-3-8B MLP shapes (fused), ragged rows / M > 8 / M > 32 (unfused fallbacks):
    all must agree with two separate matmuls followed by silu * up computed in fp32."""
    rng = np.random.default_rng(4200)
    N, K = shape
    qtype = "sym_int4"
    pg = oracle.quantize((rng.standard_normal((N, K)) * 0.03).astype(np.float32), qtype)
    pu = oracle.quantize((rng.standard_normal((N, K)) * 0.03).astype(np.float32), qtype)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half().to(cuda_device)
    tg, tu = to_tiled(lbl, cuda_device, pg, shape, qtype), to_tiled(lbl, cuda_device, pu, shape, qtype)
    out = torch.empty((M, N), dtype=torch.float32, device=cuda_device)
    lbl.mlp_forward(x_t, tg, tu, M, K, N, lbl.SILU, qid(qtype), out=out)
    g = lbl.lowbit_matmul(x_t, tg, shape, qid(qtype), out_dtype=torch.float32)
    u = lbl.lowbit_matmul(x_t, tu, shape, qid(qtype), out_dtype=torch.float32)
    want = torch.nn.functional.silu(g.double()) * u.double()
    scale = float(want.abs().max())
    if lbl.SKINNY_MIN_M <= M <= lbl.SKINNY_MAX_M and lbl.takes_tensor_core_path(M, N, K, torch.float16):
        # both matrices in ONE pass of the skinny kernel + its reduction: the stream-K ranges over twice the tiles cut K at
        # other places than the two separate matmuls do -- same products, another fp32 summation order
        assert lbl.ggml.last_launch_count() == 2
        xa = x_t.double().abs()
        mg = xa @ torch.from_numpy(np.abs(oracle.dequantize(pg, shape, qtype))).to(cuda_device).double().T
        mu = xa @ torch.from_numpy(np.abs(oracle.dequantize(pu, shape, qtype))).to(cuda_device).double().T
        tol = 4e-6 * (1.13 * mg * u.double().abs() + torch.nn.functional.silu(g.double()).abs() * mu) + 1e-6 * want.abs() + 1e-9
        assert bool(((out.double() - want).abs() <= tol).all()), float(((out.double() - want).abs() / tol).max())
    else:
        assert float((out.double() - want).abs().max()) <= 2e-6 * scale + 1e-6 * float(u.abs().max())
    # default output dtype = activation dtype, same values rounded once
    out16 = lbl.mlp_forward(x_t, tg, tu, M, K, N, lbl.SILU, qid(qtype))
    assert out16.dtype == torch.float16
    if lbl.takes_tensor_core_path(M, N, K, torch.float16):
        # behind the tensor-core kernels the two intermediates are kept in the output's half precision (their contract
        # rounds the weights to it anyway): each of gate and up carries one more rounding of 2^-11
        g16, u16 = g.half().double(), u.half().double()
        want16 = torch.nn.functional.silu(g16) * u16
        assert float((out16.double() - want16).abs().max()) <= 2e-3 * scale
    else:
        assert torch.equal(out16, out.half())


def test_mlp_deterministic_and_single_launch(lbl, cuda_device):
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
    rng = np.random.default_rng(4300)
    N, K = 11008, 4096
    pg = oracle.quantize((rng.standard_normal((N, K)) * 0.03).astype(np.float32), "sym_int4")
    tg = to_tiled(lbl, cuda_device, pg, (N, K), "sym_int4")
    x_t = torch.from_numpy(rng.standard_normal((1, K)).astype(np.float32)).half().to(cuda_device)
    outs = [lbl.mlp_forward(x_t, tg, tg, 1, K, N, lbl.SILU, qid("sym_int4")) for _ in range(4)]
    assert ggml.last_launch_count() == 1
    for o in outs[1:]:
        assert torch.equal(o, outs[0])


def test_mlp_rejects_bad_arguments(lbl, cuda_device):
    x = torch.zeros((1, 64), dtype=torch.float16, device=cuda_device)
    w = torch.zeros(34 * 64, dtype=torch.uint8, device=cuda_device)
    with pytest.raises(Exception):
        lbl.mlp_forward(x, w, w, 1, 64, 64, 7, qid("sym_int4"))
    with pytest.raises(Exception):
        lbl.mlp_forward(x, w, w[:100], 1, 64, 64, 0, qid("sym_int4"))


# ---------------------------------------------------------------- mixture-of-experts, one token
class Experts:
    """n_experts random experts quantised by the oracle, uploaded in the device layout, plus the
    address tables the reference builds (qwen3_moe.py:90-98)."""

    def __init__(self, lbl, dev, rng, n_experts, hidden, inter, qtype):
        self.qtype, self.hidden, self.inter = qtype, hidden, inter
        self.packed, self.tiled = [], []
        for _ in range(n_experts):
            trio = [oracle.quantize((rng.standard_normal(s) * 0.05).astype(np.float32), qtype)
                    for s in ((inter, hidden), (inter, hidden), (hidden, inter))]
            self.packed.append(trio)
            self.tiled.append([to_tiled(lbl, dev, p, s, qtype)
                               for p, s in zip(trio, ((inter, hidden), (inter, hidden), (hidden, inter)))])
        self.gates, self.ups, self.downs = (torch.tensor([t[i].data_ptr() for t in self.tiled], dtype=torch.uint64, device=dev)
                                            for i in range(3))

    def expect(self, x, ids, weights, act, h_dtype):
        """float64 result with the intermediate rounded to h_dtype like the kernel stores it, and the
        tolerance: one rounding of h (half an ulp, relative) + the summation-order bounds."""
        eps = {torch.float16: 2.0 ** -11, torch.bfloat16: 2.0 ** -8, torch.float32: 2.0 ** -22}[h_dtype]
        want = np.zeros(self.hidden)
        tol = np.zeros(self.hidden)
        for e, wz in zip(ids, weights):
            if not 0 <= e < len(self.packed):
                continue
            pg, pu, pd = self.packed[e]
            h, htol = mlp_expect(pg, pu, (self.inter, self.hidden), self.qtype, x, act, h_dtype)
            hq = torch.from_numpy(h).to(h_dtype).double().numpy()
            wd = oracle.dequantize(pd, (self.hidden, self.inter), self.qtype).astype(np.float64)
            want += wz * (hq @ wd.T)[0]
            rel = 2e-6 if self.qtype != "nf4" else (2.6e-4 if h_dtype == torch.float16 else 1e-5)
            tol += abs(wz) * (((htol + eps * np.abs(h)) @ np.abs(wd).T)[0] + rel * (np.abs(hq) @ np.abs(wd).T)[0])
        return want, tol + 1e-7 * np.abs(want) + 1e-30


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"])
@pytest.mark.parametrize("xdt", [torch.float16, torch.bfloat16, torch.float32])
def test_moe_matches_oracle(lbl, cuda_device, qtype, xdt):
    rng = np.random.default_rng(5100)
    hidden, inter, n_exp = 256, 192, 6
    ex = Experts(lbl, cuda_device, rng, n_exp, hidden, inter, qtype)
    x_t = torch.from_numpy(rng.standard_normal((1, hidden)).astype(np.float32)).to(xdt)
    for top_k in (1, 2, 3, 6):
        ids = rng.permutation(n_exp)[:top_k]
        wts = rng.random(top_k).astype(np.float32)
        wts_t = torch.from_numpy(wts).to(xdt)
        out = torch.empty((1, hidden), dtype=torch.float32, device=cuda_device)
        lbl.moe_forward_vec(x_t.to(cuda_device), torch.from_numpy(ids)[None].to(cuda_device), wts_t[None].to(cuda_device),
                            ex.gates, ex.ups, ex.downs, hidden, inter, qid(qtype), out=out)
        want, tol = ex.expect(x_t.float().numpy(), ids, wts_t.float().numpy(), 0, xdt)
        got = out.cpu().numpy()[0].astype(np.float64)
        assert (np.abs(got - want) <= tol).all(), (top_k, float((np.abs(got - want) / tol).max()))


@pytest.mark.parametrize("hidden,inter,top_k", [(2048, 768, 8), (2048, 1408, 6), (4096, 14336, 2), (320, 64, 4), (64, 64, 1)])
def test_moe_shapes(lbl, cuda_device, hidden, inter, top_k):
    """Qwen3-30B-A3B (2048/768, top 8), DeepSeek-V2-Lite (2048/1408, top 6), Mixtral (4096/14336, top 2)
    expert shapes and two small ragged-split ones, against the unfused path built from the same ops."""
    rng = np.random.default_rng(5200)
    n_exp = top_k + 2
    qtype = "sym_int4"
    q = qid(qtype)
    dev = cuda_device
    gen = torch.Generator(device=dev).manual_seed(7)
    tiled = []
    for _ in range(n_exp):
        trio = []
        for shape in ((inter, hidden), (inter, hidden), (hidden, inter)):
            w = torch.empty(shape, device=dev).normal_(0, 0.03, generator=gen)
            trio.append(lbl.FP4Params(w, qtype=q).to(dev).data)
        tiled.append(trio)
    gates, ups, downs = (torch.tensor([t[i].data_ptr() for t in tiled], dtype=torch.uint64, device=dev) for i in range(3))
    x = torch.randn((1, hidden), device=dev, generator=gen).half()
    ids = torch.from_numpy(rng.permutation(n_exp)[:top_k].astype(np.int32))[None].to(dev)          # int32 ids
    wts = torch.softmax(torch.randn(top_k, device=dev, generator=gen), 0)[None].half()
    got = lbl.moe_forward_vec(x, ids, wts, gates, ups, downs, hidden, inter, q)
    assert got.dtype == torch.float16 and got.shape == (1, hidden)
    want = torch.zeros((1, hidden), dtype=torch.float64, device=dev)
    mag = torch.zeros((1, hidden), dtype=torch.float64, device=dev)
    for z in range(top_k):
        g, u, d = tiled[int(ids[0, z])]
        h = lbl.mlp_forward(x, g, u, 1, hidden, inter, lbl.SILU, q)
        o = lbl.lowbit_matmul(h, d, (hidden, inter), q, out_dtype=torch.float32).double()
        want += wts[0, z].double() * o
        mag += wts[0, z].double() * o.abs()
    # fp16 output rounding + possible 1-ulp differences of h between the two paths
    err = (got.double() - want).abs()
    assert bool((err <= 1.5e-3 * mag.abs().max() + 1e-6).all()), float(err.max() / mag.abs().max())


def test_moe_bad_ids_accumulate_determinism_launches(lbl, cuda_device):
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
    rng = np.random.default_rng(5300)
    hidden, inter, n_exp = 128, 128, 4
    ex = Experts(lbl, cuda_device, rng, n_exp, hidden, inter, "sym_int4")
    x_t = torch.from_numpy(rng.standard_normal((1, hidden)).astype(np.float32)).half().to(cuda_device)
    wts = torch.tensor([[0.5, 0.25, 0.125]], dtype=torch.float32, device=cuda_device)      # fp32 weights with fp16 x
    ids = torch.tensor([[2, 0, 3]], device=cuda_device)
    q = qid("sym_int4")
    outs = [lbl.moe_forward_vec(x_t, ids, wts, ex.gates, ex.ups, ex.downs, hidden, inter, q) for _ in range(3)]
    assert ggml.last_launch_count() == 2
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    # an id outside the table contributes zero (never dereferenced)
    bad = torch.tensor([[2, 99, 3]], device=cuda_device)
    ref = lbl.moe_forward_vec(x_t, torch.tensor([[2, 3]], device=cuda_device), wts[:, [0, 2]].contiguous(), ex.gates, ex.ups,
                              ex.downs, hidden, inter, q, out=torch.empty((1, hidden), dtype=torch.float32, device=cuda_device))
    got = lbl.moe_forward_vec(x_t, bad, wts, ex.gates, ex.ups, ex.downs, hidden, inter, q,
                              out=torch.empty((1, hidden), dtype=torch.float32, device=cuda_device))
    torch.testing.assert_close(got, ref, rtol=1e-5, atol=1e-6)
    neg = torch.tensor([[2, -1, 3]], device=cuda_device)
    got = lbl.moe_forward_vec(x_t, neg, wts, ex.gates, ex.ups, ex.downs, hidden, inter, q,
                              out=torch.empty((1, hidden), dtype=torch.float32, device=cuda_device))
    torch.testing.assert_close(got, ref, rtol=1e-5, atol=1e-6)
    # accumulate: the shared-expert sum of DeepSeek (final_out + shared_experts(identity))
    base = torch.from_numpy(rng.standard_normal((1, hidden)).astype(np.float32)).to(cuda_device)
    acc = base.clone()
    lbl.moe_forward_vec(x_t, ids, wts, ex.gates, ex.ups, ex.downs, hidden, inter, q, out=acc, accumulate=True)
    plain = lbl.moe_forward_vec(x_t, ids, wts, ex.gates, ex.ups, ex.downs, hidden, inter, q,
                                out=torch.empty((1, hidden), dtype=torch.float32, device=cuda_device))
    torch.testing.assert_close(acc, base + plain, rtol=0, atol=1e-6)
    # unsupported problems are refused, not approximated
    assert not lbl.use_fuse_moe(hidden, 96, 2, q)
    assert not lbl.use_fuse_moe(hidden, inter, 2, 999)
    with pytest.raises(Exception):
        lbl.moe_forward_vec(x_t, ids, wts, ex.gates, ex.ups, ex.downs, hidden, 96, q)
