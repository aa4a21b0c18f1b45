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


def bound64(packed, shape, qtype, x):
    wd = np.abs(oracle.dequantize(packed, shape, qtype)).astype(np.float64)
    return 2e-6 * (np.abs(x).astype(np.float64) @ wd.T) + 1e-30


def mlp_expect(pg, pu, shape, qtype, x, act):
    g = oracle.matmul_f64(pg, shape, qtype, x)
    u = oracle.matmul_f64(pu, shape, qtype, x)
    bg, bu = bound64(pg, shape, qtype, x), bound64(pu, shape, qtype, x)
    a = act64(g, act)
    want = a * u
    # |d(act(g) u)| <= L |dg| |u| + |act(g)| |du| + |dg||du| L, plus fp32 rounding of act and the product
    tol = 1.13 * bg * (np.abs(u) + bu) + np.abs(a) * bu + 4e-7 * np.abs(want) + 4e-7 * np.abs(u) * (np.abs(g) + 1) + 1e-30
    return want, tol


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "fp8_e4m3", "fp8_e5m2", "fp6", "nf4"])
@pytest.mark.parametrize("xdt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,act", [(1, 0), (4, 1), (8, 0)])
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
    want, tol = mlp_expect(pg, pu, (N, K), qtype, x, act)
    got = out.cpu().numpy().astype(np.float64)
    assert (np.abs(got - want) <= tol).all(), float((np.abs(got - want) / tol).max())


@pytest.mark.parametrize("shape,M", [((11008, 4096), 1), ((14336, 4096), 2), ((1000, 704), 3), ((256, 2048), 12),
                                     ((256, 2048), 48), ((64, 64), 1)])
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
    assert float((out.double() - want).abs().max()) <= 2e-6 * scale + 1e-6 * float(u.abs().max())
    # default output dtype = activation dtype, same values rounded once
    out16 = lbl.mlp_forward(x_t, tg, tu, M, K, N, lbl.SILU, qid(qtype))
    assert out16.dtype == torch.float16 and torch.equal(out16, out.half())


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
