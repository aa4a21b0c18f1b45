"""BASELINE.json's configurations at their FULL layer sizes (the oracle finishes only the small cases
in seconds, so these use size-independent checks):

  * the packed bytes are produced by kernel (a) on the GPU (bit-exact vs the oracle: test_gpu_parity);
  * the reference value of every matmul is  x.double() @ dequant(W).double().T  on the GPU, with
    dequant = the tiled-layout dequantise kernel (bit-exact vs the oracle: test_gpu_parity);
  * bound: decode kernels (M <= 32) |diff| <= rel * sum|x w| with rel = 2e-6 (summation order only;
    nf4 through its fp16 pair table: 2.6e-4); prefill GEMM (M > 32) rel = 1e-3 (weights rounded once
    to fp16).

Config 1 Llama-2-7B sym_int4 bs1; 2 Llama-3-8B asym_int4 / nf4, prefill 2048 + decode 1/4/16;
3 Mistral-7B fp8_e4m3 / fp6, decode 1..32; 4 Llama-2-70B sym_int4 TP shards; 5 Mixtral-8x7B MoE.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lbl(native_lib, cuda_device):
    from ipex_llm_b200.transformers import low_bit_linear
    return low_bit_linear


def qid(name):
    from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
    return ggml_tensor_qtype[name]


def make_weight(lbl, dev, n, k, qtype, seed):
    """random-init [n, k] weight (HF initializer_range 0.02) quantised on the GPU -> (tiled, dense fp64)"""
    gen = torch.Generator(device=dev).manual_seed(seed)
    w = torch.empty(n, k, device=dev, dtype=torch.float32).normal_(0.0, 0.02, generator=gen)
    tiled = lbl.FP4Params(w, qtype=qid(qtype)).to(dev).data
    del w
    dense = lbl.dequantize_device_layout(tiled, (n, k), qid(qtype)).double()
    return tiled, dense


def check_matmul(lbl, dev, tiled, dense, shape, qtype, M, xdt=torch.float16, seed=0):
    gen = torch.Generator(device=dev).manual_seed(100 + seed)
    x = torch.randn(M, shape[1], device=dev, generator=gen).to(xdt)
    y = lbl.lowbit_matmul(x, tiled, shape, qid(qtype), out_dtype=torch.float32).double()
    xd = x.double()
    want = xd @ dense.T
    mag = xd.abs() @ dense.abs().T
    gemm = lbl.takes_tensor_core_path(M, shape[0], shape[1], xdt)
    rel = 1e-3 if gemm else (2.6e-4 if qtype == "nf4" else 2e-6)
    if gemm and xdt != torch.float16:
        rel = 8e-3
    err = (y - want).abs()
    assert bool((err <= rel * mag + 1e-30).all()), (shape, qtype, M, float((err / (rel * mag + 1e-30)).max()))
    return y


# ---------------------------------------------------------------- config 1: Llama-2-7B sym_int4, batch 1
@pytest.mark.parametrize("shape", [(12288, 4096), (4096, 4096), (11008, 4096), (4096, 11008), (32000, 4096)])
def test_config1_llama2_7b_sym_int4_decode(lbl, cuda_device, shape):
    tiled, dense = make_weight(lbl, cuda_device, shape[0], shape[1], "sym_int4", seed=shape[0])
    check_matmul(lbl, cuda_device, tiled, dense, shape, "sym_int4", 1)


def test_config1_fused_mlp_full_size(lbl, cuda_device):
    dev = cuda_device
    tg, dg = make_weight(lbl, dev, 11008, 4096, "sym_int4", 1)
    tu, du = make_weight(lbl, dev, 11008, 4096, "sym_int4", 2)
    x = torch.randn(1, 4096, device=dev, generator=torch.Generator(device=dev).manual_seed(3)).half()
    out = torch.empty((1, 11008), dtype=torch.float32, device=dev)
    lbl.mlp_forward(x, tg, tu, 1, 4096, 11008, lbl.SILU, qid("sym_int4"), out=out)
    g, u = x.double() @ dg.T, x.double() @ du.T
    want = torch.nn.functional.silu(g) * u
    mg, mu = x.double().abs() @ dg.abs().T, x.double().abs() @ du.abs().T
    tol = 2e-6 * (1.13 * mg * u.abs() + torch.nn.functional.silu(g).abs() * mu) + 1e-6 * want.abs() + 1e-9
    assert bool(((out.double() - want).abs() <= tol).all())


# ---------------------------------------------------------------- config 2: Llama-3-8B asym_int4 / nf4
@pytest.mark.parametrize("qtype", ["asym_int4", "nf4"])
@pytest.mark.parametrize("shape", [(6144, 4096), (14336, 4096), (4096, 14336)])
def test_config2_llama3_8b_decode_and_prefill(lbl, cuda_device, qtype, shape):
    tiled, dense = make_weight(lbl, cuda_device, shape[0], shape[1], qtype, seed=shape[1])
    for M in (1, 4, 16):                      # decode batch 1 / 4 / 16
        check_matmul(lbl, cuda_device, tiled, dense, shape, qtype, M, seed=M)
    check_matmul(lbl, cuda_device, tiled, dense, shape, qtype, 2048, seed=9)          # 2048-token prefill (tcgen05 GEMM)
    check_matmul(lbl, cuda_device, tiled, dense, shape, qtype, 2048, xdt=torch.bfloat16, seed=10)


def test_config2_llama3_lm_head(lbl, cuda_device):
    shape = (128256, 4096)                    # 525 M weights: quantised and checked on the GPU only
    tiled, dense = make_weight(lbl, cuda_device, shape[0], shape[1], "asym_int4", seed=5)
    check_matmul(lbl, cuda_device, tiled, dense, shape, "asym_int4", 1)


# ---------------------------------------------------------------- config 3: Mistral-7B fp8_e4m3 / fp6, batch sweep
@pytest.mark.parametrize("qtype", ["fp8_e4m3", "fp6"])
def test_config3_mistral_batch_sweep(lbl, cuda_device, qtype):
    for shape in ((14336, 4096), (4096, 14336)):
        tiled, dense = make_weight(lbl, cuda_device, shape[0], shape[1], qtype, seed=7)
        for M in (1, 2, 4, 8, 16, 32):
            check_matmul(lbl, cuda_device, tiled, dense, shape, qtype, M, seed=M)


# ---------------------------------------------------------------- config 4: Llama-2-70B sym_int4, tensor parallel
@pytest.mark.parametrize("tp", [2, 8])
def test_config4_llama2_70b_tp_shards(lbl, cuda_device, tp):
    """column-parallel gate_proj (28672 x 8192) and row-parallel down_proj (8192 x 28672), sharded the way
    tensor_parallel.py does it, every rank's shard run on this one GPU: column shards concatenate to the full
    output, row-shard partial sums add up to it (what the all-reduce computes)."""
    from ipex_llm_b200.transformers import tensor_parallel as tpm
    dev = cuda_device
    q = qid("sym_int4")
    gen = torch.Generator(device=dev).manual_seed(40)
    x_h = torch.randn(1, 8192, device=dev, generator=gen).half()
    x_i = torch.randn(1, 28672, device=dev, generator=gen).half()
    for mode, shape, x in (("column", (28672, 8192), x_h), ("row", (8192, 28672), x_i)):
        w = torch.empty(shape, device=dev, dtype=torch.float32).normal_(0.0, 0.02, generator=gen)
        packed = lbl.ggml_convert_qtype(w, q)                          # canonical bytes, on the GPU
        del w
        full_tiled = lbl.ggml_q_format_convet_cpu2xpu(packed, shape[0] * shape[1], q, shape=shape)
        y_full = lbl.lowbit_matmul(x, full_tiled, shape, q, out_dtype=torch.float32)
        parts = []
        for rank in range(tp):
            if mode == "column":
                p, shp = tpm.shard_packed_column(packed, shape, q, rank, tp)
                xs = x
            else:
                p, shp = tpm.shard_packed_row(packed, shape, q, rank, tp)
                k0, k1 = tpm.shard_bounds(shape[1], tp, rank, unit=64)
                xs = x[:, k0:k1].contiguous()
            tiled = lbl.ggml_q_format_convet_cpu2xpu(p, shp[0] * shp[1], q, shape=shp)
            parts.append(lbl.lowbit_matmul(xs, tiled, shp, q, out_dtype=torch.float32))
        dense = lbl.dequantize_device_layout(full_tiled, shape, q).double()
        mag = x.double().abs() @ dense.abs().T
        if mode == "column":
            # same rows and weights; the K split across the cluster follows the row count, so only the
            # summation order may differ
            assert bool(((torch.cat(parts, dim=1).double() - y_full.double()).abs() <= 4e-6 * mag).all())
        else:
            total = torch.stack(parts).double().sum(0)
            assert bool(((total - y_full.double()).abs() <= 4e-6 * mag).all())


# ---------------------------------------------------------------- config 5: Mixtral-8x7B sym_int4 MoE
def test_config5_mixtral_moe_decode(lbl, cuda_device):
    """8 experts (w1/w3 14336 x 4096, w2 4096 x 14336), top-2: the fused one-token MoE against the routed
    experts evaluated in float64 from the dequantised weights."""
    dev = cuda_device
    q = qid("sym_int4")
    hidden, inter, n_exp, top_k = 4096, 14336, 8, 2
    experts = [[make_weight(lbl, dev, *shape, "sym_int4", seed=50 + 3 * e + i)
                for i, shape in enumerate(((inter, hidden), (inter, hidden), (hidden, inter)))] for e in range(n_exp)]
    gates, ups, downs = (torch.tensor([ex[i][0].data_ptr() for ex in experts], dtype=torch.uint64, device=dev) for i in range(3))
    gen = torch.Generator(device=dev).manual_seed(60)
    x = torch.randn(1, hidden, device=dev, generator=gen).half()
    logits = torch.randn(1, n_exp, device=dev, generator=gen)
    wts, ids = torch.topk(torch.softmax(logits, dim=-1), top_k, dim=-1)
    wts = (wts / wts.sum(-1, keepdim=True)).half()
    assert lbl.use_fuse_moe(hidden, inter, top_k, q)
    out = torch.empty((1, hidden), dtype=torch.float32, device=dev)
    lbl.moe_forward_vec(x, ids, wts, gates, ups, downs, hidden, inter, q, out=out)
    want = torch.zeros((1, hidden), dtype=torch.float64, device=dev)
    mag = torch.zeros((1, hidden), dtype=torch.float64, device=dev)
    for z in range(top_k):
        (_, dg), (_, du), (_, dd) = experts[int(ids[0, z])]
        h = (torch.nn.functional.silu(x.double() @ dg.T) * (x.double() @ du.T)).half().double()   # kernel keeps h in fp16
        want += wts[0, z].double() * (h @ dd.T)
        mag += wts[0, z].double() * (h.abs() @ dd.abs().T)
    # one fp16 rounding of h may differ by an ulp (2^-11) where the fused value sits next to a rounding boundary
    assert bool(((out.double() - want).abs() <= 6e-4 * mag + 1e-9).all())


def test_config5_mixtral_moe_tp_inside_experts(lbl, cuda_device):
    """SURVEY 8e, MoE sharding (1): tensor parallelism inside every expert -- w1 / w3 column-sharded, w2 row-sharded
    (14336 / 8 = 1792 = 28 blocks), the ranks' outputs summed by the same all-reduce as the dense layers.  Every
    rank's fused one-token MoE is run on this one GPU; the sum over ranks must equal the unsharded MoE up to the
    summation order and the fp16 rounding of the intermediate activations."""
    from ipex_llm_b200.transformers import tensor_parallel as tpm
    dev = cuda_device
    q = qid("sym_int4")
    hidden, inter, n_exp, top_k, tp = 4096, 14336, 4, 2, 8
    gen = torch.Generator(device=dev).manual_seed(70)
    packed = []                                          # canonical packed weights per expert: w1, w3 [inter, hidden], w2 [hidden, inter]
    for e in range(n_exp):
        ws = []
        for shape in ((inter, hidden), (inter, hidden), (hidden, inter)):
            w = torch.empty(shape, device=dev, dtype=torch.float32).normal_(0.0, 0.02, generator=gen)
            ws.append(lbl.ggml_convert_qtype(w, q))
            del w
        packed.append(ws)
    x = torch.randn(1, hidden, device=dev, generator=gen).half()
    logits = torch.randn(1, n_exp, device=dev, generator=gen)
    wts, ids = torch.topk(torch.softmax(logits, dim=-1), top_k, dim=-1)
    wts = (wts / wts.sum(-1, keepdim=True)).half()

    def run(shards, inter_local):
        keep = [[lbl.ggml_q_format_convet_cpu2xpu(p, shp[0] * shp[1], q, shape=shp) for p, shp in ex] for ex in shards]
        tabs = [torch.tensor([ex[i].data_ptr() for ex in keep], dtype=torch.uint64, device=dev) for i in range(3)]
        assert lbl.use_fuse_moe(hidden, inter_local, top_k, q)
        out = torch.empty((1, hidden), dtype=torch.float32, device=dev)
        lbl.moe_forward_vec(x, ids, wts, tabs[0], tabs[1], tabs[2], hidden, inter_local, q, out=out)
        torch.cuda.synchronize()
        return out

    full = run([[(p, s) for p, s in zip(ex, ((inter, hidden), (inter, hidden), (hidden, inter)))] for ex in packed], inter)
    total = torch.zeros_like(full, dtype=torch.float64)
    for rank in range(tp):
        shards = []
        for ex in packed:
            g = tpm.shard_packed_column(ex[0], (inter, hidden), q, rank, tp)
            u = tpm.shard_packed_column(ex[1], (inter, hidden), q, rank, tp)
            d = tpm.shard_packed_row(ex[2], (hidden, inter), q, rank, tp)
            shards.append([g, u, d])
        assert shards[0][0][1] == (inter // tp, hidden) and shards[0][2][1] == (hidden, inter // tp)
        total += run(shards, inter // tp).double()
    scale = float(full.abs().max())
    assert float((total - full.double()).abs().max()) <= 2e-3 * scale
