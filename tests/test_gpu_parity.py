"""GPU parity tests: the sm_100a kernels behind the C ABI against the CPU oracle.

Bars:
  * packed bytes (quantise, relayout round trip): bit-exact;
  * dequantise: bit-exact (same fp32 formula, no contraction);
  * matmul vs mode-B oracle: |diff| <= 2e-6 * sum_k |x_k w_k| + 1e-30 for every element when the
    activations are exactly representable (the only difference is fp32 summation order), checked
    against the float64 oracle with the same bound for the oracle's own serial fp32 sum.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

QTYPES = ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"]
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def lbl(native_lib, cuda_device):
    from ipex_llm_b200.transformers import low_bit_linear
    return low_bit_linear


def qid(name):
    from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
    return ggml_tensor_qtype[name]


def rand_w(rng, rows, k, kind="normal"):
    if kind == "normal":
        return (rng.standard_normal((rows, k)) * 0.02).astype(np.float32)
    if kind == "wide":          # wide dynamic range across blocks, outliers inside blocks
        w = rng.standard_normal((rows, k)) * np.exp(rng.standard_normal((rows, k // 64, 1)).repeat(64, 2).reshape(rows, k) * 2)
        w[rng.random((rows, k)) < 0.01] *= 20
        return w.astype(np.float32)
    raise ValueError(kind)


def as_np_u8(t):
    return (t.view(torch.uint8) if t.dtype != torch.uint8 else t).cpu().numpy()


# ---------------------------------------------------------------- (a) quantise / pack
@pytest.mark.parametrize("qtype", QTYPES)
@pytest.mark.parametrize("kind", ["normal", "wide"])
def test_quantize_bit_exact_host_abi(lbl, qtype, kind):
    rng = np.random.default_rng(100 + len(qtype))
    w = rand_w(rng, 96, 512, kind)
    w[0, :64] = 0.0
    w[1, 64:128] = -0.0
    w[2, :64] = 0.017
    w[3, 5] = 3.0
    w[4, 70] = -3.0
    w[5, :64] = np.linspace(-1, 1, 64)          # exact ties at .5 steps for some formats
    got = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w), qid(qtype), device="cpu"))
    want = oracle.quantize(w, qtype)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("qtype", QTYPES)
def test_quantize_bit_exact_device_abi(lbl, cuda_device, qtype):
    rng = np.random.default_rng(200 + len(qtype))
    w = rand_w(rng, 257, 1024, "wide")           # ragged row count
    got = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w).to(cuda_device), qid(qtype)))
    np.testing.assert_array_equal(got, oracle.quantize(w, qtype))


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"])
@pytest.mark.parametrize("rows", [1, 7, 9, 33])
def test_quantize_four_lanes_per_block_edge_cases(lbl, cuda_device, qtype, rows):
    """the device path of the 4-bit formats and of fp8 without scale search (four lanes per block, eight blocks per warp instruction): ragged block counts
    (tail groups), all-zero and -0 blocks, the extremum in every lane position and half, equal |x| of both signs (the first
    one decides the scale), exact .5 ties, a constant block, one huge element; and an unaligned source (the one-warp-per-block
    kernel takes over) -- bytes equal to the oracle's"""
    rng = np.random.default_rng(250 + rows)
    k = 192
    w = rand_w(rng, rows, k, "wide")
    blocks = w.reshape(-1, 64)
    nb = blocks.shape[0]
    blocks[0 % nb] = 0.0
    blocks[1 % nb] = -0.0
    for j in range(2, nb):                                   # the maximum walks over all 64 positions, both signs
        pos = (5 * j) % 64
        blocks[j, pos] = 4.0 * (-1.0) ** j
        if j % 3 == 0:                                       # ... with an equal magnitude of the other sign later or earlier
            blocks[j, (pos + 17) % 64] = -blocks[j, pos]
        if j % 4 == 1:
            blocks[j, (pos + 32) % 64] = blocks[j, pos]
    if nb > 5:
        blocks[5] = np.linspace(-1, 1, 64)                   # exact ties at .5 steps for some formats
    if nb > 6:
        blocks[6] = 0.25
    if nb > 12:                                              # a zero extremum: the first of -0 / +0 in element order is kept
        blocks[9] = np.abs(blocks[9]) + 0.1; blocks[9, 3] = 0.0; blocks[9, 40] = -0.0
        blocks[10] = np.abs(blocks[10]) + 0.1; blocks[10, 35] = -0.0; blocks[10, 50] = 0.0
        blocks[11] = -np.abs(blocks[11]) - 0.1; blocks[11, 7] = -0.0; blocks[11, 8] = 0.0
        blocks[12] = -np.abs(blocks[12]) - 0.1; blocks[12, 60] = 0.0; blocks[12, 61] = -0.0
    w = blocks.reshape(rows, k).copy()
    want = oracle.quantize(w, qtype)
    got = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w).to(cuda_device), qid(qtype)))
    np.testing.assert_array_equal(got, want)
    # source not 16-byte aligned
    pad = torch.zeros(rows * k + 1, device=cuda_device)
    pad[1:] = torch.from_numpy(w).to(cuda_device).flatten()
    got2 = as_np_u8(lbl.ggml_convert_qtype(pad[1:].view(rows, k), qid(qtype)))
    np.testing.assert_array_equal(got2, want)


@pytest.mark.parametrize("amax", [1.0, 3.0, 0.7, 2.0 ** -100, 2.0 ** -125, 1e-40, 2.0 ** 100, 3.0e38])
def test_quantize_nf4_thresholds(lbl, cuda_device, amax):
    """nf4 codes come from a 32-bin table on the device and from fifteen compares in the oracle: values at, one ulp below
    and one ulp above every midpoint between table entries and every bin edge (k/16), scaled by block maxima over the
    whole exponent range (subnormal scale, overflowing reciprocal), plus NaN / inf elements -- same bytes"""
    tab = np.array([-1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453, -0.28444138169288635,
                    -0.18477343022823334, -0.09105003625154495, 0.0, 0.07958029955625534, 0.16093020141124725,
                    0.24611230194568634, 0.33791524171829224, 0.44070982933044434, 0.5626170039176941,
                    0.7229568362236023, 1.0], dtype=np.float32)
    mids = ((tab[:-1] + tab[1:]).astype(np.float32) * np.float32(0.5)).astype(np.float32)
    pts = np.concatenate([mids, np.arange(-16, 17, dtype=np.float32) / np.float32(16)])
    pts = np.concatenate([pts, np.nextafter(pts, np.float32(-2)), np.nextafter(pts, np.float32(2)),
                          np.nextafter(np.nextafter(pts, np.float32(2)), np.float32(2))]).astype(np.float32)
    rng = np.random.default_rng(77)
    with np.errstate(over="ignore", under="ignore"):
        vals = (pts.astype(np.float64) * amax).astype(np.float32)
        vals = np.concatenate([vals, (rng.uniform(-1, 1, 4096) * amax).astype(np.float32)])
    vals = np.clip(vals, -np.float32(amax), np.float32(amax))
    nb = (vals.size + 61) // 62
    blocks = np.zeros((nb + 3, 64), dtype=np.float32)
    flat = np.resize(vals, nb * 62).reshape(nb, 62)
    blocks[:nb, 1:63] = flat
    blocks[:nb, 0] = np.float32(amax)                        # the block maximum, first and last so that both halves hold one
    blocks[:nb, 63] = -np.float32(amax)
    blocks[nb, :] = flat[0, 0]; blocks[nb, 5] = np.nan       # a NaN element, an inf element, an all-inf block
    blocks[nb + 1, :] = flat[0, 1]; blocks[nb + 1, 40] = np.inf
    blocks[nb + 2, :] = -np.inf
    rows = blocks.shape[0]
    w = blocks.reshape(rows, 64).copy()
    want = oracle.quantize(w, "nf4")
    got = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w).to(cuda_device), qid("nf4")))
    np.testing.assert_array_equal(got, want)
    pad = torch.zeros(rows * 64 + 1, device=cuda_device)      # unaligned source: the one-warp-per-block kernel (compare chain)
    pad[1:] = torch.from_numpy(w).to(cuda_device).flatten()
    got2 = as_np_u8(lbl.ggml_convert_qtype(pad[1:].view(rows, 64), qid("nf4")))
    np.testing.assert_array_equal(got2, want)


@pytest.mark.parametrize("qtype", ["fp8_e4m3", "fp8_e5m2", "fp6"])
def test_quantize_scale_search_bit_exact(lbl, cuda_device, qtype):
    rng = np.random.default_rng(300)
    w = rand_w(rng, 64, 512, "wide")
    got = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w).to(cuda_device), qid(qtype), enable_scale_search=True))
    np.testing.assert_array_equal(got, oracle.quantize(w, qtype, scale_search=True))


def test_quantize_reference_golden_vectors(lbl):
    """the committed vectors made by running the reference's own Python quantisers"""
    g = np.load(os.path.join(GOLD, "ref_python_q4.npz"))
    w = g["w"]
    got0 = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w), qid("sym_int4"), device="cpu")).reshape(-1, 34)
    want0 = g["q4_0_bytes"].reshape(-1, 34)
    nz = w.reshape(-1, 64).any(axis=1)
    np.testing.assert_array_equal(got0[nz], want0[nz])       # all-zero block: see test_oracle_golden
    got1 = as_np_u8(lbl.ggml_convert_qtype(torch.from_numpy(w), qid("asym_int4"), device="cpu"))
    np.testing.assert_array_equal(got1, g["q4_1_bytes"])


def test_quantize_histogram(native_lib, cuda_device):
    rng = np.random.default_rng(5)
    w = rng.standard_normal((32, 256)).astype(np.float32)
    dst = np.empty(w.size // 64 * 34, np.uint8)
    hist = (ctypes.c_int64 * 16)()
    n = native_lib.ggml_quantize_tensor(ctypes.cast(w.ctypes.data, ctypes.POINTER(ctypes.c_float)),
                                        ctypes.c_void_p(dst.ctypes.data), 2, w.size, 256, hist, False)
    assert n == dst.size
    _, want = oracle.quantize(w, "sym_int4", return_hist=True)
    np.testing.assert_array_equal(np.array(list(hist)), want)


def test_quantize_rejects_bad_shapes(lbl):
    with pytest.raises(RuntimeError):
        lbl.ggml_convert_qtype(torch.zeros(4, 96), qid("sym_int4"), device="cpu")
    with pytest.raises(RuntimeError):
        lbl.ggml_convert_qtype(torch.zeros(4, 128, dtype=torch.float16), qid("sym_int4"), device="cpu")
    assert lbl.ggml_convert_qtype(torch.zeros(0, 128), qid("sym_int4"), device="cpu").numel() == 0


# ---------------------------------------------------------------- dequantise + relayout
@pytest.mark.parametrize("qtype", QTYPES)
def test_dequantize_bit_exact(lbl, cuda_device, qtype):
    rng = np.random.default_rng(400)
    w = rand_w(rng, 40, 320, "wide")
    packed = oracle.quantize(w, qtype)
    want = oracle.dequantize(packed, w.shape, qtype)
    got_host = lbl.ggml_convert_fp32(torch.from_numpy(packed), w.shape, w.size, qid(qtype)).numpy()
    np.testing.assert_array_equal(got_host, want)
    got_dev = lbl.ggml_convert_fp32(torch.from_numpy(packed).to(cuda_device), w.shape, w.size, qid(qtype)).cpu().numpy()
    np.testing.assert_array_equal(got_dev, want)
    if qtype == "sym_int4":
        np.testing.assert_array_equal(lbl.ggml_int4_convert_fp32(torch.from_numpy(packed), w.shape, w.size).numpy(), want)


@pytest.mark.parametrize("qtype", QTYPES)
@pytest.mark.parametrize("rows", [16, 40, 7, 256])
def test_relayout_roundtrip_and_device_dequant(lbl, cuda_device, qtype, rows):
    rng = np.random.default_rng(500 + rows)
    k = 384
    w = rand_w(rng, rows, k, "wide")
    packed = torch.from_numpy(oracle.quantize(w, qtype))
    if qtype == "nf4":
        packed = packed.view(torch.bfloat16)
    n = rows * k
    # host ABI (reference-shaped symbols) and device ABI give the same tiled bytes
    tiled_h = lbl.ggml_q_format_convet_cpu2xpu(packed, n, qid(qtype), shape=(rows, k))
    tiled_d = lbl.ggml_q_format_convet_cpu2xpu(packed.to(cuda_device), n, qid(qtype), shape=(rows, k))
    assert tiled_h.shape == packed.shape and tiled_h.dtype == packed.dtype
    np.testing.assert_array_equal(as_np_u8(tiled_h), as_np_u8(tiled_d))
    back = lbl.ggml_q_format_convet_xpu2cpu(tiled_d, n, qid(qtype), shape=(rows, k))
    np.testing.assert_array_equal(as_np_u8(back), as_np_u8(packed))
    back_h = lbl.ggml_q_format_convet_xpu2cpu(tiled_h, n, qid(qtype), shape=(rows, k))
    np.testing.assert_array_equal(as_np_u8(back_h), as_np_u8(packed))
    # dequantising the tiled layout gives the oracle's values bit for bit
    deq = lbl.dequantize_device_layout(tiled_d, (rows, k), qid(qtype)).cpu().numpy()
    np.testing.assert_array_equal(deq, oracle.dequantize(as_np_u8(packed), (rows, k), qtype))


@pytest.mark.parametrize("qtype", QTYPES)
@pytest.mark.parametrize("rows,k", [(200, 1024), (64, 512), (136, 256), (320, 128), (192, 192), (128, 2048 + 1024)])
def test_relayout_staged_equals_general(lbl, cuda_device, monkeypatch, qtype, rows, k):
    """whole 64-row groups with an even block count go through the shared-memory staged relayout kernels (16 / 8 / 4 / 2
    k-blocks per CTA by divisibility), tail rows and odd block counts through the byte-per-thread kernels: both give the
    same tiled bytes, the way back restores the records, and the 8-elements-per-thread dequantise of either layout
    equals the oracle"""
    rng = np.random.default_rng(rows + k)
    w = rand_w(rng, rows, k, "wide")
    packed_np = oracle.quantize(w, qtype)
    packed = torch.from_numpy(packed_np).to(cuda_device)
    if qtype == "nf4":
        packed = packed.view(torch.bfloat16)
    n = rows * k
    tiled = lbl.ggml_q_format_convet_cpu2xpu(packed, n, qid(qtype), shape=(rows, k))
    back = lbl.ggml_q_format_convet_xpu2cpu(tiled, n, qid(qtype), shape=(rows, k))
    monkeypatch.setenv("LOWBIT_RELAYOUT_GENERAL", "1")
    tiled_g = lbl.ggml_q_format_convet_cpu2xpu(packed, n, qid(qtype), shape=(rows, k))
    back_g = lbl.ggml_q_format_convet_xpu2cpu(tiled_g, n, qid(qtype), shape=(rows, k))
    monkeypatch.delenv("LOWBIT_RELAYOUT_GENERAL")
    np.testing.assert_array_equal(as_np_u8(tiled), as_np_u8(tiled_g))
    np.testing.assert_array_equal(as_np_u8(back), packed_np.reshape(-1))
    np.testing.assert_array_equal(as_np_u8(back_g), packed_np.reshape(-1))
    want = oracle.dequantize(packed_np, (rows, k), qtype)
    for dt in (torch.float32, torch.float16, torch.bfloat16):
        deq = lbl.dequantize_device_layout(tiled, (rows, k), qid(qtype), dtype=dt)
        assert torch.equal(deq.cpu(), torch.from_numpy(want).to(dt))
    got = lbl.ggml_convert_fp32(packed, w.shape, w.size, qid(qtype)).cpu().numpy()
    np.testing.assert_array_equal(got, want)


def test_relayout_needs_row_length(native_lib):
    src = np.zeros(34 * 16, np.uint8); dst = np.zeros_like(src)
    native_lib._lib.b200_lowbit_set_row_length(0)
    with pytest.raises(RuntimeError, match="row"):
        native_lib.ggml_q_format_convet_cpu2xpu(ctypes.c_void_p(src.ctypes.data), ctypes.c_void_p(dst.ctypes.data), 16 * 64, 2)


# ---------------------------------------------------------------- (b) matmul
def matmul_bound(packed, shape, qtype, x, xdt=None):
    """per-element bound for fp32 summation-order differences: 2e-6 * sum |x w| (+ tiny).
    nf4 on full 64-row groups goes through a table of fp16 pairs (fp16 activations: every table value
    rounded to fp16, 2^-12 relative) or of two bf16 parts (bf16 / fp32 activations: 2^-17)."""
    wd = np.abs(oracle.dequantize(packed, shape, qtype)).astype(np.float64)
    rel = 2e-6
    if qtype == "nf4" and xdt is not None:
        rel = 2.6e-4 if xdt == torch.float16 else 1e-5
    return rel * (np.abs(x).astype(np.float64) @ wd.T) + 1e-30


def run_matmul(lbl, dev, packed, shape, qtype, x_t, **kw):
    p = torch.from_numpy(packed)
    if qtype == "nf4":
        p = p.view(torch.bfloat16)
    tiled = lbl.ggml_q_format_convet_cpu2xpu(p.to(dev), shape[0] * shape[1], qid(qtype), shape=shape)
    return lbl.lowbit_matmul(x_t.to(dev), tiled, shape, qid(qtype), **kw)


@pytest.mark.parametrize("qtype", QTYPES)
@pytest.mark.parametrize("xdt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M", [1, 3, 8])
def test_gemv_matches_oracle(lbl, cuda_device, qtype, xdt, M):
    rng = np.random.default_rng(600 + M)
    N, K = 208, 1024
    w = rand_w(rng, N, K)
    packed = oracle.quantize(w, qtype)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(xdt)
    x = x_t.float().numpy()                                        # the values the kernel sees
    y = run_matmul(lbl, cuda_device, packed, (N, K), qtype, x_t, out_dtype=torch.float32).cpu().numpy()
    y64 = oracle.matmul_f64(packed, (N, K), qtype, x)
    bound = matmul_bound(packed, (N, K), qtype, x, xdt)
    assert (np.abs(y - y64) <= bound).all(), float((np.abs(y - y64) / bound).max())
    # the oracle's own serial fp32 sum sits inside the plain summation-order bound
    assert (np.abs(oracle.matmul(packed, (N, K), qtype, x) - y64) <= matmul_bound(packed, (N, K), qtype, x)).all()


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "fp8_e4m3"])
@pytest.mark.parametrize("shape,M", [((16, 64), 1), ((4096, 4096), 1), ((11008, 4096), 2), ((4096, 11008), 1),
                                      ((1000, 704), 5), ((8, 4096), 1), ((250, 128), 12), ((512, 2048), 32),
                                      ((512, 2048), 40), ((1000, 704), 70)])
def test_gemv_shapes(lbl, cuda_device, qtype, shape, M):
    rng = np.random.default_rng(700)
    N, K = shape
    w = rand_w(rng, N, K)
    packed = oracle.quantize(w, qtype)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half()
    x = x_t.float().numpy()
    y = run_matmul(lbl, cuda_device, packed, shape, qtype, x_t, out_dtype=torch.float32).cpu().numpy()
    y64 = oracle.matmul_f64(packed, shape, qtype, x)
    bound = matmul_bound(packed, shape, qtype, x)
    if M >= 17 and N % 128 == 0:
        bound = bound * 500          # tensor-core paths (17..64 rows: skinny kernel, beyond: prefill GEMM): weights rounded once to fp16 (1e-3 * sum|x w|)
    assert (np.abs(y - y64) <= bound).all(), float((np.abs(y - y64) / bound).max())


def test_gemv_deterministic_and_counters_reset(lbl, cuda_device):
    rng = np.random.default_rng(800)
    N, K = 4096, 4096
    packed = oracle.quantize(rand_w(rng, N, K), "sym_int4")
    x_t = torch.from_numpy(rng.standard_normal((1, K)).astype(np.float32)).half()
    ys = [run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32) for _ in range(5)]
    for y in ys[1:]:
        assert torch.equal(y, ys[0])


def test_gemv_epilogue_bias_accumulate_dtypes(lbl, cuda_device):
    rng = np.random.default_rng(900)
    N, K, M = 320, 512, 4
    packed = oracle.quantize(rand_w(rng, N, K), "sym_int4")
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half()
    base = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32)
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(cuda_device)
    yb = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32, bias=bias)
    torch.testing.assert_close(yb, base + bias, rtol=0, atol=1e-6)
    res = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).to(cuda_device)
    out = res.clone()
    run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out=out, accumulate=True)
    torch.testing.assert_close(out, res + base, rtol=0, atol=1e-6)
    for dt in (torch.float16, torch.bfloat16):
        y = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=dt)
        assert y.dtype == dt
        torch.testing.assert_close(y.float(), base.to(dt).float(), rtol=0, atol=0)


@pytest.mark.parametrize("qtype", QTYPES)
def test_reference_host_abi_matmul(lbl, qtype):
    """ggml_compute_forward_mul_mat_q_fp32 with host pointers, as LowBitLinear's CPU branch calls it"""
    rng = np.random.default_rng(1000)
    N, K, M = 96, 256, 3
    packed = oracle.quantize(rand_w(rng, N, K), qtype)
    x = rng.standard_normal((M, K)).astype(np.float32)
    p = torch.from_numpy(packed)
    y = lbl.ggml_matmul_src1_x_src0_t(p.view(torch.bfloat16) if qtype == "nf4" else p, torch.from_numpy(x), (N, K), qid(qtype)).numpy()
    y64 = oracle.matmul_f64(packed, (N, K), qtype, x)
    # fp32 activations go through the 3-way bf16 split: exact to ~2^-24 relative per term
    bound = matmul_bound(packed, (N, K), qtype, x, torch.float32)
    assert (np.abs(y - y64) <= bound).all(), float((np.abs(y - y64) / bound).max())


# ---------------------------------------------------------------- module level
@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e5m2", "fp6"])
def test_lowbit_linear_module(lbl, cuda_device, qtype):
    torch.manual_seed(0)
    lin = torch.nn.Linear(512, 272, bias=True)
    q = lbl.LowBitLinear(512, 272, qid(qtype), bias=True)
    q.weight = lbl.FP4Params(lin.weight.data.clone(), qtype=qid(qtype))
    q.bias = torch.nn.Parameter(lin.bias.data.clone())
    q = q.to(cuda_device)
    assert q.weight.quantized and q.weight.data.device.type == "cuda" and tuple(q.weight._shape) == (272, 512)
    assert q.weight.data.dtype == (torch.bfloat16 if qtype == "nf4" else torch.uint8)
    packed = oracle.quantize(lin.weight.data.numpy(), qtype)
    x = torch.randn(2, 3, 512)
    want = torch.from_numpy(oracle.matmul(packed, (272, 512), qtype, x.half().float().view(-1, 512).numpy())).view(2, 3, 272) + lin.bias.data
    y = q(x.half().to(cuda_device))
    assert y.dtype == torch.float16 and y.shape == (2, 3, 272)
    torch.testing.assert_close(y.float().cpu(), want, rtol=2e-3, atol=2e-3)
    # dtype casts leave the packed weight alone; .to('cpu') restores the canonical bytes
    before = q.weight.data.clone()
    q = q.to(torch.float16)
    assert torch.equal(q.weight.data.view(torch.uint8), before.view(torch.uint8))
    qc = q.to("cpu")
    np.testing.assert_array_equal(as_np_u8(qc.weight.data), packed)
    # CPU-resident weight: the reference's CPU call path still computes (on the GPU, via the host ABI)
    if qtype in ("sym_int4", "asym_int4", "fp6"):
        yc = qc(x)
        torch.testing.assert_close(yc, torch.from_numpy(oracle.matmul(packed, (272, 512), qtype, x.view(-1, 512).numpy())).view(2, 3, 272) + lin.bias.data.float(), rtol=1e-4, atol=1e-4)


# ---------------------------------------------------------------- tensor parallel shards (one GPU)
@pytest.mark.parametrize("qtype", ["sym_int4", "fp8_e5m2"])
def test_tp_shards_reproduce_full_result(lbl, cuda_device, qtype):
    """column shards concatenate, row shards sum to the unsharded output (the all-reduce the
    row-parallel LowBitLinear issues over NCCL is emulated by a local sum of the rank partials)"""
    from ipex_llm_b200.transformers import tensor_parallel as tp
    rng = np.random.default_rng(1100)
    N, K, M, world = 256, 1024, 2, 4
    w = rand_w(rng, N, K)
    packed = torch.from_numpy(oracle.quantize(w, qtype))
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half()
    full = run_matmul(lbl, cuda_device, oracle.quantize(w, qtype), (N, K), qtype, x_t, out_dtype=torch.float32)
    cols, rows = [], torch.zeros(M, N, device=cuda_device)
    for rank in range(world):
        lin = tp.make_parallel_linear(packed, (N, K), qid(qtype), "column", rank, world, device=cuda_device)
        cols.append(lin(x_t.to(cuda_device)).float())
        lin = tp.make_parallel_linear(packed, (N, K), qid(qtype), "row", rank, world, device=cuda_device)
        k0, k1 = rank * K // world, (rank + 1) * K // world
        rows += lbl.lowbit_matmul(x_t[:, k0:k1].contiguous().to(cuda_device), lin.weight.data, lin.weight_shape,
                                  lin.qtype, out_dtype=torch.float32)
    torch.testing.assert_close(torch.cat(cols, dim=1), full.half().float(), rtol=0, atol=0)
    torch.testing.assert_close(rows, full, rtol=1e-5, atol=1e-6)


# ---------------------------------------------------------------- (c) prefill GEMM on tcgen05
@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"])
@pytest.mark.parametrize("xdt,tol", [(torch.float16, 1e-3), (torch.bfloat16, 8e-3), (torch.float32, 8e-3)])
@pytest.mark.parametrize("shape,M", [((256, 512), 64), ((128, 64), 33), ((384, 1024), 300)])
def test_gemm_prefill_matches_oracle(lbl, cuda_device, qtype, xdt, tol, shape, M):
    """M > 32 goes to the tensor-core kernel: dequantised weights are rounded once to fp16 (bf16 for
    bf16 / fp32 activations, which are themselves taken in bf16): |diff| <= tol * sum|x w|"""
    rng = np.random.default_rng(1200 + M)
    N, K = shape
    w = rand_w(rng, N, K)
    packed = oracle.quantize(w, qtype)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(xdt)
    x = x_t.float().numpy()
    y = run_matmul(lbl, cuda_device, packed, shape, qtype, x_t, out_dtype=torch.float32).cpu().numpy()
    y64 = oracle.matmul_f64(packed, shape, qtype, x)
    wd = np.abs(oracle.dequantize(packed, shape, qtype)).astype(np.float64)
    bound = tol * (np.abs(x).astype(np.float64) @ wd.T) + 1e-30
    assert np.isfinite(y).all()
    assert (np.abs(y - y64) <= bound).all(), float((np.abs(y - y64) / bound).max())


def test_gemm_prefill_persistent_many_tiles(lbl, cuda_device):
    """more tiles than SMs: every persistent CTA walks several tiles (both TMEM accumulator buffers, ragged
    last token tile), checked against the fp32 product of the oracle-dequantised weights on the GPU"""
    rng = np.random.default_rng(1300)
    N, K, M = 128 * 44, 192, 1000                    # 44 feature tiles x 4 token tiles = 176 tiles of 256 tokens
    w = rand_w(rng, N, K)
    packed = oracle.quantize(w, "asym_int4")
    wd = torch.from_numpy(oracle.dequantize(packed, (N, K), "asym_int4")).to(cuda_device)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half()
    y = run_matmul(lbl, cuda_device, packed, (N, K), "asym_int4", x_t, out_dtype=torch.float32)
    want = x_t.to(cuda_device).float() @ wd.T
    bound = 1e-3 * (x_t.to(cuda_device).float().abs() @ wd.abs().T) + 1e-30
    assert bool(((y - want).abs() <= bound).all()), float(((y - want).abs() / bound).max())
    # bias + accumulate + fp16 output through the vector epilogue
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).half().to(cuda_device)
    res = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).half().to(cuda_device)
    out = res.clone()
    run_matmul(lbl, cuda_device, packed, (N, K), "asym_int4", x_t, out=out, bias=bias, accumulate=True)
    want2 = (want + bias.float() + res.float())
    assert bool(((out.float() - want2).abs() <= bound + 2e-3 * want2.abs() + 1e-3).all())


@pytest.mark.parametrize("shape,M,qtype", [((4096, 11008), 100, "asym_int4"), ((4096, 4096), 33, "sym_int4"), ((11008, 4096), 128, "fp8_e4m3"),
                                           ((1024, 2048 + 64), 57, "nf4")])
def test_gemm_small_batch_split_k(lbl, cuda_device, shape, M, qtype):
    """one token tile and few feature tiles: the K range of every output tile is split over several CTAs (fp32
    partial sums + a reduction launch); uneven splits (172 and 33 k-blocks), bias, accumulate, fp16 output"""
    rng = np.random.default_rng(1400 + M)
    N, K = shape
    dev = cuda_device
    w = torch.from_numpy(rand_w(rng, N, K)).to(dev)
    tiled = lbl.FP4Params(w, qtype=qid(qtype)).to(dev).data
    wd = lbl.dequantize_device_layout(tiled, (N, K), qid(qtype)).float()
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half().to(dev)
    y = lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out_dtype=torch.float32)
    if shape == (4096, 11008):
        assert lbl.ggml.last_launch_count() == 2      # GEMM + reduction of the partial sums (other shapes: the cost model decides)
    want = x_t.double() @ wd.double().T
    bound = 1e-3 * (x_t.double().abs() @ wd.double().abs().T) + 1e-30
    assert bool(((y.double() - want).abs() <= bound).all()), float(((y.double() - want).abs() / bound).max())
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).half().to(dev)
    res = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).half().to(dev)
    out = res.clone()
    lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out=out, bias=bias, accumulate=True)
    want2 = y + bias.float() + res.float()
    assert bool(((out.float() - want2).abs() <= 1e-3 * want2.abs() + 1e-3).all())


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"])
@pytest.mark.parametrize("xdt,tol", [(torch.float16, 1e-3), (torch.bfloat16, 8e-3)])
@pytest.mark.parametrize("shape,M", [((256, 512), 17), ((384, 1024), 32), ((128 * 9, 128 * 5), 33), ((256, 128), 64), ((4096, 4096), 24)])
def test_gemm_skinny_matches_oracle(lbl, cuda_device, qtype, xdt, tol, shape, M):
    """17..64 rows go to the tensor-core kernel with the operands swapped (weights = the 128-row operand, tokens = N,
    stream-K ranges + a reduction launch): the prefill kernel's contract, |diff| <= tol * sum|x w|; shapes with fewer
    items than SMs, ranges that cross several tiles (9 tiles x 5 pairs over 45 CTAs... ) and a full-size layer"""
    rng = np.random.default_rng(1250 + M)
    N, K = shape
    packed = oracle.quantize(rand_w(rng, N, K), qtype)
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(xdt)
    x = x_t.float().numpy()
    y = run_matmul(lbl, cuda_device, packed, shape, qtype, x_t, out_dtype=torch.float32)
    assert lbl.ggml.last_launch_count() == 2
    y = y.cpu().numpy()
    y64 = oracle.matmul_f64(packed, shape, qtype, x)
    wd = np.abs(oracle.dequantize(packed, shape, qtype)).astype(np.float64)
    bound = tol * (np.abs(x).astype(np.float64) @ wd.T) + 1e-30
    assert np.isfinite(y).all()
    assert (np.abs(y - y64) <= bound).all(), float((np.abs(y - y64) / bound).max())


@pytest.mark.timeout(300)
def test_gemm_skinny_random_shapes(lbl, cuda_device):
    """forty seeded draws of (format, N, K, rows, dtype): N a multiple of 128, K of 128, 17..64 rows -- fewer items than SMs, one
    pair of k-blocks per tile, ranges that span many tiles, bias / accumulate on some -- against float64 on the dequantised weights"""
    rng = np.random.default_rng(1280)
    qtypes = ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"]
    for trial in range(40):
        qtype = qtypes[trial % 6]
        N = 128 * int(rng.integers(1, 20))
        K = 128 * int(rng.integers(1, 24))
        M = int(rng.integers(17, 65))
        xdt = torch.float16 if trial % 2 == 0 else torch.bfloat16
        w = torch.from_numpy(rand_w(rng, N, K)).to(cuda_device)
        tiled = lbl.FP4Params(w, qtype=qid(qtype)).to(cuda_device).data
        wd = lbl.dequantize_device_layout(tiled, (N, K), qid(qtype)).double()
        x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).to(xdt).to(cuda_device)
        want = x_t.double() @ wd.T
        bound = (1e-3 if xdt == torch.float16 else 8e-3) * (x_t.double().abs() @ wd.abs().T) + 1e-30
        if trial % 3 == 0:
            bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(cuda_device)
            out = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).to(cuda_device)
            want = want + bias.double() + out.double()
            lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out=out, bias=bias, accumulate=True)
            y = out
        else:
            y = lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out_dtype=torch.float32)
        assert lbl.ggml.last_launch_count() == 2, (trial, qtype, N, K, M)
        err = (y.double() - want).abs()
        assert bool((err <= bound + 1e-6 * want.abs()).all()), (trial, qtype, N, K, M, float((err / bound).max()))


def test_gemm_skinny_epilogue_window_and_determinism(lbl, cuda_device):
    rng = np.random.default_rng(1260)
    N, K, M = 1024, 2048, 40
    packed = oracle.quantize(rand_w(rng, N, K), "sym_int4")
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half()
    base = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32)
    again = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32)
    assert torch.equal(base, again)                                  # partial sums are added in CTA order
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(cuda_device)
    yb = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32, bias=bias)
    torch.testing.assert_close(yb, base + bias, rtol=0, atol=1e-5)
    res = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).to(cuda_device)
    out = res.clone()
    run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out=out, accumulate=True)
    torch.testing.assert_close(out, res + base, rtol=0, atol=1e-5)
    for dt in (torch.float16, torch.bfloat16):
        yh = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=dt)
        torch.testing.assert_close(yh.float(), base.to(dt).float(), rtol=0, atol=0)
    # the window: 16 rows stay on the GEMV (one launch, exact), 17 take this kernel, fp32 rows never do, nor does a K that
    # is not a whole number of k-block pairs
    run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t[:16], out_dtype=torch.float32)
    assert lbl.ggml.last_launch_count() == 1
    small = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t[:17], out_dtype=torch.float32)
    assert lbl.ggml.last_launch_count() == 2
    torch.testing.assert_close(small, base[:17], rtol=1e-6, atol=1e-6)     # a row's result does not depend on the batch around it
    y32 = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t[:24].float(), out_dtype=torch.float32)
    torch.testing.assert_close(y32, base[:24], rtol=2e-3, atol=2e-3)
    odd = oracle.quantize(rand_w(rng, 256, 192), "sym_int4")
    run_matmul(lbl, cuda_device, odd, (256, 192), "sym_int4", x_t[:24, :192].contiguous(), out_dtype=torch.float32)
    assert lbl.ggml.last_launch_count() == 1


@pytest.mark.timeout(300)
@pytest.mark.parametrize("qtype,M", [("fp8_e4m3", 1024), ("sym_int4", 512), ("fp6", 300)])
def test_gemm_prefill_repeatable_with_odd_ring_depth(lbl, cuda_device, qtype, M):
    """8-bit weights at 256-token tiles leave room for three stages: a stage then alternates between the two dequant groups,
    and a warp that waited only for its own items could mistake an old phase of the stage's barrier for the new one
    (parity waits are safe one phase behind, no more).  Fifty launches over a long K must give the same bits every time."""
    rng = np.random.default_rng(1350)
    N, K = 1024, 11008
    tiled = lbl.FP4Params(torch.from_numpy(rand_w(rng, N, K)).to(cuda_device), qtype=qid(qtype)).to(cuda_device).data
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half().to(cuda_device)
    first = lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out_dtype=torch.float32)
    wd = lbl.dequantize_device_layout(tiled, (N, K), qid(qtype)).double()
    want = x_t.double() @ wd.T
    bound = 1e-3 * (x_t.double().abs() @ wd.abs().T) + 1e-30
    assert bool(((first.double() - want).abs() <= bound).all())
    for _ in range(50):
        again = lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out_dtype=torch.float32)
        assert torch.equal(first, again)


@pytest.mark.parametrize("qtype,N,M", [("sym_int4", 19200, 300), ("sym_int4", 9600, 512), ("asym_int4", 9472, 700),
                                        ("fp8_e4m3", 19456, 258), ("nf4", 9600, 385)])
def test_gemm_prefill_last_round_in_halves(lbl, cuda_device, qtype, N, M):
    """more 256-token tiles than SMs with a partial last round: those tiles run as two 128-token units each (a half that lies
    wholly beyond M included), with bias and accumulate in the epilogue -- the same bits as the products of whole tiles
    (the halves do the same MMAs on the same operands), checked against float64 on the dequantised weights.  Shapes (148 SMs,
    the host picks 256-token tiles for all of them): 300 tiles + 4 left over, one of whose halves starts at token 384 > M;
    150 + 2; 222 = 148 + 74 (exactly 148 half units); 304 + 8 with two valid tokens in the second tile; token 384 alone in a half"""
    rng = np.random.default_rng(N + M)
    K = 256
    tiled = lbl.FP4Params(torch.from_numpy(rand_w(rng, N, K)).to(cuda_device), qtype=qid(qtype)).to(cuda_device).data
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half().to(cuda_device)
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(cuda_device)
    res = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).to(cuda_device)
    out = res.clone()
    lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out=out, bias=bias, accumulate=True)
    wd = lbl.dequantize_device_layout(tiled, (N, K), qid(qtype)).double()
    want = x_t.double() @ wd.T + bias.double() + res.double()
    bound = 1e-3 * (x_t.double().abs() @ wd.abs().T) + 1e-5 * (bias.double().abs() + res.double().abs()) + 1e-30
    assert bool(((out.double() - want).abs() <= bound).all())
    plain = lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype), out_dtype=torch.float32)
    half_out = lbl.lowbit_matmul(x_t, tiled, (N, K), qid(qtype))
    assert half_out.dtype == torch.float16 and torch.equal(half_out, plain.half())


def test_gemm_prefill_epilogue_and_module(lbl, cuda_device):
    rng = np.random.default_rng(1300)
    N, K, M = 256, 512, 96
    packed = oracle.quantize(rand_w(rng, N, K), "sym_int4")
    x_t = torch.from_numpy(rng.standard_normal((M, K)).astype(np.float32)).half()
    base = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32)
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(cuda_device)
    yb = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float32, bias=bias)
    torch.testing.assert_close(yb, base + bias, rtol=0, atol=1e-5)
    res = torch.from_numpy(rng.standard_normal((M, N)).astype(np.float32)).to(cuda_device)
    out = res.clone()
    run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out=out, accumulate=True)
    torch.testing.assert_close(out, res + base, rtol=0, atol=1e-5)
    yh = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t, out_dtype=torch.float16)
    torch.testing.assert_close(yh.float(), base.half().float(), rtol=0, atol=0)
    # the GEMV (M <= 32, exact) and the GEMM (fp16-rounded weights) agree to the GEMM's tolerance
    small = run_matmul(lbl, cuda_device, packed, (N, K), "sym_int4", x_t[:32], out_dtype=torch.float32)
    torch.testing.assert_close(base[:32], small, rtol=2e-3, atol=2e-3)


# ---------------------------------------------------------------- autograd wrappers (LBL:532-602)
@pytest.mark.parametrize("qtype", ["sym_int4", "nf4", "fp8_e4m3"])
def test_matmul_lowbit_autograd(lbl, cuda_device, qtype):
    """forward = the device kernels, backward = grad @ dequant(W): both against the oracle-dequantised weight"""
    rng = np.random.default_rng(1500)
    N, K, M = 192, 256, 5
    packed = oracle.quantize(rand_w(rng, N, K), qtype)
    wd = torch.from_numpy(oracle.dequantize(packed, (N, K), qtype)).to(cuda_device)
    p = torch.from_numpy(packed)
    if qtype == "nf4":
        p = p.view(torch.bfloat16)
    weight = lbl.FP4Params(p, quantized=True, _shape=torch.Size((N, K)), qtype=qid(qtype)).to(cuda_device)
    x = torch.from_numpy(rng.standard_normal((2, M, K)).astype(np.float32)).to(cuda_device).requires_grad_(True)
    y = lbl.MatMulLowBit.apply(x, weight, N)
    assert y.shape == (2, M, N)
    torch.testing.assert_close(y, x.detach() @ wd.T, rtol=2e-5, atol=2e-5)
    g = torch.from_numpy(rng.standard_normal((2, M, N)).astype(np.float32)).to(cuda_device)
    y.backward(g)
    torch.testing.assert_close(x.grad, g @ wd, rtol=1e-5, atol=1e-5)
    # host-pointer variant on the canonical CPU weight
    wcpu = lbl.FP4Params(p, quantized=True, _shape=torch.Size((N, K)), qtype=qid(qtype))
    xc = x.detach().cpu().requires_grad_(True)
    yc = lbl.MatMulLowBitCPU.apply(xc, wcpu)
    torch.testing.assert_close(yc, xc.detach() @ wd.cpu().T, rtol=1e-5, atol=1e-5)
    yc.backward(g.cpu())
    torch.testing.assert_close(xc.grad, g.cpu() @ wd.cpu(), rtol=1e-5, atol=1e-5)


# ---------------------------------------------------------------- LowBitEmbedding (reference embedding.py:179-241)
@pytest.mark.parametrize("qtype", QTYPES)
def test_lowbit_embedding_rows_bit_exact(cuda_device, native_lib, qtype):
    from ipex_llm_b200.transformers.embedding import LowBitEmbedding
    rng = np.random.default_rng(1600)
    V, D = 203, 256                                   # ragged row count: the last 11 rows live in the canonical tail
    table = rand_w(rng, V, D, "wide")
    emb = torch.nn.Embedding(V, D)
    emb.weight.data = torch.from_numpy(table)
    low = LowBitEmbedding.from_embedding(emb, False, qid(qtype)).to(cuda_device)
    want = oracle.dequantize(oracle.quantize(table, qtype), (V, D), qtype)
    ids = torch.from_numpy(rng.integers(0, V, size=(3, 17))).to(cuda_device)
    ids[0, 0], ids[0, 1], ids[2, 16] = 0, V - 1, 192
    got = low(ids)
    assert got.shape == (3, 17, D) and got.dtype == torch.float32
    np.testing.assert_array_equal(got.cpu().numpy(), want[ids.cpu().numpy()])
    got32 = low(ids.int())                            # int32 indices
    assert torch.equal(got32, got)
    bad = torch.tensor([[V, -1, 5]], device=cuda_device)
    out = low(bad)
    assert float(out[0, 0].abs().max()) == 0.0 and float(out[0, 1].abs().max()) == 0.0
    np.testing.assert_array_equal(out[0, 2].cpu().numpy(), want[5])
    # fp16 model: the fp16 table is what gets quantised, and the output is rounded once to fp16
    emb16 = torch.nn.Embedding(V, D)
    emb16.weight.data = torch.from_numpy(table).half()
    t16 = emb16.weight.data.float().numpy()
    want16 = oracle.dequantize(oracle.quantize(t16, qtype), (V, D), qtype)
    low16 = LowBitEmbedding.from_embedding(emb16, False, qid(qtype)).to(cuda_device)
    got16 = low16(ids)
    assert got16.dtype == torch.float16
    np.testing.assert_array_equal(got16.float().cpu().numpy(), want16[ids.cpu().numpy()].astype(np.float16).astype(np.float32))
