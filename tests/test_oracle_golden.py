"""The CPU oracle against every pinned vector we have for this path (CPU only).

 * tests/golden/ref_python_q4.npz -- produced by RUNNING the reference's own Python statements of
   the formats (make_golden.py): convert_chatglm.quantize_q4_0/q4_1 (QK=64), the GGUF loader's
   Q4_0/Q4_1 dequantisation, convert_gptq's asym_int4 records.
 * tests/golden/gguf_q4_qk32.npz -- the installed gguf package (upstream ggml arithmetic, QK=32).
 * hand-computed single-block cases.
"""
import os

import numpy as np
import pytest

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def ref():
    return np.load(os.path.join(GOLD, "ref_python_q4.npz"))


def test_table_sizes():
    assert [oracle.qk_size(q) for q in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6")] == [64] * 6
    assert [oracle.type_size(q) for q in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6")] == \
        [34, 36, 34, 66, 66, 50]
    assert oracle.type_size("woq_int4") == 34 and oracle.qk_size(8) == 0


def test_sym_int4_matches_reference_python_qk64(ref):
    w = ref["w"]
    got = oracle.quantize(w, "sym_int4").reshape(-1, 34)
    want = ref["q4_0_bytes"].reshape(-1, 34)
    zero_blocks = ~w.reshape(-1, 64).any(axis=1)
    # the reference's torch statement divides by a zero scale on an all-zero block (NaN -> 0
    # codes); ggml's id = d ? 1/d : 0 yields code 8.  Both dequantise to exactly 0.
    assert zero_blocks.sum() == 1
    np.testing.assert_array_equal(got[~zero_blocks], want[~zero_blocks])
    np.testing.assert_array_equal(got[zero_blocks][:, :2], want[zero_blocks][:, :2])
    assert (got[zero_blocks][:, 2:] == 0x88).all()


def test_asym_int4_matches_reference_python_qk64(ref):
    np.testing.assert_array_equal(oracle.quantize(ref["w"], "asym_int4"), ref["q4_1_bytes"])


def test_dequant_matches_reference_gguf_loader(ref):
    # gguf.py:281-282 multiplies int8 codes by the fp16 scale IN fp16; upstream ggml (and the
    # oracle) do it in fp32: agreement to fp16 rounding of the product.
    w = ref["w"]
    for name, key in (("sym_int4", "q4_0"), ("asym_int4", "q4_1")):
        d = oracle.dequantize(ref[key + "_bytes"], w.shape, name)
        want = ref[key + "_dequant"]
        np.testing.assert_allclose(d, want, rtol=2e-3, atol=2 ** -11 * np.abs(want).max())


def test_gptq_records_decode_to_gptq_weights(ref):
    # convert.py:448-453: record = d, m = -(zero*scale), qs ; w = q*d + m
    W = ref["gptq_weight"]
    d = oracle.dequantize(ref["gptq_bytes"], W.shape, "asym_int4")
    np.testing.assert_allclose(d, W, rtol=0, atol=2e-3 * np.abs(W).max())
    # and the nibble order: low nibble = element j, high nibble = element j + QK/2
    rec = ref["gptq_bytes"].reshape(W.shape[0], -1, 36)
    ints = ref["gptq_ints"].T                                   # [out, in]
    qs = rec[:, :, 4:]
    lo, hi = qs & 0x0f, qs >> 4
    blocks = ints.reshape(W.shape[0], -1, 64)
    np.testing.assert_array_equal(lo, blocks[:, :, :32])
    np.testing.assert_array_equal(hi, blocks[:, :, 32:])


def test_qk32_mode_matches_gguf_package():
    g = np.load(os.path.join(GOLD, "gguf_q4_qk32.npz"))
    oracle.set_qk32(True)
    try:
        assert oracle.type_size("sym_int4") == 18 and oracle.type_size("asym_int4") == 20
        np.testing.assert_array_equal(oracle.quantize(g["w"], "sym_int4"), g["q4_0_bytes"])
        np.testing.assert_array_equal(oracle.quantize(g["w"], "asym_int4"), g["q4_1_bytes"])
    finally:
        oracle.set_qk32(False)


def test_qk32_live_against_gguf_package():
    gq = pytest.importorskip("gguf.quants")
    rng = np.random.default_rng(7)
    w = (rng.standard_normal((10000, 32)) * rng.lognormal(size=(10000, 1))).astype(np.float32)
    oracle.set_qk32(True)
    try:
        np.testing.assert_array_equal(oracle.quantize(w, "sym_int4"), gq.Q4_0.quantize_blocks(w).reshape(-1))
        np.testing.assert_array_equal(oracle.quantize(w, "asym_int4"), gq.Q4_1.quantize_blocks(w).reshape(-1))
    finally:
        oracle.set_qk32(False)


def _f16(x):
    return np.array([x], dtype=np.float16).view(np.uint8)


def test_hand_computed_blocks():
    # single negative maximum: d = -1/-8 = 0.125 > 0 ; single positive maximum: d < 0
    x = np.zeros(64, np.float32); x[3] = -1.0; x[40] = 0.5
    b = oracle.quantize(x, "sym_int4")
    np.testing.assert_array_equal(b[:2], _f16(0.125))
    codes = np.concatenate([b[2:] & 15, b[2:] >> 4])
    assert codes[3] == 0 and codes[40] == 12 and (np.delete(codes, [3, 40]) == 8).all()
    x = np.zeros(64, np.float32); x[0] = 2.0
    b = oracle.quantize(x, "sym_int4")
    np.testing.assert_array_equal(b[:2], _f16(-0.25))
    assert (b[2] & 15) == 0 and (b[2] >> 4) == 8          # 2.0 / -0.25 = -8 -> code 0
    # first of two equal |max| wins, keeping ITS sign
    x = np.zeros(64, np.float32); x[5] = 1.0; x[6] = -1.0
    np.testing.assert_array_equal(oracle.quantize(x, "sym_int4")[:2], _f16(-0.125))
    # asym: ramp 0..15 repeated -> d = 1, m = 0, codes = values
    x = np.tile(np.arange(16, dtype=np.float32), 4)
    b = oracle.quantize(x, "asym_int4")
    np.testing.assert_array_equal(b[:4], np.concatenate([_f16(1.0), _f16(0.0)]))
    codes = np.concatenate([b[4:] & 15, b[4:] >> 4])
    np.testing.assert_array_equal(codes, x.astype(np.uint8))
    # rounding is floor(v + 0.5) on the scaled value, clamp at 15
    x = np.zeros(64, np.float32); x[0] = 15.0; x[1] = 7.5; x[2] = 7.49
    codes = oracle.quantize(x, "asym_int4")[4:] & 15
    assert codes[0] == 15 and codes[1] == 8 and codes[2] == 7
    # nf4: absmax scale, +-1 ends of the table
    x = np.zeros(64, np.float32); x[0] = 3.0; x[1] = -3.0; x[2] = 1.5
    b = oracle.quantize(x, "nf4")
    np.testing.assert_array_equal(b[:2], _f16(3.0))
    assert (b[2] & 15) == 15 and (b[3] & 15) == 0 and (b[5] & 15) == 7
    # fp8 e4m3: d = absmax / 448 ; code of the maximum = 0x7e ; zero = 0
    x = np.zeros(64, np.float32); x[0] = 448.0; x[1] = -224.0; x[2] = 1.0
    b = oracle.quantize(x, "fp8_e4m3")
    np.testing.assert_array_equal(b[:2], _f16(1.0))
    assert b[2] == 0x7e and b[3] == (0x80 | 0x76) and b[4] == 0x38 and b[5] == 0
    # fp8 e5m2: 57344 -> 0x7b
    x = np.zeros(64, np.float32); x[0] = 57344.0; x[1] = 1.0
    b = oracle.quantize(x, "fp8_e5m2")
    assert b[2] == 0x7b and b[3] == 0x3c
    # fp6 e3m2: max 28 -> 0x1f ; 1.0 -> e=3,m=0 = 0b01100 ; 4 codes -> 3 bytes
    x = np.zeros(64, np.float32); x[0] = 28.0; x[1] = 1.0; x[2] = -1.0; x[3] = 0.0625
    b = oracle.quantize(x, "fp6")
    word = int(b[2]) | int(b[3]) << 8 | int(b[4]) << 16
    assert [(word >> (6 * i)) & 63 for i in range(4)] == [0x1f, 0x0c, 0x2c, 0x01]


def test_mini_float_codecs_roundtrip_and_rne():
    L = oracle.lib()
    for which, ncodes, skip in ((0, 256, (0x7f, 0xff)), (1, 256, tuple(range(0x7c, 0x80)) + tuple(range(0xfc, 0x100))), (2, 64, ())):
        vals = {}
        for c in range(ncodes):
            if c in skip:
                continue
            v = L.oracle_mini_decode(which, c)
            vals[c] = v
            if v == 0.0:
                continue
            assert L.oracle_mini_encode(which, v) == c, (which, c, v)
        pos = sorted(v for v in vals.values() if v > 0)
        for a, b in zip(pos[:-1], pos[1:]):
            mid = (a + b) / 2.0
            ca, cb = L.oracle_mini_encode(which, a), L.oracle_mini_encode(which, b)
            even = ca if ca % 2 == 0 else cb            # ties go to the even code
            assert L.oracle_mini_encode(which, mid) == even
            assert L.oracle_mini_encode(which, np.nextafter(np.float32(mid), np.float32(0))) == ca
            assert L.oracle_mini_encode(which, np.nextafter(np.float32(mid), np.float32(1e9))) == cb
    # e4m3 agrees with torch's float8_e4m3fn / e5m2 on random finite inputs
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(3)
    x = (rng.standard_normal(4000) * rng.choice([1e-3, 1, 50, 300], 4000)).astype(np.float32)
    for which, dt, mx in ((0, torch.float8_e4m3fn, 448.0), (1, torch.float8_e5m2, 57344.0)):
        xc = np.clip(x, -mx, mx)
        want = torch.from_numpy(xc).to(dt).view(torch.uint8).numpy()
        got = np.array([L.oracle_mini_encode(which, float(v)) for v in xc], dtype=np.uint8)
        np.testing.assert_array_equal(got, want)


def test_fp16_conversion_matches_numpy():
    L = oracle.lib()
    rng = np.random.default_rng(5)
    x = np.concatenate([(rng.standard_normal(5000) * rng.choice([1e-8, 1e-5, 1e-3, 1, 1e3, 6e4], 5000)),
                        [0.0, -0.0, 65504.0, 65519.9, 65520.0, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-8]]).astype(np.float32)
    want = x.astype(np.float16).view(np.uint16)
    got = np.array([L.oracle_f32_to_f16(float(v)) for v in x], dtype=np.uint16)
    np.testing.assert_array_equal(got, want)
    hs = rng.integers(0, 0x7c00, 3000).astype(np.uint16)
    np.testing.assert_array_equal(np.array([L.oracle_f16_to_f32(int(h)) for h in hs], dtype=np.float32),
                                  hs.view(np.float16).astype(np.float32))


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"])
def test_roundtrip_error_bounds(qtype):
    rng = np.random.default_rng(11)
    w = (rng.standard_normal((64, 256)) * 0.02).astype(np.float32)
    packed = oracle.quantize(w, qtype)
    assert packed.size == w.size // 64 * oracle.type_size(qtype)
    d = oracle.dequantize(packed, w.shape, qtype)
    blocks, db = w.reshape(-1, 64), d.reshape(-1, 64)
    amax = np.abs(blocks).max(axis=1, keepdims=True)
    bound = {"sym_int4": 1 / 8, "asym_int4": 2 / 15 * 0.5, "nf4": 0.16, "fp8_e4m3": 1 / 16,
             "fp8_e5m2": 1 / 8, "fp6": 1 / 8}[qtype]
    assert (np.abs(db - blocks) <= amax * bound * 1.02 + 1e-9).all()
    # idempotence: quantising the dequantised weights reproduces the codes (scales may move by
    # an fp16 ulp for the absmax-derived formats, so compare dequantised values)
    d2 = oracle.dequantize(oracle.quantize(d, qtype), w.shape, qtype)
    np.testing.assert_allclose(d2, d, rtol=2e-3, atol=1e-6)


@pytest.mark.parametrize("qtype", ["fp8_e4m3", "fp6"])
def test_scale_search_never_worse(qtype):
    rng = np.random.default_rng(13)
    w = (rng.standard_normal((32, 256)) * 0.02).astype(np.float32)
    e0 = ((oracle.dequantize(oracle.quantize(w, qtype), w.shape, qtype) - w).reshape(-1, 64) ** 2).sum(1)
    e1 = ((oracle.dequantize(oracle.quantize(w, qtype, scale_search=True), w.shape, qtype) - w).reshape(-1, 64) ** 2).sum(1)
    assert (e1 <= e0 * (1 + 1e-5)).all() and (e1 < e0).any()


def test_histogram_counts_codes():
    rng = np.random.default_rng(17)
    w = rng.standard_normal((8, 128)).astype(np.float32)
    packed, hist = oracle.quantize(w, "sym_int4", return_hist=True)
    qs = packed.reshape(-1, 34)[:, 2:]
    want = np.bincount(np.concatenate([(qs & 15).ravel(), (qs >> 4).ravel()]), minlength=16)
    np.testing.assert_array_equal(hist, want)


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"])
def test_matmul_modes_consistent(qtype):
    rng = np.random.default_rng(19)
    N, K, M = 48, 256, 3
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float32)
    x = rng.standard_normal((M, K)).astype(np.float32)
    packed = oracle.quantize(w, qtype)
    yb = oracle.matmul(packed, (N, K), qtype, x)
    want = x.astype(np.float64) @ oracle.dequantize(packed, (N, K), qtype).astype(np.float64).T
    np.testing.assert_allclose(yb, want, rtol=0, atol=1e-5)
    np.testing.assert_allclose(oracle.matmul_f64(packed, (N, K), qtype, x), want, rtol=0, atol=1e-12)
    ya = oracle.matmul_mode_a(packed, (N, K), qtype, x)          # int8 activations: ~1e-2 noise
    assert np.abs(ya - yb).max() <= 0.03 * np.abs(yb).max() + 1e-4


def test_empty_inputs():
    assert oracle.quantize(np.zeros((0, 64), np.float32), "sym_int4").size == 0
    with pytest.raises(ValueError):
        oracle.quantize(np.zeros((2, 48), np.float32), "sym_int4")
