/*
 * lowbit_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the arithmetic behind ipex-llm's low-bit linear hot path
 * (the closed `libllama-api.so` symbols bound in
 *  python/llm/src/ipex_llm/ggml/model/llama/llama_cpp.py:945-1185).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this file's shared object.  The product
 * (ipex_llm_b200/csrc) never includes, links or calls anything from here.
 *
 * PARITY STATUS (see DESIGN.md "Oracle"):
 *   - the reference's own binary for this path is NOT in /root/reference and the
 *     reference has no golden vectors for it, so byte-level parity with the
 *     closed binary is UNPINNED;
 *   - sym_int4 / asym_int4 arithmetic follows public upstream ggml
 *     quantize_row_q4_0_ref / quantize_row_q4_1_ref and is pinned against
 *     (a) the installed `gguf.quants` numpy restatement (QK=32 mode, byte
 *         equality), and
 *     (b) golden vectors produced by running the reference's OWN in-repo Python
 *         statements of these formats here:
 *         utils/convert_chatglm.py:101-133 (quantize_q4_0 / quantize_q4_1, QK=64),
 *         transformers/gguf/gguf.py:273-296 (Q4_0 / Q4_1 dequantisation) and
 *         transformers/convert.py:382-456 (convert_gptq -> asym_int4 records);
 *   - nf4 / fp8_e4m3 / fp8_e5m2 / fp6 have no statement of layout or arithmetic
 *     anywhere in the reference: they are BUILD-DEFINED here (FORMAT section of
 *     DESIGN.md) and the oracle is the specification the CUDA kernels must meet.
 *
 * Block record (canonical / CPU / on-disk layout), row-major [rows][k/QK]:
 *   d (fp16) [, m (fp16)] , codes            -- convert.py:448-453 (d, m, qs order)
 *   4-bit codes: byte j = code[j] | code[j+QK/2] << 4   -- convert.py:421-426,
 *                                                          gguf.py:281,294
 *
 * Compile: see oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp).
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

/* qtype enum: python/llm/src/ipex_llm/ggml/quantize.py:28-60 */
enum {
    QT_SYM_INT4 = 2, QT_ASYM_INT4 = 3, QT_NF4 = 10, QT_FP8_E4M3 = 15,
    QT_FP8_E5M2 = 19, QT_FP6 = 29, QT_WOQ_INT4 = 34
};

/* Block geometry.  Default QK=64 for every in-scope qtype (in-repo evidence:
 * utils/convert_chatglm.py:60-62 and transformers/convert.py:615-616).  g_qk32
 * switches sym_int4/asym_int4 to upstream ggml's QK=32 so the arithmetic can be
 * checked byte-for-byte against gguf.quants. */
static int g_qk32 = 0;
void oracle_set_qk32(int on) { g_qk32 = on ? 1 : 0; }

static int canon(int qtype) { return qtype == QT_WOQ_INT4 ? QT_SYM_INT4 : qtype; }

/* low_bit_linear.py:98-103 get_block_size / get_qk_size -> ggml_qk_size */
int oracle_qk_size(int qtype) {
    switch (canon(qtype)) {
        case QT_SYM_INT4: case QT_ASYM_INT4: return g_qk32 ? 32 : 64;
        case QT_NF4: case QT_FP8_E4M3: case QT_FP8_E5M2: case QT_FP6: return 64;
        default: return 0;
    }
}

/* low_bit_linear.py:117 ggml_type_size: bytes per block record */
size_t oracle_type_size(int qtype) {
    int qk = oracle_qk_size(qtype);
    switch (canon(qtype)) {
        case QT_SYM_INT4:  return 2 + qk / 2;
        case QT_ASYM_INT4: return 4 + qk / 2;
        case QT_NF4:       return 2 + 32;
        case QT_FP8_E4M3: case QT_FP8_E5M2: return 2 + 64;
        case QT_FP6:       return 2 + 48;
        default: return 0;
    }
}

/* ---- fp16 <-> fp32 (round to nearest even), written out so that the oracle
 * does not depend on compiler _Float16 support ---- */
static uint16_t f32_to_f16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t absx = x & 0x7fffffffu;
    if (absx >= 0x7f800000u)                     /* inf / nan */
        return (uint16_t)(sign | 0x7c00u | ((absx > 0x7f800000u) ? 0x200u : 0));
    if (absx >= 0x477ff000u) {                   /* >= 65520 rounds to inf */
        return (uint16_t)(sign | 0x7c00u);
    }
    if (absx < 0x38800000u) {                    /* result is subnormal or zero */
        if (absx < 0x33000000u) return (uint16_t)sign;   /* < 2^-25 -> 0 */
        int e = (int)(absx >> 23);               /* biased fp32 exponent */
        uint32_t mant = (absx & 0x7fffffu) | 0x800000u;
        int shift = 126 - e;                     /* 14..24 */
        uint32_t q = mant >> shift;
        uint32_t rem = mant & ((1u << shift) - 1u);
        uint32_t half = 1u << (shift - 1);
        if (rem > half || (rem == half && (q & 1u))) q++;
        return (uint16_t)(sign | q);
    }
    uint32_t e = (absx >> 23) - 112u;            /* rebias 127 -> 15 */
    uint32_t mant = absx & 0x7fffffu;
    uint32_t q = (e << 10) | (mant >> 13);
    uint32_t rem = mant & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (q & 1u))) q++;
    return (uint16_t)(sign | q);
}

static float f16_to_f32(uint16_t h) {
    uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
    uint32_t e = (h >> 10) & 0x1fu, m = h & 0x3ffu, x;
    if (e == 0) {
        if (m == 0) x = sign;
        else {
            int s = 0;
            while (!(m & 0x400u)) { m <<= 1; s++; }
            x = sign | ((uint32_t)(113 - s) << 23) | ((m & 0x3ffu) << 13);
        }
    } else if (e == 31) x = sign | 0x7f800000u | (m << 13);
    else x = sign | ((e + 112u) << 23) | (m << 13);
    float f; memcpy(&f, &x, 4); return f;
}

/* exported for tests */
uint16_t oracle_f32_to_f16(float f) { return f32_to_f16(f); }
float oracle_f16_to_f32(uint16_t h) { return f16_to_f32(h); }

/* ---- small float formats: generic sign / E exponent bits / M mantissa bits,
 * round-to-nearest-even, saturate to the largest finite code (build-defined,
 * matches OCP e4m3fn / e5m2 / e3m2 finite encodings) ---- */
typedef struct { int ebits, mbits, bias; uint32_t max_code; } minifmt;
static const minifmt FMT_E4M3 = {4, 3, 7, 0x7e};   /* max 448, 0x7f = NaN (never produced) */
static const minifmt FMT_E5M2 = {5, 2, 15, 0x7b};  /* max 57344, 0x7c = inf (never produced) */
static const minifmt FMT_E3M2 = {3, 2, 3, 0x1f};   /* max 28, no inf / nan */

static float mini_decode(const minifmt* f, uint32_t code) {
    int sbit = f->ebits + f->mbits;
    uint32_t sign = (code >> sbit) & 1u;
    uint32_t e = (code >> f->mbits) & ((1u << f->ebits) - 1u);
    uint32_t m = code & ((1u << f->mbits) - 1u);
    float v;
    if (e == 0) v = ldexpf((float)m, 1 - f->bias - f->mbits);
    else        v = ldexpf((float)(m | (1u << f->mbits)), (int)e - f->bias - f->mbits);
    return sign ? -v : v;
}

static uint32_t mini_encode(const minifmt* f, float x) {
    int sbit = f->ebits + f->mbits;
    uint32_t sign = signbit(x) ? (1u << sbit) : 0u;
    float a = fabsf(x);
    if (!(a == a)) return sign | f->max_code;           /* NaN -> saturate (inputs are finite) */
    float maxv = mini_decode(f, f->max_code);
    if (a >= maxv) return sign | f->max_code;
    if (a == 0.0f) return sign;
    int e; (void)frexpf(a, &e);                         /* a = frac * 2^e, frac in [0.5,1) */
    int ue = e - 1;                                     /* a in [2^ue, 2^(ue+1)) */
    int emin = 1 - f->bias;
    if (ue < emin) ue = emin;                           /* subnormal range shares the min exponent */
    float ulp = ldexpf(1.0f, ue - f->mbits);
    float q = a / ulp;                                  /* exact: power-of-two scaling */
    float r = nearbyintf(q);                            /* RNE in the default rounding mode */
    uint32_t code;
    uint32_t qi = (uint32_t)r;
    if (ue == emin && qi < (1u << f->mbits)) code = qi; /* subnormal */
    else {
        /* qi in [2^mbits, 2^(mbits+1)] ; qi == 2^(mbits+1) carries into the next exponent */
        uint32_t eb = (uint32_t)(ue + f->bias);
        if (qi == (2u << f->mbits)) { eb += 1; qi = (1u << f->mbits); }
        code = (eb << f->mbits) | (qi & ((1u << f->mbits) - 1u));
    }
    if (code > f->max_code) code = f->max_code;
    return sign | code;
}

uint32_t oracle_mini_encode(int which, float x) {
    return mini_encode(which == 0 ? &FMT_E4M3 : which == 1 ? &FMT_E5M2 : &FMT_E3M2, x);
}
float oracle_mini_decode(int which, uint32_t c) {
    return mini_decode(which == 0 ? &FMT_E4M3 : which == 1 ? &FMT_E5M2 : &FMT_E3M2, c);
}

/* QLoRA NormalFloat4 table (build-defined choice for nf4; the reference holds no table) */
static const float NF4_TABLE[16] = {
    -1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f,
    -0.28444138169288635f, -0.18477343022823334f, -0.09105003625154495f, 0.0f,
    0.07958029955625534f, 0.16093020141124725f, 0.24611230194568634f, 0.33791524171829224f,
    0.44070982933044434f, 0.5626170039176941f, 0.7229568362236023f, 1.0f };

static uint32_t nf4_encode(float v) {
    uint32_t c = 0;
    for (int i = 0; i < 15; i++) {
        float mid = (NF4_TABLE[i] + NF4_TABLE[i + 1]) * 0.5f;
        if (v > mid) c++;            /* tie -> lower index */
    }
    return c;
}

/* ---------------------------------------------------------------- quantize */

/* upstream ggml quantize_row_q4_0_ref restated; generic in qk (32 or 64).
 * scale = signed absmax / -8 (first max wins), id = 1/d,
 * code = min(15, (int8)(x*id + 8.5f))   -- cf. utils/convert_chatglm.py:101-115 */
static void quant_block_sym_int4(const float* x, uint8_t* out, int qk, int64_t* hist) {
    float amax = 0.0f, max = 0.0f;
    for (int j = 0; j < qk; j++) {
        float v = x[j];
        if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
    }
    const float d = max / -8.0f;
    const float id = d != 0.0f ? 1.0f / d : 0.0f;
    uint16_t dh = f32_to_f16(d);
    memcpy(out, &dh, 2);
    for (int j = 0; j < qk / 2; j++) {
        const float x0 = x[j] * id;
        const float x1 = x[qk / 2 + j] * id;
        int q0 = (int)(int8_t)(x0 + 8.5f); if (q0 > 15) q0 = 15;
        int q1 = (int)(int8_t)(x1 + 8.5f); if (q1 > 15) q1 = 15;
        out[2 + j] = (uint8_t)(q0 | (q1 << 4));
        if (hist) { hist[q0]++; hist[q1]++; }
    }
}

/* upstream ggml quantize_row_q4_1_ref restated: d = (max-min)/15, m = min,
 * code = min(15, (int8)((x-min)*id + 0.5f))  -- cf. utils/convert_chatglm.py:118-133 */
static void quant_block_asym_int4(const float* x, uint8_t* out, int qk, int64_t* hist) {
    float mn = INFINITY, mx = -INFINITY;
    for (int j = 0; j < qk; j++) {
        float v = x[j];
        if (v < mn) mn = v;
        if (v > mx) mx = v;
    }
    const float d = (mx - mn) / 15.0f;
    const float id = d != 0.0f ? 1.0f / d : 0.0f;
    uint16_t dh = f32_to_f16(d), mh = f32_to_f16(mn);
    memcpy(out, &dh, 2); memcpy(out + 2, &mh, 2);
    for (int j = 0; j < qk / 2; j++) {
        const float x0 = (x[j] - mn) * id;
        const float x1 = (x[qk / 2 + j] - mn) * id;
        int q0 = (int)(int8_t)(x0 + 0.5f); if (q0 > 15) q0 = 15;
        int q1 = (int)(int8_t)(x1 + 0.5f); if (q1 > 15) q1 = 15;
        out[4 + j] = (uint8_t)(q0 | (q1 << 4));
        if (hist) { hist[q0]++; hist[q1]++; }
    }
}

static float block_absmax(const float* x, int qk) {
    float amax = 0.0f;
    for (int j = 0; j < qk; j++) { float a = fabsf(x[j]); if (a > amax) amax = a; }
    return amax;
}

/* nf4 (build-defined): d = absmax (fp16), code = nearest NF4 entry of x * (1/d) */
static void quant_block_nf4(const float* x, uint8_t* out, int64_t* hist) {
    const float d = block_absmax(x, 64);
    const float id = d != 0.0f ? 1.0f / d : 0.0f;
    uint16_t dh = f32_to_f16(d); memcpy(out, &dh, 2);
    for (int j = 0; j < 32; j++) {
        uint32_t q0 = nf4_encode(x[j] * id), q1 = nf4_encode(x[32 + j] * id);
        out[2 + j] = (uint8_t)(q0 | (q1 << 4));
        if (hist) { hist[q0]++; hist[q1]++; }
    }
}

/* scale-search candidates (build-defined; use_scale_search convert.py:459-469 only
 * says WHEN the flag is set): multipliers applied to absmax/MAX, tried in this
 * order, first strictly-smaller squared error wins. */
static const float SEARCH_FACTORS[9] = {1.0f, 0.95f, 1.05f, 0.9f, 1.1f, 0.85f, 1.15f, 0.8f, 1.2f};

/* fp8 / fp6 (build-defined): d = absmax / MAXV (fp16), code = RNE(x * (1/d)) saturating */
static void quant_block_mini(const float* x, uint8_t* out, const minifmt* f, int scale_search,
                             uint8_t* codes /* 64 scratch */) {
    const float amax = block_absmax(x, 64);
    const float maxv = mini_decode(f, f->max_code);
    const float d0 = amax / maxv;
    float best_d = d0;
    if (scale_search && d0 != 0.0f) {
        float best_err = INFINITY;
        for (int c = 0; c < 9; c++) {
            const float dc = d0 * SEARCH_FACTORS[c];
            const float idc = 1.0f / dc;
            const float d16 = f16_to_f32(f32_to_f16(dc));
            float err = 0.0f;
            for (int j = 0; j < 64; j++) {
                float r = mini_decode(f, mini_encode(f, x[j] * idc)) * d16;
                float e = r - x[j];
                err = err + e * e;          /* -ffp-contract=off: mul then add, sequential */
            }
            if (err < best_err) { best_err = err; best_d = dc; }
        }
    }
    const float d = best_d;
    const float id = d != 0.0f ? 1.0f / d : 0.0f;
    uint16_t dh = f32_to_f16(d); memcpy(out, &dh, 2);
    for (int j = 0; j < 64; j++) codes[j] = (uint8_t)mini_encode(f, x[j] * id);
}

static void quant_block_fp8(const float* x, uint8_t* out, const minifmt* f, int scale_search) {
    uint8_t codes[64];
    quant_block_mini(x, out, f, scale_search, codes);
    memcpy(out + 2, codes, 64);               /* element j -> byte 2 + j */
}

/* fp6 e3m2: 4 consecutive codes -> 3 bytes little-endian: c0 | c1<<6 | c2<<12 | c3<<18 */
static void quant_block_fp6(const float* x, uint8_t* out, int scale_search) {
    uint8_t codes[64];
    quant_block_mini(x, out, &FMT_E3M2, scale_search, codes);
    for (int g = 0; g < 16; g++) {
        uint32_t w = (uint32_t)codes[4 * g] | ((uint32_t)codes[4 * g + 1] << 6) |
                     ((uint32_t)codes[4 * g + 2] << 12) | ((uint32_t)codes[4 * g + 3] << 18);
        out[2 + 3 * g] = (uint8_t)(w & 0xff);
        out[2 + 3 * g + 1] = (uint8_t)((w >> 8) & 0xff);
        out[2 + 3 * g + 2] = (uint8_t)((w >> 16) & 0xff);
    }
}

/* ggml_quantize_tensor(src, dst, qtype, n, k, hist, scale_search)
 * llama_cpp.py:946-967; called from low_bit_linear.py:180 with n = numel,
 * k = last dim; blocks are laid out row-major so k only has to divide n and be a
 * multiple of QK (low_bit_linear.py:125).  Returns bytes written. */
size_t oracle_quantize_tensor(const float* src, void* dst, int qtype, size_t n, int k,
                              int64_t* hist, int scale_search) {
    const int qt = canon(qtype);
    const int qk = oracle_qk_size(qt);
    const size_t bs = oracle_type_size(qt);
    if (qk == 0 || k <= 0 || (k % qk) != 0 || (n % (size_t)qk) != 0) return 0;
    const size_t nblk = n / (size_t)qk;
    uint8_t* out = (uint8_t*)dst;
    if (hist) memset(hist, 0, 16 * sizeof(int64_t));
    for (size_t b = 0; b < nblk; b++) {
        const float* x = src + b * (size_t)qk;
        uint8_t* o = out + b * bs;
        switch (qt) {
            case QT_SYM_INT4:  quant_block_sym_int4(x, o, qk, hist); break;
            case QT_ASYM_INT4: quant_block_asym_int4(x, o, qk, hist); break;
            case QT_NF4:       quant_block_nf4(x, o, hist); break;
            case QT_FP8_E4M3:  quant_block_fp8(x, o, &FMT_E4M3, scale_search); break;
            case QT_FP8_E5M2:  quant_block_fp8(x, o, &FMT_E5M2, scale_search); break;
            case QT_FP6:       quant_block_fp6(x, o, scale_search); break;
            default: return 0;
        }
    }
    return nblk * bs;
}

/* -------------------------------------------------------------- dequantize */

/* one block -> qk floats.  sym: (code-8)*d ; asym: code*d + m
 * (transformers/gguf/gguf.py:273-296, fp32 arithmetic as in upstream ggml) */
static void dequant_block(const uint8_t* in, float* y, int qt, int qk) {
    uint16_t dh; memcpy(&dh, in, 2);
    const float d = f16_to_f32(dh);
    switch (qt) {
        case QT_SYM_INT4:
            for (int j = 0; j < qk / 2; j++) {
                y[j] = (float)((int)(in[2 + j] & 0x0f) - 8) * d;
                y[j + qk / 2] = (float)((int)(in[2 + j] >> 4) - 8) * d;
            }
            break;
        case QT_ASYM_INT4: {
            uint16_t mh; memcpy(&mh, in + 2, 2);
            const float m = f16_to_f32(mh);
            for (int j = 0; j < qk / 2; j++) {
                y[j] = (float)(in[4 + j] & 0x0f) * d + m;
                y[j + qk / 2] = (float)(in[4 + j] >> 4) * d + m;
            }
            break;
        }
        case QT_NF4:
            for (int j = 0; j < 32; j++) {
                y[j] = NF4_TABLE[in[2 + j] & 0x0f] * d;
                y[j + 32] = NF4_TABLE[in[2 + j] >> 4] * d;
            }
            break;
        case QT_FP8_E4M3:
            for (int j = 0; j < 64; j++) y[j] = mini_decode(&FMT_E4M3, in[2 + j]) * d;
            break;
        case QT_FP8_E5M2:
            for (int j = 0; j < 64; j++) y[j] = mini_decode(&FMT_E5M2, in[2 + j]) * d;
            break;
        case QT_FP6:
            for (int g = 0; g < 16; g++) {
                uint32_t w = (uint32_t)in[2 + 3 * g] | ((uint32_t)in[2 + 3 * g + 1] << 8) |
                             ((uint32_t)in[2 + 3 * g + 2] << 16);
                for (int i = 0; i < 4; i++)
                    y[4 * g + i] = mini_decode(&FMT_E3M2, (w >> (6 * i)) & 0x3f) * d;
            }
            break;
        default: break;
    }
}

/* ggml_dequantize(src, dst, k, qtype)  llama_cpp.py:1084-1099 ; k = element count
 * (low_bit_linear.py:269-279).  ggml_dequantize_q4_0 (:1068-1080) == qtype 2. */
void oracle_dequantize(const void* src, float* dst, size_t k, int qtype) {
    const int qt = canon(qtype);
    const int qk = oracle_qk_size(qt);
    const size_t bs = oracle_type_size(qt);
    if (qk == 0) return;
    const uint8_t* in = (const uint8_t*)src;
    const size_t nblk = k / (size_t)qk;
#pragma omp parallel for schedule(static)
    for (size_t b = 0; b < nblk; b++) dequant_block(in + b * bs, dst + b * (size_t)qk, qt, qk);
}

/* ---------------------------------------------------------------- matmul */

/* MODE B (primary parity target): y[m][n] = sum_k x[m][k] * dequant(W)[n][k], fp32, fixed
 * k order -- the semantics the reference itself states at
 * low_bit_linear.py:581-583 (MatMulLowBitCPU.forward) and :741-742.
 * Argument convention of ggml_compute_forward_mul_mat_q_fp32 (llama_cpp.py:1154-1185,
 * low_bit_linear.py:509-516): ne0 = (K, N), ne1 = (K, M); y is [M][N] row-major fp32. */
void oracle_mul_mat_q_fp32(const int64_t* ne0, const void* w, int qtype, const int64_t* ne1,
                           const float* x, float* y) {
    const int qt = canon(qtype);
    const int qk = oracle_qk_size(qt);
    const size_t bs = oracle_type_size(qt);
    const int64_t K = ne0[0], N = ne0[1], M = ne1[1];
    if (qk == 0 || ne1[0] != K || K % qk) return;
    const int64_t nb = K / qk;
    const uint8_t* wb = (const uint8_t*)w;
#pragma omp parallel
    {
        float* row = (float*)malloc((size_t)K * sizeof(float));
#pragma omp for schedule(static)
        for (int64_t n = 0; n < N; n++) {
            for (int64_t b = 0; b < nb; b++)
                dequant_block(wb + ((size_t)n * nb + b) * bs, row + b * qk, qt, qk);
            for (int64_t m = 0; m < M; m++) {
                const float* xr = x + (size_t)m * K;
                float acc = 0.0f;
                for (int64_t k = 0; k < K; k++) acc = acc + row[k] * xr[k];
                y[(size_t)m * N + n] = acc;
            }
        }
        free(row);
    }
}

/* double-precision variant of mode B: used by tests to bound the error of any
 * fp32 summation order (GPU split-K trees differ from the serial order above). */
void oracle_mul_mat_q_f64(const int64_t* ne0, const void* w, int qtype, const int64_t* ne1,
                          const float* x, double* y) {
    const int qt = canon(qtype);
    const int qk = oracle_qk_size(qt);
    const size_t bs = oracle_type_size(qt);
    const int64_t K = ne0[0], N = ne0[1], M = ne1[1];
    if (qk == 0 || ne1[0] != K || K % qk) return;
    const int64_t nb = K / qk;
    const uint8_t* wb = (const uint8_t*)w;
#pragma omp parallel
    {
        float* row = (float*)malloc((size_t)K * sizeof(float));
#pragma omp for schedule(static)
        for (int64_t n = 0; n < N; n++) {
            for (int64_t b = 0; b < nb; b++)
                dequant_block(wb + ((size_t)n * nb + b) * bs, row + b * qk, qt, qk);
            for (int64_t m = 0; m < M; m++) {
                const float* xr = x + (size_t)m * K;
                double acc = 0.0;
                for (int64_t k = 0; k < K; k++) acc += (double)row[k] * (double)xr[k];
                y[(size_t)m * N + n] = acc;
            }
        }
        free(row);
    }
}

/* MODE A (ggml-faithful CPU arithmetic, the timed CPU baseline): activations are
 * quantised per QK block to int8 (Q8_0 style: dx = absmax/127, q = roundf(x/dx)),
 * integer dot with the 4-bit codes, then * d_w * d_x, fp32 accumulate over blocks.
 * This is the presumed behaviour of the closed CPU kernel (ggml lineage); it cannot
 * be verified (binary absent) and is reported, not used as the parity target.
 * sym_int4 / asym_int4 only; other qtypes fall back to mode B. */
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx512f,avx512bw,avx512vnni")))
static int32_t dot_u4_s8_64_vnni(const uint8_t* qs, const int8_t* xq) {
    /* 64 codes: low nibbles = elements 0..31, high nibbles = elements 32..63 */
    __m256i raw = _mm256_loadu_si256((const __m256i*)qs);
    __m256i lo = _mm256_and_si256(raw, _mm256_set1_epi8(0x0f));
    __m256i hi = _mm256_and_si256(_mm256_srli_epi16(raw, 4), _mm256_set1_epi8(0x0f));
    __m512i u = _mm512_inserti64x4(_mm512_castsi256_si512(lo), hi, 1);
    __m512i s = _mm512_loadu_si512((const void*)xq);
    __m512i acc = _mm512_dpbusd_epi32(_mm512_setzero_si512(), u, s);
    return _mm512_reduce_add_epi32(acc);
}
static int have_vnni(void) {
    static int v = -1;
    if (v < 0) v = __builtin_cpu_supports("avx512vnni") && __builtin_cpu_supports("avx512bw");
    return v;
}
#else
static int have_vnni(void) { return 0; }
#endif

static int32_t dot_u4_s8_scalar(const uint8_t* qs, const int8_t* xq, int qk) {
    int32_t s = 0;
    for (int j = 0; j < qk / 2; j++) {
        s += (int32_t)(qs[j] & 0x0f) * xq[j];
        s += (int32_t)(qs[j] >> 4) * xq[j + qk / 2];
    }
    return s;
}

int oracle_modeA_uses_vnni(void) { return have_vnni(); }

void oracle_mul_mat_q_fp32_modeA(const int64_t* ne0, const void* w, int qtype, const int64_t* ne1,
                                 const float* x, float* y) {
    const int qt = canon(qtype);
    if (qt != QT_SYM_INT4 && qt != QT_ASYM_INT4) { oracle_mul_mat_q_fp32(ne0, w, qtype, ne1, x, y); return; }
    const int qk = oracle_qk_size(qt);
    const size_t bs = oracle_type_size(qt);
    const int64_t K = ne0[0], N = ne0[1], M = ne1[1];
    if (ne1[0] != K || K % qk) return;
    const int64_t nb = K / qk;
    const uint8_t* wb = (const uint8_t*)w;
    int8_t* xq = (int8_t*)malloc((size_t)M * K);
    float* xd = (float*)malloc((size_t)M * nb * sizeof(float));   /* per-block activation scale */
    float* xs = (float*)malloc((size_t)M * nb * sizeof(float));   /* dx * sum(q) (Q8_1 style) */
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < M * nb; i++) {
        const float* xb = x + (size_t)i * qk;
        float amax = 0.0f;
        for (int j = 0; j < qk; j++) { float a = fabsf(xb[j]); if (a > amax) amax = a; }
        const float dx = amax / 127.0f;
        const float idx = dx != 0.0f ? 1.0f / dx : 0.0f;
        int32_t sum = 0;
        for (int j = 0; j < qk; j++) {
            int q = (int)roundf(xb[j] * idx);
            xq[(size_t)i * qk + j] = (int8_t)q; sum += q;
        }
        xd[i] = dx; xs[i] = dx * (float)sum;
    }
    const int vnni = have_vnni() && qk == 64;
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; n++) {
        for (int64_t m = 0; m < M; m++) {
            float acc = 0.0f;
            for (int64_t b = 0; b < nb; b++) {
                const uint8_t* blk = wb + ((size_t)n * nb + b) * bs;
                uint16_t dh; memcpy(&dh, blk, 2);
                const float d = f16_to_f32(dh);
                const uint8_t* qs = blk + (qt == QT_ASYM_INT4 ? 4 : 2);
                const int8_t* xb = xq + ((size_t)m * nb + b) * qk;
                int32_t dot;
#if defined(__x86_64__)
                dot = vnni ? dot_u4_s8_64_vnni(qs, xb) : dot_u4_s8_scalar(qs, xb, qk);
#else
                dot = dot_u4_s8_scalar(qs, xb, qk);
#endif
                const float dx = xd[(size_t)m * nb + b], sx = xs[(size_t)m * nb + b];
                if (qt == QT_SYM_INT4) {
                    /* sum (c-8) q = dot - 8*sum(q) ; dx*sum(q) = sx */
                    acc += d * (dx * (float)dot - 8.0f * sx);
                } else {
                    uint16_t mh; memcpy(&mh, blk + 2, 2);
                    acc += d * dx * (float)dot + f16_to_f32(mh) * sx;
                }
            }
            y[(size_t)m * N + n] = acc;
        }
    }
    free(xq); free(xd); free(xs);
}
