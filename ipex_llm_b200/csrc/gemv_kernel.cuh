// gemv_kernel.cuh -- kernel (b): HBM-bound low-bit dequant-GEMV for decode (M = 1..32 rows of x).
//
// Replaces the matmul behind LowBitLinear.forward for small batches: the CPU symbol
// ggml_compute_forward_mul_mat_q_fp32 (reference low_bit_linear.py:494-529) and the XPU ops
// xe_linear.forward_new / xe_batch.batch_forward (low_bit_linear.py:696-708).
//
// Design (B200-first):
//  * Weights live in the tiled device layout (lowbit_common.cuh): a 16-row x 64-element tile is
//    512 contiguous bytes (4-bit) in which lane l's 16 bytes are exactly the codes of its
//    mma.m16n8k16 A-fragment rows.  Tiles are numbered row-tile-major, so ANY run of tiles is one
//    contiguous byte range (codes) plus one contiguous range of fp16 scales.
//  * Staging: every warp streams its own span through a private shared-memory ring with TMA bulk
//    copies (cp.async.bulk global->shared, mbarrier complete_tx).  Lane 0 issues 2-4 KB copies
//    several stages ahead; memory-level parallelism no longer depends on registers or on the
//    6-entry load scoreboard (the first version kept a register ring of LDG.128 and stalled on
//    long_scoreboard at ~1.2 TB/s).  Consumers read the stage with conflict-free LDS.128.
//  * In-register dequant is the instruction budget of this kernel (B200 moves ~23 bytes per SM
//    per cycle: at most ~95 issue slots per 544-byte tile).  fp16 activations: a masked nibble
//    pair (w & 0x000f000f) IS a pair of fp16 SUBNORMALS n * 2^-24 -- no magic constant, no
//    subtract: one LOP3 per two elements.  The 16x16 fragment of raw codes times the x fragment
//    (8 activation rows per mma) goes through mma.sync with fp32 accumulation per 64-block;
//    the block scale d * 2^24 and the code offset (-8 d for sym_int4, m for asym_int4, times
//    sum(x) of the block, obtained once per super-tile from a ones-row mma) are applied in fp32.
//    All products are exact in fp32, so the result differs from the fp32 serial-k statement only
//    by summation order.  bf16 activations use the 0x4300 magic (128 + n) and one subtract;
//    fp32 activations are split into three bf16 planes (hi/mid/lo).  fp8/fp6 use cvt.
//  * One x fragment, one sum(x) and one 16-byte scale load serve the four row tiles of a
//    super-tile (64 rows x 64 k).
//  * Work split ("stream-K"): warp w of W owns the contiguous tile span [w*T/W, (w+1)*T/W):
//    balanced over 148 SMs whatever N and K are.  A warp that does not cover a whole row tile
//    writes its fp32 partial fragment to the workspace and bumps that row tile's arrival counter;
//    the last arriver adds the partials in warp order (deterministic, no float atomics) and
//    stores.  Counters return to zero.
//  * griddepcontrol.wait sits between the first weight copies and the first read of x, so under
//    programmatic dependent launch the weight stream of call i+1 overlaps the tail of call i.
#pragma once
#include <cuda_fp8.h>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"
#include "gemv_args.h"

namespace lowbit {

// ------------------------------------------------------------------ small PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint4 lds16(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ uint2 lds8(uint32_t a) {
    uint2 r;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a));
    return r;
}
__device__ __forceinline__ uint32_t lds4(uint32_t a) {
    uint32_t r;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a));
    return r;
}
__device__ __forceinline__ uint4 ldg_cached16(const void* p) {   // plain cached load (activations)
    uint4 r;
    asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
template <bool BF>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if (BF)
        asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// (a & b) | c in one LOP3 (nvcc otherwise emits AND and OR separately when it hoists the ANDs)
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t h2_as_u32(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 u32_as_h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t bf2_as_u32(__nv_bfloat162 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __nv_bfloat162 u32_as_bf2(uint32_t u) { return *reinterpret_cast<__nv_bfloat162*>(&u); }

// fp16x2 (exactly representable small values) -> bf16x2
__device__ __forceinline__ uint32_t h2_to_bf2(uint32_t h) {
    const float2 f = __half22float2(u32_as_h2(h));
    return bf2_as_u32(__floats2bfloat162_rn(f.x, f.y));
}

// ------------------------------------------------------------------ per-qtype codecs
// kTileBytes : code bytes of one 16x64 tile in the Q region
// kOffset    : 0 none, 1 value = (code - 8) * d, 2 value = code * d + m (m from the M region)
// kLo / kHi  : fp32 factors that undo the scaling of fragment rows g (regs a0,a2) / g+8 (a1,a3)
// Tile       : the registers one lane holds for one tile
// load       : read the lane's part of tile `r` of a staged super-tile from shared memory
// frag       : A fragment (4 x b32) of k-step s: code values, UNSCALED, offset NOT applied
template <int QT, bool BF> struct Codec;

template <bool BF, int OFFSET> struct Codec4 {
    struct Tile { uint4 q; };
    static constexpr int kTileBytes = 512;
    static constexpr int kOffset = OFFSET;
    static constexpr float kLo = BF ? 1.0f : 16777216.0f;    // fp16 path: a0,a2 hold n * 2^-24
    static constexpr float kHi = BF ? 1.0f : 1048576.0f;     //            a1,a3 hold n * 2^-20
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, int r, int lane) {
        t.q = lds16(codes + r * 512 + lane * 16);
    }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[4]) {
        const uint32_t w = s == 0 ? t.q.x : s == 1 ? t.q.y : s == 2 ? t.q.z : t.q.w;
        if (!BF) {
            // fp16 subnormals: bits 0x000n = n * 2^-24, bits 0x00n0 = n * 2^-20 (exact, no offset)
            const uint32_t w2 = w >> 8;
            a[0] = w & 0x000f000fu; a[1] = w & 0x00f000f0u;
            a[2] = w2 & 0x000f000fu; a[3] = w2 & 0x00f000f0u;
        } else {
            // bf16: 0x4300 | n = 128 + n (7 mantissa bits: exact), then subtract 128
            const __nv_bfloat162 sub = __floats2bfloat162_rn(128.0f, 128.0f);
#pragma unroll
            for (int i = 0; i < 4; i++)
                a[i] = bf2_as_u32(__hsub2(u32_as_bf2(and_or(w >> (4 * i), 0x000f000fu, 0x43004300u)), sub));
        }
    }
};
template <bool BF> struct Codec<QT_SYM_INT4, BF> : Codec4<BF, 1> {};
template <bool BF> struct Codec<QT_ASYM_INT4, BF> : Codec4<BF, 2> {};

template <int QT, bool BF> struct Codec8 {
    struct Tile { uint4 q0, q1; };
    static constexpr int kTileBytes = 1024;
    static constexpr int kOffset = 0;
    static constexpr float kLo = 1.0f, kHi = 1.0f;
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, int r, int lane) {
        t.q0 = lds16(codes + r * 1024 + lane * 16);
        t.q1 = lds16(codes + r * 1024 + 512 + lane * 16);
    }
    __device__ static __forceinline__ uint32_t pair(uint32_t w, int hi) {   // two codes -> 2 x fp16/bf16
        uint32_t h;
        if (QT == QT_FP8_E5M2) {
            h = __byte_perm(w, 0u, hi ? 0x3424u : 0x1404u);                  // e5m2 = top byte of fp16
        } else {
            const unsigned short two = (unsigned short)(hi ? (w >> 16) : (w & 0xffffu));
            asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h) : "h"(two));
        }
        return BF ? h2_to_bf2(h) : h;
    }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[4]) {
        const uint32_t w0 = s == 0 ? t.q0.x : s == 1 ? t.q0.y : s == 2 ? t.q0.z : t.q0.w;
        const uint32_t w1 = s == 0 ? t.q1.x : s == 1 ? t.q1.y : s == 2 ? t.q1.z : t.q1.w;
        a[0] = pair(w0, 0); a[1] = pair(w1, 0); a[2] = pair(w0, 1); a[3] = pair(w1, 1);
    }
};
template <bool BF> struct Codec<QT_FP8_E4M3, BF> : Codec8<QT_FP8_E4M3, BF> {};
template <bool BF> struct Codec<QT_FP8_E5M2, BF> : Codec8<QT_FP8_E5M2, BF> {};

template <bool BF> struct Codec<QT_FP6, BF> {
    struct Tile { uint4 lo; uint2 hi; };
    static constexpr int kTileBytes = 768;
    static constexpr int kOffset = 0;
    static constexpr float kLo = 1.0f, kHi = 1.0f;
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, int r, int lane) {
        t.lo = lds16(codes + r * 768 + lane * 16);
        t.hi = lds8(codes + r * 768 + 512 + lane * 8);
    }
    __device__ static __forceinline__ uint32_t cvt2(unsigned short two) {
        uint32_t h;
        asm("cvt.rn.f16x2.e3m2x2 %0, %1;" : "=r"(h) : "h"(two));
        return BF ? h2_to_bf2(h) : h;
    }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[4]) {
        const int u = s >> 1;
        const uint32_t l0 = u ? t.lo.y : t.lo.x, l1 = u ? t.lo.w : t.lo.z;
        const uint32_t sh = (s & 1) ? 4u : 0u;
        const uint32_t e0 = ((l0 >> sh) & 0x0f0f0f0fu) | (((t.hi.x >> (2 * s)) & 0x03030303u) << 4);
        const uint32_t e1 = ((l1 >> sh) & 0x0f0f0f0fu) | (((t.hi.y >> (2 * s)) & 0x03030303u) << 4);
        a[0] = cvt2((unsigned short)(e0 & 0xffffu)); a[1] = cvt2((unsigned short)(e1 & 0xffffu));
        a[2] = cvt2((unsigned short)(e0 >> 16));     a[3] = cvt2((unsigned short)(e1 >> 16));
    }
};

// staging geometry: a chunk = kChunk super-tiles; per warp kStages chunks of [codes | scales | mins]
template <int QT, bool BF> struct Staging {
    using C = Codec<QT, BF>;
    static constexpr int kSuperCodes = RG_TILES * C::kTileBytes;                  // 2048 / 4096 / 3072
    static constexpr int kChunk = kSuperCodes <= 2048 ? 2 : 1;                    // ~4 KB of codes per bulk copy
    static constexpr int kStages = 3;
    static constexpr int kMinBytes = (C::kOffset == 2) ? 128 : 0;
    static constexpr int kStageBytes = kChunk * (kSuperCodes + 128 + kMinBytes);
    static constexpr int kSmemBytes = GEMV_WARPS * (kStages * kStageBytes + kStages * 8);
};

// ------------------------------------------------------------------ the kernel
// ACC = accumulators per (row tile, n-tile): 4 (c0..c3), or 2 when M == 1 (only column 0 is live:
// c0 = row g, c2 = row g+8).
//
// Row group `rg` is finished for this warp, which covered its k-blocks [kb_a, kb_b).  Rare path
// (once or twice per warp), kept out of line so the main loop stays small.
template <int NT, int ACC>
__device__ __noinline__ void flush_row_group(const GemvArgs& A, const float (&acc)[RG_TILES][NT][ACC], int64_t wid,
                                             int64_t u0, int rg, int kb_a, int kb_b, int lane) {
    constexpr int NV = RG_TILES * NT * ACC;            // floats per lane
    const int64_t T = A.G.nsuper, W = A.total_warps, nb = A.G.nb;
    const int64_t r_begin = (int64_t)rg * nb, r_end = r_begin + nb;
    float tot[RG_TILES][NT][ACC];
    bool do_store = false;
    if (kb_a == 0 && kb_b == (int)nb) {
        do_store = true;
#pragma unroll
        for (int r = 0; r < RG_TILES; r++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int c = 0; c < ACC; c++) tot[r][j][c] = acc[r][j][c];
    } else {
        const int which = (u0 >= r_begin) ? 0 : 1;
        float* P = A.partials + (wid * 2 + which) * (NV * 32);
#pragma unroll
        for (int r = 0; r < RG_TILES; r++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int c = 0; c < ACC; c++) __stcg(P + ((r * NT + j) * ACC + c) * 32 + lane, acc[r][j][c]);
        __threadfence();
        // owners of the first and last super-tile of this row group: owner(u) = ((u+1)*W - 1) / T
        const int64_t w_first = ((r_begin + 1) * W - 1) / T;
        const int64_t w_last = (r_end * W - 1) / T;
        const unsigned need = (unsigned)(w_last - w_first + 1);
        unsigned old = 0;
        if (lane == 0) old = atomicAdd(A.counters + rg, 1u);
        old = __shfl_sync(0xffffffffu, old, 0);
        if (old == need - 1) {
            __threadfence();
            do_store = true;
#pragma unroll
            for (int r = 0; r < RG_TILES; r++)
#pragma unroll
                for (int j = 0; j < NT; j++)
#pragma unroll
                    for (int c = 0; c < ACC; c++) tot[r][j][c] = 0.0f;
            // every contributor but the first starts inside this row group (slot 0); the first one
            // uses slot 1 iff its span began in an earlier row group
            const int first_slot = (w_first * T / W >= r_begin) ? 0 : 1;
            for (int64_t wc = w_first; wc <= w_last; wc++) {
                const float* Q = A.partials + (wc * 2 + (wc == w_first ? first_slot : 0)) * (NV * 32);
                float v[NV];
#pragma unroll
                for (int i = 0; i < NV; i++) v[i] = __ldcg(Q + i * 32 + lane);
#pragma unroll
                for (int r = 0; r < RG_TILES; r++)
#pragma unroll
                    for (int j = 0; j < NT; j++)
#pragma unroll
                        for (int c = 0; c < ACC; c++) tot[r][j][c] += v[(r * NT + j) * ACC + c];
            }
            if (lane == 0) A.counters[rg] = 0u;
        }
    }
    if (do_store) {
        const int g = lane >> 2, t = lane & 3;
        const int64_t N = A.G.rows;
#pragma unroll
        for (int r = 0; r < RG_TILES; r++) {
            const int64_t r0 = (int64_t)rg * RG_ROWS + r * 16 + g, r1 = r0 + 8;
#pragma unroll
            for (int j = 0; j < NT; j++) {
                const int m0 = 8 * j + 2 * t;
                if (ACC == 2) {
                    if (t == 0) {
                        store_out(A.y, A.y_dtype, r0, tot[r][j][0], A.bias, r0, A.accumulate);
                        store_out(A.y, A.y_dtype, r1, tot[r][j][ACC - 1], A.bias, r1, A.accumulate);
                    }
                } else {
                    if (m0 < A.M) {
                        store_out(A.y, A.y_dtype, (int64_t)m0 * N + r0, tot[r][j][0], A.bias, r0, A.accumulate);
                        store_out(A.y, A.y_dtype, (int64_t)m0 * N + r1, tot[r][j][2 % ACC], A.bias, r1, A.accumulate);
                    }
                    if (m0 + 1 < A.M) {
                        store_out(A.y, A.y_dtype, (int64_t)(m0 + 1) * N + r0, tot[r][j][1], A.bias, r0, A.accumulate);
                        store_out(A.y, A.y_dtype, (int64_t)(m0 + 1) * N + r1, tot[r][j][3 % ACC], A.bias, r1, A.accumulate);
                    }
                }
            }
        }
    }
}

template <int QT, bool BF, int NS, int NT, int ACC>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
gemv_kernel(const GemvArgs A) {
    using C = Codec<QT, BF>;
    using S = Staging<QT, BF>;
    using Tile = typename C::Tile;
    constexpr int CH = S::kChunk, STAGES = S::kStages, STAGE_BYTES = S::kStageBytes;
    constexpr int SUPER = S::kSuperCodes;
    extern __shared__ __align__(128) uint8_t smem[];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int64_t wid = (int64_t)blockIdx.x * GEMV_WARPS + warp;
    if (A.pdl) asm volatile("griddepcontrol.launch_dependents;");
    if (wid >= A.total_warps) return;                 // warps are independent: no CTA-wide barrier below
    const int64_t u0 = wid * A.G.nsuper / A.total_warps, u1 = (wid + 1) * A.G.nsuper / A.total_warps;
    const int n_super = (int)(u1 - u0);               // >= 1 because total_warps <= nsuper
    const int nb = (int)A.G.nb;
    const int n_chunks = (n_super + CH - 1) / CH;

    // this warp's private ring: [STAGES][codes | scales | mins], then STAGES mbarriers
    const uint32_t ring = smem_u32(smem) + warp * (STAGES * STAGE_BYTES);
    const uint32_t bars = smem_u32(smem) + GEMV_WARPS * (STAGES * STAGE_BYTES) + warp * (STAGES * 8);
    const uint8_t* gq = A.w + u0 * SUPER;
    const uint8_t* gs = A.w + A.G.off_s + u0 * 128;
    const uint8_t* gm = A.w + A.G.off_m + u0 * 128;

    auto issue = [&](int c) {                         // lane 0: copy chunk c into stage c % STAGES
        const int st = c % STAGES;
        const int ns = min(CH, n_super - c * CH);
        const uint32_t dst = ring + st * STAGE_BYTES, bar = bars + st * 8;
        mbar_expect_tx(bar, ns * (SUPER + 128 + S::kMinBytes));
        bulk_g2s(dst, gq + (size_t)c * (CH * SUPER), ns * SUPER, bar);
        bulk_g2s(dst + CH * SUPER, gs + (size_t)c * (CH * 128), ns * 128, bar);
        if (S::kMinBytes) bulk_g2s(dst + CH * (SUPER + 128), gm + (size_t)c * (CH * 128), ns * 128, bar);
    };

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(bars + s * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        // weights do not depend on the previous kernel: start streaming before the PDL wait
        for (int c = 0; c < STAGES && c < n_chunks; c++) issue(c);
    }
    __syncwarp();
    if (A.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");

    float acc[RG_TILES][NT][ACC];
#pragma unroll
    for (int r = 0; r < RG_TILES; r++)
#pragma unroll
        for (int j = 0; j < NT; j++)
#pragma unroll
            for (int c = 0; c < ACC; c++) acc[r][j][c] = 0.0f;
    int rg = (int)(u0 / nb);
    int kb = (int)(u0 - (int64_t)rg * nb);
    int kb_first = kb;              // first k-block of the current row group handled by this warp

    // lane (g,t) reads x[8j+g][kb*64 + 16t .. +15] for n-tile j (plane p); rows >= M read as zero
    const char* xbase[NT][NS];
    bool xvalid[NT];
    uint4 xb[NT][NS][2];            // lanes without a valid row keep zeros
#pragma unroll
    for (int j = 0; j < NT; j++) {
        xvalid[j] = (8 * j + g) < A.M;
#pragma unroll
        for (int p = 0; p < NS; p++) {
            xbase[j][p] = reinterpret_cast<const char*>(A.x + p * A.plane_stride + (int64_t)(8 * j + g) * A.G.k + t * 16);
            xb[j][p][0] = make_uint4(0, 0, 0, 0); xb[j][p][1] = make_uint4(0, 0, 0, 0);
        }
    }
    unsigned xoff = (unsigned)kb * 128u;   // byte offset of the current k-block inside an x row

    uint32_t ones[4];
    ones[0] = ones[1] = ones[2] = ones[3] = BF ? 0x3f803f80u : 0x3c003c00u;

    // hand the accumulators to the out-of-line flush through a by-value copy so that `acc` itself
    // never has to live in local memory
    auto flush_acc = [&]() {
        float tmp[RG_TILES][NT][ACC];
#pragma unroll
        for (int r = 0; r < RG_TILES; r++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int c = 0; c < ACC; c++) { tmp[r][j][c] = acc[r][j][c]; acc[r][j][c] = 0.0f; }
        flush_row_group<NT, ACC>(A, tmp, wid, u0, rg, kb_first, kb, lane);
    };

    for (int c = 0; c < n_chunks; c++) {
        const int st = c % STAGES;
        const uint32_t stage = ring + st * STAGE_BYTES;
        const int ns = min(CH, n_super - c * CH);
        mbar_wait(bars + st * 8, (uint32_t)((c / STAGES) & 1));
#pragma unroll
        for (int i = 0; i < CH; i++) {
            if (i < ns) {
                const uint32_t codes = stage + i * SUPER;
                // ---- once per super-tile: x fragment, sum(x), scales of the lane's 8 rows
#pragma unroll
                for (int j = 0; j < NT; j++)
#pragma unroll
                    for (int p = 0; p < NS; p++)
                        if (xvalid[j]) {                        // L1-cached: x is re-read for every row group
                            xb[j][p][0] = ldg_cached16(xbase[j][p] + xoff);
                            xb[j][p][1] = ldg_cached16(xbase[j][p] + xoff + 16);
                        }
                float xs[NT][4];                                // sum over the k-block of x (all 16 "rows" equal)
                if (C::kOffset) {
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        xs[j][0] = xs[j][1] = xs[j][2] = xs[j][3] = 0.0f;
#pragma unroll
                        for (int p = 0; p < NS; p++) {
                            mma16816<BF>(xs[j], ones, xb[j][p][0].x, xb[j][p][0].y);
                            mma16816<BF>(xs[j], ones, xb[j][p][0].z, xb[j][p][0].w);
                            mma16816<BF>(xs[j], ones, xb[j][p][1].x, xb[j][p][1].y);
                            mma16816<BF>(xs[j], ones, xb[j][p][1].z, xb[j][p][1].w);
                        }
                    }
                }
                const uint4 dq = lds16(stage + CH * SUPER + i * 128 + g * 16);       // [r][h] fp16 d
                uint4 mq = make_uint4(0, 0, 0, 0);
                if (C::kOffset == 2) mq = lds16(stage + CH * (SUPER + 128) + i * 128 + g * 16);
#pragma unroll
                for (int r = 0; r < RG_TILES; r++) {
                    Tile tile;
                    C::load(tile, codes, r, lane);
                    float cc[NT][4];
#pragma unroll
                    for (int j = 0; j < NT; j++) cc[j][0] = cc[j][1] = cc[j][2] = cc[j][3] = 0.0f;
#pragma unroll
                    for (int s = 0; s < 4; s++) {
                        uint32_t a[4];
                        C::frag(tile, s, a);
#pragma unroll
                        for (int j = 0; j < NT; j++)
#pragma unroll
                            for (int p = 0; p < NS; p++) {
                                const uint4 v = xb[j][p][s >> 1];
                                mma16816<BF>(cc[j], a, (s & 1) ? v.z : v.x, (s & 1) ? v.w : v.y);
                            }
                    }
                    const uint32_t dr = r == 0 ? dq.x : r == 1 ? dq.y : r == 2 ? dq.z : dq.w;
                    const float2 d = __half22float2(u32_as_h2(dr));
                    const float dlo = d.x * C::kLo, dhi = d.y * C::kHi;
                    float olo = 0.0f, ohi = 0.0f;
                    if (C::kOffset == 1) { olo = -8.0f * d.x; ohi = -8.0f * d.y; }
                    if (C::kOffset == 2) {
                        const uint32_t mr = r == 0 ? mq.x : r == 1 ? mq.y : r == 2 ? mq.z : mq.w;
                        const float2 mn = __half22float2(u32_as_h2(mr));
                        olo = mn.x; ohi = mn.y;
                    }
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        if (ACC == 2) {
                            acc[r][j][0] = fmaf(dlo, cc[j][0], acc[r][j][0]);
                            acc[r][j][1] = fmaf(dhi, cc[j][2], acc[r][j][1]);
                            if (C::kOffset) {
                                acc[r][j][0] = fmaf(olo, xs[j][0], acc[r][j][0]);
                                acc[r][j][1] = fmaf(ohi, xs[j][0], acc[r][j][1]);
                            }
                        } else {
                            acc[r][j][0] = fmaf(dlo, cc[j][0], acc[r][j][0]); acc[r][j][1] = fmaf(dlo, cc[j][1], acc[r][j][1]);
                            acc[r][j][2 % ACC] = fmaf(dhi, cc[j][2], acc[r][j][2 % ACC]);
                            acc[r][j][3 % ACC] = fmaf(dhi, cc[j][3], acc[r][j][3 % ACC]);
                            if (C::kOffset) {
                                acc[r][j][0] = fmaf(olo, xs[j][0], acc[r][j][0]); acc[r][j][1] = fmaf(olo, xs[j][1], acc[r][j][1]);
                                acc[r][j][2 % ACC] = fmaf(ohi, xs[j][0], acc[r][j][2 % ACC]);
                                acc[r][j][3 % ACC] = fmaf(ohi, xs[j][1], acc[r][j][3 % ACC]);
                            }
                        }
                    }
                }
                xoff += 128u;
                if (++kb == nb) {                     // end of the row group: flush and wrap
                    flush_acc();
                    rg++; kb = 0; kb_first = 0; xoff = 0u;
                }
            }
        }
        // every lane has consumed the stage (its data sits in registers already used): refill it
        __syncwarp();
        if (lane == 0 && c + STAGES < n_chunks) issue(c + STAGES);
    }
    if (kb != kb_first) flush_acc();
}

template <int QT, bool BF, int NS>
static void launch_nt(int m, int grid, const GemvArgs& A, cudaStream_t st) {
    constexpr int SMEM = Staging<QT, BF>::kSmemBytes;
    static bool configured = false;     // per instantiation: raise the dynamic shared memory limit once
    if (!configured) {
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        configured = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(GEMV_WARPS * 32); cfg.dynamicSmemBytes = SMEM; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = A.pdl ? 1 : 0;
    if (m == 1) cudaLaunchKernelEx(&cfg, gemv_kernel<QT, BF, NS, 1, 2>, A);
    else if (m <= 8) cudaLaunchKernelEx(&cfg, gemv_kernel<QT, BF, NS, 1, 4>, A);
    else if (m <= 16) cudaLaunchKernelEx(&cfg, gemv_kernel<QT, BF, NS, 2, 4>, A);
    else cudaLaunchKernelEx(&cfg, gemv_kernel<QT, BF, NS, 4, 4>, A);
}
template <int QT>
static void launch_qt(int x_dtype, int m, int grid, const GemvArgs& A, cudaStream_t st) {
    if (x_dtype == DT_F16) launch_nt<QT, false, 1>(m, grid, A, st);
    else if (x_dtype == DT_BF16) launch_nt<QT, true, 1>(m, grid, A, st);
    else launch_nt<QT, true, 3>(m, grid, A, st);
}

// resident warps per SM this qtype's kernel is sized for (shared memory bound)
template <int QT> static int warps_per_sm_qt() {
    const int per_cta = Staging<QT, false>::kSmemBytes + 1024;      // + driver reservation per CTA
    int ctas = (227 * 1024) / per_cta;
    if (ctas * GEMV_WARPS > GEMV_MAX_WARPS_PER_SM) ctas = GEMV_MAX_WARPS_PER_SM / GEMV_WARPS;
    return ctas * GEMV_WARPS;
}

// one translation unit per qtype instantiates this (keeps nvcc times parallel)
#define LOWBIT_DEFINE_GEMV_QT(NAME, QT)                                                         \
    void gemv_launch_##NAME(int x_dtype, int m, int grid, const GemvArgs& A, cudaStream_t st) { \
        launch_qt<QT>(x_dtype, m, grid, A, st);                                                 \
    }                                                                                           \
    int gemv_warps_per_sm_##NAME() { return warps_per_sm_qt<QT>(); }

}  // namespace lowbit
