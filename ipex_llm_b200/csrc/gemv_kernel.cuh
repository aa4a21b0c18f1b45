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
//  * In-register dequant: 4-bit codes become exact fp16/bf16 integers with the magic-number trick
//    (one LOP3 + one HSUB2/HFMA2 per two elements), fp8/fp6 with cvt; codes stay UNscaled.  The
//    16x16 fragment times the x fragment (8 activation rows per mma) goes through mma.sync with
//    fp32 accumulation per 64-block; the fp16 block scale d (and m for asym_int4, through a
//    ones-row mma that yields sum(x)) is applied in fp32.  Products of small integers and
//    fp16/bf16 activations are exact in fp32, so the result differs from the fp32 serial-k
//    statement only by summation order.  fp32 activations are split into three bf16 planes
//    (hi/mid/lo): two extra mma per step.
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
// Tile       : the registers one lane holds for one tile
// load       : read the lane's part of tile `i` of a staged chunk from shared memory
// frag       : A fragment (4 x b32) of k-step s, UNSCALED code values in fp16 (BF=false) / bf16
template <int QT, bool BF> struct Codec;

// value = code - OFF  (OFF = 8 for sym_int4, 0 for asym_int4)
template <bool BF, int OFF> __device__ __forceinline__ void frag4(uint32_t w, uint32_t (&a)[4]) {
    if (!BF) {
        // fp16: 0x6400 | n = 1024 + n ; 0x6400 | (n << 4) = 1024 + 16 n
        const __half2 sub_lo = __floats2half2_rn(1024.0f + OFF, 1024.0f + OFF);
        const __half2 mul_hi = __floats2half2_rn(0.0625f, 0.0625f);
        const __half2 add_hi = __floats2half2_rn(-64.0f - OFF, -64.0f - OFF);
        const uint32_t w2 = w >> 8;
        a[0] = h2_as_u32(__hsub2(u32_as_h2(and_or(w, 0x000f000fu, 0x64006400u)), sub_lo));
        a[1] = h2_as_u32(__hfma2(u32_as_h2(and_or(w, 0x00f000f0u, 0x64006400u)), mul_hi, add_hi));
        a[2] = h2_as_u32(__hsub2(u32_as_h2(and_or(w2, 0x000f000fu, 0x64006400u)), sub_lo));
        a[3] = h2_as_u32(__hfma2(u32_as_h2(and_or(w2, 0x00f000f0u, 0x64006400u)), mul_hi, add_hi));
    } else {
        // bf16: 0x4300 | n = 128 + n (7 mantissa bits: exact for n < 128)
        const __nv_bfloat162 sub = __floats2bfloat162_rn(128.0f + OFF, 128.0f + OFF);
#pragma unroll
        for (int i = 0; i < 4; i++)
            a[i] = bf2_as_u32(__hsub2(u32_as_bf2(and_or(w >> (4 * i), 0x000f000fu, 0x43004300u)), sub));
    }
}

template <bool BF, int OFF, bool HAS_MIN> struct Codec4 {
    struct Tile { uint4 q; uint32_t d; uint32_t m; };
    static constexpr int kTileBytes = 512;
    static constexpr bool kHasMin = HAS_MIN;
    // codes/scales/mins: shared-memory addresses of the staged chunk; i = tile inside the chunk
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, uint32_t scales, uint32_t mins, int i, int lane) {
        t.q = lds16(codes + i * 512 + lane * 16);
        t.d = lds4(scales + i * 32 + (lane >> 2) * 4);
        if (HAS_MIN) t.m = lds4(mins + i * 32 + (lane >> 2) * 4);
    }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[4]) {
        const uint32_t w = s == 0 ? t.q.x : s == 1 ? t.q.y : s == 2 ? t.q.z : t.q.w;
        frag4<BF, OFF>(w, a);
    }
};
template <bool BF> struct Codec<QT_SYM_INT4, BF> : Codec4<BF, 8, false> {};
template <bool BF> struct Codec<QT_ASYM_INT4, BF> : Codec4<BF, 0, true> {};

template <int QT, bool BF> struct Codec8 {
    struct Tile { uint4 q0, q1; uint32_t d; uint32_t m; };
    static constexpr int kTileBytes = 1024;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, uint32_t scales, uint32_t, int i, int lane) {
        t.q0 = lds16(codes + i * 1024 + lane * 16);
        t.q1 = lds16(codes + i * 1024 + 512 + lane * 16);
        t.d = lds4(scales + i * 32 + (lane >> 2) * 4);
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
    struct Tile { uint4 lo; uint2 hi; uint32_t d; uint32_t m; };
    static constexpr int kTileBytes = 768;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, uint32_t scales, uint32_t, int i, int lane) {
        t.lo = lds16(codes + i * 768 + lane * 16);
        t.hi = lds8(codes + i * 768 + 512 + lane * 8);
        t.d = lds4(scales + i * 32 + (lane >> 2) * 4);
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

// shared-memory bytes one warp needs: kStages x (TPC tiles of codes + scales (+ mins))
template <int QT, bool BF> __host__ __device__ constexpr int gemv_stage_bytes() {
    return GEMV_TPC * (Codec<QT, BF>::kTileBytes + 32 + (Codec<QT, BF>::kHasMin ? 32 : 0));
}
template <int QT, bool BF> __host__ __device__ constexpr int gemv_smem_bytes() {
    return GEMV_WARPS * (GEMV_STAGES * gemv_stage_bytes<QT, BF>() + GEMV_STAGES * 8);
}

// ------------------------------------------------------------------ the kernel
// Row tile `rt` is finished for this warp, which covered its k-blocks [kb_a, kb_b).  Rare path
// (once or twice per warp), kept out of line so the tile loop stays small.
template <int NT>
__device__ __noinline__ void flush_row_tile(const GemvArgs& A, const float (&acc)[NT][4], int64_t wid, int64_t f0,
                                            int rt, int kb_a, int kb_b, int lane) {
    const int64_t T = A.G.ntiles, W = A.total_warps, nb = A.G.nb;
    const int64_t r_begin = (int64_t)rt * nb, r_end = r_begin + nb;
    float tot[NT][4];
    bool do_store = false;
    if (kb_a == 0 && kb_b == (int)nb) {
        do_store = true;
#pragma unroll
        for (int j = 0; j < NT; j++) { tot[j][0] = acc[j][0]; tot[j][1] = acc[j][1]; tot[j][2] = acc[j][2]; tot[j][3] = acc[j][3]; }
    } else {
        const int which = (f0 >= r_begin) ? 0 : 1;
        float* P = A.partials + ((wid * 2 + which) * NT) * 128;
#pragma unroll
        for (int j = 0; j < NT; j++)
#pragma unroll
            for (int r = 0; r < 4; r++) __stcg(P + (j * 4 + r) * 32 + lane, acc[j][r]);
        __threadfence();
        // owners of the first and last tile of this row tile: owner(f) = ((f+1)*W - 1) / T
        const int64_t w_first = ((r_begin + 1) * W - 1) / T;
        const int64_t w_last = (r_end * W - 1) / T;
        const unsigned need = (unsigned)(w_last - w_first + 1);
        unsigned old = 0;
        if (lane == 0) old = atomicAdd(A.counters + rt, 1u);
        old = __shfl_sync(0xffffffffu, old, 0);
        if (old == need - 1) {
            __threadfence();
            do_store = true;
#pragma unroll
            for (int j = 0; j < NT; j++) { tot[j][0] = tot[j][1] = tot[j][2] = tot[j][3] = 0.0f; }
            for (int64_t wc = w_first; wc <= w_last; wc++) {
                const int wh = (wc * T / W >= r_begin) ? 0 : 1;
                const float* Q = A.partials + ((wc * 2 + wh) * NT) * 128;
#pragma unroll
                for (int j = 0; j < NT; j++)
#pragma unroll
                    for (int r = 0; r < 4; r++) tot[j][r] += __ldcg(Q + (j * 4 + r) * 32 + lane);
            }
            if (lane == 0) A.counters[rt] = 0u;
        }
    }
    if (do_store) {
        const int g = lane >> 2, t = lane & 3;
        const int64_t r0 = (int64_t)rt * 16 + g, r1 = r0 + 8, N = A.G.rows;
#pragma unroll
        for (int j = 0; j < NT; j++) {
            const int m0 = 8 * j + 2 * t;
            if (m0 < A.M) {
                store_out(A.y, A.y_dtype, (int64_t)m0 * N + r0, tot[j][0], A.bias, r0, A.accumulate);
                store_out(A.y, A.y_dtype, (int64_t)m0 * N + r1, tot[j][2], A.bias, r1, A.accumulate);
            }
            if (m0 + 1 < A.M) {
                store_out(A.y, A.y_dtype, (int64_t)(m0 + 1) * N + r0, tot[j][1], A.bias, r0, A.accumulate);
                store_out(A.y, A.y_dtype, (int64_t)(m0 + 1) * N + r1, tot[j][3], A.bias, r1, A.accumulate);
            }
        }
    }
}

template <int QT, bool BF, int NS, int NT>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
gemv_kernel(const GemvArgs A) {
    using C = Codec<QT, BF>;
    using Tile = typename C::Tile;
    constexpr int TPC = GEMV_TPC, STAGES = GEMV_STAGES;
    constexpr int TILE = C::kTileBytes;
    constexpr int STAGE_BYTES = gemv_stage_bytes<QT, BF>();
    extern __shared__ __align__(128) uint8_t smem[];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int64_t wid = (int64_t)blockIdx.x * GEMV_WARPS + warp;
    const int64_t f0 = wid * A.G.ntiles / A.total_warps, f1 = (wid + 1) * A.G.ntiles / A.total_warps;
    const int n_tiles = (int)(f1 - f0);
    if (n_tiles <= 0) return;                         // warps are independent: no CTA-wide barrier below
    const int nb = (int)A.G.nb;
    const int n_chunks = (n_tiles + TPC - 1) / TPC;

    // this warp's private ring: [STAGES][codes | scales | mins], then STAGES mbarriers
    const uint32_t ring = smem_u32(smem) + warp * (STAGES * STAGE_BYTES);
    const uint32_t bars = smem_u32(smem) + GEMV_WARPS * (STAGES * STAGE_BYTES) + warp * (STAGES * 8);
    const uint8_t* gq = A.w + f0 * TILE;
    const uint8_t* gs = A.w + A.G.off_s + f0 * 32;
    const uint8_t* gm = A.w + A.G.off_m + f0 * 32;

    auto issue = [&](int c) {                         // lane 0: copy chunk c into stage c % STAGES
        const int st = c % STAGES;
        const int nt_c = min(TPC, n_tiles - c * TPC);
        const uint32_t dst = ring + st * STAGE_BYTES, bar = bars + st * 8;
        const uint32_t tx = nt_c * (TILE + 32 + (C::kHasMin ? 32 : 0));
        mbar_expect_tx(bar, tx);
        bulk_g2s(dst, gq + (size_t)c * (TPC * TILE), nt_c * TILE, bar);
        bulk_g2s(dst + TPC * TILE, gs + (size_t)c * (TPC * 32), nt_c * 32, bar);
        if (C::kHasMin) bulk_g2s(dst + TPC * (TILE + 32), gm + (size_t)c * (TPC * 32), nt_c * 32, bar);
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
    asm volatile("griddepcontrol.wait;" ::: "memory");

    float acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f; }
    int rt = (int)(f0 / nb);
    int kb = (int)(f0 - (int64_t)rt * nb);
    int kb_first = kb;              // first k-block of the current row tile handled by this warp

    // lane (g,t) reads x[8j+g][kb*64 + 16t .. +15] for n-tile j (plane p); rows >= M read as zero
    const char* xbase[NT][NS];
    bool xvalid[NT];
    uint4 xb[NT][NS][2];            // persistent: lanes without a valid row keep zeros
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
        float tmp[NT][4];
#pragma unroll
        for (int j = 0; j < NT; j++) {
#pragma unroll
            for (int r = 0; r < 4; r++) { tmp[j][r] = acc[j][r]; acc[j][r] = 0.0f; }
        }
        if (!(A.dbg & 1)) flush_row_tile<NT>(A, tmp, wid, f0, rt, kb_first, kb, lane);
        else if (tmp[0][0] == 123.456f) A.partials[0] = tmp[0][1] + tmp[0][2] + tmp[0][3];
    };

    for (int c = 0; c < n_chunks; c++) {
        const int st = c % STAGES;
        const uint32_t codes = ring + st * STAGE_BYTES;
        const uint32_t scales = codes + TPC * TILE, mins = scales + TPC * 32;
        const int nt_c = min(TPC, n_tiles - c * TPC);
        mbar_wait(bars + st * 8, (uint32_t)((c / STAGES) & 1));
#pragma unroll
        for (int i = 0; i < TPC; i++) {
            if (i < nt_c) {
                Tile tile;
                C::load(tile, codes, scales, mins, i, lane);
#pragma unroll
                for (int j = 0; j < NT; j++)
#pragma unroll
                    for (int p = 0; p < NS; p++)
                        if (xvalid[j] && !(A.dbg & 2)) {                        // L1-cached: x is re-read for every row tile
                            xb[j][p][0] = ldg_cached16(xbase[j][p] + xoff);
                            xb[j][p][1] = ldg_cached16(xbase[j][p] + xoff + 16);
                        }
                float cc[NT][4], cs[NT][4];
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    cc[j][0] = cc[j][1] = cc[j][2] = cc[j][3] = 0.0f;
                    cs[j][0] = cs[j][1] = cs[j][2] = cs[j][3] = 0.0f;
                }
#pragma unroll
                for (int s = 0; s < 4; s++) {
                    uint32_t a[4];
                    C::frag(tile, s, a);
#pragma unroll
                    for (int j = 0; j < NT; j++)
#pragma unroll
                        for (int p = 0; p < NS; p++) {
                            const uint4 v = xb[j][p][s >> 1];
                            const uint32_t b0 = (s & 1) ? v.z : v.x, b1 = (s & 1) ? v.w : v.y;
                            if (!(A.dbg & 4)) mma16816<BF>(cc[j], a, b0, b1); else { cc[j][0] += __uint_as_float(a[0] ^ a[1] ^ b0); cc[j][2] += __uint_as_float(a[2] ^ a[3] ^ b1); }
                            if (C::kHasMin) mma16816<BF>(cs[j], ones, b0, b1);
                        }
                }
                const float2 d = __half22float2(u32_as_h2(tile.d));
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    acc[j][0] = fmaf(d.x, cc[j][0], acc[j][0]); acc[j][1] = fmaf(d.x, cc[j][1], acc[j][1]);
                    acc[j][2] = fmaf(d.y, cc[j][2], acc[j][2]); acc[j][3] = fmaf(d.y, cc[j][3], acc[j][3]);
                }
                if (C::kHasMin) {
                    const float2 mn = __half22float2(u32_as_h2(tile.m));
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        acc[j][0] = fmaf(mn.x, cs[j][0], acc[j][0]); acc[j][1] = fmaf(mn.x, cs[j][1], acc[j][1]);
                        acc[j][2] = fmaf(mn.y, cs[j][2], acc[j][2]); acc[j][3] = fmaf(mn.y, cs[j][3], acc[j][3]);
                    }
                }
                xoff += 128u;
                if (++kb == nb) {                     // end of the row: flush and wrap
                    flush_acc();
                    rt++; kb = 0; kb_first = 0; xoff = 0u;
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
static void launch_nt(int nt, int grid, const GemvArgs& A, cudaStream_t st) {
    constexpr int SMEM = gemv_smem_bytes<QT, BF>();
    static bool configured = false;     // per instantiation: raise the dynamic shared memory limit once
    if (!configured) {
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        configured = true;
    }
    switch (nt) {
        case 1: gemv_kernel<QT, BF, NS, 1><<<grid, GEMV_WARPS * 32, SMEM, st>>>(A); break;
        case 2: gemv_kernel<QT, BF, NS, 2><<<grid, GEMV_WARPS * 32, SMEM, st>>>(A); break;
        default: gemv_kernel<QT, BF, NS, 4><<<grid, GEMV_WARPS * 32, SMEM, st>>>(A); break;
    }
}
template <int QT>
static void launch_qt(int x_dtype, int nt, int grid, const GemvArgs& A, cudaStream_t st) {
    if (x_dtype == DT_F16) launch_nt<QT, false, 1>(nt, grid, A, st);
    else if (x_dtype == DT_BF16) launch_nt<QT, true, 1>(nt, grid, A, st);
    else launch_nt<QT, true, 3>(nt, grid, A, st);
}

// resident warps per SM this qtype's kernel is sized for (shared memory bound)
template <int QT> static int warps_per_sm_qt() {
    const int per_cta = gemv_smem_bytes<QT, false>() + 1024;      // + driver reservation per CTA
    int ctas = (227 * 1024) / per_cta;
    if (ctas * GEMV_WARPS > GEMV_MAX_WARPS_PER_SM) ctas = GEMV_MAX_WARPS_PER_SM / GEMV_WARPS;
    return ctas * GEMV_WARPS;
}

// one translation unit per qtype instantiates this (keeps nvcc times parallel)
#define LOWBIT_DEFINE_GEMV_QT(NAME, QT)                                                          \
    void gemv_launch_##NAME(int x_dtype, int nt, int grid, const GemvArgs& A, cudaStream_t st) { \
        launch_qt<QT>(x_dtype, nt, grid, A, st);                                                \
    }                                                                                            \
    int gemv_warps_per_sm_##NAME() { return warps_per_sm_qt<QT>(); }

}  // namespace lowbit
