// gemv_kernel.cuh -- kernel (b): HBM-bound low-bit dequant-GEMV for decode (M = 1..32 rows of x).
//
// Replaces the matmul behind LowBitLinear.forward for small batches: the CPU symbol
// ggml_compute_forward_mul_mat_q_fp32 (reference low_bit_linear.py:494-529) and the XPU ops
// xe_linear.forward_new / xe_batch.batch_forward (low_bit_linear.py:696-708).
//
// Design (B200-first):
//  * Weights live in the tiled device layout (lowbit_common.cuh): one warp-wide 128-bit load
//    (512 contiguous bytes) fetches a 16-row x 64-element tile with every lane receiving
//    exactly the codes of its mma.m16n8k16 A-fragment rows.  Weights are read once, with
//    ld.global.nc.L1::no_allocate, several tiles in flight per warp (register ring).
//  * In-register dequant: 4-bit codes become exact fp16/bf16 integers with the magic-number
//    trick (one LOP3 + one HSUB2/HFMA2 per two elements), fp8/fp6 with cvt; the codes stay
//    UNscaled.  The 16x16 fragment times the x fragment (M <= 8 columns per mma) goes through
//    mma.sync with fp32 accumulation per 64-block; the fp16 block scale d (and m for
//    asym_int4, through a ones-row mma that yields sum(x)) is applied in fp32.  Products of
//    small integers and fp16/bf16 activations are exact in fp32, so the result differs from
//    the fp32 serial-k statement only by summation order.  fp32 activations are split into
//    three bf16 planes (hi/mid/lo) and cost two extra mma per step.
//  * Work split ("stream-K"): tiles are numbered row-tile-major; warp w of W owns the
//    contiguous span [w*T/W, (w+1)*T/W): perfectly balanced over 148 SMs whatever N and K are.
//    A warp that does not cover a whole row tile writes its fp32 partial fragment to the
//    workspace and bumps that row tile's arrival counter; the last arriver adds the partials
//    in warp order (deterministic, no float atomics) and stores.  Counters return to zero.
//  * griddepcontrol.wait sits between the first weight loads and the first read of x, so under
//    programmatic dependent launch the weight stream of call i+1 overlaps the tail of call i.
#pragma once
#include <cuda_fp8.h>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"
#include "gemv_args.h"

namespace lowbit {

// ------------------------------------------------------------------ small PTX wrappers
__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ldg_stream8(const void* p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream4(const void* p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
template <bool BF>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if (BF)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
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
// Tile  : the registers one lane holds for one 16x64 tile (codes + scales)
// load  : issue the global loads of tile f
// frag  : A fragment (4 x b32) of k-step s, UNSCALED code values in fp16 (BF=false) / bf16
template <int QT, bool BF> struct Codec;

template <bool BF> struct Codec4Base {
    struct Tile { uint4 q; uint32_t d; uint32_t m; };
    static constexpr int kDepth = 8;
    static constexpr int kTileBytes = 512;
};

// value = code - OFF  (OFF = 8 for sym_int4, 0 for asym_int4)
template <bool BF, int OFF> __device__ __forceinline__ void frag4(uint32_t w, uint32_t (&a)[4]) {
    if (!BF) {
        // fp16: 0x6400 | n = 1024 + n ; 0x6400 | (n << 4) = 1024 + 16 n
        const __half2 sub_lo = __floats2half2_rn(1024.0f + OFF, 1024.0f + OFF);
        const __half2 mul_hi = __floats2half2_rn(0.0625f, 0.0625f);
        const __half2 add_hi = __floats2half2_rn(-64.0f - OFF, -64.0f - OFF);
        const uint32_t w2 = w >> 8;
        a[0] = h2_as_u32(__hsub2(u32_as_h2((w & 0x000f000fu) | 0x64006400u), sub_lo));
        a[1] = h2_as_u32(__hfma2(u32_as_h2((w & 0x00f000f0u) | 0x64006400u), mul_hi, add_hi));
        a[2] = h2_as_u32(__hsub2(u32_as_h2((w2 & 0x000f000fu) | 0x64006400u), sub_lo));
        a[3] = h2_as_u32(__hfma2(u32_as_h2((w2 & 0x00f000f0u) | 0x64006400u), mul_hi, add_hi));
    } else {
        // bf16: 0x4300 | n = 128 + n (7 mantissa bits: exact for n < 128)
        const __nv_bfloat162 sub = __floats2bfloat162_rn(128.0f + OFF, 128.0f + OFF);
#pragma unroll
        for (int i = 0; i < 4; i++)
            a[i] = bf2_as_u32(__hsub2(u32_as_bf2(((w >> (4 * i)) & 0x000f000fu) | 0x43004300u), sub));
    }
}

template <bool BF> struct Codec<QT_SYM_INT4, BF> : Codec4Base<BF> {
    using Tile = typename Codec4Base<BF>::Tile;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ void load(Tile& t, const uint8_t* w, const TiledGeom& G, int64_t f, int lane) {
        t.q = ldg_stream16(w + f * 512 + lane * 16);
        t.d = ldg_stream4(w + G.off_s + f * 32 + (lane >> 2) * 4);
    }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[4]) {
        const uint32_t w = s == 0 ? t.q.x : s == 1 ? t.q.y : s == 2 ? t.q.z : t.q.w;
        frag4<BF, 8>(w, a);
    }
};
template <bool BF> struct Codec<QT_ASYM_INT4, BF> : Codec4Base<BF> {
    using Tile = typename Codec4Base<BF>::Tile;
    static constexpr bool kHasMin = true;
    __device__ static __forceinline__ void load(Tile& t, const uint8_t* w, const TiledGeom& G, int64_t f, int lane) {
        t.q = ldg_stream16(w + f * 512 + lane * 16);
        t.d = ldg_stream4(w + G.off_s + f * 32 + (lane >> 2) * 4);
        t.m = ldg_stream4(w + G.off_m + f * 32 + (lane >> 2) * 4);
    }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[4]) {
        const uint32_t w = s == 0 ? t.q.x : s == 1 ? t.q.y : s == 2 ? t.q.z : t.q.w;
        frag4<BF, 0>(w, a);
    }
};

template <int QT, bool BF> struct Codec8 {
    struct Tile { uint4 q0, q1; uint32_t d; uint32_t m; };
    static constexpr int kDepth = 4;
    static constexpr int kTileBytes = 1024;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ void load(Tile& t, const uint8_t* w, const TiledGeom& G, int64_t f, int lane) {
        t.q0 = ldg_stream16(w + f * 1024 + lane * 16);
        t.q1 = ldg_stream16(w + f * 1024 + 512 + lane * 16);
        t.d = ldg_stream4(w + G.off_s + f * 32 + (lane >> 2) * 4);
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
    static constexpr int kDepth = 5;
    static constexpr int kTileBytes = 768;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ void load(Tile& t, const uint8_t* w, const TiledGeom& G, int64_t f, int lane) {
        t.lo = ldg_stream16(w + f * 768 + lane * 16);
        t.hi = ldg_stream8(w + f * 768 + 512 + lane * 8);
        t.d = ldg_stream4(w + G.off_s + f * 32 + (lane >> 2) * 4);
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

// ------------------------------------------------------------------ the kernel
template <int QT, bool BF, int NS, int NT>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
gemv_kernel(const GemvArgs A) {
    using C = Codec<QT, BF>;
    using Tile = typename C::Tile;
    constexpr int D = C::kDepth;
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int64_t wid = (int64_t)blockIdx.x * GEMV_WARPS + (threadIdx.x >> 5);
    const int64_t T = A.G.ntiles, W = A.total_warps, nb = A.G.nb;
    const int64_t f0 = wid * T / W, f1 = (wid + 1) * T / W;

    // weights do not depend on the previous kernel: start streaming before the PDL wait
    Tile ring[D];
#pragma unroll
    for (int i = 0; i < D; i++)
        if (f0 + i < f1) C::load(ring[i], A.w, A.G, f0 + i, lane);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (f0 >= f1) return;

    float acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f; }
    int64_t cur_rt = f0 / nb;
    int64_t span_begin = f0;       // first tile of cur_rt handled by this warp

    auto flush = [&](int64_t rt, int64_t a, int64_t b) {
        const int64_t r_begin = rt * nb, r_end = r_begin + nb;
        float tot[NT][4];
        bool do_store = false;
        if (a == r_begin && b == r_end) {
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
            const int64_t r0 = rt * 16 + g, r1 = r0 + 8, N = A.G.rows;
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
    };

    uint32_t ones[4];
    ones[0] = ones[1] = ones[2] = ones[3] = BF ? 0x3f803f80u : 0x3c003c00u;

    for (int64_t fb = f0; fb < f1; fb += D) {
#pragma unroll
        for (int i = 0; i < D; i++) {
            const int64_t f = fb + i;
            if (f < f1) {
                const Tile cur = ring[i];
                if (f + D < f1) C::load(ring[i], A.w, A.G, f + D, lane);
                const int64_t rt = f / nb, kb = f - rt * nb;
                if (rt != cur_rt) {
                    flush(cur_rt, span_begin, f);
#pragma unroll
                    for (int j = 0; j < NT; j++) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f; }
                    cur_rt = rt; span_begin = f;
                }
                // x fragments of this k-block: lane (g,t) reads x[8j+g][kb*64 + 16t .. +15]
                uint4 xb[NT][NS][2];
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const int m = 8 * j + g;
#pragma unroll
                    for (int p = 0; p < NS; p++) {
                        if (m < A.M) {
                            const uint4* xp = reinterpret_cast<const uint4*>(
                                A.x + p * A.plane_stride + (int64_t)m * A.G.k + kb * 64 + t * 16);
                            xb[j][p][0] = xp[0]; xb[j][p][1] = xp[1];   // L1-cached: x is re-read per row tile
                        } else {
                            xb[j][p][0] = make_uint4(0, 0, 0, 0); xb[j][p][1] = make_uint4(0, 0, 0, 0);
                        }
                    }
                }
                float c[NT][4], cs[NT][4];
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.0f;
                    cs[j][0] = cs[j][1] = cs[j][2] = cs[j][3] = 0.0f;
                }
#pragma unroll
                for (int s = 0; s < 4; s++) {
                    uint32_t a[4];
                    C::frag(cur, s, a);
#pragma unroll
                    for (int j = 0; j < NT; j++)
#pragma unroll
                        for (int p = 0; p < NS; p++) {
                            const uint4 v = xb[j][p][s >> 1];
                            const uint32_t b0 = (s & 1) ? v.z : v.x, b1 = (s & 1) ? v.w : v.y;
                            mma16816<BF>(c[j], a, b0, b1);
                            if (C::kHasMin) mma16816<BF>(cs[j], ones, b0, b1);
                        }
                }
                const float2 d = __half22float2(u32_as_h2(cur.d));
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    acc[j][0] = fmaf(d.x, c[j][0], acc[j][0]); acc[j][1] = fmaf(d.x, c[j][1], acc[j][1]);
                    acc[j][2] = fmaf(d.y, c[j][2], acc[j][2]); acc[j][3] = fmaf(d.y, c[j][3], acc[j][3]);
                }
                if (C::kHasMin) {
                    const float2 mn = __half22float2(u32_as_h2(cur.m));
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        acc[j][0] = fmaf(mn.x, cs[j][0], acc[j][0]); acc[j][1] = fmaf(mn.x, cs[j][1], acc[j][1]);
                        acc[j][2] = fmaf(mn.y, cs[j][2], acc[j][2]); acc[j][3] = fmaf(mn.y, cs[j][3], acc[j][3]);
                    }
                }
            }
        }
    }
    flush(cur_rt, span_begin, f1);
}

template <int QT, bool BF, int NS>
static void launch_nt(int nt, int grid, const GemvArgs& A, cudaStream_t st) {
    switch (nt) {
        case 1: gemv_kernel<QT, BF, NS, 1><<<grid, GEMV_WARPS * 32, 0, st>>>(A); break;
        case 2: gemv_kernel<QT, BF, NS, 2><<<grid, GEMV_WARPS * 32, 0, st>>>(A); break;
        default: gemv_kernel<QT, BF, NS, 4><<<grid, GEMV_WARPS * 32, 0, st>>>(A); break;
    }
}
template <int QT>
static void launch_qt(int x_dtype, int nt, int grid, const GemvArgs& A, cudaStream_t st) {
    if (x_dtype == DT_F16) launch_nt<QT, false, 1>(nt, grid, A, st);
    else if (x_dtype == DT_BF16) launch_nt<QT, true, 1>(nt, grid, A, st);
    else launch_nt<QT, true, 3>(nt, grid, A, st);
}


// one translation unit per qtype instantiates this (keeps nvcc times parallel)
#define LOWBIT_DEFINE_GEMV_QT(NAME, QT)                                                          \
    void gemv_launch_##NAME(int x_dtype, int nt, int grid, const GemvArgs& A, cudaStream_t st) { \
        launch_qt<QT>(x_dtype, nt, grid, A, st);                                                \
    }

}  // namespace lowbit
