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
//  * Work split: one thread-block CLUSTER per 64-row group; the S CTAs of the cluster split K,
//    the 4 warps of a CTA split the CTA's K slice, each warp streams one contiguous run of
//    super-tiles.  Partial sums meet in shared memory (warps of a CTA) and then in the leader
//    CTA through distributed shared memory (ld.shared::cluster), always in the same order:
//    deterministic, no atomics, no global scratch, no fences.  (Measured: a global-memory
//    "last arriver" fix-up cost 3-5 us of serial L2 round trips per launch -- more than the
//    3.7 us the 24 MB of a gate_proj take at HBM speed.)  S is chosen so that groups x S fills
//    the ~4 CTA slots of every SM.
//  * griddepcontrol.wait sits between the first weight copies and the first read of x, so under
//    programmatic dependent launch the weight stream of call i+1 overlaps the tail of call i.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <cuda_fp8.h>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"
#include "gemv_args.h"

namespace lowbit {

// ------------------------------------------------------------------ small PTX wrappers
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#ifdef LOWBIT_GEMV_TIMESTAMPS     // development builds only: the checks cost issue slots in every warp
#define TS(i) do { if (A.ts && lane == 0) { A.ts[wid * 8 + (i)] = gtime(); if ((i) == 0) { unsigned sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); A.ts[wid * 8 + 6] = sm + 1; } } } while (0)
#else
#define TS(i) do { } while (0)
#endif
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
__device__ __forceinline__ bool elect_one() {       // one lane of the (converged) warp
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
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
// frag       : the A fragments (kPlanes x 4 x b32) of k-step s: code values, UNSCALED, offset NOT applied
// kPlanes    : fragments per k-step whose products add up to the weight (2 for nf4 x bf16: the table
//              value split into two bf16), kLutBytes: shared-memory table the codec reads through `lut`
//              at the lane's `lane_slot`
template <int QT, bool BF> struct Codec;

template <bool BF, int OFFSET> struct Codec4 {
    struct Tile { uint4 q; };
    static constexpr int kTileBytes = 512;
    static constexpr int kOffset = OFFSET;
    static constexpr float kLo = BF ? 1.0f : 16777216.0f;    // fp16 path: a0,a2 hold n * 2^-24
    static constexpr float kHi = BF ? 1.0f : 1048576.0f;     //            a1,a3 hold n * 2^-20
    static constexpr int kPlanes = 1, kLutBytes = 0;
    static constexpr bool kRowSplit = false;
    __device__ static __forceinline__ void fill_lut(uint8_t*, int) {}
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, int r, int lane) {
        t.q = lds16(codes + r * 512 + lane * 16);
    }
    __device__ static __forceinline__ uint32_t lane_slot(int) { return 0u; }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&ap)[1][4], uint32_t, uint32_t) {
        uint32_t (&a)[4] = ap[0];
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

// nf4: w = T[code] * d with the 16-entry NormalFloat table.  Byte b of the lane's word s is the
// code pair of fragment register a_b (tilednf4_pos), looked up in a 256-entry table of
// (T[lo], T[hi]) pairs: fp16x2 for fp16 activations (table values rounded to fp16: 2^-12 relative),
// two bf16x2 planes (value = hi + lo part, 2^-17) for bf16.  An entry is 128 bytes, one slot per lane, so that no
// two lanes of a load ever meet in a bank (8 slots per entry cost 3.1 wavefronts per LDS, and the kernel was bound
// by them): fp16 = 32 slots of 4 bytes; bf16 = 16 slots of {hi, lo} read with one LDS.64, which the hardware
// serves half-warp by half-warp -- lanes l and l + 16 share a slot and never collide.  32 KB per CTA.
template <bool BF> struct Codec<QT_NF4, BF> {
    struct Tile { uint4 q; };
    static constexpr int kTileBytes = 512;
    static constexpr int kOffset = 0;
    static constexpr float kLo = 1.0f, kHi = 1.0f;
    static constexpr int kPlanes = BF ? 2 : 1, kLutBytes = 256 * 128;
    static constexpr bool kRowSplit = BF;      // see frag
    __device__ static __forceinline__ uint32_t entry(int e, int plane) {
        const float lo = kNF4Table[e & 15], hi = kNF4Table[e >> 4];
        if (!BF) return h2_as_u32(__floats2half2_rn(lo, hi));
        const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
        if (plane == 0) return bf2_as_u32(h);
        const float2 hf = __bfloat1622float2(h);
        return bf2_as_u32(__floats2bfloat162_rn(lo - hf.x, hi - hf.y));
    }
    __device__ static __forceinline__ void fill_lut(uint8_t* lut, int tid) {   // 128 threads, 256 entries
        for (int e = tid; e < 256; e += GEMV_WARPS * 32) {
            const uint32_t v0 = entry(e, 0), v1 = BF ? entry(e, 1) : v0;
            uint4* dst = reinterpret_cast<uint4*>(lut + e * 128);
#pragma unroll
            for (int c = 0; c < 8; c++) dst[c] = make_uint4(v0, v1, v0, v1);
        }
    }
    __device__ static __forceinline__ uint32_t lane_slot(int lane) { return BF ? 8u * (lane & 15) : 4u * lane; }
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, int r, int lane) {
        t.q = lds16(codes + r * 512 + lane * 16);
    }
    // lut: the table (warp-uniform); slot = lane_slot(lane)
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&a)[kPlanes][4], uint32_t lut, uint32_t slot) {
        const uint32_t w = s == 0 ? t.q.x : s == 1 ? t.q.y : s == 2 ? t.q.z : t.q.w;
        // byte b -> entry address: one PRMT (extract) and one IMAD (x 128 + the lane's slot in the table)
        const uint32_t base = lut + slot;
        const uint32_t i0 = __byte_perm(w, 0u, 0x4440u) * 128u, i1 = __byte_perm(w, 0u, 0x4441u) * 128u;
        const uint32_t i2 = __byte_perm(w, 0u, 0x4442u) * 128u, i3 = __byte_perm(w, 0u, 0x4443u) * 128u;
        if (BF) {
            const uint2 v0 = lds8(base + i0), v1 = lds8(base + i1), v2 = lds8(base + i2), v3 = lds8(base + i3);
            // an LDS.64 leaves {hi, lo} in a register pair and the MMA wants its A quad in four consecutive registers:
            // fragment 0 = {hi, lo} of bytes 0 and 2 (weight row g), fragment 1 = those of bytes 1 and 3 (row g + 8).
            // The MMA's "rows 8..15" are then the lo parts of the same weight row (kRowSplit): the kernel keeps one
            // accumulator quad per fragment and adds c0 + c2, c1 + c3 afterwards.  No register moves.
            a[0][0] = v0.x; a[0][1] = v0.y; a[0][2] = v2.x; a[0][3] = v2.y;
            a[kPlanes - 1][0] = v1.x; a[kPlanes - 1][1] = v1.y; a[kPlanes - 1][2] = v3.x; a[kPlanes - 1][3] = v3.y;
        } else {
            a[0][0] = lds4(base + i0); a[0][1] = lds4(base + i1); a[0][2] = lds4(base + i2); a[0][3] = lds4(base + i3);
        }
    }
};
template <bool BF> struct Codec<QT_ASYM_INT4, BF> : Codec4<BF, 2> {};

template <int QT, bool BF> struct Codec8 {
    struct Tile { uint4 q0, q1; };
    static constexpr int kTileBytes = 1024;
    static constexpr int kOffset = 0;
    static constexpr float kLo = 1.0f, kHi = 1.0f;
    static constexpr int kPlanes = 1, kLutBytes = 0;
    static constexpr bool kRowSplit = false;
    __device__ static __forceinline__ void fill_lut(uint8_t*, int) {}
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
    __device__ static __forceinline__ uint32_t lane_slot(int) { return 0u; }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&ap)[1][4], uint32_t, uint32_t) {
        uint32_t (&a)[4] = ap[0];
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
    static constexpr int kPlanes = 1, kLutBytes = 0;
    static constexpr bool kRowSplit = false;
    __device__ static __forceinline__ void fill_lut(uint8_t*, int) {}
    __device__ static __forceinline__ void load(Tile& t, uint32_t codes, int r, int lane) {
        t.lo = lds16(codes + r * 768 + lane * 16);
        t.hi = lds8(codes + r * 768 + 512 + lane * 8);
    }
    __device__ static __forceinline__ uint32_t cvt2(unsigned short two) {
        uint32_t h;
        asm("cvt.rn.f16x2.e3m2x2 %0, %1;" : "=r"(h) : "h"(two));
        return BF ? h2_to_bf2(h) : h;
    }
    __device__ static __forceinline__ uint32_t lane_slot(int) { return 0u; }
    __device__ static __forceinline__ void frag(const Tile& t, int s, uint32_t (&ap)[1][4], uint32_t, uint32_t) {
        uint32_t (&a)[4] = ap[0];
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
#ifndef LOWBIT_GEMV_CH
    static constexpr int kChunk = kSuperCodes <= 2048 ? 2 : 1;                    // ~4 KB of codes per bulk copy
#else
    static constexpr int kChunk = LOWBIT_GEMV_CH;
#endif
#ifndef LOWBIT_GEMV_STAGES
    static constexpr int kStages = 2;
#else
    static constexpr int kStages = LOWBIT_GEMV_STAGES;
#endif
    static constexpr int kMinBytes = (C::kOffset == 2) ? 128 : 0;
    static constexpr int kStageBytes = kChunk * (kSuperCodes + 128 + kMinBytes);
    static constexpr int kRingBytes = kStages * kStageBytes;
    static constexpr int kBarBytes = GEMV_WARPS * kStages * 8 + 8;                // + the leader's receive barrier
    // bytes of one matrix' accumulators of one warp in the reduction slot (M == 1: 8 live lanes per value)
    // live lanes: only the lanes that hold existing activation rows take part in the reduction: the
    // t-lanes 0 .. tl-1 with tl = ceil(M / 2) for M <= 8 (M == 1: tl = 1), all four beyond
    __host__ __device__ static constexpr int live_t(int nt, int acc, int m) {
        return acc == 2 ? 1 : (nt > 1 ? 4 : ((m < 8 ? m : 8) + 1) / 2);
    }
    __host__ __device__ static constexpr int slot_bytes(int nt, int acc, int tl = 4) { return RG_TILES * nt * acc * 32 * (acc == 2 ? 1 : tl); }
    // a warp's region: its ring, which later holds its reduction slot
    __host__ __device__ static constexpr int warp_bytes(int nt, int acc, int nm) {
        // M = 2..8 kernels: the four rings together also receive the peers' sums (RINGRECV in the kernel)
        const int need = (nt <= 2 && acc == 4) ? (GEMV_MAX_SPLIT - 1) * nm * slot_bytes(nt, acc) / GEMV_WARPS : nm * slot_bytes(nt, acc);
        return ((kRingBytes > need ? kRingBytes : need) + 127) / 128 * 128;
    }
    // + the leader CTA's receive buffer for its peers' CTA sums (push reduction, M <= 8 kernels)
    static constexpr int smem_bytes(int nt, int acc, int nm, int tl = 4, bool ring = false) {
        return lut_offset(nt, acc, nm, tl, ring) + C::kLutBytes;
    }
    // the codec's lookup table (if any) sits last, 32-byte aligned
    // ring: the M = 2..8 kernel receives into its rings (no dedicated buffer)
    __host__ __device__ static constexpr int lut_offset(int nt, int acc, int nm, int tl = 4, bool ring = false) {
        return (GEMV_WARPS * warp_bytes(nt, acc, nm) + kBarBytes + 8 + ((nt == 1 && !(acc == 4 && ring)) ? (GEMV_MAX_SPLIT - 1) * nm * slot_bytes(nt, acc, tl) : 0) + 31) / 32 * 32;
    }
};

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t local_addr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    return remote;
}
// asynchronous remote store that counts its 4 bytes on the destination CTA's mbarrier: no fences, and
// the sending CTA may exit as soon as it has issued its stores
__device__ __forceinline__ void st_async_f32(uint32_t remote_addr, float v, uint32_t remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
                 ::"r"(remote_addr), "r"(__float_as_uint(v)), "r"(remote_bar) : "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t rank) {
    uint32_t remote; float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote));
    return v;
}

// ------------------------------------------------------------------ the kernel
// ACC = accumulators per (row tile, n-tile): 4 (c0..c3), or 2 when M == 1 (only column 0 is live:
// c0 = row g, c2 = row g+8).
// grid = (S, groups): blockIdx.y = 64-row group, blockIdx.x = K slice = rank in the cluster (S,1,1).
// NM = 2 is the fused MLP (reference fuse_mlp_base, models/common.py:146-161: xe_linear.mlp_forward_xpu):
// the same rows of a second matrix (A.w2, "up") are streamed after the first ("gate") and the epilogue
// stores act(gate . x) * (up . x).
// MODE 1 / 2 are the two launches of the mixture-of-experts decode (see GemvArgs): 1 = one routed
// expert per blockIdx.z (used with NM = 2 for gate/up), 2 = the routed experts' down projections
// concatenated along K, each slot's sums weighted by its routing weight.
// resident CTAs per SM the register budget allows: 6 (80 registers) for one token; the 2..8-token kernels carry twice the
// accumulators and spilled at 80 -- 5 (102 registers) measured 7-9 % faster at M = 4 / 8 than 6, and than 4
#ifndef LOWBIT_GEMV_ACC4_CTAS
#define LOWBIT_GEMV_ACC4_CTAS 5
#endif
template <int QT, bool BF, int NS, int NT, int ACC, int NM, int MODE>
__global__ void __launch_bounds__(GEMV_WARPS * 32, (NT == 1 && NS == 1) ? (Codec<QT, BF>::kLutBytes ? 4 : (ACC == 4 ? LOWBIT_GEMV_ACC4_CTAS : 6)) : 1)
gemv_kernel(const GemvArgs A) {
    using C = Codec<QT, BF>;
    using S = Staging<QT, BF>;
    using Tile = typename C::Tile;
    constexpr int CH = S::kChunk, STAGES = S::kStages, STAGE_BYTES = S::kStageBytes;
    constexpr int SUPER = S::kSuperCodes;
    constexpr int NV = RG_TILES * NT * ACC;            // accumulator floats per lane and matrix
    constexpr int WB = S::warp_bytes(NT, ACC, NM);     // bytes of one warp's ring / reduction slot
    extern __shared__ __align__(128) uint8_t smem[];

    // the warp index through a shuffle: the compiler then knows it is warp-uniform and keeps everything
    // derived from it (ring addresses, copy sources) in uniform registers
    const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int g = lane >> 2, t = lane & 3;
    const int nsplit = gridDim.x, slice = blockIdx.x, rg = blockIdx.y;
    const int nb = (int)A.G.nb;
    const int nbt = (MODE == 2) ? nb * A.n_seg : nb;   // k-blocks along the (concatenated) K axis
    if (A.pdl) asm volatile("griddepcontrol.launch_dependents;");
    // the routed expert ids come from the previous kernel: nothing can be fetched before it is done
    if (MODE != 0 && A.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");

    auto expert_of = [&](int z, bool& ok) -> int {
        const long long e = A.sel_i64 ? reinterpret_cast<const long long*>(A.sel)[z]
                                      : (long long)reinterpret_cast<const int*>(A.sel)[z];
        ok = e >= 0 && e < A.n_experts;
        return ok ? (int)e : 0;
    };
    const uint8_t* wb0 = A.w;
    const uint8_t* wb1 = A.w2;
    float out_scale = 1.0f;
    if (MODE == 1) {
        bool ok;
        const int e = expert_of(blockIdx.z, ok);
        wb0 = reinterpret_cast<const uint8_t*>(A.tab_w[e]);
        wb1 = reinterpret_cast<const uint8_t*>(A.tab_w2[e]);
        out_scale = ok ? 1.0f : 0.0f;
    }

    // this CTA's k-blocks [ck0, ck1), this warp's [kb0, kb1)
    // (the host passes nbt / nsplit and nbt % nsplit: the first `rem` slices take one block more)
    const int ck0 = slice * A.slice_q + min(slice, A.slice_r);
    const int ck1 = ck0 + A.slice_q + (slice < A.slice_r ? 1 : 0);
    const int kb0 = ck0 + (int)((unsigned)warp * (unsigned)(ck1 - ck0) / (unsigned)GEMV_WARPS);
    const int kb1 = ck0 + (int)((unsigned)(warp + 1) * (unsigned)(ck1 - ck0) / (unsigned)GEMV_WARPS);
    const int n_super = kb1 - kb0;
    // the warp's work is a list of segments: one per matrix (NM) and, in MODE 2, per routed slot its
    // range touches; a segment is `n` consecutive super-tiles starting at k-block `a` of matrix `base`
    const int z_first = (MODE == 2) ? kb0 / nb : 0;
    const int z_last = (MODE == 2) ? (n_super > 0 ? (kb1 - 1) / nb : z_first - 1) : 0;
    auto seg_load = [&](int mt, int z, int& a, int& n, const uint8_t*& base) {
        if (MODE == 2) {
            const int lo = max(kb0, z * nb), hi = min(kb1, (z + 1) * nb);
            bool ok;
            a = lo - z * nb; n = hi - lo;
            base = reinterpret_cast<const uint8_t*>(A.tab_w[expert_of(z, ok)]);
        } else {
            a = kb0; n = n_super; base = mt ? wb1 : wb0;
        }
    };
    const int64_t wid = ((int64_t)rg * nsplit + slice) * GEMV_WARPS + warp;   // timestamps only
    TS(0);

    // this warp's private ring: [STAGES][codes | scales | mins]; the mbarriers sit after all rings
    const uint32_t ring = smem_u32(smem) + warp * WB;
    const uint32_t bars = smem_u32(smem) + GEMV_WARPS * WB + warp * (STAGES * 8);
    // the codec's lookup table (nf4): built once per CTA, read-only afterwards
    const int tl = S::live_t(NT, ACC, A.M);            // live t-lanes of the reduction (see Staging)
    const uint32_t lut = smem_u32(smem) + S::lut_offset(NT, ACC, NM, tl, A.ring_recv != 0);      // warp-uniform
    const uint32_t lut_slot = C::lane_slot(lane);
    if (C::kLutBytes) {
        C::fill_lut(smem + S::lut_offset(NT, ACC, NM, tl, A.ring_recv != 0), threadIdx.x);
        __syncthreads();
    }
    // issue cursor: walks the segment list one chunk at a time.  Warp-uniform state, called by the whole
    // warp; one elected lane issues the copies.
    int i_mt = 0, i_z = z_first, i_left = 0;
    uint32_t i_st = 0;                                 // stage the next chunk goes to
    const uint8_t* i_q = nullptr;                      // next chunk's codes; scales / mins follow at fixed offsets
    const uint8_t* i_s = nullptr;
    bool i_live = n_super > 0;
    auto seg_open = [&](int mt, int z) {
        int a;
        const uint8_t* base;
        seg_load(mt, z, a, i_left, base);
        const int64_t u = (int64_t)rg * nb + a;
        i_q = base + u * SUPER;
        i_s = base + A.G.off_s + u * 128;
    };
    if (i_live) seg_open(0, z_first);
    const int64_t s_to_m = A.G.off_m - A.G.off_s;
    auto issue_next = [&]() {                          // copy the cursor's chunk into stage i_c % STAGES, advance
        if (!i_live) return;
        const int ns = min(CH, i_left);
        const uint32_t dst = ring + i_st * STAGE_BYTES, bar = bars + i_st * 8;
        if (elect_one()) {
            mbar_expect_tx(bar, ns * (SUPER + 128 + S::kMinBytes));
            bulk_g2s(dst, i_q, ns * SUPER, bar);
            bulk_g2s(dst + CH * SUPER, i_s, ns * 128, bar);
            if (S::kMinBytes) bulk_g2s(dst + CH * (SUPER + 128), i_s + s_to_m, ns * 128, bar);
        }
        i_st = (i_st + 1 == STAGES) ? 0 : i_st + 1;
        i_left -= CH;
        i_q += CH * SUPER; i_s += CH * 128;
        if (i_left <= 0) {
            if (NM == 1 && MODE != 2) i_live = false;
            else {
                if (i_z < z_last) i_z++; else { i_z = z_first; i_mt++; }
                if (i_mt == NM) i_live = false; else seg_open(i_mt, i_z);
            }
        }
    };

    constexpr bool PUSH = (NT <= 2);                   // peers push their sums into the leader (else the leader pulls)
    // M == 1: the leader has a small dedicated receive buffer, and the cluster barrier that publishes its
    // mbarrier is split around the main loop (arrive here, wait before the tail: free).  M = 2..8: up to
    // 28 KB would cost a resident CTA per SM, so the peers push into the leader's idle rings instead, after
    // a cluster barrier that follows the leader's own CTA sums (RINGRECV).  That late barrier costs about
    // a microsecond, so the host asks for it (A.ring_recv) only when it buys a larger cluster.
    const bool RINGRECV = PUSH && (ACC == 4) && (NT == 2 || A.ring_recv != 0);       // M = 9..16: always into the rings
    static_assert(!(PUSH && ACC == 4) || (GEMV_MAX_SPLIT - 1) * NM * S::slot_bytes(NT, ACC, 4) <= GEMV_WARPS * WB, "receive area must fit in the rings");
    const uint32_t rbar = smem_u32(smem) + GEMV_WARPS * WB + GEMV_WARPS * STAGES * 8;
    float* recv = reinterpret_cast<float*>(smem + GEMV_WARPS * WB + S::kBarBytes + 8);
    if (PUSH && nsplit > 1) {
        if (slice == 0 && threadIdx.x == 0) {
            mbar_init(rbar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        if (!RINGRECV) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");   // matched by the wait before the tail
    }
    // lane s owns stage s during start-up (the barrier inits and first copies go out in parallel);
    // weights do not depend on the previous kernel: they start streaming before the PDL wait
    if (lane < STAGES) {
        mbar_init(bars + lane * 8, 1);
        // make the init visible to the async proxy (the TMA engine completes bytes on this barrier);
        // the cluster-scoped fence.mbarrier_init costs ~1 us here (it also invalidates L1)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
#pragma unroll
    for (int s = 0; s < STAGES; s++) issue_next();
    TS(1);
    // HBM is idle during a kernel's start-up, its reduction tail and the launch gap: about a third of a
    // 24 MB GEMV.  If the host told us which weights the NEXT matmul on this stream reads, every warp asks
    // the L2 (126 MB) to fetch its share of them now, behind our own demand loads.
    if (A.next_w && lane == 0) {
        const unsigned long long nw = (unsigned long long)gridDim.x * gridDim.y * GEMV_WARPS;
        const unsigned long long me = ((unsigned long long)blockIdx.y * gridDim.x + blockIdx.x) * GEMV_WARPS + warp;
        unsigned long long share = ((A.next_bytes + nw - 1) / nw + 127ull) & ~127ull;
        const unsigned long long off = me * share;
        if (off < A.next_bytes) {
            if (off + share > A.next_bytes) share = (A.next_bytes - off) & ~15ull;
            if (share)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(A.next_w + off), "r"((uint32_t)share) : "memory");
        }
    }

    float accm[NM][RG_TILES][NT][ACC];
#pragma unroll
    for (int mt = 0; mt < NM; mt++)
#pragma unroll
        for (int r = 0; r < RG_TILES; r++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int c = 0; c < ACC; c++) accm[mt][r][j][c] = 0.0f;

    // lane (g,t) reads x[8j+g][kb*64 + 16t .. +15] for n-tile j (plane p); rows >= M read as zero
    int64_t xptr[NT][NS];
    // per-lane address of the fragment of n-tile 0 / plane 0 that is fetched next (advances 128 B per k-block)
    const char* const xbase = reinterpret_cast<const char*>(A.x + (int64_t)g * A.G.k + t * 16) + (size_t)kb0 * 128;
    const char* xrun = xbase;
    bool xvalid[NT];
    constexpr bool XPF = (NT * NS <= 4);
    constexpr int XPF_BUFS = XPF ? 2 : 1;
    // two fragment buffers used alternately by even / odd super-tiles (no copies between them, so the
    // compiler has no reason to delay the prefetch); lanes without a valid row keep zeros
    uint4 xq[XPF_BUFS][NT][NS][2];
#pragma unroll
    for (int j = 0; j < NT; j++) {
        xvalid[j] = (8 * j + g) < A.M;
#pragma unroll
        for (int p = 0; p < NS; p++) {
            xptr[j][p] = (int64_t)2 * (p * A.plane_stride + (int64_t)(8 * j) * A.G.k);   // warp-uniform byte offset
#pragma unroll
            for (int h = 0; h < XPF_BUFS; h++) { xq[h][j][p][0] = make_uint4(0, 0, 0, 0); xq[h][j][p][1] = make_uint4(0, 0, 0, 0); }
        }
    }
    // x comes from L2 (each warp reads its own K slice; L1 is cold): the fragment of the NEXT k-block is
    // requested one super-tile ahead so that the ~0.4 us round trip hides behind the current one
    int xleft = 0;
    auto fetch_x = [&](uint4 (&dst)[NT][NS][2]) {      // fetch the k-block at xrun, advance
#pragma unroll
        for (int j = 0; j < NT; j++)
#pragma unroll
            for (int p = 0; p < NS; p++)
                if (xvalid[j]) {                        // L1/L2-cached: x is shared by every row group
                    dst[j][p][0] = ldg_cached16(xrun + xptr[j][p]);
                    dst[j][p][1] = ldg_cached16(xrun + xptr[j][p] + 16);
                }
        xrun += 128;
    };
    // everything above is independent of the previous kernel's output: the wait sits right before the first
    // read of x, so the set-up instructions of all resident warps run while the SM would be idle anyway
    if (MODE == 0 && A.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    TS(2);
    if (XPF && n_super > 0) fetch_x(xq[0]);

    uint32_t ones[4];
    ones[0] = ones[1] = ones[2] = ones[3] = BF ? 0x3f803f80u : 0x3c003c00u;

    float tot[MODE == 2 ? RG_TILES : 1][NT][ACC];      // MODE 2: sum over finished slots of scale * acc
#pragma unroll
    for (int r = 0; r < (MODE == 2 ? RG_TILES : 1); r++)
#pragma unroll
        for (int j = 0; j < NT; j++)
#pragma unroll
            for (int q = 0; q < ACC; q++) tot[r][j][q] = 0.0f;

    uint32_t c_st = 0, c_ph = 0;                       // stage and mbarrier phase of the chunk being consumed
    bool first_chunk = true;
#pragma unroll
    for (int mt = 0; mt < NM; mt++) {
    float (&acc)[RG_TILES][NT][ACC] = accm[mt];
    xleft = n_super;
    if (!XPF) xrun = xbase;
    for (int z = z_first; z <= z_last; z++) {
    int s_a, s_n;
    const uint8_t* s_base;
    seg_load(mt, z, s_a, s_n, s_base);
    uint32_t stage = 0;
    for (int sidx = 0; sidx < s_n; sidx += 2) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (h == 0 || sidx + 1 < s_n) {
                const int i = (CH == 2) ? h : 0;                  // super-tile index inside its stage
                if (i == 0) {
                    stage = ring + c_st * STAGE_BYTES;
                    mbar_wait(bars + c_st * 8, c_ph);
                    if (first_chunk) { TS(3); first_chunk = false; }
                }
                const bool last_of_stage = (i == CH - 1) || (sidx + h == s_n - 1);
                const uint32_t codes = stage + i * SUPER;
                uint4 (&xb)[NT][NS][2] = xq[XPF ? h : 0];
                // ---- once per super-tile: x fragment, sum(x), scales of the lane's 8 rows
                if (XPF) {
                    xleft--;
                    if (xleft > 0) fetch_x(xq[XPF ? (h ^ 1) : 0]);
                    else if (mt + 1 < NM) {            // the second matrix starts over at the slice's first block
                        xrun = xbase;
                        fetch_x(xq[XPF ? (h ^ 1) : 0]);
                    }
                } else {
                    fetch_x(xb);
                }
                // -8 * sum over the k-block of x (sym) / sum (asym) for columns 2t, 2t+1: a ones-row mma,
                // two independent chains (the tensor pipe is ~13 % busy in this kernel; issue slots are scarce)
                float xs[NT][2];
                if (C::kOffset) {
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        float s1[4] = {0.0f, 0.0f, 0.0f, 0.0f}, s2[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                        for (int p = 0; p < NS; p++) {
                            mma16816<BF>(s1, ones, xb[j][p][0].x, xb[j][p][0].y);
                            mma16816<BF>(s2, ones, xb[j][p][0].z, xb[j][p][0].w);
                            mma16816<BF>(s1, ones, xb[j][p][1].x, xb[j][p][1].y);
                            mma16816<BF>(s2, ones, xb[j][p][1].z, xb[j][p][1].w);
                        }
                        const float k = (C::kOffset == 1) ? -8.0f : 1.0f;
                        xs[j][0] = k * (s1[0] + s2[0]); xs[j][1] = k * (s1[1] + s2[1]);
                    }
                }
                const uint4 dq = lds16(stage + CH * SUPER + i * 128 + g * 16);       // [r][h] fp16 d
                uint4 mq = make_uint4(0, 0, 0, 0);
                if (C::kOffset == 2) mq = lds16(stage + CH * (SUPER + 128) + i * 128 + g * 16);
                // the four row tiles advance in lock step so that four independent HMMA chains are in flight
                Tile tile[RG_TILES];
                float cc[RG_TILES][NT][4];
#pragma unroll
                for (int r = 0; r < RG_TILES; r++) {
                    C::load(tile[r], codes, r, lane);
#pragma unroll
                    for (int j = 0; j < NT; j++) cc[r][j][0] = cc[r][j][1] = cc[r][j][2] = cc[r][j][3] = 0.0f;
                }
                // row-split codecs (nf4 x bf16): fragment 1 = weight row g + 8, accumulated apart and folded in below
                constexpr bool RS = C::kRowSplit;
                float cc2[RS ? RG_TILES : 1][NT][4];
                if (RS) {
#pragma unroll
                    for (int r = 0; r < RG_TILES; r++)
#pragma unroll
                        for (int j = 0; j < NT; j++) cc2[r][j][0] = cc2[r][j][1] = cc2[r][j][2] = cc2[r][j][3] = 0.0f;
                }
#pragma unroll
                for (int s = 0; s < 4; s++) {
#pragma unroll
                    for (int r = 0; r < RG_TILES; r++) {
#pragma unroll
                        uint32_t af[C::kPlanes][4];
                        C::frag(tile[r], s, af, lut, lut_slot);
#pragma unroll
                        for (int pw = 0; pw < C::kPlanes; pw++) {
                            uint32_t (&a)[4] = af[pw];
#pragma unroll
                            for (int j = 0; j < NT; j++)
#pragma unroll
                                for (int p = 0; p < NS; p++) {
                                    const uint4 v = xb[j][p][s >> 1];
                                    mma16816<BF>((RS && pw == 1) ? cc2[RS ? r : 0][j] : cc[r][j], a, (s & 1) ? v.z : v.x, (s & 1) ? v.w : v.y);
                                }
                        }
                    }
                }
                if (RS) {
#pragma unroll
                    for (int r = 0; r < RG_TILES; r++)
#pragma unroll
                        for (int j = 0; j < NT; j++) {
                            const float g0 = cc[r][j][0] + cc[r][j][2], g1 = cc[r][j][1] + cc[r][j][3];
                            const float h0 = cc2[r][j][0] + cc2[r][j][2], h1 = cc2[r][j][1] + cc2[r][j][3];
                            cc[r][j][0] = g0; cc[r][j][1] = g1; cc[r][j][2] = h0; cc[r][j][3] = h1;
                        }
                }
#pragma unroll
                for (int r = 0; r < RG_TILES; r++) {
                    const uint32_t dr = r == 0 ? dq.x : r == 1 ? dq.y : r == 2 ? dq.z : dq.w;
                    const float2 d = __half22float2(u32_as_h2(dr));
                    float2 mn = make_float2(0.0f, 0.0f);
                    if (C::kOffset == 2) {
                        const uint32_t mr = r == 0 ? mq.x : r == 1 ? mq.y : r == 2 ? mq.z : mq.w;
                        mn = __half22float2(u32_as_h2(mr));
                    }
#pragma unroll
                    for (int j = 0; j < NT; j++) {
#pragma unroll
                        for (int c = 0; c < ACC; c++) {
                            const int ci = (ACC == 2) ? 2 * c : c;             // cc register feeding accumulator c
                            const bool hi = (ACC == 2) ? (c == 1) : (c >= 2);   // row g+8
                            const float dv = hi ? d.y : d.x;
                            const float col = xs[j][(ACC == 2) ? 0 : (c & 1)];
                            if (C::kOffset == 1)        // d * (code_sum * 2^s - 8 * sum(x))
                                acc[r][j][c] = fmaf(dv, fmaf(cc[r][j][ci], hi ? C::kHi : C::kLo, col), acc[r][j][c]);
                            else if (C::kOffset == 2)   // d * code_sum * 2^s + m * sum(x)
                                acc[r][j][c] = fmaf(hi ? mn.y : mn.x, col, fmaf(dv * (hi ? C::kHi : C::kLo), cc[r][j][ci], acc[r][j][c]));
                            else
                                acc[r][j][c] = fmaf(dv, cc[r][j][ci], acc[r][j][c]);
                        }
                    }
                }
                if (last_of_stage) {
                    // every lane has consumed the stage (its data sits in registers already used): refill it
                    __syncwarp();
                    issue_next();
                    if (++c_st == STAGES) { c_st = 0; c_ph ^= 1u; }
                }
            }
        }
    }
    if (XPF && (s_n & 1)) {                            // odd segment: the next one starts on buffer 0 again
#pragma unroll
        for (int j = 0; j < NT; j++)
#pragma unroll
            for (int p = 0; p < NS; p++) { xq[0][j][p][0] = xq[XPF_BUFS - 1][j][p][0]; xq[0][j][p][1] = xq[XPF_BUFS - 1][j][p][1]; }
    }
    if (MODE == 2) {                                   // slot z is done: fold it in with its routing weight
        bool ok;
        expert_of(z, ok);
        float sc = 0.0f;
        if (ok) {
            if (A.seg_scale_dtype == DT_F32) sc = reinterpret_cast<const float*>(A.seg_scale)[z];
            else if (A.seg_scale_dtype == DT_F16) sc = __half2float(reinterpret_cast<const __half*>(A.seg_scale)[z]);
            else sc = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(A.seg_scale)[z]);
        }
#pragma unroll
        for (int r = 0; r < (MODE == 2 ? RG_TILES : 0); r++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int q = 0; q < ACC; q++) { tot[r][j][q] = fmaf(sc, acc[r][j][q], tot[r][j][q]); acc[r][j][q] = 0.0f; }
    }
    }
    }
    if (MODE == 2) {
#pragma unroll
        for (int r = 0; r < (MODE == 2 ? RG_TILES : 0); r++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int q = 0; q < ACC; q++) accm[0][r][j][q] = tot[r][j][q];
    }
    TS(4);

    // ---- reduction.  Slot layout: value v = (r*NT + j)*ACC + c of live lane l at float index v*LW + l.
    // Only the t-lanes below tl hold existing activation rows (M == 1: tl = 1, LW = 8 lanes per value).
    const int LW = 8 * tl;
    const int NVL = NV * LW;                           // floats per matrix and warp
    // 1) every warp parks its fragment in its own (now idle) ring
    float* slot = reinterpret_cast<float*>(smem + warp * WB);
    if (t < tl) {
        const int l = g * tl + t;
#pragma unroll
        for (int mt = 0; mt < NM; mt++)
#pragma unroll
            for (int r = 0; r < RG_TILES; r++)
#pragma unroll
                for (int j = 0; j < NT; j++)
#pragma unroll
                    for (int c = 0; c < ACC; c++) slot[mt * NVL + ((r * NT + j) * ACC + c) * LW + l] = accm[mt][r][j][c];
    }
    __syncthreads();
    const int64_t N = A.G.rows;
    // call number of the communicator: its counter is advanced by the receive kernel of the previous call, which this
    // grid has waited for (stream order, or griddepcontrol.wait above)
    unsigned int push_epoch = 0, push_dead = 0;
    if (MODE == 0 && A.push_world > 0) {
        asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(push_epoch) : "l"(A.push_epochs) : "memory");
        asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(push_dead) : "l"(A.push_epochs + 64) : "memory");
        push_epoch += 1;
        if (push_epoch == 0u) push_epoch = 2u;         // 0 = "never written": skipped when the counter wraps (comm.cu next_epoch)
    }
    auto store_value = [&](int i, float v) {
        const int l = i % LW, val = i / LW, c = val % ACC, j = (val / ACC) % NT, r = val / (ACC * NT);
        const int gg = l / tl, tt = l - gg * tl;
        const int hi = (ACC == 2) ? c : (c >> 1);                     // row g (0) or g+8 (1)
        const int col = 8 * j + 2 * tt + ((ACC == 2) ? 0 : (c & 1));
        const int64_t row = (int64_t)rg * RG_ROWS + r * 16 + gg + 8 * hi;
        if (col < A.M) {
            if (MODE == 0 && A.push_world > 0) {
                // tensor parallel: word (col, row) of this rank's partial result goes to every peer's inbox ...
                const unsigned long long off = (unsigned long long)(push_epoch & 1u) * A.push_slot_bytes + (unsigned long long)((int64_t)col * N + row) * 8ull;
                bool bad = push_dead != 0;             // an earlier wait on this communicator ran out: NaN, no waiting
#pragma unroll
                for (int pr = 0; pr < 8; pr++)
                    if (!bad && pr < A.push_world && pr != A.push_rank)
                        asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" :: "l"(A.push_inbox[pr] + off), "r"(__float_as_uint(v)), "r"(push_epoch) : "memory");
                // ... and the peers' words of the same element arrive in the own buffer: poll in rounds (every round loads
                // all words still missing at once -- one memory round trip whatever the world size), bounded
                float got[8];
                unsigned int missing = 0;
#pragma unroll
                for (int pr = 0; pr < 8; pr++) {
                    got[pr] = 0.0f;
                    if (!bad && pr < A.push_world && pr != A.push_rank) missing |= 1u << pr;
                }
                for (unsigned int polls = 0; missing != 0; polls++) {
                    if (polls > A.push_spin_limit) { bad = true; break; }
                    if (polls > 16) __nanosleep(polls > 1024 ? 256 : 32);
                    unsigned int tag[8];
#pragma unroll
                    for (int pr = 0; pr < 8; pr++)
                        if (missing & (1u << pr)) {
                            unsigned int data;
                            asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(data), "=r"(tag[pr])
                                         : "l"(A.push_own + (unsigned long long)pr * A.push_inbox_bytes + off) : "memory");
                            got[pr] = __uint_as_float(data);
                        }
#pragma unroll
                    for (int pr = 0; pr < 8; pr++)
                        if ((missing & (1u << pr)) && tag[pr] == push_epoch) missing &= ~(1u << pr);
                }
                float sum = 0.0f;                      // rank order: the same additions, the same bits, on every rank
#pragma unroll
                for (int pr = 0; pr < 8; pr++)
                    if (pr < A.push_world) sum += (pr == A.push_rank) ? v : got[pr];
                if (bad) {
                    sum = __uint_as_float(0x7fc00000u);
                    if (!push_dead) atomicExch(A.push_epochs + 64, 1u);
                }
                store_out(A.y, A.y_dtype, (int64_t)col * N + row, sum, A.bias, row, A.accumulate);
            } else
                store_out(A.y, A.y_dtype, (MODE == 1 ? (int64_t)blockIdx.z * N : (int64_t)col * N) + row, v, A.bias, row, A.accumulate);
        }
    };
    // tensor parallel: the last leader CTA to finish advances the communicator's call counter (every leader has read it
    // by then; the next call reads it after griddepcontrol.wait / stream order on this grid)
    auto comm_done = [&]() {
        if (MODE == 0 && A.push_world > 0 && slice == 0) {
            volatile int* s_last = reinterpret_cast<volatile int*>(smem);   // the rings are idle by now
            __syncthreads();
            if (threadIdx.x == 0) {
                __threadfence();
                *s_last = atomicAdd(A.push_epochs + 65, 1u) == gridDim.y - 1;
            }
            __syncthreads();
            if (*s_last) {
                if (threadIdx.x < 64) A.push_epochs[threadIdx.x] = push_epoch;
                if (threadIdx.x == 0) A.push_epochs[65] = 0u;
            }
        }
    };
    auto finish = [&](int i, const float (&vm)[NM]) {  // both matrices' sums of output i are in one thread
        const float v = NM == 2 ? apply_act(vm[0], A.act) * vm[NM - 1] : vm[0];
        store_value(i, MODE == 1 ? v * out_scale : v);
    };
    // 2) the CTA's sums (fixed warp order): thread i owns output i of every matrix from here on
    // 3) across the cluster, always in rank order.  M <= 8: every peer pushes its CTA sums into the
    //    leader's receive buffer with st.async (completion counted on the leader's mbarrier) and exits;
    //    the leader waits for the bytes, adds and stores.  Larger M: the leader pulls through
    //    ld.shared::cluster between two cluster barriers.
    float* cta = reinterpret_cast<float*>(smem);
    if (ACC == 4 && RINGRECV && nsplit > 1) {
        constexpr int ITERS = NV * 32 / (GEMV_WARPS * 32);            // NVL <= NV * 32
        float vmall[ITERS][NM];
#pragma unroll
        for (int it = 0; it < ITERS; it++) {
            const int i = threadIdx.x + it * GEMV_WARPS * 32;
#pragma unroll
            for (int mt = 0; mt < NM; mt++) {
                float v = 0.0f;
                if (i < NVL) {
                    v = cta[mt * NVL + i];
#pragma unroll
                    for (int w = 1; w < GEMV_WARPS; w++) v += reinterpret_cast<const float*>(smem + w * WB)[mt * NVL + i];
                }
                vmall[it][mt] = v;
            }
        }
        __syncthreads();                                              // every slot has been read: the rings are free
        asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");   // ... in every CTA of the cluster
        float* rrecv = cta;                                           // leader's receive area = its rings
        if (slice != 0) {
            const uint32_t rb = map_to_rank(rbar, 0);
#pragma unroll
            for (int it = 0; it < ITERS; it++) {
                const int i = threadIdx.x + it * GEMV_WARPS * 32;
                if (i < NVL) {
#pragma unroll
                    for (int mt = 0; mt < NM; mt++)
                        st_async_f32(map_to_rank(smem_u32(rrecv + (slice - 1) * (NM * NVL) + mt * NVL + i), 0), vmall[it][mt], rb);
                }
            }
        } else {
            if (threadIdx.x == 0) mbar_expect_tx(rbar, (uint32_t)((nsplit - 1) * NM * NVL * 4));
            mbar_wait(rbar, 0);
#pragma unroll
            for (int it = 0; it < ITERS; it++) {
                const int i = threadIdx.x + it * GEMV_WARPS * 32;
                if (i < NVL) {
#pragma unroll
                    for (int mt = 0; mt < NM; mt++)
                        for (int rk = 1; rk < nsplit; rk++) vmall[it][mt] += rrecv[(rk - 1) * (NM * NVL) + mt * NVL + i];
                    finish(i, vmall[it]);
                }
            }
        }
        comm_done();
        TS(5);
        return;
    }
    if (PUSH && nsplit > 1) {
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");   // the leader's barrier init is visible
        if (slice == 0 && threadIdx.x == 0) mbar_expect_tx(rbar, (uint32_t)((nsplit - 1) * NM * NVL * 4));
    }
    for (int i = threadIdx.x; i < NVL; i += GEMV_WARPS * 32) {
        float vm[NM];
#pragma unroll
        for (int mt = 0; mt < NM; mt++) {
            float v = cta[mt * NVL + i];
#pragma unroll
            for (int w = 1; w < GEMV_WARPS; w++) v += reinterpret_cast<const float*>(smem + w * WB)[mt * NVL + i];
            vm[mt] = v;
        }
        if (nsplit == 1) {
            finish(i, vm);
        } else if (PUSH) {
            if (slice != 0) {
                const uint32_t rb = map_to_rank(rbar, 0);
#pragma unroll
                for (int mt = 0; mt < NM; mt++)
                    st_async_f32(map_to_rank(smem_u32(recv + (slice - 1) * (NM * NVL) + mt * NVL + i), 0), vm[mt], rb);
            } else {
                mbar_wait(rbar, 0);
#pragma unroll
                for (int mt = 0; mt < NM; mt++)
                    for (int rk = 1; rk < nsplit; rk++) vm[mt] += recv[(rk - 1) * (NM * NVL) + mt * NVL + i];
                finish(i, vm);
            }
        } else {
#pragma unroll
            for (int mt = 0; mt < NM; mt++) cta[mt * NVL + i] = vm[mt];
        }
    }
    if (!PUSH && nsplit > 1) {
        cluster_sync_all();
        if (slice == 0) {
            for (int i = threadIdx.x; i < NVL; i += GEMV_WARPS * 32) {
                float vm[NM];
#pragma unroll
                for (int mt = 0; mt < NM; mt++) {
                    float pv[GEMV_MAX_SPLIT - 1];
#pragma unroll
                    for (int rk = 1; rk < GEMV_MAX_SPLIT; rk++)      // all remote loads in flight before the adds
                        pv[rk - 1] = rk < nsplit ? ld_dsmem_f32(smem_u32(cta + mt * NVL + i), (uint32_t)rk) : 0.0f;
                    float v = cta[mt * NVL + i];
#pragma unroll
                    for (int rk = 1; rk < GEMV_MAX_SPLIT; rk++) v += pv[rk - 1];
                    vm[mt] = v;
                }
                finish(i, vm);
            }
        }
        cluster_sync_all();                                      // peers' smem must outlive the reads
    }
    comm_done();
    TS(5);
}

// K slices per 64-row group = cluster size: the largest of 8..1 for which every group's cluster
// is resident at once (cudaOccupancyMaxActiveClusters: cluster placement strands some SMs, e.g.
// size 4 leaves 132 of 148 usable) and every warp still gets at least one k-block.
// Fraction of the co-resident cluster slots one launch may take.  Below 1 the NEXT launch's CTAs
// find free slots while this one still runs, so that under programmatic dependent launch they
// stream their first chunks during this launch's compute and reduction tail.
static float occupancy_fraction() {
    static float f = -1.0f;
    if (f < 0.0f) {
        const char* e = getenv("LOWBIT_GEMV_OCC");
        f = e ? (float)atof(e) : 1.0f;
        if (!(f > 0.05f && f <= 1.0f)) f = 1.0f;
    }
    return f;
}

template <typename K>
static int pick_split(K kernel, int smem, int64_t groups, int64_t nb, int (&cache)[8]) {
    static const int sizes[8] = {8, 7, 6, 5, 4, 3, 2, 1};
    static int forced = -1;                              // development: LOWBIT_GEMV_SPLIT pins the cluster size
    if (forced < 0) { const char* e = getenv("LOWBIT_GEMV_SPLIT"); forced = e ? atoi(e) : 0; }
    if (forced >= 1 && forced <= 8 && (int64_t)forced * GEMV_WARPS <= nb) return forced;
    for (int i = 0; i < 8; i++) {
        const int sz = sizes[i];
        if (sz > 1 && (int64_t)sz * GEMV_WARPS > nb) continue;
        if (cache[i] < 0) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(sz, 64, 1); cfg.blockDim = dim3(GEMV_WARPS * 32); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = sz; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
            cache[i] = n;
        }
        if (groups <= (int64_t)(cache[i] * occupancy_fraction()) || sz == 1) return sz;
    }
    return 1;
}

// co-resident clusters of `kernel` at cluster size sz with `smem` bytes of dynamic shared memory (memoised)
template <typename K>
static int max_clusters(K kernel, int sz, int smem) {
    static std::mutex mu;
    static std::unordered_map<unsigned long long, int> memo;
    const unsigned long long key = ((unsigned long long)sz << 32) | (unsigned)smem;     // one map per kernel instantiation
    std::lock_guard<std::mutex> lk(mu);
    auto it = memo.find(key);
    if (it != memo.end()) return it->second;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(sz, 64, 1); cfg.blockDim = dim3(GEMV_WARPS * 32); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = sz; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
    memo[key] = n;
    return n;
}

// The hardware fills an SM up to its resident-CTA limit before it moves on, and under PDL the CTAs of the next
// launch take whatever slots free up first: with 6 slots per SM a 512-CTA launch ends up with 2 to 6 CTAs per SM,
// and the launch lasts as long as its fullest SM (measured: finish time proportional to the CTA count, 2.7 us at
// 3 CTAs, 5.1 us at 6).  So the launch asks for as much dynamic shared memory as makes ceil(CTAs / SMs) the
// per-SM limit -- provided every cluster is still co-resident.  Returns the padded size (or `smem`).
template <typename K>
static int balance_smem(K kernel, int smem, int split, int64_t clusters) {
    const int64_t ctas = clusters * split;
    const int sms = sm_count();
    for (int per_sm = (int)((ctas + sms - 1) / sms); per_sm <= 8; per_sm++) {
        const int want = 232448 / per_sm - 1024 - 64;          // 1 KB per CTA is reserved by the system
        if (want <= smem) break;                                // the natural footprint already allows no more than this
        // 5 % slack: a launch that fills every slot leaves none for the next launch's early CTAs
        if ((int64_t)max_clusters(kernel, split, want) * 20 >= clusters * 21) return want;
    }
    return smem;
}

// smem_ring / cache_ring (optional): the same kernel receiving into its rings (GemvArgs::ring_recv): less shared
// memory, a late cluster barrier; taken only if it allows a larger cluster.
template <typename K>
static void launch_kernel(K kernel, int smem, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st,
                          int (&cache)[8], int grid_z = 1, int smem_ring = 0, int (*cache_ring)[8] = nullptr) {
    // nb = k-blocks one cluster splits; the cluster count that has to fit in one wave is groups * grid_z
    int split = pick_split(kernel, smem, groups * grid_z, nb, cache);
    GemvArgs B = A;
    B.ring_recv = 0;
    if (cache_ring) {
        const int split_ring = pick_split(kernel, smem_ring, groups * grid_z, nb, *cache_ring);
        if (split_ring > split) { split = split_ring; smem = smem_ring; B.ring_recv = 1; }
    }
    B.slice_q = (int)(nb / split); B.slice_r = (int)(nb % split);
    {
        static int padl = -1;                            // development: LOWBIT_GEMV_PAD_L = resident CTAs per SM to allow
        if (padl < 0) { const char* e = getenv("LOWBIT_GEMV_PAD_L"); padl = e ? atoi(e) : 0; }
        if (padl >= 1 && padl <= 8) { const int want = 232448 / padl - 1024 - 64; if (want > smem) smem = want; }
        else if (padl == 0) smem = balance_smem(kernel, smem, split, groups * grid_z);
    }

    {
        static int dbg = -1;                             // development: LOWBIT_GEMV_DEBUG=1 prints every launch geometry
        if (dbg < 0) { const char* e = getenv("LOWBIT_GEMV_DEBUG"); dbg = e ? atoi(e) : 0; }
        if (dbg)
            fprintf(stderr, "[gemv] groups %lld x z %d, nb %lld: cluster %d, smem %d B (ring_recv %d), co-resident clusters at this size %d\n",
                    (long long)groups, grid_z, (long long)nb, split, smem, B.ring_recv, max_clusters(kernel, split, smem));
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)split, (unsigned)groups, (unsigned)grid_z);
    cfg.blockDim = dim3(GEMV_WARPS * 32); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[3];
    int na = 0;
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = split; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
    na++;
    {
        // load balancing: the default policy left 6 of 148 SMs without a single CTA of a 172 x 4 launch
        static int policy = -1;                          // development: LOWBIT_GEMV_CLUSTER_POLICY = 0 default, 1 spread, 2 load balancing
        if (policy < 0) { const char* e = getenv("LOWBIT_GEMV_CLUSTER_POLICY"); policy = e ? atoi(e) : 2; }
        if (policy == 1 || policy == 2) {
            at[na].id = cudaLaunchAttributeClusterSchedulingPolicyPreference;
            at[na].val.clusterSchedulingPolicyPreference = policy == 1 ? cudaClusterSchedulingPolicySpread : cudaClusterSchedulingPolicyLoadBalancing;
            na++;
        }
    }
    if (A.pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = at; cfg.numAttrs = na;
    cudaLaunchKernelEx(&cfg, kernel, B);
}

template <int QT, bool BF, int NS>
static void launch_nt(int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st) {
    using SG = Staging<QT, BF>;
    // per instantiation and device: raise the dynamic shared memory limit once (the attribute is per context)
    static bool configured_dev[64] = {};
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    bool& configured = configured_dev[dev_id & 63];
    static bool caches_ready = false;
    static int cache[8][8];
    static int cache_tl[2][4][8];        // [fused MLP][live t-lanes - 1]: the M <= 8 kernels' shared memory depends on M
    static int cache_ring[2][8];         // ... receiving into the rings
    if (!caches_ready) {
        for (auto& row : cache) for (int& v : row) v = -1;
        for (auto& a : cache_tl) for (auto& row : a) for (int& v : row) v = -1;
        for (auto& row : cache_ring) for (int& v : row) v = -1;
        caches_ready = true;
    }
    if (!configured) {
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 2, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 4, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 2, 4, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 4, 4, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 2, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 4, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        if constexpr (NS == 1) cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 2, 4, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 2, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        cudaFuncSetAttribute(gemv_kernel<QT, BF, NS, 1, 2, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        configured = true;
    }
    if (A.mode == 1) {                   // routed experts' gate/up: one slot per blockIdx.z (M == 1)
        launch_kernel(gemv_kernel<QT, BF, NS, 1, 2, 2, 1>, SG::smem_bytes(1, 2, 2), groups, nb, A, st, cache[6], A.n_seg);
        return;
    }
    if (A.mode == 2) {                   // routed experts' down projections, concatenated along K (M == 1)
        launch_kernel(gemv_kernel<QT, BF, NS, 1, 2, 1, 2>, SG::smem_bytes(1, 2, 1), groups, nb * A.n_seg, A, st, cache[7]);
        return;
    }
    if (A.w2) {                          // fused MLP: M <= 8, and 9..16 half-precision rows (the host falls back otherwise)
        if constexpr (NS == 1) {
            if (m > 8) {
                static int cache_nt2[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
                launch_kernel(gemv_kernel<QT, BF, NS, 2, 4, 2, 0>, SG::smem_bytes(2, 4, 2), groups, nb, A, st, cache_nt2);
                return;
            }
        }
        if (m == 1) launch_kernel(gemv_kernel<QT, BF, NS, 1, 2, 2, 0>, SG::smem_bytes(1, 2, 2), groups, nb, A, st, cache[4]);
        else launch_kernel(gemv_kernel<QT, BF, NS, 1, 4, 2, 0>, SG::smem_bytes(1, 4, 2, SG::live_t(1, 4, m)), groups, nb, A, st,
                           cache_tl[1][SG::live_t(1, 4, m) - 1], 1, SG::smem_bytes(1, 4, 2, 4, true), &cache_ring[1]);
        return;
    }
    if (m == 1) launch_kernel(gemv_kernel<QT, BF, NS, 1, 2, 1, 0>, SG::smem_bytes(1, 2, 1), groups, nb, A, st, cache[0]);
    else if (m <= 8) launch_kernel(gemv_kernel<QT, BF, NS, 1, 4, 1, 0>, SG::smem_bytes(1, 4, 1, SG::live_t(1, 4, m)), groups, nb, A, st,
                                   cache_tl[0][SG::live_t(1, 4, m) - 1], 1, SG::smem_bytes(1, 4, 1, 4, true), &cache_ring[0]);
    else if (m <= 16) launch_kernel(gemv_kernel<QT, BF, NS, 2, 4, 1, 0>, SG::smem_bytes(2, 4, 1), groups, nb, A, st, cache[2]);
    else launch_kernel(gemv_kernel<QT, BF, NS, 4, 4, 1, 0>, SG::smem_bytes(4, 4, 1), groups, nb, A, st, cache[3]);
}
template <int QT>
static void launch_qt(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st) {
    if (x_dtype == DT_F16) launch_nt<QT, false, 1>(m, groups, nb, A, st);
    else if (x_dtype == DT_BF16) launch_nt<QT, true, 1>(m, groups, nb, A, st);
    else launch_nt<QT, true, 3>(m, groups, nb, A, st);
}

// one translation unit per qtype instantiates this (keeps nvcc times parallel)
#define LOWBIT_DEFINE_GEMV_QT(NAME, QT)                                                                         \
    void gemv_launch_##NAME(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st) { \
        launch_qt<QT>(x_dtype, m, groups, nb, A, st);                                                           \
    }

}  // namespace lowbit
