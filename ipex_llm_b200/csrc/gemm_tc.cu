// gemm_tc.cu -- kernel (c): prefill dequant-GEMM on the 5th-generation tensor cores.
//
// y[M][N] (+)= x[M][K] . dequant(W[N][K])^T for M > 32 activation rows: what LowBitLinear.forward
// asks of xe_linear.forward_new during prefill (reference low_bit_linear.py:700-708) and of
// ggml_compute_forward_mul_mat_q_fp32 / the dequantise + F.linear fallback on CPU (:739-745).
//
// One CTA computes a [128*MT tokens] x [128 features] tile (MT = 1 or 2), looping over K in
// 64-element k-blocks through a 3-stage shared-memory pipeline:
//   warp 0   producer : TMA 2-D tiled loads (128B swizzle) of the activation tile(s) and 1-D TMA
//                       bulk copies of the PACKED weights of the tile's two 64-row groups (codes +
//                       fp16 scales [+ minima]) -- weights cross HBM->SM at 4.25-8.25 bits each;
//   warps 2-17 dequant: two groups of 8 take alternate k-blocks and expand the packed codes to fp16/bf16
//                       (magic-number trick / cvt, scale applied with one multiply per pair) straight into
//                       the K-major 128B-swizzled B operand tile in shared memory;
//   warps 18-21       : epilogue (tcgen05.ld of their 32 TMEM lanes -> bias/accumulate -> 16-byte stores);
//   warp 1   MMA      : the whole warp runs the issue loop in warp-uniform control flow and one elected lane
//                       issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=128, K=16; 4 per k-block per token
//                       sub-tile), accumulators in TMEM (128*MT columns), tcgen05.commit releases the stage and
//                       finally signals the epilogue.  (Round 2: with the loop inside a single-lane branch every
//                       operand was moved into a uniform register per instruction, ~45 ns per MMA: the issue
//                       thread, not shared-memory bandwidth, set the pace -- 1017 -> 1262 TFLOP/s.)
// The expanded weights never touch HBM; both token sub-tiles reuse one expanded B tile.
// Weight rounding: (code - offset) * d is rounded ONCE to fp16 (11 bits) -- or to bf16 when the
// activations are bf16 -- which is the precision contract of this path (tests: 2e-3 * sum|x w|).
#include <cuda.h>
#include <cuda_fp8.h>
#include <mutex>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"
#include "gemv_args.h"       // apply_act

namespace lowbit {

namespace gemmtc {

constexpr int kMaxStages = 6;        // pipeline depth = as many stages as fit in 227 KB, at most this
constexpr int kDeqWarps = 16;        // dequant warps in two groups of 8 that take alternate k-blocks; a warp expands one 16x64 tile of the
                                     // 128x64 B stage (one group alone: 0.4 us per k-block, the pace of the whole kernel at <= 128-token tiles)
constexpr int kEpiWarps = 4;         // epilogue warps: one per TMEM lane quarter (warp index % 4 must be 0..3 in order)
constexpr int kThreads = 64 + 32 * kDeqWarps + 32 * kEpiWarps;
constexpr int kTileN = 128;          // features per CTA (two 64-row groups)
constexpr int kTileBytesAB = 128 * 128;   // one 128-row x 64-element 2-byte operand tile

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
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
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
// K-major, 128B-swizzled operand tile whose rows are 128 bytes: LBO = 1 (unused), SBO = 1024 B
// (eight rows), descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)64 << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// The same two for a whole warp in warp-uniform control flow: `leader` is true in one elected lane.  With every operand
// computed identically by all lanes the compiler keeps descriptors and addresses in uniform registers; a single-lane
// branch made it move every operand into a uniform register per instruction (~23 instructions and 45 ns per MMA).
__device__ __forceinline__ void umma_f16_w(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate, uint32_t leader) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "setp.ne.b32 q, %5, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(leader) : "memory");
}
__device__ __forceinline__ void umma_commit_w(uint32_t bar, uint32_t leader) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.ne.b32 q, %1, 0;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar), "r"(leader) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_w(uint32_t bar, uint32_t bytes, uint32_t leader) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}"
                 ::"r"(bar), "r"(bytes), "r"(leader) : "memory");
}
__device__ __forceinline__ void bulk_g2s_w(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint32_t leader) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\t"
                 "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n\t}"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "r"(leader) : "memory");
}
__device__ __forceinline__ void tma_load_2d_w(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint32_t leader) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\t"
                 "@q cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n\t}"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar), "r"(leader) : "memory");
}
__device__ __forceinline__ uint32_t elect_leader() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t h2_to_bf2(uint32_t h) {
    const float2 f = __half22float2(u2h(h));
    __nv_bfloat162 b = __floats2bfloat162_rn(f.x, f.y);
    return *reinterpret_cast<uint32_t*>(&b);
}
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Scaled fp16 values of k-step s of one 16x64 tile for lane (g,t): out[0] = row g (4s,4s+1),
// out[1] = row g+8 (4s,4s+1), out[2] = row g (4s+2,4s+3), out[3] = row g+8 (4s+2,4s+3); the same
// register order as the mma.sync fragments of the GEMV kernel (the tile layout is shared).
template <int QT> struct Expand;

template <int OFFSET> struct Expand4 {          // sym_int4 (OFFSET 1) / asym_int4 (OFFSET 2)
    static constexpr int kTileBytes = 512;
    static constexpr bool kHasMin = OFFSET == 2;
    static constexpr int kLutBytes = 0;
    // the lane's 16 bytes of the tile in one conflict-free LDS.128 (four LDS.32 at a 16-byte lane stride cost 4 wavefronts each)
    struct Codes { uint4 q; };
    __device__ static __forceinline__ void load(uint32_t codes, int lane, Codes& c) { c.q = lds128(codes + lane * 16); }
    __device__ static __forceinline__ void run(const Codes& c, int s, __half2 dlo, __half2 dhi, __half2 mlo,
                                               __half2 mhi, uint32_t, uint32_t (&out)[4]) {
        const uint32_t w = s == 0 ? c.q.x : s == 1 ? c.q.y : s == 2 ? c.q.z : c.q.w;
        const uint32_t w2 = w >> 8;
        const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
        // 0x6400 | n = 1024 + n ; 0x6400 | (n << 4) = 1024 + 16 n  (exact in fp16)
        if (OFFSET == 1) {
            // n - 8 straight from the magic number (exact): one subtraction / one fma per pair, then the product, rounded once
            const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f), km72 = __floats2half2_rn(-72.0f, -72.0f);
            const __half2 v0 = __hsub2(u2h(and_or(w, 0x000f000fu, 0x64006400u)), k1032);
            const __half2 v1 = __hfma2(u2h(and_or(w, 0x00f000f0u, 0x64006400u)), k16th, km72);
            const __half2 v2 = __hsub2(u2h(and_or(w2, 0x000f000fu, 0x64006400u)), k1032);
            const __half2 v3 = __hfma2(u2h(and_or(w2, 0x00f000f0u, 0x64006400u)), k16th, km72);
            out[0] = h2u(__hmul2(v0, dlo)); out[1] = h2u(__hmul2(v1, dhi));
            out[2] = h2u(__hmul2(v2, dlo)); out[3] = h2u(__hmul2(v3, dhi));
        } else {
            const __half2 k1024 = __floats2half2_rn(1024.0f, 1024.0f), km64 = __floats2half2_rn(-64.0f, -64.0f);
            const __half2 v0 = __hsub2(u2h(and_or(w, 0x000f000fu, 0x64006400u)), k1024);                 // n
            const __half2 v1 = __hfma2(u2h(and_or(w, 0x00f000f0u, 0x64006400u)), k16th, km64);           // n
            const __half2 v2 = __hsub2(u2h(and_or(w2, 0x000f000fu, 0x64006400u)), k1024);
            const __half2 v3 = __hfma2(u2h(and_or(w2, 0x00f000f0u, 0x64006400u)), k16th, km64);
            out[0] = h2u(__hfma2(v0, dlo, mlo)); out[1] = h2u(__hfma2(v1, dhi, mhi));
            out[2] = h2u(__hfma2(v2, dlo, mlo)); out[3] = h2u(__hfma2(v3, dhi, mhi));
        }
    }
};
template <> struct Expand<QT_SYM_INT4> : Expand4<1> {};
template <> struct Expand<QT_ASYM_INT4> : Expand4<2> {};

template <int QT> struct Expand8 {
    static constexpr int kTileBytes = 1024;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ uint32_t pair(uint32_t w, int hi) {
        uint32_t h;
        if (QT == QT_FP8_E5M2) h = __byte_perm(w, 0u, hi ? 0x3424u : 0x1404u);
        else {
            const unsigned short two = (unsigned short)(hi ? (w >> 16) : (w & 0xffffu));
            asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h) : "h"(two));
        }
        return h;
    }
    static constexpr int kLutBytes = 0;
    struct Codes { uint4 q0, q1; };
    __device__ static __forceinline__ void load(uint32_t codes, int lane, Codes& c) {
        c.q0 = lds128(codes + lane * 16); c.q1 = lds128(codes + 512 + lane * 16);
    }
    __device__ static __forceinline__ void run(const Codes& c, int s, __half2 dlo, __half2 dhi, __half2, __half2,
                                               uint32_t, uint32_t (&out)[4]) {
        const uint32_t w0 = s == 0 ? c.q0.x : s == 1 ? c.q0.y : s == 2 ? c.q0.z : c.q0.w;
        const uint32_t w1 = s == 0 ? c.q1.x : s == 1 ? c.q1.y : s == 2 ? c.q1.z : c.q1.w;
        out[0] = h2u(__hmul2(u2h(pair(w0, 0)), dlo)); out[1] = h2u(__hmul2(u2h(pair(w1, 0)), dhi));
        out[2] = h2u(__hmul2(u2h(pair(w0, 1)), dlo)); out[3] = h2u(__hmul2(u2h(pair(w1, 1)), dhi));
    }
};
template <> struct Expand<QT_FP8_E4M3> : Expand8<QT_FP8_E4M3> {};
template <> struct Expand<QT_FP8_E5M2> : Expand8<QT_FP8_E5M2> {};

template <> struct Expand<QT_FP6> {
    static constexpr int kTileBytes = 768;
    static constexpr bool kHasMin = false;
    __device__ static __forceinline__ uint32_t cvt2(unsigned short two) {
        uint32_t h;
        asm("cvt.rn.f16x2.e3m2x2 %0, %1;" : "=r"(h) : "h"(two));
        return h;
    }
    static constexpr int kLutBytes = 0;
    struct Codes { uint4 lo; uint2 hi; };
    __device__ static __forceinline__ void load(uint32_t codes, int lane, Codes& c) {
        c.lo = lds128(codes + lane * 16);
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(c.hi.x), "=r"(c.hi.y) : "r"(codes + 512 + lane * 8));
    }
    __device__ static __forceinline__ void run(const Codes& c, int s, __half2 dlo, __half2 dhi, __half2, __half2,
                                               uint32_t, uint32_t (&out)[4]) {
        const int u = s >> 1;
        const uint32_t l0 = u ? c.lo.y : c.lo.x, l1 = u ? c.lo.w : c.lo.z, h0 = c.hi.x, h1 = c.hi.y;
        const uint32_t sh = (s & 1) ? 4u : 0u;
        const uint32_t e0 = ((l0 >> sh) & 0x0f0f0f0fu) | (((h0 >> (2 * s)) & 0x03030303u) << 4);
        const uint32_t e1 = ((l1 >> sh) & 0x0f0f0f0fu) | (((h1 >> (2 * s)) & 0x03030303u) << 4);
        out[0] = h2u(__hmul2(u2h(cvt2((unsigned short)(e0 & 0xffffu))), dlo));
        out[1] = h2u(__hmul2(u2h(cvt2((unsigned short)(e1 & 0xffffu))), dhi));
        out[2] = h2u(__hmul2(u2h(cvt2((unsigned short)(e0 >> 16))), dlo));
        out[3] = h2u(__hmul2(u2h(cvt2((unsigned short)(e1 >> 16))), dhi));
    }
};

// nf4: byte b of the lane's word s is the code pair of out[b] (tilednf4_pos); a 256-entry table of
// fp16 pairs of NormalFloat values sits in shared memory, 128 bytes per entry = one 4-byte slot per lane, so that no two
// lanes of a lookup meet in a bank (the first layout, 8 slots per entry, cost 3.1 wavefronts per LDS -- see the GEMV codec).
template <> struct Expand<QT_NF4> {
    static constexpr int kTileBytes = 512;
    static constexpr bool kHasMin = false;
    static constexpr int kLutBytes = 256 * 128;
    __device__ static __forceinline__ void fill_lut(uint8_t* lut, int tid, int nthreads) {
        for (int e = tid; e < 256; e += nthreads) {
            const uint32_t v = h2u(__floats2half2_rn(kNF4Table[e & 15], kNF4Table[e >> 4]));
            uint4* dst = reinterpret_cast<uint4*>(lut + e * 128);
#pragma unroll
            for (int c = 0; c < 8; c++) dst[c] = make_uint4(v, v, v, v);
        }
    }
    struct Codes { uint4 q; };
    __device__ static __forceinline__ void load(uint32_t codes, int lane, Codes& c) { c.q = lds128(codes + lane * 16); }
    // lut = table address + 4 * lane
    __device__ static __forceinline__ void run(const Codes& c, int s, __half2 dlo, __half2 dhi, __half2, __half2,
                                               uint32_t lut, uint32_t (&out)[4]) {
        const uint32_t w = s == 0 ? c.q.x : s == 1 ? c.q.y : s == 2 ? c.q.z : c.q.w;
        uint32_t v0, v1, v2, v3;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v0) : "r"(lut + __byte_perm(w, 0u, 0x4440u) * 128u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v1) : "r"(lut + __byte_perm(w, 0u, 0x4441u) * 128u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v2) : "r"(lut + __byte_perm(w, 0u, 0x4442u) * 128u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v3) : "r"(lut + __byte_perm(w, 0u, 0x4443u) * 128u));
        out[0] = h2u(__hmul2(u2h(v0), dlo)); out[1] = h2u(__hmul2(u2h(v1), dhi));
        out[2] = h2u(__hmul2(u2h(v2), dlo)); out[3] = h2u(__hmul2(u2h(v3), dhi));
    }
};

struct GemmArgs {
    const uint8_t* w;
    const uint8_t* w2;       // a second matrix of the same shape / qtype (gate and up of an MLP) or NULL: its feature tiles
    void* y2;                // follow w's in the tile order and go to y2 -- one grid, one set of rounds over the SMs
    TiledGeom G;
    int M;
    void* y; int y_dtype;
    const void* bias;
    int accumulate;
    int half0;               // 256-token tiles only: tiles from this index on (the grid's last, partial round) run as two
                             // 128-token units each on twice as many SMs; >= the tile count: none
    int ksplit;              // K splits per output tile (1: the epilogue writes y)
    float* part;             // ksplit > 1: fp32 partial sums [ksplit][M][N]
};

template <int QT> struct Smem {
    static constexpr int kSuper = RG_TILES * Expand<QT>::kTileBytes;
    static constexpr int kPackedGroup = kSuper + 128 + (Expand<QT>::kHasMin ? 128 : 0);
    static constexpr int kPacked = ((2 * kPackedGroup + 127) / 128) * 128;
    // every stage starts 1024-byte aligned (the swizzled operand tiles sit at its start)
    __host__ __device__ static constexpr int stage_bytes(int mt) { return ((mt * kTileBytesAB + kTileBytesAB + kPacked + 1023) / 1024) * 1024; }
    // [align slack | stages | barriers (256) | codec lookup table]
    __host__ __device__ static constexpr int stages(int mt) {
        return (232448 - 1024 - 256 - Expand<QT>::kLutBytes) / stage_bytes(mt) < kMaxStages
                   ? (232448 - 1024 - 256 - Expand<QT>::kLutBytes) / stage_bytes(mt) : kMaxStages;
    }
    __host__ __device__ static constexpr int total(int mt) { return 1024 + stages(mt) * stage_bytes(mt) + 256 + Expand<QT>::kLutBytes; }
};

// 32 consecutive features of one token: 64 bytes (fp16 / bf16) or 128 bytes (fp32), written with
// 16-byte stores (n0 is a multiple of 32 and N of 128, so the run is 64-byte aligned).  Scalar
// 2-byte stores cost one 32-byte sector transaction each: 16x the traffic.
__device__ __forceinline__ void store_vals(const GemmArgs& A, void* ybase, int64_t token, int64_t n0, float (&v)[32]) {
    const int64_t N = A.G.rows;
    const int64_t idx = token * N + n0;
    if (A.y_dtype == DT_F32) {
        float4* p = reinterpret_cast<float4*>((float*)ybase + idx);
        const float4* bp = reinterpret_cast<const float4*>((const float*)A.bias + n0);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            float4 o = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            if (A.bias) { const float4 b = bp[i]; o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w; }
            if (A.accumulate) { const float4 c = p[i]; o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w; }
            p[i] = o;
        }
        return;
    }
    const bool f16 = A.y_dtype == DT_F16;
    auto unpack = [&](uint32_t u, float& lo, float& hi) {
        if (f16) { const float2 f = __half22float2(u2h(u)); lo = f.x; hi = f.y; }
        else { lo = __uint_as_float(u << 16); hi = __uint_as_float(u & 0xffff0000u); }
    };
    auto pack = [&](float lo, float hi) -> uint32_t {
        if (f16) return h2u(__floats2half2_rn(lo, hi));
        __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&b);
    };
    uint4* p = reinterpret_cast<uint4*>((uint16_t*)ybase + idx);
    const uint4* bp = reinterpret_cast<const uint4*>((const uint16_t*)A.bias + n0);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (A.bias) {
            const uint4 b = bp[i];
            const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int j = 0; j < 4; j++) { float lo, hi; unpack(bw[j], lo, hi); v[8 * i + 2 * j] += lo; v[8 * i + 2 * j + 1] += hi; }
        }
        if (A.accumulate) {
            const uint4 c = p[i];
            const uint32_t cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int j = 0; j < 4; j++) { float lo, hi; unpack(cw[j], lo, hi); v[8 * i + 2 * j] += lo; v[8 * i + 2 * j + 1] += hi; }
        }
        p[i] = make_uint4(pack(v[8 * i], v[8 * i + 1]), pack(v[8 * i + 2], v[8 * i + 3]), pack(v[8 * i + 4], v[8 * i + 5]),
                          pack(v[8 * i + 6], v[8 * i + 7]));
    }
}

// grid = (N / 128 feature tiles, ceil(M / (128*MT)) token tiles)
// Persistent: grid = min(units, SMs) CTAs; CTA c takes units c, c + grid, ... in token-tile-major order (the
// CTAs running at one moment share the token tile's activations in L2 and stream different weights).
// Warp roles: 0 TMA producer, 1 MMA issuer (owns TMEM), 2..17 dequant, 18..21 epilogue.  The accumulators
// are double buffered in TMEM (2 x 128*MT columns), so the epilogue of tile i overlaps the main loop of i+1.
template <int QT, bool BF, int MT>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap xmap, const GemmArgs A) {
    using E = Expand<QT>;
    using SM = Smem<QT>;
    constexpr int STAGE = SM::stage_bytes(MT);
    constexpr int kStages = SM::stages(MT);
    static_assert(kStages >= 2 && 3 * kStages * 8 + 48 <= 256, "barrier region is 256 bytes");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // 128B swizzle wants 1024-byte alignment
    const uint32_t bars = base + kStages * STAGE;                      // full[s], deq[s], empty[s], tmem_full[2], tmem_empty[2], tmem slot
    auto full_bar = [&](int s) { return bars + s * 8; };
    auto deq_bar = [&](int s) { return bars + (kStages + s) * 8; };
    auto empty_bar = [&](int s) { return bars + (2 * kStages + s) * 8; };
    auto tmem_full = [&](int b) { return bars + (3 * kStages + b) * 8; };
    auto tmem_empty = [&](int b) { return bars + (3 * kStages + 2 + b) * 8; };
    const uint32_t tmem_slot = bars + (3 * kStages + 4) * 8;

    // the warp index through a shuffle: the compiler then treats everything derived from it as warp-uniform
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const uint32_t lut = bars + 256 + 4 * lane;
    if constexpr (E::kLutBytes > 0) E::fill_lut(smem_raw + (bars + 256 - smem_u32(smem_raw)), threadIdx.x, kThreads);   // visible after the __syncthreads below
    const int nb = (int)A.G.nb;
    const int n_ft1 = (int)(A.G.rows / kTileN);
    const int n_ftiles = A.w2 ? 2 * n_ft1 : n_ft1;         // the second matrix's feature tiles follow the first's
    const int n_tiles = n_ftiles * ((A.M + 128 * MT - 1) / (128 * MT));
    // work units = tiles x K splits (A.ksplit > 1 when the tiles alone would leave most SMs idle); split ks covers the
    // k-blocks [kb_begin(ks), kb_begin(ks + 1))
    const bool halves = MT == 2 && A.half0 < n_tiles;     // (host: only with ksplit == 1)
    const int n_units = halves ? A.half0 + 2 * (n_tiles - A.half0) : n_tiles * A.ksplit;
    auto kb_begin = [&](int ks) { return A.ksplit == 1 ? ks * nb : (ks * nb) / A.ksplit; };      // ks * nb < 2^31
    // unit -> output tile, K split, first 128-token sub-tile and the number of sub-tiles it covers
    auto decode_unit = [&](int unit, int& tile, int& ks, int& mt0, int& mt_n) {
        if (halves && unit >= A.half0) { const int hu = unit - A.half0; tile = A.half0 + (hu >> 1); ks = 0; mt0 = hu & 1; mt_n = 1; }
        else { tile = unit % n_tiles; ks = unit / n_tiles; mt0 = 0; mt_n = MT; }
    };
    constexpr uint32_t ACC_COLS = 128 * MT;            // one accumulator buffer
    constexpr uint32_t TMEM_COLS = 2 * ACC_COLS;       // 256 or 512 (powers of two)

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { mbar_init(full_bar(s), 1); mbar_init(deq_bar(s), 32 * (kDeqWarps / 2)); mbar_init(empty_bar(s), 1); }
        for (int b = 0; b < 2; b++) { mbar_init(tmem_full(b), 1); mbar_init(tmem_empty(b), 32 * kEpiWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp == 0) {
        // ================= producer: the whole warp runs the loop, one elected lane issues the copies (see umma_f16_w) =================
        {
            const uint32_t leader = elect_leader();
            int it = 0;                                              // k-iterations issued so far (all tiles)
            for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
                int tile, ks, mt0, mt_n;
                decode_unit(unit, tile, ks, mt0, mt_n);
                const int ftile = tile % n_ftiles, m0 = (tile / n_ftiles) * (128 * MT) + mt0 * 128;
                const uint8_t* wb = ftile < n_ft1 ? A.w : A.w2;
                const int64_t rg0 = (int64_t)(ftile < n_ft1 ? ftile : ftile - n_ft1) * 2;
                const int kb1 = kb_begin(ks + 1);
                for (int kb = kb_begin(ks); kb < kb1; kb++, it++) {
                    const int s = it % kStages, round = it / kStages;
                    mbar_wait(empty_bar(s), (uint32_t)((round & 1) ^ 1));
                    const uint32_t st = base + s * STAGE;
                    mbar_expect_tx_w(full_bar(s), mt_n * kTileBytesAB + 2 * SM::kPackedGroup, leader);
#pragma unroll
                    for (int mt = 0; mt < MT; mt++)
                        if (mt < mt_n) tma_load_2d_w(st + mt * kTileBytesAB, &xmap, kb * 64, m0 + mt * 128, full_bar(s), leader);
                    const uint32_t pk = st + (MT + 1) * kTileBytesAB;
#pragma unroll
                    for (int gi = 0; gi < 2; gi++) {
                        const int64_t u = (rg0 + gi) * nb + kb;
                        const uint32_t dst = pk + gi * SM::kPackedGroup;
                        bulk_g2s_w(dst, wb + u * SM::kSuper, SM::kSuper, full_bar(s), leader);
                        bulk_g2s_w(dst + SM::kSuper, wb + A.G.off_s + u * 128, 128, full_bar(s), leader);
                        if (E::kHasMin) bulk_g2s_w(dst + SM::kSuper + 128, wb + A.G.off_m + u * 128, 128, full_bar(s), leader);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer: the whole warp runs the loop, one elected lane issues (see umma_f16_w) =================
        {
            const uint32_t leader = elect_leader();
            // c_format F32 (1<<4), a/b format (0 f16, 1 bf16), both K-major, N = 128 (>>3 at bit 17), M = 128 (>>4 at bit 24)
            const uint32_t idesc = (1u << 4) | ((BF ? 1u : 0u) << 7) | ((BF ? 1u : 0u) << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
            int it = 0, ti = 0;
            for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x, ti++) {
                const int buf = ti & 1;
                mbar_wait(tmem_empty(buf), (uint32_t)(((ti >> 1) & 1) ^ 1));      // the epilogue has drained this buffer
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc = tmem_base + buf * ACC_COLS;
                int tile_, ks, mt0_, mt_n;
                decode_unit(unit, tile_, ks, mt0_, mt_n);
                const int kb0 = kb_begin(ks), kb1 = kb_begin(ks + 1);
                for (int kb = kb0; kb < kb1; kb++, it++) {
                    const int s = it % kStages, round = it / kStages;
                    mbar_wait(full_bar(s), (uint32_t)(round & 1));
                    mbar_wait(deq_bar(s), (uint32_t)(round & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t st = base + s * STAGE;
                    const uint64_t ad0 = make_desc(st), bd0 = make_desc(st + MT * kTileBytesAB);
#pragma unroll
                    for (int mt = 0; mt < MT; mt++) {
                        if (mt >= mt_n) break;
#pragma unroll
                        for (int k4 = 0; k4 < 4; k4++)       // the address field counts 16-byte units
                            umma_f16_w(acc + mt * 128, ad0 + (uint64_t)((mt * kTileBytesAB + k4 * 32) >> 4), bd0 + (uint64_t)((k4 * 32) >> 4), idesc,
                                       (kb > kb0 || k4 > 0) ? 1u : 0u, leader);
                    }
                    umma_commit_w(empty_bar(s), leader);        // stage reusable once these MMAs have read it
                }
                umma_commit_w(tmem_full(buf), leader);           // this tile's accumulators are complete
            }
        }
    } else if (warp < 2 + kDeqWarps) {
        // ================= dequant warps: group (warp - 2) / 8 takes the k-iterations of its parity =================
        const int grp_id = (warp - 2) >> 3;
        const int dw = (warp - 2) & 3;                       // tile r = dw of row group gi
        const int gi = ((warp - 2) >> 2) & 1;
        const int g = lane >> 2, t = lane & 3;
        int it = 0;
        for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
            int tile_, ks, mt0_, mt_n_;
            decode_unit(unit, tile_, ks, mt0_, mt_n_);
            const int kb0 = kb_begin(ks), kb1 = kb_begin(ks + 1);
            for (int kb = kb0; kb < kb1; kb++, it++) {
                const int s = it % kStages, round = it / kStages;
                // both groups wait for every stage, also the other group's: with an odd ring depth a stage alternates
                // between the groups, and a warp that skipped a phase of full_bar(s) could take the parity of the phase
                // before last for the one it is waiting for (parity waits are only safe one phase behind)
                mbar_wait(full_bar(s), (uint32_t)(round & 1));
                if ((it & 1) != grp_id) continue;
                const uint32_t st = base + s * STAGE;
                const uint32_t btile = st + MT * kTileBytesAB;
                const uint32_t pk = st + (MT + 1) * kTileBytesAB;
                const uint32_t grp = pk + gi * SM::kPackedGroup;
                const uint32_t codes = grp + dw * E::kTileBytes;
                uint32_t dq, mq = 0;
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(dq) : "r"(grp + SM::kSuper + g * 16 + dw * 4));
                if (E::kHasMin) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(mq) : "r"(grp + SM::kSuper + 128 + g * 16 + dw * 4));
                const __half2 d2 = u2h(dq), m2 = u2h(mq);
                const __half2 dlo = __low2half2(d2), dhi = __high2half2(d2);
                const __half2 mlo = __low2half2(m2), mhi = __high2half2(m2);
                // feature rows of this lane inside the 128-row B tile: f = 64*gi + 16*dw + g (+8)
                const int f_lo = 64 * gi + 16 * dw + g;
                const uint32_t row_lo = btile + (f_lo >> 3) * 1024 + (f_lo & 7) * 128;
                const uint32_t row_hi = row_lo + 1024;      // f + 8: next 8-row group, same row inside it
                typename E::Codes cw;
                E::load(codes, lane, cw);
#pragma unroll
                for (int q = 0; q < 2; q++) {               // k-steps 2q, 2q+1 -> one 16-byte chunk per row
                    uint32_t a[4], b[4];
                    E::run(cw, 2 * q, dlo, dhi, mlo, mhi, lut, a);
                    E::run(cw, 2 * q + 1, dlo, dhi, mlo, mhi, lut, b);
                    if (BF) {
#pragma unroll
                        for (int i = 0; i < 4; i++) { a[i] = h2_to_bf2(a[i]); b[i] = h2_to_bf2(b[i]); }
                    }
                    const uint32_t chunk = (uint32_t)((2 * t + q) ^ g) * 16;      // 128B swizzle: chunk ^ (row & 7)
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row_lo + chunk), "r"(a[0]), "r"(a[2]), "r"(b[0]), "r"(b[2]) : "memory");
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row_hi + chunk), "r"(a[1]), "r"(a[3]), "r"(b[1]), "r"(b[3]) : "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> async-proxy (UMMA) reads
                mbar_arrive(deq_bar(s));
            }
        }
    } else {
        // ================= epilogue warps: warp w may touch TMEM lanes 32*(w%4) .. +31 =================
        const int quarter = warp & 3;
        int ti = 0;
        for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x, ti++) {
            int tile, ks, mt0, mt_n;
            decode_unit(unit, tile, ks, mt0, mt_n);
            const int ftile_all = tile % n_ftiles, m0 = (tile / n_ftiles) * (128 * MT) + mt0 * 128;
            const int ftile = ftile_all < n_ft1 ? ftile_all : ftile_all - n_ft1;
            void* ybase = ftile_all < n_ft1 ? A.y : A.y2;
            const int buf = ti & 1;
            mbar_wait(tmem_full(buf), (uint32_t)((ti >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
                if (mt >= mt_n) break;
                const int64_t token = (int64_t)m0 + mt * 128 + quarter * 32 + lane;
#pragma unroll 1
                for (int c = 0; c < 128; c += 32) {
                    float v[32];
                    tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * ACC_COLS + mt * 128 + c), v);
                    if (token < A.M) {
                        if (A.ksplit > 1) {             // K split: fp32 partial sums, added up by reduce_partials_kernel
                            float4* pp = reinterpret_cast<float4*>(A.part + ((int64_t)ks * A.M + token) * A.G.rows + (int64_t)ftile * kTileN + c);
#pragma unroll
                            for (int i = 0; i < 8; i++) pp[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                        } else store_vals(A, ybase, token, (int64_t)ftile * kTileN + c, v);
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(tmem_empty(buf));                    // the MMA warp may overwrite this buffer
        }
    }
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// fp32 activations -> bf16 (prefill precision contract) into the workspace
__global__ void f32_to_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = __float2bfloat16_rn(x[i]);
}

// K split: y[token][n] = sum over splits (in split order) of the fp32 partials, + bias, + y when accumulating
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float* __restrict__ part, int ksplit, int64_t mn, int64_t n_cols,
                                                              void* y, int y_dtype, const void* bias, int accumulate) {
    for (int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * 4; i < mn; i += (int64_t)gridDim.x * blockDim.x * 4) {
        float4 a = *reinterpret_cast<const float4*>(part + i);
        for (int ks = 1; ks < ksplit; ks++) {
            const float4 b = *reinterpret_cast<const float4*>(part + ks * mn + i);
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        float v[4] = {a.x, a.y, a.z, a.w};
        const int64_t col = i % n_cols;                           // n_cols % 128 == 0: the four values share a row
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (bias) v[j] += y_dtype == DT_F32 ? ((const float*)bias)[col + j]
                            : (y_dtype == DT_F16 ? __half2float(((const __half*)bias)[col + j]) : __bfloat162float(((const __nv_bfloat16*)bias)[col + j]));
            if (y_dtype == DT_F32) { float* p = (float*)y + i + j; *p = accumulate ? *p + v[j] : v[j]; }
            else if (y_dtype == DT_F16) { __half* p = (__half*)y + i + j; *p = __float2half_rn(accumulate ? __half2float(*p) + v[j] : v[j]); }
            else { __nv_bfloat16* p = (__nv_bfloat16*)y + i + j; *p = __float2bfloat16_rn(accumulate ? __bfloat162float(*p) + v[j] : v[j]); }
        }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
        else cudaGetLastError();
    });
    return fn;
}

template <int QT, bool BF, int MT>
static int launch_one(const CUtensorMap& map, const GemmArgs& A, dim3 grid, cudaStream_t st) {
    constexpr int SMEM = Smem<QT>::total(MT);
    static bool configured_dev[64] = {};          // the attribute is per device context
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    bool& configured = configured_dev[dev_id & 63];
    if (!configured) {
        cudaFuncSetAttribute(gemm_kernel<QT, BF, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        configured = true;
    }
    gemm_kernel<QT, BF, MT><<<grid, kThreads, SMEM, st>>>(map, A);
    return 0;
}
template <int QT>
static int launch_qt(bool bf, int mt, const CUtensorMap& map, const GemmArgs& A, dim3 grid, cudaStream_t st) {
    if (bf) return mt == 2 ? launch_one<QT, true, 2>(map, A, grid, st) : launch_one<QT, true, 1>(map, A, grid, st);
    return mt == 2 ? launch_one<QT, false, 2>(map, A, grid, st) : launch_one<QT, false, 1>(map, A, grid, st);
}


// ====================================================================================================================
// Skinny variant: 9 <= M <= 128 activation rows (batched decode, short prefills).  The roles of the operands are swapped:
// the EXPANDED WEIGHTS are the 128-row A operand and the tokens are the N dimension (16 / 32 / 64 / 128), so nothing is
// padded to a 128-token tile, an activation stage is NTOK x 256 bytes and the accumulator is NTOK TMEM columns.
//   * work = the linear sequence of (128-feature tile, PAIR of k-blocks) items, cut into gridDim.x contiguous ranges of
//     equal length (stream-K): every SM expands the same number of blocks whatever the matrix shape; a range that crosses
//     a tile boundary produces one partial result per tile it touches (fp32, in the workspace) and a second,
//     programmatic-dependent launch adds a tile's partials in CTA order (deterministic) and applies bias / accumulate;
//   * an item is 128 k: the barrier handshakes, the MMA thread's loop and the bulk copies are per item, and measured they,
//     not the arithmetic, set the pace (0.34 us per handshake round with nothing else in the kernel);
//   * warp 0 streams PACKED weights + the activation k-slices through a ring of stages (the bytes in flight keep HBM
//     busy; the weights of the first round are requested before griddepcontrol.wait);
//   * 16 dequant warps in two groups take alternate items; a group owns one 128 x 128 K-major 128B-swizzled A tile
//     (two 16 KB atoms) and every warp expands its 16 rows of both k-blocks -- one group's barrier waits, LDS / STS
//     latency and proxy fence overlap the other group's arithmetic;
//   * warp 1 issues tcgen05.mma M = 128, N = NTOK, K = 16 (8 per item) into one of two TMEM accumulators; ONE commit per
//     item frees both the stage (for the producer) and the group's A tile;  4 epilogue warps read TMEM (lane = feature,
//     column = token).
constexpr int kSkStagesMax = 12;
constexpr int kSkDeqWarps = 16;
constexpr int kSkThreads = 64 + 32 * kSkDeqWarps + 128;

#ifdef LOWBIT_SK_TIMESTAMPS
__device__ unsigned long long g_sk_ts[6][64];
#define SKTS(row, it) do { if (blockIdx.x == 0 && lane == 0 && (it) < 64) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_sk_ts[row][it] = t_; } } while (0)
#else
#define SKTS(row, it) do { } while (0)
#endif

struct SkinnyArgs {
    const uint8_t* w;
    const uint8_t* w2;       // fused MLP: a second matrix of the same shape; its feature tiles follow w's in the item sequence
    TiledGeom G;
    int M;
    float* part;             // [gridDim.x * max_seg][NTOK][128] fp32 partial tiles
    int max_seg;
    int pdl;
    int dbg;                 // development (LOWBIT_SK_DBG): 1 no MMAs, 2 no expansion, 4 no activation loads, 8 no reduction
};

template <int QT, int NTOK> struct SkSmem {
    static constexpr int kSuper = Smem<QT>::kSuper;
    static constexpr int kAux = Expand<QT>::kHasMin ? 512 : 256;                // scales (+ minima) of the two k-blocks
    static constexpr int kGroup = 2 * kSuper + kAux;                            // one 64-row group, two k-blocks
    static constexpr int kX = NTOK * 256;                                       // two [NTOK][64] atoms
    static constexpr int kStage = ((kX + 2 * kGroup + 1023) / 1024) * 1024;     // the x atoms sit first (1024-byte aligned)
    static constexpr int kA = 2 * 2 * kTileBytesAB;                             // two groups x two atoms
    static constexpr int kFree = 232448 - 1024 - kA - 512 - Expand<QT>::kLutBytes;
    static constexpr int kFit = kFree / kStage < kSkStagesMax ? kFree / kStage : kSkStagesMax;
    static constexpr int kStages = kFit & ~1;                                   // even: a stage always serves the same group
    static constexpr int kTotal = 1024 + kA + kStages * kStage + 512 + Expand<QT>::kLutBytes;
};

template <int N> __device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, float (&v)[N]);
template <> __device__ __forceinline__ void tmem_ld_cols<32>(uint32_t taddr, float (&v)[32]) { tmem_ld32(taddr, v); }
template <> __device__ __forceinline__ void tmem_ld_cols<16>(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}

template <int QT, bool BF, int NTOK>
__global__ void __launch_bounds__(kSkThreads, 1)
skinny_kernel(const __grid_constant__ CUtensorMap xmap, const SkinnyArgs A) {
    using E = Expand<QT>;
    using SM = SkSmem<QT, NTOK>;
    constexpr int NS = SM::kStages, STAGE = SM::kStage;
    static_assert(NS >= 2 && (NS & 1) == 0, "ring depth");
    constexpr uint32_t ACC_COLS = NTOK, TMEM_COLS = 2 * NTOK < 32 ? 32 : 2 * NTOK;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t atiles = base, ring = base + SM::kA, bars = ring + NS * STAGE;
    auto full_bar = [&](int s) { return bars + s * 8; };
    auto done_bar = [&](int s) { return bars + (kSkStagesMax + s) * 8; };
    auto afull_bar = [&](int q) { return bars + (2 * kSkStagesMax + q) * 8; };
    auto tmem_full = [&](int b) { return bars + (2 * kSkStagesMax + 2 + b) * 8; };
    auto tmem_empty = [&](int b) { return bars + (2 * kSkStagesMax + 4 + b) * 8; };
    const uint32_t tmem_slot = bars + (2 * kSkStagesMax + 6) * 8;

    // the warp index through a shuffle: the compiler then treats everything derived from it as warp-uniform
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    if (A.pdl) asm volatile("griddepcontrol.launch_dependents;");
    const uint32_t lut = bars + 512 + 4 * lane;
    if constexpr (E::kLutBytes > 0) E::fill_lut(smem_raw + (bars + 512 - smem_u32(smem_raw)), threadIdx.x, kSkThreads);
    const int nb = (int)A.G.nb, np = nb >> 1;                // k-blocks / pairs of k-blocks per tile (nb is even)
    const int n_ft1 = (int)(A.G.rows / kTileN);                    // feature tiles of one matrix
    const int64_t total = (int64_t)(A.w2 ? 2 * n_ft1 : n_ft1) * np;
    const int64_t i_begin = (int64_t)blockIdx.x * total / gridDim.x, i_end = (int64_t)(blockIdx.x + 1) * total / gridDim.x;
    const int n_items = (int)(i_end - i_begin);
    const int ft0 = (int)(i_begin / np), p_first = (int)(i_begin - (int64_t)ft0 * np);

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; s++) { mbar_init(full_bar(s), 1); mbar_init(done_bar(s), 1); }
        for (int q = 0; q < 2; q++) mbar_init(afull_bar(q), 32 * (kSkDeqWarps / 2));
        for (int b = 0; b < 2; b++) { mbar_init(tmem_full(b), 1); mbar_init(tmem_empty(b), 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp == 0) {
        // ================= producer: the whole warp runs the loop, one elected lane issues the copies =================
        {
            const uint32_t leader = elect_leader();
            auto locate = [&](int it, int& ft, int& kb) {
                int p = p_first + it;
                ft = ft0;
                while (p >= np) { p -= np; ft++; }
                kb = 2 * p;
            };
            auto weights = [&](int it) {               // packed codes / scales / minima of both k-blocks into the stage
                int ft, kb;
                locate(it, ft, kb);
                const int s = it % NS;
                const uint32_t pk = ring + s * STAGE + SM::kX;
                mbar_expect_tx_w(full_bar(s), ((A.dbg & 4) ? 0 : SM::kX) + 2 * SM::kGroup, leader);
#pragma unroll
                for (int gi = 0; gi < 2; gi++) {
                    const uint8_t* wb = ft < n_ft1 ? A.w : A.w2;
                    const int64_t u = ((int64_t)(ft < n_ft1 ? ft : ft - n_ft1) * 2 + gi) * nb + kb;   // units kb and kb + 1 are adjacent in memory
                    const uint32_t dst = pk + gi * SM::kGroup;
                    bulk_g2s_w(dst, wb + u * SM::kSuper, 2 * SM::kSuper, full_bar(s), leader);
                    bulk_g2s_w(dst + 2 * SM::kSuper, wb + A.G.off_s + u * 128, 256, full_bar(s), leader);
                    if (E::kHasMin) bulk_g2s_w(dst + 2 * SM::kSuper + 256, wb + A.G.off_m + u * 128, 256, full_bar(s), leader);
                }
            };
            auto acts = [&](int it) {                  // the k-slices of the activations: 2 x [NTOK tokens] x [64]
                if (A.dbg & 4) return;
                int ft, kb;
                locate(it, ft, kb);
                const int s = it % NS;
                tma_load_2d_w(ring + s * STAGE, &xmap, kb * 64, 0, full_bar(s), leader);
                tma_load_2d_w(ring + s * STAGE + NTOK * 128, &xmap, kb * 64 + 64, 0, full_bar(s), leader);
            };
            // the first round's weights do not depend on the previous kernel: they are on their way before the wait
            const int first = n_items < NS ? n_items : NS;
            for (int it = 0; it < first; it++) weights(it);
            if (A.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
            for (int it = 0; it < first; it++) acts(it);
            for (int it = first; it < n_items; it++) {
                const int s = it % NS;
                mbar_wait(done_bar(s), (uint32_t)((it / NS - 1) & 1));
                SKTS(0, it);
                weights(it);
                acts(it);
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer: the whole warp runs the loop, one elected lane issues =================
        {
            const uint32_t leader = elect_leader();
            // c F32 (1<<4), a/b format (0 f16, 1 bf16), both K-major, N = NTOK (>>3 at bit 17), M = 128 (>>4 at bit 24)
            const uint32_t idesc = (1u << 4) | ((BF ? 1u : 0u) << 7) | ((BF ? 1u : 0u) << 10) | (((uint32_t)NTOK >> 3) << 17) | ((128u >> 4) << 24);
            int p = p_first, seg = 0;
            bool fresh = true;
            if (n_items > 0) {
                mbar_wait(tmem_empty(0), 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            for (int it = 0; it < n_items; it++) {
                const int s = it % NS, q = it & 1;
                mbar_wait(full_bar(s), (uint32_t)((it / NS) & 1));
                mbar_wait(afull_bar(q), (uint32_t)((it >> 1) & 1));
                SKTS(1, it);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc = tmem_base + (uint32_t)(seg & 1) * ACC_COLS;
                const uint64_t ad = make_desc(atiles + q * (2 * kTileBytesAB)), xd = make_desc(ring + s * STAGE);
                if (!(A.dbg & 1)) {
#pragma unroll
                    for (int kk = 0; kk < 2; kk++)
#pragma unroll
                        for (int k4 = 0; k4 < 4; k4++)       // the address field counts 16-byte units
                            umma_f16_w(acc, ad + (uint64_t)((kk * kTileBytesAB + k4 * 32) >> 4), xd + (uint64_t)((kk * NTOK * 128 + k4 * 32) >> 4),
                                       idesc, (!fresh || kk > 0 || k4 > 0) ? 1u : 0u, leader);
                }
                fresh = false;
                umma_commit_w(done_bar(s), leader);          // frees the stage (producer) and the group's A tile (dequant warps)
                SKTS(2, it);
                if (++p == np || it + 1 == n_items) {        // the tile (or this CTA's part of it) is complete
                    umma_commit_w(tmem_full(seg & 1), leader);
                    p = 0; seg++; fresh = true;
                    if (it + 1 < n_items) {
                        mbar_wait(tmem_empty(seg & 1), (uint32_t)(((seg >> 1) & 1) ^ 1));
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    }
                }
            }
        }
    } else if (warp < 2 + kSkDeqWarps) {
        // ================= dequant warps: group q = (warp - 2) / 8 takes the items of its parity =================
        const int q = (warp - 2) >> 3, w8 = (warp - 2) & 7;
        const int dw = w8 & 3, gi = w8 >> 2;                 // tile r = dw of row group gi, both k-blocks
        const int g = lane >> 2, t = lane & 3;
        const int f_lo = 64 * gi + 16 * dw + g;
        const uint32_t atile = atiles + q * (2 * kTileBytesAB);
        const uint32_t row_lo0 = atile + (f_lo >> 3) * 1024 + (f_lo & 7) * 128;
        for (int it = q; it < n_items; it += 2) {
            const int s = it % NS;
            mbar_wait(full_bar(s), (uint32_t)((it / NS) & 1));
            if (it >= 2) mbar_wait(done_bar((it - 2) % NS), (uint32_t)(((it - 2) / NS) & 1));     // the MMAs that read the tile last
            if (w8 == 0) SKTS(3, it);
            const uint32_t grp = ring + s * STAGE + SM::kX + gi * SM::kGroup;
            uint32_t dq[2], mq[2] = {0u, 0u};
#pragma unroll
            for (int kk = 0; kk < 2; kk++) {
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(dq[kk]) : "r"(grp + 2 * SM::kSuper + kk * 128 + g * 16 + dw * 4));
                if (E::kHasMin) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(mq[kk]) : "r"(grp + 2 * SM::kSuper + 256 + kk * 128 + g * 16 + dw * 4));
            }
#pragma unroll
            for (int kk = 0; kk < 2; kk++) {
                if (A.dbg & 2) break;
                const uint32_t codes = grp + kk * SM::kSuper + dw * E::kTileBytes;
                const __half2 d2 = u2h(dq[kk]), m2 = u2h(mq[kk]);
                const __half2 dlo = __low2half2(d2), dhi = __high2half2(d2);
                const __half2 mlo = __low2half2(m2), mhi = __high2half2(m2);
                const uint32_t row_lo = row_lo0 + kk * kTileBytesAB, row_hi = row_lo + 1024;
                typename E::Codes cw;
                E::load(codes, lane, cw);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    uint32_t va[4], vb[4];
                    E::run(cw, 2 * h, dlo, dhi, mlo, mhi, lut, va);
                    E::run(cw, 2 * h + 1, dlo, dhi, mlo, mhi, lut, vb);
                    if (BF) {
#pragma unroll
                        for (int i = 0; i < 4; i++) { va[i] = h2_to_bf2(va[i]); vb[i] = h2_to_bf2(vb[i]); }
                    }
                    const uint32_t chunk = (uint32_t)((2 * t + h) ^ g) * 16;
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row_lo + chunk), "r"(va[0]), "r"(va[2]), "r"(vb[0]), "r"(vb[2]) : "memory");
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row_hi + chunk), "r"(va[1]), "r"(va[3]), "r"(vb[1]), "r"(vb[3]) : "memory");
                }
            }
            if (w8 == 0) SKTS(4, it);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(afull_bar(q));
            if (w8 == 0) SKTS(5, it);
        }
    } else {
        // ================= epilogue warps: TMEM lanes 32 * (warp % 4) .. + 31 = features, columns = tokens =================
        const int quarter = warp & 3;
        const int n_seg = n_items > 0 ? (int)((i_end - 1) / np) - ft0 + 1 : 0;
        if (A.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");      // the workspace may still be read by the previous reduction
        for (int seg = 0; seg < n_seg; seg++) {
            mbar_wait(tmem_full(seg & 1), (uint32_t)((seg >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float* dst = A.part + ((int64_t)blockIdx.x * A.max_seg + seg) * (NTOK * 128) + quarter * 32 + lane;
            constexpr int CH = NTOK < 32 ? 16 : 32;
#pragma unroll 1
            for (int c = 0; c < NTOK; c += CH) {
                float v[CH];
                tmem_ld_cols<CH>(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)((seg & 1) * ACC_COLS + c), v);
#pragma unroll
                for (int j = 0; j < CH; j++)
                    if (c + j < A.M) dst[(c + j) * 128] = v[j];          // a warp writes 128 contiguous bytes per token
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(tmem_empty(seg & 1));
        }
    }
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// y[token][feature] = sum over the CTAs whose range touches the feature's tile, in CTA order, + bias (+ y).
// grid = (feature tiles, ceil(M / 8)); a block adds eight tokens of one tile, thread = feature.
struct SkReduceArgs {
    const float* part; int max_seg, grid, np, ntok, M;
    int64_t total, n_cols;
    void* y; int y_dtype; const void* bias; int accumulate; int pdl;
    int mlp_act, n_ft1;      // mlp_act != 0: y = act(tile ft) * (tile ft + n_ft1), the fused MLP's two matrices (1 SiLU, 2 GELU)
};
__global__ void __launch_bounds__(128) skinny_reduce_kernel(const SkReduceArgs A) {
    if (A.pdl) {
        asm volatile("griddepcontrol.launch_dependents;");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    const int ft = blockIdx.x, f = threadIdx.x;
    // CTA of item i = the largest c with c * total / grid <= i
    auto cta_of = [&](int64_t i) { return (int)(((i + 1) * A.grid + A.total - 1) / A.total - 1); };
    const int t0 = blockIdx.y * 8;
    auto tile_sum = [&](int tile, float (&v)[8]) {       // the partial tiles of `tile`, in CTA order
        const int c_lo = cta_of((int64_t)tile * A.np), c_hi = cta_of((int64_t)(tile + 1) * A.np - 1);
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = 0.0f;
        for (int c = c_lo; c <= c_hi; c++) {
            const int seg = tile - (int)(((int64_t)c * A.total / A.grid) / A.np);
            const float* src = A.part + ((int64_t)c * A.max_seg + seg) * (A.ntok * 128) + f;
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (t0 + j < A.M) v[j] += src[(t0 + j) * 128];
        }
    };
    float v[8];
    tile_sum(ft, v);
    if (A.mlp_act) {                                     // fused MLP: act(gate) * up
        float u[8];
        tile_sum(ft + A.n_ft1, u);
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = apply_act(v[j], A.mlp_act) * u[j];
    }
    const int64_t col = (int64_t)ft * kTileN + f;
    float b = 0.0f;
    if (A.bias) b = A.y_dtype == DT_F32 ? ((const float*)A.bias)[col]
                  : A.y_dtype == DT_F16 ? __half2float(((const __half*)A.bias)[col]) : __bfloat162float(((const __nv_bfloat16*)A.bias)[col]);
#pragma unroll
    for (int j = 0; j < 8; j++) {
        if (t0 + j >= A.M) break;
        const int64_t idx = (int64_t)(t0 + j) * A.n_cols + col;
        const float r = v[j] + b;
        if (A.y_dtype == DT_F32) { float* p = (float*)A.y + idx; *p = A.accumulate ? *p + r : r; }
        else if (A.y_dtype == DT_F16) { __half* p = (__half*)A.y + idx; *p = __float2half_rn(A.accumulate ? __half2float(*p) + r : r); }
        else { __nv_bfloat16* p = (__nv_bfloat16*)A.y + idx; *p = __float2bfloat16_rn(A.accumulate ? __bfloat162float(*p) + r : r); }
    }
}

template <int QT, bool BF, int NTOK>
static int sk_launch_one(const CUtensorMap& map, const SkinnyArgs& A, int grid, cudaStream_t st) {
    constexpr int SMEM = SkSmem<QT, NTOK>::kTotal;
    static bool configured_dev[64] = {};
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    bool& configured = configured_dev[dev_id & 63];
    if (!configured) {
        cudaFuncSetAttribute(skinny_kernel<QT, BF, NTOK>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        configured = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(kSkThreads); cfg.dynamicSmemBytes = SMEM; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = A.pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, skinny_kernel<QT, BF, NTOK>, map, A);
    return 0;
}
template <int QT>
static int sk_launch_qt(bool bf, int ntok, const CUtensorMap& map, const SkinnyArgs& A, int grid, cudaStream_t st) {
#define SK_CASE(N) case N: return bf ? sk_launch_one<QT, true, N>(map, A, grid, st) : sk_launch_one<QT, false, N>(map, A, grid, st);
    switch (ntok) { SK_CASE(16) SK_CASE(32) SK_CASE(64) SK_CASE(128) }
#undef SK_CASE
    return LB_ERR_ARG;
}
static int sk_ntok(int64_t m) { return m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : 128; }
static int sk_grid(int64_t total) { const int sms = sm_count(); return (int)(total < sms ? total : sms); }
static int sk_max_seg(int64_t total, int grid, int64_t np) { return (int)(((total + grid - 1) / grid + np - 1) / np + 1); }
}  // namespace gemmtc

bool gemm_tc_supported(int qt, int64_t n_rows, int64_t k) {
    qt = canon_qtype(qt);
    const bool q = qt == QT_SYM_INT4 || qt == QT_ASYM_INT4 || qt == QT_NF4 || qt == QT_FP8_E4M3 || qt == QT_FP8_E5M2 || qt == QT_FP6;
    return q && n_rows >= 128 && (n_rows % 128) == 0 && k >= 64 && (k % 64) == 0;
}

// K splits for one token tile (m <= 128) whose feature tiles alone leave SMs idle: the split count that minimises a
// cost model in microseconds -- rounds of the grid x k-blocks per unit x 0.36 us (what one CTA's dequant warps need per
// 128 x 64 block), plus, when split, the partial sums' round trip through memory at ~4 TB/s and the second launch.
static int pick_ksplit(int64_t n_tiles, int64_t nb, int64_t m, int64_t n_main) {
    static const int forced = [] { const char* e = getenv("LOWBIT_GEMM_KSPLIT"); return e ? atoi(e) : 0; }();   // development
    const int64_t sms = sm_count();
    const int64_t cap = nb < 16 ? nb : 16;
    if (forced > 0) return (int)(forced < cap ? forced : cap);
    if (n_tiles >= sms) return 1;
    int best = 1; double best_cost = 1e30;
    for (int64_t ks = 1; ks <= cap; ks++) {
        const double rounds = (double)((n_tiles * ks + sms - 1) / sms);
        double cost = rounds * (double)((nb + ks - 1) / ks) * 0.36;
        if (ks > 1) cost += (double)ks * (double)m * (double)n_main * 8.0 / 4.0e6 + 2.0;
        if (cost < best_cost - 1e-9) { best_cost = cost; best = (int)ks; }
    }
    return best;
}

static size_t x_convert_bytes(int64_t k, int64_t m, int x_dtype) { return x_dtype == DT_F32 ? (size_t)((m * k * 2 + 255) / 256 * 256) : 0; }

size_t gemm_tc_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    const int64_t n_main = n_rows & ~int64_t(127);
    const int64_t n_tiles = (n_main / 128) * ((m + 127) / 128);           // as gemm_tc_launch counts them for m <= 128
    const int ks = m <= 128 ? pick_ksplit(n_tiles, k / 64, m, n_main) : 1;
    return x_convert_bytes(k, m, x_dtype) + (ks > 1 ? (size_t)ks * m * n_main * sizeof(float) : 0);
}

// rows [0, n_main) with n_main = n_rows & ~127 are computed here; the caller handles the rest.  w2 / y2 (or NULL): a second
// matrix of the same shape and qtype over the same activations in the same grid (m > 128, no bias, no accumulate).
static int gemm_tc_launch_impl(const void* w, const void* w2, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m,
                               void* y, void* y2, int y_dtype, const void* bias, int accumulate, void* workspace,
                               size_t workspace_bytes, cudaStream_t st, int64_t* n_done) {
    using namespace gemmtc;
    qt = canon_qtype(qt);
    *n_done = 0;
    if (!gemm_tc_supported(qt, n_rows, k)) return LB_ERR_ARG;
    if (w2 && (!y2 || m <= 128 || bias || accumulate || ((uintptr_t)w2 & 15))) return LB_ERR_ARG;
    EncodeTiledFn enc = encode_fn();
    if (!enc) { set_error(LB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver"); return LB_ERR_CUDA; }
    const int64_t n_main = n_rows & ~int64_t(127);
    const void* xa = x;
    bool bf = (x_dtype == DT_BF16);
    if (x_dtype == DT_F32) {
        const size_t need = gemm_tc_workspace_size(n_rows, k, m, x_dtype);
        if (!workspace || workspace_bytes < need) return LB_ERR_WORKSPACE;
        const int64_t n = m * k;
        int sg = (int)((n + 255) / 256); if (sg > sm_count() * 16) sg = sm_count() * 16;
        f32_to_bf16_kernel<<<sg, 256, 0, st>>>((const float*)x, (__nv_bfloat16*)workspace, n);
        note_launch();
        xa = workspace; bf = true;
    }
    if (((uintptr_t)xa & 15) || ((uintptr_t)w & 15)) return LB_ERR_ARG;
    CUtensorMap map;
    const cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)m};
    const cuuint64_t gstride[1] = {(cuuint64_t)k * 2};
    const cuuint32_t box[2] = {64, 128};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&map, bf ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(xa), gdim,
                     gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error(LB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r); return LB_ERR_CUDA; }
    GemmArgs A;
    A.w = (const uint8_t*)w; A.w2 = (const uint8_t*)w2; A.y2 = y2; A.G = make_geom(n_rows, k, qt); A.M = (int)m;
    A.y = y; A.y_dtype = y_dtype; A.bias = bias; A.accumulate = accumulate;
    const int64_t n_mats = w2 ? 2 : 1;
    // token tile: 256 (two accumulator tiles per CTA, half the weight traffic) unless 128 gives fewer
    // rounds of the 1-CTA-per-SM grid over the SMs
    int mt = 1;
    if (m > 128) {
        const int64_t sms = sm_count();
        const int64_t c2 = n_mats * (n_main / 128) * ((m + 255) / 256), c1 = n_mats * (n_main / 128) * ((m + 127) / 128);
        // rounds of the grid x time per k-block of a unit: measured 0.46 us for a 256-token unit (one expansion serves two
        // token sub-tiles) against 0.33 us for a 128-token one
        // a partial last round of 256-token tiles that covers at most half the SMs runs as 128-token units (half0 below)
        const int64_t rem2 = c2 > sms ? c2 % sms : 0;
        const int64_t t2 = (rem2 > 0 && 2 * rem2 <= sms) ? (c2 / sms) * 14 + 10 : (c2 + sms - 1) / sms * 14;
        const int64_t t1 = (c1 + sms - 1) / sms * 10;
        mt = t2 <= t1 ? 2 : 1;
    }
    const int64_t n_tiles = n_mats * (n_main / 128) * ((m + 128 * mt - 1) / (128 * mt));
    A.ksplit = m <= 128 ? pick_ksplit(n_tiles, k / 64, m, n_main) : 1;      // one token tile: few output tiles
    A.part = nullptr;
    if (A.ksplit > 1) {
        const size_t off = x_convert_bytes(k, m, x_dtype);
        if (!workspace || workspace_bytes < off + (size_t)A.ksplit * m * n_main * sizeof(float)) return LB_ERR_WORKSPACE;
        A.part = reinterpret_cast<float*>((uint8_t*)workspace + off);
    }
    int64_t n_units = n_tiles * A.ksplit;
    A.half0 = 0x7fffffff;
    static const bool no_halves = getenv("LOWBIT_GEMM_NO_HALVES") && atoi(getenv("LOWBIT_GEMM_NO_HALVES")) != 0;   // development
    if (mt == 2 && A.ksplit == 1 && n_tiles > sm_count() && !no_halves) {
        const int64_t rem = n_tiles % sm_count();
        if (rem > 0 && 2 * rem <= sm_count()) { A.half0 = (int)(n_tiles - rem); n_units = n_tiles + rem; }
    }
    const dim3 grid((unsigned)(n_units < sm_count() ? n_units : sm_count()), 1, 1);       // persistent CTAs
    switch (qt) {
        case QT_SYM_INT4:  launch_qt<QT_SYM_INT4>(bf, mt, map, A, grid, st); break;
        case QT_NF4:       launch_qt<QT_NF4>(bf, mt, map, A, grid, st); break;
        case QT_ASYM_INT4: launch_qt<QT_ASYM_INT4>(bf, mt, map, A, grid, st); break;
        case QT_FP8_E4M3:  launch_qt<QT_FP8_E4M3>(bf, mt, map, A, grid, st); break;
        case QT_FP8_E5M2:  launch_qt<QT_FP8_E5M2>(bf, mt, map, A, grid, st); break;
        case QT_FP6:       launch_qt<QT_FP6>(bf, mt, map, A, grid, st); break;
        default: return LB_ERR_ARG;
    }
    note_launch();
    if (A.ksplit > 1) {
        const int64_t mn = m * n_main;
        int rg = (int)((mn / 4 + 255) / 256); if (rg > sm_count() * 8) rg = sm_count() * 8;
        reduce_partials_kernel<<<rg, 256, 0, st>>>(A.part, A.ksplit, mn, n_main, y, y_dtype, bias, accumulate);
        note_launch();
    }
    *n_done = n_main;
    return check_launch();
}

int gemm_tc_launch(const void* w, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m, void* y,
                   int y_dtype, const void* bias, int accumulate, void* workspace, size_t workspace_bytes, cudaStream_t st,
                   int64_t* n_done) {
    return gemm_tc_launch_impl(w, nullptr, qt, n_rows, k, x, x_dtype, m, y, nullptr, y_dtype, bias, accumulate, workspace,
                               workspace_bytes, st, n_done);
}

// gate and up of an MLP at prefill sizes: y = x W^T and y2 = x W2^T from ONE grid.  Two launches of 172 tiles each
// (11008 x 4096 at 512 tokens) are two rounds over 148 SMs EACH; 344 tiles are three rounds together.
int gemm_tc_pair_launch(const void* w, const void* w2, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m,
                        void* y, void* y2, int y_dtype, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    int64_t done = 0;
    if (!w2 || !y2 || (n_rows % 128)) return LB_ERR_ARG;
    return gemm_tc_launch_impl(w, w2, qt, n_rows, k, x, x_dtype, m, y, y2, y_dtype, nullptr, 0, workspace, workspace_bytes, st, &done);
}

#ifdef LOWBIT_SK_TIMESTAMPS
extern "C" int lowbit_debug_skinny_timestamps(unsigned long long* out) {
    return (int)cudaMemcpyFromSymbol(out, gemmtc::g_sk_ts, sizeof(gemmtc::g_sk_ts));
}
#endif
// ---- skinny variant (9 <= m <= 128, fp16 / bf16 activations): see gemmtc::skinny_kernel
bool gemm_skinny_supported(int qt, int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    return (x_dtype == DT_F16 || x_dtype == DT_BF16) && m >= 1 && m <= 128 && gemm_tc_supported(qt, n_rows, k) && (n_rows % 128) == 0 &&
           (k % 128) == 0;                                 // an item is a pair of k-blocks
}
size_t gemm_skinny_workspace_size(int64_t n_rows, int64_t k, int64_t m) {
    using namespace gemmtc;
    const int64_t np = k / 128, total = (n_rows / 128) * np;
    const int grid = sk_grid(total);
    return (size_t)grid * sk_max_seg(total, grid, np) * sk_ntok(m) * 128 * sizeof(float);
}
static int skinny_launch_impl(const void* w, const void* w2, int mlp_act, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype,
                              int64_t m, void* y, int y_dtype, const void* bias, int accumulate, void* workspace, size_t workspace_bytes,
                              cudaStream_t st, int pdl);
int gemm_skinny_launch(const void* w, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m, void* y, int y_dtype,
                       const void* bias, int accumulate, void* workspace, size_t workspace_bytes, cudaStream_t st, int pdl) {
    return skinny_launch_impl(w, nullptr, 0, qt, n_rows, k, x, x_dtype, m, y, y_dtype, bias, accumulate, workspace, workspace_bytes, st, pdl);
}
// fused MLP front half for 17..64 rows: y = act(x . Wgate^T) * (x . Wup^T); both matrices in ONE stream-K pass (their feature
// tiles follow each other in the item sequence), the activation and the product in the reduction launch.  act: 1 SiLU, 2 GELU
size_t gemm_skinny_mlp_workspace_size(int64_t n_rows, int64_t k, int64_t m) {
    using namespace gemmtc;
    const int64_t np = k / 128, total = 2 * (n_rows / 128) * np;
    const int grid = sk_grid(total);
    return (size_t)grid * sk_max_seg(total, grid, np) * sk_ntok(m) * 128 * sizeof(float);
}
int gemm_skinny_mlp_launch(const void* w_gate, const void* w_up, int act, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype,
                           int64_t m, void* y, int y_dtype, void* workspace, size_t workspace_bytes, cudaStream_t st, int pdl) {
    if (!w_up || act < 1 || act > 2 || ((uintptr_t)w_up & 15)) return LB_ERR_ARG;
    return skinny_launch_impl(w_gate, w_up, act, qt, n_rows, k, x, x_dtype, m, y, y_dtype, nullptr, 0, workspace, workspace_bytes, st, pdl);
}
static int skinny_launch_impl(const void* w, const void* w2, int mlp_act, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype,
                              int64_t m, void* y, int y_dtype, const void* bias, int accumulate, void* workspace, size_t workspace_bytes,
                              cudaStream_t st, int pdl) {
    using namespace gemmtc;
    qt = canon_qtype(qt);
    if (!gemm_skinny_supported(qt, n_rows, k, m, x_dtype)) return LB_ERR_ARG;
    EncodeTiledFn enc = encode_fn();
    if (!enc) { set_error(LB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver"); return LB_ERR_CUDA; }
    if (((uintptr_t)x & 15) || ((uintptr_t)w & 15)) return LB_ERR_ARG;
    if (!workspace || workspace_bytes < (w2 ? gemm_skinny_mlp_workspace_size(n_rows, k, m) : gemm_skinny_workspace_size(n_rows, k, m))) return LB_ERR_WORKSPACE;
    const bool bf = x_dtype == DT_BF16;
    const int ntok = sk_ntok(m);
    CUtensorMap map;
    const cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)m};
    const cuuint64_t gstride[1] = {(cuuint64_t)k * 2};
    const cuuint32_t box[2] = {64, (cuuint32_t)ntok};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&map, bf ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(x), gdim,
                     gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error(LB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r); return LB_ERR_CUDA; }
    SkinnyArgs A;
    A.w = (const uint8_t*)w; A.w2 = (const uint8_t*)w2; A.G = make_geom(n_rows, k, qt); A.M = (int)m;
    const int64_t np = k / 128, total = (w2 ? 2 : 1) * (n_rows / 128) * np;
    const int grid = sk_grid(total);
    A.part = (float*)workspace; A.max_seg = sk_max_seg(total, grid, np); A.pdl = pdl;
    { static const int dbg = [] { const char* e = getenv("LOWBIT_SK_DBG"); return e ? atoi(e) : 0; }(); A.dbg = dbg; }
    int rc;
    switch (qt) {
        case QT_SYM_INT4:  rc = sk_launch_qt<QT_SYM_INT4>(bf, ntok, map, A, grid, st); break;
        case QT_NF4:       rc = sk_launch_qt<QT_NF4>(bf, ntok, map, A, grid, st); break;
        case QT_ASYM_INT4: rc = sk_launch_qt<QT_ASYM_INT4>(bf, ntok, map, A, grid, st); break;
        case QT_FP8_E4M3:  rc = sk_launch_qt<QT_FP8_E4M3>(bf, ntok, map, A, grid, st); break;
        case QT_FP8_E5M2:  rc = sk_launch_qt<QT_FP8_E5M2>(bf, ntok, map, A, grid, st); break;
        case QT_FP6:       rc = sk_launch_qt<QT_FP6>(bf, ntok, map, A, grid, st); break;
        default: return LB_ERR_ARG;
    }
    if (rc) return rc;
    note_launch();
    SkReduceArgs R;
    R.part = A.part; R.max_seg = A.max_seg; R.grid = grid; R.np = (int)np; R.ntok = ntok; R.M = (int)m;
    R.total = total; R.n_cols = n_rows; R.y = y; R.y_dtype = y_dtype; R.bias = bias; R.accumulate = accumulate; R.pdl = pdl;
    R.mlp_act = w2 ? mlp_act : 0; R.n_ft1 = (int)(n_rows / 128);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(n_rows / 128), (unsigned)((m + 7) / 8)); cfg.blockDim = dim3(128); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (!(A.dbg & 8)) cudaLaunchKernelEx(&cfg, skinny_reduce_kernel, R);
    note_launch();
    return check_launch();
}

}  // namespace lowbit
