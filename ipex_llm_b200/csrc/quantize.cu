// quantize.cu -- kernel (a): fp32 -> canonical low-bit block records on the GPU.
//
// Replaces ggml_quantize_tensor (reference llama_cpp.py:946-967, called from
// low_bit_linear.py:180).  One warp quantises one 64-element block: lane l owns elements l
// and l+32, which are exactly the two codes packed into byte l of a 4-bit record
// (convert.py:421-426).  All arithmetic that decides a code is done with explicit
// __fmul_rn/__fadd_rn/__fdiv_rn so that nvcc cannot contract it into FMAs: the result must
// be bit-identical to the sequential CPU statement of the same formulas.
//   sym_int4 : d = signed-absmax / -8 (first max wins), code = min(15, (int8)(x*(1/d) + 8.5))
//   asym_int4: d = (max-min)/15, m = min,               code = min(15, (int8)((x-min)*(1/d) + 0.5))
//   nf4      : d = absmax, code = nearest NormalFloat4 entry (ties to the lower index)
//   fp8/fp6  : d = absmax / MAX, code = RNE-saturating e4m3 / e5m2 / e3m2 of x*(1/d);
//              scale_search tries 9 multipliers of d and keeps the least squared error.
// 4 bytes read per element, ~0.53-1.03 written.  The 4-bit formats take four lanes per block (quantize4_kernel: 0.73 of the
// copy peak for sym_int4); the others, and every call that wants the code histogram, one warp per block with 8 consecutive
// blocks per pass (all loads first, redux.sync reductions, records staged in shared memory, 16-byte stores): quantize_kernel.
#include <cuda_fp8.h>
#include <cuda_fp6.h>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {

// fp32 -> fp16 (RNE) bit pattern, little-endian bytes (records are only byte-aligned)
__device__ __forceinline__ void store_half_bits(uint8_t* p, float v) {
    // via a packed half2 in a 32-bit register: ptxas 12.9 mis-fuses cvt.rn.f16.f32 + st.u8 of the
    // 16-bit result into F2I.U8.F16 (a numeric half->u8 conversion), which zeroes the low byte
    const __half2 h2 = __floats2half2_rn(v, 0.0f);
    const unsigned int bits = *reinterpret_cast<const unsigned int*>(&h2);
    p[0] = (uint8_t)(bits & 0xffu);
    p[1] = (uint8_t)(bits >> 8);
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// (key value, index) reductions that reproduce "first occurrence wins" of a serial scan
__device__ __forceinline__ void warp_argbest(float& key, int& idx, float& val, bool want_max) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const float ok = __shfl_xor_sync(0xffffffffu, key, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        const float ov = __shfl_xor_sync(0xffffffffu, val, o);
        const bool better = want_max ? (ok > key) : (ok < key);
        if (better || (ok == key && oi < idx)) { key = ok; idx = oi; val = ov; }
    }
}

template <int QT> __device__ __forceinline__ float fmt_max() {
    return QT == QT_FP8_E4M3 ? 448.0f : QT == QT_FP8_E5M2 ? 57344.0f : 28.0f;
}

template <int QT> __device__ __forceinline__ uint32_t mini_encode(float v) {
    if (QT == QT_FP8_E4M3) return __nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
    if (QT == QT_FP8_E5M2) return __nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E5M2);
    return __nv_cvt_float_to_fp6(v, __NV_E3M2, cudaRoundNearest);
}
template <int QT> __device__ __forceinline__ float mini_decode(uint32_t c) {
    if (QT == QT_FP8_E4M3) return decode_e4m3(c);
    if (QT == QT_FP8_E5M2) return decode_e5m2(c);
    return decode_e3m2(c);
}

__device__ __forceinline__ uint32_t nf4_encode(float v) {
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 15; i++) {
        const float mid = __fmul_rn(__fadd_rn(kNF4Table[i], kNF4Table[i + 1]), 0.5f);
        c += (v > mid) ? 1u : 0u;
    }
    return c;
}

// nf4_encode through a table of 32 bins of width 1/16 over [-1, 1): the 15 midpoints are more than 0.08 apart, so a bin holds
// at most one.  Entry b = { 16 * (the midpoint inside [b/16 - 1, (b+1)/16 - 1), or +inf), number of midpoints below the
// bin }.  v16 = 16 * v (exact: the caller multiplies by 16 / d's reciprocal, a power-of-two scaling of the same rounded
// product), bin = floor(v16) clamped (NaN -> bin 0 -> code 0, like fifteen false compares), code = below + (v16 > mid16):
// one exact compare decides, the same one the chain makes.  8 instructions per element instead of 31.
__device__ __forceinline__ void nf4_bins_init(float2* bins, int b) {           // b = 0..31
    const float lo = (float)(b - 16), hi = (float)(b - 15);
    float m16 = __int_as_float(0x7f800000);
    uint32_t below = 0;
#pragma unroll
    for (int i = 0; i < 15; i++) {
        const float mid16 = __fmul_rn(__fmul_rn(__fadd_rn(kNF4Table[i], kNF4Table[i + 1]), 0.5f), 16.0f);
        below += (mid16 < lo) ? 1u : 0u;
        if (mid16 >= lo && mid16 < hi) m16 = mid16;
    }
    bins[b] = make_float2(m16, __uint_as_float(below));
}
__device__ __forceinline__ uint32_t nf4_encode_bins(float v16, const float2* bins) {
    const float t = fminf(fmaxf(v16, -16.0f), 15.5f);                          // fmaxf drops a NaN
    const float2 e = bins[__float2int_rd(t) + 16];
    return __float_as_uint(e.y) + ((v16 > e.x) ? 1u : 0u);
}

static __device__ __constant__ float kSearchFactors[9] = {1.0f, 0.95f, 1.05f, 0.9f, 1.1f, 0.85f, 1.15f, 0.8f, 1.2f};

constexpr int QUANT_WARPS = 8;
constexpr int QUANT_GROUP = 8;      // consecutive blocks a warp takes at once: 8 records are a multiple of 16 bytes

// |v| as bits (ordered like |v|); a NaN counts as 0: the serial scans' `amax < fabsf(v)` never takes one
__device__ __forceinline__ uint32_t abs_bits(float v) {
    const uint32_t u = __float_as_uint(v) & 0x7fffffffu;
    return u > 0x7f800000u ? 0u : u;
}

// monotonic unsigned key of a float (a < b  <=>  key(a) < key(b); -0 < +0)
__device__ __forceinline__ uint32_t ordered_key(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
// the element (index in 0..63, value) a serial scan "if (x better than best) best = x" over the block ends with, given
// that `hit0` / `hit1` say whether this lane's element l / l + 32 compares EQUAL to the block's best: the first hit
__device__ __forceinline__ float first_hit_value(bool hit0, bool hit1, float x0, float x1) {
    const uint32_t b0 = __ballot_sync(0xffffffffu, hit0), b1 = __ballot_sync(0xffffffffu, hit1);
    const int src = b0 ? (__ffs(b0) - 1) : (__ffs(b1) - 1);
    const float v0 = __shfl_sync(0xffffffffu, x0, src), v1 = __shfl_sync(0xffffffffu, x1, src);
    return b0 ? v0 : v1;
}

// One warp quantises QUANT_GROUP consecutive blocks per pass: all of the group's elements are requested first (2 KB per
// warp in flight: the kernel is HBM-bound and needs the memory-level parallelism), the block extrema come from redux.sync
// (one instruction per reduction) + a ballot that finds the first element equal to the extremum (the serial scan's
// "first wins"), the records are assembled in shared memory and leave as 16-byte stores.
template <int QT>
__global__ void __launch_bounds__(QUANT_WARPS * 32)
quantize_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst, size_t nblk,
                unsigned long long* __restrict__ hist, int scale_search) {
    constexpr int BS = (QT == QT_ASYM_INT4) ? 36 : (QT == QT_FP8_E4M3 || QT == QT_FP8_E5M2) ? 66
                     : (QT == QT_FP6) ? 50 : 34;
    constexpr bool IS4 = (QT == QT_SYM_INT4 || QT == QT_ASYM_INT4 || QT == QT_NF4);
    constexpr bool SEARCH = (QT == QT_FP8_E4M3 || QT == QT_FP6 || QT == QT_FP8_E5M2);
    static_assert((QUANT_GROUP * BS) % 16 == 0, "a group of records must be a whole number of 16-byte pieces");
    __shared__ unsigned int s_hist[16];
    __shared__ float s_err[SEARCH ? QUANT_WARPS : 1][9][64];   // scale search only (fp8 / fp6)
    __shared__ __align__(16) uint8_t s_rec[QUANT_WARPS][QUANT_GROUP * BS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (IS4 && hist) { if (threadIdx.x < 16) s_hist[threadIdx.x] = 0; __syncthreads(); }

    const size_t ngroups = (nblk + QUANT_GROUP - 1) / QUANT_GROUP;
    const size_t wstride = (size_t)gridDim.x * QUANT_WARPS;
    for (size_t grp = (size_t)blockIdx.x * QUANT_WARPS + warp; grp < ngroups; grp += wstride) {
        const size_t b_first = grp * QUANT_GROUP;
        const int nb = (int)((nblk - b_first) < (size_t)QUANT_GROUP ? (nblk - b_first) : (size_t)QUANT_GROUP);
        float xa[QUANT_GROUP], xb[QUANT_GROUP];
#pragma unroll
        for (int j = 0; j < QUANT_GROUP; j++) {
            xa[j] = 0.0f; xb[j] = 0.0f;
            if (j < nb) {
                xa[j] = __ldcs(src + (b_first + j) * 64 + lane);          // streamed once: evict-first
                xb[j] = __ldcs(src + (b_first + j) * 64 + 32 + lane);
            }
        }
#pragma unroll
        for (int j = 0; j < QUANT_GROUP; j++) {
        if (j < nb) {
        const float x0 = xa[j], x1 = xb[j];
        uint8_t* out = s_rec[warp] + j * BS;
        uint32_t q0 = 0, q1 = 0;
        if (QT == QT_SYM_INT4) {
            const uint32_t a0 = abs_bits(x0), a1 = abs_bits(x1);
            const uint32_t amax = __reduce_max_sync(0xffffffffu, max(a0, a1));        // |x| bits order like |x|
            const float val = first_hit_value(a0 == amax, a1 == amax, x0, x1);        // the first element with |x| == max
            const float d = (amax != 0u) ? __fmul_rn(val, -0.125f) : -0.0f;      // x / -8: exact as a product
            const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
            int i0 = (int)(signed char)(int)__fadd_rn(__fmul_rn(x0, id), 8.5f);
            int i1 = (int)(signed char)(int)__fadd_rn(__fmul_rn(x1, id), 8.5f);
            q0 = min(15, i0); q1 = min(15, i1);
            if (lane == 0) store_half_bits(out, d);
            out[2 + lane] = (uint8_t)(q0 | (q1 << 4));
        } else if (QT == QT_ASYM_INT4) {
            const uint32_t k0 = ordered_key(x0), k1 = ordered_key(x1);
            const uint32_t lo = __reduce_min_sync(0xffffffffu, min(k0, k1)), hi = __reduce_max_sync(0xffffffffu, max(k0, k1));
            // back to floats; the serial scan keeps the FIRST element that compares equal to the extremum (-0 == +0)
            const float lo_f = __uint_as_float((lo & 0x80000000u) ? (lo & 0x7fffffffu) : ~lo);
            const float hi_f = __uint_as_float((hi & 0x80000000u) ? (hi & 0x7fffffffu) : ~hi);
            const float mn = first_hit_value(x0 == lo_f, x1 == lo_f, x0, x1);
            const float mx = first_hit_value(x0 == hi_f, x1 == hi_f, x0, x1);
            const float d = __fdiv_rn(__fsub_rn(mx, mn), 15.0f);
            const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
            int i0 = (int)(signed char)(int)__fadd_rn(__fmul_rn(__fsub_rn(x0, mn), id), 0.5f);
            int i1 = (int)(signed char)(int)__fadd_rn(__fmul_rn(__fsub_rn(x1, mn), id), 0.5f);
            q0 = min(15, i0); q1 = min(15, i1);
            if (lane == 0) {
                store_half_bits(out, d); store_half_bits(out + 2, mn);
            }
            out[4 + lane] = (uint8_t)(q0 | (q1 << 4));
        } else if (QT == QT_NF4) {
            const uint32_t a0 = abs_bits(x0), a1 = abs_bits(x1);
            const float d = __uint_as_float(__reduce_max_sync(0xffffffffu, max(a0, a1)));
            const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
            q0 = nf4_encode(__fmul_rn(x0, id)); q1 = nf4_encode(__fmul_rn(x1, id));
            if (lane == 0) store_half_bits(out, d);
            out[2 + lane] = (uint8_t)(q0 | (q1 << 4));
        } else {
            const uint32_t a0 = abs_bits(x0), a1 = abs_bits(x1);
            const float amax = __uint_as_float(__reduce_max_sync(0xffffffffu, max(a0, a1)));
            const float d0 = __fdiv_rn(amax, fmt_max<QT>());
            float d = d0;
            if (scale_search && d0 != 0.0f) {
                // every lane evaluates its two elements for the 9 candidate scales; lanes 0..8
                // then add up one candidate each in element order (the order of the sequential
                // statement), and lane 0 keeps the first strictly-smallest error.
#pragma unroll 1
                for (int c = 0; c < 9; c++) {
                    const float dc = __fmul_rn(d0, kSearchFactors[c]);
                    const float idc = __fdiv_rn(1.0f, dc);
                    const float d16 = __half2float(__float2half_rn(dc));
                    const float e0 = __fsub_rn(__fmul_rn(mini_decode<QT>(mini_encode<QT>(__fmul_rn(x0, idc))), d16), x0);
                    const float e1 = __fsub_rn(__fmul_rn(mini_decode<QT>(mini_encode<QT>(__fmul_rn(x1, idc))), d16), x1);
                    s_err[SEARCH ? warp : 0][c][lane] = __fmul_rn(e0, e0);
                    s_err[SEARCH ? warp : 0][c][lane + 32] = __fmul_rn(e1, e1);
                }
                __syncwarp();
                float err = 0.0f;
                if (lane < 9) for (int e = 0; e < 64; e++) err = __fadd_rn(err, s_err[SEARCH ? warp : 0][lane][e]);
                float best = __int_as_float(0x7f800000); int bc = 0;
                for (int c = 0; c < 9; c++) {
                    const float ec = __shfl_sync(0xffffffffu, err, c);
                    if (ec < best) { best = ec; bc = c; }
                }
                d = __fmul_rn(d0, kSearchFactors[bc]);
                __syncwarp();
            }
            const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
            q0 = mini_encode<QT>(__fmul_rn(x0, id)); q1 = mini_encode<QT>(__fmul_rn(x1, id));
            if (lane == 0) store_half_bits(out, d);
            if (QT == QT_FP6) {
                // 4 consecutive 6-bit codes -> 3 bytes, little-endian (c0 | c1<<6 | c2<<12 | c3<<18)
                const int base = lane & ~3;
                uint32_t w0 = 0, w1 = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    w0 |= (__shfl_sync(0xffffffffu, q0, base + i) & 0x3f) << (6 * i);
                    w1 |= (__shfl_sync(0xffffffffu, q1, base + i) & 0x3f) << (6 * i);
                }
                const int byte = lane & 3, g4 = lane >> 2;
                if (byte < 3) {
                    out[2 + 3 * g4 + byte] = (uint8_t)(w0 >> (8 * byte));
                    out[2 + 3 * (g4 + 8) + byte] = (uint8_t)(w1 >> (8 * byte));
                }
            } else {
                out[2 + lane] = (uint8_t)q0;
                out[2 + 32 + lane] = (uint8_t)q1;
            }
        }
        if (IS4 && hist) { atomicAdd(&s_hist[q0], 1u); atomicAdd(&s_hist[q1], 1u); }
        }
        }
        __syncwarp();
        // the group's records: 16-byte pieces when the group is complete (its offset, 8 * BS * grp, is 16-byte aligned
        // whenever dst is), bytes for the tail group and unaligned destinations
        uint8_t* gdst = dst + b_first * BS;
        if (nb == QUANT_GROUP && !((uintptr_t)gdst & 15)) {
            constexpr int NV = QUANT_GROUP * BS / 16;
            for (int v = lane; v < NV; v += 32)
                __stcs(reinterpret_cast<uint4*>(gdst) + v, reinterpret_cast<const uint4*>(s_rec[warp])[v]);
        } else {
            for (int v = lane; v < nb * BS; v += 32) gdst[v] = s_rec[warp][v];
        }
        __syncwarp();
    }
    if (IS4 && hist) {
        __syncthreads();
        if (threadIdx.x < 16 && s_hist[threadIdx.x]) atomicAdd(&hist[threadIdx.x], (unsigned long long)s_hist[threadIdx.x]);
    }
}

// The 4-bit formats without a histogram request: FOUR lanes per block, eight blocks per warp instruction.  A block's scalar
// work (the extremum's reduction, "first wins", 1 / d, the record header) costs the same warp instructions whether 32 or 4
// lanes share it, so eight blocks per instruction cut the instructions per element from 0.8 (one block per warp) to about
// 0.4 -- the kernel was bound by them (0.48 of the copy peak; 0.68 with eight lanes per block).  Lane l of a block's four
// owns elements 8l..8l+7 and 32+8l..32+8l+7: exactly the code pairs of bytes 8l..8l+7 of the record
// (byte j = code[j] | code[j+32] << 4), loaded as four float4.  "First wins" in element order = the lowest lane with a
// hit among the low halves, else among the high halves (two ballots).
#ifndef LOWBIT_QUANT4_GROUP
#define LOWBIT_QUANT4_GROUP 8
#endif
constexpr int QUANT4_GROUP = LOWBIT_QUANT4_GROUP;
constexpr int QUANT4_LPB = 4;                      // lanes per block
template <int QT>
__global__ void __launch_bounds__(QUANT_WARPS * 32)
quantize4_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst, size_t nblk) {
    constexpr int LPB = QUANT4_LPB, EL = 32 / LPB;          // elements per lane and half (8)
    constexpr int BPP = 32 / LPB;                            // blocks per pass (8)
    constexpr int QG = QUANT4_GROUP, NP = QG / BPP;          // blocks per warp and group; passes
    static_assert(QG % BPP == 0 && NP >= 1, "group = whole passes");
    constexpr int BS = (QT == QT_ASYM_INT4) ? 36 : 34;
    constexpr int HDR = (QT == QT_ASYM_INT4) ? 4 : 2;
    static_assert(QT == QT_SYM_INT4 || QT == QT_ASYM_INT4 || QT == QT_NF4, "4-bit formats");
    __shared__ __align__(16) uint8_t s_rec[QUANT_WARPS][QG * BS];
    __shared__ float2 s_bins[QT == QT_NF4 ? 32 : 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane / LPB, ls = lane % LPB;
    if (QT == QT_NF4) {
        if (threadIdx.x < 32) nf4_bins_init(s_bins, threadIdx.x);
        __syncthreads();
    }
    const size_t ngroups = (nblk + QG - 1) / QG;
    const size_t wstride = (size_t)gridDim.x * QUANT_WARPS;
    for (size_t grp = (size_t)blockIdx.x * QUANT_WARPS + warp; grp < ngroups; grp += wstride) {
        const size_t b_first = grp * QG;
        const int nb = (int)((nblk - b_first) < (size_t)QG ? (nblk - b_first) : (size_t)QG);
        float4 raw[NP][EL / 2];                        // [pass][lo 0..EL/4-1 | hi]
#pragma unroll
        for (int p = 0; p < NP; p++) {                 // all loads first: 64 bytes per lane and pass in flight
            const int j = p * BPP + sub;
#pragma unroll
            for (int v = 0; v < EL / 2; v++) raw[p][v] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (j < nb) {
                const float4* bp = reinterpret_cast<const float4*>(src + (b_first + j) * 64);
#pragma unroll
                for (int v = 0; v < EL / 4; v++) {
                    raw[p][v] = __ldcs(bp + ls * (EL / 4) + v);
                    raw[p][EL / 4 + v] = __ldcs(bp + 8 + ls * (EL / 4) + v);
                }
            }
        }
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const int j = p * BPP + sub;
            float x[2 * EL];                           // [0, EL): elements EL*ls.. ; [EL, 2 EL): 32 + EL*ls..
#pragma unroll
            for (int v = 0; v < EL / 2; v++) { x[4 * v] = raw[p][v].x; x[4 * v + 1] = raw[p][v].y; x[4 * v + 2] = raw[p][v].z; x[4 * v + 3] = raw[p][v].w; }
            uint8_t* out = s_rec[warp] + j * BS;
            uint32_t q[2 * EL];
            float d = 0.0f, mn = 0.0f;
            // the first of the block's 64 elements (index order) for which hit[] holds, given the lane's own flags
            auto first_hit = [&](const bool (&hit)[2 * EL]) -> float {
                bool hl = false, hh = false;
                float vl = x[EL - 1], vh = x[2 * EL - 1];
#pragma unroll
                for (int i = EL - 1; i >= 0; i--) {    // descending: the lowest index is assigned last
                    hl |= hit[i]; hh |= hit[EL + i];
                    vl = hit[i] ? x[i] : vl; vh = hit[EL + i] ? x[EL + i] : vh;
                }
                const uint32_t bl = (__ballot_sync(0xffffffffu, hl) >> (LPB * sub)) & ((1u << LPB) - 1u);
                const uint32_t bh = (__ballot_sync(0xffffffffu, hh) >> (LPB * sub)) & ((1u << LPB) - 1u);
                const int srcl = LPB * sub + (bl ? (__ffs(bl) - 1) : bh ? (__ffs(bh) - 1) : 0);
                return __shfl_sync(0xffffffffu, bl ? vl : vh, srcl);
            };
            if (QT == QT_SYM_INT4) {
                // the signed value of the first element with the largest |x|.  Largest positive and largest negated value
                // (two FMNMX chains, NaNs drop out as in the serial `amax < fabsf(v)` scan): when they differ the winner's
                // sign is known; only a block whose extremes are +a and -a needs the scan for the first of them (rare)
                float pm = x[0], nm = -x[0];
#pragma unroll
                for (int i = 1; i < 2 * EL; i++) { pm = fmaxf(pm, x[i]); nm = fmaxf(nm, -x[i]); }
#pragma unroll
                for (int o = 1; o < LPB; o <<= 1) {
                    pm = fmaxf(pm, __shfl_xor_sync(0xffffffffu, pm, o));
                    nm = fmaxf(nm, __shfl_xor_sync(0xffffffffu, nm, o));
                }
                const float am = fmaxf(pm, nm);
                float val = pm > nm ? pm : -nm;
                if (__any_sync(0xffffffffu, pm == nm && am > 0.0f)) {
                    bool hit[2 * EL];
#pragma unroll
                    for (int i = 0; i < 2 * EL; i++) hit[i] = fabsf(x[i]) == am;
                    const float first = first_hit(hit);
                    if (pm == nm) val = first;
                }
                d = (am > 0.0f) ? __fmul_rn(val, -0.125f) : -0.0f;                         // x / -8: exact as a product
                const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
#pragma unroll
                for (int i = 0; i < 2 * EL; i++) q[i] = (uint32_t)min(15, (int)(signed char)(int)__fadd_rn(__fmul_rn(x[i], id), 8.5f));
            } else if (QT == QT_NF4) {
                d = 0.0f;                                                                  // fmaxf drops NaNs like `a > amax`
#pragma unroll
                for (int i = 0; i < 2 * EL; i++) d = fmaxf(d, fabsf(x[i]));
#pragma unroll
                for (int o = 1; o < LPB; o <<= 1) d = fmaxf(d, __shfl_xor_sync(0xffffffffu, d, o));
                const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
                // 16 * (x * id) == x * (16 * id) bit for bit unless x * id is subnormal (both products then lie deep inside
                // the bin around zero: same code) or 16 * id overflows (d below 2^-124: those warps take two products)
                const float id16 = __fmul_rn(id, 16.0f);
                if (__all_sync(0xffffffffu, id16 < 3.0e38f)) {
#pragma unroll
                    for (int i = 0; i < 2 * EL; i++) q[i] = nf4_encode_bins(__fmul_rn(x[i], id16), s_bins);
                } else {
#pragma unroll
                    for (int i = 0; i < 2 * EL; i++) q[i] = nf4_encode_bins(__fmul_rn(__fmul_rn(x[i], id), 16.0f), s_bins);
                }
            } else {
                // min and max by value (FMNMX chains; NaNs drop out as in the serial `v < min` / `v > max` scans).  The serial
                // scan keeps the FIRST element that compares equal to the extremum, which only matters for a zero extremum
                // (-0 == +0, and the record stores min's sign bit): those blocks take the scan
                float lo_f = x[0], hi_f = x[0];
#pragma unroll
                for (int i = 1; i < 2 * EL; i++) { lo_f = fminf(lo_f, x[i]); hi_f = fmaxf(hi_f, x[i]); }
#pragma unroll
                for (int o = 1; o < LPB; o <<= 1) {
                    lo_f = fminf(lo_f, __shfl_xor_sync(0xffffffffu, lo_f, o));
                    hi_f = fmaxf(hi_f, __shfl_xor_sync(0xffffffffu, hi_f, o));
                }
                mn = lo_f;
                float mx = hi_f;
                if (__any_sync(0xffffffffu, lo_f == 0.0f || hi_f == 0.0f)) {
                    bool hit[2 * EL];
#pragma unroll
                    for (int i = 0; i < 2 * EL; i++) hit[i] = x[i] == lo_f;
                    const float first_lo = first_hit(hit);
#pragma unroll
                    for (int i = 0; i < 2 * EL; i++) hit[i] = x[i] == hi_f;
                    const float first_hi = first_hit(hit);
                    if (lo_f == 0.0f) mn = first_lo;
                    if (hi_f == 0.0f) mx = first_hi;
                }
                d = __fdiv_rn(__fsub_rn(mx, mn), 15.0f);
                const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
#pragma unroll
                for (int i = 0; i < 2 * EL; i++)
                    q[i] = (uint32_t)min(15, (int)(signed char)(int)__fadd_rn(__fmul_rn(__fsub_rn(x[i], mn), id), 0.5f));
            }
            if (j < nb) {
                uint8_t* o8 = out + HDR + EL * ls;     // 34 j + 2 + 8 l (2-byte aligned) / 36 j + 4 + 8 l (4-byte aligned)
#pragma unroll
                for (int w = 0; w < EL / 4; w++) {
                    const int i = 4 * w;
                    const uint32_t word = (q[i] | (q[EL + i] << 4)) | ((q[i + 1] | (q[EL + i + 1] << 4)) << 8) |
                                          ((q[i + 2] | (q[EL + i + 2] << 4)) << 16) | ((q[i + 3] | (q[EL + i + 3] << 4)) << 24);
                    if (QT == QT_ASYM_INT4) *reinterpret_cast<uint32_t*>(o8 + 4 * w) = word;
                    else {
                        *reinterpret_cast<unsigned short*>(o8 + 4 * w) = (unsigned short)(word & 0xffffu);
                        *reinterpret_cast<unsigned short*>(o8 + 4 * w + 2) = (unsigned short)(word >> 16);
                    }
                }
                if (ls == 0) {
                    store_half_bits(out, d);
                    if (QT == QT_ASYM_INT4) store_half_bits(out + 2, mn);
                }
            }
        }
        __syncwarp();
        uint8_t* gdst = dst + b_first * BS;
        if (nb == QG && !((uintptr_t)gdst & 15)) {
            constexpr int NV = QG * BS / 16;
            for (int v = lane; v < NV; v += 32)
                __stcs(reinterpret_cast<uint4*>(gdst) + v, reinterpret_cast<const uint4*>(s_rec[warp])[v]);
        } else {
            for (int v = lane; v < nb * BS; v += 32) gdst[v] = s_rec[warp][v];
        }
        __syncwarp();
    }
}

// fp8 (e4m3 / e5m2) and fp6 without scale search: the same four-lanes-per-block mapping; lane l owns elements 16l..16l+15 =
// bytes 16l..16l+15 of the record (fp8: one code per byte) or 12l..12l+11 (fp6), four float4 loads, d = absmax / MAX.
template <int QT>
__global__ void __launch_bounds__(QUANT_WARPS * 32)
quantize8_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst, size_t nblk) {
    static_assert(QT == QT_FP8_E4M3 || QT == QT_FP8_E5M2 || QT == QT_FP6, "fp8 / fp6 formats");
    constexpr int LPB = 4, BPP = 32 / LPB, QG = 8, BS = (QT == QT_FP6) ? 50 : 66;
    __shared__ __align__(16) uint8_t s_rec[QUANT_WARPS][QG * BS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane / LPB, ls = lane % LPB;
    const size_t ngroups = (nblk + QG - 1) / QG;
    const size_t wstride = (size_t)gridDim.x * QUANT_WARPS;
    for (size_t grp = (size_t)blockIdx.x * QUANT_WARPS + warp; grp < ngroups; grp += wstride) {
        const size_t b_first = grp * QG;
        const int nb = (int)((nblk - b_first) < (size_t)QG ? (nblk - b_first) : (size_t)QG);
        const int j = sub;                             // one pass: eight blocks, one per sub-group
        float4 raw[4];
#pragma unroll
        for (int v = 0; v < 4; v++) raw[v] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (j < nb) {
            const float4* bp = reinterpret_cast<const float4*>(src + (b_first + j) * 64);
#pragma unroll
            for (int v = 0; v < 4; v++) raw[v] = __ldcs(bp + ls * 4 + v);
        }
        float x[16];
#pragma unroll
        for (int v = 0; v < 4; v++) { x[4 * v] = raw[v].x; x[4 * v + 1] = raw[v].y; x[4 * v + 2] = raw[v].z; x[4 * v + 3] = raw[v].w; }
        float amax = 0.0f;                                                                 // fmaxf drops NaNs like `a > amax`
#pragma unroll
        for (int i = 0; i < 16; i++) amax = fmaxf(amax, fabsf(x[i]));
#pragma unroll
        for (int o = 1; o < LPB; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float d = __fdiv_rn(amax, fmt_max<QT>());
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        if (j < nb) {
            uint8_t* out = s_rec[warp] + j * BS;
            if (QT == QT_FP6) {
                // elements 16l..16l+15 = code groups 4l..4l+3 (4 codes -> 3 bytes, c0 | c1<<6 | c2<<12 | c3<<18) = bytes
                // 12l..12l+11 of the payload; 50 j + 2 + 12 l is even: six 2-byte stores
                uint8_t* o12 = out + 2 + 12 * ls;
                uint32_t g[4];
#pragma unroll
                for (int v = 0; v < 4; v++) {
                    g[v] = 0;
#pragma unroll
                    for (int i = 0; i < 4; i++) g[v] |= (mini_encode<QT>(__fmul_rn(x[4 * v + i], id)) & 0x3f) << (6 * i);
                }
                const uint32_t w0 = g[0] | (g[1] << 24), w1 = (g[1] >> 8) | (g[2] << 16), w2 = (g[2] >> 16) | (g[3] << 8);
                unsigned short* o = reinterpret_cast<unsigned short*>(o12);
                o[0] = (unsigned short)w0; o[1] = (unsigned short)(w0 >> 16);
                o[2] = (unsigned short)w1; o[3] = (unsigned short)(w1 >> 16);
                o[4] = (unsigned short)w2; o[5] = (unsigned short)(w2 >> 16);
            } else {
                uint8_t* o16 = out + 2 + 16 * ls;      // 66 j + 2 + 16 l: 2-byte aligned
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    const uint32_t c0 = mini_encode<QT>(__fmul_rn(x[i], id)), c1 = mini_encode<QT>(__fmul_rn(x[i + 1], id));
                    *reinterpret_cast<unsigned short*>(o16 + i) = (unsigned short)(c0 | (c1 << 8));
                }
            }
            if (ls == 0) store_half_bits(out, d);
        }
        __syncwarp();
        uint8_t* gdst = dst + b_first * BS;
        if (nb == QG && !((uintptr_t)gdst & 15)) {
            constexpr int NV = QG * BS / 16;
            for (int v = lane; v < NV; v += 32)
                __stcs(reinterpret_cast<uint4*>(gdst) + v, reinterpret_cast<const uint4*>(s_rec[warp])[v]);
        } else {
            for (int v = lane; v < nb * BS; v += 32) gdst[v] = s_rec[warp][v];
        }
        __syncwarp();
    }
}

template <typename K>
static int resident_ctas(K kernel, int& cache) {
    if (!cache) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, QUANT_WARPS * 32, 0) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 4; }
        cache = n;
    }
    return cache;
}

int quantize_launch(const float* src, void* dst, int qt, size_t n, int k, int64_t* hist, bool scale_search,
                    cudaStream_t st) {
    qt = canon_qtype(qt);
    if (!supported(qt) || k <= 0 || (k % QK) || (n % QK)) return LB_ERR_ARG;
    const size_t nblk = n / QK;
    if (nblk == 0) return 0;
    if (hist) {
        if (cudaMemsetAsync(hist, 0, 16 * sizeof(int64_t), st) != cudaSuccess) return check_launch();
    }
    size_t blocks = ((nblk + QUANT_GROUP - 1) / QUANT_GROUP + QUANT_WARPS - 1) / QUANT_WARPS;
    const size_t blocks4 = ((nblk + QUANT4_GROUP - 1) / QUANT4_GROUP + QUANT_WARPS - 1) / QUANT_WARPS;
    if (!scale_search && !((uintptr_t)src & 15) && (qt == QT_FP8_E4M3 || qt == QT_FP8_E5M2 || qt == QT_FP6)) {
        static int c8[3] = {0, 0, 0};
        const int per = qt == QT_FP8_E4M3 ? resident_ctas(quantize8_kernel<QT_FP8_E4M3>, c8[0])
                      : qt == QT_FP8_E5M2 ? resident_ctas(quantize8_kernel<QT_FP8_E5M2>, c8[1]) : resident_ctas(quantize8_kernel<QT_FP6>, c8[2]);
        const size_t cap8 = (size_t)sm_count() * per;
        const int grid8 = (int)(blocks > cap8 ? cap8 : blocks);
        if (qt == QT_FP8_E4M3) quantize8_kernel<QT_FP8_E4M3><<<grid8, QUANT_WARPS * 32, 0, st>>>(src, (uint8_t*)dst, nblk);
        else if (qt == QT_FP8_E5M2) quantize8_kernel<QT_FP8_E5M2><<<grid8, QUANT_WARPS * 32, 0, st>>>(src, (uint8_t*)dst, nblk);
        else quantize8_kernel<QT_FP6><<<grid8, QUANT_WARPS * 32, 0, st>>>(src, (uint8_t*)dst, nblk);
        note_launch();
        return check_launch();
    }
    if (!hist && !((uintptr_t)src & 15) && (qt == QT_SYM_INT4 || qt == QT_ASYM_INT4 || qt == QT_NF4)) {
        // eight lanes per block (see quantize4_kernel); one resident wave of CTAs, each looping over its groups
        static int c4[3] = {0, 0, 0};
        const int per = qt == QT_SYM_INT4 ? resident_ctas(quantize4_kernel<QT_SYM_INT4>, c4[0])
                      : qt == QT_ASYM_INT4 ? resident_ctas(quantize4_kernel<QT_ASYM_INT4>, c4[1]) : resident_ctas(quantize4_kernel<QT_NF4>, c4[2]);
        const size_t cap4 = (size_t)sm_count() * per;
        const int grid4 = (int)(blocks4 > cap4 ? cap4 : blocks4);
        if (qt == QT_SYM_INT4) quantize4_kernel<QT_SYM_INT4><<<grid4, QUANT_WARPS * 32, 0, st>>>(src, (uint8_t*)dst, nblk);
        else if (qt == QT_ASYM_INT4) quantize4_kernel<QT_ASYM_INT4><<<grid4, QUANT_WARPS * 32, 0, st>>>(src, (uint8_t*)dst, nblk);
        else quantize4_kernel<QT_NF4><<<grid4, QUANT_WARPS * 32, 0, st>>>(src, (uint8_t*)dst, nblk);
        note_launch();
        return check_launch();
    }
    // one wave: as many CTAs as are resident at once (the kernel loops over its groups)
    static int per_sm[6] = {0, 0, 0, 0, 0, 0};
    const int qi = qt == QT_SYM_INT4 ? 0 : qt == QT_ASYM_INT4 ? 1 : qt == QT_NF4 ? 2 : qt == QT_FP8_E4M3 ? 3 : qt == QT_FP8_E5M2 ? 4 : 5;
    if (!per_sm[qi]) {
        int n = 0;
        cudaError_t e = cudaErrorUnknown;
        switch (qt) {
            case QT_SYM_INT4:  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, quantize_kernel<QT_SYM_INT4>, QUANT_WARPS * 32, 0); break;
            case QT_ASYM_INT4: e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, quantize_kernel<QT_ASYM_INT4>, QUANT_WARPS * 32, 0); break;
            case QT_NF4:       e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, quantize_kernel<QT_NF4>, QUANT_WARPS * 32, 0); break;
            case QT_FP8_E4M3:  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, quantize_kernel<QT_FP8_E4M3>, QUANT_WARPS * 32, 0); break;
            case QT_FP8_E5M2:  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, quantize_kernel<QT_FP8_E5M2>, QUANT_WARPS * 32, 0); break;
            default:           e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, quantize_kernel<QT_FP6>, QUANT_WARPS * 32, 0); break;
        }
        if (e != cudaSuccess || n <= 0) { cudaGetLastError(); n = 4; }
        per_sm[qi] = n;
    }
    const size_t cap = (size_t)sm_count() * per_sm[qi];
    const int grid = (int)(blocks > cap ? cap : blocks);
    unsigned long long* h = (unsigned long long*)hist;
    uint8_t* o = (uint8_t*)dst;
    const int ss = scale_search ? 1 : 0;
    switch (qt) {
        case QT_SYM_INT4:  quantize_kernel<QT_SYM_INT4><<<grid, QUANT_WARPS * 32, 0, st>>>(src, o, nblk, h, ss); break;
        case QT_ASYM_INT4: quantize_kernel<QT_ASYM_INT4><<<grid, QUANT_WARPS * 32, 0, st>>>(src, o, nblk, h, ss); break;
        case QT_NF4:       quantize_kernel<QT_NF4><<<grid, QUANT_WARPS * 32, 0, st>>>(src, o, nblk, h, ss); break;
        case QT_FP8_E4M3:  quantize_kernel<QT_FP8_E4M3><<<grid, QUANT_WARPS * 32, 0, st>>>(src, o, nblk, h, ss); break;
        case QT_FP8_E5M2:  quantize_kernel<QT_FP8_E5M2><<<grid, QUANT_WARPS * 32, 0, st>>>(src, o, nblk, h, ss); break;
        case QT_FP6:       quantize_kernel<QT_FP6><<<grid, QUANT_WARPS * 32, 0, st>>>(src, o, nblk, h, ss); break;
        default: return LB_ERR_ARG;
    }
    note_launch();
    return check_launch();
}

}  // namespace lowbit
