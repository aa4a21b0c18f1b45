// kvcache.cu -- FP8 key/value cache and the decode attention that reads it.
//
// Reference: DynamicFp8Cache (transformers/kv.py:33-77) keeps K and V as uint8 [batch, kv_heads, len, head_dim]
// (models/utils.py:103-139: init_fp8_kv_cache / append_fp8_kv_cache -> xe_addons.quantize_key_value, restore ->
// xe_addons.dequantize_key_value) and scaled_dot_product_attention (models/common.py:250-306) calls
// xe_addons.sdp_fp8* when key.dtype == uint8.  The xe_addons kernels are closed binaries; the byte format is
// BUILD-DEFINED here: one e5m2 code per element (round-to-nearest-even, saturating to +-57344; no scales --
// the cache tensor has exactly the shape of the fp16 one), i.e. the high byte of the fp16 value after rounding.
//
// Kernels:
//   kv_quantize / kv_dequantize : elementwise, strided 4-D views (the cache is an as_strided view of a larger buffer)
//   sdp_fp8_decode              : q rows <= 32 per (batch, kv head) incl. the GQA group; split-KV over blockIdx.x,
//                                 one lane per key for q.k, one lane per 4 head-dim elements for p.v, online
//                                 softmax in fp32; partial (m, l, o) per split -> sdp_fp8_combine
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <math_constants.h>
#include "../../include/b200_lowbit.h"
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {
namespace {

struct View4 {                 // [b][h][s][d] with d contiguous; strides in elements
    int64_t sb, sh, ss;
};

__device__ __forceinline__ uint8_t to_e5m2(float v) {
    return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E5M2);
}
__device__ __forceinline__ float from_e5m2(uint8_t c) {
    __half_raw hr; hr.x = (unsigned short)((unsigned short)c << 8);
    return __half2float(__half(hr));
}

template <typename T>
__global__ void kv_quantize_kernel(const T* __restrict__ src, View4 vs, uint8_t* __restrict__ dst, View4 vd,
                                   int64_t B, int64_t H, int64_t S, int64_t D) {
    const int64_t total = B * H * S * D;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t d = i % D, s = (i / D) % S, h = (i / (D * S)) % H, b = i / (D * S * H);
        dst[b * vd.sb + h * vd.sh + s * vd.ss + d] = to_e5m2(to_float<T>(src[b * vs.sb + h * vs.sh + s * vs.ss + d]));
    }
}

template <typename T>
__global__ void kv_dequantize_kernel(const uint8_t* __restrict__ src, View4 vs, T* __restrict__ dst, View4 vd,
                                     int64_t B, int64_t H, int64_t S, int64_t D) {
    const int64_t total = B * H * S * D;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t d = i % D, s = (i / D) % S, h = (i / (D * S)) % H, b = i / (D * S * H);
        dst[b * vd.sb + h * vd.sh + s * vd.ss + d] = from_float<T>(from_e5m2(src[b * vs.sb + h * vs.sh + s * vs.ss + d]));
    }
}

// 8 consecutive elements of a row per thread (head_dim % 8 == 0, rows 16- / 8-byte aligned): one 16-byte load (two
// for fp32) and one 8-byte store -- the prefill writes and restores 4 M elements per layer through these kernels
template <typename T>
__global__ void __launch_bounds__(256) kv_quantize_vec_kernel(const T* __restrict__ src, View4 vs, uint8_t* __restrict__ dst, View4 vd,
                                                              int64_t B, int64_t H, int64_t S, int64_t D) {
    const int64_t D8 = D / 8, total = B * H * S * D8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t d = (i % D8) * 8, s = (i / D8) % S, h = (i / (D8 * S)) % H, b = i / (D8 * S * H);
        const T* p = src + b * vs.sb + h * vs.sh + s * vs.ss + d;
        T v[8];
        if (sizeof(T) == 4) { *reinterpret_cast<uint4*>(&v[0]) = *reinterpret_cast<const uint4*>(p); *reinterpret_cast<uint4*>(&v[4]) = *reinterpret_cast<const uint4*>(p + 4); }
        else *reinterpret_cast<uint4*>(&v[0]) = *reinterpret_cast<const uint4*>(p);
        uint32_t w0 = 0, w1 = 0;
#pragma unroll
        for (int e = 0; e < 4; e++) {
            w0 |= (uint32_t)to_e5m2(to_float<T>(v[e])) << (8 * e);
            w1 |= (uint32_t)to_e5m2(to_float<T>(v[4 + e])) << (8 * e);
        }
        *reinterpret_cast<uint2*>(dst + b * vd.sb + h * vd.sh + s * vd.ss + d) = make_uint2(w0, w1);
    }
}
template <typename T>
__global__ void __launch_bounds__(256) kv_dequantize_vec_kernel(const uint8_t* __restrict__ src, View4 vs, T* __restrict__ dst, View4 vd,
                                                                int64_t B, int64_t H, int64_t S, int64_t D) {
    const int64_t D8 = D / 8, total = B * H * S * D8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t d = (i % D8) * 8, s = (i / D8) % S, h = (i / (D8 * S)) % H, b = i / (D8 * S * H);
        const uint2 w = *reinterpret_cast<const uint2*>(src + b * vs.sb + h * vs.sh + s * vs.ss + d);
        T v[8];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            v[e] = from_float<T>(from_e5m2((uint8_t)(w.x >> (8 * e))));
            v[4 + e] = from_float<T>(from_e5m2((uint8_t)(w.y >> (8 * e))));
        }
        T* q = dst + b * vd.sb + h * vd.sh + s * vd.ss + d;
        if (sizeof(T) == 4) { *reinterpret_cast<uint4*>(q) = *reinterpret_cast<const uint4*>(&v[0]); *reinterpret_cast<uint4*>(q + 4) = *reinterpret_cast<const uint4*>(&v[4]); }
        else *reinterpret_cast<uint4*>(q) = *reinterpret_cast<const uint4*>(&v[0]);
    }
}
// both views allow the 8-element path: D % 8 == 0, every stride a multiple of 8 elements, 16-byte aligned bases
static bool kv_vec_ok(const void* fp, int fp_esz, const View4& vf, const void* u8, const View4& vu, int64_t D) {
    return D % 8 == 0 && !((vf.sb | vf.sh | vf.ss | vu.sb | vu.sh | vu.ss) & 7) && !((uintptr_t)fp & 15) && !((uintptr_t)u8 & 7) && fp_esz > 0;
}

int grid_for(int64_t total) {
    int64_t b = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}


// ------------------------------------------------------------------ decode attention over the FP8 cache
constexpr int SDP_WARPS = 4;
constexpr int SDP_RT = 4;                  // most query rows per CTA

struct SdpArgs {
    const void* q; int q_dtype; int64_t q_sb, q_sh, q_ss;        // [B][Hq][S][D]
    const uint8_t* k; int64_t k_sb, k_sh, k_ss;                   // [B][Hkv][L][D] e5m2
    const uint8_t* v; int64_t v_sb, v_sh, v_ss;
    const void* mask; int mask_dtype; int64_t m_sb, m_sh, m_ss;   // additive [B][1|Hq][S][L] or NULL
    int causal;                                                    // row s sees keys <= L - S + s
    float scale_log2e;                                             // scale * log2(e)
    int B, Hq, Hkv, S, L, D;
    int n_split, keys_per_split;
    const int* len_ptr; int len_add;                               // non-NULL: the key count is *len_ptr + len_add (<= L), read on the device
    // fused decode step (sdp_mma.cuh, FUSED): one new token per sequence; q, the new key and the new value come straight
    // from the merged qkv projection's output, the rotary embedding is applied here and the new row is written to the cache
    const void* qkv; int64_t qkv_sb;                               // [B][(Hq + 2 Hkv) * D] of q_dtype
    const void* cos; const void* sin; int cs_dtype; int64_t cs_row; int n_pos;    // [positions][>= D/2]
    uint8_t* kw; uint8_t* vw;                                      // the caches, writable
    int early;                                                     // FUSED: cached rows and the length are complete: prefetch before griddepcontrol.wait
    int rt;                                                        // query rows per CTA (1, 2 or 4)
    float* part;                                                   // [B][Hkv][row tiles][n_split][RT * (D + 2)]
    void* out; int out_dtype;                                      // [B][Hq][S][D] contiguous
    int64_t o_sb, o_sh, o_ss;                                      // prefill kernel only: element strides of out (rows of D contiguous)
};

// key count and keys per split of this launch: the host's values, or -- when the cache length lives in device memory
// (a decode step replayed from a CUDA graph) -- derived from it with the host's split count
__device__ __forceinline__ void sdp_lengths(const SdpArgs& A, int& L, int& keys_per_split) {
    L = A.L; keys_per_split = A.keys_per_split;
    if (A.len_ptr) {
        L = min(max(*A.len_ptr + A.len_add, 0), A.L);
        keys_per_split = ((L + A.n_split - 1) / A.n_split + 31) / 32 * 32;
    }
}

__device__ __forceinline__ float load_as_float(const void* p, int dt, int64_t i) {
    if (dt == DT_F32) return ((const float*)p)[i];
    if (dt == DT_F16) return __half2float(((const __half*)p)[i]);
    return __bfloat162float(((const __nv_bfloat16*)p)[i]);
}

// two e5m2 codes (bytes b0, b1 of `w` selected by sel) -> float2: e5m2 is the high byte of fp16
__device__ __forceinline__ float2 e5m2x2(uint32_t w, uint32_t sel) {
    const uint32_t h = __byte_perm(w, 0u, sel);
    return __half22float2(*reinterpret_cast<const __half2*>(&h));
}

// DPL consecutive e5m2 bytes of a value row -> up to two words, and back to floats
template <int DPL> __device__ __forceinline__ void load_v(const uint8_t* p, uint32_t (&w)[2]) {
    w[0] = 0; w[1] = 0;
    if (DPL == 4) w[0] = *reinterpret_cast<const uint32_t*>(p);
    else if (DPL == 8) { const uint2 t = *reinterpret_cast<const uint2*>(p); w[0] = t.x; w[1] = t.y; }
    else if (DPL == 2) w[0] = *reinterpret_cast<const uint16_t*>(p);
    else if (DPL == 6) {
        const uint16_t* q = reinterpret_cast<const uint16_t*>(p);
        w[0] = (uint32_t)q[0] | ((uint32_t)q[1] << 16); w[1] = q[2];
    } else {
#pragma unroll
        for (int e = 0; e < DPL; e++) w[e >> 2] |= (uint32_t)p[e] << (8 * (e & 3));
    }
}
template <int DPL> __device__ __forceinline__ void unpack_v(const uint32_t (&w)[2], float (&vv)[DPL]) {
#pragma unroll
    for (int e = 0; e < DPL; e += 2) {
        const float2 f = e5m2x2(w[e >> 2], (e & 2) ? 0x3424u : 0x1404u);
        vv[e] = f.x;
        if (e + 1 < DPL) vv[e + 1] = f.y;
    }
}

// DPL = head_dim / 32: elements of the output each lane owns
// RT = query rows per CTA (1, 2 or 4): the GQA group x new tokens, tiled by blockIdx.z
template <int DPL, int RT>
__global__ void __launch_bounds__(SDP_WARPS * 32)
sdp_fp8_decode_kernel(const SdpArgs A) {
    constexpr int D = DPL * 32;
    __shared__ float q_s[RT][D];
    __shared__ float red_m[SDP_WARPS][RT], red_l[SDP_WARPS][RT];
    __shared__ float red_o[SDP_WARPS][RT][D];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int split = blockIdx.x, hkv = blockIdx.y % A.Hkv, b = blockIdx.y / A.Hkv, tile = blockIdx.z;
    const int n_rep = A.Hq / A.Hkv, rows_total = n_rep * A.S;
    const int row0 = tile * RT;
    const int n_rows = min(RT, rows_total - row0);
    // row r of the tile = (q head hkv * n_rep + rr / S, token rr % S) with rr = row0 + r
    for (int i = threadIdx.x; i < RT * D; i += blockDim.x) {
        const int r = i / D, d = i % D;
        float v = 0.0f;
        if (r < n_rows) {
            const int rr = row0 + r, hq = hkv * n_rep + rr / A.S, st = rr % A.S;
            v = load_as_float(A.q, A.q_dtype, b * A.q_sb + hq * A.q_sh + st * A.q_ss + d) * A.scale_log2e;
        }
        q_s[r][d] = v;
    }
    __syncthreads();

    int L, keys_per_split;
    sdp_lengths(A, L, keys_per_split);
    const int key_begin = split * keys_per_split;
    const int key_end = min(L, key_begin + keys_per_split);
    const uint8_t* kb = A.k + b * A.k_sb + hkv * A.k_sh;
    const uint8_t* vb = A.v + b * A.v_sb + hkv * A.v_sh;

    float m[RT], l[RT], acc[RT][DPL];
#pragma unroll
    for (int r = 0; r < RT; r++) {
        m[r] = -CUDART_INF_F; l[r] = 0.0f;
#pragma unroll
        for (int e = 0; e < DPL; e++) acc[r][e] = 0.0f;
    }

    for (int c0 = key_begin + warp * 32; c0 < key_end; c0 += SDP_WARPS * 32) {
        const int key = c0 + lane;
        const bool live = key < key_end;
        // ---- scores: one lane per key
        float sc[RT];
#pragma unroll
        for (int r = 0; r < RT; r++) sc[r] = 0.0f;
        if (live) {
            const uint4* kr = reinterpret_cast<const uint4*>(kb + (int64_t)key * A.k_ss);
#pragma unroll
            for (int i = 0; i < D / 16; i++) {
                const uint4 w = kr[i];
                const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float2 a = e5m2x2(ws[j], 0x1404u), c = e5m2x2(ws[j], 0x3424u);   // bytes (0,1), (2,3)
#pragma unroll
                    for (int r = 0; r < RT; r++) {
                        const float4 qv = *reinterpret_cast<const float4*>(&q_s[r][16 * i + 4 * j]);
                        sc[r] = fmaf(a.x, qv.x, fmaf(a.y, qv.y, fmaf(c.x, qv.z, fmaf(c.y, qv.w, sc[r]))));
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RT; r++) {
            const int rr = row0 + r;
            bool ok = live && r < n_rows;
            if (ok && A.causal) ok = key <= L - A.S + (rr % A.S);
            if (ok && A.mask) {
                const int hq = hkv * n_rep + rr / A.S, st = rr % A.S;
                sc[r] += load_as_float(A.mask, A.mask_dtype, b * A.m_sb + hq * A.m_sh + st * A.m_ss + key) * 1.4426950408889634f;
            }
            if (!ok) sc[r] = -CUDART_INF_F;
        }
        // ---- online softmax: new running maximum per row (the only cross-lane step of the chunk)
        float p[RT];
#pragma unroll
        for (int r = 0; r < RT; r++) {
            float mx = sc[r];
#pragma unroll
            for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            const float m_new = fmaxf(m[r], mx);
            const float corr = (m_new == -CUDART_INF_F) ? 1.0f : exp2f(m[r] - m_new);
            p[r] = (sc[r] == -CUDART_INF_F) ? 0.0f : exp2f(sc[r] - m_new);
            l[r] = l[r] * corr + p[r];                   // per-lane partial sum; lanes are added up at the end
#pragma unroll
            for (int e = 0; e < DPL; e++) acc[r][e] *= corr;
            m[r] = m_new;
        }
        // ---- p . V: one lane per DPL output elements.  The value rows of 8 keys are requested together (a row
        // is one coalesced DPL*32-byte read of the warp) before they are used: a key-by-key loop exposed one
        // memory round trip per key.  Keys beyond the range are clamped to the last one; their p is 0.
#pragma unroll
        for (int j0 = 0; j0 < 32; j0 += 8) {
            uint32_t vw[8][2];
#pragma unroll
            for (int jj = 0; jj < 8; jj++) {
                const int kj = min(c0 + j0 + jj, key_end - 1);
                load_v<DPL>(vb + (int64_t)kj * A.v_ss + lane * DPL, vw[jj]);
            }
#pragma unroll
            for (int jj = 0; jj < 8; jj++) {
                float vv[DPL];
                unpack_v<DPL>(vw[jj], vv);
#pragma unroll
                for (int r = 0; r < RT; r++) {
                    const float pj = __shfl_sync(0xffffffffu, p[r], j0 + jj);
#pragma unroll
                    for (int e = 0; e < DPL; e++) acc[r][e] = fmaf(pj, vv[e], acc[r][e]);
                }
            }
        }
    }
    // ---- combine: lanes (l), then warps (m, l, acc), fixed order
#pragma unroll
    for (int r = 0; r < RT; r++) {
#pragma unroll
        for (int o = 16; o; o >>= 1) l[r] += __shfl_xor_sync(0xffffffffu, l[r], o);
        if (lane == 0) { red_m[warp][r] = m[r]; red_l[warp][r] = l[r]; }
#pragma unroll
        for (int e = 0; e < DPL; e++) red_o[warp][r][lane * DPL + e] = acc[r][e];
    }
    __syncthreads();
    float* part = A.part + ((((int64_t)b * A.Hkv + hkv) * gridDim.z + tile) * A.n_split + split) * (RT * (D + 2));
    for (int i = threadIdx.x; i < RT * D; i += blockDim.x) {
        const int r = i / D, d = i % D;
        float M = red_m[0][r];
#pragma unroll
        for (int w = 1; w < SDP_WARPS; w++) M = fmaxf(M, red_m[w][r]);
        float o = 0.0f, lsum = 0.0f;
#pragma unroll
        for (int w = 0; w < SDP_WARPS; w++) {
            const float f = (red_m[w][r] == -CUDART_INF_F) ? 0.0f : exp2f(red_m[w][r] - M);
            o = fmaf(red_o[w][r][d], f, o);
            lsum = fmaf(red_l[w][r], f, lsum);
        }
        part[r * (D + 2) + 2 + d] = o;
        if (d == 0) { part[r * (D + 2)] = M; part[r * (D + 2) + 1] = lsum; }
    }
}

// one CTA per output row: merge the splits' (m, l, o).  The (m, l) of the splits sit one per lane (two for more than
// 32 splits) and are reduced with shuffles; warp w then takes splits w, w + 4, ..., four at a time with all their loads
// issued before the first use (a split-by-split loop exposed one memory round trip per split), every lane owning
// columns lane, lane + 32, ...; the four warps' sums meet in shared memory.  Fixed order throughout.
constexpr int SDP_COMBINE_WARPS = 4;
__global__ void __launch_bounds__(SDP_COMBINE_WARPS * 32) sdp_fp8_combine_kernel(const SdpArgs A, int n_tiles) {
    __shared__ float red[SDP_COMBINE_WARPS][256];                                 // D <= 256
    const int D = A.D, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = blockIdx.x;
    const int n_rep = A.Hq / A.Hkv;
    const int st = (int)(row % A.S), hq = (int)((row / A.S) % A.Hq), b = (int)(row / ((int64_t)A.S * A.Hq));
    const int hkv = hq / n_rep, rr = (hq % n_rep) * A.S + st, tile = rr / A.rt, r = rr % A.rt;
    const int64_t ps = (int64_t)A.rt * (D + 2);                                   // floats between two splits
    const float* part = A.part + (((int64_t)b * A.Hkv + hkv) * n_tiles + tile) * A.n_split * ps + r * (D + 2);
    asm volatile("griddepcontrol.wait;" ::: "memory");                           // the partial results are complete
    float m2[2], l2[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int sp = lane + 32 * h;
        m2[h] = sp < A.n_split ? part[sp * ps] : -CUDART_INF_F;
        l2[h] = sp < A.n_split ? part[sp * ps + 1] : 0.0f;
    }
    float M = fmaxf(m2[0], m2[1]);
#pragma unroll
    for (int o = 16; o; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
    float f2[2];
#pragma unroll
    for (int h = 0; h < 2; h++) f2[h] = (m2[h] == -CUDART_INF_F) ? 0.0f : exp2f(m2[h] - M);
    float l = fmaf(l2[0], f2[0], l2[1] * f2[1]);
#pragma unroll
    for (int o = 16; o; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; e++) acc[e] = 0.0f;
    for (int sp0 = warp; sp0 < A.n_split; sp0 += 4 * SDP_COMBINE_WARPS) {
        float v[4][8];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float* po = part + (int64_t)min(sp0 + SDP_COMBINE_WARPS * k, A.n_split - 1) * ps + 2 + lane;
#pragma unroll
            for (int e = 0; e < 8; e++)
                if (32 * e < D) v[k][e] = po[32 * e];
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int sp = sp0 + SDP_COMBINE_WARPS * k;
            float f = __shfl_sync(0xffffffffu, sp < 32 ? f2[0] : f2[1], sp & 31);
            if (sp >= A.n_split) f = 0.0f;
#pragma unroll
            for (int e = 0; e < 8; e++)
                if (32 * e < D) acc[e] = fmaf(v[k][e], f, acc[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; e++)
        if (32 * e < D) red[warp][32 * e + lane] = acc[e];
    __syncthreads();
    const float inv = l > 0.0f ? 1.0f / l : 0.0f;                                 // a row with every key masked gives zeros
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float y = red[0][d];
#pragma unroll
        for (int w = 1; w < SDP_COMBINE_WARPS; w++) y += red[w][d];
        y *= inv;
        const int64_t i = row * D + d;
        if (A.out_dtype == DT_F32) ((float*)A.out)[i] = y;
        else if (A.out_dtype == DT_F16) ((__half*)A.out)[i] = __float2half_rn(y);
        else ((__nv_bfloat16*)A.out)[i] = __float2bfloat16_rn(y);
    }
}

#include "sdp_mma.cuh"
#include "sdp_prefill.cuh"

// rows per CTA: the whole GQA group x new tokens if it is 1, 2 or 4 rows (8 on the tensor-core path), else tiles
int sdp_rt(int Hq, int Hkv, int S, bool mma) {
    const int rows = Hq / Hkv * S;
    return rows <= 1 ? 1 : (rows <= 2 ? 2 : (rows <= 4 || !mma ? SDP_RT : SDP_MMA_ROWS));
}
int sdp_tiles(int Hq, int Hkv, int S, bool mma) { const int rt = sdp_rt(Hq, Hkv, S, mma); return (Hq / Hkv * S + rt - 1) / rt; }

// the tensor-core kernel: fp16 / bf16 queries, head_dim 64 or 128 (LOWBIT_SDP_MMA=0 switches it off: development)
bool sdp_use_mma(int q_dtype, int D) {
    static const bool off = [] { const char* e = getenv("LOWBIT_SDP_MMA"); return e && e[0] == '0'; }();
    return !off && q_dtype != DT_F32 && (D == 64 || D == 128);
}

// splits of the key axis, at least 128 keys each.
// CUDA-core kernel: enough CTAs for ~8 per SM (its loads are latency bound).
// Tensor-core kernel: the split count that minimises  waves x (fixed + keys per CTA)  with waves = CTAs / resident
// CTAs rounded up and the fixed cost of a CTA (query fragments, first chunk's latency, merge) worth ~384 keys:
// one full wave of long CTAs beats several waves of short ones.
int sdp_splits(int B, int Hq, int Hkv, int S, int L, int D, bool mma) {
    static const int per_sm = [] { const char* e = getenv("LOWBIT_SDP_CTAS_PER_SM"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 8; }();
    static const int forced = [] { const char* e = getenv("LOWBIT_SDP_SPLITS"); return e ? atoi(e) : 0; }();
    const int base = B * Hkv * sdp_tiles(Hq, Hkv, S, mma);
    int max_by_len = (L + 127) / 128;
    if (max_by_len > 64) max_by_len = 64;
    if (forced > 0) return forced < max_by_len ? forced : max_by_len;
    if (mma) {
        const int64_t cap = (int64_t)sm_count() * (D == 128 ? 2 : 4);
        int best = 1; int64_t best_cost = INT64_MAX;
        for (int n = 1; n <= max_by_len; n++) {
            const int64_t waves = ((int64_t)base * n + cap - 1) / cap;
            const int64_t keys = (((L + n - 1) / n + 127) / 128) * 128;
            const int64_t cost = waves * (384 + keys);
            if (cost < best_cost) { best_cost = cost; best = n; }
        }
        return best;
    }
    int want = (per_sm * sm_count() + base - 1) / base;
    if (want > max_by_len) want = max_by_len;
    return want < 1 ? 1 : want;
}
size_t sdp_workspace(int B, int Hq, int Hkv, int S, int L, int D, bool mma) {
    return (size_t)B * Hkv * sdp_tiles(Hq, Hkv, S, mma) * sdp_splits(B, Hq, Hkv, S, L, D, mma) * sdp_rt(Hq, Hkv, S, mma) * (D + 2) * sizeof(float);
}
}  // namespace
}  // namespace lowbit

// ------------------------------------------------------------------ C ABI (declared in include/b200_lowbit.h)
using namespace lowbit;

static View4 view_of(const int64_t* st) { View4 v; v.sb = st[0]; v.sh = st[1]; v.ss = st[2]; return v; }

extern "C" int lowbit_kv_quantize_cuda(const void* d_src, int src_dtype, const int64_t* src_strides, void* d_dst,
                                       const int64_t* dst_strides, int64_t B, int64_t H, int64_t S, int64_t D, void* stream) {
    reset_launch_count();
    if (!d_src || !d_dst || !src_strides || !dst_strides || B < 0 || H < 0 || S < 0 || D <= 0 || src_dtype < 0 || src_dtype > 2) {
        set_error(LB_ERR_ARG, "lowbit_kv_quantize_cuda: bad argument");
        return LB_ERR_ARG;
    }
    const int64_t total = B * H * S * D;
    if (total == 0) return LB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const View4 vs = view_of(src_strides), vd = view_of(dst_strides);
    if (kv_vec_ok(d_src, src_dtype == DT_F32 ? 4 : 2, vs, d_dst, vd, D)) {
        const int grid = grid_for(total / 8);
        if (src_dtype == DT_F32) kv_quantize_vec_kernel<float><<<grid, 256, 0, st>>>((const float*)d_src, vs, (uint8_t*)d_dst, vd, B, H, S, D);
        else if (src_dtype == DT_F16) kv_quantize_vec_kernel<__half><<<grid, 256, 0, st>>>((const __half*)d_src, vs, (uint8_t*)d_dst, vd, B, H, S, D);
        else kv_quantize_vec_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)d_src, vs, (uint8_t*)d_dst, vd, B, H, S, D);
        note_launch();
        return check_launch();
    }
    const int grid = grid_for(total);
    if (src_dtype == DT_F32) kv_quantize_kernel<float><<<grid, 256, 0, st>>>((const float*)d_src, vs, (uint8_t*)d_dst, vd, B, H, S, D);
    else if (src_dtype == DT_F16) kv_quantize_kernel<__half><<<grid, 256, 0, st>>>((const __half*)d_src, vs, (uint8_t*)d_dst, vd, B, H, S, D);
    else kv_quantize_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)d_src, vs, (uint8_t*)d_dst, vd, B, H, S, D);
    note_launch();
    return check_launch();
}

extern "C" int lowbit_kv_dequantize_cuda(const void* d_src, const int64_t* src_strides, void* d_dst, int dst_dtype,
                                         const int64_t* dst_strides, int64_t B, int64_t H, int64_t S, int64_t D, void* stream) {
    reset_launch_count();
    if (!d_src || !d_dst || !src_strides || !dst_strides || B < 0 || H < 0 || S < 0 || D <= 0 || dst_dtype < 0 || dst_dtype > 2) {
        set_error(LB_ERR_ARG, "lowbit_kv_dequantize_cuda: bad argument");
        return LB_ERR_ARG;
    }
    const int64_t total = B * H * S * D;
    if (total == 0) return LB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const View4 vs = view_of(src_strides), vd = view_of(dst_strides);
    if (kv_vec_ok(d_dst, dst_dtype == DT_F32 ? 4 : 2, vd, d_src, vs, D)) {
        const int grid = grid_for(total / 8);
        if (dst_dtype == DT_F32) kv_dequantize_vec_kernel<float><<<grid, 256, 0, st>>>((const uint8_t*)d_src, vs, (float*)d_dst, vd, B, H, S, D);
        else if (dst_dtype == DT_F16) kv_dequantize_vec_kernel<__half><<<grid, 256, 0, st>>>((const uint8_t*)d_src, vs, (__half*)d_dst, vd, B, H, S, D);
        else kv_dequantize_vec_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const uint8_t*)d_src, vs, (__nv_bfloat16*)d_dst, vd, B, H, S, D);
        note_launch();
        return check_launch();
    }
    const int grid = grid_for(total);
    if (dst_dtype == DT_F32) kv_dequantize_kernel<float><<<grid, 256, 0, st>>>((const uint8_t*)d_src, vs, (float*)d_dst, vd, B, H, S, D);
    else if (dst_dtype == DT_F16) kv_dequantize_kernel<__half><<<grid, 256, 0, st>>>((const uint8_t*)d_src, vs, (__half*)d_dst, vd, B, H, S, D);
    else kv_dequantize_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const uint8_t*)d_src, vs, (__nv_bfloat16*)d_dst, vd, B, H, S, D);
    note_launch();
    return check_launch();
}

extern "C" int lowbit_sdp_fp8_supported(int Hq, int Hkv, int S, int D) {
    return Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && S >= 1 && S <= 32 && (D == 64 || D == 96 || D == 128 || D == 192 || D == 256);
}

extern "C" size_t lowbit_sdp_fp8_workspace_size(int B, int Hq, int Hkv, int S, int L, int D) {
    if (!lowbit_sdp_fp8_supported(Hq, Hkv, S, D) || B <= 0 || L <= 0) return 0;
    // the query dtype is not known here: room for either kernel
    const size_t a = sdp_workspace(B, Hq, Hkv, S, L, D, false), m = sdp_workspace(B, Hq, Hkv, S, L, D, true);
    return a > m ? a : m;
}

static int sdp_launch(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                      const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                      int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq,
                      int Hkv, int S, int L, int D, void* d_out, int out_dtype, void* workspace,
                      size_t workspace_bytes, void* stream, const int* d_len, int len_add) {
    reset_launch_count();
    if (!d_q || !d_k || !d_v || !d_out || !q_strides || !k_strides || !v_strides || (d_mask && !mask_strides) ||
        q_dtype < 0 || q_dtype > 2 || out_dtype < 0 || out_dtype > 2 || (d_mask && (mask_dtype < 0 || mask_dtype > 2)) ||
        B <= 0 || L <= 0 || !lowbit_sdp_fp8_supported(Hq, Hkv, S, D)) {
        set_error(LB_ERR_ARG, "lowbit_sdp_fp8_cuda: bad argument (B %d, Hq %d, Hkv %d, S %d, L %d, D %d)", B, Hq, Hkv, S, L, D);
        return LB_ERR_ARG;
    }
    if (((uintptr_t)d_k & 15) || ((uintptr_t)d_v & 3) || (k_strides[0] & 15) || (k_strides[1] & 15) || (k_strides[2] & 15) ||
        (v_strides[0] & 3) || (v_strides[1] & 3) || (v_strides[2] & 3)) {
        set_error(LB_ERR_ARG, "lowbit_sdp_fp8_cuda: the key cache must be 16-byte aligned (pointer and strides), the value cache 4-byte");
        return LB_ERR_ARG;
    }
    const size_t need = lowbit_sdp_fp8_workspace_size(B, Hq, Hkv, S, L, D);
    if (!workspace || workspace_bytes < need) {
        set_error(LB_ERR_WORKSPACE, "lowbit_sdp_fp8_cuda: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
        return LB_ERR_WORKSPACE;
    }
    SdpArgs A;
    A.q = d_q; A.q_dtype = q_dtype; A.q_sb = q_strides[0]; A.q_sh = q_strides[1]; A.q_ss = q_strides[2];
    A.k = (const uint8_t*)d_k; A.k_sb = k_strides[0]; A.k_sh = k_strides[1]; A.k_ss = k_strides[2];
    A.v = (const uint8_t*)d_v; A.v_sb = v_strides[0]; A.v_sh = v_strides[1]; A.v_ss = v_strides[2];
    A.mask = d_mask; A.mask_dtype = mask_dtype;
    A.m_sb = d_mask ? mask_strides[0] : 0; A.m_sh = d_mask ? mask_strides[1] : 0; A.m_ss = d_mask ? mask_strides[2] : 0;
    A.causal = causal; A.scale_log2e = scale * 1.4426950408889634f;
    A.B = B; A.Hq = Hq; A.Hkv = Hkv; A.S = S; A.L = L; A.D = D;
    A.len_ptr = d_len; A.len_add = len_add;
    // the tensor-core kernel reads value rows in 16-byte (head_dim 128) / 8-byte (64) pieces
    // ... and query rows as 16-byte pieces
    const bool mma = sdp_use_mma(q_dtype, D) && scale > 0.0f && !(((uintptr_t)d_v | (uintptr_t)v_strides[0] | (uintptr_t)v_strides[1] | (uintptr_t)v_strides[2]) & 15) &&
                     !((uintptr_t)d_q & 15) && !((q_strides[0] | q_strides[1] | q_strides[2]) & 7);
    A.n_split = sdp_splits(B, Hq, Hkv, S, L, D, mma);
    A.keys_per_split = ((L + A.n_split - 1) / A.n_split + 31) / 32 * 32;
    A.part = (float*)workspace; A.out = d_out; A.out_dtype = out_dtype;
    const int tiles = sdp_tiles(Hq, Hkv, S, mma);
    const dim3 grid((unsigned)A.n_split, (unsigned)(B * Hkv), (unsigned)tiles);
    cudaStream_t st = (cudaStream_t)stream;
    A.rt = sdp_rt(Hq, Hkv, S, mma);
#define SDP_LAUNCH(DPL) \
    do { \
        if (A.rt == 1) sdp_fp8_decode_kernel<DPL, 1><<<grid, SDP_WARPS * 32, 0, st>>>(A); \
        else if (A.rt == 2) sdp_fp8_decode_kernel<DPL, 2><<<grid, SDP_WARPS * 32, 0, st>>>(A); \
        else sdp_fp8_decode_kernel<DPL, 4><<<grid, SDP_WARPS * 32, 0, st>>>(A); \
    } while (0)
    if (mma) {
        // head_dim 128: chunks staged through a per-warp cp.async ring (LOWBIT_SDP_RING = stages; 0 = loads into
        // registers, the only variant for head_dim 64)
        static const int nst = [] { const char* e = getenv("LOWBIT_SDP_RING"); return e ? atoi(e) : 3; }();
        if (D == 64) sdp_fp8_decode_mma_kernel<64, 4, 0, 0><<<grid, SDP_WARPS * 32, 0, st>>>(A);
        else if (nst == 0) sdp_fp8_decode_mma_kernel<128, 2, 0, 0><<<grid, SDP_WARPS * 32, 0, st>>>(A);
        else {
            auto kern = nst == 2 ? sdp_fp8_decode_mma_kernel<128, 3, 2, 0> : (nst == 4 ? sdp_fp8_decode_mma_kernel<128, 1, 4, 0> : sdp_fp8_decode_mma_kernel<128, 2, 3, 0>);
            const int smem = SDP_WARPS * (nst == 2 ? 2 : (nst == 4 ? 4 : 3)) * 8192;
            static bool configured[64][3] = {};
            int dev = 0; cudaGetDevice(&dev);
            const int vi = nst == 2 ? 0 : (nst == 4 ? 2 : 1);
            if (dev >= 0 && dev < 64 && !configured[dev][vi]) {
                if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
                    set_error(LB_ERR_CUDA, "lowbit_sdp_fp8_cuda: cannot reserve %d bytes of shared memory", smem);
                    return LB_ERR_CUDA;
                }
                configured[dev][vi] = true;
            }
            kern<<<grid, SDP_WARPS * 32, smem, st>>>(A);
        }
    } else switch (D / 32) {
        case 2: SDP_LAUNCH(2); break;
        case 3: SDP_LAUNCH(3); break;
        case 4: SDP_LAUNCH(4); break;
        case 6: SDP_LAUNCH(6); break;
        default: SDP_LAUNCH(8); break;
    }
#undef SDP_LAUNCH
    note_launch();
    // the merge is launched as a programmatic dependent of the kernel above: its CTAs are scheduled while that grid
    // drains and wait (griddepcontrol.wait) until its partial results are complete and visible
    const int64_t out_rows = (int64_t)B * Hq * S;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)out_rows); cfg.blockDim = dim3(SDP_COMBINE_WARPS * 32); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, sdp_fp8_combine_kernel, A, tiles);
    note_launch();
    return check_launch();
}

extern "C" int lowbit_sdp_fp8_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                   const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                                   int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq,
                                   int Hkv, int S, int L, int D, void* d_out, int out_dtype, void* workspace,
                                   size_t workspace_bytes, void* stream) {
    return sdp_launch(d_q, q_dtype, q_strides, d_k, k_strides, d_v, v_strides, d_mask, mask_dtype, mask_strides, causal, scale,
                      B, Hq, Hkv, S, L, D, d_out, out_dtype, workspace, workspace_bytes, stream, nullptr, 0);
}

extern "C" int lowbit_sdp_fp8_devlen_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                          const int64_t* k_strides, const void* d_v, const int64_t* v_strides, int causal,
                                          float scale, int B, int Hq, int Hkv, int S, int L_max, const int32_t* d_len,
                                          int len_add, int D, void* d_out, int out_dtype, void* workspace,
                                          size_t workspace_bytes, void* stream) {
    if (!d_len) { set_error(LB_ERR_ARG, "lowbit_sdp_fp8_devlen_cuda: d_len is NULL"); return LB_ERR_ARG; }
    return sdp_launch(d_q, q_dtype, q_strides, d_k, k_strides, d_v, v_strides, nullptr, 0, nullptr, causal, scale,
                      B, Hq, Hkv, S, L_max, D, d_out, out_dtype, workspace, workspace_bytes, stream, d_len, len_add);
}

// ------------------------------------------------------------------ prefill attention over the e5m2 cache
extern "C" int lowbit_sdp_fp8_prefill_supported(int Hq, int Hkv, int D, int q_dtype) {
    return Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && D == 128 && (q_dtype == DT_F16 || q_dtype == DT_BF16);
}

extern "C" int lowbit_sdp_fp8_prefill_strided_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                                   const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                                                   int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq,
                                                   int Hkv, int S, int L, int D, void* d_out, int out_dtype, const int64_t* out_strides,
                                                   void* stream);
extern "C" int lowbit_sdp_fp8_prefill_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                           const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                                           int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq,
                                           int Hkv, int S, int L, int D, void* d_out, int out_dtype, void* stream) {
    const int64_t os[3] = {(int64_t)Hq * S * D, (int64_t)S * D, (int64_t)D};
    return lowbit_sdp_fp8_prefill_strided_cuda(d_q, q_dtype, q_strides, d_k, k_strides, d_v, v_strides, d_mask, mask_dtype, mask_strides,
                                               causal, scale, B, Hq, Hkv, S, L, D, d_out, out_dtype, os, stream);
}

extern "C" int lowbit_sdp_fp8_prefill_strided_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                                   const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                                                   int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq,
                                                   int Hkv, int S, int L, int D, void* d_out, int out_dtype, const int64_t* out_strides,
                                                   void* stream) {
    reset_launch_count();
    if (!d_q || !d_k || !d_v || !d_out || !q_strides || !k_strides || !v_strides || !out_strides || (d_mask && !mask_strides) || out_dtype < 0 ||
        out_dtype > 2 || (d_mask && (mask_dtype < 0 || mask_dtype > 2)) || B <= 0 || S <= 0 || L < S || !(scale > 0.0f) ||
        !lowbit_sdp_fp8_prefill_supported(Hq, Hkv, D, q_dtype)) {
        set_error(LB_ERR_ARG, "lowbit_sdp_fp8_prefill_cuda: bad argument (B %d, Hq %d, Hkv %d, S %d, L %d, D %d, q_dtype %d)", B, Hq, Hkv, S, L, D, q_dtype);
        return LB_ERR_ARG;
    }
    if ((((uintptr_t)d_k | (uintptr_t)d_v | (uintptr_t)d_q | (uintptr_t)d_out) & 15) ||
        ((k_strides[0] | k_strides[1] | k_strides[2] | v_strides[0] | v_strides[1] | v_strides[2]) & 15) ||
        ((q_strides[0] | q_strides[1] | q_strides[2]) & 7) || ((out_strides[0] | out_strides[1] | out_strides[2]) & 7)) {
        set_error(LB_ERR_ARG, "lowbit_sdp_fp8_prefill_cuda: caches, query rows and output must be 16-byte aligned (pointers and strides)");
        return LB_ERR_ARG;
    }
    SdpArgs A = {};
    A.q = d_q; A.q_dtype = q_dtype; A.q_sb = q_strides[0]; A.q_sh = q_strides[1]; A.q_ss = q_strides[2];
    A.k = (const uint8_t*)d_k; A.k_sb = k_strides[0]; A.k_sh = k_strides[1]; A.k_ss = k_strides[2];
    A.v = (const uint8_t*)d_v; A.v_sb = v_strides[0]; A.v_sh = v_strides[1]; A.v_ss = v_strides[2];
    A.mask = d_mask; A.mask_dtype = mask_dtype;
    A.m_sb = d_mask ? mask_strides[0] : 0; A.m_sh = d_mask ? mask_strides[1] : 0; A.m_ss = d_mask ? mask_strides[2] : 0;
    A.causal = causal; A.scale_log2e = scale * 1.4426950408889634f;
    A.B = B; A.Hq = Hq; A.Hkv = Hkv; A.S = S; A.L = L; A.D = D;
    A.out = d_out; A.out_dtype = out_dtype; A.o_sb = out_strides[0]; A.o_sh = out_strides[1]; A.o_ss = out_strides[2];
    const dim3 grid((unsigned)((S + PF_ROWS - 1) / PF_ROWS), (unsigned)(B * Hq), 1);
    sdp_fp8_prefill_kernel<<<grid, PF_WARPS * 32, 0, (cudaStream_t)stream>>>(A);
    note_launch();
    return check_launch();
}

// ------------------------------------------------------------------ the attention of one decode step, one call
extern "C" int lowbit_rope_kv_append_cuda(void* d_qkv, int dtype, int64_t batch_stride, int64_t token_stride, int B, int S,
                                          int Hq, int Hkv, int D, const void* d_cos, const void* d_sin, int cs_dtype,
                                          int64_t cs_row_stride, int n_positions, const int32_t* d_pos, int pos_add,
                                          void* d_k_cache, const int64_t* k_strides, void* d_v_cache,
                                          const int64_t* v_strides, int capacity, void* stream);

// key splits of the fused kernel: the merge runs inside a thread-block cluster, so at most 8; same cost model as sdp_splits
static int fused_splits(int B, int Hq, int Hkv, int L, int D) {
    const int base = B * Hkv * sdp_tiles(Hq, Hkv, 1, true);
    int max_by_len = (L + 127) / 128;
    if (max_by_len > 8) max_by_len = 8;
    const char* forced = getenv("LOWBIT_DECODE_ATTN_SPLITS");          // development: fixed split count
    if (forced && atoi(forced) > 0) return atoi(forced) < max_by_len ? atoi(forced) : max_by_len;
    const int64_t cap = (int64_t)sm_count() * 2;
    int best = 1; int64_t best_cost = INT64_MAX;
    for (int n = 1; n <= max_by_len; n++) {
        const int64_t waves = ((int64_t)base * n + cap - 1) / cap;
        const int64_t keys = (((L + n - 1) / n + 127) / 128) * 128;
        const int64_t cost = waves * (384 + keys);
        if (cost < best_cost) { best_cost = cost; best = n; }
    }
    return best;
}

extern "C" size_t lowbit_decode_attention_workspace_size(int B, int Hq, int Hkv, int capacity, int D) {
    return lowbit_sdp_fp8_workspace_size(B, Hq, Hkv, 1, capacity, D);
}

extern "C" int lowbit_decode_attention_cuda(void* d_qkv, int dtype, int64_t batch_stride, int B, int Hq, int Hkv, int D,
                                            const void* d_cos, const void* d_sin, int cs_dtype, int64_t cs_row_stride,
                                            int n_positions, const int32_t* d_len, void* d_k_cache, const int64_t* k_strides,
                                            void* d_v_cache, const int64_t* v_strides, int capacity, float scale, void* d_out,
                                            int out_dtype, int cache_complete, void* workspace, size_t workspace_bytes, void* stream) {
    reset_launch_count();
    if (!d_qkv || !d_cos || !d_sin || !d_len || !d_k_cache || !d_v_cache || !k_strides || !v_strides || !d_out || dtype < 0 ||
        dtype > 2 || cs_dtype < 0 || cs_dtype > 2 || out_dtype < 0 || out_dtype > 2 || B <= 0 || capacity <= 0 || n_positions <= 0 ||
        !lowbit_sdp_fp8_supported(Hq, Hkv, 1, D) || batch_stride < (int64_t)(Hq + 2 * Hkv) * D || cs_row_stride < D / 2) {
        set_error(LB_ERR_ARG, "lowbit_decode_attention_cuda: bad argument (B %d, Hq %d, Hkv %d, D %d, capacity %d)", B, Hq, Hkv, D, capacity);
        return LB_ERR_ARG;
    }
    static const bool fused_off = [] { const char* e = getenv("LOWBIT_DECODE_ATTN_FUSED"); return e && e[0] == '0'; }();   // development
    const int esz = 2;
    // the one-launch kernel: half-precision activations, head_dim 128, 16-byte aligned rows everywhere, a split count
    // that fits a cluster.  Long caches with few heads want more than 8 splits: they take the separate kernels below.
    const bool aligned = !(((uintptr_t)d_qkv | (uintptr_t)d_k_cache | (uintptr_t)d_v_cache) & 15) && !((batch_stride * esz) & 15) &&
                         !((k_strides[0] | k_strides[1] | k_strides[2] | v_strides[0] | v_strides[1] | v_strides[2]) & 15) &&
                         (cs_dtype == DT_F32 || (!(((uintptr_t)d_cos | (uintptr_t)d_sin) & 15) && !((cs_row_stride * 2) & 15)));
    const int want = sdp_splits(B, Hq, Hkv, 1, capacity, D, true);
    if (!fused_off && dtype != DT_F32 && D == 128 && scale > 0.0f && aligned && want <= 8) {
        SdpArgs A = {};
        A.q = nullptr; A.q_dtype = dtype;
        A.k = (const uint8_t*)d_k_cache; A.k_sb = k_strides[0]; A.k_sh = k_strides[1]; A.k_ss = k_strides[2];
        A.v = (const uint8_t*)d_v_cache; A.v_sb = v_strides[0]; A.v_sh = v_strides[1]; A.v_ss = v_strides[2];
        A.kw = (uint8_t*)d_k_cache; A.vw = (uint8_t*)d_v_cache;
        A.mask = nullptr; A.causal = 0; A.scale_log2e = scale * 1.4426950408889634f;
        A.B = B; A.Hq = Hq; A.Hkv = Hkv; A.S = 1; A.L = capacity; A.D = D;
        A.len_ptr = d_len; A.len_add = 1; A.early = cache_complete ? 1 : 0;
        A.qkv = d_qkv; A.qkv_sb = batch_stride; A.cos = d_cos; A.sin = d_sin; A.cs_dtype = cs_dtype; A.cs_row = cs_row_stride;
        A.n_pos = n_positions;
        A.n_split = fused_splits(B, Hq, Hkv, capacity, D);
        A.keys_per_split = ((capacity + A.n_split - 1) / A.n_split + 31) / 32 * 32;
        A.rt = sdp_rt(Hq, Hkv, 1, true);
        A.part = nullptr; A.out = d_out; A.out_dtype = out_dtype;
        const int tiles = sdp_tiles(Hq, Hkv, 1, true);
        auto kern = sdp_fp8_decode_mma_kernel<128, 2, 3, 1>;
        const int smem = SDP_WARPS * 3 * 8192 + 16 + (A.n_split - 1) * A.rt * (D + 2) * (int)sizeof(float);
        static int configured[64] = {};
        int dev = 0; cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && configured[dev] < smem) {
            const int ask = SDP_WARPS * 3 * 8192 + 16 + 7 * SDP_MMA_ROWS * (128 + 2) * (int)sizeof(float);
            if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, ask) != cudaSuccess) {
                set_error(LB_ERR_CUDA, "lowbit_decode_attention_cuda: cannot reserve %d bytes of shared memory", ask);
                return LB_ERR_CUDA;
            }
            configured[dev] = ask;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)A.n_split, (unsigned)(B * Hkv), (unsigned)tiles);
        cfg.blockDim = dim3(SDP_WARPS * 32); cfg.dynamicSmemBytes = (size_t)smem; cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute at[2];
        int na = 0;
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = (unsigned)A.n_split; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1; na++;
        if (gemv_get_pdl()) {
            at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[na].val.programmaticStreamSerializationAllowed = 1; na++;
        }
        cfg.attrs = at; cfg.numAttrs = na;
        if (cudaLaunchKernelEx(&cfg, kern, A) != cudaSuccess) return check_launch();
        note_launch();
        return check_launch();
    }
    // separate kernels: rotary + cache write, attention with the device-side length, merge
    int rc = lowbit_rope_kv_append_cuda(d_qkv, dtype, batch_stride, batch_stride, B, 1, Hq, Hkv, D, d_cos, d_sin, cs_dtype, cs_row_stride,
                                        n_positions, d_len, 0, d_k_cache, k_strides, d_v_cache, v_strides, capacity, stream);
    if (rc) return rc;
    const int64_t q_strides[3] = {batch_stride, (int64_t)D, batch_stride};
    rc = sdp_launch(d_qkv, dtype, q_strides, d_k_cache, k_strides, d_v_cache, v_strides, nullptr, 0, nullptr, 0, scale, B, Hq, Hkv, 1,
                    capacity, D, d_out, out_dtype, workspace, workspace_bytes, stream, d_len, 1);
    return rc;
}
