// rope.cu -- in-place rotary position embedding of the query and key heads that the merged qkv projection produced
// (second half: reference models/llama.py:151-165 and 40 other decoders call
// xe_addons.rotary_half_inplaced / rotary_two_inplaced(inv_freq, position_ids, q, k) and
// rotary_half_with_cache_inplaced / rotary_two_with_cache_inplaced(q, k, cos, sin[, half_layout]),
// models/common.py:354-367).
//
// "half" pairs element i with i + R/2 (rotate_half, Llama / Mistral / Qwen), "two" pairs 2i with 2i + 1 (GPT-J,
// ChatGLM); R = rotary width (<= head_dim, elements beyond R are left as they are).  The angle of pair i of a token
// is position * inv_freq[i] (fp32, sincosf with full range reduction) or comes from cos / sin tables.  The rotation
// is computed in fp32 and rounded once to the tensors' dtype:
//     x0' = x0 cos - x1 sin,   x1' = x1 cos + x0 sin.
// One thread per pair; q and k heads form one index space, so one launch covers both tensors.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/b200_lowbit.h"
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {
namespace {

struct RopeArgs {
    void* q; int64_t q_sb, q_sh, q_ss; int Hq;       // [B][Hq][S][D], D contiguous
    void* k; int64_t k_sb, k_sh, k_ss; int Hkv;      // [B][Hkv][S][D] or NULL
    int B, S, D, R;                                   // R = rotary width (even)
    int two;                                          // 0: pairs (i, i + R/2), 1: pairs (2i, 2i + 1)
    const float* inv_freq;                            // [R/2] or NULL
    const int64_t* pos; int64_t p_sb, p_ss;          // position ids [B][S] (p_sb = 0 broadcasts over the batch)
    const void* cos; const void* sin; int cs_dtype;  // tables [B|1][S][>= R/2 ...] or NULL
    int64_t c_sb, c_ss; int cs_step;                  // table index of pair i = cs_step * i (1: half layout, 2: interleaved)
};

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float table_at(const void* p, int dt, int64_t i) {
    if (dt == DT_F32) return ((const float*)p)[i];
    if (dt == DT_F16) return __half2float(((const __half*)p)[i]);
    return __bfloat162float(((const __nv_bfloat16*)p)[i]);
}

template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(const RopeArgs A) {
    const int half = A.R / 2, heads = A.Hq + A.Hkv;
    const int64_t total = (int64_t)A.B * heads * A.S * half;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx % half), s = (int)((idx / half) % A.S), h = (int)((idx / ((int64_t)half * A.S)) % heads);
        const int b = (int)(idx / ((int64_t)half * A.S * heads));
        float c, sn;
        if (A.inv_freq) {
            const float angle = (float)A.pos[b * A.p_sb + s * A.p_ss] * A.inv_freq[i];
            sincosf(angle, &sn, &c);
        } else {
            const int64_t ci = b * A.c_sb + s * A.c_ss + (int64_t)A.cs_step * i;
            c = table_at(A.cos, A.cs_dtype, ci);
            sn = table_at(A.sin, A.cs_dtype, ci);
        }
        T* row = h < A.Hq ? (T*)A.q + b * A.q_sb + h * A.q_sh + s * A.q_ss
                          : (T*)A.k + b * A.k_sb + (h - A.Hq) * A.k_sh + s * A.k_ss;
        const int d0 = A.two ? 2 * i : i, d1 = A.two ? 2 * i + 1 : i + half;
        const float x0 = to_f<T>(row[d0]), x1 = to_f<T>(row[d1]);
        row[d0] = from_f<T>(fmaf(x0, c, -x1 * sn));
        row[d1] = from_f<T>(fmaf(x1, c, x0 * sn));
    }
}
// rotate_half pairing with tables, 8 pairs per thread (R/2 % 8 == 0, 2-byte dtypes for tensors and tables, 16-byte
// aligned rows): the prefill rotates S x (Hq + Hkv) x R/2 pairs per layer; the one-pair-per-thread kernel above spent its
// time in 64-bit index arithmetic and 2-byte accesses
template <typename T>
__global__ void __launch_bounds__(256) rope_half_table_vec_kernel(const RopeArgs A) {
    const int half = A.R / 2, heads = A.Hq + A.Hkv, H8 = half / 8;
    const int64_t total = (int64_t)A.B * heads * A.S * H8;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx % H8) * 8, h = (int)((idx / H8) % heads), s = (int)((idx / ((int64_t)H8 * heads)) % A.S);
        const int b = (int)(idx / ((int64_t)H8 * heads * A.S));                    // heads inner: a token's rows share cos / sin
        const int64_t ci = b * A.c_sb + s * A.c_ss + i;
        T c[8], sn[8], x0[8], x1[8];
        *reinterpret_cast<uint4*>(c) = *reinterpret_cast<const uint4*>((const T*)A.cos + ci);
        *reinterpret_cast<uint4*>(sn) = *reinterpret_cast<const uint4*>((const T*)A.sin + ci);
        T* row = h < A.Hq ? (T*)A.q + b * A.q_sb + h * A.q_sh + s * A.q_ss
                          : (T*)A.k + b * A.k_sb + (h - A.Hq) * A.k_sh + s * A.k_ss;
        *reinterpret_cast<uint4*>(x0) = *reinterpret_cast<const uint4*>(row + i);
        *reinterpret_cast<uint4*>(x1) = *reinterpret_cast<const uint4*>(row + i + half);
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const float a = to_f<T>(x0[e]), d = to_f<T>(x1[e]), cf = to_f<T>(c[e]), sf = to_f<T>(sn[e]);
            x0[e] = from_f<T>(fmaf(a, cf, -d * sf));
            x1[e] = from_f<T>(fmaf(d, cf, a * sf));
        }
        *reinterpret_cast<uint4*>(row + i) = *reinterpret_cast<const uint4*>(x0);
        *reinterpret_cast<uint4*>(row + i + half) = *reinterpret_cast<const uint4*>(x1);
    }
}
}  // namespace
}  // namespace lowbit

using namespace lowbit;

extern "C" int lowbit_rotary_inplaced_cuda(void* d_q, const int64_t* q_strides, int Hq, void* d_k, const int64_t* k_strides,
                                           int Hkv, int dtype, int B, int S, int D, int R, int two, const float* d_inv_freq,
                                           const int64_t* d_position_ids, const int64_t* pos_strides, const void* d_cos,
                                           const void* d_sin, int cs_dtype, const int64_t* cs_strides, int cs_interleaved,
                                           void* stream) {
    reset_launch_count();
    const bool by_freq = d_inv_freq != nullptr;
    if (!d_q || !q_strides || Hq <= 0 || (d_k && (!k_strides || Hkv <= 0)) || dtype < 0 || dtype > 2 || B < 0 || S < 0 || D <= 0 ||
        R <= 0 || R > D || (R & 1) || (by_freq && (!d_position_ids || !pos_strides)) ||
        (!by_freq && (!d_cos || !d_sin || !cs_strides || cs_dtype < 0 || cs_dtype > 2))) {
        set_error(LB_ERR_ARG, "lowbit_rotary_inplaced_cuda: bad argument (Hq %d, Hkv %d, B %d, S %d, D %d, R %d)", Hq, Hkv, B, S, D, R);
        return LB_ERR_ARG;
    }
    RopeArgs A;
    A.q = d_q; A.q_sb = q_strides[0]; A.q_sh = q_strides[1]; A.q_ss = q_strides[2]; A.Hq = Hq;
    A.k = d_k; A.Hkv = d_k ? Hkv : 0;
    A.k_sb = d_k ? k_strides[0] : 0; A.k_sh = d_k ? k_strides[1] : 0; A.k_ss = d_k ? k_strides[2] : 0;
    A.B = B; A.S = S; A.D = D; A.R = R; A.two = two ? 1 : 0;
    A.inv_freq = d_inv_freq; A.pos = d_position_ids;
    A.p_sb = by_freq ? pos_strides[0] : 0; A.p_ss = by_freq ? pos_strides[1] : 0;
    A.cos = d_cos; A.sin = d_sin; A.cs_dtype = cs_dtype;
    A.c_sb = by_freq ? 0 : cs_strides[0]; A.c_ss = by_freq ? 0 : cs_strides[1]; A.cs_step = cs_interleaved ? 2 : 1;
    const int64_t total = (int64_t)B * (A.Hq + A.Hkv) * S * (R / 2);
    if (total == 0) return LB_OK;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = !by_freq && !A.two && A.cs_step == 1 && dtype != DT_F32 && cs_dtype == dtype && (R / 2) % 8 == 0 &&
                     !((A.q_sb | A.q_sh | A.q_ss | A.k_sb | A.k_sh | A.k_ss | A.c_sb | A.c_ss) & 7) &&
                     !(((uintptr_t)d_q | (uintptr_t)d_k | (uintptr_t)d_cos | (uintptr_t)d_sin) & 15);
    if (vec) {
        int64_t vb = (total / 8 + 255) / 256;
        if (vb > cap) vb = cap;
        if (vb < 1) vb = 1;
        if (dtype == DT_F16) rope_half_table_vec_kernel<__half><<<(unsigned)vb, 256, 0, st>>>(A);
        else rope_half_table_vec_kernel<__nv_bfloat16><<<(unsigned)vb, 256, 0, st>>>(A);
        note_launch();
        return check_launch();
    }
    if (dtype == DT_F32) rope_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(A);
    else if (dtype == DT_F16) rope_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>(A);
    else rope_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(A);
    note_launch();
    return check_launch();
}
