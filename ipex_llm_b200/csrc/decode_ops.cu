// decode_ops.cu -- the small operators between the low-bit linears of a decoder layer, written so that ONE decode
// step can be recorded in a CUDA graph and replayed: every per-token quantity (cache length, position, the token
// itself) is read from device memory, never baked into a launch.
//
// Reference call sites (closed xe_addons kernels; S = /root/reference/python/llm/src/ipex_llm/):
//   rms_norm        xe_addons.rms_norm(weight, x_2d, eps)                 S/transformers/models/common.py:187-200
//   rope + KV write xe_addons.rotary_half_with_cache_inplaced(q, k, cos, sin) followed by
//                   DynamicFp8Cache.update -> xe_addons.quantize_key_value  S/transformers/models/llama.py:146-183,
//                                                                           S/transformers/models/utils.py:119-139
//   greedy pick     torch.argmax on the lm_head logits in the reference's generate loop
//                                                                           S/utils/benchmark_util_4_47.py:3340-3380
// Arithmetic follows the torch statements the reference falls back to on CPU (same roundings), so that the graph
// path, the eager path and a CPU run of the oracle-dequantised model differ only by summation order.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <math_constants.h>
#include <stdint.h>
#include "../../include/b200_lowbit.h"
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {
namespace {

__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ float ld_any(const void* p, int dt, int64_t i) {
    if (dt == DT_F32) return ((const float*)p)[i];
    if (dt == DT_F16) return __half2float(((const __half*)p)[i]);
    return __bfloat162float(((const __nv_bfloat16*)p)[i]);
}
__device__ __forceinline__ float round_to(float v, int dt) {
    if (dt == DT_F32) return v;
    if (dt == DT_F16) return __half2float(__float2half_rn(v));
    return __bfloat162float(__float2bfloat16_rn(v));
}
__device__ __forceinline__ void st_any(void* p, int dt, int64_t i, float v) {
    if (dt == DT_F32) ((float*)p)[i] = v;
    else if (dt == DT_F16) ((__half*)p)[i] = __float2half_rn(v);
    else ((__nv_bfloat16*)p)[i] = __float2bfloat16_rn(v);
}

// launch with programmatic stream serialization when the library-wide switch is on: the kernels below release
// their dependents at entry and read their inputs only after griddepcontrol.wait
template <typename... KArgs, typename... Args>
cudaError_t launch(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_get_pdl() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ------------------------------------------------------------------ RMSNorm
// y = weight * round_x((x * rsqrt(mean(x^2) + eps)))  -- transformers' LlamaRMSNorm.forward, the statement the
// reference's rms_norm_forward falls back to (models/common.py:201-205): statistics in fp32, the normalised value
// rounded to the activation dtype before the weight is applied.  One CTA per row; the row is read twice (the second
// read hits L1/L2).  Sum of squares: per-thread serial, lanes by shuffle, warps in index order -- deterministic.
constexpr int RMS_THREADS = 256;
struct RmsArgs {
    const void* x; int x_dtype; int64_t x_stride;     // [rows][hidden], row stride in elements
    const void* w; int w_dtype;                       // [hidden]
    void* y; int y_dtype; int64_t y_stride;
    int hidden; float eps;
};
__global__ void __launch_bounds__(RMS_THREADS) rms_norm_kernel(const RmsArgs A) {
    __shared__ float red[RMS_THREADS / 32];
    pdl_release();
    pdl_wait();
    const int64_t row = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float ss = 0.0f;
    for (int i = threadIdx.x; i < A.hidden; i += RMS_THREADS) {
        const float v = ld_any(A.x, A.x_dtype, row * A.x_stride + i);
        ss = fmaf(v, v, ss);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0) red[warp] = ss;
    __syncthreads();
    float tot = 0.0f;
#pragma unroll
    for (int w = 0; w < RMS_THREADS / 32; w++) tot += red[w];
    const float r = 1.0f / sqrtf(tot / (float)A.hidden + A.eps);
    for (int i = threadIdx.x; i < A.hidden; i += RMS_THREADS) {
        const float v = ld_any(A.x, A.x_dtype, row * A.x_stride + i);
        const float n = round_to(v * r, A.x_dtype);
        st_any(A.y, A.y_dtype, row * A.y_stride + i, ld_any(A.w, A.w_dtype, i) * n);
    }
}

// The same statement for the common case -- x, weight and y of one 2- or 4-byte type, rows and weight 16-byte aligned,
// hidden a multiple of the vector width and at most RMSV_THREADS * RMSV_MAX vectors: every thread keeps its part of the
// row in registers (one read of x), the weight vectors are requested BEFORE griddepcontrol.wait (they do not depend on
// the previous kernel), and the row leaves in 16-byte stores.  Latency, not bandwidth, is what this kernel costs a
// decode step (65 launches per Llama-2-7B token): the generic kernel above took ~10 us per launch in the chain.
constexpr int RMSV_THREADS = 512;
constexpr int RMSV_MAX = 4;
template <typename T> struct Vec16 { static constexpr int N = 16 / sizeof(T); };
template <typename T>
__global__ void __launch_bounds__(RMSV_THREADS) rms_norm_vec_kernel(const RmsArgs A) {
    constexpr int N = Vec16<T>::N;
    __shared__ float red[RMSV_THREADS / 32];
    pdl_release();
    const int64_t row = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nvec = A.hidden / N;
    uint4 wv[RMSV_MAX], xv[RMSV_MAX];
    const uint4* wp = reinterpret_cast<const uint4*>(A.w);
#pragma unroll
    for (int j = 0; j < RMSV_MAX; j++) {
        const int v = threadIdx.x + j * RMSV_THREADS;
        if (v < nvec) wv[j] = __ldg(wp + v);
    }
    pdl_wait();
    const uint4* xp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(A.x) + row * A.x_stride);
    float ss = 0.0f;
#pragma unroll
    for (int j = 0; j < RMSV_MAX; j++) {
        const int v = threadIdx.x + j * RMSV_THREADS;
        if (v < nvec) {
            xv[j] = xp[v];
            const T* e = reinterpret_cast<const T*>(&xv[j]);
#pragma unroll
            for (int i = 0; i < N; i++) { const float f = to_float<T>(e[i]); ss = fmaf(f, f, ss); }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0) red[warp] = ss;
    __syncthreads();
    float tot = 0.0f;
#pragma unroll
    for (int w = 0; w < RMSV_THREADS / 32; w++) tot += red[w];
    const float r = 1.0f / sqrtf(tot / (float)A.hidden + A.eps);
    uint4* yp = reinterpret_cast<uint4*>(reinterpret_cast<T*>(A.y) + row * A.y_stride);
#pragma unroll
    for (int j = 0; j < RMSV_MAX; j++) {
        const int v = threadIdx.x + j * RMSV_THREADS;
        if (v < nvec) {
            uint4 out;
            const T* e = reinterpret_cast<const T*>(&xv[j]);
            const T* we = reinterpret_cast<const T*>(&wv[j]);
            T* oe = reinterpret_cast<T*>(&out);
#pragma unroll
            for (int i = 0; i < N; i++) {
                const float n = to_float<T>(from_float<T>(to_float<T>(e[i]) * r));      // rounded to the activation dtype
                oe[i] = from_float<T>(to_float<T>(we[i]) * n);
            }
            yp[v] = out;
        }
    }
}

// ------------------------------------------------------------------ rotary embedding + FP8 cache write
// qkv: the merged projection's output [B][S][(Hq + 2 Hkv) * D].  Per token: q heads rotated in place; k heads
// rotated, rounded to the activation dtype (what rotary_half_with_cache_inplaced leaves in memory) and then to
// e5m2 (quantize_key_value) at cache position pos0 + s; v heads to e5m2 likewise.  cos / sin: tables indexed by
// absolute position, rotate_half layout (entry i serves the pair (i, i + D/2)).  One thread per pair.
struct RopeKvArgs {
    void* qkv; int dtype; int64_t sb, ss;             // batch / token strides in elements
    int B, S, Hq, Hkv, D;
    const void* cos; const void* sin; int cs_dtype; int64_t cs_row;   // [positions][>= D/2], row stride in elements
    int n_pos;                                        // rows of the tables
    const int* pos_ptr; int pos_add;                  // first position = (pos_ptr ? *pos_ptr : 0) + pos_add
    uint8_t* kc; int64_t kc_sb, kc_sh, kc_ss;         // [B][Hkv][capacity][D] e5m2
    uint8_t* vc; int64_t vc_sb, vc_sh, vc_ss;
    int capacity;
};
__device__ __forceinline__ uint8_t f_to_e5m2(float v) {
    return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E5M2);
}
__global__ void __launch_bounds__(256) rope_kv_append_kernel(const RopeKvArgs A) {
    pdl_release();
    pdl_wait();
    const int half = A.D / 2, heads = A.Hq + 2 * A.Hkv;
    const int pos0 = (A.pos_ptr ? *A.pos_ptr : 0) + A.pos_add;
    const int64_t total = (int64_t)A.B * A.S * heads * half;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx % half), h = (int)((idx / half) % heads);
        const int s = (int)((idx / ((int64_t)half * heads)) % A.S), b = (int)(idx / ((int64_t)half * heads * A.S));
        const int64_t base = b * A.sb + s * A.ss + (int64_t)h * A.D;
        const int pos = pos0 + s;
        float x0 = ld_any(A.qkv, A.dtype, base + i), x1 = ld_any(A.qkv, A.dtype, base + i + half);
        if (h < A.Hq + A.Hkv) {
            const int pr = pos < A.n_pos ? pos : A.n_pos - 1;
            const float c = ld_any(A.cos, A.cs_dtype, pr * A.cs_row + i), sn = ld_any(A.sin, A.cs_dtype, pr * A.cs_row + i);
            const float y0 = round_to(fmaf(x0, c, -x1 * sn), A.dtype), y1 = round_to(fmaf(x1, c, x0 * sn), A.dtype);
            x0 = y0; x1 = y1;
            if (h < A.Hq) {
                st_any(A.qkv, A.dtype, base + i, x0);
                st_any(A.qkv, A.dtype, base + i + half, x1);
                continue;
            }
        }
        if (pos >= A.capacity) continue;               // the host sized the cache; never write beyond it
        const bool is_k = h < A.Hq + A.Hkv;
        const int hk = h - A.Hq - (is_k ? 0 : A.Hkv);
        uint8_t* dst = is_k ? A.kc + b * A.kc_sb + hk * A.kc_sh + pos * A.kc_ss
                            : A.vc + b * A.vc_sb + hk * A.vc_sh + pos * A.vc_ss;
        dst[i] = f_to_e5m2(x0);
        dst[i + half] = f_to_e5m2(x1);
    }
}

// The same for many tokens (a prompt): 8 pairs per thread, 16-byte loads of the row halves and of the table rows, 8-byte cache
// stores -- the one-pair kernel above spends its time in index arithmetic and 1- / 2-byte accesses (34 us for 512 tokens of a
// 7B layer).  Half-precision qkv and tables of the same dtype, D % 16 == 0, 16-byte aligned rows; same roundings.
template <typename T>
__global__ void __launch_bounds__(256) rope_kv_append_vec_kernel(const RopeKvArgs A) {
    pdl_release();
    pdl_wait();
    const int half = A.D / 2, heads = A.Hq + 2 * A.Hkv, H8 = half / 8;
    const int pos0 = (A.pos_ptr ? *A.pos_ptr : 0) + A.pos_add;
    const int64_t total = (int64_t)A.B * A.S * heads * H8;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx % H8) * 8, h = (int)((idx / H8) % heads);
        const int s = (int)((idx / ((int64_t)H8 * heads)) % A.S), b = (int)(idx / ((int64_t)H8 * heads * A.S));
        T* row = (T*)A.qkv + b * A.sb + s * A.ss + (int64_t)h * A.D;
        const int pos = pos0 + s;
        T x0[8], x1[8];
        *reinterpret_cast<uint4*>(x0) = *reinterpret_cast<const uint4*>(row + i);
        *reinterpret_cast<uint4*>(x1) = *reinterpret_cast<const uint4*>(row + i + half);
        if (h < A.Hq + A.Hkv) {
            const int pr = pos < A.n_pos ? pos : A.n_pos - 1;
            T c[8], sn[8];
            *reinterpret_cast<uint4*>(c) = *reinterpret_cast<const uint4*>((const T*)A.cos + (int64_t)pr * A.cs_row + i);
            *reinterpret_cast<uint4*>(sn) = *reinterpret_cast<const uint4*>((const T*)A.sin + (int64_t)pr * A.cs_row + i);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float a = to_float<T>(x0[e]), d = to_float<T>(x1[e]), cf = to_float<T>(c[e]), sf = to_float<T>(sn[e]);
                x0[e] = from_float<T>(fmaf(a, cf, -d * sf));
                x1[e] = from_float<T>(fmaf(d, cf, a * sf));
            }
            if (h < A.Hq) {
                *reinterpret_cast<uint4*>(row + i) = *reinterpret_cast<const uint4*>(x0);
                *reinterpret_cast<uint4*>(row + i + half) = *reinterpret_cast<const uint4*>(x1);
                continue;
            }
        }
        if (pos >= A.capacity) continue;               // the host sized the cache; never write beyond it
        const bool is_k = h < A.Hq + A.Hkv;
        const int hk = h - A.Hq - (is_k ? 0 : A.Hkv);
        uint8_t* dst = is_k ? A.kc + b * A.kc_sb + hk * A.kc_sh + pos * A.kc_ss
                            : A.vc + b * A.vc_sb + hk * A.vc_sh + pos * A.vc_ss;
        uint8_t q0[8], q1[8];
#pragma unroll
        for (int e = 0; e < 8; e++) { q0[e] = f_to_e5m2(to_float<T>(x0[e])); q1[e] = f_to_e5m2(to_float<T>(x1[e])); }
        *reinterpret_cast<uint2*>(dst + i) = *reinterpret_cast<const uint2*>(q0);
        *reinterpret_cast<uint2*>(dst + i + half) = *reinterpret_cast<const uint2*>(q1);
    }
}


// ------------------------------------------------------------------ greedy pick + bookkeeping of one decode step
// One CTA per sequence: index of the largest logit (the lowest index among equals, torch.argmax's rule), written to
// next_ids[b] (the next step's input) and to tokens[b][state[1]].  The last CTA to finish advances the step's device
// state: state[0] (tokens in the cache = next position) and state[1] (tokens generated) += 1; state[2] is the
// arrival counter and returns to 0.  While state[1] < state[3], the state[4] (<= 8) ids in `suppress` are left out of the
// pick -- transformers' MinNewTokensLength rule, which keeps the end-of-sequence ids out of the first min_new_tokens.
constexpr int ARGMAX_THREADS = 1024;
struct GreedyArgs {
    const void* logits; int dtype; int64_t row_stride; int V;
    int64_t* next_ids;
    int64_t* tokens; int64_t tokens_stride; int tokens_cap;
    int* state;
    const int64_t* suppress;
};
__global__ void __launch_bounds__(ARGMAX_THREADS) greedy_advance_kernel(const GreedyArgs A) {
    __shared__ float s_v[ARGMAX_THREADS / 32];
    __shared__ int s_i[ARGMAX_THREADS / 32];
    pdl_release();
    pdl_wait();
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int step = A.state[1];
    int n_sup = (A.suppress && step < A.state[3]) ? min(max(A.state[4], 0), 8) : 0;
    int64_t sup[8];
#pragma unroll
    for (int j = 0; j < 8; j++) sup[j] = j < n_sup ? A.suppress[j] : -1;
    float best = -CUDART_INF_F; int bi = 0x7fffffff;
    auto consider = [&](float v, int i) {
        if (n_sup) {
#pragma unroll
            for (int j = 0; j < 8; j++) if (sup[j] == i) v = -CUDART_INF_F;
        }
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    };
    const int esz = A.dtype == DT_F32 ? 4 : 2, nper = 16 / esz;
    const char* rowp = (const char*)A.logits + (int64_t)b * A.row_stride * esz;
    int done = 0;
    if (!((uintptr_t)rowp & 15)) {                    // 16-byte pieces: a 64 KB row is a handful of loads per thread
        const int nvec = A.V / nper;
        for (int v = threadIdx.x; v < nvec; v += ARGMAX_THREADS) {
            const uint4 q = reinterpret_cast<const uint4*>(rowp)[v];
            if (A.dtype == DT_F32) {
                const float* e = reinterpret_cast<const float*>(&q);
#pragma unroll
                for (int i = 0; i < 4; i++) consider(e[i], v * 4 + i);
            } else if (A.dtype == DT_F16) {
                const __half* e = reinterpret_cast<const __half*>(&q);
#pragma unroll
                for (int i = 0; i < 8; i++) consider(__half2float(e[i]), v * 8 + i);
            } else {
                const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
                for (int i = 0; i < 8; i++) consider(__bfloat162float(e[i]), v * 8 + i);
            }
        }
        done = nvec * nper;
    }
    for (int i = done + threadIdx.x; i < A.V; i += ARGMAX_THREADS) consider(ld_any(A.logits, A.dtype, b * A.row_stride + i), i);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (lane == 0) { s_v[warp] = best; s_i[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < ARGMAX_THREADS / 32; w++)
            if (s_v[w] > best || (s_v[w] == best && s_i[w] < bi)) { best = s_v[w]; bi = s_i[w]; }
        if (bi == 0x7fffffff) bi = 0;                  // a row of NaNs / -inf
        A.next_ids[b] = bi;
        if (step >= 0 && step < A.tokens_cap) A.tokens[b * A.tokens_stride + step] = bi;
        __threadfence();
        if (atomicAdd(&A.state[2], 1) == (int)gridDim.x - 1) {
            A.state[0] += 1; A.state[1] = step + 1; A.state[2] = 0;
        }
    }
}

}  // namespace
}  // namespace lowbit

using namespace lowbit;

extern "C" int lowbit_rms_norm_cuda(const void* d_x, int x_dtype, int64_t x_row_stride, const void* d_weight, int w_dtype,
                                    float eps, int64_t rows, int64_t hidden, void* d_y, int y_dtype, int64_t y_row_stride,
                                    void* stream) {
    reset_launch_count();
    if (!d_x || !d_weight || !d_y || x_dtype < 0 || x_dtype > 2 || w_dtype < 0 || w_dtype > 2 || y_dtype < 0 || y_dtype > 2 ||
        rows < 0 || hidden <= 0 || hidden > INT32_MAX || x_row_stride < hidden || y_row_stride < hidden) {
        set_error(LB_ERR_ARG, "lowbit_rms_norm_cuda: bad argument (rows %lld, hidden %lld)", (long long)rows, (long long)hidden);
        return LB_ERR_ARG;
    }
    if (rows == 0) return LB_OK;
    RmsArgs A;
    A.x = d_x; A.x_dtype = x_dtype; A.x_stride = x_row_stride; A.w = d_weight; A.w_dtype = w_dtype;
    A.y = d_y; A.y_dtype = y_dtype; A.y_stride = y_row_stride; A.hidden = (int)hidden; A.eps = eps;
    const int esz = x_dtype == DT_F32 ? 4 : 2, nper = 16 / esz;
    const bool vec = x_dtype == w_dtype && x_dtype == y_dtype && hidden % nper == 0 && hidden / nper <= RMSV_THREADS * RMSV_MAX &&
                     !(((uintptr_t)d_x | (uintptr_t)d_weight | (uintptr_t)d_y) & 15) && x_row_stride % nper == 0 && y_row_stride % nper == 0;
    cudaError_t e;
    if (!vec) e = launch(rms_norm_kernel, dim3((unsigned)rows), dim3(RMS_THREADS), (cudaStream_t)stream, A);
    else if (x_dtype == DT_F32) e = launch(rms_norm_vec_kernel<float>, dim3((unsigned)rows), dim3(RMSV_THREADS), (cudaStream_t)stream, A);
    else if (x_dtype == DT_F16) e = launch(rms_norm_vec_kernel<__half>, dim3((unsigned)rows), dim3(RMSV_THREADS), (cudaStream_t)stream, A);
    else e = launch(rms_norm_vec_kernel<__nv_bfloat16>, dim3((unsigned)rows), dim3(RMSV_THREADS), (cudaStream_t)stream, A);
    if (e != cudaSuccess) return check_launch();
    note_launch();
    return check_launch();
}

extern "C" int lowbit_rope_kv_append_cuda(void* d_qkv, int dtype, int64_t batch_stride, int64_t token_stride, int B, int S,
                                          int Hq, int Hkv, int D, const void* d_cos, const void* d_sin, int cs_dtype,
                                          int64_t cs_row_stride, int n_positions, const int32_t* d_pos, int pos_add,
                                          void* d_k_cache, const int64_t* k_strides, void* d_v_cache,
                                          const int64_t* v_strides, int capacity, void* stream) {
    reset_launch_count();
    if (!d_qkv || !d_cos || !d_sin || !d_k_cache || !d_v_cache || !k_strides || !v_strides || dtype < 0 || dtype > 2 ||
        cs_dtype < 0 || cs_dtype > 2 || B < 0 || S < 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || (D & 1) || cs_row_stride < D / 2 ||
        n_positions <= 0 || capacity <= 0 || token_stride < (int64_t)(Hq + 2 * Hkv) * D) {
        set_error(LB_ERR_ARG, "lowbit_rope_kv_append_cuda: bad argument (B %d, S %d, Hq %d, Hkv %d, D %d)", B, S, Hq, Hkv, D);
        return LB_ERR_ARG;
    }
    const int64_t total = (int64_t)B * S * (Hq + 2 * Hkv) * (D / 2);
    if (total == 0) return LB_OK;
    RopeKvArgs A;
    A.qkv = d_qkv; A.dtype = dtype; A.sb = batch_stride; A.ss = token_stride; A.B = B; A.S = S; A.Hq = Hq; A.Hkv = Hkv; A.D = D;
    A.cos = d_cos; A.sin = d_sin; A.cs_dtype = cs_dtype; A.cs_row = cs_row_stride; A.n_pos = n_positions;
    A.pos_ptr = d_pos; A.pos_add = pos_add;
    A.kc = (uint8_t*)d_k_cache; A.kc_sb = k_strides[0]; A.kc_sh = k_strides[1]; A.kc_ss = k_strides[2];
    A.vc = (uint8_t*)d_v_cache; A.vc_sb = v_strides[0]; A.vc_sh = v_strides[1]; A.vc_ss = v_strides[2];
    A.capacity = capacity;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    const bool vec = S > 1 && dtype != DT_F32 && cs_dtype == dtype && D % 16 == 0 && token_stride % 8 == 0 && batch_stride % 8 == 0 &&
                     cs_row_stride % 8 == 0 && !(((uintptr_t)d_qkv | (uintptr_t)d_cos | (uintptr_t)d_sin) & 15) &&
                     !(((uintptr_t)d_k_cache | (uintptr_t)d_v_cache) & 7) &&
                     !((k_strides[0] | k_strides[1] | k_strides[2] | v_strides[0] | v_strides[1] | v_strides[2]) & 7);
    if (vec) {
        int64_t vb = (total / 8 + 255) / 256;
        if (vb > cap) vb = cap;
        if (vb < 1) vb = 1;
        const cudaError_t e = dtype == DT_F16 ? launch(rope_kv_append_vec_kernel<__half>, dim3((unsigned)vb), dim3(256), (cudaStream_t)stream, A)
                                              : launch(rope_kv_append_vec_kernel<__nv_bfloat16>, dim3((unsigned)vb), dim3(256), (cudaStream_t)stream, A);
        if (e != cudaSuccess) return check_launch();
        note_launch();
        return check_launch();
    }
    if (launch(rope_kv_append_kernel, dim3((unsigned)blocks), dim3(256), (cudaStream_t)stream, A) != cudaSuccess) return check_launch();
    note_launch();
    return check_launch();
}

extern "C" int lowbit_greedy_advance_cuda(const void* d_logits, int dtype, int64_t row_stride, int B, int V, int64_t* d_next_ids,
                                          int64_t* d_tokens, int64_t tokens_row_stride, int tokens_capacity, int32_t* d_state,
                                          const int64_t* d_suppress_ids, void* stream) {
    reset_launch_count();
    if (!d_logits || !d_next_ids || !d_tokens || !d_state || dtype < 0 || dtype > 2 || B <= 0 || V <= 0 || row_stride < V ||
        tokens_capacity <= 0 || tokens_row_stride < tokens_capacity) {
        set_error(LB_ERR_ARG, "lowbit_greedy_advance_cuda: bad argument (B %d, V %d)", B, V);
        return LB_ERR_ARG;
    }
    GreedyArgs A;
    A.logits = d_logits; A.dtype = dtype; A.row_stride = row_stride; A.V = V; A.next_ids = d_next_ids;
    A.tokens = d_tokens; A.tokens_stride = tokens_row_stride; A.tokens_cap = tokens_capacity; A.state = d_state; A.suppress = d_suppress_ids;
    if (launch(greedy_advance_kernel, dim3((unsigned)B), dim3(ARGMAX_THREADS), (cudaStream_t)stream, A) != cudaSuccess) return check_launch();
    note_launch();
    return check_launch();
}
