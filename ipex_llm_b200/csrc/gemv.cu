// gemv.cu -- host side of kernel (b): workspace layout, grid sizing, qtype dispatch, plus the
// scalar generic kernel (nf4 until its mma codec lands; rows beyond the last 16-row tile) and
// the fp32 -> 3 x bf16 activation split.  The mma kernel itself is in gemv_kernel.cuh.
#include <cstdlib>
#include <mutex>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"
#include "gemv_args.h"

namespace lowbit {

void gemv_launch_sym_int4(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st);
void gemv_launch_asym_int4(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st);
void gemv_launch_fp8_e4m3(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st);
void gemv_launch_fp8_e5m2(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st);
void gemv_launch_fp6(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st);
void gemv_launch_nf4(int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st);

// ------------------------------------------------------------------ generic (any qtype, tail rows)
// one warp per output row, scalar decode; used for nf4 until its mma codec lands, and for the
// rows beyond the last full 16-row tile (kept in canonical records in the tail region).
template <typename XT>
__global__ void gemv_generic_kernel(const uint8_t* __restrict__ w, TiledGeom G, int qt, int64_t row_begin,
                                    int64_t row_end, const XT* __restrict__ x, int M, void* y, int y_dtype,
                                    const void* bias, int accumulate) {
    const int lane = threadIdx.x & 31;
    const int64_t row = row_begin + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= row_end) return;
    const int hdr = header_bytes(qt);
    for (int m0 = 0; m0 < M; m0 += 8) {
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = 0.0f;
        for (int64_t kb = 0; kb < G.nb; kb++) {
            float d, mn = 0.0f; uint32_t c0, c1;
            if (row < G.r64) {
                const int r = (int)(row % 16);
                const uint8_t* tile = w + tile_code_offset(G, row, kb);
                c0 = tiled_code(qt, tile, r, lane); c1 = tiled_code(qt, tile, r, lane + 32);
                d = half_bits_to_float(*(const uint16_t*)(w + G.off_s + scale_offset(G, row, kb)));
                if (G.has_m) mn = half_bits_to_float(*(const uint16_t*)(w + G.off_m + scale_offset(G, row, kb)));
            } else {
                const uint8_t* rec = w + G.off_tail + ((row - G.r64) * G.nb + kb) * G.blk_bytes;
                d = half_bits_to_float((uint16_t)rec[0] | ((uint16_t)rec[1] << 8));
                if (G.has_m) mn = half_bits_to_float((uint16_t)rec[2] | ((uint16_t)rec[3] << 8));
                c0 = canonical_code(qt, rec + hdr, lane); c1 = canonical_code(qt, rec + hdr, lane + 32);
            }
            float w0 = __fmul_rn(code_value(qt, c0), d), w1 = __fmul_rn(code_value(qt, c1), d);
            if (G.has_m) { w0 = __fadd_rn(w0, mn); w1 = __fadd_rn(w1, mn); }
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (m0 + i < M) {
                    const XT* xr = x + (int64_t)(m0 + i) * G.k + kb * 64;
                    acc[i] = fmaf(w0, to_float<XT>(xr[lane]), acc[i]);
                    acc[i] = fmaf(w1, to_float<XT>(xr[lane + 32]), acc[i]);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
#pragma unroll
            for (int o = 16; o; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
            if (lane == 0 && m0 + i < M)
                store_out(y, y_dtype, (int64_t)(m0 + i) * G.rows + row, acc[i], bias, row, accumulate);
        }
    }
}

// fp32 activations -> three bf16 planes (hi, mid, lo) whose sum reproduces x to ~2^-24
__global__ void split_f32_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = x[i];
        const __nv_bfloat16 hi = __float2bfloat16_rn(v);
        const float r1 = v - __bfloat162float(hi);
        const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
        const float r2 = r1 - __bfloat162float(mid);
        out[i] = hi; out[n + i] = mid; out[2 * n + i] = __float2bfloat16_rn(r2);
    }
}

// ------------------------------------------------------------------ host side
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static unsigned long long* g_ts = nullptr;   // development: phase timestamps
void gemv_set_debug_timestamps(void* p) { g_ts = (unsigned long long*)p; }
static thread_local const uint8_t* t_next_w = nullptr;     // L2 prefetch hint for the next launch (one-shot)
static thread_local unsigned long long t_next_bytes = 0;
void gemv_set_next_weights(const void* w, size_t bytes) { t_next_w = (const uint8_t*)w; t_next_bytes = bytes; }
static int g_pdl = 0;     // programmatic dependent launch for the GEMV kernels (set by the host engine)
void gemv_set_pdl(int on) { g_pdl = on ? 1 : 0; }
int gemv_get_pdl() { return g_pdl; }
static thread_local GemvPush t_push = {};
void gemv_set_push(const GemvPush* p) { if (p) t_push = *p; else t_push = GemvPush{}; }

// The only scratch the matmul needs: three bf16 planes of an fp32 activation chunk.
size_t gemv_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    (void)n_rows;
    const int64_t mm = m < 8 * GEMV_MAX_NT ? m : 8 * GEMV_MAX_NT;
    return x_dtype == DT_F32 ? align_up((size_t)3 * mm * k * 2, 256) : 0;
}

static bool has_mma_codec(int qt) {
    return qt == QT_SYM_INT4 || qt == QT_ASYM_INT4 || qt == QT_NF4 || qt == QT_FP8_E4M3 || qt == QT_FP8_E5M2 || qt == QT_FP6;
}

template <typename XT>
static void launch_generic(const void* w, const TiledGeom& G, int qt, int64_t r0, int64_t r1, const void* x, int M,
                           void* y, int y_dtype, const void* bias, int accumulate, cudaStream_t st) {
    const int warps = 4;
    const int grid = (int)((r1 - r0 + warps - 1) / warps);
    gemv_generic_kernel<XT><<<grid, warps * 32, 0, st>>>((const uint8_t*)w, G, qt, r0, r1, (const XT*)x, M, y, y_dtype,
                                                        bias, accumulate);
}

// y = act(g) * u elementwise (the unfused tail of the MLP when the fused GEMV does not apply).  gate / up: fp32, or the
// activation dtype when they come from the tensor-core kernels (whose precision contract rounds there anyway).
__device__ __forceinline__ float in_at(const void* p, int dt, int64_t i) {
    return dt == DT_F32 ? ((const float*)p)[i] : dt == DT_F16 ? __half2float(((const __half*)p)[i]) : __bfloat162float(((const __nv_bfloat16*)p)[i]);
}
__global__ void act_mul_kernel(const void* __restrict__ gate, const void* __restrict__ up, int in_dtype, void* y, int y_dtype, int64_t n,
                               int act) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        store_out(y, y_dtype, i, apply_act(in_at(gate, in_dtype, i), act) * in_at(up, in_dtype, i), nullptr, 0, 0);
}
// eight elements per thread, 16-byte loads and stores: half-precision in and out (the prefill route)
template <bool BF>
__global__ void __launch_bounds__(256) act_mul_vec_kernel(const uint4* __restrict__ gate, const uint4* __restrict__ up, uint4* __restrict__ y,
                                                          int64_t n8, int act) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const uint4 g = __ldcs(gate + i), u = __ldcs(up + i);
        const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w};
        uint32_t ow[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float g0, g1, u0, u1;
            if (BF) {
                g0 = __uint_as_float(gw[j] << 16); g1 = __uint_as_float(gw[j] & 0xffff0000u);
                u0 = __uint_as_float(uw[j] << 16); u1 = __uint_as_float(uw[j] & 0xffff0000u);
            } else {
                const float2 gf = __half22float2(*reinterpret_cast<const __half2*>(&gw[j])), uf = __half22float2(*reinterpret_cast<const __half2*>(&uw[j]));
                g0 = gf.x; g1 = gf.y; u0 = uf.x; u1 = uf.y;
            }
            const float r0 = apply_act(g0, act) * u0, r1 = apply_act(g1, act) * u1;
            if (BF) { const __nv_bfloat162 b = __floats2bfloat162_rn(r0, r1); ow[j] = *reinterpret_cast<const uint32_t*>(&b); }
            else { const __half2 h = __floats2half2_rn(r0, r1); ow[j] = *reinterpret_cast<const uint32_t*>(&h); }
        }
        y[i] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
}

int act_mul_launch(const void* gate, const void* up, int in_dtype, void* y, int y_dtype, int64_t n, int act, cudaStream_t st) {
    if (in_dtype != DT_F32 && in_dtype == y_dtype && n % 8 == 0 && !(((uintptr_t)gate | (uintptr_t)up | (uintptr_t)y) & 15)) {
        const int64_t n8 = n / 8;
        int g = (int)((n8 + 255) / 256); if (g > sm_count() * 8) g = sm_count() * 8;
        if (in_dtype == DT_BF16) act_mul_vec_kernel<true><<<g, 256, 0, st>>>((const uint4*)gate, (const uint4*)up, (uint4*)y, n8, act);
        else act_mul_vec_kernel<false><<<g, 256, 0, st>>>((const uint4*)gate, (const uint4*)up, (uint4*)y, n8, act);
        note_launch();
        return check_launch();
    }
    int g = (int)((n + 255) / 256); if (g > sm_count() * 8) g = sm_count() * 8;
    act_mul_kernel<<<g, 256, 0, st>>>(gate, up, in_dtype, y, y_dtype, n, act);
    note_launch();
    return check_launch();
}

// up to 8 rows of any activation dtype; 9..16 rows in half precision (two n-tiles; fp32 rows would need three planes of each)
bool gemv_mlp_fused_supported(int qt, int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    qt = canon_qtype(qt);
    return supported(qt) && has_mma_codec(qt) && n_rows > 0 && n_rows % RG_ROWS == 0 && k > 0 && k % QK == 0 && m >= 1 &&
           (m <= 8 || (m <= 16 && x_dtype != DT_F32));
}

int gemv_launch(const void* w, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m,
                void* y, int y_dtype, const void* bias, int accumulate, void* workspace,
                size_t workspace_bytes, cudaStream_t st) {
    return gemv_launch_ex(w, nullptr, 0, qt, n_rows, k, x, x_dtype, m, y, y_dtype, bias, accumulate, workspace, workspace_bytes, st);
}

// w2 != NULL: y = act(x . w^T) * (x . w2^T) in one launch (caller checked gemv_mlp_fused_supported)
int gemv_launch_ex(const void* w, const void* w2, int act, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype,
                   int64_t m, void* y, int y_dtype, const void* bias, int accumulate, void* workspace,
                   size_t workspace_bytes, cudaStream_t st) {
    qt = canon_qtype(qt);
    const GemvPush push = t_push;                      // one-shot
    t_push = GemvPush{};
    // fused all-reduce: one pass (m <= 32), every row in the tensor-core part
    if (push.world > 0 && (w2 || m > 8 * GEMV_MAX_NT || !has_mma_codec(qt) || n_rows % RG_ROWS)) return LB_ERR_ARG;
    if (w2 && (!gemv_mlp_fused_supported(qt, n_rows, k, m, x_dtype) || ((uintptr_t)w2 & 15) || bias || accumulate)) return LB_ERR_ARG;
    if (!supported(qt) || k <= 0 || k % QK || n_rows <= 0 || m <= 0) return LB_ERR_ARG;
    if (x_dtype < 0 || x_dtype > 2 || y_dtype < 0 || y_dtype > 2) return LB_ERR_ARG;
    if (((uintptr_t)w & 15) || ((uintptr_t)x & 15)) return LB_ERR_ARG;        // TMA bulk / 128-bit loads
    const TiledGeom G = make_geom(n_rows, k, qt);
    const int esz_x = x_dtype == DT_F32 ? 4 : 2, esz_y = y_dtype == DT_F32 ? 4 : 2;

    // M beyond one pass: chunks of 32 activation rows (the tensor-core GEMM takes over for large M)
    for (int64_t mb = 0; mb < m; mb += 8 * GEMV_MAX_NT) {
        const int64_t mc = (m - mb) < 8 * GEMV_MAX_NT ? (m - mb) : 8 * GEMV_MAX_NT;
        const char* xc = (const char*)x + (size_t)mb * k * esz_x;
        char* yc = (char*)y + (size_t)mb * n_rows * esz_y;
        int64_t mma_rows = has_mma_codec(qt) ? G.r64 : 0;
        if (mma_rows > 0) {
            GemvArgs A = {};
            A.w = (const uint8_t*)w; A.w2 = (const uint8_t*)w2; A.act = act; A.G = G; A.M = (int)mc;
            A.y = yc; A.y_dtype = y_dtype; A.bias = bias; A.accumulate = accumulate;
            A.plane_stride = mc * k;
            A.pdl = g_pdl; A.ts = g_ts;
            A.next_w = t_next_w; A.next_bytes = t_next_bytes;
            t_next_w = nullptr; t_next_bytes = 0;            // one-shot hint
            A.push_world = push.world; A.push_rank = push.rank; A.push_slot_bytes = push.slot_bytes; A.push_epochs = push.epochs;
            A.push_own = push.own; A.push_inbox_bytes = push.inbox_bytes; A.push_spin_limit = push.spin_limit;
            for (int r = 0; r < 8; r++) A.push_inbox[r] = push.inbox[r];
            if (x_dtype == DT_F32) {
                const size_t need = gemv_workspace_size(n_rows, k, mc, x_dtype);
                if (!workspace || workspace_bytes < need) return LB_ERR_WORKSPACE;
                __nv_bfloat16* planes = (__nv_bfloat16*)workspace;
                const int64_t n = mc * k;
                int sg = (int)((n + 255) / 256); if (sg > sm_count() * 8) sg = sm_count() * 8;
                split_f32_kernel<<<sg, 256, 0, st>>>((const float*)xc, planes, n);
                note_launch();
                A.x = (const uint16_t*)planes;
            } else A.x = (const uint16_t*)xc;
            const int64_t groups = G.r64 / RG_ROWS;
            switch (qt) {
                case QT_SYM_INT4:  gemv_launch_sym_int4(x_dtype, (int)mc, groups, G.nb, A, st); break;
                case QT_ASYM_INT4: gemv_launch_asym_int4(x_dtype, (int)mc, groups, G.nb, A, st); break;
                case QT_FP8_E4M3:  gemv_launch_fp8_e4m3(x_dtype, (int)mc, groups, G.nb, A, st); break;
                case QT_FP8_E5M2:  gemv_launch_fp8_e5m2(x_dtype, (int)mc, groups, G.nb, A, st); break;
                case QT_FP6:       gemv_launch_fp6(x_dtype, (int)mc, groups, G.nb, A, st); break;
                case QT_NF4:       gemv_launch_nf4(x_dtype, (int)mc, groups, G.nb, A, st); break;
                default: return LB_ERR_ARG;
            }
            note_launch();
        }
        if (mma_rows < n_rows) {
            if (x_dtype == DT_F32) launch_generic<float>(w, G, qt, mma_rows, n_rows, xc, (int)mc, yc, y_dtype, bias, accumulate, st);
            else if (x_dtype == DT_F16) launch_generic<__half>(w, G, qt, mma_rows, n_rows, xc, (int)mc, yc, y_dtype, bias, accumulate, st);
            else launch_generic<__nv_bfloat16>(w, G, qt, mma_rows, n_rows, xc, (int)mc, yc, y_dtype, bias, accumulate, st);
            note_launch();
        }
    }
    return check_launch();
}

// ------------------------------------------------------------------ mixture-of-experts decode (one token)
static int launch_by_qt(int qt, int x_dtype, int m, int64_t groups, int64_t nb, const GemvArgs& A, cudaStream_t st) {
    switch (qt) {
        case QT_SYM_INT4:  gemv_launch_sym_int4(x_dtype, m, groups, nb, A, st); break;
        case QT_ASYM_INT4: gemv_launch_asym_int4(x_dtype, m, groups, nb, A, st); break;
        case QT_FP8_E4M3:  gemv_launch_fp8_e4m3(x_dtype, m, groups, nb, A, st); break;
        case QT_FP8_E5M2:  gemv_launch_fp8_e5m2(x_dtype, m, groups, nb, A, st); break;
        case QT_FP6:       gemv_launch_fp6(x_dtype, m, groups, nb, A, st); break;
        case QT_NF4:       gemv_launch_nf4(x_dtype, m, groups, nb, A, st); break;
        default: return LB_ERR_ARG;
    }
    note_launch();
    return LB_OK;
}

bool moe_supported(int qt, int64_t hidden, int64_t inter, int top_k) {
    qt = canon_qtype(qt);
    return supported(qt) && has_mma_codec(qt) && hidden > 0 && inter > 0 && hidden % RG_ROWS == 0 && inter % RG_ROWS == 0 &&
           top_k >= 1 && top_k <= 1024;
}

// [planes of x | routed intermediates | planes of the intermediates]; the planes only for fp32 activations
size_t moe_workspace_size(int64_t hidden, int64_t inter, int top_k, int x_dtype) {
    const size_t esz = x_dtype == DT_F32 ? 4 : 2;
    size_t n = align_up((size_t)top_k * inter * esz, 256);
    if (x_dtype == DT_F32) n += align_up((size_t)3 * hidden * 2, 256) + align_up((size_t)3 * top_k * inter * 2, 256);
    return n;
}

static void split_planes(const float* x, __nv_bfloat16* planes, int64_t n, cudaStream_t st) {
    int sg = (int)((n + 255) / 256); if (sg > sm_count() * 8) sg = sm_count() * 8;
    split_f32_kernel<<<sg, 256, 0, st>>>(x, planes, n);
    note_launch();
}

int moe_launch(const MoeParams& P, cudaStream_t st) {
    const int qt = canon_qtype(P.qtype);
    if (!moe_supported(qt, P.hidden, P.inter, P.top_k)) return LB_ERR_ARG;
    if (P.x_dtype < 0 || P.x_dtype > 2 || P.y_dtype < 0 || P.y_dtype > 2 || P.w_dtype < 0 || P.w_dtype > 2) return LB_ERR_ARG;
    if (((uintptr_t)P.x & 15) || ((uintptr_t)P.workspace & 255)) return LB_ERR_ARG;
    if (!P.workspace || P.workspace_bytes < moe_workspace_size(P.hidden, P.inter, P.top_k, P.x_dtype)) return LB_ERR_WORKSPACE;
    const bool f32 = P.x_dtype == DT_F32;
    const size_t esz = f32 ? 4 : 2;
    char* ws = (char*)P.workspace;
    char* inter_buf = ws;
    char* planes1 = ws + align_up((size_t)P.top_k * P.inter * esz, 256);
    char* planes2 = planes1 + align_up((size_t)3 * P.hidden * 2, 256);

    // launch 1: h[z] = act(x . Wgate[e_z]^T) * (x . Wup[e_z]^T) for every routed slot z
    GemvArgs A = {};
    A.mode = 1; A.M = 1; A.act = P.act;
    A.G = make_geom(P.inter, P.hidden, qt);
    A.tab_w = P.gates; A.tab_w2 = P.ups; A.sel = P.ids; A.sel_i64 = P.ids_i64; A.n_experts = P.n_experts; A.n_seg = P.top_k;
    A.y = inter_buf; A.y_dtype = P.x_dtype;
    A.plane_stride = P.hidden; A.pdl = g_pdl; A.ts = nullptr;
    if (f32) { split_planes((const float*)P.x, (__nv_bfloat16*)planes1, P.hidden, st); A.x = (const uint16_t*)planes1; }
    else A.x = (const uint16_t*)P.x;
    int rc = launch_by_qt(qt, P.x_dtype, 1, A.G.r64 / RG_ROWS, A.G.nb, A, st);
    if (rc) return rc;

    // launch 2: y (+)= sum_z weight[z] * (h[z] . Wdown[e_z]^T): one GEMV over the concatenated K axis
    GemvArgs B = {};
    B.mode = 2; B.M = 1;
    B.G = make_geom(P.hidden, P.inter, qt);
    B.tab_w = P.downs; B.sel = P.ids; B.sel_i64 = P.ids_i64; B.n_experts = P.n_experts; B.n_seg = P.top_k;
    B.seg_scale = P.weights; B.seg_scale_dtype = P.w_dtype;
    B.y = P.y; B.y_dtype = P.y_dtype; B.accumulate = P.accumulate;
    B.plane_stride = (int64_t)P.top_k * P.inter; B.pdl = g_pdl;
    if (f32) { split_planes((const float*)inter_buf, (__nv_bfloat16*)planes2, (int64_t)P.top_k * P.inter, st); B.x = (const uint16_t*)planes2; }
    else B.x = (const uint16_t*)inter_buf;
    rc = launch_by_qt(qt, P.x_dtype, 1, B.G.r64 / RG_ROWS, B.G.nb, B, st);
    if (rc) return rc;
    return check_launch();
}

}  // namespace lowbit

// ------------------------------------------------------------------ allocation-free fp32 matmul
// lowbit_matmul_cuda with fp32 activations needs scratch for the three bf16 planes it splits the
// activations into.  Normally the caller allocates that (or lets the library cudaMalloc it).  Some
// hosts can do neither: a serving runtime that handed every byte of device memory to its own arena
// at start-up and forbids later cudaMalloc calls, a process running under a fixed device-memory
// budget, or a latency-critical thread that must not take the allocator's implicit device
// synchronisation.  For them the scratch lives in the library image itself.
namespace {

constexpr size_t kNoAllocMaxElems = 262144;   // activation elements (m * k) one call may carry
// three bf16 planes of kNoAllocMaxElems elements = 1.5 MiB, resident on every GPU that loads the
// library until the process exits
static __device__ __align__(256) unsigned char g_noalloc_scratch[3 * kNoAllocMaxElems * 2];

std::mutex g_noalloc_mu;
void* g_noalloc_ptr[64] = {nullptr};          // device address of the array, per device ordinal

// In host code the symbol above is only a stand-in: the device address has to be asked for, once per
// device, with that device current.
int noalloc_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        lowbit::set_error(lowbit::LB_ERR_NO_DEVICE, "lowbit_matmul_f32_noalloc_cuda: no current CUDA device (%s)",
                          e == cudaSuccess ? "ordinal out of range" : cudaGetErrorString(e));
        cudaGetLastError();
        return lowbit::LB_ERR_NO_DEVICE;
    }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            lowbit::set_error(lowbit::LB_ERR_CUDA, "lowbit_matmul_f32_noalloc_cuda: cudaGetSymbolAddress failed: %s",
                              cudaGetErrorString(e));
            cudaGetLastError();
            return lowbit::LB_ERR_CUDA;
        }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return lowbit::LB_OK;
}

}  // namespace

extern "C" int lowbit_matmul_f32_noalloc_cuda(const void* d_w, int qtype, int64_t n_rows, int64_t k, const float* d_x,
                                              int64_t m, void* d_y, int y_dtype, const void* d_bias, int accumulate,
                                              void* stream) {
    using namespace lowbit;
    reset_launch_count();
    if (m == 0 || n_rows == 0) return LB_OK;
    if (m < 0 || k <= 0 || (uint64_t)m > kNoAllocMaxElems / (uint64_t)k) {       // no m * k product: it could wrap
        set_error(LB_ERR_ARG, "lowbit_matmul_f32_noalloc_cuda: m * k = %lld x %lld exceeds the built-in scratch (%zu activation "
                  "elements); use lowbit_matmul_cuda with a caller workspace", (long long)m, (long long)k, kNoAllocMaxElems);
        return LB_ERR_ARG;
    }
    void* scratch = nullptr;
    int rc = noalloc_scratch(&scratch);          // first call on a device resolves the symbol address
    if (rc) return rc;
    // every element of the planes this call reads is written by its own split pass first, so nothing in
    // the shared array has to start at zero and nothing left by an earlier call is ever read
    rc = gemv_launch(d_w, qtype, n_rows, k, d_x, DT_F32, m, d_y, y_dtype, d_bias, accumulate, scratch,
                     sizeof(g_noalloc_scratch), (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_matmul_f32_noalloc_cuda: bad argument (qtype %d, N %lld, K %lld, M %lld)", qtype,
                                    (long long)n_rows, (long long)k, (long long)m);
    else if (rc == LB_ERR_WORKSPACE) set_error(rc, "lowbit_matmul_f32_noalloc_cuda: built-in scratch too small for M %lld, K %lld",
                                               (long long)m, (long long)k);
    return rc;
}
