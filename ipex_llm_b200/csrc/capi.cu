// capi.cu -- the extern "C" surface declared in include/b200_lowbit.h.
//
// Family (1): reference-named symbols with host pointers (reference
// ggml/model/llama/llama_cpp.py:945-1185).  They stage through library-owned device buffers,
// run the same CUDA kernels as family (2) and copy the result back.  No CPU arithmetic: if
// CUDA is unusable the call records LB_ERR_NO_DEVICE and returns without writing.
// Family (2): *_cuda symbols on device pointers + stream.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include "../../include/b200_lowbit.h"
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {

static thread_local int t_err = 0;
static thread_local char t_msg[512] = "";
static thread_local int t_launches = 0;
static thread_local int t_row_len = 0;

void set_error(int code, const char* fmt, ...) {
    t_err = code;
    va_list ap; va_start(ap, fmt);
    vsnprintf(t_msg, sizeof(t_msg), fmt, ap);
    va_end(ap);
}
void note_launch() { t_launches++; }
void reset_launch_count() { t_launches = 0; }
int check_launch() {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error(LB_ERR_CUDA, "CUDA error: %s", cudaGetErrorString(e)); return LB_ERR_CUDA; }
    return LB_OK;
}
int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return 148; }
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 148; }
        cached[dev] = n;
    }
    return cached[dev];
}

static bool have_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return false; }
    return true;
}

// ---- library-owned per-device buffers: staging for family (1), default matmul workspace ----
struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    int ensure(size_t bytes, bool zero_new) {
        if (bytes <= cap) return LB_OK;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { set_error(LB_ERR_CUDA, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); cudaGetLastError(); return LB_ERR_CUDA; }
        cap = want;
        if (zero_new) cudaMemset(p, 0, want);
        return LB_OK;
    }
};
struct DevState { DevBuf in, out, aux, ws; };
static std::mutex g_mu;
static DevState g_dev[64];
static DevState* dev_state() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return nullptr; }
    return &g_dev[dev];
}

#define HOST_PROLOGUE(retval)                                                              \
    t_launches = 0;                                                                        \
    if (!have_device()) {                                                                  \
        set_error(LB_ERR_NO_DEVICE, "%s: no CUDA device (this library has no CPU path)", __func__); \
        return retval;                                                                     \
    }                                                                                      \
    std::lock_guard<std::mutex> lk(g_mu);                                                  \
    DevState* S = dev_state();                                                             \
    if (!S) { set_error(LB_ERR_NO_DEVICE, "%s: cudaGetDevice failed", __func__); return retval; }

static bool cuda_ok(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    set_error(LB_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
    cudaGetLastError();
    return false;
}

}  // namespace lowbit

using namespace lowbit;

extern "C" {

void llama_init_backend(void) {}

size_t ggml_type_size(int qtype) { return (size_t)type_size(qtype); }
int ggml_qk_size(int qtype) { return supported(qtype) ? QK : 0; }

int b200_lowbit_last_error(void) { return t_err; }
const char* b200_lowbit_last_error_string(void) { return t_msg; }
void b200_lowbit_clear_error(void) { t_err = 0; t_msg[0] = 0; }
int b200_lowbit_last_launch_count(void) { return t_launches; }
int b200_lowbit_sm_count(void) { return have_device() ? sm_count() : 0; }
void b200_lowbit_set_row_length(int k) { t_row_len = k; }
void lowbit_set_programmatic_dependent_launch(int on) { gemv_set_pdl(on); }
int lowbit_get_programmatic_dependent_launch(void) { return gemv_get_pdl(); }
void lowbit_hint_next_weights(const void* d_w, size_t bytes) { gemv_set_next_weights(d_w, bytes); }
/* development aid, not in the public header: per-warp phase timestamps of the GEMV kernel */
void lowbit_debug_set_timestamps(void* dev_ptr) { gemv_set_debug_timestamps(dev_ptr); }

// ------------------------------------------------------------------ family (2)
int ggml_quantize_tensor_cuda(const float* d_src, void* d_dst, int qtype, size_t n, int k, int64_t* d_hist,
                              bool scale_search, void* stream) {
    t_launches = 0;
    int rc = quantize_launch(d_src, d_dst, qtype, n, k, d_hist, scale_search, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "ggml_quantize_tensor_cuda: bad argument (qtype %d, n %zu, k %d)", qtype, n, k);
    return rc;
}

int ggml_dequantize_cuda(const void* d_src, void* d_dst, size_t k, int qtype, int out_dtype, void* stream) {
    t_launches = 0;
    int rc = dequant_canonical_launch(d_src, d_dst, k, qtype, out_dtype, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "ggml_dequantize_cuda: bad argument (qtype %d, k %zu)", qtype, k);
    return rc;
}

int lowbit_dequantize_device_layout_cuda(const void* d_src, void* d_dst, int64_t rows, int64_t k, int qtype,
                                         int out_dtype, void* stream) {
    t_launches = 0;
    int rc = dequant_tiled_launch(d_src, d_dst, rows, k, qtype, out_dtype, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_dequantize_device_layout_cuda: bad argument");
    return rc;
}

int lowbit_dequantize_rows_cuda(const void* d_w, int qtype, int64_t rows, int64_t k, const void* d_ids, int ids_dtype,
                                int64_t n_ids, void* d_out, int out_dtype, void* stream) {
    t_launches = 0;
    if (!d_w || (n_ids > 0 && (!d_ids || !d_out)) || (ids_dtype != 0 && ids_dtype != 1)) {
        set_error(LB_ERR_ARG, "lowbit_dequantize_rows_cuda: null pointer or bad ids_dtype");
        return LB_ERR_ARG;
    }
    int rc = dequant_rows_launch(d_w, d_out, d_ids, ids_dtype, n_ids, rows, k, qtype, out_dtype, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_dequantize_rows_cuda: bad argument (rows %lld, k %lld, qtype %d)", (long long)rows, (long long)k, qtype);
    return rc;
}

int lowbit_relayout_cuda(const void* d_src, void* d_dst, int64_t rows, int64_t k, int qtype, int to_device,
                         void* stream) {
    t_launches = 0;
    int rc = relayout_launch(d_src, d_dst, rows, k, qtype, to_device != 0, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_relayout_cuda: bad argument (rows %lld, k %lld, qtype %d)", (long long)rows, (long long)k, qtype);
    return rc;
}

// fp32 activations never take the tensor-core GEMM: it rounds them to one bf16 plane (2^-9), while the GEMV keeps
// them to 2^-24 (three planes) -- the fp32-in / fp32-out contract of ggml_compute_forward_mul_mat_q_fp32
// (LBL:494-529).  Larger fp32 batches run as chunks of the GEMV; a caller that accepts half precision casts first,
// as the reference's XPU path does (LBL:700-705, conver_to_half).
static bool use_gemm(int qtype, int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    static const int64_t max_m = [] { const char* e = getenv("LOWBIT_GEMV_MAX_M"); return e ? (int64_t)atoi(e) : kGemvMaxM; }();   // development
    return x_dtype != DT_F32 && m > max_m && gemm_tc_supported(qtype, n_rows, k);
}

// Batches between the GEMV's best range and a full 128-token tile take the tensor-core kernel with the operands swapped
// (gemm_skinny_launch): weights rounded once to the activation dtype, the prefill kernel's precision contract.
// Measured on 11008x4096 sym_int4: GEMV 12 us at M = 16 and 22 us at M = 32, this kernel 15.5 us for 9 <= M <= 32, 17-20 us
// up to 64 (the split-K prefill kernel: 22-25 us), 29 us at 96-128 -- hence the window 17..64.
// LOWBIT_SKINNY_MIN_M / LOWBIT_SKINNY_MAX_M (development) move the window; MIN_M above MAX_M switches the kernel off.
static bool use_skinny(int qtype, int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    static const int64_t min_m = [] { const char* e = getenv("LOWBIT_SKINNY_MIN_M"); return e ? (int64_t)atoi(e) : (int64_t)17; }();
    static const int64_t max_m = [] { const char* e = getenv("LOWBIT_SKINNY_MAX_M"); return e ? (int64_t)atoi(e) : (int64_t)64; }();
    return m >= min_m && m <= max_m && gemm_skinny_supported(qtype, n_rows, k, m, x_dtype);
}

size_t lowbit_matmul_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    if (use_skinny(B200_QT_SYM_INT4, n_rows, k, m, x_dtype)) return gemm_skinny_workspace_size(n_rows, k, m);
    if (use_gemm(B200_QT_SYM_INT4, n_rows, k, m, x_dtype)) return gemm_tc_workspace_size(n_rows, k, m, x_dtype);
    return gemv_workspace_size(n_rows, k, m, x_dtype);
}

static int matmul_dispatch(const void* d_w, int qtype, int64_t n_rows, int64_t k, const void* d_x, int x_dtype, int64_t m,
                           void* d_y, int y_dtype, const void* d_bias, int accumulate, void* workspace, size_t workspace_bytes,
                           cudaStream_t st) {
    if (use_skinny(qtype, n_rows, k, m, x_dtype))
        return gemm_skinny_launch(d_w, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias, accumulate, workspace,
                                  workspace_bytes, st, gemv_get_pdl());
    if (use_gemm(qtype, n_rows, k, m, x_dtype)) {
        int64_t done = 0;
        return gemm_tc_launch(d_w, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias, accumulate, workspace,
                              workspace_bytes, st, &done);
    }
    return gemv_launch(d_w, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias, accumulate, workspace, workspace_bytes, st);
}

int lowbit_matmul_cuda(const void* d_w, int qtype, int64_t n_rows, int64_t k, const void* d_x, int x_dtype,
                       int64_t m, void* d_y, int y_dtype, const void* d_bias, int accumulate, void* workspace,
                       size_t workspace_bytes, void* stream) {
    t_launches = 0;
    if (m == 0 || n_rows == 0) return LB_OK;
    std::unique_lock<std::mutex> lk(g_mu, std::defer_lock);
    if (!workspace) {
        lk.lock();
        DevState* S = dev_state();
        if (!S) { set_error(LB_ERR_NO_DEVICE, "lowbit_matmul_cuda: no CUDA device"); return LB_ERR_NO_DEVICE; }
        const size_t need = lowbit_matmul_workspace_size(n_rows, k, m, x_dtype);
        if (need > S->ws.cap) {
            // growing frees the old buffer: wait for work that may still use it
            cudaDeviceSynchronize();
            int rc = S->ws.ensure(need, true);
            if (rc) return rc;
        }
        workspace = S->ws.p; workspace_bytes = S->ws.cap;
    }
    int rc = matmul_dispatch(d_w, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias, accumulate, workspace,
                             workspace_bytes, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_matmul_cuda: bad argument (qtype %d, N %lld, K %lld, M %lld)", qtype, (long long)n_rows, (long long)k, (long long)m);
    if (rc == LB_ERR_WORKSPACE) set_error(rc, "lowbit_matmul_cuda: workspace too small (%zu bytes)", workspace_bytes);
    return rc;
}

size_t lowbit_mlp_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype) {
    const size_t base = (lowbit_matmul_workspace_size(n_rows, k, m, x_dtype) + 255) / 256 * 256;
    const size_t two = base + 2 * (size_t)m * n_rows * sizeof(float);
    const size_t sk = use_skinny(B200_QT_SYM_INT4, n_rows, k, m, x_dtype) ? gemm_skinny_mlp_workspace_size(n_rows, k, m) : 0;
    return two > sk ? two : sk;
}

int lowbit_mlp_cuda(const void* d_w_gate, const void* d_w_up, int qtype, int64_t n_rows, int64_t k, const void* d_x,
                    int x_dtype, int64_t m, void* d_y, int y_dtype, int act, void* workspace, size_t workspace_bytes,
                    void* stream) {
    t_launches = 0;
    if (m == 0 || n_rows == 0) return LB_OK;
    if (act != 0 && act != 1) { set_error(LB_ERR_ARG, "lowbit_mlp_cuda: act must be 0 (SILU) or 1 (GELU), got %d", act); return LB_ERR_ARG; }
    if (!d_w_gate || !d_w_up || !d_x || !d_y || m < 0 || n_rows < 0) { set_error(LB_ERR_ARG, "lowbit_mlp_cuda: null pointer or negative size"); return LB_ERR_ARG; }
    const bool fused = gemv_mlp_fused_supported(qtype, n_rows, k, m, x_dtype);
    const size_t base = (lowbit_matmul_workspace_size(n_rows, k, m, x_dtype) + 255) / 256 * 256;
    const size_t need = fused ? base : lowbit_mlp_workspace_size(n_rows, k, m, x_dtype);
    std::unique_lock<std::mutex> lk(g_mu, std::defer_lock);
    if (!workspace) {
        lk.lock();
        DevState* S = dev_state();
        if (!S) { set_error(LB_ERR_NO_DEVICE, "lowbit_mlp_cuda: no CUDA device"); return LB_ERR_NO_DEVICE; }
        if (need > S->ws.cap) {
            cudaDeviceSynchronize();
            int rc = S->ws.ensure(need, true);
            if (rc) return rc;
        }
        workspace = S->ws.p; workspace_bytes = S->ws.cap;
    } else if (workspace_bytes < need) {
        set_error(LB_ERR_WORKSPACE, "lowbit_mlp_cuda: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
        return LB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (fused) {
        rc = gemv_launch_ex(d_w_gate, d_w_up, act + 1, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, nullptr, 0, workspace,
                            workspace_bytes, st);
    } else if (use_skinny(qtype, n_rows, k, m, x_dtype) && workspace_bytes >= gemm_skinny_mlp_workspace_size(n_rows, k, m)) {
        // 17..64 half-precision rows: both matrices in one pass of the skinny tensor-core kernel, act * up in its reduction launch
        rc = gemm_skinny_mlp_launch(d_w_gate, d_w_up, act + 1, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, workspace, workspace_bytes,
                                    st, gemv_get_pdl());
    } else {
        // the two intermediates: fp32 behind the exact GEMV; behind the tensor-core kernels (whose contract rounds the weights to
        // the activation dtype anyway) in y's half-precision dtype -- half the bytes through the GEMM epilogue and the activation
        const bool tc = use_skinny(qtype, n_rows, k, m, x_dtype) || use_gemm(qtype, n_rows, k, m, x_dtype);
        const int t_dtype = (tc && y_dtype != DT_F32) ? y_dtype : DT_F32;
        char* tg = (char*)workspace + base;
        char* tu = tg + (size_t)m * n_rows * (t_dtype == DT_F32 ? 4 : 2);
        if (m > 128 && (n_rows % 128) == 0 && use_gemm(qtype, n_rows, k, m, x_dtype)) {
            // prefill sizes: gate and up from one grid of the tensor-core kernel (one set of rounds over the SMs instead of two)
            rc = gemm_tc_pair_launch(d_w_gate, d_w_up, qtype, n_rows, k, d_x, x_dtype, m, tg, tu, t_dtype, workspace, base, st);
        } else {
            rc = matmul_dispatch(d_w_gate, qtype, n_rows, k, d_x, x_dtype, m, tg, t_dtype, nullptr, 0, workspace, base, st);
            if (!rc) rc = matmul_dispatch(d_w_up, qtype, n_rows, k, d_x, x_dtype, m, tu, t_dtype, nullptr, 0, workspace, base, st);
        }
        if (!rc) rc = act_mul_launch(tg, tu, t_dtype, d_y, y_dtype, m * n_rows, act + 1, st);
    }
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_mlp_cuda: bad argument (qtype %d, N %lld, K %lld, M %lld)", qtype, (long long)n_rows, (long long)k, (long long)m);
    if (rc == LB_ERR_WORKSPACE) set_error(rc, "lowbit_mlp_cuda: workspace too small (%zu bytes)", workspace_bytes);
    return rc;
}

int lowbit_moe_supported(int qtype, int64_t hidden, int64_t inter, int top_k) {
    return moe_supported(qtype, hidden, inter, top_k) ? 1 : 0;
}

size_t lowbit_moe_workspace_size(int64_t hidden, int64_t inter, int top_k, int x_dtype) {
    return moe_workspace_size(hidden, inter, top_k, x_dtype);
}

int lowbit_moe_cuda(const void* d_x, int x_dtype, const void* d_topk_ids, int ids_dtype, const void* d_topk_weight,
                    int weight_dtype, int top_k, const uint64_t* d_gates, const uint64_t* d_ups, const uint64_t* d_downs,
                    int n_experts, int64_t hidden, int64_t inter, int qtype, int act, void* d_y, int y_dtype, int accumulate,
                    void* workspace, size_t workspace_bytes, void* stream) {
    t_launches = 0;
    if (!d_x || !d_topk_ids || !d_topk_weight || !d_gates || !d_ups || !d_downs || !d_y || n_experts <= 0 ||
        (ids_dtype != 0 && ids_dtype != 1) || (act != 0 && act != 1)) {
        set_error(LB_ERR_ARG, "lowbit_moe_cuda: null pointer, bad ids_dtype / act or n_experts <= 0");
        return LB_ERR_ARG;
    }
    if (!moe_supported(qtype, hidden, inter, top_k)) {
        set_error(LB_ERR_ARG, "lowbit_moe_cuda: unsupported problem (qtype %d, hidden %lld, inter %lld, top_k %d)", qtype,
                  (long long)hidden, (long long)inter, top_k);
        return LB_ERR_ARG;
    }
    const size_t need = moe_workspace_size(hidden, inter, top_k, x_dtype);
    std::unique_lock<std::mutex> lk(g_mu, std::defer_lock);
    if (!workspace) {
        lk.lock();
        DevState* S = dev_state();
        if (!S) { set_error(LB_ERR_NO_DEVICE, "lowbit_moe_cuda: no CUDA device"); return LB_ERR_NO_DEVICE; }
        if (need > S->ws.cap) {
            cudaDeviceSynchronize();
            int rc = S->ws.ensure(need, true);
            if (rc) return rc;
        }
        workspace = S->ws.p; workspace_bytes = S->ws.cap;
    }
    MoeParams P;
    P.x = d_x; P.x_dtype = x_dtype; P.ids = d_topk_ids; P.ids_i64 = ids_dtype; P.weights = d_topk_weight; P.w_dtype = weight_dtype;
    P.top_k = top_k; P.gates = (const unsigned long long*)d_gates; P.ups = (const unsigned long long*)d_ups;
    P.downs = (const unsigned long long*)d_downs; P.n_experts = n_experts; P.hidden = hidden; P.inter = inter;
    P.qtype = qtype; P.act = act + 1; P.y = d_y; P.y_dtype = y_dtype; P.accumulate = accumulate;
    P.workspace = workspace; P.workspace_bytes = workspace_bytes;
    int rc = moe_launch(P, (cudaStream_t)stream);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_moe_cuda: bad argument (dtype codes or alignment)");
    if (rc == LB_ERR_WORKSPACE) set_error(rc, "lowbit_moe_cuda: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
    return rc;
}

// ------------------------------------------------------------------ family (1)
size_t ggml_quantize_tensor(const float* src, void* dst, int qtype, size_t n, int k, int64_t* hist,
                            bool scale_search) {
    HOST_PROLOGUE(0)
    const int bs = type_size(qtype);
    if (!bs || k <= 0 || k % QK || n % QK) { set_error(LB_ERR_ARG, "ggml_quantize_tensor: bad argument (qtype %d, n %zu, k %d)", qtype, n, k); return 0; }
    const size_t out_bytes = n / QK * bs;
    if (n == 0) return 0;
    if (S->in.ensure(n * sizeof(float), false) || S->out.ensure(out_bytes, false) || S->aux.ensure(256, false)) return 0;
    if (!cuda_ok(cudaMemcpy(S->in.p, src, n * sizeof(float), cudaMemcpyHostToDevice), "H2D")) return 0;
    if (quantize_launch((const float*)S->in.p, S->out.p, qtype, n, k, hist ? (int64_t*)S->aux.p : nullptr, scale_search, 0)) return 0;
    if (!cuda_ok(cudaMemcpy(dst, S->out.p, out_bytes, cudaMemcpyDeviceToHost), "D2H")) return 0;
    if (hist && !cuda_ok(cudaMemcpy(hist, S->aux.p, 16 * sizeof(int64_t), cudaMemcpyDeviceToHost), "D2H hist")) return 0;
    return out_bytes;
}

void ggml_dequantize(const void* src, void* dst, size_t k, int qtype) {
    HOST_PROLOGUE()
    const int bs = type_size(qtype);
    if (!bs || k % QK) { set_error(LB_ERR_ARG, "ggml_dequantize: bad argument (qtype %d, k %zu)", qtype, k); return; }
    if (k == 0) return;
    const size_t in_bytes = k / QK * bs;
    if (S->in.ensure(in_bytes, false) || S->out.ensure(k * sizeof(float), false)) return;
    if (!cuda_ok(cudaMemcpy(S->in.p, src, in_bytes, cudaMemcpyHostToDevice), "H2D")) return;
    if (dequant_canonical_launch(S->in.p, S->out.p, k, qtype, DT_F32, 0)) return;
    cuda_ok(cudaMemcpy(dst, S->out.p, k * sizeof(float), cudaMemcpyDeviceToHost), "D2H");
}

void ggml_dequantize_q4_0(const void* src, void* dst, size_t k) { ggml_dequantize(src, dst, k, QT_SYM_INT4); }

static int relayout_host(const void* src, void* dst, int64_t rows, int64_t k, int qtype, bool to_device, const char* who) {
    HOST_PROLOGUE(LB_ERR_NO_DEVICE)
    const int bs = type_size(qtype);
    if (!bs || k <= 0 || k % QK || rows <= 0) { set_error(LB_ERR_ARG, "%s: bad argument (rows %lld, k %lld, qtype %d)", who, (long long)rows, (long long)k, qtype); return LB_ERR_ARG; }
    const size_t bytes = (size_t)rows * (k / QK) * bs;
    if (S->in.ensure(bytes, false) || S->out.ensure(bytes, false)) return LB_ERR_CUDA;
    if (!cuda_ok(cudaMemcpy(S->in.p, src, bytes, cudaMemcpyHostToDevice), "H2D")) return LB_ERR_CUDA;
    int rc = relayout_launch(S->in.p, S->out.p, rows, k, qtype, to_device, 0);
    if (rc) return rc;
    if (!cuda_ok(cudaMemcpy(dst, S->out.p, bytes, cudaMemcpyDeviceToHost), "D2H")) return LB_ERR_CUDA;
    return LB_OK;
}

int ggml_q_format_convert_cpu2cuda_2d(const void* src, void* dst, int64_t rows, int64_t k, int qtype) {
    return relayout_host(src, dst, rows, k, qtype, true, "ggml_q_format_convert_cpu2cuda_2d");
}
int ggml_q_format_convert_cuda2cpu_2d(const void* src, void* dst, int64_t rows, int64_t k, int qtype) {
    return relayout_host(src, dst, rows, k, qtype, false, "ggml_q_format_convert_cuda2cpu_2d");
}

static void relayout_shapeless(const void* src, void* dst, size_t n, int qtype, bool to_device, const char* who) {
    const int k = t_row_len;
    if (k <= 0 || n % (size_t)k) {
        set_error(LB_ERR_ARG, "%s: the B200 device layout is tiled by rows; call b200_lowbit_set_row_length(k) first "
                  "(n = %zu, registered row length = %d)", who, n, k);
        return;
    }
    relayout_host(src, dst, (int64_t)(n / (size_t)k), k, qtype, to_device, who);
}
void ggml_q_format_convet_cpu2xpu(const void* src, void* dst, size_t n, int qtype) {
    relayout_shapeless(src, dst, n, qtype, true, "ggml_q_format_convet_cpu2xpu");
}
void ggml_q_format_convet_xpu2cpu(const void* src, void* dst, size_t n, int qtype) {
    relayout_shapeless(src, dst, n, qtype, false, "ggml_q_format_convet_xpu2cpu");
}

void ggml_compute_forward_mul_mat_q_fp32(const int64_t* src0_ne, const void* w, int qtype, const int64_t* src1_ne,
                                         const void* x, void* result) {
    HOST_PROLOGUE()
    const int64_t K = src0_ne[0], N = src0_ne[1], M = src1_ne[1];
    const int bs = type_size(qtype);
    if (!bs || K <= 0 || K % QK || src1_ne[0] != K || N <= 0 || M < 0) {
        set_error(LB_ERR_ARG, "ggml_compute_forward_mul_mat_q_fp32: bad shapes (K %lld, N %lld, M %lld, qtype %d)", (long long)K, (long long)N, (long long)M, qtype);
        return;
    }
    if (M == 0) return;
    const size_t w_bytes = (size_t)N * (K / QK) * bs, x_bytes = (size_t)M * K * 4, y_bytes = (size_t)M * N * 4;
    const size_t ws_bytes = lowbit_matmul_workspace_size(N, K, M, DT_F32);
    // in: [canonical w][x]   out: [tiled w]   aux: y
    const size_t w_al = (w_bytes + 255) / 256 * 256;
    if (S->in.ensure(w_al + x_bytes, false) || S->out.ensure(w_bytes, false) || S->aux.ensure(y_bytes, false)) return;
    if (ws_bytes > S->ws.cap) { cudaDeviceSynchronize(); if (S->ws.ensure(ws_bytes, true)) return; }
    if (!cuda_ok(cudaMemcpy(S->in.p, w, w_bytes, cudaMemcpyHostToDevice), "H2D w")) return;
    if (!cuda_ok(cudaMemcpy((char*)S->in.p + w_al, x, x_bytes, cudaMemcpyHostToDevice), "H2D x")) return;
    if (relayout_launch(S->in.p, S->out.p, N, K, qtype, true, 0)) return;
    if (matmul_dispatch(S->out.p, qtype, N, K, (char*)S->in.p + w_al, DT_F32, M, S->aux.p, DT_F32, nullptr, 0, S->ws.p, S->ws.cap, 0)) {
        if (!t_err) set_error(LB_ERR_CUDA, "ggml_compute_forward_mul_mat_q_fp32: launch failed");
        return;
    }
    cuda_ok(cudaMemcpy(result, S->aux.p, y_bytes, cudaMemcpyDeviceToHost), "D2H y");
}

}  // extern "C"
