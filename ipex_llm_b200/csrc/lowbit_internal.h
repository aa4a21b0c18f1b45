// lowbit_internal.h -- declarations shared between the .cu translation units of
// libb200lowbit.so (not installed; the public ABI is include/b200_lowbit.h).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace lowbit {

enum {
    LB_OK = 0,
    LB_ERR_ARG = 1,        // bad qtype / shape / pointer combination
    LB_ERR_CUDA = 2,       // a CUDA runtime call failed (see last_error_string)
    LB_ERR_NO_DEVICE = 3,  // no usable CUDA device: there is no CPU fallback
    LB_ERR_WORKSPACE = 4,  // caller workspace too small
};

// per-thread error / launch bookkeeping (capi.cu)
void set_error(int code, const char* fmt, ...);
void note_launch();
void reset_launch_count();
int check_launch();   // cudaGetLastError() -> LB_OK / LB_ERR_CUDA (records the message)
int sm_count();       // SMs of the current device (cached per device)

// layout.cu
int relayout_launch(const void* src, void* dst, int64_t rows, int64_t k, int qt, bool to_device,
                    cudaStream_t st);
int dequant_canonical_launch(const void* src, void* dst, size_t n, int qt, int out_dtype, cudaStream_t st);
int dequant_tiled_launch(const void* src, void* dst, int64_t rows, int64_t k, int qt, int out_dtype,
                         cudaStream_t st);

int dequant_rows_launch(const void* src, void* dst, const void* ids, int ids_i64, int64_t n_ids, int64_t rows, int64_t k,
                        int qt, int out_dtype, cudaStream_t st);

// quantize.cu
int quantize_launch(const float* src, void* dst, int qt, size_t n, int k, int64_t* hist, bool scale_search,
                    cudaStream_t st);

// gemv.cu
void gemv_set_pdl(int on);
int gemv_get_pdl();
struct GemvPush;
void gemv_set_push(const GemvPush* p);   // the next gemv_launch_ex on this thread pushes its result to the peers (comm.cu)
void gemv_set_next_weights(const void* w, size_t bytes);
void gemv_set_debug_timestamps(void* p);
size_t gemv_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype);
int gemv_launch(const void* w, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m,
                void* y, int y_dtype, const void* bias, int accumulate, void* workspace,
                size_t workspace_bytes, cudaStream_t st);

int gemv_launch_ex(const void* w, const void* w2, int act, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype,
                   int64_t m, void* y, int y_dtype, const void* bias, int accumulate, void* workspace,
                   size_t workspace_bytes, cudaStream_t st);
bool gemv_mlp_fused_supported(int qt, int64_t n_rows, int64_t k, int64_t m, int x_dtype);
int act_mul_launch(const void* gate, const void* up, int in_dtype, void* y, int y_dtype, int64_t n, int act, cudaStream_t st);

// one-token mixture-of-experts (reference xe_linear.moe_forward_vec)
struct MoeParams {
    const void* x; int x_dtype;                  // [hidden]
    const void* ids; int ids_i64;                // [top_k] routed expert ids (device)
    const void* weights; int w_dtype;            // [top_k] routing weights (device)
    int top_k;
    const unsigned long long* gates;             // [n_experts] device addresses of the tiled [inter][hidden] weights
    const unsigned long long* ups;
    const unsigned long long* downs;             // ... of the tiled [hidden][inter] weights
    int n_experts;
    int64_t hidden, inter;
    int qtype, act;
    void* y; int y_dtype; int accumulate;        // [hidden]
    void* workspace; size_t workspace_bytes;
};
bool moe_supported(int qt, int64_t hidden, int64_t inter, int top_k);
size_t moe_workspace_size(int64_t hidden, int64_t inter, int top_k, int x_dtype);
int moe_launch(const MoeParams& P, cudaStream_t st);

// gemm_tc.cu (tcgen05 prefill GEMM)
bool gemm_tc_supported(int qt, int64_t n_rows, int64_t k);
size_t gemm_tc_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype);
int gemm_tc_launch(const void* w, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m, void* y,
                   int y_dtype, const void* bias, int accumulate, void* workspace, size_t workspace_bytes, cudaStream_t st,
                   int64_t* n_done);
// gate and up of an MLP in one grid (m > 128, n_rows % 128 == 0)
int gemm_tc_pair_launch(const void* w, const void* w2, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m,
                        void* y, void* y2, int y_dtype, void* workspace, size_t workspace_bytes, cudaStream_t st);

// M above this goes to the tensor-core GEMM when the shape allows, else to chunks of the GEMV
constexpr int64_t kGemvMaxM = 32;
// the tensor-core GEMM with the operands swapped (weights = the 128-row operand, tokens = N): 9 <= m <= 128
bool gemm_skinny_supported(int qt, int64_t n_rows, int64_t k, int64_t m, int x_dtype);
size_t gemm_skinny_workspace_size(int64_t n_rows, int64_t k, int64_t m);
size_t gemm_skinny_mlp_workspace_size(int64_t n_rows, int64_t k, int64_t m);
int gemm_skinny_mlp_launch(const void* w_gate, const void* w_up, int act, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype,
                           int64_t m, void* y, int y_dtype, void* workspace, size_t workspace_bytes, cudaStream_t st, int pdl);
int gemm_skinny_launch(const void* w, int qt, int64_t n_rows, int64_t k, const void* x, int x_dtype, int64_t m, void* y, int y_dtype,
                       const void* bias, int accumulate, void* workspace, size_t workspace_bytes, cudaStream_t st, int pdl);

}  // namespace lowbit
