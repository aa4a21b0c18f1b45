/*
 * b200_lowbit.h -- C ABI of libb200lowbit.so: the B200 (sm_100a) replacement for the
 * low-bit linear entry points ipex-llm binds from its closed `libllama-api.so`.
 *
 * Reference paths are relative to /root/reference/python/llm/src/ipex_llm/ :
 *   LC = ggml/model/llama/llama_cpp.py      LBL = transformers/low_bit_linear.py
 *
 * Two families:
 *  (1) reference-named symbols with the reference's exact signatures.  They take HOST
 *      pointers (that is what the reference passes), copy to the GPU, run the CUDA
 *      kernels and copy back.  There is NO CPU implementation behind them: without a
 *      usable CUDA device they record an error (b200_lowbit_last_error) and write nothing.
 *  (2) `*_cuda` symbols taking DEVICE pointers and a cudaStream_t (as void*); they only
 *      enqueue work on that stream and return 0 or a non-zero error code.  The Python
 *      host (ipex_llm_b200/transformers/low_bit_linear.py) uses these.
 *
 * Ownership: the caller allocates every buffer it passes; the library keeps no pointer
 * after the call returns (LBL:113-148, :506).  Library-owned state is a per-device
 * staging/workspace cache used only by family (1) and by lowbit_matmul_cuda when the
 * caller passes workspace == NULL.
 */
#ifndef B200_LOWBIT_H
#define B200_LOWBIT_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

/* qtype values: ggml/quantize.py:28-60.  In scope: */
enum {
    B200_QT_SYM_INT4 = 2,   /* "sym_int4"  (q4_0 lineage), 34-byte blocks of 64 */
    B200_QT_ASYM_INT4 = 3,  /* "asym_int4" (q4_1 lineage), 36-byte blocks of 64 */
    B200_QT_NF4 = 10,       /* "nf4",      34-byte blocks of 64 */
    B200_QT_FP8_E4M3 = 15,  /* "fp8_e4m3", 66-byte blocks of 64 */
    B200_QT_FP8_E5M2 = 19,  /* "fp8_e5m2" / "fp8", 66-byte blocks of 64 */
    B200_QT_FP6 = 29,       /* "fp6" (e3m2), 50-byte blocks of 64 */
    B200_QT_WOQ_INT4 = 34   /* alias of sym_int4 (convert.py:1101-1109 on torch >= 2.6 string compare) */
};

/* element types of activations / outputs for the *_cuda matmul entry points */
enum { B200_DT_F32 = 0, B200_DT_F16 = 1, B200_DT_BF16 = 2 };

/* ------------------------------------------------------------------ family (1) */

/* replaces LC:1189-1193 llama_init_backend(): selects nothing, allocates nothing. */
void llama_init_backend(void);

/* replaces LC:1050-1056 ggml_type_size: bytes per block record; 0 for unsupported qtypes */
size_t ggml_type_size(int qtype);
/* replaces LC:1059-1065 ggml_qk_size: elements per block; 0 for unsupported qtypes */
int ggml_qk_size(int qtype);

/* replaces LC:946-967 ggml_quantize_tensor (called at LBL:180).
 * src: n fp32 values, row length k (k % QK == 0); dst: (n/QK)*type_size bytes in the
 * canonical block layout; hist: 16 x int64 code histogram for the 4-bit types (may be NULL);
 * returns bytes written, 0 on error. */
size_t ggml_quantize_tensor(const float* src, void* dst, int qtype, size_t n, int k,
                            int64_t* hist, bool scale_search);

/* replaces LC:1068-1080 ggml_dequantize_q4_0 (LBL:256-266): sym_int4 blocks -> k fp32 */
void ggml_dequantize_q4_0(const void* src, void* dst, size_t k);
/* replaces LC:1084-1099 ggml_dequantize (LBL:269-279): any in-scope qtype -> k fp32 */
void ggml_dequantize(const void* src, void* dst, size_t k, int qtype);

/* replace LC:1102-1135 ggml_q_format_convet_cpu2xpu / xpu2cpu (LBL:198-253): same-size
 * relayout canonical <-> device layout.  The B200 device layout is tiled by 16 rows, so the
 * row length must be known: these reference-shaped symbols use the row length registered
 * by b200_lowbit_set_row_length() on the calling thread (error if unset or if n is not a
 * multiple of it).  New code should call the *_2d forms below. */
void ggml_q_format_convet_cpu2xpu(const void* src, void* dst, size_t n, int qtype);
void ggml_q_format_convet_xpu2cpu(const void* src, void* dst, size_t n, int qtype);
void b200_lowbit_set_row_length(int k);
int ggml_q_format_convert_cpu2cuda_2d(const void* src, void* dst, int64_t rows, int64_t k, int qtype);
int ggml_q_format_convert_cuda2cpu_2d(const void* src, void* dst, int64_t rows, int64_t k, int qtype);

/* replaces LC:1154-1185 ggml_compute_forward_mul_mat_q_fp32 (LBL:494-529).
 * src0_ne = (K, N), src1_ne = (K, M) (LBL:509-516); w = canonical packed [N][K/QK] blocks;
 * x = fp32 [M][K]; result = fp32 [M][N].  All host pointers. */
void ggml_compute_forward_mul_mat_q_fp32(const int64_t* src0_ne, const void* w, int qtype,
                                         const int64_t* src1_ne, const void* x, void* result);

/* error channel for the void / size_t symbols above (the reference ABI has none, LBL checks
 * preconditions in Python).  0 = ok.  The string is valid until the next call on the thread. */
int b200_lowbit_last_error(void);
const char* b200_lowbit_last_error_string(void);
void b200_lowbit_clear_error(void);

/* ------------------------------------------------------------------ family (2) */

/* device-pointer twin of ggml_quantize_tensor; d_hist may be NULL. */
int ggml_quantize_tensor_cuda(const float* d_src, void* d_dst, int qtype, size_t n, int k,
                              int64_t* d_hist, bool scale_search, void* stream);
/* device-pointer twin of ggml_dequantize over the CANONICAL layout; out_dtype B200_DT_* */
int ggml_dequantize_cuda(const void* d_src, void* d_dst, size_t k, int qtype, int out_dtype,
                         void* stream);
/* dequantize a [rows][k] weight held in the DEVICE (tiled) layout */
int lowbit_dequantize_device_layout_cuda(const void* d_src, void* d_dst, int64_t rows, int64_t k,
                                         int qtype, int out_dtype, void* stream);
/* rows of a device-layout [rows][k] weight picked by an index vector and dequantised -- the XPU op
 * `xe_linear.dequantize_rows(ids, qweight, qtype, embedding_dim, num_embeddings)` behind LowBitEmbedding.forward
 * (transformers/embedding.py:214-224).  d_ids: n_ids indices (ids_dtype 0 = int32, 1 = int64) read on the
 * device; an index outside [0, rows) yields a zero row.  d_out: [n_ids][k] of out_dtype.  Same arithmetic as
 * the dequantise kernels (bit-exact against the oracle). */
int lowbit_dequantize_rows_cuda(const void* d_w, int qtype, int64_t rows, int64_t k, const void* d_ids,
                                int ids_dtype, int64_t n_ids, void* d_out, int out_dtype, void* stream);
/* ---- FP8 key/value cache: DynamicFp8Cache (transformers/kv.py:33-77) keeps K and V as uint8
 * [batch, kv_heads, len, head_dim].  The reference's xe_addons kernels are closed; the byte format here is
 * BUILD-DEFINED: one e5m2 code per element (round to nearest even, saturating), no scales.
 * Tensors are 4-D views [B][H][S][D] with D contiguous; strides[3] = element strides of B, H, S (host array).
 * lowbit_kv_quantize_cuda   = xe_addons.quantize_key_value   (models/utils.py:132-134), one call per tensor
 * lowbit_kv_dequantize_cuda = xe_addons.dequantize_key_value (models/utils.py:181-187) */
int lowbit_kv_quantize_cuda(const void* d_src, int src_dtype, const int64_t* src_strides, void* d_dst,
                            const int64_t* dst_strides, int64_t B, int64_t H, int64_t S, int64_t D, void* stream);
int lowbit_kv_dequantize_cuda(const void* d_src, const int64_t* src_strides, void* d_dst, int dst_dtype,
                              const int64_t* dst_strides, int64_t B, int64_t H, int64_t S, int64_t D, void* stream);
/* All-reduce(sum) of the [M, hidden] results of the row-parallel linears in tensor-parallel decode -- the
 * torch.distributed / DeepSpeed all_reduce of low_bit_linear.py:715-725 (after o_proj and down_proj), for one process
 * per GPU on one node.  Every rank creates one buffer (lowbit_comm_create: device memory + a 64-byte IPC handle),
 * the handles are exchanged by the host (torch.distributed all_gather), every rank maps the peers' buffers
 * (lowbit_comm_open) and passes the world's pointers, own buffer at [rank], to every call.  The kernel writes this
 * rank's vector into its inbox in every peer's buffer (8-byte stores over NVLink carrying 4 bytes of data and the call
 * number), polls (bounded) its own inboxes for the peers' words and sums all vectors in rank order: every rank gets the
 * same bits.  n * sizeof(dtype) must be a multiple of 16 and at most max_bytes (<= lowbit_comm_max_bytes() = 256 KiB:
 * decode-sized vectors; larger tensors belong to NCCL); d_in == d_out is allowed.  Every rank issues the same sequence
 * of calls.  Call counters live in device memory (they skip 0, the fresh buffer's state, when they wrap): a captured
 * CUDA graph replays.  A wait is bounded (LOWBIT_COMM_TIMEOUT_S seconds, default 20): if a peer never calls, the words
 * waited for become NaN, the communicator's status word is set for good and every later call on it writes NaN without
 * waiting -- a stalled rank fails loudly instead of yielding a plausible wrong sum.  lowbit_comm_status reports 1 then
 * (it synchronises the device); destroy the communicator and create a new one to go on. */
size_t lowbit_comm_max_bytes(void);
int lowbit_comm_create(size_t max_bytes, void** d_buffer, void* ipc_handle_64);
int lowbit_comm_open(const void* ipc_handle_64, void** d_peer);
int lowbit_comm_close(void* d_peer);
int lowbit_comm_destroy(void* d_buffer);
int lowbit_comm_status(const void* d_buffer, int* status);
int lowbit_allreduce_sum_cuda(void* const* peer_buffers, int rank, int world, size_t max_bytes, const void* d_in, void* d_out,
                              int64_t n, int dtype, void* stream);
/* Row-parallel linear and its all-reduce in ONE kernel (M <= 32, n_rows % 64 == 0, M * n_rows * 4 <= max_bytes): the
 * decode GEMV's leader CTAs send their fp32 partial sums -- tagged with the communicator's call number -- into every
 * peer's inbox as each 64-row group finishes, wait (bounded) for the peers' words of the same rows in the own buffer, add
 * the ranks' values in rank order, then bias / accumulate, and round once to y_dtype; the last leader advances the call
 * counter.  Replaces lowbit_matmul_cuda + lowbit_allreduce_sum_cuda for the reference's
 * `result = matmul(...); all_reduce(result); result += bias` (low_bit_linear.py:704-725).  The sum is taken over the
 * fp32 partials (the separate all-reduce sums values already rounded to y_dtype); same bits on every rank. */
int lowbit_matmul_allreduce_cuda(const void* d_w, int qtype, int64_t n_rows, int64_t k, const void* d_x, int x_dtype, int64_t m,
                                 void* d_y, int y_dtype, const void* d_bias, int accumulate, void* const* peer_buffers, int rank,
                                 int world, size_t max_bytes, void* workspace, size_t workspace_bytes, void* stream);
/* In-place rotary position embedding of the q and k heads of the merged qkv projection --
 * xe_addons.rotary_half_inplaced / rotary_two_inplaced(inv_freq, position_ids, q, k) (models/llama.py:151-159,
 * models/chatglm2.py:251) and rotary_half_with_cache_inplaced / rotary_two_with_cache_inplaced(q, k, cos, sin
 * [, half_layout]) (models/common.py:354-367).  q [B][Hq][S][D], k [B][Hkv][S][D] (NULL: q only), same dtype,
 * D contiguous, strides in elements.  R = rotary width (even, <= D; elements beyond stay).  two = 0 pairs element i
 * with i + R/2, two != 0 pairs 2i with 2i + 1.  Angles: d_inv_freq (fp32 [R/2]) x d_position_ids (int64 [B][S],
 * pos_strides = {batch, token}, batch stride 0 broadcasts) when d_inv_freq != NULL, else the tables d_cos / d_sin
 * (cs_dtype, cs_strides = {batch, token}, last dimension contiguous; pair i reads entry i, or 2i when
 * cs_interleaved != 0).  fp32 arithmetic, one rounding to the tensors' dtype. */
int lowbit_rotary_inplaced_cuda(void* d_q, const int64_t* q_strides, int Hq, void* d_k, const int64_t* k_strides, int Hkv,
                                int dtype, int B, int S, int D, int R, int two, const float* d_inv_freq,
                                const int64_t* d_position_ids, const int64_t* pos_strides, const void* d_cos,
                                const void* d_sin, int cs_dtype, const int64_t* cs_strides, int cs_interleaved, void* stream);
/* Decode attention over the FP8 cache -- xe_addons.sdp_fp8(query, key, value, mask, scale)
 * (models/common.py:279-283: key.dtype == uint8, seq_length <= 32).  q [B][Hq][S][D] (q_dtype), k / v uint8
 * [B][Hkv][L][D] (GQA: Hq a multiple of Hkv), mask NULL or additive [B][1|Hq][S][L] (mask_strides[1] = 0
 * broadcasts over heads), causal != 0: query token s sees keys <= L - S + s.  out [B][Hq][S][D] contiguous.
 * Softmax and accumulation in fp32; split over the key axis with a deterministic merge (two launches).
 * Key cache pointer and strides 16-byte aligned, value cache 4-byte.  D in {64, 96, 128, 192, 256}, S <= 32.
 * fp16 / bf16 queries with D = 64 or 128, a positive scale, a 16-byte aligned value cache (pointer and strides) and
 * 16-byte aligned query rows run on the tensor cores (q.k and p.v as mma.m16n8k16 from the fp8 bytes, P and Q as
 * fp16); everything else takes a CUDA-core kernel with fp32 products. */
int lowbit_sdp_fp8_supported(int Hq, int Hkv, int S, int D);
size_t lowbit_sdp_fp8_workspace_size(int B, int Hq, int Hkv, int S, int L, int D);
int lowbit_sdp_fp8_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                        const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                        int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq, int Hkv,
                        int S, int L, int D, void* d_out, int out_dtype, void* workspace, size_t workspace_bytes,
                        void* stream);

/* Attention of MANY new tokens over the FP8 cache -- xe_addons.sdp_fp8_causal / sdp_fp8_non_causal(query, key, value, mask,
 * scale) (models/common.py:264-286: the prefill, and any call with more than 32 query tokens).  Arguments as
 * lowbit_sdp_fp8_cuda (causal != 0: query token s sees keys <= L - S + s; additive mask optional), any S >= 1, no
 * workspace.  fp16 / bf16 queries with head_dim 128 (lowbit_sdp_fp8_prefill_supported); 16-byte aligned caches, query rows
 * and output.  One CTA per 64 query rows of a query head; the 32-key chunks are staged once per CTA through a cp.async ring
 * and both products run on the tensor cores from the e5m2 bytes (csrc/sdp_prefill.cuh).  Same tolerance as the decode kernel. */
int lowbit_sdp_fp8_prefill_supported(int Hq, int Hkv, int D, int q_dtype);
int lowbit_sdp_fp8_prefill_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                                int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq, int Hkv,
                                int S, int L, int D, void* d_out, int out_dtype, void* stream);
/* ... with the output's element strides {batch, head, token} (rows of D contiguous, multiples of 8): writing [B][S][Hq][D]
 * directly saves the transpose copy that follows attention in the model forwards (models/llama.py:198). */
int lowbit_sdp_fp8_prefill_strided_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                                const int64_t* k_strides, const void* d_v, const int64_t* v_strides, const void* d_mask,
                                int mask_dtype, const int64_t* mask_strides, int causal, float scale, int B, int Hq, int Hkv,
                                int S, int L, int D, void* d_out, int out_dtype, const int64_t* out_strides, void* stream);

/* lowbit_sdp_fp8_cuda for a decode step that is replayed from a CUDA graph: the key count is read on the device,
 * L = min(*d_len + len_add, L_max), so one recorded launch serves every step.  L_max (the cache capacity the step
 * was recorded for) fixes the grid and the workspace size (lowbit_sdp_fp8_workspace_size with L = L_max); no mask;
 * causal as above.  No counterpart in the reference, whose XPU path issues every op from Python with host-side
 * shapes (models/common.py:279-283). */
int lowbit_sdp_fp8_devlen_cuda(const void* d_q, int q_dtype, const int64_t* q_strides, const void* d_k,
                               const int64_t* k_strides, const void* d_v, const int64_t* v_strides, int causal,
                               float scale, int B, int Hq, int Hkv, int S, int L_max, const int32_t* d_len,
                               int len_add, int D, void* d_out, int out_dtype, void* workspace,
                               size_t workspace_bytes, void* stream);

/* RMSNorm -- xe_addons.rms_norm(weight, x_2d, eps) (models/common.py:187-200), with the arithmetic of the torch
 * statement the reference uses off-XPU (models/common.py:201-205 = transformers' LlamaRMSNorm): fp32 statistics,
 * x * rsqrt(mean(x^2) + eps) rounded to x_dtype, then times weight, rounded to y_dtype.  x [rows][hidden] and
 * y [rows][hidden] with row strides in elements; weight [hidden]. */
int lowbit_rms_norm_cuda(const void* d_x, int x_dtype, int64_t x_row_stride, const void* d_weight, int w_dtype,
                         float eps, int64_t rows, int64_t hidden, void* d_y, int y_dtype, int64_t y_row_stride,
                         void* stream);

/* Rotary embedding of the merged qkv projection's output and the FP8 cache write in one launch -- the pair
 * rotary_half_with_cache_inplaced(q, k, cos, sin) + DynamicFp8Cache.update -> xe_addons.quantize_key_value of the
 * reference's attention forward (models/llama.py:146-183, models/utils.py:119-139).  d_qkv [B][S][(Hq + 2 Hkv) * D]
 * (strides in elements): q heads are rotated in place; k heads are rotated, rounded to `dtype` and stored as e5m2 at
 * cache position p0 + s; v heads are stored as e5m2.  p0 = (d_pos ? *d_pos : 0) + pos_add, read on the device.
 * d_cos / d_sin: [n_positions][>= D/2] tables indexed by absolute position, rotate_half layout (full-width rotary,
 * R == D).  Caches: uint8 [B][Hkv][capacity][D], strides {batch, head, position} in bytes; positions >= capacity are
 * not written.  Same bytes as lowbit_rotary_inplaced_cuda followed by lowbit_kv_quantize_cuda. */
int lowbit_rope_kv_append_cuda(void* d_qkv, int dtype, int64_t batch_stride, int64_t token_stride, int B, int S,
                               int Hq, int Hkv, int D, const void* d_cos, const void* d_sin, int cs_dtype,
                               int64_t cs_row_stride, int n_positions, const int32_t* d_pos, int pos_add,
                               void* d_k_cache, const int64_t* k_strides, void* d_v_cache, const int64_t* v_strides,
                               int capacity, void* stream);

/* The whole attention of one decode step (one new token per sequence) -- the middle of the reference's attention forward
 * (models/llama.py:146-196): rotary embedding of q and k, DynamicFp8Cache.update, scaled_dot_product_attention -> sdp_fp8.
 * d_qkv [B][(Hq + 2 Hkv) * D] = the merged projection's output (batch stride in elements); *d_len = tokens already in
 * the cache = position of the new token (read on the device: the call can be replayed from a CUDA graph); tables and
 * caches as in lowbit_rope_kv_append_cuda; d_out [B][Hq][D] contiguous.  For fp16 / bf16 activations with head_dim 128
 * this is ONE launch: every CTA rotates its query rows itself, the CTA that owns the new position computes the new key /
 * value row (and tile 0 stores it in the cache), and the key splits of a (sequence, kv head) form a thread-block cluster
 * whose members push their partial softmax sums into the leader's shared memory (deterministic merge, no second launch,
 * no global scratch).  Other shapes, and caches long enough to want more than 8 splits, run the separate kernels
 * (lowbit_rope_kv_append_cuda, lowbit_sdp_fp8_devlen_cuda) and need the workspace.  Same results as the separate calls.
 * cache_complete != 0: the caller guarantees that *d_len and the cached rows below it were written by work that has
 * COMPLETED when this call's kernel starts -- under programmatic dependent launch that means an ordinarily launched
 * kernel (or a stream synchronisation) lies between their producers and this call, as between two decode steps; the
 * kernel then stages its first chunks of the cache while the previous kernel (the qkv projection) still drains.  0 is
 * always safe. */
size_t lowbit_decode_attention_workspace_size(int B, int Hq, int Hkv, int capacity, int D);
int lowbit_decode_attention_cuda(void* d_qkv, int dtype, int64_t batch_stride, int B, int Hq, int Hkv, int D,
                                 const void* d_cos, const void* d_sin, int cs_dtype, int64_t cs_row_stride, int n_positions,
                                 const int32_t* d_len, void* d_k_cache, const int64_t* k_strides, void* d_v_cache,
                                 const int64_t* v_strides, int capacity, float scale, void* d_out, int out_dtype,
                                 int cache_complete, void* workspace, size_t workspace_bytes, void* stream);

/* Greedy pick and step bookkeeping of a recorded decode step -- the torch.argmax + append of the reference's
 * generate loop (utils/benchmark_util_4_47.py:3340-3380), done on the device so that steps replay without a host
 * round trip.  logits [B][V]: d_next_ids[b] = index of the largest value (lowest index among equals);
 * d_tokens[b][d_state[1]] = the same; then d_state[0] (tokens in the cache) and d_state[1] (tokens generated) += 1.
 * d_state: int32[8] {cache length, tokens generated, 0 (the kernel's arrival counter), suppress_below, n_suppress, ...}.
 * d_suppress_ids (may be NULL): while d_state[1] < suppress_below, the first n_suppress (<= 8) ids in it cannot be
 * picked -- transformers' min_new_tokens rule for end-of-sequence ids. */
int lowbit_greedy_advance_cuda(const void* d_logits, int dtype, int64_t row_stride, int B, int V, int64_t* d_next_ids,
                               int64_t* d_tokens, int64_t tokens_row_stride, int tokens_capacity, int32_t* d_state,
                               const int64_t* d_suppress_ids, void* stream);

/* canonical <-> device layout on device buffers (same byte size, src != dst).
 * to_device != 0: canonical -> tiled ; 0: tiled -> canonical. */
int lowbit_relayout_cuda(const void* d_src, void* d_dst, int64_t rows, int64_t k, int qtype,
                         int to_device, void* stream);

/* bytes of workspace lowbit_matmul_cuda needs for this problem: the three bf16 planes an fp32 activation
 * chunk is split into (GEMV), or the fp32 partial sums of the K-split small-batch GEMM; 0 for fp16 / bf16
 * activations on the GEMV.  Nothing in it has to be initialised; it may be reused by any number of calls
 * that do not overlap in time. */
size_t lowbit_matmul_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype);

/* y[M][N] (+)= x[M][K] . dequant(W[N][K])^T  -- the XPU-side op `xe_linear.forward_new` /
 * `xe_batch.batch_forward` (LBL:696-708) and the CPU-side ggml_compute_forward_mul_mat_q_fp32.
 * d_w: weight in the DEVICE (tiled) layout.  d_x: [M][K] of x_dtype, row stride k.
 * d_y: [M][N] of y_dtype, row stride n_rows.  d_bias: NULL or [N] of y_dtype (added once).
 * accumulate != 0: y += result (residual add), else y = result.
 * workspace: a buffer of at least lowbit_matmul_workspace_size() bytes, or NULL = one library-owned buffer per
 * device.  The NULL mode is for ONE stream per device and must not be used under stream capture: calls are
 * serialised on the host only (two streams would race on the buffer on the GPU), and growing the buffer
 * synchronises the device and reallocates.
 * Any M >= 1.  fp16 / bf16 activations: M <= 16 runs the HBM-bound GEMV kernel (so does 17 <= M <= 32 when n_rows is
 * not a multiple of 128 or k not of 128), larger M the tensor-core kernels -- 17..64 rows the "skinny" one with the
 * weights as the 128-row operand, beyond that the prefill GEMM (activations and expanded weights enter the MMA as
 * fp16 / bf16, fp32 accumulation: about 1e-3 * sum|x w|); both may use a second, reduction launch.
 * fp32 activations: always the GEMV (chunks of 32 rows), which keeps them to 2^-24 -- the fp32-in / fp32-out
 * contract of ggml_compute_forward_mul_mat_q_fp32; cast to fp16 first to trade that precision for the GEMM,
 * as the reference's XPU path does (LBL:700-705). */
int lowbit_matmul_cuda(const void* d_w, int qtype, int64_t n_rows, int64_t k,
                       const void* d_x, int x_dtype, int64_t m,
                       void* d_y, int y_dtype, const void* d_bias, int accumulate,
                       void* workspace, size_t workspace_bytes, void* stream);

/* Fused MLP front half: y[M][N] = act(x . Wgate^T) * (x . Wup^T) -- the XPU op
 * `xe_linear.mlp_forward_xpu(x_2d, gate_w, up_w, M, K, N, act, qtype)` called from fuse_mlp_base
 * (transformers/models/common.py:146-161; act ids from models/utils.py: SILU 0, GELU 1).
 * d_w_gate / d_w_up: two [N][K] weights of the SAME qtype in the device layout.  For M <= 8 (M <= 16 with fp16 /
 * bf16 activations), N a multiple of 64 and a tensor-core qtype this is ONE kernel launch that streams both matrices once
 * and never writes the two intermediate vectors.  17 <= M <= 64 half-precision rows: both matrices in one pass of the skinny
 * tensor-core kernel, act * up in its reduction launch.  M > 128 half-precision rows (prefill): gate and up as two outputs of
 * ONE grid of the tensor-core GEMM (half-precision scratch) + one elementwise kernel.  Otherwise two matmuls into scratch
 * (fp32 behind the exact GEMV) and the elementwise kernel.  Same result in every case up to the rounding of the intermediates.
 * workspace: NULL (library-owned) or lowbit_mlp_workspace_size() bytes. */
size_t lowbit_mlp_workspace_size(int64_t n_rows, int64_t k, int64_t m, int x_dtype);
int lowbit_mlp_cuda(const void* d_w_gate, const void* d_w_up, int qtype, int64_t n_rows, int64_t k,
                    const void* d_x, int x_dtype, int64_t m, void* d_y, int y_dtype, int act,
                    void* workspace, size_t workspace_bytes, void* stream);

/* One-token mixture-of-experts MLP -- the XPU op `xe_linear.moe_forward_vec(x, topk_ids, topk_weight,
 * gates, ups, downs, hidden, intermediate, qtype)` (transformers/models/qwen3_moe.py:88-105,
 * deepseek.py:296-312):  y[hidden] (+)= sum_z weight[z] * Wdown[e_z] . (act(Wgate[e_z] . x) * (Wup[e_z] . x)),
 * e_z = topk_ids[z].  d_gates / d_ups / d_downs: DEVICE arrays of n_experts 64-bit device addresses
 * (what the reference builds from `expert.*_proj.weight.data_ptr()`), each pointing at a weight of
 * `qtype` in the device layout ([inter][hidden] for gate/up, [hidden][inter] for down).  topk ids are
 * read on the device (ids_dtype 0 = int32, 1 = int64), so the call needs no host synchronisation and
 * can be recorded in a CUDA graph; an id outside [0, n_experts) contributes zero.  Two launches:
 * all routed gate/up pairs, then ONE GEMV over the routed down projections concatenated along K
 * (deterministic: no atomics).  Requires a tensor-core qtype and hidden, inter multiples of 64
 * (lowbit_moe_supported() says so); act as in lowbit_mlp_cuda.
 * workspace: NULL (library-owned) or lowbit_moe_workspace_size() bytes, 256-byte aligned. */
int lowbit_moe_supported(int qtype, int64_t hidden, int64_t inter, int top_k);
size_t lowbit_moe_workspace_size(int64_t hidden, int64_t inter, int top_k, int x_dtype);
int lowbit_moe_cuda(const void* d_x, int x_dtype, const void* d_topk_ids, int ids_dtype,
                    const void* d_topk_weight, int weight_dtype, int top_k,
                    const uint64_t* d_gates, const uint64_t* d_ups, const uint64_t* d_downs, int n_experts,
                    int64_t hidden, int64_t inter, int qtype, int act,
                    void* d_y, int y_dtype, int accumulate,
                    void* workspace, size_t workspace_bytes, void* stream);

/* No counterpart in the reference.  The fp32-activation matmul of lowbit_matmul_cuda, for hosts that
 * cannot give the library memory: an inference server whose arena owns all device memory and bans
 * cudaMalloc after start-up, a process held to a fixed device-memory budget, a thread that must not
 * hit the allocator's implicit synchronisation.  The scratch the kernel needs (three bf16 planes of
 * the activations) is a fixed array inside the library: 1.5 MiB on every GPU that loads it, kept
 * until the process exits.  Limit: m * k <= 262144 activation elements, larger calls return an
 * error and launch nothing.  x is fp32 [m][k]; the other arguments are those of lowbit_matmul_cuda.
 * The call allocates nothing.  Its first use on a device also looks up the array's device address
 * (that device must be current); make that first call outside any stream capture.  All calls share
 * the one array: two calls whose kernels may overlap on the GPU (different streams) race on it.
 * Nothing in it needs clearing between calls - each call overwrites what it reads. */
int lowbit_matmul_f32_noalloc_cuda(const void* d_w, int qtype, int64_t n_rows, int64_t k,
                                   const float* d_x, int64_t m, void* d_y, int y_dtype,
                                   const void* d_bias, int accumulate, void* stream);

/* on != 0: the matmul kernels are launched with programmatic stream serialization (PDL): a
 * kernel may start streaming its weights while the previous kernel on the stream drains; it
 * reads activations only after griddepcontrol.wait.  Process-wide switch, default off. */
void lowbit_set_programmatic_dependent_launch(int on);
int lowbit_get_programmatic_dependent_launch(void);

/* One-shot hint consumed by the next lowbit_matmul_cuda on the calling thread: the device
 * pointer and size of the weights the FOLLOWING matmul on the same stream will read (in a decoder
 * the order of the linears is static).  The kernel asks the L2 to prefetch them
 * (cp.async.bulk.prefetch.L2) while it runs, which keeps HBM busy during launch gaps and
 * reduction tails.  Purely a performance hint: NULL / 0 or never calling it is always correct. */
void lowbit_hint_next_weights(const void* d_w, size_t bytes);

/* number of kernels the last lowbit_* / ggml_*_cuda call on this thread launched */
int b200_lowbit_last_launch_count(void);
/* SM count of the current device (0 if no device) */
int b200_lowbit_sm_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_LOWBIT_H */
