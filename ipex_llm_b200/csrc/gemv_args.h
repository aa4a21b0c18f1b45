// gemv_args.h -- launch-parameter block and constants shared by gemv.cu (host) and
// gemv_kernel.cuh (device) of the dequant-GEMV.
#pragma once
#include "lowbit_common.cuh"

namespace lowbit {

#ifndef LOWBIT_GEMV_WARPS
#define LOWBIT_GEMV_WARPS 4
#endif
constexpr int GEMV_WARPS = LOWBIT_GEMV_WARPS;   // warps per CTA (each warp is an independent pipeline)
constexpr int GEMV_MAX_SPLIT = 8;           // CTAs per cluster = K slices of one 64-row group
constexpr int GEMV_MAX_NT = 4;             // n-tiles of 8 activation rows -> M <= 32 per pass

// activation of the fused MLP (reference xe_linear.mlp_forward_xpu act ids: SILU 0 / GELU 1 are
// mapped by the host to 1 / 2 here; 0 = identity).  expf, not __expf: the result must not depend
// on fast-math flags.
__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == 1) return v / (1.0f + expf(-v));                                    // SiLU
    if (act == 2) return 0.5f * v * (1.0f + erff(v * 0.70710678118654752f));       // GELU (erf)
    return v;
}

struct GemvArgs {
    const uint8_t* w;
    const uint8_t* w2;        // second matrix of the fused MLP (same shape / qtype) or NULL
    int act;                  // fused MLP activation on w's output: 1 SiLU, 2 GELU(erf)
    TiledGeom G;
    const uint16_t* x;        // NS planes of [M][K] fp16 or bf16
    int64_t plane_stride;     // elements between planes
    int M;
    void* y; int y_dtype;     // [M][rows]
    const void* bias;         // NULL or [rows] of y_dtype
    int accumulate;
    int pdl;                  // launched with programmatic stream serialization
    const uint8_t* next_w;    // hint: weights of the NEXT matmul on this stream, prefetched into L2 (or NULL)
    unsigned long long next_bytes;
    unsigned long long* ts;   // development: per-warp phase timestamps (8 per warp) or NULL
    // mixture-of-experts decode (reference xe_linear.moe_forward_vec), M == 1:
    //   mode 1: blockIdx.z = routed slot z; w / w2 = tab_w[sel[z]] / tab_w2[sel[z]]; y row z
    //   mode 2: the K axis is the concatenation of n_seg matrices tab_w[sel[z]] ([rows][k] each) and
    //           x is [n_seg][k]; slot z's partial sums are multiplied by seg_scale[z]
    int ring_recv;            // M = 2..8: the leader receives its peers' sums in its rings (set at launch)
    int slice_q, slice_r;     // k-blocks per cluster rank: (blocks along K) / split and the remainder (set at launch)
    int mode;
    const unsigned long long* tab_w;
    const unsigned long long* tab_w2;
    const void* sel; int sel_i64;          // routed expert ids, int32 or int64
    int n_experts;                         // ids outside [0, n_experts) contribute zero
    int n_seg;
    const void* seg_scale; int seg_scale_dtype;
    // row-parallel linear under tensor parallelism (comm.cu): the matmul and its all-reduce are ONE kernel.  The leader
    // CTA of every 64-row group sends its fp32 sums, each with the communicator's call number, into this rank's inbox in
    // every PEER's buffer (8-byte stores over NVLink), then waits for the peers' words of the same rows in its own buffer,
    // adds the ranks' values in rank order (every rank gets the same bits), applies bias / accumulate and stores y.  The
    // last leader to finish advances the call counter.  push_world = 0: an ordinary matmul.
    int push_world, push_rank;
    unsigned long long push_inbox[8];      // address of this rank's inbox (slot 0) inside rank r's buffer
    unsigned long long push_own;           // slot 0 of rank 0's inbox inside the OWN buffer (the words to wait for)
    unsigned long long push_inbox_bytes;   // rank r's inbox = rank 0's + r * push_inbox_bytes
    unsigned long long push_slot_bytes;    // slot 1 = slot 0 + push_slot_bytes
    unsigned int* push_epochs;             // own buffer's header: [0..63] call counter (kept equal), [64] status, [65] finished leaders
    unsigned int push_spin_limit;          // bound of the wait, in polls
};

// one-shot request consumed by the next gemv_launch_ex on this thread (see lowbit_matmul_allreduce_cuda)
struct GemvPush {
    int world, rank;
    unsigned long long inbox[8];
    unsigned long long own, inbox_bytes;
    unsigned long long slot_bytes;
    unsigned int* epochs;
    unsigned int spin_limit;
};

__device__ __forceinline__ void store_out(void* y, int dt, int64_t idx, float v, const void* bias, int64_t row, int acc) {
    if (dt == DT_F32) {
        float* p = (float*)y + idx;
        if (bias) v += ((const float*)bias)[row];
        if (acc) v += __ldcg(p);
        *p = v;
    } else if (dt == DT_F16) {
        __half* p = (__half*)y + idx;
        if (bias) v += __half2float(((const __half*)bias)[row]);
        if (acc) v += __half2float(__ldcg(p));
        *p = __float2half_rn(v);
    } else {
        __nv_bfloat16* p = (__nv_bfloat16*)y + idx;
        if (bias) v += __bfloat162float(((const __nv_bfloat16*)bias)[row]);
        if (acc) v += __bfloat162float(__ldcg(p));
        *p = __float2bfloat16_rn(v);
    }
}

}  // namespace lowbit
