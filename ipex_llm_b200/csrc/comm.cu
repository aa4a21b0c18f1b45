// comm.cu -- all-reduce(sum) of the small [M, hidden] results of the row-parallel linears in tensor-parallel decode
// (reference low_bit_linear.py:715-725 calls torch.distributed / the DeepSpeed group all_reduce after
// o_proj and down_proj: 2 per layer, 16 - 32 KiB at M = 1).  One process per GPU; every rank owns one buffer
// (cudaMalloc, exported with cudaIpcGetMemHandle, mapped by the peers with cudaIpcOpenMemHandle):
//     [epochs: COMM_MAX_CTAS x u32][status u32] ... [slot 0: world x inbox][slot 1: world x inbox]
// One-shot, push: every rank writes its vector into its inbox in every PEER's buffer with stores over NVLink, then sums
// the world's vectors from its OWN memory in rank order (all ranks get bit-identical results).  The stores are 8 bytes
// wide, {4 bytes of data, 4-byte call number}: an aligned 8-byte store is single-copy atomic, so the receiver needs no
// separate flag and no fence -- it polls each word until the call number is the current one (the latency is one NVLink
// crossing).  An inbox therefore takes twice the payload.
//
// Per thread (no thread waits for another thread of its grid, there is no barrier):
//   1. epoch = epochs[cta] + 1 (device-resident, so a captured CUDA graph replays); slot = epoch & 1;
//   2. for its 4-byte words w: read in[w]; store {in[w], epoch} to peer r's slot[rank][w] for every r != rank;
//   3. for every r != rank: poll own slot[r][w] until its call number is epoch.  The poll is bounded (spin_limit polls
//      with nanosleep back-off: about LOWBIT_COMM_TIMEOUT_S = 20 seconds, far above any host-side skew between the ranks'
//      processes) so that a missing peer cannot hang the GPU.  Expiry is LOUD: the words that were waited for become NaN,
//      the communicator's status word is set and stays set, and every later call on the communicator writes NaN at once
//      (it is unusable: a rank that gave up may be calls ahead of its peers, the two-slot argument below no longer holds);
//   4. out[w] = sum over ranks in rank order (fp32 accumulate); last: epochs[cta] = epoch.
// Two slots are enough: a rank can only start call e + 2 (same slot as e) after it received every peer's call e + 1
// words, which a peer sends after its call-e kernel has finished reading.
// Every rank must issue the same sequence of calls with the same n and dtype (as with any collective).  The grid
// depends on the communicator's max_bytes only, so every call runs every CTA and the CTAs' epochs stay equal.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include "../../include/b200_lowbit.h"
#include "lowbit_common.cuh"
#include "lowbit_internal.h"
#include "gemv_args.h"

namespace lowbit {
namespace {

constexpr int COMM_MAX_WORLD = 8;
constexpr int COMM_MAX_CTAS = 64;
constexpr int COMM_THREADS = 256;
constexpr int COMM_CHUNK_BYTES = COMM_THREADS * 8;              // two 4-byte words per thread
constexpr size_t COMM_HEADER_BYTES = 1024;                      // epochs 256 B, status
// polls per second of waiting, roughly (a poll = one volatile load round trip + a 256 ns sleep once the wait is long)
constexpr double COMM_POLLS_PER_SECOND = 1.5e6;
static_assert(COMM_MAX_CTAS * 4 + 8 <= COMM_HEADER_BYTES, "header too small");   // epochs, status, finished-leader count

struct CommArgs {
    uint8_t* peer[COMM_MAX_WORLD];       // every rank's buffer as mapped into this process (peer[rank] = own)
    int rank, world;
    size_t inbox_bytes;                  // 2 x max_bytes (rounded): one rank's words with their call numbers
    const void* in; void* out;
    int64_t n; int dtype;
    int64_t bytes_per_cta;               // multiple of 16
    int pdl;
    uint32_t spin_limit;
};

// call numbers: 1, 2, 3, ... ; 0 is the "never written" state of a fresh buffer, so the counter skips it when it wraps
// (after 2^32 calls; 0xffffffff is odd, 2 is even: the two slots keep alternating)
__device__ __forceinline__ uint32_t next_epoch(uint32_t prev) { const uint32_t e = prev + 1u; return e == 0u ? 2u : e; }
template <int DT> __device__ __forceinline__ uint32_t nan_word() { return DT == DT_F32 ? 0x7fc00000u : (DT == DT_F16 ? 0x7e007e00u : 0x7fc07fc0u); }


__device__ __forceinline__ void st_word(uint8_t* p, uint32_t data, uint32_t epoch) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(data), "r"(epoch) : "memory");
}
__device__ __forceinline__ uint2 ld_word(const uint8_t* p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

template <int DT> __device__ __forceinline__ void add_word(float (&acc)[2], uint32_t w);
template <> __device__ __forceinline__ void add_word<DT_F32>(float (&acc)[2], uint32_t w) { acc[0] += __uint_as_float(w); }
template <> __device__ __forceinline__ void add_word<DT_F16>(float (&acc)[2], uint32_t w) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    acc[0] += f.x; acc[1] += f.y;
}
template <> __device__ __forceinline__ void add_word<DT_BF16>(float (&acc)[2], uint32_t w) {
    acc[0] += __uint_as_float(w << 16); acc[1] += __uint_as_float(w & 0xffff0000u);
}
template <int DT> __device__ __forceinline__ uint32_t pack_word(const float (&acc)[2]);
template <> __device__ __forceinline__ uint32_t pack_word<DT_F32>(const float (&acc)[2]) { return __float_as_uint(acc[0]); }
template <> __device__ __forceinline__ uint32_t pack_word<DT_F16>(const float (&acc)[2]) {
    const __half2 h = __floats2half2_rn(acc[0], acc[1]);
    return *reinterpret_cast<const uint32_t*>(&h);
}
template <> __device__ __forceinline__ uint32_t pack_word<DT_BF16>(const float (&acc)[2]) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(acc[0], acc[1]);
    return *reinterpret_cast<const uint32_t*>(&h);
}

constexpr int COMM_WPT = 4;              // words a thread sends before it starts to poll

template <int DT>
__global__ void __launch_bounds__(COMM_THREADS) allreduce_kernel(const CommArgs A) {
    const int c = blockIdx.x, tid = threadIdx.x;
    uint8_t* own = A.peer[A.rank];
    uint32_t* epochs = reinterpret_cast<uint32_t*>(own);
    uint32_t* status = epochs + COMM_MAX_CTAS;
    const int64_t total_words = A.n * (DT == DT_F32 ? 4 : 2) / 4;
    const int64_t begin = (int64_t)c * (A.bytes_per_cta / 4);
    const int64_t end = begin + A.bytes_per_cta / 4 < total_words ? begin + A.bytes_per_cta / 4 : total_words;
    // Under programmatic dependent launch (lowbit_set_programmatic_dependent_launch): the next kernel of the stream
    // -- the following linear, which streams its weights before it looks at its input -- may start now; this grid
    // itself waits here for the kernel that produces `in`.  Both are no-ops for ordinary launches.  The call counter
    // is read AFTER the wait: chains of early launches can start this grid before the previous call's grid has
    // finished and advanced it.
    if (A.pdl) {
        asm volatile("griddepcontrol.launch_dependents;");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    uint32_t epoch, dead;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(epoch) : "l"(epochs + c) : "memory");
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(dead) : "l"(status) : "memory");
    epoch = next_epoch(epoch);
    const size_t slot_off = COMM_HEADER_BYTES + (epoch & 1) * (A.inbox_bytes * COMM_MAX_WORLD);
    const uint32_t* in = reinterpret_cast<const uint32_t*>(A.in);
    uint32_t* out = reinterpret_cast<uint32_t*>(A.out);
    bool expired = dead != 0;                         // an earlier wait ran out: no sends, no waits, NaN results
    for (int64_t w0 = begin + tid; w0 < end; w0 += COMM_THREADS * COMM_WPT) {
        uint32_t mine[COMM_WPT];
#pragma unroll
        for (int j = 0; j < COMM_WPT; j++) {
            const int64_t w = w0 + j * COMM_THREADS;
            if (w < end) {
                mine[j] = in[w];
#pragma unroll
                for (int r = 0; r < COMM_MAX_WORLD; r++)
                    if (!expired && r < A.world && r != A.rank) st_word(A.peer[r] + slot_off + A.rank * A.inbox_bytes + w * 8, mine[j], epoch);
            }
        }
        // poll in rounds: every round loads all words that are still missing at once (one memory round trip per
        // round, whatever the world size) and then looks at their call numbers
        uint2 got[COMM_WPT][COMM_MAX_WORLD];
        uint32_t missing = 0;                         // bit j * 8 + r
#pragma unroll
        for (int j = 0; j < COMM_WPT; j++)
#pragma unroll
            for (int r = 0; r < COMM_MAX_WORLD; r++)
                if (r < A.world && r != A.rank && w0 + j * COMM_THREADS < end) missing |= 1u << (j * 8 + r);
        // (first a plain spin on one word -- a single load in flight while nothing has arrived yet keeps the polling
        // traffic out of the way of the incoming stores -- then everything else, which has usually arrived with it)
        if (missing != 0 && !expired) {
            const int sb = __ffs(missing) - 1;
            const uint8_t* p = own + slot_off + (sb & 7) * A.inbox_bytes + (w0 + (sb >> 3) * COMM_THREADS) * 8;
            uint32_t polls = 0;
            while (ld_word(p).y != epoch) {
                if (++polls > A.spin_limit) { expired = true; break; }
                if (polls > 64) __nanosleep(polls > 4096 ? 256 : 32);
            }
        }
        for (uint32_t polls = 0; missing != 0 && !expired; polls++) {
            if (polls > A.spin_limit) { expired = true; break; }
            if (polls > 16) __nanosleep(polls > 1024 ? 256 : 32);
#pragma unroll
            for (int j = 0; j < COMM_WPT; j++)
#pragma unroll
                for (int r = 0; r < COMM_MAX_WORLD; r++)
                    if (missing & (1u << (j * 8 + r)))
                        got[j][r] = ld_word(own + slot_off + r * A.inbox_bytes + (w0 + j * COMM_THREADS) * 8);
#pragma unroll
            for (int j = 0; j < COMM_WPT; j++)
#pragma unroll
                for (int r = 0; r < COMM_MAX_WORLD; r++)
                    if ((missing & (1u << (j * 8 + r))) && got[j][r].y == epoch) missing &= ~(1u << (j * 8 + r));
        }
#pragma unroll
        for (int j = 0; j < COMM_WPT; j++) {
            const int64_t w = w0 + j * COMM_THREADS;
            if (w < end) {
                float acc[2] = {0.0f, 0.0f};
#pragma unroll
                for (int r = 0; r < COMM_MAX_WORLD; r++)
                    if (r < A.world) {
                        uint32_t v = mine[j];
                        if (r != A.rank) v = got[j][r].x;
                        add_word<DT>(acc, v);
                    }
                out[w] = expired ? nan_word<DT>() : pack_word<DT>(acc);
            }
        }
    }
    if (expired) atomicExch(status, 1u);
    // every thread has read epochs[c] before it is advanced
    __syncthreads();
    if (tid == 0) epochs[c] = epoch;
}

}  // namespace
}  // namespace lowbit

using namespace lowbit;

// bound of a wait inside the kernels, from LOWBIT_COMM_TIMEOUT_S (seconds, default 20; clamped to [0.05, 1800])
static uint32_t comm_spin_limit() {
    static const uint32_t lim = [] {
        const char* e = getenv("LOWBIT_COMM_TIMEOUT_S");
        double sec = e ? atof(e) : 20.0;
        if (!(sec >= 0.05)) sec = 0.05;
        if (sec > 1800.0) sec = 1800.0;
        const double polls = sec * COMM_POLLS_PER_SECOND;
        return polls > 4.0e9 ? 4000000000u : (uint32_t)polls;
    }();
    return lim;
}

extern "C" size_t lowbit_comm_max_bytes(void) { return (size_t)256 << 10; }     // decode-sized vectors (M <= 16 x hidden 8192 x fp16)

static size_t comm_inbox_bytes(size_t max_bytes) { return (2 * max_bytes + 255) / 256 * 256; }

extern "C" int lowbit_comm_create(size_t max_bytes, void** d_buffer, void* ipc_handle_64) {
    if (!d_buffer || !ipc_handle_64 || max_bytes == 0 || max_bytes > lowbit_comm_max_bytes()) {
        set_error(LB_ERR_ARG, "lowbit_comm_create: bad argument (max_bytes %zu, limit %zu)", max_bytes, lowbit_comm_max_bytes());
        return LB_ERR_ARG;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    const size_t total = COMM_HEADER_BYTES + 2 * COMM_MAX_WORLD * comm_inbox_bytes(max_bytes);      // two slots of `world` inboxes
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, total);
    if (e == cudaSuccess) e = cudaMemset(p, 0, total);                 // call number 0 is never used
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        if (p) cudaFree(p);
        set_error(LB_ERR_CUDA, "lowbit_comm_create: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return LB_ERR_CUDA;
    }
    memcpy(ipc_handle_64, &h, 64);
    *d_buffer = p;
    return LB_OK;
}

extern "C" int lowbit_comm_open(const void* ipc_handle_64, void** d_peer) {
    if (!ipc_handle_64 || !d_peer) { set_error(LB_ERR_ARG, "lowbit_comm_open: bad argument"); return LB_ERR_ARG; }
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle_64, 64);
    const cudaError_t e = cudaIpcOpenMemHandle(d_peer, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        set_error(LB_ERR_CUDA, "lowbit_comm_open: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return LB_ERR_CUDA;
    }
    return LB_OK;
}

extern "C" int lowbit_comm_close(void* d_peer) {
    if (d_peer && cudaIpcCloseMemHandle(d_peer) != cudaSuccess) { cudaGetLastError(); set_error(LB_ERR_CUDA, "lowbit_comm_close failed"); return LB_ERR_CUDA; }
    return LB_OK;
}

extern "C" int lowbit_comm_destroy(void* d_buffer) {
    if (d_buffer && cudaFree(d_buffer) != cudaSuccess) { cudaGetLastError(); set_error(LB_ERR_CUDA, "lowbit_comm_destroy failed"); return LB_ERR_CUDA; }
    return LB_OK;
}

// 0: every wait so far was answered; 1: some wait ran into its bound (a peer did not take part).  Synchronises.
extern "C" int lowbit_comm_status(const void* d_buffer, int* status) {
    if (!d_buffer || !status) { set_error(LB_ERR_ARG, "lowbit_comm_status: bad argument"); return LB_ERR_ARG; }
    uint32_t v = 0;
    const cudaError_t e = cudaMemcpy(&v, (const uint8_t*)d_buffer + COMM_MAX_CTAS * 4, 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { set_error(LB_ERR_CUDA, "lowbit_comm_status: %s", cudaGetErrorString(e)); cudaGetLastError(); return LB_ERR_CUDA; }
    *status = (int)v;
    return LB_OK;
}

extern "C" int lowbit_allreduce_sum_cuda(void* const* peer_buffers, int rank, int world, size_t max_bytes, const void* d_in,
                                         void* d_out, int64_t n, int dtype, void* stream) {
    reset_launch_count();
    const int64_t bytes = n * (dtype == DT_F32 ? 4 : 2);
    if (!peer_buffers || world < 1 || world > COMM_MAX_WORLD || rank < 0 || rank >= world || !d_in || !d_out || n < 0 || dtype < 0 ||
        dtype > 2 || (bytes & 15) || max_bytes == 0 || (size_t)bytes > max_bytes || max_bytes > lowbit_comm_max_bytes() || ((uintptr_t)d_in & 15) ||
        ((uintptr_t)d_out & 15)) {
        set_error(LB_ERR_ARG, "lowbit_allreduce_sum_cuda: bad argument (rank %d of %d, n %lld, dtype %d; n * size must be a multiple of 16 "
                  "bytes and at most the size given to lowbit_comm_create; pointers 16-byte aligned)", rank, world, (long long)n, dtype);
        return LB_ERR_ARG;
    }
    if (n == 0) return LB_OK;
    CommArgs A;
    for (int r = 0; r < COMM_MAX_WORLD; r++) A.peer[r] = r < world ? (uint8_t*)peer_buffers[r] : nullptr;
    for (int r = 0; r < world; r++)
        if (!A.peer[r]) { set_error(LB_ERR_ARG, "lowbit_allreduce_sum_cuda: peer buffer %d is NULL", r); return LB_ERR_ARG; }
    A.rank = rank; A.world = world; A.inbox_bytes = comm_inbox_bytes(max_bytes);
    A.in = d_in; A.out = d_out; A.n = n; A.dtype = dtype;
    // The grid and the chunk of every CTA follow from max_bytes alone (2 KiB per CTA until COMM_MAX_CTAS CTAs, then
    // larger chunks): every call of a communicator runs the same CTAs over the same byte ranges, whatever n is, so
    // the per-CTA call counters stay equal and the two-slot argument above holds chunk by chunk.  Create the
    // communicator with the size that is actually reduced (decode: M x hidden) to keep the grid small.
    int ctas = (int)((max_bytes + COMM_CHUNK_BYTES - 1) / COMM_CHUNK_BYTES);
    if (ctas > COMM_MAX_CTAS) ctas = COMM_MAX_CTAS;
    A.bytes_per_cta = (int64_t)(((max_bytes + ctas - 1) / ctas + 15) / 16 * 16);
    A.pdl = gemv_get_pdl();
    A.spin_limit = comm_spin_limit();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(COMM_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = A.pdl ? 1 : 0;
    if (dtype == DT_F32) cudaLaunchKernelEx(&cfg, allreduce_kernel<DT_F32>, A);
    else if (dtype == DT_F16) cudaLaunchKernelEx(&cfg, allreduce_kernel<DT_F16>, A);
    else cudaLaunchKernelEx(&cfg, allreduce_kernel<DT_BF16>, A);
    note_launch();
    return check_launch();
}

// Row-parallel linear + all-reduce in ONE kernel: the GEMV's leader CTAs send their fp32 sums to every peer as they
// finish, wait for the peers' words of the same rows in the own buffer, add the ranks' values in rank order and store y
// with bias / accumulate applied once after the sum (low_bit_linear.py:715-725 adds the bias after the all-reduce).  No
// send pass, no receive kernel, no kernel boundary between the matmul and the collective: a decode-sized all-reduce costs
// one NVLink crossing at the end of the row groups' tails.  m <= 32, n_rows % 64 == 0, m * n_rows * 4 <= max_bytes.
extern "C" int lowbit_matmul_allreduce_cuda(const void* d_w, int qtype, int64_t n_rows, int64_t k, const void* d_x, int x_dtype,
                                            int64_t m, void* d_y, int y_dtype, const void* d_bias, int accumulate,
                                            void* const* peer_buffers, int rank, int world, size_t max_bytes, void* workspace,
                                            size_t workspace_bytes, void* stream) {
    reset_launch_count();
    const int64_t n_words = m * n_rows;
    if (!peer_buffers || world < 1 || world > COMM_MAX_WORLD || rank < 0 || rank >= world || !d_w || !d_x || !d_y || m <= 0 || m > 32 ||
        n_rows <= 0 || (n_rows % 64) || y_dtype < 0 || y_dtype > 2 || max_bytes == 0 || max_bytes > lowbit_comm_max_bytes() ||
        (size_t)n_words * 4 > max_bytes) {
        set_error(LB_ERR_ARG, "lowbit_matmul_allreduce_cuda: bad argument (rank %d of %d, M %lld, N %lld: M <= 32, N %% 64 == 0 and "
                  "M * N * 4 bytes within the communicator's size)", rank, world, (long long)m, (long long)n_rows);
        return LB_ERR_ARG;
    }
    for (int r = 0; r < world; r++)
        if (!peer_buffers[r]) { set_error(LB_ERR_ARG, "lowbit_matmul_allreduce_cuda: peer buffer %d is NULL", r); return LB_ERR_ARG; }
    const size_t inbox = comm_inbox_bytes(max_bytes);
    GemvPush push = {};
    push.world = world; push.rank = rank; push.slot_bytes = inbox * COMM_MAX_WORLD; push.inbox_bytes = inbox;
    push.epochs = reinterpret_cast<unsigned int*>(peer_buffers[rank]);
    push.own = (unsigned long long)((uint8_t*)peer_buffers[rank] + COMM_HEADER_BYTES);
    push.spin_limit = comm_spin_limit();
    for (int r = 0; r < world; r++) push.inbox[r] = (unsigned long long)((uint8_t*)peer_buffers[r] + COMM_HEADER_BYTES + (size_t)rank * inbox);
    gemv_set_push(&push);
    int rc = gemv_launch_ex(d_w, nullptr, 0, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias, accumulate, workspace, workspace_bytes,
                            (cudaStream_t)stream);
    gemv_set_push(nullptr);
    if (rc == LB_ERR_ARG) set_error(rc, "lowbit_matmul_allreduce_cuda: the GEMV does not take this shape / qtype (qtype %d, N %lld, K %lld, M %lld)",
                                    qtype, (long long)n_rows, (long long)k, (long long)m);
    return rc;
}
