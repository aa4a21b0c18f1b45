// comm.cu -- all-reduce(sum) of the small [M, hidden] results of the row-parallel linears in tensor-parallel decode
// (reference low_bit_linear.py:715-725 calls torch.distributed / the DeepSpeed group all_reduce after
// o_proj and down_proj: 2 per layer, 16 - 32 KiB at M = 1).  One process per GPU; every rank owns one buffer
// (cudaMalloc, exported with cudaIpcGetMemHandle, mapped by the peers with cudaIpcOpenMemHandle) holding
//     [flags: world x COMM_MAX_CTAS x u32][epochs: COMM_MAX_CTAS x u32][status u32] ... [slot 0][slot 1]
// and the kernel reads the peers' slots over NVLink with plain loads (one-shot: every rank sums all partials itself,
// in rank order, so all ranks get bit-identical results).
//
// Protocol, per CTA c (a CTA owns one chunk of the vector on every rank; no CTA waits for another CTA of its grid):
//   1. epoch = epochs[c] + 1 (device-resident, so a CUDA graph replays correctly); slot = epoch & 1;
//   2. copy chunk c of the input into the own slot; __syncthreads; thread r < world: system fence, then
//      st.release.sys  peer r's flags[my_rank][c] = epoch;
//   3. thread r < world: ld.acquire.sys own flags[r][c] until it reaches epoch -- rank r's CTA c has published its
//      chunk.  The spin is bounded (COMM_SPIN_LIMIT polls with nanosleep back-off, about a second): on expiry the
//      status word is set and the kernel carries on, so a missing peer cannot hang the GPU;
//   4. out[chunk c] = sum over ranks 0..world-1 of their slot's chunk c (fp32 accumulate); epochs[c] = epoch.
// Two slots are enough: a rank can only start epoch e + 2 (same slot as e) after every peer published e + 1, which a
// peer does after its epoch-e kernel has finished reading.
// Every rank must issue the same sequence of calls with the same n and dtype (as with any collective).  The grid
// depends on the communicator's max_bytes only, so every call runs every CTA and the CTAs' epochs stay equal.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>
#include "../../include/b200_lowbit.h"
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {
namespace {

constexpr int COMM_MAX_WORLD = 8;
constexpr int COMM_MAX_CTAS = 64;
constexpr int COMM_THREADS = 256;
constexpr int COMM_CHUNK_BYTES = COMM_THREADS * 16;             // one 16-byte vector per thread and pass
constexpr size_t COMM_HEADER_BYTES = 4096;                      // flags 8 x 64 x 4 = 2048, epochs 256, status
constexpr uint32_t COMM_SPIN_LIMIT = 1u << 22;
static_assert(COMM_MAX_WORLD * COMM_MAX_CTAS * 4 + COMM_MAX_CTAS * 4 + 4 <= COMM_HEADER_BYTES, "header too small");

struct CommArgs {
    uint8_t* peer[COMM_MAX_WORLD];       // every rank's buffer as mapped into this process (peer[rank] = own)
    int rank, world;
    size_t slot_bytes;
    const void* in; void* out;
    int64_t n; int dtype;
    int64_t bytes_per_cta;               // multiple of 16
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// a slot is rewritten every other call: its loads must not be served from a stale L1 line
__device__ __forceinline__ uint4 ld_volatile16(const uint8_t* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

template <int DT> __device__ __forceinline__ void accumulate16(float (&acc)[8], const uint4& w);
template <> __device__ __forceinline__ void accumulate16<DT_F32>(float (&acc)[8], const uint4& w) {
    acc[0] += __uint_as_float(w.x); acc[1] += __uint_as_float(w.y); acc[2] += __uint_as_float(w.z); acc[3] += __uint_as_float(w.w);
}
template <> __device__ __forceinline__ void accumulate16<DT_F16>(float (&acc)[8], const uint4& w) {
    const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&ws[i]));
        acc[2 * i] += f.x; acc[2 * i + 1] += f.y;
    }
}
template <> __device__ __forceinline__ void accumulate16<DT_BF16>(float (&acc)[8], const uint4& w) {
    const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int i = 0; i < 4; i++) { acc[2 * i] += __uint_as_float(ws[i] << 16); acc[2 * i + 1] += __uint_as_float(ws[i] & 0xffff0000u); }
}
template <int DT> __device__ __forceinline__ uint4 pack16(const float (&acc)[8]);
template <> __device__ __forceinline__ uint4 pack16<DT_F32>(const float (&acc)[8]) {
    return make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
}
template <> __device__ __forceinline__ uint4 pack16<DT_F16>(const float (&acc)[8]) {
    uint32_t ws[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { const __half2 h = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]); ws[i] = *reinterpret_cast<const uint32_t*>(&h); }
    return make_uint4(ws[0], ws[1], ws[2], ws[3]);
}
template <> __device__ __forceinline__ uint4 pack16<DT_BF16>(const float (&acc)[8]) {
    uint32_t ws[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { const __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]); ws[i] = *reinterpret_cast<const uint32_t*>(&h); }
    return make_uint4(ws[0], ws[1], ws[2], ws[3]);
}

template <int DT>
__global__ void __launch_bounds__(COMM_THREADS) allreduce_kernel(const CommArgs A) {
    const int c = blockIdx.x, tid = threadIdx.x;
    uint8_t* own = A.peer[A.rank];
    uint32_t* flags = reinterpret_cast<uint32_t*>(own);                                   // [world][COMM_MAX_CTAS]
    uint32_t* epochs = flags + COMM_MAX_WORLD * COMM_MAX_CTAS;
    uint32_t* status = epochs + COMM_MAX_CTAS;
    const uint32_t epoch = epochs[c] + 1;
    const size_t slot_off = COMM_HEADER_BYTES + (epoch & 1) * A.slot_bytes;
    const int64_t total_bytes = A.n * (DT == DT_F32 ? 4 : 2);
    const int64_t begin = (int64_t)c * A.bytes_per_cta;
    const int64_t end = begin + A.bytes_per_cta < total_bytes ? begin + A.bytes_per_cta : total_bytes;   // total_bytes % 16 == 0

    // 1-2. publish this rank's chunk
    for (int64_t o = begin + tid * 16; o < end; o += COMM_THREADS * 16)
        *reinterpret_cast<uint4*>(own + slot_off + o) = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(A.in) + o);
    __syncthreads();
    if (tid < A.world) {
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t*>(A.peer[tid]) + A.rank * COMM_MAX_CTAS + c, epoch);
        // 3. wait for rank `tid`'s chunk
        uint32_t polls = 0;
        while ((int32_t)(ld_acquire_sys(flags + tid * COMM_MAX_CTAS + c) - epoch) < 0) {
            if (++polls > COMM_SPIN_LIMIT) { atomicExch(status, 1u); break; }
            if (polls > 64) __nanosleep(polls > 4096 ? 256 : 32);
        }
    }
    __syncthreads();
    // 4. sum in rank order
    for (int64_t o = begin + tid * 16; o < end; o += COMM_THREADS * 16) {
        float acc[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
        uint4 w[COMM_MAX_WORLD];
#pragma unroll
        for (int r = 0; r < COMM_MAX_WORLD; r++)
            if (r < A.world) w[r] = ld_volatile16(A.peer[r] + slot_off + o);
#pragma unroll
        for (int r = 0; r < COMM_MAX_WORLD; r++)
            if (r < A.world) accumulate16<DT>(acc, w[r]);
        *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(A.out) + o) = pack16<DT>(acc);
    }
    if (tid == 0) epochs[c] = epoch;
}
}  // namespace
}  // namespace lowbit

using namespace lowbit;

extern "C" size_t lowbit_comm_max_bytes(void) { return (size_t)COMM_MAX_CTAS * COMM_CHUNK_BYTES * 4; }     // 1 MiB per call

extern "C" int lowbit_comm_create(size_t max_bytes, void** d_buffer, void* ipc_handle_64) {
    if (!d_buffer || !ipc_handle_64 || max_bytes == 0 || max_bytes > lowbit_comm_max_bytes()) {
        set_error(LB_ERR_ARG, "lowbit_comm_create: bad argument (max_bytes %zu, limit %zu)", max_bytes, lowbit_comm_max_bytes());
        return LB_ERR_ARG;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    const size_t slot = (max_bytes + 255) / 256 * 256;
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, COMM_HEADER_BYTES + 2 * slot);
    if (e == cudaSuccess) e = cudaMemset(p, 0, COMM_HEADER_BYTES + 2 * slot);
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
    const cudaError_t e = cudaMemcpy(&v, (const uint8_t*)d_buffer + (COMM_MAX_WORLD * COMM_MAX_CTAS + COMM_MAX_CTAS) * 4, 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { set_error(LB_ERR_CUDA, "lowbit_comm_status: %s", cudaGetErrorString(e)); cudaGetLastError(); return LB_ERR_CUDA; }
    *status = (int)v;
    return LB_OK;
}

extern "C" int lowbit_allreduce_sum_cuda(void* const* peer_buffers, int rank, int world, size_t max_bytes, const void* d_in,
                                         void* d_out, int64_t n, int dtype, void* stream) {
    reset_launch_count();
    const int64_t bytes = n * (dtype == DT_F32 ? 4 : 2);
    if (!peer_buffers || world < 1 || world > COMM_MAX_WORLD || rank < 0 || rank >= world || !d_in || !d_out || n < 0 || dtype < 0 ||
        dtype > 2 || (bytes & 15) || (size_t)bytes > max_bytes || max_bytes > lowbit_comm_max_bytes() || ((uintptr_t)d_in & 15) ||
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
    A.rank = rank; A.world = world; A.slot_bytes = (max_bytes + 255) / 256 * 256;
    A.in = d_in; A.out = d_out; A.n = n; A.dtype = dtype;
    // The grid and the chunk of every CTA follow from max_bytes alone (4 KiB per CTA until COMM_MAX_CTAS CTAs, then
    // larger chunks): every call of a communicator runs the same CTAs over the same byte ranges, whatever n is, so
    // the per-CTA call counters stay equal and the two-slot argument above holds chunk by chunk.  Create the
    // communicator with the size that is actually reduced (decode: M x hidden) to keep the grid small.
    int ctas = (int)((max_bytes + COMM_CHUNK_BYTES - 1) / COMM_CHUNK_BYTES);
    if (ctas > COMM_MAX_CTAS) ctas = COMM_MAX_CTAS;
    A.bytes_per_cta = (int64_t)(((max_bytes + ctas - 1) / ctas + 15) / 16 * 16);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == DT_F32) allreduce_kernel<DT_F32><<<ctas, COMM_THREADS, 0, st>>>(A);
    else if (dtype == DT_F16) allreduce_kernel<DT_F16><<<ctas, COMM_THREADS, 0, st>>>(A);
    else allreduce_kernel<DT_BF16><<<ctas, COMM_THREADS, 0, st>>>(A);
    note_launch();
    return check_launch();
}
