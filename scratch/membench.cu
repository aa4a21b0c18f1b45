// membench.cu -- what read bandwidth can ONE kernel of a given size reach on B200, launched
// back to back (with and without programmatic dependent launch)?  Calibrates the GEMV ceiling.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// A: batched LDG.128: each warp owns a contiguous span; U loads in flight per thread, batch-consume
template <int U>
__global__ void __launch_bounds__(256) read_ldg(const uint4* __restrict__ src, size_t n16, unsigned* out, int pdl) {
    if (pdl) asm volatile("griddepcontrol.launch_dependents;");
    const size_t nwarps = (size_t)gridDim.x * (blockDim.x / 32);
    const size_t wid = (size_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    const int lane = threadIdx.x & 31;
    const size_t per = n16 / nwarps;          // uint4 per warp (assume divisible enough)
    const uint4* p = src + wid * per + lane;
    unsigned acc = 0;
    size_t i = 0;
    for (; i + U * 32 <= per; i += U * 32) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ldg_stream16(p + i + u * 32);
#pragma unroll
        for (int u = 0; u < U; u++) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    for (; i + 32 <= per; i += 32) { uint4 v = ldg_stream16(p + i); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (acc == 0x12345678u) out[wid] = acc;
}

// B: per-warp TMA bulk ring
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int CH, int STAGES>
__global__ void __launch_bounds__(128) read_tma(const uint8_t* __restrict__ src, size_t bytes, unsigned* out, int pdl) {
    extern __shared__ __align__(128) uint8_t smem[];
    if (pdl) asm volatile("griddepcontrol.launch_dependents;");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t nwarps = (size_t)gridDim.x * 4;
    const size_t wid = (size_t)blockIdx.x * 4 + warp;
    const size_t per = bytes / nwarps / CH * CH;
    const uint8_t* g = src + wid * per;
    const int n_chunks = (int)(per / CH);
    const uint32_t ring = smem_u32(smem) + warp * (STAGES * CH);
    const uint32_t bars = smem_u32(smem) + 4 * STAGES * CH + warp * STAGES * 8;
    auto issue = [&](int c) {
        const uint32_t bar = bars + (c % STAGES) * 8, dst = ring + (c % STAGES) * CH;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(CH) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(g + (size_t)c * CH), "r"(CH), "r"(bar) : "memory");
    };
    if (lane == 0) {
        for (int s = 0; s < STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + s * 8) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int c = 0; c < STAGES && c < n_chunks; c++) issue(c);
    }
    __syncwarp();
    unsigned acc = 0;
    for (int c = 0; c < n_chunks; c++) {
        const uint32_t bar = bars + (c % STAGES) * 8, base = ring + (c % STAGES) * CH;
        const uint32_t parity = (c / STAGES) & 1;
        asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
#pragma unroll
        for (int i = 0; i < CH / 512; i++) {
            uint4 v;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(base + i * 512 + lane * 16));
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
        __syncwarp();
        if (lane == 0 && c + STAGES < n_chunks) issue(c + STAGES);
    }
    if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (acc == 0x12345678u) out[wid] = acc;
}

template <typename F>
float time_graph(F launch, int nbuf, int reps, cudaStream_t st) {
    cudaGraph_t graph; cudaGraphExec_t exec;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
    for (int i = 0; i < nbuf; i++) launch(i);
    CK(cudaStreamEndCapture(st, &graph));
    CK(cudaGraphInstantiate(&exec, graph, 0));
    CK(cudaGraphLaunch(exec, st)); CK(cudaStreamSynchronize(st));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    CK(cudaEventRecord(e0, st));
    for (int r = 0; r < reps; r++) CK(cudaGraphLaunch(exec, st));
    CK(cudaEventRecord(e1, st)); CK(cudaStreamSynchronize(st));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaGraphExecDestroy(exec); cudaGraphDestroy(graph);
    return ms / (reps * nbuf);
}

template <typename K, typename... Args>
void launch_k(K kernel, int grid, int block, size_t smem, cudaStream_t st, int pdl, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, kernel, args...));
}

int main() {
    cudaStream_t st; CK(cudaStreamCreate(&st));
    const size_t total = 1200ull << 20;
    uint8_t* pool; CK(cudaMalloc(&pool, total)); CK(cudaMemset(pool, 1, total));
    unsigned* out; CK(cudaMalloc(&out, 1 << 20));
    CK(cudaFuncSetAttribute(read_tma<2048, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 4 * 2048 + 128));
    CK(cudaFuncSetAttribute(read_tma<4096, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 4 * 4096 + 128));
    CK(cudaFuncSetAttribute(read_tma<8192, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 3 * 8192 + 128));
    const double sizes_mb[] = {8.9, 24, 70, 280};
    for (double mb : sizes_mb) {
        size_t bytes = ((size_t)(mb * 1e6) / (148 * 32 * 8192)) * (148 * 32 * 8192);   // divisible for every config
        if (bytes == 0) bytes = 148 * 32 * 8192;
        int nbuf = (int)(total / bytes); if (nbuf > 24) nbuf = 24;
        printf("---- %.1f MB per kernel (%zu bytes), %d distinct buffers\n", bytes / 1e6, bytes, nbuf);
        for (int pdl = 0; pdl < 2; pdl++) {
            for (int cps : {2, 4, 8}) {
                float ms = time_graph([&](int i) { launch_k(read_ldg<8>, 148 * cps, 256, 0, st, pdl, (const uint4*)(pool + i * bytes), bytes / 16, out, pdl); }, nbuf, 10, st);
                printf("ldg U=8  ctas/sm=%d pdl=%d : %7.2f us  %7.0f GB/s\n", cps, pdl, ms * 1e3, bytes / ms / 1e6);
            }
            {
                float ms = time_graph([&](int i) { launch_k(read_ldg<4>, 148 * 8, 256, 0, st, pdl, (const uint4*)(pool + i * bytes), bytes / 16, out, pdl); }, nbuf, 10, st);
                printf("ldg U=4  ctas/sm=8 pdl=%d : %7.2f us  %7.0f GB/s\n", pdl, ms * 1e3, bytes / ms / 1e6);
            }
            for (int cps : {2, 4, 6}) {
                float ms = time_graph([&](int i) { launch_k(read_tma<2048, 4>, 148 * cps, 128, 4 * 4 * 2048 + 128, st, pdl, (const uint8_t*)(pool + i * bytes), bytes, out, pdl); }, nbuf, 10, st);
                printf("tma 2K x4 ctas/sm=%d pdl=%d : %7.2f us  %7.0f GB/s\n", cps, pdl, ms * 1e3, bytes / ms / 1e6);
            }
            for (int cps : {2, 3}) {
                float ms = time_graph([&](int i) { launch_k(read_tma<4096, 4>, 148 * cps, 128, 4 * 4 * 4096 + 128, st, pdl, (const uint8_t*)(pool + i * bytes), bytes, out, pdl); }, nbuf, 10, st);
                printf("tma 4K x4 ctas/sm=%d pdl=%d : %7.2f us  %7.0f GB/s\n", cps, pdl, ms * 1e3, bytes / ms / 1e6);
            }
            {
                float ms = time_graph([&](int i) { launch_k(read_tma<8192, 3>, 148 * 2, 128, 4 * 3 * 8192 + 128, st, pdl, (const uint8_t*)(pool + i * bytes), bytes, out, pdl); }, nbuf, 10, st);
                printf("tma 8K x3 ctas/sm=2 pdl=%d : %7.2f us  %7.0f GB/s\n", pdl, ms * 1e3, bytes / ms / 1e6);
            }
        }
    }
    return 0;
}
