// sdp_mma.cuh -- included by kvcache.cu (inside namespace lowbit::{anonymous}, after SdpArgs and the helpers).
// Decode attention over the e5m2 cache with q.k on the tensor cores (fp16 / bf16 queries, head_dim % 64 == 0).
//
// Per warp and chunk of 32 keys = four 8-key tiles:  S[rows(16, padded) x keys(8)] = Q . K^T as mma.m16n8k16 with
//   A = Q (rows g < RT live; rows g+8 are zero padding), fp16, kept in registers for the whole kernel,
//   B = K^T: lane (g,t) supplies key g of the tile; e5m2 -> fp16 is a byte permute (e5m2 = high byte of fp16).
// The head dimension is consumed in a permuted order so that every lane reads 16 contiguous bytes of its key row:
// k-step (c, s), c = 0..D/64-1, s = 0..3 covers d = 64c + 16t + 4s + {0,1 | 2,3} for lane t (the A fragment of Q
// uses the same order; a dot product does not care).
// Scores stay in the accumulator layout (lane (g,t): row g, keys 2t, 2t+1 of each tile); softmax runs there, and
// p.V keeps the lane-per-DPL-outputs layout of the CUDA-core kernel, fetching p with one shuffle per key and row.
// bf16 queries are converted to fp16 (exact inside fp16's normal range, which attention queries live in).

__device__ __forceinline__ void mma16816_f16(float (&c)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}

template <int DPL, int RT>
__global__ void __launch_bounds__(SDP_WARPS * 32)
sdp_fp8_decode_mma_kernel(const SdpArgs A) {
    constexpr int D = DPL * 32;
    constexpr int NC = D / 64;                       // 64-element groups of the head dimension
    static_assert(D % 64 == 0, "head_dim must be a multiple of 64");
    __shared__ float red_m[SDP_WARPS][RT], red_l[SDP_WARPS][RT];
    __shared__ float red_o[SDP_WARPS][RT][D];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int split = blockIdx.x, hkv = blockIdx.y % A.Hkv, b = blockIdx.y / A.Hkv, tile = blockIdx.z;
    const int n_rep = A.Hq / A.Hkv, rows_total = n_rep * A.S;
    const int row0 = tile * RT;
    const int n_rows = min(RT, rows_total - row0);

    // A fragments of Q: lane (g,t) holds row g, elements 64c + 16t .. +15 as 8 fp16 pairs (zeros for rows >= n_rows)
    uint32_t qa[NC][8];
#pragma unroll
    for (int c = 0; c < NC; c++)
#pragma unroll
        for (int i = 0; i < 8; i++) qa[c][i] = 0u;
    int my_hq = 0, my_st = 0;                         // query head / token of this lane's row
    if (g < n_rows) {
        const int rr = row0 + g;
        my_hq = hkv * n_rep + rr / A.S; my_st = rr % A.S;
        const int64_t base = b * A.q_sb + my_hq * A.q_sh + my_st * A.q_ss;
#pragma unroll
        for (int c = 0; c < NC; c++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int d = 64 * c + 16 * t + 2 * i;
                const __half2 h = __floats2half2_rn(load_as_float(A.q, A.q_dtype, base + d), load_as_float(A.q, A.q_dtype, base + d + 1));
                qa[c][i] = *reinterpret_cast<const uint32_t*>(&h);
            }
    }
