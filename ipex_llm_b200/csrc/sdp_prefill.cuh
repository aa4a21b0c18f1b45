// sdp_prefill.cuh -- included by kvcache.cu (inside namespace lowbit::{anonymous}, after sdp_mma.cuh).
// Attention of MANY query tokens over the e5m2 cache -- xe_addons.sdp_fp8_causal / sdp_fp8_non_causal of the reference
// (models/common.py:264-286: the prefill, where key.dtype == uint8 because DynamicFp8Cache.update returned the cache).
//
// fp16 / bf16 queries, head_dim 128.  One CTA = 64 query rows of one (sequence, query head): 4 warps x one full m16 tile
// (rows g and g + 8 per lane -- the decode kernel in sdp_mma.cuh uses the same fragments with rows 8..15 zero).  All four
// warps walk the same keys, so the 32-key chunks (4 KB of K + 4 KB of V) are staged ONCE per CTA through a 3-stage
// cp.async ring with the decode kernel's XOR swizzles, and read by every warp:
//   S = Q . K^T : mma.m16n8k16, e5m2 -> fp16 by byte permute, head dimension in the permuted order of sdp_mma.cuh;
//   online softmax per row in the accumulator layout (a row's 32 scores of a chunk sit in the 4 lanes of a quad);
//   O += P . V  : the score accumulators are the A fragments of P; a tile of ones accumulates the denominator from the
//                 same rounded P.
// Causal: query token s (0-based among the S new ones) sees keys <= L - S + s; chunks beyond a tile's last row are
// never loaded, chunks beyond a warp's last row are skipped by that warp.  An additive mask [B][1|Hq][S][L] is honoured.
// Softmax and accumulation in fp32, P and Q enter the MMAs as fp16: the decode kernel's tolerance statement applies.

__device__ __forceinline__ void mma16816_f16_full(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int PF_WARPS = 4;
constexpr int PF_ROWS = 16 * PF_WARPS;           // query rows per CTA
constexpr int PF_STAGES = 3;

__global__ void __launch_bounds__(PF_WARPS * 32, 1)
sdp_fp8_prefill_kernel(const SdpArgs A) {
    constexpr int D = 128, NC = 2, BPL = 16, VW = 4;
    __shared__ __align__(128) uint8_t ring[PF_STAGES * 8192];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int hq = blockIdx.y % A.Hq, b = blockIdx.y / A.Hq;
    const int hkv = hq / (A.Hq / A.Hkv);
    const int q0 = blockIdx.x * PF_ROWS;                              // first query token of the CTA
    const int row_a = q0 + 16 * warp + g, row_b = row_a + 8;          // this lane's two query tokens
    const int L = A.L, S = A.S;

    // A fragments of Q (fp16): rows g (qa) and g + 8 (qb), elements 64c + 16t .. +15 as 8 pairs each
    uint32_t qa[NC][8], qb[NC][8];
    auto load_q = [&](int row, uint32_t (&dst)[NC][8]) {
#pragma unroll
        for (int c = 0; c < NC; c++)
#pragma unroll
            for (int i = 0; i < 8; i++) dst[c][i] = 0u;
        if (row < S) {
            const uint16_t* qrow = reinterpret_cast<const uint16_t*>(A.q) + b * A.q_sb + hq * A.q_sh + (int64_t)row * A.q_ss + 16 * t;
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const uint4 w0 = *reinterpret_cast<const uint4*>(qrow + 64 * c), w1 = *reinterpret_cast<const uint4*>(qrow + 64 * c + 8);
                const uint32_t ws[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (A.q_dtype == DT_F16) dst[c][i] = ws[i];
                    else {
                        const __half2 h = __floats2half2_rn(__uint_as_float(ws[i] << 16), __uint_as_float(ws[i] & 0xffff0000u));
                        dst[c][i] = *reinterpret_cast<const uint32_t*>(&h);
                    }
                }
            }
        }
    };
    load_q(row_a, qa);
    load_q(row_b, qb);

    // keys this CTA / this warp / each row needs
    const int last_row_cta = min(q0 + PF_ROWS, S) - 1, last_row_warp = min(q0 + 16 * warp + 15, S - 1);
    const int key_end = A.causal ? min(L, L - S + last_row_cta + 1) : L;           // exclusive
    const int key_end_warp = A.causal ? min(L, L - S + last_row_warp + 1) : L;
    const int lim_a = A.causal ? L - S + row_a : L - 1, lim_b = A.causal ? L - S + row_b : L - 1;   // last visible key
    const int first_full = A.causal ? L - S + q0 + 16 * warp : L;                  // keys below are visible to every row of the warp
    const int64_t mbase_a = b * A.m_sb + hq * A.m_sh + (int64_t)row_a * A.m_ss, mbase_b = b * A.m_sb + hq * A.m_sh + (int64_t)row_b * A.m_ss;

    const uint8_t* const kb0 = A.k + b * A.k_sb + hkv * A.k_sh;
    const uint8_t* const vb0 = A.v + b * A.v_sb + hkv * A.v_sh;
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
    const int cr = threadIdx.x >> 3, unit = threadIdx.x & 7;         // thread copies piece `unit` of rows cr and cr + 16
    auto stage_chunk = [&](int c0, int slot) {
        if (c0 < key_end) {
            const uint32_t dst = ring_s + slot * 8192;
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const int r = cr + 16 * i;
                const int64_t key = min(c0 + r, key_end - 1);        // rows beyond the range repeat the last key (p = 0)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst + r * 128 + ((unit ^ ((r & 1) << 2)) << 4)),
                             "l"(kb0 + key * A.k_ss + 16 * unit) : "memory");
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst + 4096 + r * 128 + ((unit ^ (((r >> 1) & 3) << 1)) << 4)),
                             "l"(vb0 + key * A.v_ss + 16 * unit) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    float m_a = -CUDART_INF_F, m_b = -CUDART_INF_F;
    float o[BPL][4], ol[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int u = 0; u < BPL; u++) o[u][0] = o[u][1] = o[u][2] = o[u][3] = 0.0f;

    stage_chunk(0, 0);
    stage_chunk(32, 1);
    int slot = 0;
    for (int c0 = 0; c0 < key_end; c0 += 32) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();                              // chunk c0 has landed for everybody; everybody is done with chunk c0 - 32
        {
            int ahead = slot + 2; if (ahead >= PF_STAGES) ahead -= PF_STAGES;
            stage_chunk(c0 + 64, ahead);
        }
        const uint8_t* stage = ring + slot * 8192;
        if (++slot == PF_STAGES) slot = 0;
        if (c0 >= key_end_warp) continue;             // nothing in this chunk is visible to this warp's rows
        // ---- scores of 4 tiles x 8 keys, rows g and g + 8
        float sc[4][4];
#pragma unroll
        for (int n = 0; n < 4; n++) {
            sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.0f;
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const uint4 kw = *reinterpret_cast<const uint4*>(stage + (8 * n + g) * 128 + (((4 * c + t) ^ ((g & 1) << 2)) << 4));
                const uint32_t ws[4] = {kw.x, kw.y, kw.z, kw.w};
#pragma unroll
                for (int s = 0; s < 4; s++)
                    mma16816_f16_full(sc[n], qa[c][2 * s], qb[c][2 * s], qa[c][2 * s + 1], qb[c][2 * s + 1],
                                      __byte_perm(ws[s], 0u, 0x1404u), __byte_perm(ws[s], 0u, 0x3424u));
            }
        }
        // ---- scale, mask, online softmax of the two rows; this lane's values are keys c0 + 8n + 2t + i
        const bool plain = c0 + 32 <= key_end && c0 + 31 <= first_full && !A.mask;     // every key visible to every row
        float mx_a = -CUDART_INF_F, mx_b = -CUDART_INF_F;
#pragma unroll
        for (int n = 0; n < 4; n++)
#pragma unroll
            for (int i = 0; i < 2; i++) {
                float sa = sc[n][i] * A.scale_log2e, sb = sc[n][2 + i] * A.scale_log2e;
                if (!plain) {
                    const int key = c0 + 8 * n + 2 * t + i;
                    const bool oka = key <= lim_a && key < key_end && row_a < S, okb = key <= lim_b && key < key_end && row_b < S;
                    if (A.mask) {
                        if (oka) sa += load_as_float(A.mask, A.mask_dtype, mbase_a + key) * 1.4426950408889634f;
                        if (okb) sb += load_as_float(A.mask, A.mask_dtype, mbase_b + key) * 1.4426950408889634f;
                    }
                    if (!oka) sa = -CUDART_INF_F;
                    if (!okb) sb = -CUDART_INF_F;
                }
                sc[n][i] = sa; sc[n][2 + i] = sb;
                mx_a = fmaxf(mx_a, sa); mx_b = fmaxf(mx_b, sb);
            }
        mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 1)); mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 2));
        mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 1)); mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 2));
        const float mn_a = fmaxf(m_a, mx_a), mn_b = fmaxf(m_b, mx_b);
        const float ref_a = (mn_a == -CUDART_INF_F) ? 0.0f : mn_a, ref_b = (mn_b == -CUDART_INF_F) ? 0.0f : mn_b;   // all masked so far: p = 0
        const float corr_a = ex2_approx(m_a - ref_a), corr_b = ex2_approx(m_b - ref_b);
        m_a = mn_a; m_b = mn_b;
        uint32_t pa[4], pb[4];                        // A fragments of P (fp16 pairs), rows g and g + 8
#pragma unroll
        for (int n = 0; n < 4; n++) {
            const __half2 ha = __floats2half2_rn(ex2_approx(sc[n][0] - ref_a), ex2_approx(sc[n][1] - ref_a));
            const __half2 hb = __floats2half2_rn(ex2_approx(sc[n][2] - ref_b), ex2_approx(sc[n][3] - ref_b));
            pa[n] = *reinterpret_cast<const uint32_t*>(&ha);
            pb[n] = *reinterpret_cast<const uint32_t*>(&hb);
        }
        ol[0] *= corr_a; ol[1] *= corr_a; ol[2] *= corr_b; ol[3] *= corr_b;
#pragma unroll
        for (int u = 0; u < BPL; u++) { o[u][0] *= corr_a; o[u][1] *= corr_a; o[u][2] *= corr_b; o[u][3] *= corr_b; }
        // ---- O += P . V (and the denominators through a tile of ones)
#pragma unroll
        for (int j = 0; j < 2; j++) {
            uint32_t vw[4][VW];
#pragma unroll
            for (int h = 0; h < 4; h++) {
                const uint4 w = *reinterpret_cast<const uint4*>(stage + 4096 + (16 * j + 2 * t + (h & 1) + 8 * (h >> 1)) * 128 + ((g ^ (2 * t)) << 4));
                vw[h][0] = w.x; vw[h][1] = w.y; vw[h][2] = w.z; vw[h][3] = w.w;
            }
            mma16816_f16_full(ol, pa[2 * j], pb[2 * j], pa[2 * j + 1], pb[2 * j + 1], 0x3c003c00u, 0x3c003c00u);
#pragma unroll
            for (int q = 0; q < VW; q++) {
                // bytes of keys (2t, 2t+1) and (2t+8, 2t+9) interleaved: [a0 b0 a1 b1], [a2 b2 a3 b3]
                const uint32_t lo01 = __byte_perm(vw[0][q], vw[1][q], 0x5140u), lo23 = __byte_perm(vw[0][q], vw[1][q], 0x7362u);
                const uint32_t hi01 = __byte_perm(vw[2][q], vw[3][q], 0x5140u), hi23 = __byte_perm(vw[2][q], vw[3][q], 0x7362u);
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const uint32_t sel = (e & 1) ? 0x3424u : 0x1404u;
                    mma16816_f16_full(o[4 * q + e], pa[2 * j], pb[2 * j], pa[2 * j + 1], pb[2 * j + 1],
                                      __byte_perm(e < 2 ? lo01 : lo23, 0u, sel), __byte_perm(e < 2 ? hi01 : hi23, 0u, sel));
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ---- normalise and store: row g holds columns 32t .. 32t + 31 (o[u][i] = column 16 (2t + i) + u), row g + 8 likewise
    auto store_row = [&](int row, int base, float denom) {
        if (row >= S) return;
        const float inv = denom > 0.0f ? 1.0f / denom : 0.0f;         // a row with every key masked gives zeros
        const int64_t oi = (int64_t)b * A.o_sb + (int64_t)hq * A.o_sh + (int64_t)row * A.o_ss + 32 * t;
#pragma unroll
        for (int i = 0; i < 2; i++) {
            if (A.out_dtype == DT_F32) {
                float* p = (float*)A.out + oi + 16 * i;
#pragma unroll
                for (int u = 0; u < BPL; u += 4)
                    *reinterpret_cast<float4*>(p + u) = make_float4(o[u][base + i] * inv, o[u + 1][base + i] * inv, o[u + 2][base + i] * inv, o[u + 3][base + i] * inv);
            } else {
                uint32_t w[8];
#pragma unroll
                for (int u = 0; u < BPL; u += 2) {
                    const float x = o[u][base + i] * inv, y = o[u + 1][base + i] * inv;
                    if (A.out_dtype == DT_F16) { const __half2 h = __floats2half2_rn(x, y); w[u / 2] = *reinterpret_cast<const uint32_t*>(&h); }
                    else { const __nv_bfloat162 h = __floats2bfloat162_rn(x, y); w[u / 2] = *reinterpret_cast<const uint32_t*>(&h); }
                }
                uint16_t* p = (uint16_t*)A.out + oi + 16 * i;
                *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
                *reinterpret_cast<uint4*>(p + 8) = make_uint4(w[4], w[5], w[6], w[7]);
            }
        }
    };
    store_row(row_a, 0, ol[0]);
    store_row(row_b, 2, ol[2]);
}
