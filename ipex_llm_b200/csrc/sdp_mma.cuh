// sdp_mma.cuh -- included by kvcache.cu (inside namespace lowbit::{anonymous}, after SdpArgs and the helpers).
// Decode attention over the e5m2 cache with q.k AND p.v on the tensor cores (fp16 / bf16 queries, head_dim 64 or
// 128, up to 8 query rows per CTA = the GQA group x new tokens).
//
// Per warp and chunk of 32 keys, both products are mma.m16n8k16 (fp16 in, fp32 accumulate) with M = query rows
// (rows 0..7 live, rows 8..15 zero padding: a1 = a3 = 0 and the c2/c3 outputs are write-only dummies):
//   S[rows x 8 keys] = Q . K^T, four 8-key tiles.  A = Q, converted to fp16 once and kept in registers;
//     B = K^T: lane (g,t) supplies key g of the tile.  e5m2 -> fp16 is a byte permute (e5m2 = high byte of fp16).
//     The head dimension is consumed in a permuted order so that every lane reads 16 contiguous bytes per 64
//     elements of its key row: k-step (c,s) covers d = 64c + 16t + 4s + {0,1 | 2,3} for lane t; Q's fragments use
//     the same order (a dot product does not care).
//   O[rows x D] += P . V, two 16-key steps x D/8 column tiles.  The score accumulators ARE the A fragments of P
//     (lane (g,t): row g, keys 2t, 2t+1 of each tile), rounded to fp16.  B = V: lane (g,t) reads D/8 contiguous
//     bytes at column (D/8) g of the four value rows 2t, 2t+1, 2t+8, 2t+9 of the step; column tile u takes byte u
//     of each, i.e. the tile's columns are d = (D/8) n + u, n = 0..7 -- again a permutation, undone when the
//     accumulators are written out.  One extra column tile of ones accumulates the softmax denominator from the
//     SAME rounded p.
// Online softmax runs in the accumulator layout: a row's 32 scores of the chunk sit in the 4 lanes of a quad.
// bf16 queries are converted to fp16 (exact inside fp16's normal range, which attention queries live in).

__device__ __forceinline__ void mma16816_f16(float& c0, float& c1, float& z0, float& z1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c0), "+f"(c1), "+f"(z0), "+f"(z1)
        : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}

// 2^x, x <= 0 or -inf (-> 0): one MUFU, no range fix-up
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr int SDP_MMA_ROWS = 8;

// ---- pieces of the fused decode step (FUSED = 1): mbarrier / cluster primitives for the in-cluster merge
__device__ __forceinline__ uint32_t sdp_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sdp_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void sdp_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sdp_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t sdp_map_to_rank(uint32_t local_addr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    return remote;
}
// remote store whose 4 bytes are counted on the destination CTA's mbarrier: the sender needs no fence and may exit
__device__ __forceinline__ void sdp_st_async_f32(uint32_t remote_addr, float v, uint32_t remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
                 ::"r"(remote_addr), "r"(__float_as_uint(v)), "r"(remote_bar) : "memory");
}
__device__ __forceinline__ float sdp_round_to(float v, int dt) {
    if (dt == DT_F16) return __half2float(__float2half_rn(v));
    if (dt == DT_BF16) return __bfloat162float(__float2bfloat16_rn(v));
    return v;
}

// FUSED = 1 is the whole attention of ONE decode step in one launch (S = 1, D = 128, NST > 0): the query rows are
// read from the merged qkv projection's output and rotated here; the new key (rotated, rounded to the activation dtype,
// then e5m2) and the new value (e5m2) are computed by every CTA whose key range holds the new position -- it patches them
// into its staged chunk, so nobody depends on another CTA's write -- and written to the cache by tile 0; the key splits of a
// (sequence, kv head) form a thread-block cluster (<= 8 CTAs) whose members PUSH their partial (m, l, o) into the leader's
// shared memory (st.async + mbarrier transaction count, as the GEMV's split-K reduction does); the leader merges in rank
// order and writes the output.  Same arithmetic as the separate kernels (rope.cu, kv_quantize, this kernel + the merge
// kernel): rotate -> round -> quantise, the same split-local softmax, the same merge formula.
template <int D, int OCC, int NST, int FUSED>
__global__ void __launch_bounds__(SDP_WARPS * 32, OCC)
sdp_fp8_decode_mma_kernel(const SdpArgs A) {
    constexpr int NC = D / 64;                       // 16-byte pieces of a key row per lane
    constexpr int BPL = D / 8;                       // bytes of a value row per lane = output column tiles
    constexpr int VW = BPL / 4;                      // ... as 32-bit words
    static_assert(D == 64 || D == 128, "head_dim 64 or 128");
    static_assert(NST == 0 || D == 128, "the staged variant is written for 128-byte rows");
    static_assert(!FUSED || (NST > 0 && D == 128), "the fused decode step is written for head_dim 128 and staged chunks");
    extern __shared__ __align__(128) uint8_t sdp_ring[];            // NST > 0: [warp][stage][K 4 KB | V 4 KB]
    __shared__ float red_m[SDP_WARPS][SDP_MMA_ROWS], red_l[SDP_WARPS][SDP_MMA_ROWS];
    constexpr int RO = SDP_MMA_ROWS * (D + 16);       // floats of one warp's staged accumulators (NST > 0: in its ring)
    __shared__ __align__(16) float red_o_static[NST > 0 ? 4 : SDP_WARPS * RO];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int split = blockIdx.x, hkv = blockIdx.y % A.Hkv, b = blockIdx.y / A.Hkv, tile = blockIdx.z;
    const int n_rep = A.Hq / A.Hkv, rows_total = n_rep * A.S;
    const int row0 = tile * A.rt;
    const int n_rows = min(A.rt, rows_total - row0);

    asm volatile("griddepcontrol.launch_dependents;");               // the merge kernel may be scheduled as CTAs here retire
    // FUSED: leader's receive barrier; the cluster barrier that publishes it is split around the main loop (free)
    const uint32_t rbar = sdp_smem_u32(sdp_ring) + SDP_WARPS * NST * 8192;
    float* const recv = reinterpret_cast<float*>(sdp_ring + SDP_WARPS * NST * 8192 + 16);
    if (FUSED) {
        if (A.n_split > 1) {
            if (split == 0 && threadIdx.x == 0) {
                sdp_mbar_init(rbar, 1);
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            }
            asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
        }
        // qkv comes from the previous kernel.  The cache length and the cached rows come from earlier steps: a caller that
        // knows them complete (A.early: an ordinarily launched kernel lies between their producers and this call) lets the
        // first chunks stream in while the previous kernel drains; otherwise nothing is read before the wait.
        if (!A.early) asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    int L, keys_per_split;
    sdp_lengths(A, L, keys_per_split);
    const int new_pos = L - 1;                        // FUSED: position of the token being decoded
    const int key_begin = split * keys_per_split;
    const int key_end = min(L, key_begin + keys_per_split);
    // NST > 0: every warp streams its chunks through a ring of NST stages filled by cp.async (16 bytes per lane and
    // instruction, rows coalesced), NST - 1 chunks ahead of the one being used.  Inside a stage the 16-byte pieces of
    // a row are XOR-swizzled so that the fragment reads below are bank-conflict free: key rows by (row & 1) << 2
    // (a quarter warp reads 4 pieces of 2 consecutive rows), value rows by ((row >> 1) & 3) << 1 (a quarter warp
    // reads 2 pieces of rows 2t).
    uint8_t* const ring = sdp_ring + warp * (NST * 8192);
    const uint8_t* const kb0 = A.k + b * A.k_sb + hkv * A.k_sh + 16 * (lane & 7);
    const uint8_t* const vb0 = A.v + b * A.v_sb + hkv * A.v_sh + 16 * (lane & 7);
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
    const int r0 = lane >> 3, unit = lane & 7;        // lane copies piece `unit` of rows r0, r0 + 4, ...
    const uint32_t k_off = r0 * 128 + ((unit ^ ((r0 & 1) << 2)) << 4);
    const uint32_t v_off0 = 4096 + r0 * 128 + ((unit ^ (((r0 >> 1) & 3) << 1)) << 4);          // rows r0 + 8i'
    const uint32_t v_off1 = 4096 + (r0 + 4) * 128 + ((unit ^ ((((r0 >> 1) + 2) & 3) << 1)) << 4);   // rows r0 + 4 + 8i'
    const bool rows_128 = A.k_ss == 128 && A.v_ss == 128;
    auto stage_chunk = [&](int c0, int slot) {
        if (c0 < key_end) {
            const uint32_t dst = ring_s + slot * 8192;
            if (c0 + 32 <= key_end && rows_128) {     // a whole chunk of densely stored rows: constant address steps
                const uint8_t* kp = kb0 + (int64_t)(c0 + r0) * 128;
                const uint8_t* vp = vb0 + (int64_t)(c0 + r0) * 128;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst + k_off + i * 512), "l"(kp + i * 512) : "memory");
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst + ((i & 1) ? v_off1 : v_off0) + (i >> 1) * 1024), "l"(vp + i * 512) : "memory");
                }
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int64_t key = min(c0 + 4 * i + r0, key_end - 1);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst + k_off + i * 512), "l"(kb0 + key * A.k_ss) : "memory");
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst + ((i & 1) ? v_off1 : v_off0) + (i >> 1) * 1024), "l"(vb0 + key * A.v_ss) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (FUSED && A.early) {
#pragma unroll
        for (int s = 0; s < NST - 1; s++) stage_chunk(key_begin + (warp + SDP_WARPS * s) * 32, s);
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }

    // A fragments of Q: lane (g,t) holds row g, elements 64c + 16t .. +15 as 8 fp16 pairs (zeros for rows >= n_rows)
    uint32_t qa[NC][8];
#pragma unroll
    for (int c = 0; c < NC; c++)
#pragma unroll
        for (int i = 0; i < 8; i++) qa[c][i] = 0u;
    int my_hq = 0, my_st = 0;                         // query head / token of this lane's row
    if (FUSED) {
        if (g < n_rows) {
            my_hq = hkv * n_rep + row0 + g;
            const uint16_t* qrow = reinterpret_cast<const uint16_t*>(A.qkv) + b * A.qkv_sb + (int64_t)my_hq * D + 16 * t;
            const int pr = min(max(new_pos, 0), A.n_pos - 1);
            uint4 xw[2][2], cw[2], sw[2];             // x: [c][half of 16], cos / sin: 16 entries of the row
#pragma unroll
            for (int c = 0; c < 2; c++) { xw[c][0] = *reinterpret_cast<const uint4*>(qrow + 64 * c); xw[c][1] = *reinterpret_cast<const uint4*>(qrow + 64 * c + 8); }
            float cf[16], sf[16];
            if (A.cs_dtype == DT_F32) {
#pragma unroll
                for (int e = 0; e < 16; e++) { cf[e] = ((const float*)A.cos)[pr * A.cs_row + 16 * t + e]; sf[e] = ((const float*)A.sin)[pr * A.cs_row + 16 * t + e]; }
            } else {
                const uint16_t* cr = reinterpret_cast<const uint16_t*>(A.cos) + pr * A.cs_row + 16 * t;
                const uint16_t* sr = reinterpret_cast<const uint16_t*>(A.sin) + pr * A.cs_row + 16 * t;
                cw[0] = *reinterpret_cast<const uint4*>(cr); cw[1] = *reinterpret_cast<const uint4*>(cr + 8);
                sw[0] = *reinterpret_cast<const uint4*>(sr); sw[1] = *reinterpret_cast<const uint4*>(sr + 8);
                const uint16_t* ce = reinterpret_cast<const uint16_t*>(cw);
                const uint16_t* se = reinterpret_cast<const uint16_t*>(sw);
#pragma unroll
                for (int e = 0; e < 16; e++) {
                    if (A.cs_dtype == DT_F16) { __half_raw a, c2; a.x = ce[e]; c2.x = se[e]; cf[e] = __half2float(__half(a)); sf[e] = __half2float(__half(c2)); }
                    else { cf[e] = __uint_as_float((uint32_t)ce[e] << 16); sf[e] = __uint_as_float((uint32_t)se[e] << 16); }
                }
            }
            const uint16_t* x0e = reinterpret_cast<const uint16_t*>(xw[0]);
            const uint16_t* x1e = reinterpret_cast<const uint16_t*>(xw[1]);
            float y0[16], y1[16];
#pragma unroll
            for (int e = 0; e < 16; e++) {
                float x0, x1;
                if (A.q_dtype == DT_F16) { __half_raw a, c2; a.x = x0e[e]; c2.x = x1e[e]; x0 = __half2float(__half(a)); x1 = __half2float(__half(c2)); }
                else { x0 = __uint_as_float((uint32_t)x0e[e] << 16); x1 = __uint_as_float((uint32_t)x1e[e] << 16); }
                y0[e] = sdp_round_to(fmaf(x0, cf[e], -x1 * sf[e]), A.q_dtype);
                y1[e] = sdp_round_to(fmaf(x1, cf[e], x0 * sf[e]), A.q_dtype);
            }
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const __half2 h0 = __floats2half2_rn(y0[2 * i], y0[2 * i + 1]), h1 = __floats2half2_rn(y1[2 * i], y1[2 * i + 1]);
                qa[0][i] = *reinterpret_cast<const uint32_t*>(&h0);
                qa[1 % NC][i] = *reinterpret_cast<const uint32_t*>(&h1);
            }
        }
    } else
    if (g < n_rows) {
        const int rr = row0 + g;
        my_hq = hkv * n_rep + rr / A.S; my_st = rr % A.S;
        const uint16_t* qrow = reinterpret_cast<const uint16_t*>(A.q) + b * A.q_sb + my_hq * A.q_sh + my_st * A.q_ss + 16 * t;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const uint4 w0 = *reinterpret_cast<const uint4*>(qrow + 64 * c), w1 = *reinterpret_cast<const uint4*>(qrow + 64 * c + 8);
            const uint32_t ws[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (A.q_dtype == DT_F16) qa[c][i] = ws[i];
                else {                                  // bf16 pair -> fp16 pair
                    const __half2 h = __floats2half2_rn(__uint_as_float(ws[i] << 16), __uint_as_float(ws[i] & 0xffff0000u));
                    qa[c][i] = *reinterpret_cast<const uint32_t*>(&h);
                }
            }
        }
    }

    // FUSED: the new row of this kv head -- lane l owns key elements l, l + 32, l + 64, l + 96 (two rotary pairs) and
    // value elements 4l .. 4l + 3 -- for every CTA whose range holds the new position; tile 0 also writes it to the cache
    const bool has_new = FUSED && new_pos >= key_begin && new_pos < key_end;
    uint32_t knew = 0, vnew = 0;
    if (has_new) {
        const int64_t kbase = b * A.qkv_sb + (int64_t)(A.Hq + hkv) * D, vbase = kbase + (int64_t)A.Hkv * D;
        const int pr = min(new_pos, A.n_pos - 1);
        float kx[4];
#pragma unroll
        for (int j = 0; j < 4; j++) kx[j] = load_as_float(A.qkv, A.q_dtype, kbase + lane + 32 * j);
#pragma unroll
        for (int j = 0; j < 2; j++) {                  // pairs (lane + 32 j, lane + 32 j + 64)
            const float c = load_as_float(A.cos, A.cs_dtype, pr * A.cs_row + lane + 32 * j), sn = load_as_float(A.sin, A.cs_dtype, pr * A.cs_row + lane + 32 * j);
            const float y0 = sdp_round_to(fmaf(kx[j], c, -kx[j + 2] * sn), A.q_dtype), y1 = sdp_round_to(fmaf(kx[j + 2], c, kx[j] * sn), A.q_dtype);
            knew |= (uint32_t)to_e5m2(y0) << (8 * j);
            knew |= (uint32_t)to_e5m2(y1) << (8 * (j + 2));
        }
#pragma unroll
        for (int j = 0; j < 4; j++) vnew |= (uint32_t)to_e5m2(load_as_float(A.qkv, A.q_dtype, vbase + 4 * lane + j)) << (8 * j);
        if (tile == 0 && warp == 0) {
            uint8_t* kd = A.kw + b * A.k_sb + hkv * A.k_sh + (int64_t)new_pos * A.k_ss;
#pragma unroll
            for (int j = 0; j < 4; j++) kd[lane + 32 * j] = (uint8_t)(knew >> (8 * j));
            *reinterpret_cast<uint32_t*>(A.vw + b * A.v_sb + hkv * A.v_sh + (int64_t)new_pos * A.v_ss + 4 * lane) = vnew;
        }
    }
    const uint8_t* kb = A.k + b * A.k_sb + hkv * A.k_sh + 16 * t;
    const uint8_t* vb = A.v + b * A.v_sb + hkv * A.v_sh + BPL * g;
    const int64_t mask_base = b * A.m_sb + my_hq * A.m_sh + my_st * A.m_ss;
    const int causal_last = A.causal ? L - A.S + my_st : L;           // last key this lane's row sees
    const int causal_first = A.causal ? L - A.S : L;                  // ... and the first row of the tile at least

    float m_run = -CUDART_INF_F;                      // of row g, the same in the quad's four lanes
    float o[BPL][2], ol[2] = {0.0f, 0.0f};            // row g: columns BPL (2t + i) + u;  ol = softmax denominator
    float z[4][2];                                    // rows 8..15 of every accumulator tile: always zero
#pragma unroll
    for (int u = 0; u < BPL; u++) o[u][0] = o[u][1] = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i++) z[i][0] = z[i][1] = 0.0f;

    if (NST > 0 && !(FUSED && A.early)) {
#pragma unroll
        for (int s = 0; s < NST - 1; s++) stage_chunk(key_begin + (warp + SDP_WARPS * s) * 32, s);
    }
    int slot = 0;
    for (int c0 = key_begin + warp * 32; c0 < key_end; c0 += SDP_WARPS * 32) {
        uint4 kw[4][NC];
        uint32_t vw[2][4][VW];                        // [16-key step][row 2t, 2t+1, 2t+8, 2t+9][word]
        const uint8_t* stage = ring + slot * 8192;
        if (NST > 0) {
            int ahead = slot + NST - 1; if (ahead >= NST) ahead -= NST;
            stage_chunk(c0 + (NST - 1) * SDP_WARPS * 32, ahead);
            asm volatile("cp.async.wait_group %0;" :: "n"(NST > 0 ? NST - 1 : 0) : "memory");
            if (has_new && new_pos >= c0 && new_pos < c0 + 32) {
                // the chunk holds the new position: its row comes from this CTA's own registers, not from the cache
                __syncwarp();
                const int r = new_pos - c0;
                uint8_t* st8 = ring + slot * 8192;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int d = lane + 32 * j;
                    st8[r * 128 + ((((d >> 4) ^ ((r & 1) << 2))) << 4) + (d & 15)] = (uint8_t)(knew >> (8 * j));
                }
                const int dv = 4 * lane;
                *reinterpret_cast<uint32_t*>(st8 + 4096 + r * 128 + ((((dv >> 4) ^ (((r >> 1) & 3) << 1))) << 4) + (dv & 15)) = vnew;
            }
            __syncwarp();
        } else {
        // ---- all loads of the chunk first (one memory round trip per chunk).  Keys beyond the range are clamped
        // to the last one; their p is 0.
#pragma unroll
        for (int n = 0; n < 4; n++) {
            const uint4* kr = reinterpret_cast<const uint4*>(kb + (int64_t)min(c0 + 8 * n + g, key_end - 1) * A.k_ss);
#pragma unroll
            for (int c = 0; c < NC; c++) kw[n][c] = kr[4 * c];
        }
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int h = 0; h < 4; h++) {
                const int key = min(c0 + 16 * j + 2 * t + (h & 1) + 8 * (h >> 1), key_end - 1);
                const uint8_t* vr = vb + (int64_t)key * A.v_ss;
                if (VW == 4) {
                    const uint4 w = *reinterpret_cast<const uint4*>(vr);
                    vw[j][h][0] = w.x; vw[j][h][1] = w.y; vw[j][h][VW - 2] = w.z; vw[j][h][VW - 1] = w.w;
                } else {
                    const uint2 w = *reinterpret_cast<const uint2*>(vr);
                    vw[j][h][0] = w.x; vw[j][h][1] = w.y;
                }
            }
        }
        // ---- scores of 4 tiles x 8 keys
        float sc[4][2];
#pragma unroll
        for (int n = 0; n < 4; n++) {
            sc[n][0] = sc[n][1] = 0.0f;
#pragma unroll
            for (int c = 0; c < NC; c++) {
                if (NST > 0) kw[n][c] = *reinterpret_cast<const uint4*>(stage + (8 * n + g) * 128 + (((4 * c + t) ^ ((g & 1) << 2)) << 4));
                const uint32_t ws[4] = {kw[n][c].x, kw[n][c].y, kw[n][c].z, kw[n][c].w};
#pragma unroll
                for (int s = 0; s < 4; s++)
                    mma16816_f16(sc[n][0], sc[n][1], z[n][0], z[n][1], qa[c][2 * s], qa[c][2 * s + 1],
                                 __byte_perm(ws[s], 0u, 0x1404u), __byte_perm(ws[s], 0u, 0x3424u));
            }
        }
        // ---- scale, mask, online softmax of row g (its 32 scores of the chunk live in the 4 lanes of the quad); this
        // lane's values are keys c0 + 8n + 2t + i.  p = 2^(s - m_run) with m_run a reference point, not necessarily
        // the exact running maximum.
        uint32_t pa[4];                               // A fragments of P (fp16 pairs)
        float corr = 1.0f;
        if (c0 + 32 <= key_end && c0 + 31 <= causal_first && !A.mask) {
            // every key of the chunk is visible to every row (the common case): no predicates, the scale folded into
            // the exponent, and the accumulators are rescaled only when some row's maximum grew by more than 2^8
            // (p <= 256 then; fp16 holds it with the same relative precision)
            float mx = fmaxf(fmaxf(fmaxf(sc[0][0], sc[0][1]), fmaxf(sc[1][0], sc[1][1])), fmaxf(fmaxf(sc[2][0], sc[2][1]), fmaxf(sc[3][0], sc[3][1])));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            const float m_new = mx * A.scale_log2e;   // scale > 0 (checked by the launcher)
            const bool grow = m_new > m_run + 8.0f;
            if (__any_sync(0xffffffffu, grow)) {
                if (grow) { corr = ex2_approx(m_run - m_new); m_run = m_new; }
                ol[0] *= corr; ol[1] *= corr;
#pragma unroll
                for (int u = 0; u < BPL; u++) { o[u][0] *= corr; o[u][1] *= corr; }
            }
#pragma unroll
            for (int n = 0; n < 4; n++) {
                const __half2 h = __floats2half2_rn(ex2_approx(fmaf(sc[n][0], A.scale_log2e, -m_run)), ex2_approx(fmaf(sc[n][1], A.scale_log2e, -m_run)));
                pa[n] = *reinterpret_cast<const uint32_t*>(&h);
            }
        } else {
            float mx = -CUDART_INF_F;
#pragma unroll
            for (int n = 0; n < 4; n++)
#pragma unroll
                for (int i = 0; i < 2; i++) {
                    const int key = c0 + 8 * n + 2 * t + i;
                    float s = sc[n][i] * A.scale_log2e;
                    const bool ok = key < key_end && g < n_rows && key <= causal_last;
                    if (ok && A.mask) s += load_as_float(A.mask, A.mask_dtype, mask_base + key) * 1.4426950408889634f;
                    if (!ok) s = -CUDART_INF_F;
                    sc[n][i] = s;
                    mx = fmaxf(mx, s);
                }
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            const float m_new = fmaxf(m_run, mx);
            const float m_ref = (m_new == -CUDART_INF_F) ? 0.0f : m_new;          // all keys masked so far: p = 2^-inf = 0
            corr = ex2_approx(m_run - m_ref);
            m_run = m_new;
#pragma unroll
            for (int n = 0; n < 4; n++) {
                const __half2 h = __floats2half2_rn(ex2_approx(sc[n][0] - m_ref), ex2_approx(sc[n][1] - m_ref));
                pa[n] = *reinterpret_cast<const uint32_t*>(&h);
            }
            ol[0] *= corr; ol[1] *= corr;
#pragma unroll
            for (int u = 0; u < BPL; u++) { o[u][0] *= corr; o[u][1] *= corr; }
        }
        // ---- O += P . V (and the denominator through a tile of ones)
#pragma unroll
        for (int j = 0; j < 2; j++) {
            if (NST > 0) {
#pragma unroll
                for (int h = 0; h < 4; h++) {
                    const uint4 w = *reinterpret_cast<const uint4*>(stage + 4096 + (16 * j + 2 * t + (h & 1) + 8 * (h >> 1)) * 128 + ((g ^ (2 * t)) << 4));
                    vw[j][h][0] = w.x; vw[j][h][1] = w.y; vw[j][h][VW - 2] = w.z; vw[j][h][VW - 1] = w.w;
                }
            }
            mma16816_f16(ol[0], ol[1], z[j][0], z[j][1], pa[2 * j], pa[2 * j + 1], 0x3c003c00u, 0x3c003c00u);
#pragma unroll
            for (int q = 0; q < VW; q++) {
                // bytes of keys (2t, 2t+1) and (2t+8, 2t+9) interleaved: [a0 b0 a1 b1], [a2 b2 a3 b3]
                const uint32_t lo01 = __byte_perm(vw[j][0][q], vw[j][1][q], 0x5140u), lo23 = __byte_perm(vw[j][0][q], vw[j][1][q], 0x7362u);
                const uint32_t hi01 = __byte_perm(vw[j][2][q], vw[j][3][q], 0x5140u), hi23 = __byte_perm(vw[j][2][q], vw[j][3][q], 0x7362u);
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const uint32_t sel = (e & 1) ? 0x3424u : 0x1404u;
                    const int u = 4 * q + e;
                    mma16816_f16(o[u][0], o[u][1], z[e][0], z[e][1], pa[2 * j], pa[2 * j + 1],
                                 __byte_perm(e < 2 ? lo01 : lo23, 0u, sel), __byte_perm(e < 2 ? hi01 : hi23, 0u, sel));
                }
            }
        }
        if (NST > 0) {
            __syncwarp();                             // the stage is refilled by the next iteration's cp.async
            if (++slot == NST) slot = 0;
        }
    }
    // ---- combine the warps (m, l, o) in a fixed order; the accumulator columns go back to their places
    // (each lane owns runs of BPL consecutive columns; for D = 128 the 16-byte pieces of a run are XOR-swizzled
    // by t and rows are padded by 16 floats so that the float4 stores of a quarter warp cover all 32 banks)
    if (NST > 0) { asm volatile("cp.async.wait_group 0;" ::: "memory"); __syncwarp(); }
    auto red_o = [&](int w) -> float* { return NST > 0 ? reinterpret_cast<float*>(sdp_ring + w * (NST * 8192)) : red_o_static + w * RO; };
    if (t == 0) { red_m[warp][g] = m_run; red_l[warp][g] = ol[0]; }
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int q = 0; q < VW; q++)
            *reinterpret_cast<float4*>(red_o(warp) + g * (D + 16) + BPL * (2 * t + i) + 4 * (D == 128 ? (q ^ t) : q)) =
                make_float4(o[4 * q][i], o[4 * q + 1][i], o[4 * q + 2][i], o[4 * q + 3][i]);
    __syncthreads();
    float* part = FUSED ? nullptr : A.part + ((((int64_t)b * A.Hkv + hkv) * gridDim.z + tile) * A.n_split + split) * (A.rt * (D + 2));
    const int ps = A.rt * (D + 2);                                    // floats of one CTA's partial result
    if (FUSED && A.n_split > 1) {
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // the leader's barrier init is visible
        if (split == 0 && threadIdx.x == 0) sdp_mbar_expect_tx(rbar, (uint32_t)((A.n_split - 1) * ps * 4));
    }
    const uint32_t rb = FUSED ? sdp_map_to_rank(rbar, 0) : 0u;
    for (int i = threadIdx.x; i < A.rt * D; i += blockDim.x) {
        const int r = i / D, d = i % D;
        const int pos = D == 128 ? ((d & ~12) | ((((d >> 2) ^ (d >> 5)) & 3) << 2)) : d;       // where column d was staged
        float M = red_m[0][r];
#pragma unroll
        for (int w = 1; w < SDP_WARPS; w++) M = fmaxf(M, red_m[w][r]);
        float acc = 0.0f, lsum = 0.0f;
#pragma unroll
        for (int w = 0; w < SDP_WARPS; w++) {
            const float f = (red_m[w][r] == -CUDART_INF_F) ? 0.0f : exp2f(red_m[w][r] - M);
            acc = fmaf(red_o(w)[r * (D + 16) + pos], f, acc);
            lsum = fmaf(red_l[w][r], f, lsum);
        }
        if (!FUSED) {
            part[r * (D + 2) + 2 + d] = acc;
            if (d == 0) { part[r * (D + 2)] = M; part[r * (D + 2) + 1] = lsum; }
        } else if (split != 0) {
            // a peer: push (m, l, o) of the row into the leader's receive area and be done
            const uint32_t dst = sdp_smem_u32(recv + (split - 1) * ps + r * (D + 2));
            sdp_st_async_f32(sdp_map_to_rank(dst + 4 * (2 + d), 0), acc, rb);
            if (d == 0) {
                sdp_st_async_f32(sdp_map_to_rank(dst, 0), M, rb);
                sdp_st_async_f32(sdp_map_to_rank(dst + 4, 0), lsum, rb);
            }
        } else {
            // the leader: merge the splits in rank order (the merge kernel's formula), normalise, store
            if (A.n_split > 1) sdp_mbar_wait(rbar, 0);
            float Mall = M;
            for (int k = 1; k < A.n_split; k++) Mall = fmaxf(Mall, recv[(k - 1) * ps + r * (D + 2)]);
            const float f0 = (M == -CUDART_INF_F) ? 0.0f : exp2f(M - Mall);
            float o = acc * f0, l = lsum * f0;
            for (int k = 1; k < A.n_split; k++) {
                const float* pk = recv + (k - 1) * ps + r * (D + 2);
                const float f = (pk[0] == -CUDART_INF_F) ? 0.0f : exp2f(pk[0] - Mall);
                o = fmaf(pk[2 + d], f, o);
                l = fmaf(pk[1], f, l);
            }
            const int rr = row0 + r;
            if (rr < rows_total) {
                const float y = l > 0.0f ? o / l : 0.0f;
                const int64_t oi = ((int64_t)b * A.Hq + (hkv * n_rep + rr)) * D + d;          // S == 1
                if (A.out_dtype == DT_F32) ((float*)A.out)[oi] = y;
                else if (A.out_dtype == DT_F16) ((__half*)A.out)[oi] = __float2half_rn(y);
                else ((__nv_bfloat16*)A.out)[oi] = __float2bfloat16_rn(y);
            }
        }
    }
}
