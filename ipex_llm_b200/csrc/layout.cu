// layout.cu -- canonical <-> tiled relayout and dequantisation kernels (load-time / helper
// paths; HBM-bound byte shuffles, one output byte or element per thread).
//
// Replaces ggml_q_format_convet_cpu2xpu / xpu2cpu (reference low_bit_linear.py:198-253,
// llama_cpp.py:1102-1135) and ggml_dequantize / ggml_dequantize_q4_0
// (low_bit_linear.py:256-279, llama_cpp.py:1068-1099).
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {

// ---------------------------------------------------------------- canonical -> tiled
// one thread per output byte of the Q region (gathers its codes from canonical records),
// then one thread per fp16 of the S / M regions, then a straight copy of the tail records.
__global__ void relayout_to_tiled_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                         TiledGeom G, int qt) {
    const int64_t q_bytes = G.off_s;
    const int64_t sm_halves = (G.off_tail - G.off_s) / 2;
    const int64_t tail_bytes = (G.rows - G.r64) * G.nb * G.blk_bytes;
    const int64_t total = q_bytes + sm_halves + tail_bytes;
    const int hdr = header_bytes(qt);
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        if (idx < q_bytes) {
            const int64_t u = idx / G.super_code_bytes;
            const int rem = (int)(idx % G.super_code_bytes);
            const int tr = rem / G.tile_code_bytes, off = rem % G.tile_code_bytes;
            const int64_t rg = u / G.nb, kb = u % G.nb;
            // canonical record of (row, kb): src + (row*nb + kb)*blk ; r = row inside the 16-row tile
            auto rec = [&](int r) { return src + ((rg * RG_ROWS + tr * 16 + r) * G.nb + kb) * G.blk_bytes + hdr; };
            uint32_t out = 0;
            if (qt == QT_FP8_E4M3 || qt == QT_FP8_E5M2) {
                const int h = off >> 9, l = (off & 511) >> 4, e = off & 15;
                const int g = l >> 2, t = l & 3;
                out = canonical_code(qt, rec(g + 8 * h), 16 * t + e);
            } else if (qt == QT_FP6) {
                if (off < 512) {
                    const int l = off >> 4, rem = off & 15, h = rem >> 3, u = (rem >> 2) & 1, pb = rem & 3;
                    const int g = l >> 2, t = l & 3;
                    for (int sub = 0; sub < 2; sub++) {
                        const int p = 2 * pb + sub, i = p >> 1, sodd = p & 1;
                        const int e = 8 * u + 4 * sodd + i;
                        out |= (canonical_code(qt, rec(g + 8 * h), 16 * t + e) & 0x0f) << (4 * sub);
                    }
                } else {
                    const int o = off - 512, l = o >> 3, rem = o & 7, h = rem >> 2, i = rem & 3;
                    const int g = l >> 2, t = l & 3;
                    for (int s = 0; s < 4; s++)
                        out |= ((canonical_code(qt, rec(g + 8 * h), 16 * t + 4 * s + i) >> 4) & 0x03) << (2 * s);
                }
            } else if (qt == QT_NF4) {  // byte = fragment register: pair of codes (tilednf4_pos)
                const int l = off >> 4, s = (off & 15) >> 2, pb = off & 3;
                const int g = l >> 2, t = l & 3, h = pb & 1, ihi = pb >> 1;
                for (int sub = 0; sub < 2; sub++)
                    out |= canonical_code(qt, rec(g + 8 * h), 16 * t + 4 * s + 2 * ihi + sub) << (4 * sub);
            } else {  // 4-bit integer types
                const int l = off >> 4, s = (off & 15) >> 2, pb = off & 3;
                const int g = l >> 2, t = l & 3;
                for (int sub = 0; sub < 2; sub++) {
                    const int p = 2 * pb + sub, h = p & 1, i = ((p >> 1) & 1) * 2 + (p >> 2);
                    out |= canonical_code(qt, rec(g + 8 * h), 16 * t + 4 * s + i) << (4 * sub);
                }
            }
            dst[idx] = (uint8_t)out;
        } else if (idx < q_bytes + sm_halves) {
            int64_t hidx = idx - q_bytes;                 // index of the fp16 in S (then M)
            const int which = hidx >= G.nsuper * 64 ? 1 : 0;   // 0: d, 1: m
            hidx -= which ? G.nsuper * 64 : 0;
            const int64_t u = hidx / 64;
            const int e = (int)(hidx % 64), g = e >> 3, tr = (e & 7) >> 1, h = e & 1;   // [g][r][h]
            const int64_t rg = u / G.nb, kb = u % G.nb;
            const uint8_t* r = src + ((rg * RG_ROWS + tr * 16 + g + 8 * h) * G.nb + kb) * G.blk_bytes + 2 * which;
            uint8_t* o = dst + G.off_s + (which ? G.nsuper * 128 : 0) + hidx * 2;
            o[0] = r[0]; o[1] = r[1];
        } else {
            const int64_t b = idx - q_bytes - sm_halves;
            dst[G.off_tail + b] = src[G.r64 * G.nb * G.blk_bytes + b];
        }
    }
}

// ---------------------------------------------------------------- tiled -> canonical
// one thread per output canonical byte.
__global__ void relayout_to_canonical_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                             TiledGeom G, int qt) {
    const int64_t total = G.rows * G.nb * G.blk_bytes;
    const int64_t main_bytes = G.r64 * G.nb * G.blk_bytes;
    const int hdr = header_bytes(qt);
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        if (idx >= main_bytes) { dst[idx] = src[G.off_tail + (idx - main_bytes)]; continue; }
        const int64_t blk = idx / G.blk_bytes;
        const int b = (int)(idx % G.blk_bytes);
        const int64_t row = blk / G.nb, kb = blk % G.nb;
        const int r = (int)(row % 16);
        if (b < hdr) {
            const int which = b >> 1;
            dst[idx] = src[(which ? G.off_m : G.off_s) + scale_offset(G, row, kb) + (b & 1)];
            continue;
        }
        const uint8_t* tile = src + tile_code_offset(G, row, kb);
        const int cb = b - hdr;
        uint32_t out;
        if (qt == QT_FP8_E4M3 || qt == QT_FP8_E5M2) out = tiled_code(qt, tile, r, cb);
        else if (qt == QT_FP6) {
            const int grp = cb / 3, byte = cb % 3;
            uint32_t w = 0;
            for (int i = 0; i < 4; i++) w |= tiled_code(qt, tile, r, 4 * grp + i) << (6 * i);
            out = (w >> (8 * byte)) & 0xff;
        } else out = tiled_code(qt, tile, r, cb) | (tiled_code(qt, tile, r, cb + 32) << 4);
        dst[idx] = (uint8_t)out;
    }
}

// ---------------------------------------------------------------- dequantise
template <typename OutT, bool TILED>
__global__ void dequant_kernel(const uint8_t* __restrict__ src, OutT* __restrict__ dst, int64_t n_elems,
                               TiledGeom G, int qt) {
    const int hdr = header_bytes(qt);
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < n_elems;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t blk = idx / QK;
        const int j = (int)(idx % QK);
        uint32_t code; uint16_t dh, mh = 0;
        bool tiled_here = false;
        if (TILED) {
            const int64_t row = blk / G.nb, kb = blk % G.nb;
            if (row < G.r64) {
                tiled_here = true;
                code = tiled_code(qt, src + tile_code_offset(G, row, kb), (int)(row % 16), j);
                dh = *(const uint16_t*)(src + G.off_s + scale_offset(G, row, kb));
                if (G.has_m) mh = *(const uint16_t*)(src + G.off_m + scale_offset(G, row, kb));
            }
        }
        if (!tiled_here) {
            const uint8_t* rec = TILED ? src + G.off_tail + (blk - G.r64 * G.nb) * G.blk_bytes
                                       : src + blk * G.blk_bytes;
            dh = (uint16_t)rec[0] | ((uint16_t)rec[1] << 8);
            if (G.has_m) mh = (uint16_t)rec[2] | ((uint16_t)rec[3] << 8);
            code = canonical_code(qt, rec + hdr, j);
        }
        const float d = half_bits_to_float(dh);
        float v = __fmul_rn(code_value(qt, code), d);
        if (G.has_m) v = __fadd_rn(v, half_bits_to_float(mh));
        dst[idx] = from_float<OutT>(v);
    }
}

// rows of a TILED-layout weight picked by an index vector (LowBitEmbedding, reference embedding.py:214-224:
// xe_linear.dequantize_rows): out[t][:] = dequant(W[ids[t]][:]); an index outside [0, rows) gives a zero row.
template <typename OutT>
__global__ void dequant_rows_kernel(const uint8_t* __restrict__ src, OutT* __restrict__ dst, const void* __restrict__ ids,
                                    int ids_i64, int64_t n_ids, TiledGeom G, int qt) {
    const int hdr = header_bytes(qt);
    const int64_t n_elems = n_ids * G.k;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < n_elems;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = idx / G.k, col = idx % G.k;
        const int64_t row = ids_i64 ? ((const long long*)ids)[t] : (long long)((const int*)ids)[t];
        if (row < 0 || row >= G.rows) { dst[idx] = from_float<OutT>(0.0f); continue; }
        const int64_t kb = col / QK;
        const int j = (int)(col % QK);
        uint32_t code; uint16_t dh, mh = 0;
        if (row < G.r64) {
            code = tiled_code(qt, src + tile_code_offset(G, row, kb), (int)(row % 16), j);
            dh = *(const uint16_t*)(src + G.off_s + scale_offset(G, row, kb));
            if (G.has_m) mh = *(const uint16_t*)(src + G.off_m + scale_offset(G, row, kb));
        } else {
            const uint8_t* rec = src + G.off_tail + ((row - G.r64) * G.nb + kb) * G.blk_bytes;
            dh = (uint16_t)rec[0] | ((uint16_t)rec[1] << 8);
            if (G.has_m) mh = (uint16_t)rec[2] | ((uint16_t)rec[3] << 8);
            code = canonical_code(qt, rec + hdr, j);
        }
        float v = __fmul_rn(code_value(qt, code), half_bits_to_float(dh));
        if (G.has_m) v = __fadd_rn(v, half_bits_to_float(mh));
        dst[idx] = from_float<OutT>(v);
    }
}

static int grid_for(int64_t total, int threads) {
    int64_t b = (total + threads - 1) / threads;
    const int64_t cap = 148 * 32;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

int relayout_launch(const void* src, void* dst, int64_t rows, int64_t k, int qt, bool to_device,
                    cudaStream_t st) {
    qt = canon_qtype(qt);
    if (!supported(qt) || k <= 0 || k % QK || rows <= 0) return LB_ERR_ARG;
    if (src == dst) return LB_ERR_ARG;
    TiledGeom G = make_geom(rows, k, qt);
    const int64_t total = rows * G.nb * G.blk_bytes;
    if (to_device) relayout_to_tiled_kernel<<<grid_for(total, 256), 256, 0, st>>>((const uint8_t*)src, (uint8_t*)dst, G, qt);
    else relayout_to_canonical_kernel<<<grid_for(total, 256), 256, 0, st>>>((const uint8_t*)src, (uint8_t*)dst, G, qt);
    note_launch();
    return check_launch();
}

template <bool TILED>
static int dequant_dispatch(const void* src, void* dst, int64_t n, TiledGeom G, int qt, int dt, cudaStream_t st) {
    const int grid = grid_for(n, 256);
    switch (dt) {
        case DT_F32: dequant_kernel<float, TILED><<<grid, 256, 0, st>>>((const uint8_t*)src, (float*)dst, n, G, qt); break;
        case DT_F16: dequant_kernel<__half, TILED><<<grid, 256, 0, st>>>((const uint8_t*)src, (__half*)dst, n, G, qt); break;
        case DT_BF16: dequant_kernel<__nv_bfloat16, TILED><<<grid, 256, 0, st>>>((const uint8_t*)src, (__nv_bfloat16*)dst, n, G, qt); break;
        default: return LB_ERR_ARG;
    }
    note_launch();
    return check_launch();
}

int dequant_canonical_launch(const void* src, void* dst, size_t n, int qt, int out_dtype, cudaStream_t st) {
    qt = canon_qtype(qt);
    if (!supported(qt) || n % QK) return LB_ERR_ARG;
    if (n == 0) return 0;
    TiledGeom G = make_geom(1, (int64_t)n, qt);   // geometry only supplies nb / blk_bytes / has_m
    return dequant_dispatch<false>(src, dst, (int64_t)n, G, qt, out_dtype, st);
}

int dequant_tiled_launch(const void* src, void* dst, int64_t rows, int64_t k, int qt, int out_dtype, cudaStream_t st) {
    qt = canon_qtype(qt);
    if (!supported(qt) || k <= 0 || k % QK || rows <= 0) return LB_ERR_ARG;
    TiledGeom G = make_geom(rows, k, qt);
    return dequant_dispatch<true>(src, dst, rows * k, G, qt, out_dtype, st);
}

int dequant_rows_launch(const void* src, void* dst, const void* ids, int ids_i64, int64_t n_ids, int64_t rows, int64_t k,
                        int qt, int out_dtype, cudaStream_t st) {
    qt = canon_qtype(qt);
    if (!supported(qt) || k <= 0 || k % QK || rows <= 0 || n_ids < 0) return LB_ERR_ARG;
    if (n_ids == 0) return LB_OK;
    const TiledGeom G = make_geom(rows, k, qt);
    const int grid = grid_for(n_ids * k, 256);
    switch (out_dtype) {
        case DT_F32: dequant_rows_kernel<float><<<grid, 256, 0, st>>>((const uint8_t*)src, (float*)dst, ids, ids_i64, n_ids, G, qt); break;
        case DT_F16: dequant_rows_kernel<__half><<<grid, 256, 0, st>>>((const uint8_t*)src, (__half*)dst, ids, ids_i64, n_ids, G, qt); break;
        case DT_BF16: dequant_rows_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const uint8_t*)src, (__nv_bfloat16*)dst, ids, ids_i64, n_ids, G, qt); break;
        default: return LB_ERR_ARG;
    }
    note_launch();
    return check_launch();
}

}  // namespace lowbit
