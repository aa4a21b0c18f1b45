// layout.cu -- canonical <-> tiled relayout and dequantisation kernels (load-time / helper
// paths; HBM-bound byte shuffles, one output byte or element per thread).
//
// Replaces ggml_q_format_convet_cpu2xpu / xpu2cpu (reference low_bit_linear.py:198-253,
// llama_cpp.py:1102-1135) and ggml_dequantize / ggml_dequantize_q4_0
// (low_bit_linear.py:256-279, llama_cpp.py:1068-1099).
#include <cstdlib>
#include "lowbit_common.cuh"
#include "lowbit_internal.h"

namespace lowbit {

static int grid_for(int64_t total, int threads) {
    int64_t b = (total + threads - 1) / threads;
    const int64_t cap = 148 * 32;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

// ---------------------------------------------------------------- canonical -> tiled
// byte `off` of the 16-row tile of the tiled Q region, gathered from the canonical code bytes of the tile's rows
// (rec(r) = the code bytes of row r = 0..15 of the tile, k-block fixed)
template <typename Rec>
__device__ __forceinline__ uint32_t tiled_q_byte(int qt, int off, Rec rec) {
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
    return out;
}

// The general kernel (any shape; also the tail rows and odd block counts of the staged kernels below): one thread per
// output byte of the Q region (gathers its codes from canonical records), then one thread per fp16 of the S / M regions,
// then a straight copy of the tail records.  idx0 = first index served (the staged kernel leaves it the tail records).
__global__ void relayout_to_tiled_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                         TiledGeom G, int qt, int64_t idx0) {
    const int64_t q_bytes = G.off_s;
    const int64_t sm_halves = (G.off_tail - G.off_s) / 2;
    const int64_t tail_bytes = (G.rows - G.r64) * G.nb * G.blk_bytes;
    const int64_t total = q_bytes + sm_halves + tail_bytes;
    const int hdr = header_bytes(qt);
    for (int64_t idx = idx0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        if (idx < q_bytes) {
            const int64_t u = idx / G.super_code_bytes;
            const int rem = (int)(idx % G.super_code_bytes);
            const int tr = rem / G.tile_code_bytes, off = rem % G.tile_code_bytes;
            const int64_t rg = u / G.nb, kb = u % G.nb;
            // canonical record of (row, kb): src + (row*nb + kb)*blk ; r = row inside the 16-row tile
            auto rec = [&](int r) { return src + ((rg * RG_ROWS + tr * 16 + r) * G.nb + kb) * G.blk_bytes + hdr; };
            const uint32_t out = tiled_q_byte(qt, off, rec);
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
                                             TiledGeom G, int qt, int64_t idx0) {
    const int64_t total = G.rows * G.nb * G.blk_bytes;
    const int64_t main_bytes = G.r64 * G.nb * G.blk_bytes;
    const int hdr = header_bytes(qt);
    for (int64_t idx = idx0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
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


// ---------------------------------------------------------------- staged relayout (whole matrices at load time)
// One CTA = one 64-row group x KB k-blocks.  Canonical side: 64 row segments of KB records (KB * blk contiguous bytes each,
// moved as coalesced 4-byte words); tiled side: KB super-tiles = ONE contiguous range of codes plus one of scales (16-byte
// pieces).  Shared memory holds the canonical segments split into code bytes [row][KB * cb + 32] (the 32 staggers the banks
// of the eight rows a quarter-warp reads) and header halves [row][kb][d, m], so that the 16 code bytes a lane of the tiled
// layout owns are ONE aligned 16-byte read, and the nibble shuffle of the 4-bit layouts is two byte permutes per word.
// The general kernels above did one byte per thread with two 64-bit divisions each: 0.3 TB/s on 12288x4096.
constexpr int RL_THREADS = 256;
template <int QT> struct RlFmt {
    static constexpr bool IS8 = (QT == QT_FP8_E4M3 || QT == QT_FP8_E5M2);
    static constexpr bool IS4 = (QT == QT_SYM_INT4 || QT == QT_ASYM_INT4 || QT == QT_NF4);
    static constexpr int CB = IS8 ? 64 : (QT == QT_FP6 ? 48 : 32);
    static constexpr int HDR = (QT == QT_ASYM_INT4) ? 4 : 2;
    static constexpr int HH = HDR / 2;                          // header halves per record
    static constexpr int BLK = CB + HDR;
    static constexpr int TCB = 16 * CB, SCB = 4 * TCB;
    static constexpr int KB_MAX = IS4 ? 16 : 8;
};
// index of (row of the group, kb) in a super-tile's [g][r][h] scale block
__device__ __forceinline__ int rl_scale_slot(int row) { return (row & 7) * 8 + (row >> 4) * 2 + ((row >> 3) & 1); }

__device__ __forceinline__ void rl_cp_async4(uint32_t smem_dst, const void* gsrc) {      // LDGSTS.32: no register staging
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void rl_cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// canonical -> tiled.  The 64 row segments go to shared memory AS THEY ARE with 4-byte asynchronous copies (every word of
// the CTA in flight at once, any 4-byte alignment: k = 11008 gives 5848-byte rows); a lane's 16 code bytes start 2 bytes
// off a word boundary in every other record, so they are read as five words and funnel-shifted.  Row stride = 3 words
// (mod 32): the 16 (row, half) pairs a warp reads per instruction fall into 16 different banks.
template <int QT>
__global__ void __launch_bounds__(RL_THREADS, 6)
relayout_to_tiled_staged_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, TiledGeom G, int KB, int RSW) {
    using F = RlFmt<QT>;
    extern __shared__ __align__(16) uint8_t rl_smem[];
    uint32_t* raw = reinterpret_cast<uint32_t*>(rl_smem);                           // [64][RSW] words
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t rg = blockIdx.x;
    const int kb0 = blockIdx.y * KB;
    const int SW = KB * F::BLK / 4;                                                 // words per row segment
    const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(raw);
    for (int row = warp; row < RG_ROWS; row += RL_THREADS / 32) {
        const uint32_t* seg = reinterpret_cast<const uint32_t*>(src + ((rg * RG_ROWS + row) * G.nb + kb0) * F::BLK);
        for (int w = lane; w < SW; w += 32) rl_cp_async4(raw_s + 4 * (row * RSW + w), seg + w);
    }
    rl_cp_async_wait_all();
    __syncthreads();
    const int64_t u0 = rg * G.nb + kb0;
    uint8_t* q_out = dst + u0 * (int64_t)F::SCB;
    // 16 code bytes at byte `byte` of row `row`'s segment
    auto codes16 = [&](int row, int byte) -> uint4 {
        const uint32_t* p = raw + row * RSW + (byte >> 2);
        const uint32_t fs = (byte & 3) * 8;
        const uint32_t w0 = p[0], w1 = p[1], w2 = p[2], w3 = p[3], w4 = p[4];       // p[4]: at worst the stride's padding
        return make_uint4(__funnelshift_r(w0, w1, fs), __funnelshift_r(w1, w2, fs), __funnelshift_r(w2, w3, fs), __funnelshift_r(w3, w4, fs));
    };
    if (F::IS4) {
        for (int item = tid; item < KB * 128; item += RL_THREADS) {                 // item = (kb, tile, lane): 16 output bytes
            const int kb = item >> 7, tr = (item >> 5) & 3, l = item & 31, g = l >> 2, t = l & 3;
            const int byte = kb * F::BLK + F::HDR + (t & 1) * 16;
            const uint4 a = codes16(tr * 16 + g, byte), b = codes16(tr * 16 + g + 8, byte);
            const int sh = (t >> 1) * 4;                                            // elements 32.. are the high nibbles
            auto word = [&](uint32_t wa, uint32_t wb) -> uint32_t {
                const uint32_t x = (wa >> sh) & 0x0f0f0f0fu, y = (wb >> sh) & 0x0f0f0f0fu;   // codes of rows g / g+8, one per byte
                if (QT == QT_NF4) return __byte_perm(x | (x >> 4), y | (y >> 4), 0x6240);
                return __byte_perm(x, 0, 0x3120) | (__byte_perm(y, 0, 0x3120) << 4);
            };
            uint4 o;
            o.x = word(a.x, b.x); o.y = word(a.y, b.y); o.z = word(a.z, b.z); o.w = word(a.w, b.w);
            __stcs(reinterpret_cast<uint4*>(q_out) + item, o);
        }
    } else if (F::IS8) {
        for (int item = tid; item < KB * 256; item += RL_THREADS) {                 // item = (kb, tile, half, lane)
            const int kb = item >> 8, tr = (item >> 6) & 3, h = (item >> 5) & 1, l = item & 31, g = l >> 2, t = l & 3;
            __stcs(reinterpret_cast<uint4*>(q_out) + item, codes16(tr * 16 + g + 8 * h, kb * F::BLK + F::HDR + 16 * t));
        }
    } else {
        const uint8_t* raw8 = rl_smem;
        for (int item = tid; item < KB * F::SCB / 4; item += RL_THREADS) {          // one output word
            const int o = item * 4, kb = o / F::SCB, rem = o % F::SCB, tr = rem / F::TCB, off = rem % F::TCB;
            auto rec = [&](int r) { return raw8 + (tr * 16 + r) * RSW * 4 + kb * F::BLK + F::HDR; };
            uint32_t w = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) w |= tiled_q_byte(QT, off + i, rec) << (8 * i);
            __stcs(reinterpret_cast<uint32_t*>(q_out) + item, w);
        }
    }
    const unsigned short* raw16 = reinterpret_cast<const unsigned short*>(rl_smem);
    unsigned short* s_out = reinterpret_cast<unsigned short*>(dst + G.off_s) + u0 * 64;
    unsigned short* m_out = reinterpret_cast<unsigned short*>(dst + G.off_m) + u0 * 64;
    for (int item = tid; item < KB * 64; item += RL_THREADS) {
        const int kb = item >> 6, e = item & 63, g = e >> 3, r = (e & 7) >> 1, h = e & 1;
        const int row = r * 16 + g + 8 * h;
        const unsigned short* hp = raw16 + row * RSW * 2 + kb * (F::BLK / 2);
        s_out[item] = hp[0];
        if (F::HH == 2) m_out[item] = hp[1];
    }
}

// tiled -> canonical.  4-bit and 8-bit codes go from global memory through registers (16-byte loads, four items in
// flight per thread) into canonical code bytes in shared memory; fp6 stages the super-tiles first and uses the general
// position functions.  The rows then leave as coalesced 4-byte words put together from header and code halves.
template <int QT>
__global__ void __launch_bounds__(RL_THREADS, RlFmt<QT>::IS4 || RlFmt<QT>::IS8 ? 4 : 3)
relayout_to_canonical_staged_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, TiledGeom G, int KB) {
    using F = RlFmt<QT>;
    constexpr bool DIRECT = F::IS4 || F::IS8;
    extern __shared__ __align__(16) uint8_t rl_smem[];
    const int RS = KB * F::CB + 32;
    uint8_t* s_codes = rl_smem;                                                     // [64][RS] canonical code bytes
    unsigned short* s_sc = reinterpret_cast<unsigned short*>(rl_smem + 64 * RS);    // [HH][KB][64] scale blocks as stored
    uint8_t* s_til = rl_smem + 64 * RS + KB * 128 * F::HH;                          // fp6: KB super-tiles as stored
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t rg = blockIdx.x;
    const int kb0 = blockIdx.y * KB;
    const int64_t u0 = rg * G.nb + kb0;
    const uint8_t* q_in = src + u0 * (int64_t)F::SCB;
    {
        const uint4* s_in = reinterpret_cast<const uint4*>(src + G.off_s + u0 * 128);
        const uint4* m_in = reinterpret_cast<const uint4*>(src + G.off_m + u0 * 128);
        for (int i = tid; i < KB * 8; i += RL_THREADS) {
            reinterpret_cast<uint4*>(s_sc)[i] = __ldcs(s_in + i);
            if (F::HH == 2) reinterpret_cast<uint4*>(s_sc + KB * 64)[i] = __ldcs(m_in + i);
        }
    }
    if (F::IS4) {
        const int n_items = KB * 64;                                                // item = (kb, tile, g, c): lanes (g, c) and (g, c + 2)
        for (int i0 = tid; i0 < n_items; i0 += 4 * RL_THREADS) {
            uint4 lo[4], hi[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int item = i0 + j * RL_THREADS;
                lo[j] = hi[j] = make_uint4(0, 0, 0, 0);
                if (item < n_items) {
                    const int kb = item >> 6, tr = (item >> 4) & 3, g = (item >> 1) & 7, c = item & 1;
                    const uint8_t* tile = q_in + kb * F::SCB + tr * F::TCB;
                    lo[j] = __ldcs(reinterpret_cast<const uint4*>(tile + 16 * (4 * g + c)));
                    hi[j] = __ldcs(reinterpret_cast<const uint4*>(tile + 16 * (4 * g + c + 2)));
                }
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int item = i0 + j * RL_THREADS;
                if (item >= n_items) break;
                const int kb = item >> 6, tr = (item >> 4) & 3, g = (item >> 1) & 7, c = item & 1;
                // per-byte codes of row g (rowsel 0) / g + 8 (rowsel 1) held by one word of the tiled layout
                auto codes = [&](uint32_t w, int rowsel) -> uint32_t {
                    if (QT == QT_NF4) {
                        const uint32_t t1 = __byte_perm(w, 0, rowsel ? 0x4341 : 0x4240);
                        return (t1 | (t1 << 4)) & 0x0f0f0f0fu;
                    }
                    return __byte_perm((w >> (4 * rowsel)) & 0x0f0f0f0fu, 0, 0x3120);
                };
                uint4 ra, rb;
                ra.x = codes(lo[j].x, 0) | (codes(hi[j].x, 0) << 4); rb.x = codes(lo[j].x, 1) | (codes(hi[j].x, 1) << 4);
                ra.y = codes(lo[j].y, 0) | (codes(hi[j].y, 0) << 4); rb.y = codes(lo[j].y, 1) | (codes(hi[j].y, 1) << 4);
                ra.z = codes(lo[j].z, 0) | (codes(hi[j].z, 0) << 4); rb.z = codes(lo[j].z, 1) | (codes(hi[j].z, 1) << 4);
                ra.w = codes(lo[j].w, 0) | (codes(hi[j].w, 0) << 4); rb.w = codes(lo[j].w, 1) | (codes(hi[j].w, 1) << 4);
                uint8_t* pa = s_codes + (tr * 16 + g) * RS + kb * 32 + 16 * c;
                *reinterpret_cast<uint4*>(pa) = ra;
                *reinterpret_cast<uint4*>(pa + 8 * RS) = rb;
            }
        }
    } else if (F::IS8) {
        const int n_items = KB * 256;                                               // item = (kb, tile, half, lane)
        for (int i0 = tid; i0 < n_items; i0 += 4 * RL_THREADS) {
            uint4 v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int item = i0 + j * RL_THREADS;
                v[j] = item < n_items ? __ldcs(reinterpret_cast<const uint4*>(q_in) + item) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int item = i0 + j * RL_THREADS;
                if (item >= n_items) break;
                const int kb = item >> 8, tr = (item >> 6) & 3, h = (item >> 5) & 1, l = item & 31, g = l >> 2, t = l & 3;
                *reinterpret_cast<uint4*>(s_codes + (tr * 16 + g + 8 * h) * RS + kb * 64 + 16 * t) = v[j];
            }
        }
    } else {
        const int n16 = KB * F::SCB / 16;
        for (int i0 = tid; i0 < n16; i0 += 4 * RL_THREADS) {                        // four loads in flight per thread
            uint4 v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) { const int i = i0 + j * RL_THREADS; v[j] = i < n16 ? __ldcs(reinterpret_cast<const uint4*>(q_in) + i) : make_uint4(0, 0, 0, 0); }
#pragma unroll
            for (int j = 0; j < 4; j++) { const int i = i0 + j * RL_THREADS; if (i < n16) reinterpret_cast<uint4*>(s_til)[i] = v[j]; }
        }
        __syncthreads();
        for (int item = tid; item < RG_ROWS * KB * (F::CB / 4); item += RL_THREADS) {   // one canonical code word (fp6: 4 bytes = part of 3-byte groups)
            const int wi = item % (F::CB / 4), kb = (item / (F::CB / 4)) % KB, row = item / (F::CB / 4) / KB;
            const uint8_t* tile = s_til + kb * F::SCB + (row >> 4) * F::TCB;
            uint32_t w = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int cb = 4 * wi + i, grp = cb / 3, byte = cb % 3;
                uint32_t g3 = 0;
                for (int e = 0; e < 4; e++) g3 |= tiled_code(QT, tile, row & 15, 4 * grp + e) << (6 * e);
                w |= ((g3 >> (8 * byte)) & 0xffu) << (8 * i);
            }
            *reinterpret_cast<uint32_t*>(s_codes + row * RS + kb * F::CB + 4 * wi) = w;
        }
    }
    __syncthreads();
    const int SW = KB * F::BLK / 4;
    for (int row = warp; row < RG_ROWS; row += RL_THREADS / 32) {
        uint32_t* seg = reinterpret_cast<uint32_t*>(dst + ((rg * RG_ROWS + row) * G.nb + kb0) * F::BLK);
        const int slot = rl_scale_slot(row);
        for (int w = lane; w < SW; w += 32) {
            uint32_t v = 0;
#pragma unroll
            for (int part = 0; part < 2; part++) {
                const int u = 2 * w + part, kb = u / (F::BLK / 2), r = u % (F::BLK / 2);
                const unsigned short hv = r < F::HH ? s_sc[(r * KB + kb) * 64 + slot]
                                                    : *reinterpret_cast<const unsigned short*>(s_codes + row * RS + kb * F::CB + 2 * (r - F::HH));
                v |= (uint32_t)hv << (16 * part);
            }
            __stcs(seg + w, v);
        }
    }
    (void)DIRECT;
}

// the staged kernels take: whole 64-row groups, an even number of blocks per row, 16-byte aligned buffers
template <int QT>
static bool relayout_staged_launch(const uint8_t* src, uint8_t* dst, const TiledGeom& G, bool to_device, cudaStream_t st) {
    using F = RlFmt<QT>;
    int KB = F::KB_MAX;
    while (KB > 1 && (G.nb % KB)) KB >>= 1;
    if (KB < 2 || G.r64 == 0 || G.nb / KB > 65535 || (((uintptr_t)src | (uintptr_t)dst) & 15)) return false;
    const dim3 grid((unsigned)(G.r64 / RG_ROWS), (unsigned)(G.nb / KB));
    const int RS = KB * F::CB + 32;
    if (to_device) {
        const int SW = KB * F::BLK / 4;
        // row stride = 3 words (mod 32): the banks 3 g + 4 c of the 16 (row, 16-byte half) pairs a warp reads per LDS.32 are
        // all different (stride 8 left rows g and g + 4 on the same bank: 43 % of the wavefronts were conflicts)
        const int RSW = SW + ((3 - SW % 32) + 32) % 32;
        const int smem = 64 * RSW * 4 + 16;
        relayout_to_tiled_staged_kernel<QT><<<grid, RL_THREADS, smem, st>>>(src, dst, G, KB, RSW);
    } else {
        const int smem = 64 * RS + KB * 128 * F::HH + ((F::IS4 || F::IS8) ? 0 : KB * F::SCB);
        // per device, and this is a load / save-time path: set on every call
        if (cudaFuncSetAttribute(relayout_to_canonical_staged_kernel<QT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        relayout_to_canonical_staged_kernel<QT><<<grid, RL_THREADS, smem, st>>>(src, dst, G, KB);
    }
    note_launch();
    return true;
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

// Eight consecutive elements of a block per thread (one 16-byte store for fp16 / bf16, two for fp32), the block's scale read
// once, 32-bit index arithmetic when the block count allows: the whole-matrix dequantise of both layouts.  The kernel above
// (one element per thread, two 64-bit divisions each) stays for destinations that are not 16-byte aligned.
template <int qt, typename OutT, bool TILED>
__global__ void __launch_bounds__(256) dequant8_kernel(const uint8_t* __restrict__ src, OutT* __restrict__ dst, int64_t chunk0,
                                                       int64_t n_chunks, TiledGeom G) {
    const int hdr = header_bytes(qt);
    __shared__ float s_nf4[16];                    // the constant-memory table serialises on divergent codes
    if (qt == QT_NF4) {
        if (threadIdx.x < 16) s_nf4[threadIdx.x] = kNF4Table[threadIdx.x];
        __syncthreads();
    }
    const bool small = G.rows * G.nb < (int64_t)0x7fffffff;
    for (int64_t chunk = chunk0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; chunk < n_chunks;
         chunk += (int64_t)gridDim.x * blockDim.x) {
        const int64_t blk = chunk >> 3;
        const int j0 = (int)(chunk & 7) * 8;
        const uint8_t* rec = nullptr;      // canonical record, or
        const uint8_t* tile = nullptr;     // the 16-row tile of the tiled layout
        int r16 = 0;
        uint16_t dh, mh = 0;
        if (TILED) {
            int64_t row, kb;
            if (small) { const uint32_t b = (uint32_t)blk, q = b / (uint32_t)G.nb; row = q; kb = b - q * (uint32_t)G.nb; }
            else { row = blk / G.nb; kb = blk % G.nb; }
            if (row < G.r64) {
                tile = src + tile_code_offset(G, row, kb);
                r16 = (int)(row & 15);
                const int64_t so = scale_offset(G, row, kb);
                dh = *(const uint16_t*)(src + G.off_s + so);
                if (G.has_m) mh = *(const uint16_t*)(src + G.off_m + so);
            } else {
                rec = src + G.off_tail + (blk - G.r64 * G.nb) * G.blk_bytes;
            }
        } else {
            rec = src + blk * G.blk_bytes;
        }
        if (rec) {
            dh = (uint16_t)rec[0] | ((uint16_t)rec[1] << 8);
            if (G.has_m) mh = (uint16_t)rec[2] | ((uint16_t)rec[3] << 8);
        }
        const float d = half_bits_to_float(dh), m = half_bits_to_float(mh);
        OutT out[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t code = rec ? canonical_code(qt, rec + hdr, j0 + i) : tiled_code(qt, tile, r16, j0 + i);
            float v = __fmul_rn(qt == QT_NF4 ? s_nf4[code] : code_value(qt, code), d);
            if (G.has_m) v = __fadd_rn(v, m);
            out[i] = from_float<OutT>(v);
        }
        uint4* o = reinterpret_cast<uint4*>(dst + chunk * 8);
#pragma unroll
        for (int i = 0; i < (int)(8 * sizeof(OutT) / 16); i++) __stcs(o + i, reinterpret_cast<const uint4*>(out)[i]);
    }
}

// The tiled layout of the 4-bit and 8-bit formats, whole 64-row groups: one thread per (16-row tile, lane) = the 16 code
// bytes that lane owns in the GEMV (4-bit: 16 elements of rows g and g + 8; 8-bit: 16 elements of one row), one 16-byte
// load, the two scales of the lane's rows from one 4-byte load, 32-byte (fp16) / 64-byte (fp32) runs per row that four
// neighbouring lanes complete to 128 / 256 contiguous bytes.
template <int QT, typename OutT>
__global__ void __launch_bounds__(256) dequant_tiled_frag_kernel(const uint8_t* __restrict__ src, OutT* __restrict__ dst, TiledGeom G) {
    using F = RlFmt<QT>;
    __shared__ float s_nf4[16];
    if (QT == QT_NF4) {
        if (threadIdx.x < 16) s_nf4[threadIdx.x] = kNF4Table[threadIdx.x];
        __syncthreads();
    }
    constexpr int PER_TILE = F::IS8 ? 64 : 32;                 // threads per tile
    const int64_t n_items = G.nsuper * 4 * PER_TILE;
    for (int64_t item = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; item < n_items; item += (int64_t)gridDim.x * blockDim.x) {
        const int l = (int)(item & 31), g = l >> 2, t = l & 3;
        const int h8 = F::IS8 ? (int)((item >> 5) & 1) : 0;
        const int64_t tile_idx = item / PER_TILE;              // u * 4 + tr
        const int tr = (int)(tile_idx & 3);
        const int64_t u = tile_idx >> 2;
        int64_t rg, kb;
        if (G.nsuper < (int64_t)0x7fffffff) { const uint32_t q = (uint32_t)u / (uint32_t)G.nb; rg = q; kb = (uint32_t)u - q * (uint32_t)G.nb; }
        else { rg = u / G.nb; kb = u % G.nb; }
        const uint4 c = __ldcs(reinterpret_cast<const uint4*>(src + tile_idx * F::TCB + (F::IS8 ? 512 * h8 : 0) + 16 * l));
        const uint32_t dd = *reinterpret_cast<const uint32_t*>(src + G.off_s + u * 128 + g * 16 + tr * 4);   // d of rows g, g + 8
        uint32_t mm = 0;
        if (F::HH == 2) mm = *reinterpret_cast<const uint32_t*>(src + G.off_m + u * 128 + g * 16 + tr * 4);
        const uint32_t cw[4] = {c.x, c.y, c.z, c.w};
        constexpr int NR = F::IS8 ? 1 : 2;                     // rows this thread writes
#pragma unroll
        for (int rr = 0; rr < NR; rr++) {
            const int hsel = F::IS8 ? h8 : rr;
            const float d = half_bits_to_float((uint16_t)(dd >> (16 * hsel)));
            const float m = half_bits_to_float((uint16_t)(mm >> (16 * hsel)));
            OutT out[16];
#pragma unroll
            for (int e = 0; e < 16; e++) {
                uint32_t code;
                if (F::IS8) code = (cw[e >> 2] >> (8 * (e & 3))) & 0xffu;
                else if (QT == QT_NF4) code = (cw[e >> 2] >> (8 * (rr + 2 * ((e & 3) >> 1)) + 4 * (e & 1))) & 0x0fu;
                else { const int i = e & 3, p = ((i >> 1) << 1) + rr + ((i & 1) << 2); code = (cw[e >> 2] >> (4 * p)) & 0x0fu; }
                float v = __fmul_rn(QT == QT_NF4 ? s_nf4[code] : code_value(QT, code), d);
                if (F::HH == 2) v = __fadd_rn(v, m);
                out[e] = from_float<OutT>(v);
            }
            const int64_t row = rg * RG_ROWS + tr * 16 + g + 8 * hsel;
            uint4* o = reinterpret_cast<uint4*>(dst + row * G.k + kb * QK + 16 * t);
#pragma unroll
            for (int i = 0; i < (int)(16 * sizeof(OutT) / 16); i++) __stcs(o + i, reinterpret_cast<const uint4*>(out)[i]);
        }
    }
}

// rows of a TILED-layout weight picked by an index vector (LowBitEmbedding, reference embedding.py:214-224:
// xe_linear.dequantize_rows): out[t][:] = dequant(W[ids[t]][:]); an index outside [0, rows) gives a zero row.
template <typename OutT>
__global__ void dequant_rows_kernel(const uint8_t* __restrict__ src, OutT* __restrict__ dst, const void* __restrict__ ids,
                                    int ids_i64, int64_t n_ids, TiledGeom G, int qt, int vec) {
    // eight consecutive elements of one block per thread (16-byte stores when `vec`: the destination is 16-byte aligned)
    const int hdr = header_bytes(qt);
    const int64_t cpr = G.k / 8;                                   // chunks per row
    const int64_t n_chunks = n_ids * cpr;
    for (int64_t chunk = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; chunk < n_chunks;
         chunk += (int64_t)gridDim.x * blockDim.x) {
        int64_t t, c;
        if (n_chunks < (int64_t)0x7fffffff) { const uint32_t q = (uint32_t)chunk / (uint32_t)cpr; t = q; c = (uint32_t)chunk - q * (uint32_t)cpr; }
        else { t = chunk / cpr; c = chunk % cpr; }
        const int64_t row = ids_i64 ? ((const long long*)ids)[t] : (long long)((const int*)ids)[t];
        OutT out[8];
        if (row < 0 || row >= G.rows) {
#pragma unroll
            for (int i = 0; i < 8; i++) out[i] = from_float<OutT>(0.0f);
        } else {
            const int64_t kb = c >> 3;
            const int j0 = (int)(c & 7) * 8;
            const uint8_t* rec = nullptr;
            const uint8_t* tile = nullptr;
            uint16_t dh, mh = 0;
            if (row < G.r64) {
                tile = src + tile_code_offset(G, row, kb);
                const int64_t so = scale_offset(G, row, kb);
                dh = *(const uint16_t*)(src + G.off_s + so);
                if (G.has_m) mh = *(const uint16_t*)(src + G.off_m + so);
            } else {
                rec = src + G.off_tail + ((row - G.r64) * G.nb + kb) * G.blk_bytes;
                dh = (uint16_t)rec[0] | ((uint16_t)rec[1] << 8);
                if (G.has_m) mh = (uint16_t)rec[2] | ((uint16_t)rec[3] << 8);
            }
            const float d = half_bits_to_float(dh), m = half_bits_to_float(mh);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t code = rec ? canonical_code(qt, rec + hdr, j0 + i) : tiled_code(qt, tile, (int)(row & 15), j0 + i);
                float v = __fmul_rn(code_value(qt, code), d);
                if (G.has_m) v = __fadd_rn(v, m);
                out[i] = from_float<OutT>(v);
            }
        }
        OutT* o = dst + chunk * 8;
        if (vec) {
#pragma unroll
            for (int i = 0; i < (int)(8 * sizeof(OutT) / 16); i++) reinterpret_cast<uint4*>(o)[i] = reinterpret_cast<const uint4*>(out)[i];
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) o[i] = out[i];
        }
    }
}

int relayout_launch(const void* src, void* dst, int64_t rows, int64_t k, int qt, bool to_device,
                    cudaStream_t st) {
    qt = canon_qtype(qt);
    if (!supported(qt) || k <= 0 || k % QK || rows <= 0) return LB_ERR_ARG;
    if (src == dst) return LB_ERR_ARG;
    TiledGeom G = make_geom(rows, k, qt);
    const int64_t total = rows * G.nb * G.blk_bytes;
    const uint8_t* s8 = (const uint8_t*)src;
    uint8_t* d8 = (uint8_t*)dst;
    const char* env = getenv("LOWBIT_RELAYOUT_GENERAL");                 // development / tests; read per call (load-time path)
    const bool general_only = env && atoi(env) != 0;
    bool staged = false;
    if (!general_only) {
        switch (qt) {
            case QT_SYM_INT4: staged = relayout_staged_launch<QT_SYM_INT4>(s8, d8, G, to_device, st); break;
            case QT_ASYM_INT4: staged = relayout_staged_launch<QT_ASYM_INT4>(s8, d8, G, to_device, st); break;
            case QT_NF4: staged = relayout_staged_launch<QT_NF4>(s8, d8, G, to_device, st); break;
            case QT_FP8_E4M3: staged = relayout_staged_launch<QT_FP8_E4M3>(s8, d8, G, to_device, st); break;
            case QT_FP8_E5M2: staged = relayout_staged_launch<QT_FP8_E5M2>(s8, d8, G, to_device, st); break;
            case QT_FP6: staged = relayout_staged_launch<QT_FP6>(s8, d8, G, to_device, st); break;
            default: break;
        }
    }
    // the general kernel: everything, or only the tail rows (rows % 64) the staged kernel leaves
    const int64_t tail_bytes = (G.rows - G.r64) * G.nb * G.blk_bytes;
    if (!staged || tail_bytes > 0) {
        const int64_t idx0 = !staged ? 0 : (to_device ? G.off_s + (G.off_tail - G.off_s) / 2 : G.r64 * G.nb * G.blk_bytes);
        const int64_t work = !staged ? total : tail_bytes;
        if (to_device) relayout_to_tiled_kernel<<<grid_for(work, 256), 256, 0, st>>>(s8, d8, G, qt, idx0);
        else relayout_to_canonical_kernel<<<grid_for(work, 256), 256, 0, st>>>(s8, d8, G, qt, idx0);
        note_launch();
    }
    return check_launch();
}

template <bool TILED>
static int dequant_dispatch(const void* src, void* dst, int64_t n, TiledGeom G, int qt, int dt, cudaStream_t st) {
    if (!((uintptr_t)dst & 15)) {                          // n is a multiple of 64
        if (dt != DT_F32 && dt != DT_F16 && dt != DT_BF16) return LB_ERR_ARG;
        // the tiled layout of 4-bit / 8-bit codes: whole row groups by the fragment kernel, tail rows by the chunk kernel
        int64_t chunk0 = 0;
        if (TILED && qt != QT_FP6 && G.nsuper > 0 && !((uintptr_t)src & 15)) {
            const int per_tile = (qt == QT_FP8_E4M3 || qt == QT_FP8_E5M2) ? 64 : 32;
            const int grid = grid_for(G.nsuper * 4 * per_tile, 256);
#define LOWBIT_DEQF(QT)                                                                                                          \
    case QT:                                                                                                                     \
        if (dt == DT_F32) dequant_tiled_frag_kernel<QT, float><<<grid, 256, 0, st>>>((const uint8_t*)src, (float*)dst, G);         \
        else if (dt == DT_F16) dequant_tiled_frag_kernel<QT, __half><<<grid, 256, 0, st>>>((const uint8_t*)src, (__half*)dst, G);  \
        else dequant_tiled_frag_kernel<QT, __nv_bfloat16><<<grid, 256, 0, st>>>((const uint8_t*)src, (__nv_bfloat16*)dst, G);      \
        break;
            switch (qt) {
                LOWBIT_DEQF(QT_SYM_INT4) LOWBIT_DEQF(QT_ASYM_INT4) LOWBIT_DEQF(QT_NF4) LOWBIT_DEQF(QT_FP8_E4M3) LOWBIT_DEQF(QT_FP8_E5M2)
                default: return LB_ERR_ARG;
            }
#undef LOWBIT_DEQF
            note_launch();
            chunk0 = G.r64 * G.k / 8;
            if (chunk0 >= n / 8) return check_launch();
        }
        const int grid = grid_for(n / 8 - chunk0, 256);
#define LOWBIT_DEQ8(QT)                                                                                                          \
    case QT:                                                                                                                     \
        if (dt == DT_F32) dequant8_kernel<QT, float, TILED><<<grid, 256, 0, st>>>((const uint8_t*)src, (float*)dst, chunk0, n / 8, G);    \
        else if (dt == DT_F16) dequant8_kernel<QT, __half, TILED><<<grid, 256, 0, st>>>((const uint8_t*)src, (__half*)dst, chunk0, n / 8, G); \
        else dequant8_kernel<QT, __nv_bfloat16, TILED><<<grid, 256, 0, st>>>((const uint8_t*)src, (__nv_bfloat16*)dst, chunk0, n / 8, G);  \
        break;
        switch (qt) {
            LOWBIT_DEQ8(QT_SYM_INT4) LOWBIT_DEQ8(QT_ASYM_INT4) LOWBIT_DEQ8(QT_NF4)
            LOWBIT_DEQ8(QT_FP8_E4M3) LOWBIT_DEQ8(QT_FP8_E5M2) LOWBIT_DEQ8(QT_FP6)
            default: return LB_ERR_ARG;
        }
#undef LOWBIT_DEQ8
        note_launch();
        return check_launch();
    }
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
    const int grid = grid_for(n_ids * (k / 8), 256);
    const int vec = ((uintptr_t)dst & 15) ? 0 : 1;
    switch (out_dtype) {
        case DT_F32: dequant_rows_kernel<float><<<grid, 256, 0, st>>>((const uint8_t*)src, (float*)dst, ids, ids_i64, n_ids, G, qt, vec); break;
        case DT_F16: dequant_rows_kernel<__half><<<grid, 256, 0, st>>>((const uint8_t*)src, (__half*)dst, ids, ids_i64, n_ids, G, qt, vec); break;
        case DT_BF16: dequant_rows_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const uint8_t*)src, (__nv_bfloat16*)dst, ids, ids_i64, n_ids, G, qt, vec); break;
        default: return LB_ERR_ARG;
    }
    note_launch();
    return check_launch();
}

}  // namespace lowbit
