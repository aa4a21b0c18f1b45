// lowbit_common.cuh -- formats, the tiled device layout and small device helpers shared by
// the sm_100a kernels of libb200lowbit.so.
//
// Canonical (CPU / on-disk / C-ABI) layout: row-major [rows][k/64] block records
//   d (fp16) [, m (fp16)], codes          (reference: transformers/convert.py:448-453)
//   4-bit codes: byte j = code[j] | code[j+32] << 4   (convert.py:421-426, gguf/gguf.py:281)
// qtype numbers: ggml/quantize.py:28-60.
//
// Device ("tiled") layout, same byte size as canonical, for a [rows][k] weight:
//   R64 = rows & ~63 rows live in 64-row groups, nb = k/64 blocks per row.  Super-tile
//   u = rg*nb + kb is the 64 rows of group rg x the 64 elements of k-block kb = four 16-row
//   tiles r = 0..3 (rows 64rg + 16r + 0..15).
//   region Q : u * 4*TB + r*TB   codes of tile r (TB = 512 / 1024 / 768 bytes for 4/8/6-bit),
//              arranged so that lane l = 4*g + t of a warp owns rows (g, g+8) of the tile x
//              elements 16t..16t+15 -- exactly the mma.m16n8k16 A-fragment rows of that lane
//              -- in 16-byte pieces at 16*l.
//   region S : after Q, u*128 bytes + fp16 d as [g][r][h]  (row 64rg + 16r + g + 8h), so lane
//              (g,*) gets the scales of its 8 rows of the super-tile with one 16-byte load
//   region M : (asym_int4 only) after S, same shape, fp16 m
//   tail     : rows R64..rows-1 stay in canonical records after the planes.
//   A run of super-tiles is ONE contiguous byte range of codes (+ one of scales): a warp streams
//   its share of the matrix with plain 1-D TMA bulk copies.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <stdint.h>

namespace lowbit {

enum QType : int {
    QT_SYM_INT4 = 2, QT_ASYM_INT4 = 3, QT_NF4 = 10, QT_FP8_E4M3 = 15, QT_FP8_E5M2 = 19,
    QT_FP6 = 29, QT_WOQ_INT4 = 34
};
enum DType : int { DT_F32 = 0, DT_F16 = 1, DT_BF16 = 2 };

constexpr int QK = 64;  // elements per block for every in-scope qtype

__host__ __device__ inline int canon_qtype(int qt) { return qt == QT_WOQ_INT4 ? QT_SYM_INT4 : qt; }

// bytes of codes per block
__host__ __device__ inline int code_bytes(int qt) {
    switch (canon_qtype(qt)) {
        case QT_SYM_INT4: case QT_ASYM_INT4: case QT_NF4: return 32;
        case QT_FP8_E4M3: case QT_FP8_E5M2: return 64;
        case QT_FP6: return 48;
        default: return 0;
    }
}
__host__ __device__ inline int header_bytes(int qt) { return canon_qtype(qt) == QT_ASYM_INT4 ? 4 : 2; }
__host__ __device__ inline int type_size(int qt) {
    int c = code_bytes(qt);
    return c ? c + header_bytes(qt) : 0;
}
__host__ __device__ inline bool supported(int qt) { return code_bytes(qt) != 0; }

// ---- tiled layout geometry -------------------------------------------------------------
constexpr int RG_TILES = 4;     // 16-row tiles per row group
constexpr int RG_ROWS = 64;     // rows per row group

struct TiledGeom {
    int64_t rows, k, nb, r64, nsuper;   // nsuper = (r64/64) * nb super-tiles
    int tile_code_bytes;                // 16 rows * code_bytes
    int super_code_bytes;               // 4 tiles
    int64_t off_s, off_m, off_tail;     // byte offsets of the regions
    int blk_bytes;
    bool has_m;
};
__host__ __device__ inline TiledGeom make_geom(int64_t rows, int64_t k, int qt) {
    TiledGeom g;
    g.rows = rows; g.k = k; g.nb = k / QK; g.r64 = rows & ~int64_t(RG_ROWS - 1);
    g.nsuper = (g.r64 / RG_ROWS) * g.nb;
    g.tile_code_bytes = 16 * code_bytes(qt);
    g.super_code_bytes = RG_TILES * g.tile_code_bytes;
    g.has_m = canon_qtype(qt) == QT_ASYM_INT4;
    g.blk_bytes = type_size(qt);
    g.off_s = g.nsuper * g.super_code_bytes;
    g.off_m = g.off_s + g.nsuper * 128;
    g.off_tail = g.off_m + (g.has_m ? g.nsuper * 128 : 0);
    return g;
}
// byte offset of the codes of the 16-row tile holding (row, kb), row < r64
__host__ __device__ inline int64_t tile_code_offset(const TiledGeom& G, int64_t row, int64_t kb) {
    const int64_t u = (row / RG_ROWS) * G.nb + kb;
    return u * G.super_code_bytes + ((row % RG_ROWS) / 16) * G.tile_code_bytes;
}
// byte offset (from off_s, or from off_m for the minimum) of the fp16 scale of (row, kb)
__host__ __device__ inline int64_t scale_offset(const TiledGeom& G, int64_t row, int64_t kb) {
    const int64_t u = (row / RG_ROWS) * G.nb + kb;
    const int rr = (int)(row % RG_ROWS), r = rr >> 4, g = rr & 7, h = (rr >> 3) & 1;
    return u * 128 + (g * 8 + r * 2 + h) * 2;
}

// ---- scalar decode helpers (used by dequant / generic kernels; not the GEMV hot loop) ----
static __device__ __constant__ float kNF4Table[16] = {
    -1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f,
    -0.28444138169288635f, -0.18477343022823334f, -0.09105003625154495f, 0.0f,
    0.07958029955625534f, 0.16093020141124725f, 0.24611230194568634f, 0.33791524171829224f,
    0.44070982933044434f, 0.5626170039176941f, 0.7229568362236023f, 1.0f};

__device__ __forceinline__ float decode_e4m3(uint32_t c) {
    __half_raw hr = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)c, __NV_E4M3);
    return __half2float(__half(hr));
}
__device__ __forceinline__ float decode_e5m2(uint32_t c) {
    __half_raw hr; hr.x = (unsigned short)(c << 8);   // e5m2 is the top byte of fp16
    return __half2float(__half(hr));
}
__device__ __forceinline__ float decode_e3m2(uint32_t c) {
    // 1-3-2, bias 3, no inf/nan.  value = (-1)^s * 2^(e-3) * (1 + m/4), subnormal m/4 * 2^-2
    uint32_t s = (c >> 5) & 1u, e = (c >> 2) & 7u, m = c & 3u;
    float v = e == 0 ? (float)m * 0.0625f : __int2float_rn((int)(m | 4u)) * __int_as_float((int)(e + 122u) << 23);
    return s ? -v : v;
}

// value of code `c` of element-in-block (before scale/min), per qtype
__device__ __forceinline__ float code_value(int qt, uint32_t c) {
    switch (qt) {
        case QT_SYM_INT4: return (float)((int)c - 8);
        case QT_ASYM_INT4: return (float)c;
        case QT_NF4: return kNF4Table[c];
        case QT_FP8_E4M3: return decode_e4m3(c);
        case QT_FP8_E5M2: return decode_e5m2(c);
        case QT_FP6: return decode_e3m2(c);
        default: return 0.0f;
    }
}

// code of element j (0..63) of a canonical block's code bytes
__device__ __forceinline__ uint32_t canonical_code(int qt, const uint8_t* codes, int j) {
    switch (qt) {
        case QT_SYM_INT4: case QT_ASYM_INT4: case QT_NF4: {
            uint32_t b = codes[j & 31];
            return (j < 32) ? (b & 0x0f) : (b >> 4);
        }
        case QT_FP8_E4M3: case QT_FP8_E5M2: return codes[j];
        case QT_FP6: {
            const uint8_t* p = codes + 3 * (j >> 2);
            uint32_t w = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
            return (w >> (6 * (j & 3))) & 0x3f;
        }
        default: return 0;
    }
}

__device__ __forceinline__ float half_bits_to_float(uint16_t h) {
    __half_raw hr; hr.x = h; return __half2float(__half(hr));
}

// store helpers
template <typename T> __device__ __forceinline__ T from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_float<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <typename T> __device__ __forceinline__ float to_float(T v);
template <> __device__ __forceinline__ float to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_float<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// ---- position of element (row-in-tile r 0..15, element-in-block j 0..63) in the tiled Q
// region of a 4-bit type: returns byte offset within the 512-byte tile and the nibble shift.
// lane l = 4*g + t owns rows (g, g+8), elements 16t..16t+15; word s (0..3) of the lane's 16
// bytes holds elements 4s..4s+3 of both rows; nibble slot p of the word:
//   p=0: (g, 4s)  p=4: (g, 4s+1)  p=1: (g+8, 4s)  p=5: (g+8, 4s+1)
//   p=2: (g, 4s+2) p=6: (g, 4s+3) p=3: (g+8, 4s+2) p=7: (g+8, 4s+3)
// so that (w & 0x000f000f), (w & 0x00f000f0), ((w>>8) & 0x000f000f), ((w>>8) & 0x00f000f0)
// are directly the fp16 magic-number fragments a0, a1, a2, a3 of mma.m16n8k16.
__host__ __device__ inline void tiled4_pos(int r, int j, int& byte_off, int& shift) {
    const int g = r & 7, h = r >> 3, t = j >> 4, e = j & 15, s = e >> 2, i = e & 3;
    const int p = ((i >> 1) << 1) + h + ((i & 1) << 2);
    byte_off = 16 * (4 * g + t) + 4 * s + (p >> 1);
    shift = (p & 1) * 4;
}
// nf4 (table codes): same lane ownership and word s = elements 4s..4s+3 of both rows, but byte b of
// the word is fragment register a_b of mma.m16n8k16 as a PAIR of codes (low nibble = first element):
//   byte 0: (g, 4s), (g, 4s+1)   byte 1: (g+8, 4s), (g+8, 4s+1)
//   byte 2: (g, 4s+2), (g, 4s+3) byte 3: (g+8, 4s+2), (g+8, 4s+3)
// so that one 256-entry table lookup per byte yields the fp16x2 / bf16x2 register.
__host__ __device__ inline void tilednf4_pos(int r, int j, int& byte_off, int& shift) {
    const int g = r & 7, h = r >> 3, t = j >> 4, e = j & 15, s = e >> 2, i = e & 3;
    byte_off = 16 * (4 * g + t) + 4 * s + h + 2 * (i >> 1);
    shift = (i & 1) * 4;
}
// 8-bit codes: tile = 1024 bytes; half h of the tile (512 bytes) holds rows g+8h; lane's 16
// bytes hold elements 16t..16t+15 of that row in natural order.
__host__ __device__ inline int tiled8_pos(int r, int j) {
    const int g = r & 7, h = r >> 3, t = j >> 4, e = j & 15;
    return 512 * h + 16 * (4 * g + t) + e;
}
// fp6 (6-bit codes split 4+2): tile = 768 bytes = 512 (low-nibble plane) + 256 (high-2-bit
// plane).  Lane l = 4g+t, element e = 0..15 of its 16, row half h:
//   low plane : 16 bytes at 16*l = words [h=0,u=0][h=0,u=1][h=1,u=0][h=1,u=1]; word u holds
//               elements 8u..8u+7; nibble slot p = 2*(e&3) + ((e>>2)&1), so
//               (w & 0x0f0f0f0f) / ((w>>4) & 0x0f0f0f0f) are the bytes of k-steps 2u / 2u+1.
//   high plane: 8 bytes at 512 + 8*l = words [h=0][h=1]; the 2-bit field of element e sits at
//               bit 8*(e&3) + 2*(e>>2), so ((w >> 2s) & 0x03030303) << 4 lines up with step s.
__host__ __device__ inline void tiled6_lo_pos(int r, int j, int& byte_off, int& shift) {
    const int g = r & 7, h = r >> 3, t = j >> 4, e = j & 15, u = e >> 3;
    const int p = 2 * (e & 3) + ((e >> 2) & 1);
    byte_off = 16 * (4 * g + t) + 8 * h + 4 * u + (p >> 1);
    shift = (p & 1) * 4;
}
__host__ __device__ inline void tiled6_hi_pos(int r, int j, int& byte_off, int& shift) {
    const int g = r & 7, h = r >> 3, t = j >> 4, e = j & 15;
    const int bit = 8 * (e & 3) + 2 * (e >> 2);
    byte_off = 512 + 8 * (4 * g + t) + 4 * h + (bit >> 3);
    shift = bit & 7;
}

// code of (row-in-tile r, element j) read from one tile of the tiled Q region
__device__ __forceinline__ uint32_t tiled_code(int qt, const uint8_t* tile, int r, int j) {
    int off, sh;
    switch (qt) {
        case QT_SYM_INT4: case QT_ASYM_INT4:
            tiled4_pos(r, j, off, sh);
            return (tile[off] >> sh) & 0x0f;
        case QT_NF4:
            tilednf4_pos(r, j, off, sh);
            return (tile[off] >> sh) & 0x0f;
        case QT_FP8_E4M3: case QT_FP8_E5M2:
            return tile[tiled8_pos(r, j)];
        case QT_FP6: {
            tiled6_lo_pos(r, j, off, sh);
            uint32_t lo = (tile[off] >> sh) & 0x0f;
            tiled6_hi_pos(r, j, off, sh);
            return lo | (((tile[off] >> sh) & 0x03u) << 4);
        }
        default: return 0;
    }
}

}  // namespace lowbit
