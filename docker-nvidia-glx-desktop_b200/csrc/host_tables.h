// host_tables.h — the numbers both host files (api.cu, noalloc.cu) turn negotiated caps into: the
// 8.8 colour matrix in dp4a form and videoscale's 2-tap positions. One copy,
// derived from the spec constants of include/b200vc_spec.h; the tap functions also compile for the
// device (the no-allocation entry point evaluates them in its kernel instead of reading tables).
#pragma once

#include "../../include/b200vc.h"
#include "../../include/b200vc_spec.h"
#include "kernels.cuh"

namespace b200vc {

#ifdef __CUDACC__
#define B200VC_HD __host__ __device__ inline
#else
#define B200VC_HD inline
#endif

// SURVEY A.2 table; rows Y,U,V of (cR, cG, cB, off). The oracle derives the same numbers from
// Kr/Kb in double precision; tests compare the two.
inline const int (*matrix_rows(int matrix, int range))[4]
{
    static const int k[3][2][3][4] = {
        { { { 66, 129, 25, 4096 }, { -38, -74, 112, 32768 }, { 112, -94, -18, 32768 } },      // 601 limited
          { { 77, 150, 29, 0 }, { -43, -85, 128, 32768 }, { 128, -107, -21, 32768 } } },      // 601 full
        { { { 47, 157, 16, 4096 }, { -26, -87, 112, 32768 }, { 112, -102, -10, 32768 } },     // 709 limited
          { { 54, 183, 18, 0 }, { -29, -99, 128, 32768 }, { 128, -116, -12, 32768 } } },      // 709 full
        { { { 58, 149, 13, 4096 }, { -31, -81, 112, 32768 }, { 112, -103, -9, 32768 } },      // 2020 limited
          { { 67, 174, 15, 0 }, { -36, -92, 128, 32768 }, { 128, -118, -10, 32768 } } },      // 2020 full
    };
    return k[matrix][range];
}

// byte position of B, G, R in one pixel, per B200VC_IN_* (unpack_BGRA / _RGBA / _ARGB / _ABGR). The 16-bit
// formats reach the matrix unpacked by the kernel into one word per pixel, low field in byte 0, high field in
// byte 2: RGB16 then reads like BGRx, BGR16 like RGBx.
inline const int *byte_pos(int in_format)
{
    static const int k[6][3] = { { 0, 1, 2 }, { 2, 1, 0 }, { 3, 2, 1 }, { 1, 2, 3 }, { 0, 1, 2 }, { 2, 1, 0 } };
    return k[in_format];
}

inline int in_bytes_per_pixel(int in_format) { return in_format >= B200VC_IN_RGB16 ? 2 : 4; }
inline int out_planes(int fmt) { return fmt == B200VC_FMT_NV12 ? 2 : 3; }
inline int chroma_vsub(int fmt) { return (fmt == B200VC_FMT_I420 || fmt == B200VC_FMT_NV12) ? 2 : 1; }   // lines per chroma line
inline int chroma_hsub(int fmt) { return fmt == B200VC_FMT_Y444 ? 1 : 2; }                                 // pixels per chroma sample
// samples per line and lines of one chroma plane of a w x h frame (odd sizes round up, video-info.c)
inline void chroma_dims(int fmt, int w, int h, int &cw, int &ch)
{
    cw = chroma_hsub(fmt) == 2 ? (w + 1) / 2 : w;
    ch = chroma_vsub(fmt) == 2 ? (h + 1) / 2 : h;
}

inline uint32_t pack_coef_at(const int pos[3], int cb, int cg, int cr)
{
    return ((uint32_t)(cb & 0xff) << (8 * pos[0])) | ((uint32_t)(cg & 0xff) << (8 * pos[1])) |
           ((uint32_t)(cr & 0xff) << (8 * pos[2]));
}

// The matrix as the kernels consume it (Matrix, kernels.cuh). B200VC_SPEC_MATRIX_ROUND adds the
// rounding term to every offset here and nowhere else.
inline Matrix make_matrix(int matrix, int range, int in_format)
{
    const int (*m)[4] = matrix_rows(matrix, range);
    const int *pos = byte_pos(in_format);
    const int rnd = B200VC_SPEC_MATRIX_ROUND ? 128 : 0;
    Matrix r;
    // (damage: the three colour bytes of a 4-byte pixel; every bit of a 16-bit one, compared before unpacking)
    r.rgb_mask = in_bytes_per_pixel(in_format) == 2 ? 0xffffffffu : pack_coef_at(pos, 0xff, 0xff, 0xff);
    r.cy = pack_coef_at(pos, m[0][2], m[0][1], m[0][0]);
    r.oy = (uint32_t)(m[0][3] + rnd);
    bool neg = false;                      // a +128 coefficient does not fit a signed byte: evaluate negated
    for (int row = 1; row < 3; row++)
        for (int c = 0; c < 3; c++)
            if (m[row][c] > 127) neg = true;
    r.neg = neg ? 1 : 0;
    const int s = neg ? -1 : 1;
    r.cu = pack_coef_at(pos, s * m[1][2], s * m[1][1], s * m[1][0]);
    r.cv = pack_coef_at(pos, s * m[2][2], s * m[2][1], s * m[2][0]);
    r.ou = (uint32_t)(neg ? 65535 - (m[1][3] + rnd) : m[1][3] + rnd);
    r.ov = (uint32_t)(neg ? 65535 - (m[2][3] + rnd) : m[2][3] + rnd);
    return r;
}

// ---- videoscale 2-tap positions, integer arithmetic only --------------------------------

// horizontal: edge-aligned 16.16 step
B200VC_HD long long scale_h_inc(int in_n, int out_n)
{
    if (out_n <= 1) return 0;
    const long long inc = (((long long)(in_n - 1)) << 16) / (out_n - 1) - 1;
    return inc < 0 ? 0 : inc;
}

// horizontal tap of output x: first index (unclamped), 8-bit fraction
B200VC_HD void scale_h_tap(long long inc, int x, int &i, int &f)
{
    const long long p = (long long)x * inc;
    i = (int)(p >> 16);
    f = (int)((p >> 8) & 0xff);
}

// vertical tap of output y: centre-aligned position y' = (y + 0.5) * in / out - 0.5; first index
// (unclamped, may be -1) and the tap p1 in 0 .. 2^B200VC_SPEC_VTAP_BITS, rounded or floored per the spec
B200VC_HD void scale_v_tap(int y, int in_n, int out_n, long long &i, int &p1)
{
    const long long den = 2LL * out_n, num = (2LL * y + 1) * in_n - out_n;
    i = num / den;
    long long rem = num - i * den;
    if (rem < 0) { i -= 1; rem += den; }
    p1 = (int)((2LL * B200VC_SPEC_VTAP_ONE * rem + (B200VC_SPEC_VTAP_ROUND ? den : 0)) / (2 * den));
}

// method=nearest: the source index of output sample x (B200VC_SPEC_NEAREST_CENTRE)
B200VC_HD int scale_nearest_index(int x, int in_n, int out_n)
{
    if (B200VC_SPEC_NEAREST_CENTRE) return (int)(((2LL * x + 1) * in_n) / (2LL * out_n));
    return (int)(((long long)x * scale_h_inc(in_n, out_n)) >> 16);
}

B200VC_HD int clamp_index(long long i, int n) { return (int)(i < 0 ? 0 : (i > n - 1 ? n - 1 : i)); }

// The vertical tap as dp4a weights against [a,a,b,b] (the two lines' samples, each twice):
// wa1+wa2 = (ONE-p1) * 256/ONE, wb1+wb2 = p1 * 256/ONE, every byte <= 255;
// ([a,a,b,b] . w + 128) >> 8 == (a*(ONE-p1) + b*p1 + ONE/2) >> BITS == a + (((b-a)*p1 + ONE/2) >> BITS)
inline uint32_t v_weight_word(int p1)
{
    const int sh = 8 - B200VC_SPEC_VTAP_BITS;
    const int a = (B200VC_SPEC_VTAP_ONE - p1) << sh, b = p1 << sh;
    const int a1 = a < 255 ? a : 255, b1 = b < 255 ? b : 255;
    return (uint32_t)a1 | ((uint32_t)(a - a1) << 8) | ((uint32_t)b1 << 16) | ((uint32_t)(b - b1) << 24);
}

}  // namespace b200vc
