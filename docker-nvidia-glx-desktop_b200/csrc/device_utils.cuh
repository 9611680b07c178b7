// device_utils.cuh — per-pixel arithmetic shared by the kernels (sm_100a).
//
// Bit-exact restatement, for the GPU, of the integer steps in SURVEY.md Appendix A:
//   A.2  (c . [R,G,B] + off) >> 8 with no rounding term     -> dp4a, result taken from byte 1
//   A.4  chroma 4:4:4 -> 4:2:0, horizontal then vertical, two roundings -> 2x16-bit SIMD in a word
#pragma once

#include <stdint.h>
#include "kernels.cuh"
#include "../../include/b200vc_spec.h"

namespace b200vc {

__device__ __forceinline__ uint32_t dp4a_uu(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// a: four unsigned bytes (the pixel), b: four signed bytes (coefficients)
__device__ __forceinline__ uint32_t dp4a_us(uint32_t a, uint32_t b, uint32_t c)
{
    int32_t d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"((int32_t)c));
    return (uint32_t)d;
}

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    return __byte_perm(a, b, sel);
}

// streaming 16-byte load: read once, do not keep in L1
__device__ __forceinline__ uint4 ldg_stream_v4(const void *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ uint2 ldg_stream_v2(const void *p)
{
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}

__device__ __forceinline__ uint32_t ldg_stream_u32(const void *p)
{
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

// The same loads for a frame that is still being written while the kernel runs (row gate): ordinary
// coherent loads ordered behind the gate's acquire, where .nc would assume memory that is read-only
// for the kernel's lifetime.
__device__ __forceinline__ uint4 ldg_live_v4(const void *p)
{
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}

__device__ __forceinline__ uint32_t ldg_live_u32(const void *p)
{
    uint32_t r;
    asm volatile("ld.global.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}

__device__ __forceinline__ uint32_t ldg_live_u32_if(const void *p, bool pred)
{
    uint32_t r;
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\tmov.u32 %0, 0;\n\t"
                 "@q ld.global.L1::no_allocate.u32 %0, [%1];\n\t}"
                 : "=r"(r) : "l"(p), "r"((uint32_t)pred) : "memory");
    return r;
}

// Row gate: wait until the chunk holding line `last_row` has arrived. Lane 0 of each warp spins (with
// back-off, so that a thousand waiting warps do not hammer one L2 line); then every thread performs the
// acquire itself, so that its own later loads are ordered behind it.
__device__ __forceinline__ uint32_t gate_load(const uint32_t *f)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
    return v;
}

__device__ __forceinline__ void gate_spin(const RowGate &g, const uint32_t *f)
{
    const long long t0 = clock64();
    unsigned ns = 64;
    while (gate_load(f) != g.epoch) {
        __nanosleep(ns);
        if (ns < 1024) ns *= 2;
        if (clock64() - t0 > 6000000000LL) {          // ~3 s at 2 GHz: the copy is not coming
            *g.timeout = 1u;
            break;
        }
    }
}

// called by all 32 lanes of a warp
__device__ __forceinline__ void gate_wait_warp(const RowGate &g, int last_row)
{
    const uint32_t *f = g.flags + last_row / g.rows;
    if (__all_sync(0xffffffffu, gate_load(f) == g.epoch)) return;
    if ((threadIdx.x & 31) == 0) gate_spin(g, f);
    __syncwarp();
    (void)gate_load(f);
}

// called by one thread (the one that issues a tile's TMA load)
__device__ __forceinline__ void gate_wait_thread(const RowGate &g, int last_row)
{
    const uint32_t *f = g.flags + last_row / g.rows;
    if (gate_load(f) != g.epoch) gate_spin(g, f);
}

// predicated 4-byte streaming load (no branch): 0 when pred is false
__device__ __forceinline__ uint32_t ldg_stream_u32_if(const void *p, bool pred)
{
    uint32_t r;
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\tmov.u32 %0, 0;\n\t"
                 "@q ld.global.nc.L1::no_allocate.u32 %0, [%1];\n\t}"
                 : "=r"(r) : "l"(p), "r"((uint32_t)pred));
    return r;
}

// Y sum of one pixel; the 8-bit result is byte 1 of the return value (sum is in 0..65535).
__device__ __forceinline__ uint32_t luma_sum(uint32_t pix, const Matrix &m)
{
    return dp4a_uu(pix, m.cy, m.oy);
}

// Four luma sums -> four Y bytes in pixel order.
__device__ __forceinline__ uint32_t pack_luma4(uint32_t s0, uint32_t s1, uint32_t s2, uint32_t s3)
{
    return prmt(prmt(s0, s1, 0x0051), prmt(s2, s3, 0x0051), 0x5410);
}

// (u,v) of one pixel as two 16-bit lanes: u in bits 0..7, v in bits 16..23.
template <bool NEG>
__device__ __forceinline__ uint32_t chroma_pair(uint32_t pix, const Matrix &m)
{
    uint32_t su = dp4a_us(pix, m.cu, m.ou);
    uint32_t sv = dp4a_us(pix, m.cv, m.ov);
    if (B200VC_SPEC_MATRIX_ROUND && NEG) {
        // with the rounding term a saturated primary reaches 65536 (negated: -1): clamp, i.e. 255
        su = (uint32_t)max((int32_t)su, 0);
        sv = (uint32_t)max((int32_t)sv, 0);
    }
    uint32_t p = prmt(su, sv, 0x6521);      // bytes: su.1, su.2 (=0), sv.1, sv.2 (=0)
    return NEG ? (p ^ 0x00FF00FFu) : p;
}

constexpr uint32_t LANES8 = 0x00FF00FFu;

// video_chroma_down_h2_u8 / v2: (a + b + 1) >> 1 on both lanes
__device__ __forceinline__ uint32_t avg2(uint32_t a, uint32_t b)
{
    return ((a + b + 0x00010001u) >> 1) & LANES8;
}

// video_chroma_down_h2_cs_u8: (l + 2c + r + 2) >> 2 on both lanes
__device__ __forceinline__ uint32_t filt121(uint32_t l, uint32_t c, uint32_t r)
{
    return ((l + 2u * c + r + 0x00020002u) >> 2) & LANES8;
}

// The same two without the final mask: the u lane's top bits hold the bits the shift dragged out of the
// v lane (the v lane is exact). For callers that add a masked operand and use bytes 0 and 2 only.
__device__ __forceinline__ uint32_t avg2_raw(uint32_t a, uint32_t b)
{
    return (a + b + 0x00010001u) >> 1;
}
__device__ __forceinline__ uint32_t filt121_raw(uint32_t l, uint32_t c, uint32_t r)
{
    return (l + 2u * c + r + 0x00020002u) >> 2;
}

// One 5-6-5 pixel (low 16 bits of v) as a 4-byte pixel: low field in byte 0, middle field in byte 1, high
// field in byte 2 (RGB16 reads like BGRx then, BGR16 like RGBx), each widened to 8 bits the way video-format.c
// unpacks them: high bits repeated into the low ones (B200VC_SPEC_UNPACK16_REPLICATE; 0: plain shift).
__device__ __forceinline__ uint32_t unpack565(uint32_t v)
{
    uint32_t t = ((v & 0xF800u) << 8) | ((v & 0x07E0u) << 5) | ((v & 0x001Fu) << 3);
    if (B200VC_SPEC_UNPACK16_REPLICATE) t |= ((t >> 5) & 0x00070007u) | ((t >> 6) & 0x00000300u);
    return t;
}

// Two 5-6-5 pixels (the halves of w) at once, on 2x16-bit lanes: a 5-bit field f widened with its high bits
// repeated is (f << 3) | (f >> 2) = (33 * f) >> 2, a 6-bit one (65 * f) >> 4 (no carry between the lanes: 33 * 31
// and 65 * 63 fit 16 bits); the shifts drag the upper lane's low bits into the top of the lower lane, where no
// byte is taken from — except into byte 3 of the first pixel, the unused x byte, whose matrix coefficient is 0.
__device__ __forceinline__ uint2 unpack565x2(uint32_t w)
{
    uint32_t lo = w & 0x001F001Fu, mid = (w >> 5) & 0x003F003Fu, hi = (w >> 11) & 0x001F001Fu;
    if (B200VC_SPEC_UNPACK16_REPLICATE) {
        lo = (lo * 33u) >> 2;  mid = (mid * 65u) >> 4;  hi = (hi * 33u) >> 2;
    } else {
        lo <<= 3;  mid <<= 2;  hi <<= 3;
    }
    const uint32_t lm = prmt(lo, mid, 0x6240);                       // lo0 mid0 lo1 mid1
    return make_uint2(prmt(lm, hi, 0x5410), prmt(lm, hi, 0x7632));   // {lo, mid, hi, x} of each pixel
}

__device__ __forceinline__ uint4 unpack565x4(uint2 raw)
{
    const uint2 a = unpack565x2(raw.x), b = unpack565x2(raw.y);
    return make_uint4(a.x, a.y, b.x, b.y);
}

// Scaler helpers (phase 2 of the general fused kernel, the tiled plane scaler of the two-pass path). A line's horizontal pass leaves two samples that still
// carry their 8 fraction bits (the sample is byte 1 of each sum); the vertical tap takes [A,A,B,B] from
// the sums of the two lines (v_weight_word, host_tables.h) and leaves the result in byte 1 again.
//
// luma, two adjacent output columns per lane: their four taps lie within 4 bytes of the first tap
// (step <= 2), which one funnel shift of two aligned words brings into a register
__device__ __forceinline__ uint2 h_pass_luma_pair(const uint8_t *line, uint32_t sh, uint32_t sel1, uint32_t wh0, uint32_t wh1)
{
    const uint32_t w0 = *reinterpret_cast<const uint32_t *>(line), w1 = *reinterpret_cast<const uint32_t *>(line + 4);
    const uint32_t v = __funnelshift_r(w0, w1, sh);                        // bytes i0 .. i0 + 3 of the line
    return make_uint2(dp4a_uu(prmt(v, 0u, 0x3100), wh0, 0u),               // [a,a,b,_] . (255-f, 1, f, 0)
                      dp4a_uu(prmt(v, 0u, sel1), wh1, 0u));
}

// chroma, one (u,v) sample per lane: the plane holds u,v byte pairs, the two taps are two 16-bit loads
__device__ __forceinline__ uint2 h_pass_chroma(const uint8_t *smp, uint32_t wh)
{
    const uint32_t lo = *reinterpret_cast<const uint16_t *>(smp), hi = *reinterpret_cast<const uint16_t *>(smp + 2);
    return make_uint2(dp4a_uu(prmt(lo, hi, 0x2400), wh, 0u),               // [u0,u0,u1,0]
                      dp4a_uu(prmt(lo, hi, 0x2511), wh, 0u));              // [v0,v0,v1,0]
}

// vertical tap on the two lines' sums of both samples; the samples land in bytes 0 and 1
__device__ __forceinline__ uint32_t v_pass_pair(uint2 top, uint2 bot, uint32_t wv)
{
    const uint32_t d0 = dp4a_uu(prmt(top.x, bot.x, 0x5511), wv, 128u), d1 = dp4a_uu(prmt(top.y, bot.y, 0x5511), wv, 128u);
    return prmt(d0, d1, 0x0051);
}

}  // namespace b200vc
