// k1_convert.cu — K1: BGRx -> I420/NV12, and K3: 16x16 macroblock damage map, fused (sm_100a).
//
// Replaces the per-frame body of GstVideoConvert::transform_frame -> gst_video_converter_frame on
// the reference's x264enc fallback (/root/reference Dockerfile:467-471, README.md:21); the line
// chain it restates is SURVEY.md §8 (a2)-(a5) / Appendix A, the damage map is §8 (a7).
//
// Layout of the work. A warp owns one 128 x 16 pixel stripe = 8 macroblocks of one macroblock row,
// and walks it in two steps of 8 lines. Lane l owns pixels 4l..4l+3 of every line: one coalesced
// 16-byte streaming load per line (512 contiguous bytes per warp request), eight in flight per
// step. Everything per pixel happens in registers:
//   Y      one dp4a per pixel, four results packed to a word with PRMT
//   U,V    two dp4a per pixel, kept as 2x16-bit lanes in one word; the horizontal filter
//          (h2 or cosited 1-2-1, neighbours across lanes by one shuffle, stripe halo by one
//          extra 4-byte load in lane 0) and the vertical rounded mean run on both lanes at once
//   damage XOR against the previous frame (x byte masked) accumulated over the 16 lines, one warp
//          ballot at the end, 4 lanes = 1 macroblock, 8 map bytes per warp
// Each step's planar output is staged in the warp's own 1.5 KB of shared memory (no block-level
// barrier anywhere) and leaves the SM as 16-byte stores of contiguous 128-byte row segments.
// A CTA is four independent warps covering 512 x 16 pixels.
//
// HBM roofline: algorithmic bytes per pixel = 4 (read) + 1.5 (write) [+ 4 previous frame].
#include "device_utils.cuh"
#include "../../include/b200vc_spec.h"

namespace b200vc {

namespace {

constexpr int TW = K1_TILE_W;            // 128 pixels per warp stripe
constexpr int TH = K1_TILE_H;            // 16 lines per warp stripe
constexpr int WARPS = K1_WARPS_PER_CTA;  // 4
constexpr uint32_t FULL = 0xffffffffu;

// HR = lines a warp loads, converts and stores per step (4 or 8); a stripe takes 16 / HR steps.
// FMT as in b200vc.h: 0 I420, 1 NV12 (one chroma row per line pair), 2 Y42B (one per line, half width),
// 3 Y444 (one per line, full width).
template <int HR, int FMT>
struct WarpStage {
    static constexpr int CROWS = FMT <= 1 ? HR / 2 : HR;
    static constexpr int CBYTES = FMT == 3 ? 2 * TW : TW;
    uint8_t y[HR][TW];                   // HR x 128 B
    uint8_t c[CROWS][CBYTES];            // planar: U in the first half of a row, V in the second; NV12: UVUV..
};

__device__ __forceinline__ uint32_t ldg_stream_u16(const void *p)
{
    uint16_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return r;
}

__device__ __noinline__ void store_tail(uint8_t *dst, const uint8_t *src, int valid)
{
    for (int i = 0; i < valid; i++) dst[i] = src[i];
}

// 16 bytes from shared to global, or the leading `valid` bytes when the row ends inside them
// Output that nothing on the GPU reads again goes out with the streaming hint (st.global.cs): measured
// +2-4 % on every K1 variant (the planes stop displacing the input stream in L2). The intermediate frame of
// the two-pass path is read back by pass 2 and keeps plain stores.
__device__ __forceinline__ void store16(uint8_t *dst, const uint8_t *src, int valid, bool streaming, bool vec)
{
    if (valid >= 16 && vec) {
        if (streaming)
            __stcs(reinterpret_cast<uint4 *>(dst), *reinterpret_cast<const uint4 *>(src));
        else
            *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(src);
    } else
        store_tail(dst, src, min(valid, 16));
}

// MODE 0: the planes are read again on the GPU (two-pass intermediate): plain stores. MODE 1: final
// output: streaming stores. MODE 2: final output of a frame that is still arriving (row gate): every
// stripe first waits for its lines, and the frame is read with coherent loads.
// IN16: the source holds 2-byte 5-6-5 pixels (RGB16/BGR16), unpacked in registers right after the load.
template <int FMT, int SITE, bool NEG, bool DMG, int HR, int MODE, bool IN16>
__global__ void __launch_bounds__(WARPS * 32) k1_convert_kernel(const ConvertParams p)
{
    constexpr bool STREAM = MODE != 0;
    constexpr bool LIVE = MODE == 2;
    static_assert(!(LIVE && IN16), "the row gate is built for 4-byte pixels");
    __shared__ __align__(16) WarpStage<HR, FMT> s_stage[WARPS];
    constexpr int PAIRS = HR / 2;
    constexpr int BPP = IN16 ? 2 : 4;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int x0 = (blockIdx.x * WARPS + warp) * TW;      // first column of this warp's stripe
    const int y0 = blockIdx.y * TH;
    const int frame = blockIdx.z;
    if (x0 >= p.w) return;                                 // warp-uniform: whole warp leaves
    if (LIVE) gate_wait_warp(p.gate, min(y0 + TH, p.h) - 1);

    WarpStage<HR, FMT> &st = s_stage[warp];
    const Matrix &m = p.mx;
    const int x = x0 + lane * 4;                           // first pixel of this thread
    const int remain = p.w - x;                            // pixels of this lane inside the frame: <= 0, 1..3, >= 4
    const bool interior_x = x0 + TW <= p.w;                // warp-uniform
    const bool halo = SITE == 1 && FMT != 3 && lane == 0 && x0 > 0;    // lane 0 fetches the stripe's left neighbour
    // Right frame edge, cosited: the last chroma site (column w-2 of an even width, w-1 of an odd
    // one; w > 2) takes itself as right neighbour (FILT_1_3). (Jpeg siting needs no flag: the edge
    // loader duplicates an unpaired last pixel, and avg2(p, p) = p.)
    // (B200VC_SPEC_COSITED_LAST_1_3 = 0: only an odd width's unpaired last pixel does; an even
    // width's last site gets the plain 1-2-1 filter.)
    const bool edge1 = SITE == 1 && p.w > 2 && remain <= (B200VC_SPEC_COSITED_LAST_1_3 ? 2 : 1);   // site at pixel 0
    const bool edge3 = SITE == 1 && p.w > 2 && remain <= (B200VC_SPEC_COSITED_LAST_1_3 ? 4 : 3);   // site at pixel 2

    const uint8_t *src = p.src + (long long)frame * p.src_frame_stride + (long long)y0 * p.src_pitch + (long long)x * BPP;
    const uint8_t *prv = DMG ? p.prev + (long long)frame * p.prev_frame_stride + (long long)y0 * p.prev_pitch + (long long)x * BPP
                             : nullptr;
    const long long fo = (long long)frame * p.out.frame_stride;
    const int seg = lane & 7, gx = x0 + seg * 16;          // this lane's 16-byte segment in the store phase
    const bool vec = p.out.aligned16 != 0;
    uint32_t diff = 0;

#pragma unroll 1
    for (int step = 0; step < TH / HR; step++) {
        const int yh = y0 + step * HR;
        if (yh >= p.h) break;                              // warp-uniform
        const int rows = min(HR, p.h - yh);                // even

        // ---- loads: HR lines x 16 bytes per thread in flight ------------------------------------
        uint4 px[HR];
        uint32_t hl[HR];
        if (rows == HR && interior_x) {                    // warp-uniform: straight-line loads
#pragma unroll
            for (int r = 0; r < HR; r++) {
                if (IN16) {
                    const uint2 raw = ldg_stream_v2(src);
                    px[r] = unpack565x4(raw);
                    hl[r] = (SITE == 1 && FMT != 3 && halo) ? unpack565(ldg_stream_u16(src - 2)) : 0;   // (lane 0 only: one short branch)
                    if (DMG) {
                        const uint2 q = ldg_stream_v2(prv);
                        diff |= (raw.x ^ q.x) | (raw.y ^ q.y);
                        prv += p.prev_pitch;
                    }
                } else {
                    px[r] = LIVE ? ldg_live_v4(src) : ldg_stream_v4(src);
                    hl[r] = (SITE == 1 && FMT != 3) ? (LIVE ? ldg_live_u32_if(src - 4, halo) : ldg_stream_u32_if(src - 4, halo)) : 0;
                    if (DMG) {
                        const uint4 q = ldg_stream_v4(prv);
                        diff |= (px[r].x ^ q.x) | (px[r].y ^ q.y) | (px[r].z ^ q.z) | (px[r].w ^ q.w);
                        prv += p.prev_pitch;
                    }
                }
                src += p.src_pitch;
            }
        } else {                                           // right/bottom frame edge: per-lane validity
#pragma unroll
            for (int r = 0; r < HR; r++) {
                px[r] = make_uint4(0, 0, 0, 0);            // pixels past the edge: never stored, never damage
                hl[r] = 0;
                if (r < rows && IN16) {
                    // pixel by pixel: 2-byte loads are always aligned; raw values side by side for the damage diff
                    uint32_t v[4] = { 0, 0, 0, 0 }, q[4] = { 0, 0, 0, 0 };
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        if (i < remain) {
                            v[i] = ldg_stream_u16(src + 2 * i);
                            if (DMG) q[i] = ldg_stream_u16(prv + 2 * i);
                        }
                    }
                    px[r] = make_uint4(remain > 0 ? unpack565(v[0]) : 0u, remain > 1 ? unpack565(v[1]) : 0u,
                                       remain > 2 ? unpack565(v[2]) : 0u, remain > 3 ? unpack565(v[3]) : 0u);
                    if (SITE == 1 && FMT != 3 && halo) hl[r] = unpack565(ldg_stream_u16(src - 2));
                    if (DMG) diff |= (v[0] ^ q[0]) | (v[1] ^ q[1]) | (v[2] ^ q[2]) | (v[3] ^ q[3]);
                    if (SITE == 0) {                       // odd width: the unpaired last pixel pairs with itself
                        if (remain == 1) px[r].y = px[r].x;
                        if (remain == 3) px[r].w = px[r].z;
                    }
                } else if (r < rows) {
                    uint4 q = make_uint4(0, 0, 0, 0);
                    if (remain >= 4) {
                        px[r] = LIVE ? ldg_live_v4(src) : ldg_stream_v4(src);
                        if (DMG) q = ldg_stream_v4(prv);
                    } else if (remain > 0) {               // 1..3 pixels: odd widths and widths % 4 != 0
                        px[r].x = LIVE ? ldg_live_u32(src) : ldg_stream_u32(src);
                        if (remain > 1) px[r].y = LIVE ? ldg_live_u32(src + 4) : ldg_stream_u32(src + 4);
                        if (remain > 2) px[r].z = LIVE ? ldg_live_u32(src + 8) : ldg_stream_u32(src + 8);
                        if (DMG) {
                            q.x = ldg_stream_u32(prv);
                            if (remain > 1) q.y = ldg_stream_u32(prv + 4);
                            if (remain > 2) q.z = ldg_stream_u32(prv + 8);
                        }
                    }
                    if (SITE == 1 && FMT != 3) hl[r] = LIVE ? ldg_live_u32_if(src - 4, halo) : ldg_stream_u32_if(src - 4, halo);
                    if (DMG) diff |= (px[r].x ^ q.x) | (px[r].y ^ q.y) | (px[r].z ^ q.z) | (px[r].w ^ q.w);
                    if (SITE == 0) {                       // odd width: the unpaired last pixel pairs with itself
                        if (remain == 1) px[r].y = px[r].x;
                        if (remain == 3) px[r].w = px[r].z;
                    }
                }
                src += p.src_pitch;
                if (DMG) prv += p.prev_pitch;
            }
            if (rows & 1) {                                // odd height: the last line pairs with itself
#pragma unroll
                for (int r = 1; r < HR; r += 2)
                    if (r == rows) { px[r] = px[r - 1]; hl[r] = hl[r - 1]; }
            }
        }

        // ---- per line pair: luma, chroma ---------------------------------------------------------
#pragma unroll
        for (int k = 0; k < PAIRS; k++) {
            const uint4 a = px[2 * k], b = px[2 * k + 1];
            *reinterpret_cast<uint32_t *>(&st.y[2 * k][lane * 4]) =
                pack_luma4(luma_sum(a.x, m), luma_sum(a.y, m), luma_sum(a.z, m), luma_sum(a.w, m));
            *reinterpret_cast<uint32_t *>(&st.y[2 * k + 1][lane * 4]) =
                pack_luma4(luma_sum(b.x, m), luma_sum(b.y, m), luma_sum(b.z, m), luma_sum(b.w, m));

            const uint32_t pa0 = chroma_pair<NEG>(a.x, m), pa1 = chroma_pair<NEG>(a.y, m);
            const uint32_t pa2 = chroma_pair<NEG>(a.z, m), pa3 = chroma_pair<NEG>(a.w, m);
            const uint32_t pb0 = chroma_pair<NEG>(b.x, m), pb1 = chroma_pair<NEG>(b.y, m);
            const uint32_t pb2 = chroma_pair<NEG>(b.z, m), pb3 = chroma_pair<NEG>(b.w, m);
            if (FMT == 3) {
                // 4:4:4: every pixel's u and v, planar: four u bytes and four v bytes per lane and line
                *reinterpret_cast<uint32_t *>(&st.c[2 * k][lane * 4]) = prmt(prmt(pa0, pa1, 0x0040), prmt(pa2, pa3, 0x0040), 0x5410);
                *reinterpret_cast<uint32_t *>(&st.c[2 * k][TW + lane * 4]) = prmt(prmt(pa0, pa1, 0x0062), prmt(pa2, pa3, 0x0062), 0x5410);
                *reinterpret_cast<uint32_t *>(&st.c[2 * k + 1][lane * 4]) = prmt(prmt(pb0, pb1, 0x0040), prmt(pb2, pb3, 0x0040), 0x5410);
                *reinterpret_cast<uint32_t *>(&st.c[2 * k + 1][TW + lane * 4]) = prmt(prmt(pb0, pb1, 0x0062), prmt(pb2, pb3, 0x0062), 0x5410);
                continue;
            }
            // horizontal filter per line: two sites per lane (pixels 0 and 2)
            uint32_t ha0, ha1, hb0, hb1;
            if (SITE == 0) {
                ha0 = avg2(pa0, pa1);  ha1 = avg2(pa2, pa3);
                hb0 = avg2(pb0, pb1);  hb1 = avg2(pb2, pb3);
            } else {
                // left neighbour of pixel 0: previous lane's pixel 3, the stripe halo, or (column 0) itself
                uint32_t la = __shfl_up_sync(FULL, pa3, 1);
                uint32_t lb = __shfl_up_sync(FULL, pb3, 1);
                if (lane == 0) {
                    la = halo ? chroma_pair<NEG>(hl[2 * k], m) : pa0;
                    lb = halo ? chroma_pair<NEG>(hl[2 * k + 1], m) : pb0;
                }
                const uint32_t ra1 = edge1 ? pa0 : pa1, rb1 = edge1 ? pb0 : pb1;
                const uint32_t ra3 = edge3 ? pa2 : pa3, rb3 = edge3 ? pb2 : pb3;
                ha0 = filt121(la, pa0, ra1);  ha1 = filt121(pa1, pa2, ra3);
                hb0 = filt121(lb, pb0, rb1);  hb1 = filt121(pb1, pb2, rb3);
            }
            if (FMT == 2) {
                // 4:2:2: the horizontally filtered sites of each line are the samples
                const uint32_t a = prmt(ha0, ha1, 0x6240), b = prmt(hb0, hb1, 0x6240);   // u0 u1 v0 v1
                *reinterpret_cast<uint16_t *>(&st.c[2 * k][lane * 2]) = (uint16_t)(a & 0xffffu);
                *reinterpret_cast<uint16_t *>(&st.c[2 * k][64 + lane * 2]) = (uint16_t)(a >> 16);
                *reinterpret_cast<uint16_t *>(&st.c[2 * k + 1][lane * 2]) = (uint16_t)(b & 0xffffu);
                *reinterpret_cast<uint16_t *>(&st.c[2 * k + 1][64 + lane * 2]) = (uint16_t)(b >> 16);
                continue;
            }
            // 4:2:0: then the rounded mean of the two lines
            const uint32_t c0 = avg2(ha0, hb0), c1 = avg2(ha1, hb1);
            if (FMT == 0) {
                const uint32_t uuvv = prmt(c0, c1, 0x6240);                 // u0 u1 v0 v1
                *reinterpret_cast<uint16_t *>(&st.c[k][lane * 2]) = (uint16_t)(uuvv & 0xffffu);
                *reinterpret_cast<uint16_t *>(&st.c[k][64 + lane * 2]) = (uint16_t)(uuvv >> 16);
            } else {
                *reinterpret_cast<uint32_t *>(&st.c[k][lane * 4]) = prmt(c0, c1, 0x6420);   // u0 v0 u1 v1
            }
        }
        __syncwarp();

        // ---- planar stores: 16 bytes per lane, 128-byte row segments -----------------------------
#pragma unroll
        for (int i = 0; i < HR / 4; i++) {                // Y: HR rows x 8 segments
            const int r = i * 4 + (lane >> 3);
            if (r < rows && gx < p.w)
                store16(p.out.p[0] + fo + (long long)(yh + r) * p.out.pitch[0] + gx, &st.y[r][seg * 16], p.w - gx, STREAM, vec);
        }
        if (FMT <= 1) {
            const int r = lane >> 3;                       // chroma: HR/2 rows x 8 segments
            if (r < PAIRS && 2 * r < rows) {
                const long long cy = (yh >> 1) + r;
                const int cw = (p.w + 1) >> 1;             // chroma samples per line (odd widths round up)
                if (FMT == 0) {
                    const int plane = 1 + (seg >> 2), cx = (x0 >> 1) + (seg & 3) * 16;
                    if (cx < cw)
                        store16(p.out.p[plane] + fo + cy * p.out.pitch[plane] + cx, &st.c[r][seg * 16], cw - cx, STREAM, vec);
                } else if (gx < 2 * cw) {
                    store16(p.out.p[1] + fo + cy * p.out.pitch[1] + gx, &st.c[r][seg * 16], 2 * cw - gx, STREAM, vec);
                }
            }
        } else if (FMT == 2) {
            const int cw = (p.w + 1) >> 1;                 // Y42B: HR rows x (4 U + 4 V) segments
            const int plane = 1 + (seg >> 2), cx = (x0 >> 1) + (seg & 3) * 16;
#pragma unroll
            for (int i = 0; i < HR / 4; i++) {
                const int r = i * 4 + (lane >> 3);
                if (r < rows && cx < cw)
                    store16(p.out.p[plane] + fo + (long long)(yh + r) * p.out.pitch[plane] + cx, &st.c[r][seg * 16], cw - cx, STREAM, vec);
            }
        } else {
#pragma unroll
            for (int i = 0; i < HR / 2; i++) {             // Y444: HR rows x (8 U + 8 V) segments
                const int idx = i * 32 + lane, r = idx >> 4, sg = idx & 15;
                const int plane = 1 + (sg >> 3), cx = x0 + (sg & 7) * 16;
                if (r < rows && cx < p.w)
                    store16(p.out.p[plane] + fo + (long long)(yh + r) * p.out.pitch[plane] + cx, &st.c[r][sg * 16], p.w - cx, STREAM, vec);
            }
        }
        __syncwarp();                                      // the stage is reused by the next step
    }

    // ---- K3: one ballot per stripe, 4 lanes = 1 macroblock ---------------------------------------
    if (DMG) {
        const uint32_t ballot = __ballot_sync(FULL, (diff & p.mx.rgb_mask) != 0);
        const int mbx = (x0 >> 4) + lane;
        if (lane < 8 && mbx < p.mb_w)
            p.damage[(long long)frame * p.damage_frame_stride + (long long)(y0 >> 4) * p.mb_w + mbx] =
                ((ballot >> (4 * lane)) & 0xFu) ? 1 : 0;
    }
}

// Measured on B200 (profiles/README.md): 8 lines per step (8 x 16 bytes in flight per thread,
// 63-96 registers) runs at 0.99 of the copy roofline; 4 lines per step (48-56 registers) at 0.93-0.99.
constexpr int K1_ROWS_PER_STEP = 8;

template <int FMT, int SITE, bool NEG, bool DMG>
cudaError_t launch4(const ConvertParams &p, dim3 grid, cudaStream_t st)
{
    if (p.in16) {                                          // (no row gate for 2-byte pixels: launch_convert refuses it)
        if (p.stream_out)
            k1_convert_kernel<FMT, SITE, NEG, DMG, K1_ROWS_PER_STEP, 1, true><<<grid, WARPS * 32, 0, st>>>(p);
        else
            k1_convert_kernel<FMT, SITE, NEG, DMG, K1_ROWS_PER_STEP, 0, true><<<grid, WARPS * 32, 0, st>>>(p);
    } else if (p.gate.flags) {
        if constexpr (FMT <= 1)
            k1_convert_kernel<FMT, SITE, NEG, DMG, K1_ROWS_PER_STEP, 2, false><<<grid, WARPS * 32, 0, st>>>(p);
        else
            return cudaErrorInvalidValue;                  // (the row gate serves the 4:2:0 formats)
    } else if (p.stream_out) {
        k1_convert_kernel<FMT, SITE, NEG, DMG, K1_ROWS_PER_STEP, 1, false><<<grid, WARPS * 32, 0, st>>>(p);
    } else {
        k1_convert_kernel<FMT, SITE, NEG, DMG, K1_ROWS_PER_STEP, 0, false><<<grid, WARPS * 32, 0, st>>>(p);
    }
    return cudaGetLastError();
}

template <int FMT, int SITE, bool NEG>
cudaError_t launch3(const ConvertParams &p, dim3 grid, cudaStream_t st)
{
    if constexpr (FMT <= 1) {
        if (p.prev && p.damage) return launch4<FMT, SITE, NEG, true>(p, grid, st);
    } else {
        if (p.prev && p.damage) return cudaErrorInvalidValue;      // (the damage map goes with the 4:2:0 formats)
    }
    return launch4<FMT, SITE, NEG, false>(p, grid, st);
}

template <int FMT, int SITE>
cudaError_t launch2(const ConvertParams &p, dim3 grid, cudaStream_t st)
{
    return p.mx.neg ? launch3<FMT, SITE, true>(p, grid, st) : launch3<FMT, SITE, false>(p, grid, st);
}

}  // namespace

cudaError_t launch_convert(const ConvertParams &p, int fmt, int site, cudaStream_t st)
{
    if (p.w <= 0 || p.h <= 0 || p.n_frames <= 0 || (p.gate.flags && (p.n_frames != 1 || p.gate.rows <= 0 || !p.gate.timeout || p.in16)))
        return cudaErrorInvalidValue;
    dim3 grid((p.w + TW * WARPS - 1) / (TW * WARPS), (p.h + TH - 1) / TH, p.n_frames);
    switch (fmt) {
    case 0: return site ? launch2<0, 1>(p, grid, st) : launch2<0, 0>(p, grid, st);
    case 1: return site ? launch2<1, 1>(p, grid, st) : launch2<1, 0>(p, grid, st);
    case 2: return site ? launch2<2, 1>(p, grid, st) : launch2<2, 0>(p, grid, st);
    case 3: return launch2<3, 0>(p, grid, st);             // 4:4:4: no chroma filter, the site does not matter
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace b200vc
