// k1_convert.cu — K1: BGRx -> I420/NV12, and K3: 16x16 macroblock damage map, fused (sm_100a).
//
// Replaces the per-frame body of GstVideoConvert::transform_frame -> gst_video_converter_frame on
// the reference's x264enc fallback (/root/reference Dockerfile:467-471, README.md:21); the line
// chain it restates is SURVEY.md §8 (a2)-(a5) / Appendix A, the damage map is §8 (a7).
//
// Layout of the work. One CTA (8 warps) owns a 128 x 16 pixel tile = 8 macroblocks of one
// macroblock row. Warp w owns the line pair (2w, 2w+1); lane l owns pixels 4l..4l+3 of both lines,
// fetched as one coalesced 16-byte streaming load per line (512 contiguous bytes per warp
// request). Everything per pixel happens in registers:
//   Y      one dp4a per pixel, four results packed to a word with PRMT
//   U,V    two dp4a per pixel, kept as 2x16-bit lanes in one word; the horizontal filter
//          (h2 or cosited 1-2-1, neighbours across lanes by one shuffle, tile halo by one
//          extra 4-byte load in lane 0) and the vertical rounded mean run on both lanes at once
//   damage XOR against the previous frame (x byte masked), warp ballot, 4 lanes = 1 macroblock
// The tile's planar output is staged in 3 KB of shared memory so that every plane leaves the SM
// as full 16-byte stores of contiguous row segments.
//
// HBM roofline: algorithmic bytes per pixel = 4 (read) + 1.5 (write) [+ 4 previous frame].
#include "device_utils.cuh"

namespace b200vc {

namespace {

constexpr int TW = K1_TILE_W;
constexpr int TH = K1_TILE_H;
constexpr uint32_t FULL = 0xffffffffu;

template <int FMT, int SITE, bool NEG, bool DMG>
__global__ void __launch_bounds__(256) k1_convert_kernel(const ConvertParams p)
{
    __shared__ __align__(16) uint8_t s_y[TH][TW];
    __shared__ __align__(16) uint8_t s_c[TH / 2][TW];   // I420: [0..63] U, [64..127] V; NV12: UVUV..
    __shared__ uint32_t s_dmg[8];

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int x0 = blockIdx.x * TW;
    const int y0 = blockIdx.y * TH;
    const int frame = blockIdx.z;

    const int x = x0 + lane * 4;          // first pixel of this thread
    const int ya = y0 + warp * 2;         // even line of this warp's pair
    const bool row_ok = ya < p.h;         // h is even, so ya + 1 < h as well
    int nvalid = p.w - x;                 // w even, x % 4 == 0  ->  <=0, 2, or >= 4
    nvalid = nvalid < 0 ? 0 : (nvalid > 4 ? 4 : nvalid);

    const uint8_t *src = p.src + (long long)frame * p.src_frame_stride + (long long)ya * p.src_pitch;
    uint4 a = make_uint4(0, 0, 0, 0), b = make_uint4(0, 0, 0, 0);
    uint32_t halo_a = 0, halo_b = 0;

    if (row_ok) {
        if (nvalid == 4) {
            a = ldg_stream_v4(src + (long long)x * 4);
            b = ldg_stream_v4(src + p.src_pitch + (long long)x * 4);
        } else if (nvalid == 2) {
            uint2 t = ldg_stream_v2(src + (long long)x * 4);
            a = make_uint4(t.x, t.y, t.y, t.y);
            t = ldg_stream_v2(src + p.src_pitch + (long long)x * 4);
            b = make_uint4(t.x, t.y, t.y, t.y);
        }
        if (SITE == 1 && lane == 0 && x > 0 && nvalid) {       // left halo pixel of the tile
            halo_a = ldg_stream_u32(src + (long long)(x - 1) * 4);
            halo_b = ldg_stream_u32(src + p.src_pitch + (long long)(x - 1) * 4);
        }
    }

    // ---- K3: damage -----------------------------------------------------------------------
    if (DMG) {
        uint32_t diff = 0;
        if (row_ok && nvalid) {
            const uint8_t *prv = p.prev + (long long)frame * p.prev_frame_stride + (long long)ya * p.prev_pitch;
            uint4 pa, pb;
            if (nvalid == 4) {
                pa = ldg_stream_v4(prv + (long long)x * 4);
                pb = ldg_stream_v4(prv + p.prev_pitch + (long long)x * 4);
            } else {
                uint2 t = ldg_stream_v2(prv + (long long)x * 4);
                pa = make_uint4(t.x, t.y, t.y, t.y);
                t = ldg_stream_v2(prv + p.prev_pitch + (long long)x * 4);
                pb = make_uint4(t.x, t.y, t.y, t.y);
            }
            diff = ((a.x ^ pa.x) | (a.y ^ pa.y) | (a.z ^ pa.z) | (a.w ^ pa.w) |
                    (b.x ^ pb.x) | (b.y ^ pb.y) | (b.z ^ pb.z) | (b.w ^ pb.w)) & 0x00FFFFFFu;
        }
        const uint32_t ballot = __ballot_sync(FULL, diff != 0);
        if (lane == 0)
            s_dmg[warp] = ballot;
    }

    // ---- luma --------------------------------------------------------------------------------
    const Matrix &m = p.mx;
    {
        uint32_t ya4 = pack_luma4(luma_sum(a.x, m), luma_sum(a.y, m), luma_sum(a.z, m), luma_sum(a.w, m));
        uint32_t yb4 = pack_luma4(luma_sum(b.x, m), luma_sum(b.y, m), luma_sum(b.z, m), luma_sum(b.w, m));
        *reinterpret_cast<uint32_t *>(&s_y[warp * 2][lane * 4]) = ya4;
        *reinterpret_cast<uint32_t *>(&s_y[warp * 2 + 1][lane * 4]) = yb4;
    }

    // ---- chroma: 4:4:4 (u,v) pairs, horizontal filter per line, vertical mean ------------------
    uint32_t pa0 = chroma_pair<NEG>(a.x, m), pa1 = chroma_pair<NEG>(a.y, m);
    uint32_t pa2 = chroma_pair<NEG>(a.z, m), pa3 = chroma_pair<NEG>(a.w, m);
    uint32_t pb0 = chroma_pair<NEG>(b.x, m), pb1 = chroma_pair<NEG>(b.y, m);
    uint32_t pb2 = chroma_pair<NEG>(b.z, m), pb3 = chroma_pair<NEG>(b.w, m);
    uint32_t c0, c1;
    if (SITE == 0) {
        c0 = avg2(avg2(pa0, pa1), avg2(pb0, pb1));
        c1 = avg2(avg2(pa2, pa3), avg2(pb2, pb3));
    } else {
        // left neighbour of pixel 0: previous lane's pixel 3, the tile halo, or (column 0) itself
        uint32_t la = __shfl_up_sync(FULL, pa3, 1);
        uint32_t lb = __shfl_up_sync(FULL, pb3, 1);
        if (lane == 0) {
            la = (x > 0) ? chroma_pair<NEG>(halo_a, m) : pa0;
            lb = (x > 0) ? chroma_pair<NEG>(halo_b, m) : pb0;
        }
        // last even column (w-2, when w > 2) takes itself as right neighbour: FILT_1_3
        uint32_t ra1 = pa1, rb1 = pb1, ra3 = pa3, rb3 = pb3;
        if (p.w > 2) {
            if (x + 4 == p.w) { ra3 = pa2; rb3 = pb2; }
            if (x + 2 == p.w) { ra1 = pa0; rb1 = pb0; }
        }
        c0 = avg2(filt121(la, pa0, ra1), filt121(lb, pb0, rb1));
        c1 = avg2(filt121(pa1, pa2, ra3), filt121(pb1, pb2, rb3));
    }
    if (FMT == 0) {
        uint32_t uuvv = prmt(c0, c1, 0x6240);                 // u0 u1 v0 v1
        *reinterpret_cast<uint16_t *>(&s_c[warp][lane * 2]) = (uint16_t)(uuvv & 0xffffu);
        *reinterpret_cast<uint16_t *>(&s_c[warp][64 + lane * 2]) = (uint16_t)(uuvv >> 16);
    } else {
        *reinterpret_cast<uint32_t *>(&s_c[warp][lane * 4]) = prmt(c0, c1, 0x6420);   // u0 v0 u1 v1
    }

    __syncthreads();

    // ---- planar stores, 16 bytes per thread -------------------------------------------------------
    const int t = threadIdx.x;
    const long long fo = (long long)frame * p.out.frame_stride;
    if (t < 128) {                                         // Y: 16 rows x 8 segments
        const int row = t >> 3, seg = t & 7;
        const int gy = y0 + row, gx = x0 + seg * 16;
        if (gy < p.h && gx < p.w) {
            uint8_t *dst = p.out.p[0] + fo + (long long)gy * p.out.pitch[0] + gx;
            if (gx + 16 <= p.w) {
                *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(&s_y[row][seg * 16]);
            } else {
                for (int i = 0; i < p.w - gx; i++) dst[i] = s_y[row][seg * 16 + i];
            }
        }
    } else if (t < 192) {                                  // chroma: 8 rows x 8 segments
        const int idx = t - 128, row = idx >> 3, s = idx & 7;
        const int gy = (y0 >> 1) + row;
        if (gy < (p.h >> 1)) {
            if (FMT == 0) {
                const int plane = 1 + (s >> 2), seg = s & 3;
                const int gx = (x0 >> 1) + seg * 16, cw = p.w >> 1;
                if (gx < cw) {
                    uint8_t *dst = p.out.p[plane] + fo + (long long)gy * p.out.pitch[plane] + gx;
                    const uint8_t *sp = &s_c[row][(plane - 1) * 64 + seg * 16];
                    if (gx + 16 <= cw) {
                        *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(sp);
                    } else {
                        for (int i = 0; i < cw - gx; i++) dst[i] = sp[i];
                    }
                }
            } else {
                const int gx = x0 + s * 16;
                if (gx < p.w) {
                    uint8_t *dst = p.out.p[1] + fo + (long long)gy * p.out.pitch[1] + gx;
                    const uint8_t *sp = &s_c[row][s * 16];
                    if (gx + 16 <= p.w) {
                        *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(sp);
                    } else {
                        for (int i = 0; i < p.w - gx; i++) dst[i] = sp[i];
                    }
                }
            }
        }
    } else if (DMG && t < 200) {                           // damage: 8 macroblocks of this tile
        const int k = t - 192;
        const int mbx = (x0 >> 4) + k;
        if (mbx < p.mb_w) {
            uint32_t any = 0;
#pragma unroll
            for (int w = 0; w < 8; w++) any |= (s_dmg[w] >> (4 * k)) & 0xFu;
            p.damage[(long long)frame * p.damage_frame_stride + (long long)(y0 >> 4) * p.mb_w + mbx] = any ? 1 : 0;
        }
    }
}

template <int FMT, int SITE, bool NEG>
cudaError_t launch3(const ConvertParams &p, dim3 grid, cudaStream_t st)
{
    if (p.prev && p.damage)
        k1_convert_kernel<FMT, SITE, NEG, true><<<grid, 256, 0, st>>>(p);
    else
        k1_convert_kernel<FMT, SITE, NEG, false><<<grid, 256, 0, st>>>(p);
    return cudaGetLastError();
}

template <int FMT, int SITE>
cudaError_t launch2(const ConvertParams &p, dim3 grid, cudaStream_t st)
{
    return p.mx.neg ? launch3<FMT, SITE, true>(p, grid, st) : launch3<FMT, SITE, false>(p, grid, st);
}

}  // namespace

cudaError_t launch_convert(const ConvertParams &p, int fmt, int site, cudaStream_t st)
{
    if (p.w <= 0 || p.h <= 0 || p.n_frames <= 0)
        return cudaErrorInvalidValue;
    dim3 grid((p.w + TW - 1) / TW, (p.h + TH - 1) / TH, p.n_frames);
    if (fmt == 0)
        return site ? launch2<0, 1>(p, grid, st) : launch2<0, 0>(p, grid, st);
    return site ? launch2<1, 1>(p, grid, st) : launch2<1, 0>(p, grid, st);
}

}  // namespace b200vc
