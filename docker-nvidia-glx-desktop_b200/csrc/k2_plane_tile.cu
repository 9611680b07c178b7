// k2_plane_tile.cu — two-pass scale, pass 2, tiled: scales the planes of the full-size intermediate
// I420 frame (written by K1) to the output size, videoscale's 2-tap "lq" arithmetic:
// horizontal (a*(256-f) + b*f) >> 8 on each of two lines, then vertical (a*(64-p1) + b*p1 + 32) >> 6.
// This is the path of geometries the fused kernels do not take (odd sizes, windows larger than a TMA
// box, B200VC_FLAG_FORCE_TWO_PASS); the gather kernel in k2_scale.cu stays as the fallback for windows
// that do not fit shared memory.
//
// A CTA makes a 128 x 64 tile of one output plane (for NV12 chroma: of U and V, interleaved on the way
// out). Its source window is copied into shared memory with coalesced 16-byte loads, taps come ready-made
// from host-built per-tile records (PlaneTileKind, kernels.cuh), a lane owns two adjacent output columns
// (word loads; or one column and byte gathers free of bank conflicts where a pair's taps do not fit 4 bytes),
// and the tile leaves through shared memory as 16-byte stores — the structure measured on the general
// fused kernel (profiles/README.md).
#include "device_utils.cuh"

#include <algorithm>
#include "kernels.cuh"
#include "../../include/b200vc_spec.h"

namespace b200vc {
namespace {

constexpr int TW = KPT_TILE_W;
constexpr int TH = KPT_TILE_H;
constexpr int CG = TW / 32;           // column groups of 32 lanes
constexpr int RPT = TH / (8 / CG);   // rows per thread: 8 warps = CG column groups x 8/CG row groups

// Copies a window (bh + 1 lines of bwp bytes, bwp and sx multiples of 16; lines past the plane's last
// one repeat it) into s_win with 16-byte loads: thread t takes chunk t % 16 of lines t / 16, t / 16 + 16,
// ...; six lines' loads are in flight per thread before the first is stored (one load per pass of the
// loop would serialise on DRAM latency).
__device__ __forceinline__ void load_window(uint8_t *s_win, const uint8_t *plane, long long pitch, int sx, int sy, int bwp,
                                            int bh, int src_h, int tid)
{
    constexpr int U = 6;
    const int cpl = bwp >> 4;                                      // 16-byte chunks per window line
    const int cmax = min(cpl, (int)((pitch - sx) >> 4));           // never past the plane's line
    for (int c = tid & 15; c < cmax; c += 16) {
        for (int r0 = tid >> 4; r0 <= bh; r0 += 16 * U) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int r = r0 + 16 * u;
                if (r <= bh)
                    v[u] = __ldg(reinterpret_cast<const uint4 *>(plane + (long long)min(sy + r, src_h - 1) * pitch + sx) + c);
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int r = r0 + 16 * u;
                if (r <= bh) *reinterpret_cast<uint4 *>(s_win + r * bwp + c * 16) = v[u];
            }
        }
    }
}

// One output sample from ready-made taps: both lines' horizontal taps as [a,a,b,_] . (255-f, 1, f, 0),
// then the vertical tap on the two sums' high bytes as [A,A,B,B] . (wa1,wa2,wb1,wb2) + 128; the sample
// is byte 1 of the result (as a byte: result >> 8).
__device__ __forceinline__ uint32_t tile_sample(const uint8_t *r0, int bwp, uint32_t wh, uint32_t wv, int box)
{
    const uint8_t *r1 = r0 + bwp;
    if (box)
        return (uint32_t)(r0[0] + r0[1] + r1[0] + r1[1] + 2) << 6;
    if (!B200VC_SPEC_SCALE_H_FIRST) {                    // vertical pass first (spec variant)
        const uint32_t v0 = dp4a_uu(prmt(r0[0], r1[0], 0x4400), wv, 128u) >> 8;
        const uint32_t v1 = dp4a_uu(prmt(r0[1], r1[1], 0x4400), wv, 128u) >> 8;
        return dp4a_uu(prmt(v0, v1, 0x1400), wh, 0u);
    }
    const uint32_t s0 = dp4a_uu(prmt(r0[0], r0[1], 0x1400), wh, 0u);
    const uint32_t s1 = dp4a_uu(prmt(r1[0], r1[1], 0x1400), wh, 0u);
    return dp4a_uu(prmt(s0, s1, 0x5511), wv, 128u);
}

template <int FMT>
__global__ void __launch_bounds__(256) k2_plane_tile_kernel(const PlaneTileParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ __align__(16) uint32_t s_col[2 * TW];         // offsets[TW] | horizontal weight words[TW], or TW/2 pair records
    __shared__ __align__(8) uint2 s_row[TH];                 // {line offset in the window, vertical weight word}
    __shared__ __align__(16) uint8_t s_out[TW * TH * 2];     // output tile (NV12 chroma: 128-byte rows)

    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int bx = blockIdx.x, frame = blockIdx.z;

    // ---- which tile of which plane (uniform per CTA) ------------------------------------------------
    int kind = 0, pl = 0, trow = blockIdx.y;
    if (trow >= p.k[0].tiles_y) {
        kind = 1;
        trow -= p.k[0].tiles_y;
        pl = 1;
        if (FMT == 0 && trow >= p.k[1].tiles_y) { trow -= p.k[1].tiles_y;  pl = 2; }
    }
    const PlaneTileKind &k = p.k[kind];
    if (bx >= k.tiles_x) return;
    const bool both = FMT == 1 && kind == 1;                 // NV12 chroma: U and V in one tile
    const int bwp = k.bwp, bh = k.bh;
    const int sx = k.tile_x[bx], sy = k.tile_y[trow];

    // ---- taps and window into shared memory ------------------------------------------------------------
    if (tid < 2 * TW) s_col[tid] = k.col_rec[bx * (2 * TW) + tid];
    if (tid < 2 * TH) reinterpret_cast<uint32_t *>(s_row)[tid] = k.row_rec[trow * (2 * TH) + tid];
    const long long fi = (long long)frame * p.src_frame_stride;
    const int win_bytes = bwp * (bh + 1);
    load_window(smem, p.src[pl] + fi, p.src_pitch[pl], sx, sy, bwp, bh, k.src_h, tid);
    if (both)
        load_window(smem + win_bytes, p.src[2] + fi, p.src_pitch[2], sx, sy, bwp, bh, k.src_h, tid);
    __syncthreads();

    // ---- scale -------------------------------------------------------------------------------------------
    // Two adjacent output columns per lane where every pair's taps lie within 4 bytes (ratios up to 2; the form
    // the general fused kernel's phase 2 took in round 2: word loads and a funnel shift instead of byte loads,
    // a quarter of the shared-memory wavefronts); a warp is half a tile row wide, TH/4 rows per thread.
    if (k.pair && !p.box2x2 && B200VC_SPEC_SCALE_H_FIRST) {
        const int cg = wrp & 1, rb = (wrp >> 1) * (TH / 4);
        const uint4 rec = reinterpret_cast<const uint4 *>(s_col)[cg * 32 + lane];
        const uint8_t *col = smem + (rec.x & 0xffffu);
        const uint32_t sh = rec.x >> 16;
#pragma unroll 4
        for (int i = 0; i < TH / 4; i++) {
            const uint2 tv = s_row[rb + i];
            const uint8_t *l0 = col + tv.x;
            const uint32_t d = v_pass_pair(h_pass_luma_pair(l0, sh, rec.y, rec.z, rec.w), h_pass_luma_pair(l0 + bwp, sh, rec.y, rec.z, rec.w), tv.y);
            if (both) {
                const uint8_t *m0 = l0 + win_bytes;
                const uint32_t e = v_pass_pair(h_pass_luma_pair(m0, sh, rec.y, rec.z, rec.w), h_pass_luma_pair(m0 + bwp, sh, rec.y, rec.z, rec.w), tv.y);
                reinterpret_cast<uint32_t *>(s_out)[(rb + i) * (TW / 2) + cg * 32 + lane] = prmt(d, e, 0x5140);   // u0 v0 u1 v1
            } else {
                reinterpret_cast<uint16_t *>(s_out)[(rb + i) * (TW / 2) + cg * 32 + lane] = (uint16_t)d;
            }
        }
    } else {
        // one output column x RPT rows per thread, byte loads (ratios above 2, box mode, the vertical-first variant)
        const int c = (wrp % CG) * 32 + lane, rb = (wrp / CG) * RPT;
        const uint32_t wh = s_col[TW + c];
        const uint8_t *col = smem + s_col[c];
#pragma unroll
        for (int i = 0; i < RPT; i++) {
            const uint2 tv = s_row[rb + i];
            const uint32_t d = tile_sample(col + tv.x, bwp, wh, tv.y, p.box2x2);
            if (both) {
                const uint32_t e = tile_sample(col + win_bytes + tv.x, bwp, wh, tv.y, p.box2x2);
                reinterpret_cast<uint16_t *>(s_out)[(rb + i) * TW + c] = (uint16_t)prmt(d, e, 0x4451);   // u, v
            } else {
                s_out[(rb + i) * TW + c] = (uint8_t)(d >> 8);
            }
        }
    }
    __syncthreads();

    // ---- the tile leaves as 16-byte stores ---------------------------------------------------------------
    const int row_bytes = both ? 2 * TW : TW;                // bytes of a full tile row
    const int cpr = row_bytes >> 4;                          // 16-byte chunks per row: 4 or 8
    const int valid = (both ? 2 : 1) * (k.out_w - bx * TW);  // bytes of this tile's rows inside the plane
    const int y0 = trow * TH;
    uint8_t *plane = p.out.p[pl] + (long long)frame * p.out.frame_stride + (long long)bx * row_bytes;
    for (int ch = tid; ch < TH * cpr; ch += 256) {
        const int r = ch / cpr, q = ch & (cpr - 1);
        if (y0 + r >= k.out_h) break;
        const int n = valid - q * 16;
        const uint8_t *src = s_out + r * row_bytes + q * 16;
        uint8_t *dst = plane + (long long)(y0 + r) * p.out.pitch[pl] + q * 16;
        if (n >= 16 && p.out.aligned16) {
            *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(src);
        } else {
            for (int i = 0; i < min(n, 16); i++) dst[i] = src[i];
        }
    }
}

}  // namespace

namespace {

// add-borders: a thread looks at 16 bytes of one row of one plane and writes what lies outside the picture
__global__ void __launch_bounds__(256) border_fill_kernel(const BorderParams p)
{
    const int pl = blockIdx.y, frame = blockIdx.z;
    if (pl >= p.n_planes) return;
    const int chunks = (p.w[pl] + 15) >> 4;
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    const int row = (int)(idx / chunks), x0 = (int)(idx % chunks) * 16;
    if (row >= p.h[pl]) return;
    uint8_t *dst = p.out.p[pl] + (long long)frame * p.out.frame_stride + (long long)row * p.out.pitch[pl];
    const bool row_inside = row >= p.cy[pl] && row < p.cy[pl] + p.ch[pl];
    const int x1 = min(x0 + 16, p.w[pl]);
    const uint32_t v = p.val[pl] * 0x01010101u;
    if ((!row_inside || x1 <= p.cx[pl] || x0 >= p.cx[pl] + p.cw[pl]) && x1 - x0 == 16 && ((p.out.pitch[pl] | (uintptr_t)p.out.p[pl]) & 15) == 0 &&
        (p.out.frame_stride & 15) == 0) {
        *reinterpret_cast<uint4 *>(dst + x0) = make_uint4(v, v, v, v);
        return;
    }
    for (int x = x0; x < x1; x++)
        if (!row_inside || x < p.cx[pl] || x >= p.cx[pl] + p.cw[pl]) dst[x] = (uint8_t)v;
}

}  // namespace

cudaError_t launch_border_fill(const BorderParams &p, cudaStream_t st)
{
    if (p.n_planes < 1 || p.n_planes > 3 || p.n_frames < 1) return cudaErrorInvalidValue;
    long long most = 0;
    for (int i = 0; i < p.n_planes; i++) most = std::max(most, (long long)((p.w[i] + 15) >> 4) * p.h[i]);
    dim3 grid((unsigned)((most + 255) / 256), p.n_planes, p.n_frames);
    border_fill_kernel<<<grid, 256, 0, st>>>(p);
    return cudaGetLastError();
}

size_t plane_tile_smem_bytes(const PlaneTileParams &p, int fmt)
{
    const size_t luma = (size_t)p.k[0].bwp * (p.k[0].bh + 1);
    const size_t chroma = (size_t)p.k[1].bwp * (p.k[1].bh + 1) * (fmt == 1 ? 2 : 1);
    return (luma > chroma ? luma : chroma) + 16;             // a clamped tap reads one byte past its window line
}

cudaError_t plane_tile_prepare(size_t optin_bytes)
{
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, k2_plane_tile_kernel<0>);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k2_plane_tile_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(optin_bytes - a.sharedSizeBytes));
    if (e != cudaSuccess) return e;
    e = cudaFuncGetAttributes(&a, k2_plane_tile_kernel<1>);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k2_plane_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(optin_bytes - a.sharedSizeBytes));
}

cudaError_t launch_plane_tile(const PlaneTileParams &p, int fmt, cudaStream_t st)
{
    if (p.n_frames <= 0 || p.k[0].tiles_x <= 0 || p.k[0].tiles_y <= 0)
        return cudaErrorInvalidValue;
    const int tiles_x = p.k[0].tiles_x > p.k[1].tiles_x ? p.k[0].tiles_x : p.k[1].tiles_x;
    dim3 grid(tiles_x, p.k[0].tiles_y + (fmt == 0 ? 2 : 1) * p.k[1].tiles_y, p.n_frames);
    const size_t smem = plane_tile_smem_bytes(p, fmt);
    if (fmt == 0)
        k2_plane_tile_kernel<0><<<grid, 256, smem, st>>>(p);
    else
        k2_plane_tile_kernel<1><<<grid, 256, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace b200vc
