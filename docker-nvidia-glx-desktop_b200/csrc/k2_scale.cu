// k2_scale.cu — K2: BGRx -> I420/NV12 with the bilinear scale fused into the same pass (sm_100a).
//
// Replaces `videoconvert ! videoscale` (GstVideoScale -> GstVideoScaler 2-tap "lq" path) on the
// reference's software-encoder fallback (/root/reference Dockerfile:467-471, README.md:21).
// Semantics, SURVEY.md Appendix B.1: conversion and chroma downsampling happen at the INPUT size,
// then every plane is scaled on its own (horizontal 2-tap with an 8-bit fraction, then vertical
// 2-tap with a 6-bit tap). The fused kernel reproduces that two-stage result bit for bit without
// ever writing the full-size I420 intermediate to HBM:
//
//   phase 0  one thread issues a TMA tiled load (cp.async.bulk.tensor.3d) of the tile's source
//            window — the rows and columns its output pixels depend on, halo included; pixels
//            outside the frame are zero-filled by the TMA unit — and bulk copies of the tile's
//            host-made tap records; the CTA then waits on one mbarrier for all the bytes
//   phase 1  convert the window: Y per pixel and the 4:2:0 (u,v) pair per 2x2 block, kept in
//            shared memory (this is the "intermediate frame", tile-sized)
//   phase 2  scale out of shared memory (a lane owns two adjacent output columns of luma or one
//            (u,v) sample, word loads, all taps are dp4a weight words) into an output tile, which
//            takes the place of the window's pixels in shared memory
//   phase 3  the tile leaves by TMA stores (device planes whose rows and pitches are multiples of
//            16 bytes) or as 16-byte stores
// (k2_fused_scale_kernel, any ratio whose window fits a TMA box). When the ratio is exactly 2:1,
// k2_fused_2x_kernel does all of it in registers straight from the TMA tile. The file also holds the
// gather form of the two-pass path's plane scaler (its tiled form is k2_plane_tile.cu).
//
// HBM roofline: algorithmic bytes per output frame = in_w*in_h*4 + out_w*out_h*1.5.
#include "device_utils.cuh"

#include <type_traits>
#include "../../include/b200vc_spec.h"

namespace b200vc {

namespace {

constexpr int TW = K2_TILE_W;

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

// Scaler arithmetic used below. Horizontal, video_scale_h_2tap_1u8 (lq):
// (a*(256-f) + b*f) >> 8, applied to two lines at once on 16-bit lanes. Vertical,
// video_scale_v_2tap_u8 (lq): a + (((b-a)*p1 + 32) >> 6), which equals (a*(64-p1) + b*p1 + 32) >> 6
// (no clamp needed for 0 <= p1 <= 64) and is evaluated as one dp4a on the lanes' high bytes.

// mbarrier init + TMA tiled load of the tile's source window + bulk copies of the tile's tap records (one
// per tile column, one per tile row; both 16-byte multiples at 16-byte aligned addresses), all completing
// on the one mbarrier, by one thread. Kept out of line: inlined, the sequence's uniform-datapath
// instructions are stepped through by all eight warps (~40 issue slots each); as a call only the calling
// warp pays.
struct TileLoad {
    void *window;  const CUtensorMap *map;  int x, y, z;  uint32_t window_bytes;
    void *col_dst;  const void *col_src;  uint32_t col_bytes;
    void *row_dst;  const void *row_src;  uint32_t row_bytes;
};
__device__ __noinline__ void issue_tile_load(uint64_t *s_bar, const TileLoad t)
{
    const uint32_t bar = smem_u32(s_bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(t.window_bytes + t.col_bytes + t.row_bytes) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_u32(t.window)), "l"(reinterpret_cast<uint64_t>(t.map)), "r"(t.x), "r"(t.y), "r"(t.z), "r"(bar)
        : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(t.col_dst)), "l"(reinterpret_cast<uint64_t>(t.col_src)), "r"(t.col_bytes), "r"(bar) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(t.row_dst)), "l"(reinterpret_cast<uint64_t>(t.row_src)), "r"(t.row_bytes), "r"(bar) : "memory");
}

// The general fused kernel's output tile, shared memory -> the planes, as TMA tiled stores by one thread
// (out of line for the same reason). The caller has made the tile visible to the async proxy. The thread
// waits until the unit has read the tile: shared memory must outlive the reads.
__device__ __noinline__ void store_tile(const FusedOutMaps *om, const uint8_t *s_out, int fmt, int tile_h, int x0, int y0, int z)
{
    const uint32_t sy = smem_u32(s_out), sc1 = sy + TW * tile_h, sc2 = sc1 + (TW / 2) * (tile_h / 2);
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                 ::"l"(reinterpret_cast<uint64_t>(&om->y)), "r"(x0), "r"(y0), "r"(z), "r"(sy) : "memory");
    if (fmt == 0) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                     ::"l"(reinterpret_cast<uint64_t>(&om->c1)), "r"(x0 >> 1), "r"(y0 >> 1), "r"(z), "r"(sc1) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                     ::"l"(reinterpret_cast<uint64_t>(&om->c2)), "r"(x0 >> 1), "r"(y0 >> 1), "r"(z), "r"(sc2) : "memory");
    } else {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                     ::"l"(reinterpret_cast<uint64_t>(&om->c1)), "r"(x0), "r"(y0 >> 1), "r"(z), "r"(sc1) : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// Odd frames, edge tiles of the general fused kernel: window column dupx / line dupy (the ones just past
// the frame, zero-filled by the TMA unit) become copies of the column / line before them.
template <typename PX>
__device__ __forceinline__ void repeat_last_column_and_line(PX *s_src, int bw, int bh, int dupx, int dupy, int odd_w,
                                                         int odd_h, int tid)
{
    const bool fixx = odd_w && dupx > 0 && dupx < bw, fixy = odd_h && dupy > 0 && dupy < bh;
    if (fixx)
        for (int r = tid; r < bh; r += 256)
            s_src[r * bw + dupx] = s_src[((fixy && r == dupy) ? dupy - 1 : r) * bw + dupx - 1];
    if (fixy)
        for (int c = tid; c < bw; c += 256)
            if (!(fixx && c == dupx)) s_src[dupy * bw + c] = s_src[(dupy - 1) * bw + c];
}

// Phase 1 of the general fused kernel: one item = 4 columns x 2 lines of the window, pixels at ra (second
// line bw pixels on; PX = uint32_t: 4-byte pixels, uint16_t: 5-6-5 pixels, unpacked here) -> Y bytes at ya and
// two filtered, vertically averaged (u,v) byte pairs at cf. `left`: the group has a left chroma neighbour in
// the window (not frame column 0, not the window's first group); e0 / e2: its first / second even column is
// the frame's last one (B200VC_SPEC_COSITED_LAST_1_3).
__device__ __forceinline__ uint4 load_px4(const uint32_t *p) { return *reinterpret_cast<const uint4 *>(p); }
__device__ __forceinline__ uint4 load_px4(const uint16_t *p) { return unpack565x4(*reinterpret_cast<const uint2 *>(p)); }
__device__ __forceinline__ uint32_t load_px(const uint32_t *p) { return *p; }
__device__ __forceinline__ uint32_t load_px(const uint16_t *p) { return unpack565(*p); }

template <int SITE, bool NEG, bool EDGE, typename PX>
__device__ __forceinline__ void convert_item(const PX *ra, uint8_t *ya, uint8_t *cf, int bw, bool left, bool e0, bool e2,
                                             const Matrix &m)
{
    const PX *rb = ra + bw;
    const uint4 a = load_px4(ra);
    const uint4 b = load_px4(rb);
    *reinterpret_cast<uint32_t *>(ya) = pack_luma4(luma_sum(a.x, m), luma_sum(a.y, m), luma_sum(a.z, m), luma_sum(a.w, m));
    *reinterpret_cast<uint32_t *>(ya + bw) = pack_luma4(luma_sum(b.x, m), luma_sum(b.y, m), luma_sum(b.z, m), luma_sum(b.w, m));

    const uint32_t pa0 = chroma_pair<NEG>(a.x, m), pa1 = chroma_pair<NEG>(a.y, m);
    const uint32_t pa2 = chroma_pair<NEG>(a.z, m), pa3 = chroma_pair<NEG>(a.w, m);
    const uint32_t pb0 = chroma_pair<NEG>(b.x, m), pb1 = chroma_pair<NEG>(b.y, m);
    const uint32_t pb2 = chroma_pair<NEG>(b.z, m), pb3 = chroma_pair<NEG>(b.w, m);
    // The 2x16-bit lane results below keep whatever the shifts drag from the v lane into the top bits
    // of the u lane wherever it cannot reach a byte that is used: one operand of the vertical mean is
    // masked (so the sum cannot carry into the v lane), the other and the mean itself are not, because
    // only bytes 0 and 2 of the mean are stored.
    uint32_t c0, c1;
    if (SITE == 0) {
        c0 = (avg2_raw(pa0, pa1) + avg2(pb0, pb1) + 0x00010001u) >> 1;
        c1 = (avg2_raw(pa2, pa3) + avg2(pb2, pb3) + 0x00010001u) >> 1;
    } else {
        // frame column 0 takes itself; the window's first group has no left neighbour (its first
        // chroma column is halo and is never consumed)
        uint32_t la = pa0, lb = pb0;
        if (left) {
            la = chroma_pair<NEG>(load_px(ra - 1), m);
            lb = chroma_pair<NEG>(load_px(rb - 1), m);
        }
        // right neighbours: the frame's last even column (in_w - 2, in_w > 2) takes itself
        // (B200VC_SPEC_COSITED_LAST_1_3 = 0: the plain 1-2-1 filter runs to the end; an odd frame's
        // last pixel has been duplicated into the window by then, which is the same thing)
        c0 = (filt121_raw(la, pa0, (EDGE && e0) ? pa0 : pa1) + filt121(lb, pb0, (EDGE && e0) ? pb0 : pb1) + 0x00010001u) >> 1;
        c1 = (filt121_raw(pa1, pa2, (EDGE && e2) ? pa2 : pa3) + filt121(pb1, pb2, (EDGE && e2) ? pb2 : pb3) + 0x00010001u) >> 1;
    }
    *reinterpret_cast<uint32_t *>(cf) = prmt(c0, c1, 0x6420);              // u0 v0 u1 v1
}

template <int FMT, int SITE, bool NEG, int TILE_H, bool ODD, bool IN16>
__global__ void __launch_bounds__(256, 8)
k2_fused_scale_kernel(const __grid_constant__ CUtensorMap src_map, const __grid_constant__ FusedOutMaps out_maps,
                      const FusedScaleParams p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t s_bar;
    // this tile's taps, ready to use (FusedTaps): bulk copies of the host-made records, in flight with the
    // window's TMA load and completing on the same mbarrier
    __shared__ __align__(16) uint32_t s_col[4 * TW];           // luma records (two forms, FusedTaps) | hc[TW/2] quads
    __shared__ __align__(16) uint32_t s_row[3 * TILE_H];       // vy[TILE_H] pairs | vc[TILE_H/2] pairs
    const uint4 *s_hc = reinterpret_cast<const uint4 *>(s_col + 2 * TW);
    const uint2 *s_vy = reinterpret_cast<const uint2 *>(s_row), *s_vc = s_vy + TILE_H;

    const int bw = p.box_w, bh = p.box_h;
    using PX = typename std::conditional<IN16, uint16_t, uint32_t>::type;
    PX *s_src = reinterpret_cast<PX *>(smem);                             // [bh][bw] pixels (4-byte, or 5-6-5)
    uint8_t *s_out = smem;                                                // output tile: Y, then U and V (or UV) tiles; written
                                                                          // in phase 2, when the pixels are dead
    uint8_t *s_yf = smem + p.planes_off;                                  // [bh][bw] luma: after the larger of the two
    uint8_t *s_cf = s_yf + bw * bh;                                       // [bh/2][bw/2] (u,v) byte pairs

    const int tid = threadIdx.x;

    // ---- phase 0: source window origin (host-made, per tile column / tile row), TMA load --------
    const int2 org_x = p.ft.tile_x[blockIdx.x];          // window's frame column (multiple of 4), kfirst
    const int2 org_y = p.ft.tile_y[blockIdx.y];          // window's frame line (even), line pairs in use
    const int sx0 = org_x.x, kfirst = org_x.y, sy0 = org_y.x;
    if (tid == 0) {
        if (p.gate.flags) {                               // frame still arriving: wait for the window's last line
            gate_wait_thread(p.gate, min(max(sy0 + bh, 1), p.in_h) - 1);
            asm volatile("fence.proxy.async;" ::: "memory");
        }
        TileLoad t;
        t.window = s_src;  t.map = &src_map;  t.x = sx0;  t.y = sy0;  t.z = (int)blockIdx.z;  t.window_bytes = (uint32_t)bw * bh * (uint32_t)sizeof(PX);
        t.col_dst = s_col;  t.col_src = p.ft.col_rec + blockIdx.x * (4 * TW);  t.col_bytes = 4 * TW * 4;
        t.row_dst = s_row;  t.row_src = p.ft.row_rec + blockIdx.y * (3 * TILE_H);  t.row_bytes = 3 * TILE_H * 4;
        issue_tile_load(&s_bar, t);
    }
    __syncthreads();                                       // (the mbarrier exists)
    {
        const uint32_t bar = smem_u32(&s_bar);
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred q;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, q;\n\t}"
                : "=r"(done) : "r"(bar), "r"(0u) : "memory");
        }
    }

    // ---- odd frames: the pixel column / line just past the frame repeats the last one --------------
    // (an odd frame's last pixel and last line are chroma sites of their own: h2 averages it with itself,
    // the cosited filter gives it l + 3c, the vertical mean pairs the line with itself — SURVEY A.4.
    // The TMA unit zero-filled those cells; after this, phase 1 needs no special case.)
    if (ODD) {                                         // instantiated apart: even frames pay nothing
        repeat_last_column_and_line(s_src, bw, bh, p.in_w - sx0, p.in_h - sy0, p.in_w & 1, p.in_h & 1, tid);
        __syncthreads();
    }

    // ---- phase 1: convert the window (4 columns x 2 lines per item) ---------------------------------
    // A thread keeps its 4-column group and walks down the window: the 256 threads form a block of
    // (groups per line pair) x (256 / groups) items that sweeps the window top to bottom, so everything
    // but three addresses is loop-invariant (no divisions: host-made reciprocal of the group count).
    // Only the line pairs this tile's taps reach are converted (the box is the tallest window of any
    // tile), and a first group that holds nothing but halo is skipped. Only tiles whose window reaches
    // the frame's last column pay for the right-edge rule of the cosited filter (uniform branch).
    {
        const int items_x = (bw >> 2) - kfirst, pairs = org_y.y;
        const uint32_t inv = kfirst ? p.inv_items_x1 : p.inv_items_x;
        const int rps = (int)((256u * inv) >> 16);                          // line pairs per sweep: 256 / items_x
        const int ky0 = (int)(((uint32_t)tid * inv) >> 16), kx = tid - ky0 * items_x + kfirst;
        if (ky0 < rps) {
            const int ac = sx0 + 4 * kx;                                    // frame column of the group (multiple of 4)
            const bool left = SITE != 0 && ac != 0 && kx > 0;
            const bool redge = SITE != 0 && B200VC_SPEC_COSITED_LAST_1_3 && p.in_w > 2 && sx0 + bw >= p.in_w;
            const bool e0 = redge && ac == p.in_w - 2, e2 = redge && ac + 2 == p.in_w - 2;
            const int o = 2 * ky0 * bw + 4 * kx, step = rps * bw;
            const PX *ra = s_src + o;
            uint8_t *ya = s_yf + o;
            uint8_t *cf = s_cf + ky0 * bw + 4 * kx;                         // two (u,v) byte pairs per item
            // (not unrolled: the unrolled form computes its trip count with a division, per warp and tile
            // more than an item's worth of instructions)
            if (redge) {
#pragma unroll 1
                for (int ky = ky0; ky < pairs; ky += rps, ra += 2 * step, ya += 2 * step, cf += step)
                    convert_item<SITE, NEG, true, PX>(ra, ya, cf, bw, left, e0, e2, p.mx);
            } else {
#pragma unroll 1
                for (int ky = ky0; ky < pairs; ky += rps, ra += 2 * step, ya += 2 * step, cf += step)
                    convert_item<SITE, NEG, false, PX>(ra, ya, cf, bw, left, false, false, p.mx);
            }
        }
    }
    __syncthreads();

    // ---- phase 2: scale out of the planes into an output tile in shared memory ----------------
    // Taps come ready to use (FusedTaps, staged above as offsets into the planes): nothing is decoded
    // here. Both taps are always read (a clamped one has weight 0; the byte or line after the window
    // is inside this CTA's shared memory, see fused_scale_smem_bytes). Every sample of the tile is
    // computed (the staged taps of columns and rows past the frame edge are clamped copies); the frame
    // edge is applied when the tile is stored. A warp works on whole rows (the row's vertical tap is a
    // broadcast load). Round 1's form of this phase (four byte loads and a byte store per luma sample)
    // kept the LSU data pipe at 89 % (profiles/r1_k2general_4to3_ncu_full_raw_v4.csv); word loads shared
    // by a lane's two columns halve that. Keeping the lower line's horizontal pass for the next row, which
    // at ratios below 2 usually starts on it, saves loads but costs more issue slots in branches than
    // it saves in arithmetic (measured, profiles/r2_general_scaler_experiments.md) and is not done.
    const int lane = tid & 31, wrp = tid >> 5;
    if (p.pair_luma) {
        // luma: two adjacent output columns per lane, rows wrp * RPW ...
        constexpr int RPW = TILE_H / 8;
        const uint4 rec = reinterpret_cast<const uint4 *>(s_col)[lane];   // {window byte of the first word | shift << 16, sel1, wh0, wh1}
        const uint8_t *col = s_yf + (rec.x & 0xffffu);
        const uint32_t sh = rec.x >> 16;
        uint16_t *dst = reinterpret_cast<uint16_t *>(s_out) + wrp * RPW * (TW / 2) + lane;
#pragma unroll(RPW < 4 ? RPW : 4)
        for (int k = 0; k < RPW; k++) {
            const uint2 tv = s_vy[wrp * RPW + k]; // {byte offset of the upper line, vertical weights}
            const uint8_t *l0 = col + (int)tv.x;
            const uint2 top = h_pass_luma_pair(l0, sh, rec.y, rec.z, rec.w);
            const uint2 bot = h_pass_luma_pair(l0 + bw, sh, rec.y, rec.z, rec.w);
            dst[k * (TW / 2)] = (uint16_t)v_pass_pair(top, bot, tv.y);
        }
    } else {
        // luma, any step (ratios above 2, BOX): one output column x TILE_H/4 rows per thread, 8 warps =
        // 2 column halves x 4 row groups; byte loads
        const uint32_t *s_hyo = s_col, *s_hyw = s_col + TW;
        constexpr int RPT = TILE_H / 4;
        const int c = (wrp & 1) * 32 + lane, rb = (wrp >> 1) * RPT;
        const uint32_t wh = s_hyw[c];
        const uint8_t *col = s_yf + (int)s_hyo[c];
#pragma unroll
        for (int r = rb; r < rb + RPT; r++) {
            const uint2 tv = s_vy[r];
            const uint8_t *r0 = col + (int)tv.x, *r1 = r0 + bw;
            uint32_t d;
            if (p.box2x2) {
                d = (uint32_t)(r0[0] + r0[1] + r1[0] + r1[1] + 2) << 6;
            } else {
                // [a,a,b,_] . (255-f, 1, f, 0) per line, then [A,A,B,B] . (wa1,wa2,wb1,wb2) + 128 on
                // the two sums' high bytes; the sample is byte 1 of the result
                const uint32_t s0 = dp4a_uu(prmt(r0[0], r0[1], 0x1400), wh, 0u);
                const uint32_t s1 = dp4a_uu(prmt(r1[0], r1[1], 0x1400), wh, 0u);
                d = dp4a_uu(prmt(s0, s1, 0x5511), tv.y, 128u);
            }
            s_out[r * TW + c] = (uint8_t)(d >> 8);
        }
    }
    {
        // chroma: one (u,v) sample per lane, warp w does chroma rows w * CPW ... (while there are rows)
        constexpr int CPW = (TILE_H / 2 + 7) / 8;
        const uint4 th = s_hc[lane];              // {byte offset of the first tap in a (u,v) line, wh, -, -}
        const uint8_t *col = s_cf + (int)th.x;
        uint8_t *oc = s_out + TW * TILE_H;
#pragma unroll(CPW < 4 ? CPW : 4)
        for (int r = wrp * CPW; r < wrp * CPW + CPW; r++) {
            if (r >= TILE_H / 2) break;
            const uint2 tv = s_vc[r];
            const uint8_t *l0 = col + (int)tv.x;
            uint32_t uv;                          // u in byte 0, v in byte 1
            if (p.box2x2) {
                const uint32_t a0 = *reinterpret_cast<const uint16_t *>(l0), a1 = *reinterpret_cast<const uint16_t *>(l0 + 2);
                const uint32_t b0 = *reinterpret_cast<const uint16_t *>(l0 + bw), b1 = *reinterpret_cast<const uint16_t *>(l0 + bw + 2);
                const uint32_t q = (a0 & 0xffu) + (a1 & 0xffu) + (b0 & 0xffu) + (b1 & 0xffu) + 2u;          // u
                const uint32_t w = (a0 >> 8) + (a1 >> 8) + (b0 >> 8) + (b1 >> 8) + 2u;                      // v
                uv = (q >> 2) | ((w >> 2) << 8);
            } else {
                uv = v_pass_pair(h_pass_chroma(l0, th.y), h_pass_chroma(l0 + bw, th.y), tv.y);
            }
            if (FMT == 0) {
                oc[r * (TW / 2) + lane] = (uint8_t)uv;                                   // U tile, then V tile
                oc[(TW / 2) * (TILE_H / 2) + r * (TW / 2) + lane] = (uint8_t)(uv >> 8);
            } else {
                reinterpret_cast<uint16_t *>(oc)[r * (TW / 2) + lane] = (uint16_t)uv;    // u, v
            }
        }
    }
    // ---- phase 3: the tile leaves --------------------------------------------------------------------
    // by TMA stores where the planes allow it (device memory, every pitch a multiple of 16): one thread
    // issues them, the unit clips tiles that overhang the frame. Otherwise as 16-byte stores, 4 chunks
    // per luma row, 4 per chroma row pair.
    if (p.tma_out) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // this thread's part of the tile, for the TMA unit
        __syncthreads();
        if (tid == 0)
            store_tile(&out_maps, s_out, FMT, TILE_H, (int)blockIdx.x * TW, (int)blockIdx.y * TILE_H, (int)blockIdx.z);
        return;
    }
    __syncthreads();
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TILE_H;
    const long long fo = (long long)blockIdx.z * p.out.frame_stride;
#pragma unroll
    for (int c = tid; c < TILE_H * 6; c += 256) {
        const uint8_t *src = s_out + c * 16;
        uint8_t *dst;
        int n;                                    // valid bytes of this chunk, <= 0: none
        bool row_ok;
        if (c < TILE_H * 4) {
            const int r = c >> 2, xb = x0 + (c & 3) * 16;
            row_ok = y0 + r < p.out_h;
            n = p.out_w - xb;
            dst = p.out.p[0] + fo + (long long)(y0 + r) * p.out.pitch[0] + xb;
        } else if (FMT == 0) {
            const int i = c - TILE_H * 4;
            const int pl = i / TILE_H, r = (i - pl * TILE_H) >> 1, xb = (x0 >> 1) + (c & 1) * 16;
            row_ok = (y0 >> 1) + r < ((p.out_h + 1) >> 1);
            n = ((p.out_w + 1) >> 1) - xb;
            dst = p.out.p[1 + pl] + fo + (long long)((y0 >> 1) + r) * p.out.pitch[1 + pl] + xb;
        } else {
            const int r = (c - TILE_H * 4) >> 2, xb = x0 + (c & 3) * 16;
            row_ok = (y0 >> 1) + r < ((p.out_h + 1) >> 1);
            n = 2 * ((p.out_w + 1) >> 1) - xb;
            dst = p.out.p[1] + fo + (long long)((y0 >> 1) + r) * p.out.pitch[1] + xb;
        }
        if (row_ok) {
            if (n >= 16 && p.out.aligned16) {
                *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(src);
            } else {
                for (int k = 0; k < min(n, 16); k++) dst[k] = src[k];
            }
        }
    }
}


// ---- K2, exact 2:1 ------------------------------------------------------------------------------
// When every horizontal tap starts at source index 2x and every vertical tap is the rounded mean of
// lines 2y, 2y+1 (an exact 2:1 downscale up to ~3970 output columns; 4K -> 1080p is the case that
// matters), one thread's 4x4 source block plus one halo column holds everything two output line
// pairs need: 2x2 luma outputs and one (u,v) output. The tile still arrives by TMA (132 x 32 box,
// halo columns included, zero fill outside the frame), but the intermediate frame never exists
// even in shared memory: conversion, both chroma roundings, the fractional horizontal tap and the
// vertical mean all run in registers on 2x16-bit lanes.
template <int FMT, int SITE, bool NEG, bool BOX>
__global__ void __launch_bounds__(256)
k2_fused_2x_kernel(const __grid_constant__ CUtensorMap src_map, const Fused2xParams p)
{
    constexpr int BW = K2X_SRC_W + K2X_HALO;               // 132
    constexpr int NT = K2X_TILES_PER_CTA;                  // vertically adjacent tiles per CTA, all loads issued up front
    __shared__ __align__(128) uint32_t s_tiles[NT][K2X_SRC_H][BW];
    __shared__ __align__(8) uint64_t s_bar[NT];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int frame = blockIdx.z;
    const int tx = blockIdx.x * K2X_SRC_W, ty0 = blockIdx.y * (K2X_SRC_H * NT);   // source origin of the first tile

    if (threadIdx.x == 0) {
#pragma unroll
        for (int t = 0; t < NT; t++)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[t])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
        for (int t = 0; t < NT; t++) {
            if (ty0 + t * K2X_SRC_H < p.in_h) {
                if (p.gate.flags) {                       // frame still arriving: wait for the tile's last line
                    gate_wait_thread(p.gate, min(ty0 + (t + 1) * K2X_SRC_H, p.in_h) - 1);
                    asm volatile("fence.proxy.async;" ::: "memory");
                }
                const uint32_t bar = smem_u32(&s_bar[t]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                             ::"r"(bar), "r"((uint32_t)sizeof(s_tiles[0])) : "memory");
                asm volatile(
                    "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
                    " [%0], [%1, {%2, %3, %4}], [%5];"
                    ::"r"(smem_u32(&s_tiles[t][0][0])), "l"(reinterpret_cast<uint64_t>(&src_map)), "r"(tx - K2X_HALO),
                      "r"(ty0 + t * K2X_SRC_H), "r"(frame), "r"(bar) : "memory");
            }
        }
    }
    __syncthreads();
#pragma unroll 1
    for (int t = 0; t < NT; t++) {
    const int ty = ty0 + t * K2X_SRC_H;
    if (ty >= p.in_h) break;
    const uint32_t (*s_src)[BW] = s_tiles[t];
    {
        const uint32_t bar = smem_u32(&s_bar[t]);
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred q;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, q;\n\t}"
                : "=r"(done) : "r"(bar), "r"(0u) : "memory");
        }
    }

    const Matrix &m = p.mx;
    const int c = tx + lane * 4;                           // absolute source column of this thread's pixel 0
    const int r = ty + warp * 4;                           // absolute source line of this thread's line 0
    uint4 a[4];
#pragma unroll
    for (int i = 0; i < 4; i++)
        a[i] = *reinterpret_cast<const uint4 *>(&s_src[warp * 4 + i][K2X_HALO + lane * 4]);

    // ---- chroma: 4:4:4 pairs, horizontal filter per line, vertical means, 2:1 scale --------------
    uint32_t h0[4], h1[4];                                 // filtered chroma at columns c and c+2, per line
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t q0 = chroma_pair<NEG>(a[i].x, m), q1 = chroma_pair<NEG>(a[i].y, m);
        const uint32_t q2 = chroma_pair<NEG>(a[i].z, m), q3 = chroma_pair<NEG>(a[i].w, m);
        if (SITE == 0) {
            h0[i] = avg2(q0, q1);
            h1[i] = avg2(q2, q3);
        } else {
            uint32_t l = __shfl_up_sync(0xffffffffu, q3, 1);
            if (lane == 0)                                 // tile halo column, or (frame column 0) itself
                l = c > 0 ? chroma_pair<NEG>(s_src[warp * 4 + i][K2X_HALO - 1], m) : q0;
            const uint32_t rr = (B200VC_SPEC_COSITED_LAST_1_3 && c + 4 == p.in_w) ? q2 : q3;   // last even column takes itself (FILT_1_3)
            h0[i] = filt121(l, q0, q1);
            h1[i] = filt121(q1, q2, rr);
        }
    }
    // intermediate chroma plane samples: lines (r/2, r/2+1) x columns (c/2, c/2+1)
    const uint32_t v00 = avg2(h0[0], h0[1]), v01 = avg2(h1[0], h1[1]);
    const uint32_t v10 = avg2(h0[2], h0[3]), v11 = avg2(h1[2], h1[3]);
    uint32_t cuv;
    if (BOX) {
        cuv = ((v00 + v01 + v10 + v11 + 0x00020002u) >> 2) & LANES8;
    } else {
        const uint32_t f = (((uint32_t)(c >> 2) * p.inc_c) >> 8) & 0xffu;
        const uint32_t ha = v00 * (256u - f) + v01 * f;    // high bytes of the lanes: scaled u, v of line 0
        const uint32_t hb = v10 * (256u - f) + v11 * f;    // ... of line 1
        const uint32_t uv = prmt(ha, hb, 0x7351);          // u0 u1 v0 v1 (line 0 / line 1)
        cuv = (dp4a_uu(uv, 0x00000101u, 1u) >> 1) | ((dp4a_uu(uv, 0x01010000u, 1u) >> 1) << 16);
    }

    // ---- luma: lines (2i, 2i+1) ride in the two 16-bit lanes --------------------------------------
    uint32_t yout[2];
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        const uint4 t = a[2 * ii], b = a[2 * ii + 1];
        const uint32_t w0 = prmt(luma_sum(t.x, m), luma_sum(b.x, m), 0x6521);
        const uint32_t w1 = prmt(luma_sum(t.y, m), luma_sum(b.y, m), 0x6521);
        const uint32_t w2 = prmt(luma_sum(t.z, m), luma_sum(b.z, m), 0x6521);
        const uint32_t w3 = prmt(luma_sum(t.w, m), luma_sum(b.w, m), 0x6521);
        uint32_t d0, d1;
        if (BOX) {
            const uint32_t s0 = w0 + w1, s1 = w2 + w3;
            d0 = ((s0 & 0xffffu) + (s0 >> 16) + 2u) >> 2;
            d1 = ((s1 & 0xffffu) + (s1 >> 16) + 2u) >> 2;
        } else {
            const uint32_t ox = (uint32_t)(c >> 1);
            const uint32_t f0 = ((ox * p.inc_y) >> 8) & 0xffu, f1 = (((ox + 1u) * p.inc_y) >> 8) & 0xffu;
            // per 16-bit lane the weighted sum is < 65536 and its high byte is the horizontally
            // scaled pixel (the >> 8); dp4a with weights on bytes 1 and 3 adds the two lines + 1
            const uint32_t s0 = w0 * (256u - f0) + w1 * f0;
            const uint32_t s1 = w2 * (256u - f1) + w3 * f1;
            d0 = dp4a_uu(s0, 0x01000100u, 1u) >> 1;
            d1 = dp4a_uu(s1, 0x01000100u, 1u) >> 1;
        }
        yout[ii] = d0 | (d1 << 8);
    }

    // ---- stores ---------------------------------------------------------------------------------------
    if (c < p.in_w && r < p.in_h) {
        const long long fo = (long long)frame * p.out.frame_stride;
        const int ox = c >> 1, oy = r >> 1, cx = c >> 2, cy = r >> 2;
        *reinterpret_cast<uint16_t *>(p.out.p[0] + fo + (long long)oy * p.out.pitch[0] + ox) = (uint16_t)yout[0];
        *reinterpret_cast<uint16_t *>(p.out.p[0] + fo + (long long)(oy + 1) * p.out.pitch[0] + ox) = (uint16_t)yout[1];
        if (FMT == 0) {
            p.out.p[1][fo + (long long)cy * p.out.pitch[1] + cx] = (uint8_t)(cuv & 0xffu);
            p.out.p[2][fo + (long long)cy * p.out.pitch[2] + cx] = (uint8_t)(cuv >> 16);
        } else {
            *reinterpret_cast<uint16_t *>(p.out.p[1] + fo + (long long)cy * p.out.pitch[1] + 2 * cx) =
                (uint16_t)((cuv & 0xffu) | ((cuv >> 16) << 8));
        }
    }
    }   // tiles of this CTA
}

// ---- two-pass, pass 2: scale the planes of a full-size intermediate I420 -----------------------
// The intermediate was written by K1 a moment ago and is read out of L2. A thread owns four
// consecutive output samples of one line of one plane; the two source lines ride in the two 16-bit
// lanes of a word, so one multiply-add pair applies the horizontal tap to both.
// vertical 2-tap on a word holding line 0's sample in byte 1 and line 1's in byte 3:
// a + (((b-a)*p1 + half) >> bits) == (a*(one-p1) + b*p1 + half) >> bits, one dp4a
__device__ __forceinline__ uint32_t scale_v2(uint32_t hs, int p1)
{
    constexpr int ONE = B200VC_SPEC_VTAP_ONE, HALF = B200VC_SPEC_VTAP_HALF, BITS = B200VC_SPEC_VTAP_BITS;
    if (ONE < 256)                                            // weights fit a byte: bytes 1 and 3 as they are
        return dp4a_uu(hs, ((uint32_t)(ONE - p1) << 8) | ((uint32_t)p1 << 24), (uint32_t)HALF) >> BITS;
    // an 8-bit tap can be 256: each sample twice, the weight split over the two bytes
    const int a = ONE - p1, b = p1, a1 = min(a, 255), b1 = min(b, 255);
    const uint32_t w = (uint32_t)a1 << 8 | (uint32_t)(a - a1) | (uint32_t)b1 << 24 | (uint32_t)(b - b1) << 16;
    return dp4a_uu(prmt(hs, hs, 0x3311), w, (uint32_t)HALF) >> BITS;
}

__device__ __forceinline__ uint32_t scale4(const uint8_t *r0, const uint8_t *r1, const uint4 th4, int p1, int box)
{
    const Tap th[4] = { th4.x, th4.y, th4.z, th4.w };
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int j0 = tap_idx(th[k]), j1 = j0 + tap_step(th[k]);
        const uint32_t f = (uint32_t)tap_frac(th[k]);
        const uint32_t w0 = (uint32_t)__ldg(r0 + j0) | ((uint32_t)__ldg(r1 + j0) << 16);
        const uint32_t w1 = (uint32_t)__ldg(r0 + j1) | ((uint32_t)__ldg(r1 + j1) << 16);
        uint32_t d;
        if (box) {
            const uint32_t q = w0 + w1;
            d = ((q & 0xffffu) + (q >> 16) + 2u) >> 2;
        } else {
            if (B200VC_SPEC_SCALE_H_FIRST) {
                // per lane the sum is < 65536 and its high byte is the horizontally scaled pixel
                d = scale_v2(w0 * (256u - f) + w1 * f, p1);
            } else {                                          // vertical pass first (spec variant)
                d = (scale_v2(w0 << 8, p1) * (256u - f) + scale_v2(w1 << 8, p1) * f) >> 8;
            }
        }
        out |= d << (8 * k);
    }
    return out;
}

template <int FMT>
__global__ void __launch_bounds__(128) k2_plane_scale_kernel(const PlaneScaleParams p)
{
    const int x = (blockIdx.x * 128 + threadIdx.x) * 4;
    const int row = blockIdx.y;
    const long long fi = (long long)blockIdx.z * p.src_frame_stride;
    const long long fo = (long long)blockIdx.z * p.out.frame_stride;
    const int ch = p.c_out_h, cw = p.c_out_w;             // by format; odd frames: chroma planes round up

    if (row < p.out_h) {                                   // luma line
        if (x >= p.out_w) return;
        const Tap tv = p.tab.vy[row];
        const uint8_t *r0 = p.src[0] + fi + (long long)tap_idx(tv) * p.src_pitch[0];
        const uint8_t *r1 = r0 + (long long)tap_step(tv) * p.src_pitch[0];
        const uint32_t out = scale4(r0, r1, *reinterpret_cast<const uint4 *>(p.tab.hy + x), tap_frac(tv), p.box2x2);
        uint8_t *dst = p.out.p[0] + fo + (long long)row * p.out.pitch[0] + x;
        if (x + 4 <= p.out_w && p.out.aligned16) {
            *reinterpret_cast<uint32_t *>(dst) = out;
        } else {
            for (int k = 0; k < min(4, p.out_w - x); k++) dst[k] = (uint8_t)(out >> (8 * k));
        }
        return;
    }
    if (FMT == 0) {                                        // I420: lines of U, then lines of V
        const int pl = row < p.out_h + ch ? 1 : 2, cy = row - p.out_h - (pl - 1) * ch;
        if (cy >= ch || x >= cw) return;
        const Tap tv = p.tab.vc[cy];
        const uint8_t *r0 = p.src[pl] + fi + (long long)tap_idx(tv) * p.src_pitch[pl];
        const uint8_t *r1 = r0 + (long long)tap_step(tv) * p.src_pitch[pl];
        const uint32_t out = scale4(r0, r1, *reinterpret_cast<const uint4 *>(p.tab.hc + x), tap_frac(tv), p.box2x2);
        uint8_t *dst = p.out.p[pl] + fo + (long long)cy * p.out.pitch[pl] + x;
        if (x + 4 <= cw && p.out.aligned16) {
            *reinterpret_cast<uint32_t *>(dst) = out;
        } else {
            for (int k = 0; k < min(4, cw - x); k++) dst[k] = (uint8_t)(out >> (8 * k));
        }
    } else {                                               // NV12: U and V of four samples, interleaved
        const int cy = row - p.out_h;
        if (cy >= ch || x >= cw) return;
        const Tap tv = p.tab.vc[cy];
        const uint4 th4 = *reinterpret_cast<const uint4 *>(p.tab.hc + x);
        uint32_t o[2];
#pragma unroll
        for (int pl = 1; pl <= 2; pl++) {
            const uint8_t *r0 = p.src[pl] + fi + (long long)tap_idx(tv) * p.src_pitch[pl];
            const uint8_t *r1 = r0 + (long long)tap_step(tv) * p.src_pitch[pl];
            o[pl - 1] = scale4(r0, r1, th4, tap_frac(tv), p.box2x2);
        }
        const uint32_t lo = prmt(o[0], o[1], 0x5140), hi = prmt(o[0], o[1], 0x7362);   // u0 v0 u1 v1 | u2 v2 u3 v3
        uint8_t *dst = p.out.p[1] + fo + (long long)cy * p.out.pitch[1] + 2 * x;
        if (x + 4 <= cw && p.out.aligned16) {
            *reinterpret_cast<uint2 *>(dst) = make_uint2(lo, hi);
        } else {
            const unsigned long long both = (unsigned long long)lo | ((unsigned long long)hi << 32);
            for (int k = 0; k < 2 * min(4, cw - x); k++) dst[k] = (uint8_t)(both >> (8 * k));
        }
    }
}

template <int FMT, int SITE, bool NEG, bool ODD, bool IN16>
cudaError_t launch_fused5(const CUtensorMap &map, const FusedOutMaps &om, const FusedScaleParams &p, dim3 grid, size_t smem, cudaStream_t st)
{
    switch (p.tile_h) {                                   // K2_TILE_HEIGHTS
    case 8:  k2_fused_scale_kernel<FMT, SITE, NEG, 8, ODD, IN16><<<grid, 256, smem, st>>>(map, om, p);  break;
    case 16: k2_fused_scale_kernel<FMT, SITE, NEG, 16, ODD, IN16><<<grid, 256, smem, st>>>(map, om, p);  break;
    case 32: k2_fused_scale_kernel<FMT, SITE, NEG, 32, ODD, IN16><<<grid, 256, smem, st>>>(map, om, p);  break;
    case 48: k2_fused_scale_kernel<FMT, SITE, NEG, 48, ODD, IN16><<<grid, 256, smem, st>>>(map, om, p);  break;
    case 64: k2_fused_scale_kernel<FMT, SITE, NEG, 64, ODD, IN16><<<grid, 256, smem, st>>>(map, om, p);  break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

template <int FMT, int SITE, bool NEG, bool ODD>
cudaError_t launch_fused4(const CUtensorMap &map, const FusedOutMaps &om, const FusedScaleParams &p, dim3 grid, size_t smem, cudaStream_t st)
{
    return p.in16 ? launch_fused5<FMT, SITE, NEG, ODD, true>(map, om, p, grid, smem, st)
                  : launch_fused5<FMT, SITE, NEG, ODD, false>(map, om, p, grid, smem, st);
}

template <int FMT, int SITE>
cudaError_t launch_fused2(const CUtensorMap &map, const FusedOutMaps &om, const FusedScaleParams &p, dim3 grid, size_t smem, cudaStream_t st)
{
    const bool odd = (p.in_w | p.in_h) & 1;                // odd frames: the variant that repeats the last column / line
    if (p.mx.neg)
        return odd ? launch_fused4<FMT, SITE, true, true>(map, om, p, grid, smem, st) : launch_fused4<FMT, SITE, true, false>(map, om, p, grid, smem, st);
    return odd ? launch_fused4<FMT, SITE, false, true>(map, om, p, grid, smem, st) : launch_fused4<FMT, SITE, false, false>(map, om, p, grid, smem, st);
}

// Lets a kernel ask for dynamic shared memory up to the device's opt-in limit less its own static arrays.
template <typename K>
cudaError_t set_smem(K kernel, size_t optin_bytes)
{
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, kernel);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(optin_bytes - a.sharedSizeBytes));
}


template <int FMT, int SITE, bool NEG>
cudaError_t launch_2x3(const CUtensorMap &map, const Fused2xParams &p, int box, dim3 grid, cudaStream_t st)
{
    if (box)
        k2_fused_2x_kernel<FMT, SITE, NEG, true><<<grid, 256, 0, st>>>(map, p);
    else
        k2_fused_2x_kernel<FMT, SITE, NEG, false><<<grid, 256, 0, st>>>(map, p);
    return cudaGetLastError();
}

template <int FMT, int SITE>
cudaError_t launch_2x2(const CUtensorMap &map, const Fused2xParams &p, int box, dim3 grid, cudaStream_t st)
{
    return p.mx.neg ? launch_2x3<FMT, SITE, true>(map, p, box, grid, st) : launch_2x3<FMT, SITE, false>(map, p, box, grid, st);
}

}  // namespace

cudaError_t launch_fused_2x(const CUtensorMap &src_map, const Fused2xParams &p, int fmt, int site, int box2x2,
                            cudaStream_t st)
{
    if (p.in_w <= 0 || p.in_h <= 0 || p.n_frames <= 0 || (p.in_w & 3) || (p.in_h & 3))
        return cudaErrorInvalidValue;
    dim3 grid((p.in_w + K2X_SRC_W - 1) / K2X_SRC_W, (p.in_h + K2X_SRC_H * K2X_TILES_PER_CTA - 1) / (K2X_SRC_H * K2X_TILES_PER_CTA), p.n_frames);
    if (fmt == 0)
        return site ? launch_2x2<0, 1>(src_map, p, box2x2, grid, st) : launch_2x2<0, 0>(src_map, p, box2x2, grid, st);
    return site ? launch_2x2<1, 1>(src_map, p, box2x2, grid, st) : launch_2x2<1, 0>(src_map, p, box2x2, grid, st);
}

// offset of the Y plane in the kernel's dynamic shared memory: behind the window's pixels and behind
// the output tile, which takes the pixels' place in phase 2 (an upscale's window can be the smaller one)
size_t fused_scale_planes_off(int box_w, int box_h, int tile_h, int px_bytes)
{
    const size_t window = (size_t)box_w * box_h * px_bytes, tile = (size_t)TW * tile_h * 3 / 2;
    return ((window > tile ? window : tile) + 15) & ~(size_t)15;
}

size_t fused_scale_smem_bytes(int box_w, int box_h, int tile_h, int px_bytes)
{
    // source pixels (4 bytes) + Y plane (1) + (u,v) plane (a byte pair per 2x2 pixels), and one spare
    // chroma line: a clamped second tap still reads index+1 (with weight 0), which for the last chroma
    // line lies just behind the plane
    return fused_scale_planes_off(box_w, box_h, tile_h, px_bytes) + (size_t)box_w * box_h * 3 / 2 + (size_t)box_w * 2 + 16;
}

// the kernel's static shared memory for tiles of tile_h rows: tap records, mbarrier
size_t fused_scale_static_bytes(int tile_h)
{
    return 12 * (size_t)tile_h + 4 * 4 * TW + 16 + 64;      // (+ alignment gaps)
}

cudaError_t fused_scale_prepare(size_t optin_bytes)
{
    cudaError_t e;
#define B200VC_SET1(F, S, N, H) \
    if ((e = set_smem(k2_fused_scale_kernel<F, S, N, H, false, false>, optin_bytes)) != cudaSuccess) return e; \
    if ((e = set_smem(k2_fused_scale_kernel<F, S, N, H, true, false>, optin_bytes)) != cudaSuccess) return e; \
    if ((e = set_smem(k2_fused_scale_kernel<F, S, N, H, false, true>, optin_bytes)) != cudaSuccess) return e; \
    if ((e = set_smem(k2_fused_scale_kernel<F, S, N, H, true, true>, optin_bytes)) != cudaSuccess) return e;
#define B200VC_SET(F, S, N) B200VC_SET1(F, S, N, 8) B200VC_SET1(F, S, N, 16) B200VC_SET1(F, S, N, 32) B200VC_SET1(F, S, N, 48) \
    B200VC_SET1(F, S, N, 64)
    B200VC_SET(0, 0, false) B200VC_SET(0, 0, true) B200VC_SET(0, 1, false) B200VC_SET(0, 1, true)
    B200VC_SET(1, 0, false) B200VC_SET(1, 0, true) B200VC_SET(1, 1, false) B200VC_SET(1, 1, true)
#undef B200VC_SET
#undef B200VC_SET1
    return cudaSuccess;
}

cudaError_t launch_fused_scale(const CUtensorMap &src_map, const FusedOutMaps &out_maps, const FusedScaleParams &p, int fmt, int site,
                               cudaStream_t st)
{
    if (p.out_w <= 0 || p.out_h <= 0 || p.n_frames <= 0)
        return cudaErrorInvalidValue;
    if (p.tile_h < 8 || p.tile_h > K2_MAX_TILE_H || (p.tile_h & 7))
        return cudaErrorInvalidValue;
    if (p.box_h > 256 || p.box_w > 256)                    // TMA box limit; and phase 1 divides by up to 64 groups
        return cudaErrorInvalidValue;
    dim3 grid((p.out_w + TW - 1) / TW, (p.out_h + p.tile_h - 1) / p.tile_h, p.n_frames);
    const int px_bytes = p.in16 ? 2 : 4;
    if ((p.box_w * px_bytes) & 15)                         // a TMA box line is a multiple of 16 bytes
        return cudaErrorInvalidValue;
    const size_t smem = fused_scale_smem_bytes(p.box_w, p.box_h, p.tile_h, px_bytes);
    if ((size_t)p.planes_off != fused_scale_planes_off(p.box_w, p.box_h, p.tile_h, px_bytes))
        return cudaErrorInvalidValue;
    if (fmt == 0)
        return site ? launch_fused2<0, 1>(src_map, out_maps, p, grid, smem, st) : launch_fused2<0, 0>(src_map, out_maps, p, grid, smem, st);
    return site ? launch_fused2<1, 1>(src_map, out_maps, p, grid, smem, st) : launch_fused2<1, 0>(src_map, out_maps, p, grid, smem, st);
}

cudaError_t launch_plane_scale(const PlaneScaleParams &p, int fmt, cudaStream_t st)
{
    if (p.out_w <= 0 || p.out_h <= 0 || p.n_frames <= 0)
        return cudaErrorInvalidValue;
    // one block row per output line: luma lines, then chroma lines (I420: U lines then V lines)
    const int rows = p.out_h + (fmt != 1 ? 2 : 1) * p.c_out_h;
    dim3 grid((((p.out_w > p.c_out_w ? p.out_w : p.c_out_w) + 3) / 4 + 127) / 128, rows, p.n_frames);
    if (fmt != 1)
        k2_plane_scale_kernel<0><<<grid, 128, 0, st>>>(p);
    else
        k2_plane_scale_kernel<1><<<grid, 128, 0, st>>>(p);
    return cudaGetLastError();
}

}  // namespace b200vc
