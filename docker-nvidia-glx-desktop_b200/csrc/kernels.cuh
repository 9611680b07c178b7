// kernels.cuh — parameter blocks and launchers shared by the sm_100a kernels and the C-ABI layer.
//
// Path replaced: GStreamer videoconvert ! videoscale on the reference's x264enc fallback
// (/root/reference Dockerfile:467-471, README.md:21); arithmetic per SURVEY.md Appendix A/B.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200vc {

// Matrix in the form the kernels consume. Coefficients are packed as four bytes in
// the pixel's byte order (B,G,R,x for BGRx; any of the four sink formats) for dp4a; x's coefficient
// is 0 so the pad byte never matters.
//   Y = dp4a.u32.u32(pix, cy, oy) >> 8
//   U = dp4a.u32.s32(pix, cu, ou) >> 8            (neg == 0)
// Full-range rows hold a +128 coefficient, which does not fit a signed byte. Those matrices are
// evaluated negated (neg == 1): T = dp4a(pix, -c, 65535 - o) = 65535 - S, whose byte 1 is the
// one's complement of byte 1 of S; the kernel undoes it with one XOR on the packed (u,v) word.
struct Matrix {
    uint32_t cy, cu, cv;
    uint32_t oy, ou, ov;
    int32_t neg;
    uint32_t rgb_mask;      // the three bytes of a pixel word that are B, G, R (damage ignores the 4th)
};

struct PlaneSet {
    uint8_t *p[3];          // Y,U,V (I420) or Y,UV (NV12)
    long long pitch[3];
    long long frame_stride; // bytes between consecutive frames, same for every plane pointer
    int aligned16;          // every plane pointer, pitch and the frame stride are multiples of 16: rows may leave as
                            // 16-byte stores (a picture placed at an odd multiple of 8 inside a bordered frame may not)
};

// Row gate of the synchronous host path (b200vc_convert): the kernel is launched while its frame is
// still arriving over PCIe in chunks of `rows` lines; chunk k is complete once flags[k] == epoch (a
// 32-bit word the copy stream writes behind the chunk's copy). A warp / CTA waits for the chunk that
// holds the last line it reads, so conversion runs under the transfer and only the last chunk's work
// is left when the transfer ends. flags == nullptr: no gate (the frame is resident). A wait that
// sees nothing for seconds gives up, sets *timeout and lets the kernel end (the host reports it).
struct RowGate {
    const uint32_t *flags;
    uint32_t epoch;
    int rows;
    uint32_t *timeout;
};

// K1 (+K3): convert a band of rows of n_frames frames.
struct ConvertParams {
    const uint8_t *src;  long long src_pitch;  long long src_frame_stride;
    const uint8_t *prev; long long prev_pitch; long long prev_frame_stride;   // K3 only
    PlaneSet out;
    uint8_t *damage; long long damage_frame_stride; int mb_w;                  // K3 only
    int w, h;               // pixels covered by this launch (h = rows of the band, even)
    int n_frames;
    int stream_out;         // 1: the planes are final output (streaming stores); 0: read again on the GPU (two-pass intermediate)
    int in16;               // 1: 2-byte 5-6-5 pixels (RGB16/BGR16), unpacked in registers; src/prev pitches are in bytes as ever
    RowGate gate;           // n_frames == 1 only
    Matrix mx;
};

// Packed scaler tap: bits 0..14 first source index (clamped), bit 15 set when the second tap is
// index+1 (clear when clamped onto the first), bits 16..24 fraction (H: 0..255, V: 0..2^VTAP_BITS,
// at most 256; include/b200vc_spec.h).
typedef uint32_t Tap;
#ifdef __CUDACC__
#define B200VC_TAP_FN __host__ __device__ __forceinline__
#else
#define B200VC_TAP_FN inline
#endif
B200VC_TAP_FN Tap make_tap(int idx, int frac, int step) { return (Tap)idx | ((Tap)(step & 1) << 15) | ((Tap)frac << 16); }
B200VC_TAP_FN int tap_idx(Tap t) { return (int)(t & 0x7fffu); }
B200VC_TAP_FN int tap_step(Tap t) { return (int)((t >> 15) & 1u); }
B200VC_TAP_FN int tap_frac(Tap t) { return (int)((t >> 16) & 0x1ffu); }

struct ScaleTables {
    const Tap *hy, *vy;     // luma: out_w / out_h entries
    const Tap *hc, *vc;     // chroma: out_w/2, out_h/2 entries
};

// The taps in the form the general fused kernel consumes without decoding, as one record per tile
// column and one per tile row (host-made, api.cu build_fused_taps). Indices are offsets into the
// tile's source window; a clamped second tap (step 0) has weight 0, so the kernel always reads index+1.
//   tile_x[bx] = {window's first frame column (multiple of 4, may be -4), 1 if its first 4 columns
//                 are pure halo}
//   tile_y[by] = {window's first frame line (even), line pairs the tile's taps reach}
//   col_rec[bx][256], luma part, pair_luma == 0:
//                     64 x hyo (byte offset of the first tap in a Y line of the window)
//                     64 x hyw = (255-f) | 1<<8 | f<<16 : [a,a,b,_] . hyw = a*(256-f) + b*f
//                   pair_luma == 1 (every column pair's taps lie within 4 bytes of the first):
//                     32 x {(hyo of column 2l) & ~3 | 8 * (hyo & 3) << 16, selector of column 2l+1's
//                           [a,a,b,_] in the 4 bytes from column 2l's first tap, hyw of 2l, hyw of 2l+1}
//                   chroma part: 32 x {byte offset of the first tap in a (u,v) line, hyw, 0, 0}
//   row_rec[by][3*tile_h]: tile_h x {byte offset of the first line in the Y plane, wv}
//                          tile_h/2 x {byte offset of the first line in the (u,v) plane, wv}
//       wv = wa1 | wa2<<8 | wb1<<16 | wb2<<24 (host_tables.h v_weight_word), for the 6-bit tap
//       wa1+wa2 = 256-4*p1, wb1+wb2 = 4*p1: ([a,a,b,b] . wv + 128) >> 8 == (a*(64-p1) + b*p1 + 32) >> 6
// Columns and rows past the frame edge repeat the last valid entry.
struct FusedTaps {
    const int2 *tile_x;
    const int2 *tile_y;
    const uint32_t *col_rec, *row_rec;
};

// Tensor maps of the output planes (Y; U and V, or UV), for the general fused kernel's tile stores: boxes
// of 64 x tile_h, 32 x tile_h/2 (I420) or 64 x tile_h/2 bytes (NV12), UINT8 elements, dims = the plane's
// samples, so tiles that overhang the frame are clipped by the TMA unit.
struct FusedOutMaps { CUtensorMap y, c1, c2; };

// K2 fused: TMA-staged source window -> convert -> chroma down -> 2-tap scale, one pass.
struct FusedScaleParams {
    PlaneSet out;
    FusedTaps ft;
    int in_w, in_h, out_w, out_h;
    int box_w, box_h;       // source window per tile (pixels); a line of it a multiple of 16 bytes, box_h even
    int tile_h;             // output rows per tile: 8, 16, 32, 48 or 64 (the cheapest per frame, api.cu fused_frame_cost)
    int n_frames;
    int box2x2;             // scale_mode == BOX_2X2
    int pair_luma;          // form of the luma column records (FusedTaps); never with box2x2
    int in16;               // 5-6-5 input pixels (2 bytes; box_w % 8 == 0), unpacked in phase 1
    int planes_off;         // fused_scale_planes_off(box_w, box_h, tile_h, bytes per pixel)
    int tma_out;            // the tile leaves by TMA stores through FusedOutMaps (planes in device memory, 16-byte pitches)
    uint32_t inv_items_x;   // ceil(65536 / items per line pair), for box_w/4 and box_w/4 - 1 groups:
    uint32_t inv_items_x1;  // tid / items_x without a division (exact for tid <= 256, up to 64 items)
    RowGate gate;           // n_frames == 1 only
    Matrix mx;
};

// K2 exact 2:1: TMA-staged 128x32 source tile (+4 halo columns), everything else in registers.
struct Fused2xParams {
    PlaneSet out;
    int in_w, in_h;         // in = 2 * out exactly; in_w % 4 == 0, in_h % 4 == 0
    uint32_t inc_y, inc_c;  // 16.16 horizontal steps of the luma / chroma plane scalers
    int n_frames;
    RowGate gate;           // n_frames == 1 only
    Matrix mx;
};

// Two-pass scale, pass 2: plane scaler reading the full-size intermediate I420.
struct PlaneScaleParams {
    const uint8_t *src[3]; long long src_pitch[3]; long long src_frame_stride;
    PlaneSet out;
    ScaleTables tab;
    int in_w, in_h, out_w, out_h;
    int c_out_w, c_out_h;   // samples per line / lines of a chroma plane of the output (by format)
    int n_frames;
    int box2x2;
};

// Two-pass scale, pass 2, tiled form: KPT_TILE_W x KPT_TILE_H (128 x 64) output tiles of one plane (NV12 chroma: of U and V
// together), the source window staged in shared memory by coalesced 16-byte loads, taps ready-made in
// per-tile records like FusedTaps. Kind 0 = luma plane, kind 1 = chroma planes.
//   tile_x[bx] = window's first source column (multiple of 16), tile_y[by] = its first source line
//   col_rec[bx][2*KPT_TILE_W]: KPT_TILE_W x byte offset of the first tap in a window line,
//                              KPT_TILE_W x (255-f) | 1<<8 | f<<16
//                          or, pair == 1: KPT_TILE_W/2 x the four words of FusedTaps' paired luma records
//   row_rec[by][2*KPT_TILE_H]: KPT_TILE_H x {byte offset of the first line in the window, wv}   (wv as in FusedTaps)
constexpr int KPT_TILE_W = 128;
constexpr int KPT_TILE_H = 64;
struct PlaneTileKind {
    const int *tile_x, *tile_y;
    const uint32_t *col_rec, *row_rec;
    int tiles_x, tiles_y;
    int bwp, bh;            // window pitch in shared memory (bytes, multiple of 16) and lines (without the spare one)
    int src_h;              // lines of the source plane (window lines past the last one repeat it)
    int out_w, out_h;       // samples of the output plane
    int pair;               // col_rec in the paired form (two adjacent columns per lane, as FusedTaps' luma records)
};
struct PlaneTileParams {
    const uint8_t *src[3]; long long src_pitch[3]; long long src_frame_stride;
    PlaneSet out;
    PlaneTileKind k[2];
    int n_frames;
    int box2x2;
};
size_t plane_tile_smem_bytes(const PlaneTileParams &p, int fmt);
cudaError_t plane_tile_prepare(size_t optin_bytes);   // once per device: dynamic shared memory up to the opt-in limit
cudaError_t launch_plane_tile(const PlaneTileParams &p, int fmt, cudaStream_t st);

constexpr int K1_TILE_W = 128;   // pixels per warp stripe (one LDG.128 per lane per line)
constexpr int K1_TILE_H = 16;    // one macroblock row per warp stripe
constexpr int K1_WARPS_PER_CTA = 4;
constexpr int K2_TILE_W = 64;    // output pixels per CTA tile
constexpr int K2_TILE_H = 16;    // exact-2:1 kernel; the general kernel's is FusedScaleParams::tile_h
constexpr int K2_MAX_TILE_H = 64;
constexpr int K2X_SRC_W = 128;   // exact-2:1 kernel: source tile per CTA (8 warps x 4 lines, 32 lanes x 4 pixels)
constexpr int K2X_SRC_H = 32;
constexpr int K2X_TILES_PER_CTA = 2;   // exact-2:1 kernel: tiles per CTA (the later loads hide behind the first tile's work)
constexpr int K2X_HALO = 4;      // columns loaded left of the tile (TMA boxes start 16-byte aligned)

// add-borders: fill everything of n_frames output frames that lies outside the picture's rectangle with
// black (val[i] = the byte for plane i). w/h: bytes per row and rows of each plane; cx/cy/cw/ch: the
// picture's rectangle in each plane, in bytes and rows.
struct BorderParams {
    PlaneSet out;
    int n_planes, n_frames;
    int w[3], h[3];
    int cx[3], cy[3], cw[3], ch[3];
    uint32_t val[3];
};
cudaError_t launch_border_fill(const BorderParams &p, cudaStream_t st);

// Launchers return the cudaError_t of the launch. fmt/site as in b200vc.h.
cudaError_t launch_convert(const ConvertParams &p, int fmt, int site, cudaStream_t st);
cudaError_t launch_fused_scale(const CUtensorMap &src_map, const FusedOutMaps &out_maps, const FusedScaleParams &p, int fmt, int site,
                               cudaStream_t st);
cudaError_t launch_fused_2x(const CUtensorMap &src_map, const Fused2xParams &p, int fmt, int site, int box2x2,
                            cudaStream_t st);
cudaError_t launch_plane_scale(const PlaneScaleParams &p, int fmt, cudaStream_t st);
size_t fused_scale_smem_bytes(int box_w, int box_h, int tile_h, int px_bytes);
size_t fused_scale_planes_off(int box_w, int box_h, int tile_h, int px_bytes);
size_t fused_scale_static_bytes(int tile_h);
cudaError_t fused_scale_prepare(size_t optin_bytes);   // once per device: dynamic shared memory up to the opt-in limit

}  // namespace b200vc
