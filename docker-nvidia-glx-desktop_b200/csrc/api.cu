// api.cu — the C ABI of libb200vc.so (include/b200vc.h) and the host runtime behind it:
// negotiation checks, per-session device frames, pinned staging, the PCIe pipeline (copy streams,
// kernel stream, events; row-gated kernels for the synchronous call), tap tables and TMA
// descriptors for the fused scaler.
//
// Mirrors what GstVideoConvert/GstVideoScale do around gst_video_converter_frame on the
// reference's x264enc fallback (/root/reference Dockerfile:467-471, README.md:21): set_info builds
// the converter once, transform_frame runs it per buffer (b). No CPU fallback
// exists anywhere in this file: without a usable sm_100 device every call fails.
//
// Locking. A context has one mutex. It is held while a call books a slot and enqueues work, and it
// is released around everything that takes long on the host: waiting for a frame's completion
// event, and copying pageable caller buffers to or from the pinned staging buffers. Session threads
// sharing a context therefore overlap their waits and their staging copies; only the enqueueing is
// serialised.
#include "../../include/b200vc.h"
#include "kernels.cuh"
#include "host_tables.h"

#include <cudaTypedefs.h>

#include <algorithm>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace b200vc;

namespace {

thread_local std::string g_create_error;

constexpr int kMaxChunks = 64;                 // row-gate flags per slot (n_bands <= 64)

struct Session {
    uint8_t *d_in[2] = { nullptr, nullptr };   // ping-pong: current and previous frame stay in HBM
    int cur = 0;
    bool have_prev = false;
    cudaEvent_t ev_kernel = nullptr;          // last kernel that read this session's frames
    bool ev_valid = false;
};

// FREE -> SUBMITTING (booked, work still being enqueued: nobody may collect it) -> BUSY (its completion
// event is recorded) -> COLLECTING (a thread waits for it with the mutex released) -> FREE
enum SlotState { SLOT_FREE = 0, SLOT_SUBMITTING = 1, SLOT_BUSY = 2, SLOT_COLLECTING = 3 };

struct Slot {
    uint8_t *d_out = nullptr;                 // device output frame (+ damage map at d_dmg)
    uint8_t *d_dmg = nullptr;
    uint32_t *d_flags = nullptr;              // row gate: kMaxChunks flags + 1 time-out word, zero at create
    uint32_t *h_flag_src = nullptr;           // pinned, mapped: word kMaxChunks is the gate's time-out word
    uint32_t epoch = 0;
    bool gated = false;                       // this frame ran row-gated: look at the time-out word when collecting
    uint8_t *h_in = nullptr;                  // pinned staging, allocated on first pageable frame
    uint8_t *h_out = nullptr;
    uint8_t *h_map = nullptr;                 // pinned staging for a pageable damage map (damage contexts)
    cudaEvent_t done = nullptr;
    int state = SLOT_FREE;
    uint64_t ticket = 0;
    bool staged_out = false;                  // deferred host copy out of h_out (pageable planes)
    bool staged_map = false;                  // deferred host copy out of h_map (pageable map)
    b200vc_frame frame;
};

struct HostReg { const uint8_t *base; size_t bytes; uint64_t last_use; };
struct PtrKind { const void *p; int kind; const uint8_t *lo, *hi; };   // hi/lo: extent of the page-locked allocation
struct MapEntry { const void *base; size_t pitch, fstride; int n_frames; CUtensorMap map; };
// the output planes' maps of the general fused kernel; usable: planes in device memory, everything 16-byte aligned
struct OutMapEntry { PlaneSet planes; int n_frames; bool usable; FusedOutMaps maps; };

typedef CUresult (*WriteValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

}  // namespace

struct b200vc_ctx {
    b200vc_config cfg;
    b200vc_config kcfg;                        // cfg as the scaler sees it: out_w x out_h = the picture's rectangle (add-borders)
    bool borders = false;                      // the picture does not fill the output frame
    int cont_x = 0, cont_y = 0;                // where it sits
    int ccw = 0, cch = 0;                      // its chroma plane's samples per line / lines
    bool scaling = false;
    int path = 0;                              // 0 convert, 1 fused scale, 2 two-pass scale
    Matrix mx;
    cudaDeviceProp prop;
    cudaStream_t s_h2d = nullptr, s_k = nullptr, s_d2h = nullptr;
    cudaStream_t s_up[4] = { nullptr, nullptr, nullptr, nullptr };   // row gate: chunks alternate over these (s_up[0] == s_h2d)
    int n_up = 1;
    bool auto_bands = false;                   // n_bands was left 0: chosen per frame
    std::vector<cudaEvent_t> ev_band_h2d, ev_band_k;

    // device geometry (library-owned buffers)
    size_t d_in_pitch = 0, d_in_size = 0;
    size_t d_off[3] = { 0, 0, 0 }, d_pitch[3] = { 0, 0, 0 }, d_out_size = 0;
    bool out_compact = false;                  // device output frame = b200vc_default_layout: a contiguous caller frame leaves in one copy
    int mb_w = 0, mb_h = 0;
    // by format: bytes of an input row; planes, bytes per row and rows of each output plane; chroma plane
    // dimensions of the input-sized intermediate and of the output; lines per chroma line (band arithmetic)
    size_t row_bytes = 0;
    int n_planes = 3, vsub = 2;
    size_t prow_bytes[3] = { 0, 0, 0 };
    int prows[3] = { 0, 0, 0 };
    int cw_in = 0, ch_in = 0, cw_out = 0, ch_out = 0;

    std::vector<Session> sessions;
    std::vector<Slot> slots;
    uint64_t next_ticket = 1;
    int next_slot = 0;

    // scaling
    Tap *d_taps = nullptr;
    ScaleTables tab = { nullptr, nullptr, nullptr, nullptr };
    uint32_t *d_ftaps = nullptr;               // the general fused kernel's per-tile tap records
    FusedTaps ft = { nullptr, nullptr, nullptr, nullptr };
    int tile_h = K2_TILE_H;
    bool pair_luma = false;
    bool no_tma_out = false;
    bool equal_bands = false;                  // B200VC_EQUAL_BANDS: the synchronous call's bands all the same size (measurements only)
    uint32_t *d_ptaps = nullptr;               // tiled plane scaler's records (two-pass path)
    PlaneTileKind pt[2] = {};
    bool plane_tiled = false;                  // two-pass path: pass 2 is the tiled kernel, not the gather kernel
    int box_w = 0, box_h = 0;
    size_t smem_bytes = 0;
    bool fused2x = false;                      // path 1 runs the exact-2:1 register kernel
    uint32_t inc_y = 0, inc_c = 0;
    uint8_t *d_mid = nullptr;                  // two-pass: full-size intermediate I420
    size_t mid_off[3] = { 0, 0, 0 }, mid_pitch[3] = { 0, 0, 0 }, mid_size = 0;
    int mid_frames = 0;
    cudaEvent_t ev_mid = nullptr;              // last pass 2 that read d_mid, on whichever stream it ran
    bool ev_mid_valid = false;
    PFN_cuTensorMapEncodeTiled_v12000 encode_tiled = nullptr;
    std::vector<MapEntry> maps;                // tensor maps seen so far (a session has two input buffers)
    std::vector<OutMapEntry> out_maps;         // and of output planes (general fused kernel)
    size_t next_out_map = 0;
    size_t next_map = 0;
    WriteValue32Fn write_value = nullptr;      // cuStreamWriteValue32, for the row gate's flags

    std::vector<HostReg> registered;           // B200VC_FLAG_REGISTER_HOST: ranges this context page-locked
    uint64_t reg_clock = 0;
    PtrKind kinds[256] = {};                   // what cudaPointerGetAttributes said about recent caller pointers
    b200vc_stats stats;
    mutable std::mutex mu;
    std::condition_variable cv;                // a slot left SLOT_COLLECTING
    std::string err;
};

namespace {

typedef std::unique_lock<std::mutex> Lock;
int fail(b200vc_ctx *c, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}

#define CU_TRY(c, call)                                                                          \
    do {                                                                                         \
        cudaError_t e_ = (call);                                                                 \
        if (e_ != cudaSuccess)                                                                   \
            return fail((c), e_ == cudaErrorMemoryAllocation ? B200VC_ERR_NOMEM : B200VC_ERR_CUDA, \
                        "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

size_t ru(size_t v, size_t a) { return (v + a - 1) / a * a; }
bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

// Dynamic shared memory the general fused kernel may ask for: the device's opt-in limit less room for
// its static arrays. The kernels' function attribute is set to the device limit (*_prepare), never to one
// geometry's need: the kernels are shared by every context of the process, and a later, smaller
// negotiation must not shrink an earlier one's.

// ---- scaler taps as tables; positions and spec constants from host_tables.h ------------
void build_h_taps(int in_n, int out_n, std::vector<Tap> &t)
{
    const long long inc = scale_h_inc(in_n, out_n);
    t.resize(ru(out_n, 4));
    for (int x = 0; x < out_n; x++) {
        int i, f;
        scale_h_tap(inc, x, i, f);
        const int i0 = clamp_index(i, in_n), i1 = clamp_index((long long)i + 1, in_n);
        t[x] = make_tap(i0, f, i1 - i0);
    }
    for (size_t x = out_n; x < t.size(); x++) t[x] = t[out_n - 1];
}

void build_v_taps(int in_n, int out_n, std::vector<Tap> &t)
{
    t.resize(ru(out_n, 4));
    for (int y = 0; y < out_n; y++) {
        long long i;
        int p1;
        scale_v_tap(y, in_n, out_n, i, p1);
        const int i0 = clamp_index(i, in_n), i1 = clamp_index(i + 1, in_n);
        t[y] = make_tap(i0, p1, i1 - i0);
    }
    for (size_t y = out_n; y < t.size(); y++) t[y] = t[out_n - 1];
}

// method=nearest as taps: all the weight on one sample (fraction 0; the second tap exists so that the kernels,
// which always read two, stay inside the plane).
void build_nearest_taps(int in_n, int out_n, std::vector<Tap> &t)
{
    t.resize(ru(out_n, 4));
    for (int x = 0; x < out_n; x++) {
        const int i0 = clamp_index(scale_nearest_index(x, in_n, out_n), in_n), i1 = clamp_index((long long)i0 + 1, in_n);
        t[x] = make_tap(i0, 0, i1 - i0);
    }
    for (size_t x = out_n; x < t.size(); x++) t[x] = t[out_n - 1];
}

// BOX_2X2 reuses the tables as plain index maps: entry k = source index 2k.
void build_box_taps(int out_n, std::vector<Tap> &t)
{
    t.resize(ru(out_n, 4));
    for (int k = 0; k < out_n; k++) t[k] = make_tap(2 * k, 0, 1);
    for (size_t k = out_n; k < t.size(); k++) t[k] = t[out_n - 1];
}

inline int tidx(Tap t) { return tap_idx(t); }
inline int tstep(Tap t) { return tap_step(t); }
inline int tfrac(Tap t) { return tap_frac(t); }

// The general fused kernel's form of the tables (FusedTaps, kernels.cuh) as one blob of 32-bit words:
// [tile_x pairs | tile_y | col_rec | row_rec], every section padded to a multiple of 4 words.
struct FusedTapBlob { std::vector<uint32_t> w; size_t tile_x, tile_y, col_rec, row_rec; };

uint32_t v_weights(Tap t)
{
    return v_weight_word(tstep(t) ? tfrac(t) : 0);   // clamped second line: all weight on the first
}

// Window origin of a tile column / tile row: the first source column (line) any of its taps reads,
// the cosited chroma filter's left neighbour included, rounded down to what the TMA box (16-byte
// aligned start) and the 2x2 chroma sites need.
inline int tile_org_x(const b200vc_config &c, const std::vector<Tap> &hy, const std::vector<Tap> &hc, int x0, int site)
{
    const int align = 16 / in_bytes_per_pixel(c.in_format);              // pixels per 16 bytes: 4, or 8 (5-6-5)
    return std::min(tidx(hy[x0]), 2 * tidx(hc[x0 >> 1]) - site) & ~(align - 1);
}
inline int tile_org_y(const std::vector<Tap> &vy, const std::vector<Tap> &vc, int y0)
{
    return std::min(tidx(vy[y0]), 2 * tidx(vc[y0 >> 1])) & ~1;
}

// The part of its window a tile's taps reach, as phase 1 of the general fused kernel walks it: the
// window's first frame column (16-byte aligned), whether its first 4-column group is pure halo, and
// the number of groups after that / the first frame line (even) and the number of line pairs.
struct FusedSpanX { int sx, kfirst, groups; };
struct FusedSpanY { int sy, pairs; };
inline FusedSpanX fused_span_x(const b200vc_config &c, const std::vector<Tap> &hy, const std::vector<Tap> &hc, int x0, int site)
{
    const int x1 = std::min(x0 + K2_TILE_W, c.out_w) - 1;
    const int sx = tile_org_x(c, hy, hc, x0, site);
    const int first = std::min(tidx(hy[x0]), 2 * tidx(hc[x0 >> 1]));
    const int last = std::max(tidx(hy[x1]) + tstep(hy[x1]), 2 * (tidx(hc[x1 >> 1]) + tstep(hc[x1 >> 1])) + 1);
    const int kfirst = std::min(first - sx, 4) >> 2;
    return { sx, kfirst, (last - sx) / 4 + 1 - kfirst };
}
inline FusedSpanY fused_span_y(const b200vc_config &c, const std::vector<Tap> &vy, const std::vector<Tap> &vc, int y0, int tile_h)
{
    const int y1 = std::min(y0 + tile_h, c.out_h) - 1;
    const int sy = tile_org_y(vy, vc, y0);
    const int last = std::max(tidx(vy[y1]) + tstep(vy[y1]), 2 * (tidx(vc[y1 >> 1]) + tstep(vc[y1 >> 1])) + 1);
    return { sy, (last - sy) / 2 + 1 };
}

// True when the two taps of every output column pair (2l, 2l+1) lie within 4 bytes of the pair's first
// tap, which is what the kernel's paired form of the luma pass needs (any ratio up to 2).
bool luma_pairs_fit(const b200vc_config &c, const std::vector<Tap> &hy)
{
    for (int x = 0; x + 1 < c.out_w; x += 2)
        if (tidx(hy[x + 1]) - tidx(hy[x]) > 2 || tidx(hy[x + 1]) < tidx(hy[x])) return false;
    return true;
}

void build_fused_taps(const b200vc_config &c, const std::vector<Tap> &hy, const std::vector<Tap> &vy,
                      const std::vector<Tap> &hc, const std::vector<Tap> &vc, int tile_h, int box_w, bool pair_luma,
                      FusedTapBlob &o)
{
    const int site = (c.chroma_site & B200VC_SITE_H_COSITED) ? 1 : 0;
    const int tiles_x = (c.out_w + K2_TILE_W - 1) / K2_TILE_W, tiles_y = (c.out_h + tile_h - 1) / tile_h;
    const int cw_out = (c.out_w + 1) / 2, ch_out = (c.out_h + 1) / 2;          // odd frames: chroma planes round up
    auto pad4 = [&] { while (o.w.size() % 4) o.w.push_back(0u); };
    o.w.clear();
    o.tile_x = o.w.size();
    for (int bx = 0; bx < tiles_x; bx++) {
        const FusedSpanX sp = fused_span_x(c, hy, hc, bx * K2_TILE_W, site);
        o.w.push_back((uint32_t)sp.sx);
        o.w.push_back((uint32_t)sp.kfirst);
    }
    pad4();
    o.tile_y = o.w.size();
    for (int by = 0; by < tiles_y; by++) {
        const FusedSpanY sp = fused_span_y(c, vy, vc, by * tile_h, tile_h);
        o.w.push_back((uint32_t)sp.sy);
        o.w.push_back((uint32_t)sp.pairs);
    }
    pad4();
    o.col_rec = o.w.size();
    for (int bx = 0; bx < tiles_x; bx++) {
        const int x0 = bx * K2_TILE_W, sx = tile_org_x(c, hy, hc, x0, site);
        auto h_weights = [](Tap t) {
            const uint32_t f = tstep(t) ? (uint32_t)tfrac(t) : 0u;
            return (255u - f) | (1u << 8) | (f << 16);
        };
        if (pair_luma) {
            for (int l = 0; l < K2_TILE_W / 2; l++) {
                const Tap t0 = hy[std::min(x0 + 2 * l, c.out_w - 1)], t1 = hy[std::min(x0 + 2 * l + 1, c.out_w - 1)];
                const uint32_t o0 = (uint32_t)(tidx(t0) - sx), d = (uint32_t)(tidx(t1) - tidx(t0));
                o.w.push_back((o0 & ~3u) | ((8u * (o0 & 3u)) << 16));
                o.w.push_back(d | (d << 4) | ((d + 1u) << 8) | (3u << 12));
                o.w.push_back(h_weights(t0));  o.w.push_back(h_weights(t1));
            }
        } else {
            for (int k = 0; k < K2_TILE_W; k++)
                o.w.push_back((uint32_t)(tidx(hy[std::min(x0 + k, c.out_w - 1)]) - sx));
            for (int k = 0; k < K2_TILE_W; k++)
                o.w.push_back(h_weights(hy[std::min(x0 + k, c.out_w - 1)]));
        }
        for (int k = 0; k < K2_TILE_W / 2; k++) {
            const Tap t = hc[std::min(x0 / 2 + k, cw_out - 1)];
            o.w.push_back((uint32_t)(2 * (tidx(t) - sx / 2)));  o.w.push_back(h_weights(t));  o.w.push_back(0u);  o.w.push_back(0u);
        }
    }
    o.row_rec = o.w.size();
    for (int by = 0; by < tiles_y; by++) {
        const int y0 = by * tile_h, sy = tile_org_y(vy, vc, y0);
        for (int k = 0; k < tile_h; k++) {
            const Tap t = vy[std::min(y0 + k, c.out_h - 1)];
            o.w.push_back((uint32_t)((tidx(t) - sy) * box_w));  o.w.push_back(v_weights(t));
        }
        for (int k = 0; k < tile_h / 2; k++) {
            const Tap t = vc[std::min(y0 / 2 + k, ch_out - 1)];
            o.w.push_back((uint32_t)((tidx(t) - sy / 2) * box_w));  o.w.push_back(v_weights(t));
        }
    }
}

// Per-tile tap records of the tiled plane scaler (PlaneTileKind, kernels.cuh) for one plane kind:
// in_w x in_h -> out_w x out_h through the taps th (horizontal) and tv (vertical). Returns the blob
// [tile_x | tile_y | col_rec | row_rec] (sections padded to 4 words) and fills everything of `k` but
// the pointers; `off` receives the sections' word offsets.
void build_plane_tile_taps(const std::vector<Tap> &th, const std::vector<Tap> &tv, int in_h, int out_w, int out_h, bool allow_pair,
                           PlaneTileKind &k, std::vector<uint32_t> &w, size_t off[4])
{
    const int tiles_x = (out_w + KPT_TILE_W - 1) / KPT_TILE_W, tiles_y = (out_h + KPT_TILE_H - 1) / KPT_TILE_H;
    int bw = 0, bh = 0;
    std::vector<int> sx(tiles_x), sy(tiles_y);
    for (int bx = 0; bx < tiles_x; bx++) {
        const int x0 = bx * KPT_TILE_W, x1 = std::min(x0 + KPT_TILE_W, out_w) - 1;
        sx[bx] = tidx(th[x0]) & ~15;                                       // 16-byte loads
        bw = std::max(bw, tidx(th[x1]) + 1 - sx[bx] + 1);                  // index+1 is always read
    }
    for (int by = 0; by < tiles_y; by++) {
        const int y0 = by * KPT_TILE_H, y1 = std::min(y0 + KPT_TILE_H, out_h) - 1;
        sy[by] = tidx(tv[y0]);
        bh = std::max(bh, tidx(tv[y1]) + 1 - sy[by] + 1);
    }
    k.tiles_x = tiles_x;  k.tiles_y = tiles_y;
    k.bwp = (int)ru(bw, 16);  k.bh = bh;
    k.src_h = in_h;  k.out_w = out_w;  k.out_h = out_h;
    auto pad4 = [&] { while (w.size() % 4) w.push_back(0u); };
    off[0] = w.size();
    for (int v : sx) w.push_back((uint32_t)v);
    pad4();
    off[1] = w.size();
    for (int v : sy) w.push_back((uint32_t)v);
    pad4();
    off[2] = w.size();
    // the paired form (two adjacent columns per lane) when every pair's taps lie within 4 bytes of the pair's first
    // (not in box mode and not with the vertical-first variant of the spec: the kernel's byte-load form handles those)
    k.pair = allow_pair ? 1 : 0;
    for (int x = 0; x + 1 < out_w; x += 2)
        if (tidx(th[x + 1]) - tidx(th[x]) > 2 || tidx(th[x + 1]) < tidx(th[x])) k.pair = 0;
    auto h_weights = [](Tap t) {
        const uint32_t f = tstep(t) ? (uint32_t)tfrac(t) : 0u;
        return (255u - f) | (1u << 8) | (f << 16);
    };
    for (int bx = 0; bx < tiles_x; bx++) {
        const int x0 = bx * KPT_TILE_W;
        if (k.pair) {
            for (int l = 0; l < KPT_TILE_W / 2; l++) {
                const Tap t0 = th[std::min(x0 + 2 * l, out_w - 1)], t1 = th[std::min(x0 + 2 * l + 1, out_w - 1)];
                const uint32_t o0 = (uint32_t)(tidx(t0) - sx[bx]), d = (uint32_t)(tidx(t1) - tidx(t0));
                w.push_back((o0 & ~3u) | ((8u * (o0 & 3u)) << 16));
                w.push_back(d | (d << 4) | ((d + 1u) << 8) | (3u << 12));
                w.push_back(h_weights(t0));  w.push_back(h_weights(t1));
            }
            continue;
        }
        for (int i = 0; i < KPT_TILE_W; i++) w.push_back((uint32_t)(tidx(th[std::min(x0 + i, out_w - 1)]) - sx[bx]));
        for (int i = 0; i < KPT_TILE_W; i++) w.push_back(h_weights(th[std::min(x0 + i, out_w - 1)]));
    }
    off[3] = w.size();
    for (int by = 0; by < tiles_y; by++) {
        const int y0 = by * KPT_TILE_H;
        for (int i = 0; i < KPT_TILE_H; i++) {
            const Tap t = tv[std::min(y0 + i, out_h - 1)];
            w.push_back((uint32_t)((tidx(t) - sy[by]) * k.bwp));  w.push_back(v_weights(t));
        }
    }
    pad4();
}

// Source window per output tile, mirrored from k2_fused_scale_kernel phase 0; returns the box that
// covers every tile.
void fused_box(const b200vc_config &c, const std::vector<Tap> &hy, const std::vector<Tap> &vy,
               const std::vector<Tap> &hc, const std::vector<Tap> &vc, int tile_h, int &box_w, int &box_h)
{
    const int site = (c.chroma_site & B200VC_SITE_H_COSITED) ? 1 : 0;
    box_w = box_h = 0;
    for (int x0 = 0; x0 < c.out_w; x0 += K2_TILE_W) {
        const FusedSpanX sp = fused_span_x(c, hy, hc, x0, site);
        box_w = std::max(box_w, 4 * (sp.kfirst + sp.groups));
    }
    for (int y0 = 0; y0 < c.out_h; y0 += tile_h)
        box_h = std::max(box_h, 2 * fused_span_y(c, vy, vc, y0, tile_h).pairs);
    box_w = (int)ru((size_t)box_w, 16 / in_bytes_per_pixel(c.in_format));   // a box line is a multiple of 16 bytes
}

// Issue slots the general fused kernel spends on one frame with tiles of tile_h rows, in warp
// instructions (the kernel is issue-bound; counts per phase from its ncu source page joined with the
// line table, tools/ncu_phases.py, profiles/r2_general_scaler_experiments.md): per tile and warp a fixed
// part (set-up, the prologues of the phases), phase 1 per warp with work in each sweep of the 256
// threads down the window (items of 4 columns x 2 lines, 256 / groups line pairs at a time), phase 2
// per row of luma / chroma, phase 3 per sweep over the tile's 16-byte chunks. Tall tiles amortise the
// fixed part and the halo lines, and the height decides how full phase 1's last sweep is.
double fused_frame_cost(const b200vc_config &c, const std::vector<Tap> &hy, const std::vector<Tap> &vy,
                        const std::vector<Tap> &hc, const std::vector<Tap> &vc, int tile_h)
{
    const int site = (c.chroma_site & B200VC_SITE_H_COSITED) ? 1 : 0;
    int box_w, box_h;
    fused_box(c, hy, vy, hc, vc, tile_h, box_w, box_h);
    if (box_w > 256 || box_h > 256) return 1e30;           // (a window the kernel does not take)
    std::vector<int> gx, py;
    for (int x0 = 0; x0 < c.out_w; x0 += K2_TILE_W) gx.push_back(box_w / 4 - fused_span_x(c, hy, hc, x0, site).kfirst);
    for (int y0 = 0; y0 < c.out_h; y0 += tile_h) py.push_back(fused_span_y(c, vy, vc, y0, tile_h).pairs);
    double phase1 = 0;
    for (int g : gx) {
        const int rps = 256 / g;                           // line pairs per sweep
        for (int p : py)
            for (int done = 0; done < p; done += rps)
                phase1 += 90.0 * ((g * std::min(rps, p - done) + 31) / 32);
    }
    const double tiles = (double)gx.size() * py.size();
    const double per_tile = 8.0 * (65 + 45 + 28 + 15) + 8.0 * 23 * (tile_h / 8) + 8.0 * 22 * ((tile_h / 2 + 7) / 8) +
                            8.0 * 30 * ((tile_h * 6 + 255) / 256);
    return tiles * per_tile + phase1;
}

// True when every tap of an (in_n -> out_n) scaler starts at 2k with the second tap at 2k+1 and,
// for vertical taps, weighs both lines equally (p1 == half): the pattern k2_fused_2x_kernel assumes.
bool taps_are_2to1(const std::vector<Tap> &t, int out_n, bool vertical)
{
    for (int k = 0; k < out_n; k++) {
        if (tidx(t[k]) != 2 * k || tstep(t[k]) != 1) return false;
        if (vertical && tfrac(t[k]) != B200VC_SPEC_VTAP_HALF) return false;
    }
    return true;
}

uint32_t h_inc(int in_n, int out_n) { return (uint32_t)scale_h_inc(in_n, out_n); }

int check_config(const b200vc_config *c, std::string *why)
{
    auto bad = [&](int code, const char *msg) { if (why) *why = msg; return code; };
    if (!c) return bad(B200VC_ERR_INVALID_ARG, "config is NULL");
    if (c->in_w < 2 || c->in_h < 2 || c->out_w < 2 || c->out_h < 2)
        return bad(B200VC_ERR_INVALID_ARG, "width/height must be >= 2");
    if (c->in_w > 16384 || c->in_h > 16384 || c->out_w > 16384 || c->out_h > 16384)
        return bad(B200VC_ERR_UNSUPPORTED, "width/height above 16384");
    if (c->scale_mode == B200VC_SCALE_BOX_2X2 && ((c->in_w | c->in_h | c->out_w | c->out_h) & 1) &&
        (c->in_w != c->out_w || c->in_h != c->out_h || c->content_w))
        return bad(B200VC_ERR_UNSUPPORTED, "BOX_2X2 needs even widths and heights");
    if (c->out_format < B200VC_FMT_I420 || c->out_format > B200VC_FMT_Y444)
        return bad(B200VC_ERR_UNSUPPORTED, "output format must be I420, NV12, Y42B or Y444");
    if (c->in_format < B200VC_IN_BGRX || c->in_format > B200VC_IN_BGR16)
        return bad(B200VC_ERR_UNSUPPORTED, "input format must be BGRx/BGRA, RGBx/RGBA, xRGB/ARGB, xBGR/ABGR, RGB16 or BGR16");
    if (c->matrix < B200VC_MATRIX_BT601 || c->matrix > B200VC_MATRIX_BT2020)
        return bad(B200VC_ERR_UNSUPPORTED, "colorimetry matrix must be bt601, bt709 or bt2020");
    if (c->range != B200VC_RANGE_LIMITED && c->range != B200VC_RANGE_FULL)
        return bad(B200VC_ERR_UNSUPPORTED, "range must be limited (16-235) or full (0-255)");
    if (c->chroma_site < 0 || c->chroma_site > B200VC_SITE_DV)
        return bad(B200VC_ERR_UNSUPPORTED, "chroma-site must be a combination of h-cosited, v-cosited and alt-line (jpeg, mpeg2, cosited, dv)");
    if (c->scale_mode < B200VC_SCALE_GST_LQ || c->scale_mode > B200VC_SCALE_NEAREST)
        return bad(B200VC_ERR_UNSUPPORTED, "scale method must be bilinear, nearest (or the 2x2 box)");
    int pw = c->out_w, ph = c->out_h;                      // the picture inside the output frame (add-borders)
    if (c->content_x || c->content_y || c->content_w || c->content_h) {
        if (c->content_x < 0 || c->content_y < 0 || c->content_w < 2 || c->content_h < 2 ||
            c->content_x + c->content_w > c->out_w || c->content_y + c->content_h > c->out_h)
            return bad(B200VC_ERR_INVALID_ARG, "the content rectangle does not lie inside the output frame");
        if ((c->content_x | c->content_y | c->content_w | c->content_h) & 1)
            return bad(B200VC_ERR_UNSUPPORTED, "the content rectangle's position and size must be even");
        pw = c->content_w;  ph = c->content_h;
    }
    const bool scaling = c->in_w != pw || c->in_h != ph;
    if (scaling && c->scale_mode == B200VC_SCALE_BOX_2X2 && (c->in_w != 2 * pw || c->in_h != 2 * ph))
        return bad(B200VC_ERR_UNSUPPORTED, "BOX_2X2 needs an exact 2:1 downscale");
    if (scaling && (c->flags & B200VC_FLAG_DAMAGE_MAP))
        return bad(B200VC_ERR_UNSUPPORTED, "the damage map is produced on the unscaled path only");
    if ((c->flags & B200VC_FLAG_DAMAGE_MAP) && c->out_format > B200VC_FMT_NV12)
        return bad(B200VC_ERR_UNSUPPORTED, "the damage map is produced with I420 and NV12 output (what the encoders take)");
    if (c->max_sessions < 0 || c->max_sessions > 1024 || c->n_inflight < 0 || c->n_inflight > 64 ||
        c->n_bands < 0 || c->n_bands > 64)
        return bad(B200VC_ERR_INVALID_ARG, "max_sessions/n_inflight/n_bands out of range");
    return B200VC_OK;
}

// What a caller pointer is: 1 = page-locked host memory (cudaMallocHost / cudaHostRegister), 0 = pageable
// host memory, -1 = device or managed memory (belongs to b200vc_convert_device, not to the host entry
// points). For page-locked memory also the extent [lo, hi) of the page-locked allocation it lies in: a
// buffer is page-locked only if it lies in ONE such extent as a whole. (A buffer on the heap can begin and
// end in pages its neighbours had locked while its middle is pageable; the copy engine refuses that.)
typedef CUresult (*PointerGetAttributeFn)(void *, CUpointer_attribute, CUdeviceptr);

PointerGetAttributeFn pointer_get_attribute()
{
    static PointerGetAttributeFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            f = nullptr;
        }
        return (PointerGetAttributeFn)f;
    }();
    return fn;
}

PtrKind classify_pointer(const void *p)
{
    PtrKind k = { p, 0, nullptr, nullptr };
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return k;
    }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) { k.kind = -1;  return k; }
    if (a.type != cudaMemoryTypeHost) return k;
    // extent of the page-locked allocation; without the driver's answer only the byte asked about is vouched for
    k.kind = 1;
    k.lo = (const uint8_t *)p;  k.hi = k.lo + 1;
    if (PointerGetAttributeFn get = pointer_get_attribute()) {
        CUdeviceptr start = 0;
        size_t size = 0;
        if (get(&start, CU_POINTER_ATTRIBUTE_RANGE_START_ADDR, (CUdeviceptr)(uintptr_t)p) == CUDA_SUCCESS &&
            get(&size, CU_POINTER_ATTRIBUTE_RANGE_SIZE, (CUdeviceptr)(uintptr_t)p) == CUDA_SUCCESS && size) {
            k.lo = (const uint8_t *)(uintptr_t)start;
            k.hi = k.lo + size;
        }
    }
    return k;
}

// A session hands in the same few buffers frame after frame (an XShm pool, a buffer pool): the
// driver is asked once per pointer, not five times per frame. A stale answer is harmless: cudaMemcpy
// copes with either kind of host memory, and a mapping that is gone fails the zero-copy set-up,
// which falls back to copies.
const PtrKind &pointer_kind(b200vc_ctx *c, const void *p)
{
    PtrKind &k = c->kinds[((uintptr_t)p >> 4) % (sizeof c->kinds / sizeof c->kinds[0])];
    if (k.p != p) k = classify_pointer(p);
    return k;
}

int host_kind(b200vc_ctx *c, const void *p) { return pointer_kind(c, p).kind; }

// the whole buffer [p, p+bytes)
int host_kind(b200vc_ctx *c, const void *p, size_t bytes)
{
    const PtrKind &k = pointer_kind(c, p);
    if (k.kind != 1) return k.kind;
    return ((const uint8_t *)p >= k.lo && (const uint8_t *)p + bytes <= k.hi) ? 1 : 0;
}

void forget_kinds(b200vc_ctx *c)
{
    for (PtrKind &k : c->kinds) k = { nullptr, 0, nullptr, nullptr };
}

// B200VC_FLAG_REGISTER_HOST: page-lock the caller's range [p, p+bytes) once, whole pages; false =
// fall back to staging. At most kMaxRegistered ranges are kept; the least recently used one is
// released (after the work in flight has drained) to make room.
constexpr size_t kMaxRegistered = 16;

bool ensure_registered(b200vc_ctx *c, const void *p, size_t bytes)
{
    const uint8_t *lo = (const uint8_t *)((uintptr_t)p & ~(uintptr_t)4095);
    const uint8_t *hi = (const uint8_t *)(((uintptr_t)p + bytes + 4095) & ~(uintptr_t)4095);
    for (HostReg &r : c->registered)
        if (lo >= r.base && hi <= r.base + r.bytes) { r.last_use = ++c->reg_clock; return true; }
    if (c->registered.size() >= kMaxRegistered) {
        size_t victim = 0;
        for (size_t i = 1; i < c->registered.size(); i++)
            if (c->registered[i].last_use < c->registered[victim].last_use) victim = i;
        cudaStreamSynchronize(c->s_h2d);  cudaStreamSynchronize(c->s_k);  cudaStreamSynchronize(c->s_d2h);
        cudaHostUnregister(const_cast<uint8_t *>(c->registered[victim].base));
        cudaGetLastError();
        c->registered.erase(c->registered.begin() + victim);
        forget_kinds(c);
    }
    cudaError_t e = cudaHostRegister(const_cast<uint8_t *>(lo), (size_t)(hi - lo), cudaHostRegisterDefault);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        // Buffers that sit next to each other on the heap share their boundary pages, which an earlier
        // range of ours already holds: merge — release every range of ours that touches this one and
        // page-lock their union as one piece. (Pages someone else locked stay a failure: a range that is
        // page-locked only in parts cannot be handed to the copy engine.)
        cudaGetLastError();
        bool merged = false;
        for (size_t i = 0; i < c->registered.size();) {
            const HostReg r = c->registered[i];
            if (r.base < hi && lo < r.base + r.bytes) {
                if (!merged) { cudaStreamSynchronize(c->s_h2d);  cudaStreamSynchronize(c->s_k);  cudaStreamSynchronize(c->s_d2h); }
                merged = true;
                cudaHostUnregister(const_cast<uint8_t *>(r.base));
                cudaGetLastError();
                lo = std::min(lo, r.base);
                hi = std::max(hi, r.base + r.bytes);
                c->registered.erase(c->registered.begin() + i);
            } else {
                i++;
            }
        }
        if (merged) {
            forget_kinds(c);
            e = cudaHostRegister(const_cast<uint8_t *>(lo), (size_t)(hi - lo), cudaHostRegisterDefault);
        }
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    c->registered.push_back({ lo, (size_t)(hi - lo), ++c->reg_clock });
    forget_kinds(c);
    return true;
}

// TMA descriptor of a source buffer. A session has two input buffers and a device ring is reused launch
// after launch, so descriptors are kept (cuTensorMapEncodeTiled costs microseconds on the streaming thread).
int get_tensor_map(b200vc_ctx *c, const void *base, size_t pitch, size_t frame_stride, int n_frames, const CUtensorMap **out)
{
    for (const MapEntry &m : c->maps)
        if (m.base == base && m.pitch == pitch && m.fstride == frame_stride && m.n_frames == n_frames) { *out = &m.map; return B200VC_OK; }
    MapEntry e;
    e.base = base;  e.pitch = pitch;  e.fstride = frame_stride;  e.n_frames = n_frames;
    cuuint64_t dims[3] = { (cuuint64_t)c->cfg.in_w, (cuuint64_t)c->cfg.in_h, (cuuint64_t)n_frames };
    cuuint64_t strides[2] = { (cuuint64_t)pitch, (cuuint64_t)(frame_stride ? frame_stride : pitch * c->cfg.in_h) };
    cuuint32_t box[3] = { (cuuint32_t)c->box_w, (cuuint32_t)c->box_h, 1 };
    cuuint32_t estr[3] = { 1, 1, 1 };
    const bool in16 = in_bytes_per_pixel(c->cfg.in_format) == 2;
    CUresult r = c->encode_tiled(&e.map, in16 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void *>(base), dims, strides,
                                 box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(c, B200VC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    constexpr size_t kMaxMaps = 64;
    if (c->maps.size() < kMaxMaps) {
        c->maps.push_back(e);                  // (reserved at create: entries never move)
        *out = &c->maps.back().map;
    } else {
        c->maps[c->next_map] = e;
        *out = &c->maps[c->next_map].map;
        c->next_map = (c->next_map + 1) % kMaxMaps;
    }
    return B200VC_OK;
}

// TMA descriptors of the output planes (general fused kernel). Kept like the sources': a session's planes
// are the same few buffers launch after launch. `usable` false: the kernel stores the tile itself.
const OutMapEntry *get_out_maps(b200vc_ctx *c, const PlaneSet &out, int n_frames)
{
    for (const OutMapEntry &m : c->out_maps)
        if (m.n_frames == n_frames && m.planes.frame_stride == out.frame_stride && m.planes.p[0] == out.p[0] && m.planes.p[1] == out.p[1] &&
            m.planes.p[2] == out.p[2] && m.planes.pitch[0] == out.pitch[0] && m.planes.pitch[1] == out.pitch[1] && m.planes.pitch[2] == out.pitch[2])
            return &m;
    const b200vc_config &g = c->kcfg;
    OutMapEntry e;
    memset(&e, 0, sizeof e);
    e.planes = out;  e.n_frames = n_frames;
    cudaPointerAttributes a;
    // (the unit clips a tile at the plane's last line exactly but at its last column in units of 16 bytes —
    // measured: a 94-byte row got 96 — so rows must be whole multiples of 16 bytes: 1080p, 720p, 1440p, 1536x864 are)
    e.usable = out.aligned16 && c->encode_tiled && g.out_format <= B200VC_FMT_NV12 && g.out_w % 16 == 0 &&
               (g.out_format == B200VC_FMT_NV12 || (g.out_w / 2) % 16 == 0);
    if (e.usable && (cudaPointerGetAttributes(&a, out.p[0]) != cudaSuccess || a.type != cudaMemoryTypeDevice)) {
        cudaGetLastError();
        e.usable = false;                      // (mapped host memory: the kernel's own stores are what was measured there)
    }
    if (e.usable) {
        const int cw = (g.out_w + 1) / 2, ch = (g.out_h + 1) / 2, nv12 = g.out_format == B200VC_FMT_NV12;
        auto encode = [&](CUtensorMap *m, uint8_t *base, size_t pitch, int w, int h, int bw, int bh) {
            cuuint64_t dims[3] = { (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_frames };
            cuuint64_t strides[2] = { (cuuint64_t)pitch, (cuuint64_t)(out.frame_stride ? out.frame_stride : ru(pitch * h, 16)) };
            cuuint32_t box[3] = { (cuuint32_t)bw, (cuuint32_t)bh, 1 };
            cuuint32_t estr[3] = { 1, 1, 1 };
            return c->encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
        };
        e.usable = encode(&e.maps.y, out.p[0], out.pitch[0], g.out_w, g.out_h, K2_TILE_W, c->tile_h);
        if (e.usable && nv12) e.usable = encode(&e.maps.c1, out.p[1], out.pitch[1], 2 * cw, ch, K2_TILE_W, c->tile_h / 2);
        if (e.usable && !nv12)
            e.usable = encode(&e.maps.c1, out.p[1], out.pitch[1], cw, ch, K2_TILE_W / 2, c->tile_h / 2) &&
                       encode(&e.maps.c2, out.p[2], out.pitch[2], cw, ch, K2_TILE_W / 2, c->tile_h / 2);
    }
    constexpr size_t kMaxMaps = 64;
    if (c->out_maps.size() < kMaxMaps) {
        c->out_maps.push_back(e);              // (reserved at create: entries never move)
        return &c->out_maps.back();
    }
    c->out_maps[c->next_out_map] = e;
    const OutMapEntry *r = &c->out_maps[c->next_out_map];
    c->next_out_map = (c->next_out_map + 1) % kMaxMaps;
    return r;
}

void set_alignment(PlaneSet &ps)
{
    uintptr_t bits = (uintptr_t)ps.frame_stride;
    for (int i = 0; i < 3; i++)
        if (ps.p[i]) bits |= (uintptr_t)ps.p[i] | (uintptr_t)ps.pitch[i];
    ps.aligned16 = (bits & 15) == 0;
}

PlaneSet plane_set(uint8_t *base, const size_t off[3], const size_t pitch[3], size_t frame_stride, int fmt)
{
    PlaneSet ps;
    for (int i = 0; i < 3; i++) {
        ps.p[i] = i < out_planes(fmt) ? base + off[i] : nullptr;
        ps.pitch[i] = (long long)pitch[i];
    }
    ps.frame_stride = (long long)frame_stride;
    set_alignment(ps);
    return ps;
}

// add-borders: black (in the output's colorimetry: the matrix's answer to RGB 0,0,0) everywhere outside the
// picture's rectangle of n_frames output frames.
int fill_borders(b200vc_ctx *c, const PlaneSet &out, int n_frames, cudaStream_t st)
{
    const b200vc_config &g = c->cfg;
    const int (*m)[4] = matrix_rows(g.matrix, g.range);
    const int rnd = B200VC_SPEC_MATRIX_ROUND ? 128 : 0;
    const int nv12 = g.out_format == B200VC_FMT_NV12 ? 2 : 1, hs = chroma_hsub(g.out_format);
    BorderParams b;
    memset(&b, 0, sizeof b);
    b.out = out;  b.n_planes = c->n_planes;  b.n_frames = n_frames;
    for (int i = 0; i < c->n_planes; i++) {
        b.w[i] = (int)c->prow_bytes[i];  b.h[i] = c->prows[i];
        b.val[i] = (uint32_t)((m[i ? 1 : 0][3] + rnd) >> 8);
        b.cx[i] = i ? c->cont_x / hs * nv12 : c->cont_x;
        b.cy[i] = i ? c->cont_y / c->vsub : c->cont_y;
        b.cw[i] = i ? c->ccw * nv12 : c->kcfg.out_w;
        b.ch[i] = i ? c->cch : c->kcfg.out_h;
    }
    CU_TRY(c, launch_border_fill(b, st));
    c->stats.kernel_launches += 1;
    return B200VC_OK;
}

const RowGate kNoGate = { nullptr, 0, 0, nullptr };

// Enqueue the kernels for n_frames frames on `st`. Input rows [row0,row0+rows) (unscaled path) or
// whole frames (scaled paths). `gate` (one frame only): the frame is still arriving, see RowGate.
int enqueue_kernels(b200vc_ctx *c, const uint8_t *d_src, size_t src_pitch, size_t src_fstride,
                    const uint8_t *d_prev, size_t prev_pitch, size_t prev_fstride,
                    const PlaneSet &out_frame, uint8_t *d_dmg, size_t dmg_fstride,
                    int row0, int rows, int n_frames, const RowGate &gate, cudaStream_t st)
{
    const b200vc_config &g = c->kcfg;                  // output size = the picture's rectangle inside the frame
    PlaneSet out = out_frame;
    if (c->borders) {                                  // whose origin is where the planes start, for the kernels
        const int cx = c->cont_x / chroma_hsub(g.out_format) * (g.out_format == B200VC_FMT_NV12 ? 2 : 1);
        out.p[0] += (long long)c->cont_y * out.pitch[0] + c->cont_x;
        for (int i = 1; i < c->n_planes; i++) out.p[i] += (long long)(c->cont_y / c->vsub) * out.pitch[i] + cx;
    }
    set_alignment(out);
    const int site = (g.chroma_site & B200VC_SITE_H_COSITED) ? 1 : 0;
    if (c->path == 0) {
        ConvertParams p;
        memset(&p, 0, sizeof p);
        p.src = d_src + (size_t)row0 * src_pitch;  p.src_pitch = (long long)src_pitch;  p.src_frame_stride = (long long)src_fstride;
        if (d_prev && d_dmg) {
            p.prev = d_prev + (size_t)row0 * prev_pitch;  p.prev_pitch = (long long)prev_pitch;  p.prev_frame_stride = (long long)prev_fstride;
            p.damage = d_dmg + (size_t)(row0 / 16) * c->mb_w;  p.damage_frame_stride = (long long)dmg_fstride;
        }
        p.mb_w = c->mb_w;
        p.out = out;
        p.out.p[0] += (size_t)row0 * out.pitch[0];
        p.out.p[1] += (size_t)(row0 / c->vsub) * out.pitch[1];
        if (p.out.p[2]) p.out.p[2] += (size_t)(row0 / c->vsub) * out.pitch[2];
        p.w = g.in_w;  p.h = rows;  p.n_frames = n_frames;  p.mx = c->mx;  p.stream_out = 1;  p.gate = gate;
        p.in16 = in_bytes_per_pixel(g.in_format) == 2;
        CU_TRY(c, launch_convert(p, g.out_format, site, st));
        c->stats.kernel_launches += 1;
        return B200VC_OK;
    }
    if (c->path == 1) {
        const CUtensorMap *map = nullptr;
        int rc = get_tensor_map(c, d_src, src_pitch, src_fstride, n_frames, &map);
        if (rc) return rc;
        if (c->fused2x) {
            Fused2xParams q;
            memset(&q, 0, sizeof q);
            q.out = out;  q.in_w = g.in_w;  q.in_h = g.in_h;  q.inc_y = c->inc_y;  q.inc_c = c->inc_c;
            q.n_frames = n_frames;  q.mx = c->mx;  q.gate = gate;
            CU_TRY(c, launch_fused_2x(*map, q, g.out_format, site, g.scale_mode == B200VC_SCALE_BOX_2X2, st));
            c->stats.kernel_launches += 1;
            return B200VC_OK;
        }
        FusedScaleParams p;
        memset(&p, 0, sizeof p);
        p.out = out;  p.ft = c->ft;
        p.in_w = g.in_w;  p.in_h = g.in_h;  p.out_w = g.out_w;  p.out_h = g.out_h;
        p.box_w = c->box_w;  p.box_h = c->box_h;  p.tile_h = c->tile_h;  p.n_frames = n_frames;
        p.in16 = in_bytes_per_pixel(g.in_format) == 2;
        p.planes_off = (int)fused_scale_planes_off(c->box_w, c->box_h, c->tile_h, in_bytes_per_pixel(g.in_format));
        p.box2x2 = g.scale_mode == B200VC_SCALE_BOX_2X2;  p.pair_luma = c->pair_luma ? 1 : 0;  p.mx = c->mx;  p.gate = gate;
        // items per line pair of the window: the box's 4-column groups, or one group fewer when a
        // tile's first group is pure halo
        const uint32_t groups_x = (uint32_t)c->box_w / 4;
        const uint32_t items_x = groups_x, items_x1 = std::max(groups_x - 1, 1u);
        p.inv_items_x = (65536u + items_x - 1) / items_x;
        p.inv_items_x1 = (65536u + items_x1 - 1) / items_x1;
        const OutMapEntry *om = get_out_maps(c, out, n_frames);
        p.tma_out = om->usable && !c->no_tma_out;
        CU_TRY(c, launch_fused_scale(*map, om->maps, p, g.out_format, site, st));
        c->stats.kernel_launches += 1;
        return B200VC_OK;
    }
    // two-pass: K1 into the full-size intermediate (always planar), then the plane scaler. The
    // intermediate belongs to the context, the stream to the caller: whoever used it last, on whatever
    // stream, must be done before pass 1 overwrites it.
    if (c->ev_mid_valid) CU_TRY(c, cudaStreamWaitEvent(st, c->ev_mid, 0));
    for (int f0 = 0; f0 < n_frames; f0 += c->mid_frames) {
        const int nf = std::min(c->mid_frames, n_frames - f0);
        ConvertParams p;
        memset(&p, 0, sizeof p);
        p.src = d_src + (size_t)f0 * src_fstride;  p.src_pitch = (long long)src_pitch;  p.src_frame_stride = (long long)src_fstride;
        // (the intermediate has the output's chroma format, planar: NV12 is interleaved by pass 2)
        const int mid_fmt = g.out_format == B200VC_FMT_NV12 ? B200VC_FMT_I420 : g.out_format;
        p.out = plane_set(c->d_mid, c->mid_off, c->mid_pitch, c->mid_size, mid_fmt);
        p.w = g.in_w;  p.h = g.in_h;  p.n_frames = nf;  p.mx = c->mx;  p.gate = gate;
        p.in16 = in_bytes_per_pixel(g.in_format) == 2;
        CU_TRY(c, launch_convert(p, mid_fmt, site, st));
        if (c->plane_tiled) {
            PlaneTileParams t;
            memset(&t, 0, sizeof t);
            for (int i = 0; i < 3; i++) { t.src[i] = c->d_mid + c->mid_off[i];  t.src_pitch[i] = (long long)c->mid_pitch[i]; }
            t.src_frame_stride = (long long)c->mid_size;
            t.out = out;
            for (int i = 0; i < 3; i++) if (t.out.p[i]) t.out.p[i] += (size_t)f0 * out.frame_stride;
            t.k[0] = c->pt[0];  t.k[1] = c->pt[1];
            t.n_frames = nf;  t.box2x2 = g.scale_mode == B200VC_SCALE_BOX_2X2;
            CU_TRY(c, launch_plane_tile(t, g.out_format == B200VC_FMT_NV12 ? 1 : 0, st));   // planar, or interleaving U and V
            c->stats.kernel_launches += 2;
            continue;
        }
        PlaneScaleParams q;
        memset(&q, 0, sizeof q);
        for (int i = 0; i < 3; i++) { q.src[i] = c->d_mid + c->mid_off[i];  q.src_pitch[i] = (long long)c->mid_pitch[i]; }
        q.src_frame_stride = (long long)c->mid_size;
        q.out = out;
        for (int i = 0; i < 3; i++) if (q.out.p[i]) q.out.p[i] += (size_t)f0 * out.frame_stride;
        q.tab = c->tab;
        q.in_w = g.in_w;  q.in_h = g.in_h;  q.out_w = g.out_w;  q.out_h = g.out_h;  q.n_frames = nf;
        q.c_out_w = c->ccw;  q.c_out_h = c->cch;             // (of the picture, not of a bordered frame)
        q.box2x2 = g.scale_mode == B200VC_SCALE_BOX_2X2;
        CU_TRY(c, launch_plane_scale(q, g.out_format == B200VC_FMT_NV12 ? 1 : 0, st));
        c->stats.kernel_launches += 2;
    }
    CU_TRY(c, cudaEventRecord(c->ev_mid, st));
    c->ev_mid_valid = true;
    return B200VC_OK;
}

void destroy_locked(b200vc_ctx *c)
{
    cudaSetDevice(c->cfg.device);
    if (c->s_h2d) cudaStreamSynchronize(c->s_h2d);
    for (int i = 1; i < 4; i++) if (c->s_up[i]) { cudaStreamSynchronize(c->s_up[i]);  cudaStreamDestroy(c->s_up[i]); }
    if (c->s_k) cudaStreamSynchronize(c->s_k);
    if (c->s_d2h) cudaStreamSynchronize(c->s_d2h);
    if (c->ev_mid_valid) cudaEventSynchronize(c->ev_mid);     // a caller's stream may still be in pass 2
    for (Session &s : c->sessions) {
        cudaFree(s.d_in[0]);  cudaFree(s.d_in[1]);
        if (s.ev_kernel) cudaEventDestroy(s.ev_kernel);
    }
    for (Slot &s : c->slots) {
        cudaFree(s.d_out);
        cudaFree(s.d_flags);
        if (s.h_flag_src) cudaFreeHost(s.h_flag_src);
        if (s.h_in) cudaFreeHost(s.h_in);
        if (s.h_out) cudaFreeHost(s.h_out);
        if (s.h_map) cudaFreeHost(s.h_map);
        if (s.done) cudaEventDestroy(s.done);
    }
    for (cudaEvent_t e : c->ev_band_h2d) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_band_k) cudaEventDestroy(e);
    if (c->ev_mid) cudaEventDestroy(c->ev_mid);
    // (a range the caller has unmapped meanwhile makes this fail; there is nothing left to release then)
    for (const HostReg &r : c->registered) cudaHostUnregister(const_cast<uint8_t *>(r.base));
    cudaFree(c->d_taps);
    cudaFree(c->d_ftaps);
    cudaFree(c->d_ptaps);
    cudaFree(c->d_mid);
    if (c->s_h2d) cudaStreamDestroy(c->s_h2d);
    if (c->s_k) cudaStreamDestroy(c->s_k);
    if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
    cudaGetLastError();
}

int create_impl(b200vc_ctx *c)
{
    b200vc_config &g = c->cfg;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(c, B200VC_ERR_NO_DEVICE, "no CUDA device is visible; libb200vc has no CPU fallback");
    }
    if (g.device < 0 || g.device >= ndev)
        return fail(c, B200VC_ERR_INVALID_ARG, "device %d out of range (0..%d)", g.device, ndev - 1);
    CU_TRY(c, cudaSetDevice(g.device));
    CU_TRY(c, cudaGetDeviceProperties(&c->prop, g.device));
    if (c->prop.major != 10)
        return fail(c, B200VC_ERR_NO_DEVICE, "device %d (%s, sm_%d%d) is not sm_100: this library carries sm_100a code only",
                    g.device, c->prop.name, c->prop.major, c->prop.minor);

    c->mx = make_matrix(g.matrix, g.range, g.in_format);
    c->kcfg = g;
    if (g.content_w) {
        c->kcfg.out_w = g.content_w;  c->kcfg.out_h = g.content_h;
        c->cont_x = g.content_x;  c->cont_y = g.content_y;
        c->borders = g.content_w != g.out_w || g.content_h != g.out_h;
    }
    chroma_dims(g.out_format, c->kcfg.out_w, c->kcfg.out_h, c->ccw, c->cch);
    c->scaling = g.in_w != c->kcfg.out_w || g.in_h != c->kcfg.out_h;
    c->mb_w = (g.in_w + 15) / 16;
    c->mb_h = (g.in_h + 15) / 16;
    c->maps.reserve(64);
    c->out_maps.reserve(64);
    c->no_tma_out = getenv("B200VC_NO_TMA_STORE") != nullptr;          // measurements and tests only
    c->equal_bands = getenv("B200VC_EQUAL_BANDS") != nullptr;

    CU_TRY(c, cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
    CU_TRY(c, cudaStreamCreateWithFlags(&c->s_k, cudaStreamNonBlocking));
    CU_TRY(c, cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
    // Row-gated uploads alternate their chunks over two copy streams: consecutive copies of one stream
    // leave the link idle for ~9 us between them, copies of different streams ~6 us (measured; more streams
    // do not help further).
    c->n_up = (g.flags & B200VC_FLAG_ROW_GATE) ? 2 : 1;
    c->s_up[0] = c->s_h2d;
    for (int i = 1; i < c->n_up; i++) CU_TRY(c, cudaStreamCreateWithFlags(&c->s_up[i], cudaStreamNonBlocking));
    c->ev_band_h2d.resize(g.n_bands);
    c->ev_band_k.resize(g.n_bands);
    for (int i = 0; i < g.n_bands; i++) {
        CU_TRY(c, cudaEventCreateWithFlags(&c->ev_band_h2d[i], cudaEventDisableTiming));
        CU_TRY(c, cudaEventCreateWithFlags(&c->ev_band_k[i], cudaEventDisableTiming));
    }
    CU_TRY(c, cudaEventCreateWithFlags(&c->ev_mid, cudaEventDisableTiming));
    {
        // the row gate's flags are written by the copy stream itself where the driver offers that
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &fn, cudaEnableDefault, &qres) == cudaSuccess && fn &&
            qres == cudaDriverEntryPointSuccess)
            c->write_value = (WriteValue32Fn)fn;
        else
            cudaGetLastError();
    }

    // Library-owned device frames. Input rows are as long as the caller's when that keeps them 16-byte
    // aligned (a frame with the default pitch then arrives in one linear copy), else padded to 256 bytes.
    c->row_bytes = (size_t)g.in_w * in_bytes_per_pixel(g.in_format);
    c->n_planes = out_planes(g.out_format);
    c->vsub = chroma_vsub(g.out_format);
    chroma_dims(g.out_format, g.in_w, g.in_h, c->cw_in, c->ch_in);
    chroma_dims(g.out_format, g.out_w, g.out_h, c->cw_out, c->ch_out);
    c->prow_bytes[0] = (size_t)g.out_w;  c->prows[0] = g.out_h;
    c->prow_bytes[1] = g.out_format == B200VC_FMT_NV12 ? 2 * (size_t)c->cw_out : (size_t)c->cw_out;  c->prows[1] = c->ch_out;
    c->prow_bytes[2] = c->n_planes == 3 ? (size_t)c->cw_out : 0;  c->prows[2] = c->n_planes == 3 ? c->ch_out : 0;
    c->d_in_pitch = c->row_bytes % 16 == 0 ? c->row_bytes : ru(c->row_bytes, 256);
    c->d_in_size = ru(c->d_in_pitch * g.in_h, 256);
    {
        // Output frame: GstVideoInfo's default layout (b200vc_default_layout) when every pitch and plane
        // offset of it is a multiple of 16 — 1080p, 1440p, 720p, 4K are — so that a caller frame laid out
        // that way leaves in one linear copy; else rows padded to 256 bytes, copied plane by plane.
        size_t off[3], pit[3];
        const size_t total = b200vc_default_layout(g.out_format, g.out_w, g.out_h, off, pit);
        const int np = c->n_planes;
        bool compact = total != 0;
        for (int i = 0; i < np; i++) compact = compact && pit[i] % 16 == 0 && off[i] % 16 == 0;
        c->out_compact = compact;
        if (compact) {
            for (int i = 0; i < 3; i++) { c->d_off[i] = i < np ? off[i] : 0;  c->d_pitch[i] = i < np ? pit[i] : 0; }
            c->d_out_size = ru(total, 256);
        } else {
            const size_t ch = (size_t)c->ch_out;                                           // chroma lines
            const size_t yp = ru(g.out_w + 1, 256), cp = ru(c->prow_bytes[1], 256);
            c->d_pitch[0] = yp;  c->d_pitch[1] = cp;  c->d_pitch[2] = np == 3 ? cp : 0;
            c->d_off[0] = 0;  c->d_off[1] = yp * g.out_h;
            c->d_off[2] = np == 3 ? c->d_off[1] + cp * ch : 0;
            c->d_out_size = ru(c->d_off[1] + cp * ch * (np == 3 ? 2 : 1), 256);
        }
    }
    const size_t dmg_bytes = ru((size_t)c->mb_w * c->mb_h, 256);

    c->sessions.resize(g.max_sessions);
    for (Session &s : c->sessions) {
        CU_TRY(c, cudaMalloc(&s.d_in[0], c->d_in_size));
        CU_TRY(c, cudaMalloc(&s.d_in[1], c->d_in_size));
        CU_TRY(c, cudaEventCreateWithFlags(&s.ev_kernel, cudaEventDisableTiming));
    }
    c->slots.resize(std::max(g.n_inflight, g.max_sessions));
    for (Slot &s : c->slots) {
        CU_TRY(c, cudaMalloc(&s.d_out, c->d_out_size + dmg_bytes));
        s.d_dmg = s.d_out + c->d_out_size;
        CU_TRY(c, cudaMalloc(&s.d_flags, kMaxChunks * sizeof(uint32_t)));
        CU_TRY(c, cudaMemset(s.d_flags, 0, kMaxChunks * sizeof(uint32_t)));
        // pinned and mapped: the gate's time-out word (the last one), which a kernel writes over PCIe only
        // if a wait gave up
        CU_TRY(c, cudaHostAlloc(&s.h_flag_src, (kMaxChunks + 1) * sizeof(uint32_t), cudaHostAllocMapped));
        memset(s.h_flag_src, 0, (kMaxChunks + 1) * sizeof(uint32_t));
        if (g.flags & B200VC_FLAG_DAMAGE_MAP) CU_TRY(c, cudaMallocHost(&s.h_map, dmg_bytes));
        CU_TRY(c, cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        // add-borders: the library's own output frames get their black border once; the kernels only ever
        // write the picture's rectangle, and every download carries the whole frame
        if (c->borders) {
            int rc = fill_borders(c, plane_set(s.d_out, c->d_off, c->d_pitch, 0, g.out_format), 1, nullptr);
            if (rc) return rc;
        }
    }

    if (c->scaling) {
        const b200vc_config &k = c->kcfg;              // the scaler's geometry: source -> the picture's rectangle
        std::vector<Tap> hy, vy, hc, vc;
        if (g.scale_mode == B200VC_SCALE_BOX_2X2) {
            build_box_taps(k.out_w, hy);  build_box_taps(k.out_h, vy);
            build_box_taps(c->ccw, hc);  build_box_taps(c->cch, vc);       // BOX: even sizes only; chroma planes by format
        } else if (g.scale_mode == B200VC_SCALE_NEAREST) {
            build_nearest_taps(g.in_w, k.out_w, hy);  build_nearest_taps(g.in_h, k.out_h, vy);
            build_nearest_taps(c->cw_in, c->ccw, hc);  build_nearest_taps(c->ch_in, c->cch, vc);
        } else {
            build_h_taps(g.in_w, k.out_w, hy);  build_v_taps(g.in_h, k.out_h, vy);
            // chroma planes of odd frames round up, like the intermediate frame's own layout
            build_h_taps(c->cw_in, c->ccw, hc);  build_v_taps(c->ch_in, c->cch, vc);
        }
        const size_t n = hy.size() + vy.size() + hc.size() + vc.size();       // every size a multiple of 4
        CU_TRY(c, cudaMalloc(&c->d_taps, n * sizeof(Tap)));
        Tap *d = c->d_taps;
        c->tab.hy = d;  CU_TRY(c, cudaMemcpy(d, hy.data(), hy.size() * sizeof(Tap), cudaMemcpyHostToDevice));  d += hy.size();
        c->tab.vy = d;  CU_TRY(c, cudaMemcpy(d, vy.data(), vy.size() * sizeof(Tap), cudaMemcpyHostToDevice));  d += vy.size();
        c->tab.hc = d;  CU_TRY(c, cudaMemcpy(d, hc.data(), hc.size() * sizeof(Tap), cudaMemcpyHostToDevice));  d += hc.size();
        c->tab.vc = d;  CU_TRY(c, cudaMemcpy(d, vc.data(), vc.size() * sizeof(Tap), cudaMemcpyHostToDevice));

        // Output rows per tile of the general fused kernel: the height (one instantiation each) that
        // costs the fewest issue slots per frame (fused_frame_cost) among those whose window leaves room
        // for 6 CTAs per SM. With fewer the TMA waits and the barriers between the phases are no longer
        // hidden, and that outweighs what the count promises. Measured with forced heights
        // (tools/run_general.sh, fraction of the HBM roofline at 16 / 32 / 48 / 64 rows; CTAs per SM in
        // brackets): 4K -> 1440p 0.61 / 0.82 [8] / 0.81 [5] / 0.79; 1440p -> 1080p 0.53 / 0.72 [8] / 0.76 [6] / 0.71 [4];
        // 1080p -> 1536x864 0.51 / 0.70 / 0.75 [7] / 0.72 [5]; 4K -> 1080p 0.84 [8] / 0.83 [4] / 0.78 / 0.66;
        // 4K -> 720p 0.85 (8 rows) / 0.86 (16) [4] / 0.66 (32); the 2:3 upscale, whose window is small,
        // 0.26 / 0.38 / 0.42 / 0.46 [8].
        {
            static const int heights[] = { 8, 16, 32, 48, 64 };       // K2_TILE_HEIGHTS, k2_scale.cu
            auto ctas_per_sm = [&](int th) {
                int bw, bh;
                fused_box(k, hy, vy, hc, vc, th, bw, bh);
                const size_t cta = fused_scale_smem_bytes(bw, bh, th, in_bytes_per_pixel(k.in_format)) + fused_scale_static_bytes(th);
                if (bw > 256 || bh > 256 || cta > (size_t)c->prop.sharedMemPerBlockOptin) return 0;
                return (int)std::min<size_t>(8, (size_t)c->prop.sharedMemPerMultiprocessor / (cta + 1024));
            };
            auto usable = [&](int th) { return ctas_per_sm(th) > 0; };
            int best = 8;
            double best_cost = 0;
            for (int th : heights) {
                if (th > 8 && ctas_per_sm(th) < 6) continue;
                const double cost = fused_frame_cost(k, hy, vy, hc, vc, th);
                if (th == 8 || cost < best_cost) { best = th;  best_cost = cost; }
            }
            const char *force = getenv("B200VC_TILE_H");                       // measurements and tests only
            if (force)
                for (int th : heights)
                    if (atoi(force) == th && usable(th)) best = th;
            c->tile_h = best;
            fused_box(k, hy, vy, hc, vc, c->tile_h, c->box_w, c->box_h);
            c->smem_bytes = fused_scale_smem_bytes(c->box_w, c->box_h, c->tile_h, in_bytes_per_pixel(k.in_format));
        }
        const bool fits = c->box_w <= 256 && c->box_h <= 256 &&
                          c->smem_bytes + fused_scale_static_bytes(c->tile_h) <= (size_t)c->prop.sharedMemPerBlockOptin;
        // (odd sizes run the general fused kernel; the exact-2:1 kernel needs multiples of 4, BOX even sizes)
        // (the fused kernels scale horizontally first; a build with the other order of the spec takes the two-pass path)
        // (and they make 4:2:0 planes: Y42B/Y444 output takes the two-pass path too; 5-6-5 input is unpacked in the
        // general fused kernel's phase 1, the exact-2:1 kernel reads 4-byte pixels only)
        const bool fusable = g.out_format <= B200VC_FMT_NV12;
        c->path = (fits && fusable && !(g.flags & B200VC_FLAG_FORCE_TWO_PASS) && B200VC_SPEC_SCALE_H_FIRST) ? 1 : 2;
        const bool exact2 = g.in_w == 2 * k.out_w && g.in_h == 2 * k.out_h && (g.in_w % 4) == 0 && (g.in_h % 4) == 0;
        c->fused2x = c->path == 1 && exact2 && !(g.flags & B200VC_FLAG_FORCE_GENERIC_FUSED) && in_bytes_per_pixel(g.in_format) == 4 &&
                     (g.scale_mode == B200VC_SCALE_BOX_2X2 ||
                      (taps_are_2to1(hy, k.out_w, false) && taps_are_2to1(vy, k.out_h, true) &&
                       taps_are_2to1(hc, k.out_w / 2, false) && taps_are_2to1(vc, k.out_h / 2, true)));
        if (c->fused2x) {
            c->box_w = K2X_SRC_W + K2X_HALO;
            c->box_h = K2X_SRC_H;
            c->smem_bytes = (size_t)c->box_w * c->box_h * 4;
            c->inc_y = h_inc(g.in_w, k.out_w);
            c->inc_c = h_inc(g.in_w / 2, k.out_w / 2);
        }
        if (c->path == 1 && !c->fused2x) {
            FusedTapBlob blob;
            c->pair_luma = g.scale_mode != B200VC_SCALE_BOX_2X2 && luma_pairs_fit(k, hy);
            build_fused_taps(k, hy, vy, hc, vc, c->tile_h, c->box_w, c->pair_luma, blob);
            CU_TRY(c, cudaMalloc(&c->d_ftaps, blob.w.size() * sizeof(uint32_t)));
            CU_TRY(c, cudaMemcpy(c->d_ftaps, blob.w.data(), blob.w.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
            c->ft.tile_x = reinterpret_cast<const int2 *>(c->d_ftaps + blob.tile_x);
            c->ft.tile_y = reinterpret_cast<const int2 *>(c->d_ftaps + blob.tile_y);
            c->ft.col_rec = c->d_ftaps + blob.col_rec;
            c->ft.row_rec = c->d_ftaps + blob.row_rec;
        }
        if (c->path == 1) {
            void *fn = nullptr;
            cudaDriverEntryPointQueryResult qres;
            CU_TRY(c, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
            if (!fn || qres != cudaDriverEntryPointSuccess)
                return fail(c, B200VC_ERR_CUDA, "the driver does not export cuTensorMapEncodeTiled");
            c->encode_tiled = (PFN_cuTensorMapEncodeTiled_v12000)fn;
            if (!c->fused2x) CU_TRY(c, fused_scale_prepare((size_t)c->prop.sharedMemPerBlockOptin));
        } else {
            const size_t yp = ru(g.in_w + 1, 256), cp = ru((size_t)c->cw_in, 256), mch = (size_t)c->ch_in;
            c->mid_pitch[0] = yp;  c->mid_pitch[1] = cp;  c->mid_pitch[2] = cp;
            c->mid_off[0] = 0;  c->mid_off[1] = yp * g.in_h;  c->mid_off[2] = c->mid_off[1] + cp * mch;
            c->mid_size = ru(c->mid_off[2] + cp * mch, 256);
            c->mid_frames = std::max(1, std::min(g.max_sessions, 8));   // measured: groups of 8 beat groups of 2
            CU_TRY(c, cudaMalloc(&c->d_mid, c->mid_size * c->mid_frames));
            // pass 2 as 64x32 tiles out of shared memory when every tile's window fits (any ratio up to
            // about 8:1); larger ratios keep the gather kernel
            std::vector<uint32_t> w;
            size_t o0[4], o1[4];
            const bool allow_pair = g.scale_mode != B200VC_SCALE_BOX_2X2 && B200VC_SPEC_SCALE_H_FIRST;
            build_plane_tile_taps(hy, vy, g.in_h, k.out_w, k.out_h, allow_pair, c->pt[0], w, o0);
            build_plane_tile_taps(hc, vc, c->ch_in, c->ccw, c->cch, allow_pair, c->pt[1], w, o1);
            PlaneTileParams probe;
            memset(&probe, 0, sizeof probe);
            probe.k[0] = c->pt[0];  probe.k[1] = c->pt[1];
            const size_t pt_smem = plane_tile_smem_bytes(probe, g.out_format == B200VC_FMT_NV12 ? 1 : 0);
            if (pt_smem <= 64 * 1024) {
                CU_TRY(c, cudaMalloc(&c->d_ptaps, w.size() * sizeof(uint32_t)));
                CU_TRY(c, cudaMemcpy(c->d_ptaps, w.data(), w.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
                const size_t *o[2] = { o0, o1 };
                for (int i = 0; i < 2; i++) {
                    c->pt[i].tile_x = reinterpret_cast<const int *>(c->d_ptaps + o[i][0]);
                    c->pt[i].tile_y = reinterpret_cast<const int *>(c->d_ptaps + o[i][1]);
                    c->pt[i].col_rec = c->d_ptaps + o[i][2];
                    c->pt[i].row_rec = c->d_ptaps + o[i][3];
                }
                CU_TRY(c, plane_tile_prepare((size_t)c->prop.sharedMemPerBlockOptin));
                c->plane_tiled = true;
            }
        }
    }
    CU_TRY(c, cudaDeviceSynchronize());
    return B200VC_OK;
}

// Copy `rows` rows of `bytes` each between host buffers with different pitches.
void copy_rows(uint8_t *dst, size_t dpitch, const uint8_t *src, size_t spitch, size_t bytes, int rows)
{
    if (dpitch == bytes && spitch == bytes) {
        memcpy(dst, src, bytes * rows);
        return;
    }
    for (int r = 0; r < rows; r++) memcpy(dst + r * dpitch, src + r * spitch, bytes);
}

// One copy between host and device rows: linear when both sides are dense, 2-D otherwise.
cudaError_t copy_rows_async(void *dst, size_t dpitch, const void *src, size_t spitch, size_t bytes, int rows,
                            cudaMemcpyKind kind, cudaStream_t st)
{
    if (dpitch == bytes && spitch == bytes)
        return cudaMemcpyAsync(dst, src, bytes * rows, kind, st);
    return cudaMemcpy2DAsync(dst, dpitch, src, spitch, bytes, rows, kind, st);
}

int validate_frame(b200vc_ctx *c, const b200vc_frame *f)
{
    const b200vc_config &g = c->cfg;
    if (!f || !f->bgrx || !f->planes[0] || !f->planes[1] || (c->n_planes == 3 && !f->planes[2]))
        return fail(c, B200VC_ERR_INVALID_ARG, "frame, input or plane pointer is NULL");
    if (f->session < 0 || f->session >= g.max_sessions)
        return fail(c, B200VC_ERR_INVALID_ARG, "session %d out of range (max_sessions %d)", f->session, g.max_sessions);
    if (f->in_pitch < c->row_bytes)
        return fail(c, B200VC_ERR_INVALID_ARG, "input pitch %zu is smaller than a row (%zu bytes)", f->in_pitch, c->row_bytes);
    for (int i = 0; i < c->n_planes; i++)
        if (f->pitches[i] < c->prow_bytes[i])
            return fail(c, B200VC_ERR_INVALID_ARG, "an output pitch is smaller than its row");
    if (f->damage && !(g.flags & B200VC_FLAG_DAMAGE_MAP))
        return fail(c, B200VC_ERR_INVALID_ARG, "a damage map was requested but the context was created without B200VC_FLAG_DAMAGE_MAP");
    return B200VC_OK;
}

// Wait for a submitted frame and hand its slot back. The mutex is released while this thread waits
// for the frame's completion event and while it copies staged output to pageable caller buffers, so
// other sessions' threads submit and collect meanwhile; the slot stays SLOT_COLLECTING until done.
int collect_slot(b200vc_ctx *c, Lock &lk, Slot &s)
{
    if (s.state == SLOT_FREE) return B200VC_OK;
    if (s.state != SLOT_BUSY) {                           // another thread is on it: wait for that
        const uint64_t t = s.ticket;
        c->cv.wait(lk, [&] { return s.state == SLOT_FREE || s.state == SLOT_BUSY || s.ticket != t; });
        if (s.state != SLOT_BUSY || s.ticket != t) return B200VC_OK;
    }
    s.state = SLOT_COLLECTING;
    lk.unlock();
    const cudaError_t e = cudaEventSynchronize(s.done);
    size_t staged = 0;
    bool timed_out = false;
    if (e == cudaSuccess) {
        if (s.gated && s.h_flag_src[kMaxChunks]) { timed_out = true;  s.h_flag_src[kMaxChunks] = 0; }
        if (s.staged_out) {
            for (int i = 0; i < c->n_planes; i++) {
                copy_rows((uint8_t *)s.frame.planes[i], s.frame.pitches[i], s.h_out + c->d_off[i], c->d_pitch[i], c->prow_bytes[i], c->prows[i]);
                staged += c->prow_bytes[i] * c->prows[i];
            }
        }
        if (s.staged_map && s.frame.damage) {
            memcpy(s.frame.damage, s.h_map, (size_t)c->mb_w * c->mb_h);
            staged += (size_t)c->mb_w * c->mb_h;
        }
    }
    lk.lock();
    s.state = SLOT_FREE;
    c->stats.staged_out_bytes += staged;
    c->cv.notify_all();
    if (e != cudaSuccess)
        return fail(c, B200VC_ERR_CUDA, "waiting for a frame failed: %s", cudaGetErrorString(e));
    if (timed_out)
        return fail(c, B200VC_ERR_CUDA, "the row-gated kernel gave up waiting for its frame's copy (a host-to-device copy did not complete)");
    return B200VC_OK;
}

int collect_ticket(b200vc_ctx *c, Lock &lk, uint64_t ticket)
{
    for (Slot &s : c->slots)
        if (s.state != SLOT_FREE && s.ticket == ticket) return collect_slot(c, lk, s);
    return B200VC_OK;   // already collected (its slot was reused or collected before)
}

// A free slot for a new frame, in ring order; with none free, the oldest frame in flight is collected
// first (its owner's later collect finds it done).
int acquire_slot(b200vc_ctx *c, Lock &lk, Slot **out)
{
    for (;;) {
        const int n = (int)c->slots.size();
        for (int i = 0; i < n; i++) {
            Slot &s = c->slots[(c->next_slot + i) % n];
            if (s.state == SLOT_FREE) {
                c->next_slot = (c->next_slot + i + 1) % n;
                *out = &s;
                return B200VC_OK;
            }
        }
        Slot *oldest = nullptr;
        for (Slot &s : c->slots)
            if (s.state == SLOT_BUSY && (!oldest || s.ticket < oldest->ticket)) oldest = &s;
        if (oldest) {
            const int rc = collect_slot(c, lk, *oldest);
            if (rc) return rc;
        } else {
            c->cv.wait(lk);
        }
    }
}

void release_slot(b200vc_ctx *c, Slot &s)
{
    s.state = SLOT_FREE;
    c->cv.notify_all();
}

// A frame's rows [row0, row0+rows) from the caller's buffer to the session's device frame on the copy
// stream; pageable input goes through the slot's pinned staging buffer first (mutex released for that copy).
int upload_rows(b200vc_ctx *c, Lock &lk, Slot &slot, const b200vc_frame *f, uint8_t *d_cur, bool in_pinned, int row0, int rows,
                cudaStream_t st)
{
    const size_t row_bytes = c->row_bytes;
    const uint8_t *hsrc = (const uint8_t *)f->bgrx + (size_t)row0 * f->in_pitch;
    size_t hpitch = f->in_pitch;
    if (!in_pinned) {
        uint8_t *stage = slot.h_in + (size_t)row0 * row_bytes;
        lk.unlock();
        copy_rows(stage, row_bytes, hsrc, f->in_pitch, row_bytes, rows);
        lk.lock();
        hsrc = stage;
        hpitch = row_bytes;
        c->stats.staged_in_bytes += row_bytes * rows;
    }
    CU_TRY(c, copy_rows_async(d_cur + (size_t)row0 * c->d_in_pitch, c->d_in_pitch, hsrc, hpitch, row_bytes, rows,
                              cudaMemcpyHostToDevice, st));
    c->stats.h2d_bytes += row_bytes * rows;
    return B200VC_OK;
}

// Device frame -> host on the D2H stream, rows [pr0, pr0+prs) of plane i counted in that plane's rows.
int download_plane_rows(b200vc_ctx *c, Slot &slot, const b200vc_frame *f, bool out_pinned, int i, int pr0, int prs, size_t row_bytes)
{
    uint8_t *hdst = out_pinned ? (uint8_t *)f->planes[i] + (size_t)pr0 * f->pitches[i]
                               : slot.h_out + c->d_off[i] + (size_t)pr0 * c->d_pitch[i];
    const size_t hp = out_pinned ? f->pitches[i] : c->d_pitch[i];
    CU_TRY(c, copy_rows_async(hdst, hp, slot.d_out + c->d_off[i] + (size_t)pr0 * c->d_pitch[i], c->d_pitch[i], row_bytes, prs,
                              cudaMemcpyDeviceToHost, c->s_d2h));
    c->stats.d2h_bytes += row_bytes * prs;
    return B200VC_OK;
}

// Enqueue one frame. sync == false: the frame goes up whole, then the kernel, then the planes come
// back (batches and submit/collect: consecutive frames overlap on the three streams). sync == true
// (b200vc_convert, one frame, nothing to overlap with): on the unscaled path the frame goes up in bands
// of rows, each band's kernel launch chained to its copy by an event, and with page-locked output the
// kernels store the planes into the caller's buffers themselves, so that only the last band's work is
// left when the transfer ends. Measured on B200 (tools/sync_latency.py, 1080p): every extra piece a
// host-to-device copy is cut into costs 6-9 us of idle link, a whole frame's kernel only 6 us and its
// planes 60 us by copy engine or 80 us as the kernel's own stores — so four bands with kernel stores is
// the optimum (204 us; 233 us unpipelined; the link alone needs 153 us), two bands without them.
// B200VC_FLAG_ROW_GATE replaces the launch per band by ONE launch whose stripes wait for their chunk's
// flag (set by the copy stream behind each chunk); it reaches the scaled paths too, but the flag writes
// make it slower than events on this machine (216 us at its best), so it is not the default.
int submit_frame(b200vc_ctx *c, Lock &lk, const b200vc_frame *f, bool sync, uint64_t *ticket)
{
    int rc = validate_frame(c, f);
    if (rc) return rc;
    const b200vc_config &g = c->cfg;
    const size_t row_bytes = c->row_bytes;
    const int n_planes = c->n_planes;
    const size_t *out_row_bytes = c->prow_bytes;
    const int *out_rows = c->prows;
    const size_t map_bytes = (size_t)c->mb_w * c->mb_h;

    // where do the host sides live?
    if (host_kind(c, f->bgrx) < 0 || host_kind(c, f->planes[0]) < 0)
        return fail(c, B200VC_ERR_INVALID_ARG, "a device pointer was passed to a host-buffer entry point; "
                                               "frames already on the GPU go through b200vc_convert_device");
    const size_t in_span = f->in_pitch * (g.in_h - 1) + row_bytes;
    bool in_pinned = host_kind(c, f->bgrx, in_span) == 1;
    if (!in_pinned && (g.flags & B200VC_FLAG_REGISTER_HOST))
        in_pinned = ensure_registered(c, f->bgrx, in_span);
    bool out_pinned = true;
    for (int i = 0; i < n_planes; i++)
        out_pinned = out_pinned && host_kind(c, f->planes[i], f->pitches[i] * (out_rows[i] - 1) + out_row_bytes[i]) == 1;
    if (!out_pinned && (g.flags & B200VC_FLAG_REGISTER_HOST)) {
        // the planes of one buffer share pages: page-lock the range they span as one piece when they are
        // close together (one allocation), else plane by plane
        const uint8_t *lo = (const uint8_t *)f->planes[0], *hi = lo;
        size_t sum = 0;
        for (int i = 0; i < n_planes; i++) {
            const uint8_t *p = (const uint8_t *)f->planes[i];
            const size_t bytes = f->pitches[i] * (out_rows[i] - 1) + out_row_bytes[i];
            lo = std::min(lo, p);  hi = std::max(hi, p + bytes);  sum += bytes;
        }
        if ((size_t)(hi - lo) <= 2 * sum) {
            out_pinned = ensure_registered(c, lo, (size_t)(hi - lo));
        } else {
            out_pinned = true;
            for (int i = 0; i < n_planes; i++)
                out_pinned = ensure_registered(c, f->planes[i], f->pitches[i] * (out_rows[i] - 1) + out_row_bytes[i]) && out_pinned;
        }
    }
    const bool map_pinned = !f->damage || host_kind(c, f->damage, map_bytes) == 1;

    Slot *slot_p = nullptr;
    if ((rc = acquire_slot(c, lk, &slot_p)) != 0) return rc;
    Slot &slot = *slot_p;
    if (!in_pinned && !slot.h_in) CU_TRY(c, cudaMallocHost(&slot.h_in, c->d_in_size));
    if (!out_pinned && !slot.h_out) CU_TRY(c, cudaMallocHost(&slot.h_out, c->d_out_size));
    slot.state = SLOT_SUBMITTING;
    slot.ticket = c->next_ticket++;
    slot.frame = *f;
    slot.staged_out = false;
    slot.staged_map = false;
    slot.gated = false;
    if (ticket) *ticket = slot.ticket;
    // from here on an error hands the slot back
#define SUBMIT_TRY(expr) do { int rc_ = (expr); if (rc_) { release_slot(c, slot); return rc_; } } while (0)
#define SUBMIT_CU(call)                                                                                    \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            release_slot(c, slot);                                                                         \
            return fail(c, e_ == cudaErrorMemoryAllocation ? B200VC_ERR_NOMEM : B200VC_ERR_CUDA,           \
                        "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);       \
        }                                                                                                  \
    } while (0)

    Session &ss = c->sessions[f->session];

    // B200VC_FLAG_ZERO_COPY: K1 reads the page-locked frame over PCIe and writes the planes back itself
    // (not on contexts that keep previous frames for the damage map: those need the frame in HBM)
    if ((g.flags & B200VC_FLAG_ZERO_COPY) && !(g.flags & B200VC_FLAG_DAMAGE_MAP) && c->path == 0 && in_pinned && out_pinned && !c->borders) {
        void *d_src = nullptr, *d_pl[3] = { nullptr, nullptr, nullptr };
        bool ok = cudaHostGetDevicePointer(&d_src, const_cast<void *>(f->bgrx), 0) == cudaSuccess;
        for (int i = 0; i < n_planes && ok; i++) ok = cudaHostGetDevicePointer(&d_pl[i], f->planes[i], 0) == cudaSuccess;
        if (!ok) cudaGetLastError();
        ok = ok && aligned16(d_src) && f->in_pitch % 16 == 0;
        for (int i = 0; i < n_planes; i++) ok = ok && aligned16(d_pl[i]) && f->pitches[i] % 16 == 0;
        if (ok) {
            PlaneSet zout;
            for (int i = 0; i < 3; i++) { zout.p[i] = (uint8_t *)d_pl[i];  zout.pitch[i] = i < n_planes ? (long long)f->pitches[i] : 0; }
            zout.frame_stride = 0;
            SUBMIT_TRY(enqueue_kernels(c, (const uint8_t *)d_src, f->in_pitch, 0, nullptr, 0, 0, zout, nullptr, 0, 0, g.in_h, 1, kNoGate, c->s_k));
            SUBMIT_CU(cudaEventRecord(slot.done, c->s_k));
            c->stats.h2d_bytes += row_bytes * g.in_h;
            for (int i = 0; i < n_planes; i++) c->stats.d2h_bytes += out_row_bytes[i] * out_rows[i];
            c->stats.frames += 1;
            slot.state = SLOT_BUSY;
            c->cv.notify_all();
            return B200VC_OK;
        }
    }

    // (no stream write in the driver: bands. The gated kernels read 4-byte pixels and make 4:2:0 planes.)
    const bool gated = sync && (g.flags & B200VC_FLAG_ROW_GATE) && c->write_value &&
                       in_bytes_per_pixel(g.in_format) == 4 && g.out_format <= B200VC_FMT_NV12;

    uint8_t *d_cur = ss.d_in[ss.cur];
    const uint8_t *d_prev = (f->damage && ss.have_prev) ? ss.d_in[ss.cur ^ 1] : nullptr;
    PlaneSet out = plane_set(slot.d_out, c->d_off, c->d_pitch, 0, g.out_format);
    uint8_t *d_dmg = slot.d_dmg;

    // The kernels store planes (and map) straight into the caller's page-locked memory: on request
    // (B200VC_FLAG_ZERO_COPY_OUT), and by itself for the row-gated unscaled call, whose 16-byte row-segment
    // stores cross PCIe well and which has no later frame to hide a device-to-host copy behind.
    bool direct_out = false;
    // (Not with add-borders: the caller's buffer would need its border painted every frame; the library's
    // own frames carry theirs from set-up, and the download takes the whole frame.)
    if (((g.flags & B200VC_FLAG_ZERO_COPY_OUT) || (sync && c->path == 0)) && out_pinned && map_pinned && !c->borders) {
        void *d_pl[3] = { nullptr, nullptr, nullptr }, *d_map = nullptr;
        bool ok = true;
        for (int i = 0; i < n_planes && ok; i++)
            ok = cudaHostGetDevicePointer(&d_pl[i], f->planes[i], 0) == cudaSuccess && aligned16(d_pl[i]) && f->pitches[i] % 16 == 0;
        if (ok && f->damage) ok = cudaHostGetDevicePointer(&d_map, f->damage, 0) == cudaSuccess;
        if (!ok) cudaGetLastError();
        if (ok) {
            for (int i = 0; i < n_planes; i++) { out.p[i] = (uint8_t *)d_pl[i];  out.pitch[i] = (long long)f->pitches[i]; }
            if (f->damage) d_dmg = (uint8_t *)d_map;
            direct_out = true;
        }
    }

    // pieces of a synchronous frame: as negotiated, or (n_bands 0) what was measured best
    const int want = !sync ? 1 : c->auto_bands ? (direct_out ? 4 : 2) : g.n_bands;
    int bands = (sync && !gated && c->path == 0) ? want : 1;                      // launches chained by events
    bands = std::min(bands, (int)c->ev_band_h2d.size());
    bands = std::min(bands, std::max(1, c->mb_h));
    if (c->borders) bands = 1;                         // (the download is the whole frame, border included)

    // the copy engine must not overwrite a frame the previous kernel of this session still reads
    if (ss.ev_valid) {
        SUBMIT_CU(cudaStreamWaitEvent(c->s_h2d, ss.ev_kernel, 0));
        for (int i = 1; i < c->n_up && gated; i++) SUBMIT_CU(cudaStreamWaitEvent(c->s_up[i], ss.ev_kernel, 0));
    }
    if (f->damage && !d_prev)
        SUBMIT_CU(cudaMemsetAsync(d_dmg, 1, map_bytes, c->s_k));

    if (gated) {
        // ---- row gate: chunks of rows and their flags on the copy stream, then one launch ----------
        int chunks = std::min(std::min(want, kMaxChunks), std::max(1, c->mb_h));
        const int rpc = (int)ru((size_t)(g.in_h + chunks - 1) / chunks, 16);
        chunks = (g.in_h + rpc - 1) / rpc;
        if (++slot.epoch == 0) slot.epoch = 1;
        slot.h_flag_src[kMaxChunks] = 0;
        for (int k = 0; k < chunks; k++) {
            const int row0 = k * rpc, rows = std::min(rpc, g.in_h - row0);
            cudaStream_t up = c->s_up[k % c->n_up];
            SUBMIT_TRY(upload_rows(c, lk, slot, f, d_cur, in_pinned, row0, rows, up));
            // (a stream write carries a system-wide fence behind the copy before it; a small copy as the flag
            // gives a running kernel no such ordering — tried, and it let stripes start early)
            const CUresult r = c->write_value((CUstream)up, (CUdeviceptr)(uintptr_t)(slot.d_flags + k), slot.epoch, 0);
            if (r != CUDA_SUCCESS) {
                release_slot(c, slot);
                return fail(c, B200VC_ERR_CUDA, "cuStreamWriteValue32 failed with CUresult %d", (int)r);
            }
        }
        // every copy and every flag is in the queue: the kernel below cannot wait for one that never comes
        void *d_timeout = nullptr;
        SUBMIT_CU(cudaHostGetDevicePointer(&d_timeout, slot.h_flag_src + kMaxChunks, 0));
        const RowGate gate = { slot.d_flags, slot.epoch, rpc, (uint32_t *)d_timeout };
        slot.gated = true;
        SUBMIT_TRY(enqueue_kernels(c, d_cur, c->d_in_pitch, 0, d_prev, c->d_in_pitch, 0, out, d_prev ? d_dmg : nullptr, 0,
                                   0, g.in_h, 1, gate, c->s_k));
    } else {
        // ---- bands chained by events (one band: the whole frame) ---------------------------------------
        // Each band is half the one before it (36, 18, 9, 5 macroblock rows of 1080p's 68): what the call
        // waits for after the last byte has come up is the last band's kernel and its way back, so that band
        // is small, and a band's way back (less than half the bytes of its way up) still fits under the
        // upload of the next one. Measured against equal bands: profiles/r2_sync_latency_v2.txt.
        int mb_left = c->mb_h, mb_row0 = 0;
        for (int b = 0; b < bands && mb_left > 0; b++) {
            const int shares_left = (1 << std::min(bands - b, 16)) - 1;          // 2^(bands-b) - 1
            const int mb_rows = (c->equal_bands || bands > 8) ? (mb_left + (bands - b) - 1) / (bands - b)
                                               : std::max(1, (int)(((long long)mb_left << (bands - b - 1)) + shares_left - 1) / shares_left);
            const int row0 = mb_row0 * 16;
            const int rows = std::min(mb_rows * 16, g.in_h - row0);
            mb_row0 += mb_rows;  mb_left -= mb_rows;
            SUBMIT_TRY(upload_rows(c, lk, slot, f, d_cur, in_pinned, row0, rows, c->s_h2d));
            SUBMIT_CU(cudaEventRecord(c->ev_band_h2d[b], c->s_h2d));
            SUBMIT_CU(cudaStreamWaitEvent(c->s_k, c->ev_band_h2d[b], 0));
            if (c->path != 0) {
                SUBMIT_TRY(enqueue_kernels(c, d_cur, c->d_in_pitch, 0, nullptr, 0, 0, out, nullptr, 0, 0, g.in_h, 1, kNoGate, c->s_k));
                break;
            }
            SUBMIT_TRY(enqueue_kernels(c, d_cur, c->d_in_pitch, 0, d_prev, c->d_in_pitch, 0, out, d_prev ? d_dmg : nullptr, 0,
                                       row0, rows, 1, kNoGate, c->s_k));
            if (direct_out || bands == 1) continue;
            // this band's rows of every plane go back while the next band is still coming up
            SUBMIT_CU(cudaEventRecord(c->ev_band_k[b], c->s_k));
            SUBMIT_CU(cudaStreamWaitEvent(c->s_d2h, c->ev_band_k[b], 0));
            for (int i = 0; i < n_planes; i++)
                SUBMIT_TRY(download_plane_rows(c, slot, f, out_pinned, i, i ? row0 / c->vsub : row0,
                                               i ? (rows + c->vsub - 1) / c->vsub : rows, out_row_bytes[i]));
        }
    }
    SUBMIT_CU(cudaEventRecord(ss.ev_kernel, c->s_k));
    ss.ev_valid = true;

    // ---- the planes (unless the kernels stored them into host memory, or the bands took them) and the map
    SUBMIT_CU(cudaStreamWaitEvent(c->s_d2h, ss.ev_kernel, 0));      // "done" is recorded on the D2H stream
    if (direct_out) {
        for (int i = 0; i < n_planes; i++) c->stats.d2h_bytes += out_row_bytes[i] * out_rows[i];
        if (f->damage) c->stats.d2h_bytes += map_bytes;
    } else {
        if (gated || bands == 1) {
            // whole frame: one linear copy when the caller's frame is laid out like the device frame
            bool same_layout = c->out_compact && out_pinned;
            for (int i = 0; i < n_planes && same_layout; i++)
                same_layout = f->pitches[i] == c->d_pitch[i] && (uint8_t *)f->planes[i] == (uint8_t *)f->planes[0] + c->d_off[i];
            const size_t dense = c->d_off[n_planes - 1] + c->d_pitch[n_planes - 1] * (size_t)(out_rows[n_planes - 1] - 1) + out_row_bytes[n_planes - 1];
            if (same_layout) {
                SUBMIT_CU(cudaMemcpyAsync(f->planes[0], slot.d_out, dense, cudaMemcpyDeviceToHost, c->s_d2h));
                c->stats.d2h_bytes += dense;
            } else if (!out_pinned) {
                SUBMIT_CU(cudaMemcpyAsync(slot.h_out, slot.d_out, dense, cudaMemcpyDeviceToHost, c->s_d2h));
                c->stats.d2h_bytes += dense;
            } else {
                for (int i = 0; i < n_planes; i++)
                    SUBMIT_TRY(download_plane_rows(c, slot, f, true, i, 0, out_rows[i], out_row_bytes[i]));
            }
        }
        if (f->damage) {
            SUBMIT_CU(cudaMemcpyAsync(map_pinned ? f->damage : slot.h_map, slot.d_dmg, map_bytes, cudaMemcpyDeviceToHost, c->s_d2h));
            c->stats.d2h_bytes += map_bytes;
        }
        slot.staged_out = !out_pinned;
        slot.staged_map = f->damage && !map_pinned;
    }
    SUBMIT_CU(cudaEventRecord(slot.done, c->s_d2h));
#undef SUBMIT_TRY
#undef SUBMIT_CU

    if (g.flags & B200VC_FLAG_DAMAGE_MAP) {
        ss.cur ^= 1;               // this frame becomes the next one's "previous"; no copy
        ss.have_prev = true;
    }
    c->stats.frames += 1;
    slot.state = SLOT_BUSY;
    c->cv.notify_all();
    return B200VC_OK;
}

}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

int b200vc_version(void) { return B200VC_VERSION_MAJOR * 100 + B200VC_VERSION_MINOR; }

const char *b200vc_spec_string(void) { return B200VC_SPEC_STRING; }

const char *b200vc_status_string(int s)
{
    switch (s) {
    case B200VC_OK: return "ok";
    case B200VC_ERR_INVALID_ARG: return "invalid argument";
    case B200VC_ERR_UNSUPPORTED: return "caps not supported";
    case B200VC_ERR_NOT_NEGOTIATED: return "not negotiated";
    case B200VC_ERR_CUDA: return "CUDA error";
    case B200VC_ERR_NO_DEVICE: return "no usable sm_100 device";
    case B200VC_ERR_ALIGNMENT: return "pointer or pitch not 16-byte aligned";
    case B200VC_ERR_NOMEM: return "out of memory";
    case B200VC_ERR_BAD_TICKET: return "unknown ticket";
    default: return "unknown status";
    }
}

int b200vc_default_colorimetry(int out_h, int *matrix, int *range, int *chroma_site)
{
    if (out_h <= 0) return B200VC_ERR_INVALID_ARG;
    // >= 2160 lines -> bt2020 (GStreamer >= 1.18; the bundle the reference downloads is
    // recent), > 576 lines -> bt709, else bt601; limited range; chroma-site mpeg2 above 576 lines, else jpeg.
    const bool hd = out_h > 576;
    if (matrix) *matrix = out_h >= 2160 ? B200VC_MATRIX_BT2020 : hd ? B200VC_MATRIX_BT709 : B200VC_MATRIX_BT601;
    if (range) *range = B200VC_RANGE_LIMITED;
    if (chroma_site) *chroma_site = hd ? B200VC_SITE_H_COSITED : B200VC_SITE_NONE;
    return B200VC_OK;
}

size_t b200vc_default_layout(int format, int w, int h, size_t offsets[3], size_t pitches[3])
{
    if (w <= 0 || h <= 0 || !offsets || !pitches) return 0;
    const size_t h2 = ru(h, 2), yp = ru(w, 4);
    offsets[0] = 0;  pitches[0] = yp;  offsets[1] = yp * h2;
    if (format == B200VC_FMT_NV12) {
        pitches[1] = yp;  offsets[2] = 0;  pitches[2] = 0;
        return offsets[1] + yp * (h2 / 2);
    }
    if (format == B200VC_FMT_Y42B) {                       // chroma halved horizontally only; heights as they are
        const size_t cp = ru(w, 8) / 2;
        pitches[1] = pitches[2] = cp;
        offsets[1] = yp * h;  offsets[2] = offsets[1] + cp * h;
        return offsets[2] + cp * h;
    }
    if (format == B200VC_FMT_Y444) {
        pitches[1] = pitches[2] = yp;
        offsets[1] = yp * h;  offsets[2] = 2 * yp * h;
        return 3 * yp * h;
    }
    if (format != B200VC_FMT_I420) return 0;
    const size_t cp = ru(ru(w, 2) / 2, 4);
    pitches[1] = pitches[2] = cp;
    offsets[2] = offsets[1] + cp * (h2 / 2);
    return offsets[2] + cp * (h2 / 2);
}

int b200vc_fit_borders(int in_w, int in_h, int in_par_n, int in_par_d, int out_w, int out_h, int out_par_n, int out_par_d,
                       int *x, int *y, int *w, int *h)
{
    if (in_w < 1 || in_h < 1 || out_w < 2 || out_h < 2 || in_par_n < 1 || in_par_d < 1 || out_par_n < 1 || out_par_d < 1 ||
        !x || !y || !w || !h)
        return B200VC_ERR_INVALID_ARG;
    // gstvideoscale.c set_info, add-borders: the source's display aspect ratio in output pixels is num : den
    const unsigned long long num = (unsigned long long)in_w * in_par_n * out_par_d, den = (unsigned long long)in_h * in_par_d * out_par_n;
    *x = *y = 0;  *w = out_w;  *h = out_h;
    if (num * out_h == den * out_w) return B200VC_OK;                    // same shape: no border
    const unsigned long long to_h = (unsigned long long)out_w * den / num;
    if (to_h <= (unsigned long long)out_h) {                             // bars above and below
        *h = std::max(2, (int)to_h & ~1);
        *y = ((out_h - *h) / 2) & ~1;
        *w = out_w & ~1;                                                 // (an odd frame keeps its last column as border)
    } else {
        *h = out_h & ~1;                                                             // bars left and right
        const unsigned long long to_w = (unsigned long long)out_h * num / den;
        *w = std::max(2, (int)std::min<unsigned long long>(to_w, (unsigned long long)out_w) & ~1);
        *x = ((out_w - *w) / 2) & ~1;
    }
    return B200VC_OK;
}

int b200vc_damage_to_rects(const uint8_t *map, int frame_w, int frame_h, b200vc_rect *rects, int max_rects)
{
    if (!map || frame_w < 1 || frame_h < 1 || max_rects < 0 || (max_rects > 0 && !rects)) return -1;
    const int mb_w = (frame_w + 15) / 16, mb_h = (frame_h + 15) / 16;
    struct Open { int x0, x1, y0; };                       // runs of the previous row still growing downwards
    std::vector<Open> open, next;
    int n = 0;
    auto emit = [&](const Open &o, int y1) {
        if (n < max_rects) {
            b200vc_rect &r = rects[n];
            r.x = o.x0 * 16;  r.y = o.y0 * 16;
            r.w = std::min(o.x1 * 16, frame_w) - r.x;
            r.h = std::min(y1 * 16, frame_h) - r.y;
        }
        n++;
    };
    for (int y = 0; y <= mb_h; y++) {
        next.clear();
        if (y < mb_h) {
            const uint8_t *row = map + (size_t)y * mb_w;
            for (int x = 0; x < mb_w;) {
                if (!row[x]) { x++; continue; }
                int x1 = x;
                while (x1 < mb_w && row[x1]) x1++;
                int y0 = y;
                for (const Open &o : open)
                    if (o.x0 == x && o.x1 == x1) y0 = o.y0;   // same run as in the row above: keep growing
                next.push_back({ x, x1, y0 });
                x = x1;
            }
        }
        for (const Open &o : open) {                       // runs that did not continue are complete
            bool cont = false;
            for (const Open &q : next) cont = cont || (q.x0 == o.x0 && q.x1 == o.x1 && q.y0 == o.y0);
            if (!cont) emit(o, y);
        }
        open.swap(next);
    }
    return n;
}

int b200vc_check_config(const b200vc_config *cfg)
{
    std::string why;
    int rc = check_config(cfg, &why);
    if (rc) g_create_error = why;
    return rc;
}

int b200vc_create(const b200vc_config *cfg, b200vc_ctx **out)
{
    if (!out) return fail(nullptr, B200VC_ERR_INVALID_ARG, "out pointer is NULL");
    *out = nullptr;
    std::string why;
    int rc = check_config(cfg, &why);
    if (rc) return fail(nullptr, rc, "%s", why.c_str());
    b200vc_ctx *c = new (std::nothrow) b200vc_ctx();
    if (!c) return fail(nullptr, B200VC_ERR_NOMEM, "out of host memory");
    c->cfg = *cfg;
    if (c->cfg.max_sessions == 0) c->cfg.max_sessions = 1;
    if (c->cfg.n_inflight == 0) c->cfg.n_inflight = 3;
    // pieces of a synchronous convert: 0 = what was measured best for the frame at hand (submit_frame)
    c->auto_bands = c->cfg.n_bands == 0;
    if (c->auto_bands) c->cfg.n_bands = 4;
    memset(&c->stats, 0, sizeof c->stats);
    rc = create_impl(c);
    if (rc) {
        g_create_error = c->err;
        destroy_locked(c);
        delete c;
        return rc;
    }
    *out = c;
    return B200VC_OK;
}

int b200vc_destroy(b200vc_ctx *c)
{
    if (!c) return B200VC_ERR_INVALID_ARG;
    {
        Lock lk(c->mu);
        cudaSetDevice(c->cfg.device);
        for (Slot &s : c->slots) collect_slot(c, lk, s);
        destroy_locked(c);
    }
    delete c;
    return B200VC_OK;
}

const char *b200vc_last_error(const b200vc_ctx *c)
{
    return c ? c->err.c_str() : g_create_error.c_str();
}

int b200vc_get_info(const b200vc_ctx *c, b200vc_info *info)
{
    if (!c || !info) return B200VC_ERR_INVALID_ARG;
    memset(info, 0, sizeof *info);
    info->path = c->path;
    info->tile_w = c->path == 1 ? K2_TILE_W : K1_TILE_W;
    info->tile_h = c->path == 1 ? (c->fused2x ? K2_TILE_H : c->tile_h) : K1_TILE_H;
    info->fused_2to1 = c->fused2x ? 1 : 0;
    info->box_w = c->box_w;  info->box_h = c->box_h;
    info->smem_bytes = c->path == 1 ? (int)c->smem_bytes : 3 * 1024 + 32;
    info->sm_count = c->prop.multiProcessorCount;
    info->cc_major = c->prop.major;  info->cc_minor = c->prop.minor;
    snprintf(info->device_name, sizeof info->device_name, "%.63s", c->prop.name);
    return B200VC_OK;
}

int b200vc_get_stats(const b200vc_ctx *c, b200vc_stats *st)
{
    if (!c || !st) return B200VC_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    *st = c->stats;
    return B200VC_OK;
}

int b200vc_reset(b200vc_ctx *c, int session)
{
    if (!c) return B200VC_ERR_NOT_NEGOTIATED;
    std::lock_guard<std::mutex> lk(c->mu);
    if (session >= (int)c->sessions.size()) return fail(c, B200VC_ERR_INVALID_ARG, "session %d out of range", session);
    for (int i = 0; i < (int)c->sessions.size(); i++)
        if (session < 0 || session == i) c->sessions[i].have_prev = false;
    return B200VC_OK;
}

int b200vc_submit(b200vc_ctx *c, const b200vc_frame *f, uint64_t *ticket)
{
    if (!c) return B200VC_ERR_NOT_NEGOTIATED;
    Lock lk(c->mu);
    CU_TRY(c, cudaSetDevice(c->cfg.device));
    return submit_frame(c, lk, f, false, ticket);
}

int b200vc_collect(b200vc_ctx *c, uint64_t ticket)
{
    if (!c) return B200VC_ERR_NOT_NEGOTIATED;
    Lock lk(c->mu);
    CU_TRY(c, cudaSetDevice(c->cfg.device));
    if (ticket == 0 || ticket >= c->next_ticket) return fail(c, B200VC_ERR_BAD_TICKET, "ticket %llu was never issued", (unsigned long long)ticket);
    return collect_ticket(c, lk, ticket);
}

int b200vc_convert(b200vc_ctx *c, const void *bgrx, size_t in_pitch, void *const planes[3], const size_t pitches[3],
                   uint8_t *damage)
{
    return b200vc_convert_session(c, 0, bgrx, in_pitch, planes, pitches, damage);
}

int b200vc_convert_session(b200vc_ctx *c, int session, const void *bgrx, size_t in_pitch, void *const planes[3],
                           const size_t pitches[3], uint8_t *damage)
{
    if (!c) return B200VC_ERR_NOT_NEGOTIATED;
    Lock lk(c->mu);
    if (!planes || !pitches) return fail(c, B200VC_ERR_INVALID_ARG, "planes/pitches is NULL");
    CU_TRY(c, cudaSetDevice(c->cfg.device));
    b200vc_frame f;
    memset(&f, 0, sizeof f);
    f.bgrx = bgrx;  f.in_pitch = in_pitch;
    for (int i = 0; i < c->n_planes; i++) { f.planes[i] = planes[i];  f.pitches[i] = pitches[i]; }
    f.damage = damage;  f.session = session;
    uint64_t ticket = 0;
    int rc = submit_frame(c, lk, &f, true, &ticket);
    if (rc) return rc;
    return collect_ticket(c, lk, ticket);          // this frame only: other sessions' frames stay in flight
}

int b200vc_convert_batch(b200vc_ctx *c, const b200vc_frame *frames, int n)
{
    if (!c) return B200VC_ERR_NOT_NEGOTIATED;
    Lock lk(c->mu);
    if (!frames || n < 0) return fail(c, B200VC_ERR_INVALID_ARG, "frames is NULL or n < 0");
    CU_TRY(c, cudaSetDevice(c->cfg.device));
    int rc;
    std::vector<uint64_t> tickets((size_t)n, 0);
    for (int i = 0; i < n; i++)
        if ((rc = submit_frame(c, lk, &frames[i], false, &tickets[i])) != 0) return rc;
    for (int i = 0; i < n; i++)
        if ((rc = collect_ticket(c, lk, tickets[i])) != 0) return rc;
    return B200VC_OK;
}

int b200vc_host_alloc(size_t bytes, void **out)
{
    if (!out || !bytes) return B200VC_ERR_INVALID_ARG;
    *out = nullptr;
    cudaError_t e = cudaMallocHost(out, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_create_error = std::string("cudaMallocHost failed: ") + cudaGetErrorString(e);
        return e == cudaErrorMemoryAllocation ? B200VC_ERR_NOMEM : B200VC_ERR_CUDA;
    }
    return B200VC_OK;
}

int b200vc_host_free(void *p)
{
    if (!p) return B200VC_OK;
    return cudaFreeHost(p) == cudaSuccess ? B200VC_OK : B200VC_ERR_CUDA;
}

int b200vc_convert_device(b200vc_ctx *c, const b200vc_device_frames *d, void *cuda_stream)
{
    if (!c) return B200VC_ERR_NOT_NEGOTIATED;
    std::lock_guard<std::mutex> lk(c->mu);
    const b200vc_config &g = c->cfg;
    if (!d || !d->bgrx || !d->planes[0] || !d->planes[1] || (c->n_planes == 3 && !d->planes[2]) || d->n_frames < 1)
        return fail(c, B200VC_ERR_INVALID_ARG, "device frame set is NULL, has a NULL plane or n_frames < 1");
    if (d->n_frames > 65535) return fail(c, B200VC_ERR_INVALID_ARG, "at most 65535 frames per launch");
    if (d->damage && (!(g.flags & B200VC_FLAG_DAMAGE_MAP) || c->path != 0))
        return fail(c, B200VC_ERR_INVALID_ARG, "a damage map was requested but the context was created without B200VC_FLAG_DAMAGE_MAP");
    const int np = c->n_planes;
    bool ok = aligned16(d->bgrx) && d->in_pitch % 16 == 0 && d->in_frame_stride % 16 == 0 && d->out_frame_stride % 16 == 0;
    for (int i = 0; i < np; i++) ok = ok && aligned16(d->planes[i]) && d->pitches[i] % 16 == 0;
    if (d->prev) ok = ok && aligned16(d->prev) && d->prev_pitch % 16 == 0 && d->prev_frame_stride % 16 == 0;
    if (!ok) return fail(c, B200VC_ERR_ALIGNMENT, "device pointers, pitches and frame strides must be multiples of 16 bytes");
    if (d->in_pitch < c->row_bytes || d->pitches[0] < c->prow_bytes[0] || d->pitches[1] < c->prow_bytes[1] ||
        (np == 3 && d->pitches[2] < c->prow_bytes[2]))
        return fail(c, B200VC_ERR_INVALID_ARG, "a pitch is smaller than its row");
    if (d->n_frames > 1 && (d->in_frame_stride == 0 || d->out_frame_stride == 0))
        return fail(c, B200VC_ERR_INVALID_ARG, "frame strides are required for n_frames > 1");
    if (d->damage && d->n_frames > 1 && d->damage_frame_stride < (size_t)c->mb_w * c->mb_h)
        return fail(c, B200VC_ERR_INVALID_ARG, "damage_frame_stride %zu is smaller than a map (%d bytes): the frames' maps would overlap",
                    d->damage_frame_stride, c->mb_w * c->mb_h);
    if (d->damage && d->prev && d->n_frames > 1 && d->prev_frame_stride == 0)
        return fail(c, B200VC_ERR_INVALID_ARG, "frame strides are required for n_frames > 1");

    CU_TRY(c, cudaSetDevice(g.device));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    PlaneSet out;
    for (int i = 0; i < 3; i++) { out.p[i] = i < np ? (uint8_t *)d->planes[i] : nullptr;  out.pitch[i] = i < np ? (long long)d->pitches[i] : 0; }
    out.frame_stride = (long long)d->out_frame_stride;
    if (d->damage && !d->prev) {
        if (d->n_frames > 1 && d->damage_frame_stride != (size_t)c->mb_w * c->mb_h) {
            for (int i = 0; i < d->n_frames; i++)
                CU_TRY(c, cudaMemsetAsync(d->damage + (size_t)i * d->damage_frame_stride, 1, (size_t)c->mb_w * c->mb_h, st));
        } else {
            CU_TRY(c, cudaMemsetAsync(d->damage, 1, (size_t)c->mb_w * c->mb_h * d->n_frames, st));
        }
    }
    if (c->borders) {                                  // the caller's frames: their border is painted with every call
        int rcb = fill_borders(c, out, d->n_frames, st);
        if (rcb) return rcb;
    }
    int rc = enqueue_kernels(c, (const uint8_t *)d->bgrx, d->in_pitch, d->in_frame_stride,
                             (const uint8_t *)d->prev, d->prev_pitch, d->prev_frame_stride, out,
                             d->prev ? d->damage : nullptr, d->damage_frame_stride, 0, g.in_h, d->n_frames, kNoGate, st);
    if (rc) return rc;
    c->stats.frames += d->n_frames;
    return B200VC_OK;
}

}  // extern "C"
