// noalloc.cu — b200vc_convert_scale_device_static: convert + downscale without allocating anything.
//
// Who needs this. Every other scaling entry point works on a context, and creating a context
// allocates (device frames, tap tables, the intermediate frame). Some callers must not allocate:
//   * a streaming thread that has to follow a caps change in the middle of the stream (the browser
//     window was resized): cudaMalloc and cudaFree serialise against all work in flight on the
//     device, so one session renegotiating would stall every other session sharing that GPU;
//   * a process whose GPU memory has been handed, whole, to another component's pool at start-up,
//     so that a late cudaMalloc simply fails.
// For them this file carries the two-pass scaler (K1 into a full-size intermediate I420, then the
// plane scaler) with the intermediate in a fixed array that is part of the library image.
//
// The scratch array. `g_mid` below is a file-scope __device__ array of B200VC_STATIC_SCRATCH_BYTES
// (3 133 440 bytes: an I420 frame of up to 1920x1088 source pixels). Every GPU on which this
// library runs a kernel keeps its own copy resident until the process exits. In host code the name
// `g_mid` is only a handle, not a device address, so the address is looked up once per device with
// cudaGetSymbolAddress while that device is current, and the looked-up pointer is what the kernels
// receive. That look-up happens inside the first call made on each device; make that first call
// before starting any stream capture.
//
// Sharing. All calls on one device use the same array, and it keeps whatever the previous call
// left in it. Nothing in it has to start at zero: pass 1 writes every byte of the intermediate that
// pass 2 reads (the taps are clamped to the source frame), and frames of one call are processed one
// after another on the caller's stream. Two calls whose work overlaps on the GPU (different
// streams, or different threads on the same device) race on the array and corrupt each other's
// output; callers must order them.
//
// No cudaMalloc, cudaFree, cudaMallocHost or any other allocating routine is called here, and no
// tap tables are needed: the scaler below computes its taps from the frame sizes in registers, with
// the same functions (host_tables.h) the context-based path builds its tables from.
#include "../../include/b200vc.h"
#include "device_utils.cuh"
#include "host_tables.h"

#include <cstdarg>
#include <cstdio>
#include <mutex>

using namespace b200vc;

static __device__ __align__(256) uint8_t g_mid[B200VC_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint8_t *g_mid_ptr[kMaxDevices];            // per device: address of that device's copy of g_mid

struct StaticScaleParams {
    const uint8_t *src[3]; long long src_pitch[3];
    PlaneSet out;
    int in_w, in_h, out_w, out_h;
    long long inc_y, inc_c;                   // 16.16 horizontal steps, luma and chroma planes
    int box2x2;
};

// SURVEY B.2 taps computed in place (scale_h_tap / scale_v_tap of host_tables.h, indices clamped).
__device__ __forceinline__ void h_tap(long long inc, int x, int in_n, int &i0, int &i1, int &f)
{
    int i;
    scale_h_tap(inc, x, i, f);
    i0 = clamp_index(i, in_n);
    i1 = clamp_index((long long)i + 1, in_n);
}

__device__ __forceinline__ void v_tap(int y, int in_n, int out_n, int &i0, int &i1, int &p1)
{
    long long i;
    scale_v_tap(y, in_n, out_n, i, p1);
    i0 = clamp_index(i, in_n);
    i1 = clamp_index(i + 1, in_n);
}

__device__ __forceinline__ int v2(int a, int b, int p1)
{
    return a + (((b - a) * p1 + B200VC_SPEC_VTAP_HALF) >> B200VC_SPEC_VTAP_BITS);
}

__device__ __forceinline__ int h2(int a, int b, int f) { return (a * (256 - f) + b * f) >> 8; }

__device__ __forceinline__ int scale_px(const uint8_t *r0, const uint8_t *r1, int j0, int j1, int f, int p1, int box)
{
    if (box)
        return (r0[j0] + r0[j0 + 1] + r1[j0] + r1[j0 + 1] + 2) >> 2;
    if (B200VC_SPEC_SCALE_H_FIRST)
        return v2(h2(r0[j0], r0[j1], f), h2(r1[j0], r1[j1], f), p1);
    return h2(v2(r0[j0], r1[j0], p1), v2(r0[j1], r1[j1], p1), f);
}

// Pass 2 on the static intermediate: one thread per 4 output samples of a row; rows 0..out_h-1 are
// luma, the following out_h/2 rows are chroma (U and V by the same thread).
template <int FMT>
__global__ void __launch_bounds__(256) static_plane_scale_kernel(const StaticScaleParams p)
{
    const int x = (blockIdx.x * 64 + (threadIdx.x & 63)) * 4;
    const int ry = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (ry < p.out_h) {
        if (x >= p.out_w) return;
        int i0, i1, p1;
        if (p.box2x2) { i0 = 2 * ry; i1 = i0 + 1; p1 = 0; } else v_tap(ry, p.in_h, p.out_h, i0, i1, p1);
        const uint8_t *r0 = p.src[0] + (long long)i0 * p.src_pitch[0];
        const uint8_t *r1 = p.src[0] + (long long)i1 * p.src_pitch[0];
        uint8_t *dst = p.out.p[0] + (long long)ry * p.out.pitch[0] + x;
        for (int k = 0; k < 4 && x + k < p.out_w; k++) {
            int j0, j1, f;
            if (p.box2x2) { j0 = 2 * (x + k); j1 = j0 + 1; f = 0; } else h_tap(p.inc_y, x + k, p.in_w, j0, j1, f);
            dst[k] = (uint8_t)scale_px(r0, r1, j0, j1, f, p1, p.box2x2);
        }
        return;
    }
    const int cy = ry - p.out_h, cw = p.out_w >> 1;
    if (cy >= (p.out_h >> 1) || x >= cw) return;
    int i0, i1, p1;
    if (p.box2x2) { i0 = 2 * cy; i1 = i0 + 1; p1 = 0; } else v_tap(cy, p.in_h >> 1, p.out_h >> 1, i0, i1, p1);
#pragma unroll
    for (int pl = 1; pl <= 2; pl++) {
        const uint8_t *r0 = p.src[pl] + (long long)i0 * p.src_pitch[pl];
        const uint8_t *r1 = p.src[pl] + (long long)i1 * p.src_pitch[pl];
        for (int k = 0; k < 4 && x + k < cw; k++) {
            int j0, j1, f;
            if (p.box2x2) { j0 = 2 * (x + k); j1 = j0 + 1; f = 0; } else h_tap(p.inc_c, x + k, p.in_w >> 1, j0, j1, f);
            const uint8_t d = (uint8_t)scale_px(r0, r1, j0, j1, f, p1, p.box2x2);
            if (FMT == 0)
                p.out.p[pl][(long long)cy * p.out.pitch[pl] + x + k] = d;
            else
                p.out.p[1][(long long)cy * p.out.pitch[1] + 2 * (x + k) + (pl - 1)] = d;
        }
    }
}

int report(char *buf, size_t len, int code, const char *fmt, ...)
{
    if (buf && len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int b200vc_convert_scale_device_static(const b200vc_config *cfg, const b200vc_device_frames *d,
                                                  void *cuda_stream, char *errbuf, size_t errbuf_len)
{
    if (errbuf && errbuf_len) errbuf[0] = 0;
    if (!cfg || !d)
        return report(errbuf, errbuf_len, B200VC_ERR_INVALID_ARG, "config or frame set is NULL");
    if (b200vc_check_config(cfg) != B200VC_OK)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED, "caps refused: %s", b200vc_last_error(nullptr));
    if (cfg->in_w == cfg->out_w && cfg->in_h == cfg->out_h)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED,
                      "no scaling requested: the unscaled path needs no scratch, use b200vc_convert_device");
    if (cfg->out_format > B200VC_FMT_NV12 || cfg->in_format > B200VC_IN_XBGR)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED,
                      "this entry point converts 4-byte pixels to I420 or NV12; other formats go through a context");
    if (cfg->scale_mode == B200VC_SCALE_NEAREST)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED, "method=nearest goes through a context");
    if (cfg->content_w || cfg->content_h || cfg->content_x || cfg->content_y)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED, "borders (a content rectangle) go through a context");
    if (cfg->flags & B200VC_FLAG_DAMAGE_MAP)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED, "the damage map is produced on the unscaled path only");
    if ((cfg->in_w | cfg->in_h | cfg->out_w | cfg->out_h) & 1)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED, "this entry point scales even widths and heights only");

    // intermediate I420 at the source size, rows padded to 16 bytes for K1's vector stores
    const size_t yp = ((size_t)cfg->in_w + 15) / 16 * 16, cp = ((size_t)cfg->in_w / 2 + 15) / 16 * 16;
    const size_t need = yp * cfg->in_h + 2 * cp * (cfg->in_h / 2);
    if (need > (size_t)B200VC_STATIC_SCRATCH_BYTES)
        return report(errbuf, errbuf_len, B200VC_ERR_UNSUPPORTED,
                      "source %dx%d needs %zu bytes of intermediate frame; the static scratch holds %d bytes "
                      "(sources up to 1920x1088). Use a context (b200vc_create) for larger frames",
                      cfg->in_w, cfg->in_h, need, B200VC_STATIC_SCRATCH_BYTES);

    const int np = cfg->out_format == B200VC_FMT_NV12 ? 2 : 3;
    if (!d->bgrx || d->n_frames < 1 || !d->planes[0] || !d->planes[1] || (np == 3 && !d->planes[2]))
        return report(errbuf, errbuf_len, B200VC_ERR_INVALID_ARG, "NULL input or plane pointer, or n_frames < 1");
    if (d->damage || d->prev)
        return report(errbuf, errbuf_len, B200VC_ERR_INVALID_ARG, "prev/damage are not used on the scaled path");
    bool ok = ((uintptr_t)d->bgrx & 15) == 0 && d->in_pitch % 16 == 0 && d->in_frame_stride % 16 == 0;
    for (int i = 0; i < np; i++) ok = ok && ((uintptr_t)d->planes[i] & 3) == 0 && d->pitches[i] % 4 == 0;
    if (!ok)
        return report(errbuf, errbuf_len, B200VC_ERR_ALIGNMENT,
                      "input pointer/pitch/stride must be multiples of 16 bytes, output pointers/pitches of 4");
    const size_t cw = np == 2 ? (size_t)cfg->out_w : (size_t)cfg->out_w / 2;
    if (d->in_pitch < (size_t)cfg->in_w * 4 || d->pitches[0] < (size_t)cfg->out_w || d->pitches[1] < cw ||
        (np == 3 && d->pitches[2] < cw))
        return report(errbuf, errbuf_len, B200VC_ERR_INVALID_ARG, "a pitch is smaller than its row");
    if (d->n_frames > 1 && (d->in_frame_stride == 0 || d->out_frame_stride == 0))
        return report(errbuf, errbuf_len, B200VC_ERR_INVALID_ARG, "frame strides are required for n_frames > 1");

    // the calling thread's current device is the one we run on; cfg->device must name it
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return report(errbuf, errbuf_len, B200VC_ERR_NO_DEVICE, "cudaGetDevice failed: %s", cudaGetErrorString(e));
    if (dev != cfg->device || dev < 0 || dev >= kMaxDevices)
        return report(errbuf, errbuf_len, B200VC_ERR_INVALID_ARG,
                      "config names device %d but the calling thread's current device is %d", cfg->device, dev);

    // one-time set-up for this device: resolve the device address of g_mid (first call on a device)
    uint8_t *mid;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_mid_ptr[dev]) {
            void *sym = nullptr;
            e = cudaGetSymbolAddress(&sym, g_mid);
            if (e != cudaSuccess || !sym)
                return report(errbuf, errbuf_len, B200VC_ERR_CUDA, "cudaGetSymbolAddress(g_mid) failed on device %d: %s",
                              dev, cudaGetErrorString(e));
            g_mid_ptr[dev] = static_cast<uint8_t *>(sym);
        }
        mid = g_mid_ptr[dev];
    }

    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int site = (cfg->chroma_site & B200VC_SITE_H_COSITED) ? 1 : 0;
    const Matrix mx = make_matrix(cfg->matrix, cfg->range, cfg->in_format);
    for (int f = 0; f < d->n_frames; f++) {
        // pass 1: K1 at the source size into the static intermediate (always planar)
        ConvertParams p1 = {};
        p1.src = (const uint8_t *)d->bgrx + (size_t)f * d->in_frame_stride;
        p1.src_pitch = (long long)d->in_pitch;
        p1.out.p[0] = mid;  p1.out.p[1] = mid + yp * cfg->in_h;  p1.out.p[2] = p1.out.p[1] + cp * (cfg->in_h / 2);
        p1.out.pitch[0] = (long long)yp;  p1.out.pitch[1] = p1.out.pitch[2] = (long long)cp;
        p1.out.aligned16 = 1;                        // the scratch array and its 16-byte-padded rows
        p1.w = cfg->in_w;  p1.h = cfg->in_h;  p1.n_frames = 1;  p1.mx = mx;
        if ((e = launch_convert(p1, B200VC_FMT_I420, site, st)) != cudaSuccess)
            return report(errbuf, errbuf_len, B200VC_ERR_CUDA, "launching the conversion kernel failed: %s", cudaGetErrorString(e));
        // pass 2: scale the three planes out of the intermediate
        StaticScaleParams p2 = {};
        for (int i = 0; i < 3; i++) { p2.src[i] = p1.out.p[i];  p2.src_pitch[i] = p1.out.pitch[i]; }
        for (int i = 0; i < 3; i++) {
            p2.out.p[i] = i < np ? (uint8_t *)d->planes[i] + (size_t)f * d->out_frame_stride : nullptr;
            p2.out.pitch[i] = i < np ? (long long)d->pitches[i] : 0;
        }
        p2.in_w = cfg->in_w;  p2.in_h = cfg->in_h;  p2.out_w = cfg->out_w;  p2.out_h = cfg->out_h;
        p2.inc_y = scale_h_inc(cfg->in_w, cfg->out_w);  p2.inc_c = scale_h_inc(cfg->in_w / 2, cfg->out_w / 2);
        p2.box2x2 = cfg->scale_mode == B200VC_SCALE_BOX_2X2;
        const int rows = cfg->out_h + cfg->out_h / 2;
        dim3 grid(((cfg->out_w + 3) / 4 + 63) / 64, (rows + 3) / 4, 1);
        if (np == 3)
            static_plane_scale_kernel<0><<<grid, 256, 0, st>>>(p2);
        else
            static_plane_scale_kernel<1><<<grid, 256, 0, st>>>(p2);
        if ((e = cudaGetLastError()) != cudaSuccess)
            return report(errbuf, errbuf_len, B200VC_ERR_CUDA, "launching the scaling kernel failed: %s", cudaGetErrorString(e));
    }
    return B200VC_OK;
}
