/*
 * b200vc.h — C ABI of libb200vc.so: BGRx -> I420/NV12 conversion, fused bilinear downscale and
 * 16x16 damage map on NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for the one data-parallel hot path of
 * COx2/docker-nvidia-glx-desktop: the `videoconvert ! videoscale` stage of the selkies-gstreamer
 * pipeline on the WEBRTC_ENCODER=x264enc fallback. The reference holds no source for that stage;
 * it installs it as a binary (Dockerfile:467-471), sources its environment
 * (selkies-gstreamer-entrypoint.sh:8), starts the application that instantiates it
 * (selkies-gstreamer-entrypoint.sh:44-47) and selects the software-encoder path that uses it
 * (README.md:21,35,58; xgl.yml:61-63). Each entry point below names the GStreamer
 * GstBaseTransform/GstVideoFilter virtual method (SURVEY.md §8 b) whose work it carries; the
 * element that binds them is docker-nvidia-glx-desktop_b200/gst/gstb200convert.c and the binding
 * is shown in INTEGRATION.md.
 *
 * Plain C types only: no GLib, no CUDA, no torch types cross this boundary. There is no CPU
 * fallback: every entry point that does work fails with B200VC_ERR_NO_DEVICE / B200VC_ERR_CUDA
 * when no sm_100 device can be used.
 *
 * Threading: a context is internally locked; entry points may be called from any thread (the
 * GStreamer streaming thread is not ours), one frame stream per session index. The lock covers
 * booking and enqueueing only: waits for a frame and copies of pageable buffers run outside it.
 * b200vc_destroy must not run while another call on the context is in progress.
 *
 * b200vc_convert_device on the two-pass path (b200vc_info.path == 2) uses an intermediate frame that
 * belongs to the context; calls on different streams are ordered against each other on the GPU by the
 * library (each waits for the previous user of the intermediate), so they are safe but do not overlap.
 */
#ifndef B200VC_H
#define B200VC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200VC_VERSION_MAJOR 0
#define B200VC_VERSION_MINOR 2

/* ---- status codes ------------------------------------------------------------------------- */
enum {
    B200VC_OK = 0,
    B200VC_ERR_INVALID_ARG = 1,   /* NULL pointer, bad enum, size out of range, ...               */
    B200VC_ERR_UNSUPPORTED = 2,   /* caps this element does not accept (refuse at negotiation)    */
    B200VC_ERR_NOT_NEGOTIATED = 3,/* called before a successful create (GST_FLOW_NOT_NEGOTIATED)  */
    B200VC_ERR_CUDA = 4,          /* a CUDA call failed (GST_FLOW_ERROR); see b200vc_last_error   */
    B200VC_ERR_NO_DEVICE = 5,     /* no usable sm_100 GPU                                          */
    B200VC_ERR_ALIGNMENT = 6,     /* device pointers/pitches must be 16-byte multiples             */
    B200VC_ERR_NOMEM = 7,
    B200VC_ERR_BAD_TICKET = 8
};

/* ---- negotiated caps, as plain enums ------------------------------------------------------- */
/* src pad format: 4:2:0 planar and semi-planar (what the reference's encoders take), and the planar
 * 4:2:2 and 4:4:4 formats videoconvert also offers (chroma halved horizontally only / not at all) */
enum { B200VC_FMT_I420 = 0, B200VC_FMT_NV12 = 1, B200VC_FMT_Y42B = 2, B200VC_FMT_Y444 = 3 };
enum { B200VC_MATRIX_BT601 = 0, B200VC_MATRIX_BT709 = 1, B200VC_MATRIX_BT2020 = 2 };   /* colorimetry matrix */
enum { B200VC_RANGE_LIMITED = 0, B200VC_RANGE_FULL = 1 };           /* colorimetry range       */
/* chroma-site, as GstVideoChromaSite's flags: "jpeg"/none 0, "mpeg2" 1 (h-cosited), "cosited" 3 (h and v),
 * "dv" 7 (cosited, alternate lines). The horizontal flag selects the 1-2-1 filter; the vertical step of
 * videoconvert's 4:2:0 downsampling is the rounded mean of the two lines for every site (its v-cosited
 * variants point at the same function upstream, as recalled: SURVEY A.4), and ALT_LINE only matters for
 * interlaced video. So every value negotiates; bit 0 is what changes the output. */
enum { B200VC_SITE_NONE = 0, B200VC_SITE_H_COSITED = 1, B200VC_SITE_V_COSITED = 2, B200VC_SITE_ALT_LINE = 4,
       B200VC_SITE_COSITED = 3, B200VC_SITE_DV = 7 };
/* videoscale method: 0 = bilinear (its default), 2 = nearest-neighbour (method=nearest: the sample
 * under the centre of the output pixel; b200vc_spec.h holds the alignment convention). 1 is not a videoscale
 * method: the plain rounded 2x2 mean, kept as the survey's alternative reading of an exact 2:1 downscale. */
enum { B200VC_SCALE_GST_LQ = 0, B200VC_SCALE_BOX_2X2 = 1, B200VC_SCALE_NEAREST = 2 };
/* sink pad format: byte order of one pixel in memory; x is a pad or alpha byte and never matters
 * (BGRA, RGBA, ARGB, ABGR negotiate as their x twins: I420/NV12 carry no alpha). BGRx is what
 * ximagesrc delivers at the reference's default depth 24 (Dockerfile:206, entrypoint.sh:102). */
enum { B200VC_IN_BGRX = 0, B200VC_IN_RGBX = 1, B200VC_IN_XRGB = 2, B200VC_IN_XBGR = 3,
       /* 16 bits per pixel, little endian, 5-6-5: what ximagesrc delivers at the reference's CDEPTH=16
        * (Dockerfile:206, entrypoint.sh:102). RGB16: red in bits 11..15; BGR16: blue there. Unpacked to 8 bits
        * per channel as video-format.c does (high bits repeated into the low ones, b200vc_spec.h). These run
        * the conversion kernel; with scaling, the two-pass path. */
       B200VC_IN_RGB16 = 4, B200VC_IN_BGR16 = 5 };

/* flags */
#define B200VC_FLAG_DAMAGE_MAP      0x1u  /* keep the previous frame per session and emit the map */
#define B200VC_FLAG_FORCE_TWO_PASS  0x2u  /* scaling: convert to a full-size intermediate, then
                                             scale the planes (the general path), never the fused
                                             kernel                                               */
#define B200VC_FLAG_FORCE_GENERIC_FUSED 0x8u /* scaling: never the exact-2:1 kernel, always the
                                             general fused kernel (any ratio that fits a tile)    */
#define B200VC_FLAG_ZERO_COPY       0x10u /* unscaled path, page-locked buffers, no damage map: the
                                             kernel reads the frame from host memory and writes the
                                             planes to host memory itself (one launch, no copy
                                             engine, both PCIe directions busy at once)           */
#define B200VC_FLAG_ZERO_COPY_OUT   0x20u /* page-locked output planes: the kernels store the planes
                                             (and the damage map) straight into host memory; the
                                             input still arrives through the copy engine. Saves the
                                             D2H copies and their per-band set-up cost            */
#define B200VC_FLAG_REGISTER_HOST   0x4u  /* page-lock caller buffers on first sight instead of
                                             staging them through a pinned ring. For buffers that
                                             come from a small, long-lived pool (ximagesrc's XShm
                                             images): a range stays page-locked until 16 others
                                             have been seen or the context is destroyed, so memory
                                             that is freed and whose address is handed out again
                                             for something else must NOT be used with this flag
                                             (the GPU would still copy the old pages)             */
#define B200VC_FLAG_ROW_GATE        0x40u /* b200vc_convert (the synchronous call) launches ONE
                                             kernel while the frame is still crossing PCIe in
                                             n_bands chunks; every stripe waits for the flag the
                                             copy stream sets behind its chunk. Reaches the scaled
                                             paths too. Measured slower than the default (a launch
                                             per band, chained by events) on B200 hosts so far     */

typedef struct b200vc_config {
    int32_t device;        /* CUDA device ordinal                                                  */
    int32_t in_w, in_h;    /* sink caps: video/x-raw,format=BGRx (or in_format below)              */
    int32_t out_w, out_h;  /* src caps; equal to in_* = videoconvert only                          */
    int32_t out_format;    /* B200VC_FMT_*                                                         */
    int32_t matrix;        /* B200VC_MATRIX_*                                                      */
    int32_t range;         /* B200VC_RANGE_*                                                       */
    int32_t chroma_site;   /* B200VC_SITE_*                                                        */
    int32_t scale_mode;    /* B200VC_SCALE_*                                                       */
    uint32_t flags;        /* B200VC_FLAG_*                                                        */
    int32_t max_sessions;  /* independent desktop sessions served by this context (>=1; 0 -> 1)   */
    int32_t n_inflight;    /* frames that may be in flight between submit and collect (0 -> 3)    */
    int32_t n_bands;       /* bands of rows a synchronous convert is pipelined in (0 -> measured
                              best: 4 when the kernels store into page-locked planes, else 2)      */
    int32_t in_format;     /* B200VC_IN_* (0 = BGRx)                                               */
    /* videoscale add-borders: where the scaled picture sits inside the out_w x out_h frame when the
     * display aspect ratios of the two sides differ (letterbox / pillarbox; b200vc_fit_borders computes
     * it from the pixel-aspect-ratios the way videoscale's set_info does). The rest of the frame is
     * black in the output's colorimetry. All four 0 = the picture fills the frame. x, y, w, h even. */
    int32_t content_x, content_y, content_w, content_h;
} b200vc_config;

typedef struct b200vc_ctx b200vc_ctx;

/* One frame crossing the boundary in host memory. Mirrors what transform_frame receives: the
 * mapped input GstVideoFrame (data[0], stride[0]) and the mapped output frame (data[p],
 * stride[p]); strides are always honoured. planes[2]/pitches[2] unused for NV12. The
 * member is called bgrx after the reference's default format; it holds pixels of the negotiated in_format. */
typedef struct b200vc_frame {
    const void *bgrx;
    size_t in_pitch;
    void *planes[3];
    size_t pitches[3];
    uint8_t *damage;       /* ceil(in_w/16)*ceil(in_h/16) bytes, or NULL                           */
    int32_t session;       /* 0 .. max_sessions-1                                                  */
    int32_t reserved;
} b200vc_frame;

/* A batch of frames already resident in device memory (kernel-only path; also what a
 * memory:CUDAMemory caps feature would hand over). Frame i lives at base + i*stride. All
 * pointers, pitches and strides must be multiples of 16 bytes. */
typedef struct b200vc_device_frames {
    const void *bgrx;  size_t in_pitch;   size_t in_frame_stride;
    const void *prev;  size_t prev_pitch; size_t prev_frame_stride;   /* NULL: map is all ones    */
    void *planes[3];   size_t pitches[3]; size_t out_frame_stride;
    uint8_t *damage;   size_t damage_frame_stride;                    /* NULL: no map             */
    int32_t n_frames;
    int32_t reserved;
} b200vc_device_frames;

typedef struct b200vc_stats {
    uint64_t frames;            /* frames converted                                                */
    uint64_t kernel_launches;   /* kernels launched by this context                                */
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t staged_in_bytes;   /* bytes copied host->pinned because the caller's buffer was pageable */
    uint64_t staged_out_bytes;
} b200vc_stats;

typedef struct b200vc_info {
    int32_t path;               /* 0 convert only, 1 fused scale (TMA), 2 two-pass scale           */
    int32_t tile_w, tile_h;     /* output tile of the main kernel                                   */
    int32_t box_w, box_h;       /* fused scale: source window loaded per tile (pixels)              */
    int32_t smem_bytes;         /* dynamic + static shared memory of the main kernel                */
    int32_t sm_count;
    int32_t cc_major, cc_minor;
    int32_t fused_2to1;         /* path 1 only: the exact-2:1 register kernel is in use              */
    char device_name[64];
} b200vc_info;

/* ---- negotiation helpers (no GPU needed) -------------------------------------------------- */

/* GstVideoConvert/GstVideoScale fixate_caps + video-info.c set_default_colorimetry:
 * colorimetry and chroma-site implied by the output height when the caps leave them open. */
int b200vc_default_colorimetry(int out_h, int *matrix, int *range, int *chroma_site);

/* video-info.c default strides/offsets for I420/NV12. Returns the frame size. */
size_t b200vc_default_layout(int format, int w, int h, size_t offsets[3], size_t pitches[3]);

/* GstVideoScale::set_info with add-borders=true (its default): the rectangle of the out_w x out_h frame
 * the picture is scaled into so that its display aspect ratio (width x par_n : height x par_d) is kept,
 * centred, the rest being border. Sizes are rounded down to even and positions to even (4:2:0 chroma
 * sites; videoconvert aligns the position the same way, the size it leaves odd — a deviation, noted in
 * DESIGN.md). Equal aspect ratios give the whole frame. */
int b200vc_fit_borders(int in_w, int in_h, int in_par_n, int in_par_d, int out_w, int out_h, int out_par_n, int out_par_d,
                       int *x, int *y, int *w, int *h);

/* transform_caps/set_caps acceptance test: B200VC_OK, or the reason the caps are refused. */
int b200vc_check_config(const b200vc_config *cfg);

/* Damage-map consumer (no GPU needed): merge the dirty 16x16 macroblocks of a map into pixel
 * rectangles — runs of dirty macroblocks along a row, stacked while consecutive rows hold the same
 * run — clipped to the frame. What the element attaches downstream as GstVideoRegionOfInterestMeta
 * (x264enc does not read it today). Returns the number of rectangles the map needs;
 * at most max_rects of them are written. */
typedef struct b200vc_rect { int32_t x, y, w, h; } b200vc_rect;
int b200vc_damage_to_rects(const uint8_t *map, int frame_w, int frame_h, b200vc_rect *rects, int max_rects);

const char *b200vc_status_string(int status);
int b200vc_version(void);

/* The values of include/b200vc_spec.h this library was built with, as one token ("m0c1h1b6r1u1n1"): the
 * details of videoconvert/videoscale's arithmetic that SURVEY.md Appendix A/B could not verify against
 * GStreamer (matrix rounding term, last column of the cosited filter, order of the scaler's passes,
 * bits and rounding of its vertical tap). The oracle reports its own the same way; tests compare. */
const char *b200vc_spec_string(void);

/* ---- lifecycle ----------------------------------------------------------------------------- */

/* GstBaseTransform::start + GstVideoFilter::set_info: build every CUDA resource for the
 * negotiated caps. Renegotiation = destroy + create. */
int b200vc_create(const b200vc_config *cfg, b200vc_ctx **out);

/* GstBaseTransform::stop. Waits for work in flight. */
int b200vc_destroy(b200vc_ctx *ctx);

/* Text of the last failure on this context (ctx == NULL: of the last failed create on this
 * thread). What the element posts through GST_ELEMENT_ERROR. */
const char *b200vc_last_error(const b200vc_ctx *ctx);

int b200vc_get_info(const b200vc_ctx *ctx, b200vc_info *info);
int b200vc_get_stats(const b200vc_ctx *ctx, b200vc_stats *stats);

/* Stream discontinuity (flush, seek, caps change on the same geometry): forget the previous
 * frame of `session` (-1: all); its next damage map is all ones. */
int b200vc_reset(b200vc_ctx *ctx, int session);

/* ---- the hot path, host buffers ------------------------------------------------------------ */

/* GstVideoFilter::transform_frame -> gst_video_converter_frame: one frame of session 0,
 * synchronous; the planes are complete on return. damage may be NULL. Inside the call (unscaled
 * path) the frame goes to the GPU in n_bands bands of rows, each band's kernel launch chained to its
 * copy, and with page-locked output the kernels store the planes into the caller's buffers
 * themselves, so that when the last band has crossed PCIe only that band's work is left. The call
 * waits for its own frame only; other sessions' frames stay in flight. */
int b200vc_convert(b200vc_ctx *ctx, const void *bgrx, size_t in_pitch,
                   void *const planes[3], const size_t pitches[3], uint8_t *damage_or_null);

/* The same for session `session` (0 .. max_sessions-1): what each of several elements sharing one
 * context (one GPU, several desktop sessions) calls from its own streaming thread. The context's
 * lock is held only while work is enqueued, not while a thread waits for its frame or copies
 * pageable buffers, so the sessions' calls overlap. */
int b200vc_convert_session(b200vc_ctx *ctx, int session, const void *bgrx, size_t in_pitch,
                           void *const planes[3], const size_t pitches[3], uint8_t *damage_or_null);

/* One tick of n sessions (config "64 sessions, 8 per GPU"): frames are pipelined against each
 * other (H2D of frame i+1 under kernel and D2H of frame i); all complete on return. */
int b200vc_convert_batch(b200vc_ctx *ctx, const b200vc_frame *frames, int n);

/* One-frame-latency mode (GstBaseTransform::submit_input_buffer / generate_output): enqueue a
 * frame and return at once; b200vc_collect blocks until that frame's planes are complete.
 * The caller's buffers must stay valid and untouched until collect returns. */
int b200vc_submit(b200vc_ctx *ctx, const b200vc_frame *frame, uint64_t *ticket);
int b200vc_collect(b200vc_ctx *ctx, uint64_t ticket);

/* propose_allocation/decide_allocation: page-locked host memory for buffer pools either side of
 * the element, so frames cross PCIe without a staging copy. */
int b200vc_host_alloc(size_t bytes, void **out);
int b200vc_host_free(void *p);

/* ---- the hot path, device buffers ---------------------------------------------------------- */

/* Same transform on frames already in HBM, one batched launch on `cuda_stream` (a cudaStream_t
 * passed as void*; NULL = the default stream). Asynchronous: returns after enqueueing. */
int b200vc_convert_device(b200vc_ctx *ctx, const b200vc_device_frames *frames, void *cuda_stream);

/* ---- no-allocation variant ----------------------------------------------------------------- */

/* Maximum intermediate frame the entry point below can hold: an I420 frame of a 1920x1088 source. */
#define B200VC_STATIC_SCRATCH_BYTES 3133440

/* This entry point has no counterpart in the reference.
 *
 * Convert + scale (the general two-pass path: conversion at the source size into an intermediate
 * I420 frame, then the 2-tap plane scaler) for callers that must not allocate memory: a streaming
 * thread following a caps change mid-stream cannot afford cudaMalloc/cudaFree, which serialise
 * against everything in flight on the device and would stall the other sessions on that GPU; and a
 * process that gave all GPU memory to another pool at start-up has none left to ask for. It takes
 * no context (creating one allocates): `cfg` carries the negotiated caps (max_sessions, n_inflight,
 * n_bands are ignored; in and out sizes must differ and be even; no damage map), `frames` the device buffers as
 * for b200vc_convert_device (input pointer/pitch/stride multiples of 16 bytes, output pointers and
 * pitches multiples of 4), `cuda_stream` the stream to enqueue on. It calls no cudaMalloc, cudaFree
 * or any other allocating routine.
 *
 * The intermediate frame lives in a fixed device array inside the library, B200VC_STATIC_SCRATCH_BYTES
 * long, which every GPU that runs this library keeps until the process exits. Sources whose
 * intermediate frame would not fit (anything above 1920x1088 pixels) are refused with
 * B200VC_ERR_UNSUPPORTED and a message in `errbuf`. The work runs on the calling thread's current
 * device, which cfg->device must name. The first call on each device also looks up that device's
 * address of the array; make that first call before any stream capture begins. All calls on a device
 * share the one array: nothing in it needs clearing between calls (each call overwrites what it
 * reads), but two calls that overlap on the GPU race on it, so order them (one stream, or events).
 *
 * errbuf (may be NULL) receives the reason when the result is not B200VC_OK. */
int b200vc_convert_scale_device_static(const b200vc_config *cfg, const b200vc_device_frames *frames,
                                       void *cuda_stream, char *errbuf, size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* B200VC_H */
