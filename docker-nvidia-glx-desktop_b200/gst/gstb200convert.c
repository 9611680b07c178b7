/*
 * gstb200convert.c — GStreamer element "b200convert": drop-in for `videoconvert ! videoscale` on
 * the BGRx -> I420/NV12 path of the selkies-gstreamer pipeline (reference call sites:
 * /root/reference Dockerfile:467-471, selkies-gstreamer-entrypoint.sh:8,44-47, README.md:21).
 *
 * NEVER COMPILED HERE. GStreamer, GLib and their headers are absent from the build image and the
 * GPU box, so this file has not been through a compiler. It is a thin shim: every
 * piece of work is a call into libb200vc.so (include/b200vc.h), which is what is tested and timed.
 * Build, on a machine that has GStreamer:
 *   gcc -shared -fPIC gstb200convert.c -o libgstb200convert.so \
 *       $(pkg-config --cflags --libs gstreamer-video-1.0) -I../../include -L.. -lb200vc
 * then put it on GST_PLUGIN_PATH (the reference wipes the registry cache at every start,
 * selkies-gstreamer-entrypoint.sh:41, so it is picked up) and replace `videoconvert ! videoscale`
 * by `b200convert` in the pipeline string.
 */
#include <gst/gst.h>
#include <gst/video/video.h>
#include <gst/video/gstvideofilter.h>

#include "b200vc.h"

#define GST_TYPE_B200_CONVERT (gst_b200_convert_get_type())
G_DECLARE_FINAL_TYPE(GstB200Convert, gst_b200_convert, GST, B200_CONVERT, GstVideoFilter)

/* One frame in flight in one-frame-latency mode: both buffers stay mapped until the frame is collected. */
typedef struct {
    gboolean valid;
    GstBuffer *in, *out;
    GstVideoFrame in_frame, out_frame;
    guint64 ticket;
    guint8 *map;                  /* which of the element's two damage maps this frame writes, or NULL */
} GstB200Pending;

struct _GstB200Convert {
    GstVideoFilter parent;
    b200vc_ctx *ctx;
    GstAllocator *allocator;      /* page-locked host memory (b200vc_host_alloc) for the output pool */
    /* properties (object lock); read once per negotiation into the latched copies below */
    gint device_id;
    gboolean damage_map;
    gint n_bands;
    gboolean one_frame_latency;   /* submit frame N, hand out frame N-1: its D2H overlaps frame N's H2D */
    gboolean register_host;       /* page-lock upstream's buffers on first sight (B200VC_FLAG_REGISTER_HOST) */
    gint method;                  /* videoscale's method property: 0 nearest, 1 bilinear (its default); others are not offered */
    /* what the context in use was negotiated with: a property set in mid-stream takes effect at the next set_info */
    gboolean cur_damage_map, cur_one_frame_latency;
    gint frame_w, frame_h;
    guint8 *maps[2];              /* page-locked damage maps (b200vc_host_alloc), one per frame that can be in flight */
    guint next_map;
    GstB200Pending pending;
};

/* ---- allocator over b200vc_host_alloc: buffers the copy engines reach without staging -------------
 * Output buffers come from a GstVideoBufferPool configured with this allocator (decide_allocation), so
 * the planes x264enc reads are page-locked and the library copies device -> host straight into them.
 * The same allocator and a pool are offered upstream (propose_allocation); ximagesrc ignores the offer
 * (it wraps XShm images), which is what the register-host property (B200VC_FLAG_REGISTER_HOST) is for. */
typedef struct { GstMemory mem; gpointer data; } GstB200PinnedMemory;
typedef struct { GstAllocator parent; } GstB200PinnedAllocator;
typedef struct { GstAllocatorClass parent_class; } GstB200PinnedAllocatorClass;
GType gst_b200_pinned_allocator_get_type(void);
G_DEFINE_TYPE(GstB200PinnedAllocator, gst_b200_pinned_allocator, GST_TYPE_ALLOCATOR)

static GstMemory *gst_b200_pinned_alloc(GstAllocator *allocator, gsize size, GstAllocationParams *params)
{
    const gsize maxsize = size + params->prefix + params->padding;
    void *data = NULL;
    if (b200vc_host_alloc(maxsize, &data) != B200VC_OK)     /* page-aligned: satisfies any params->align */
        return NULL;
    GstB200PinnedMemory *m = g_slice_new0(GstB200PinnedMemory);
    gst_memory_init(GST_MEMORY_CAST(m), params->flags, allocator, NULL, maxsize, params->align, params->prefix, size);
    m->data = data;
    return GST_MEMORY_CAST(m);
}

static void gst_b200_pinned_free(GstAllocator *allocator, GstMemory *mem)
{
    GstB200PinnedMemory *m = (GstB200PinnedMemory *)mem;
    (void)allocator;
    b200vc_host_free(m->data);
    g_slice_free(GstB200PinnedMemory, m);
}

static gpointer gst_b200_pinned_map(GstMemory *mem, gsize maxsize, GstMapFlags flags)
{
    (void)maxsize; (void)flags;
    return ((GstB200PinnedMemory *)mem)->data;
}

static void gst_b200_pinned_unmap(GstMemory *mem) { (void)mem; }

static void gst_b200_pinned_allocator_class_init(GstB200PinnedAllocatorClass *klass)
{
    GstAllocatorClass *ac = GST_ALLOCATOR_CLASS(klass);
    ac->alloc = gst_b200_pinned_alloc;
    ac->free = gst_b200_pinned_free;
}

static void gst_b200_pinned_allocator_init(GstB200PinnedAllocator *self)
{
    GstAllocator *a = GST_ALLOCATOR_CAST(self);
    a->mem_type = "B200PinnedMemory";
    a->mem_map = gst_b200_pinned_map;
    a->mem_unmap = gst_b200_pinned_unmap;
    /* no mem_share/mem_copy: the defaults copy into system memory, which is always correct */
    GST_OBJECT_FLAG_SET(a, GST_ALLOCATOR_FLAG_CUSTOM_ALLOC);
}

G_DEFINE_TYPE(GstB200Convert, gst_b200_convert, GST_TYPE_VIDEO_FILTER)

GST_DEBUG_CATEGORY_STATIC(b200convert_debug);
#define GST_CAT_DEFAULT b200convert_debug

enum { PROP_0, PROP_DEVICE_ID, PROP_DAMAGE_MAP, PROP_N_BANDS, PROP_ONE_FRAME_LATENCY, PROP_REGISTER_HOST, PROP_METHOD };

static GstStaticPadTemplate sink_tmpl = GST_STATIC_PAD_TEMPLATE("sink", GST_PAD_SINK, GST_PAD_ALWAYS,
    GST_STATIC_CAPS("video/x-raw, format=(string){BGRx,BGRA,RGBx,RGBA,xRGB,ARGB,xBGR,ABGR,RGB16,BGR16}, width=(int)[2,16384], height=(int)[2,16384], "
                    "framerate=(fraction)[0/1,2147483647/1]"));
static GstStaticPadTemplate src_tmpl = GST_STATIC_PAD_TEMPLATE("src", GST_PAD_SRC, GST_PAD_ALWAYS,
    GST_STATIC_CAPS("video/x-raw, format=(string){I420,NV12,Y42B,Y444}, width=(int)[2,16384], height=(int)[2,16384], "
                    "framerate=(fraction)[0/1,2147483647/1]"));

/* transform_caps: what videoconvert and videoscale do combined — drop format, colorimetry and
 * chroma-site, open width/height/pixel-aspect-ratio to ranges; the pad templates then restrict. */
static GstCaps *
gst_b200_convert_transform_caps(GstBaseTransform *trans, GstPadDirection direction, GstCaps *caps, GstCaps *filter)
{
    GstCaps *out = gst_caps_new_empty();
    for (guint i = 0; i < gst_caps_get_size(caps); i++) {
        GstStructure *s = gst_structure_copy(gst_caps_get_structure(caps, i));
        gst_structure_remove_fields(s, "format", "colorimetry", "chroma-site", NULL);
        gst_structure_set(s, "width", GST_TYPE_INT_RANGE, 2, 16384, "height", GST_TYPE_INT_RANGE, 2, 16384, NULL);
        if (gst_structure_has_field(s, "pixel-aspect-ratio"))
            gst_structure_set(s, "pixel-aspect-ratio", GST_TYPE_FRACTION_RANGE, 1, G_MAXINT, G_MAXINT, 1, NULL);
        gst_caps_append_structure(out, s);
    }
    GstCaps *tmpl = gst_pad_get_pad_template_caps(direction == GST_PAD_SINK ? trans->srcpad : trans->sinkpad);
    GstCaps *res = gst_caps_intersect_full(out, tmpl, GST_CAPS_INTERSECT_FIRST);
    gst_caps_unref(out);
    gst_caps_unref(tmpl);
    if (filter) {
        GstCaps *f = gst_caps_intersect_full(filter, res, GST_CAPS_INTERSECT_FIRST);
        gst_caps_unref(res);
        res = f;
    }
    return res;
}

/* fixate_caps: takes ownership of othercaps; prefer the input size (no scaling unless asked). */
static GstCaps *
gst_b200_convert_fixate_caps(GstBaseTransform *trans, GstPadDirection direction, GstCaps *caps, GstCaps *othercaps)
{
    othercaps = gst_caps_truncate(gst_caps_make_writable(othercaps));
    GstStructure *in = gst_caps_get_structure(caps, 0), *out = gst_caps_get_structure(othercaps, 0);
    gint w, h;
    if (gst_structure_get_int(in, "width", &w)) gst_structure_fixate_field_nearest_int(out, "width", w);
    if (gst_structure_get_int(in, "height", &h)) gst_structure_fixate_field_nearest_int(out, "height", h);
    if (direction == GST_PAD_SINK) gst_structure_fixate_field_string(out, "format", "I420");
    /* pixel-aspect-ratio: keep the input's where the other side leaves it open (what videoscale prefers); a
     * different fixed one is honoured, and set_info then letterboxes (add-borders) */
    if (gst_structure_has_field(out, "pixel-aspect-ratio")) {
        gint pn = 1, pd = 1;
        gst_structure_get_fraction(in, "pixel-aspect-ratio", &pn, &pd);
        gst_structure_fixate_field_nearest_fraction(out, "pixel-aspect-ratio", pn, pd);
    }
    return gst_caps_fixate(othercaps);
}

static int in_format_of(const GstVideoInfo *i)
{
    switch (GST_VIDEO_INFO_FORMAT(i)) {
    case GST_VIDEO_FORMAT_RGBx: case GST_VIDEO_FORMAT_RGBA: return B200VC_IN_RGBX;
    case GST_VIDEO_FORMAT_xRGB: case GST_VIDEO_FORMAT_ARGB: return B200VC_IN_XRGB;
    case GST_VIDEO_FORMAT_xBGR: case GST_VIDEO_FORMAT_ABGR: return B200VC_IN_XBGR;
    case GST_VIDEO_FORMAT_RGB16: return B200VC_IN_RGB16;      /* ximagesrc at CDEPTH=16 */
    case GST_VIDEO_FORMAT_BGR16: return B200VC_IN_BGR16;
    default: return B200VC_IN_BGRX;   /* BGRx, BGRA: what ximagesrc delivers at depth 24/32 */
    }
}

static int matrix_of(const GstVideoInfo *i)
{
    switch (i->colorimetry.matrix) {
    case GST_VIDEO_COLOR_MATRIX_BT601: return B200VC_MATRIX_BT601;
    case GST_VIDEO_COLOR_MATRIX_BT2020: return B200VC_MATRIX_BT2020;
    default: return B200VC_MATRIX_BT709;
    }
}

/* Dirty macroblocks travel downstream as regions of interest on the output buffer. */
static void gst_b200_convert_attach_damage(GstB200Convert *self, GstBuffer *out, const guint8 *map)
{
    b200vc_rect rects[256];
    int n = b200vc_damage_to_rects(map, self->frame_w, self->frame_h, rects, 256);
    for (int i = 0; i < n && i < 256; i++)
        gst_buffer_add_video_region_of_interest_meta(out, "damage", rects[i].x, rects[i].y, rects[i].w, rects[i].h);
}

static void gst_b200_convert_free_maps(GstB200Convert *self)
{
    for (int i = 0; i < 2; i++) {
        if (self->maps[i]) b200vc_host_free(self->maps[i]);
        self->maps[i] = NULL;
    }
}

/* One-frame-latency mode: waits for the frame in flight, releases its input and returns its output
 * buffer (transfer full), or NULL if there is none or it failed. */
static GstBuffer *gst_b200_convert_finish_pending(GstB200Convert *self)
{
    GstB200Pending *p = &self->pending;
    GstBuffer *out;
    int rc;
    if (!p->valid)
        return NULL;
    rc = self->ctx ? b200vc_collect(self->ctx, p->ticket) : B200VC_ERR_NOT_NEGOTIATED;
    gst_video_frame_unmap(&p->out_frame);
    gst_video_frame_unmap(&p->in_frame);
    gst_buffer_unref(p->in);
    out = p->out;
    if (rc == B200VC_OK && p->map)
        gst_b200_convert_attach_damage(self, out, p->map);
    p->in = p->out = NULL;
    p->map = NULL;
    p->valid = FALSE;
    if (rc != B200VC_OK) {
        GST_ELEMENT_ERROR(self, LIBRARY, FAILED, ("b200vc_collect: %s", b200vc_status_string(rc)), ("%s", b200vc_last_error(self->ctx)));
        gst_buffer_unref(out);
        return NULL;
    }
    return out;
}

static void gst_b200_convert_drop_pending(GstB200Convert *self)
{
    GstBuffer *out = gst_b200_convert_finish_pending(self);
    if (out)
        gst_buffer_unref(out);
}

/* set_info: (re)build the converter for the negotiated caps. FALSE = not negotiated. */
static gboolean
gst_b200_convert_set_info(GstVideoFilter *filter, GstCaps *incaps, GstVideoInfo *in, GstCaps *outcaps, GstVideoInfo *out)
{
    GstB200Convert *self = GST_B200_CONVERT(filter);
    b200vc_config cfg = { 0 };
    int m, r, s;
    gboolean want_damage, want_register;
    b200vc_default_colorimetry(GST_VIDEO_INFO_HEIGHT(out), &m, &r, &s);
    GST_OBJECT_LOCK(self);                       /* property setters run on application threads */
    cfg.device = self->device_id;
    cfg.n_bands = self->n_bands;
    want_damage = self->damage_map;
    want_register = self->register_host;
    cfg.scale_mode = self->method == 0 ? B200VC_SCALE_NEAREST : B200VC_SCALE_GST_LQ;
    self->cur_one_frame_latency = self->one_frame_latency;
    GST_OBJECT_UNLOCK(self);
    cfg.in_w = GST_VIDEO_INFO_WIDTH(in);    cfg.in_h = GST_VIDEO_INFO_HEIGHT(in);
    cfg.out_w = GST_VIDEO_INFO_WIDTH(out);  cfg.out_h = GST_VIDEO_INFO_HEIGHT(out);
    cfg.in_format = in_format_of(in);
    /* videoscale add-borders (its default): keep the display aspect ratio, black bars around the picture */
    if (b200vc_fit_borders(cfg.in_w, cfg.in_h, GST_VIDEO_INFO_PAR_N(in), GST_VIDEO_INFO_PAR_D(in), cfg.out_w, cfg.out_h,
                           GST_VIDEO_INFO_PAR_N(out), GST_VIDEO_INFO_PAR_D(out), &cfg.content_x, &cfg.content_y,
                           &cfg.content_w, &cfg.content_h) != B200VC_OK ||
        (cfg.content_w == cfg.out_w && cfg.content_h == cfg.out_h))
        cfg.content_x = cfg.content_y = cfg.content_w = cfg.content_h = 0;
    switch (GST_VIDEO_INFO_FORMAT(out)) {
    case GST_VIDEO_FORMAT_NV12: cfg.out_format = B200VC_FMT_NV12; break;
    case GST_VIDEO_FORMAT_Y42B: cfg.out_format = B200VC_FMT_Y42B; break;
    case GST_VIDEO_FORMAT_Y444: cfg.out_format = B200VC_FMT_Y444; break;
    default: cfg.out_format = B200VC_FMT_I420; break;
    }
    /* honour the caps, do not assume */
    cfg.matrix = out->colorimetry.matrix == GST_VIDEO_COLOR_MATRIX_UNKNOWN ? m : matrix_of(out);
    cfg.range = out->colorimetry.range == GST_VIDEO_COLOR_RANGE_UNKNOWN ? r
              : (out->colorimetry.range == GST_VIDEO_COLOR_RANGE_0_255 ? B200VC_RANGE_FULL : B200VC_RANGE_LIMITED);
    cfg.chroma_site = out->chroma_site == GST_VIDEO_CHROMA_SITE_UNKNOWN ? s
                    : (((out->chroma_site & GST_VIDEO_CHROMA_SITE_H_COSITED) ? B200VC_SITE_H_COSITED : 0) |
                       ((out->chroma_site & GST_VIDEO_CHROMA_SITE_V_COSITED) ? B200VC_SITE_V_COSITED : 0) |
                       ((out->chroma_site & GST_VIDEO_CHROMA_SITE_ALT_LINE) ? B200VC_SITE_ALT_LINE : 0));
    self->cur_damage_map = want_damage && cfg.in_w == (cfg.content_w ? cfg.content_w : cfg.out_w) &&
                           cfg.in_h == (cfg.content_h ? cfg.content_h : cfg.out_h);          /* the map exists on the unscaled path */
    cfg.flags = self->cur_damage_map ? B200VC_FLAG_DAMAGE_MAP : 0;
    /* Off by default: page-locking on first sight is for upstream buffers that come from a small long-lived
     * pool (ximagesrc with XShm). Buffers malloc'd per frame would stay locked after they are freed, and an
     * address handed out again would be copied from stale pages; those are staged through a pinned ring. */
    if (want_register) cfg.flags |= B200VC_FLAG_REGISTER_HOST;
    cfg.n_inflight = self->cur_one_frame_latency ? 2 : 1;
    gst_b200_convert_drop_pending(self);          /* normally gone already: the CAPS event pushed it */
    if (self->ctx) { b200vc_destroy(self->ctx); self->ctx = NULL; }
    gst_b200_convert_free_maps(self);
    self->frame_w = cfg.in_w;  self->frame_h = cfg.in_h;
    if (self->cur_damage_map) {
        /* page-locked maps, one per frame that can be in flight: a pageable map would be staged */
        const size_t mb = (size_t)((cfg.in_w + 15) / 16) * ((cfg.in_h + 15) / 16);
        for (int i = 0; i < 2; i++) {
            void *p = NULL;
            if (b200vc_host_alloc(mb, &p) != B200VC_OK) {
                GST_ELEMENT_ERROR(self, RESOURCE, NO_SPACE_LEFT, ("page-locked damage map"), ("%s", b200vc_last_error(NULL)));
                gst_b200_convert_free_maps(self);
                return FALSE;
            }
            self->maps[i] = p;
        }
    }
    int rc = b200vc_create(&cfg, &self->ctx);
    if (rc != B200VC_OK) {
        GST_ELEMENT_ERROR(self, LIBRARY, INIT, ("b200vc_create: %s", b200vc_status_string(rc)), ("%s", b200vc_last_error(NULL)));
        return FALSE;
    }
    return TRUE;
}

/* transform_frame: one synchronous conversion; `out` is complete on return. */
static GstFlowReturn
gst_b200_convert_transform_frame(GstVideoFilter *filter, GstVideoFrame *in, GstVideoFrame *out)
{
    GstB200Convert *self = GST_B200_CONVERT(filter);
    if (!self->ctx) return GST_FLOW_NOT_NEGOTIATED;
    void *planes[3] = { GST_VIDEO_FRAME_PLANE_DATA(out, 0), GST_VIDEO_FRAME_PLANE_DATA(out, 1),
                        GST_VIDEO_FRAME_N_PLANES(out) > 2 ? GST_VIDEO_FRAME_PLANE_DATA(out, 2) : NULL };
    size_t pitches[3] = { (size_t)GST_VIDEO_FRAME_PLANE_STRIDE(out, 0), (size_t)GST_VIDEO_FRAME_PLANE_STRIDE(out, 1),
                          GST_VIDEO_FRAME_N_PLANES(out) > 2 ? (size_t)GST_VIDEO_FRAME_PLANE_STRIDE(out, 2) : 0 };
    guint8 *map = self->cur_damage_map ? self->maps[0] : NULL;     /* as negotiated, not as the property reads now */
    int rc = b200vc_convert(self->ctx, GST_VIDEO_FRAME_PLANE_DATA(in, 0), (size_t)GST_VIDEO_FRAME_PLANE_STRIDE(in, 0),
                            planes, pitches, map);
    if (rc == B200VC_OK && map)
        gst_b200_convert_attach_damage(self, out->buffer, map);
    if (rc != B200VC_OK) {
        GST_ELEMENT_ERROR(self, LIBRARY, FAILED, ("b200vc_convert: %s", b200vc_status_string(rc)), ("%s", b200vc_last_error(self->ctx)));
        return GST_FLOW_ERROR;
    }
    return GST_FLOW_OK;
}

/* generate_output, one-frame-latency mode: the queued input is enqueued (H2D, kernel, D2H on the
 * library's streams) and the call returns the frame before it, which has had a whole frame period to
 * finish. Every output buffer carries its own input's timestamps (prepare_output_buffer copied them).
 * Without the property this is the base class's synchronous path through transform_frame. */
static GstFlowReturn gst_b200_convert_generate_output(GstBaseTransform *trans, GstBuffer **outbuf)
{
    GstB200Convert *self = GST_B200_CONVERT(trans);
    GstVideoFilter *vf = (GstVideoFilter *)trans;
    GstBaseTransformClass *bclass = GST_BASE_TRANSFORM_CLASS(gst_b200_convert_parent_class);
    GstB200Pending next = { 0 };
    GstFlowReturn ret;
    b200vc_frame f = { 0 };
    int rc;
    if (!self->cur_one_frame_latency)
        return bclass->generate_output(trans, outbuf);
    *outbuf = NULL;
    next.in = trans->queued_buf;
    if (next.in == NULL)                     /* second call of the chain loop: nothing more to give */
        return GST_FLOW_OK;
    trans->queued_buf = NULL;
    if (!self->ctx) {
        gst_buffer_unref(next.in);
        return GST_FLOW_NOT_NEGOTIATED;
    }
    ret = bclass->prepare_output_buffer(trans, next.in, &next.out);
    if (ret != GST_FLOW_OK || next.out == NULL) {
        gst_buffer_unref(next.in);
        return ret != GST_FLOW_OK ? ret : GST_FLOW_ERROR;
    }
    if (!gst_video_frame_map(&next.in_frame, &vf->in_info, next.in, GST_MAP_READ)) {
        gst_buffer_unref(next.in);  gst_buffer_unref(next.out);
        return GST_FLOW_ERROR;
    }
    if (!gst_video_frame_map(&next.out_frame, &vf->out_info, next.out, GST_MAP_WRITE)) {
        gst_video_frame_unmap(&next.in_frame);
        gst_buffer_unref(next.in);  gst_buffer_unref(next.out);
        return GST_FLOW_ERROR;
    }
    f.bgrx = GST_VIDEO_FRAME_PLANE_DATA(&next.in_frame, 0);
    f.in_pitch = (size_t)GST_VIDEO_FRAME_PLANE_STRIDE(&next.in_frame, 0);
    for (guint i = 0; i < GST_VIDEO_FRAME_N_PLANES(&next.out_frame) && i < 3; i++) {
        f.planes[i] = GST_VIDEO_FRAME_PLANE_DATA(&next.out_frame, i);
        f.pitches[i] = (size_t)GST_VIDEO_FRAME_PLANE_STRIDE(&next.out_frame, i);
    }
    if (self->cur_damage_map) {                  /* frame N writes one map while frame N-1's is still being read */
        next.map = self->maps[self->next_map];
        self->next_map ^= 1;
        f.damage = next.map;
    }
    rc = b200vc_submit(self->ctx, &f, &next.ticket);
    if (rc != B200VC_OK) {
        GST_ELEMENT_ERROR(self, LIBRARY, FAILED, ("b200vc_submit: %s", b200vc_status_string(rc)), ("%s", b200vc_last_error(self->ctx)));
        gst_video_frame_unmap(&next.out_frame);
        gst_video_frame_unmap(&next.in_frame);
        gst_buffer_unref(next.in);  gst_buffer_unref(next.out);
        return GST_FLOW_ERROR;
    }
    next.valid = TRUE;
    *outbuf = gst_b200_convert_finish_pending(self);      /* frame N-1, or NULL on the first call */
    self->pending = next;
    return GST_FLOW_OK;
}

/* The frame in flight must leave before anything that ends or changes the stream. */
static gboolean gst_b200_convert_sink_event(GstBaseTransform *trans, GstEvent *event)
{
    GstB200Convert *self = GST_B200_CONVERT(trans);
    GstBuffer *out;
    switch (GST_EVENT_TYPE(event)) {
    case GST_EVENT_EOS:
    case GST_EVENT_CAPS:
    case GST_EVENT_SEGMENT:
        out = gst_b200_convert_finish_pending(self);
        if (out)
            gst_pad_push(trans->srcpad, out);
        break;
    case GST_EVENT_FLUSH_STOP:
        gst_b200_convert_drop_pending(self);
        break;
    default:
        break;
    }
    return GST_BASE_TRANSFORM_CLASS(gst_b200_convert_parent_class)->sink_event(trans, event);
}

/* The mode adds one frame period to the pipeline's latency. */
static gboolean gst_b200_convert_query(GstBaseTransform *trans, GstPadDirection direction, GstQuery *query)
{
    GstB200Convert *self = GST_B200_CONVERT(trans);
    GstVideoFilter *vf = (GstVideoFilter *)trans;
    gboolean ok = GST_BASE_TRANSFORM_CLASS(gst_b200_convert_parent_class)->query(trans, direction, query);
    if (ok && self->cur_one_frame_latency && GST_QUERY_TYPE(query) == GST_QUERY_LATENCY && GST_VIDEO_INFO_FPS_N(&vf->in_info) > 0) {
        gboolean live;
        GstClockTime min, max;
        const GstClockTime frame = gst_util_uint64_scale_int(GST_SECOND, GST_VIDEO_INFO_FPS_D(&vf->in_info),
                                                             GST_VIDEO_INFO_FPS_N(&vf->in_info));
        gst_query_parse_latency(query, &live, &min, &max);
        min += frame;
        if (max != GST_CLOCK_TIME_NONE)
            max += frame;
        gst_query_set_latency(query, live, min, max);
    }
    return ok;
}

/* A video buffer pool over the pinned allocator for `caps` (transfer full), or NULL. */
static GstBufferPool *gst_b200_convert_new_pool(GstB200Convert *self, GstCaps *caps, guint size, guint min, guint max,
                                                gboolean video_meta)
{
    GstBufferPool *pool = gst_video_buffer_pool_new();
    GstStructure *cfg = gst_buffer_pool_get_config(pool);
    GstAllocationParams params;
    gst_allocation_params_init(&params);
    gst_buffer_pool_config_set_params(cfg, caps, size, min, max);
    gst_buffer_pool_config_set_allocator(cfg, self->allocator, &params);
    if (video_meta)
        gst_buffer_pool_config_add_option(cfg, GST_BUFFER_POOL_OPTION_VIDEO_META);
    if (!gst_buffer_pool_set_config(pool, cfg)) {
        gst_object_unref(pool);
        return NULL;
    }
    return pool;
}

/* decide_allocation: where our output buffers come from. A pool proposed by downstream is kept
 * (x264enc proposes none); otherwise the output pool is ours, in page-locked memory. */
static gboolean gst_b200_convert_decide_allocation(GstBaseTransform *trans, GstQuery *query)
{
    GstB200Convert *self = GST_B200_CONVERT(trans);
    GstCaps *caps = NULL;
    GstVideoInfo info;
    gst_query_parse_allocation(query, &caps, NULL);
    if (caps && gst_video_info_from_caps(&info, caps) && gst_query_get_n_allocation_pools(query) == 0) {
        const gboolean vmeta = gst_query_find_allocation_meta(query, GST_VIDEO_META_API_TYPE, NULL);
        GstBufferPool *pool = gst_b200_convert_new_pool(self, caps, (guint)GST_VIDEO_INFO_SIZE(&info), 2, 0, vmeta);
        if (pool) {
            gst_query_add_allocation_pool(query, pool, (guint)GST_VIDEO_INFO_SIZE(&info), 2, 0);
            gst_object_unref(pool);
        }
    }
    return GST_BASE_TRANSFORM_CLASS(gst_b200_convert_parent_class)->decide_allocation(trans, query);
}

/* propose_allocation: tell upstream we read strides from GstVideoMeta, and offer it page-locked buffers. */
static gboolean gst_b200_convert_propose_allocation(GstBaseTransform *trans, GstQuery *decide_query, GstQuery *query)
{
    GstB200Convert *self = GST_B200_CONVERT(trans);
    GstCaps *caps = NULL;
    gboolean need_pool = FALSE;
    GstVideoInfo info;
    if (!GST_BASE_TRANSFORM_CLASS(gst_b200_convert_parent_class)->propose_allocation(trans, decide_query, query))
        return FALSE;
    if (decide_query == NULL)          /* passthrough: the base class forwarded the query */
        return TRUE;
    gst_query_add_allocation_meta(query, GST_VIDEO_META_API_TYPE, NULL);
    gst_query_parse_allocation(query, &caps, &need_pool);
    if (caps && gst_video_info_from_caps(&info, caps)) {
        GstAllocationParams params;
        gst_allocation_params_init(&params);
        gst_query_add_allocation_param(query, self->allocator, &params);
        if (need_pool) {
            GstBufferPool *pool = gst_b200_convert_new_pool(self, caps, (guint)GST_VIDEO_INFO_SIZE(&info), 2, 0, TRUE);
            if (pool) {
                gst_query_add_allocation_pool(query, pool, (guint)GST_VIDEO_INFO_SIZE(&info), 2, 0);
                gst_object_unref(pool);
            }
        }
    }
    return TRUE;
}

static void gst_b200_convert_finalize(GObject *obj)
{
    GstB200Convert *self = GST_B200_CONVERT(obj);
    gst_b200_convert_drop_pending(self);
    if (self->ctx) { b200vc_destroy(self->ctx); self->ctx = NULL; }
    gst_b200_convert_free_maps(self);
    gst_clear_object(&self->allocator);
    G_OBJECT_CLASS(gst_b200_convert_parent_class)->finalize(obj);
}

static gboolean gst_b200_convert_stop(GstBaseTransform *trans)
{
    GstB200Convert *self = GST_B200_CONVERT(trans);
    gst_b200_convert_drop_pending(self);
    if (self->ctx) { b200vc_destroy(self->ctx); self->ctx = NULL; }
    gst_b200_convert_free_maps(self);
    return TRUE;
}

static void gst_b200_convert_set_property(GObject *obj, guint id, const GValue *v, GParamSpec *ps)
{
    GstB200Convert *self = GST_B200_CONVERT(obj);
    GST_OBJECT_LOCK(self);
    switch (id) {
    case PROP_DEVICE_ID: self->device_id = g_value_get_int(v); break;
    case PROP_DAMAGE_MAP: self->damage_map = g_value_get_boolean(v); break;
    case PROP_N_BANDS: self->n_bands = g_value_get_int(v); break;
    case PROP_ONE_FRAME_LATENCY: self->one_frame_latency = g_value_get_boolean(v); break;   /* read at the next negotiation */
    case PROP_REGISTER_HOST: self->register_host = g_value_get_boolean(v); break;
    case PROP_METHOD: self->method = g_value_get_int(v); break;
    default: G_OBJECT_WARN_INVALID_PROPERTY_ID(obj, id, ps);
    }
    GST_OBJECT_UNLOCK(self);
}

static void gst_b200_convert_get_property(GObject *obj, guint id, GValue *v, GParamSpec *ps)
{
    GstB200Convert *self = GST_B200_CONVERT(obj);
    GST_OBJECT_LOCK(self);
    switch (id) {
    case PROP_DEVICE_ID: g_value_set_int(v, self->device_id); break;
    case PROP_DAMAGE_MAP: g_value_set_boolean(v, self->damage_map); break;
    case PROP_N_BANDS: g_value_set_int(v, self->n_bands); break;
    case PROP_ONE_FRAME_LATENCY: g_value_set_boolean(v, self->one_frame_latency); break;
    case PROP_REGISTER_HOST: g_value_set_boolean(v, self->register_host); break;
    case PROP_METHOD: g_value_set_int(v, self->method); break;
    default: G_OBJECT_WARN_INVALID_PROPERTY_ID(obj, id, ps);
    }
    GST_OBJECT_UNLOCK(self);
}

static void gst_b200_convert_class_init(GstB200ConvertClass *klass)
{
    GObjectClass *go = G_OBJECT_CLASS(klass);
    GstElementClass *el = GST_ELEMENT_CLASS(klass);
    GstBaseTransformClass *bt = GST_BASE_TRANSFORM_CLASS(klass);
    GstVideoFilterClass *vf = GST_VIDEO_FILTER_CLASS(klass);
    go->set_property = gst_b200_convert_set_property;
    go->get_property = gst_b200_convert_get_property;
    go->finalize = gst_b200_convert_finalize;
    g_object_class_install_property(go, PROP_DEVICE_ID, g_param_spec_int("device-id", "Device", "CUDA device ordinal", 0, 63, 0, G_PARAM_READWRITE));
    g_object_class_install_property(go, PROP_DAMAGE_MAP, g_param_spec_boolean("damage-map", "Damage map", "Emit the 16x16 damage map", FALSE, G_PARAM_READWRITE));
    g_object_class_install_property(go, PROP_N_BANDS, g_param_spec_int("n-bands", "Bands", "Row bands a frame is pipelined in", 0, 64, 0, G_PARAM_READWRITE));
    g_object_class_install_property(go, PROP_ONE_FRAME_LATENCY, g_param_spec_boolean("one-frame-latency", "One frame of latency",
        "Return frame N-1 while frame N is in flight (copies of consecutive frames overlap)", FALSE, G_PARAM_READWRITE));
    g_object_class_install_property(go, PROP_REGISTER_HOST, g_param_spec_boolean("register-host", "Page-lock upstream buffers",
        "Page-lock upstream's buffers on first sight instead of staging them; only for a small long-lived pool (ximagesrc XShm images)",
        FALSE, G_PARAM_READWRITE));
    g_object_class_install_property(go, PROP_METHOD, g_param_spec_int("method", "Scale method",
        "videoscale's method: 0 nearest-neighbour, 1 bilinear (2-tap)", 0, 1, 1, G_PARAM_READWRITE));
    gst_element_class_add_static_pad_template(el, &sink_tmpl);
    gst_element_class_add_static_pad_template(el, &src_tmpl);
    gst_element_class_set_static_metadata(el, "B200 BGRx converter/scaler", "Filter/Converter/Video/Scaler",
                                          "BGRx to I420/NV12 with bilinear scale on NVIDIA B200", "b200vc");
    bt->transform_caps = GST_DEBUG_FUNCPTR(gst_b200_convert_transform_caps);
    bt->fixate_caps = GST_DEBUG_FUNCPTR(gst_b200_convert_fixate_caps);
    bt->stop = GST_DEBUG_FUNCPTR(gst_b200_convert_stop);
    bt->generate_output = GST_DEBUG_FUNCPTR(gst_b200_convert_generate_output);
    bt->sink_event = GST_DEBUG_FUNCPTR(gst_b200_convert_sink_event);
    bt->query = GST_DEBUG_FUNCPTR(gst_b200_convert_query);
    bt->decide_allocation = GST_DEBUG_FUNCPTR(gst_b200_convert_decide_allocation);
    bt->propose_allocation = GST_DEBUG_FUNCPTR(gst_b200_convert_propose_allocation);
    bt->passthrough_on_same_caps = FALSE;
    vf->set_info = GST_DEBUG_FUNCPTR(gst_b200_convert_set_info);
    vf->transform_frame = GST_DEBUG_FUNCPTR(gst_b200_convert_transform_frame);
}

static void gst_b200_convert_init(GstB200Convert *self)
{
    self->ctx = NULL;
    self->allocator = g_object_new(gst_b200_pinned_allocator_get_type(), NULL);
    gst_object_ref_sink(self->allocator);
    self->device_id = 0;
    self->damage_map = FALSE;
    self->n_bands = 0;
    self->one_frame_latency = FALSE;
    self->register_host = FALSE;
    self->method = 1;
    self->cur_damage_map = self->cur_one_frame_latency = FALSE;
    self->frame_w = self->frame_h = 0;
    self->maps[0] = self->maps[1] = NULL;
    self->next_map = 0;
    self->pending.valid = FALSE;
    self->pending.map = NULL;
    self->pending.in = self->pending.out = NULL;
}

static gboolean plugin_init(GstPlugin *plugin)
{
    GST_DEBUG_CATEGORY_INIT(b200convert_debug, "b200convert", 0, "B200 BGRx converter");
    return gst_element_register(plugin, "b200convert", GST_RANK_NONE, GST_TYPE_B200_CONVERT);
}

#ifndef PACKAGE
#define PACKAGE "b200vc"
#endif
GST_PLUGIN_DEFINE(GST_VERSION_MAJOR, GST_VERSION_MINOR, b200convert, "BGRx to I420/NV12 on NVIDIA B200",
                  plugin_init, "0.1", "MPL", PACKAGE, "https://example.invalid/b200vc")
