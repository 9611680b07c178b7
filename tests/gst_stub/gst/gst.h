/* tests/gst_stub/gst/gst.h — NOT GStreamer. A hand-written stand-in for the slice of the GLib/GStreamer
 * 1.x C API that docker-nvidia-glx-desktop_b200/gst/gstb200convert.c touches, so that the shim can at
 * least go through a C compiler's front end here (tests/test_abi.py: gcc -fsyntax-only). Declarations
 * follow the public 1.x headers from memory; this checks the shim's C and its use of libb200vc, not its
 * conformance to the real headers — that still needs a machine with GStreamer. */
#ifndef B200VC_GST_STUB_H
#define B200VC_GST_STUB_H
#include <stddef.h>
#include <stdint.h>

typedef int gint;  typedef unsigned int guint;  typedef int gboolean;  typedef char gchar;
typedef void *gpointer;  typedef size_t gsize;  typedef uint8_t guint8;  typedef unsigned long GType;
#define TRUE 1
#define FALSE 0
#define G_MAXINT 2147483647
typedef struct _GValue GValue;  typedef struct _GParamSpec GParamSpec;
typedef struct _GObject { int ref; } GObject;
typedef struct _GObjectClass {
    void (*set_property)(GObject *, guint, const GValue *, GParamSpec *);
    void (*get_property)(GObject *, guint, GValue *, GParamSpec *);
    void (*finalize)(GObject *);
} GObjectClass;
typedef enum { G_PARAM_READWRITE = 3 } GParamFlags;
gpointer g_malloc(gsize n);  void g_free(gpointer p);
gpointer g_object_new(GType type, const gchar *first, ...);
void g_object_class_install_property(GObjectClass *c, guint id, GParamSpec *ps);
GParamSpec *g_param_spec_int(const gchar *, const gchar *, const gchar *, gint, gint, gint, GParamFlags);
GParamSpec *g_param_spec_boolean(const gchar *, const gchar *, const gchar *, gboolean, GParamFlags);
gint g_value_get_int(const GValue *);  gboolean g_value_get_boolean(const GValue *);
void g_value_set_int(GValue *, gint);  void g_value_set_boolean(GValue *, gboolean);
gpointer g_slice_alloc0(gsize n);  void g_slice_free1(gsize n, gpointer p);
#define g_slice_new0(T) ((T *)g_slice_alloc0(sizeof(T)))
#define g_slice_free(T, p) g_slice_free1(sizeof(T), (p))
#define G_OBJECT_CLASS(k) ((GObjectClass *)(k))
#define G_OBJECT_WARN_INVALID_PROPERTY_ID(o, id, ps) ((void)(o), (void)(id), (void)(ps))
/* GObject boilerplate, reduced to what a final type needs */
#define G_DECLARE_FINAL_TYPE(T, t, MOD, NAME, Parent) \
    GType t##_get_type(void); typedef struct _##T T; typedef struct { Parent##Class parent_class; } T##Class; \
    static inline T *MOD##_##NAME(gpointer p) { return (T *)p; }
#define G_DEFINE_TYPE(T, t, PARENT_TYPE) \
    static void t##_class_init(T##Class *klass); static void t##_init(T *self); \
    static gpointer t##_parent_class = NULL; \
    GType t##_get_type(void) { (void)t##_class_init; (void)t##_init; (void)t##_parent_class; return (GType)(PARENT_TYPE); }

/* ---- core GStreamer ------------------------------------------------------------------------------ */
typedef struct _GstObject { GObject object; guint flags; } GstObject;
typedef struct _GstCaps GstCaps;  typedef struct _GstStructure GstStructure;  typedef struct _GstPad GstPad;
typedef struct _GstQuery GstQuery;  typedef struct _GstBuffer GstBuffer;  typedef struct _GstPlugin GstPlugin;
typedef struct _GstBufferPool GstBufferPool;  typedef struct _GstEvent { int type; } GstEvent;
typedef uint64_t guint64;  typedef guint64 GstClockTime;
#define GST_SECOND ((GstClockTime)1000000000)
#define GST_CLOCK_TIME_NONE ((GstClockTime)-1)
typedef enum { GST_EVENT_UNKNOWN, GST_EVENT_FLUSH_STOP, GST_EVENT_CAPS, GST_EVENT_SEGMENT, GST_EVENT_EOS } GstEventType;
#define GST_EVENT_TYPE(e) ((GstEventType)(e)->type)
typedef enum { GST_QUERY_UNKNOWN, GST_QUERY_LATENCY } GstQueryType;
GstQueryType gst_stub_query_type(GstQuery *q);
#define GST_QUERY_TYPE(q) gst_stub_query_type(q)
void gst_query_parse_latency(GstQuery *q, gboolean *live, GstClockTime *min, GstClockTime *max);
void gst_query_set_latency(GstQuery *q, gboolean live, GstClockTime min, GstClockTime max);
guint64 gst_util_uint64_scale_int(guint64 val, gint num, gint denom);
void gst_buffer_unref(GstBuffer *b);
typedef struct _GstElement { GstObject object; } GstElement;
typedef struct _GstElementClass { GObjectClass parent_class; } GstElementClass;
typedef enum { GST_PAD_UNKNOWN, GST_PAD_SRC, GST_PAD_SINK } GstPadDirection;
typedef enum { GST_PAD_ALWAYS } GstPadPresence;
typedef enum { GST_FLOW_OK = 0, GST_FLOW_NOT_NEGOTIATED = -4, GST_FLOW_ERROR = -5 } GstFlowReturn;
typedef enum { GST_CAPS_INTERSECT_ZIG_ZAG, GST_CAPS_INTERSECT_FIRST } GstCapsIntersectMode;
GstFlowReturn gst_pad_push(GstPad *pad, GstBuffer *buffer);
typedef enum { GST_MAP_READ = 1, GST_MAP_WRITE = 2 } GstMapFlags;
typedef struct { const gchar *string; } GstStaticCaps;
typedef struct { const gchar *name_template; GstPadDirection direction; GstPadPresence presence; GstStaticCaps static_caps; } GstStaticPadTemplate;
#define GST_STATIC_CAPS(s) { (s) }
#define GST_STATIC_PAD_TEMPLATE(n, d, p, c) { (n), (d), (p), c }
#define GST_RANK_NONE 0
#define GST_VERSION_MAJOR 1
#define GST_VERSION_MINOR 22
#define GST_TYPE_INT_RANGE ((GType)1)
#define GST_TYPE_FRACTION_RANGE ((GType)2)
#define GST_ELEMENT_CLASS(k) ((GstElementClass *)(k))
#define GST_DEBUG_CATEGORY_STATIC(c) static int c
#define GST_DEBUG_CATEGORY_INIT(c, n, col, d) ((c) = 0)
#define GST_DEBUG_FUNCPTR(f) (f)
#define GST_ELEMENT_ERROR(el, dom, code, text, dbg) do { (void)(el); gst_stub_msg text; gst_stub_msg dbg; } while (0)
void gst_stub_msg(const gchar *fmt, ...);
#define GST_OBJECT_LOCK(o) ((void)(o))
#define GST_OBJECT_UNLOCK(o) ((void)(o))
#define GST_OBJECT_FLAG_SET(o, f) (((GstObject *)(o))->flags |= (f))
gpointer gst_object_ref_sink(gpointer o);  void gst_object_unref(gpointer o);
#define gst_clear_object(pp) do { if (*(pp)) { gst_object_unref(*(pp)); *(pp) = NULL; } } while (0)
GstCaps *gst_caps_new_empty(void);  guint gst_caps_get_size(const GstCaps *);  void gst_caps_unref(GstCaps *);
GstStructure *gst_caps_get_structure(const GstCaps *, guint);  void gst_caps_append_structure(GstCaps *, GstStructure *);
GstCaps *gst_caps_intersect_full(GstCaps *, GstCaps *, GstCapsIntersectMode);
GstCaps *gst_caps_truncate(GstCaps *);  GstCaps *gst_caps_make_writable(GstCaps *);  GstCaps *gst_caps_fixate(GstCaps *);
GstStructure *gst_structure_copy(const GstStructure *);
void gst_structure_remove_fields(GstStructure *, const gchar *, ...);
void gst_structure_set(GstStructure *, const gchar *, ...);
gboolean gst_structure_has_field(const GstStructure *, const gchar *);
gboolean gst_structure_get_int(const GstStructure *, const gchar *, gint *);
gboolean gst_structure_fixate_field_nearest_int(GstStructure *, const gchar *, gint);
gboolean gst_structure_fixate_field_string(GstStructure *, const gchar *, const gchar *);
gboolean gst_structure_get_fraction(const GstStructure *, const gchar *, gint *, gint *);
gboolean gst_structure_fixate_field_nearest_fraction(GstStructure *, const gchar *, gint, gint);
GstCaps *gst_pad_get_pad_template_caps(GstPad *);
void gst_element_class_add_static_pad_template(GstElementClass *, GstStaticPadTemplate *);
void gst_element_class_set_static_metadata(GstElementClass *, const gchar *, const gchar *, const gchar *, const gchar *);
gboolean gst_element_register(GstPlugin *, const gchar *, guint, GType);
#define GST_PLUGIN_DEFINE(maj, min, name, desc, init, ver, lic, pkg, origin) \
    gboolean gst_plugin_##name##_register(GstPlugin *p) { return init(p); }

/* ---- memory, allocators, pools, allocation query ----------------------------------------------------- */
typedef struct _GstAllocator GstAllocator;
typedef struct _GstMemory { GstAllocator *allocator; gsize maxsize, align, offset, size; } GstMemory;
typedef struct { guint flags; gsize align, prefix, padding; } GstAllocationParams;
struct _GstAllocator {
    GstObject object;
    const gchar *mem_type;
    gpointer (*mem_map)(GstMemory *mem, gsize maxsize, GstMapFlags flags);
    void (*mem_unmap)(GstMemory *mem);
};
typedef struct _GstAllocatorClass {
    GObjectClass object_class;
    GstMemory *(*alloc)(GstAllocator *allocator, gsize size, GstAllocationParams *params);
    void (*free)(GstAllocator *allocator, GstMemory *memory);
} GstAllocatorClass;
#define GST_TYPE_ALLOCATOR ((GType)3)
#define GST_ALLOCATOR_CAST(o) ((GstAllocator *)(o))
#define GST_ALLOCATOR_CLASS(k) ((GstAllocatorClass *)(k))
#define GST_ALLOCATOR_FLAG_CUSTOM_ALLOC (1u << 4)
#define GST_MEMORY_CAST(m) ((GstMemory *)(m))
void gst_memory_init(GstMemory *mem, guint flags, GstAllocator *allocator, GstMemory *parent, gsize maxsize, gsize align,
                     gsize offset, gsize size);
void gst_allocation_params_init(GstAllocationParams *params);
GstStructure *gst_buffer_pool_get_config(GstBufferPool *pool);
gboolean gst_buffer_pool_set_config(GstBufferPool *pool, GstStructure *config);
void gst_buffer_pool_config_set_params(GstStructure *config, GstCaps *caps, guint size, guint min, guint max);
void gst_buffer_pool_config_set_allocator(GstStructure *config, GstAllocator *allocator, const GstAllocationParams *params);
void gst_buffer_pool_config_add_option(GstStructure *config, const gchar *option);
void gst_query_parse_allocation(GstQuery *query, GstCaps **caps, gboolean *need_pool);
guint gst_query_get_n_allocation_pools(GstQuery *query);
void gst_query_add_allocation_pool(GstQuery *query, GstBufferPool *pool, guint size, guint min, guint max);
void gst_query_add_allocation_param(GstQuery *query, GstAllocator *allocator, const GstAllocationParams *params);
void gst_query_add_allocation_meta(GstQuery *query, GType api, const GstStructure *params);
gboolean gst_query_find_allocation_meta(GstQuery *query, GType api, guint *index);

/* ---- base transform ------------------------------------------------------------------------------------ */
typedef struct _GstBaseTransform { GstElement element; GstPad *sinkpad, *srcpad; GstBuffer *queued_buf; } GstBaseTransform;
typedef struct _GstBaseTransformClass {
    GstElementClass parent_class;
    gboolean passthrough_on_same_caps;
    GstCaps *(*transform_caps)(GstBaseTransform *, GstPadDirection, GstCaps *, GstCaps *);
    GstCaps *(*fixate_caps)(GstBaseTransform *, GstPadDirection, GstCaps *, GstCaps *);
    gboolean (*decide_allocation)(GstBaseTransform *, GstQuery *);
    gboolean (*propose_allocation)(GstBaseTransform *, GstQuery *, GstQuery *);
    GstFlowReturn (*prepare_output_buffer)(GstBaseTransform *, GstBuffer *, GstBuffer **);
    GstFlowReturn (*generate_output)(GstBaseTransform *, GstBuffer **);
    gboolean (*sink_event)(GstBaseTransform *, GstEvent *);
    gboolean (*query)(GstBaseTransform *, GstPadDirection, GstQuery *);
    gboolean (*start)(GstBaseTransform *);
    gboolean (*stop)(GstBaseTransform *);
} GstBaseTransformClass;
#define GST_BASE_TRANSFORM_CLASS(k) ((GstBaseTransformClass *)(k))

/* ---- video --------------------------------------------------------------------------------------------- */
typedef enum { GST_VIDEO_FORMAT_UNKNOWN, GST_VIDEO_FORMAT_I420 = 2, GST_VIDEO_FORMAT_RGBx = 7, GST_VIDEO_FORMAT_BGRx,
               GST_VIDEO_FORMAT_xRGB, GST_VIDEO_FORMAT_xBGR, GST_VIDEO_FORMAT_RGBA, GST_VIDEO_FORMAT_BGRA,
               GST_VIDEO_FORMAT_ARGB, GST_VIDEO_FORMAT_ABGR, GST_VIDEO_FORMAT_Y42B = 19, GST_VIDEO_FORMAT_Y444 = 20,
               GST_VIDEO_FORMAT_NV12 = 23, GST_VIDEO_FORMAT_RGB16 = 29, GST_VIDEO_FORMAT_BGR16 = 30 } GstVideoFormat;
typedef enum { GST_VIDEO_COLOR_RANGE_UNKNOWN, GST_VIDEO_COLOR_RANGE_0_255, GST_VIDEO_COLOR_RANGE_16_235 } GstVideoColorRange;
typedef enum { GST_VIDEO_COLOR_MATRIX_UNKNOWN, GST_VIDEO_COLOR_MATRIX_RGB, GST_VIDEO_COLOR_MATRIX_FCC, GST_VIDEO_COLOR_MATRIX_BT709,
               GST_VIDEO_COLOR_MATRIX_BT601, GST_VIDEO_COLOR_MATRIX_SMPTE240M, GST_VIDEO_COLOR_MATRIX_BT2020 } GstVideoColorMatrix;
typedef enum { GST_VIDEO_CHROMA_SITE_UNKNOWN = 0, GST_VIDEO_CHROMA_SITE_NONE = 1, GST_VIDEO_CHROMA_SITE_H_COSITED = 2,
               GST_VIDEO_CHROMA_SITE_V_COSITED = 4, GST_VIDEO_CHROMA_SITE_ALT_LINE = 8 } GstVideoChromaSite;
typedef struct { GstVideoColorRange range; GstVideoColorMatrix matrix; int transfer, primaries; } GstVideoColorimetry;
typedef struct _GstVideoInfo {
    GstVideoFormat format; gint width, height; gsize size; GstVideoChromaSite chroma_site; GstVideoColorimetry colorimetry;
    gint stride[4];
    gint fps_n, fps_d;
    gint par_n, par_d;
} GstVideoInfo;
typedef struct _GstVideoFrame { GstVideoInfo info; GstBuffer *buffer; gpointer data[4]; gint n_planes; } GstVideoFrame;
#define GST_VIDEO_INFO_FORMAT(i) ((i)->format)
#define GST_VIDEO_INFO_WIDTH(i) ((i)->width)
#define GST_VIDEO_INFO_HEIGHT(i) ((i)->height)
#define GST_VIDEO_INFO_SIZE(i) ((i)->size)
#define GST_VIDEO_INFO_PAR_N(i) ((i)->par_n)
#define GST_VIDEO_INFO_PAR_D(i) ((i)->par_d)
#define GST_VIDEO_INFO_FPS_N(i) ((i)->fps_n)
#define GST_VIDEO_INFO_FPS_D(i) ((i)->fps_d)
#define GST_VIDEO_FRAME_WIDTH(f) ((f)->info.width)
#define GST_VIDEO_FRAME_HEIGHT(f) ((f)->info.height)
#define GST_VIDEO_FRAME_N_PLANES(f) ((f)->n_planes)
#define GST_VIDEO_FRAME_PLANE_DATA(f, p) ((f)->data[p])
#define GST_VIDEO_FRAME_PLANE_STRIDE(f, p) ((f)->info.stride[p])
#define GST_VIDEO_META_API_TYPE ((GType)4)
#define GST_BUFFER_POOL_OPTION_VIDEO_META "GstBufferPoolOptionVideoMeta"
gboolean gst_video_info_from_caps(GstVideoInfo *info, const GstCaps *caps);
gboolean gst_video_frame_map(GstVideoFrame *frame, const GstVideoInfo *info, GstBuffer *buffer, GstMapFlags flags);
void gst_video_frame_unmap(GstVideoFrame *frame);
GstBufferPool *gst_video_buffer_pool_new(void);
gpointer gst_buffer_add_video_region_of_interest_meta(GstBuffer *buffer, const gchar *roi_type, guint x, guint y, guint w, guint h);
typedef struct _GstVideoFilter { GstBaseTransform element; gboolean negotiated; GstVideoInfo in_info, out_info; } GstVideoFilter;
typedef struct _GstVideoFilterClass {
    GstBaseTransformClass parent_class;
    gboolean (*set_info)(GstVideoFilter *, GstCaps *, GstVideoInfo *, GstCaps *, GstVideoInfo *);
    GstFlowReturn (*transform_frame)(GstVideoFilter *, GstVideoFrame *, GstVideoFrame *);
} GstVideoFilterClass;
#define GST_TYPE_VIDEO_FILTER ((GType)5)
#define GST_VIDEO_FILTER_CLASS(k) ((GstVideoFilterClass *)(k))
#endif
