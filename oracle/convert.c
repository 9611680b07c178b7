/*
 * convert.c — CPU restatement of GStreamer videoconvert for BGRx -> I420/NV12.
 * TEST INFRASTRUCTURE ONLY; PARITY UNPINNED (see oracle.h).
 *
 * Follows SURVEY.md Appendix A, which restates (from recollection; upstream source is not on
 * this machine) gst-libs/gst/video/video-converter.c, video-chroma.c and video-format.c of the
 * gst-plugins-base bundled by the reference at Dockerfile:467-471. The line chain is
 *   unpack_BGRA -> video_converter_matrix8_table -> video_chroma_down_{h2,h2_cs}+v2 -> pack_I420/NV12
 * run per output line pair, exactly the structure described in SURVEY A.1 / §8(a2-a5).
 */
#include "oracle.h"
#include "../include/b200vc_spec.h"   /* the unverified details of SURVEY A/B, one constant each */

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- A.2: matrix ------------------------------------------------------------------------ */

void orc_build_matrix(int matrix, int range, orc_matrix *out)
{
    /* gst_video_color_matrix_get_Kr_Kb */
    const double Kr = (matrix == ORC_MATRIX_BT2020) ? 0.2627 : (matrix == ORC_MATRIX_BT709) ? 0.2126 : 0.299;
    const double Kb = (matrix == ORC_MATRIX_BT2020) ? 0.0593 : (matrix == ORC_MATRIX_BT709) ? 0.0722 : 0.114;
    const double Kg = 1.0 - Kr - Kb;
    double m[3][4];

    /* RGB is full range: scale 1/255, offset 0 (compute_matrix_to_RGB side is identity here). */
    m[0][0] = Kr;                          m[0][1] = Kg;                          m[0][2] = Kb;
    m[1][0] = -Kr / (2.0 * (1.0 - Kb));    m[1][1] = -Kg / (2.0 * (1.0 - Kb));    m[1][2] = 0.5;
    m[2][0] = 0.5;                         m[2][1] = -Kg / (2.0 * (1.0 - Kr));    m[2][2] = -Kb / (2.0 * (1.0 - Kr));

    /* gst_video_color_range_offsets for 8-bit YUV */
    const double sy = (range == ORC_RANGE_FULL) ? 255.0 : 219.0;
    const double sc = (range == ORC_RANGE_FULL) ? 255.0 : 224.0;
    const double oy = (range == ORC_RANGE_FULL) ? 0.0 : 16.0;
    const double scale[3] = { sy, sc, sc };
    const double off[3] = { oy, 128.0, 128.0 };

    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++)
            out->m[r][c] = (int32_t)rint(256.0 * m[r][c] * scale[r] / 255.0);   /* SCALE = 8 */
        /* A.2 warning item: a rounding term, if upstream has one, is +128 on every offset */
        out->m[r][3] = (int32_t)rint(256.0 * off[r]) + (B200VC_SPEC_MATRIX_ROUND ? 128 : 0);
    }
}

const char *orc_spec_string(void) { return B200VC_SPEC_STRING; }

/* ---- A.5: default layout ------------------------------------------------------------------ */

#define RU2(x) (((x) + 1) & ~1)
#define RU4(x) (((x) + 3) & ~3)

size_t orc_default_layout(int format, int w, int h, size_t offsets[3], size_t pitches[3])
{
    size_t ypitch = (size_t)RU4(w);
    offsets[0] = 0;
    pitches[0] = ypitch;
    if (format == ORC_FMT_Y42B) {          /* fill_planes: stride[1] = RU8(w)/2, heights as they are */
        size_t cp = (size_t)(((w + 7) & ~7) / 2);
        offsets[1] = ypitch * (size_t)h;  pitches[1] = cp;
        offsets[2] = offsets[1] + cp * (size_t)h;  pitches[2] = cp;
        return offsets[2] + cp * (size_t)h;
    }
    if (format == ORC_FMT_Y444) {
        offsets[1] = ypitch * (size_t)h;  pitches[1] = ypitch;
        offsets[2] = 2 * ypitch * (size_t)h;  pitches[2] = ypitch;
        return 3 * ypitch * (size_t)h;
    }
    if (format == ORC_FMT_NV12) {
        offsets[1] = ypitch * (size_t)RU2(h);
        pitches[1] = ypitch;
        offsets[2] = 0;
        pitches[2] = 0;
        return offsets[1] + pitches[1] * (size_t)(RU2(h) / 2);
    }
    size_t cpitch = (size_t)RU4(RU2(w) / 2);
    offsets[1] = ypitch * (size_t)RU2(h);
    pitches[1] = cpitch;
    offsets[2] = offsets[1] + cpitch * (size_t)(RU2(h) / 2);
    pitches[2] = cpitch;
    return offsets[2] + cpitch * (size_t)(RU2(h) / 2);
}

/* ---- line stages ---------------------------------------------------------------------------- */

/* unpack_BGRA / unpack_RGBA / unpack_ARGB / unpack_ABGR: 4-byte pixel -> A,R,G,B temp line
 * (A forced to 0xff like the x-format unpacks; it is dropped again by the pack stage). */
static void unpack_line(const uint8_t *s, uint8_t *d, int w, int in_format)
{
    if (in_format == ORC_IN_RGB16 || in_format == ORC_IN_BGR16) {
        /* unpack_RGB16 / unpack_BGR16: little-endian 5-6-5, widened to 8 bits with the high bits repeated
         * into the low ones (video_orc_unpack_RGB16; _trunc, a plain shift, under TRUNCATE_RANGE — the
         * B200VC_SPEC_UNPACK16_REPLICATE constant) */
        for (int i = 0; i < w; i++) {
            const unsigned v = s[2 * i] | ((unsigned)s[2 * i + 1] << 8);
            unsigned hi = v >> 11, g = (v >> 5) & 63, lo = v & 31;
            unsigned hi8 = hi << 3, g8 = g << 2, lo8 = lo << 3;
            if (B200VC_SPEC_UNPACK16_REPLICATE) { hi8 |= hi >> 2;  g8 |= g >> 4;  lo8 |= lo >> 2; }
            d[4 * i + 0] = 0xff;
            d[4 * i + 1] = (uint8_t)(in_format == ORC_IN_RGB16 ? hi8 : lo8);
            d[4 * i + 2] = (uint8_t)g8;
            d[4 * i + 3] = (uint8_t)(in_format == ORC_IN_RGB16 ? lo8 : hi8);
        }
        return;
    }
    static const int pos[4][3] = { { 2, 1, 0 }, { 0, 1, 2 }, { 1, 2, 3 }, { 3, 2, 1 } };   /* byte of R, G, B */
    const int pr = pos[in_format][0], pg = pos[in_format][1], pb = pos[in_format][2];
    for (int i = 0; i < w; i++) {
        d[4 * i + 0] = 0xff;
        d[4 * i + 1] = s[4 * i + pr];
        d[4 * i + 2] = s[4 * i + pg];
        d[4 * i + 3] = s[4 * i + pb];
    }
}

/* videoconvert_convert_init_tables: three 256-entry int64 tables, one 16-bit lane per row. */
typedef struct {
    int64_t t_r[256], t_g[256], t_b[256];
    int64_t t_c;
    orc_matrix mx;
} matrix_tables;

static void init_tables(const orc_matrix *mx, matrix_tables *t)
{
    t->mx = *mx;
    for (int i = 0; i < 256; i++) {
        int64_t r = 0, g = 0, b = 0;
        for (int j = 0; j < 3; j++) {
            r = r * 65536 + (int64_t)mx->m[j][0] * i;
            g = g * 65536 + (int64_t)mx->m[j][1] * i;
            b = b * 65536 + (int64_t)mx->m[j][2] * i;
        }
        t->t_r[i] = r;
        t->t_g[i] = g;
        t->t_b[i] = b;
    }
    t->t_c = (int64_t)mx->m[0][3] * 65536 * 65536 + (int64_t)mx->m[1][3] * 65536 + (int64_t)mx->m[2][3];
}

/* video_converter_matrix8_table: ARGB -> AYUV in place, (sum) >> 8 with no rounding term. */
static void matrix8_table(const matrix_tables *t, uint8_t *p, int w)
{
#if B200VC_SPEC_MATRIX_ROUND
    /* with a rounding term a lane can reach 65536 (full-range U/V of a saturated primary) and would
     * carry into its neighbour: evaluate the rows apart and clamp, as a clipping matrix stage would */
    for (int i = 0; i < 4 * w; i += 4) {
        const int r = p[i + 1], g = p[i + 2], b = p[i + 3];
        for (int j = 0; j < 3; j++) {
            int v = t->mx.m[j][0] * r + t->mx.m[j][1] * g + t->mx.m[j][2] * b + t->mx.m[j][3];
            v = v < 0 ? 0 : (v > 65535 ? 65535 : v);
            p[i + 1 + j] = (uint8_t)(v >> 8);
        }
    }
    return;
#endif
    for (int i = 0; i < 4 * w; i += 4) {
        int64_t x = t->t_r[p[i + 1]] + t->t_g[p[i + 2]] + t->t_b[p[i + 3]] + t->t_c;
        p[i + 1] = (uint8_t)(x >> (32 + 8));
        p[i + 2] = (uint8_t)(x >> (16 + 8));
        p[i + 3] = (uint8_t)(x >> (0 + 8));
    }
}

#define PU(i) p[4 * (i) + 2]
#define PV(i) p[4 * (i) + 3]

/* video_chroma_down_h2_u8: site NONE/JPEG */
static void chroma_down_h2(uint8_t *p, int w)
{
    for (int i = 0; i + 1 < w; i += 2) {
        PU(i) = (uint8_t)((PU(i) + PU(i + 1) + 1) >> 1);
        PV(i) = (uint8_t)((PV(i) + PV(i + 1) + 1) >> 1);
    }
}

/* video_chroma_down_h2_cs_u8: site H_COSITED. In place; odd pixels are never overwritten. */
static void chroma_down_h2_cs(uint8_t *p, int w)
{
    int i;
    if (w < 2)
        return;
    PU(0) = (uint8_t)((3 * PU(0) + PU(1) + 2) >> 2);
    PV(0) = (uint8_t)((3 * PV(0) + PV(1) + 2) >> 2);
    /* A.4 warning item: where the 1-2-1 loop stops. As recalled it stops before the last even column,
     * which then ignores u[W-1]; the other candidate runs it while a right neighbour exists. */
    const int stop = B200VC_SPEC_COSITED_LAST_1_3 ? w - 2 : w - 1;
    for (i = 2; i < stop; i += 2) {
        PU(i) = (uint8_t)((PU(i - 1) + 2 * PU(i) + PU(i + 1) + 2) >> 2);
        PV(i) = (uint8_t)((PV(i - 1) + 2 * PV(i) + PV(i + 1) + 2) >> 2);
    }
    if (i < w) {
        PU(i) = (uint8_t)((PU(i - 1) + 3 * PU(i) + 2) >> 2);
        PV(i) = (uint8_t)((PV(i - 1) + 3 * PV(i) + 2) >> 2);
    }
}

/* video_chroma_down_v2_u8: horizontal on each line first, then rounded vertical mean into l0. */
/* 4:2:2 (Y42B): horizontal only, every line */
static void chroma_down_h(uint8_t *l, int w, int site)
{
    if (site & ORC_SITE_H_COSITED)
        chroma_down_h2_cs(l, w);
    else
        chroma_down_h2(l, w);
}

static void chroma_down_v2(uint8_t *l0, uint8_t *l1, int w, int site)
{
    if (site & ORC_SITE_H_COSITED) {
        chroma_down_h2_cs(l0, w);
        chroma_down_h2_cs(l1, w);
    } else {
        chroma_down_h2(l0, w);
        chroma_down_h2(l1, w);
    }
    for (int i = 0; i < w; i++) {
        l0[4 * i + 2] = (uint8_t)((l0[4 * i + 2] + l1[4 * i + 2] + 1) >> 1);
        l0[4 * i + 3] = (uint8_t)((l0[4 * i + 3] + l1[4 * i + 3] + 1) >> 1);
    }
}

/* pack_I420 / pack_NV12: Y of every pixel; U,V of even pixels of the even line. */
static void pack_y(const uint8_t *p, uint8_t *y, int w)
{
    for (int i = 0; i < w; i++)
        y[i] = p[4 * i + 1];
}

/* odd widths: the last pixel is a chroma site of its own (pack_I420's `if (width & 1)` tail) */
static void pack_uv_planar(const uint8_t *p, uint8_t *u, uint8_t *v, int w)
{
    for (int i = 0; i < (w + 1) / 2; i++) {
        u[i] = p[8 * i + 2];
        v[i] = p[8 * i + 3];
    }
}

/* pack_Y444: U and V of every pixel */
static void pack_uv_full(const uint8_t *p, uint8_t *u, uint8_t *v, int w)
{
    for (int i = 0; i < w; i++) {
        u[i] = p[4 * i + 2];
        v[i] = p[4 * i + 3];
    }
}

static void pack_uv_interleaved(const uint8_t *p, uint8_t *uv, int w)
{
    for (int i = 0; i < (w + 1) / 2; i++) {
        uv[2 * i + 0] = p[8 * i + 2];
        uv[2 * i + 1] = p[8 * i + 3];
    }
}

/* ---- frame entry points ------------------------------------------------------------------- */

static int convert_rows_ex(const uint8_t *bgrx, size_t in_pitch, int in_format, int w, int h,
                           int format, int matrix, int range, int site,
                           uint8_t *const planes[3], const size_t pitches[3],
                           int row0, int row1)
{
    if (in_format < ORC_IN_BGRX || in_format > ORC_IN_BGR16 || matrix < ORC_MATRIX_BT601 || matrix > ORC_MATRIX_BT2020)
        return -1;
    if (!bgrx || !planes || !pitches || w < 2 || h < 2)
        return -1;
    if (row0 < 0 || row1 > h || row0 > row1 || (row0 & 1) || ((row1 & 1) && row1 != h))
        return -1;
    if (format < ORC_FMT_I420 || format > ORC_FMT_Y444)
        return -1;
    if (site < 0 || site > 7)
        return -1;

    orc_matrix mx;
    matrix_tables *tab = (matrix_tables *)malloc(sizeof(*tab));
    uint8_t *l0 = (uint8_t *)malloc((size_t)w * 4);
    uint8_t *l1 = (uint8_t *)malloc((size_t)w * 4);
    if (!tab || !l0 || !l1) {
        free(tab); free(l0); free(l1);
        return -1;
    }
    orc_build_matrix(matrix, range, &mx);
    init_tables(&mx, tab);

    for (int y = row0; y < row1; y += 2) {
        /* do_unpack_lines clamps the line index: below an odd frame's last line comes that line again */
        const int y1 = (y + 1 < h) ? y + 1 : y;
        unpack_line(bgrx + (size_t)y * in_pitch, l0, w, in_format);
        unpack_line(bgrx + (size_t)y1 * in_pitch, l1, w, in_format);
        matrix8_table(tab, l0, w);
        matrix8_table(tab, l1, w);
        pack_y(l0, planes[0] + (size_t)y * pitches[0], w);
        if (y1 != y)
            pack_y(l1, planes[0] + (size_t)y1 * pitches[0], w);
        if (format == ORC_FMT_Y444 || format == ORC_FMT_Y42B) {
            /* no vertical subsampling: each line on its own (4:2:2: horizontal filter first) */
            for (int k = 0; k < (y1 != y ? 2 : 1); k++) {
                uint8_t *l = k ? l1 : l0;
                const size_t yy = (size_t)(k ? y1 : y);
                if (format == ORC_FMT_Y444) {
                    pack_uv_full(l, planes[1] + yy * pitches[1], planes[2] + yy * pitches[2], w);
                } else {
                    chroma_down_h(l, w, site);
                    pack_uv_planar(l, planes[1] + yy * pitches[1], planes[2] + yy * pitches[2], w);
                }
            }
            continue;
        }
        chroma_down_v2(l0, l1, w, site);
        if (format == ORC_FMT_NV12)
            pack_uv_interleaved(l0, planes[1] + (size_t)(y / 2) * pitches[1], w);
        else
            pack_uv_planar(l0, planes[1] + (size_t)(y / 2) * pitches[1],
                           planes[2] + (size_t)(y / 2) * pitches[2], w);
    }
    free(tab); free(l0); free(l1);
    return 0;
}

int orc_convert_rows(const uint8_t *bgrx, size_t in_pitch, int w, int h,
                     int format, int matrix, int range, int site,
                     uint8_t *const planes[3], const size_t pitches[3],
                     int row0, int row1)
{
    return convert_rows_ex(bgrx, in_pitch, ORC_IN_BGRX, w, h, format, matrix, range, site, planes, pitches, row0, row1);
}

int orc_convert(const uint8_t *bgrx, size_t in_pitch, int w, int h,
                int format, int matrix, int range, int site,
                uint8_t *const planes[3], const size_t pitches[3])
{
    return orc_convert_rows(bgrx, in_pitch, w, h, format, matrix, range, site, planes, pitches, 0, h);
}

typedef struct {
    const uint8_t *bgrx; size_t in_pitch; int in_format, w, h, format, matrix, range, site;
    uint8_t *const *planes; const size_t *pitches; int row0, row1, rc;
} band_job;

static void *band_main(void *arg)
{
    band_job *j = (band_job *)arg;
    j->rc = convert_rows_ex(j->bgrx, j->in_pitch, j->in_format, j->w, j->h, j->format, j->matrix, j->range,
                            j->site, j->planes, j->pitches, j->row0, j->row1);
    return NULL;
}

/* n-threads > 1: horizontal bands aligned to 2 lines. */
int orc_convert_mt_ex(const uint8_t *bgrx, size_t in_pitch, int in_format, int w, int h,
                      int format, int matrix, int range, int site,
                      uint8_t *const planes[3], const size_t pitches[3], int n_threads);

int orc_convert_mt(const uint8_t *bgrx, size_t in_pitch, int w, int h,
                   int format, int matrix, int range, int site,
                   uint8_t *const planes[3], const size_t pitches[3], int n_threads)
{
    return orc_convert_mt_ex(bgrx, in_pitch, ORC_IN_BGRX, w, h, format, matrix, range, site, planes, pitches, n_threads);
}

int orc_convert_mt_ex(const uint8_t *bgrx, size_t in_pitch, int in_format, int w, int h,
                      int format, int matrix, int range, int site,
                      uint8_t *const planes[3], const size_t pitches[3], int n_threads)
{
    if (n_threads <= 1)
        return convert_rows_ex(bgrx, in_pitch, in_format, w, h, format, matrix, range, site, planes, pitches, 0, h);
    if (n_threads > 64)
        n_threads = 64;
    pthread_t th[64];
    int started[64];
    band_job jobs[64];
    int pairs = (h + 1) / 2, rc = 0;
    for (int t = 0; t < n_threads; t++) {
        band_job *j = &jobs[t];
        j->bgrx = bgrx; j->in_pitch = in_pitch; j->in_format = in_format; j->w = w; j->h = h; j->format = format;
        j->matrix = matrix; j->range = range; j->site = site; j->planes = planes; j->pitches = pitches;
        j->row0 = 2 * (int)((long)pairs * t / n_threads);
        j->row1 = 2 * (int)((long)pairs * (t + 1) / n_threads);
        if (j->row1 > h) j->row1 = h;
        j->rc = 0;
        started[t] = (pthread_create(&th[t], NULL, band_main, j) == 0);
        if (!started[t])
            band_main(j);
    }
    for (int t = 0; t < n_threads; t++) {
        if (started[t])
            pthread_join(th[t], NULL);
        if (jobs[t].rc)
            rc = jobs[t].rc;
    }
    return rc;
}
