/*
 * scale.c — CPU restatement of GStreamer videoscale (method=bilinear, 2-tap "lq" path) and of
 * the two-element chain `videoconvert ! videoscale`.
 * TEST INFRASTRUCTURE ONLY; PARITY UNPINNED (see oracle.h).
 *
 * Follows SURVEY.md Appendix B, which restates (from recollection) video-scaler.c /
 * video-resampler.c / video-orc.orc of the gst-plugins-base bundled by the reference at
 * Dockerfile:467-471. Spec constants: horizontal pass first, then vertical (B.2); conversion runs
 * at input resolution and each output plane is scaled independently (B.1).
 * All tap arithmetic is integer so that every restatement of it agrees bit for bit.
 */
#include "oracle.h"
#include "../include/b200vc_spec.h"   /* the unverified details of SURVEY A/B, one constant each */

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* B.2 horizontal: edge-aligned 16.16 stepping with an 8-bit fraction. */
void orc_scale_h_taps(int in_w, int out_w, int32_t *idx, int32_t *frac)
{
    int64_t inc = 0;
    if (out_w > 1) {
        inc = (((int64_t)(in_w - 1)) << 16) / (out_w - 1) - 1;
        if (inc < 0)
            inc = 0;
    }
    for (int x = 0; x < out_w; x++) {
        int64_t p = (int64_t)x * inc;
        idx[x] = (int32_t)(p >> 16);
        frac[x] = (int32_t)((p >> 8) & 0xff);
    }
}

/* B.2 vertical: centre-aligned tap: y' = (y+0.5)*in_h/out_h - 0.5, p1 = round(frac * 2^bits) with
 * bits = 6 as recalled (B200VC_SPEC_VTAP_BITS; B200VC_SPEC_VTAP_ROUND = 0 floors instead). */
void orc_scale_v_taps(int in_h, int out_h, int32_t *idx, int32_t *p1)
{
    const int64_t den = 2 * (int64_t)out_h;
    for (int y = 0; y < out_h; y++) {
        int64_t num = (2 * (int64_t)y + 1) * in_h - out_h;
        int64_t i = num / den;
        int64_t rem = num - i * den;
        if (rem < 0) {            /* C division truncates toward zero; want floor */
            i -= 1;
            rem += den;
        }
        idx[y] = (int32_t)i;
        p1[y] = (int32_t)((2 * (int64_t)B200VC_SPEC_VTAP_ONE * rem + (B200VC_SPEC_VTAP_ROUND ? den : 0)) / (2 * den));
    }
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* video_scale_h_2tap_1u8 (lq): d = (s[i]*(256-f) + s[i+1]*f) >> 8 */
static void scale_h_line(const uint8_t *s, int in_w, uint8_t *d, int out_w, int comps,
                         const int32_t *idx, const int32_t *frac)
{
    for (int x = 0; x < out_w; x++) {
        int i0 = clampi(idx[x], 0, in_w - 1);
        int i1 = clampi(idx[x] + 1, 0, in_w - 1);
        int f = frac[x];
        for (int c = 0; c < comps; c++)
            d[x * comps + c] = (uint8_t)((s[i0 * comps + c] * (256 - f) + s[i1 * comps + c] * f) >> 8);
    }
}

int orc_scale_plane(const uint8_t *src, size_t src_pitch, int in_w, int in_h,
                    uint8_t *dst, size_t dst_pitch, int out_w, int out_h,
                    int comps, int scale_mode)
{
    if (!src || !dst || in_w < 1 || in_h < 1 || out_w < 1 || out_h < 1 || comps < 1 || comps > 2)
        return -1;

    if (scale_mode == ORC_SCALE_BOX_2X2) {
        if (in_w != 2 * out_w || in_h != 2 * out_h)
            return -1;
        for (int y = 0; y < out_h; y++) {
            const uint8_t *a = src + (size_t)(2 * y) * src_pitch;
            const uint8_t *b = a + src_pitch;
            uint8_t *d = dst + (size_t)y * dst_pitch;
            for (int x = 0; x < out_w; x++)
                for (int c = 0; c < comps; c++)
                    d[x * comps + c] = (uint8_t)((a[(2 * x) * comps + c] + a[(2 * x + 1) * comps + c] +
                                                  b[(2 * x) * comps + c] + b[(2 * x + 1) * comps + c] + 2) >> 2);
        }
        return 0;
    }
    if (scale_mode == ORC_SCALE_NEAREST) {
        /* method=nearest (video_scale_h_near_u8 / v_near): one source sample per output sample, the one under the
         * output pixel's centre — or, the other candidate, at the bilinear path's edge-aligned 16.16 position
         * (B200VC_SPEC_NEAREST_CENTRE) */
        for (int y = 0; y < out_h; y++) {
            int64_t inc_v = out_h > 1 ? (((int64_t)(in_h - 1)) << 16) / (out_h - 1) - 1 : 0;
            int sy = B200VC_SPEC_NEAREST_CENTRE ? (int)(((2 * (int64_t)y + 1) * in_h) / (2 * (int64_t)out_h))
                                                : (int)(((int64_t)y * (inc_v < 0 ? 0 : inc_v)) >> 16);
            const uint8_t *s = src + (size_t)clampi(sy, 0, in_h - 1) * src_pitch;
            uint8_t *d = dst + (size_t)y * dst_pitch;
            int64_t inc_h = out_w > 1 ? (((int64_t)(in_w - 1)) << 16) / (out_w - 1) - 1 : 0;
            for (int x = 0; x < out_w; x++) {
                int sx = B200VC_SPEC_NEAREST_CENTRE ? (int)(((2 * (int64_t)x + 1) * in_w) / (2 * (int64_t)out_w))
                                                    : (int)(((int64_t)x * (inc_h < 0 ? 0 : inc_h)) >> 16);
                sx = clampi(sx, 0, in_w - 1);
                for (int c = 0; c < comps; c++)
                    d[x * comps + c] = s[sx * comps + c];
            }
        }
        return 0;
    }
    if (scale_mode != ORC_SCALE_GST_LQ)
        return -1;

    int32_t *hi = (int32_t *)malloc(sizeof(int32_t) * (size_t)out_w * 2);
    int32_t *vi = (int32_t *)malloc(sizeof(int32_t) * (size_t)out_h * 2);
    uint8_t *la = (uint8_t *)malloc((size_t)out_w * comps);
    uint8_t *lb = (uint8_t *)malloc((size_t)out_w * comps);
    if (!hi || !vi || !la || !lb) {
        free(hi); free(vi); free(la); free(lb);
        return -1;
    }
    int32_t *hf = hi + out_w, *vp = vi + out_h;
    orc_scale_h_taps(in_w, out_w, hi, hf);
    orc_scale_v_taps(in_h, out_h, vi, vp);

#if !B200VC_SPEC_SCALE_H_FIRST
    /* the other order of B.2: vertical 2-tap over the source lines first, horizontal after it */
    uint8_t *lv = (uint8_t *)malloc((size_t)in_w * comps);
    if (!lv) { free(hi); free(vi); free(la); free(lb); return -1; }
    for (int y = 0; y < out_h; y++) {
        const uint8_t *sa = src + (size_t)clampi(vi[y], 0, in_h - 1) * src_pitch;
        const uint8_t *sb = src + (size_t)clampi(vi[y] + 1, 0, in_h - 1) * src_pitch;
        for (int x = 0; x < in_w * comps; x++)
            lv[x] = (uint8_t)clampi(sa[x] + (((sb[x] - sa[x]) * vp[y] + B200VC_SPEC_VTAP_HALF) >> B200VC_SPEC_VTAP_BITS), 0, 255);
        scale_h_line(lv, in_w, dst + (size_t)y * dst_pitch, out_w, comps, hi, hf);
    }
    free(lv);
#else
    int have_a = -1, have_b = -1;
    for (int y = 0; y < out_h; y++) {
        int ia = clampi(vi[y], 0, in_h - 1);
        int ib = clampi(vi[y] + 1, 0, in_h - 1);
        /* horizontal pass on the two source lines (reuse the line scaled for the previous row) */
        if (have_b == ia) {
            uint8_t *t = la; la = lb; lb = t;
            have_a = ia; have_b = -1;
        }
        if (have_a != ia) {
            scale_h_line(src + (size_t)ia * src_pitch, in_w, la, out_w, comps, hi, hf);
            have_a = ia;
        }
        if (have_b != ib) {
            scale_h_line(src + (size_t)ib * src_pitch, in_w, lb, out_w, comps, hi, hf);
            have_b = ib;
        }
        /* video_scale_v_2tap_u8 (lq): d = a + (((b-a)*p1 + half) >> bits), saturated. The ORC kernel
         * does this in 16-bit lanes, which a 6- or 7-bit tap cannot overflow (255*128 + 64 < 32768);
         * plain int here so that an 8-bit tap is evaluated as wider arithmetic would. */
        uint8_t *d = dst + (size_t)y * dst_pitch;
        int p1 = vp[y];
        for (int x = 0; x < out_w * comps; x++) {
            int t = ((lb[x] - la[x]) * p1 + B200VC_SPEC_VTAP_HALF) >> B200VC_SPEC_VTAP_BITS;
            d[x] = (uint8_t)clampi(la[x] + t, 0, 255);
        }
    }
#endif
    free(hi); free(vi); free(la); free(lb);
    return 0;
}

int orc_convert_mt_ex(const uint8_t *bgrx, size_t in_pitch, int in_format, int w, int h,
                      int format, int matrix, int range, int site,
                      uint8_t *const planes[3], const size_t pitches[3], int n_threads);

static int convert_scale_impl(const uint8_t *bgrx, size_t in_pitch, int in_format, int in_w, int in_h,
                              int out_w, int out_h, int format, int matrix, int range, int site,
                              int scale_mode, uint8_t *const planes[3], const size_t pitches[3],
                              int n_threads)
{
    if (in_w == out_w && in_h == out_h)
        return orc_convert_mt_ex(bgrx, in_pitch, in_format, in_w, in_h, format, matrix, range, site, planes, pitches, n_threads);
    if (out_w < 2 || out_h < 2)
        return -1;
    if (scale_mode == ORC_SCALE_BOX_2X2 && ((in_w | in_h | out_w | out_h) & 1))
        return -1;
    /* odd sizes: the chroma planes of both frames round up (video-info.c), and are scaled as such */
    const int hs = format == ORC_FMT_Y444 ? 1 : 2, vs = (format == ORC_FMT_I420 || format == ORC_FMT_NV12) ? 2 : 1;
    const int icw = (in_w + hs - 1) / hs, ich = (in_h + vs - 1) / vs, ocw = (out_w + hs - 1) / hs, och = (out_h + vs - 1) / vs;

    /* stage 1: videoconvert at input size into a default-layout intermediate frame */
    size_t off[3], pit[3];
    size_t total = orc_default_layout(format, in_w, in_h, off, pit);
    uint8_t *tmp = (uint8_t *)malloc(total);
    if (!tmp)
        return -1;
    uint8_t *tp[3] = { tmp + off[0], tmp + off[1], tmp + off[2] };
    int rc = orc_convert_mt_ex(bgrx, in_pitch, in_format, in_w, in_h, format, matrix, range, site, tp, pit, n_threads);

    /* stage 2: videoscale, each plane independently */
    if (!rc)
        rc = orc_scale_plane(tp[0], pit[0], in_w, in_h, planes[0], pitches[0], out_w, out_h, 1, scale_mode);
    if (!rc && format == ORC_FMT_NV12) {
        rc = orc_scale_plane(tp[1], pit[1], icw, ich, planes[1], pitches[1], ocw, och, 2, scale_mode);
    } else if (!rc) {
        rc = orc_scale_plane(tp[1], pit[1], icw, ich, planes[1], pitches[1], ocw, och, 1, scale_mode);
        if (!rc)
            rc = orc_scale_plane(tp[2], pit[2], icw, ich, planes[2], pitches[2], ocw, och, 1, scale_mode);
    }
    free(tmp);
    return rc;
}

int orc_convert_scale(const uint8_t *bgrx, size_t in_pitch, int in_w, int in_h,
                      int out_w, int out_h, int format, int matrix, int range, int site,
                      int scale_mode, uint8_t *const planes[3], const size_t pitches[3])
{
    return convert_scale_impl(bgrx, in_pitch, ORC_IN_BGRX, in_w, in_h, out_w, out_h, format, matrix, range, site,
                              scale_mode, planes, pitches, 1);
}

/* videoscale itself runs n-threads=1 by default; only stage 1 is banded. */
int orc_convert_scale_mt(const uint8_t *bgrx, size_t in_pitch, int in_w, int in_h,
                         int out_w, int out_h, int format, int matrix, int range, int site,
                         int scale_mode, uint8_t *const planes[3], const size_t pitches[3],
                         int n_threads)
{
    return convert_scale_impl(bgrx, in_pitch, ORC_IN_BGRX, in_w, in_h, out_w, out_h, format, matrix, range, site,
                              scale_mode, planes, pitches, n_threads);
}

int orc_convert_scale_mt_ex(const uint8_t *src, size_t in_pitch, int in_format, int in_w, int in_h,
                            int out_w, int out_h, int format, int matrix, int range, int site,
                            int scale_mode, uint8_t *const planes[3], const size_t pitches[3],
                            int n_threads)
{
    return convert_scale_impl(src, in_pitch, in_format, in_w, in_h, out_w, out_h, format, matrix, range, site,
                              scale_mode, planes, pitches, n_threads);
}
