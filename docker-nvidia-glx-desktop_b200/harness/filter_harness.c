/*
 * filter_harness.c — headless stand-in for the GStreamer base classes: drives libb200vc.so in the
 * order GstVideoFilter would (set_info once -> transform_frame per buffer -> renegotiation in
 * mid-stream -> stop), because GStreamer itself is not available on this machine.
 * Plain C, links only against libb200vc.so; the frames are generated here (splitmix64, the
 * generator SURVEY §8 d defines), the outputs are written to a file the test compares with the
 * oracle.
 *
 *   filter_harness W H OUT_W OUT_H N_FRAMES OUTFILE
 * Writes, for each frame k: first format I420 for k < N/2, then (renegotiated) NV12.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200vc.h"

static uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static void synth_rand(uint8_t *dst, size_t pitch, int w, int h, uint64_t frame)
{
    for (int y = 0; y < h; y++) {
        uint64_t st = 0x9E3779B97F4A7C15ull + frame * 0x100000001B3ull + (uint64_t)y * 0xD6E8FEB86659FD93ull;
        uint8_t *row = dst + (size_t)y * pitch;
        int n = w * 4, i = 0;
        for (; i + 8 <= n; i += 8) { uint64_t v = splitmix64(&st); memcpy(row + i, &v, 8); }
        if (i < n) { uint64_t v = splitmix64(&st); memcpy(row + i, &v, (size_t)(n - i)); }
    }
}

static int negotiate(b200vc_ctx **ctx, int w, int h, int ow, int oh, int fmt)
{
    b200vc_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.in_w = w; cfg.in_h = h; cfg.out_w = ow; cfg.out_h = oh; cfg.out_format = fmt;
    b200vc_default_colorimetry(oh, &cfg.matrix, &cfg.range, &cfg.chroma_site);   /* caps left open */
    if (*ctx) { b200vc_destroy(*ctx); *ctx = NULL; }                             /* renegotiation */
    int rc = b200vc_create(&cfg, ctx);
    if (rc) fprintf(stderr, "set_info failed: %s: %s\n", b200vc_status_string(rc), b200vc_last_error(NULL));
    return rc;
}

int main(int argc, char **argv)
{
    if (argc < 7) { fprintf(stderr, "usage: %s W H OUT_W OUT_H N_FRAMES OUTFILE\n", argv[0]); return 2; }
    const int w = atoi(argv[1]), h = atoi(argv[2]), ow = atoi(argv[3]), oh = atoi(argv[4]), n = atoi(argv[5]);
    FILE *fo = fopen(argv[6], "wb");
    if (!fo) { perror("outfile"); return 2; }
    b200vc_ctx *ctx = NULL;
    uint8_t *planes_dummy[3] = { 0, 0, 0 };
    size_t pit[3] = { 0, 0, 0 };
    /* transform_frame before set_info must be refused, like GST_FLOW_NOT_NEGOTIATED */
    if (b200vc_convert(NULL, NULL, 0, (void *const *)planes_dummy, pit, NULL) != B200VC_ERR_NOT_NEGOTIATED) return 3;

    uint8_t *in = (uint8_t *)malloc((size_t)w * h * 4);
    int fmt = -1;
    for (int k = 0; k < n; k++) {
        const int want = k < n / 2 ? B200VC_FMT_I420 : B200VC_FMT_NV12;
        if (want != fmt) {
            fmt = want;
            if (negotiate(&ctx, w, h, ow, oh, fmt)) return 4;
        }
        size_t off[3], total = b200vc_default_layout(fmt, ow, oh, off, pit);
        uint8_t *out = (uint8_t *)calloc(1, total);
        void *planes[3] = { out + off[0], out + off[1], fmt == B200VC_FMT_I420 ? out + off[2] : NULL };
        synth_rand(in, (size_t)w * 4, w, h, (uint64_t)k);
        int rc = b200vc_convert(ctx, in, (size_t)w * 4, planes, pit, NULL);
        if (rc) { fprintf(stderr, "transform_frame %d failed: %s\n", k, b200vc_last_error(ctx)); return 5; }
        fwrite(out, 1, total, fo);
        free(out);
    }
    b200vc_stats st;
    b200vc_get_stats(ctx, &st);
    printf("frames %llu kernels %llu h2d %llu d2h %llu\n", (unsigned long long)st.frames, (unsigned long long)st.kernel_launches,
           (unsigned long long)st.h2d_bytes, (unsigned long long)st.d2h_bytes);
    b200vc_destroy(ctx);   /* stop */
    fclose(fo);
    free(in);
    return 0;
}
