/*
 * oracle.h — CPU restatement of the BGRx -> I420/NV12 hot path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker or as the timed CPU arm.
 *
 * PARITY UNPINNED. The algorithm restated here lives in GStreamer gst-plugins-base
 * (videoconvert / videoscale), which the reference installs as an unpinned binary bundle
 * (/root/reference Dockerfile:467-471, selkies-gstreamer-entrypoint.sh:8,44-47) and which is
 * absent from this machine. The restatement follows SURVEY.md Appendix A (convert) and
 * Appendix B (scale); the reference holds no test or golden vector for this path, so the
 * vectors under tests/golden/ pin this restatement, not GStreamer itself.
 */
#ifndef B200VC_ORACLE_H
#define B200VC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_FMT_I420 = 0, ORC_FMT_NV12 = 1, ORC_FMT_Y42B = 2, ORC_FMT_Y444 = 3 };
enum { ORC_MATRIX_BT601 = 0, ORC_MATRIX_BT709 = 1, ORC_MATRIX_BT2020 = 2 };
enum { ORC_RANGE_LIMITED = 0, ORC_RANGE_FULL = 1 };
/* GstVideoChromaSite flags; only H_COSITED changes videoconvert's output (see b200vc.h) */
enum { ORC_SITE_NONE = 0 /* "jpeg" */, ORC_SITE_H_COSITED = 1 /* "mpeg2" */, ORC_SITE_V_COSITED = 2, ORC_SITE_ALT_LINE = 4 };
enum { ORC_SCALE_GST_LQ = 0, ORC_SCALE_BOX_2X2 = 1, ORC_SCALE_NEAREST = 2 /* videoscale method=nearest */ };
/* input byte orders: unpack_BGRA, unpack_RGBA, unpack_ARGB, unpack_ABGR (alpha/pad byte ignored) */
enum { ORC_IN_BGRX = 0, ORC_IN_RGBX = 1, ORC_IN_XRGB = 2, ORC_IN_XBGR = 3,
       ORC_IN_RGB16 = 4, ORC_IN_BGR16 = 5 /* unpack_RGB16 / unpack_BGR16: 2 bytes per pixel, 5-6-5, little endian */ };

/* 3x4 matrix in 8.8 fixed point: rows Y,U,V; columns cR,cG,cB,offset. */
typedef struct {
    int32_t m[3][4];
} orc_matrix;

/* The values of include/b200vc_spec.h this oracle was built with ("m0c1h1b6r1u1n1"), see there. */
const char *orc_spec_string(void);

/* compute_matrix_to_YUV restated: double-precision build, x256, rint() per entry. */
void orc_build_matrix(int matrix, int range, orc_matrix *out);

/* Default plane layout (video-info.c restated). offsets/pitches have 3 entries;
 * NV12 uses entries 0 and 1. Returns total size in bytes. */
size_t orc_default_layout(int format, int w, int h, size_t offsets[3], size_t pitches[3]);

/* videoconvert: BGRx (pitch in_pitch) -> planes. planes[2]/pitches[2] unused for NV12.
 * Rows [row0,row1) of the output luma are produced (row0,row1 even or row1==h): this is the
 * horizontal-band split upstream uses for n-threads > 1. Returns 0, or -1 on bad arguments. */
int orc_convert_rows(const uint8_t *bgrx, size_t in_pitch, int w, int h,
                     int format, int matrix, int range, int site,
                     uint8_t *const planes[3], const size_t pitches[3],
                     int row0, int row1);

int orc_convert(const uint8_t *bgrx, size_t in_pitch, int w, int h,
                int format, int matrix, int range, int site,
                uint8_t *const planes[3], const size_t pitches[3]);

/* Same, split over n_threads horizontal bands (pthreads). */
int orc_convert_mt(const uint8_t *bgrx, size_t in_pitch, int w, int h,
                   int format, int matrix, int range, int site,
                   uint8_t *const planes[3], const size_t pitches[3], int n_threads);

/* videoscale, one 8-bit plane with `comps` interleaved components (1 = Y/U/V, 2 = NV12 UV). */
int orc_scale_plane(const uint8_t *src, size_t src_pitch, int in_w, int in_h,
                    uint8_t *dst, size_t dst_pitch, int out_w, int out_h,
                    int comps, int scale_mode);

/* videoconvert ! videoscale: convert at input size, then scale each plane. */
int orc_convert_scale(const uint8_t *bgrx, size_t in_pitch, int in_w, int in_h,
                      int out_w, int out_h, int format, int matrix, int range, int site,
                      int scale_mode, uint8_t *const planes[3], const size_t pitches[3]);

int orc_convert_scale_mt(const uint8_t *bgrx, size_t in_pitch, int in_w, int in_h,
                         int out_w, int out_h, int format, int matrix, int range, int site,
                         int scale_mode, uint8_t *const planes[3], const size_t pitches[3],
                         int n_threads);

/* The same entry points for the other 4-byte input formats (in_format = ORC_IN_*). */
int orc_convert_scale_mt_ex(const uint8_t *src, size_t in_pitch, int in_format, int in_w, int in_h,
                            int out_w, int out_h, int format, int matrix, int range, int site,
                            int scale_mode, uint8_t *const planes[3], const size_t pitches[3],
                            int n_threads);
int orc_damage_map_ex(const uint8_t *cur, size_t cur_pitch, const uint8_t *prev, size_t prev_pitch,
                      int in_format, int w, int h, uint8_t *map);

/* Scaler taps, exposed so tests can pin them. */
void orc_scale_h_taps(int in_w, int out_w, int32_t *idx, int32_t *frac);   /* frac 0..255 */
void orc_scale_v_taps(int in_h, int out_h, int32_t *idx, int32_t *p1);     /* p1 0..64   */

/* Damage map: map[my*mb_w+mx] = 1 iff any pixel of macroblock (mx,my) differs
 * from prev in B, G or R. prev == NULL: all 1. mb_w = ceil(w/16), mb_h = ceil(h/16). */
int orc_damage_map(const uint8_t *cur, size_t cur_pitch, const uint8_t *prev, size_t prev_pitch,
                   int w, int h, uint8_t *map);

/* Synthetic frames (SURVEY §8 d). */
uint64_t orc_splitmix64(uint64_t *state);
void orc_synth_rand(uint8_t *dst, size_t pitch, int w, int h, uint64_t frame_index);
void orc_synth_bars(uint8_t *dst, size_t pitch, int w, int h);
void orc_synth_desk(uint8_t *dst, size_t pitch, int w, int h, uint64_t frame_index);

#ifdef __cplusplus
}
#endif
#endif
