/*
 * b200vc_spec.h — the arithmetic details of `videoconvert ! videoscale` that this repository could
 * not check against GStreamer, one constant each.
 *
 * The reference (COx2/docker-nvidia-glx-desktop) holds no source for this path: it installs
 * GStreamer as an unpinned binary (Dockerfile:467-471). The arithmetic was restated from
 * SURVEY.md Appendix A/B, where the points below carry a warning sign. Everything that evaluates
 * one of them reads the constant from this header and nothing else:
 *
 *   product   csrc/host_tables.h (matrix offsets, scaler taps, dp4a weight words), the kernels
 *             (k1_convert.cu, k2_scale.cu, k2_plane_tile.cu, noalloc.cu)
 *   oracle    oracle/convert.c, oracle/scale.c; oracle/spec.py parses this file for np_oracle.py
 *
 * so the day someone runs tools/gst_validate.py on a machine with GStreamer and one of these turns
 * out wrong, the correction is one edited default below (or a -D on the compiler line) and nothing
 * else. `make variant NAME=... SPEC="-D..."` and `oracle.build(spec=...)` build library and oracle with
 * another value; tests/test_spec_variants.py runs the parity sweep on every value listed in
 * oracle/spec.py VARIANTS.
 *
 * Plain C preprocessor only: this file is included by C (gcc), C++/CUDA (nvcc) and read by Python.
 */
#ifndef B200VC_SPEC_H
#define B200VC_SPEC_H

/* SURVEY A.2 — the 8.8 matrix stage. 0: Y = (c.RGB + off) >> 8, no rounding term (what
 * video_converter_matrix8_table does as recalled). 1: every offset grows by 128 (round to nearest);
 * sums that then reach 65536 (full-range U/V of a saturated primary) clamp to 255. */
#ifndef B200VC_SPEC_MATRIX_ROUND
#define B200VC_SPEC_MATRIX_ROUND 0
#endif

/* SURVEY A.4 — video_chroma_down_h2_cs_u8, last even column of an even-width line. 1: it takes
 * itself as right neighbour, (u[W-3] + 3*u[W-2] + 2) >> 2, u[W-1] is ignored (as recalled).
 * 0: the plain 1-2-1 filter runs to the end, (u[W-3] + 2*u[W-2] + u[W-1] + 2) >> 2. */
#ifndef B200VC_SPEC_COSITED_LAST_1_3
#define B200VC_SPEC_COSITED_LAST_1_3 1
#endif

/* SURVEY B.2 — order of the two 2-tap passes of videoscale. 1: horizontal, then vertical over the
 * horizontally scaled lines (as recalled for downscales). 0: vertical first. The two do not
 * commute bit for bit. (With 0 every scale negotiates onto the two-pass path: the fused kernels
 * implement the horizontal-first order only.) */
#ifndef B200VC_SPEC_SCALE_H_FIRST
#define B200VC_SPEC_SCALE_H_FIRST 1
#endif

/* SURVEY B.2 — precision of the vertical tap in bits (6: the "lq" ORC path as recalled; 8 is the
 * other candidate) and how the fractional position becomes the tap: 1 round to nearest, 0 floor.
 * d = a + (((b - a) * p1 + 2^(bits-1)) >> bits), p1 in 0 .. 2^bits. */
#ifndef B200VC_SPEC_VTAP_BITS
#define B200VC_SPEC_VTAP_BITS 6
#endif
#ifndef B200VC_SPEC_VTAP_ROUND
#define B200VC_SPEC_VTAP_ROUND 1
#endif

/* video-format.c — unpacking 5- and 6-bit channels (RGB16/BGR16 input) to 8 bits. 1: the high bits are
 * repeated into the low ones, (v << 3) | (v >> 2) and (v << 2) | (v >> 4), so that full scale maps to 255
 * (as recalled, the default). 0: plain shift (what GST_VIDEO_PACK_FLAG_TRUNCATE_RANGE selects). */
#ifndef B200VC_SPEC_UNPACK16_REPLICATE
#define B200VC_SPEC_UNPACK16_REPLICATE 1
#endif

/* videoscale method=nearest — which source sample an output sample takes. 1: the one under the output
 * pixel's centre, floor((x + 0.5) * in / out), both directions (the resampler's centre alignment, as
 * recalled). 0: edge-aligned, the bilinear path's 16.16 step without its fraction, (x * inc) >> 16, horizontally
 * and vertically alike. */
#ifndef B200VC_SPEC_NEAREST_CENTRE
#define B200VC_SPEC_NEAREST_CENTRE 1
#endif

#if B200VC_SPEC_VTAP_BITS != 6 && B200VC_SPEC_VTAP_BITS != 7 && B200VC_SPEC_VTAP_BITS != 8
#error "B200VC_SPEC_VTAP_BITS must be 6, 7 or 8 (the kernels evaluate the tap as 8-bit dp4a weights)"
#endif

#define B200VC_SPEC_VTAP_ONE  (1 << B200VC_SPEC_VTAP_BITS)
#define B200VC_SPEC_VTAP_HALF (1 << (B200VC_SPEC_VTAP_BITS - 1))

/* "m0c1h1b6r1u1n1": the values above in one token; b200vc_spec_string() and orc_spec_string()
 * return it so that a test can see that library and oracle were built from the same settings. */
#define B200VC_SPEC_STR2(x) #x
#define B200VC_SPEC_STR(x) B200VC_SPEC_STR2(x)
#define B200VC_SPEC_STRING "m" B200VC_SPEC_STR(B200VC_SPEC_MATRIX_ROUND) "c" B200VC_SPEC_STR(B200VC_SPEC_COSITED_LAST_1_3) \
                           "h" B200VC_SPEC_STR(B200VC_SPEC_SCALE_H_FIRST) "b" B200VC_SPEC_STR(B200VC_SPEC_VTAP_BITS)       \
                           "r" B200VC_SPEC_STR(B200VC_SPEC_VTAP_ROUND) "u" B200VC_SPEC_STR(B200VC_SPEC_UNPACK16_REPLICATE) \
                           "n" B200VC_SPEC_STR(B200VC_SPEC_NEAREST_CENTRE)

#endif /* B200VC_SPEC_H */
