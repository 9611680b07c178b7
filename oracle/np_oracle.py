"""Independent NumPy restatement of SURVEY.md Appendix A/B (vectorised, plain formulas).

TEST INFRASTRUCTURE ONLY — PARITY UNPINNED (see oracle.h). It exists so that two separately
written implementations of one spec (this file and oracle/*.c) can be compared: agreement catches
transcription slips in either, not errors in the spec itself. It shares no code with the C oracle:
the matrix comes from the table printed in SURVEY A.2 (the C side derives it from Kr/Kb), the
conversion uses the plain per-pixel formula (the C side uses upstream's packed int64 tables).
"""
from __future__ import annotations

import numpy as np

from . import spec as _spec

# the unverified details of SURVEY A/B (include/b200vc_spec.h, overridable through $B200VC_SPEC)
SPEC = _spec.current()

# (matrix, range) -> rows Y,U,V of (cR, cG, cB, off)
MATRICES = {
    (0, 0): [(66, 129, 25, 4096), (-38, -74, 112, 32768), (112, -94, -18, 32768)],
    (0, 1): [(77, 150, 29, 0), (-43, -85, 128, 32768), (128, -107, -21, 32768)],
    (1, 0): [(47, 157, 16, 4096), (-26, -87, 112, 32768), (112, -102, -10, 32768)],
    (1, 1): [(54, 183, 18, 0), (-29, -99, 128, 32768), (128, -116, -12, 32768)],
    # BT.2020 non-constant luminance (Kr 0.2627, Kb 0.0593), same procedure; what GStreamer >= 1.18
    # picks by default for >= 2160 lines
    (2, 0): [(58, 149, 13, 4096), (-31, -81, 112, 32768), (112, -103, -9, 32768)],
    (2, 1): [(67, 174, 15, 0), (-36, -92, 128, 32768), (128, -118, -10, 32768)],
}


# byte index of (R, G, B) in a pixel: BGRx, RGBx, xRGB, xBGR
CHANNELS = {0: (2, 1, 0), 1: (0, 1, 2), 2: (1, 2, 3), 3: (3, 2, 1)}


def unpack16(frame: np.ndarray, in_format: int) -> np.ndarray:
    """(h,w,2) little-endian 5-6-5 pixels (4 RGB16: red in the top bits; 5 BGR16: blue there) -> (h,w,4) BGRx,
    each channel widened to 8 bits by repeating its high bits (or by a plain shift, SPEC UNPACK16_REPLICATE)."""
    v = frame[..., 0].astype(np.uint32) | (frame[..., 1].astype(np.uint32) << 8)
    hi, g, lo = v >> 11, (v >> 5) & 63, v & 31
    hi8, g8, lo8 = hi << 3, g << 2, lo << 3
    if SPEC["UNPACK16_REPLICATE"]:
        hi8, g8, lo8 = hi8 | (hi >> 2), g8 | (g >> 4), lo8 | (lo >> 2)
    r, b = (hi8, lo8) if in_format == 4 else (lo8, hi8)
    out = np.zeros(frame.shape[:2] + (4,), np.uint8)
    out[..., 0], out[..., 1], out[..., 2], out[..., 3] = b, g8, r, 0xFF
    return out


def yuv444(frame: np.ndarray, matrix: int, rng: int, in_format: int = 0):
    """(h,w,4) pixels -> three (h,w) uint8 arrays, (c.RGB + off) >> 8 with no rounding term."""
    if in_format >= 4:
        frame, in_format = unpack16(frame, in_format), 0
    ir, ig, ib = CHANNELS[in_format]
    b = frame[..., ib].astype(np.int32)
    g = frame[..., ig].astype(np.int32)
    r = frame[..., ir].astype(np.int32)
    out = []
    for cr, cg, cb, off in MATRICES[(matrix, rng)]:
        s = cr * r + cg * g + cb * b + off
        if SPEC["MATRIX_ROUND"]:                      # A.2 warning item: +128, clamped like a clipping stage
            s = np.clip(s + 128, 0, 65535)
        assert s.min() >= 0 and s.max() <= 65535
        out.append((s >> 8).astype(np.uint8))
    return out


def chroma_down(c: np.ndarray, site: int, vertical: bool = True) -> np.ndarray:
    """(h,w) 4:4:4 chroma -> (ceil(h/2),ceil(w/2)): horizontal (rounded) on each line, then vertical
    (rounded). Odd widths: the last pixel is a site of its own; odd heights: the last line pairs with
    itself. vertical=False (4:2:2): the horizontal step alone. Of the chroma-site flags only H_COSITED
    (bit 0) matters."""
    site &= 1
    h, w = c.shape
    c = c.astype(np.int32)
    n = (w + 1) // 2
    hz = np.empty((h, n), np.int32)
    if site == 0:  # NONE / JPEG: h2 on full pairs, an unpaired last pixel stays as it is
        hz[:, : w // 2] = (c[:, 0 : 2 * (w // 2) : 2] + c[:, 1 : 2 * (w // 2) : 2] + 1) >> 1
        if w & 1:
            hz[:, n - 1] = c[:, w - 1]
    else:  # H_COSITED: h2_cs
        hz[:, 0] = (3 * c[:, 0] + c[:, 1] + 2) >> 2
        i = 2
        stop = w - 2 if SPEC["COSITED_LAST_1_3"] else w - 1      # A.4 warning item: where the 1-2-1 loop stops
        while i < stop:
            hz[:, i // 2] = (c[:, i - 1] + 2 * c[:, i] + c[:, i + 1] + 2) >> 2
            i += 2
        if i < w:
            hz[:, i // 2] = (c[:, i - 1] + 3 * c[:, i] + 2) >> 2
    if not vertical:
        return hz.astype(np.uint8)
    if h & 1:
        hz = np.concatenate([hz, hz[-1:]], axis=0)
    return ((hz[0::2] + hz[1::2] + 1) >> 1).astype(np.uint8)


def convert(frame: np.ndarray, fmt: int, matrix: int, rng: int, site: int, in_format: int = 0):
    """Returns (Y, U, V) planes for I420 or (Y, UV) for NV12 at the input size."""
    y, u, v = yuv444(frame, matrix, rng, in_format)
    if fmt == 3:                                   # Y444
        return y, u, v
    if fmt == 2:                                   # Y42B
        return y, chroma_down(u, site, vertical=False), chroma_down(v, site, vertical=False)
    u2, v2 = chroma_down(u, site), chroma_down(v, site)
    if fmt == 1:
        uv = np.empty((u2.shape[0], u2.shape[1] * 2), np.uint8)
        uv[:, 0::2], uv[:, 1::2] = u2, v2
        return y, uv
    return y, u2, v2


def h_taps(in_w: int, out_w: int):
    inc = max(0, ((in_w - 1) << 16) // (out_w - 1) - 1) if out_w > 1 else 0
    p = np.arange(out_w, dtype=np.int64) * inc
    return (p >> 16).astype(np.int64), ((p >> 8) & 0xFF).astype(np.int64)


def v_taps(in_h: int, out_h: int):
    y = np.arange(out_h, dtype=np.int64)
    num, den = (2 * y + 1) * in_h - out_h, 2 * out_h
    i = np.floor_divide(num, den)
    rem = num - i * den
    one = 1 << SPEC["VTAP_BITS"]                                 # B.2 warning items: bits and rounding of the tap
    return i, (2 * one * rem + (den if SPEC["VTAP_ROUND"] else 0)) // (2 * den)


def scale_plane(src: np.ndarray, out_w: int, out_h: int, comps: int = 1, mode: int = 0) -> np.ndarray:
    in_h, in_w = src.shape[0], src.shape[1] // comps
    s = src.reshape(in_h, in_w, comps).astype(np.int64)
    if mode == 1:
        assert in_w == 2 * out_w and in_h == 2 * out_h
        d = (s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2
        return d.astype(np.uint8).reshape(out_h, out_w * comps)
    if mode == 2:                                                # method=nearest: one source sample per output sample
        def idx(n_in, n_out):
            k = np.arange(n_out, dtype=np.int64)
            if SPEC["NEAREST_CENTRE"]:
                return np.clip(((2 * k + 1) * n_in) // (2 * n_out), 0, n_in - 1)
            inc = max(0, ((n_in - 1) << 16) // (n_out - 1) - 1) if n_out > 1 else 0
            return np.clip((k * inc) >> 16, 0, n_in - 1)
        return s[idx(in_h, out_h)][:, idx(in_w, out_w)].astype(np.uint8).reshape(out_h, out_w * comps)
    hi, hf = h_taps(in_w, out_w)
    i0, i1 = np.clip(hi, 0, in_w - 1), np.clip(hi + 1, 0, in_w - 1)
    f = hf[None, :, None]
    vi, vp = v_taps(in_h, out_h)
    bits = SPEC["VTAP_BITS"]
    half = 1 << (bits - 1)
    if SPEC["SCALE_H_FIRST"]:                                    # B.2 warning item: order of the passes
        hz = (s[:, i0] * (256 - f) + s[:, i1] * f) >> 8
        a, b = hz[np.clip(vi, 0, in_h - 1)], hz[np.clip(vi + 1, 0, in_h - 1)]
        d = a + (((b - a) * vp[:, None, None] + half) >> bits)
    else:
        a, b = s[np.clip(vi, 0, in_h - 1)], s[np.clip(vi + 1, 0, in_h - 1)]
        vz = np.clip(a + (((b - a) * vp[:, None, None] + half) >> bits), 0, 255)
        d = (vz[:, i0] * (256 - f) + vz[:, i1] * f) >> 8
    return np.clip(d, 0, 255).astype(np.uint8).reshape(out_h, out_w * comps)


def convert_scale(frame, out_w, out_h, fmt, matrix, rng, site, mode=0, in_format=0):
    planes = convert(frame, fmt, matrix, rng, site, in_format)
    if out_w == frame.shape[1] and out_h == frame.shape[0]:
        return planes
    out = [scale_plane(planes[0], out_w, out_h, 1, mode)]
    ocw = out_w if fmt == 3 else (out_w + 1) // 2        # chroma planes of odd frames round up
    och = out_h if fmt >= 2 else (out_h + 1) // 2
    if fmt == 1:
        out.append(scale_plane(planes[1], ocw, och, 2, mode))
    else:
        out.append(scale_plane(planes[1], ocw, och, 1, mode))
        out.append(scale_plane(planes[2], ocw, och, 1, mode))
    return tuple(out)


def convert_scale_borders(frame, out_w, out_h, content, fmt, matrix, rng, site, mode=0, in_format=0):
    """videoscale add-borders: the picture converted and scaled into the rectangle content = (x, y, w, h) of an
    out_w x out_h frame that is otherwise black in the output's colorimetry ((off) >> 8 of the matrix rows)."""
    cx, cy, cw, ch = content
    pic = convert_scale(frame, cw, ch, fmt, matrix, rng, site, mode, in_format)
    rows = MATRICES[(matrix, rng)]
    rnd = 128 if SPEC["MATRIX_ROUND"] else 0
    by, bc = (rows[0][3] + rnd) >> 8, (rows[1][3] + rnd) >> 8
    hs, vs = (1 if fmt == 3 else 2), (2 if fmt <= 1 else 1)
    ocw, och = (out_w + hs - 1) // hs, (out_h + vs - 1) // vs
    y = np.full((out_h, out_w), by, np.uint8)
    y[cy:cy + ch, cx:cx + cw] = pic[0]
    if fmt == 1:
        uv = np.full((och, 2 * ocw), bc, np.uint8)
        uv[cy // 2:cy // 2 + pic[1].shape[0], cx:cx + pic[1].shape[1]] = pic[1]
        return y, uv
    out = [y]
    for p in pic[1:]:
        c = np.full((och, ocw), bc, np.uint8)
        c[cy // vs:cy // vs + p.shape[0], cx // hs:cx // hs + p.shape[1]] = p
        out.append(c)
    return tuple(out)


def damage_map(cur: np.ndarray, prev: np.ndarray | None, in_format: int = 0) -> np.ndarray:
    h, w = cur.shape[:2]
    mh, mw = (h + 15) // 16, (w + 15) // 16
    if prev is None:
        return np.ones((mh, mw), np.uint8)
    ch = [0, 1] if in_format >= 4 else list(CHANNELS[in_format])     # 16-bit pixels: both bytes
    diff = np.any(cur[..., ch] != prev[..., ch], axis=2)
    pad = np.zeros((mh * 16, mw * 16), bool)
    pad[:h, :w] = diff
    return pad.reshape(mh, 16, mw, 16).any(axis=(1, 3)).astype(np.uint8)


def splitmix64(state: int):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def synth_rand(w: int, h: int, frame_index: int) -> np.ndarray:
    """Python mirror of orc_synth_rand (small frames only: pure-Python generator)."""
    out = np.zeros((h, w * 4), np.uint8)
    for y in range(h):
        st = (0x9E3779B97F4A7C15 + frame_index * 0x100000001B3 + y * 0xD6E8FEB86659FD93) & 0xFFFFFFFFFFFFFFFF
        row = bytearray()
        while len(row) < w * 4:
            st, v = splitmix64(st)
            row += v.to_bytes(8, "little")
        out[y] = np.frombuffer(bytes(row[: w * 4]), np.uint8)
    return out.reshape(h, w, 4)
