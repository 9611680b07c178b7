"""ctypes loader for the CPU oracle (oracle/*.c).

TEST INFRASTRUCTURE ONLY — PARITY UNPINNED. Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package; the product
(docker-nvidia-glx-desktop_b200/) never does. The oracle restates GStreamer videoconvert /
videoscale from SURVEY.md Appendix A/B: the reference (/root/reference Dockerfile:467-471)
installs GStreamer as an unpinned binary and holds neither its source nor any test vector.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import spec as _spec

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRCS = ["convert.c", "scale.c", "damage_synth.c"]

FMT_I420, FMT_NV12, FMT_Y42B, FMT_Y444 = 0, 1, 2, 3
MATRIX_BT601, MATRIX_BT709, MATRIX_BT2020 = 0, 1, 2
RANGE_LIMITED, RANGE_FULL = 0, 1
SITE_NONE, SITE_H_COSITED = 0, 1
SCALE_GST_LQ, SCALE_BOX_2X2, SCALE_NEAREST = 0, 1, 2
IN_BGRX, IN_RGBX, IN_XRGB, IN_XBGR, IN_RGB16, IN_BGR16 = 0, 1, 2, 3, 4, 5


def build(native: bool = False, out_dir: str | None = None, spec: dict | None = None) -> str:
    """Compile the oracle with gcc into oracle/_build/ (x86-64-v3: the file travels to the GPU box), or
    with native=True into oracle/_build_native/ (-march=native, built on the box that runs it: the
    timed CPU arm gets the host's full ISA). `spec` (default: include/b200vc_spec.h, then $B200VC_SPEC)
    selects the values of the spec constants; other values than the header's build into a
    subdirectory named after them."""
    sp = _spec.current(spec)
    march = "native" if native else "x86-64-v3"
    if out_dir is None:
        out_dir = os.path.join(_HERE, "_build_native" if native else "_build")
        if not _spec.is_default(sp):
            out_dir = os.path.join(out_dir, _spec.name(sp))
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "liboracle.so")
    srcs = [os.path.join(_HERE, s) for s in _SRCS]
    deps = srcs + [os.path.join(_HERE, "oracle.h"), os.path.join(os.path.dirname(_HERE), "include", "b200vc_spec.h")]
    newest = max(os.path.getmtime(p) for p in deps)
    if os.path.exists(out) and os.path.getmtime(out) >= newest:
        return out
    tmp = out + f".{os.getpid()}.tmp"                       # ranks of one job may build at the same time
    cmd = [os.environ.get("CC", "gcc"), "-O3", f"-march={march}", "-fPIC", "-std=c11", "-D_GNU_SOURCE",
           "-shared", "-o", tmp] + _spec.cflags(sp) + srcs + ["-lm", "-lpthread"]
    subprocess.run(cmd, check=True)
    os.replace(tmp, out)
    return out


class Oracle:
    """Thin typed wrapper over liboracle.so working on numpy uint8 arrays."""

    def __init__(self, path: str | None = None, native: bool = False, spec: dict | None = None):
        self.path = path or build(native=native, spec=spec)
        L = self.lib = C.CDLL(self.path)
        L.orc_spec_string.restype = C.c_char_p
        self.spec = _spec.current(spec)
        if not path and L.orc_spec_string().decode() != _spec.name(self.spec):
            raise RuntimeError(f"{self.path} was built as {L.orc_spec_string().decode()}, wanted {_spec.name(self.spec)}")
        u8p, sz, i = C.c_void_p, C.c_size_t, C.c_int
        pl, pi = C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)
        L.orc_build_matrix.argtypes = [i, i, C.POINTER(C.c_int32 * 12)]
        L.orc_default_layout.argtypes = [i, i, i, pi, pi]
        L.orc_default_layout.restype = sz
        L.orc_convert_mt.argtypes = [u8p, sz, i, i, i, i, i, i, pl, pi, i]
        L.orc_convert_rows.argtypes = [u8p, sz, i, i, i, i, i, i, pl, pi, i, i]
        L.orc_scale_plane.argtypes = [u8p, sz, i, i, u8p, sz, i, i, i, i]
        L.orc_convert_scale_mt.argtypes = [u8p, sz, i, i, i, i, i, i, i, i, i, pl, pi, i]
        L.orc_convert_scale_mt_ex.argtypes = [u8p, sz, i, i, i, i, i, i, i, i, i, i, pl, pi, i]
        L.orc_damage_map_ex.argtypes = [u8p, sz, u8p, sz, i, i, i, u8p]
        L.orc_scale_h_taps.argtypes = [i, i, u8p, u8p]
        L.orc_scale_v_taps.argtypes = [i, i, u8p, u8p]
        L.orc_damage_map.argtypes = [u8p, sz, u8p, sz, i, i, u8p]
        L.orc_synth_rand.argtypes = [u8p, sz, i, i, C.c_uint64]
        L.orc_synth_bars.argtypes = [u8p, sz, i, i]
        L.orc_synth_desk.argtypes = [u8p, sz, i, i, C.c_uint64]

    # -- spec tables ---------------------------------------------------------------------------
    def matrix(self, matrix: int, rng: int) -> np.ndarray:
        m = (C.c_int32 * 12)()
        self.lib.orc_build_matrix(matrix, rng, C.byref(m))
        return np.array(list(m), dtype=np.int32).reshape(3, 4)

    def default_layout(self, fmt: int, w: int, h: int):
        off, pit = (C.c_size_t * 3)(), (C.c_size_t * 3)()
        total = self.lib.orc_default_layout(fmt, w, h, off, pit)
        return int(total), [int(x) for x in off], [int(x) for x in pit]

    def h_taps(self, in_w: int, out_w: int):
        idx, fr = np.zeros(out_w, np.int32), np.zeros(out_w, np.int32)
        self.lib.orc_scale_h_taps(in_w, out_w, idx.ctypes.data, fr.ctypes.data)
        return idx, fr

    def v_taps(self, in_h: int, out_h: int):
        idx, p1 = np.zeros(out_h, np.int32), np.zeros(out_h, np.int32)
        self.lib.orc_scale_v_taps(in_h, out_h, idx.ctypes.data, p1.ctypes.data)
        return idx, p1

    # -- the path ------------------------------------------------------------------------------
    def _planes(self, fmt, w, h, pitches):
        if pitches is None:
            total, off, pit = self.default_layout(fmt, w, h)
        else:
            pit = list(pitches) + [0] * (3 - len(pitches))
            ch = h if fmt >= FMT_Y42B else (h + 1) // 2
            rows = [h, ch, ch]
            off, total = [], 0
            for p in range(2 if fmt == FMT_NV12 else 3):
                off.append(total)
                total += pit[p] * rows[p]
            off += [0] * (3 - len(off))
        buf = np.zeros(total, np.uint8)
        ptrs = (C.c_void_p * 3)(*[buf.ctypes.data + o for o in off])
        pits = (C.c_size_t * 3)(*pit)
        return buf, off, pit, ptrs, pits

    def convert(self, frame: np.ndarray, fmt=FMT_I420, matrix=MATRIX_BT601, rng=RANGE_LIMITED,
                site=SITE_NONE, out_size=None, scale_mode=SCALE_GST_LQ, pitches=None, threads=1, in_format=IN_BGRX,
                content=None):
        """frame: (h, w, 4) uint8 pixels, or (h, w, 2) for the 16-bit formats (may be a strided view).
        Returns (buf, offsets, pitches)."""
        h, w = frame.shape[:2]
        bpp = 2 if in_format >= IN_RGB16 else 4
        assert frame.dtype == np.uint8 and frame.shape[2] == bpp and frame.strides[2] == 1 and frame.strides[1] == bpp
        ow, oh = out_size if out_size else (w, h)
        buf, off, pit, ptrs, pits = self._planes(fmt, ow, oh, pitches)
        if content and tuple(content) != (0, 0, ow, oh):
            # videoscale add-borders (gst_video_converter with a destination rectangle and fill-border): the
            # frame is black in the output's colorimetry — the matrix's answer to RGB 0,0,0 — and the picture is
            # converted and scaled into its rectangle (plane pointers moved to the rectangle's origin)
            cx, cy, cw, ch = content
            hs, vs = (1 if fmt == FMT_Y444 else 2), (2 if fmt <= FMT_NV12 else 1)
            black = self.matrix(matrix, rng)[:, 3] >> 8
            rows = [oh, oh if fmt >= FMT_Y42B else (oh + 1) // 2]
            buf[off[0]:off[0] + pit[0] * rows[0]] = black[0]
            for p in range(1, 2 if fmt == FMT_NV12 else 3):
                buf[off[p]:off[p] + pit[p] * rows[1]] = black[1]
            shifted = [buf.ctypes.data + off[0] + cy * pit[0] + cx]
            for p in range(1, 2 if fmt == FMT_NV12 else 3):
                shifted.append(buf.ctypes.data + off[p] + (cy // vs) * pit[p] + (cx // hs) * (2 if fmt == FMT_NV12 else 1))
            ptrs = (C.c_void_p * 3)(*(shifted + [None] * (3 - len(shifted))))
            ow, oh = cw, ch
        rc = self.lib.orc_convert_scale_mt_ex(frame.ctypes.data, frame.strides[0], in_format, w, h, ow, oh, fmt, matrix,
                                              rng, site, scale_mode, ptrs, pits, threads)
        if rc:
            raise ValueError("oracle rejected the arguments")
        return buf, off, pit

    def convert_into(self, frame_ptr: int, in_pitch: int, w: int, h: int, ow: int, oh: int, fmt, matrix, rng,
                     site, scale_mode, ptrs, pits, threads: int) -> int:
        """Raw-pointer variant for the timed CPU arm (no allocation inside the timed region)."""
        return self.lib.orc_convert_scale_mt(frame_ptr, in_pitch, w, h, ow, oh, fmt, matrix, rng, site,
                                             scale_mode, ptrs, pits, threads)

    def scale_plane(self, src: np.ndarray, out_w: int, out_h: int, comps=1, scale_mode=SCALE_GST_LQ):
        in_h, in_w = src.shape[0], src.shape[1] // comps
        dst = np.zeros((out_h, out_w * comps), np.uint8)
        rc = self.lib.orc_scale_plane(src.ctypes.data, src.strides[0], in_w, in_h, dst.ctypes.data,
                                      dst.strides[0], out_w, out_h, comps, scale_mode)
        if rc:
            raise ValueError("oracle rejected the arguments")
        return dst

    def damage_map(self, cur: np.ndarray, prev: np.ndarray | None, in_format=IN_BGRX) -> np.ndarray:
        h, w = cur.shape[:2]
        m = np.zeros(((h + 15) // 16, (w + 15) // 16), np.uint8)
        self.lib.orc_damage_map_ex(cur.ctypes.data, cur.strides[0], prev.ctypes.data if prev is not None else None,
                                   prev.strides[0] if prev is not None else 0, in_format, w, h, m.ctypes.data)
        return m

    # -- synthetic frames ------------------------------------------------------------------------
    def synth(self, kind: str, w: int, h: int, frame_index: int = 0, pitch: int | None = None,
              out: np.ndarray | None = None) -> np.ndarray:
        pitch = pitch or w * 4
        if out is None:
            out = np.zeros((h, pitch), np.uint8)
        ptr = out.ctypes.data
        if kind == "rand":
            self.lib.orc_synth_rand(ptr, pitch, w, h, frame_index)
        elif kind == "bars":
            self.lib.orc_synth_bars(ptr, pitch, w, h)
        elif kind == "desk":
            self.lib.orc_synth_desk(ptr, pitch, w, h, frame_index)
        else:
            raise ValueError(kind)
        return np.ndarray((h, w, 4), np.uint8, buffer=out, strides=(pitch, 4, 1))


_default: Oracle | None = None


def get() -> Oracle:
    global _default
    if _default is None:
        _default = Oracle()
    return _default
