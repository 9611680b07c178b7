"""ctypes binding of libb200vc.so, shaped like the element it stands in for.

The reference runs this path as a GStreamer GstVideoFilter (`videoconvert ! videoscale`, installed
at /root/reference Dockerfile:467-471 and selected by WEBRTC_ENCODER=x264enc, README.md:21). The
class below keeps that life cycle and its names — set_info / transform_frame / stop, caps given as
format, colorimetry and chroma-site — so the parity tests read like tests of the element.

There is no fallback of any kind here: if libb200vc.so is missing or no sm_100 GPU is usable, the
constructor or set_info raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200VC_LIB") or os.path.join(_HERE, "libb200vc.so")   # (override: A/B builds of the library)

OK = 0
ERR_INVALID_ARG, ERR_UNSUPPORTED, ERR_NOT_NEGOTIATED, ERR_CUDA, ERR_NO_DEVICE, ERR_ALIGNMENT, ERR_NOMEM, ERR_BAD_TICKET = range(1, 9)
FMT_I420, FMT_NV12, FMT_Y42B, FMT_Y444 = 0, 1, 2, 3
MATRIX_BT601, MATRIX_BT709, MATRIX_BT2020 = 0, 1, 2
RANGE_LIMITED, RANGE_FULL = 0, 1
SITE_NONE, SITE_H_COSITED, SITE_V_COSITED, SITE_ALT_LINE, SITE_COSITED, SITE_DV = 0, 1, 2, 4, 3, 7
SCALE_GST_LQ, SCALE_BOX_2X2, SCALE_NEAREST = 0, 1, 2
IN_BGRX, IN_RGBX, IN_XRGB, IN_XBGR, IN_RGB16, IN_BGR16 = 0, 1, 2, 3, 4, 5
FLAG_DAMAGE_MAP, FLAG_FORCE_TWO_PASS, FLAG_REGISTER_HOST, FLAG_FORCE_GENERIC_FUSED, FLAG_ZERO_COPY, FLAG_ZERO_COPY_OUT = 1, 2, 4, 8, 16, 32
FLAG_ROW_GATE = 64

# every symbol include/b200vc.h declares
SYMBOLS = [
    "b200vc_default_colorimetry", "b200vc_default_layout", "b200vc_fit_borders", "b200vc_check_config", "b200vc_status_string",
    "b200vc_version", "b200vc_spec_string", "b200vc_create", "b200vc_destroy", "b200vc_last_error", "b200vc_get_info",
    "b200vc_get_stats", "b200vc_reset", "b200vc_convert", "b200vc_convert_session", "b200vc_convert_batch", "b200vc_submit",
    "b200vc_collect", "b200vc_host_alloc", "b200vc_host_free", "b200vc_convert_device",
    "b200vc_convert_scale_device_static", "b200vc_damage_to_rects",
]
STATIC_SCRATCH_BYTES = 3133440


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("in_w", C.c_int32), ("in_h", C.c_int32), ("out_w", C.c_int32),
                ("out_h", C.c_int32), ("out_format", C.c_int32), ("matrix", C.c_int32), ("range", C.c_int32),
                ("chroma_site", C.c_int32), ("scale_mode", C.c_int32), ("flags", C.c_uint32),
                ("max_sessions", C.c_int32), ("n_inflight", C.c_int32), ("n_bands", C.c_int32), ("in_format", C.c_int32),
                ("content_x", C.c_int32), ("content_y", C.c_int32), ("content_w", C.c_int32), ("content_h", C.c_int32)]


class Frame(C.Structure):
    _fields_ = [("bgrx", C.c_void_p), ("in_pitch", C.c_size_t), ("planes", C.c_void_p * 3),
                ("pitches", C.c_size_t * 3), ("damage", C.c_void_p), ("session", C.c_int32), ("reserved", C.c_int32)]


class DeviceFrames(C.Structure):
    _fields_ = [("bgrx", C.c_void_p), ("in_pitch", C.c_size_t), ("in_frame_stride", C.c_size_t),
                ("prev", C.c_void_p), ("prev_pitch", C.c_size_t), ("prev_frame_stride", C.c_size_t),
                ("planes", C.c_void_p * 3), ("pitches", C.c_size_t * 3), ("out_frame_stride", C.c_size_t),
                ("damage", C.c_void_p), ("damage_frame_stride", C.c_size_t), ("n_frames", C.c_int32),
                ("reserved", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("frames", C.c_uint64), ("kernel_launches", C.c_uint64), ("h2d_bytes", C.c_uint64),
                ("d2h_bytes", C.c_uint64), ("staged_in_bytes", C.c_uint64), ("staged_out_bytes", C.c_uint64)]


class Info(C.Structure):
    _fields_ = [("path", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32), ("box_w", C.c_int32),
                ("box_h", C.c_int32), ("smem_bytes", C.c_int32), ("sm_count", C.c_int32), ("cc_major", C.c_int32),
                ("cc_minor", C.c_int32), ("fused_2to1", C.c_int32), ("device_name", C.c_char * 64)]


class Rect(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("w", C.c_int32), ("h", C.c_int32)]


class B200vcError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"b200vc status {status}: {message}")
        self.status = status


_lib = None


def load() -> C.CDLL:
    """dlopen libb200vc.so from the tree. Raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() or `make -C {_HERE}`")
    L = C.CDLL(LIB_PATH)
    vp, sz, i = C.c_void_p, C.c_size_t, C.c_int
    L.b200vc_default_colorimetry.argtypes = [i, C.POINTER(i), C.POINTER(i), C.POINTER(i)]
    L.b200vc_default_layout.argtypes = [i, i, i, C.POINTER(sz), C.POINTER(sz)]
    L.b200vc_default_layout.restype = sz
    L.b200vc_fit_borders.argtypes = [i] * 8 + [C.POINTER(i)] * 4
    L.b200vc_check_config.argtypes = [C.POINTER(Config)]
    L.b200vc_status_string.argtypes = [i]
    L.b200vc_status_string.restype = C.c_char_p
    L.b200vc_spec_string.restype = C.c_char_p
    L.b200vc_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.b200vc_destroy.argtypes = [vp]
    L.b200vc_last_error.argtypes = [vp]
    L.b200vc_last_error.restype = C.c_char_p
    L.b200vc_get_info.argtypes = [vp, C.POINTER(Info)]
    L.b200vc_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.b200vc_reset.argtypes = [vp, i]
    L.b200vc_convert.argtypes = [vp, vp, sz, C.POINTER(vp), C.POINTER(sz), vp]
    L.b200vc_convert_session.argtypes = [vp, i, vp, sz, C.POINTER(vp), C.POINTER(sz), vp]
    L.b200vc_convert_batch.argtypes = [vp, C.POINTER(Frame), i]
    L.b200vc_submit.argtypes = [vp, C.POINTER(Frame), C.POINTER(C.c_uint64)]
    L.b200vc_collect.argtypes = [vp, C.c_uint64]
    L.b200vc_host_alloc.argtypes = [sz, C.POINTER(vp)]
    L.b200vc_host_free.argtypes = [vp]
    L.b200vc_convert_device.argtypes = [vp, C.POINTER(DeviceFrames), vp]
    L.b200vc_convert_scale_device_static.argtypes = [C.POINTER(Config), C.POINTER(DeviceFrames), vp, C.c_char_p, sz]
    L.b200vc_damage_to_rects.argtypes = [vp, i, i, C.POINTER(Rect), i]
    _lib = L
    return L


def default_colorimetry(out_h: int):
    m, r, s = C.c_int(), C.c_int(), C.c_int()
    rc = load().b200vc_default_colorimetry(out_h, C.byref(m), C.byref(r), C.byref(s))
    if rc:
        raise B200vcError(rc, "bad height")
    return m.value, r.value, s.value


def default_layout(fmt: int, w: int, h: int):
    off, pit = (C.c_size_t * 3)(), (C.c_size_t * 3)()
    total = load().b200vc_default_layout(fmt, w, h, off, pit)
    return int(total), [int(x) for x in off], [int(x) for x in pit]


def fit_borders(in_w: int, in_h: int, out_w: int, out_h: int, in_par=(1, 1), out_par=(1, 1)):
    """videoscale add-borders: (x, y, w, h) of the picture inside the out_w x out_h frame."""
    v = [C.c_int() for _ in range(4)]
    rc = load().b200vc_fit_borders(in_w, in_h, in_par[0], in_par[1], out_w, out_h, out_par[0], out_par[1], *[C.byref(a) for a in v])
    if rc:
        raise B200vcError(rc, "bad sizes or pixel-aspect-ratios")
    return tuple(a.value for a in v)


def damage_rects(dmg: np.ndarray, frame_w: int, frame_h: int, max_rects: int = 4096):
    """Dirty macroblocks of a damage map as a list of (x, y, w, h) pixel rectangles."""
    dmg = np.ascontiguousarray(dmg, np.uint8)
    arr = (Rect * max_rects)()
    n = load().b200vc_damage_to_rects(dmg.ctypes.data, frame_w, frame_h, arr, max_rects)
    if n < 0:
        raise B200vcError(ERR_INVALID_ARG, "bad damage map arguments")
    return [(r.x, r.y, r.w, r.h) for r in arr[: min(n, max_rects)]], n


class PinnedBuffer:
    """Page-locked host memory from b200vc_host_alloc, viewed as a numpy uint8 array."""

    def __init__(self, nbytes: int):
        p = C.c_void_p()
        rc = load().b200vc_host_alloc(nbytes, C.byref(p))
        if rc:
            raise B200vcError(rc, load().b200vc_last_error(None).decode())
        self.ptr, self.nbytes = p.value, nbytes
        self.array = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,))

    def free(self):
        if self.ptr:
            load().b200vc_host_free(self.ptr)
            self.ptr, self.array = None, None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Converter:
    """The element: set_info() negotiates, transform_frame() converts, stop() releases."""

    def __init__(self, device: int = 0):
        self.lib = load()
        self.device = device
        self.ctx = C.c_void_p()
        self.cfg: Config | None = None

    # -- GstVideoFilter::set_info ----------------------------------------------------------------
    def set_info(self, in_w: int, in_h: int, out_w: int | None = None, out_h: int | None = None,
                 out_format: int = FMT_I420, matrix: int | None = None, range_: int | None = None,
                 chroma_site: int | None = None, scale_mode: int = SCALE_GST_LQ, flags: int = 0,
                 max_sessions: int = 1, n_inflight: int = 0, n_bands: int = 0, in_format: int = IN_BGRX,
                 content: tuple | None = None) -> None:
        out_w, out_h = out_w or in_w, out_h or in_h
        if matrix is None or range_ is None or chroma_site is None:
            dm, dr, ds = default_colorimetry(out_h)   # caps left colorimetry open: defaults by height
            matrix = dm if matrix is None else matrix
            range_ = dr if range_ is None else range_
            chroma_site = ds if chroma_site is None else chroma_site
        self.stop()
        cfg = Config(self.device, in_w, in_h, out_w, out_h, out_format, matrix, range_, chroma_site, scale_mode,
                     flags, max_sessions, n_inflight, n_bands, in_format, *(content or (0, 0, 0, 0)))
        ctx = C.c_void_p()
        rc = self.lib.b200vc_create(C.byref(cfg), C.byref(ctx))
        if rc:
            raise B200vcError(rc, self.lib.b200vc_last_error(None).decode())
        self.ctx, self.cfg = ctx, cfg

    def _check(self, rc: int) -> None:
        if rc:
            raise B200vcError(rc, self.lib.b200vc_last_error(self.ctx).decode() if self.ctx else "not negotiated")

    @property
    def mb_shape(self):
        return ((self.cfg.in_h + 15) // 16, (self.cfg.in_w + 15) // 16)

    def out_layout(self):
        return default_layout(self.cfg.out_format, self.cfg.out_w, self.cfg.out_h)

    def info(self) -> Info:
        inf = Info()
        self._check(self.lib.b200vc_get_info(self.ctx, C.byref(inf)))
        return inf

    def stats(self) -> Stats:
        st = Stats()
        self._check(self.lib.b200vc_get_stats(self.ctx, C.byref(st)))
        return st

    def reset(self, session: int = -1) -> None:
        self._check(self.lib.b200vc_reset(self.ctx, session))

    # -- GstVideoFilter::transform_frame ---------------------------------------------------------
    def transform_frame(self, frame: np.ndarray, out: np.ndarray | None = None, pitches=None, damage: bool = False,
                        session: int = 0):
        """frame: (h, w, 4) uint8 pixels of the negotiated in_format ((h, w, 2) for RGB16/BGR16), row stride
        free. Returns (buf, offsets, pitches[, map])."""
        if not self.ctx:
            raise B200vcError(ERR_NOT_NEGOTIATED, "set_info has not succeeded")
        g = self.cfg
        if pitches is None:
            total, off, pit = self.out_layout()
        else:
            pit = list(pitches) + [0] * (3 - len(pitches))
            ch = g.out_h if g.out_format >= FMT_Y42B else (g.out_h + 1) // 2
            rows = [g.out_h, ch, ch]
            off, total = [], 0
            for p in range(2 if g.out_format == FMT_NV12 else 3):
                off.append(total)
                total += pit[p] * rows[p]
            off += [0] * (3 - len(off))
        if out is None:
            out = np.zeros(total, np.uint8)
        planes = (C.c_void_p * 3)(*[out.ctypes.data + o for o in off])
        pits = (C.c_size_t * 3)(*pit)
        dmg = np.zeros(self.mb_shape, np.uint8) if damage else None
        self._check(self.lib.b200vc_convert_session(self.ctx, session, frame.ctypes.data, frame.strides[0], planes, pits,
                                                    dmg.ctypes.data if damage else None))
        return (out, off, pit, dmg) if damage else (out, off, pit)

    def make_frame(self, in_ptr: int, in_pitch: int, out_ptr: int, off, pit, session: int = 0,
                   damage_ptr: int | None = None) -> Frame:
        f = Frame()
        f.bgrx, f.in_pitch = in_ptr, in_pitch
        for i in range(3):
            f.planes[i] = (out_ptr + off[i]) if (i < 2 or self.cfg.out_format != FMT_NV12) else None
            f.pitches[i] = pit[i]
        f.damage, f.session = damage_ptr, session
        return f

    def convert_batch(self, frames) -> None:
        arr = (Frame * len(frames))(*frames)
        self._check(self.lib.b200vc_convert_batch(self.ctx, arr, len(frames)))

    def submit(self, frame: Frame) -> int:
        t = C.c_uint64()
        self._check(self.lib.b200vc_submit(self.ctx, C.byref(frame), C.byref(t)))
        return t.value

    def collect(self, ticket: int) -> None:
        self._check(self.lib.b200vc_collect(self.ctx, ticket))

    def convert_device(self, df: DeviceFrames, stream: int = 0) -> None:
        self._check(self.lib.b200vc_convert_device(self.ctx, C.byref(df), stream))

    # -- GstBaseTransform::stop --------------------------------------------------------------------
    def stop(self) -> None:
        if self.ctx:
            self.lib.b200vc_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.stop()
        except Exception:
            pass
