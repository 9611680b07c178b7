"""CPU tests of the C-ABI boundary: the library loads, exports what include/b200vc.h declares,
negotiates like `videoconvert ! videoscale` for these formats, and refuses to work without a GPU.
No compute entry point is called here."""
import ctypes as C
import os
import re

import pytest

import b200vc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_and_library_agree(lib):
    hdr = open(os.path.join(ROOT, "include", "b200vc.h")).read()
    declared = sorted(set(re.findall(r"\b(b200vc_[a-z_0-9]+)\s*\(", hdr)))
    assert declared == sorted(b200vc.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), f"libb200vc.so does not export {name}"


def test_struct_sizes_match_the_header():
    # the header's structs, laid out by hand (LP64): catches a drifting ctypes mirror
    assert C.sizeof(b200vc.Config) == 19 * 4
    assert C.sizeof(b200vc.Frame) == 8 + 8 + 24 + 24 + 8 + 4 + 4
    assert C.sizeof(b200vc.DeviceFrames) == 3 * 8 + 3 * 8 + 24 + 24 + 8 + 8 + 8 + 4 + 4
    assert C.sizeof(b200vc.Stats) == 6 * 8
    assert C.sizeof(b200vc.Info) == 10 * 4 + 64


def test_version_and_status_strings(lib):
    assert lib.b200vc_version() == 2
    assert lib.b200vc_status_string(0) == b"ok"
    assert b"sm_100" in lib.b200vc_status_string(b200vc.ERR_NO_DEVICE)


def test_default_colorimetry_by_height():
    # >= 2160 lines -> bt2020, > 576 lines -> bt709 limited + mpeg2 siting; else bt601 limited + jpeg
    assert b200vc.default_colorimetry(1080) == (b200vc.MATRIX_BT709, b200vc.RANGE_LIMITED, b200vc.SITE_H_COSITED)
    assert b200vc.default_colorimetry(2160) == (b200vc.MATRIX_BT2020, b200vc.RANGE_LIMITED, b200vc.SITE_H_COSITED)
    assert b200vc.default_colorimetry(2159)[0] == b200vc.MATRIX_BT709
    assert b200vc.default_colorimetry(577)[0] == b200vc.MATRIX_BT709
    assert b200vc.default_colorimetry(576) == (b200vc.MATRIX_BT601, b200vc.RANGE_LIMITED, b200vc.SITE_NONE)
    assert b200vc.default_colorimetry(480)[2] == b200vc.SITE_NONE


@pytest.mark.parametrize("fmt,w,h", [(0, 1920, 1080), (1, 2560, 1440), (0, 1366, 768), (1, 1366, 768), (0, 2, 2),
                                     (2, 1920, 1080), (3, 1920, 1080), (2, 1366, 767), (3, 1365, 767)])
def test_default_layout_matches_oracle(orc, fmt, w, h):
    assert b200vc.default_layout(fmt, w, h) == orc.default_layout(fmt, w, h)


def _cfg(**kw):
    base = dict(device=0, in_w=1920, in_h=1080, out_w=1920, out_h=1080, out_format=0, matrix=0, range=0,
                chroma_site=0, scale_mode=0, flags=0, max_sessions=1, n_inflight=0, n_bands=0, in_format=0,
                content_x=0, content_y=0, content_w=0, content_h=0)
    base.update(kw)
    return b200vc.Config(**base)


def test_caps_acceptance(lib):
    ok = [_cfg(), _cfg(out_format=1, matrix=1, range=1, chroma_site=1, in_w=2560, in_h=1440, out_w=2560, out_h=1440),
          _cfg(in_w=3840, in_h=2160), _cfg(in_w=3840, in_h=2160, scale_mode=1), _cfg(flags=b200vc.FLAG_DAMAGE_MAP),
          _cfg(in_w=1366, in_h=768, out_w=1366, out_h=768), _cfg(max_sessions=8, n_inflight=4), _cfg(in_format=3),
          _cfg(in_w=1365, in_h=767, out_w=1365, out_h=767), _cfg(in_w=3841, in_h=2161),
          _cfg(chroma_site=3), _cfg(chroma_site=7), _cfg(chroma_site=2),      # cosited, dv, v-cosited: all negotiate
          _cfg(out_format=2), _cfg(out_format=3, in_w=3840, in_h=2160),       # Y42B, Y444 (scaled: two-pass)
          _cfg(in_format=4), _cfg(in_format=5, flags=b200vc.FLAG_DAMAGE_MAP)]  # RGB16, BGR16 (CDEPTH=16)
    for c in ok:
        assert lib.b200vc_check_config(C.byref(c)) == b200vc.OK
    refused = [(_cfg(in_w=3842, in_h=2162, out_w=1921, out_h=1081, scale_mode=1), b200vc.ERR_UNSUPPORTED),   # BOX, odd
               (_cfg(out_format=4), b200vc.ERR_UNSUPPORTED),               # not I420/NV12/Y42B/Y444
               (_cfg(in_format=6), b200vc.ERR_UNSUPPORTED),                # not one of the six sink formats
               (_cfg(matrix=3), b200vc.ERR_UNSUPPORTED),                   # not bt601/bt709/bt2020
               (_cfg(chroma_site=8), b200vc.ERR_UNSUPPORTED),              # not a combination of the site flags
               (_cfg(out_format=3, flags=b200vc.FLAG_DAMAGE_MAP), b200vc.ERR_UNSUPPORTED),   # the map goes with 4:2:0 output
               (_cfg(in_w=3840, in_h=2160, out_w=1280, out_h=720, scale_mode=1), b200vc.ERR_UNSUPPORTED),
               (_cfg(in_w=3840, in_h=2160, flags=b200vc.FLAG_DAMAGE_MAP), b200vc.ERR_UNSUPPORTED),
               (_cfg(in_w=0), b200vc.ERR_INVALID_ARG), (_cfg(max_sessions=-1), b200vc.ERR_INVALID_ARG),
               (_cfg(in_w=32768, out_w=32768), b200vc.ERR_UNSUPPORTED)]
    for c, want in refused:
        assert lib.b200vc_check_config(C.byref(c)) == want
        assert lib.b200vc_last_error(None)          # a reason is always given
    assert lib.b200vc_check_config(None) == b200vc.ERR_INVALID_ARG


def test_fit_borders_like_videoscale_add_borders():
    # 16:9 into 5:4: bars above and below
    assert b200vc.fit_borders(1920, 1080, 1280, 1024) == (0, 152, 1280, 720)
    # 4:3 into 16:9: bars left and right
    assert b200vc.fit_borders(1024, 768, 1920, 1080) == (240, 0, 1440, 1080)
    # same shape (any scale): the whole frame
    assert b200vc.fit_borders(3840, 2160, 1920, 1080) == (0, 0, 1920, 1080)
    # anamorphic source (PAR 4:3 on 1440x1080 = 16:9 display) into square-pixel 1280x720: fills it
    assert b200vc.fit_borders(1440, 1080, 1280, 720, in_par=(4, 3)) == (0, 0, 1280, 720)
    # and into a 4:3 frame of square pixels: letterboxed; sizes and positions come out even
    x, y, w, h = b200vc.fit_borders(1440, 1080, 1024, 768, in_par=(4, 3))
    assert (x, w) == (0, 1024) and h == 576 and y == 96
    x, y, w, h = b200vc.fit_borders(1366, 768, 1001, 700)
    assert x % 2 == 0 and y % 2 == 0 and w % 2 == 0 and h % 2 == 0 and x + w <= 1001 and y + h <= 700


def test_content_rectangle_checks(lib):
    ok = _cfg(in_w=1920, in_h=1080, out_w=1280, out_h=1024, content_x=0, content_y=152, content_w=1280, content_h=720)
    assert lib.b200vc_check_config(C.byref(ok)) == b200vc.OK
    # unscaled picture inside a bigger frame, with the damage map: the scaler is not involved
    ok = _cfg(in_w=1280, in_h=720, out_w=1280, out_h=1024, content_y=152, content_w=1280, content_h=720, flags=b200vc.FLAG_DAMAGE_MAP)
    assert lib.b200vc_check_config(C.byref(ok)) == b200vc.OK
    for bad, want in [(_cfg(out_w=1280, out_h=1024, content_y=400, content_w=1280, content_h=720), b200vc.ERR_INVALID_ARG),   # sticks out
                      (_cfg(out_w=1280, out_h=1024, content_x=1, content_w=1278, content_h=720), b200vc.ERR_UNSUPPORTED),     # odd position
                      (_cfg(out_w=1280, out_h=1024, content_w=1280, content_h=0), b200vc.ERR_INVALID_ARG)]:
        assert lib.b200vc_check_config(C.byref(bad)) == want


def test_calls_before_negotiation_fail(lib):
    assert lib.b200vc_convert(None, None, 0, None, None, None) == b200vc.ERR_NOT_NEGOTIATED
    assert lib.b200vc_convert_batch(None, None, 0) == b200vc.ERR_NOT_NEGOTIATED
    assert lib.b200vc_submit(None, None, None) == b200vc.ERR_NOT_NEGOTIATED
    assert lib.b200vc_convert_device(None, None, None) == b200vc.ERR_NOT_NEGOTIATED
    assert lib.b200vc_destroy(None) == b200vc.ERR_INVALID_ARG
    with pytest.raises(b200vc.B200vcError):
        b200vc.Converter().transform_frame(None)


@pytest.mark.skipif(_has_gpu(), reason="this box has a GPU")
def test_no_gpu_means_loud_failure_not_fallback(lib):
    ctx = C.c_void_p()
    c = _cfg()
    assert lib.b200vc_create(C.byref(c), C.byref(ctx)) == b200vc.ERR_NO_DEVICE
    assert not ctx.value
    assert b"no CPU fallback" in lib.b200vc_last_error(None)
    with pytest.raises(b200vc.B200vcError) as e:
        b200vc.Converter().set_info(1920, 1080)
    assert e.value.status == b200vc.ERR_NO_DEVICE


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "docker-nvidia-glx-desktop_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "oracle.h" not in src and "liboracle" not in src, f


def test_damage_map_to_rectangles(orc):
    """The damage-map consumer: rectangles cover exactly the dirty macroblocks, clipped to the frame."""
    import numpy as np
    w, h = 1920, 1080
    a, b = orc.synth("desk", w, h, 29), orc.synth("desk", w, h, 30)
    dm = orc.damage_map(b, a)
    rects, n = b200vc.damage_rects(dm, w, h)
    assert n == len(rects) and 0 < n < 40
    cover = np.zeros_like(dm)
    for x, y, rw, rh in rects:
        assert x % 16 == 0 and y % 16 == 0 and rw > 0 and rh > 0 and x + rw <= w and y + rh <= h
        sub = cover[y // 16:(y + rh + 15) // 16, x // 16:(x + rw + 15) // 16]
        assert not sub.any()                             # rectangles do not overlap
        sub[:] = 1
    assert np.array_equal(cover, dm)
    # the last macroblock row of 1080p is 8 lines high
    full = np.ones_like(dm)
    assert b200vc.damage_rects(full, w, h)[0] == [(0, 0, 1920, 1080)]
    assert b200vc.damage_rects(np.zeros_like(dm), w, h) == ([], 0)
    lone = np.zeros_like(dm)
    lone[67, 119] = 1
    assert b200vc.damage_rects(lone, w, h)[0] == [(1904, 1072, 16, 8)]
    # more rectangles than room: the count is still reported
    chk = (np.indices(dm.shape).sum(0) & 1).astype(np.uint8)
    got, need = b200vc.damage_rects(chk, w, h, max_rects=10)
    assert len(got) == 10 and need == int(chk.sum())


def test_gstreamer_shim_parses_against_stub_headers():
    """GStreamer is not installed here, so gst/gstb200convert.c has never met the real headers. This runs
    it through gcc's front end against tests/gst_stub (a hand-written stand-in for the slice of the 1.x API
    the shim uses) with implicit declarations and pointer mismatches as errors: it pins the shim's own C and
    its calls into include/b200vc.h, not its conformance to GStreamer."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([gcc, "-std=gnu11", "-fsyntax-only", "-Wall", "-Werror=implicit-function-declaration",
                        "-Werror=incompatible-pointer-types", "-Werror=int-conversion",
                        "-I" + os.path.join(root, "tests", "gst_stub"), "-I" + os.path.join(root, "include"),
                        os.path.join(root, "docker-nvidia-glx-desktop_b200", "gst", "gstb200convert.c")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
