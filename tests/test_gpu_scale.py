"""GPU parity tests, K2: conversion with the bilinear downscale fused in (TMA-staged tiles), and the
general two-pass scaler, against the oracle's `videoconvert ! videoscale` restatement. Bit-exact."""
import numpy as np
import pytest

import b200vc
from helpers import assert_planes_equal, plane_checksums, plane_views

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def conv():
    c = b200vc.Converter(0)
    yield c
    c.stop()


def _want(orc, frame, ow, oh, fmt, m, rg, site, mode=0):
    buf, off, pit = orc.convert(frame, fmt, m, rg, site, out_size=(ow, oh), scale_mode=mode)
    return plane_views(buf, off, pit, fmt, ow, oh)


GEOS = [(3840, 2160, 1920, 1080), (64, 36, 32, 18), (128, 64, 64, 32), (260, 132, 130, 66), (4, 4, 2, 2), (96, 48, 64, 32),
        (100, 60, 46, 26), (1920, 1080, 1280, 720), (1280, 720, 854, 480), (200, 120, 150, 90), (640, 360, 212, 120)]


@pytest.mark.parametrize("geo", GEOS)
@pytest.mark.parametrize("variant", ["fused", "fused_generic", "two_pass"])
def test_scale_parity(conv, orc, geo, variant):
    """Three implementations of the same two-stage semantics: the exact-2:1 register kernel (where
    the taps allow it), the general fused kernel, and convert-then-scale through an intermediate."""
    iw, ih, ow, oh = geo
    two_pass = variant == "two_pass"
    frame = orc.synth("rand", iw, ih, 9)
    flags = {"fused": 0, "fused_generic": b200vc.FLAG_FORCE_GENERIC_FUSED, "two_pass": b200vc.FLAG_FORCE_TWO_PASS}[variant]
    for fmt, m, rg, site in [(0, 1, 0, 1), (1, 0, 1, 0), (0, 0, 0, 0), (1, 1, 1, 1)]:
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, flags=flags)
        inf = conv.info()
        assert inf.path == (2 if two_pass else 1)
        assert inf.fused_2to1 == (1 if variant == "fused" and iw == 2 * ow and ih == 2 * oh and iw % 4 == 0 and ih % 4 == 0 else 0)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), _want(orc, frame, ow, oh, fmt, m, rg, site),
                            f"{geo} fmt{fmt} m{m} r{rg} s{site} two_pass={two_pass}")


@pytest.mark.parametrize("flags", [0, b200vc.FLAG_FORCE_GENERIC_FUSED, b200vc.FLAG_FORCE_TWO_PASS])
def test_box_mode(conv, orc, flags):
    iw, ih, ow, oh = 516, 260, 258, 130
    frame = orc.synth("rand", iw, ih, 2)
    for fmt in (0, 1):
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=1, range_=0, chroma_site=1, scale_mode=1, flags=flags)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), _want(orc, frame, ow, oh, fmt, 1, 0, 1, 1))


def test_structured_frames_4k(conv, orc):
    # C3 on bars/ramp/checkerboard: the ramp exercises every horizontal fraction 0..247
    frame = orc.synth("desk", 3840, 2160, 3)
    conv.set_info(3840, 2160, 1920, 1080, matrix=1, range_=0, chroma_site=1)
    inf = conv.info()
    assert inf.path == 1 and inf.fused_2to1 == 1 and inf.box_w >= 130 and inf.box_h >= 32
    buf, off, pit = conv.transform_frame(frame)
    want = _want(orc, frame, 1920, 1080, 0, 1, 0, 1)
    assert_planes_equal(plane_views(buf, off, pit, 0, 1920, 1080), want)
    assert plane_checksums(plane_views(buf, off, pit, 0, 1920, 1080)) == plane_checksums(want)


def test_large_ratio_falls_back_to_two_pass(conv, orc):
    # 6:1 needs a source window wider than one TMA box: negotiated onto the general path
    frame = orc.synth("rand", 1920, 1080, 4)
    conv.set_info(1920, 1080, 320, 180, matrix=0, range_=0, chroma_site=0)
    assert conv.info().path == 2
    buf, off, pit = conv.transform_frame(frame)
    assert_planes_equal(plane_views(buf, off, pit, 0, 320, 180), _want(orc, frame, 320, 180, 0, 0, 0, 0))


def test_upscale_uses_the_same_taps(conv, orc):
    frame = orc.synth("rand", 160, 90, 4)
    for flags in (0, b200vc.FLAG_FORCE_TWO_PASS):
        conv.set_info(160, 90, 320, 180, matrix=0, range_=0, chroma_site=1, flags=flags)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, 0, 320, 180), _want(orc, frame, 320, 180, 0, 0, 0, 1))


def test_device_entry_point_scaled_batch(conv, orc):
    import torch
    iw, ih, ow, oh, n = 1280, 720, 640, 360, 5
    total, off, pit = b200vc.default_layout(0, ow, oh)
    host = np.stack([orc.synth("rand", iw, ih, 60 + i) for i in range(n)])
    d_in = torch.from_numpy(host.reshape(n, -1)).cuda()
    for flags in (0, b200vc.FLAG_FORCE_TWO_PASS):
        d_out = torch.zeros((n, total), dtype=torch.uint8, device="cuda")
        conv.set_info(iw, ih, ow, oh, matrix=1, range_=0, chroma_site=1, flags=flags, max_sessions=2)
        df = b200vc.DeviceFrames()
        df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * 4, iw * ih * 4
        for i in range(3):
            df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
        df.out_frame_stride, df.n_frames = total, n
        conv.convert_device(df, 0)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        for i in range(n):
            assert_planes_equal(plane_views(got[i], off, pit, 0, ow, oh), _want(orc, host[i], ow, oh, 0, 1, 0, 1),
                                f"frame {i} flags {flags}")


@pytest.mark.parametrize("geo", [(7, 9, 4, 5), (131, 35, 65, 17), (1365, 767, 683, 384), (100, 61, 47, 26), (64, 36, 33, 19),
                                 (1919, 1079, 1280, 720), (1920, 1080, 1281, 721)])
@pytest.mark.parametrize("two_pass", [False, True])
def test_odd_sizes_with_scaling(conv, orc, geo, two_pass):
    """Odd widths and heights: chroma planes round up, the last pixel / line is a chroma site of its own.
    The general fused kernel takes them (it repeats the last column / line inside its window) and so does
    the two-pass path."""
    iw, ih, ow, oh = geo
    frame = orc.synth("rand", iw, ih, 17)
    for fmt, m, rg, site in [(0, 1, 0, 1), (1, 0, 1, 0)]:
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site,
                      flags=b200vc.FLAG_FORCE_TWO_PASS if two_pass else 0)
        assert conv.info().path == (2 if two_pass else 1)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), _want(orc, frame, ow, oh, fmt, m, rg, site),
                            f"{geo} fmt{fmt} two_pass={two_pass}")


@pytest.mark.parametrize("geo,flags", [((2560, 1440, 1920, 1080), 0),
                                       ((3840, 2160, 1920, 1080), b200vc.FLAG_FORCE_GENERIC_FUSED),
                                       ((3840, 2160, 1280, 720), 0),
                                       ((1280, 720, 1920, 1080), 0)])
def test_general_kernel_tile_heights(conv, orc, geo, flags, monkeypatch):
    """The general fused kernel is instantiated for tiles of 8 to 64 rows; the library picks the cheapest
    for the geometry. The pick and every forced height (B200VC_TILE_H, read by set_info; a height whose
    window does not fit falls back to the pick) must give the oracle's frame (bars: every fraction occurs)."""
    iw, ih, ow, oh = geo
    frame = orc.synth("bars", iw, ih, 4)
    want = {}
    for force in (None, 8, 16, 32, 48, 64):
        if force is None:
            monkeypatch.delenv("B200VC_TILE_H", raising=False)
        else:
            monkeypatch.setenv("B200VC_TILE_H", str(force))
        for fmt, site in [(0, 1), (1, 0)]:
            conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=1, range_=0, chroma_site=site, flags=flags)
            inf = conv.info()
            assert inf.path == 1 and inf.fused_2to1 == 0 and inf.tile_h in (8, 16, 32, 48, 64)
            if force is not None and inf.tile_h != force:
                continue                                    # (a window too large for that height: the library fell back)
            if (fmt, site) not in want:
                want[(fmt, site)] = _want(orc, frame, ow, oh, fmt, 1, 0, site)
            buf, off, pit = conv.transform_frame(frame)
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want[(fmt, site)], f"{geo} fmt{fmt} tile_h {inf.tile_h}")
    monkeypatch.delenv("B200VC_TILE_H", raising=False)


@pytest.mark.parametrize("geo", [(1920, 1080, 1312, 730), (2560, 1440, 1920, 1080), (1280, 720, 1920, 1080)])
def test_general_kernel_tile_stores(conv, orc, geo, monkeypatch):
    """The general fused kernel hands its output tile to the TMA unit when the planes lie in device memory and
    every pitch is a multiple of 16, and stores it itself otherwise (B200VC_NO_TMA_STORE, read by set_info,
    forces that). Both must give the oracle's frame, tiles overhanging the right and bottom edges included
    (1312x730: 20.5 tile columns, rows that end inside a tile; all rows multiples of 16 bytes — the unit clips
    columns in units of 16 bytes, so other widths keep the kernel's own stores; the add-borders cases of
    test_gpu_formats.py and the fuzz run are what would show a store reaching past the picture)."""
    iw, ih, ow, oh = geo
    frame = orc.synth("bars", iw, ih, 5)
    for fmt, site in [(0, 1), (1, 0)]:
        want = _want(orc, frame, ow, oh, fmt, 1, 0, site)
        for off_switch in (False, True):
            if off_switch:
                monkeypatch.setenv("B200VC_NO_TMA_STORE", "1")
            else:
                monkeypatch.delenv("B200VC_NO_TMA_STORE", raising=False)
            conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=1, range_=0, chroma_site=site)
            assert conv.info().path == 1
            buf, off, pit = conv.transform_frame(frame)
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want, f"{geo} fmt{fmt} tma stores off={off_switch}")
    monkeypatch.delenv("B200VC_NO_TMA_STORE", raising=False)


@pytest.mark.parametrize("fmt", [0, 1])
def test_general_kernel_device_batch(conv, orc, fmt, monkeypatch):
    """Several frames in one launch of the general fused kernel on the device entry point, planes in GstVideoInfo's
    default layout (rows and pitches multiples of 16: the tiles leave by TMA stores through 3-D maps whose third
    dimension is the frame), and the same with the kernel's own stores."""
    import torch
    iw, ih, ow, oh, n = 1920, 1080, 1280, 720, 3
    total, off, pit = b200vc.default_layout(fmt, ow, oh)
    host = np.stack([orc.synth("rand", iw, ih, 80 + i) for i in range(n)])
    want = [_want(orc, host[i], ow, oh, fmt, 1, 0, 1) for i in range(n)]
    d_in = torch.from_numpy(host.reshape(n, -1)).cuda()
    for off_switch in (False, True):
        if off_switch:
            monkeypatch.setenv("B200VC_NO_TMA_STORE", "1")
        else:
            monkeypatch.delenv("B200VC_NO_TMA_STORE", raising=False)
        d_out = torch.full((n, total), 0x5A, dtype=torch.uint8, device="cuda")
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=1, range_=0, chroma_site=1, max_sessions=2)
        inf = conv.info()
        assert inf.path == 1 and inf.fused_2to1 == 0
        df = b200vc.DeviceFrames()
        df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * 4, iw * ih * 4
        for i in range(2 if fmt == 1 else 3):
            df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
        df.out_frame_stride, df.n_frames = total, n
        conv.convert_device(df, 0)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        for i in range(n):
            assert_planes_equal(plane_views(got[i], off, pit, fmt, ow, oh), want[i], f"frame {i} fmt{fmt} tma stores off={off_switch}")
    monkeypatch.delenv("B200VC_NO_TMA_STORE", raising=False)
