"""GPU parity tests of the wider caps surface: Y42B and Y444 output, RGB16/BGR16 input (what
ximagesrc delivers at the reference's CDEPTH=16, Dockerfile:206), and every chroma-site value — unscaled
(the conversion kernel) and scaled (the two-pass path), byte for byte against the oracle through the C ABI."""
import numpy as np
import pytest

import b200vc
from helpers import assert_planes_equal, plane_views

pytestmark = pytest.mark.gpu
GOLDEN_NEW = ["y42b_709l_mpeg2", "y444_601f", "y42b_scale_odd", "rgb16_709l_i420_mpeg2", "bgr16_601f_nv12_scale", "site_cosited_i420",
              "nearest_3to2_709l_i420", "nearest_up_odd_601f_nv12"]


@pytest.fixture(scope="module")
def conv():
    c = b200vc.Converter(0)
    yield c
    c.stop()


def _want(orc, frame, fmt, m, rg, site, inf, size=None):
    h, w = frame.shape[:2]
    ow, oh = size or (w, h)
    buf, off, pit = orc.convert(frame, fmt, m, rg, site, out_size=(ow, oh), in_format=inf)
    return plane_views(buf, off, pit, fmt, ow, oh)


@pytest.mark.parametrize("size", [(1920, 1080), (1366, 768), (131, 35), (126, 1082), (6, 4), (2, 2)])
@pytest.mark.parametrize("fmt", [2, 3])
def test_y42b_y444_parity(conv, orc, fmt, size):
    w, h = size
    frame = orc.synth("rand", w, h, 21)
    for m, rg in [(0, 0), (1, 1), (2, 0)]:
        for site in (0, 1):
            conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site)
            buf, off, pit = conv.transform_frame(frame)
            assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _want(orc, frame, fmt, m, rg, site, 0), f"{size} fmt{fmt} m{m} r{rg} s{site}")


@pytest.mark.parametrize("size", [(1920, 1080), (1366, 768), (133, 37), (2, 2)])
@pytest.mark.parametrize("inf", [4, 5])
def test_16_bit_input_parity(conv, orc, inf, size):
    w, h = size
    rng = np.random.default_rng(w + h + inf)
    frame = rng.integers(0, 256, (h, w, 2), dtype=np.uint8)
    for fmt in (0, 1, 2, 3):
        for m, rg, site in [(0, 0, 0), (1, 1, 1)]:
            conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site, in_format=inf)
            buf, off, pit = conv.transform_frame(frame)
            assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _want(orc, frame, fmt, m, rg, site, inf), f"{size} in{inf} fmt{fmt}")


def test_16_bit_damage_map(conv, orc):
    w, h = 1280, 720
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (h, w, 2), dtype=np.uint8)
    b = a.copy()
    b[100, 200, 0] ^= 1          # low byte of a pixel: blue/green bits
    b[719, 1279, 1] ^= 0x80      # the top bit of the last pixel
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, in_format=4, flags=b200vc.FLAG_DAMAGE_MAP)
    assert conv.transform_frame(a, damage=True)[3].all()
    buf, off, pit, dm = conv.transform_frame(b, damage=True)
    assert np.array_equal(dm, orc.damage_map(b, a, 4)) and dm.sum() == 2
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), _want(orc, b, 0, 1, 0, 1, 4))


@pytest.mark.parametrize("site", [2, 3, 7])
def test_chroma_sites_with_vertical_cositing_negotiate(conv, orc, site):
    """cosited (3), dv (7) and v-cosited alone (2): accepted; the output follows the site's h-cosited bit."""
    w, h = 322, 50
    frame = orc.synth("rand", w, h, 2)
    for fmt in (0, 1, 2):
        conv.set_info(w, h, out_format=fmt, matrix=1, range_=0, chroma_site=site)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _want(orc, frame, fmt, 1, 0, site, 0), f"site {site} fmt{fmt}")
        assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _want(orc, frame, fmt, 1, 0, site & 1, 0), f"site {site} == site {site & 1}")
    conv.set_info(640, 360, 320, 180, matrix=1, range_=0, chroma_site=site)      # and through the fused scaler
    frame = orc.synth("rand", 640, 360, 3)
    buf, off, pit = conv.transform_frame(frame)
    assert_planes_equal(plane_views(buf, off, pit, 0, 320, 180), _want(orc, frame, 0, 1, 0, site, 0, (320, 180)))


@pytest.mark.parametrize("geo", [(1920, 1080, 1280, 720), (1280, 720, 1920, 1080), (640, 360, 320, 180), (133, 37, 65, 20), (1920, 1080, 320, 180)])
def test_new_formats_scaled(conv, orc, geo):
    """Y42B/Y444 output and 16-bit input with scaling: convert at the source size, scale every plane. Y42B/Y444 take
    the two-pass path (the intermediate and the plane scaler follow the output format's chroma geometry); 5-6-5 input
    to I420/NV12 is unpacked inside the general fused kernel (the 6:1 geometry's window exceeds a TMA box: two-pass),
    and both ways are checked for it."""
    iw, ih, ow, oh = geo
    rng = np.random.default_rng(iw + oh)
    f4 = orc.synth("rand", iw, ih, 30)
    f2 = rng.integers(0, 256, (ih, iw, 2), dtype=np.uint8)
    for fmt, inf, frame in [(2, 0, f4), (3, 0, f4), (0, 4, f2), (1, 5, f2), (3, 4, f2), (2, 5, f2)]:
        want = _want(orc, frame, fmt, 1, 0, 1, inf, (ow, oh))
        for flags in ((0, b200vc.FLAG_FORCE_TWO_PASS) if fmt <= 1 else (0,)):
            conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=1, range_=0, chroma_site=1, in_format=inf, flags=flags)
            fused = fmt <= 1 and not flags and iw < 6 * ow
            assert conv.info().path == (1 if fused else 2) and conv.info().fused_2to1 == 0
            buf, off, pit = conv.transform_frame(frame)
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want, f"{geo} fmt{fmt} in{inf} flags{flags}")


def test_new_golden_vectors_on_gpu(conv):
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "vectors.npz"))
    for name in GOLDEN_NEW:
        iw, ih, ow, oh, fmt, mat, rg, site, mode, inf = [int(v) for v in z[name + "/cfg"]]
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=mat, range_=rg, chroma_site=site, scale_mode=mode, in_format=inf)
        buf, off, pit = conv.transform_frame(np.ascontiguousarray(z[name + "/in"]))
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), [z[f"{name}/p{i}"] for i in range(2 if fmt == 1 else 3)], name)


def test_device_entry_point_and_batches_with_new_formats(conv, orc):
    import torch
    w, h, n = 1280, 720, 3
    rng = np.random.default_rng(9)
    host = rng.integers(0, 256, (n, h, w, 2), dtype=np.uint8)
    total, off, pit = b200vc.default_layout(3, w, h)
    conv.set_info(w, h, out_format=3, matrix=0, range_=1, chroma_site=0, in_format=5, max_sessions=n, n_inflight=n)
    d_in = torch.from_numpy(host.reshape(n, -1)).cuda()
    d_out = torch.zeros((n, total), dtype=torch.uint8, device="cuda")
    df = b200vc.DeviceFrames()
    df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), w * 2, w * h * 2
    for i in range(3):
        df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
    df.out_frame_stride, df.n_frames = total, n
    conv.convert_device(df, 0)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy()
    for k in range(n):
        assert_planes_equal(plane_views(got[k], off, pit, 3, w, h), _want(orc, host[k], 3, 0, 1, 0, 5), f"device frame {k}")
    # the same frames as a batch of sessions through pinned host buffers
    pin_in, pin_out = b200vc.PinnedBuffer(n * w * h * 2), b200vc.PinnedBuffer(n * total)
    pin_in.array[:] = host.reshape(-1)
    conv.convert_batch([conv.make_frame(pin_in.ptr + k * w * h * 2, w * 2, pin_out.ptr + k * total, off, pit, session=k) for k in range(n)])
    for k in range(n):
        assert_planes_equal(plane_views(pin_out.array[k * total:(k + 1) * total], off, pit, 3, w, h), _want(orc, host[k], 3, 0, 1, 0, 5), f"batch frame {k}")
    assert conv.stats().d2h_bytes == n * total
    pin_in.free()
    pin_out.free()


@pytest.mark.parametrize("geo", [(1920, 1080, 1280, 1024), (1024, 768, 1920, 1080), (1280, 720, 1280, 1024), (3840, 2160, 1920, 1200),
                                 (640, 360, 322, 200)])
def test_add_borders(conv, orc, geo):
    """videoscale add-borders: the picture keeps its display aspect ratio inside the output frame (b200vc_fit_borders), the
    rest is black in the output's colorimetry. Unscaled-in-a-bigger-frame, fused and two-pass paths, all four output formats;
    synchronous call with pageable and with page-locked buffers, and the device entry point."""
    import torch
    iw, ih, ow, oh = geo
    content = b200vc.fit_borders(iw, ih, ow, oh)
    assert content != (0, 0, ow, oh)
    frame = orc.synth("rand", iw, ih, 44)
    for fmt, m, rg, site, flags in [(0, 1, 0, 1, 0), (1, 0, 1, 0, 0), (2, 1, 0, 1, 0), (3, 2, 1, 0, 0), (0, 0, 0, 1, b200vc.FLAG_FORCE_TWO_PASS)]:
        scaled = (iw, ih) != content[2:]
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, flags=flags if scaled else 0, content=content)
        obuf, ooff, opit = orc.convert(frame, fmt, m, rg, site, out_size=(ow, oh), content=content)
        want = plane_views(obuf, ooff, opit, fmt, ow, oh)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want, f"{geo} fmt{fmt} pageable")
        total = b200vc.default_layout(fmt, ow, oh)[0]
        pin_in, pin_out = b200vc.PinnedBuffer(iw * ih * 4), b200vc.PinnedBuffer(total)
        pin_in.array[:] = frame.reshape(-1)
        pin_out.array[:] = 0x5A                               # a recycled buffer: the border must be painted
        buf, off, pit = conv.transform_frame(pin_in.array.reshape(ih, iw, 4), out=pin_out.array)
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want, f"{geo} fmt{fmt} pinned")
        if all(p % 16 == 0 for p in pit[:2 if fmt == 1 else 3]) and all(o % 16 == 0 for o in off[:2 if fmt == 1 else 3]) and total % 16 == 0:
            d_in = torch.from_numpy(np.stack([frame.reshape(-1)] * 2)).cuda()
            d_out = torch.full((2, total), 0x5A, dtype=torch.uint8, device="cuda")
            df = b200vc.DeviceFrames()
            df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * 4, iw * ih * 4
            for i in range(2 if fmt == 1 else 3):
                df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
            df.out_frame_stride, df.n_frames = total, 2
            conv.convert_device(df, 0)
            torch.cuda.synchronize()
            got = d_out.cpu().numpy()
            for k in range(2):
                assert_planes_equal(plane_views(got[k], off, pit, fmt, ow, oh), want, f"{geo} fmt{fmt} device frame {k}")
        pin_in.free()
        pin_out.free()
    # the border is black: Y 16 (limited range), chroma 128
    conv.set_info(iw, ih, ow, oh, matrix=1, range_=0, chroma_site=1, content=content)
    buf, off, pit = conv.transform_frame(frame)
    y, u, v = plane_views(buf, off, pit, 0, ow, oh)
    cx, cy, cw, ch = content
    mask = np.ones((oh, ow), bool)
    mask[cy:cy + ch, cx:cx + cw] = False
    assert (y[mask] == 16).all() and (u[0 if cy else slice(None), 0] == 128).all()


@pytest.mark.parametrize("geo", [(3840, 2160, 1920, 1080), (1920, 1080, 1280, 720), (1280, 720, 1920, 1080), (131, 35, 65, 17), (1920, 1080, 320, 180),
                                 (640, 360, 641, 361)])
def test_method_nearest(conv, orc, geo):
    """videoscale method=nearest: the same kernels with all the weight on one tap — fused (the general kernel, also at
    exactly 2:1), two-pass and every output format."""
    iw, ih, ow, oh = geo
    frame = orc.synth("rand", iw, ih, 51)
    for fmt, flags in [(0, 0), (1, 0), (0, b200vc.FLAG_FORCE_TWO_PASS), (2, 0), (3, 0)]:
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=1, range_=0, chroma_site=1, scale_mode=b200vc.SCALE_NEAREST, flags=flags)
        assert conv.info().fused_2to1 == 0
        buf, off, pit = conv.transform_frame(frame)
        obuf, ooff, opit = orc.convert(frame, fmt, 1, 0, 1, out_size=(ow, oh), scale_mode=2)
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), plane_views(obuf, ooff, opit, fmt, ow, oh), f"{geo} fmt{fmt} flags{flags}")
