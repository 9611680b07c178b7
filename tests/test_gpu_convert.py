"""GPU parity tests, K1 (BGRx -> I420/NV12) and K3 (damage map): the CUDA path, called through the
C ABI exactly as the GStreamer element would call it, against the CPU oracle on the same frames.
Bar: bit-exact. (Oracle = this repo's restatement of videoconvert; GStreamer itself is not
available on any machine here — SURVEY.md §0.4.)"""
import ctypes as C
import os

import numpy as np
import pytest

import b200vc
from helpers import assert_planes_equal, plane_checksums, plane_views

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "vectors.npz")
COMBOS = [(0, 0), (0, 1), (1, 0), (1, 1), (2, 0), (2, 1)]


@pytest.fixture(scope="module")
def conv():
    c = b200vc.Converter(0)
    yield c
    c.stop()


def _oracle_planes(orc, frame, fmt, m, rg, site, **kw):
    buf, off, pit = orc.convert(frame, fmt, m, rg, site, **kw)
    ow, oh = kw.get("out_size", (frame.shape[1], frame.shape[0]))
    return plane_views(buf, off, pit, fmt, ow, oh)


def test_native_library_is_what_runs(conv):
    conv.set_info(64, 32)
    inf = conv.info()
    assert inf.cc_major == 10 and inf.sm_count >= 100 and b"B200" in inf.device_name
    maps = open("/proc/self/maps").read()
    assert "libb200vc.so" in maps


@pytest.mark.parametrize("size", [(1920, 1080), (2560, 1440), (2, 2), (4, 2), (6, 4), (130, 18), (258, 34),
                                  (1366, 768), (126, 1082)])
def test_convert_parity_all_caps(conv, orc, size):
    """Every matrix x range x chroma-site x format on random frames, including ragged sizes:
    widths that are not multiples of 4, 16 or 128 and heights that end inside a macroblock row."""
    w, h = size
    frame = orc.synth("rand", w, h, 3)
    for fmt in (0, 1):
        for m, rg in COMBOS:
            for site in (0, 1):
                conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site)
                buf, off, pit = conv.transform_frame(frame)
                assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _oracle_planes(orc, frame, fmt, m, rg, site),
                                    f"{size} fmt{fmt} m{m} r{rg} s{site}")


@pytest.mark.parametrize("kind", ["bars", "desk"])
def test_convert_parity_structured_frames(conv, orc, kind):
    # saturated bars, ramp and 1-px checkerboard: stresses truncation and both chroma filters
    w, h = 1920, 1080
    frame = orc.synth(kind, w, h, 12)
    for fmt, m, rg, site in [(0, 0, 0, 0), (0, 1, 0, 1), (1, 1, 1, 1), (1, 0, 1, 0)]:
        conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site)
        buf, off, pit = conv.transform_frame(frame)
        assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _oracle_planes(orc, frame, fmt, m, rg, site), kind)


def test_known_answers_on_gpu(conv):
    # SURVEY A.6 canaries: white under BT.709 is (235,127,128) limited and (254,128,128) full
    frame = np.full((16, 32, 4), 255, np.uint8)
    for (m, rg), (y, u, v) in {(1, 0): (235, 127, 128), (1, 1): (254, 128, 128), (0, 0): (235, 128, 128)}.items():
        conv.set_info(32, 16, matrix=m, range_=rg, chroma_site=1)
        buf, off, pit = conv.transform_frame(frame)
        pl = plane_views(buf, off, pit, 0, 32, 16)
        assert (pl[0] == y).all() and (pl[1] == u).all() and (pl[2] == v).all()


def test_golden_vectors_on_gpu(conv):
    z = np.load(GOLDEN)
    names = sorted({k.split("/")[0] for k in z.files if k.endswith("/cfg")})
    for name in names:
        iw, ih, ow, oh, fmt, mat, rg, site, mode, inf = [int(v) for v in z[name + "/cfg"]]
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=mat, range_=rg, chroma_site=site, scale_mode=mode,
                      in_format=inf)
        buf, off, pit = conv.transform_frame(np.ascontiguousarray(z[name + "/in"]))
        want = [z[f"{name}/p{i}"] for i in range(2 if fmt == 1 else 3)]
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want, name)
    conv.set_info(70, 40, flags=b200vc.FLAG_DAMAGE_MAP)
    conv.transform_frame(np.ascontiguousarray(z["damage/a"]), damage=True)
    dm = conv.transform_frame(np.ascontiguousarray(z["damage/b"]), damage=True)[3]
    assert np.array_equal(dm, z["damage/map"])


def test_pitches_are_honoured(conv, orc):
    # padded input pitch (XShm-like) and padded output pitches (GstVideoMeta)
    w, h = 322, 50
    frame = orc.synth("rand", w, h, 1, pitch=w * 4 + 64)
    conv.set_info(w, h, out_format=0, matrix=1, range_=0, chroma_site=1)
    buf, off, pit = conv.transform_frame(frame, pitches=[384, 200, 176])
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), _oracle_planes(orc, frame, 0, 1, 0, 1))
    # bytes between rows must not be touched
    y = np.ndarray((h, 384), np.uint8, buffer=buf, offset=off[0], strides=(384, 1))
    assert (y[:, w:] == 0).all()
    conv.set_info(w, h, out_format=1, matrix=0, range_=1, chroma_site=0)
    buf, off, pit = conv.transform_frame(frame, pitches=[336, 400])
    assert_planes_equal(plane_views(buf, off, pit, 1, w, h), _oracle_planes(orc, frame, 1, 0, 1, 0))


@pytest.mark.parametrize("bands", [1, 3, 4, 64])
def test_band_pipelining_does_not_change_results(conv, orc, bands):
    w, h = 640, 360
    frame = orc.synth("desk", w, h, 4)
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, n_bands=bands)
    buf, off, pit = conv.transform_frame(frame)
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), _oracle_planes(orc, frame, 0, 1, 0, 1), f"bands={bands}")


def test_pinned_and_registered_host_buffers(conv, orc):
    w, h = 1280, 720
    total, off, pit = b200vc.default_layout(0, w, h)
    pin_in, pin_out = b200vc.PinnedBuffer(w * h * 4), b200vc.PinnedBuffer(total)
    frame = orc.synth("rand", w, h, 8, out=pin_in.array.reshape(h, w * 4))
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=0)
    pin_out.array[:] = 0
    buf, off, pit = conv.transform_frame(frame, out=pin_out.array)
    want = _oracle_planes(orc, frame, 0, 0, 0, 0)
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), want, "pinned")
    st = conv.stats()
    assert st.staged_in_bytes == 0 and st.staged_out_bytes == 0 and st.h2d_bytes == w * h * 4 and st.d2h_bytes == total
    # pageable buffers: staged through the pinned ring...
    pageable = np.array(frame)
    buf, off, pit = conv.transform_frame(pageable)
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), want, "staged")
    assert conv.stats().staged_in_bytes == w * h * 4
    # ...or page-locked on first sight
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=0, flags=b200vc.FLAG_REGISTER_HOST)
    out = np.zeros(total, np.uint8)
    for _ in range(2):
        buf, off, pit = conv.transform_frame(pageable, out=out)
        assert_planes_equal(plane_views(buf, off, pit, 0, w, h), want, "registered")
    assert conv.stats().staged_in_bytes == 0
    conv.stop()
    pin_in.free()
    pin_out.free()


def test_damage_map_sequence(conv, orc):
    """C4: mostly static desktop. First frame all ones; then exactly the oracle's map; reset -> ones."""
    w, h = 1920, 1080
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=0, flags=b200vc.FLAG_DAMAGE_MAP)
    prev = None
    for k in range(5):
        frame = orc.synth("desk", w, h, 28 + k)      # crosses the caret toggle at frame 30
        buf, off, pit, dm = conv.transform_frame(frame, damage=True)
        assert np.array_equal(dm, orc.damage_map(frame, prev)), f"frame {k}"
        assert_planes_equal(plane_views(buf, off, pit, 0, w, h), _oracle_planes(orc, frame, 0, 0, 0, 0), f"frame {k}")
        if k:
            assert 0 < dm.mean() < 0.06
        prev = frame
    same = conv.transform_frame(prev, damage=True)[3]
    assert same.sum() == 0                            # identical frame: nothing damaged
    xonly = prev.copy()
    xonly[..., 3] ^= 0xFF
    assert conv.transform_frame(xonly, damage=True)[3].sum() == 0     # pad byte is not damage
    conv.reset()
    assert conv.transform_frame(prev, damage=True)[3].all()


@pytest.mark.parametrize("size", [(130, 18), (126, 1082), (1366, 768)])
def test_damage_map_ragged(conv, orc, size):
    w, h = size
    conv.set_info(w, h, flags=b200vc.FLAG_DAMAGE_MAP, chroma_site=1, matrix=1, range_=0)
    a = orc.synth("rand", w, h, 1)
    b = a.copy()
    rng = np.random.default_rng(5)
    for _ in range(12):
        b[rng.integers(0, h), rng.integers(0, w), rng.integers(0, 4)] ^= 0x10
    b[h - 1, w - 1, 2] ^= 1
    conv.transform_frame(a, damage=True)
    dm = conv.transform_frame(b, damage=True)[3]
    assert np.array_equal(dm, orc.damage_map(b, a))


def test_batch_of_sessions_and_async(conv, orc):
    """C5 on one GPU: 8 sessions per tick, each with its own previous frame."""
    w, h, ns = 640, 368, 8
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_DAMAGE_MAP, max_sessions=ns, n_inflight=3)
    total, off, pit = b200vc.default_layout(0, w, h)
    mbh, mbw = conv.mb_shape
    ins = b200vc.PinnedBuffer(ns * w * h * 4)
    outs = b200vc.PinnedBuffer(ns * (total + mbh * mbw))
    prev = [None] * ns
    for tick in range(3):
        frames, host = [], []
        for s in range(ns):
            view = ins.array[s * w * h * 4:(s + 1) * w * h * 4].reshape(h, w * 4)
            f = orc.synth("desk", w, h, 100 * s + tick, out=view)
            host.append(np.array(f))
            o = s * (total + mbh * mbw)
            frames.append(conv.make_frame(f.ctypes.data, w * 4, outs.array.ctypes.data + o, off, pit, session=s,
                                          damage_ptr=outs.array.ctypes.data + o + total))
        if tick < 2:
            conv.convert_batch(frames)
        else:                                          # same tick through submit/collect
            tickets = [conv.submit(f) for f in frames]
            for t in reversed(tickets):
                conv.collect(t)
        for s in range(ns):
            o = s * (total + mbh * mbw)
            got = plane_views(outs.array[o:o + total], off, pit, 0, w, h)
            assert_planes_equal(got, _oracle_planes(orc, host[s], 0, 1, 0, 1), f"tick {tick} session {s}")
            dm = outs.array[o + total:o + total + mbh * mbw].reshape(mbh, mbw)
            assert np.array_equal(dm, orc.damage_map(host[s], prev[s])), f"tick {tick} session {s}"
            prev[s] = host[s]
    assert conv.stats().frames == 3 * ns
    with pytest.raises(b200vc.B200vcError):
        conv.collect(10**9)
    conv.stop()
    ins.free()
    outs.free()


def test_device_entry_point_batched(conv, orc):
    """Kernel-only path: a ring of frames resident in HBM, one batched launch on the caller's stream."""
    import torch
    w, h, n = 1920, 1080, 6
    total, off, pit = b200vc.default_layout(0, w, h)
    mbh, mbw = (h + 15) // 16, (w + 15) // 16
    host = np.stack([orc.synth("desk", w, h, 40 + i) for i in range(n)])
    d_in = torch.from_numpy(host.reshape(n, -1)).cuda()
    d_out = torch.zeros((n, total), dtype=torch.uint8, device="cuda")
    d_dmg = torch.zeros((n, mbh * mbw), dtype=torch.uint8, device="cuda")
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=0, flags=b200vc.FLAG_DAMAGE_MAP)
    df = b200vc.DeviceFrames()
    df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), w * 4, w * h * 4
    # frame i diffs against frame i-1 of the same ring; frame 0 against the last one
    d_prev = torch.roll(d_in, 1, 0).contiguous()
    df.prev, df.prev_pitch, df.prev_frame_stride = d_prev.data_ptr(), w * 4, w * h * 4
    for i in range(3):
        df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
    df.out_frame_stride, df.damage, df.damage_frame_stride, df.n_frames = total, d_dmg.data_ptr(), mbh * mbw, n
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        conv.convert_device(df, stream.cuda_stream)
    stream.synchronize()
    got, dmg = d_out.cpu().numpy(), d_dmg.cpu().numpy()
    for i in range(n):
        assert_planes_equal(plane_views(got[i], off, pit, 0, w, h), _oracle_planes(orc, host[i], 0, 0, 0, 0), f"frame {i}")
        assert np.array_equal(dmg[i].reshape(mbh, mbw), orc.damage_map(host[i], host[i - 1]))
    # no previous frame: map is all ones
    df.prev = None
    conv.convert_device(df, 0)
    torch.cuda.synchronize()
    assert d_dmg.cpu().numpy().all()
    # misaligned pointer is refused, not silently mis-stored
    df.planes[0] = d_out.data_ptr() + 4
    with pytest.raises(b200vc.B200vcError) as e:
        conv.convert_device(df, 0)
    assert e.value.status == b200vc.ERR_ALIGNMENT


def test_full_size_properties(conv, orc):
    """BASELINE sizes: checksum-of-planes equality with the oracle, determinism across repeats, and
    independence from the pad byte."""
    for (w, h, fmt, m, rg, site) in [(1920, 1080, 0, 0, 0, 0), (2560, 1440, 1, 1, 1, 1)]:
        frame = orc.synth("rand", w, h, 77)
        conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site)
        buf, off, pit = conv.transform_frame(frame)
        want = plane_checksums(_oracle_planes(orc, frame, fmt, m, rg, site))
        assert plane_checksums(plane_views(buf, off, pit, fmt, w, h)) == want
        again = conv.transform_frame(frame)[0]
        assert np.array_equal(buf, again)
        nox = frame.copy()
        nox[..., 3] = 0
        assert np.array_equal(conv.transform_frame(nox)[0], buf)


def test_errors_at_run_time(conv):
    conv.set_info(64, 32)
    with pytest.raises(b200vc.B200vcError) as e:      # damage without the flag
        conv.transform_frame(np.zeros((32, 64, 4), np.uint8), damage=True)
    assert e.value.status == b200vc.ERR_INVALID_ARG
    with pytest.raises(b200vc.B200vcError):           # pitch smaller than a row
        conv.transform_frame(np.zeros((32, 64, 4), np.uint8), pitches=[32, 32, 32])
    with pytest.raises(b200vc.B200vcError) as e:      # caps the element does not offer
        conv.set_info(64, 32, out_format=5)
    assert e.value.status == b200vc.ERR_UNSUPPORTED


@pytest.mark.parametrize("inf", [1, 2, 3])
def test_input_formats_on_gpu(conv, orc, inf):
    """RGBx/RGBA, xRGB/ARGB, xBGR/ABGR sink formats: conversion, fused scale and damage map."""
    for (w, h) in [(1920, 1080), (134, 34)]:
        f = orc.synth("rand", w, h, 21)
        for fmt, m, rg, site in [(0, 1, 0, 1), (1, 0, 1, 0)]:
            conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site, in_format=inf,
                          flags=b200vc.FLAG_DAMAGE_MAP)
            buf, off, pit, dm = conv.transform_frame(f, damage=True)
            assert_planes_equal(plane_views(buf, off, pit, fmt, w, h),
                                plane_views(*orc.convert(f, fmt, m, rg, site, in_format=inf), fmt, w, h), f"in{inf} {w}x{h}")
            g = f.copy()
            g[5, 7, 0] ^= 1
            g[h - 1, w - 1, 3] ^= 1
            dm = conv.transform_frame(g, damage=True)[3]
            assert np.array_equal(dm, orc.damage_map(g, f, inf)) and dm.sum() == 1
    f = orc.synth("rand", 1280, 720, 22)
    for flags in (0, b200vc.FLAG_FORCE_GENERIC_FUSED, b200vc.FLAG_FORCE_TWO_PASS):
        conv.set_info(1280, 720, 640, 360, matrix=1, range_=0, chroma_site=1, in_format=inf, flags=flags)
        buf, off, pit = conv.transform_frame(f)
        want = orc.convert(f, 0, 1, 0, 1, in_format=inf, out_size=(640, 360))
        assert_planes_equal(plane_views(buf, off, pit, 0, 640, 360), plane_views(*want, 0, 640, 360), f"in{inf} scaled {flags}")


@pytest.mark.parametrize("size", [(3, 3), (5, 2), (2, 5), (7, 9), (131, 35), (1365, 767), (1919, 1079), (129, 17), (127, 15)])
def test_odd_sizes(conv, orc, size):
    """Odd widths and heights (unscaled path): the last pixel / last line is a chroma site of its own."""
    w, h = size
    f = orc.synth("rand", w, h, 31)
    g = f.copy()
    g[h - 1, w - 1, 1] ^= 2
    for fmt in (0, 1):
        for m, rg, site in [(0, 0, 0), (1, 0, 1), (1, 1, 1), (0, 1, 0)]:
            conv.set_info(w, h, out_format=fmt, matrix=m, range_=rg, chroma_site=site, flags=b200vc.FLAG_DAMAGE_MAP)
            buf, off, pit, dm = conv.transform_frame(f, damage=True)
            assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), _oracle_planes(orc, f, fmt, m, rg, site),
                                f"{size} fmt{fmt} m{m} r{rg} s{site}")
            dm = conv.transform_frame(g, damage=True)[3]
            assert np.array_equal(dm, orc.damage_map(g, f)) and dm.sum() == 1
    with pytest.raises(b200vc.B200vcError) as e:      # the 2x2 box mean needs even sizes
        conv.set_info(65, 37, 32, 18, scale_mode=1)
    assert e.value.status == b200vc.ERR_UNSUPPORTED


def test_zero_copy_mode(conv, orc):
    """B200VC_FLAG_ZERO_COPY: the kernel reads the page-locked frame and writes the planes itself."""
    w, h = 1920, 1080
    total, off, pit = b200vc.default_layout(1, w, h)
    pin_in, pin_out = b200vc.PinnedBuffer(w * h * 4), b200vc.PinnedBuffer(total)
    f = orc.synth("desk", w, h, 5, out=pin_in.array.reshape(h, w * 4))
    conv.set_info(w, h, out_format=1, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_ZERO_COPY)
    pin_out.array[:] = 0
    buf, off, pit = conv.transform_frame(f, out=pin_out.array)
    want = _oracle_planes(orc, f, 1, 1, 0, 1)
    assert_planes_equal(plane_views(buf, off, pit, 1, w, h), want, "zero-copy")
    st = conv.stats()
    assert st.kernel_launches == 1 and st.h2d_bytes == w * h * 4 and st.d2h_bytes == total
    # pageable buffers cannot be read by the kernel: the call falls back to the copy pipeline
    buf, off, pit = conv.transform_frame(np.array(f))
    assert_planes_equal(plane_views(buf, off, pit, 1, w, h), want, "zero-copy flag, pageable buffers")
    assert conv.stats().staged_in_bytes == w * h * 4
    conv.stop()
    pin_in.free()
    pin_out.free()


def test_zero_copy_out_mode(conv, orc):
    """B200VC_FLAG_ZERO_COPY_OUT: input through the copy engine, planes and damage map stored by the
    kernels straight into page-locked host memory — unscaled with damage, and scaled."""
    w, h = 1920, 1080
    total, off, pit = b200vc.default_layout(0, w, h)
    mbh, mbw = (h + 15) // 16, (w + 15) // 16
    pin_in, pin_out = b200vc.PinnedBuffer(w * h * 4), b200vc.PinnedBuffer(total + mbh * mbw)
    planes = (C.c_void_p * 3)(*[pin_out.ptr + o for o in off])
    pits = (C.c_size_t * 3)(*pit)
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=1, n_bands=4, flags=b200vc.FLAG_ZERO_COPY_OUT | b200vc.FLAG_DAMAGE_MAP)
    prev = None
    for k in range(3):
        f = orc.synth("desk", w, h, 70 + k, out=pin_in.array.reshape(h, w * 4))
        host = np.array(f)
        pin_out.array[:] = 0
        assert conv.lib.b200vc_convert(conv.ctx, pin_in.ptr, w * 4, planes, pits, pin_out.ptr + total) == 0
        assert_planes_equal(plane_views(pin_out.array[:total], off, pit, 0, w, h), _oracle_planes(orc, host, 0, 0, 0, 1), f"frame {k}")
        assert np.array_equal(pin_out.array[total:].reshape(mbh, mbw), orc.damage_map(host, prev)), f"frame {k} damage"
        prev = host
    st = conv.stats()
    assert st.d2h_bytes == 3 * (total + mbh * mbw) and st.h2d_bytes == 3 * w * h * 4
    conv.set_info(w, h, w // 2, h // 2, out_format=1, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_ZERO_COPY_OUT)
    t2, o2, p2 = b200vc.default_layout(1, w // 2, h // 2)
    f = orc.synth("rand", w, h, 3, out=pin_in.array.reshape(h, w * 4))
    pl = (C.c_void_p * 3)(pin_out.ptr + o2[0], pin_out.ptr + o2[1], None)
    pt = (C.c_size_t * 3)(*p2)
    assert conv.lib.b200vc_convert(conv.ctx, pin_in.ptr, w * 4, pl, pt, None) == 0
    want = orc.convert(np.array(f), 1, 1, 0, 1, out_size=(w // 2, h // 2))
    assert_planes_equal(plane_views(pin_out.array[:t2], o2, p2, 1, w // 2, h // 2), plane_views(*want, 1, w // 2, h // 2), "scaled")
    conv.stop()
    pin_in.free()
    pin_out.free()


@pytest.mark.parametrize("scaled", [False, True])
def test_one_frame_latency_call_order(conv, orc, scaled):
    """The element's one-frame-latency mode: one session, frame N submitted before frame N-1 is
    collected, each frame with its own input and output buffers (gst/gstb200convert.c generate_output)."""
    w, h = 640, 360
    ow, oh = (480, 270) if scaled else (w, h)
    conv.set_info(w, h, ow, oh, matrix=1, range_=0, chroma_site=1, n_inflight=2)
    total, off, pit = b200vc.default_layout(0, ow, oh)
    n = 6
    ins, outs = b200vc.PinnedBuffer(n * w * h * 4), b200vc.PinnedBuffer(n * total)
    host, tickets = [], []
    for k in range(n):
        view = ins.array[k * w * h * 4:(k + 1) * w * h * 4].reshape(h, w * 4)
        host.append(np.array(orc.synth("desk", w, h, k, out=view)))
        f = conv.make_frame(view.ctypes.data, w * 4, outs.array.ctypes.data + k * total, off, pit, session=0)
        tickets.append(conv.submit(f))
        if k:
            conv.collect(tickets[k - 1])
            got = plane_views(outs.array[(k - 1) * total:k * total], off, pit, 0, ow, oh)
            want, woff, wpit = orc.convert(host[k - 1], 0, 1, 0, 1, out_size=(ow, oh))
            assert_planes_equal(got, plane_views(want, woff, wpit, 0, ow, oh), f"frame {k - 1}")
    conv.collect(tickets[-1])
    got = plane_views(outs.array[(n - 1) * total:n * total], off, pit, 0, ow, oh)
    want, woff, wpit = orc.convert(host[-1], 0, 1, 0, 1, out_size=(ow, oh))
    assert_planes_equal(got, plane_views(want, woff, wpit, 0, ow, oh), "last frame")
    conv.stop()
    ins.free()
    outs.free()
