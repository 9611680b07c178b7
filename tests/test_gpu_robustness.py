"""GPU robustness tests of the C-ABI boundary: the streaming thread is not ours, so contexts are hit
from several threads; renegotiation must not leak device memory; a failed call must leave the
context usable."""
import ctypes as C
import threading

import numpy as np
import pytest

import b200vc
from helpers import assert_planes_equal, plane_views

pytestmark = pytest.mark.gpu


def test_sessions_from_many_threads(orc):
    """Eight session threads share one context (submit/collect from each); outputs stay bit-exact."""
    w, h, ns, frames = 640, 368, 8, 6
    conv = b200vc.Converter(0)
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_DAMAGE_MAP, max_sessions=ns, n_inflight=4)
    total, off, pit = b200vc.default_layout(0, w, h)
    mbh, mbw = conv.mb_shape
    errors = []

    def session(s):
        try:
            pin_in, pin_out = b200vc.PinnedBuffer(w * h * 4), b200vc.PinnedBuffer(total + mbh * mbw)
            prev = None
            for k in range(frames):
                f = orc.synth("desk", w, h, 1000 * s + k, out=pin_in.array.reshape(h, w * 4))
                host = np.array(f)
                fr = conv.make_frame(pin_in.ptr, w * 4, pin_out.ptr, off, pit, session=s, damage_ptr=pin_out.ptr + total)
                conv.collect(conv.submit(fr))
                want = plane_views(*orc.convert(host, 0, 1, 0, 1), 0, w, h)
                assert_planes_equal(plane_views(pin_out.array[:total], off, pit, 0, w, h), want, f"session {s} frame {k}")
                dm = pin_out.array[total:total + mbh * mbw].reshape(mbh, mbw)
                assert np.array_equal(dm, orc.damage_map(host, prev)), f"session {s} frame {k} damage"
                prev = host
            pin_in.free()
            pin_out.free()
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=session, args=(s,)) for s in range(ns)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[0]
    assert conv.stats().frames == ns * frames
    conv.stop()


def test_two_contexts_at_once(orc):
    a, b = b200vc.Converter(0), b200vc.Converter(0)
    a.set_info(1280, 720, out_format=0, matrix=0, range_=0, chroma_site=0)
    b.set_info(1280, 720, 640, 360, out_format=1, matrix=1, range_=1, chroma_site=1)
    f = orc.synth("rand", 1280, 720, 3)
    res = {}

    def run(name, conv, n):
        for _ in range(n):
            res[name] = conv.transform_frame(f)

    ts = [threading.Thread(target=run, args=("a", a, 5)), threading.Thread(target=run, args=("b", b, 5))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert_planes_equal(plane_views(*res["a"], 0, 1280, 720), plane_views(*orc.convert(f, 0, 0, 0, 0), 0, 1280, 720))
    assert_planes_equal(plane_views(*res["b"], 1, 640, 360),
                        plane_views(*orc.convert(f, 1, 1, 1, 1, out_size=(640, 360)), 1, 640, 360))
    a.stop()
    b.stop()


def test_renegotiation_does_not_leak():
    import torch
    torch.cuda.init()
    conv = b200vc.Converter(0)
    geos = [(1920, 1080, 1920, 1080, 0), (3840, 2160, 1920, 1080, 0), (1280, 720, 854, 480, b200vc.FLAG_FORCE_TWO_PASS),
            (2560, 1440, 2560, 1440, b200vc.FLAG_DAMAGE_MAP)]
    frame = {g[:2]: np.zeros((g[1], g[0], 4), np.uint8) for g in geos}
    for iw, ih, ow, oh, fl in geos:                       # warm every kernel image first
        conv.set_info(iw, ih, ow, oh, flags=fl, max_sessions=2)
        conv.transform_frame(frame[(iw, ih)])
    conv.stop()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(5):
        for iw, ih, ow, oh, fl in geos:
            conv.set_info(iw, ih, ow, oh, flags=fl, max_sessions=2)
            conv.transform_frame(frame[(iw, ih)])
    conv.stop()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0 - (8 << 20), "device memory shrank across 20 renegotiations"


def test_context_survives_a_refused_call(orc):
    conv = b200vc.Converter(0)
    conv.set_info(320, 180, matrix=0, range_=0, chroma_site=0)
    f = orc.synth("rand", 320, 180, 1)
    with pytest.raises(b200vc.B200vcError):
        conv.transform_frame(f, pitches=[16, 16, 16])
    planes = (C.c_void_p * 3)()
    pits = (C.c_size_t * 3)(320, 160, 160)
    assert conv.lib.b200vc_convert(conv.ctx, f.ctypes.data, 1280, planes, pits, None) == b200vc.ERR_INVALID_ARG
    assert b"NULL" in conv.lib.b200vc_last_error(conv.ctx)
    buf, off, pit = conv.transform_frame(f)               # still works
    assert_planes_equal(plane_views(buf, off, pit, 0, 320, 180), plane_views(*orc.convert(f, 0, 0, 0, 0), 0, 320, 180))
    conv.stop()


def test_device_pointer_on_the_host_entry_point_is_refused():
    import torch
    conv = b200vc.Converter(0)
    conv.set_info(64, 32)
    d = torch.zeros(64 * 32 * 4, dtype=torch.uint8, device="cuda")
    total, off, pit = b200vc.default_layout(0, 64, 32)
    out = np.zeros(total, np.uint8)
    planes = (C.c_void_p * 3)(*[out.ctypes.data + o for o in off])
    pits = (C.c_size_t * 3)(*pit)
    assert conv.lib.b200vc_convert(conv.ctx, d.data_ptr(), 256, planes, pits, None) == b200vc.ERR_INVALID_ARG
    assert b"b200vc_convert_device" in conv.lib.b200vc_last_error(conv.ctx)
    conv.stop()


def test_8k_frame(orc):
    """7680x4320 (133 MB in): offsets beyond 2^27 bytes, 480 x 270 macroblocks."""
    w, h = 7680, 4320
    f = orc.synth("rand", w, h, 2)
    conv = b200vc.Converter(0)
    conv.set_info(w, h, out_format=1, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_DAMAGE_MAP)
    buf, off, pit, dm = conv.transform_frame(f, damage=True)
    want = orc.convert(f, 1, 1, 0, 1, threads=8)
    assert_planes_equal(plane_views(buf, off, pit, 1, w, h), plane_views(*want, 1, w, h), "8K NV12")
    g = f.copy()
    g[h - 1, w - 1, 0] ^= 1
    g[0, 0, 2] ^= 1
    dm = conv.transform_frame(g, damage=True)[3]
    assert dm.shape == (270, 480) and dm.sum() == 2 and dm[0, 0] == 1 and dm[269, 479] == 1
    conv.set_info(w, h, w // 2, h // 2, matrix=1, range_=0, chroma_site=1)
    assert conv.info().fused_2to1 == 1
    buf, off, pit = conv.transform_frame(f)
    want = orc.convert(f, 0, 1, 0, 1, out_size=(w // 2, h // 2), threads=8)
    assert_planes_equal(plane_views(buf, off, pit, 0, w // 2, h // 2), plane_views(*want, 0, w // 2, h // 2), "8K -> 4K")
    conv.stop()


def test_two_contexts_with_different_windows(orc):
    """Two converters in one process share the kernels: negotiating the second (small window) must not
    take shared memory away from the first (large window)."""
    big, small = b200vc.Converter(0), b200vc.Converter(0)
    try:
        big.set_info(1920, 1080, 640, 360, matrix=1, range_=0, chroma_site=1)            # 3:1, ~30 KB window
        small.set_info(64, 36, 48, 28, matrix=1, range_=0, chroma_site=1)                # a few KB
        assert big.info().path == 1 and small.info().path == 1
        for conv, (iw, ih, ow, oh) in ((big, (1920, 1080, 640, 360)), (small, (64, 36, 48, 28)), (big, (1920, 1080, 640, 360))):
            frame = orc.synth("rand", iw, ih, 5)
            buf, off, pit = conv.transform_frame(frame)
            want, woff, wpit = orc.convert(frame, 0, 1, 0, 1, out_size=(ow, oh))
            assert_planes_equal(plane_views(buf, off, pit, 0, ow, oh), plane_views(want, woff, wpit, 0, ow, oh), f"{iw}x{ih}")
    finally:
        big.stop()
        small.stop()
