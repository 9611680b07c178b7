"""GPU tests of the host pipeline around the kernels (SURVEY §8 b "Synchrony" and "Threading" rows, and the
pinned double-buffered PCIe path of north_star): the synchronous call in all its variants (bands chained by
events, or row-gated), one-copy downloads, staging of pageable buffers, page-locking on
first sight, and several session threads on one context. Everything is compared byte for byte with the
oracle through the C ABI."""
import ctypes as C
import threading
import time

import numpy as np
import pytest

import b200vc
from helpers import assert_planes_equal, plane_views

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def conv():
    c = b200vc.Converter(0)
    yield c
    c.stop()


def _want(orc, frame, ow, oh, fmt, m, rg, site):
    buf, off, pit = orc.convert(frame, fmt, m, rg, site, out_size=(ow, oh))
    return plane_views(buf, off, pit, fmt, ow, oh)


GEOS = [(1920, 1080, 1920, 1080),      # K1 in bands (or row-gated), planes stored by the kernels when page-locked
        (1366, 768, 1366, 768),        # pitches that are not multiples of 16: padded device frame, 2-D copies
        (3840, 2160, 1920, 1080),      # exact-2:1 fused kernel behind the gate
        (1920, 1080, 1280, 720),       # general fused kernel behind the gate
        (1920, 1080, 320, 180)]        # two-pass: K1 gated into the intermediate, then the plane scaler


@pytest.mark.parametrize("geo", GEOS)
@pytest.mark.parametrize("flags", [0, b200vc.FLAG_ROW_GATE])
def test_synchronous_call_variants(conv, orc, geo, flags):
    iw, ih, ow, oh = geo
    fmt, m, rg, site = 0, 1, 0, 1
    total, off, pit = b200vc.default_layout(fmt, ow, oh)
    pin_in, pin_out = b200vc.PinnedBuffer(iw * ih * 4), b200vc.PinnedBuffer(total)
    frames = [orc.synth("rand", iw, ih, 40 + k) for k in range(2)]
    wants = [_want(orc, f, ow, oh, fmt, m, rg, site) for f in frames]
    for bands in (1, 8, 64):
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, flags=flags, n_bands=bands)
        for rnd in range(3):                       # the same slot and flags are reused frame after frame
            k = rnd & 1
            # pageable in and out (staged through the pinned buffers)
            buf, off, pit = conv.transform_frame(frames[k])
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), wants[k], f"{geo} flags {flags} bands {bands} pageable")
            # page-locked in and out
            pin_in.array[:] = frames[k].reshape(-1)
            pin_out.array[:] = 0
            buf, off, pit = conv.transform_frame(pin_in.array.reshape(ih, iw, 4), out=pin_out.array)
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), wants[k], f"{geo} flags {flags} bands {bands} pinned")
        st = conv.stats()
        assert st.frames == 6 and st.staged_in_bytes == 3 * iw * ih * 4
    pin_in.free()
    pin_out.free()


def test_gated_call_with_damage_map_and_nv12(conv, orc):
    w, h = 1280, 720
    total, off, pit = b200vc.default_layout(1, w, h)
    mb = ((w + 15) // 16) * ((h + 15) // 16)
    pin_in, pin_out = b200vc.PinnedBuffer(w * h * 4), b200vc.PinnedBuffer(total + mb)
    conv.set_info(w, h, out_format=1, matrix=0, range_=1, chroma_site=0, flags=b200vc.FLAG_DAMAGE_MAP)
    planes = (C.c_void_p * 3)(pin_out.ptr + off[0], pin_out.ptr + off[1], None)
    pits = (C.c_size_t * 3)(*pit)
    prev = None
    for k in range(4):
        frame = orc.synth("desk", w, h, k)
        pin_in.array[:] = frame.reshape(-1)
        # page-locked planes and page-locked map: nothing is staged
        conv._check(conv.lib.b200vc_convert(conv.ctx, pin_in.ptr, w * 4, planes, pits, pin_out.ptr + total))
        assert_planes_equal(plane_views(pin_out.array, off, pit, 1, w, h), _want(orc, frame, w, h, 1, 0, 1, 0), f"frame {k}")
        assert np.array_equal(pin_out.array[total:total + mb].reshape(h // 16, w // 16), orc.damage_map(frame, prev))
        prev = frame
    st = conv.stats()
    assert st.staged_in_bytes == 0 and st.staged_out_bytes == 0
    # a pageable map alone is staged alone (8 KB class), the planes still land in place
    dmg = np.zeros(mb, np.uint8)
    frame = orc.synth("desk", w, h, 9)
    pin_in.array[:] = frame.reshape(-1)
    conv._check(conv.lib.b200vc_convert(conv.ctx, pin_in.ptr, w * 4, planes, pits, dmg.ctypes.data))
    assert np.array_equal(dmg.reshape(h // 16, w // 16), orc.damage_map(frame, prev))
    assert_planes_equal(plane_views(pin_out.array, off, pit, 1, w, h), _want(orc, frame, w, h, 1, 0, 1, 0))
    assert conv.stats().staged_out_bytes == mb
    pin_in.free()
    pin_out.free()


@pytest.mark.parametrize("contiguous", [True, False])
def test_batch_download_one_copy_or_per_plane(conv, orc, contiguous):
    """A page-locked caller frame laid out like GstVideoInfo's default leaves the GPU in one copy; planes in
    separate buffers (or with other pitches) go plane by plane. Same bytes either way."""
    w, h, n = 1920, 1080, 4
    total, off, pit = b200vc.default_layout(0, w, h)
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=1, max_sessions=n, n_inflight=n)
    pin_in = b200vc.PinnedBuffer(n * w * h * 4)
    if contiguous:
        pin_out, offs, pits = b200vc.PinnedBuffer(n * total), off, pit
        stride = total
    else:
        pits = [2048, 1024, 960]
        offs = [0, pits[0] * h + 64 * 1024, pits[0] * h + 64 * 1024 + pits[1] * (h // 2) + 4096]
        stride = offs[2] + pits[2] * (h // 2) + 4096
        pin_out = b200vc.PinnedBuffer(n * stride)
    frames = [orc.synth("rand", w, h, 70 + k) for k in range(n)]
    for k in range(n):
        pin_in.array[k * w * h * 4:(k + 1) * w * h * 4] = frames[k].reshape(-1)
    pin_out.array[:] = 0
    conv.convert_batch([conv.make_frame(pin_in.ptr + k * w * h * 4, w * 4, pin_out.ptr + k * stride, offs, pits, session=k)
                        for k in range(n)])
    for k in range(n):
        assert_planes_equal(plane_views(pin_out.array[k * stride:(k + 1) * stride], offs, pits, 0, w, h),
                            _want(orc, frames[k], w, h, 0, 0, 0, 1), f"frame {k}")
    st = conv.stats()
    assert st.d2h_bytes == n * total and st.staged_out_bytes == 0
    pin_in.free()
    pin_out.free()


def test_register_host_keeps_a_bounded_set_of_ranges(conv, orc):
    """More distinct caller buffers than the library keeps page-locked: the least recently used range is
    released and the new one locked; results stay right throughout."""
    w, h = 640, 360
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_REGISTER_HOST)
    total = b200vc.default_layout(0, w, h)[0]
    bufs = [np.array(orc.synth("rand", w, h, 100 + k)) for k in range(20)]
    outs = [np.zeros(total, np.uint8) for _ in range(20)]
    for rnd in range(2):
        for k in range(20):
            buf, off, pit = conv.transform_frame(bufs[k], out=outs[k])
            assert_planes_equal(plane_views(buf, off, pit, 0, w, h), _want(orc, bufs[k], w, h, 0, 1, 0, 1), f"buffer {k} round {rnd}")
    assert conv.stats().staged_in_bytes == 0
    conv.stop()                                   # releases what is still page-locked


def test_session_threads_overlap_on_one_context(conv, orc):
    """Two desktop sessions share a context, each driven synchronously by its own streaming thread with
    pageable buffers (the staging copies are the long part). The context's lock is not held across those
    copies nor across the wait for the frame, so two threads take clearly less than twice one thread."""
    w, h, n = 1920, 1080, 60
    total = b200vc.default_layout(0, w, h)[0]
    conv.set_info(w, h, matrix=0, range_=0, chroma_site=1, max_sessions=2, n_inflight=2)
    frames = [np.array(orc.synth("rand", w, h, 200 + s)) for s in range(2)]
    outs = [np.zeros(total, np.uint8) for _ in range(2)]
    wants = [_want(orc, frames[s], w, h, 0, 0, 0, 1) for s in range(2)]
    errors = []

    def run(s, count):
        try:
            for _ in range(count):
                conv.transform_frame(frames[s], out=outs[s], session=s)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    for s in range(2):
        run(s, 3)                                  # staging buffers allocated, kernels loaded
    t0 = time.perf_counter()
    run(0, n)
    run(1, n)
    serial = time.perf_counter() - t0
    ts = [threading.Thread(target=run, args=(s, n)) for s in range(2)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    both = time.perf_counter() - t0
    assert not errors, errors
    _, off, pit = b200vc.default_layout(0, w, h)
    for s in range(2):
        assert_planes_equal(plane_views(outs[s], off, pit, 0, w, h), wants[s], f"session {s}")
    print(f"two sessions: serial {serial * 1e3:.1f} ms, two threads {both * 1e3:.1f} ms, ratio {both / serial:.2f}")
    assert both < 0.7 * serial, f"two session threads took {both / serial:.2f} of the serial time"


def test_synchronous_call_waits_for_its_own_frame_only(conv, orc):
    """A frame submitted by another session stays in flight across someone else's synchronous call."""
    w, h = 1280, 720
    total, off, pit = b200vc.default_layout(0, w, h)
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, max_sessions=2, n_inflight=4)
    pin_in, pin_out = b200vc.PinnedBuffer(2 * w * h * 4), b200vc.PinnedBuffer(2 * total)
    frames = [orc.synth("rand", w, h, 300 + s) for s in range(2)]
    for s in range(2):
        pin_in.array[s * w * h * 4:(s + 1) * w * h * 4] = frames[s].reshape(-1)
    t1 = conv.submit(conv.make_frame(pin_in.ptr + w * h * 4, w * 4, pin_out.ptr + total, off, pit, session=1))
    buf, off, pit = conv.transform_frame(pin_in.array[:w * h * 4].reshape(h, w, 4), out=pin_out.array[:total], session=0)
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), _want(orc, frames[0], w, h, 0, 1, 0, 1), "session 0")
    conv.collect(t1)                               # still collectable: the synchronous call left it alone
    assert_planes_equal(plane_views(pin_out.array[total:], off, pit, 0, w, h), _want(orc, frames[1], w, h, 0, 1, 0, 1), "session 1")
    pin_in.free()
    pin_out.free()


def test_two_pass_device_calls_on_two_streams_are_ordered(conv, orc):
    """b200vc_convert_device on the two-pass path shares the context's intermediate frame: calls on different
    streams are ordered by the library, so back-to-back launches on two streams both come out right."""
    import torch
    iw, ih, ow, oh = 1280, 720, 800, 448
    total, off, pit = b200vc.default_layout(0, ow, oh)
    conv.set_info(iw, ih, ow, oh, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_FORCE_TWO_PASS)
    assert conv.info().path == 2
    hosts = [orc.synth("rand", iw, ih, 400 + k) for k in range(2)]
    d_in = [torch.from_numpy(hh.reshape(-1)).cuda() for hh in hosts]
    d_out = [torch.zeros(total, dtype=torch.uint8, device="cuda") for _ in range(2)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    for rnd in range(8):
        for k in range(2):
            df = b200vc.DeviceFrames()
            df.bgrx, df.in_pitch, df.in_frame_stride = d_in[k].data_ptr(), iw * 4, iw * ih * 4
            for i in range(3):
                df.planes[i], df.pitches[i] = d_out[k].data_ptr() + off[i], pit[i]
            df.out_frame_stride, df.n_frames = total, 1
            conv.convert_device(df, streams[k].cuda_stream)
    torch.cuda.synchronize()
    for k in range(2):
        assert_planes_equal(plane_views(d_out[k].cpu().numpy(), off, pit, 0, ow, oh), _want(orc, hosts[k], ow, oh, 0, 1, 0, 1), f"stream {k}")


@pytest.mark.parametrize("latency_mode", [False, True])
def test_damage_map_to_rectangles_through_the_abi(conv, orc, latency_mode):
    """SURVEY §8 f-4 end to end inside the boundary: S-desk frames -> b200vc_convert (or the element's
    one-frame-latency order: submit frame N with its own page-locked map, collect frame N-1) -> the map ->
    b200vc_damage_to_rects; the rectangles must cover exactly the macroblocks the oracle's map marks."""
    w, h, n = 1920, 1080, 6
    total, off, pit = b200vc.default_layout(0, w, h)
    mb_h, mb_w = (h + 15) // 16, (w + 15) // 16
    conv.set_info(w, h, matrix=1, range_=0, chroma_site=1, flags=b200vc.FLAG_DAMAGE_MAP, n_inflight=2)
    pin_in = b200vc.PinnedBuffer(2 * w * h * 4)
    pin_out = b200vc.PinnedBuffer(2 * total)
    maps = b200vc.PinnedBuffer(2 * mb_h * mb_w)                  # one map per frame that can be in flight
    frames = [orc.synth("desk", w, h, 50 + k) for k in range(n)]

    def check(k, slot):
        want = orc.damage_map(frames[k], frames[k - 1] if k else None)
        got = maps.array[slot * mb_h * mb_w:(slot + 1) * mb_h * mb_w].reshape(mb_h, mb_w)
        assert np.array_equal(got, want), f"frame {k}: map differs from the oracle"
        rects, need = b200vc.damage_rects(got, w, h)
        assert need == len(rects)
        cover = np.zeros_like(want)
        for x, y, rw, rh in rects:
            assert x % 16 == 0 and y % 16 == 0 and x + rw <= w and y + rh <= h
            cover[y // 16:(y + rh + 15) // 16, x // 16:(x + rw + 15) // 16] += 1
        assert np.array_equal(cover, want), f"frame {k}: rectangles do not cover exactly the dirty macroblocks"
        if k:
            assert 0 < want.mean() < 0.08                          # a mostly static desktop
        assert_planes_equal(plane_views(pin_out.array[slot * total:(slot + 1) * total], off, pit, 0, w, h),
                            _want(orc, frames[k], w, h, 0, 1, 0, 1), f"frame {k}")

    pending = None
    for k in range(n):
        slot = k & 1
        pin_in.array[slot * w * h * 4:(slot + 1) * w * h * 4] = frames[k].reshape(-1)
        f = conv.make_frame(pin_in.ptr + slot * w * h * 4, w * 4, pin_out.ptr + slot * total, off, pit, session=0,
                            damage_ptr=maps.ptr + slot * mb_h * mb_w)
        if latency_mode:
            ticket = conv.submit(f)
            if pending is not None:
                conv.collect(pending[0])
                check(pending[1], pending[2])
            pending = (ticket, k, slot)
        else:
            conv.convert_batch([f])
            check(k, slot)
    if pending is not None:
        conv.collect(pending[0])
        check(pending[1], pending[2])
    assert conv.stats().staged_out_bytes == 0
    for b in (pin_in, pin_out, maps):
        b.free()
