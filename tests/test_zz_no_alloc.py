"""GPU test of b200vc_convert_scale_device_static, the convert + scale entry point for callers that
must not allocate memory.

Such callers exist: a streaming thread that has to follow a caps change in mid-stream cannot call
cudaMalloc or cudaFree, because both wait for everything in flight on the device and would stall the
other desktop sessions sharing the GPU; and a process that handed all GPU memory to another pool at
start-up has nothing left to allocate from. For them the intermediate frame of the two-pass scaler
sits in a fixed array inside the library (3 133 440 bytes, sources up to 1920x1088), and the entry
point takes the negotiated caps directly instead of a context.

The check: through the C ABI, its output must equal, byte for byte (the tolerance every parity test
here uses), what the existing context-based entry point b200vc_convert_device writes for the same
device buffers, and what the CPU oracle computes for the same frames — on calls in a row with
different geometry, so that whatever one call left in the shared array would show up in the next.
Where the box has a second GPU the same runs there too: every device holds its own copy of the
array, whose address the library looks up on the first call made on that device. The file name sorts last on purpose: a run
that stops at the first failure has then already run every other test. It passes on its own in a
fresh process as well as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import b200vc
from helpers import assert_planes_equal, plane_views

pytestmark = pytest.mark.gpu


def _device_frames(d_in, d_out, iw, ih, ow, oh, fmt, n):
    total, off, pit = b200vc.default_layout(fmt, ow, oh)
    df = b200vc.DeviceFrames()
    df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * 4, iw * ih * 4
    for i in range(2 if fmt == 1 else 3):
        df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
    df.out_frame_stride, df.n_frames = total, n
    return df


def _static(lib, cfg, df, stream=0):
    err = C.create_string_buffer(512)
    rc = lib.b200vc_convert_scale_device_static(C.byref(cfg), C.byref(df), stream, err, len(err))
    return rc, err.value.decode()


def _devices():
    import torch
    return [0, 1] if torch.cuda.device_count() > 1 else [0]


@pytest.mark.parametrize("dev", [0, 1])
def test_static_scratch_entry_point_matches_context_entry_point(lib, orc, dev):
    import torch
    if dev not in _devices():
        pytest.skip("one GPU on this box")
    torch.cuda.set_device(dev)
    gen = torch.Generator(device=f"cuda:{dev}")
    gen.manual_seed(1234)
    conv = b200vc.Converter(dev)
    # calls in a row, different geometry, format and caps; the second is larger than the first
    cases = [(1280, 720, 640, 360, 0, 1, 0, 1, 0, 2),      # I420 bt709 limited mpeg2, 2 frames
             (1920, 1080, 1280, 720, 1, 0, 1, 0, 0, 1),    # NV12 bt601 full jpeg, 3:2
             (640, 360, 320, 180, 0, 1, 0, 1, 1, 1),       # BOX_2X2, smaller again: stale bytes beyond it
             (960, 540, 1280, 720, 0, 2, 0, 1, 0, 1),      # bt2020 limited, upscale
             (1280, 720, 848, 480, 1, 2, 1, 0, 0, 1)]      # bt2020 full (the negated-matrix kernels), NV12
    # second round over the same cases: every kernel image is loaded by then, so the device's free
    # memory must not move across a call
    for rnd, (iw, ih, ow, oh, fmt, m, rg, site, mode, n) in [(r, c) for r in (0, 1) for c in cases]:
        total = b200vc.default_layout(fmt, ow, oh)[0]
        d_in = torch.randint(0, 256, (n, iw * ih * 4), dtype=torch.uint8, device=f"cuda:{dev}", generator=gen)
        want = torch.zeros((n, total), dtype=torch.uint8, device=f"cuda:{dev}")
        got = torch.full((n, total), 0xA5, dtype=torch.uint8, device=f"cuda:{dev}")
        # existing entry point (fused kernel where it applies) on a context
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, scale_mode=mode)
        conv.convert_device(_device_frames(d_in, want, iw, ih, ow, oh, fmt, n), 0)
        torch.cuda.synchronize()
        # new entry point: no context, nothing allocated between the two memory readings
        cfg = b200vc.Config(dev, iw, ih, ow, oh, fmt, m, rg, site, mode, 0, 0, 0, 0)
        df = _device_frames(d_in, got, iw, ih, ow, oh, fmt, n)
        free_before = torch.cuda.mem_get_info()[0]
        rc, msg = _static(lib, cfg, df)
        torch.cuda.synchronize()
        assert rc == b200vc.OK, msg
        if rnd == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, "the entry point changed the device's free memory"
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert np.array_equal(a, b), f"{iw}x{ih}->{ow}x{oh}: {int((a != b).sum())} bytes differ from b200vc_convert_device"
        if rnd == 0:                 # and against the oracle itself, frame by frame
            _, off, pit = b200vc.default_layout(fmt, ow, oh)
            host_in = d_in.cpu().numpy()
            for k in range(n):
                obuf, ooff, opit = orc.convert(host_in[k].reshape(ih, iw, 4), fmt, m, rg, site, out_size=(ow, oh), scale_mode=mode)
                assert_planes_equal(plane_views(a[k], off, pit, fmt, ow, oh), plane_views(obuf, ooff, opit, fmt, ow, oh),
                                    f"device {dev} {iw}x{ih}->{ow}x{oh} frame {k} vs oracle")
    conv.stop()
    torch.cuda.set_device(0)


def test_static_scratch_limits_and_errors(lib):
    import torch
    torch.cuda.set_device(0)
    d_in = torch.zeros(64, dtype=torch.uint8, device="cuda")
    d_out = torch.zeros(64, dtype=torch.uint8, device="cuda")
    # a 4K source needs a 12.4 MB intermediate: refused with the limit spelled out
    cfg = b200vc.Config(0, 3840, 2160, 1920, 1080, 0, 1, 0, 1, 0, 0, 0, 0, 0)
    rc, msg = _static(lib, cfg, _device_frames(d_in, d_out, 3840, 2160, 1920, 1080, 0, 1))
    assert rc == b200vc.ERR_UNSUPPORTED and "3133440" in msg and "1920x1088" in msg
    # the largest source that fits is accepted by the size check (1920x1088 exactly fills the array)
    assert 1920 * 1088 * 3 // 2 == b200vc.STATIC_SCRATCH_BYTES
    # same size in and out: nothing to scale, no scratch involved
    cfg = b200vc.Config(0, 640, 360, 640, 360, 0, 1, 0, 1, 0, 0, 0, 0, 0)
    rc, msg = _static(lib, cfg, _device_frames(d_in, d_out, 640, 360, 640, 360, 0, 1))
    assert rc == b200vc.ERR_UNSUPPORTED and "b200vc_convert_device" in msg
    # wrong device named in the caps
    cfg = b200vc.Config(7, 640, 360, 320, 180, 0, 1, 0, 1, 0, 0, 0, 0, 0)
    rc, msg = _static(lib, cfg, _device_frames(d_in, d_out, 640, 360, 320, 180, 0, 1))
    assert rc == b200vc.ERR_INVALID_ARG and "current device" in msg
    # odd sizes are scaled by the context-based path only
    cfg = b200vc.Config(0, 641, 360, 320, 180, 0, 1, 0, 1, 0, 0, 0, 0, 0)
    rc, msg = _static(lib, cfg, _device_frames(d_in, d_out, 641, 360, 320, 180, 0, 1))
    assert rc == b200vc.ERR_UNSUPPORTED and msg
    assert lib.b200vc_convert_scale_device_static(None, None, None, None, 0) == b200vc.ERR_INVALID_ARG
