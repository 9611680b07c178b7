"""Synchronous single-frame call (b200vc_convert, GstBaseTransform's default mode): frames/s by how the
frame is pipelined inside the call — event-chained bands (the default) or row-gated — and by the number of pieces; page-locked buffers, planes stored by the kernels. Also batches of 8 and the one-frame-latency mode.
  python tools/sync_latency.py [WxH]"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "docker-nvidia-glx-desktop_b200")]
import b200vc
W, H = (int(v) for v in sys.argv[1].split("x")) if len(sys.argv) > 1 else (1920, 1080)
total, off, pit = b200vc.default_layout(0, W, H)
h_in, h_out = b200vc.PinnedBuffer(8 * W * H * 4), b200vc.PinnedBuffer(total)
h_in.array[:] = 77
planes = (C.c_void_p * 3)(*[h_out.ptr + o for o in off]); pits = (C.c_size_t * 3)(*pit)
conv = b200vc.Converter(0)
NAMES = {0: "bands chained by events, kernel stores", b200vc.FLAG_ROW_GATE: "row gate (flags by stream write)",
         b200vc.FLAG_ZERO_COPY: "kernel reads and writes host"}
for flags in NAMES:
    for bands in (1, 2, 3, 4, 5, 6, 8, 16, 32):
        conv.set_info(W, H, matrix=0, range_=0, chroma_site=1, n_bands=bands, flags=flags)
        for i in range(20):
            conv.lib.b200vc_convert(conv.ctx, h_in.ptr + (i % 8) * W * H * 4, W * 4, planes, pits, None)
        n = 300
        t0 = time.perf_counter()
        for i in range(n):
            conv.lib.b200vc_convert(conv.ctx, h_in.ptr + (i % 8) * W * H * 4, W * 4, planes, pits, None)
        dt = time.perf_counter() - t0
        print(f"{NAMES[flags]:40s} chunks={bands:2d}: {n / dt:8.1f} frames/s  {dt / n * 1e6:7.1f} us/frame", flush=True)
        if flags == b200vc.FLAG_ZERO_COPY:
            break
# batch of 8 sessions, copy engines vs kernels storing to host
for flags in (0, b200vc.FLAG_ZERO_COPY, b200vc.FLAG_ZERO_COPY_OUT):
    conv.set_info(W, H, matrix=0, range_=0, chroma_site=1, max_sessions=8, n_inflight=8, flags=flags)
    outs = b200vc.PinnedBuffer(8 * total)
    frames = (b200vc.Frame * 8)(*[conv.make_frame(h_in.ptr + s * W * H * 4, W * 4, outs.ptr + s * total, off, pit, session=s) for s in range(8)])
    for _ in range(5):
        conv.lib.b200vc_convert_batch(conv.ctx, frames, 8)
    n = 100
    t0 = time.perf_counter()
    for _ in range(n):
        conv.lib.b200vc_convert_batch(conv.ctx, frames, 8)
    dt = time.perf_counter() - t0
    print(f"batch of 8, flags={flags:2d}: {8 * n / dt:8.1f} frames/s", flush=True)
    conv.stop()
    outs.free()
# one session, one frame of latency (the element's one-frame-latency mode): submit N, collect N-1
conv = b200vc.Converter(0)
conv.set_info(W, H, matrix=0, range_=0, chroma_site=1, n_inflight=2)
outs = b200vc.PinnedBuffer(2 * total)
fr = [conv.make_frame(h_in.ptr + k * W * H * 4, W * 4, outs.ptr + (k % 2) * total, off, pit, session=0) for k in range(8)]
prev = conv.submit(fr[0])
n = 400
t0 = time.perf_counter()
for i in range(1, n + 1):
    t = conv.submit(fr[i % 8])
    conv.collect(prev)
    prev = t
conv.collect(prev)
dt = time.perf_counter() - t0
print(f"one session, submit N / collect N-1: {n / dt:8.1f} frames/s  {dt / n * 1e6:7.1f} us/frame", flush=True)
conv.stop()
