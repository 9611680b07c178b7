"""Kernel-only timing of every BASELINE.json config (C1..C4) through b200vc_convert_device:
frames resident in HBM as a ring larger than L2, CUDA events on the launching stream.
Prints one JSON line per config; used for tuning and for the tables in DESIGN.md."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "docker-nvidia-glx-desktop_b200")]
import torch  # noqa: E402

import b200vc  # noqa: E402
from bench import ClockSampler  # noqa: E402  (every line carries the SM clock and throttle reasons sampled while it ran)

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0

CONFIGS = [
    # name, in_w, in_h, out_w, out_h, fmt, matrix, range, site, flags, damage
    ("C1 1080p I420 601L mpeg2", 1920, 1080, 1920, 1080, 0, 0, 0, 1, 0, False),
    ("C1 1080p I420 601L jpeg", 1920, 1080, 1920, 1080, 0, 0, 0, 0, 0, False),
    ("C2 1440p NV12 709F mpeg2", 2560, 1440, 2560, 1440, 1, 1, 1, 1, 0, False),
    ("C3 4K->1080p I420 709L fused", 3840, 2160, 1920, 1080, 0, 1, 0, 1, 0, False),
    ("C3 4K->1080p I420 709L two-pass", 3840, 2160, 1920, 1080, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("C4 1080p I420 601L mpeg2 +damage", 1920, 1080, 1920, 1080, 0, 0, 0, 1, b200vc.FLAG_DAMAGE_MAP, True),
    # ratios other than 2:1 run the general fused kernel
    ("G 4K->1080p generic fused", 3840, 2160, 1920, 1080, 0, 1, 0, 1, b200vc.FLAG_FORCE_GENERIC_FUSED, False),
    ("G 4K->1440p 3:2", 3840, 2160, 2560, 1440, 0, 1, 0, 1, 0, False),
    ("G 1440p->1080p 4:3", 2560, 1440, 1920, 1080, 0, 1, 0, 1, 0, False),
    ("G 1080p->720p 3:2", 1920, 1080, 1280, 720, 0, 1, 0, 1, 0, False),
    ("G 4K->720p 3:1", 3840, 2160, 1280, 720, 0, 1, 0, 1, 0, False),
    ("G 1080p->1536x864 1.25:1", 1920, 1080, 1536, 864, 0, 1, 0, 1, 0, False),
    ("G 720p->1080p 2:3 upscale", 1280, 720, 1920, 1080, 0, 1, 0, 1, 0, False),
    # 5-6-5 input (CDEPTH=16): 2 bytes per pixel, unpacked in the general fused kernel (name ends in "rgb16")
    ("G 1440p->1080p 4:3 rgb16", 2560, 1440, 1920, 1080, 0, 1, 0, 1, 0, False),
    ("G 1080p->720p 3:2 rgb16", 1920, 1080, 1280, 720, 0, 1, 0, 1, 0, False),
    ("T 1440p->1080p 4:3 rgb16 two-pass", 2560, 1440, 1920, 1080, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("T 4K->1440p 3:2 two-pass", 3840, 2160, 2560, 1440, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("T 1440p->1080p 4:3 two-pass", 2560, 1440, 1920, 1080, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("T 1080p->720p 3:2 two-pass", 1920, 1080, 1280, 720, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("T 4K->720p 3:1 two-pass", 3840, 2160, 1280, 720, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("T 1080p->1536x864 1.25:1 two-pass", 1920, 1080, 1536, 864, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
    ("T 720p->1080p 2:3 upscale two-pass", 1280, 720, 1920, 1080, 0, 1, 0, 1, b200vc.FLAG_FORCE_TWO_PASS, False),
]


def e2e(only):
    """End to end through b200vc_convert_batch (8 sessions per call, pinned host buffers) per config."""
    import ctypes as C
    import time
    conv = b200vc.Converter(0)
    sampler = ClockSampler(0)
    sampler.start()
    for name, iw, ih, ow, oh, fmt, m, rg, site, flags, damage in CONFIGS:
        if (only and only not in name) or "two-pass" in name or name.startswith("G "):
            continue
        ns = 8
        in_b, (total, off, pit) = iw * ih * 4, b200vc.default_layout(fmt, ow, oh)
        mb = ((iw + 15) // 16) * ((ih + 15) // 16)
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, flags=flags, max_sessions=ns, n_inflight=ns)
        h_in, h_out = b200vc.PinnedBuffer(2 * ns * in_b), b200vc.PinnedBuffer(ns * (total + mb))
        h_in.array[:] = 99
        h_in.array[ns * in_b::977] = 7                    # the second set of frames differs a little
        ticks = []
        for t in range(2):
            fr = [conv.make_frame(h_in.ptr + (t * ns + s) * in_b, iw * 4, h_out.ptr + s * (total + mb), off, pit, session=s,
                                  damage_ptr=(h_out.ptr + s * (total + mb) + total) if damage else None) for s in range(ns)]
            ticks.append((b200vc.Frame * ns)(*fr))
        for k in range(6):
            conv.lib.b200vc_convert_batch(conv.ctx, ticks[k & 1], ns)
        n = max(10, int(2e9 // (ns * in_b)))
        sampler.mark(name)
        t0 = time.perf_counter()
        for k in range(n):
            rc = conv.lib.b200vc_convert_batch(conv.ctx, ticks[k & 1], ns)
            assert rc == 0
        dt = time.perf_counter() - t0
        sampler.mark(None)
        fps = n * ns / dt
        print(json.dumps({"config": name, "e2e_fps": round(fps, 1), "h2d_GBs": round(fps * in_b / 1e9, 2),
                          "d2h_GBs": round(fps * total / 1e9, 2), "frac_of_pcie_gen5_x16_raw_63GBs": round(fps * in_b / 63.02e9, 4),
                          "clocks": sampler.record(name)}), flush=True)
        conv.stop()
        h_in.free()
        h_out.free()


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "e2e":
        return e2e(sys.argv[2] if len(sys.argv) > 2 else None)
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    only = sys.argv[2] if len(sys.argv) > 2 else None
    conv = b200vc.Converter(0)
    stream = torch.cuda.Stream()
    sampler = ClockSampler(0)
    sampler.start()
    for name, iw, ih, ow, oh, fmt, m, rg, site, flags, damage in CONFIGS:
        if only and only not in name:
            continue
        bpp, in_format = (2, b200vc.IN_RGB16) if "rgb16" in name else (4, b200vc.IN_BGRX)
        in_b, (total, off, pit) = iw * ih * bpp, b200vc.default_layout(fmt, ow, oh)
        n = max(8, min(64, int(600e6 // in_b)))                 # ring >= ~600 MB of input
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, flags=flags, max_sessions=8, in_format=in_format)
        d_in = torch.randint(0, 256, (n, in_b), dtype=torch.uint8, device="cuda")
        d_out = torch.zeros((n, total), dtype=torch.uint8, device="cuda")
        mb = ((iw + 15) // 16) * ((ih + 15) // 16)
        d_dmg = torch.zeros((n, mb), dtype=torch.uint8, device="cuda")
        df = b200vc.DeviceFrames()
        df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * bpp, in_b
        for i in range(3):
            df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
        df.out_frame_stride, df.n_frames = total, n
        algo = in_b + total
        if damage:
            d_prev = torch.roll(d_in, 1, 0).contiguous()
            df.prev, df.prev_pitch, df.prev_frame_stride = d_prev.data_ptr(), iw * bpp, in_b
            df.damage, df.damage_frame_stride = d_dmg.data_ptr(), mb
            algo += in_b + mb
        with torch.cuda.stream(stream):
            for _ in range(5):
                conv.convert_device(df, stream.cuda_stream)
            stream.synchronize()
            sampler.mark(name)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                conv.convert_device(df, stream.cuda_stream)
            e1.record(stream)
        stream.synchronize()
        sampler.mark(None)
        sec = e0.elapsed_time(e1) / 1e3 / reps
        # single-frame launch (L2-resident, launch-latency bound), reported separately
        df.n_frames = 1
        with torch.cuda.stream(stream):
            for _ in range(20):
                conv.convert_device(df, stream.cuda_stream)
            e0.record(stream)
            for _ in range(200):
                conv.convert_device(df, stream.cuda_stream)
            e1.record(stream)
        stream.synchronize()
        one = e0.elapsed_time(e1) / 1e3 / 200
        # one launch per frame, 1024 launches back to back over the ring (HBM traffic, launch overhead included)
        per = None
        if not damage:
            frames_df = []
            for i in range(n):
                d1 = b200vc.DeviceFrames()
                d1.bgrx, d1.in_pitch, d1.in_frame_stride = d_in.data_ptr() + i * in_b, iw * bpp, in_b
                for k in range(3):
                    d1.planes[k], d1.pitches[k] = d_out.data_ptr() + i * total + off[k], pit[k]
                d1.out_frame_stride, d1.n_frames = total, 1
                frames_df.append(d1)
            with torch.cuda.stream(stream):
                for i in range(64):
                    conv.convert_device(frames_df[i % n], stream.cuda_stream)
                e0.record(stream)
                for i in range(1024):
                    conv.convert_device(frames_df[i % n], stream.cuda_stream)
                e1.record(stream)
            stream.synchronize()
            per = e0.elapsed_time(e1) / 1e3 / 1024
        inf = conv.info()
        print(json.dumps({"config": name, "per_frame_launch_ring_us": round(per * 1e6, 2) if per else None,
                          "per_frame_launch_ring_fps": round(1 / per, 1) if per else None, "ring_frames": n, "fps": round(n / sec, 1), "ms_per_launch": round(sec * 1e3, 4),
                          "algo_GBs": round(algo * n / sec / 1e9, 1), "frac_of_measured_peak": round(algo * n / sec / 1e9 / PEAK, 4),
                          "single_frame_us": round(one * 1e6, 2), "path": inf.path, "box": [inf.box_w, inf.box_h],
                          "tile_h": inf.tile_h, "clocks": sampler.record(name)}), flush=True)
        del d_in, d_out, d_dmg
        torch.cuda.empty_cache()
    sampler.finish()
    conv.stop()


if __name__ == "__main__":
    main()
