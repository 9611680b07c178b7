#!/usr/bin/env python
"""bench.py — 1080p BGRx -> I420 frames/s on B200 (BASELINE.json metric), one JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config C1 of BASELINE.json, the configuration the metric is quoted on): 1920x1080 BGRx ->
I420, BT.601 limited range, chroma-site as negotiated for 1080 lines (mpeg2 / h-cosited), synthetic
uniformly random frames. A step is one pass of the hot path over one batch of 512 frames = 64 ticks
of 8 desktop sessions (the per-GPU share of config C5).

  value     kernel-only: the 512 frames are resident in HBM as a ring of 64 distinct frames
            (531 MB in + 199 MB out, several times the 126 MB L2) and are converted by 8 batched
            launches of b200vc_convert_device per step; CUDA events on the launching stream.
  e2e       the same 512 frames per step through the host-buffer C-ABI call the GStreamer element
            makes per tick (b200vc_convert_batch, 8 sessions), from pinned host memory to pinned host
            memory, PCIe copies inside the timed region.
  roofline  HBM: algorithmic bytes (8 294 400 read + 3 110 400 written per frame, SURVEY §8 d) per
            launch / average launch duration, against MEASURED_PEAKS.json's copy bandwidth.
  cpu_baseline  the CPU oracle (a port: this repo's C restatement of videoconvert — GStreamer is
            not available on this box) timed on the host cores on a bounded sample, rank 0, N=1.

--impl reference times that same CPU restatement with every host thread it can use (the
reference's own GStreamer/ORC build cannot be installed here: it is an unpinned binary download,
/root/reference Dockerfile:467-471).

Multi-GPU: one process per GPU under torchrun, sessions partitioned by GPU, no data-path
collective (SURVEY §8 e); torch.distributed is used only for the barrier and the max-over-ranks.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "docker-nvidia-glx-desktop_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

W, H = 1920, 1080
IN_BYTES = W * H * 4                     # 8 294 400
OUT_BYTES = W * H * 3 // 2               # 3 110 400
ALGO_BYTES_PER_FRAME = IN_BYTES + OUT_BYTES
RING = 64                                # distinct frames resident in HBM
SESSIONS = 8                             # sessions per GPU per tick (config C5)
TICKS = 64                               # ticks per step
FRAMES_PER_STEP = SESSIONS * TICKS       # 512
LAUNCHES_PER_STEP = FRAMES_PER_STEP // RING
MATRIX, RANGE, SITE, FMT = 0, 0, 1, 0    # BT.601 limited, mpeg2 siting (1080-line default), I420
METRIC = "1080p BGRx->I420 frames/s, aggregate over GPUs"
WORKLOAD = "C1: 1920x1080 BGRx->I420 BT.601 limited, chroma-site mpeg2, random frames"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def committed_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get("k1_convert_bytes_per_launch")
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
                 0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = get_reasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"], "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# CPU arm (the oracle; test infrastructure used here only as the timed CPU baseline)
# ------------------------------------------------------------------------------------------------
def cpu_arm(threads: int, frames_per_step: int, steps: int, warmup: int, split: str = "bands"):
    """split="bands": one frame at a time, cut into `threads` horizontal bands (what videoconvert's
    n-threads does). split="frames": `threads` workers convert different frames at once (the most a
    multi-session host can get out of its cores)."""
    import numpy as np

    import oracle
    orc = oracle.Oracle(native=True)            # -march=native build for the box this runs on
    ring = [orc.synth("rand", W, H, i) for i in range(4)]
    total, off, pit = orc.default_layout(FMT, W, H)
    workers = threads if split == "frames" else 1
    outs, ptrs, pits = [], [], []
    for _ in range(workers):
        o = np.zeros(total, np.uint8)
        outs.append(o)
        ptrs.append((C.c_void_p * 3)(*[o.ctypes.data + x for x in off]))
        pits.append((C.c_size_t * 3)(*pit))

    def work(wid: int, n: int, band_threads: int):
        for i in range(n):
            f = ring[(i + wid) % len(ring)]
            rc = orc.convert_into(f.ctypes.data, W * 4, W, H, W, H, FMT, MATRIX, RANGE, SITE, 0, ptrs[wid], pits[wid],
                                  band_threads)
            assert rc == 0

    def step():
        if workers == 1:
            work(0, frames_per_step, threads)
            return
        share = [frames_per_step // workers + (1 if k < frames_per_step % workers else 0) for k in range(workers)]
        ts = [threading.Thread(target=work, args=(k, share[k], 1)) for k in range(workers)]   # ctypes drops the GIL
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    return frames_per_step * steps / dt, dt


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, rank: int):
    if rank != 0:
        return
    cores = host_cores()
    fps, dt = cpu_arm(cores, 2 * cores, args.steps, args.warmup, split="frames")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(fps, 2), "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": 2 * cores},
        "cpu_baseline": {"value": round(fps, 2), "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": f"{2 * cores * args.steps} frames ({2 * cores} per step) of the same 1080p workload, "
                                   f"{cores} threads each converting whole frames; CPU restatement of videoconvert, "
                                   "not GStreamer/ORC (unavailable on this box)"},
        "e2e": {"value": round(fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_gpu(args, rank: int, local_rank: int, world: int):
    import numpy as np
    import torch

    import b200vc

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback")
    import sessions
    # before any pinned allocation: put this rank on the cores (and so the memory) next to its GPU
    numa = sessions.bind_to_gpu_numa(local_rank) if os.environ.get("B200VC_NUMA_BIND", "1") != "0" else {"bound": False, "off": True}
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        # NCCL announces its version on stdout when the communicator comes up; keep stdout for the
        # one JSON line by pointing fd 1 at stderr until the first collective has run
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(seconds: float) -> float:
        return sessions.job_throughput(FRAMES_PER_STEP * args.steps, seconds, dist, "cuda")[1]

    conv = b200vc.Converter(local_rank)
    conv.set_info(W, H, out_format=FMT, matrix=MATRIX, range_=RANGE, chroma_site=SITE, max_sessions=SESSIONS,
                  n_inflight=SESSIONS)
    total, off, pit = b200vc.default_layout(FMT, W, H)
    assert total == OUT_BYTES

    # ---- synthetic input: 64 distinct random frames, generated on the device -------------------
    gen = torch.Generator(device="cuda")
    gen.manual_seed(0x9E3779B9 + rank)
    d_in = torch.randint(0, 256, (RING, IN_BYTES), dtype=torch.uint8, device="cuda", generator=gen)
    d_out = torch.zeros((RING, OUT_BYTES), dtype=torch.uint8, device="cuda")
    df = b200vc.DeviceFrames()
    df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), W * 4, IN_BYTES
    for i in range(3):
        df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
    df.out_frame_stride, df.n_frames = OUT_BYTES, RING
    stream = torch.cuda.Stream()

    def kernel_step():
        for _ in range(LAUNCHES_PER_STEP):
            conv.convert_device(df, stream.cuda_stream)

    # ---- host side for e2e: pinned ring in, pinned ring out -------------------------------------
    h_in = b200vc.PinnedBuffer(RING * IN_BYTES)
    h_out = b200vc.PinnedBuffer(RING * OUT_BYTES)
    torch.cuda.synchronize()
    h_in_t = torch.from_numpy(h_in.array)
    h_in_t.copy_(d_in.reshape(-1))                 # same frames as the device ring
    torch.cuda.synchronize()
    ticks = []
    for t in range(RING // SESSIONS):
        frames = []
        for s in range(SESSIONS):
            k = t * SESSIONS + s
            frames.append(conv.make_frame(h_in.ptr + k * IN_BYTES, W * 4, h_out.ptr + k * OUT_BYTES, off, pit, session=s))
        ticks.append((b200vc.Frame * SESSIONS)(*frames))

    def e2e_step():
        for t in range(TICKS):
            rc = conv.lib.b200vc_convert_batch(conv.ctx, ticks[t % len(ticks)], SESSIONS)
            if rc:
                conv._check(rc)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- kernel-only timed region ------------------------------------------------------------------
    with torch.cuda.stream(stream):
        for _ in range(max(args.warmup, 3)):
            kernel_step()
    barrier()
    launches0 = conv.stats().kernel_launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.steps):
            kernel_step()
        ev1.record(stream)
    barrier()
    k_seconds = max_over_ranks(ev0.elapsed_time(ev1) / 1e3)
    gpu_launches = int(conv.stats().kernel_launches - launches0)

    # ---- end-to-end timed region ---------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        e2e_step()
    barrier()
    st0 = conv.stats()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e_seconds_local = time.perf_counter() - t0
    st1 = conv.stats()
    barrier()
    e_seconds = max_over_ranks(e_seconds_local)

    # output sanity: kernel-only and e2e outputs of the last frame of the ring must agree byte for byte
    same = bool(np.array_equal(d_out[RING - 1].cpu().numpy(), h_out.array[(RING - 1) * OUT_BYTES:RING * OUT_BYTES]))

    # synchronous single-session call (GstBaseTransform's default mode), for the record
    planes = (C.c_void_p * 3)(*[h_out.ptr + o for o in off])
    pits = (C.c_size_t * 3)(*pit)
    for i in range(8):
        conv.lib.b200vc_convert(conv.ctx, h_in.ptr + (i % RING) * IN_BYTES, W * 4, planes, pits, None)
    n_sync = 256
    t0 = time.perf_counter()
    for i in range(n_sync):
        conv.lib.b200vc_convert(conv.ctx, h_in.ptr + (i % RING) * IN_BYTES, W * 4, planes, pits, None)
    sync_fps = n_sync / (time.perf_counter() - t0)
    # the same session with one frame of latency (submit N, collect N-1: the element's one-frame-latency mode)
    lat_frames = [conv.make_frame(h_in.ptr + i * IN_BYTES, W * 4, h_out.ptr + i * OUT_BYTES, off, pit, session=0) for i in range(RING)]
    prev_ticket = conv.submit(lat_frames[0])
    t0 = time.perf_counter()
    for i in range(1, n_sync + 1):
        ticket = conv.submit(lat_frames[i % RING])
        conv.collect(prev_ticket)
        prev_ticket = ticket
    conv.collect(prev_ticket)
    latency_mode_fps = n_sync / (time.perf_counter() - t0)

    # PCIe ceiling for the e2e number: raw pinned copies of the same frames, both directions at once,
    # no kernel (what the link gives for this traffic pattern on this box)
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    h_out_t = torch.from_numpy(h_out.array)
    d_in_flat, d_out_flat = d_in.reshape(-1), d_out.reshape(-1)

    def raw_copies(n):
        for k in range(n):
            i = k % RING
            with torch.cuda.stream(s_up):
                d_in_flat[i * IN_BYTES:(i + 1) * IN_BYTES].copy_(h_in_t[i * IN_BYTES:(i + 1) * IN_BYTES], non_blocking=True)
            with torch.cuda.stream(s_dn):
                h_out_t[i * OUT_BYTES:(i + 1) * OUT_BYTES].copy_(d_out_flat[i * OUT_BYTES:(i + 1) * OUT_BYTES], non_blocking=True)

    raw_copies(32)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    raw_copies(256)
    torch.cuda.synchronize()
    pcie_fps = 256 / (time.perf_counter() - t0)

    clocks = sampler.finish() if rank == 0 else None


    if rank == 0:
        peak, peak_src = measured_peak()
        frames = FRAMES_PER_STEP * args.steps
        value = world * frames / k_seconds
        launch_s = k_seconds / (args.steps * LAUNCHES_PER_STEP)
        achieved = ALGO_BYTES_PER_FRAME * RING / launch_s / 1e9
        e2e_fps = world * frames / e_seconds
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(k_seconds / args.steps * 1e3, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "frames_per_step": FRAMES_PER_STEP, "sessions_per_gpu": SESSIONS,
                       "ring_frames": RING, "launches_per_step": LAUNCHES_PER_STEP, "partition": "sessions by GPU, no collective",
                       "l2": f"inputs larger than L2: ring of {RING} frames = {RING * ALGO_BYTES_PER_FRAME / 1e6:.0f} MB in+out per launch"},
            "e2e": {"value": round(e2e_fps, 1), "unit": "frames/s",
                    "h2d_bytes_per_step": int((st1.h2d_bytes - st0.h2d_bytes) // args.steps),
                    "d2h_bytes_per_step": int((st1.d2h_bytes - st0.d2h_bytes) // args.steps),
                    "api": "b200vc_convert_batch, 8 sessions per call, pinned host buffers",
                    "pcie_h2d_gbs": round(e2e_fps / world * IN_BYTES / 1e9, 2),
                    "pcie_ceiling_fps_per_gpu": round(pcie_fps, 1),
                    "frac_of_pcie_ceiling": round(e2e_fps / world / pcie_fps, 4),
                    "frac_of_pcie_gen5_x16_raw_63GBs": round(e2e_fps / world * IN_BYTES / 63.02e9, 4),
                    "pcie_ceiling_how": "raw pinned H2D 8.3 MB + D2H 3.1 MB copies in flight together, no kernel, rank 0",
                    "sync_single_session_fps": round(sync_fps, 1),
                    "one_frame_latency_single_session_fps": round(latency_mode_fps, 1), "numa": numa},
            "gpu_launches": gpu_launches,
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": committed_traffic(),
                         "peak_source": peak_src, "kernel": "k1_convert_kernel<I420,mpeg2>",
                         "algorithmic_bytes_per_launch": ALGO_BYTES_PER_FRAME * RING,
                         "launch_ms": round(launch_s * 1e3, 4), "frac_of_nominal_8TBs": round(achieved / 8000.0, 4)},
            "clocks": clocks, "outputs_agree": same,
        }
        if world == 1:
            cores = host_cores()
            t_sel = min(4, max(1, cores - 1))          # selkies' n-threads rule
            fps1, dt1 = cpu_arm(1, 64, 8, 1)            # 512 frames on one thread  (~3 s)
            fpsT, dtT = cpu_arm(t_sel, 128, 8, 1)       # 1024 frames on T band threads (~2 s wall, ~7 CPU-s)
            line["cpu_baseline"] = {
                "value": round(fpsT, 2), "unit": "frames/s", "cores": t_sel, "kind": "port",
                "sample": f"1024 frames of the same 1080p workload in {dtT:.1f} s, {t_sel} threads over horizontal bands "
                          f"(selkies n-threads rule) on a {cores}-core host; 1 thread: 512 frames in {dt1:.1f} s = {fps1:.1f} frames/s. "
                          "CPU restatement of videoconvert, not GStreamer/ORC (unavailable on this box)",
                "host_cores": cores, "one_thread_fps": round(fps1, 2)}
        print(json.dumps(line), flush=True)

    conv.stop()
    h_in.free()
    h_out.free()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # launched by hand: start one rank per GPU the way the driver does
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.abspath(__file__), "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup)]
        raise SystemExit(subprocess.call(cmd))
    run_gpu(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
