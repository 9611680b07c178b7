#!/usr/bin/env python
"""bench.py — 1080p BGRx -> I420 frames/s on B200 (BASELINE.json metric), one JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Headline workload (config C1 of BASELINE.json, the configuration the metric is quoted on): 1920x1080
BGRx -> I420, BT.601 limited range, chroma-site as negotiated for 1080 lines (mpeg2 / h-cosited),
synthetic uniformly random frames. A step is one pass of the hot path over one batch of 512 frames =
64 ticks of 8 desktop sessions (the per-GPU share of config C5).

  value     kernel-only: the 512 frames are resident in HBM as a ring of 64 distinct frames
            (531 MB in + 199 MB out, several times the 126 MB L2) and are converted by 8 batched
            launches of b200vc_convert_device per step; CUDA events on the launching stream.
  e2e       the same 512 frames per step through the host-buffer C-ABI call the GStreamer element
            makes per tick (b200vc_convert_batch, 8 sessions), from pinned host memory to pinned host
            memory, PCIe copies inside the timed region. Beside it: the link's own ceilings (raw pinned
            copies, H2D alone, D2H alone, both at once; all ranks copying between the same barriers), the
            host CPU seconds the region cost, and the same loop run by the native harness (no interpreter).
  roofline  HBM: algorithmic bytes (8 294 400 read + 3 110 400 written per frame, SURVEY §8 d) per
            launch / average launch duration, against MEASURED_PEAKS.json's copy bandwidth.
  configs   BASELINE.json's other single-GPU configs, each measured the same way in the same run with its
            own clock record: C2 2560x1440 -> NV12 BT.709 full, C3 3840x2160 -> 1920x1080 I420 (fused
            bilinear downscale), C4 = C1 + 16x16 damage map on mostly static frames (about 3 % dirty).
  c5        config C5 in real time: 8 sessions per GPU paced at 60 Hz through the C ABI by the native
            harness (harness/session_bench, one process per GPU, all started together), then the largest
            number of sessions per GPU that still misses no tick.
  cpu_baseline  the CPU oracle (a port: this repo's C restatement of videoconvert — GStreamer is
            not available on this box) timed on the host cores on a bounded sample, rank 0, N=1.

--impl reference times that same CPU restatement with every host thread it can use (the
reference's own GStreamer/ORC build cannot be installed here: it is an unpinned binary download,
/root/reference Dockerfile:467-471).

Multi-GPU: one process per GPU under torchrun, sessions partitioned by GPU, no data-path
collective (SURVEY §8 e); torch.distributed is used only for barriers and for max/sum over ranks.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
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
SESSION_BENCH = os.path.join(PKG, "harness", "session_bench")

# BASELINE.json configs[1..3]: name -> (in_w, in_h, out_w, out_h, fmt, matrix, range, site, damage, kernel, workload)
CONFIGS = {
    "C2": (2560, 1440, 2560, 1440, 1, 1, 1, 1, False, "k1_convert_kernel<NV12,mpeg2,neg>",
           "C2: 2560x1440 BGRx->NV12 BT.709 full range, chroma-site mpeg2, random frames"),
    "C3": (3840, 2160, 1920, 1080, 0, 1, 0, 1, False, "k2_fused_2x_kernel<I420,mpeg2>",
           "C3: 3840x2160 BGRx->I420 1920x1080, bilinear downscale fused into the conversion, BT.709 limited, random frames"),
    "C4": (1920, 1080, 1920, 1080, 0, 0, 0, 1, True, "k1_convert_kernel<I420,mpeg2,damage>",
           "C4: C1 + 16x16 damage map; static random desktop with a moving 320x180 noise window and a 16x24 cursor (about 3 % dirty macroblocks)"),
}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def committed_traffic():
    """dram bytes per launch of the dominant kernel: not measured by this run, read from the committed
    ncu --set full capture (profiles/roofline_traffic.json names the file)."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            d = json.load(f)
            return d.get("k1_convert_bytes_per_launch"), "committed ncu capture, not this run: " + d.get("source", "profiles/roofline_traffic.json")
    except Exception:
        return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while timed regions run. mark(name) closes the
    leg that was running and opens the next, so every leg of the bench gets its own clock record."""

    NAMES = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
             0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.max_mhz, self._stop_evt = index, None, threading.Event()
        self.legs: dict[str, dict] = {}
        self.leg = None
        self.lock = threading.Lock()
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

    def mark(self, name: str | None):
        with self.lock:
            self.leg = name
            if name is not None:
                self.legs.setdefault(name, {"samples": [], "reasons": set()})

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop_evt.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                mask = get_reasons(self.h)
                with self.lock:
                    if self.leg is not None:
                        leg = self.legs[self.leg]
                        leg["samples"].append(mhz)
                        leg["reasons"].update(n for bit, n in self.NAMES.items() if mask & bit)
            except Exception:
                pass
            time.sleep(0.002)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)

    def record(self, *names: str):
        """Clock record over the named legs together."""
        samples, reasons = [], set()
        for n in names:
            leg = self.legs.get(n)
            if leg:
                samples += leg["samples"]
                reasons |= leg["reasons"]
        if not samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"], "samples": 0}
        s = sorted(samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# CPU arm (the oracle; test infrastructure used here only as the timed CPU baseline)
# ------------------------------------------------------------------------------------------------
def cpu_arm(threads: int, frames_per_step: int, steps: int, warmup: int, split: str = "bands"):
    """split="bands": one frame at a time, cut into `threads` horizontal bands (what videoconvert's
    n-threads does). split="frames": `threads` workers convert different frames at once (the most a
    multi-session host can get out of its cores)."""
    import numpy as np

    import oracle
    orc = oracle.Oracle(native=True)            # -march=native build (oracle/_build_native/) for the box this runs on
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


def bench_config():
    """The `config` object of both arms (the reference arm runs a bounded sample of it, described in its cpu_baseline)."""
    return {"workload": WORKLOAD, "frames_per_step": FRAMES_PER_STEP, "sessions_per_gpu": SESSIONS,
            "ring_frames": RING, "launches_per_step": LAUNCHES_PER_STEP, "partition": "sessions by GPU, no collective",
            "l2": f"inputs larger than L2: ring of {RING} frames = {RING * ALGO_BYTES_PER_FRAME / 1e6:.0f} MB in+out per launch"}


def run_reference(args, rank: int):
    if rank != 0:
        return
    cores = host_cores()
    fps, dt = cpu_arm(cores, 2 * cores, args.steps, args.warmup, split="frames")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(fps, 2), "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": bench_config(),
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
class Job:
    """Rank bookkeeping: barrier, max and sum over ranks (torch.distributed carries nothing else)."""

    def __init__(self, torch, dist, world):
        self.torch, self.dist, self.world = torch, dist, world

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, value: float, op: str = "max") -> float:
        if self.dist is None:
            return float(value)
        t = self.torch.tensor([value], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())


def device_frames(b200vc, d_in, d_out, iw, ih, fmt, ow, oh, n):
    total, off, pit = b200vc.default_layout(fmt, ow, oh)
    df = b200vc.DeviceFrames()
    df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * 4, iw * ih * 4
    for i in range(2 if fmt == 1 else 3):
        df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
    df.out_frame_stride, df.n_frames = total, n
    return df


def time_kernel(torch, conv, df, stream, launches: int, warm: int, job: Job):
    """launches back-to-back batched launches on `stream`, CUDA events on that stream, max over ranks."""
    with torch.cuda.stream(stream):
        for _ in range(warm):
            conv.convert_device(df, stream.cuda_stream)
    job.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(launches):
            conv.convert_device(df, stream.cuda_stream)
        ev1.record(stream)
    job.barrier()
    return job.reduce(ev0.elapsed_time(ev1) / 1e3)


def time_e2e(conv, ticks, n_calls: int, warm: int, job: Job, sessions: int):
    """n_calls b200vc_convert_batch calls over the prepared ticks; wall clock around the loop (every call
    returns with its frames complete in host memory), max over ranks. Also the process CPU seconds."""
    for k in range(warm):
        conv.lib.b200vc_convert_batch(conv.ctx, ticks[k % len(ticks)], sessions)
    job.barrier()
    st0 = conv.stats()
    c0 = time.process_time()
    t0 = time.perf_counter()
    for k in range(n_calls):
        rc = conv.lib.b200vc_convert_batch(conv.ctx, ticks[k % len(ticks)], sessions)
        if rc:
            conv._check(rc)
    dt = time.perf_counter() - t0
    cpu = time.process_time() - c0
    st1 = conv.stats()
    job.barrier()
    return job.reduce(dt), cpu, st0, st1


def pcie_ceilings(torch, job: Job, h_in_t, h_out_t, d_in_flat, d_out_flat, in_bytes, out_bytes, ring, n=192):
    """What the link gives raw pinned copies of this traffic: H2D alone, D2H alone, both at once. Every rank
    copies between the same two barriers, so at N > 1 the figures include what the GPUs take from each other."""
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up: bool, dn: bool, count: int):
        for k in range(count):
            i = k % ring
            if up:
                with torch.cuda.stream(s_up):
                    d_in_flat[i * in_bytes:(i + 1) * in_bytes].copy_(h_in_t[i * in_bytes:(i + 1) * in_bytes], non_blocking=True)
            if dn:
                with torch.cuda.stream(s_dn):
                    h_out_t[i * out_bytes:(i + 1) * out_bytes].copy_(d_out_flat[i * out_bytes:(i + 1) * out_bytes], non_blocking=True)

    out = {}
    for name, up, dn in (("h2d_only", True, False), ("d2h_only", False, True), ("both", True, True)):
        run(up, dn, 16)
        job.barrier()
        t0 = time.perf_counter()
        run(up, dn, n)
        torch.cuda.synchronize()
        dt = job.reduce(time.perf_counter() - t0)
        job.barrier()
        out[name] = {"frames_per_s_per_gpu": round(n / dt, 1),
                     "h2d_gbs_per_gpu": round(n * in_bytes / dt / 1e9, 2) if up else 0.0,
                     "d2h_gbs_per_gpu": round(n * out_bytes / dt / 1e9, 2) if dn else 0.0}
        out[name]["gbs_all_gpus"] = round((out[name]["h2d_gbs_per_gpu"] + out[name]["d2h_gbs_per_gpu"]) * job.world, 2)
    return out


def run_session_bench(local_rank: int, sessions: int, seconds: float, hz: float, wait_file: str | None, extra=()):
    """harness/session_bench on this rank's GPU: plain C++ over the C ABI, no interpreter in the loop."""
    cmd = [SESSION_BENCH, "--gpus", "1", "--device", str(local_rank), "--sessions", str(sessions), "--seconds", str(seconds),
           "--hz", str(hz), *extra]
    if wait_file:
        cmd += ["--wait-file", wait_file]
    return subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)


def paced_run(job: Job, rank: int, local_rank: int, sessions: int, seconds: float, hz: float, extra=()):
    """All ranks start their harness together (a file rank 0 creates after a barrier releases them). A rank whose
    harness fails still walks through every barrier and reduction, so that the ranks fail together, not one by one."""
    gate = os.path.join(tempfile.gettempdir(), f"b200vc_gate_{os.environ.get('MASTER_PORT', '0')}_{sessions}_{int(hz)}_{int(time.time()) // 600}")
    if rank == 0 and os.path.exists(gate):
        os.unlink(gate)
    job.barrier()
    proc, err = None, ""
    try:
        proc = run_session_bench(local_rank, sessions, seconds, hz, gate, extra)
    except Exception as e:  # noqa: BLE001
        err = f"{type(e).__name__}: {e}"
    time.sleep(1.0)                              # contexts and pinned buffers are set up behind the gate
    job.barrier()
    if rank == 0:
        open(gate, "w").close()
    d = {"frames": 0, "ticks": 0, "late_ticks": 0, "aggregate_fps": 0.0, "tick_ms_p50": 0.0, "tick_ms_p99": 0.0, "tick_ms_max": 0.0}
    if proc is not None:
        try:
            out, perr = proc.communicate(timeout=seconds + 120)
            if proc.returncode != 0:
                err = f"session_bench exit {proc.returncode}: {perr[-300:]}"
            else:
                d = json.loads(out.strip().splitlines()[-1])
        except Exception as e:  # noqa: BLE001
            proc.kill()
            err = f"{type(e).__name__}: {e}"
    job.barrier()
    if rank == 0 and os.path.exists(gate):
        os.unlink(gate)
    res = {"frames": job.reduce(d["frames"], "sum"), "ticks": job.reduce(d["ticks"], "sum"),
           "late_ticks": job.reduce(d["late_ticks"], "sum"), "fps": job.reduce(d["aggregate_fps"], "sum"),
           "p50_ms": job.reduce(d["tick_ms_p50"]), "p99_ms": job.reduce(d["tick_ms_p99"]), "max_ms": job.reduce(d["tick_ms_max"])}
    if job.reduce(1.0 if err else 0.0, "sum") > 0:
        raise RuntimeError(err or "session_bench failed on another rank")
    return res


def ensure_harness(job: "Job", rank: int):
    """harness/session_bench travels with the tree; if it is missing, rank 0 builds it (gcc is on the box)."""
    if rank == 0 and not os.path.exists(SESSION_BENCH):
        subprocess.run(["make", "-C", PKG, "harness/session_bench"], capture_output=True)
    job.barrier()


def c5_record(job: "Job", rank: int, local_rank: int, world: int):
    """C5 in real time: 8 sessions per GPU at 60 Hz for 5 s, then as many as still miss no tick."""
    base = paced_run(job, rank, local_rank, SESSIONS, 5.0, 60, ("--damage",))
    c5 = {"workload": f"C5: {SESSIONS} sessions per GPU x {world} GPU(s), 1080p BGRx->I420 + damage map, paced at 60 Hz for 5 s, "
                      "harness/session_bench over the C ABI (one process per GPU, started together)",
          "sessions": SESSIONS * world, "target_fps": 60.0 * SESSIONS * world, "delivered_fps": round(base["fps"], 1),
          "ticks": int(base["ticks"]), "late_ticks": int(base["late_ticks"]), "tick_ms_p50": round(base["p50_ms"], 3),
          "tick_ms_p99": round(base["p99_ms"], 3), "tick_ms_max": round(base["max_ms"], 3), "tick_budget_ms": round(1000 / 60, 3)}
    best, sweep = (SESSIONS if base["late_ticks"] == 0 else 0), []
    for s in (16, 24, 32, 48, 64, 80, 96):
        r = paced_run(job, rank, local_rank, s, 1.5, 60, ("--damage",))
        sweep.append({"sessions_per_gpu": s, "late_ticks": int(r["late_ticks"]), "tick_ms_p99": round(r["p99_ms"], 3), "delivered_fps": round(r["fps"], 1)})
        if r["late_ticks"] > 0:
            break
        best = s
    c5["max_sessions_per_gpu_without_a_late_tick"] = best
    c5["sweep"] = sweep
    return c5


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
    job = Job(torch, dist, world)
    warm = max(args.warmup, 3)

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
    df = device_frames(b200vc, d_in, d_out, W, H, FMT, W, H, RING)
    stream = torch.cuda.Stream()

    # ---- host side for e2e: pinned ring in, pinned ring out -------------------------------------
    h_in = b200vc.PinnedBuffer(RING * IN_BYTES)
    h_out = b200vc.PinnedBuffer(RING * OUT_BYTES)
    torch.cuda.synchronize()
    h_in_t = torch.from_numpy(h_in.array)
    h_in_t.copy_(d_in.reshape(-1))                 # same frames as the device ring
    torch.cuda.synchronize()
    ticks = []
    for t in range(RING // SESSIONS):
        frames = [conv.make_frame(h_in.ptr + (t * SESSIONS + s) * IN_BYTES, W * 4, h_out.ptr + (t * SESSIONS + s) * OUT_BYTES,
                                  off, pit, session=s) for s in range(SESSIONS)]
        ticks.append((b200vc.Frame * SESSIONS)(*frames))

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- headline, kernel-only ---------------------------------------------------------------------
    sampler.mark("c1_kernel")
    launches0 = conv.stats().kernel_launches
    k_seconds = time_kernel(torch, conv, df, stream, args.steps * LAUNCHES_PER_STEP, warm * LAUNCHES_PER_STEP, job)
    gpu_launches = int(conv.stats().kernel_launches - launches0) - warm * LAUNCHES_PER_STEP

    # ---- headline, end to end -----------------------------------------------------------------------
    sampler.mark("c1_e2e")
    e_seconds, e_cpu, st0, st1 = time_e2e(conv, ticks, args.steps * TICKS, warm * TICKS, job, SESSIONS)
    e_cpu_max, e_cpu_sum = job.reduce(e_cpu), job.reduce(e_cpu, "sum")
    sampler.mark(None)

    # output sanity: kernel-only and e2e outputs of the last frame of the ring must agree byte for byte
    same = bool(np.array_equal(d_out[RING - 1].cpu().numpy(), h_out.array[(RING - 1) * OUT_BYTES:RING * OUT_BYTES]))

    # synchronous single-session call (GstBaseTransform's default mode), for the record
    sampler.mark("c1_sync")
    planes = (C.c_void_p * 3)(*[h_out.ptr + o for o in off])
    pits = (C.c_size_t * 3)(*pit)
    for i in range(8):
        conv.lib.b200vc_convert(conv.ctx, h_in.ptr + (i % RING) * IN_BYTES, W * 4, planes, pits, None)
    n_sync = 256
    job.barrier()
    t0 = time.perf_counter()
    for i in range(n_sync):
        conv.lib.b200vc_convert(conv.ctx, h_in.ptr + (i % RING) * IN_BYTES, W * 4, planes, pits, None)
    sync_fps = n_sync / job.reduce(time.perf_counter() - t0)
    # the same session with one frame of latency (submit N, collect N-1: the element's one-frame-latency mode)
    lat_frames = [conv.make_frame(h_in.ptr + i * IN_BYTES, W * 4, h_out.ptr + i * OUT_BYTES, off, pit, session=0) for i in range(RING)]
    prev_ticket = conv.submit(lat_frames[0])
    job.barrier()
    t0 = time.perf_counter()
    for i in range(1, n_sync + 1):
        ticket = conv.submit(lat_frames[i % RING])
        conv.collect(prev_ticket)
        prev_ticket = ticket
    conv.collect(prev_ticket)
    latency_mode_fps = n_sync / job.reduce(time.perf_counter() - t0)
    sampler.mark(None)

    # ---- the link's own ceilings, all ranks copying at once ------------------------------------------
    pcie = pcie_ceilings(torch, job, h_in_t, torch.from_numpy(h_out.array), d_in.reshape(-1), d_out.reshape(-1), IN_BYTES, OUT_BYTES, RING)
    conv.stop()
    del d_in, d_out, df
    h_in.free()
    h_out.free()
    torch.cuda.empty_cache()

    # ---- the same e2e loop without an interpreter: native harness, ticks back to back ----------------
    sampler.mark("c1_native")
    ensure_harness(job, rank)
    try:
        native = paced_run(job, rank, local_rank, SESSIONS, 2.0, 0)
    except Exception as e:  # noqa: BLE001 — the harness is an extra; the headline must not die with it
        native = {"fps": 0.0, "error": f"{type(e).__name__}: {e}"[:200]}
    sampler.mark(None)

    # ---- BASELINE configs C2, C3, C4 -------------------------------------------------------------------
    peak, peak_src = measured_peak()
    configs = {}
    for name, (iw, ih, ow, oh, fmt, mat, rg, site, damage, kernel, workload) in CONFIGS.items():
        in_b = iw * ih * 4
        ctotal, coff, cpit = b200vc.default_layout(fmt, ow, oh)
        mb = ((iw + 15) // 16) * ((ih + 15) // 16)
        n = max(8, min(64, int(600e6 // in_b)))                 # ring of >= ~600 MB of input: several times L2
        conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=mat, range_=rg, chroma_site=site,
                      flags=b200vc.FLAG_DAMAGE_MAP if damage else 0, max_sessions=SESSIONS, n_inflight=SESSIONS)
        c_in = torch.randint(0, 256, (n, in_b), dtype=torch.uint8, device="cuda", generator=gen)
        algo = in_b + ctotal
        dirty = None
        if damage:
            # a static desktop: every frame is frame 0 but for a noise window that moves and a cursor block
            v = c_in.view(n, ih, iw * 4)
            v[1:] = v[0]
            for k in range(1, n):
                x0, y0 = (64 + 40 * k) % (iw - 320), (48 + 24 * k) % (ih - 180)
                v[k, y0:y0 + 180, 4 * x0:4 * (x0 + 320)] = torch.randint(0, 256, (180, 1280), dtype=torch.uint8, device="cuda", generator=gen)
                cx, cy = (7 * k) % (iw - 16), (300 + 5 * k) % (ih - 24)
                v[k, cy:cy + 24, 4 * cx:4 * (cx + 16)] = 255
            algo += in_b + mb
        c_out = torch.zeros((n, ctotal), dtype=torch.uint8, device="cuda")
        cdf = device_frames(b200vc, c_in, c_out, iw, ih, fmt, ow, oh, n)
        if damage:
            c_prev = torch.roll(c_in, 1, 0).contiguous()
            c_dmg = torch.zeros((n, mb), dtype=torch.uint8, device="cuda")
            cdf.prev, cdf.prev_pitch, cdf.prev_frame_stride = c_prev.data_ptr(), iw * 4, in_b
            cdf.damage, cdf.damage_frame_stride = c_dmg.data_ptr(), mb
        sampler.mark(name + "_kernel")
        reps = max(20, args.steps)
        l0 = conv.stats().kernel_launches
        ksec = time_kernel(torch, conv, cdf, stream, reps, 5, job)
        per_call = (conv.stats().kernel_launches - l0) // (reps + 5)
        sampler.mark(None)
        if damage:
            dirty = float(c_dmg[1:].float().mean().item())
        # end to end: 8 sessions per call, pinned host buffers, two ticks of distinct frames
        ch_in, ch_out = b200vc.PinnedBuffer(2 * SESSIONS * in_b), b200vc.PinnedBuffer(SESSIONS * (ctotal + mb))
        torch.from_numpy(ch_in.array).copy_(c_in[:2 * SESSIONS].reshape(-1) if n >= 2 * SESSIONS else c_in.repeat((2 * SESSIONS + n - 1) // n, 1)[:2 * SESSIONS].reshape(-1))
        torch.cuda.synchronize()
        cticks = []
        for t in range(2):
            fr = [conv.make_frame(ch_in.ptr + (t * SESSIONS + s) * in_b, iw * 4, ch_out.ptr + s * (ctotal + mb), coff, cpit, session=s,
                                  damage_ptr=(ch_out.ptr + s * (ctotal + mb) + ctotal) if damage else None) for s in range(SESSIONS)]
            cticks.append((b200vc.Frame * SESSIONS)(*fr))
        n_calls = max(8, int(3e9 // (SESSIONS * in_b)))        # about 3 GB through the link
        sampler.mark(name + "_e2e")
        esec, ecpu, s0, s1 = time_e2e(conv, cticks, n_calls, 4, job, SESSIONS)
        sampler.mark(None)
        achieved = algo * n / (ksec / reps) / 1e9
        efps = world * n_calls * SESSIONS / esec
        configs[name] = {
            "workload": workload, "value": round(world * n * reps / ksec, 1), "unit": "frames/s",
            "ms_per_launch": round(ksec / reps * 1e3, 4), "ring_frames": n, "launches_timed": reps, "kernels_per_launch": int(per_call),
            "l2": f"inputs larger than L2: ring of {n} frames = {n * algo / 1e6:.0f} MB in+out per launch",
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                         "traffic": None, "peak_source": peak_src, "kernel": kernel, "algorithmic_bytes_per_frame": algo,
                         "algorithmic_bytes_per_launch": algo * n, "frac_of_nominal_8TBs": round(achieved / 8000.0, 4)},
            "e2e": {"value": round(efps, 1), "unit": "frames/s", "api": "b200vc_convert_batch, 8 sessions per call, pinned host buffers",
                    "h2d_bytes_per_call": int((s1.h2d_bytes - s0.h2d_bytes) // n_calls), "d2h_bytes_per_call": int((s1.d2h_bytes - s0.d2h_bytes) // n_calls),
                    "calls_timed": n_calls, "pcie_h2d_gbs_per_gpu": round(efps / world * in_b / 1e9, 2),
                    "frac_of_pcie_gen5_x16_raw_63GBs": round(efps / world * in_b / 63.02e9, 4),
                    "staged_bytes": int(s1.staged_in_bytes - s0.staged_in_bytes + s1.staged_out_bytes - s0.staged_out_bytes)},
            "clocks": sampler.record(name + "_kernel", name + "_e2e") if rank == 0 else None,
        }
        if damage:
            configs[name]["dirty_macroblock_fraction"] = round(dirty, 4)
        conv.stop()
        ch_in.free()
        ch_out.free()
        del c_in, c_out, cdf
        torch.cuda.empty_cache()

    # ---- C5: 8 sessions per GPU at 60 Hz, then as many as still miss no tick ---------------------------
    sampler.mark("c5")
    try:
        c5 = c5_record(job, rank, local_rank, world)
    except Exception as e:  # noqa: BLE001
        c5 = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    sampler.mark(None)
    sampler.finish()

    if rank == 0:
        frames = FRAMES_PER_STEP * args.steps
        value = world * frames / k_seconds
        launch_s = k_seconds / (args.steps * LAUNCHES_PER_STEP)
        achieved = ALGO_BYTES_PER_FRAME * RING / launch_s / 1e9
        e2e_fps = world * frames / e_seconds
        both = pcie["both"]["frames_per_s_per_gpu"]
        traffic, traffic_src = committed_traffic()
        c5["clocks"] = sampler.record("c5")
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": round(k_seconds / args.steps * 1e3, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": bench_config(),
            "e2e": {"value": round(e2e_fps, 1), "unit": "frames/s",
                    "h2d_bytes_per_step": int((st1.h2d_bytes - st0.h2d_bytes) // args.steps),
                    "d2h_bytes_per_step": int((st1.d2h_bytes - st0.d2h_bytes) // args.steps),
                    "api": "b200vc_convert_batch, 8 sessions per call, pinned host buffers",
                    "pcie_h2d_gbs": round(e2e_fps / world * IN_BYTES / 1e9, 2),
                    "pcie_ceiling_fps_per_gpu": both,
                    "frac_of_pcie_ceiling": round(e2e_fps / world / both, 4),
                    "frac_of_pcie_gen5_x16_raw_63GBs": round(e2e_fps / world * IN_BYTES / 63.02e9, 4),
                    "pcie_ceiling_how": "raw pinned copies of the same frames (8.3 MB up, 3.1 MB down), no kernel; every rank copies "
                                        "between the same two barriers, slowest rank's time; 'both' is the ceiling of this traffic",
                    "pcie": pcie,
                    "host_cpu_seconds": {"max_rank": round(e_cpu_max, 3), "all_ranks": round(e_cpu_sum, 3), "wall": round(e_seconds, 3),
                                         "host_cores": host_cores()},
                    "native_harness_fps": round(native["fps"], 1), **({"native_harness_error": native["error"]} if "error" in native else {}),
                    "native_harness_how": "harness/session_bench --hz 0: the same 8-session ticks back to back from C++, one process per GPU",
                    "sync_single_session_fps": round(sync_fps * world, 1), "sync_single_session_fps_per_gpu": round(sync_fps, 1),
                    "one_frame_latency_single_session_fps_per_gpu": round(latency_mode_fps, 1), "numa": numa},
            "gpu_launches": gpu_launches,
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "kernel": "k1_convert_kernel<I420,mpeg2>",
                         "algorithmic_bytes_per_launch": ALGO_BYTES_PER_FRAME * RING,
                         "launch_ms": round(launch_s * 1e3, 4), "frac_of_nominal_8TBs": round(achieved / 8000.0, 4)},
            "clocks": sampler.record("c1_kernel", "c1_e2e"), "outputs_agree": same,
            "configs": configs, "c5": c5,
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
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.abspath(__file__), "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup)]
        raise SystemExit(subprocess.call(cmd))
    run_gpu(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
