"""CPU tests of bench.py's contract with the driver: the reference arm runs here (it is the CPU oracle), prints one
JSON line with the keys the contract names, and carries the same `config` object as the GPU arm; the clock sampler
keeps one record per leg; a rank that is not rank 0 prints nothing in the reference arm."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600,
                          env=dict(os.environ, **(env or {})))


def test_reference_arm_line_and_config():
    sys.path.insert(0, ROOT)
    import bench
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == "frames/s" and d["higher_is_better"] is True
    assert d["config"] == bench.bench_config()                      # the two arms name the same workload, key for key
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and "restatement" in cb["sample"] and cb["value"] == d["value"]
    assert d["vs_baseline"] is None and d["dtype"] == "u8" and d["gpu_launches"] == 0
    # the native oracle is built in-tree (visible to whoever lists the loaded libraries), not in a temp dir
    assert os.path.exists(os.path.join(ROOT, "oracle", "_build_native", "liboracle.so"))


def test_reference_arm_other_ranks_stay_silent():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--gpus", "2"], env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_clock_sampler_keeps_a_record_per_leg():
    sys.path.insert(0, ROOT)
    import bench
    s = bench.ClockSampler(0)                 # no NVML here: the records say so instead of inventing clocks
    s.mark("a")
    s.mark(None)
    rec = s.record("a", "b")
    assert set(rec) == {"sm_mhz", "sm_max_mhz", "reasons", "samples"}
    if not s.ok:
        assert rec["samples"] == 0 and rec["reasons"] == ["unavailable"]
    s.finish()
