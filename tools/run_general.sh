timeout 500 python -m pytest tests/test_gpu_scale.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/fuzz_parity.py 1200 77 > gpurun_out/fuzz.log 2>&1; echo "fuzz exit $?"; tail -3 gpurun_out/fuzz.log
timeout 200 python tools/bench_configs.py 30 "G " > gpurun_out/r2_general_v2.jsonl 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/r2_general_v2.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print(d["config"], d["tile_h"], d["box"], d["frac_of_measured_peak"])
PY
