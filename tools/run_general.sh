#!/bin/bash
# GPU box: parity of the scale paths, a fuzz run, then the general fused kernel's timings (tools/bench_configs.py "G " lines);
# with arguments, the timings again for each forced tile height (B200VC_TILE_H).
timeout 500 python -m pytest tests/test_gpu_scale.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/fuzz_parity.py 1200 77 > gpurun_out/fuzz.log 2>&1; echo "fuzz exit $?"; tail -3 gpurun_out/fuzz.log
show() {
python - "$1" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print(d["config"], d["tile_h"], d["box"], d["frac_of_measured_peak"])
PY
}
timeout 200 python tools/bench_configs.py 30 "G " > gpurun_out/r2_general_v2.jsonl 2>&1; show gpurun_out/r2_general_v2.jsonl
for h in "$@"; do
  echo "--- forced tile height $h"
  B200VC_TILE_H=$h timeout 200 python tools/bench_configs.py 30 "G " > gpurun_out/r2_general_h$h.jsonl 2>&1; show gpurun_out/r2_general_h$h.jsonl
done
