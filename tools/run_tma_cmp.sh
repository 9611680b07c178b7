#!/bin/bash
# GPU box: tools/run_general.sh, then the same timings with the tile stores done by the kernel (B200VC_NO_TMA_STORE).
bash tools/run_general.sh
echo "--- without TMA stores"
B200VC_NO_TMA_STORE=1 timeout 200 python tools/bench_configs.py 30 "G " > gpurun_out/r2_general_notma.jsonl 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/r2_general_notma.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print(d["config"], d["tile_h"], d["frac_of_measured_peak"])
PY
