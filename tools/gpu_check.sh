#!/bin/bash
# Runs on the GPU box (through gpurun): GPU tests, smoke, bench, then — only if all of that exited 0 —
# the ncu launch list and one full capture of the dominant kernel. Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > gpurun_out/nvsmi.csv 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
PT=$?
echo "pytest exit $PT"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
SM=$?
echo "smoke exit $SM"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
BE=$?
echo "bench exit $BE"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "bench ref exit $?"; cat gpurun_out/bench_ref.json
if [ "${1:-}" = "ncu" ] && [ $PT -eq 0 ] && [ $SM -eq 0 ] && [ $BE -eq 0 ]; then
  timeout 300 python bench.py --steps 2 --warmup 1 > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv \
      --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launches.log 2>&1
  echo "ncu launches exit $?"
  timeout 300 python bench.py --steps 2 --warmup 1 > gpurun_out/plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k1_convert -s 10 -c 2 \
      -o gpurun_out/prof_k1 -f python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
exit 0
