#!/bin/bash
# Multi-GPU host topology + the effect of binding each rank to its GPU's NUMA node on the e2e number.
# usage (on the GPU box): bash tools/numa_probe.sh N
N=${1:-8}
mkdir -p gpurun_out
{
  nvidia-smi topo -m
  echo "--- nodes"; ls /sys/devices/system/node/ | grep node
  for n in /sys/devices/system/node/node*; do echo "$n $(cat $n/cpulist) $(grep MemTotal $n/meminfo)"; done
  echo "--- allowed cpus"; python -c "import os; print(len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:4], '...')"
  echo "--- gpus"; for d in /sys/bus/pci/drivers/nvidia/0000*; do echo "$(basename $d) numa $(cat $d/numa_node)"; done
  cat /sys/fs/cgroup/cpu.max 2>/dev/null
} > gpurun_out/topo_${N}gpu.txt 2>&1
for bind in 0 1; do
  B200VC_NUMA_BIND=$bind python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$bind \
      bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu_bind$bind.json 2> gpurun_out/bench_${N}gpu_bind$bind.err
  echo "bind=$bind rc=$?"; cat gpurun_out/bench_${N}gpu_bind$bind.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])"
done
tail -30 gpurun_out/topo_${N}gpu.txt
