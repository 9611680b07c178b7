"""world_size-2 gloo test (CPU) of the N>1 path's host logic: session partitioning and the
whole-job aggregation bench.py uses. The ranks convert their own sessions (with the CPU oracle
standing in for the GPU, which this container does not have) and exchange nothing but the barrier,
the max elapsed time and the unit count — exactly what the multi-GPU bench exchanges."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

import sessions


def test_partition_sessions():
    parts = sessions.partition_sessions(64, 8)
    assert [len(p) for p in parts] == [8] * 8
    assert sorted(s for p in parts for s in p) == list(range(64))
    assert parts[3] == [3, 11, 19, 27, 35, 43, 51, 59]
    assert sessions.partition_sessions(5, 2) == [[0, 2, 4], [1, 3]]
    assert sessions.partition_sessions(0, 3) == [[], [], []]
    with pytest.raises(ValueError):
        sessions.partition_sessions(4, 0)


def test_job_throughput_single_process():
    assert sessions.job_throughput(100, 2.0) == (50.0, 2.0, 100)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, root, q):
    sys.path[:0] = [root, os.path.join(root, "docker-nvidia-glx-desktop_b200"), os.path.join(root, "tests")]
    import torch.distributed as dist

    import oracle
    import sessions as S
    from helpers import plane_checksums, plane_views
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = oracle.get()
    mine = S.partition_sessions(6, world)[rank]
    sums = {}
    for s in mine:
        f = orc.synth("desk", 64, 32, 100 * s)
        buf, off, pit = orc.convert(f, 0, 0, 0, 1)
        sums[s] = plane_checksums(plane_views(buf, off, pit, 0, 64, 32))
    dist.barrier()
    seconds = 1.0 + rank          # rank 1 is "slower": the job time must be its time
    rate, tmax, total = S.job_throughput(len(mine), seconds, dist)
    q.put((rank, sums, rate, tmax, total))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gloo(orc):
    from helpers import plane_checksums, plane_views
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, root, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    merged = {}
    for rank, sums, rate, tmax, total in results:
        assert total == 6 and tmax == 2.0 and rate == 3.0      # all units / slowest rank
        merged.update(sums)
    # sharding changes nothing: every session's output equals the single-process result
    for s in range(6):
        f = orc.synth("desk", 64, 32, 100 * s)
        buf, off, pit = orc.convert(f, 0, 0, 0, 1)
        assert merged[s] == plane_checksums(plane_views(buf, off, pit, 0, 64, 32))


def test_cpulist_and_numa_lookup(tmp_path):
    assert sessions.parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert sessions.parse_cpulist("") == []
    d = tmp_path / "bus/pci/devices/0000:1b:00.0"
    d.mkdir(parents=True)
    (d / "numa_node").write_text("1\n")
    assert sessions.gpu_numa_node("00000000:1B:00.0", str(tmp_path)) == 1
    assert sessions.gpu_numa_node("00000000:9F:00.0", str(tmp_path)) == -1
    # no GPU / no NVML here: the binder reports instead of raising
    info = sessions.bind_to_gpu_numa(0, str(tmp_path))
    assert info["bound"] is False
