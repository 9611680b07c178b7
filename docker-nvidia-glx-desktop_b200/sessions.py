"""Session -> GPU partitioning and whole-job aggregation for the multi-GPU driver.

The reference runs one desktop session per container per GPU (/root/reference xgl.yml:10,120;
README.md:24,182): sessions share no state, so the multi-GPU path has no collective on the data
path (SURVEY.md §8 e). What the ranks do exchange is bookkeeping: a barrier around the timed
region and the max of the per-rank elapsed times, through whatever torch.distributed backend the
launcher initialised (nccl on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations


def partition_sessions(n_sessions: int, n_gpus: int) -> list[list[int]]:
    """session s runs on GPU s % n_gpus (config C5: 64 sessions -> 8 per GPU on 8 GPUs)."""
    if n_sessions < 0 or n_gpus < 1:
        raise ValueError("n_sessions >= 0 and n_gpus >= 1 required")
    return [[s for s in range(n_sessions) if s % n_gpus == g] for g in range(n_gpus)]


def job_throughput(units_local: int, seconds_local: float, dist=None, device=None) -> tuple[float, float, int]:
    """Whole-job rate = units processed by all ranks / the slowest rank's time.
    Returns (rate, max_seconds, total_units). dist=None means a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_local / seconds_local, seconds_local, units_local
    import torch
    t = torch.tensor([seconds_local], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    u = torch.tensor([units_local], dtype=torch.int64, device=device)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return int(u.item()) / float(t.item()), float(t.item()), int(u.item())


def parse_cpulist(text: str) -> list[int]:
    """'0-3,8,10-11' (the kernel's cpulist format) -> [0, 1, 2, 3, 8, 10, 11]."""
    cpus: list[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.extend(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(pci_bus_id: str, sysfs: str = "/sys") -> int:
    """NUMA node of the GPU at a PCI address as NVML prints it ('00000000:1B:00.0'); -1 if unknown."""
    dom, bus, rest = pci_bus_id.lower().split(":")
    path = f"{sysfs}/bus/pci/devices/{dom[-4:]}:{bus}:{rest}/numa_node"
    try:
        with open(path) as f:
            return int(f.read().strip())
    except (OSError, ValueError):
        return -1


def bind_to_gpu_numa(device_index: int, sysfs: str = "/sys") -> dict:
    """Run this process on the cores of the GPU's NUMA node, so that the pinned frame buffers it
    allocates afterwards (first touch, local policy) are on the memory the GPU's PCIe root is
    attached to. Each session's frames cross PCIe once in each direction; on a two-socket host a
    buffer on the far socket also crosses the socket interconnect, which all the GPUs share.
    Call before the first pinned allocation. Never fails: returns what it did."""
    import os
    info = {"device": device_index, "node": -1, "bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        bus_id = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(device_index)).busId
        if isinstance(bus_id, bytes):
            bus_id = bus_id.decode()
        node = gpu_numa_node(bus_id, sysfs)
        info["node"] = node
        if node < 0:
            return info
        with open(f"{sysfs}/devices/system/node/node{node}/cpulist") as f:
            cpus = set(parse_cpulist(f.read()))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        info["cpus_node"], info["cpus_allowed"] = len(cpus), len(allowed)
        if use:
            os.sched_setaffinity(0, use)
            info["bound"], info["cpus"] = True, len(use)
    except Exception as e:  # noqa: BLE001 — a missing sysfs entry or NVML must not stop the run
        info["error"] = f"{type(e).__name__}: {e}"
    return info
