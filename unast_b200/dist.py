"""torch.distributed plumbing for the multi-GPU runs.

Utterances are independent, so ranks never exchange data on the hot path; torch.distributed is used
for exactly two things: agreeing on who owns which utterances (deterministic, no communication -- see
sharding.py) and reducing the per-rank device time / unit counts of a run (max / sum over ranks).
Backend: NCCL on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from .sharding import deal_longest_first, length_buckets


def env_world():
    """(rank, local_rank, world_size) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_from_env(backend=None):
    """Initialise the default process group when launched by torchrun; returns (rank, local_rank, world)."""
    rank, local_rank, world = env_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kw["device_id"] = torch.device("cuda", local_rank)
            # NCCL's version / debug banner goes to stdout by default; bench.py's stdout is one JSON line
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
                os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, local_rank, world


def pin_to_local_cores(local_rank, local_world):
    """Give this rank a disjoint set of host cores on its GPU's NUMA node before it allocates pinned buffers or starts its
    reader / writer threads: ranks that share one affinity mask (every GPU of the round-1 box reported CPU affinity 0-31,
    NUMA 0) otherwise migrate across each other's cores and stage through each other's memory.  Best effort; returns what
    was done for the bench record."""
    info = {"cores_before": len(os.sched_getaffinity(0))}
    try:
        cores = sorted(os.sched_getaffinity(0))
        node_cores = None
        try:
            import pynvml
            pynvml.nvmlInit()
            bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
            bus = bus.decode() if isinstance(bus, bytes) else bus
            bus = bus.lower()
            if len(bus.split(":")[0]) == 8:                            # nvml prints a 32-bit domain, sysfs a 16-bit one
                bus = bus[4:]
            with open("/sys/bus/pci/devices/%s/numa_node" % bus) as fh:
                node = int(fh.read().strip())
            info["numa_node"] = node
            if node >= 0:
                with open("/sys/devices/system/node/node%d/cpulist" % node) as fh:
                    lst = set()
                    for part in fh.read().strip().split(","):
                        a, _, b = part.partition("-")
                        lst.update(range(int(a), int(b or a) + 1))
                node_cores = [c for c in cores if c in lst]
        except Exception as e:                                          # no nvml / sysfs: fall back to the affinity mask
            info["numa_lookup"] = repr(e)[:80]
        use = node_cores if node_cores else cores
        per = len(use) // max(1, local_world)
        if per >= 1 and local_world > 1:
            mine = use[local_rank * per:(local_rank + 1) * per]
            os.sched_setaffinity(0, mine)
            info["cores"] = [mine[0], mine[-1]]
        info["cores_after"] = len(os.sched_getaffinity(0))
    except Exception as e:
        info["error"] = repr(e)[:80]
    return info


def _device_for_collectives():
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def max_over_ranks(value):
    """Max of a python float over all ranks (device time of a step: the slowest rank defines it)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=_device_for_collectives())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value):
    """Sum of a python float over all ranks (units processed by the whole job)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=_device_for_collectives())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def barrier():
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def my_batches(sizes, rank, world, batch_size):
    """This rank's utterances (longest-first deal on `sizes`), cut into length-homogeneous batches."""
    mine = deal_longest_first(sizes, world)[rank]
    return length_buckets(mine, sizes, batch_size)
