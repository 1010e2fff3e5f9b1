"""Multi-GPU host logic: one process per GPU, disjoint contiguous segment
ranges per rank, no data-path collective; one small reduction of the counters
at the end.  Reference precedent for work splitting: the shared
block cursor of utils/ssd2ram_test.c:180 and pgsql/nvme_strom.c:997-998."""
import ctypes as C

import torch
import torch.distributed as dist

from . import abi
from .abi import lib


def partition_segments(n_segments, n_parts, rank):
    first, count = C.c_uint64(), C.c_uint64()
    abi.check(lib.strom_partition_segments(n_segments, n_parts, rank, C.byref(first), C.byref(count)),
              "strom_partition_segments")
    return first.value, count.value


def parse_cpu_list(text):
    """'0-3,8,10-11' -> [0, 1, 2, 3, 8, 10, 11]"""
    out = []
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out.extend(range(int(a), int(b or a) + 1))
    return out


def gpu_local_cpus(index):
    """the cores next to GPU `index` as sysfs reports them (local_cpulist of its PCI function),
    or None when the platform does not say (a VM without NUMA information lists every core)"""
    try:
        p = torch.cuda.get_device_properties(index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            cpus = parse_cpu_list(f.read())
        return cpus or None
    except Exception:
        return None


def rank_cpus(local_rank, local_world, allowed=None, local_lists=None):
    """The cores one rank of `local_world` ranks (one GPU each, GPU index = local rank) should keep
    to: ranks whose GPUs report the same neighbourhood split it evenly, in rank order, so that every
    rank's reader threads, its CUDA and NCCL helper threads and the pages they touch stay on the
    cores (and the NUMA node) next to its GPU, and eight ranks on one host do not fight over the
    same cores.  Reference precedent: utils/ssd2ram_test.c:66-119 (setup_cpu_affinity binds the
    worker to the device's node), pgsql/nvme_strom.c:300-378.  `allowed`: cores the job may use at
    all (default: the current affinity mask); `local_lists`: per-rank neighbourhoods, for tests."""
    import os
    allowed = sorted(allowed if allowed is not None else os.sched_getaffinity(0))
    if local_world <= 1 or not allowed:
        return allowed
    if local_lists is None:
        local_lists = [gpu_local_cpus(r) for r in range(local_world)]
    key = lambda r: tuple(local_lists[r]) if local_lists[r] else None
    mine = key(local_rank)
    group = [r for r in range(local_world) if key(r) == mine]
    cpus = [c for c in (mine or allowed) if c in set(allowed)] or allowed
    k, g = group.index(local_rank), len(group)
    if len(cpus) < g:                                   # fewer cores than ranks: share them all
        return cpus
    return cpus[k * len(cpus) // g:(k + 1) * len(cpus) // g]


def merge_totals(parts):
    """host-side merge of per-rank StromScanTotals lists (sum / xor / min)"""
    acc = abi.StromScanTotals()
    acc.v[abi.T_FIRST_MISMATCH] = 0xFFFFFFFFFFFFFFFF
    for p in parts:
        cur = abi.StromScanTotals()
        for i, x in enumerate(p):
            cur.v[i] = x
        lib.strom_totals_merge(C.byref(acc), C.byref(cur))
    return acc.as_list()


def reduce_totals(totals, device):
    """all ranks contribute 16 counters; everyone gets the merged list.  One
    all_gather of 128 bytes per rank (NCCL on GPUs, gloo on CPU); the merge
    rule (sum, XOR witness, min offset) is the library's strom_totals_merge."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(totals)
    world = dist.get_world_size()
    # int64 carrier: reinterpret the u64 bit patterns
    mine = torch.tensor([x - (1 << 64) if x >= (1 << 63) else x for x in totals], dtype=torch.int64,
                        device=device)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    parts = [[int(v) & 0xFFFFFFFFFFFFFFFF for v in g.cpu().tolist()] for g in gathered]
    return merge_totals(parts)
