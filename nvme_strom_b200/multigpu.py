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
