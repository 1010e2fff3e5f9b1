"""The two terms of the transfer roofline, measured on the box in the same run as the load they
are held against (north_star: "the slower of storage or page-cache read bandwidth and PCIe Gen5
H2D").  Harness code shared by bench.py and tools/run_configs.py; torch is used for device memory,
events and the gather only."""
import os
import time

import torch
import torch.distributed as dist

from . import abi


def gather_floats(x, dev):
    """[x of rank 0, x of rank 1, ...] on every rank (a list of one without a process group)"""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(x)]
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def _barrier():
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
    torch.cuda.synchronize()


def h2d_aggregate(dev, dst, nbytes=1 << 29, trials=8):
    """Pinned host -> device copies of `nbytes` per rank into `dst` (a device uint8 tensor of at
    least that size), all ranks at once.  A trial's aggregate is all bytes over the slowest rank's
    time; the best trial counts -- not a sum of per-rank bests, which no single moment ever saw.
    Returns (aggregate GB/s, per-rank GB/s of that trial)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    best, per_rank = 0.0, []
    for _ in range(trials):
        _barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        dst[:nbytes].copy_(host, non_blocking=True)
        b.record()
        torch.cuda.synchronize()
        times = gather_floats(a.elapsed_time(b) * 1e-3, dev)
        agg = world * nbytes / max(times) / 1e9
        if agg > best:
            best, per_rank = agg, [round(nbytes / t / 1e9, 2) for t in times]
    del host
    return best, per_rank


def pagecache_aggregate(dev, path, offset, length, n_threads, passes=3, dest_bytes_per_thread=0):
    """Page-cache read rate of this rank's range of `path`, every rank at once, by `n_threads`
    readers per rank copying into cache-resident scratch (strom_debug_pagecache_read: no staging
    ring, no device).  All bytes over the slowest rank's time, best pass.  GB/s.
    dest_bytes_per_thread > 0: the host's plain copy rate instead (every reader walks a destination
    of that size, which does not stay in cache) -- what a staged path falls to when its ring does
    not stay cache-resident."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    fd = os.open(path, os.O_RDONLY)
    best = 0.0
    try:
        for _ in range(passes):
            _barrier()
            t0 = time.perf_counter()
            abi.pagecache_read_GBps(fd, offset, length, n_threads, dest_bytes_per_thread)
            times = gather_floats(time.perf_counter() - t0, dev)
            best = max(best, world * length / max(times) / 1e9)
    finally:
        os.close(fd)
    return best


def filesystem_of(path):
    """(mount point, fs type) of the file system holding `path`, from /proc/mounts"""
    path = os.path.realpath(path)
    best = ("/", "?")
    try:
        with open("/proc/mounts") as f:
            for line in f:
                _, mnt, fstype = line.split()[:3]
                if (path == mnt or path.startswith(mnt.rstrip("/") + "/")) and len(mnt) >= len(best[0]):
                    best = (mnt, fstype)
    except OSError:
        pass
    return best
