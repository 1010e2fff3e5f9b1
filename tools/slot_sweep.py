"""Staging-geometry sweep on an N-GPU box (tuning aid, not a bench line).

One child process per GPU, each loading its own 1 GiB shard of one ext4 file through
MEMCPY_SSD2GPU, all at once; for every (readers, slot KB, slots) setting prints the summed
source-only rate (device copy switched off) and the summed load+scan rate.
Usage: python tools/slot_sweep.py N [dir] [config ...]   config = readers:slot_kb:slots"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import os, sys, time
sys.path.insert(0, %r)
import torch
from nvme_strom_b200.abi import lib
from nvme_strom_b200.loader import SegmentPipeline
path, rank, world, tag = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
torch.cuda.set_device(rank)
pipe = SegmentPipeline(rank, 32 << 20, 6)
n_seg = 32
def barrier(name):
    open("%%s.%%s.%%s.%%d" %% (path, tag, name, rank), "w").close()
    for r in range(world):
        while not os.path.exists("%%s.%%s.%%s.%%d" %% (path, tag, name, r)):
            time.sleep(0.0005)
pipe.run(path, rank * n_seg, n_seg)
out = []
for mode in ("src", "e2e"):
    lib.strom_debug_skip_h2d(1 if mode == "src" else 0)
    pipe.scan = mode == "e2e"
    pipe.run(path, rank * n_seg, n_seg)
    barrier(mode)
    t0 = time.perf_counter()
    for _ in range(5):
        pipe.run(path, rank * n_seg, n_seg)
    torch.cuda.synchronize()
    out.append((time.perf_counter() - t0) / 5)
print("%%.6f %%.6f" %% tuple(out))
''' % ROOT


def main():
    n = int(sys.argv[1])
    d = sys.argv[2] if len(sys.argv) > 2 else "/tmp"
    configs = sys.argv[3:] or ["8:4096:2"]
    from nvme_strom_b200.loader import heapgen_file
    path = os.path.join(d, "slot_sweep.heap")
    heapgen_file(path, 0x5EED0001, n * 131072)
    shared = True        # (the private-slot engine of sections A of profiles/r01_slot_sweep.txt is gone)
    print(f"# {n} GPUs, {os.cpu_count()} cpus, file {path}, shared ring", flush=True)
    try:
        for k, cfg in enumerate(configs):
            readers, slot_kb, slots = cfg.split(":")
            env = dict(os.environ, NVME_STROM_READERS=readers, NVME_STROM_SLOT_KB=slot_kb, NVME_STROM_SLOTS=slots)
            procs = [subprocess.Popen([sys.executable, "-c", CHILD, path, str(r), str(n), str(k)], env=env,
                                      stdout=subprocess.PIPE, text=True) for r in range(n)]
            res = []
            for p in procs:
                try:
                    o, _ = p.communicate(timeout=120)
                    res.append([float(x) for x in o.split()])
                except Exception as e:                      # noqa: BLE001
                    p.kill()
                    res.append(None)
            for f in os.listdir(d):
                if f.startswith("slot_sweep.heap."):
                    os.unlink(os.path.join(d, f))
            if any(r is None or len(r) != 2 for r in res):
                print(f"readers={readers} slot={slot_kb}KB x{slots}: failed {res}", flush=True)
                continue
            src = n * (1 << 30) / max(r[0] for r in res) / 1e9
            e2e = n * (1 << 30) / max(r[1] for r in res) / 1e9
            print(f"readers={readers:>2} slot={slot_kb:>5}KB x{slots} staging/GPU={(1 if shared else int(readers)) * int(slot_kb) * int(slots) >> 10:>4}MB: "
                  f"source {src:6.1f} GB/s  load+scan {e2e:6.1f} GB/s", flush=True)
    finally:
        os.unlink(path)


if __name__ == "__main__":
    main()
