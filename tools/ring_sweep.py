"""Staging sweep on an N-GPU box in ONE set of processes (tuning aid, not a bench line).

One child per GPU, each loading its own 1 GiB shard of one file through MEMCPY_SSD2GPU, all at
once.  Between settings the children rebuild their engines (strom_debug_reconfigure), so a setting
costs seconds instead of a fresh torch import.  Per setting: the pipeline's source-only rate (device
copy switched off), the load+scan rate, per-GPU rates, and what the engine decided for itself.

  python tools/ring_sweep.py N [dir] setting ...
  setting = readers:slot_kb:slots[:copy[:aff[:src]]]
     readers 0 = self-adjusting; slot_kb 0 = the engine's default for its cores
     copy = memcpy | nta | ntstore          aff = split (each rank keeps to its share of the cores) | none
     src  = file (ext4 under dir, staged) | shm (tmpfs, page-locked: zero-copy)
     further fields NAME=VALUE set other NVME_STROM_* knobs for that setting (NVME_STROM_SUB_KB=128 ...)"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import json, os, sys, time
sys.path.insert(0, %r)
import torch
from nvme_strom_b200 import abi, multigpu
from nvme_strom_b200.abi import lib
from nvme_strom_b200.loader import SegmentPipeline
path, shm_path, rank, world = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
settings = sys.argv[5:]
all_cpus = sorted(os.sched_getaffinity(0))
torch.cuda.set_device(rank)
pipe = SegmentPipeline(rank, 32 << 20, 6)
n_seg = 32
def barrier(name):
    open("%%s.b.%%s.%%d" %% (path, name, rank), "w").close()
    deadline = time.time() + 60
    for r in range(world):
        while not os.path.exists("%%s.b.%%s.%%d" %% (path, name, r)):
            if time.time() > deadline:
                sys.exit("rank %%d: rank %%d never reached barrier %%s" %% (rank, r, name))
            time.sleep(0.0005)
pinned = False
for k, text in enumerate(settings):
    f = text.split(":")
    f += ["memcpy", "split", "file"][len(f) - 3:] if len(f) < 6 else []
    readers, slot_kb, slots, copy, aff, src = f[:6]
    for name in ("NVME_STROM_SUB_KB", "NVME_STROM_SPIN_US", "NVME_STROM_ZC_PIECE_MB", "NVME_STROM_ADAPTIVE", "NVME_STROM_WC"):
        os.environ.pop(name, None)
    for extra in f[6:]:                                   # NAME=VALUE: any other NVME_STROM_* knob, this setting only
        name, _, val = extra.partition("=")
        os.environ[name] = val
    for name, val in (("NVME_STROM_READERS", readers), ("NVME_STROM_SLOT_KB", slot_kb)):
        if int(val) > 0: os.environ[name] = val
        else: os.environ.pop(name, None)
    os.environ["NVME_STROM_SLOTS"] = slots
    os.environ["NVME_STROM_COPY"] = copy
    os.sched_setaffinity(0, multigpu.rank_cpus(rank, world, allowed=all_cpus) if aff == "split" else all_cpus)
    abi.check(lib.strom_debug_reconfigure(), "reconfigure")
    p = shm_path if src == "shm" else path
    if src == "shm" and not pinned:
        fd = os.open(p, os.O_RDONLY)
        abi.check(lib.nvme_strom_pin_file_range(fd, rank * n_seg * (32 << 20), n_seg * (32 << 20)), "pin")
        os.close(fd)
        pinned = True
    out = {"k": k, "rank": rank}
    for mode in (("e2e",) if src == "shm" else ("src", "e2e")):
        lib.strom_debug_skip_h2d(1 if mode == "src" else 0)
        pipe.scan = mode == "e2e"
        pipe.run(p, rank * n_seg, n_seg)
        pipe.run(p, rank * n_seg, n_seg)
        barrier("%%d.%%s" %% (k, mode))
        t0 = time.perf_counter()
        for _ in range(5):
            pipe.run(p, rank * n_seg, n_seg)
        torch.cuda.synchronize()
        out[mode] = (time.perf_counter() - t0) / 5
    lib.strom_debug_skip_h2d(0)
    out["state"] = abi.engine_state(rank)
    print(json.dumps(out), flush=True)
''' % ROOT


def main():
    n = int(sys.argv[1])
    d = sys.argv[2] if len(sys.argv) > 2 else "/tmp"
    settings = sys.argv[3:] or ["0:0:4"]
    from nvme_strom_b200.loader import heapgen_file
    path = os.path.join(d, "ring_sweep.heap")
    shm_path = "/dev/shm/ring_sweep.heap"
    heapgen_file(path, 0x5EED0001, n * 131072)
    want_shm = any(s.split(":")[5:6] == ["shm"] for s in settings)
    print("# " + " ".join(sys.argv), flush=True)
    if want_shm:
        heapgen_file(shm_path, 0x5EED0001, n * 131072)
    print(f"# {n} GPUs, {os.cpu_count()} cpus, file {path}", flush=True)
    try:
        procs = [subprocess.Popen([sys.executable, "-c", CHILD, path, shm_path, str(r), str(n)] + settings,
                                  stdout=subprocess.PIPE, text=True) for r in range(n)]
        rows = {}
        for p in procs:
            try:
                o, _ = p.communicate(timeout=90 + 12 * len(settings))
            except Exception:                                # noqa: BLE001
                p.kill()
                o = ""
            for line in o.splitlines():
                if line.startswith("{"):
                    r = json.loads(line)
                    rows.setdefault(r["k"], []).append(r)
        for k, text in enumerate(settings):
            rs = sorted(rows.get(k, []), key=lambda r: r["rank"])
            if len(rs) != n:
                print(f"{text}: failed ({len(rs)} of {n} ranks reported)", flush=True)
                continue
            gib = (1 << 30) / 1e9
            e2e = n * gib / max(r["e2e"] for r in rs)
            src = n * gib / max(r["src"] for r in rs) if "src" in rs[0] else float("nan")
            st = rs[0]["state"]
            per = " ".join(f"{gib / r['e2e']:.1f}" for r in rs)
            print(f"{text:34s} pipeline-source {src:6.1f}  load+scan {e2e:6.1f} GB/s  per-GPU [{per}]  "
                  f"rank0: cpus {st['allowed_cpus']} slot {st['slot_bytes'] >> 10}KB x{st['slots']} "
                  f"readers {st['readers_active']}/{st['readers']} grown {st['grown']} shrunk {st['shrunk']} "
                  f"starved {st['slots_starved']}/{st['slots_staged']}", flush=True)
    finally:
        for f in os.listdir(d):
            if f.startswith("ring_sweep.heap"):
                os.unlink(os.path.join(d, f))
        if want_shm and os.path.exists(shm_path):
            os.unlink(shm_path)


if __name__ == "__main__":
    main()
