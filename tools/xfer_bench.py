"""Transfer-side calibration on the GPU box (tuning aid, not a bench line):
  1. pinned->device cudaMemcpyAsync rate (the H2D roofline of this box)
  2. file -> pinned host read rate at several thread counts (the source roofline)
  3. the library's MEMCPY_SSD2GPU pipeline at several reader/slot settings
Usage: python tools/xfer_bench.py [dir] [GiB]"""
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import os, sys, time
sys.path.insert(0, %r)
import torch
from nvme_strom_b200.loader import SegmentPipeline
path, scan = sys.argv[1], sys.argv[2] == "1"
pipe = SegmentPipeline(0, 32 << 20, 6, scan=scan)
best = 0
for it in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    tot, cnt = pipe.run(path)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    if it: best = max(best, cnt["n_segments"] * (32 << 20) / dt / 1e9)
print("%%.2f" %% best)
''' % ROOT


def main():
    d = sys.argv[1] if len(sys.argv) > 1 else "/tmp"
    gib = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    import numpy as np
    import torch
    from nvme_strom_b200.loader import heapgen_file
    path = os.path.join(d, "xfer.heap")
    n_pages = int(gib * (1 << 30)) // 8192
    t0 = time.perf_counter()
    heapgen_file(path, 0x5EED0001, n_pages)
    print(f"# generated {gib} GiB at {path} in {time.perf_counter() - t0:.2f}s ({os.cpu_count()} cpus)")
    nbytes = n_pages * 8192

    # 1. H2D calibration
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    best = 0
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); dev.copy_(host, non_blocking=True); b.record(); torch.cuda.synchronize()
        best = max(best, nbytes / a.elapsed_time(b) / 1e6)
    print(f"H2D pinned->device, one 1GiB copy: {best:.1f} GB/s")
    for piece in (1 << 20, 4 << 20):
        s = torch.cuda.Stream()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s):
            a.record(s)
            for off in range(0, nbytes, piece):
                dev[off:off + piece].copy_(host[off:off + piece], non_blocking=True)
            b.record(s)
        torch.cuda.synchronize()
        print(f"H2D in {piece >> 20}MB pieces on one stream: {nbytes / a.elapsed_time(b) / 1e6:.1f} GB/s")
    del dev

    # 2. source read rate into pinned memory (warm page cache after the first pass)
    view = memoryview(host.numpy())
    fd = os.open(path, os.O_RDONLY)
    for nth in (1, 2, 4, 8, 16, 32, 64):
        if nth > 2 * (os.cpu_count() or 1):
            break
        piece = 1 << 20
        cursor = [0]
        lock = threading.Lock()

        def work():
            while True:
                with lock:
                    off = cursor[0]; cursor[0] += piece
                if off >= nbytes:
                    return
                os.preadv(fd, [view[off:off + piece]], off)
        best = 0
        for _ in range(3):
            cursor[0] = 0
            ts = [threading.Thread(target=work) for _ in range(nth)]
            t0 = time.perf_counter(); [t.start() for t in ts]; [t.join() for t in ts]
            best = max(best, nbytes / (time.perf_counter() - t0) / 1e9)
        print(f"pread -> pinned, {nth:2d} threads, 1MB pieces: {best:.1f} GB/s")
    os.close(fd)
    del view, host

    # 3. the pipeline
    for readers, slot_kb, slots in [(6, 4096, 2), (8, 4096, 2), (10, 4096, 2), (12, 4096, 2), (8, 4096, 3),
                                    (10, 2048, 3), (8, 2048, 3), (12, 2048, 2)]:
        if readers > 2 * (os.cpu_count() or 1):
            continue
        env = dict(os.environ, NVME_STROM_READERS=str(readers), NVME_STROM_SLOT_KB=str(slot_kb),
                   NVME_STROM_SLOTS=str(slots))
        out = []
        for scan in ("0", "1"):
            r = subprocess.run([sys.executable, "-c", CHILD, path, scan], env=env, capture_output=True, text=True)
            out.append(r.stdout.strip() or ("ERR " + r.stderr[-200:]))
        print(f"pipeline readers={readers:2d} slot={slot_kb:4d}KB x{slots}: load-only {out[0]} GB/s, load+scan {out[1]} GB/s")
    os.unlink(path)


if __name__ == "__main__":
    main()
