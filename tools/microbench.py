"""Kernel micro-benchmark: every scan-kernel variant, the verify kernel and the
fused kernel over a resident buffer (>> L2), CUDA-event timed through the C ABI.
Not a bench line -- a tuning aid.  Usage: python tools/microbench.py [GiB] [reps]"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[0], ts[len(ts) // 2]


def main():
    gib = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    n_pages = int(gib * (1 << 30)) // 8192
    nbytes = n_pages * 8192
    peak = 6545.9
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    # synthetic heap pages: generate 64 MiB with the library, tile it (blkno mismatch -> checksum
    # failures are counted, same work per byte), then stamp nothing: content does not change the cost
    base_pages = 8192
    host = np.empty(base_pages * 8192, dtype=np.uint8)
    for i in range(base_pages):
        lib.strom_heapgen_page(0x5EED0001, i, 0, host[i * 8192:].ctypes.data)
    base = torch.from_numpy(host).cuda()
    buf = base.repeat((n_pages + base_pages - 1) // base_pages)[:nbytes].contiguous()
    ref = buf.clone()
    tot = torch.zeros(abi.T_COUNT, dtype=torch.int64, device="cuda")
    res = torch.zeros(n_pages * 4, dtype=torch.int16, device="cuda")
    pt, pr = C.c_void_p(tot.data_ptr()), C.c_void_p(res.data_ptr())
    pb, pf = C.c_void_p(buf.data_ptr()), C.c_void_p(ref.data_ptr())
    print(f"# {gib} GiB resident, {n_pages} pages, reps {reps}, peak {peak} GB/s (measured copy)")

    # a plain torch read-only reduction and copy for orientation
    best, med = timed(lambda: ref.copy_(buf), reps)
    print(f"torch copy_ (r+w)         best {2 * nbytes / best / 1e6:8.1f} GB/s  median {2 * nbytes / med / 1e6:8.1f}")

    variants = ["ring:512:16", "ring:256:16", "ring:384:32", "lean:8:2", "ldg:8:2", "ldg:4:3", "roll:16:2",
                "col:12", "tma:16:1", "tma:8:3"]
    rows = []
    for v in variants:
        os.environ["NVME_STROM_SCAN_VARIANT"] = v
        fn = lambda: abi.check(abi.scan_lib().strom_scan_pages_dev(pb, n_pages, 0, 0, pr, pt, None), "scan")
        best, med = timed(fn, reps)
        rows.append((v, nbytes / best / 1e6, nbytes / med / 1e6))
        print(f"scan {v:10s}           best {rows[-1][1]:8.1f} GB/s  median {rows[-1][2]:8.1f}  frac {rows[-1][2] / peak:.3f}")
    os.environ.pop("NVME_STROM_SCAN_VARIANT")
    best, med = timed(lambda: abi.check(lib.strom_verify_pages_dev(pb, pf, n_pages, None, 0, pt, None), "v"), reps)
    print(f"verify (2 buffers)        best {2 * nbytes / best / 1e6:8.1f} GB/s  median {2 * nbytes / med / 1e6:8.1f}  frac {2 * nbytes / med / 1e6 / peak:.3f}")
    for r in ("0", "384", "192", "448", "4"):
        os.environ["NVME_STROM_FUSED_R"] = r
        best, med = timed(lambda: abi.check(abi.scan_lib().strom_scan_verify_pages_dev(pb, pf, n_pages, 0, 0, 0, pr, pt, None), "f"), reps)
        print(f"fused scan+verify R={r:3s}   best {2 * nbytes / best / 1e6:8.1f} GB/s  median {2 * nbytes / med / 1e6:8.1f}  frac {2 * nbytes / med / 1e6 / peak:.3f}")
    # small launches: one 32MB segment (what the pipeline launches per slot)
    seg_pages = 4096
    for v in ("ring:256:16", "ring:512:16", "ldg:8:2"):
        os.environ["NVME_STROM_SCAN_VARIANT"] = v
        best, med = timed(lambda: abi.check(abi.scan_lib().strom_scan_pages_dev(pb, seg_pages, 0, 0, pr, pt, None), "s"), reps)
        print(f"scan {v:10s} 32MB seg   best {best * 1e3:7.1f} us  median {med * 1e3:7.1f} us  ({seg_pages * 8192 / med / 1e6:.0f} GB/s, L2-resident)")


if __name__ == "__main__":
    main()
