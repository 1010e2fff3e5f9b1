"""Fixed cost vs per-byte cost of the scan kernel: K back-to-back launches over
buffers of several sizes, for the static and the dynamic-ticket schedule.
Tuning aid.  Usage: python tools/scan_sizes.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib


def main():
    base_pages = 8192
    host = np.empty(base_pages * 8192, dtype=np.uint8)
    for i in range(base_pages):
        lib.strom_heapgen_page(0x5EED0001, i, 1, host[i * 8192:].ctypes.data)
    base = torch.from_numpy(host).cuda()
    big = base.repeat(64)                       # 4 GiB
    tot = torch.zeros(abi.T_COUNT, dtype=torch.int64, device="cuda")
    pt = C.c_void_p(tot.data_ptr())
    variants = sys.argv[1:] or ["ldg:8:2", "ldgdyn", "ldg:4:3"]
    ref_tot = None
    for v in variants:
        os.environ["NVME_STROM_SCAN_VARIANT"] = v
        abi.check(lib.strom_totals_reset_dev(pt, None), "r")
        abi.check(abi.scan_lib().strom_scan_pages_dev(C.c_void_p(big.data_ptr()), 131072, 0, 0, None, pt, None), "s")
        torch.cuda.synchronize()
        t = tot.cpu().tolist()
        if ref_tot is None:
            ref_tot = t
        print(f"# {v}: totals {'match' if t == ref_tot else 'DIFFER'}")
        for mib in (32, 128, 256, 512, 1024, 2048, 4096):
            n_pages = mib * 128
            k = max(10, min(200, 8192 // mib))
            for _ in range(3):
                abi.scan_lib().strom_scan_pages_dev(C.c_void_p(big.data_ptr()), n_pages, 0, 0, None, pt, None)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(k):
                abi.scan_lib().strom_scan_pages_dev(C.c_void_p(big.data_ptr()), n_pages, 0, 0, None, pt, None)
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) / k * 1e3
            print(f"{v:10s} {mib:5d} MiB  {us:9.1f} us/launch  {mib * 1.048576 / us * 1e3:8.1f} GB/s")


if __name__ == "__main__":
    main()
