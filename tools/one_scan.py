"""One scan-kernel variant over one resident buffer, a few launches (ncu target).
Usage: python tools/one_scan.py <variant|default> [MiB] [launches]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib

variant = sys.argv[1]
mib = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
if variant in ("", "default"):            # the shipped library's own choice
    os.environ.pop("NVME_STROM_SCAN_VARIANT", None)
else:
    os.environ["NVME_STROM_SCAN_VARIANT"] = variant
base_pages = 8192
host = np.empty(base_pages * 8192, dtype=np.uint8)
for i in range(base_pages):
    lib.strom_heapgen_page(0x5EED0001, i, 1, host[i * 8192:].ctypes.data)
buf = torch.from_numpy(host).cuda().repeat(mib // 64)
tot = torch.zeros(abi.T_COUNT, dtype=torch.int64, device="cuda")
for _ in range(reps):
    abi.check(abi.scan_lib().strom_scan_pages_dev(C.c_void_p(buf.data_ptr()), mib * 128, 0, 0, None, C.c_void_p(tot.data_ptr()), None), "scan")
torch.cuda.synchronize()
print(variant, tot.cpu().tolist()[:3])
