"""Every kernel of the library once, on small inputs, for compute-sanitizer:
  compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
(one tool per gpurun call, see the profiling guide)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib

n = 300
host = np.empty(n * 8192, dtype=np.uint8)
for i in range(n):
    lib.strom_heapgen_page(0x5EED0001, i, 1, host[i * 8192:].ctypes.data)
a = torch.from_numpy(host).cuda()
b = a.clone()
b[77 * 8192 + 5] ^= 1
tot = torch.zeros(abi.T_COUNT, dtype=torch.int64, device="cuda")
res = torch.zeros(n * 4, dtype=torch.int16, device="cuda")
pa, pb, pt, pr = (C.c_void_p(x.data_ptr()) for x in (a, b, tot, res))
ref = None
for v in ("ldg:8:2", "ldg:4:3", "ldgdyn", "col:12", "col:8", "tma:8:3", "tma:16:1", "tma:12:2"):
    os.environ["NVME_STROM_SCAN_VARIANT"] = v
    abi.check(lib.strom_totals_reset_dev(pt, None), "reset")
    abi.check(abi.scan_lib().strom_scan_pages_dev(pa, n, 0, 0, pr, pt, None), v)
    torch.cuda.synchronize()
    t = tot.cpu().tolist()
    ref = ref or t
    assert t == ref, v
os.environ.pop("NVME_STROM_SCAN_VARIANT")
abi.check(lib.strom_verify_pages_dev(pa, pb, n, None, 0, pt, None), "verify")
abi.check(abi.scan_lib().strom_scan_verify_pages_dev(pa, pb, n, 0, 0, 0, pr, pt, None), "fused")
items = torch.zeros(n * 2042 * 3, dtype=torch.int32, device="cuda")
offs = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
abi.check(lib.strom_extract_items_dev(pa, n, 0, C.c_void_p(items.data_ptr()), n * 2042, C.c_void_p(offs.data_ptr()),
                                      C.c_void_p(cnt.data_ptr()), None), "extract")
hres = np.zeros((n, 4), dtype=np.uint16)
htot = abi.StromScanTotals()
abi.check(lib.strom_scan_pages_static(pa, n, 0, hres.ctypes.data, C.byref(htot), None), "static")
torch.cuda.synchronize()
print("sanitize smoke ok", int(cnt.item()), tot.cpu().tolist()[10:12])
