"""Why does the resident-scan time per launch differ between a lone process and a torchrun rank?
Times K back-to-back scan launches over a resident 1 GiB buffer for several K, optionally after
NCCL initialisation and a dist.barrier() (as bench.py does at N > 1).  Tuning aid.
Usage: python tools/value_probe.py [device]      or under torchrun (uses LOCAL_RANK, NCCL)"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib

dist_mode = "RANK" in os.environ
dev_index = int(os.environ.get("LOCAL_RANK", sys.argv[1] if len(sys.argv) > 1 else 0))
torch.cuda.set_device(dev_index)
dev = torch.device("cuda", dev_index)
if dist_mode:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)

base_pages = 8192
host = np.empty(base_pages * 8192, dtype=np.uint8)
for i in range(base_pages):
    lib.strom_heapgen_page(0x5EED0001, i, 1, host[i * 8192:].ctypes.data)
buf = torch.from_numpy(host).to(dev).repeat(16)
tot = torch.zeros(abi.T_COUNT, dtype=torch.int64, device=dev)
stream = torch.cuda.Stream(device=dev)
p_buf, p_tot, st = C.c_void_p(buf.data_ptr()), C.c_void_p(tot.data_ptr()), C.c_void_p(stream.cuda_stream)


def launch():
    abi.check(lib.strom_scan_pages_dev(p_buf, 131072, 0, 0, None, p_tot, st), "scan")


def barrier():
    if dist_mode:
        dist.barrier()
    torch.cuda.synchronize()


for _ in range(5):
    launch()
out = []
for k in (10, 20, 100, 10):
    for pre in ("barrier", "barrier+1ms"):
        barrier()
        if pre != "barrier":
            time.sleep(0.001)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        per = [torch.cuda.Event(enable_timing=True) for _ in range(k + 1)]
        per[0].record(stream)
        for i in range(k):
            launch()
            per[i + 1].record(stream)
        torch.cuda.synchronize()
        each = [per[i].elapsed_time(per[i + 1]) * 1e3 for i in range(k)]
        out.append(f"K={k:<3} after {pre:<11}: avg {sum(each) / k:6.1f} us  first {each[0]:6.1f}  median {sorted(each)[k // 2]:6.1f}  max {max(each):6.1f}")
tag = f"[{'rank' if dist_mode else 'solo'} dev {dev_index}]"
print("\n".join(tag + " " + o for o in out), flush=True)
if dist_mode:
    dist.destroy_process_group()
