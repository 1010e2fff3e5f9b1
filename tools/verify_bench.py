"""Config C2's second half on one GPU: load + scan + byte-compare end to end (every segment crosses
PCIe twice: MEMCPY_SSD2GPU, and MEMCPY_SSD2RAM + H2D for the comparison copy), for several sizes of
the comparison copy's H2D pieces and several engine settings.  Tuning aid.
  python tools/verify_bench.py [dir]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib
from nvme_strom_b200.loader import SegmentPipeline, heapgen_file

d = sys.argv[1] if len(sys.argv) > 1 else "/tmp"
path = os.path.join(d, "verify_bench.heap")
heapgen_file(path, 0x5EED0001, 131072)
GIB = 1 << 30
try:
    for env in ({}, {"NVME_STROM_HOST_COPY": "stream"}, {"NVME_STROM_HOST_COPY": "ntstore"},
                {"NVME_STROM_HOST_COPY": "stream", "NVME_STROM_HOST_READERS": "4"},
                {"NVME_STROM_HOST_COPY": "stream", "NVME_STROM_READERS": "6", "NVME_STROM_HOST_READERS": "6"},
                {"NVME_STROM_HOST_COPY": "stream", "NVME_STROM_COPY": "stream"}):
        for k in ("NVME_STROM_READERS", "NVME_STROM_HOST_READERS", "NVME_STROM_HOST_COPY", "NVME_STROM_COPY"):
            os.environ.pop(k, None)
        os.environ.update(env)
        abi.check(lib.strom_debug_reconfigure(), "reconfigure")
        for chunk_mb, slots in ((32, 6), (4, 6), (32, 6)):
            vp = SegmentPipeline(0, 32 << 20, slots, verify=True, ram_copy_chunk=chunk_mb << 20)
            vp.run(path)
            vp.run(path)
            t0 = time.perf_counter()
            for _ in range(5):
                tot, _ = vp.run(path)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / 5
            assert tot[abi.T_MISMATCH_BLOCKS] == 0 and tot[abi.T_PAGES] == 131072
            st = abi.engine_state(0)
            vp.close()
            print(f"{str(env or 'default'):68} ref H2D pieces {chunk_mb:2d} MB, {slots:2d} slots: {GIB / dt / 1e9:6.2f} GB/s "
                  f"(readers {st['readers_active']}/{st['readers']})", flush=True)
finally:
    os.unlink(path)
