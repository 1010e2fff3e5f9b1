"""A consumer with its own (non-primary) CUDA context, as the reference's nvme_test has
(cuCtxCreate, utils/nvme_test.c:850): windows remember the owning context, engine threads bind
to it, and the kernels -- including their > 48KB shared-memory opt-in -- work in it."""
import ctypes as C
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

from conftest import ROOT

CODE = r'''
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
from cuda.bindings import driver as drv
from conftest import Oracle
from nvme_strom_b200 import abi
lib = abi.lib

def ok(r):
    assert r[0] == drv.CUresult.CUDA_SUCCESS, r[0]
    return r[1:] if len(r) > 2 else (r[1] if len(r) == 2 else None)

ok(drv.cuInit(0))
dev = ok(drv.cuDeviceGet(0))
try:
    ctx = ok(drv.cuCtxCreate(0, dev))
except TypeError:                                   # newer bindings: (ctxCreateParams, flags, dev)
    ctx = ok(drv.cuCtxCreate(None, 0, dev))
oracle = Oracle()
path = sys.argv[1]
n = 3000
data = oracle.gen_pages(0x5EED0001, 0, n, True)
data.tofile(path)
dptr = ok(drv.cuMemAlloc(n * 8192))
dtot = ok(drv.cuMemAlloc(128))
dres = ok(drv.cuMemAlloc(n * 8))
ok(drv.cuMemsetD8(dptr, 0x41, n * 8192))

# load through the ioctl API into memory owned by this context
m = abi.StromCmd__MapGpuMemory(); m.vaddress, m.length = int(dptr), n * 8192
abi.nvme_strom_ioctl(abi.STROM_IOCTL__MAP_GPU_MEMORY, m)
fd = os.open(path, os.O_RDONLY)
ids = (C.c_uint32 * n)(*range(n))
a = abi.StromCmd__MemCopySsdToGpu()
a.handle, a.file_desc, a.nr_chunks, a.chunk_sz = m.handle, fd, n, 8192
a.chunk_ids = C.cast(ids, C.POINTER(C.c_uint32))
abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_SSD2GPU, a)
w = abi.StromCmd__MemCopyWait(); w.dma_task_id = a.dma_task_id
abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_WAIT, w)
back = np.empty(n * 8192, dtype=np.uint8)
ok(drv.cuMemcpyDtoH(back.ctypes.data, dptr, n * 8192))
assert (back == data).all()

# kernels in this context: device-pointer level and handle level
want_res, want_tot = oracle.scan(data, 0)
abi.check(lib.strom_totals_reset_dev(C.c_void_p(int(dtot)), None), "reset")
abi.check(lib.strom_scan_pages_dev(C.c_void_p(int(dptr)), n, 0, 0, C.c_void_p(int(dres)), C.c_void_p(int(dtot)), None), "scan")
ok(drv.cuCtxSynchronize())
tot = np.empty(16, dtype=np.uint64); res = np.empty((n, 4), dtype=np.uint16)
ok(drv.cuMemcpyDtoH(tot.ctypes.data, dtot, 128)); ok(drv.cuMemcpyDtoH(res.ctypes.data, dres, n * 8))
assert [int(x) for x in tot] == want_tot and (res == want_res).all()
htot = abi.StromScanTotals()
abi.check(lib.strom_scan_pages(m.handle, 0, n, 0, 0, res.ctypes.data, C.byref(htot)), "scan handle")
assert htot.as_list() == want_tot and (res == want_res).all()
abi.check(lib.strom_verify_pages(m.handle, 0, m.handle, 0, n, None, C.byref(htot)), "verify")
assert htot.v[abi.T_MISMATCH_BLOCKS] == 0
u = abi.StromCmd__UnmapGpuMemory(); u.handle = m.handle
abi.nvme_strom_ioctl(abi.STROM_IOCTL__UNMAP_GPU_MEMORY, u)
print("context ok")
'''


def test_non_primary_context(tmp_path):
    code = CODE % (ROOT, ROOT)
    r = subprocess.run([sys.executable, "-c", code, str(tmp_path / "ctx.heap")], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0 and "context ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
