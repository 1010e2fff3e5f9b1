"""Two CUDA contexts on one GPU in one process -- the primary one (what any runtime-API user gets)
and one made with cuCtxCreate (what the reference's nvme_test makes, utils/nvme_test.c:850).
Streams, allocations, symbol addresses and shared-memory opt-ins belong to a context: the
library's scratch, its static array's address and its kernels' attributes must follow the context
a call is made in, not the device.  Calls alternate between the two contexts and each result is
checked against the CPU oracle."""
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

from conftest import ROOT

CODE = r"""
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
primary = ok(drv.cuDevicePrimaryCtxRetain(dev))
try:
    second = ok(drv.cuCtxCreate(0, dev))
except TypeError:
    second = ok(drv.cuCtxCreate(None, 0, dev))
oracle = Oracle()
n = 2500
state = {}
for name, ctx, seed in (("primary", primary, 0x5EED0001), ("second", second, 0x5EED0040)):
    ok(drv.cuCtxSetCurrent(ctx))
    data = oracle.gen_pages(seed, 11, n, True)
    dptr = ok(drv.cuMemAlloc(n * 8192))
    dcopy = ok(drv.cuMemAlloc(n * 8192))
    ok(drv.cuMemcpyHtoD(dptr, data.ctypes.data, n * 8192))
    other = data.copy(); other[1234 * 8192 + 99] ^= 4
    ok(drv.cuMemcpyHtoD(dcopy, other.ctypes.data, n * 8192))
    m = abi.StromCmd__MapGpuMemory(); m.vaddress, m.length = int(dptr), n * 8192
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__MAP_GPU_MEMORY, m)
    m2 = abi.StromCmd__MapGpuMemory(); m2.vaddress, m2.length = int(dcopy), n * 8192
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__MAP_GPU_MEMORY, m2)
    state[name] = dict(ctx=ctx, data=data, dptr=dptr, h=m.handle, h2=m2.handle, want=oracle.scan(data, 11))

def one_round(name, k):
    s = state[name]
    ok(drv.cuCtxSetCurrent(s["ctx"]))
    want_res, want_tot = s["want"]
    res = np.zeros((n, 4), dtype=np.uint16); tot = abi.StromScanTotals()
    # handle level: library-owned stream + scratch of THIS context
    abi.check(lib.strom_scan_pages(s["h"], 0, n, 11, 0, res.ctypes.data, C.byref(tot)), "scan_pages " + name)
    assert tot.as_list() == want_tot and (res == want_res).all(), (name, k, "handle")
    # static array: this context's instance, this context's address
    res[:] = 0; tot = abi.StromScanTotals()
    abi.check(lib.strom_scan_pages_static(C.c_void_p(int(s["dptr"])), n, 11, res.ctypes.data, C.byref(tot), None),
              "scan_pages_static " + name)
    assert tot.as_list() == want_tot and (res == want_res).all(), (name, k, "static")
    # byte-compare of the two windows of this context
    vt = abi.StromScanTotals()
    abi.check(lib.strom_verify_pages(s["h"], 0, s["h2"], 0, n, None, C.byref(vt)), "verify " + name)
    assert (vt.v[abi.T_MISMATCH_BLOCKS], vt.v[abi.T_FIRST_MISMATCH]) == (1, 1234 * 8192 + 96), (name, k)

for k in range(3):
    one_round("primary", k)
    one_round("second", k)
# windows of different contexts cannot be compared with each other
vt = abi.StromScanTotals()
assert lib.strom_verify_pages(state["primary"]["h"], 0, state["second"]["h"], 0, n, None, C.byref(vt)) == -1
print("two contexts ok")
"""


def test_calls_alternate_between_two_contexts_on_one_device(tmp_path):
    code = CODE % (ROOT, ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "two contexts ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
