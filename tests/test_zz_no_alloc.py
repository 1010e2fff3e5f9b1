"""strom_scan_pages_static(): the scan for callers that must not allocate.

Who needs it: a process whose GPU memory has already been handed out in full
(a database backend with a device-memory quota, a job living next to a caching
allocator that owns the rest of the card) cannot let a library call cudaMalloc
behind its back, and a caller on a latency-critical path cannot take the
device-wide synchronisation that cudaMalloc / cudaFree bring.  The regular
strom_scan_pages() allocates its counters and its per-page staging on demand;
this variant keeps them in a fixed 1 MiB array inside the library.

The array is shared by every call and keeps what the last call left in it, so
the test makes two calls in a row with different inputs and sizes and checks
each against the existing entry points, bit for bit (the tolerance every
integer test here uses).  Collected last (test_zz_*) so that a run stopping at
the first failure has already been through everything else; it also passes
when run alone in a fresh process.
"""
import ctypes as C
import errno

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def scan_static(abi, pages_dev, n_pages, first_blkno, want_results=True):
    res = np.full((max(n_pages, 1), 4), 0xABCD, dtype=np.uint16)
    tot = abi.StromScanTotals()
    rc = abi.lib.strom_scan_pages_static(C.c_void_p(pages_dev.data_ptr()), n_pages, first_blkno & 0xFFFFFFFF,
                                         res.ctypes.data if want_results else None, C.byref(tot), None)
    assert rc == 0, abi.lib.strom_last_error()
    return res[:n_pages], tot.as_list()


def test_two_calls_in_a_row_match_the_existing_entry_points(oracle):
    import torch
    from gpu_util import map_gpu, scan_dev, to_dev, unmap_gpu
    from nvme_strom_b200 import abi

    # first call: 5000 mixed-validity pages, so every counter is non-zero afterwards
    a = to_dev(oracle.gen_pages(0x5EED0001, 300, 5000, mixed=True))
    # second call: fewer, clean pages at other block numbers -- anything left over from
    # the first call (counters, records 700..4999) would show up as a difference
    b = to_dev(oracle.gen_pages(0x5EED0040, 9000, 700, mixed=False))
    for dev_buf, n, first in ((a, 5000, 300), (b, 700, 9000)):
        got_res, got_tot = scan_static(abi, dev_buf, n, first)
        want_res, want_tot = scan_dev(dev_buf, n, first)              # strom_scan_pages_dev
        assert (got_res == want_res).all()
        assert got_tot == want_tot
        # the CPU oracle as well, so that this file stands on its own
        ora_res, ora_tot = oracle.scan(dev_buf.cpu().numpy(), first)
        assert (got_res == ora_res).all() and got_tot == ora_tot
        # and the allocating handle-level call, which is the one this variant stands in for
        m = map_gpu(dev_buf)
        res2 = np.zeros((n, 4), dtype=np.uint16)
        tot2 = abi.StromScanTotals()
        abi.check(abi.lib.strom_scan_pages(m.handle, 0, n, first, 0, res2.ctypes.data, C.byref(tot2)), "scan")
        unmap_gpu(m.handle)
        assert (got_res == res2).all() and got_tot == tot2.as_list()
    assert got_tot[abi.T_PAGES] == 700 and got_tot[abi.T_CHECKSUM_FAIL] == 0
    # totals only, and an empty range: the counters really are cleared each time
    _, tot_only = scan_static(abi, a, 5000, 300, want_results=False)
    assert tot_only == scan_dev(a, 5000, 300)[1]
    _, empty = scan_static(abi, a, 0, 0)
    assert empty[:11] == [0] * 11 and empty[abi.T_FIRST_MISMATCH] == 2 ** 64 - 1
    del torch


def test_device_memory_is_untouched_by_the_calls(oracle):
    """The one property this entry point exists for: after its first call on the device (set-up:
    symbol address, kernel load) no call changes the amount of free device memory, whatever the
    size of the request -- small, large, with and without per-page records, growing and shrinking.
    cudaMemGetInfo is read straight from the runtime so that torch's own caching cannot hide an
    allocation made by the library."""
    import torch
    from nvme_strom_b200 import abi
    buf = torch.from_numpy(oracle.gen_pages(0x5EED0001, 0, 4096, mixed=True)).cuda().repeat(8)   # 32768 pages
    host = buf[:4096 * 8192].cpu().numpy()
    _, want_small = oracle.scan(host[:100 * 8192], 7)
    scan_static(abi, buf, 1, 0)                                   # first call: set-up happens here
    scan_static(abi, buf, 32768, 0)                               # and the long-launch kernel gets loaded
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for n, want_results in ((100, True), (32768, True), (3, False), (20000, True), (32768, False), (100, True)):
        _, tot = scan_static(abi, buf, n, 7, want_results=want_results)
        assert torch.cuda.mem_get_info()[0] == free0, (n, want_results)
    assert tot == want_small
    # the allocating entry point, for contrast, does take device memory for a request it has not seen
    from gpu_util import map_gpu, unmap_gpu
    m = map_gpu(buf)
    res = np.zeros((32768, 4), dtype=np.uint16)
    t = abi.StromScanTotals()
    abi.check(abi.lib.strom_scan_pages(m.handle, 0, 32768, 0, 0, res.ctypes.data, C.byref(t)), "scan")
    unmap_gpu(m.handle)


def test_limit_is_enforced_with_a_clear_error(oracle):
    from gpu_util import to_dev
    from nvme_strom_b200 import abi
    small = to_dev(oracle.gen_pages(1, 0, 4))
    tot = abi.StromScanTotals()
    rc = abi.lib.strom_scan_pages_static(C.c_void_p(small.data_ptr()), abi.STATIC_SCRATCH_MAX_PAGES + 1, 0,
                                         None, C.byref(tot), None)
    assert rc == -1 and C.get_errno() == errno.EINVAL
    msg = abi.lib.strom_last_error().decode()
    assert "131072" in msg and "STROM_STATIC_SCRATCH_MAX_PAGES" in msg
    assert abi.lib.strom_scan_pages_static(C.c_void_p(small.data_ptr() + 8), 1, 0, None, C.byref(tot), None) == -1
    assert abi.lib.strom_scan_pages_static(C.c_void_p(small.data_ptr()), 1, 0, None, None, None) == -1


def test_at_the_limit(oracle):
    """exactly STROM_STATIC_SCRATCH_MAX_PAGES pages (1 GiB) go through"""
    import torch
    from gpu_util import scan_dev
    from nvme_strom_b200 import abi
    n = abi.STATIC_SCRATCH_MAX_PAGES
    base = torch.from_numpy(oracle.gen_pages(0x5EED0010, 0, 2048, mixed=True)).cuda()
    buf = base.repeat(n // 2048)
    got_res, got_tot = scan_static(abi, buf, n, 0)
    want_res, want_tot = scan_dev(buf, n, 0)
    assert got_tot == want_tot and (got_res == want_res).all()
