"""MEMCPY_SSD2GPU / MEMCPY_WAIT end to end on a real device: loaded bytes are
the file's bytes (the reference's own criterion, utils/nvme_test.c:249-266),
the post-conditions of kmod/nvme_strom.c:1487-1604 hold, errors match the
module's table, and the segment pipeline reproduces the oracle's counters."""
import ctypes as C
import errno
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from conftest import ROOT


@pytest.fixture(scope="module")
def heapfile(tmp_path_factory, oracle):
    path = str(tmp_path_factory.mktemp("gheap") / "g.heap")
    assert oracle.lib.ora_gen_file(path.encode(), 0x5EED0001, 16384, 1) == 0     # 128 MiB, mixed
    return path


@pytest.fixture(scope="module")
def filedata(heapfile):
    return np.fromfile(heapfile, dtype=np.uint8)


def err_of(fn, *a, **kw):
    from nvme_strom_b200 import abi
    try:
        fn(*a, **kw)
        return 0
    except abi.StromError as e:
        return e.errno


def test_map_list_info_unmap():
    import torch
    from gpu_util import map_gpu, unmap_gpu
    from nvme_strom_b200 import abi
    buf = torch.empty(3 << 20, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf, offset=4096, length=(2 << 20))
    assert m.handle not in (0, 2 ** 64 - 1) and m.gpu_page_sz == 65536
    vaddr = buf.data_ptr() + 4096
    assert m.gpu_npages == ((vaddr & 0xFFFF) + (2 << 20) + 65535) // 65536
    lst = abi.StromCmd__ListGpuMemory(0)
    assert err_of(abi.nvme_strom_ioctl, abi.STROM_IOCTL__LIST_GPU_MEMORY, lst) == errno.ENOBUFS
    assert lst.nitems >= 1                                 # nvme_test.c:619-627 relies on this
    lst = abi.StromCmd__ListGpuMemory(lst.nitems)
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__LIST_GPU_MEMORY, lst)
    assert m.handle in list(lst.handles)[:lst.nitems]
    info = abi.StromCmd__InfoGpuMemory(1)
    info.handle = m.handle
    assert err_of(abi.nvme_strom_ioctl, abi.STROM_IOCTL__INFO_GPU_MEMORY, info) == errno.ENOBUFS
    assert info.nitems == m.gpu_npages
    info = abi.StromCmd__InfoGpuMemory(info.nitems)
    info.handle = m.handle
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__INFO_GPU_MEMORY, info)
    assert info.map_offset == vaddr & 0xFFFF and info.map_length == info.map_offset + (2 << 20)
    assert info.owner == os.geteuid() and info.gpu_page_sz == 65536 and info.version >= 1
    assert info.paddrs[0] == vaddr & ~0xFFFF and info.paddrs[1] == info.paddrs[0] + 65536
    unmap_gpu(m.handle)
    assert err_of(unmap_gpu, m.handle) == errno.ENOENT
    # a range running past the allocation is refused
    m2 = abi.StromCmd__MapGpuMemory()
    big = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
    m2.vaddress, m2.length = big.data_ptr(), (64 << 20) + (1 << 30)
    assert err_of(abi.nvme_strom_ioctl, abi.STROM_IOCTL__MAP_GPU_MEMORY, m2) == errno.EINVAL


def test_sequential_segment_lands_file_bytes(heapfile, filedata):
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    seg = 32 << 20
    buf = torch.full((seg + 8192,), 0x41, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    ids = list(range(4096, 8192))
    a, keep = ssd2gpu(m.handle, 8192, fd, ids)
    os.close(fd)                                           # library keeps its own reference
    assert (a.nr_ram2gpu, a.nr_ssd2gpu) == (0, 4096)
    assert a.nr_dma_blocks == 4096 * 16 and 256 <= a.nr_dma_submit <= 4096
    assert list(keep) == ids
    assert wait(a.dma_task_id) == 0
    got = buf.cpu().numpy()
    assert (got[:8192] == 0x41).all()                      # nothing before the offset was touched
    assert (got[8192:] == filedata[4096 * 8192:8192 * 8192]).all()
    assert wait(a.dma_task_id) == 0                        # a finished id is "unknown": 0 again
    unmap_gpu(m.handle)


@pytest.mark.parametrize("chunk_sz", [4096, 8192, 32768, 131072])
def test_scattered_chunks_and_relseg(heapfile, filedata, chunk_sz):
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    rng = np.random.default_rng(chunk_sz)
    n_file = filedata.size // chunk_sz
    ids = rng.integers(0, n_file, 100).tolist() + list(range(10, 40)) + [n_file - 1]
    buf = torch.zeros(len(ids) * chunk_sz + 65536, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf, offset=4096)                          # window not 64KB aligned
    fd = os.open(heapfile, os.O_RDONLY)
    a, keep = ssd2gpu(m.handle, 0, fd, ids, chunk_sz)
    assert wait(a.dma_task_id) == 0
    got = buf.cpu().numpy()[4096:]
    for k, cid in enumerate(list(keep)):
        assert (got[k * chunk_sz:(k + 1) * chunk_sz] == filedata[cid * chunk_sz:(cid + 1) * chunk_sz]).all(), k
    # relseg_sz: ids of a later 1GB segment file wrap into this one (:1526-1527)
    rel = n_file // 2
    ids2 = [rel * 3 + 5, rel * 7 + 1, 2]
    a, keep = ssd2gpu(m.handle, 0, fd, ids2, chunk_sz, relseg=rel)
    assert wait(a.dma_task_id) == 0
    got = buf.cpu().numpy()[4096:]
    for k, cid in enumerate([5, 1, 2]):
        assert (got[k * chunk_sz:(k + 1) * chunk_sz] == filedata[cid * chunk_sz:(cid + 1) * chunk_sz]).all()
    assert list(keep) == ids2                              # ids are echoed, not reduced
    os.close(fd)
    unmap_gpu(m.handle)


def test_six_slots_in_flight(heapfile, filedata):
    """nvme_test's default geometry: 6 tasks to disjoint offsets of one window (:40-41)"""
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    seg = 16 << 20
    buf = torch.zeros(6 * seg, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    tasks = []
    for s in range(6):
        ids = list(range(s * 2048, (s + 1) * 2048))
        tasks.append(ssd2gpu(m.handle, s * seg, fd, ids))
    for a, _ in tasks:
        assert wait(a.dma_task_id) == 0
    assert (buf.cpu().numpy() == filedata[:6 * seg]).all()
    os.close(fd)
    unmap_gpu(m.handle)


def test_errors(heapfile):
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu
    buf = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    assert err_of(ssd2gpu, m.handle + 1, 0, fd, [0]) == errno.ENOENT
    assert err_of(ssd2gpu, m.handle, 0, fd, [0], 8192 + 512) == errno.EINVAL
    assert err_of(ssd2gpu, m.handle, 0, fd, [0], 2048) == errno.EINVAL
    assert err_of(ssd2gpu, m.handle, 0, fd, [0], 256 << 10) == errno.EINVAL
    assert err_of(ssd2gpu, m.handle, 0, fd, list(range(129))) == errno.ERANGE      # 129 x 8KB > 1MB
    assert err_of(ssd2gpu, m.handle, (1 << 20) - 4096, fd, [0]) == errno.ERANGE
    assert err_of(ssd2gpu, m.handle, 0, fd, [1 << 24]) == errno.ERANGE             # beyond the file
    assert err_of(ssd2gpu, m.handle, 0, 9999, [0]) == errno.EBADF
    wfd = os.open(heapfile, os.O_WRONLY)
    assert err_of(ssd2gpu, m.handle, 0, wfd, [0]) == errno.EACCES
    os.close(wfd)
    # a failed call writes nothing back (:1665-1677)
    ids = (C.c_uint32 * 2)(7, 1 << 24)
    from nvme_strom_b200 import abi
    a = abi.StromCmd__MemCopySsdToGpu()
    a.handle, a.file_desc, a.nr_chunks, a.chunk_sz = m.handle, fd, 2, 8192
    a.chunk_ids = C.cast(ids, C.POINTER(C.c_uint32))
    a.dma_task_id, a.nr_ssd2gpu = 0x1111, 0x2222
    assert err_of(abi.nvme_strom_ioctl, abi.STROM_IOCTL__MEMCPY_SSD2GPU, a) == errno.ERANGE
    assert (a.dma_task_id, a.nr_ssd2gpu, list(ids)) == (0x1111, 0x2222, [7, 1 << 24])
    os.close(fd)
    unmap_gpu(m.handle)


def test_handle_level_scan_and_verify(heapfile, filedata, oracle):
    """the drop-in consumer's view: window handles in, host structs out"""
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    from nvme_strom_b200 import abi
    from nvme_strom_b200.abi import lib
    n = 2048
    a_buf = torch.zeros(n * 8192, dtype=torch.uint8, device="cuda")
    b_buf = torch.from_numpy(filedata[1000 * 8192:(1000 + n) * 8192].copy()).cuda()
    ma, mb = map_gpu(a_buf), map_gpu(b_buf)
    fd = os.open(heapfile, os.O_RDONLY)
    t, _ = ssd2gpu(ma.handle, 0, fd, list(range(1000, 1000 + n)))
    assert wait(t.dma_task_id) == 0
    res = np.zeros((n, 4), dtype=np.uint16)
    tot = abi.StromScanTotals()
    abi.check(lib.strom_scan_pages(ma.handle, 0, n, 1000, 0, res.ctypes.data, C.byref(tot)), "scan")
    want_res, want_tot = oracle.scan(filedata[1000 * 8192:(1000 + n) * 8192], 1000)
    assert (res == want_res).all() and tot.as_list() == want_tot
    abi.check(lib.strom_verify_pages(ma.handle, 0, mb.handle, 0, n, None, C.byref(tot)), "verify")
    assert tot.v[abi.T_MISMATCH_BLOCKS] == 0 and tot.v[abi.T_FIRST_MISMATCH] == 2 ** 64 - 1
    b_buf[77 * 8192 + 100] ^= 1
    abi.check(lib.strom_verify_pages(ma.handle, 0, mb.handle, 0, n, None, C.byref(tot)), "verify")
    assert tot.v[abi.T_MISMATCH_BLOCKS] == 1 and tot.v[abi.T_FIRST_MISMATCH] == 77 * 8192 + 96
    # range and handle errors
    assert lib.strom_scan_pages(ma.handle, 0, n + 1, 0, 0, None, C.byref(tot)) == -1
    assert C.get_errno() == errno.ERANGE
    assert lib.strom_scan_pages(0x99, 0, 1, 0, 0, None, C.byref(tot)) == -1
    assert C.get_errno() == errno.ENOENT
    os.close(fd)
    unmap_gpu(ma.handle)
    unmap_gpu(mb.handle)


@pytest.mark.parametrize("segment_mb,slots,verify", [(4, 6, False), (32, 2, False), (8, 3, True)])
def test_segment_pipeline_totals(heapfile, filedata, oracle, segment_mb, slots, verify):
    from nvme_strom_b200.loader import SegmentPipeline
    seg = segment_mb << 20
    pipe = SegmentPipeline(0, seg, slots, verify=verify, keep_results=True)
    totals, counters = pipe.run(heapfile)
    n_seg = filedata.size // seg
    want_res, want_tot = oracle.scan(filedata[:n_seg * seg], 0)
    if verify:
        want_tot[10] = 0
    assert totals == want_tot
    assert (pipe.results_host() == want_res).all()
    assert counters["nr_ssd2gpu"] == n_seg * seg // 8192 and counters["nr_ram2gpu"] == 0
    # a sub-range, as one rank of a multi-GPU run sees it
    totals2, _ = pipe.run(heapfile, first_segment=1, n_segments=2)
    _, want2 = oracle.scan(filedata[seg:3 * seg], seg // 8192)
    if verify:
        want2[10] = 0
    assert totals2 == want2
    pipe.close()


def test_page_cache_split_mode(heapfile, filedata, tmp_path):
    """NVME_STROM_PGCACHE_SPLIT=1: chunks resident in the page cache come back through
    wb_buffer from the tail, the rest land on the device from the head, chunk_ids is
    rewritten (kmod/nvme_strom.c:1540-1572).  Runs in a child: the switch is read once."""
    code = r'''
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
from gpu_util import map_gpu, ssd2gpu, wait
path = %r
data = np.fromfile(path, dtype=np.uint8)
fd = os.open(path, os.O_RDONLY)
n = 512
os.posix_fadvise(fd, 0, 0, os.POSIX_FADV_DONTNEED)
os.pread(fd, 8192 * 100, 8192 * 50)                 # warm chunks 50..149
buf = torch.zeros(n * 8192, dtype=torch.uint8, device="cuda")
wb = torch.zeros(n * 8192, dtype=torch.uint8).pin_memory()
m = map_gpu(buf)
a, ids = ssd2gpu(m.handle, 0, fd, list(range(n)), wb=wb.data_ptr())
assert wait(a.dma_task_id) == 0
ids = list(ids)
assert a.nr_ram2gpu + a.nr_ssd2gpu == n and sorted(ids) == list(range(n))
dev, host = buf.cpu().numpy(), wb.numpy()
for k in range(a.nr_ssd2gpu):
    assert (dev[k*8192:(k+1)*8192] == data[ids[k]*8192:(ids[k]+1)*8192]).all(), k
for k in range(n - a.nr_ram2gpu, n):
    assert (host[k*8192:(k+1)*8192] == data[ids[k]*8192:(ids[k]+1)*8192]).all(), k
assert ids[:a.nr_ssd2gpu] == sorted(ids[:a.nr_ssd2gpu])
print("split", a.nr_ram2gpu, a.nr_ssd2gpu)
''' % (ROOT, ROOT, heapfile)
    env = dict(os.environ, NVME_STROM_PGCACHE_SPLIT="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "split" in r.stdout


@pytest.mark.parametrize("env", [{"NVME_STROM_IO": "direct"}, {"NVME_STROM_SRC": "pread"}, {"NVME_STROM_FIEMAP": "0"},
                                 {"NVME_STROM_READERS": "1", "NVME_STROM_SLOT_KB": "128", "NVME_STROM_SLOTS": "2"},
                                 {"NVME_STROM_WC": "1", "NVME_STROM_SUB_KB": "64", "NVME_STROM_SLOTS": "3"},
                                 {"NVME_STROM_READERS": "12", "NVME_STROM_SLOT_KB": "512", "NVME_STROM_SPIN_US": "50"},
                                 {"NVME_STROM_COPY": "stream"}, {"NVME_STROM_COPY": "ntstore", "NVME_STROM_CPUS": "0-3"}])
def test_io_modes_land_the_same_bytes(heapfile, env):
    """O_DIRECT, pread, no FIEMAP, a single reader with tiny slots: same bytes, same counters.
    The switches are read once per process, so each mode runs in a child."""
    code = r'''
import os, sys
import numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
from gpu_util import map_gpu, ssd2gpu, wait
path = %r
data = np.fromfile(path, dtype=np.uint8)
fd = os.open(path, os.O_RDONLY)
rng = np.random.default_rng(9)
ids = list(range(4096, 8192)) + rng.integers(0, 16384, 200).tolist()
buf = torch.zeros(len(ids) * 8192, dtype=torch.uint8, device="cuda")
m = map_gpu(buf)
a, keep = ssd2gpu(m.handle, 0, fd, ids)
assert wait(a.dma_task_id) == 0
got = buf.cpu().numpy().reshape(-1, 8192)
want = data.reshape(-1, 8192)[np.array(list(keep))]
assert (got == want).all()
assert a.nr_ssd2gpu == len(ids) and a.nr_dma_blocks == len(ids) * 16
print("ok", a.nr_dma_submit)
''' % (ROOT, ROOT, heapfile)
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr


def test_engine_reports_what_it_decided(heapfile, filedata):
    """the self-adjusting engine after a run of loads: it counts the cores this process may use, keeps
    its reader count inside [2, readers], and accounts for every staged slot"""
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    from nvme_strom_b200 import abi
    n = min(8192, filedata.size // 8192)
    buf = torch.zeros(n * 8192, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    before = abi.engine_state(0)
    for _ in range(6):
        a, keep = ssd2gpu(m.handle, 0, fd, list(range(n)))
        assert wait(a.dma_task_id) == 0
    os.close(fd)
    st = abi.engine_state(0)
    unmap_gpu(m.handle)
    assert (buf.cpu().numpy() == filedata[:n * 8192]).all()
    assert st["allowed_cpus"] == len(os.sched_getaffinity(0))
    assert st["readers"] == st["readers_max"] >= 2 and 2 <= st["readers_active"] <= st["readers"]
    assert st["slots_staged"] - before["slots_staged"] >= 6 * n * 8192 // st["slot_bytes"]
    assert st["slots_starved"] <= st["slots_staged"]
    if os.environ.get("NVME_STROM_READERS"):
        assert not st["adaptive"] and st["grown"] == st["shrunk"] == 0
