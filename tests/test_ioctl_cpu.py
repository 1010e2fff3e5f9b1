"""Host-side behaviour of nvme_strom_ioctl() that needs no GPU: CHECK_FILE,
ALLOC_DMA_BUFFER, MEMCPY_SSD2RAM, MEMCPY_WAIT, STAT_INFO, the window commands
on an empty registry, and the errno table of SURVEY.md 8b."""
import ctypes as C
import errno
import mmap
import os
import threading

import numpy as np
import pytest


def ioctl_errno(strom, cmd, arg):
    try:
        strom.nvme_strom_ioctl(cmd, arg)
        return 0
    except strom.StromError as e:
        return e.errno


@pytest.fixture(scope="module")
def heapfile(tmp_path_factory, oracle):
    path = str(tmp_path_factory.mktemp("heap") / "t.heap")
    assert oracle.lib.ora_gen_file(path.encode(), 0x5EED0001, 2048, 1) == 0     # 16 MiB
    return path


def alloc_dmabuf(strom, length, node=-1):
    arg = strom.StromCmd__AllocDMABuffer()
    arg.length, arg.node_id = length, node
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__ALLOC_DMA_BUFFER, arg)
    m = mmap.mmap(arg.dmabuf_fdesc, length, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
    os.close(arg.dmabuf_fdesc)        # ssd2ram_test.c:139-142: mmap, then close the fd
    return m


def addr_of(m):
    return C.addressof(C.c_char.from_buffer(m))


def ssd2ram(strom, fd, dest_addr, ids, chunk_sz=8192, relseg=0):
    arr = (C.c_uint32 * max(len(ids), 1))(*ids)
    arg = strom.StromCmd__MemCopySsdToRam()
    arg.dest_uaddr = dest_addr
    arg.file_desc = fd
    arg.nr_chunks = len(ids)
    arg.chunk_sz = chunk_sz
    arg.relseg_sz = relseg
    arg.chunk_ids = C.cast(arr, C.POINTER(C.c_uint32))
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__MEMCPY_SSD2RAM, arg)
    return arg, arr


def wait(strom, task_id):
    w = strom.StromCmd__MemCopyWait()
    w.dma_task_id = task_id
    w.status = 77
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__MEMCPY_WAIT, w)
    return w.status


def test_check_file(strom, heapfile, tmp_path):
    fd = os.open(heapfile, os.O_RDONLY)
    arg = strom.StromCmd__CheckFile()
    arg.fdesc = fd
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__CHECK_FILE, arg)
    assert arg.support_dma64 == 1 and arg.numa_node_id >= -1
    os.close(fd)
    arg.fdesc = 9999
    assert ioctl_errno(strom, strom.STROM_IOCTL__CHECK_FILE, arg) == errno.EBADF
    wfd = os.open(heapfile, os.O_WRONLY)
    arg.fdesc = wfd
    assert ioctl_errno(strom, strom.STROM_IOCTL__CHECK_FILE, arg) == errno.EACCES
    os.close(wfd)
    small = tmp_path / "small"
    small.write_bytes(b"x" * 100)
    sfd = os.open(str(small), os.O_RDONLY)
    arg.fdesc = sfd
    assert ioctl_errno(strom, strom.STROM_IOCTL__CHECK_FILE, arg) == strom.ENOTSUPP
    os.close(sfd)
    dfd = os.open(str(tmp_path), os.O_RDONLY)
    arg.fdesc = dfd
    assert ioctl_errno(strom, strom.STROM_IOCTL__CHECK_FILE, arg) == strom.ENOTSUPP
    os.close(dfd)


def test_unknown_command_and_null_arg(strom):
    arg = strom.StromCmd__CheckFile()
    assert ioctl_errno(strom, 0x53FF, arg) == errno.EINVAL
    assert strom.lib.nvme_strom_ioctl(strom.STROM_IOCTL__CHECK_FILE, None) == -1
    assert C.get_errno() == errno.EFAULT


def test_alloc_dma_buffer_rules(strom):
    arg = strom.StromCmd__AllocDMABuffer()
    for bad in (0, 100, 4096 + 512):
        arg.length, arg.node_id = bad, -1
        assert ioctl_errno(strom, strom.STROM_IOCTL__ALLOC_DMA_BUFFER, arg) == errno.EINVAL
    arg.length, arg.node_id = 1 << 20, 4096
    assert ioctl_errno(strom, strom.STROM_IOCTL__ALLOC_DMA_BUFFER, arg) == errno.EINVAL
    arg.length, arg.node_id = 1 << 20, -1
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__ALLOC_DMA_BUFFER, arg)
    assert os.fstat(arg.dmabuf_fdesc).st_size == 1 << 20
    assert os.readlink(f"/proc/self/fd/{arg.dmabuf_fdesc}").startswith("/memfd:dmabuf")
    os.close(arg.dmabuf_fdesc)


def test_ssd2ram_lands_the_file_bytes(strom, heapfile):
    """ssd2ram_test's unit: 1 MB = 128 chunks per ioctl into a ring over the buffer"""
    data = np.fromfile(heapfile, dtype=np.uint8)
    buf = alloc_dmabuf(strom, 8 << 20)
    base = addr_of(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    tasks = []
    for unit in range(8):
        ids = list(range(unit * 128, unit * 128 + 128))
        arg, keep = ssd2ram(strom, fd, base + unit * (1 << 20), ids)
        assert arg.nr_ram2ram + arg.nr_ssd2ram == 128
        assert arg.nr_dma_blocks == arg.nr_ssd2ram * 16
        assert arg.dma_task_id not in (0, 2 ** 64 - 1)
        assert list(keep) == ids                       # SSD2RAM never rewrites chunk_ids
        tasks.append(arg.dma_task_id)
    os.close(fd)                                        # the library holds its own reference
    for t in tasks:
        assert wait(strom, t) == 0
    got = np.frombuffer(buf, dtype=np.uint8)
    assert (got == data[:8 << 20]).all()
    del got
    buf.close()


def test_ssd2ram_scattered_ids_relseg_and_chunk_sizes(strom, heapfile):
    data = np.fromfile(heapfile, dtype=np.uint8)
    buf = alloc_dmabuf(strom, 4 << 20)
    base = addr_of(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    rng = np.random.default_rng(3)
    for chunk_sz in (4096, 8192, 65536, 131072):
        n_file = data.size // chunk_sz
        ids = rng.integers(0, n_file, 24).tolist()
        arg, _ = ssd2ram(strom, fd, base + 4096, ids, chunk_sz)
        assert wait(strom, arg.dma_task_id) == 0
        got = np.frombuffer(buf, dtype=np.uint8)
        for k, cid in enumerate(ids):
            assert (got[4096 + k * chunk_sz:4096 + (k + 1) * chunk_sz] ==
                    data[cid * chunk_sz:(cid + 1) * chunk_sz]).all(), (chunk_sz, k)
        del got
    # relseg_sz: block numbers of a multi-segment relation wrap into this file (:1812)
    ids = [131072 + 5, 2 * 131072 + 9, 7]
    arg, _ = ssd2ram(strom, fd, base, ids, 8192, relseg=131072)
    assert wait(strom, arg.dma_task_id) == 0
    got = np.frombuffer(buf, dtype=np.uint8)
    for k, cid in enumerate([5, 9, 7]):
        assert (got[k * 8192:(k + 1) * 8192] == data[cid * 8192:(cid + 1) * 8192]).all()
    del got
    os.close(fd)
    buf.close()


def test_ssd2ram_chunk_straddling_eof_is_zero_filled(strom, tmp_path):
    p = tmp_path / "odd"
    body = bytes(range(256)) * 40 + b"tail"              # 10244 bytes
    p.write_bytes(body)
    buf = alloc_dmabuf(strom, 1 << 20)
    fd = os.open(str(p), os.O_RDONLY)
    arg, _ = ssd2ram(strom, fd, addr_of(buf), [1], 8192)   # fpos 8192 <= size: accepted (:1814)
    assert wait(strom, arg.dma_task_id) == 0
    assert bytes(buf[:8192]) == body[8192:] + bytes(8192 - (len(body) - 8192))
    os.close(fd)
    buf.close()


def test_ssd2ram_errors(strom, heapfile):
    buf = alloc_dmabuf(strom, 1 << 20)
    base = addr_of(buf)
    fd = os.open(heapfile, os.O_RDONLY)

    def err(dest, ids, chunk_sz=8192, f=fd):
        try:
            ssd2ram(strom, f, dest, ids, chunk_sz)
            return 0
        except strom.StromError as e:
            return e.errno

    plain = np.zeros(1 << 20, dtype=np.uint8)
    assert err(plain.ctypes.data, [0]) == errno.EINVAL            # not a strom DMA buffer
    priv = mmap.mmap(-1, 1 << 20)
    assert err(addr_of(priv), [0]) == errno.EINVAL
    assert err(base, list(range(129))) == errno.ERANGE            # runs off the mapping
    assert err(base + (1 << 20) - 8192, [0, 1]) == errno.ERANGE
    assert err(base, [0], 8192 + 512) == errno.EINVAL
    assert err(base, [0], 1 << 18) == errno.EINVAL
    assert err(base + 512, [0]) == errno.EINVAL                   # unaligned destination
    assert err(base, [1 << 20]) == errno.ERANGE                   # beyond the file
    assert err(base, [0], f=9999) == errno.EBADF
    os.close(fd)
    buf.close()


def test_wait_semantics(strom):
    assert wait(strom, 0xDEADBEE5) == 0                 # unknown id == already finished (:1189)
    assert wait(strom, 2 ** 64 - 1) == 0                # PG's "no task" sentinel is harmless


def test_stat_info(strom, heapfile):
    arg = strom.StromCmd__StatInfo()
    arg.version = 2
    assert ioctl_errno(strom, strom.STROM_IOCTL__STAT_INFO, arg) == errno.EINVAL
    arg.version = 1
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__STAT_INFO, arg)
    before = (arg.nr_ssd2gpu, arg.nr_setup_prps)
    buf = alloc_dmabuf(strom, 1 << 20)
    fd = os.open(heapfile, os.O_RDONLY)
    a, _ = ssd2ram(strom, fd, addr_of(buf), list(range(128)))
    assert wait(strom, a.dma_task_id) == 0
    os.close(fd)
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__STAT_INFO, arg)
    assert arg.nr_ssd2gpu == before[0] + 1 and arg.nr_setup_prps == before[1] + 1
    assert arg.tsc > 0 and arg.max_dma_count >= 1
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__STAT_INFO, arg)
    assert arg.max_dma_count == 0                        # read-and-reset (:2009)
    buf.close()


def test_window_commands_without_windows(strom):
    un = strom.StromCmd__UnmapGpuMemory()
    un.handle = 0x1234
    assert ioctl_errno(strom, strom.STROM_IOCTL__UNMAP_GPU_MEMORY, un) == errno.ENOENT
    info = strom.StromCmd__InfoGpuMemory(4)
    info.handle = 0x1234
    assert ioctl_errno(strom, strom.STROM_IOCTL__INFO_GPU_MEMORY, info) == errno.ENOENT
    g = strom.StromCmd__MemCopySsdToGpu()
    g.handle = 0x1234
    g.nr_chunks = 0
    assert ioctl_errno(strom, strom.STROM_IOCTL__MEMCPY_SSD2GPU, g) == errno.ENOENT
    host = np.zeros(1 << 16, dtype=np.uint8)
    m = strom.StromCmd__MapGpuMemory()
    m.vaddress, m.length = host.ctypes.data, host.size
    assert ioctl_errno(strom, strom.STROM_IOCTL__MAP_GPU_MEMORY, m) == errno.EINVAL   # not device memory
    m.vaddress, m.length = 0, 0
    assert ioctl_errno(strom, strom.STROM_IOCTL__MAP_GPU_MEMORY, m) == errno.EINVAL


def test_concurrent_ssd2ram_from_many_threads(strom, heapfile):
    """ssd2ram_test runs N workers pulling units from a shared cursor (:180)"""
    data = np.fromfile(heapfile, dtype=np.uint8)
    buf = alloc_dmabuf(strom, 16 << 20)
    base = addr_of(buf)
    cursor = iter(range(16))
    lock = threading.Lock()
    errors = []

    def worker():
        fd = os.open(heapfile, os.O_RDONLY)
        try:
            while True:
                with lock:
                    unit = next(cursor, None)
                if unit is None:
                    break
                arg, _ = ssd2ram(strom, fd, base + unit * (1 << 20), list(range(unit * 128, unit * 128 + 128)))
                if wait(strom, arg.dma_task_id) != 0:
                    errors.append(unit)
        except Exception as e:          # noqa: BLE001
            errors.append(e)
        finally:
            os.close(fd)

    ts = [threading.Thread(target=worker) for _ in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors
    got = np.frombuffer(buf, dtype=np.uint8)
    assert (got == data).all()
    del got
    buf.close()


def test_sessions(strom):
    tok = strom.lib.nvme_strom_open()
    assert tok > 0
    assert strom.lib.nvme_strom_close(tok) == 0
    assert strom.lib.nvme_strom_close(-1) == -1
