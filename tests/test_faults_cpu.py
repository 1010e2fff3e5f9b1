"""Error and signal paths of the task table (no GPU): a failed piece surfaces at
MEMCPY_WAIT as -1/EIO with the status, once (kmod/nvme_strom.c:1151-1171); closing
the session discards never-waited failures (:2063-2091); a signal interrupts WAIT
with EINTR (:1191-1195).  The switches are read once per process, so each case
runs in a child process."""
import os
import subprocess
import sys
import textwrap

from conftest import ROOT

PRE = f"""
import ctypes as C, errno, mmap, os, signal, sys, threading, time
sys.path.insert(0, {ROOT!r})
from nvme_strom_b200 import abi
lib = abi.lib
def dmabuf(n):
    a = abi.StromCmd__AllocDMABuffer(); a.length, a.node_id = n, -1
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__ALLOC_DMA_BUFFER, a)
    m = mmap.mmap(a.dmabuf_fdesc, n, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE); os.close(a.dmabuf_fdesc)
    return m, C.addressof(C.c_char.from_buffer(m))
def ssd2ram(fd, dest, ids, token=0):
    arr = (C.c_uint32 * len(ids))(*ids)
    g = abi.StromCmd__MemCopySsdToRam()
    g.dest_uaddr, g.file_desc, g.nr_chunks, g.chunk_sz = dest, fd, len(ids), 8192
    g.chunk_ids = C.cast(arr, C.POINTER(C.c_uint32))
    rc = lib.nvme_strom_ioctl_on(token, abi.STROM_IOCTL__MEMCPY_SSD2RAM, C.byref(g))
    assert rc == 0, C.get_errno()
    return g.dma_task_id
def wait(tid):
    w = abi.StromCmd__MemCopyWait(); w.dma_task_id = tid
    rc = lib.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_WAIT, C.byref(w))
    return rc, C.get_errno() if rc else 0, w.status
path = sys.argv[1]
fd = os.open(path, os.O_RDONLY)
"""


def run_child(body, path, **env):
    code = PRE + textwrap.dedent(body)
    r = subprocess.run([sys.executable, "-c", code, path], env=dict(os.environ, **env), capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


def make_file(tmp_path, strom, pages=2048):
    path = str(tmp_path / "f.heap")
    assert strom.lib.strom_heapgen_file(path.encode(), 7, 0, pages, 0, 2) == 0
    return path


def test_failed_piece_reports_eio_once_with_status(tmp_path, strom):
    out = run_child("""
        m, base = dmabuf(16 << 20)
        tid = ssd2ram(fd, base, list(range(2048)))          # several pieces; every 2nd one fails
        rc, e, status = wait(tid)
        assert (rc, e, status) == (-1, errno.EIO, -errno.EIO), (rc, e, status)
        assert wait(tid) == (0, 0, 0)                         # the entry was consumed (:1165-1167)
        print("ok")
    """, make_file(tmp_path, strom), NVME_STROM_FAULT_EVERY="2", NVME_STROM_SLOT_KB="1024")
    assert "ok" in out


def test_closing_the_session_discards_unwaited_failures(tmp_path, strom):
    out = run_child("""
        m, base = dmabuf(16 << 20)
        tok = lib.nvme_strom_open()
        tid = ssd2ram(fd, base, list(range(2048)), tok)
        time.sleep(0.5)                                      # let it fail in the background
        assert lib.nvme_strom_close(tok) == 0                # strom_proc_release :2063-2091
        assert wait(tid) == (0, 0, 0)                         # nothing left to report
        print("ok")
    """, make_file(tmp_path, strom), NVME_STROM_FAULT_EVERY="1", NVME_STROM_SLOT_KB="1024")
    assert "ok" in out


def test_wait_returns_eintr_on_a_signal(tmp_path, strom):
    """holds a task open with a FIFO-slow source: a pipe-backed file cannot be used (CHECK
    rejects it), so the task is kept busy by a huge request on one reader while SIGUSR1 arrives"""
    out = run_child("""
        signal.signal(signal.SIGUSR1, lambda *a: None)
        m, base = dmabuf(256 << 20)
        tids = [ssd2ram(fd, base, list(range(32768))) for _ in range(40)]   # ~10 GiB of reads on 1 reader
        main = threading.get_ident()
        threading.Timer(0.05, lambda: signal.pthread_kill(main, signal.SIGUSR1)).start()
        rc, e, status = wait(tids[-1])
        assert (rc, e) == (-1, errno.EINTR), (rc, e)
        for t in tids:                                       # PG's loop: retry after EINTR (:1039-1040)
            while True:
                rc, e, status = wait(t)
                if rc == 0: break
                assert e == errno.EINTR
        print("ok")
    """, make_file(tmp_path, strom, 32768), NVME_STROM_HOST_READERS="1")
    assert "ok" in out
