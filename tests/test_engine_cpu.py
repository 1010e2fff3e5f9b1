"""Host-side behaviour of the transfer engine that needs no GPU: read faults on the mapped source
become the task's EIO (the module reports NVMe errors the same way, kmod/nvme_strom.c:993-1039,
:1163-1168) instead of a SIGBUS that kills the host process; the engine sizes itself by the cores
the process is allowed on (CPU placement precedent: utils/ssd2ram_test.c:66-119); the page-cache
read calibration; the cached source views can be dropped; DMA buffers take a NUMA node."""
import os

from test_faults_cpu import make_file, run_child


def test_truncated_source_reports_eio_instead_of_sigbus(tmp_path, strom):
    """one reader, a 512 MiB request, and the file is cut to nothing while the reader is at work: every
    page it touches from then on raises SIGBUS in the library's thread, which must surface as
    -1/EIO with status -EIO at MEMCPY_WAIT -- and the process must still be alive to say so"""
    out = run_child("""
        m, base = dmabuf(512 << 20)
        tid = ssd2ram(fd, base, list(range(65536)))
        os.truncate(path, 0)                                  # after the size was taken, before the copy is done
        rc, e, status = wait(tid)
        assert (rc, e, status) == (-1, errno.EIO, -errno.EIO), (rc, e, status)
        # the library keeps working afterwards, on a file that is whole
        lib.strom_heapgen_file(path.encode(), 7, 0, 256, 0, 1)
        fd2 = os.open(path, os.O_RDONLY)
        tid = ssd2ram(fd2, base, list(range(256)))
        assert wait(tid) == (0, 0, 0)
        assert m[:256 * 8192] == open(path, "rb").read()
        print("ok")
    """, make_file(tmp_path, strom, 65536), NVME_STROM_HOST_READERS="1")
    assert "ok" in out


def test_foreign_sigbus_still_reaches_the_previous_handler(tmp_path, strom):
    """the guard only answers for faults inside the library's own copies: a SIGBUS anywhere else goes
    to the handler that was installed before the library's"""
    out = run_child("""
        import faulthandler
        hits = []
        signal.signal(signal.SIGBUS, lambda *a: hits.append(1))
        m, base = dmabuf(1 << 20)
        assert wait(ssd2ram(fd, base, list(range(128)))) == (0, 0, 0)     # installs the guard (first mapped source)
        os.kill(os.getpid(), signal.SIGBUS)
        time.sleep(0.05)
        assert hits == [1], hits
        print("ok")
    """, make_file(tmp_path, strom, 256))
    assert "ok" in out


def test_engine_counts_the_cores_it_is_allowed_on(tmp_path, strom):
    cpus = sorted(os.sched_getaffinity(0))
    if len(cpus) < 4:
        import pytest
        pytest.skip("needs 4 cores")
    body = """
        st = abi.engine_state(0)
        print("state", st["allowed_cpus"], st["slot_bytes"], st["slots"], st["readers_max"], st["adaptive"])
    """
    # a rank confined to 4 cores (one of eight on a 32-core host): 4 x 2 MB ring, every core reads
    os.sched_setaffinity(0, cpus[:4])
    try:
        out = run_child(body, make_file(tmp_path, strom, 8))
    finally:
        os.sched_setaffinity(0, cpus)
    assert "state 4 2097152 4 4 1" in out
    # an explicit list wins over the mask; a fixed reader count switches self-adjustment off
    out = run_child(body, make_file(tmp_path, strom, 8), NVME_STROM_CPUS=f"{cpus[0]}-{cpus[1]},{cpus[3]}",
                    NVME_STROM_READERS="5")
    assert "state 3 1572864 4 5 0" in out
    out = run_child(body, make_file(tmp_path, strom, 8), NVME_STROM_CPUS="0-63")
    assert "state 64 4194304 4 16 1" in out
    out = run_child(body, make_file(tmp_path, strom, 8), NVME_STROM_CPUS="0-7")
    assert "state 8 4194304 4 7 1" in out


def test_pagecache_read_calibration(tmp_path, strom):
    path = make_file(tmp_path, strom, 8192)                     # 64 MiB
    fd = os.open(path, os.O_RDONLY)
    try:
        for threads in (1, 3, 0):
            assert strom.pagecache_read_GBps(fd, 0, 64 << 20, threads) > 0.05
        assert strom.pagecache_read_GBps(fd, 32 << 20, 1 << 40, 2) > 0       # clipped to the file
        assert strom.pagecache_read_GBps(fd, 0, 64 << 20, 2, dest_bytes_per_thread=16 << 20) > 0.05   # plain copy
        assert strom.lib.nvme_strom_unpin_file(fd) == 0
        assert strom.pagecache_read_GBps(fd, 0, 64 << 20, 2) > 0             # mapped again on demand
    finally:
        os.close(fd)
    assert strom.lib.nvme_strom_unpin_file(fd) == -1


def test_dma_buffer_on_a_numa_node(strom):
    import mmap
    arg = strom.StromCmd__AllocDMABuffer()
    arg.length, arg.node_id = 8 << 20, 0
    strom.nvme_strom_ioctl(strom.STROM_IOCTL__ALLOC_DMA_BUFFER, arg)
    m = mmap.mmap(arg.dmabuf_fdesc, arg.length, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
    assert os.readlink(f"/proc/self/fd/{arg.dmabuf_fdesc}").startswith("/memfd:dmabuf0:8MB")
    os.close(arg.dmabuf_fdesc)
    m[:] = b"\x5a" * arg.length                                 # every page can be allocated on the node
    assert m[arg.length - 1] == 0x5a
    m.close()


def test_gds_probe_says_why_and_the_staged_path_is_what_runs(strom, tmp_path):
    """north_star (2): cuFile only if the host exposes it.  This image has neither nvidia-fs nor a
    cuFile data path in the library: the probe answers 0 with a reason, and a load goes through the
    staged engine all the same (MEMCPY_SSD2RAM here; the GPU tests cover MEMCPY_SSD2GPU)."""
    import ctypes as C
    buf = C.create_string_buffer(256)
    strom.lib.nvme_strom_gds_status.argtypes = [C.c_char_p, C.c_size_t]
    assert strom.lib.nvme_strom_gds_status(buf, 256) == 0
    text = buf.value.decode()
    assert text.startswith("GDS ") and "staged path" in text
    if not os.path.exists("/proc/driver/nvidia-fs/version"):
        assert "nvidia-fs" in text or "libcufile" in text
    assert strom.lib.nvme_strom_gds_status(None, 0) == 0
    out = run_child("""
        m, base = dmabuf(1 << 20)
        assert wait(ssd2ram(fd, base, list(range(128)))) == (0, 0, 0)
        assert m[:128 * 8192] == open(path, "rb").read()[:128 * 8192]
        print("ok")
    """, make_file(tmp_path, strom, 256), NVME_STROM_VERBOSE="1")
    assert "ok" in out


def test_rank_cpu_shares():
    """one rank per GPU: ranks whose GPUs report the same neighbourhood split it evenly in rank order;
    the shares are disjoint and cover the cores (utils/ssd2ram_test.c:66-119 is the precedent)"""
    from nvme_strom_b200 import multigpu
    none8 = [None] * 8
    shares = [multigpu.rank_cpus(r, 8, allowed=range(32), local_lists=none8) for r in range(8)]
    assert all(len(s) == 4 for s in shares) and sorted(sum(shares, [])) == list(range(32))
    two_nodes = [list(range(0, 16))] * 4 + [list(range(16, 32))] * 4
    shares = [multigpu.rank_cpus(r, 8, allowed=range(32), local_lists=two_nodes) for r in range(8)]
    assert shares[0] == [0, 1, 2, 3] and shares[4] == [16, 17, 18, 19] and shares[7] == [28, 29, 30, 31]
    # a job confined to some cores keeps to them; uneven splits lose nothing
    shares = [multigpu.rank_cpus(r, 3, allowed=[2, 3, 5, 7, 11, 13, 17], local_lists=[None] * 3) for r in range(3)]
    assert sorted(sum(shares, [])) == [2, 3, 5, 7, 11, 13, 17] and all(shares)
    assert multigpu.rank_cpus(0, 1, allowed=range(6)) == list(range(6))
    assert multigpu.rank_cpus(3, 4, allowed=[0, 1], local_lists=[None] * 4) == [0, 1]     # fewer cores than ranks
    assert multigpu.parse_cpu_list("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]


def test_filesystem_lookup(tmp_path):
    from nvme_strom_b200 import calibrate
    mnt, fstype = calibrate.filesystem_of("/dev/shm")
    assert fstype in ("tmpfs", "?") and mnt.startswith("/")
    assert calibrate.filesystem_of(str(tmp_path))[1] != ""
