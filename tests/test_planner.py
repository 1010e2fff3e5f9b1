"""The product's chunk planner against the oracle's restatement of
do_memcpy_ssd2gpu / do_memcpy_ssd2ram / memcpy_from_nvme_ssd
(reference kmod/nvme_strom.c:1487-1604, :1766-1884, :1302-1405)."""
import ctypes as C
import errno

import numpy as np
import pytest


def run_product(strom, to_gpu, ids, chunk_sz, relseg, fsize, state, extents, dest_off, dest_len, seg, piece_max=1 << 20):
    n = len(ids)
    ids = np.asarray(ids, dtype=np.uint32)
    counters = np.zeros(4, dtype=np.uint32)
    out = np.zeros(max(n, 1), dtype=np.uint32)
    pieces = np.zeros(3 * 4096, dtype=np.uint64)
    npieces = C.c_uint32()
    ext = None
    if extents:
        ext = np.asarray([[a * 4096, s * 512, c * 4096] for (a, c, s) in extents], dtype=np.uint64).ravel()
    fn = strom.lib.strom_plan_chunks_debug
    fn.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p,
                   C.c_uint64, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32,
                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    rc = fn(int(to_gpu), n, chunk_sz, relseg, ids.ctypes.data, fsize,
            state.ctypes.data if state is not None else None, 0 if state is None else state.size,
            ext.ctypes.data if ext is not None else None, len(extents) if extents else 0,
            dest_off, dest_len, seg, piece_max, counters.ctypes.data, out.ctypes.data,
            pieces.ctypes.data, 4096, C.byref(npieces))
    return rc, counters.tolist(), out[:n].tolist(), pieces[:3 * npieces.value].reshape(-1, 3).tolist()


def run_oracle(oracle, to_gpu, ids, chunk_sz, relseg, fsize, state, extents, dest_off, dest_len, seg):
    n = len(ids)
    ids = np.asarray(ids, dtype=np.uint32)
    out = np.zeros(max(n, 1), dtype=np.uint32)
    runs = (oracle.Run * 8192)()
    plan = oracle.Plan()
    ext = None
    if extents:
        ext = (oracle.Extent * len(extents))(*[oracle.Extent(a, c, s) for (a, c, s) in extents])
    rc = oracle.lib.ora_plan_chunks(int(to_gpu), n, chunk_sz, relseg, ids.ctypes.data, fsize,
                                    state.ctypes.data if state is not None else None,
                                    0 if state is None else state.size, ext, len(extents) if extents else 0,
                                    dest_off, dest_len, seg, out.ctypes.data, runs, 8192, C.byref(plan))
    return rc, [plan.nr_cached, plan.nr_direct, plan.nr_dma_submit, plan.nr_dma_blocks], out[:n].tolist(), \
        [(runs[i].src_sector, runs[i].nr_sectors, runs[i].dest_offset) for i in range(min(plan.nr_runs, 8192))]


def test_sequential_cold_segment(oracle, strom):
    """nvme_test's case: 4096 sequential 8KB chunks, cold cache, contiguous file"""
    ids = list(range(8192, 8192 + 4096))
    args = (True, ids, 8192, 0, 1 << 30, None, None, 0, 32 << 20, 0)
    rc, cnt, out, pieces = run_product(strom, *args)
    assert (rc, cnt, out) == run_oracle(oracle, *args)[:3]
    assert cnt == [0, 4096, 256, 65536]                  # 256 x 128KB requests, 64Ki sectors
    assert out == ids
    assert len(pieces) == 32 and all(p[2] == 1 << 20 for p in pieces)
    assert pieces[0][0] == 8192 * 8192 and pieces[0][1] == 0


@pytest.mark.parametrize("to_gpu", [True, False])
def test_random_plans_match_oracle(oracle, strom, to_gpu):
    rng = np.random.default_rng(42 + int(to_gpu))
    for trial in range(300):
        chunk_sz = int(rng.choice([4096, 8192, 16384, 65536, 131072]))
        relseg = int(rng.choice([0, 0, 16, 131072]))
        n = int(rng.integers(0, 65))
        n_file_chunks = int(rng.integers(8, 200))
        fsize = n_file_chunks * chunk_sz - int(rng.integers(0, 2)) * 100
        kind = rng.integers(0, 3)
        if kind == 0:
            start = int(rng.integers(0, n_file_chunks))
            ids = [(start + i) % n_file_chunks for i in range(n)]
        elif kind == 1:
            ids = rng.integers(0, n_file_chunks, n).tolist()
        else:
            ids = sorted(rng.choice(n_file_chunks, min(n, n_file_chunks), replace=False).tolist())
        state = None
        if rng.integers(0, 2):
            state = rng.choice([0, 0, 1, 1, 2], n_file_chunks * (chunk_sz // 4096) + 3).astype(np.uint8)
            if rng.integers(0, 2):                       # runs of cached pages, as a real cache has
                state = np.repeat(rng.choice([0, 1], state.size // 8 + 1), 8)[:state.size].astype(np.uint8)
        extents = None
        if rng.integers(0, 2):
            extents, page, sector = [], 0, int(rng.integers(0, 1 << 20)) * 8
            total_pages = n_file_chunks * (chunk_sz // 4096) + 4
            while page < total_pages:
                cnt = int(rng.integers(1, 64))
                extents.append((page, cnt, sector))
                page += cnt
                sector += cnt * 8 + int(rng.choice([0, 0, 8, 4096]))
        dest_off = int(rng.integers(0, 4)) * 4096 * (1 if not to_gpu else int(rng.integers(1, 3)))
        dest_len = dest_off + len(ids) * chunk_sz + int(rng.integers(0, 3)) * 4096
        seg = 0 if to_gpu else 4 << 20
        args = (to_gpu, ids, chunk_sz, relseg, fsize, state, extents, dest_off, dest_len, seg)
        prc, pcnt, pout, ppieces = run_product(strom, *args)
        orc, ocnt, oout, oruns = run_oracle(oracle, *args)
        assert prc == orc, (trial, args)
        if prc == 0:
            assert pcnt == ocnt, (trial, args)
            if to_gpu:
                assert pout == oout, trial
            # the transfer pieces cover exactly the direct chunks, contiguously, once each
            assert sum(p[2] for p in ppieces) == pcnt[1] * chunk_sz
            assert sum(r[1] for r in oruns) * 512 == pcnt[1] * chunk_sz


def test_argument_errors_match_oracle(oracle, strom):
    E = errno
    cases = [
        (True, [0], 8192 + 512, 0, 1 << 20, None, None, 0, 1 << 20, 0, -E.EINVAL),    # not a 4KB multiple
        (True, [0], 2048, 0, 1 << 20, None, None, 0, 1 << 20, 0, -E.EINVAL),           # < 4KB
        (True, [0], 256 << 10, 0, 1 << 20, None, None, 0, 1 << 20, 0, -E.EINVAL),      # > 128KB
        (False, [0], 8192, 0, 1 << 20, None, None, 100, 1 << 20, 0, -E.EINVAL),        # unaligned host dest
        (True, [0, 1], 8192, 0, 1 << 20, None, None, 0, 8192, 0, -E.ERANGE),           # window too small
        (True, [1000], 8192, 0, 1 << 20, None, None, 0, 1 << 20, 0, -E.ERANGE),        # fpos > size
        (True, [128], 8192, 0, 1 << 20, None, None, 0, 1 << 20, 0, 0),                 # fpos == size passes (:1529)
        (True, [1000], 8192, 100, 1 << 20, None, None, 0, 1 << 20, 0, 0),              # relseg wraps the id
    ]
    for *args, want in cases:
        assert run_product(strom, *args)[0] == want, args
        assert run_oracle(oracle, *args)[0] == want, args


def test_cached_chunks_go_to_the_tail(oracle, strom):
    """kmod/nvme_strom.c:1546-1571: direct chunks packed from the head, cached ones from the tail"""
    state = np.zeros(64, dtype=np.uint8)
    state[4:6] = 1          # chunk 2 fully cached
    state[10] = 1           # chunk 5 half cached: 1 > threshold(1)? no -> direct
    state[12] = 2           # chunk 6 has one dirty page -> cached
    ids = [0, 1, 2, 3, 5, 6, 7]
    args = (True, ids, 8192, 0, 1 << 20, state, None, 0, 1 << 20, 0)
    rc, cnt, out, pieces = run_product(strom, *args)
    assert rc == 0 and cnt[:2] == [2, 5]
    assert out == [0, 1, 3, 5, 7, 6, 2]
    assert (rc, cnt, out) == run_oracle(oracle, *args)[:3]


def test_fiemap_view_of_a_real_file(strom, tmp_path):
    """the extent map the planner uses in place of ext4_get_block / xfs_get_blocks
    (kmod/nvme_strom.c:153-165): sorted, non-overlapping, covers the file; on file systems
    without FIEMAP (tmpfs) the planner falls back to an identity map (0 extents)"""
    import os
    p = tmp_path / "ext.heap"
    assert strom.lib.strom_heapgen_file(str(p).encode(), 3, 0, 4096, 0, 2) == 0      # 32 MiB
    fd = os.open(str(p), os.O_RDONLY)
    os.fsync(fd)
    buf = np.zeros(3 * 4096, dtype=np.uint64)
    fn = strom.lib.strom_file_extents_debug
    fn.argtypes = [C.c_int, C.c_void_p, C.c_uint32]
    n = fn(fd, buf.ctypes.data, 4096)
    os.close(fd)
    assert n >= 0
    if n == 0:
        pytest.skip("file system has no FIEMAP (identity map is used)")
    ext = buf[:3 * n].reshape(-1, 3)
    assert (ext[:, 2] > 0).all() and (ext[:, 0] % 4096 == 0).all()
    assert (ext[1:, 0] >= ext[:-1, 0] + ext[:-1, 2]).all()           # sorted, no overlap
    assert ext[0, 0] == 0 and ext[-1, 0] + ext[-1, 2] >= 32 << 20    # covers the file


def test_requests_break_at_extent_boundaries(oracle, strom):
    """physically discontiguous extents end a <=128KB request early (:1370-1381)"""
    ids = list(range(64))                                            # 512 KB, 8KB chunks
    contiguous = [(0, 128, 1000)]
    broken = [(0, 10, 1000), (10, 50, 90000), (60, 68, 5000)]        # 4KB pages: breaks at 40KB and 240KB
    a = run_product(strom, True, ids, 8192, 0, 1 << 20, None, contiguous, 0, 1 << 20, 0)
    b = run_product(strom, True, ids, 8192, 0, 1 << 20, None, broken, 0, 1 << 20, 0)
    assert a[1] == [0, 64, 4, 1024] and b[1][2] == 6 and b[1][3] == 1024
    assert b[:3] == run_oracle(oracle, True, ids, 8192, 0, 1 << 20, None, broken, 0, 1 << 20, 0)[:3]
    assert len(a[3]) == len(b[3]) == 1                               # one transfer piece either way: pread does not care
