"""Striped sources (second half): the product's stripe map against the oracle's
restatement of strom_raid0_map_sector (kmod/nvme_strom.c:736-820), the striped planner's request
accounting against the oracle's planner fed the same layout as an extent list, and bytes landed
from a stripe set against the un-striped file.  The zones handed to the oracle are built here, in
the test, the way md builds them (drivers/md/raid0.c:create_strip_zones), so product and checker
do not share that code either."""
import ctypes as C
import errno
import mmap
import os

import numpy as np
import pytest

from test_planner import run_oracle


def md_zones(member_sects, chunk_sects):
    """(zones [(zone_end, dev_start, nb_dev)], devlist laid out zone-major with stride = members)"""
    size = [s // chunk_sects * chunk_sects for s in member_sects]
    zones, devlist, dev_start, zone_end = [], [], 0, 0
    while True:
        devs = [i for i, s in enumerate(size) if s > dev_start]
        if not devs:
            break
        smallest = min(size[i] for i in devs)
        zone_end += (smallest - dev_start) * len(devs)
        zones.append((zone_end, dev_start, len(devs)))
        devlist.extend(devs + [-1] * (len(size) - len(devs)))
        dev_start = smallest
    return zones, devlist


def oracle_map(oracle, zones, devlist, chunk_sects, sector, nr_sects):
    Z = oracle.lib_zone_array(zones)
    dl = (C.c_int * len(devlist))(*devlist)
    sec = C.c_uint64(sector)
    dev = C.c_int(-1)
    rc = oracle.lib.ora_raid0_map_sector(Z, len(zones), dl, chunk_sects, C.byref(sec), nr_sects, C.byref(dev))
    return rc, dev.value, sec.value


def product_map(strom, member_bytes, stripe_bytes, logical, length):
    arr = (C.c_uint64 * len(member_bytes))(*member_bytes)
    member, off, total = C.c_uint32(), C.c_uint64(), C.c_uint64()
    rc = strom.lib.strom_stripe_map_debug(arr, len(member_bytes), stripe_bytes, logical, length,
                                          C.byref(member), C.byref(off), C.byref(total))
    return rc, member.value, off.value, total.value


@pytest.fixture(scope="module")
def ora(oracle):
    class Zone(C.Structure):
        _fields_ = [("zone_end", C.c_uint64), ("dev_start", C.c_uint64), ("nb_dev", C.c_uint32)]
    oracle.lib.ora_raid0_map_sector.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64,
                                                C.POINTER(C.c_uint64), C.c_uint32, C.POINTER(C.c_int)]
    oracle.lib_zone_array = lambda zones: (Zone * len(zones))(*[Zone(*z) for z in zones])
    return oracle


def test_stripe_map_matches_the_reference_arithmetic(ora, strom):
    rng = np.random.default_rng(20)
    for trial in range(200):
        m = int(rng.integers(1, 7))
        chunk = int(rng.choice([4096, 8192, 65536, 131072, 524288, 12288, 40960]))     # power of two and not
        if trial % 3 == 0:
            sizes = [chunk * int(rng.integers(3, 40))] * m                              # equal members
        else:
            sizes = [chunk * int(rng.integers(1, 40)) + int(rng.integers(0, chunk)) for _ in range(m)]
        zones, devlist = md_zones([s // 512 for s in sizes], chunk // 512)
        total = zones[-1][0] * 512
        for _ in range(60):
            logical = int(rng.integers(0, total // 4096 + 2)) * 4096
            length = int(rng.choice([4096, 4096, 8192, 512]))
            want = oracle_map(ora, zones, devlist, chunk // 512, logical // 512, length // 512)
            got = product_map(strom, sizes, chunk, logical, length)
            assert got[3] == total
            assert got[0] == want[0], (trial, sizes, chunk, logical, length, got, want)
            if want[0] == 0:
                assert (got[1], got[2]) == (want[1], want[2] * 512), (trial, sizes, chunk, logical, got, want)
    # every byte of a set belongs to exactly one member position (a bijection onto the members' whole chunks)
    sizes, chunk = [5 * 8192, 9 * 8192 + 100, 2 * 8192, 9 * 8192], 8192
    seen = set()
    total = product_map(strom, sizes, chunk, 0, 4096)[3]
    for logical in range(0, total, 4096):
        rc, member, off, _ = product_map(strom, sizes, chunk, logical, 4096)
        assert rc == 0 and off + 4096 <= sizes[member] // chunk * chunk
        seen.add((member, off))
    assert len(seen) == total // 4096 == sum(s // chunk * chunk for s in sizes) // 4096
    assert product_map(strom, sizes, chunk, total, 4096)[0] == -errno.ERANGE
    assert product_map(strom, sizes, chunk, 4096, 8192)[0] == -errno.ESPIPE
    assert product_map(strom, sizes, 1000, 0, 4096)[0] == -errno.EINVAL


def run_product_striped(strom, to_gpu, ids, chunk_sz, relseg, member_bytes, stripe_bytes, dest_off, dest_len, seg,
                        piece_max=1 << 20):
    n = len(ids)
    ids = np.asarray(ids, dtype=np.uint32)
    counters = np.zeros(4, dtype=np.uint32)
    out = np.zeros(max(n, 1), dtype=np.uint32)
    pieces = np.zeros(4 * 65536, dtype=np.uint64)
    npieces = C.c_uint32()
    fn = strom.lib.strom_plan_chunks_striped_debug
    fn.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64,
                   C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                   C.c_void_p]
    mb = np.asarray(member_bytes, dtype=np.uint64)
    rc = fn(int(to_gpu), n, chunk_sz, relseg, ids.ctypes.data, mb.ctypes.data, len(member_bytes), stripe_bytes,
            dest_off, dest_len, seg, piece_max, counters.ctypes.data, out.ctypes.data, pieces.ctypes.data, 65536,
            C.byref(npieces))
    return rc, counters.tolist(), out[:n].tolist(), pieces[:4 * npieces.value].reshape(-1, 4).tolist()


@pytest.mark.parametrize("to_gpu", [True, False])
def test_striped_plans_match_the_oracle_planner(ora, strom, to_gpu):
    """the oracle's planner knows nothing of stripes; it is given the set as an extent list in which
    every member owns a far-apart sector range, so that 'same member and contiguous there' is exactly
    'contiguous on the device' -- the module's merge test at kmod/nvme_strom.c:1370-1381"""
    rng = np.random.default_rng(7 + int(to_gpu))
    for trial in range(120):
        m = int(rng.integers(1, 5))
        stripe = int(rng.choice([4096, 8192, 65536, 131072, 262144, 24576]))
        chunk_sz = int(rng.choice([4096, 8192, 16384, 131072]))
        per_member = int(rng.integers(4, 64))
        sizes = [stripe * per_member] * m if trial % 2 else [stripe * int(rng.integers(2, 64)) for _ in range(m)]
        total = product_map(strom, sizes, stripe, 0, 4096)[3]
        n_file_chunks = total // chunk_sz
        if n_file_chunks < 2:
            continue
        n = int(rng.integers(1, 80))
        start = int(rng.integers(0, n_file_chunks))
        ids = [(start + i) % n_file_chunks for i in range(n)] if trial % 3 else \
            rng.integers(0, n_file_chunks, n).tolist()
        extents = []
        for lc in range(0, total, stripe):
            rc, member, off, _ = product_map(strom, sizes, stripe, lc, 4096)
            assert rc == 0
            extents.append((lc // 4096, stripe // 4096, (member << 36) + off // 512))
        dest_len = n * chunk_sz + (1 << 20)
        seg = 0 if to_gpu else 4 << 20
        # every fourth trial addresses the set as a relation of relseg-chunk segments (ids wrap, :1526-1527)
        relseg = max(1, n_file_chunks // 2) if trial % 4 == 3 else 0
        if relseg:
            ids = [i + relseg * int(rng.integers(0, 3)) for i in (x % relseg for x in ids)]
        got = run_product_striped(strom, to_gpu, ids, chunk_sz, relseg, sizes, stripe, 0, dest_len, seg)
        want = run_oracle(ora, to_gpu, ids, chunk_sz, relseg, total, None, extents, 0, dest_len, seg)
        assert got[0] == want[0] == 0
        assert got[1] == want[1], (trial, sizes, stripe, chunk_sz, ids[:8], got[1], want[1])
        assert got[2] == want[2]
        # pieces: each inside one stripe chunk run of one member, covering the request exactly
        assert sum(p[2] for p in got[3]) == n * chunk_sz
        for off, dest, length, member in got[3]:
            assert member < m and off + length <= sizes[member] and length <= 1 << 20
            # a piece is contiguous inside ONE member file (several of its chunks when the request
            # strides over the others' chunks in step), never across members


def make_set(tmp_path, strom, n_pages, members, stripe):
    from nvme_strom_b200.loader import StripeSet, stripe_split
    flat = str(tmp_path / "flat.heap")
    assert strom.lib.strom_heapgen_file(flat.encode(), 0x5EED0001, 0, n_pages, 1, 2) == 0
    paths = [str(tmp_path / f"member{i}") for i in range(members)]
    stripe_split(flat, paths, stripe)
    return flat, paths, StripeSet(paths, stripe)


def test_ssd2ram_from_a_stripe_set_lands_the_flat_files_bytes(strom, tmp_path):
    from test_ioctl_cpu import addr_of, alloc_dmabuf, ssd2ram, wait
    flat, paths, sset = make_set(tmp_path, strom, 3 * 512, 3, 64 << 10)           # 12 MiB over 3 members, 64 KB stripes
    want = np.fromfile(flat, dtype=np.uint8)
    try:
        assert sset.size == want.size
        cf = strom.StromCmd__CheckFile()
        cf.fdesc = sset.token
        strom.nvme_strom_ioctl(strom.STROM_IOCTL__CHECK_FILE, cf)
        m = alloc_dmabuf(strom, 16 << 20)
        for ids, chunk_sz in ((list(range(1536)), 8192), (list(range(95, -1, -1)), 131072),
                              ([5, 1535, 0, 700, 701, 702, 64, 63], 8192), (list(range(0, 3072, 7)), 4096)):
            arg, _keep = ssd2ram(strom, sset.token, addr_of(m), ids, chunk_sz)
            assert wait(strom, arg.dma_task_id) == 0
            got = np.frombuffer(m, dtype=np.uint8, count=len(ids) * chunk_sz)
            exp = np.concatenate([want[i * chunk_sz:(i + 1) * chunk_sz] for i in ids])
            assert (got == exp).all()
            assert arg.nr_ssd2ram == len(ids) and arg.nr_ram2ram == 0
            # a request never spans two members: with 64 KB stripes no request is longer than 128 sectors
            assert arg.nr_dma_blocks == len(ids) * chunk_sz // 512
            assert arg.nr_dma_submit >= arg.nr_dma_blocks // 128
        # past the end of the set: ERANGE like a plain file (:1529)
        a = strom.StromCmd__MemCopySsdToRam()
        arr = (C.c_uint32 * 1)(1537)
        a.dest_uaddr, a.file_desc, a.nr_chunks, a.chunk_sz = addr_of(m), sset.token, 1, 8192
        a.chunk_ids = C.cast(arr, C.POINTER(C.c_uint32))
        assert strom.lib.nvme_strom_ioctl(strom.STROM_IOCTL__MEMCPY_SSD2RAM, C.byref(a)) == -1
        assert C.get_errno() == errno.ERANGE
    finally:
        sset.close()
    assert strom.lib.nvme_strom_stripe_close(sset.token if sset.token >= 0 else 10 ** 6) == -1


def test_stripe_open_errors(strom, tmp_path):
    p = str(tmp_path / "m0")
    open(p, "wb").write(bytes(1 << 20))
    fd = os.open(p, os.O_RDONLY)
    wfd = os.open(p, os.O_WRONLY)
    r, w = os.pipe()
    try:
        def attempt(fds, stripe):
            arr = (C.c_int * len(fds))(*fds)
            rc = strom.lib.nvme_strom_stripe_open(arr, len(fds), stripe)
            e = C.get_errno()
            if rc >= 0:
                strom.lib.nvme_strom_stripe_close(rc)
            return rc, e
        assert attempt([fd], 1000)[0] == -1 and attempt([fd], 1000)[1] == errno.EINVAL
        assert attempt([fd, 10 ** 6], 65536) == (-1, errno.EBADF)
        assert attempt([wfd], 65536) == (-1, errno.EACCES)
        assert attempt([r], 65536) == (-1, strom.ENOTSUPP)
        assert attempt([fd, fd], 2 << 20) == (-1, errno.EINVAL)      # no member holds a whole stripe
        assert strom.lib.nvme_strom_stripe_open(None, 0, 65536) == -1
        rc, _ = attempt([fd, fd], 65536)
        assert rc >= 0
    finally:
        for f in (fd, wfd, r, w):
            os.close(f)
