"""Parity of the page-scan kernels (sm_100a) with the CPU oracle: per-page
records and global counters, bit-exact, for every kernel variant."""
import os
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

VARIANTS = ["tma:8:3", "tma:4:6", "tma:6:4", "tma:12:2", "tma:7:4", "tma:16:1", "tma:8:2", "tma:14:2",
            "ldg:8:2", "ldg:4:2", "ldg:4:3", "ldg:8:1", "ldg:4:1", "ldgdyn", "col:12", "col:10", "col:8", "col:6", "col:4",
            "roll:16:2", "roll:8:2", "roll:8:3", "roll:8:320", "roll:8:224", "roll:16:1", "roll:32:1", "roll:16:160",
            "lean:8:2", "lean:4:2", "lean:4:3", "lean:4:4", "lean:8:3", "lean:16:1",
            "ring:384:32", "ring:256:32", "ring:320:32", "ring:448:32", "ring:512:16", "ring:768:16", "ring:192:64", "ring:224:64"]


@pytest.fixture(params=VARIANTS)
def variant(request, monkeypatch):
    monkeypatch.setenv("NVME_STROM_SCAN_VARIANT", request.param)
    return request.param


def crafted_pages(oracle):
    """edge pages: zero, new-but-dirty, all-ones, every header rule, full line-pointer array"""
    pages = []
    pages.append(bytes(8192))
    p = bytearray(8192); p[8191] = 1; pages.append(bytes(p))
    p = bytearray(8192); p[8] = 1; pages.append(bytes(p))          # only the checksum field is set
    pages.append(b"\xff" * 8192)
    rng = np.random.default_rng(7)
    pages.append(rng.integers(0, 256, 8192, dtype=np.uint8).tobytes())

    def heap(lps, lower=None, upper=8000, special=8192, flags=0, version=0x2004, stamp=True, blk=0):
        q = bytearray(rng.integers(0, 256, 8192, dtype=np.uint8).tobytes())
        lower = 24 + 4 * len(lps) if lower is None else lower
        struct.pack_into("<HHHHHH", q, 8, 0, flags, lower, upper, special, version)
        for i, lp in enumerate(lps):
            struct.pack_into("<I", q, 24 + 4 * i, lp & 0xFFFFFFFF)
        return q

    # 2042 line pointers: the array fills the page up to pd_upper = 8192
    full = heap([(8184 | (1 << 15) | (8 << 17))] * 2042, upper=8192)
    pages.append(bytes(full))
    # random line-pointer words of every flag class
    for _ in range(40):
        n = int(rng.integers(0, 400))
        lps = rng.integers(0, 2 ** 32, n, dtype=np.uint64).tolist()
        pages.append(bytes(heap(lps, upper=int(rng.integers(0, 1025)) * 8,
                                special=int(rng.choice([8192, 8192, 8184, 8190, 4096])),
                                flags=int(rng.choice([0, 1, 7, 8, 0x8000])),
                                version=int(rng.choice([0x2004, 0x2004, 0x2003, 0])))))
    for lower in (0, 20, 23, 24, 25, 26, 28, 8192, 8196, 65535):
        pages.append(bytes(heap([1 << 15] * 3, lower=lower)))
    return pages


def stamp_some(oracle, buf, first_blkno):
    """gives every other page a correct pd_checksum so both outcomes occur"""
    n = buf.size // 8192
    for i in range(0, n, 2):
        c = oracle.lib.ora_pg_checksum_page(buf[i * 8192:].ctypes.data, (first_blkno + i) & 0xFFFFFFFF)
        buf[i * 8192 + 8] = c & 0xFF
        buf[i * 8192 + 9] = c >> 8
    return buf


def test_crafted_pages(oracle, variant):
    from gpu_util import scan_dev, to_dev
    buf = np.frombuffer(b"".join(crafted_pages(oracle)), dtype=np.uint8).copy()
    buf = stamp_some(oracle, buf, 1000)
    n = buf.size // 8192
    want_res, want_tot = oracle.scan(buf, 1000)
    got_res, got_tot = scan_dev(to_dev(buf), n, 1000)
    assert (got_res == want_res).all(), np.argwhere(got_res != want_res)[:5]
    assert got_tot == want_tot


@pytest.mark.parametrize("n_pages", [1, 2, 3, 4, 5, 7, 8, 31, 33, 147, 148, 149, 1183, 1184, 1185, 4096 + 13])
def test_ragged_counts(oracle, variant, n_pages):
    from gpu_util import scan_dev, to_dev
    buf = oracle.gen_pages(0x5EED0001, 77, n_pages, mixed=True)
    want_res, want_tot = oracle.scan(buf, 77)
    got_res, got_tot = scan_dev(to_dev(buf), n_pages, 77)
    assert (got_res == want_res).all()
    assert got_tot == want_tot


def test_empty_input_is_a_no_op(oracle):
    from gpu_util import scan_dev, to_dev
    buf = oracle.gen_pages(1, 0, 1)
    _, tot = scan_dev(to_dev(buf), 0, 0)
    assert tot[:11] == [0] * 11 and tot[11] == 2 ** 64 - 1


def test_blkno_wraps_at_32_bits(oracle, variant):
    from gpu_util import scan_dev, to_dev
    first = 0xFFFFFFF0
    buf = oracle.gen_pages(0x5EED0010, first, 40, mixed=True)
    want_res, want_tot = oracle.scan(buf, first)
    got_res, got_tot = scan_dev(to_dev(buf), 40, first)
    assert (got_res == want_res).all() and got_tot == want_tot
    assert want_tot[2] == 0 or want_tot[2] < 5          # the generator stamped the wrapped numbers


def test_totals_accumulate_across_launches(oracle):
    import ctypes as C
    from gpu_util import new_totals, to_dev, totals_list
    from nvme_strom_b200.abi import check, lib
    buf = oracle.gen_pages(0x5EED0001, 0, 3000, mixed=True)
    _, want = oracle.scan(buf, 0)
    dev = to_dev(buf)
    tot = new_totals()
    for first, n in [(0, 1000), (1000, 1), (1001, 1999)]:
        check(lib.strom_scan_pages_dev(C.c_void_p(dev.data_ptr() + first * 8192), n, first, 0, None,
                                       C.c_void_p(tot.data_ptr()), None), "scan")
    assert totals_list(tot) == want


def test_without_results_buffer(oracle, variant):
    from gpu_util import scan_dev, to_dev
    buf = oracle.gen_pages(0x5EED0001, 5, 600, mixed=True)
    _, want = oracle.scan(buf, 5)
    _, got = scan_dev(to_dev(buf), 600, 5, want_results=False)
    assert got == want


def test_flag_selects_kernel(oracle):
    from gpu_util import scan_dev, to_dev
    from nvme_strom_b200 import abi
    buf = oracle.gen_pages(0x5EED0001, 0, 500, mixed=True)
    want_res, want_tot = oracle.scan(buf, 0)
    for fl in (abi.SCAN_KERNEL_LDG, abi.SCAN_KERNEL_TMA, abi.SCAN_DEFAULT):
        got_res, got_tot = scan_dev(to_dev(buf), 500, 0, flags=fl)
        assert (got_res == want_res).all() and got_tot == want_tot


def test_bad_arguments(oracle):
    import ctypes as C
    import errno
    from gpu_util import new_totals, to_dev
    from nvme_strom_b200.abi import lib
    dev = to_dev(oracle.gen_pages(1, 0, 2))
    tot = new_totals()
    assert lib.strom_scan_pages_dev(None, 1, 0, 0, None, C.c_void_p(tot.data_ptr()), None) == -1
    assert C.get_errno() == errno.EINVAL
    assert lib.strom_scan_pages_dev(C.c_void_p(dev.data_ptr() + 4), 1, 0, 0, None,
                                    C.c_void_p(tot.data_ptr()), None) == -1
    assert lib.strom_scan_pages_dev(C.c_void_p(dev.data_ptr()), 1, 0, 0, None, None, None) == -1


def test_full_size_properties(oracle):
    """BASELINE size (1 GiB, 131072 pages): size-independent witnesses instead of a
    page-by-page CPU pass -- a checksum of checksums over sampled ranges, zero
    failures on a clean file, idempotence, and order independence of the totals."""
    import torch
    from gpu_util import scan_dev
    from nvme_strom_b200 import abi
    from nvme_strom_b200.abi import lib
    n = 131072
    host = np.empty(n * 8192, dtype=np.uint8)
    for i in range(n):
        lib.strom_heapgen_page(0x5EED0001, i, 0, host[i * 8192:].ctypes.data)
    dev = torch.from_numpy(host).cuda()
    res, tot = scan_dev(dev, n, 0)
    assert tot[abi.T_PAGES] == n
    assert tot[abi.T_CHECKSUM_FAIL] == tot[abi.T_HEADER_FAIL] == tot[abi.T_INVALID] == 0
    assert tot[abi.T_LP_BAD_PAGES] == tot[abi.T_N_BAD_LP] == tot[abi.T_ZERO] == 0
    assert tot[abi.T_SUM_CHECKSUM] == int(res[:, 0].astype(np.uint64).sum())
    assert tot[abi.T_XOR_CHECKSUM] == int(np.bitwise_xor.reduce(res[:, 0].astype(np.uint64)))
    assert tot[abi.T_N_NORMAL] == int(res[:, 2].astype(np.uint64).sum())
    # stored checksum == computed checksum on every page
    stored = host.reshape(n, 8192)[:, 8].astype(np.uint16) | (host.reshape(n, 8192)[:, 9].astype(np.uint16) << 8)
    assert (stored == res[:, 0]).all()
    # oracle on three sampled windows
    for first in (0, 65536 - 50, n - 100):
        w_res, _ = oracle.scan(host[first * 8192:(first + 100) * 8192], first)
        assert (w_res == res[first:first + 100]).all()
    # idempotent, and both kernels agree at full size
    res2, tot2 = scan_dev(dev, n, 0, flags=abi.SCAN_KERNEL_LDG)
    assert tot2 == tot and (res2 == res).all()
