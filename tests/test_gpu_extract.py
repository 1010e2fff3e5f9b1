"""strom_extract_items_dev against the oracle's restatement of the tuple walk of
pgsql/nvme_strom.c:1069-1081 (lines = PageGetMaxOffsetNumber; skip unless ItemIdIsNormal;
offset/length from the line pointer; t_self = (blkno, lineoff))."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def oracle_items(oracle, buf, first_blkno):
    n = buf.size // 8192
    out, offs = [], [0]
    items = np.zeros((2100, 4), dtype=np.uint16)
    for i in range(n):
        k = oracle.lib.ora_page_items(buf[i * 8192:].ctypes.data, items.ctypes.data, 2100)
        for j in range(k):
            out.append(((first_blkno + i) & 0xFFFFFFFF, int(items[j, 0]), int(items[j, 1]), int(items[j, 2])))
        offs.append(len(out))
    return out, offs


def gpu_items(buf, first_blkno, max_items=None):
    import torch
    from gpu_util import to_dev
    from nvme_strom_b200 import abi
    n = buf.size // 8192
    dev = to_dev(buf) if n else torch.zeros(16, dtype=torch.uint8, device="cuda")
    cap = n * 2042 if max_items is None else max_items
    items = torch.zeros(max(cap, 1) * 3, dtype=torch.int32, device="cuda")
    offs = torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")
    total = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    abi.check(abi.lib.strom_extract_items_dev(C.c_void_p(dev.data_ptr()), n, first_blkno & 0xFFFFFFFF,
                                              C.c_void_p(items.data_ptr()), cap, C.c_void_p(offs.data_ptr()),
                                              C.c_void_p(total.data_ptr()), None), "extract")
    torch.cuda.synchronize()
    raw = items.cpu().numpy().view(np.uint8).reshape(-1, 12)
    blk = raw[:, 0:4].copy().view(np.uint32).ravel()
    u16 = raw[:, 4:12].copy().view(np.uint16).reshape(-1, 4)
    tot = int(total.item())
    got = [(int(blk[i]), int(u16[i, 0]), int(u16[i, 1]), int(u16[i, 2])) for i in range(min(tot, cap))]
    return got, offs.cpu().tolist(), tot


@pytest.mark.parametrize("n_pages,first,mixed", [(1, 0, False), (7, 5, True), (300, 0xFFFFFF00, True),
                                                 (511, 3, True), (512, 3, True), (513, 3, True), (1025, 9, True),
                                                 (9000, 100, True), (8192 * 3 + 17, 0, False)])
def test_items_match_the_oracle(oracle, n_pages, first, mixed):
    buf = oracle.gen_pages(0x5EED0001, first, n_pages, mixed)
    want, want_offs = oracle_items(oracle, buf, first)
    got, offs, total = gpu_items(buf, first)
    assert total == len(want) and offs == want_offs
    assert got == want


def test_crafted_line_pointer_arrays(oracle):
    import struct
    rng = np.random.default_rng(11)
    pages = []
    for lower in (0, 20, 24, 25, 28, 24 + 4 * 2042, 8192, 9000, 65535):
        p = bytearray(rng.integers(0, 256, 8192, dtype=np.uint8).tobytes())
        struct.pack_into("<H", p, 12, lower)
        pages.append(bytes(p))
    p = bytearray(8192)                                       # every line pointer LP_NORMAL
    struct.pack_into("<H", p, 12, 8192)
    for i in range(2042):
        struct.pack_into("<I", p, 24 + 4 * i, 8 | (1 << 15) | (9 << 17))
    pages.append(bytes(p))
    pages.append(bytes(8192))
    buf = np.frombuffer(b"".join(pages), dtype=np.uint8).copy()
    want, want_offs = oracle_items(oracle, buf, 42)
    got, offs, total = gpu_items(buf, 42)
    assert (total, offs, got) == (len(want), want_offs, want)
    assert offs[-2] - offs[-3] == 2042


def test_capacity_limit_and_empty_input(oracle):
    buf = oracle.gen_pages(0x5EED0001, 0, 50)
    want, want_offs = oracle_items(oracle, buf, 0)
    got, offs, total = gpu_items(buf, 0, max_items=100)
    assert total == len(want) and offs == want_offs and got == want[:100]
    got, offs, total = gpu_items(np.zeros(0, dtype=np.uint8), 0)
    assert (got, offs, total) == ([], [0], 0)


def test_counts_agree_with_the_scan_kernel(oracle):
    """on valid pages n_normal of the scan == items per page of the extraction"""
    from gpu_util import scan_dev, to_dev
    buf = oracle.gen_pages(0x5EED0010, 0, 4000)
    res, tot = scan_dev(to_dev(buf), 4000, 0)
    _, offs, total = gpu_items(buf, 0)
    assert total == tot[6]
    assert (np.diff(np.array(offs)) == res[:, 2]).all()
