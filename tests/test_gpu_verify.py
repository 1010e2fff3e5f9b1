"""Parity of the byte-compare kernels with the oracle's restatement of
utils/nvme_test.c:254-266 (+ :133-194 for the first differing 16-byte unit)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_blocks", [1, 2, 31, 33, 500, 4096 + 5])
def test_equal_buffers(oracle, n_blocks):
    from gpu_util import to_dev, verify_dev
    rng = np.random.default_rng(n_blocks)
    a = rng.integers(0, 256, n_blocks * 8192, dtype=np.uint8)
    assert verify_dev(to_dev(a), to_dev(a.copy()), n_blocks) == (0, 2 ** 64 - 1)


def test_single_byte_positions(oracle):
    """every position class: first/last byte, first/last unit, first/last block"""
    from gpu_util import to_dev, verify_dev
    n = 300
    rng = np.random.default_rng(2)
    a = rng.integers(0, 256, n * 8192, dtype=np.uint8)
    da = to_dev(a)
    for pos in (0, 15, 16, 8191, 8192, 8192 + 4095, 299 * 8192, n * 8192 - 1, 150 * 8192 + 511, 150 * 8192 + 512):
        b = a.copy()
        b[pos] ^= 0x80
        assert verify_dev(da, to_dev(b), n) == oracle.verify(a, b, n), pos
        assert oracle.verify(a, b, n) == (1, pos & ~15)


def test_many_mismatches_and_base_offset(oracle):
    from gpu_util import to_dev, verify_dev
    n = 2000
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, n * 8192, dtype=np.uint8)
    b = a.copy()
    hits = rng.choice(n * 8192, 700, replace=False)
    b[hits] ^= 1
    want = oracle.verify(a, b, n)
    assert want[0] == len(set(int(h) // 8192 for h in hits))
    assert verify_dev(to_dev(a), to_dev(b), n) == want
    got = verify_dev(to_dev(a), to_dev(b), n, base_offset=1 << 40)
    assert got == (want[0], want[1] + (1 << 40))


def test_block_map(oracle):
    """slot k of the landed buffer holds file block ids[k] (kmod/nvme_strom.c:1569-1571)"""
    from gpu_util import to_dev, verify_dev
    n = 257
    rng = np.random.default_rng(4)
    a = rng.integers(0, 256, n * 8192, dtype=np.uint8)
    perm = rng.permutation(n).astype(np.uint32)
    b = np.empty_like(a)
    for i in range(n):
        b[int(perm[i]) * 8192:(int(perm[i]) + 1) * 8192] = a[i * 8192:(i + 1) * 8192]
    assert verify_dev(to_dev(a), to_dev(b), n, to_dev(perm)) == (0, 2 ** 64 - 1)
    b[int(perm[100]) * 8192 + 33] ^= 2
    want = oracle.verify(a, b, n, perm)
    assert want == (1, 100 * 8192 + 32)
    assert verify_dev(to_dev(a), to_dev(b), n, to_dev(perm)) == want


@pytest.mark.parametrize("fused_r", ["", "384", "192", "448", "4", "2", "41"])
def test_fused_scan_and_verify(oracle, monkeypatch, fused_r):
    from gpu_util import scan_verify_dev, to_dev
    if fused_r:
        monkeypatch.setenv("NVME_STROM_FUSED_R", fused_r)       # "" = the default (shared-memory rings)
    else:
        monkeypatch.delenv("NVME_STROM_FUSED_R", raising=False)
    for n in (1, 3, 4, 37, 1500):
        a = oracle.gen_pages(0x5EED0001, 9, n, mixed=True)
        b = a.copy()
        rng = np.random.default_rng(n)
        if n > 1:
            b[rng.choice(n * 8192, min(n, 40), replace=False)] ^= 0x10
        b[n * 8192 - 1] ^= 1
        want_res, want_tot = oracle.scan(a, 9)
        nm, fo = oracle.verify(a, b, n)
        want_tot[10], want_tot[11] = nm, fo + 4096
        got_res, got_tot = scan_verify_dev(to_dev(a), to_dev(b), n, 9, base_offset=4096)
        assert (got_res == want_res).all()
        assert got_tot == want_tot
    a = oracle.gen_pages(0x5EED0001, 0, 100)
    _, tot = scan_verify_dev(to_dev(a), to_dev(a.copy()), 100, 0)
    assert tot[10] == 0 and tot[11] == 2 ** 64 - 1


def test_full_size_roundtrip_property():
    """1 GiB against itself: 0 mismatches; flip one bit per 1024 blocks: exactly those blocks"""
    import torch
    from gpu_util import verify_dev
    n = 131072
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randint(0, 256, (n * 8192,), dtype=torch.uint8, device="cuda", generator=g)
    b = a.clone()
    assert verify_dev(a, b, n) == (0, 2 ** 64 - 1)
    idx = torch.arange(7, n, 1024, device="cuda") * 8192 + 4242
    b[idx] ^= 1
    assert verify_dev(a, b, n) == (idx.numel(), 7 * 8192 + (4242 & ~15))
