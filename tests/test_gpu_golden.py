"""The CUDA kernels against the committed golden vectors (no oracle call on this path, so a
drift of oracle and kernel together would still be caught): tests/golden/mixed_pages_2048.json
and tests/golden/checksum_selfpins.json."""
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _golden_pages(doc):
    from nvme_strom_b200.abi import lib
    buf = np.empty(doc["n_pages"] * 8192, dtype=np.uint8)
    for i in range(doc["n_pages"]):
        lib.strom_heapgen_page(doc["seed"], doc["first_blkno"] + i, 1, buf[i * 8192:].ctypes.data)
    assert hashlib.sha256(buf.tobytes()).hexdigest() == doc["pages_sha256"]
    return buf


@pytest.mark.parametrize("variant", ["", "ring:512:16", "ring:384:32", "lean:8:2", "ldg:8:2", "tma:16:1", "col:8", "roll:16:2"])
def test_scan_matches_frozen_records(variant, monkeypatch):
    from gpu_util import scan_dev, to_dev
    doc = json.load(open(os.path.join(GOLDEN, "mixed_pages_2048.json")))
    buf = _golden_pages(doc)
    if variant:
        monkeypatch.setenv("NVME_STROM_SCAN_VARIANT", variant)
    res, tot = scan_dev(to_dev(buf), doc["n_pages"], doc["first_blkno"])
    assert res.astype("<u2").tobytes().hex() == doc["results_hex"]
    assert tot == doc["totals"]


def test_fused_scan_compare_matches_frozen_records():
    from gpu_util import scan_verify_dev, to_dev
    doc = json.load(open(os.path.join(GOLDEN, "mixed_pages_2048.json")))
    buf = _golden_pages(doc)
    other = buf.copy()
    other[5 * 8192 + 4000] ^= 1
    other[1999 * 8192 + 16] ^= 0x80
    res, tot = scan_verify_dev(to_dev(buf), to_dev(other), doc["n_pages"], doc["first_blkno"], base_offset=0)
    assert res.astype("<u2").tobytes().hex() == doc["results_hex"]
    want = list(doc["totals"])
    want[10], want[11] = 2, 5 * 8192 + 4000 - 4000 % 16
    assert tot == want


def test_extracted_items_match_frozen_digest():
    from test_gpu_extract import gpu_items
    doc = json.load(open(os.path.join(GOLDEN, "mixed_pages_2048.json")))
    buf = _golden_pages(doc)
    got, offs, total = gpu_items(buf, doc["first_blkno"])
    assert total == doc["n_items"] and offs[-1] == total
    h = hashlib.sha256()
    for blk, lineoff, off, length in got:
        h.update(struct.pack("<IHHH", blk, lineoff, off, length))
    assert h.hexdigest() == doc["items_sha256"]


def test_checksum_selfpins_on_the_gpu():
    from gpu_util import scan_dev, to_dev
    doc = json.load(open(os.path.join(GOLDEN, "checksum_selfpins.json")))
    for pin in doc["pins"]:
        if pin["page"] == "zero":
            page = bytes(8192)
        elif pin["page"] == "ramp":
            page = bytes(k & 0xFF for k in range(8192))
        else:
            b = bytearray(8192)
            struct.pack_into("<HHHH", b, 12, 24, 8192, 8192, 0x2004)
            page = bytes(b)
        res, _ = scan_dev(to_dev(np.frombuffer(page, dtype=np.uint8).copy()), 1, pin["blkno"])
        assert int(res[0, 0]) == pin["checksum"], pin
