"""Writes tests/golden/c5_expected_totals.json: the CPU oracle's counters over the
mixed-validity synthetic heap file of config C5 (seed 0x5EED0001, corruption classes
of SURVEY.md 8d), for 1, 2, 4 and 8 GiB (131072 pages per GiB).  GPU totals at any
GPU count must equal these integers.  Run from the repo root: python tests/golden/make_c5_expected.py"""
import ctypes as C
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
from conftest import Oracle

o = Oracle()
SEED, PAGES_PER_GIB, BATCH = 0x5EED0001, 131072, 4096
import numpy as np

tot = o.Totals()
tot.v[11] = 0xFFFFFFFFFFFFFFFF
out = {}
buf = np.empty(BATCH * 8192, dtype=np.uint8)
page = 0
for gib in (1, 2, 4, 8):
    while page < gib * PAGES_PER_GIB:
        for i in range(BATCH):
            o.lib.ora_gen_page(SEED, page + i, 1, buf[i * 8192:].ctypes.data)
        o.lib.ora_scan_pages(buf.ctypes.data, BATCH, page, None, C.byref(tot))
        page += BATCH
    out[str(gib)] = [int(x) for x in tot.v]
    print(gib, out[str(gib)][:10])
json.dump({"seed": SEED, "mixed": True, "pages_per_gib": PAGES_PER_GIB,
           "fields": ["pages", "zero", "checksum_fail", "header_fail", "lp_bad_pages", "invalid", "n_normal",
                      "n_bad_lp", "xor_checksum", "sum_checksum", "mismatch_blocks", "first_mismatch", "new"],
           "totals_by_gib": out}, open(os.path.join(HERE, "c5_expected_totals.json"), "w"), indent=1)
