"""Writes tests/golden/checksum_selfpins.json.

The reference tree holds no checksum vectors, so these are the
self-pins SURVEY.md 8c lists, restated here as data: page constructor, block
number, expected block hash / checksum.  They pin the oracle's restatement of
PostgreSQL's algorithm to the survey's independent computation, NOT to
upstream PostgreSQL ("parity unpinned" there).  Run: python make_selfpins.py
"""
import json
import os

PINS = [
    {"page": "zero", "blkno": 0, "block_hash": 0x54AF71FA, "checksum": 0xC6AA},
    {"page": "zero", "blkno": 1, "block_hash": 0x54AF71FA, "checksum": 0xC6AB},
    {"page": "ramp", "blkno": 0, "block_hash": 0xC4FD2208, "checksum": 0xE706},
    {"page": "ramp", "blkno": 12345, "block_hash": 0xC4FD2208, "checksum": 0xD72F},
    {"page": "empty_heap", "blkno": 0, "block_hash": None, "checksum": 0x6560},
    {"page": "empty_heap", "blkno": 131071, "block_hash": None, "checksum": 0xA0A6},
]
DOC = {
    "source": "SURVEY.md section 8c self-pins",
    "pages": {
        "zero": "8192 zero bytes",
        "ramp": "byte k = k & 0xFF",
        "empty_heap": "pd_lower=24, pd_upper=pd_special=8192, pd_pagesize_version=0x2004, rest 0",
    },
    "pins": PINS,
}
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "checksum_selfpins.json"), "w") as f:
    json.dump(DOC, f, indent=1)
