"""Writes tests/golden/memdump_cases.json by running the REFERENCE's own
memdump_on_corruption (utils/nvme_test.c:133-194) through oracle/_ref/ref_memdump
(built by `make -C oracle ref`, needs /root/reference).  Each case: a length,
a base pattern (byte k = (k*7+3) & 0xFF) and a list of byte positions XORed
with 0x5A in the destination copy; `stdout` is what the reference printed.
"""
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "..", "..", "oracle", "_ref", "ref_memdump")
CASES = [
    {"length": 8192, "flips": [0]},
    {"length": 8192, "flips": [16]},
    {"length": 8192, "flips": [40]},
    {"length": 8192, "flips": [100, 101, 150]},
    {"length": 8192, "flips": [4096, 4112, 4200, 5000]},
    {"length": 8192, "flips": [8191]},
    {"length": 8192, "flips": [8100, 8190]},
    {"length": 256, "flips": [64, 80, 200]},
    {"length": 32 << 10, "flips": [20000, 20016, 20100]},
]


def base(n):
    return bytes((k * 7 + 3) & 0xFF for k in range(n))


out = []
for c in CASES:
    src = base(c["length"])
    dst = bytearray(src)
    for p in c["flips"]:
        dst[p] ^= 0x5A
    with tempfile.TemporaryDirectory() as d:
        a, b = os.path.join(d, "a"), os.path.join(d, "b")
        open(a, "wb").write(src)
        open(b, "wb").write(bytes(dst))
        r = subprocess.run([BIN, a, b, str(c["length"])], capture_output=True, text=True)
        assert r.returncode == 1, r
    out.append(dict(c, stdout=r.stdout))
json.dump({"source": "reference utils/nvme_test.c:133-194 run via oracle/_ref/ref_memdump",
           "pattern": "byte k = (k*7+3) & 0xFF; flips XOR 0x5A", "cases": out},
          open(os.path.join(HERE, "memdump_cases.json"), "w"), indent=1)
print("wrote", len(out), "cases")
