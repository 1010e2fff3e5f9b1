"""strom_test, the library's own nvme_test-alike (same options as utils/nvme_test.c:689-751),
with the integrity check and the page scan done by the GPU kernels."""
import json
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

from conftest import GOLDEN, ROOT

TOOL = os.path.join(ROOT, "nvme_strom_b200", "bin", "strom_test")


@pytest.fixture(scope="module")
def c5_file(tmp_path_factory):
    from nvme_strom_b200.loader import heapgen_file
    path = str(tmp_path_factory.mktemp("st") / "c5.heap")
    heapgen_file(path, 0x5EED0001, 131072, mixed=True)
    return path


def parse_scan(out):
    m = re.search(r"^scan: (.*)$", out, re.M)
    assert m, out
    return {k: int(v) for k, v in (kv.split("=") for kv in m.group(1).split())}


def golden():
    t = json.load(open(os.path.join(GOLDEN, "c5_expected_totals.json")))["totals_by_gib"]["1"]
    return dict(pages=t[0], zero=t[1], checksum_fail=t[2], header_fail=t[3], lp_bad_pages=t[4], invalid=t[5],
                n_normal=t[6], n_bad_lp=t[7], xor=t[8], sum=t[9])


@pytest.mark.parametrize("opts", [["-c"], ["-k"], ["-c", "-n", "3", "-s", "16"], ["-k", "-s", "128", "-n", "2"],
                                  ["-f", "-c"], ["-f1024", "-k"]])
def test_counters_equal_the_oracle(c5_file, opts):
    r = subprocess.run([TOOL, "-d", "0"] + opts + [c5_file], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "throughput:" in r.stdout
    assert parse_scan(r.stdout) == golden()
    if "-c" in opts:
        assert "verify: mismatch_blocks=0" in r.stdout
    if "-f" not in opts[0]:
        assert "nr_ram2gpu: 0, nr_ssd2gpu: 131072, average DMA blocks: 256.00" in r.stdout


@pytest.mark.parametrize("members,stripe_kb", [(4, 128), (2, 1024), (8, 64)])
def test_stripe_set_gives_the_flat_files_counters(c5_file, tmp_path, members, stripe_kb):
    """-S: the same relation as the members of a stripe set (RAID-0 layout,
    kmod/nvme_strom.c:736-820): identical counters, byte-compare clean, and with stripes of at
    least 128 KB the same request accounting as the flat file"""
    from nvme_strom_b200.loader import stripe_split
    paths = [str(tmp_path / f"m{i}") for i in range(members)]
    stripe_split(c5_file, paths, stripe_kb << 10)
    for opts in (["-k"], ["-c"]):
        r = subprocess.run([TOOL, "-d", "0", "-S", str(stripe_kb)] + opts + paths, capture_output=True, text=True,
                           timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        assert parse_scan(r.stdout) == golden()
        if "-c" in opts:
            assert "verify: mismatch_blocks=0" in r.stdout
        avg = 256.0 if stripe_kb >= 128 else stripe_kb * 2.0
        assert f"nr_ram2gpu: 0, nr_ssd2gpu: 131072, average DMA blocks: {avg:.2f}" in r.stdout


def test_mapping_dump_and_usage(c5_file):
    r = subprocess.run([TOOL, "-d", "0", "-p", c5_file], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "Mapped GPU Memory (handle: 0x" in r.stdout and "GPU Page: version=1, size=65536" in r.stdout
    r = subprocess.run([TOOL, "-h"], capture_output=True, text=True)
    assert r.returncode == 1 and "usage: strom_test" in r.stderr
    r = subprocess.run([TOOL, "-d", "0", "/nonexistent"], capture_output=True, text=True)
    assert r.returncode == 1


def test_read_error_is_reported(c5_file):
    r = subprocess.run([TOOL, "-d", "0", c5_file], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, NVME_STROM_FAULT_EVERY="50"))
    assert r.returncode == 1, r.stdout + r.stderr
    assert "MEMCPY_WAIT" in r.stderr and "Input/output error" in r.stderr, r.stdout + r.stderr


def test_sharded_over_several_gpus(c5_file):
    """-d 0,1: contiguous segment ranges per GPU, counters merged on the host -- the same
    integers as one GPU computes"""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    devs = ",".join(str(i) for i in range(min(n, 4)))
    for opts in (["-k"], ["-c"]):
        r = subprocess.run([TOOL, "-d", devs] + opts + [c5_file], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        assert parse_scan(r.stdout) == golden()
        assert r.stdout.count("GPU") >= 2 and "nr_ssd2gpu: 131072" in r.stdout
