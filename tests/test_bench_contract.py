"""bench.py's output contract, as far as it can be checked without a GPU: the reference arm
(the CPU path of configs C1/C2, timed on this machine's cores) prints one JSON line with the
agreed keys; under a multi-rank launch only rank 0 prints; the product arm refuses to run
without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

BENCH = os.path.join(ROOT, "bench.py")


def run_bench(*args, env=None):
    e = dict(os.environ, **(env or {}))
    return subprocess.run([sys.executable, BENCH, *args], capture_output=True, text=True, timeout=600, env=e)


def test_reference_arm_line():
    r = run_bench("--impl", "reference", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "GB/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["scaling"] == "weak" and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0
    assert set(cb["phases_per_pass_s"]) == {"read_s", "scan_s", "compare_s", "wall_s"}


def test_reference_arm_maps_no_product_library():
    """the reference arm is the reference's CPU path and nothing else: the process that times it never
    loads libnvme_strom (its file comes from the oracle's own generator)"""
    code = (f"import sys, runpy; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0'];\n"
            f"try:\n    runpy.run_path({BENCH!r}, run_name='__main__')\nexcept SystemExit: pass\n"
            "maps = open('/proc/self/maps').read()\n"
            "assert 'libstrom_oracle.so' in maps\n"
            "assert 'libnvme_strom' not in maps, [l for l in maps.splitlines() if 'libnvme_strom' in l]\n"
            "print('CLEAN')")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    # bench.py points descriptor 1 at stderr for everything but its one JSON line
    assert r.returncode == 0 and "CLEAN" in r.stdout + r.stderr, r.stdout[-1500:] + r.stderr[-3000:]
    assert len([l for l in r.stdout.splitlines() if l.strip()]) == 1


def test_reference_arm_other_ranks_stay_silent():
    r = run_bench("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                  env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""


def test_product_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = run_bench("--steps", "1", "--warmup", "3")
    assert r.returncode != 0
    assert "CUDA" in (r.stderr + r.stdout)
