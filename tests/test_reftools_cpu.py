"""The reference's ssd2ram_test and nvme_stat binaries, UNMODIFIED, running on
libnvme_strom through the LD_PRELOAD shim -- no GPU involved on this path
(MEMCPY_SSD2RAM lands in host memory).  The binaries are compiled from the
reference sources where they lie by `make -C oracle ref` and travel prebuilt."""
import os
import re
import subprocess

import pytest

from conftest import ROOT

REF = os.path.join(ROOT, "oracle", "_ref")
SHIM = os.path.join(ROOT, "nvme_strom_b200", "lib", "libnvme_strom_preload.so")


def need(tool):
    p = os.path.join(REF, tool)
    if not os.path.exists(p) and os.path.isdir("/root/reference/utils"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"])
    if not os.path.exists(p):
        pytest.skip(f"{p} not built and /root/reference is absent")
    return p


def run(args, timeout=120):
    return subprocess.run(args, env=dict(os.environ, LD_PRELOAD=SHIM), capture_output=True, text=True,
                          timeout=timeout)


@pytest.fixture(scope="module")
def heapfile(tmp_path_factory, strom):
    path = str(tmp_path_factory.mktemp("ref") / "r.heap")
    assert strom.lib.strom_heapgen_file(path.encode(), 0x5EED0001, 0, 16384, 0, 4) == 0   # 128 MiB
    return path


def test_ssd2ram_test_check_mode(heapfile):
    r = run([need("ssd2ram_test"), "-c", heapfile])
    assert r.returncode == 0, r.stdout + r.stderr
    assert "DMA64: supported" in r.stdout


@pytest.mark.parametrize("extra", [["-s", "256"], ["-n", "3", "-s", "256"]])
def test_ssd2ram_test_loads_the_file(heapfile, extra):
    # -s 256: with the default 32MB buffer the tool indexes dma_tasks[] out of bounds
    # (utils/ssd2ram_test.c:216-222, noted in SURVEY.md 3.5); 256 units keep it in range
    r = run([need("ssd2ram_test")] + extra + [heapfile])
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"nr_ram2ram: (\d+), nr_ssd2ram: (\d+), avg DMA blocks: ([\d.]+)", r.stdout)
    assert m, r.stdout
    assert int(m.group(1)) + int(m.group(2)) == 16384
    assert float(m.group(3)) == 256.0          # 128KB requests, what the module reports on a contiguous file


def test_nvme_stat_one_shot():
    r = run([need("nvme_stat")], timeout=30)
    assert r.returncode == 0, r.stdout + r.stderr
    assert re.search(r"^nr_ssd2gpu:\s+0$", r.stdout, re.M) and "max_dma_count:" in r.stdout


def test_proc_entry_banner_through_the_shim(tmp_path):
    """cat /proc/nvme-strom prints the version banner (kmod/nvme_strom.c:2036-2061)"""
    r = run(["cat", "/proc/nvme-strom"], timeout=30)
    assert r.returncode == 0 and r.stdout.startswith("version: ")


def test_shim_refuses_a_descriptor_it_cannot_track(tmp_path):
    """the shim remembers sessions per descriptor below 4096; past that it fails the open like a full
    descriptor table (EMFILE) instead of handing back a descriptor whose ioctls would end in ENOTTY"""
    import subprocess
    import sys
    from nvme_strom_b200._lib import PRELOAD_PATH
    code = r'''
import errno, os, resource
soft, hard = resource.getrlimit(resource.RLIMIT_NOFILE)
want = 6000
if hard != resource.RLIM_INFINITY and hard < want:
    print("SKIP"); raise SystemExit(0)
resource.setrlimit(resource.RLIMIT_NOFILE, (want, hard))
fd = os.open("/proc/nvme-strom", os.O_RDONLY)          # low descriptor: tracked
os.close(fd)
keep = [os.dup(0) for _ in range(4200)]
try:
    os.open("/proc/nvme-strom", os.O_RDONLY)
    print("OPENED")
except OSError as e:
    print("ERRNO", e.errno == errno.EMFILE)
'''
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, LD_PRELOAD=PRELOAD_PATH),
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "SKIP" in r.stdout or "ERRNO True" in r.stdout, r.stdout + r.stderr
