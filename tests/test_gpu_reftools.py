"""The reference's own binaries, unmodified, on top of libnvme_strom.

oracle/_ref/{nvme_test,ssd2ram_test,nvme_stat} are compiled from the reference
sources where they lie (oracle/Makefile, target ref) and travel to the GPU box
as prebuilt files.  With libnvme_strom_preload.so their open("/proc/nvme-strom")
+ ioctl() land in nvme_strom_ioctl().  `nvme_test -c` is the reference's own
integrity check of the loaded bytes (utils/nvme_test.c:225-274): exit status 0
means its memcmp of every 8KB block passed."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

from conftest import ROOT

REF = os.path.join(ROOT, "oracle", "_ref")
SHIM = os.path.join(ROOT, "nvme_strom_b200", "lib", "libnvme_strom_preload.so")


def _run(args, timeout=300, **env):
    e = dict(os.environ, LD_PRELOAD=SHIM, **env)
    return subprocess.run(args, env=e, capture_output=True, text=True, timeout=timeout)


@pytest.fixture(scope="module")
def heapfile(tmp_path_factory):
    from nvme_strom_b200.loader import heapgen_file
    path = str(tmp_path_factory.mktemp("ref") / "r.heap")
    heapgen_file(path, 0x5EED0001, 32768)                   # 256 MiB
    return path


def need(tool):
    p = os.path.join(REF, tool)
    if not os.path.exists(p):
        pytest.skip(f"{p} not built (make -C oracle ref needs /root/reference)")
    return p


def test_nvme_test_with_integrity_check(heapfile):
    r = _run([need("nvme_test"), "-d", "0", "-c", "-n", "4", "-s", "32", heapfile])
    assert r.returncode == 0, r.stdout + r.stderr
    assert "throughput" in r.stdout and "nr_ssd2gpu: 32768" in r.stdout


def test_nvme_test_default_geometry(heapfile):
    r = _run([need("nvme_test"), "-d", "0", heapfile])          # 6 slots x 32MB (:40-41)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "throughput" in r.stdout and "average DMA blocks: 256.00" in r.stdout


def test_nvme_test_mapping_dump(heapfile):
    # -p lists the windows of the calling process (:605-684); a fresh process has none.
    # Its retry loop spins on `errno != 0`, so a successful call must leave errno alone.
    r = _run([need("nvme_test"), "-d", "0", "-p", heapfile], timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr


def test_nvme_test_vfs_mode(heapfile):
    r = _run([need("nvme_test"), "-d", "0", "-c", "-f", heapfile])
    assert r.returncode == 0, r.stdout + r.stderr
