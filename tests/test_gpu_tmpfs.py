"""Files on tmpfs: after nvme_strom_pin_file() the library copies straight out of the page
cache (page-locked mapping, no staging copy).  Same bytes, same counters, before and after."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def shm_file(oracle):
    if not os.path.isdir("/dev/shm"):
        pytest.skip("no /dev/shm")
    path = f"/dev/shm/strom_test_{os.getpid()}.heap"
    assert oracle.lib.ora_gen_file(path.encode(), 0x5EED0001, 20000, 1) == 0     # ~156 MiB: 3 regions, ragged tail
    yield path
    os.unlink(path)


def test_pin_file_is_refused_where_pages_cannot_be_pinned(tmp_path):
    import errno, ctypes as C
    from nvme_strom_b200 import abi
    from nvme_strom_b200.loader import heapgen_file
    p = str(tmp_path / "ext4.heap")
    heapgen_file(p, 1, 1024)
    fd = os.open(p, os.O_RDONLY)
    rc = abi.lib.nvme_strom_pin_file(fd)
    e = C.get_errno()
    os.close(fd)
    import subprocess
    fstype = subprocess.run(["stat", "-f", "-c", "%T", p], capture_output=True, text=True).stdout.strip()
    if fstype == "tmpfs":
        assert rc == 0
    else:
        assert rc == -1 and e == errno.ENOTSUP, (fstype, rc, e)


@pytest.mark.parametrize("pin", [False, True])
def test_loaded_bytes_and_counters(shm_file, oracle, pin):
    import torch
    from nvme_strom_b200 import abi
    if pin:
        pfd = os.open(shm_file, os.O_RDONLY)
        assert abi.lib.nvme_strom_pin_file(pfd) == 0
        assert abi.lib.nvme_strom_pin_file(pfd) == 0           # idempotent
        os.close(pfd)
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    from nvme_strom_b200.loader import SegmentPipeline
    data = np.fromfile(shm_file, dtype=np.uint8)
    rng = np.random.default_rng(5)
    ids = rng.integers(0, 20000, 300).tolist() + list(range(19990, 20000)) + list(range(8185, 8200))
    buf = torch.zeros(len(ids) * 8192, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf)
    fd = os.open(shm_file, os.O_RDONLY)
    a, keep = ssd2gpu(m.handle, 0, fd, ids)
    assert wait(a.dma_task_id) == 0
    got = buf.cpu().numpy()
    for k, cid in enumerate(list(keep)):
        assert (got[k * 8192:(k + 1) * 8192] == data[cid * 8192:(cid + 1) * 8192]).all(), k
    os.close(fd)
    unmap_gpu(m.handle)
    seg = 4 << 20
    pipe = SegmentPipeline(0, seg, 6, keep_results=True)
    tot, cnt = pipe.run(shm_file)
    n_seg = data.size // seg
    want_res, want_tot = oracle.scan(data[:n_seg * seg], 0)
    assert tot == want_tot and (pipe.results_host() == want_res).all()
    tot2, _ = pipe.run(shm_file)                       # second pass: everything already page-locked
    assert tot2 == want_tot
    pipe.close()
