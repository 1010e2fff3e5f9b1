"""Several windows, several devices, one process: the registry is global like the module's
(kmod/pmemmap.c:77-78), one transfer engine per device; UNMAP waits for tasks that still land
in the window (callback_release_mapped_gpu_memory, kmod/pmemmap.c:149-208)."""
import os
import threading
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def heapfile(tmp_path_factory):
    from nvme_strom_b200.loader import heapgen_file
    path = str(tmp_path_factory.mktemp("multi") / "m.heap")
    heapgen_file(path, 0x5EED0001, 32768, mixed=True)         # 256 MiB
    return path


def test_unmap_waits_for_tasks_in_flight(heapfile):
    import torch
    from gpu_util import map_gpu, ssd2gpu, unmap_gpu, wait
    data = np.fromfile(heapfile, dtype=np.uint8)
    buf = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
    m = map_gpu(buf)
    fd = os.open(heapfile, os.O_RDONLY)
    tasks = [ssd2gpu(m.handle, k * (32 << 20), fd, list(range(k * 4096, (k + 1) * 4096))) for k in range(8)]
    unmap_gpu(m.handle)                                        # returns only when all 8 tasks have landed
    torch.cuda.synchronize()
    assert (buf.cpu().numpy() == data).all()
    for a, _ in tasks:
        assert wait(a.dma_task_id) == 0
    os.close(fd)


def test_many_windows_listed(heapfile):
    import torch
    from gpu_util import map_gpu, unmap_gpu
    from nvme_strom_b200 import abi
    bufs = [torch.zeros(1 << 20, dtype=torch.uint8, device="cuda") for _ in range(40)]
    maps = [map_gpu(b) for b in bufs]
    assert len({m.handle for m in maps}) == 40
    lst = abi.StromCmd__ListGpuMemory(64)
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__LIST_GPU_MEMORY, lst)
    listed = set(list(lst.handles)[:lst.nitems])
    assert {m.handle for m in maps} <= listed
    for m in maps:
        unmap_gpu(m.handle)


def test_two_devices_in_one_process(heapfile, oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from nvme_strom_b200.loader import SegmentPipeline
    data = np.fromfile(heapfile, dtype=np.uint8)
    seg = 32 << 20
    results = {}

    def run(dev, first, n):
        pipe = SegmentPipeline(dev, seg, 3)
        results[dev] = pipe.run(heapfile, first, n)[0]
        pipe.close()

    ts = [threading.Thread(target=run, args=(0, 0, 4)), threading.Thread(target=run, args=(1, 4, 4))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    from nvme_strom_b200 import multigpu
    merged = multigpu.merge_totals([results[0], results[1]])
    _, want = oracle.scan(data, 0)
    assert merged == want
    torch.cuda.set_device(0)
