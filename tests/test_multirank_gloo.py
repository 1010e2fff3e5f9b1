"""N>1 host logic on CPU: two gloo ranks shard the segments of one file with
strom_partition_segments, each computes its shard's totals (the oracle stands in
for the GPU scan here -- this test is about the sharding and the reduction, the
GPU parity is in test_gpu_*.py), and the reduced totals equal the whole-file
totals.  Mirrors what bench.py does over NCCL at --gpus N."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, path, seg_pages, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from conftest import Oracle
    from nvme_strom_b200 import multigpu
    oracle = Oracle()
    n_seg = os.path.getsize(path) // (seg_pages * 8192)
    first, count = multigpu.partition_segments(n_seg, world, rank)
    data = np.fromfile(path, dtype=np.uint8, offset=first * seg_pages * 8192,
                       count=count * seg_pages * 8192)
    _, tot = oracle.scan(data, first * seg_pages)
    merged = multigpu.reduce_totals(tot, device="cpu")
    # every rank keeps to its share of the cores, and an engine made afterwards counts that share
    # (what bench.py does before its first MAP_GPU_MEMORY); the shares of all ranks are disjoint
    from nvme_strom_b200 import abi
    everyone = sorted(os.sched_getaffinity(0))
    mine = multigpu.rank_cpus(rank, world, allowed=everyone, local_lists=[None] * world)
    os.sched_setaffinity(0, mine)
    seen = torch.zeros(max(everyone) + 1, dtype=torch.int64)
    seen[mine] = 1
    dist.all_reduce(seen)
    assert int(seen.max()) == 1 or len(everyone) < world
    assert abi.engine_state(0)["allowed_cpus"] == len(mine)
    if rank == 0:
        np.save(out_path, np.array(merged, dtype=np.uint64))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_pages,seg_pages", [(2, 1600, 64), (2, 448, 64), (3, 640, 128)])
def test_sharded_totals_equal_whole_file(tmp_path, oracle, world, n_pages, seg_pages):
    path = str(tmp_path / "w.heap")
    assert oracle.lib.ora_gen_file(path.encode(), 0x5EED0040, n_pages, 1) == 0
    out = str(tmp_path / "merged.npy")
    mp.spawn(_rank_main, args=(world, _free_port(), path, seg_pages, out), nprocs=world, join=True)
    whole = np.fromfile(path, dtype=np.uint8)
    whole = whole[:(n_pages // seg_pages) * seg_pages * 8192]
    _, want = oracle.scan(whole, 0)
    got = np.load(out).tolist()
    assert got == want


def test_partition_is_a_disjoint_cover(strom):
    import ctypes as C
    for n_seg in (0, 1, 7, 32, 2048, 2049):
        for parts in (1, 2, 3, 4, 8):
            cover = []
            for r in range(parts):
                f, c = C.c_uint64(), C.c_uint64()
                assert strom.lib.strom_partition_segments(n_seg, parts, r, C.byref(f), C.byref(c)) == 0
                cover.append((f.value, c.value))
            assert cover[0][0] == 0 and sum(c for _, c in cover) == n_seg
            for (f0, c0), (f1, _) in zip(cover, cover[1:]):
                assert f0 + c0 == f1
            sizes = [c for _, c in cover]
            assert max(sizes) - min(sizes) <= 1
    f, c = C.c_uint64(), C.c_uint64()
    assert strom.lib.strom_partition_segments(10, 0, 0, C.byref(f), C.byref(c)) == -1
    assert strom.lib.strom_partition_segments(10, 2, 2, C.byref(f), C.byref(c)) == -1
