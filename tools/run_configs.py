"""Runs BASELINE.json's parity/measurement configs C3, C4, C5 on the GPUs at hand and
prints one JSON object per config (rank 0).  Launch plain for 1 GPU or under torchrun for N.
  C3  N x 8 GiB heap file sharded over N GPUs, segment size sweep 4 / 32 / 128 MB
  C4  MEMCPY_SSD2RAM (+H2D) vs MEMCPY_SSD2GPU, 16 GiB, 1 GPU
  C5  mixed-validity file, N x 1 GiB: merged counters must equal tests/golden/c5_expected_totals.json
Sizes can be scaled with --c3-gib / --c4-gib (what was run is printed, never silently shrunk).
Every row carries the file system, the box's H2D and page-cache read rates calibrated in the same
run (nvme_strom_b200/calibrate.py) and the row's fraction of each."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from nvme_strom_b200 import abi, calibrate, multigpu
from nvme_strom_b200.loader import SegmentPipeline, heapgen_file

GIB = 1 << 30


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--c3-gib", type=int, default=8, help="GiB per GPU for C3")
    ap.add_argument("--c4-gib", type=int, default=16)
    ap.add_argument("--dir", default=os.environ.get("NVME_STROM_BENCH_DIR", "/tmp"))
    ap.add_argument("--skip", default="")
    ap.add_argument("--pin", action="store_true", help="page-lock each rank's share of the file first (tmpfs only)")
    args = ap.parse_args()
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # each rank keeps to its share of the cores; the engine sizes itself by what it is allowed on
        os.sched_setaffinity(0, multigpu.rank_cpus(local, int(os.environ.get("LOCAL_WORLD_SIZE", world))))
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")    # keep stdout to the JSON lines
        dist.init_process_group("nccl", device_id=dev)
    fs = calibrate.filesystem_of(args.dir)[1]
    cal = {}

    def calibrated(path, offset, length):
        """H2D and page-cache read rates of this box, once per run (the first file serves)"""
        if not cal:
            scratch = torch.empty(1 << 29, dtype=torch.uint8, device=dev)
            cal["h2d_GBps"], cal["h2d_per_rank_GBps"] = calibrate.h2d_aggregate(dev, scratch, 1 << 29, 6)
            del scratch
            threads = abi.engine_state(local)["readers_max"]
            cal["page_cache_read_GBps"] = calibrate.pagecache_aggregate(dev, path, offset, min(length, 4 * GIB), threads, 2)
        return cal

    def fractions(gbps, c, zero_copy=False):
        roof = c["h2d_GBps"] if zero_copy else min(c["h2d_GBps"], c["page_cache_read_GBps"])
        return {"file_system": fs, "h2d_GBps": round(c["h2d_GBps"], 1),
                "page_cache_read_GBps": round(c["page_cache_read_GBps"], 1),
                "frac_of_h2d": round(gbps / c["h2d_GBps"], 3), "frac_of_transfer_roofline": round(gbps / roof, 3)}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def tmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def make_file(name, gib_per_rank, mixed):
        path = os.path.join(args.dir, f"strom_cfg_{name}_{os.environ.get('MASTER_PORT', os.getpid())}.heap")
        pages = gib_per_rank * GIB // 8192
        heapgen_file(path, 0x5EED0001, pages, mixed=mixed, first_page=rank * pages)
        barrier()
        if args.pin:
            fd = os.open(path, os.O_RDONLY)
            abi.check(abi.lib.nvme_strom_pin_file_range(fd, rank * pages * 8192, pages * 8192), "nvme_strom_pin_file_range")
            os.close(fd)
            barrier()
        return path

    def timed_run(pipe, path, first_seg, n_seg, reps=2):
        pipe.run(path, first_seg, n_seg)                       # warm-up (and page-cache warm)
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            tot, cnt = pipe.run(path, first_seg, n_seg)
        torch.cuda.synchronize()
        dt = tmax((time.perf_counter() - t0) / reps)
        return tot, cnt, dt

    out = []
    # ---------------- C5: exact counters on the mixed-validity file -------------------------
    if "c5" not in args.skip:
        path = make_file("c5", 1, True)
        seg = 32 << 20
        first, n = multigpu.partition_segments(world * GIB // seg, world, rank)
        pipe = SegmentPipeline(local, seg, 6)
        tot, cnt, dt = timed_run(pipe, path, first, n)
        pipe.close()
        merged = multigpu.reduce_totals(tot, dev)
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", "c5_expected_totals.json")))["totals_by_gib"]
        want = gold.get(str(world))
        c = calibrated(path, first * seg, n * seg)
        gbps = world * GIB / dt / 1e9
        res = {"config": "C5 mixed-validity", "n_gpus": world, "file_gib": world, "GBps": round(gbps, 2),
               "totals": merged[:13], "expected": want[:13] if want else None,
               "bit_exact": (merged == want) if want else None, **fractions(gbps, c, args.pin)}
        out.append(res)
        barrier()
        if rank == 0:
            os.unlink(path)
    # ---------------- C3: segment-size sweep ---------------------------------------------------
    if "c3" not in args.skip:
        path = make_file("c3", args.c3_gib, False)
        for seg_mb in (4, 32, 128):
            seg = seg_mb << 20
            first, n = multigpu.partition_segments(world * args.c3_gib * GIB // seg, world, rank)
            pipe = SegmentPipeline(local, seg, 6)
            tot, cnt, dt = timed_run(pipe, path, first, n, reps=1)
            pipe.close()
            merged = multigpu.reduce_totals(tot, dev)
            ok = merged[abi.T_PAGES] == world * args.c3_gib * GIB // 8192 and merged[abi.T_CHECKSUM_FAIL] == 0 \
                and merged[abi.T_INVALID] == 0
            c = calibrated(path, first * seg, n * seg)
            gbps = world * args.c3_gib * GIB / dt / 1e9
            out.append({"config": "C3 sharded load+scan", "n_gpus": world, "file_gib": world * args.c3_gib,
                        "segment_mb": seg_mb, "GBps": round(gbps, 2), **fractions(gbps, c, args.pin),
                        "engine": abi.engine_state(local),
                        "GBps_per_gpu": round(args.c3_gib * GIB / dt / 1e9, 2), "clean_file_no_failures": ok,
                        "page_cache": "bypassed (O_DIRECT)" if os.environ.get("NVME_STROM_IO") == "direct" else ("warm, page-locked (zero-copy)" if args.pin else "warm")})
        barrier()
        if rank == 0:
            os.unlink(path)
    # ---------------- C4: ssd2ram (+H2D) vs ssd2gpu, one GPU --------------------------------------
    if "c4" not in args.skip and world == 1:
        path = make_file("c4", args.c4_gib, False)
        seg = 32 << 20
        n = args.c4_gib * GIB // seg
        res = {"config": "C4 ssd2ram vs ssd2gpu", "n_gpus": 1, "file_gib": args.c4_gib, "segment_mb": 32}
        for name, kw in (("ssd2gpu", {}), ("ssd2ram_then_h2d", {"via_ram": True})):
            pipe = SegmentPipeline(local, seg, 6, **kw)
            tot, cnt, dt = timed_run(pipe, path, 0, n, reps=1)
            pipe.close()
            res[name + "_GBps"] = round(args.c4_gib * GIB / dt / 1e9, 2)
            res[name + "_fractions"] = fractions(args.c4_gib * GIB / dt / 1e9, calibrated(path, 0, n * seg))
            res[name + "_pages"] = tot[abi.T_PAGES]
            res[name + "_xor"] = tot[abi.T_XOR_CHECKSUM]
        res["same_result"] = res["ssd2gpu_xor"] == res["ssd2ram_then_h2d_xor"] and res["ssd2gpu_pages"] == res["ssd2ram_then_h2d_pages"]
        # cold-ish: O_DIRECT is per-process (NVME_STROM_IO=direct), run by tools/run_direct.sh
        out.append(res)
        os.unlink(path)
    if rank == 0:
        for r in out:
            print(json.dumps(r))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
