#!/bin/bash
# Third 8-GPU session of round 2 (short): sub-piece size and ring extremes, pread instead of the mapping as
# the source, then bench.py at N with the library's defaults.
N=${1:-8}
O=gpurun_out
mkdir -p $O
timeout 120 python tools/ring_sweep.py $N /tmp 0:0:4 4:2048:4:memcpy:split:file:NVME_STROM_SUB_KB=1024 \
   4:2048:4:memcpy:split:file:NVME_STROM_SUB_KB=64 4:8192:4 4:2048:2 > $O/r02_ring_n${N}c.txt 2>&1
cut -c1-250 $O/r02_ring_n${N}c.txt
NVME_STROM_SRC=pread timeout 120 python tools/ring_sweep.py $N /tmp 0:0:4 4:4096:4 3:2048:4 > $O/r02_ring_n${N}c_pread.txt 2>&1
cut -c1-250 $O/r02_ring_n${N}c_pread.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523"
timeout 150 $RUN bench.py --gpus $N --steps 10 --warmup 3 > $O/r02_bench_n${N}c.json 2> $O/r02_bench_n${N}c.err
tail -c 900 $O/r02_bench_n${N}c.json; tail -c 200 $O/r02_bench_n${N}c.err
rm -f /tmp/strom_* /tmp/ring_sweep* /dev/shm/strom_* /dev/shm/ring_sweep*
