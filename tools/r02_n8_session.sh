#!/bin/bash
# One 8-GPU box session of round 2: topology, host memory rates, staging sweep, bench at N=8, reference arm.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/r02_n8_session.sh 8'
N=${1:-8}
O=gpurun_out
mkdir -p $O
bash tools/probe_box.sh > $O/r02_box_n$N.txt 2>&1
python -c "
import sys; sys.path.insert(0,'.')
from nvme_strom_b200.loader import heapgen_file; heapgen_file('/tmp/hm.heap', 1, $N * 131072)"
gcc -O2 -mavx2 -pthread -o /tmp/hostmem_bench tools/hostmem_bench.c && /tmp/hostmem_bench /tmp/hm.heap $(nproc) $(( $(nproc) / 2 )) > $O/r02_hostmem_n$N.txt 2>&1
rm -f /tmp/hm.heap
R=$(( $(nproc) / N )); [ $R -gt 8 ] && R=8; [ $R -lt 2 ] && R=2
timeout 600 python tools/ring_sweep.py $N /tmp \
   0:0:4 0:0:4:memcpy:none $R:4096:4:memcpy:none $R:2048:4:memcpy:none $R:2048:4 $((R-1)):2048:4 \
   $R:1024:4 $R:1024:8 $R:512:8 $R:2048:4:ntstore $R:4096:4:ntstore $R:1024:4:ntstore \
   $R:2048:4:memcpy:split:file:NVME_STROM_SUB_KB=128 $R:2048:4:memcpy:split:file:NVME_STROM_SPIN_US=100 \
   0:0:4:memcpy:split:shm 0:0:4:memcpy:none:shm 0:0:4:memcpy:split:shm:NVME_STROM_ZC_PIECE_MB=8 \
   > $O/r02_ring_n$N.txt 2>&1
if [ $N -gt 1 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
     bench.py --gpus $N --steps 10 --warmup 3 > $O/r02_bench_n$N.json 2> $O/r02_bench_n$N.err
else
  timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 > $O/r02_bench_n$N.json 2> $O/r02_bench_n$N.err
fi
timeout 300 python bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $O/r02_bench_ref_n$N.json 2> $O/r02_bench_ref_n$N.err
tail -n 30 $O/r02_ring_n$N.txt; tail -c 1500 $O/r02_bench_n$N.json; tail -c 400 $O/r02_bench_n$N.err
