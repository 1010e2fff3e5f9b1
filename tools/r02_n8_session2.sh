#!/bin/bash
# Second 8-GPU session of round 2 (short): reader count x ring size sweep with per-rank core shares,
# then C3 (64 GiB on ext4, segment sweep) and bench.py at N with the best staged setting.
#   gpurun --gpus 8 --timeout 420 -- 'bash tools/r02_n8_session2.sh 8'
N=${1:-8}
O=gpurun_out
mkdir -p $O
timeout 170 python tools/ring_sweep.py $N /tmp \
   0:0:4 3:2048:4 2:2048:4 3:1024:4 3:1024:8 3:512:8 4:1024:4 3:4096:4 3:2048:8 \
   3:2048:4:ntstore 3:2048:4:memcpy:split:file:NVME_STROM_WC=1 3:2048:4:memcpy:none \
   > $O/r02_ring_n${N}b.txt 2>&1
cat $O/r02_ring_n${N}b.txt | cut -c1-260
python tools/pick_setting.py $O/r02_ring_n${N}b.txt > $O/r02_pick_n$N.sh 2>/dev/null
cat $O/r02_pick_n$N.sh
source $O/r02_pick_n$N.sh
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
timeout 130 $RUN tools/run_configs.py --c3-gib 8 --skip c5,c4 > $O/r02_configs_c3_n$N.jsonl 2> $O/r02_configs_c3_n$N.err
cut -c1-400 $O/r02_configs_c3_n$N.jsonl; tail -c 300 $O/r02_configs_c3_n$N.err
timeout 150 $RUN bench.py --gpus $N --steps 10 --warmup 3 > $O/r02_bench_n${N}b.json 2> $O/r02_bench_n${N}b.err
tail -c 1200 $O/r02_bench_n${N}b.json; tail -c 300 $O/r02_bench_n${N}b.err
rm -f /tmp/strom_* /tmp/ring_sweep* /dev/shm/strom_* /dev/shm/ring_sweep*
