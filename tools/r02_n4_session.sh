#!/bin/bash
# N-GPU box session (N = 2 or 4): the multi-device GPU tests that skip on a 1-GPU box, then bench.py at N.
N=${1:-4}
O=gpurun_out
mkdir -p $O
if [ "$2" = "full" ]; then SEL=""; else SEL=1; fi
timeout 400 python -m pytest tests -m gpu -q -rs ${SEL:+-k "multi or sharded or two or several"} > $O/r02_gputests_n$N.txt 2>&1
tail -6 $O/r02_gputests_n$N.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"
timeout 200 $RUN bench.py --gpus $N --steps 10 --warmup 3 > $O/r02_bench_n$N.json 2> $O/r02_bench_n$N.err
tail -c 1500 $O/r02_bench_n$N.json; tail -c 300 $O/r02_bench_n$N.err
timeout 100 python bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $O/r02_bench_ref_n$N.json 2> $O/r02_bench_ref_n$N.err
cat $O/r02_bench_ref_n$N.json | cut -c1-300
rm -f /tmp/strom_* /dev/shm/strom_*
