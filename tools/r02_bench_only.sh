#!/bin/bash
# bench.py at N GPUs with the library's defaults, nothing else (record for profiles/)
N=${1:-2}
O=gpurun_out
mkdir -p $O
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29527 \
  bench.py --gpus $N --steps 10 --warmup 3 > $O/r02_bench_n${N}_final.json 2> $O/r02_bench_n${N}_final.err
python -c "
import json; d=json.loads(open('$O/r02_bench_n${N}_final.json').read().strip().split(chr(10))[-1]); t=d['transfer_roofline']; print(d['value'], d['e2e']['value'], t['h2d_GBps'], t['frac'], t['host_copy_GBps'], d['engine'], d['c5_bit_exact'], d['e2e_tmpfs_zero_copy']['value'])"
tail -c 200 $O/r02_bench_n${N}_final.err; rm -f /tmp/strom_* /dev/shm/strom_*
