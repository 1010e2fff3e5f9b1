#!/bin/bash
# e2e of bench.py at N GPUs for several reader counts (tuning aid): tools/readers_sweep.sh N r1 r2 ...
N=$1; shift
port=29520
for r in "$@"; do
  port=$((port+1))
  NVME_STROM_READERS=$r NVME_STROM_HOST_READERS=$r timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --quick > gpurun_out/rs_$r.json 2> gpurun_out/rs_$r.err
  python -c "
import json,sys; d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1]); print('readers', sys.argv[2], 'e2e', d['e2e']['value'], 'tmpfs', d.get('e2e_tmpfs_zero_copy',{}).get('value'), 'source', d['transfer_roofline']['source_GBps'], 'h2d', d['transfer_roofline']['h2d_calibrated_GBps'])" gpurun_out/rs_$r.json $r
done
