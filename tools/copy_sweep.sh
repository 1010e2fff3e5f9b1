#!/bin/bash
# Copy-mode experiment left for the next GPU session: the same staging
# geometry with memcpy / non-temporal prefetch / streaming stores, at N GPUs.
#   gpurun --gpus 4 --timeout 400 -- 'bash tools/copy_sweep.sh 4'
N=${1:-1}
R=${2:-8}
for mode in memcpy nta ntstore; do
  echo "== NVME_STROM_COPY=$mode"
  NVME_STROM_COPY=$mode timeout 120 python tools/slot_sweep.py "$N" /tmp "$R:4096:4" "$R:2048:4" | grep readers
done
