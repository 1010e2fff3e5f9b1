show() { python -c "
import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], d['e2e']['value'], d['load_only_GBps'], d['e2e_with_byte_compare_GBps'], d['transfer_roofline']['source_GBps'])" $1 "$2"; }
run() { name=$1; shift; env "$@" timeout 120 python bench.py --no-cpu --steps 10 > gpurun_out/v_$name.json 2> gpurun_out/v_$name.err; show gpurun_out/v_$name.json "$name: $*"; }
run default X=1
run s2048x8 NVME_STROM_SLOT_KB=2048 NVME_STROM_SLOTS=8
run r6 NVME_STROM_READERS=6
run h4 NVME_STROM_HOST_READERS=4
run r6h6 NVME_STROM_READERS=6 NVME_STROM_HOST_READERS=6
run sub128 NVME_STROM_SUB_KB=128
run slots6 NVME_STROM_SLOTS=6
