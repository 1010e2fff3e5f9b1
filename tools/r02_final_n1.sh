#!/bin/bash
# Final 1-GPU session of round 2: GPU tests, the bench line, ncu launch list of the bench command and one
# full capture of the page-scan kernel (each ncu run only after the same command exited 0 without it).
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r02_gputests_final.txt 2>&1; tail -4 $O/r02_gputests_final.txt
timeout 400 python bench.py --steps 20 --warmup 3 > $O/r02_bench_final.json 2> $O/r02_bench_final.err; tail -c 300 $O/r02_bench_final.err
Q="python bench.py --steps 2 --warmup 1 --quick --no-cpu --no-c5"
timeout 300 $Q > $O/r02_quick_plain.log 2>&1 &&
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02_ncu_launch_list.csv $Q > $O/r02_ncu_list.log 2>&1
S="python tools/one_scan.py default 1024 4"
timeout 200 $S > $O/r02_one_scan_plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:scan_ring -s 1 -c 2 -o $O/r02_scan_ring $S > $O/r02_ncu_full.log 2>&1
ls -la $O | grep r02_ | tail -12; tail -3 $O/r02_ncu_full.log
