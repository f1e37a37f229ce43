#!/bin/bash
# round 3, 1-GPU evidence of the final tree: default bench line (probe built), ncu launch list of the same command,
# one --set full capture of the no-alloc integral image next to the pool-backed one.
set -u
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r03f_bench_n1.json 2> $O/r03f_bench_n1.err; echo rc=$? >> $O/r03f_bench_n1.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r03f_launches_bf_8k_r15.csv \
  python bench.py --steps 2 --warmup 3 --no-per-workload --no-cpu-baseline --no-probe > $O/r03f_ncu_bench.log 2>&1
timeout 300 python tools/kbench.py integral:3840x2160:7 integral:3840x2160:15 grad:3840x2160:3 grad:7680x4320:3 > $O/r03f_kbench_streaming.txt 2>&1
echo done > $O/r03f_done.txt
