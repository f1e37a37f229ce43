#!/bin/bash
# round 2, pass J: final validation of the tree (parity suite, default bench line) + ncu of BF on S-nat content
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_final.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_final.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_default_final.json 2> $O/bench_default_final.err; echo "rc=$?" >> $O/bench_default_final.err
KB_CONTENT=lenna timeout 600 ncu --set full --clock-control none -k regex:stencil_kernel -s 2 -c 1 -f -o $O/full_bf_nat python tools/kbench.py bf:7680x4320:31 > $O/ncu_full_bf_nat.log 2>&1
echo done > $O/round_done.txt
