#!/bin/bash
# round 2, pass A: parity tests (incl. the 2-process strip test), smoke, the default bench line, kernel bench, issue probes
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
timeout 300 python tools/kbench.py > $O/kbench.log 2>&1
timeout 120 python tools/peaks.py $O/issue_peaks.json > /dev/null 2> $O/peaks.err
timeout 300 python tools/btf_o1_report.py $O/btf_o1.json > $O/btf_o1.log 2>&1
echo done > $O/round_done.txt
