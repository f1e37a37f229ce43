#!/bin/bash
# round 2, pass H (final tree): parity, smoke, bench lines, kernel bench, ncu launch lists + full captures
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
timeout 300 python tools/kbench.py > $O/kbench.log 2>&1
for c in narrow nat; do timeout 300 python bench.py --content $c --no-per-workload --no-cpu-baseline --steps 10 > $O/bench_content_$c.json 2> $O/bench_content_$c.err; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_default.csv \
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-probe > $O/ncu_launch_default.log 2>&1
EXTRA="sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active,smsp__inst_issued.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active"
prof() {
  timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:"$3" -s 2 -c 1 -f -o $O/full_$1 \
    python tools/kbench.py $2 > $O/ncu_full_$1.log 2>&1
}
prof bf bf:7680x4320:31 stencil_kernel
prof abf abf:3840x2160:15 stencil_kernel
prof jbf jbf:3840x2160:17:8:1.7320508 stencil_kernel
timeout 600 ncu --set full --clock-control none -k regex:'blur_rtv|guide_kernel' -s 6 -c 2 -f -o $O/full_btf_stages python tools/kbench.py btf:3840x2160:9 > $O/ncu_full_btf.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:'gradient' -s 4 -c 1 -f -o $O/full_gradient python tools/kbench.py grad:7680x4320:3 > $O/ncu_full_grad.log 2>&1
echo done > $O/round_done.txt
