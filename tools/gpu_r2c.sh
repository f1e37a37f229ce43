#!/bin/bash
# round 2, pass C: sliding-window rows A/B (lib = runs of 8, lib_new = sliding), parity with the new library, then
# ncu --set full with source correlation on the three stencil kernels (one launch each) + pipe metrics
set -u
O=gpurun_out
mkdir -p $O
NEW=$PWD/various_image_processings_b200/lib_new/libvip_b200.so
SPECS="bf:7680x4320:31 abf:3840x2160:15 jbf:3840x2160:17:8:1.7320508 bf:3840x2160:11 bf:1920x1080:9 abf:1920x1080:9 btf:3840x2160:9 btf:1920x1080:9 bf:512x512:11"
timeout 300 python tools/kbench.py $SPECS > $O/kbench_runs8.log 2>&1
VIP_B200_LIB=$NEW timeout 300 python tools/kbench.py $SPECS > $O/kbench_sliding.log 2>&1
VIP_B200_LIB=$NEW VIP_STENCIL_PREMUL=0 timeout 300 python tools/kbench.py jbf:3840x2160:17:8:1.7320508 btf:3840x2160:9 > $O/kbench_sliding_nopremul.log 2>&1
VIP_B200_LIB=$NEW timeout 900 python -m pytest tests -m gpu -x -q -k "not bench_two_ranks" > $O/pytest_gpu_sliding.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_sliding.log
EXTRA="sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active,smsp__inst_issued.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active"
prof() {  # name spec [lib]
  timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:stencil_kernel -s 2 -c 1 -f -o $O/src_$1 \
    python tools/kbench.py $2 > $O/ncu_src_$1.log 2>&1
  ncu -i $O/src_$1.ncu-rep --page raw --csv > $O/src_$1_raw.csv 2>/dev/null
  ncu -i $O/src_$1.ncu-rep --page source --csv > $O/src_$1_source.csv 2>/dev/null
}
prof bf_runs8 bf:7680x4320:31
export VIP_B200_LIB=$NEW
prof bf bf:7680x4320:31
prof abf abf:3840x2160:15
prof jbf jbf:3840x2160:17:8:1.7320508
ls -la $O > $O/ls.txt
echo done > $O/round_done.txt
