#!/bin/bash
# round 2, pass I: integral image fast path (u8 BGR): parity, timing vs the generic kernels, ncu
set -u
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q -k "integral" > $O/pytest_integral.log 2>&1; echo "pytest rc=$?" >> $O/pytest_integral.log
SPECS="integral:3840x2160:7 integral:7680x4320:15 integral:1920x1080:4 integral:512x512:5 integral:3840x2160:1"
timeout 200 python tools/kbench.py $SPECS > $O/kbench_integral_fast.log 2>&1
VIP_INTEGRAL_FAST=0 timeout 200 python tools/kbench.py $SPECS > $O/kbench_integral_generic.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:'integral' -s 4 -c 4 -f -o $O/full_integral_fast python tools/kbench.py integral:3840x2160:7 > $O/ncu_integral_fast.log 2>&1
timeout 300 python -m pytest tests -m gpu -x -q -k "property" > $O/pytest_property.log 2>&1; echo "pytest rc=$?" >> $O/pytest_property.log
echo done > $O/round_done.txt
