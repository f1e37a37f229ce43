#!/bin/bash
# round 2, pass O: gradient, 4 vs 8 rows per thread
set -u
O=gpurun_out
mkdir -p $O
SPECS="grad:3840x2160:3 grad:7680x4320:3 grad:1920x1080:3"
timeout 120 python tools/kbench.py $SPECS > $O/kbench_grad4.log 2>&1
VIP_B200_LIB=$PWD/various_image_processings_b200/lib_g8/libvip_b200.so timeout 120 python tools/kbench.py $SPECS > $O/kbench_grad8.log 2>&1
timeout 120 python tools/kbench.py $SPECS > $O/kbench_grad4b.log 2>&1
echo done > $O/round_done.txt
