#!/bin/bash
# round 2, pass F2: 3 warp groups x 80 registers (lib) vs 4 warp groups x 64 registers (lib_g4)
set -u
O=gpurun_out
mkdir -p $O
G4=$PWD/various_image_processings_b200/lib_g4/libvip_b200.so
SPECS="bf:7680x4320:31 abf:3840x2160:15 jbf:3840x2160:17:8:1.7320508 bf:3840x2160:11 bf:1920x1080:9 abf:1920x1080:9 btf:3840x2160:9 btf:1920x1080:9 bf:512x512:11 bf:7680x540:31 abf:3840x270:15"
VIP_B200_LIB=$G4 timeout 120 python tools/kbench.py $SPECS > $O/kbench_g4.log 2>&1; echo "rc=$?" >> $O/kbench_g4.log
timeout 120 python tools/kbench.py integral:3840x2160:7 integral:7680x4320:15 >> $O/kbench_g4.log 2>&1
VIP_B200_LIB=$G4 timeout 600 python -m pytest tests -m gpu -x -q -k "not bench_two_ranks" > $O/pytest_gpu_g4.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_g4.log
echo done > $O/round_done.txt
