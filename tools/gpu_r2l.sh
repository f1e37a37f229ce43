#!/bin/bash
# round 2, pass L: packed fp32x2 A/B: everything packed (lib) vs s2/sk scalar (lib_p1) vs all scalar (lib_p0)
set -u
O=gpurun_out
mkdir -p $O
SPECS="bf:7680x4320:31 jbf:3840x2160:17:8:1.7320508 bf:3840x2160:11 bf:1920x1080:9 btf:3840x2160:9"
timeout 200 python tools/kbench.py $SPECS > $O/kbench_pack2.log 2>&1
for v in 1 0; do VIP_B200_LIB=$PWD/various_image_processings_b200/lib_p$v/libvip_b200.so timeout 200 python tools/kbench.py $SPECS > $O/kbench_pack$v.log 2>&1; done
echo done > $O/round_done.txt
