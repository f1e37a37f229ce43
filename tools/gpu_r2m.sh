#!/bin/bash
# round 2, pass M: blur/mRTV with packed magnitude-sum chains (lib_b2) vs scalar (lib)
set -u
O=gpurun_out
mkdir -p $O
B2=$PWD/various_image_processings_b200/lib_b2/libvip_b200.so
SPECS="btf:3840x2160:9 btf:1920x1080:9 btf:7680x4320:9 btf:3840x2160:5 btf:3840x2160:15"
timeout 200 python tools/kbench.py $SPECS > $O/kbench_blur_scalar.log 2>&1
VIP_B200_LIB=$B2 timeout 200 python tools/kbench.py $SPECS > $O/kbench_blur_packed.log 2>&1
VIP_B200_LIB=$B2 timeout 600 python -m pytest tests -m gpu -x -q -k "btf or blur or guide or config_full_frame or property" > $O/pytest_blur_packed.log 2>&1; echo "pytest rc=$?" >> $O/pytest_blur_packed.log
VIP_B200_LIB=$B2 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'blur_rtv|guide' -c 12 --csv --log-file $O/launches_blur_packed.csv python tools/kbench.py btf:3840x2160:9 > /dev/null 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'blur_rtv|guide' -c 12 --csv --log-file $O/launches_blur_scalar.csv python tools/kbench.py btf:3840x2160:9 > /dev/null 2>&1
echo done > $O/round_done.txt
