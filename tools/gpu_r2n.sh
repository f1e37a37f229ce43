#!/bin/bash
# round 2, pass N: ABF with the truncating add of a pair as one packed add.rz.f32x2 (lib_rz) vs scalar (lib)
set -u
O=gpurun_out
mkdir -p $O
RZ=$PWD/various_image_processings_b200/lib_rz/libvip_b200.so
SPECS="abf:3840x2160:15 abf:1920x1080:9 abf:3840x2160:31 abf:3840x2160:15:10:100 abf:3840x270:15"
timeout 200 python tools/kbench.py $SPECS > $O/kbench_abf_rz_scalar.log 2>&1
VIP_B200_LIB=$RZ timeout 200 python tools/kbench.py $SPECS > $O/kbench_abf_rz_packed.log 2>&1
timeout 200 python tools/kbench.py $SPECS > $O/kbench_abf_rz_scalar2.log 2>&1
VIP_B200_LIB=$RZ timeout 600 python -m pytest tests -m gpu -x -q -k "abf or property or strip or config_full_frame" > $O/pytest_abf_rz.log 2>&1; echo "pytest rc=$?" >> $O/pytest_abf_rz.log
echo done > $O/round_done.txt
