#!/bin/bash
# round 2, pass E: parity with the tile-parallel integral image / PDL gradient / graph replay, their timings
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_e.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_e.log
timeout 300 python tools/kbench.py grad:3840x2160:3 grad:7680x4320:3 grad:1920x1080:3 integral:3840x2160:7 integral:7680x4320:15 integral:1920x1080:4 > $O/kbench_streaming.log 2>&1
VIP_PDL=0 timeout 300 python tools/kbench.py grad:3840x2160:3 grad:7680x4320:3 grad:1920x1080:3 > $O/kbench_streaming_nopdl.log 2>&1
KB_STREAM=1 timeout 300 python tools/kbench.py btf:3840x2160:9 btf:1920x1080:9 btf:512x512:9 > $O/kbench_btf_nograph.log 2>&1
KB_STREAM=1 VIP_BTF_GRAPH=1 timeout 300 python tools/kbench.py btf:3840x2160:9 btf:1920x1080:9 btf:512x512:9 > $O/kbench_btf_graph.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:'integral' -s 4 -c 4 -f -o $O/full_integral python tools/kbench.py integral:3840x2160:7 > $O/ncu_integral.log 2>&1
timeout 120 python tools/btf_bounds_probe.py > $O/btf_bounds.log 2>&1
echo done > $O/round_done.txt
