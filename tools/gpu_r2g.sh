#!/bin/bash
# round 2, pass G: TMA (bulk-copy) tile staging A/B for the bilateral kernel
set -u
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q -k "tma_tile_staging or bf_sizes or bf_radii or unaligned" > $O/pytest_tma.log 2>&1; echo "pytest rc=$?" >> $O/pytest_tma.log
SPECS="bf:3840x2160:11 bf:1920x1080:9 bf:3840x2160:9 bf:3840x2160:19 bf:512x512:11 bf:7680x4320:31"
timeout 200 python tools/kbench.py $SPECS > $O/kbench_tma0.log 2>&1
VIP_STENCIL_TMA=1 timeout 200 python tools/kbench.py $SPECS > $O/kbench_tma1.log 2>&1; echo "rc=$?" >> $O/kbench_tma1.log
timeout 200 python tools/kbench.py $SPECS > $O/kbench_tma0_again.log 2>&1
echo done > $O/round_done.txt
