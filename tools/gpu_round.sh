#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench lines, kernel bench, ncu launch lists + full captures (outputs under gpurun_out/).
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 300 python tools/kbench.py > $O/kbench.log 2>&1
timeout 120 python tools/peaks.py $O/issue_peaks.json > /dev/null 2> $O/peaks.err
for wl in abf_4k_r7 btf_4k bf_8k_r15; do
  timeout 600 python bench.py --workload $wl > $O/bench_$wl.json 2> $O/bench_$wl.err
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
for wl in abf_4k_r7 btf_4k bf_8k_r15; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$wl.csv \
    python bench.py --workload $wl --steps 8 --warmup 3 --no-cpu-baseline --no-probe > $O/ncu_launch_$wl.log 2>&1
done
for wl in abf_4k_r7 btf_4k bf_8k_r15; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'stencil_kernel|blur_rtv|guide_kernel' -s 6 -c 3 -f -o $O/full_$wl \
    python bench.py --workload $wl --steps 4 --warmup 3 --no-cpu-baseline --no-probe > $O/ncu_full_$wl.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gradient|integral' -s 8 -c 4 -f -o $O/full_streaming \
  python tools/kbench.py grad:3840x2160:3 integral:3840x2160:7 > $O/ncu_full_streaming.log 2>&1
timeout 300 python bench.py --workload stream_1080p --steps 128 > $O/bench_stream_1080p.json 2> $O/bench_stream_1080p.err
timeout 300 python bench.py --workload bf_lenna --steps 200 > $O/bench_bf_lenna.json 2> $O/bench_bf_lenna.err
timeout 300 python bench.py --impl reference --workload bf_lenna --steps 3 --warmup 1 > $O/bench_reference_lenna.json 2> $O/bench_reference_lenna.err
echo done > $O/round_done.txt
