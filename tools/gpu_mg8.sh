#!/bin/bash
# 8-GPU pass, short (charged 8x): single-process driver bench, torchrun bench lines at N=8 (strong BF 8K strips, weak 1080p stream).
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
timeout 90 python tools/mg_bench.py --gpus $N --frames 512 > $O/mg_bench_$N.jsonl 2> $O/mg_bench_$N.err
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29517"
for wl in bf_8k_r15 stream_1080p abf_4k_r7; do
  timeout 60 $TR --nproc-per-node $N bench.py --gpus $N --workload $wl --no-cpu-baseline --steps 30 > $O/bench_${wl}_n$N.json 2> $O/bench_${wl}_n$N.err
done
echo done > $O/mg_done_$N.txt
