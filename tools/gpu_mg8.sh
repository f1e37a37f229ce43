#!/bin/bash
# 8-GPU pass, short: real-device driver tests, single-process driver bench, torchrun bench lines at N=8.
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q -k "multigpu" > $O/pytest_mg_$N.log 2>&1; echo "rc=$?" >> $O/pytest_mg_$N.log
timeout 600 python tools/mg_bench.py --gpus $N --frames 1024 > $O/mg_bench_$N.jsonl 2> $O/mg_bench_$N.err
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29517"
for wl in abf_4k_r7 bf_8k_r15 stream_1080p; do
  timeout 600 $TR --nproc-per-node $N bench.py --gpus $N --workload $wl --no-cpu-baseline > $O/bench_${wl}_n$N.json 2> $O/bench_${wl}_n$N.err
done
echo done > $O/mg_done_$N.txt
