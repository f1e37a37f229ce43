#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): real-device driver tests, single-process driver bench, torchrun bench lines.
set -u
N=${1:-2}
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/topo_$N.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q -k "multigpu or strips" > $O/pytest_mg_$N.log 2>&1; echo "rc=$?" >> $O/pytest_mg_$N.log
for n in 1 $N; do timeout 600 python tools/mg_bench.py --gpus $n > $O/mg_bench_$n.jsonl 2> $O/mg_bench_$n.err; done
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29517"
for wl in abf_4k_r7 bf_8k_r15 stream_1080p; do
  timeout 600 python bench.py --workload $wl --no-cpu-baseline > $O/bench_${wl}_n1.json 2> $O/bench_${wl}_n1.err
  timeout 600 $TR --nproc-per-node $N bench.py --gpus $N --workload $wl --no-cpu-baseline > $O/bench_${wl}_n$N.json 2> $O/bench_${wl}_n$N.err
done
echo done > $O/mg_done_$N.txt
