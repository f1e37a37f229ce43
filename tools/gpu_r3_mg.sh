#!/bin/bash
# round 3, N-GPU pass (charged N x): the default bench line under torchrun -- ordered and unordered resident legs of the
# strip split, end-to-end leg, per-workload lines.
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29533"
timeout 240 $TR --nproc-per-node $N bench.py --gpus $N --steps 20 --warmup 3 > $O/r03g_bench_n$N.json 2> $O/r03g_bench_n$N.err; echo rc=$? >> $O/r03g_bench_n$N.err
