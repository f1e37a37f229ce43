#!/bin/bash
# round 2, pass K (2 GPUs): final bench.py under torchrun
set -u
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513"
timeout 400 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2gpu_final.json 2> $O/bench_2gpu_final.err; echo "rc=$?" >> $O/bench_2gpu_final.err
timeout 200 $TR bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > $O/bench_2gpu_reference.json 2> $O/bench_2gpu_reference.err; echo "rc=$?" >> $O/bench_2gpu_reference.err
echo done > $O/round_done.txt
