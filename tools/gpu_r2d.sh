#!/bin/bash
# round 2, pass D (8 GPUs, charged 8x -- keep it short): bench.py under torchrun at N = 8 with the three halo transports,
# the single-process driver, and the multi-GPU tests on real devices
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/topo_$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR bench.py --gpus $N > $O/bench_${N}gpu_fused.json 2> $O/bench_${N}gpu_fused.err; echo "rc=$?" >> $O/bench_${N}gpu_fused.err
timeout 150 $TR bench.py --gpus $N --halo copy --no-per-workload > $O/bench_${N}gpu_copy.json 2> $O/bench_${N}gpu_copy.err; echo "rc=$?" >> $O/bench_${N}gpu_copy.err
VIP_BENCH_TRANSPORT=nccl timeout 150 $TR bench.py --gpus $N --no-per-workload > $O/bench_${N}gpu_nccl.json 2> $O/bench_${N}gpu_nccl.err; echo "rc=$?" >> $O/bench_${N}gpu_nccl.err
timeout 150 python tools/mg_bench.py --gpus $N --frames 256 > $O/mg_bench_$N.jsonl 2> $O/mg_bench_$N.err
timeout 200 python -m pytest tests -m gpu -x -q -k "multigpu or strip_peer or link" > $O/pytest_mg_$N.log 2>&1; echo "pytest rc=$?" >> $O/pytest_mg_$N.log
echo done > $O/mg_done_$N.txt
