#!/bin/bash
# round 2, pass B (2 GPUs): the strip split under torchrun (fused and copy halos), premultiplied JBF table A/B
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 > $O/bench_2gpu_fused.json 2> $O/bench_2gpu_fused.err; echo "rc=$?" >> $O/bench_2gpu_fused.err
timeout 600 $TR bench.py --gpus 2 --halo copy --no-per-workload > $O/bench_2gpu_copy.json 2> $O/bench_2gpu_copy.err; echo "rc=$?" >> $O/bench_2gpu_copy.err
timeout 300 python tools/kbench.py jbf:3840x2160:17:8:1.7320508 btf:3840x2160:9 btf:1920x1080:9 > $O/kbench_premul1.log 2>&1
VIP_STENCIL_PREMUL=0 timeout 300 python tools/kbench.py jbf:3840x2160:17:8:1.7320508 btf:3840x2160:9 btf:1920x1080:9 > $O/kbench_premul0.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -k "not config_full_frame and not bench_two_ranks" > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
echo done > $O/round_done.txt
