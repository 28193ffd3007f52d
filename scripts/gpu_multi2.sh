#!/bin/bash
# 2-GPU: NCCL all-reduce through the C ABI (test) + training bench under torchrun
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -5
echo "== train bench x2"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --workload train --gpus 2 --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/train_x2_v6.json | cut -c1-700
