#!/bin/bash
# 2-GPU: NCCL all-reduce through the C ABI (test) + training bench under torchrun, early all-reduce on / off
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -5
for e in 1 0; do
echo "== train bench x2, UVA_EARLY_ALLREDUCE=$e"
UVA_EARLY_ALLREDUCE=$e timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$e bench.py --workload train --gpus 2 --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/train_x2_early$e.json | cut -c1-330
done
