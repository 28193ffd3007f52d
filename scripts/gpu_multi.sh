#!/bin/bash
# multi-GPU check: the bench under torchrun (row sharding, no collective) and the training step with the NCCL all-reduce
N=${1:-2}
mkdir -p gpurun_out
echo "== sample bench x$N"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_x$N.json | cut -c1-900
echo "== reference arm x$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>&1 | tail -1 | cut -c1-300
echo "== train bench x$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --workload train --gpus $N --steps 5 --warmup 2 2>&1 | tail -1 | tee gpurun_out/train_x$N.json | cut -c1-600
