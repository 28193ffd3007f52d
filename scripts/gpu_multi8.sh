#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
echo "== sample bench x$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 2 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_x$N.json | cut -c1-700
echo "== train bench x$N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --workload train --gpus $N --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/train_x$N.json | cut -c1-400
