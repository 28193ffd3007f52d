#!/bin/bash
mkdir -p gpurun_out
echo "== train tests"; timeout 600 python -m pytest tests/test_gpu_train.py -q -m gpu 2>&1 | tail -3
echo "== train (tanh silu)"; python scripts/train_diag.py 2>&1 | tail -4
echo "== sample bench"; timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras 2>&1 | tail -1 | cut -c1-330
export UVA_HEAD_LIB=$PWD/build/libuva_exp.so
echo "== EXP no store wait: train"; python scripts/train_diag.py 2>&1 | tail -4
echo "== EXP sample bench"; timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras 2>&1 | tail -1 | cut -c1-330
