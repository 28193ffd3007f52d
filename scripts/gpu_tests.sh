#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -8
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4
