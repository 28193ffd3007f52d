#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py tests/test_gpu_train.py -q -k other_channel 2>&1 | tail -30
