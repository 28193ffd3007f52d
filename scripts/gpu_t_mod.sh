#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py -q -k "module_sample" 2>&1 | tail -30
