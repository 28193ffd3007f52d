#!/bin/bash
for r in 0 8 16 24; do for e in 1 0; do echo "== reserve $r eager $e"; UVA_TRAIN_RESERVE_SMS=$r UVA_EAGER_BACKWARD=$e python scripts/train_diag.py 2>&1 | sed -n 3,3p; done; done
