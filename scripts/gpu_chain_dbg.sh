#!/bin/bash
for d in 0 4 8; do echo "== UVA_CHAIN_DBG=$d"; UVA_CHAIN_DBG=$d python scripts/lat_breakdown.py 1 2>&1 | grep "lib sample"; done
