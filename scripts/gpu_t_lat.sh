#!/bin/bash
timeout 300 python scripts/lat_breakdown.py 1 2>&1 | tail -5
timeout 300 python scripts/lat_breakdown.py 56 2>&1 | tail -5
