#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "pyin or f0 or smoke or default or c3_c4" 2>&1 | tail -3
python tools/f0_hops.py 740 2>&1 | grep filter
