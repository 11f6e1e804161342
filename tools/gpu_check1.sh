#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; grep -E "knife edge|flip rate|passed|failed|Error|error" gpurun_out/pytest_gpu.log | tail -15
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
AFX_KNIFE_N=100000 timeout 900 python -m pytest tests/test_gpu_knife_edge.py -m gpu -x -q -s -k flip_rate > gpurun_out/knife_100k.log 2>&1; echo "knife100k exit $?"; grep -E "flip rate|passed|failed" gpurun_out/knife_100k.log
