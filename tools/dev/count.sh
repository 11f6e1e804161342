#!/bin/bash
# static instruction count of one interior warp step (afx_step.cuh); no GPU needed
set -e
NF=${1:-1024}; HOP=${2:-256}; ST=${3:-true}
mkdir -p /tmp/w
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -DAFX_FORCE_FAST -DONE_NFFT=$NF -DONE_HOP=$HOP -DONE_STOCK=$ST -cubin -o /tmp/w/cnt.cubin tools/dev/one_kernel.cu 2>/dev/null
nvdisasm -g -c /tmp/w/cnt.cubin > /tmp/w/cnt.dis
echo "new step:"; python tools/sass_static.py /tmp/w/cnt.dis step_only_kernel | head -2
