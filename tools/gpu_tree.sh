#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
BDPT_GPU_TREE_SAH=1 timeout 300 python -m pytest tests/test_gpu_tree.py -m gpu -q -x -p no:cacheprovider > gpurun_out/t_tree.log 2>&1; tail -3 gpurun_out/t_tree.log
BDPT_TIMING=1 timeout 300 python tools/diag_tree.py 2>&1 | grep -E "host_sah|gpu_sah|gpu_lbvh|fast tree on the GPU" | tail -12
