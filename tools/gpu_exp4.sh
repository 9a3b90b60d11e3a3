#!/bin/bash
# development aid: sweep of the batched shadow stage (BDPT_SHADOW_BATCH = pairs per batch)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
run() { TAG="$1" timeout 150 python tools/diag_step.py 6 >> gpurun_out/sweep4.log 2>&1 || echo "$1 FAILED" >> gpurun_out/sweep4.log; }
: > gpurun_out/sweep4.log
BDPT_SHADOW_BATCH=2000000 timeout 400 python -m pytest tests/test_render_gpu.py tests/test_mis_gpu.py tests/test_media.py tests/test_spheres.py -m gpu -q -x -p no:cacheprovider > gpurun_out/t_batch.log 2>&1; tail -3 gpurun_out/t_batch.log
run "batch0"
for b in 500000 1000000 2000000 4000000 8000000 16000000; do BDPT_SHADOW_BATCH=$b run "batch$b"; done
run "batch0"
STREAMS=1 run "batch0_1stream"
STREAMS=1 BDPT_SHADOW_BATCH=4000000 run "batch4M_1stream"
grep -v Triangles gpurun_out/sweep4.log
