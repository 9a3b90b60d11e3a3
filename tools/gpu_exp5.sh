#!/bin/bash
# development aid: workers x wavefront size with the batched shadow stage
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
run() { TAG="$1" timeout 150 python tools/diag_step.py 6 >> gpurun_out/sweep5.log 2>&1 || echo "$1 FAILED" >> gpurun_out/sweep5.log; }
: > gpurun_out/sweep5.log
for s in 2 3 4 5 6 8; do for b in 1048576 2097152 4194304; do STREAMS=$s BATCH=$b run "streams${s}_batch${b}"; done; done
BDPT_SHADOW_BATCH=8000000 STREAMS=6 run "streams6_sb8M"
BDPT_WALK_SLICE=2 run "slice2"; BDPT_WALK_SLICE=8 run "slice8"; BDPT_MIN_IDLE=16 run "idle16"; BDPT_MIN_IDLE=4 run "idle4"
grep -v Triangles gpurun_out/sweep5.log
