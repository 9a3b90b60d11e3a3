#!/bin/bash
# development aid: knob sweep of the walk32 kernels + one ncu capture; every step under its own timeout, no stdin
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
P=bidirectional-path-tracer_b200
run() { TAG="$1" timeout 120 python tools/diag_step.py 4 >> gpurun_out/sweep.log 2>&1 || echo "$1 FAILED" >> gpurun_out/sweep.log; }
: > gpurun_out/sweep.log
timeout 400 python -m pytest tests/test_trace_gpu.py tests/test_render_gpu.py -m gpu -q -x --deselect tests/test_render_gpu.py::test_split_walk_and_parking_keep_the_film > gpurun_out/t1.log 2>&1; tail -3 gpurun_out/t1.log
STREAMS=0 run "default"
BDPT_WALK32=0 run "old"
BDPT_W32_NSTEP=2 run "nstep2"
BDPT_W32_NSTEP=8 run "nstep8"
BDPT_W32_TLEAF=4 run "tleaf4"
BDPT_W32_TLEAF=16 run "tleaf16"
BDPT_W32_IDLE=16 run "idle16"
BDPT_W32_IDLE=16 BDPT_W32_NSTEP=8 BDPT_W32_TLEAF=12 run "idle16_nstep8_tleaf12"
cat gpurun_out/sweep.log | grep -v Triangles
# ncu: one launch of each walk kernel, streams=1
STREAMS=1 timeout 200 python tools/diag_step.py 1 > gpurun_out/plain.log 2>&1 && \
STREAMS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_trace_w32 -s 12 -c 1 -o gpurun_out/r2_trace_w32 -f python tools/diag_step.py 1 > gpurun_out/ncu1.log 2>&1
STREAMS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_shadow_w32 -s 12 -c 1 -o gpurun_out/r2_shadow_w32 -f python tools/diag_step.py 1 > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/ncu1.log; tail -n 3 gpurun_out/ncu2.log
