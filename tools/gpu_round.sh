#!/bin/bash
# development aid: new-entry-point test alone, then the whole GPU suite, then the default bench; each under a timeout
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?"; tail -n 5 gpurun_out/pytest_noalloc_alone.log
timeout ${1:-1500} python -m pytest tests -m gpu -x -q -p no:cacheprovider --durations=10 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -n 30 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
echo "bench rc=$?"; tail -n 3 gpurun_out/bench_default.err; cat gpurun_out/bench_default.json
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
nproc
