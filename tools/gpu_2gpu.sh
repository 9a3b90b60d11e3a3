#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
timeout 300 python -m pytest tests/test_scene_image.py -m gpu -q -p no:cacheprovider -k "two_gpus" > gpurun_out/pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_2gpu.log
timeout 600 python bench.py --gpus 2 --steps 8 --warmup 3 --skip-cpu --skip-micro --skip-configs > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench rc=$?"; tail -n 6 gpurun_out/bench_2gpu.err
