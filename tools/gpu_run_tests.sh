#!/bin/bash
# development aid: the whole GPU suite + a short bench, every step under its own timeout, no stdin
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
timeout ${1:-1200} python -m pytest tests -m gpu -q -p no:cacheprovider --durations=8 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -n 40 gpurun_out/pytest_gpu.log
