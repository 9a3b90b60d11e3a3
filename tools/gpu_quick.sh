#!/bin/bash
# development aid: a chosen set of GPU tests under a timeout.  usage: gpu_quick.sh <pytest args...>
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
timeout 900 python -m pytest "$@" -m gpu -q -p no:cacheprovider > gpurun_out/pytest_quick.log 2>&1
echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_quick.log
