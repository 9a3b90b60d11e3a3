#!/bin/bash
# development aid: bench.py as the driver runs it, under a timeout, no stdin.  usage: gpu_bench.sh <tag> [bench args...]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
tag=$1; shift
timeout 900 python bench.py "$@" > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
echo "bench rc=$?"; tail -n 5 gpurun_out/bench_$tag.err; wc -c gpurun_out/bench_$tag.json
