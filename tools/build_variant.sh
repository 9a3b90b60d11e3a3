#!/bin/bash
# development aid: libbdpt variant with extra -D flags for render.cu / api.cu.  usage: build_variant.sh <name> <flags...>  -> bidirectional-path-tracer_b200/variants/libbdpt_<name>.so
set -e
cd "$(dirname "$0")/../bidirectional-path-tracer_b200/csrc"
name=$1; shift
mkdir -p ../variants build/v_$name
NV="/usr/local/cuda/bin/nvcc $* -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -fmad=false -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v"
$NV -c render.cu -o build/v_$name/render.o 2> build/v_$name/render.ptxas.log &
$NV -c api.cu -o build/v_$name/api.o 2> build/v_$name/api.ptxas.log &
wait
objs=""
for o in build/*.o; do b=$(basename $o); if [ $b = render.o ] || [ $b = api.o ]; then objs="$objs build/v_$name/$b"; else objs="$objs $o"; fi; done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libbdpt_$name.so $objs -lz -ldl
echo built ../variants/libbdpt_$name.so
