#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
timeout 600 python -m pytest tests/test_media.py tests/test_render_gpu.py tests/test_mis_gpu.py tests/test_textures.py tests/test_instancing.py -m gpu -q -x -p no:cacheprovider > gpurun_out/t_media.log 2>&1; tail -12 gpurun_out/t_media.log
TAG=now timeout 150 python tools/diag_step.py 6 2>&1 | grep -v Triangles
TAG=now timeout 150 python tools/diag_step.py 6 2>&1 | grep -v Triangles
