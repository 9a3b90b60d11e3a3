#!/bin/bash
cd "$(dirname "$0")/.."
exec < /dev/null
BDPT_TIMING=1 timeout 200 python tools/diag_e2e.py 2>&1 | grep -E "bdpt_scene_upload:|upload\(build|reference tree|fast tree" | tail -12
