#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/debug_world.py > gpurun_out/r2_debug13.log 2>&1
tail -40 gpurun_out/r2_debug13.log
