#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python tools/ring_fullscale.py > gpurun_out/r2_ring_fullscale.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:"k4_usage" -s 3 -c 1 -o gpurun_out/prof_r2_ring_full python tools/ring_fullscale.py > gpurun_out/r2_ncu42.log 2>&1
cat gpurun_out/r2_ring_fullscale.log; tail -3 gpurun_out/r2_ncu42.log
