#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/profile_step.py > gpurun_out/r2_ps39.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k4_usage" -c 6 -o gpurun_out/prof_r2_k4_final python tools/profile_step.py > gpurun_out/r2_ncu39.log 2>&1
tail -2 gpurun_out/r2_ps39.log; tail -2 gpurun_out/r2_ncu39.log
