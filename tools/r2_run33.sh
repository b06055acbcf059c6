#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest33.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest33.log
timeout 300 python tools/profile_step.py > gpurun_out/r2_ps33.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k4_usage" -c 6 -o gpurun_out/prof_r2_k4_converged python tools/profile_step.py > gpurun_out/r2_ncu33.log 2>&1
tail -3 gpurun_out/r2_pytest33.log; tail -2 gpurun_out/r2_ps33.log; tail -2 gpurun_out/r2_ncu33.log
