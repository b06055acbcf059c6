#!/bin/bash
# the whole 100 M-reservation / 65 536-uuid calendar of BASELINE config 5 on ONE GPU, 12.5 M candidates, sampled parity
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/r2_k2_100M_one_gpu.log
( time TH_PROF_N=100000000 TH_PROF_U=65536 TH_PROF_M=12500000 TH_PROF_CHECK=1 timeout 900 python tools/profile_k2_cfg5.py ) >> gpurun_out/r2_k2_100M_one_gpu.log 2>&1
cat gpurun_out/r2_k2_100M_one_gpu.log
