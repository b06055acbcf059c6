#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest6.log
TH_SWEEP_X4=4,6,8 timeout 300 python tools/k2_sweep.py > gpurun_out/r2_sweep6_cfg2.log 2>&1
python tools/profile_k2_cfg5.py > gpurun_out/r2_prof6_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k2_bitmap -s 1 -c 1 -o gpurun_out/prof_r2_k2_cfg5_l2h python tools/profile_k2_cfg5.py > gpurun_out/r2_ncu6.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench6.log 2> gpurun_out/r2_bench6.err; echo "bench rc=$?" >> gpurun_out/r2_bench6.err
tail -4 gpurun_out/r2_pytest6.log; cat gpurun_out/r2_sweep6_cfg2.log gpurun_out/r2_prof6_plain.log; tail -c 400 gpurun_out/r2_bench6.err
