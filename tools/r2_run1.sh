#!/bin/bash
# round-2 GPU call 1: GPU tests, K2 grid-density sweeps at config-2 and config-5-share size, full bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_smi.log 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
TH_SWEEP_X4=3,4,6,8,12 timeout 300 python tools/k2_sweep.py > gpurun_out/r2_sweep_cfg2.log 2>&1
TH_SWEEP_N=12500000 TH_SWEEP_U=8192 TH_SWEEP_M=12500000 TH_SWEEP_X4=4,6,8 timeout 600 python tools/k2_sweep.py > gpurun_out/r2_sweep_cfg5.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench1.log 2> gpurun_out/r2_bench1.err; echo "bench rc=$?" >> gpurun_out/r2_bench1.err
tail -3 gpurun_out/r2_pytest.log; cat gpurun_out/r2_sweep_cfg2.log gpurun_out/r2_sweep_cfg5.log; tail -c 600 gpurun_out/r2_bench1.err
