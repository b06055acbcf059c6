#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$PWD/tensorhive-fixed_b200/lib/variants
L=gpurun_out/r2_pipe22.log; rm -f $L
for name in base pipe pipe3; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$V/libth_b200_$name.so; fi
  echo "== $name" >> $L
  TH_SWEEP_X4=6 timeout 300 python tools/k2_sweep.py >> $L 2>&1
  timeout 300 python tools/profile_k2_cfg5.py >> $L 2>&1
done
export TH_B200_LIB=$V/libth_b200_pipe.so
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_adversarial.py -x -q > gpurun_out/r2_pytest_pipe.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_pipe.log
grep -E "^==|x4=|k2 us" $L; tail -3 gpurun_out/r2_pytest_pipe.log
