#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=tensorhive-fixed_b200/lib/variants
timeout 900 python -m pytest tests/test_gpu_exec_vectors.py tests/test_gpu_dropin.py tests/test_gpu_delta.py -x -q > gpurun_out/r2_pytest2.log 2>&1; tail -5 gpurun_out/r2_pytest2.log
for name in base pipe min5 pipe5 pipe3; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$PWD/$V/libth_b200_$name.so; fi
  echo "== $name" >> gpurun_out/r2_variants.log
  TH_SWEEP_X4=6 timeout 300 python tools/k2_sweep.py >> gpurun_out/r2_variants.log 2>&1
  timeout 300 python tools/profile_k2_cfg5.py >> gpurun_out/r2_variants.log 2>&1
done
unset TH_B200_LIB
python tools/profile_k2_cfg5.py > gpurun_out/r2_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k2_bitmap -s 1 -c 1 -o gpurun_out/prof_r2_k2_cfg5 python tools/profile_k2_cfg5.py > gpurun_out/r2_ncu.log 2>&1
cat gpurun_out/r2_variants.log; tail -3 gpurun_out/r2_ncu.log
