#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$PWD/tensorhive-fixed_b200/lib/variants
L=gpurun_out/r2_variants7.log
rm -f $L
for name in base pipe min5 pipe5 pipe3 nohint; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$V/libth_b200_$name.so; fi
  echo "== $name" >> $L
  TH_SWEEP_X4=6 timeout 300 python tools/k2_sweep.py >> $L 2>&1
  timeout 300 python tools/profile_k2_cfg5.py >> $L 2>&1
done
grep -E "^==|x4=|k2 us" $L
