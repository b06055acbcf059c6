#!/bin/bash
# per-uuid grids: full GPU suite, then uniform and skewed calendars with the global-grid build beside it
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest12.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest12.log
V=$PWD/tensorhive-fixed_b200/lib/variants
L=gpurun_out/r2_skew12.log; rm -f $L
for name in base globalgrid; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$V/libth_b200_$name.so; fi
  for skew in 0 1.2; do
    echo "== $name skew=$skew" >> $L
    TH_SWEEP_SKEW=$skew TH_SWEEP_X4=6 timeout 300 python tools/k2_sweep.py >> $L 2>&1
  done
  echo "== $name cfg5 share" >> $L
  timeout 300 python tools/profile_k2_cfg5.py >> $L 2>&1
done
tail -3 gpurun_out/r2_pytest12.log; grep -E "^==|x4=|k2 us" $L
