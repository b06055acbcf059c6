#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$PWD/tensorhive-fixed_b200/lib/variants
L=gpurun_out/r2_small17.log; rm -f $L
for name in base nohint globalgrid; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$V/libth_b200_$name.so; fi
  echo "== $name" >> $L
  timeout 300 python tools/time_small_calls.py >> $L 2>&1
done
cat $L
