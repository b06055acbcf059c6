#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest14.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest14.log
L=gpurun_out/r2_skew14.log; rm -f $L
for plan in auto shared per_uuid; do
  if [ $plan = auto ]; then unset TH_GRID_PLAN; else export TH_GRID_PLAN=$plan; fi
  for skew in 0 1.2; do
    echo "== plan=$plan skew=$skew" >> $L
    TH_SWEEP_SKEW=$skew TH_SWEEP_X4=6 timeout 300 python tools/k2_sweep.py >> $L 2>&1
  done
done
unset TH_GRID_PLAN
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench14.json 2> gpurun_out/r2_bench14.err; echo "bench rc=$?" >> gpurun_out/r2_bench14.err
tail -3 gpurun_out/r2_pytest14.log; grep -E "^==|x4=" $L; tail -c 300 gpurun_out/r2_bench14.err
