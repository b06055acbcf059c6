#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "usage or ring or exec or fullsize or dropin" > gpurun_out/r2_pytest28.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest28.log
timeout 900 python bench.py --steps 20 --warmup 5 --skip-cpu --skip-cfg5 > gpurun_out/r2_bench28.json 2> gpurun_out/r2_bench28.err; echo "bench rc=$?" >> gpurun_out/r2_bench28.err
TH_PROF_ORDERS=group8,group32,group128,sorted timeout 600 python tools/profile_k2_cfg5.py > gpurun_out/r2_k2_orders.log 2>&1
tail -3 gpurun_out/r2_pytest28.log; tail -c 200 gpurun_out/r2_bench28.err; cat gpurun_out/r2_k2_orders.log | tail -6
