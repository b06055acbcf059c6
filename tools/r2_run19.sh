#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "usage or ring or exec or fullsize" > gpurun_out/r2_pytest19.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest19.log
timeout 900 python bench.py --steps 20 --warmup 5 --skip-cpu > gpurun_out/r2_bench19.json 2> gpurun_out/r2_bench19.err; echo "bench rc=$?" >> gpurun_out/r2_bench19.err
tail -3 gpurun_out/r2_pytest19.log; tail -c 200 gpurun_out/r2_bench19.err
