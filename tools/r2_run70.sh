#!/bin/bash
# launch list of the bench command with the final build (ncu metrics pass; only after the same command exited 0 without ncu)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/r2_plain_b70.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/r2_ncu_b70.log 2>&1
tail -c 300 gpurun_out/r2_plain_b70.log; tail -2 gpurun_out/r2_ncu_b70.log; wc -l gpurun_out/r2_launches_bench.csv
