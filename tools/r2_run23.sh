#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest23.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest23.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench23.json 2> gpurun_out/r2_bench23.err; echo "bench rc=$?" >> gpurun_out/r2_bench23.err
python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/r2_plain_b.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/r2_ncu_b.log 2>&1
python tools/profile_step.py > gpurun_out/r2_plain_ps.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k2_bitmap|k4_usage" -c 9 -o gpurun_out/prof_r2_k2_k4 python tools/profile_step.py > gpurun_out/r2_ncu_ps.log 2>&1
tail -3 gpurun_out/r2_pytest23.log; tail -c 200 gpurun_out/r2_bench23.err; tail -2 gpurun_out/r2_ncu_b.log; tail -2 gpurun_out/r2_ncu_ps.log
