#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest15.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest15.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench15.json 2> gpurun_out/r2_bench15.err; echo "bench rc=$?" >> gpurun_out/r2_bench15.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke15.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2_smoke15.log
tail -3 gpurun_out/r2_pytest15.log; tail -c 300 gpurun_out/r2_bench15.err; tail -2 gpurun_out/r2_smoke15.log
