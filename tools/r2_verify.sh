#!/bin/bash
# round-end sequence as the driver runs it: full GPU suite, smoke, reference arm, bench (defaults)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_verify_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_verify_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_verify_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2_verify_smoke.log
( time timeout 900 python bench.py --impl reference ) > gpurun_out/r2_verify_ref.json 2> gpurun_out/r2_verify_ref.err; echo "ref rc=$?" >> gpurun_out/r2_verify_ref.err
( time timeout 900 python bench.py ) > gpurun_out/r2_verify_bench.json 2> gpurun_out/r2_verify_bench.err; echo "bench rc=$?" >> gpurun_out/r2_verify_bench.err
tail -3 gpurun_out/r2_verify_pytest.log; tail -2 gpurun_out/r2_verify_smoke.log; tail -5 gpurun_out/r2_verify_ref.err; tail -5 gpurun_out/r2_verify_bench.err
