#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "usage or ring or exec or fullsize or dropin" > gpurun_out/r2_pytest44.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest44.log
tail -3 gpurun_out/r2_pytest44.log
VARIANTS="main" bash tools/r2_run32.sh
