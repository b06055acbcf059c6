#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "usage or ring or exec or fullsize or dropin" > gpurun_out/r2_pytest43.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest43.log
tail -3 gpurun_out/r2_pytest43.log
timeout 600 python tools/ring_fullscale.py > gpurun_out/r2_ring_fullscale2.log 2>&1; cat gpurun_out/r2_ring_fullscale2.log
