#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 tools/micro/probe_rate.bin > gpurun_out/r2_probe_rate.log 2>&1
cat gpurun_out/r2_probe_rate.log
