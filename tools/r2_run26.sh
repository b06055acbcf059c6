#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/exchange_trace.py > gpurun_out/r2_xtrace$N.log 2>&1; echo "rc=$?" >> gpurun_out/r2_xtrace$N.log
grep -E "^rank|rc=" gpurun_out/r2_xtrace$N.log
