#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
N=${1:-4}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/exchange_latency.py > gpurun_out/r2_xlat$N.log 2>&1; echo "rc=$?" >> gpurun_out/r2_xlat$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --skip-usage > gpurun_out/r2_bench${N}b.log 2> gpurun_out/r2_bench${N}b.err; echo "bench rc=$?" >> gpurun_out/r2_bench${N}b.err
grep "^{" gpurun_out/r2_xlat$N.log; tail -c 300 gpurun_out/r2_bench${N}b.err
