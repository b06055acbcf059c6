#!/bin/bash
# N GPUs (N = $1): multi-GPU parity tool, then the bench exactly as the driver launches it
cd "$GRAFT_REPO_ROOT" || exit 1
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py > gpurun_out/r2_dist$N.log 2>&1; echo "dist_check rc=$?" >> gpurun_out/r2_dist$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench$N.log 2> gpurun_out/r2_bench$N.err; echo "bench rc=$?" >> gpurun_out/r2_bench$N.err
tail -2 gpurun_out/r2_dist$N.log; tail -c 1500 gpurun_out/r2_bench$N.err; tail -c 600 gpurun_out/r2_bench$N.log
