#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest18.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest18.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/exchange_latency.py > gpurun_out/r2_xlat2b.log 2>&1; echo "rc=$?" >> gpurun_out/r2_xlat2b.log
timeout 900 python bench.py --steps 20 --warmup 5 --skip-cpu > gpurun_out/r2_bench18.json 2> gpurun_out/r2_bench18.err; echo "bench rc=$?" >> gpurun_out/r2_bench18.err
tail -3 gpurun_out/r2_pytest18.log; grep "^{" gpurun_out/r2_xlat2b.log; tail -c 200 gpurun_out/r2_bench18.err
