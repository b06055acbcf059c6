#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
N=${1:-2}
mkdir -p gpurun_out
V=$PWD/tensorhive-fixed_b200/lib/variants
for name in base fence; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$V/libth_b200_$name.so; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/exchange_latency.py > gpurun_out/r2_xlat_${name}_$N.log 2>&1
  echo "== $name"; grep "^{" gpurun_out/r2_xlat_${name}_$N.log
done
