#!/bin/bash
# usage: tools/gpurun_retry_n.sh <gpus> <log> <timeout> <command...>
n=$1; log=$2; to=$3; shift 3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus "$n" --timeout "$to" -- "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
