#!/bin/bash
# final multi-GPU evidence: dist_check + bench at N = 2, 4, 8 (one gpurun call each), results copied to profiles/
cd /root/repo
for n in 2 4 8; do
  tools/gpurun_retry_n.sh $n gpurun_out/r2_final_call_$n.txt 2400 "bash tools/r2_run5.sh $n"
  cp gpurun_out/r2_bench$n.log profiles/r2_bench_n$n.json 2>/dev/null
  cp gpurun_out/r2_dist$n.log profiles/r2_dist_check_n$n.log 2>/dev/null
done
