#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
B=tools/micro/probe_gran.bin
for g in 0 32 128; do
  $B $g > gpurun_out/r2_gran_plain_$g.log 2>&1 &&
  ncu --metrics dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_gran_ncu_$g.csv $B $g > gpurun_out/r2_gran_ncu_$g.log 2>&1
done
cat gpurun_out/r2_gran_plain_0.log
