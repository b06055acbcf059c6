#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$PWD/tensorhive-fixed_b200/lib/variants
L=gpurun_out/r2_variants3.log
rm -f $L
for name in base pipe min5 pipe5 pipe3 ld1 ld2 ld1pipe ld2pipe; do
  if [ $name = base ]; then unset TH_B200_LIB; else export TH_B200_LIB=$V/libth_b200_$name.so; fi
  echo "== $name" >> $L
  TH_SWEEP_X4=6 timeout 300 python tools/k2_sweep.py >> $L 2>&1
  timeout 300 python tools/profile_k2_cfg5.py >> $L 2>&1
done
for name in ld1 ld2 ld2pipe; do
  export TH_B200_LIB=$V/libth_b200_$name.so
  python tools/profile_k2_cfg5.py > gpurun_out/r2_plain_$name.log 2>&1 &&
  ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum --clock-control none -k regex:k2_bitmap -s 1 -c 1 --csv --log-file gpurun_out/r2_ncu_$name.csv python tools/profile_k2_cfg5.py > gpurun_out/r2_ncu_$name.log 2>&1
done
grep -E "^==|x4=|k2 us" $L; tail -n 5 gpurun_out/r2_ncu_*.csv
