#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$PWD/tensorhive-fixed_b200/lib/variants/libth_b200_oldk4.so
for i in 1 2; do
  TH_B200_LIB=$V timeout 600 python bench.py --steps 20 --warmup 5 --skip-cpu --skip-cfg5 --skip-sweep > gpurun_out/r2_ab_old$i.json 2>/dev/null
  timeout 600 python bench.py --steps 20 --warmup 5 --skip-cpu --skip-cfg5 --skip-sweep > gpurun_out/r2_ab_new$i.json 2>/dev/null
done
python - <<'PY'
import json
for k in ('old','new'):
    for i in (1,2):
        d=json.loads(open('gpurun_out/r2_ab_%s%d.json'%(k,i)).read().strip().splitlines()[-1]); u=d['usage_reduce']; r=u['sample_ring']
        print(k,i,'k2 %.2f us'%(1e3*d['ms_per_step']),'k4 %.3f'%u['ms_per_step'],'shuf %.3f'%u['windows_shuffled']['ms_per_step'],'u8 %.3f'%u['narrow_u8']['ms_per_step'],'ring reg %.4f plain %.4f jit %.4f'%(r['regular_stamps']['ring_ms'],r['regular_stamps']['plain_columns_ms'],r['jittered_stamps']['ring_ms']))
PY
