#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "sets or adversarial or golden or model or fullsize or config1 or exec" > gpurun_out/r2_pytest72.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest72.log
tail -3 gpurun_out/r2_pytest72.log
for i in 1 2; do timeout 600 python bench.py --steps 20 --warmup 5 --skip-cpu --skip-cfg5 --skip-usage > gpurun_out/r2_b72_$i.json 2>/dev/null; done
python - <<'PY'
import json
for i in (1,2):
    d=json.loads(open('gpurun_out/r2_b72_%d.json'%i).read().strip().splitlines()[-1]); s=d['service_sweep']
    print(i,'k2 %.2f us'%(1e3*d['ms_per_step']),'sets %.4f ms'%s['conflict_sets']['ms'],'rows',s['conflict_sets']['rows_returned'],'range %.4f'%s['calendar_range']['ms_host_call'])
PY
