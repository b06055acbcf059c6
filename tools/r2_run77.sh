#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
TH_SETS_LANES=1 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_adversarial.py -m gpu -x -q -k "sets or random_worlds or handful or int64_range" 2>&1 | tail -2
for v in 2 1; do TH_SETS_LANES=$v timeout 600 python bench.py --steps 10 --warmup 3 --skip-cpu --skip-cfg5 --skip-usage > gpurun_out/r2_b77_$v.json 2>/dev/null; done
python - <<'PY'
import json
for v in (2,1):
    d=json.loads(open('gpurun_out/r2_b77_%d.json'%v).read().strip().splitlines()[-1]); s=d['service_sweep']
    print('lanes',v,'sets %.4f ms'%s['conflict_sets']['ms'],'rows',s['conflict_sets']['rows_returned'])
PY
