#!/usr/bin/env python
"""Wall-clock time of snapshot reloads (device-resident columns) at two sizes."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th
from tensorhive_fixed_b200 import synth
dev = torch.device('cuda:0')
for n, u in ((1_000_000, 4096), (12_500_000, 8192)):
    cal = synth.make_calendar(n, u, seed=3)
    cols = [torch.from_numpy(np.ascontiguousarray(cal[k])).to(dev) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')]
    snap = th.Snapshot(0)
    for _ in range(3):
        snap.load(*cols, n_uuids=u)
    ts = []
    for _ in range(10):
        torch.cuda.synchronize(); t0 = time.perf_counter(); snap.load(*cols, n_uuids=u); ts.append(1e3 * (time.perf_counter() - t0))
    print('n=%d reload ms: median %.3f min %.3f passes %d' % (n, float(np.median(ts)), min(ts), snap.info()['radix_passes']), flush=True)
    snap.close()
