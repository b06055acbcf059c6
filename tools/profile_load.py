#!/usr/bin/env python
"""Snapshot load (K1 + index build) of a 12.5 M-row / 8192-uuid calendar from device-resident columns, twice."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import synth                 # noqa: E402

n, u = int(os.environ.get('TH_LOAD_N', '12500000')), int(os.environ.get('TH_LOAD_U', '8192'))
cal = synth.make_calendar(n, u, seed=3)
dev = torch.device('cuda:0')
cols = [torch.from_numpy(np.ascontiguousarray(cal[k])).to(dev) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')]
snap = th.Snapshot(0)
for _ in range(2):
    snap.load(*cols, n_uuids=u)
torch.cuda.synchronize()
print(snap.info())
