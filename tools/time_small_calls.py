#!/usr/bin/env python
"""Wall-clock of the small host-array calls a controller makes (calendar range view, one would_interfere) and of the
snapshot (re)load / delta on BASELINE config 2.  Used to compare library builds (TH_B200_LIB)."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import native, synth         # noqa: E402
N, U = 1_000_000, 4096
cal = synth.make_calendar(N, U, seed=20260102)
dev = torch.device('cuda:0')
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cols = [t(cal[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')]
snap = th.Snapshot(0)
ld = []
for _ in range(6):
    torch.cuda.synchronize(); a = time.perf_counter(); snap.load(*cols, n_uuids=U); ld.append(1e3 * (time.perf_counter() - a))
up = {k: cal[k][:1000].copy() for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')}
up['end'] = up['end'] + 1
dl = []
for _ in range(6):
    torch.cuda.synchronize(); a = time.perf_counter()
    snap.apply_delta(snap, U, up['id'], up['uuid_id'], up['start'], up['end'], up['owner'], up['cancelled'], np.zeros(0, np.int64))
    dl.append(1e3 * (time.perf_counter() - a))
now = int(np.median(cal['start']))
r_u = np.arange(U, dtype=np.int64); r_s = np.full(U, now, np.int64); r_e = r_s + 30 * synth.DAY_US
for mode in (1, 2):
    snap.conflict_sets(r_u, r_s, r_e, None, mode)
    reps = []
    for _ in range(20):
        a = time.perf_counter(); off, ids, canc = snap.conflict_sets(r_u, r_s, r_e, None, mode); reps.append(1e3 * (time.perf_counter() - a))
    print('calendar range mode %d: %.3f ms (min %.3f), %d rows' % (mode, np.median(reps), min(reps), len(ids)))
one = [np.ascontiguousarray(cal[k][:1]) for k in ('uuid_id', 'start', 'end')] + [np.full(1, -1, np.int64)]
for _ in range(50):
    snap.conflict_words(*one)
a = time.perf_counter()
for _ in range(500):
    snap.conflict_words(*one)
print('single would_interfere: %.1f us' % (1e6 * (time.perf_counter() - a) / 500))
print('load %.3f ms (min %.3f)  delta %.3f ms (min %.3f)  lib %s' % (np.median(ld[1:]), min(ld), np.median(dl[1:]), min(dl), native.LIB_PATH))
