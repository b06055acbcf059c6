import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import tensorhive_fixed_b200 as th
from oracle import brute
from test_gpu_adversarial import _world
from helpers import csr_to_lists
for seed in (3, 7, 11):
    cal, grid, rng = _world(seed)
    u = cal['n_uuids']
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    keys = np.sort(np.concatenate([grid - 1, grid, grid + 1])).astype(np.int64)
    for uu in range(u):
        m = len(keys)
        qu = np.full(m, uu, np.int64); qs = np.full(m, int(grid[0]) - 5, np.int64); qf = np.full(m, -1, np.int64)
        off, ids, canc = snap.conflict_sets(qu, qs, keys, qf, th.MODE_STRICT)
        got = csr_to_lists(off, ids)
        want = brute.conflict_sets(cal, qu, qs, keys, qf, brute.STRICT)
        bits = snap.conflict_bitmap(qu, qs, keys, qf)
        wb = brute.conflict_bits(cal, qu, qs, keys, qf)
        for i in range(m):
            if got[i] != want[i] or bits[i] != wb[i]:
                print('seed', seed, 'u', uu, 'qe-T0', int(keys[i]) - int(grid[0]), 'got', len(got[i]), 'want', len(want[i]), 'bit', bits[i], wb[i])
    st = np.sort(cal['start'][cal['uuid_id'] == 0]) - int(grid[0])
    print('seed', seed, 'starts of uuid 0:', np.unique(st, return_counts=True))
    snap.close()
print('debug done')
