#!/usr/bin/env python
"""Diagnostics: odd table shapes (a million uuids and ten rows, one uuid with all rows, no rows at all, ...) through the
bitmap, the row lists, the owner table, upcoming and the violation join, against the brute-force oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import synth                 # noqa: E402
from oracle import brute                                # noqa: E402
from helpers import csr_to_lists                        # noqa: E402


def run():
    bad = 0
    for plan in ('auto', 'shared', 'per_uuid'):
        if plan == 'auto':
            os.environ.pop('TH_GRID_PLAN', None)
        else:
            os.environ['TH_GRID_PLAN'] = plan
        for n, u in ((10, 1_000_000), (0, 100_000), (0, 1), (1, 1), (50_000, 1), (3, 70_000), (200_000, 65_536), (1000, 999)):
            rng = np.random.default_rng(n + u)
            k = max(1, min(u, n))
            cal = synth.make_calendar(n, k, seed=n + 3)
            if n:
                mapping = rng.choice(u, size=k, replace=False).astype(np.int64)      # the used uuids lie anywhere in [0, u)
                cal['uuid_id'] = mapping[cal['uuid_id']]
            cal['n_uuids'] = u
            snap = th.Snapshot(0)
            snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
            m = 3000
            known = cal['uuid_id'][rng.integers(0, n, size=m)] if n else np.zeros(m, np.int64)
            qu = np.where(rng.random(m) < 0.7, known, rng.integers(-2, u + 2, size=m)).astype(np.int64)
            t0 = int(cal['start'].min()) if n else synth.T0_US
            t1 = int(cal['end'].max()) if n else synth.T0_US + synth.DAY_US
            qs = rng.integers(t0 - synth.DAY_US, t1 + synth.DAY_US, size=m).astype(np.int64)
            qe = qs + rng.integers(0, 5 * synth.DAY_US, size=m)
            qf = np.where(rng.random(m) < 0.5, -1, rng.choice(cal['id'], size=m) if n else -1).astype(np.int64)
            for mode, bmode in ((th.MODE_STRICT, brute.STRICT), (th.MODE_INCLUSIVE, brute.INCLUSIVE), (th.MODE_UPCOMING, brute.UPCOMING)):
                f = qf if mode == th.MODE_STRICT else None
                fb = qf if mode == th.MODE_STRICT else np.full(m, -1, np.int64)
                if mode != th.MODE_UPCOMING:
                    got = np.asarray(snap.conflict_bitmap(qu, qs, qe, f, mode), bool)
                    want = np.asarray(brute.conflict_bits(cal, qu, qs, qe, fb, bmode), bool)
                    if not np.array_equal(got, want):
                        bad += 1
                        print(plan, n, u, 'bitmap mode', mode, int((got != want).sum()), 'mismatches')
                off, ids, _ = snap.conflict_sets(qu[:500], qs[:500], qe[:500], None if f is None else f[:500], mode)
                if csr_to_lists(off, ids) != brute.conflict_sets(cal, qu[:500], qs[:500], qe[:500], fb[:500], bmode):
                    bad += 1
                    print(plan, n, u, 'sets mode', mode, 'mismatch')
            now = int(rng.integers(t0, t1 + 1))
            res, own = snap.current_owner(now)
            eres, eown = brute.current_owner(cal, u, now)
            if not (np.array_equal(res, eres) and np.array_equal(own, eown)):
                bad += 1
                print(plan, n, u, 'current_owner mismatch')
            uu = np.unique(np.concatenate([qu[:300], [-1, u, u - 1, 0]])).astype(np.int64)
            got = snap.upcoming(now, 3 * synth.DAY_US, uu)
            want = brute.upcoming_summary(cal, uu, now, 3 * synth.DAY_US)
            if not all(np.array_equal(a, b) for a, b in zip(got, want)):
                bad += 1
                print(plan, n, u, 'upcoming mismatch')
            snap.close()
    return bad


if __name__ == '__main__':
    n_bad = run()
    print('mismatches:', n_bad)
    sys.exit(1 if n_bad else 0)
