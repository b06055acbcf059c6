#!/usr/bin/env python
"""Diagnostics: the K3 / K5 entry points (current_owner, upcoming, upcoming_owner_flags, violation_join, the UPCOMING row
lists) on a calendar whose starts lie at both ends of int64, with `now` / `period` at the extremes, against a truth
computed with Python integers (no wrap-around).  `now + period` beyond INT64_MAX is read as "no upper bound"."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import tensorhive_fixed_b200 as th                      # noqa: E402
from helpers import csr_to_lists                        # noqa: E402

I = np.iinfo(np.int64)
IMIN, IMAX = int(I.min), int(I.max)


def world(seed):
    rng = np.random.default_rng(seed)
    lo = IMIN + 1 + np.sort(rng.integers(0, 10_000, size=30))
    mid = np.sort(rng.integers(-5_000, 5_000, size=30))
    hi = IMAX - 20_000 + np.sort(rng.integers(0, 10_000, size=30))
    start = np.concatenate([lo, mid, hi, mid + 7]).astype(np.int64)
    end = (start + rng.integers(0, 3_000, size=start.size)).astype(np.int64)
    n = start.size
    return {'id': rng.permutation(np.arange(1, n + 1)).astype(np.int64),
            'uuid_id': np.concatenate([np.zeros(90, np.int64), np.ones(30, np.int64)]),
            'start': start, 'end': end, 'owner': rng.integers(1, 5, size=n).astype(np.int64),
            'cancelled': (rng.random(n) < 0.2).astype(np.uint8), 'n_uuids': 3}, rng


def rows_of(cal, u):
    return [(int(cal['id'][j]), int(cal['start'][j]), int(cal['end'][j]), int(cal['owner'][j]), int(cal['cancelled'][j]))
            for j in range(len(cal['id'])) if int(cal['uuid_id'][j]) == u]


def truth_current(cal, u, now):
    ids = sorted(r[0] for r in rows_of(cal, u) if not r[4] and r[1] <= now <= r[2])
    return ids


def truth_upcoming(cal, u, now, period):
    hi = now + period                                    # Python int: may exceed INT64_MAX = no upper bound
    rs = [r for r in rows_of(cal, u) if not r[4] and ((r[1] < now and r[2] > now) or (r[1] >= now and r[1] <= hi))]
    rs.sort(key=lambda r: (r[1], r[0]))
    return rs


def run(verbose=True):
    """-> number of mismatches over both grid plans and three worlds."""
    bad = 0
    for plan in ('shared', 'per_uuid'):
        os.environ['TH_GRID_PLAN'] = plan
        for seed in range(3):
            cal, rng = world(seed)
            own_of = {int(i): int(o) for i, o in zip(cal['id'], cal['owner'])}
            snap = th.Snapshot(0)
            snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=3)
            anchors = [int(x) for x in np.concatenate([cal['start'], cal['end']])]
            nows = [IMIN, IMIN + 1, IMAX, IMAX - 1, 0] + [int(rng.choice(anchors)) + int(rng.integers(-2, 3)) for _ in range(12)]
            nows = [min(max(x, IMIN), IMAX) for x in nows]
            periods = [0, 1, 2_500, IMAX, IMAX - 7]
            uu = np.arange(-1, 4, dtype=np.int64)
            for now in nows:
                res, own = snap.current_owner(now)
                for u in range(3):
                    cur = truth_current(cal, u, now)
                    want = cur[0] if cur else -1
                    if int(res[u]) != want or int(own[u]) != (own_of[want] if cur else -1):
                        bad += 1
                        print(plan, seed, 'current_owner', now, u, 'got', int(res[u]), 'want', want)
                pu = rng.integers(-1, 4, size=40).astype(np.int64)
                po = rng.integers(-2, 5, size=40).astype(np.int64)
                for strict in (False, True):
                    fl, rid = snap.violation_join(now, strict, pu, po)
                    for j in range(40):
                        cur = truth_current(cal, int(pu[j]), now) if 0 <= pu[j] < 3 else []
                        want = (po[j] != own_of[cur[0]], cur[0] if po[j] != own_of[cur[0]] else -1) if cur else (strict, -1)
                        if (bool(fl[j]), int(rid[j])) != (bool(want[0]), int(want[1])):
                            bad += 1
                            print(plan, seed, 'violation_join', now, strict, int(pu[j]), int(po[j]), 'got', bool(fl[j]), int(rid[j]), 'want', want)
                for period in periods:
                    fid, fst, fow, cnt = snap.upcoming(now, period, uu)
                    hi = min(now + period, IMAX)
                    offu, idu, _ = snap.conflict_sets(uu, np.full(len(uu), now), np.full(len(uu), hi), None, th.MODE_UPCOMING)
                    lists = csr_to_lists(offu, idu)
                    to = rng.integers(1, 5, size=len(uu)).astype(np.int64)
                    flags = {t: snap.upcoming_owner_flags(now, period, uu, to, t) for t in (th.OWNER_ANY, th.OWNER_OTHER, th.OWNER_SAME)}
                    for j, u in enumerate(uu):
                        rs = truth_upcoming(cal, int(u), now, period) if 0 <= u < 3 else []
                        ids = [r[0] for r in rs]
                        if lists[j] != ids:
                            bad += 1
                            print(plan, seed, 'sets UPCOMING', now, period, int(u), 'got', lists[j][:6], 'want', ids[:6])
                        w = (ids[0] if ids else -1, rs[0][1] if rs else -1, rs[0][3] if rs else -1, len(ids))
                        g = (int(fid[j]), int(fst[j]), int(fow[j]), int(cnt[j]))
                        if g != w:
                            bad += 1
                            print(plan, seed, 'upcoming', now, period, int(u), 'got', g, 'want', w)
                        o = int(to[j])
                        wf = {th.OWNER_ANY: len(ids) > 0, th.OWNER_OTHER: any(r[3] != o for r in rs), th.OWNER_SAME: any(r[3] == o for r in rs)}
                        for t in wf:
                            if bool(flags[t][j]) != wf[t]:
                                bad += 1
                                print(plan, seed, 'owner_flags', t, now, period, int(u), o, 'got', bool(flags[t][j]), 'want', wf[t])
            snap.close()
    return bad


if __name__ == '__main__':
    n_bad = run()
    print('mismatches:', n_bad)
    sys.exit(1 if n_bad else 0)
