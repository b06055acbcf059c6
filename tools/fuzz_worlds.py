#!/usr/bin/env python
"""Differential fuzzer: random calendars of several styles (tiny time grid full of ties, bursts that make long buckets,
nested intervals of wildly different lengths, plain synthetic walks), random grid plan, every entry point against the
brute-force oracle.  Self ids are drawn mostly from the candidate's OWN uuid (that is where the top-2 exclusion matters).
usage: python tools/fuzz_worlds.py [n_worlds] [first_seed]"""
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


def make_world(seed):
    rng = np.random.default_rng(seed)
    style = seed % 4
    u = int(rng.integers(1, 40))
    n = int(rng.integers(0, 4000))
    if style == 0:                                       # tiny grid: ties and touching boundaries everywhere
        grid = synth.T0_US + np.arange(int(rng.integers(2, 30))) * int(rng.choice([1, 1000, 3_600_000_000]))
        a, b = rng.choice(grid, size=n), rng.choice(grid, size=n)
        start, end = np.minimum(a, b), np.maximum(a, b)
    elif style == 1:                                     # bursts: hundreds of rows within microseconds, then long gaps
        centers = synth.T0_US + rng.integers(0, 10**12, size=max(1, n // 200 + 1))
        start = rng.choice(centers, size=n) + rng.integers(0, 50, size=n)
        end = start + rng.choice([0, 1, 10, 10**6, 10**11], size=n)
    elif style == 2:                                     # nested intervals of wildly different lengths
        start = synth.T0_US + rng.integers(0, 10**10, size=n)
        end = start + (10 ** rng.integers(0, 13, size=n)).astype(np.int64)
    else:
        cal = synth.make_calendar(n, u, seed=seed)
        start, end = cal['start'], cal['end']
    uuid = rng.integers(0, u, size=n).astype(np.int64) if style != 3 else cal['uuid_id']
    if style != 3 and n and rng.random() < 0.5:          # one hot uuid
        uuid[rng.random(n) < 0.6] = int(rng.integers(0, u))
    cal = {'id': rng.permutation(np.arange(1, 3 * n + 1))[:n].astype(np.int64), 'uuid_id': uuid,
           'start': np.asarray(start, np.int64), 'end': np.asarray(end, np.int64),
           'owner': rng.integers(1, 6, size=n).astype(np.int64),
           'cancelled': (rng.random(n) < rng.choice([0.0, 0.1, 0.5, 0.95])).astype(np.uint8), 'n_uuids': u}
    return cal, rng


def make_queries(cal, rng, m):
    n, u = len(cal['id']), cal['n_uuids']
    if n:
        pick = rng.integers(0, n, size=m)
        qu = np.where(rng.random(m) < 0.85, cal['uuid_id'][pick], rng.integers(-1, u + 1, size=m)).astype(np.int64)
        anchor_s = np.where(rng.random(m) < 0.5, cal['start'][pick], cal['end'][pick])
        anchor_e = np.where(rng.random(m) < 0.5, cal['start'][rng.integers(0, n, size=m)], cal['end'][pick])
        qs = anchor_s + rng.integers(-2, 3, size=m)
        qe = np.maximum(qs, anchor_e + rng.integers(-2, 3, size=m))
        # self ids: mostly rows of the candidate's own uuid
        own = np.where(rng.random(m) < 0.7, cal['id'][pick], cal['id'][rng.integers(0, n, size=m)])
        qf = np.where(rng.random(m) < 0.4, -1, own).astype(np.int64)
    else:
        qu = rng.integers(-1, u + 1, size=m).astype(np.int64)
        qs = synth.T0_US + rng.integers(0, 10**9, size=m)
        qe = qs + rng.integers(0, 10**9, size=m)
        qf = np.full(m, -1, np.int64)
    return qu, qs.astype(np.int64), qe.astype(np.int64), qf


def check_world(seed, verbose=True):
    cal, rng = make_world(seed)
    u = cal['n_uuids']
    plan = ['shared', 'per_uuid', None][seed % 3]
    if plan:
        os.environ['TH_GRID_PLAN'] = plan
    else:
        os.environ.pop('TH_GRID_PLAN', None)
    bad = []
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    m = 1500
    qu, qs, qe, qf = make_queries(cal, rng, m)
    none = np.full(m, -1, np.int64)
    for mode, bmode in ((th.MODE_STRICT, brute.STRICT), (th.MODE_INCLUSIVE, brute.INCLUSIVE)):
        f, fb = (qf, qf) if mode == th.MODE_STRICT else (None, none)
        got = np.asarray(snap.conflict_bitmap(qu, qs, qe, f, mode), bool)
        if not np.array_equal(got, np.asarray(brute.conflict_bits(cal, qu, qs, qe, fb, bmode), bool)):
            bad.append('bitmap mode %d' % mode)
    pu, pf = snap.pack_queries(qu, qf)
    words = snap.conflict_words_packed(pu, qs, qe, pf)
    if not np.array_equal(th.api.unpack_bits(words, m), np.asarray(brute.conflict_bits(cal, qu, qs, qe, qf, brute.STRICT), bool)):
        bad.append('packed bitmap')
    k = 300
    for mode, bmode in ((th.MODE_STRICT, brute.STRICT), (th.MODE_INCLUSIVE, brute.INCLUSIVE), (th.MODE_UPCOMING, brute.UPCOMING)):
        f, fb = (qf[:k], qf[:k]) if mode == th.MODE_STRICT else (None, none[:k])
        off, ids, canc = snap.conflict_sets(qu[:k], qs[:k], qe[:k], f, mode)
        if csr_to_lists(off, ids) != brute.conflict_sets(cal, qu[:k], qs[:k], qe[:k], fb, bmode):
            bad.append('sets mode %d' % mode)
        flag = {int(i): int(c) for i, c in zip(cal['id'], cal['cancelled'])}
        if [flag[int(i)] for i in ids] != [int(c) for c in canc]:
            bad.append('sets cancelled flags mode %d' % mode)
    uu = np.arange(-1, u + 1, dtype=np.int64)
    own_of = {int(i): int(o) for i, o in zip(cal['id'], cal['owner'])}
    for now in [int(x) for x in qs[:4]] + [int(x) for x in qe[:2]]:
        res, own = snap.current_owner(now)
        eres, eown = brute.current_owner(cal, u, now)
        if not (np.array_equal(res, eres) and np.array_equal(own, eown)):
            bad.append('current_owner')
        period = int(rng.choice([0, 1, 50, 10**6, 10**11, 10**13]))
        got = snap.upcoming(now, period, uu)
        want = brute.upcoming_summary(cal, uu, now, period)
        if not all(np.array_equal(a, b) for a, b in zip(got, want)):
            bad.append('upcoming')
        lists = brute.conflict_sets(cal, uu, np.full(len(uu), now), np.full(len(uu), now + period), np.full(len(uu), -1), brute.UPCOMING)
        to = rng.integers(1, 6, size=len(uu)).astype(np.int64)
        for test, fn in ((th.OWNER_ANY, lambda o, w: len(w) > 0), (th.OWNER_OTHER, lambda o, w: any(own_of[x] != o for x in w)),
                         (th.OWNER_SAME, lambda o, w: any(own_of[x] == o for x in w))):
            fl = snap.upcoming_owner_flags(now, period, uu, to, test)
            if [bool(x) for x in fl] != [fn(int(to[j]), lists[j]) for j in range(len(uu))]:
                bad.append('owner_flags %d' % test)
        pu_ = rng.integers(-1, u + 1, size=200).astype(np.int64)
        po_ = rng.integers(-2, 6, size=200).astype(np.int64)
        for strict in (False, True):
            fl, rid = snap.violation_join(now, strict, pu_, po_)
            for j in range(200):
                r = int(eres[pu_[j]]) if 0 <= pu_[j] < u else -1
                want = (po_[j] != own_of[r], r if po_[j] != own_of[r] else -1) if r >= 0 else (strict, -1)
                if (bool(fl[j]), int(rid[j])) != (bool(want[0]), int(want[1])):
                    bad.append('violation_join')
                    break
    snap.close()
    if bad and verbose:
        print('seed', seed, 'style', seed % 4, 'plan', plan, 'n', len(cal['id']), 'u', u, '->', sorted(set(bad)))
    return len(set(bad))


COLS = ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')


def check_delta(seed, verbose=True):
    """Three random deltas (deletes, moved / re-flagged rows, inserts taken from another world, sometimes on new uuids) on
    a world of the seed's style: after each one the snapshot must export exactly what a fresh load of the table exports."""
    cal, rng = make_world(seed)
    other, _ = make_world(seed + 7777)
    u = cal['n_uuids']
    table = {k: cal[k].copy() for k in COLS}
    cur = th.Snapshot(0)
    cur.load(*[table[k] for k in COLS], n_uuids=u)
    spare = th.Snapshot(0)
    next_id = int(table['id'].max()) + 1 if len(table['id']) else 1
    bad = []
    for step in range(3):
        n = len(table['id'])
        del_pos = rng.choice(n, size=int(rng.integers(0, n // 3 + 1)), replace=False) if n else np.zeros(0, np.int64)
        rest = np.setdiff1d(np.arange(n), del_pos)
        upd_pos = rng.choice(rest, size=int(rng.integers(0, len(rest) // 3 + 1)), replace=False) if len(rest) else np.zeros(0, np.int64)
        n_ins = int(rng.integers(0, len(other['id']) + 1))
        sel = rng.choice(len(other['id']), size=n_ins, replace=False) if n_ins else np.zeros(0, np.int64)
        ins = {k: other[k][sel].copy() for k in COLS}
        ins['id'] = np.arange(next_id, next_id + n_ins, dtype=np.int64)
        next_id += n_ins
        u = max(u, other['n_uuids']) if n_ins else u
        upd = {k: table[k][upd_pos].copy() for k in COLS}
        shift = rng.choice([0, 1, -1, 10**6, -10**9], size=len(upd_pos))
        upd['start'] = upd['start'] + shift
        upd['end'] = np.maximum(upd['end'] + rng.choice([0, 1, 10**6], size=len(upd_pos)) + shift, upd['start'])
        upd['cancelled'] = (rng.random(len(upd_pos)) < 0.4).astype(np.uint8)
        ups = {k: np.concatenate([upd[k], ins[k]]) for k in COLS}
        order = rng.permutation(len(ups['id']))
        ups = {k: v[order] for k, v in ups.items()}
        del_ids = table['id'][del_pos].copy()
        keep = np.ones(n, bool)
        keep[del_pos] = False
        keep[upd_pos] = False
        table = {k: np.concatenate([table[k][keep], ups[k]]) for k in COLS}
        in_place = bool((seed + step) & 1)
        dst = cur if in_place else spare
        dst.apply_delta(cur, u, *[ups[k] for k in COLS], del_ids)
        if not in_place:
            cur, spare = dst, cur
        ref = th.Snapshot(0)
        ref.load(*[table[k] for k in COLS], n_uuids=u)
        ea, eb = cur.export(), ref.export()
        if not all(np.array_equal(ea[k], eb[k]) for k in ea):
            bad.append('export after delta %d' % step)
        full = dict(table, n_uuids=u)
        qu, qs, qe, qf = make_queries(full, rng, 800)
        got = np.asarray(cur.conflict_bitmap(qu, qs, qe, qf), bool)
        if not np.array_equal(got, np.asarray(brute.conflict_bits(full, qu, qs, qe, qf), bool)):
            bad.append('bitmap after delta %d' % step)
        ref.close()
    cur.close()
    spare.close()
    if bad and verbose:
        print('seed', seed, 'style', seed % 4, 'delta ->', bad)
    return len(bad)


def run(n_worlds=60, first=0, verbose=True, deltas=True):
    bad = sum(check_world(s, verbose) for s in range(first, first + n_worlds))
    if deltas:
        os.environ.pop('TH_GRID_PLAN', None)
        bad += sum(check_delta(s, verbose) for s in range(first, first + n_worlds))
    return bad


if __name__ == '__main__':
    k = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    f0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    n_bad = run(k, f0)
    print('worlds', k, 'from seed', f0, 'mismatching checks:', n_bad)
    sys.exit(1 if n_bad else 0)
