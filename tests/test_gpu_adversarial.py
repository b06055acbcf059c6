"""Every entry point against the SQL oracle (the reference's WHERE clauses through sqlite3) on small, deliberately
nasty calendars: timestamps drawn from a tiny set (ties and touching boundaries everywhere), nested and duplicate
intervals, zero-length rows, NULL/0/1 cancel flags, ids with gaps and in random order, uuids without rows."""
import datetime as dt

import numpy as np
import pytest

from helpers import csr_to_lists

pytestmark = pytest.mark.gpu


def _world(seed):
    from tensorhive_fixed_b200 import synth
    rng = np.random.default_rng(seed)
    n = int(rng.integers(0, 120))
    u = int(rng.integers(1, 6))
    grid = synth.T0_US + np.arange(12) * 1000
    ids = rng.permutation(np.arange(1, 4 * n + 1))[:n].astype(np.int64)
    uu = rng.integers(0, u, size=n).astype(np.int64)
    a = rng.choice(grid, size=n)
    b = rng.choice(grid, size=n)
    s, e = np.minimum(a, b), np.maximum(a, b)          # zero-length rows allowed
    raw = rng.choice([-1, 0, 1], size=n, p=[0.4, 0.3, 0.3])
    owner = rng.integers(1, 5, size=n).astype(np.int64)
    cal = {'id': ids, 'uuid_id': uu, 'start': s.astype(np.int64), 'end': e.astype(np.int64), 'owner': owner,
           'cancelled': (raw == 1).astype(np.uint8), 'cancelled_raw': raw, 'n_uuids': u}
    return cal, grid, rng


def _sql(cal):
    from oracle.sql_ref import SqlCalendar
    from tensorhive_fixed_b200 import synth
    rows = [(int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]), int(cal['end'][i]),
             int(cal['owner'][i]), None if cal['cancelled_raw'][i] < 0 else int(cal['cancelled_raw'][i]))
            for i in range(len(cal['id']))]
    return SqlCalendar(rows)


@pytest.mark.parametrize('seed', range(40))
def test_every_entry_point_against_the_sql_text(lib_built, seed):
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle.sql_ref import EPOCH
    cal, grid, rng = _world(seed)
    sql = _sql(cal)
    u = cal['n_uuids']
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    name = lambda k: synth.uuid_string(int(k))
    owner_of = {int(i): int(o) for i, o in zip(cal['id'], cal['owner'])}
    canc_of = {int(i): int(c) for i, c in zip(cal['id'], cal['cancelled'])}

    # candidates on the same tiny time grid (+-1 us), with and without a self id, some for unknown uuids
    m = 150
    qu = rng.integers(-1, u + 1, size=m).astype(np.int64)
    qa = rng.choice(grid, size=m) + rng.integers(-1, 2, size=m)
    qb = rng.choice(grid, size=m) + rng.integers(-1, 2, size=m)
    qs, qe = np.minimum(qa, qb).astype(np.int64), np.maximum(qa, qb).astype(np.int64)
    pool = np.concatenate([cal['id'], [10_000_000]]) if len(cal['id']) else np.array([10_000_000])
    qf = np.where(rng.random(m) < 0.5, -1, rng.choice(pool, size=m)).astype(np.int64)

    bits = snap.conflict_bitmap(qu, qs, qe, qf)
    off, ids, canc = snap.conflict_sets(qu, qs, qe, qf, th.MODE_STRICT)
    lists = csr_to_lists(off, ids)
    off2, ids2, canc2 = snap.conflict_sets(qu, qs, qe, None, th.MODE_INCLUSIVE)
    lists2 = csr_to_lists(off2, ids2)
    incl_bits = snap.conflict_bitmap(qu, qs, qe, None, th.MODE_INCLUSIVE)
    for i in range(m):
        known = 0 <= qu[i] < u
        rows = sql.would_interfere_rows(name(qu[i]), int(qs[i]), int(qe[i]), None if qf[i] < 0 else int(qf[i])) if known else []
        assert lists[i] == [r[0] for r in rows], ('strict list', i)
        assert bool(bits[i]) == any(not c for _, c in rows), ('verdict', i)
        if known:
            want = sorted(sql.filter_by_uuids_and_time_range([name(qu[i])], EPOCH + dt.timedelta(microseconds=int(qs[i])),
                                                             EPOCH + dt.timedelta(microseconds=int(qe[i]))))
        else:
            want = []
        assert lists2[i] == want, ('range list', i)
        assert bool(incl_bits[i]) == bool(want)
    assert [canc_of[int(x)] for x in ids] == [int(c) for c in canc]

    uu = np.arange(-1, u + 1, dtype=np.int64)
    for now in [int(grid[0]) - 5, int(grid[3]), int(grid[3]) + 1, int(grid[7]) - 1, int(grid[-1]), int(grid[-1]) + 7]:
        res, own = snap.current_owner(now)
        for k in range(u):
            cur = sql.current_events(now, name(k))
            assert (cur[0] if cur else -1) == res[k]
            assert (owner_of[cur[0]] if cur else -1) == own[k]
        for period in (0, 1, 1000, 2500, 100_000):
            fid, fst, fow, cnt = snap.upcoming(now, period, uu)
            offu, idu, _ = snap.conflict_sets(uu, np.full(len(uu), now), np.full(len(uu), now + period), None, th.MODE_UPCOMING)
            ulists = csr_to_lists(offu, idu)
            for j, k in enumerate(uu):
                want = sql.upcoming_events_for_resource(name(k), now, period) if 0 <= k < u else []
                assert ulists[j] == want, ('upcoming', now, period, k)
                assert cnt[j] == len(want)
                assert fid[j] == (want[0] if want else -1)
            for test, fn in ((th.OWNER_ANY, lambda o, w: len(w) > 0), (th.OWNER_OTHER, lambda o, w: any(owner_of[x] != o for x in w)),
                             (th.OWNER_SAME, lambda o, w: any(owner_of[x] == o for x in w))):
                to = rng.integers(1, 5, size=len(uu)).astype(np.int64)
                flags = snap.upcoming_owner_flags(now, period, uu, to, test)
                for j, k in enumerate(uu):
                    w = sql.upcoming_events_for_resource(name(k), now, period) if 0 <= k < u else []
                    assert bool(flags[j]) == fn(int(to[j]), w)
        for strict in (False, True):
            pu = rng.integers(-1, u + 1, size=60).astype(np.int64)
            po = rng.integers(-2, 5, size=60).astype(np.int64)
            fl, rid = snap.violation_join(now, strict, pu, po)
            for j in range(60):
                cur = sql.current_events(now, name(pu[j])) if 0 <= pu[j] < u else []
                if cur:
                    want = (po[j] != owner_of[cur[0]], cur[0] if po[j] != owner_of[cur[0]] else -1)
                else:
                    want = (strict, -1)
                assert (bool(fl[j]), int(rid[j])) == (bool(want[0]), int(want[1]))
    snap.close()


def test_long_lists_take_the_radix_path(lib_built):
    """One uuid with 30 000 rows and a whole-calendar range: the id ordering of the list goes through K1."""
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle import brute
    cal = synth.make_calendar(30_000, 2, seed=12)
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=2)
    qu = np.array([0, 1, 0, 5], np.int64)
    qs = np.array([cal['start'].min(), cal['start'].min(), int(np.median(cal['start'])), 0], np.int64)
    qe = np.array([cal['end'].max(), int(np.median(cal['start'])), int(np.median(cal['start'])) + synth.DAY_US, 10], np.int64)
    for mode, bmode in ((th.MODE_INCLUSIVE, brute.INCLUSIVE), (th.MODE_STRICT, brute.STRICT)):
        off, ids, canc = snap.conflict_sets(qu, qs, qe, np.full(4, -1, np.int64) if mode == th.MODE_STRICT else None, mode)
        got = csr_to_lists(off, ids)
        want = brute.conflict_sets(cal, qu, qs, qe, np.full(4, -1, np.int64), bmode)
        assert got == want
        assert len(got[0]) > 10_000 and got[0] == sorted(got[0])
        flag = {int(i): int(c) for i, c in zip(cal['id'], cal['cancelled'])}
        assert [flag[int(i)] for i in ids] == [int(c) for c in canc]
    snap.close()


@pytest.mark.parametrize('plan', ['shared', 'per_uuid'])
def test_calendar_spanning_the_whole_int64_range(lib_built, monkeypatch, plan):
    """Starts near both ends of int64: the span of the calendar (and of one uuid) exceeds 2^63 us, which the grid sizing
    must not read as a negative number (found with a synthetic 100 M-row calendar whose generator had overflowed)."""
    import tensorhive_fixed_b200 as th
    from oracle import brute
    monkeypatch.setenv('TH_GRID_PLAN', plan)
    I = np.iinfo(np.int64)
    rng = np.random.default_rng(3)
    lo = I.min + 1 + np.sort(rng.integers(0, 10_000, size=40))
    mid = np.sort(rng.integers(-5_000, 5_000, size=40))
    hi = I.max - 20_000 + np.sort(rng.integers(0, 10_000, size=40))
    start = np.concatenate([lo, mid, hi, mid + 7]).astype(np.int64)
    end = start + rng.integers(0, 3_000, size=start.size)
    start[118:120] = [I.max - 1, I.max]                # rows that start and end AT the upper extreme
    end[118:120] = [I.max, I.max]
    n = start.size
    cal = {'id': rng.permutation(np.arange(1, n + 1)).astype(np.int64),
           'uuid_id': np.concatenate([np.zeros(120, np.int64), np.ones(40, np.int64)]),        # uuid 0 spans everything
           'start': start, 'end': end.astype(np.int64), 'owner': rng.integers(1, 5, size=n).astype(np.int64),
           'cancelled': (rng.random(n) < 0.2).astype(np.uint8), 'n_uuids': 3}
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=3)
    m = 600
    qu = rng.integers(-1, 4, size=m).astype(np.int64)
    anchor = rng.choice(np.concatenate([start, end]), size=m)
    clip = lambda x: min(max(x, int(I.min)), int(I.max))                              # Python ints: no wrap-around
    qs = np.array([clip(int(a) + int(d)) for a, d in zip(anchor, rng.integers(-2_000, 2_000, size=m))], np.int64)
    qe = np.array([clip(int(a) + int(d)) for a, d in zip(qs, rng.integers(0, 4_000, size=m))], np.int64)
    qs[:4], qe[:4] = [I.min, I.min, I.max - 5, 0], [I.max, I.min + 5, I.max, I.max]
    qe[4:60] = I.max                                   # many candidates that end at the upper extreme
    qu[:12] = 0
    qf = np.where(rng.random(m) < 0.5, -1, rng.choice(cal['id'], size=m)).astype(np.int64)
    for mode, bmode in ((th.MODE_STRICT, brute.STRICT), (th.MODE_INCLUSIVE, brute.INCLUSIVE)):
        f = qf if mode == th.MODE_STRICT else None
        bits = snap.conflict_bitmap(qu, qs, qe, f, mode)
        want = brute.conflict_bits(cal, qu, qs, qe, qf if mode == th.MODE_STRICT else np.full(m, -1, np.int64), bmode)
        assert np.array_equal(np.asarray(bits, bool), np.asarray(want, bool)), mode
        off, ids, _c = snap.conflict_sets(qu, qs, qe, f, mode)
        assert csr_to_lists(off, ids) == brute.conflict_sets(cal, qu, qs, qe, qf if mode == th.MODE_STRICT else np.full(m, -1, np.int64), bmode)
    for now in (int(lo[3]), 0, int(hi[5]) + 1, int(I.min), int(I.min) + 1, int(I.max)):
        res, own = snap.current_owner(now)
        eres, eown = brute.current_owner(cal, 3, now)
        assert np.array_equal(res, eres) and np.array_equal(own, eown)
    snap.close()


def test_lookups_at_the_int64_extremes(lib_built, monkeypatch):
    """current_owner, upcoming (head, count, owner flags, row lists) and the violation join on calendars whose starts lie at
    both ends of int64, with `now` at the extremes and horizons `now + period` beyond INT64_MAX (= no upper bound; the sum
    used to wrap into the past), against a truth computed with Python integers: tools/fuzz_extremes.py."""
    import importlib.util
    import os
    monkeypatch.setenv('TH_GRID_PLAN', 'shared')           # run() sets the plan itself; restore the variable afterwards
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools', 'fuzz_extremes.py')
    spec = importlib.util.spec_from_file_location('fuzz_extremes', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.run() == 0


@pytest.mark.parametrize('plan', ['shared', 'per_uuid'])
def test_thousands_of_rows_on_a_handful_of_instants(lib_built, monkeypatch, plan):
    """20 000 rows whose starts take only five values (every grid bucket that is not empty holds thousands of rows with
    EQUAL starts: the long-bucket bisection, the top-2 self-exclusion inside ties, the radix ordering of long lists)."""
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle import brute
    monkeypatch.setenv('TH_GRID_PLAN', plan)
    rng = np.random.default_rng(77)
    n, u = 20_000, 3
    instants = synth.T0_US + np.array([0, 1, 1000, 1001, 500_000], np.int64)
    start = rng.choice(instants, size=n)
    end = start + rng.choice([0, 1, 2, 999, 1000, 1001, 600_000], size=n)
    cal = {'id': rng.permutation(np.arange(1, n + 1)).astype(np.int64), 'uuid_id': rng.integers(0, u, size=n).astype(np.int64),
           'start': start.astype(np.int64), 'end': end.astype(np.int64), 'owner': rng.integers(1, 5, size=n).astype(np.int64),
           'cancelled': (rng.random(n) < 0.5).astype(np.uint8), 'n_uuids': u}
    # uuid 2: every row but one is cancelled, and that one is what a self-update must exclude
    sel2 = np.flatnonzero(cal['uuid_id'] == 2)
    cal['cancelled'][sel2] = 1
    cal['cancelled'][sel2[7]] = 0
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    m = 400
    qu = rng.integers(0, u, size=m).astype(np.int64)
    qs = (rng.choice(instants, size=m) + rng.integers(-2, 3, size=m)).astype(np.int64)
    qe = (qs + rng.choice([0, 1, 2, 998, 1000, 1003, 700_000], size=m)).astype(np.int64)
    qf = np.where(rng.random(m) < 0.5, -1, rng.choice(cal['id'], size=m)).astype(np.int64)
    qf[qu == 2] = np.where(rng.random(int((qu == 2).sum())) < 0.5, cal['id'][sel2[7]], qf[qu == 2])
    for mode, bmode in ((th.MODE_STRICT, brute.STRICT), (th.MODE_INCLUSIVE, brute.INCLUSIVE)):
        f = qf if mode == th.MODE_STRICT else None
        fb = qf if mode == th.MODE_STRICT else np.full(m, -1, np.int64)
        bits = snap.conflict_bitmap(qu, qs, qe, f, mode)
        assert np.array_equal(np.asarray(bits, bool), np.asarray(brute.conflict_bits(cal, qu, qs, qe, fb, bmode), bool)), mode
    off, ids, canc = snap.conflict_sets(qu[:60], qs[:60], qe[:60], qf[:60], th.MODE_STRICT)
    got = csr_to_lists(off, ids)
    assert got == brute.conflict_sets(cal, qu[:60], qs[:60], qe[:60], qf[:60], brute.STRICT)
    assert max(len(x) for x in got) > 1000
    off, ids, canc = snap.conflict_sets(qu[:60], qs[:60], qe[:60], None, th.MODE_UPCOMING)
    assert csr_to_lists(off, ids) == brute.conflict_sets(cal, qu[:60], qs[:60], qe[:60], np.full(60, -1, np.int64), brute.UPCOMING)
    for now in [int(x) for x in instants] + [int(instants[2]) + 1]:
        res, own = snap.current_owner(now)
        eres, eown = brute.current_owner(cal, u, now)
        assert np.array_equal(res, eres) and np.array_equal(own, eown)
    snap.close()


def test_odd_table_shapes(lib_built, monkeypatch):
    """A million uuids and ten rows, no rows at all, one uuid holding every row, 65 536 uuids ...: bitmap, row lists (all
    three modes), owner table and upcoming against brute force under every grid plan: tools/fuzz_shapes.py."""
    import importlib.util
    import os
    monkeypatch.setenv('TH_GRID_PLAN', 'shared')           # run() sets / clears the plan itself; restored afterwards
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools', 'fuzz_shapes.py')
    spec = importlib.util.spec_from_file_location('fuzz_shapes', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.run() == 0


def test_random_worlds_of_four_styles(lib_built, monkeypatch):
    """tools/fuzz_worlds.py on 48 seeds: tiny time grids, bursts (long buckets), nested intervals over 13 orders of
    magnitude, synthetic walks; random grid plan; EVERY entry point (packed format included) against brute force."""
    import importlib.util
    import os
    monkeypatch.setenv('TH_GRID_PLAN', 'shared')           # check_world sets / clears the plan; restored afterwards
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools', 'fuzz_worlds.py')
    spec = importlib.util.spec_from_file_location('fuzz_worlds', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.run(48, 1000) == 0
