"""Parity of the CUDA path (through the C ABI) against the oracle -- bit-exact, integer work."""
import numpy as np
import pytest

from helpers import csr_to_lists

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def th(lib_built):
    import tensorhive_fixed_b200 as th
    return th


@pytest.fixture(scope='module')
def synth():
    from tensorhive_fixed_b200 import synth
    return synth


@pytest.fixture(scope='module')
def brute():
    from oracle import brute
    return brute


def _load(th, cal):
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'],
              n_uuids=cal['n_uuids'])
    return snap


# ---- K1 --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n', [1, 2, 31, 33, 4095, 4096, 4097, 50_000, 1_000_003])
def test_sort_pairs_matches_stable_lexsort(th, n):
    rng = np.random.default_rng(n)
    hi = rng.integers(0, 300, size=n, dtype=np.int64)
    lo = rng.integers(-5_000_000_000, 5_000_000_000, size=n, dtype=np.int64)
    perm = th.sort_pairs(hi, lo)
    want = np.lexsort((lo, hi))          # lexsort is stable
    assert np.array_equal(perm, want)


@pytest.mark.parametrize('pattern', ['all_equal', 'sorted', 'reversed', 'few_values', 'wide_hi', 'one_bit', 'full_range'])
def test_sort_pairs_patterns(th, pattern):
    n = 70_000
    rng = np.random.default_rng(11)
    hi = np.zeros(n, np.int64)
    if pattern == 'all_equal':
        lo = np.full(n, 42, np.int64)
    elif pattern == 'sorted':
        lo = np.arange(n, dtype=np.int64) * 3
    elif pattern == 'reversed':
        lo = np.arange(n, dtype=np.int64)[::-1].copy()
    elif pattern == 'few_values':
        lo = rng.integers(0, 3, size=n, dtype=np.int64)
        hi = rng.integers(0, 2, size=n, dtype=np.int64)
    elif pattern == 'wide_hi':
        lo = rng.integers(0, 1 << 50, size=n, dtype=np.int64)
        hi = rng.integers(0, (1 << 32) - 1, size=n, dtype=np.int64)
    elif pattern == 'full_range':       # the low key spans all of int64 (its extremes included), the high key is signed
        I = np.iinfo(np.int64)
        lo = rng.integers(I.min, I.max, size=n, dtype=np.int64, endpoint=True)
        lo[:4] = [I.min, I.max, I.min, I.max]
        hi = rng.integers(-3, 4, size=n, dtype=np.int64)
    else:
        lo = (rng.integers(0, 2, size=n, dtype=np.int64)) << 40
    perm = th.sort_pairs(hi, lo)
    assert np.array_equal(perm, np.lexsort((lo, hi)))


# ---- snapshot load ------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n,u,shuffle,skew', [(10_000, 64, True, 0.0), (10_000, 64, False, 0.0),
                                              (200_000, 500, True, 1.2), (5000, 1, True, 0.0), (7, 3, True, 0.0)])
def test_snapshot_is_sorted_by_uuid_start_id(th, synth, n, u, shuffle, skew):
    cal = synth.make_calendar(n, u, seed=n + u, skew=skew, shuffle_ids=shuffle)
    snap = _load(th, cal)
    ex = snap.export()
    order = np.lexsort((cal['id'], cal['start'], cal['uuid_id']))
    for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled'):
        assert np.array_equal(ex[k], cal[k][order]), k
    want_off = np.searchsorted(cal['uuid_id'][order], np.arange(u + 1), side='left')
    assert np.array_equal(ex['seg_offsets'], want_off)
    assert snap.info()['n'] == n
    snap.close()


def test_equal_start_rows_keep_id_order(th):
    n = 5000
    rng = np.random.default_rng(3)
    cal = {'id': rng.permutation(np.arange(1, n + 1)).astype(np.int64), 'uuid_id': rng.integers(0, 4, n).astype(np.int64),
           'start': rng.integers(0, 5, n).astype(np.int64) * 1000, 'owner': np.ones(n, np.int64),
           'cancelled': np.zeros(n, np.uint8), 'n_uuids': 4}
    cal['end'] = cal['start'] + 10
    snap = _load(th, cal)
    ex = snap.export()
    order = np.lexsort((cal['id'], cal['start'], cal['uuid_id']))
    assert np.array_equal(ex['id'], cal['id'][order])
    snap.close()


# ---- K2 bitmap -----------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n,u,m,skew', [(10_000, 64, 10_000, 0.0), (100_000, 512, 100_000, 0.0),
                                        (100_000, 16, 50_000, 1.5), (3000, 1, 3000, 0.0), (50, 64, 1000, 0.0)])
def test_conflict_bitmap_matches_brute_force(th, synth, brute, n, u, m, skew):
    cal = synth.make_calendar(n, u, seed=20260101 + n, skew=skew)
    q = synth.make_queries(cal, m, seed=20260102 + m)
    snap = _load(th, cal)
    got = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
    want = brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'])
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, 'first mismatches at %s' % bad[:10]
    got2 = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], None, th.MODE_INCLUSIVE)
    want2 = brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'], brute.INCLUSIVE)
    assert np.array_equal(got2, want2)
    snap.close()


def test_config1_against_the_sql_oracle(th, synth):
    """BASELINE config 1 (10k reservations, 64 uuids, 10k candidates) through the reference's SQL IN FULL:
    every candidate's verdict and row list against the WHERE clause of Reservation.py:121-129 run by sqlite3."""
    from oracle.sql_ref import SqlCalendar
    cal = synth.make_calendar(10_000, 64, seed=20260101)
    q = synth.make_queries(cal, 10_000, seed=20260102)
    rows = [(int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]), int(cal['end'][i]),
             int(cal['owner'][i]), None if cal['cancelled_raw'][i] < 0 else int(cal['cancelled_raw'][i]))
            for i in range(10_000)]
    sql = SqlCalendar(rows)
    snap = _load(th, cal)
    got = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
    off, ids, canc = snap.conflict_sets(q['uuid_id'], q['start'], q['end'], q['self_id'], th.MODE_STRICT)
    lists = csr_to_lists(off, ids)
    for i in range(10_000):
        sid = None if q['self_id'][i] < 0 else int(q['self_id'][i])
        r = sql.would_interfere_rows(synth.uuid_string(int(q['uuid_id'][i])), int(q['start'][i]), int(q['end'][i]), sid)
        assert bool(got[i]) == any(not c for _, c in r), i
        assert lists[i] == [x[0] for x in r], i
    snap.close()


def test_self_exclusion_uses_second_maximum(th, brute):
    # two live rows overlap the probe; excluding either one must still report the other
    cal = {'id': np.array([10, 11, 12], np.int64), 'uuid_id': np.zeros(3, np.int64),
           'start': np.array([0, 50, 500], np.int64), 'end': np.array([100, 200, 600], np.int64),
           'owner': np.ones(3, np.int64), 'cancelled': np.zeros(3, np.uint8), 'n_uuids': 1}
    snap = _load(th, cal)
    qs = np.array([60, 60, 60, 150, 150, 0], np.int64)
    qe = np.array([70, 70, 70, 160, 160, 700], np.int64)
    qf = np.array([-1, 10, 11, 11, 10, 12], np.int64)
    qu = np.zeros(6, np.int64)
    got = snap.conflict_bitmap(qu, qs, qe, qf)
    want = brute.conflict_bits(cal, qu, qs, qe, qf)
    assert list(got) == list(want) == [True, True, True, False, True, True]
    snap.close()


def test_empty_inputs_and_edge_sizes(th, synth, brute):
    empty = synth.make_calendar(0, 8, seed=1)
    snap = _load(th, empty)
    q = synth.make_queries(empty, 100, seed=2)
    assert not snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id']).any()
    off, ids, canc = snap.conflict_sets(q['uuid_id'], q['start'], q['end'], q['self_id'])
    assert off[-1] == 0 and len(ids) == 0
    res, own = snap.current_owner(synth.T0_US)
    assert (res == -1).all() and (own == -1).all()
    snap.close()
    cal = synth.make_calendar(1000, 8, seed=3)
    snap = _load(th, cal)
    z = np.zeros(0, np.int64)
    assert snap.conflict_bitmap(z, z, z, z).size == 0
    off, ids, canc = snap.conflict_sets(z, z, z, z)
    assert list(off) == [0]
    # all rows cancelled: nothing ever interferes, but the row lists still carry them
    cal2 = dict(cal, cancelled=np.ones(1000, np.uint8))
    snap2 = _load(th, cal2)
    q = synth.make_queries(cal2, 500, seed=4)
    assert not snap2.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id']).any()
    off, ids, canc = snap2.conflict_sets(q['uuid_id'], q['start'], q['end'], q['self_id'])
    assert csr_to_lists(off, ids) == brute.conflict_sets(cal2, q['uuid_id'], q['start'], q['end'], q['self_id'], brute.STRICT)
    assert (np.asarray(canc) == 1).all()
    snap.close(); snap2.close()


def test_bad_arguments_are_reported_not_thrown_across_the_abi(th, synth):
    cal = synth.make_calendar(100, 4, seed=5)
    snap = th.Snapshot(0)
    with pytest.raises(th.NativeError) as e:
        snap.conflict_bitmap(np.zeros(1, np.int64), np.zeros(1, np.int64), np.zeros(1, np.int64), np.zeros(1, np.int64))
    assert 'not loaded' in str(e.value)
    bad = dict(cal, uuid_id=cal['uuid_id'] + 100)
    with pytest.raises(th.NativeError) as e:
        _load(th, bad)
    assert e.value.code == -1
    snap.close()


# ---- K2 row lists ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('m', [5000, 40_000, 270_000])       # one warp / four lanes (>= 32 768) / one lane (>= 262 144) per query
@pytest.mark.parametrize('mode_name', ['STRICT', 'INCLUSIVE', 'UPCOMING'])
def test_conflict_sets_match_brute_force(th, synth, brute, mode_name, m):
    cal = synth.make_calendar(30_000, 40, seed=77, skew=0.8)
    q = synth.make_queries(cal, m, seed=78)
    mode = getattr(th, 'MODE_' + mode_name)
    bmode = getattr(brute, mode_name)
    qs, qe = q['start'].copy(), q['end'].copy()
    if mode_name != 'STRICT':
        qe[::3] = qs[::3] + 40 * synth.DAY_US      # long ranges -> long lists (calendar views)
    qe[::50] = qs[::50] + 2000 * synth.DAY_US      # whole-calendar ranges: lists far longer than a warp
    snap = _load(th, cal)
    off, ids, canc = snap.conflict_sets(q['uuid_id'], qs, qe, q['self_id'] if mode_name == 'STRICT' else None, mode)
    want = brute.conflict_sets(cal, q['uuid_id'], qs, qe, q['self_id'], bmode)
    got = csr_to_lists(off, ids)
    assert got == want
    flag = {int(i): int(c) for i, c in zip(cal['id'], cal['cancelled'])}
    assert [flag[int(i)] for i in ids] == [int(c) for c in canc]
    assert max(len(x) for x in want) > 32      # lists longer than a warp are covered
    snap.close()


def test_conflict_sets_capacity_protocol(th, synth):
    import ctypes as C
    from tensorhive_fixed_b200 import native
    cal = synth.make_calendar(2000, 4, seed=9)
    q = synth.make_queries(cal, 300, seed=10)
    snap = _load(th, cal)
    off, ids, canc = snap.conflict_sets(q['uuid_id'], q['start'], q['end'] + 30 * synth.DAY_US, None, th.MODE_INCLUSIVE,
                                        capacity=1)
    assert off[-1] == len(ids) > 1
    lib = native.load()
    needed = C.c_int64(0)
    o = np.zeros(301, np.int64)
    far = np.ascontiguousarray(q['end'] + 30 * synth.DAY_US)
    rc = lib.th_conflict_sets(snap._h, 300, q['uuid_id'].ctypes.data, q['start'].ctypes.data,
                              far.ctypes.data, None, 1, o.ctypes.data, None, None, 0,
                              C.byref(needed), 0, None)
    assert rc == -5 and needed.value == len(ids) and o[-1] == len(ids)
    snap.close()


# ---- K3 / K5 -----------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n,u,skew', [(20_000, 64, 0.0), (60_000, 8, 1.0), (300, 200, 0.0)])
def test_current_owner_and_upcoming(th, synth, brute, n, u, skew):
    cal = synth.make_calendar(n, u, seed=n, skew=skew)
    snap = _load(th, cal)
    probes = [int(np.median(cal['start'])), int(cal['start'][3]), int(cal['end'][4]), int(cal['start'].min()) - 5,
              int(cal['end'].max()) + 5]
    uu = np.concatenate([np.arange(u), [-1, u, u + 9]]).astype(np.int64)
    for now in probes:
        res, own = snap.current_owner(now)
        wres, wown = brute.current_owner(cal, u, now)
        assert np.array_equal(res, wres) and np.array_equal(own, wown)
        for period in (0, 30 * synth.MIN_US, 5 * synth.DAY_US):
            fid, fst, fow, cnt = snap.upcoming(now, period, uu)
            wid, wst, wow, wcnt = brute.upcoming_summary(cal, uu, now, period)
            assert np.array_equal(cnt, wcnt)
            assert np.array_equal(fid, wid) and np.array_equal(fst, wst) and np.array_equal(fow, wow)
    snap.close()


def test_owner_flags_and_violation_join(th, synth, brute):
    cal = synth.make_calendar(40_000, 256, seed=41)
    now = int(np.median(cal['start']))
    period = 30 * synth.MIN_US
    snap = _load(th, cal)
    tasks = synth.make_job_tasks(cal, 3000, seed=42)
    sets = brute.conflict_sets(cal, tasks['uuid_id'], np.full(len(tasks['uuid_id']), now),
                               np.full(len(tasks['uuid_id']), now + period), np.full(len(tasks['uuid_id']), -1),
                               brute.UPCOMING)
    owner_of = {int(i): int(o) for i, o in zip(cal['id'], cal['owner'])}
    # make a third of the tasks owned by the reservation's owner so that SAME / OTHER both occur
    t_owner = tasks['owner'].copy()
    for k, s in enumerate(sets):
        if s and k % 3 == 0:
            t_owner[k] = owner_of[s[0]]
    for test, fn in ((th.OWNER_ANY, lambda o, s: len(s) > 0),
                     (th.OWNER_OTHER, lambda o, s: any(owner_of[i] != o for i in s)),
                     (th.OWNER_SAME, lambda o, s: any(owner_of[i] == o for i in s))):
        got = snap.upcoming_owner_flags(now, period, tasks['uuid_id'], t_owner, test)
        want = np.array([fn(int(t_owner[k]), sets[k]) for k in range(len(sets))])
        assert np.array_equal(got.astype(bool), want), test
    procs = synth.make_processes(cal, 20_000, now, seed=43)
    res, own = brute.current_owner(cal, cal['n_uuids'], now)
    for strict in (False, True):
        flags, rid = snap.violation_join(now, strict, procs['uuid_id'], procs['owner'])
        r = res[procs['uuid_id']]
        o = own[procs['uuid_id']]
        wflag = np.where(r >= 0, (o >= 0) & (procs['owner'] != o), strict)
        wrid = np.where((r >= 0) & wflag, r, -1)
        assert np.array_equal(flags.astype(bool), wflag)
        assert np.array_equal(rid, wrid)
        assert 0 < wflag.mean() < 1
    snap.close()


# ---- K4 ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('kw', [dict(n_uuids=16, samples_per_uuid=4000), dict(n_uuids=7, samples_per_uuid=1001, jitter_us=900_000),
                                dict(n_uuids=40, samples_per_uuid=300, ragged=True, cadence_us=60_000_000, jitter_us=5_000_000),
                                dict(n_uuids=3, samples_per_uuid=1), dict(n_uuids=2, samples_per_uuid=100_000)])
def test_usage_reduce_matches_oracle(th, synth, brute, kw):
    s = synth.make_samples(seed=5, **kw)
    cad = kw.get('cadence_us', 1_000_000)
    w = synth.make_windows(s, 3000, seed=6, min_len_us=20 * cad, max_len_us=180 * cad)
    sg, sm, cnt, ag, am = th.usage_reduce(s['offsets'], s['ts'], s['gpu'], s['mem'], w['uuid_id'], w['start'], w['end'])
    esg, esm, ecnt = brute.usage_reduce(s['offsets'], s['ts'], s['gpu'], s['mem'], w['uuid_id'], w['start'], w['end'])
    assert np.array_equal(cnt, ecnt)
    assert np.array_equal(sg, esg) and np.array_equal(sm, esm)
    eag = np.array([brute.floor_avg(a, c) for a, c in zip(esg, ecnt)])
    eam = np.array([brute.floor_avg(a, c) for a, c in zip(esm, ecnt)])
    # north_star tolerance for usage averages is 1e-6 relative; integer arithmetic gives exact equality
    assert np.array_equal(ag, eag) and np.array_equal(am, eam)
    assert (ecnt == 0).any() and (ecnt > 0).any()


def test_usage_reduce_touching_windows_share_the_boundary_sample(th):
    ts = np.arange(10, dtype=np.int64) * 1_000_000
    gpu = np.arange(10, dtype=np.int64) * 10
    mem = np.ones(10, np.int64)
    off = np.array([0, 10], np.int64)
    wu = np.zeros(3, np.int64)
    ws = np.array([0, 4_000_000, 9_000_001], np.int64)
    we = np.array([4_000_000, 9_000_000, 9_500_000], np.int64)
    sg, sm, cnt, ag, am = th.usage_reduce(off, ts, gpu, mem, wu, ws, we)
    assert list(cnt) == [5, 6, 0]           # sample at t = 4 s belongs to both
    assert list(sg) == [100, 390, 0]
    assert list(ag) == [20, 65, -1] and list(am) == [1, 1, -1]


@pytest.mark.parametrize('n_uuids,n_windows', [(9, 4000), (6000, 200)])     # with and without the per-uuid entry table
@pytest.mark.parametrize('narrow', [False, True])
def test_usage_reduce_irregular_logs(th, brute, n_uuids, n_windows, narrow):
    """Logs the interpolation guess cannot hit: exponential and bursty spacing, repeated timestamps, segments of
    0 / 1 / 2 / 15 / 16 / 17 samples, an odd total length, windows far outside the log and at the int64 extremes."""
    rng = np.random.default_rng(100 + n_uuids)
    sizes = [0, 1, 2, 15, 16, 17, 333, 4096, 2501][:n_uuids] if n_uuids <= 9 else list(rng.integers(0, 40, size=n_uuids))
    ts_parts = []
    for k, n in enumerate(sizes):
        kind = k % 4
        if kind == 0:
            t = np.cumsum(rng.integers(0, 3, size=n))                      # many repeated timestamps
        elif kind == 1:
            t = np.floor(np.exp(np.linspace(0, 30, n))).astype(np.int64) if n else np.zeros(0, np.int64)   # exponential
        elif kind == 2:
            t = np.cumsum(np.where(rng.random(n) < 0.02, 10**9, 1))        # bursts separated by huge gaps
        else:
            t = np.cumsum(rng.integers(1, 2_000_000, size=n))
        ts_parts.append(np.sort(t.astype(np.int64)) + 1_700_000_000_000_000)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    ts = np.concatenate(ts_parts) if ts_parts else np.zeros(0, np.int64)
    s = int(off[-1])
    assert n_uuids > 9 or s % 2 == 1
    gpu = rng.integers(0, 101, size=s, dtype=np.int64)
    mem = rng.integers(0, 101, size=s, dtype=np.int64)
    wu = rng.integers(-1, n_uuids + 1, size=n_windows, dtype=np.int64)     # includes unknown uuids on both sides
    ws = np.empty(n_windows, np.int64)
    we = np.empty(n_windows, np.int64)
    for j in range(n_windows):
        u = int(wu[j])
        a, b = (int(off[u]), int(off[u + 1])) if 0 <= u < n_uuids else (0, 0)
        c = rng.random()
        if b > a and c < 0.6:                                   # ends exactly on samples
            i, k = sorted(rng.integers(a, b, size=2))
            ws[j], we[j] = ts[i], ts[k]
        elif b > a and c < 0.8:                                 # just beside samples
            i, k = sorted(rng.integers(a, b, size=2))
            ws[j], we[j] = ts[i] + rng.integers(-1, 2), ts[k] + rng.integers(-1, 2)
        elif c < 0.85:
            ws[j], we[j] = np.iinfo(np.int64).min, np.iinfo(np.int64).max
        elif c < 0.9:
            ws[j], we[j] = np.iinfo(np.int64).max - 5, np.iinfo(np.int64).max
        elif c < 0.95:
            ws[j], we[j] = np.iinfo(np.int64).min, np.iinfo(np.int64).min + 5
        else:
            ws[j], we[j] = 1_700_000_000_000_500, 1_700_000_000_000_100       # end before start
    if narrow:
        outs = th.usage_reduce_u8(off, ts, gpu.astype(np.uint8), mem.astype(np.uint8), wu, ws, we)
    else:
        outs = th.usage_reduce(off, ts, gpu, mem, wu, ws, we)
    esg, esm, ecnt = brute.usage_reduce(off, ts, gpu, mem, wu, ws, we)
    assert np.array_equal(outs[2], ecnt)
    assert np.array_equal(outs[0], esg) and np.array_equal(outs[1], esm)
    assert np.array_equal(outs[3], np.array([brute.floor_avg(a, c) for a, c in zip(esg, ecnt)]))
    assert (ecnt == 0).any() and (ecnt > 16).any()


def test_usage_reduce_unaligned_columns_and_negative_values(th, brute):
    rng = np.random.default_rng(8)
    n = 5001
    base_g = rng.integers(-50, 100, size=n + 1, dtype=np.int64)
    base_m = rng.integers(0, 100, size=n + 1, dtype=np.int64)
    ts = np.sort(rng.integers(0, 10_000, size=n, dtype=np.int64))
    off = np.array([0, 2000, 2000, n], np.int64)
    ts[:2000].sort(); ts[2000:].sort()
    wu = rng.integers(0, 3, size=400, dtype=np.int64)
    ws = rng.integers(-100, 10_000, size=400, dtype=np.int64)
    we = ws + rng.integers(-5, 3000, size=400, dtype=np.int64)
    import torch
    dev = torch.device('cuda:0')
    g = torch.from_numpy(base_g).to(dev)[1:]        # 8-byte aligned but not 16-byte aligned view
    mm = torch.from_numpy(base_m).to(dev)[1:]
    outs = th.usage_reduce(torch.from_numpy(off).to(dev), torch.from_numpy(ts).to(dev), g, mm,
                           torch.from_numpy(wu).to(dev), torch.from_numpy(ws).to(dev), torch.from_numpy(we).to(dev))
    sg, sm, cnt, ag, am = [o.cpu().numpy() for o in outs]
    esg, esm, ecnt = brute.usage_reduce(off, ts, base_g[1:], base_m[1:], wu, ws, we)
    assert np.array_equal(sg, esg) and np.array_equal(sm, esm) and np.array_equal(cnt, ecnt)
    assert np.array_equal(ag, np.array([brute.floor_avg(a, c) for a, c in zip(esg, ecnt)]))   # floor, not truncation


# ---- device-pointer path ----------------------------------------------------------------------------------------------------
def test_device_pointer_path_equals_host_path(th, synth):
    import torch
    dev = torch.device('cuda:0')
    cal = synth.make_calendar(50_000, 128, seed=91)
    q = synth.make_queries(cal, 33_333, seed=92)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    snap_h = _load(th, cal)
    snap_d = th.Snapshot(0)
    snap_d.load(t(cal['id']), t(cal['uuid_id']), t(cal['start']), t(cal['end']), t(cal['owner']), t(cal['cancelled']),
                n_uuids=128)
    a = snap_h.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
    words = snap_d.conflict_words(t(q['uuid_id']), t(q['start']), t(q['end']), t(q['self_id']))
    torch.cuda.synchronize()
    from tensorhive_fixed_b200.api import unpack_bits
    b = unpack_bits(words, len(a))
    assert np.array_equal(a, b)
    oh = snap_h.conflict_sets(q['uuid_id'], q['start'], q['end'], q['self_id'])
    od = snap_d.conflict_sets(t(q['uuid_id']), t(q['start']), t(q['end']), t(q['self_id']))
    for x, y in zip(oh, od):
        assert np.array_equal(np.asarray(x), y.cpu().numpy())
    snap_h.close(); snap_d.close()


# ---- BASELINE config 2 at full size ---------------------------------------------------------------------------------------------
def test_pinned_host_columns_are_read_in_place(th, synth, brute, monkeypatch):
    """TH_MEM_HOST with page-locked arrays: the kernel streams the columns (and writes the bitmap) over PCIe itself.
    Same bits as pageable arrays (staged copies), as the staged path forced by TH_NO_ZERO_COPY, and as the oracle;
    a mix of pinned and pageable columns falls back to staging."""
    import torch
    cal = synth.make_calendar(50_000, 100, seed=71)
    q = synth.make_queries(cal, 70_001, seed=72)          # ragged last tile
    snap = _load(th, cal)
    cols = [q[k] for k in ('uuid_id', 'start', 'end', 'self_id')]
    want = brute.conflict_bits(cal, *cols)
    pinned = [torch.from_numpy(c.copy()).pin_memory() for c in cols]
    pin_np = [p.numpy() for p in pinned]
    lib = th.native.load()
    words = (len(want) + 31) // 32

    def run(arrs, out):
        m = len(arrs[0])
        th.native.check(lib.th_conflict_bitmap(snap._h, m, arrs[0].ctypes.data, arrs[1].ctypes.data,
                                               arrs[2].ctypes.data, arrs[3].ctypes.data, 0, out.ctypes.data, 0, None))
        return th.api.unpack_bits(out, m)

    out_pin = torch.zeros(words, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    assert np.array_equal(run(pin_np, out_pin), want)                                  # all pinned, pinned output
    assert np.array_equal(run(pin_np, np.zeros(words, np.uint32)), want)               # pinned in, pageable out
    assert np.array_equal(run(cols, np.zeros(words, np.uint32)), want)                 # pageable: staged
    assert np.array_equal(run([pin_np[0], cols[1], pin_np[2], cols[3]], np.zeros(words, np.uint32)), want)   # mixed
    off1 = [p[1:] for p in pin_np]                                                     # 8-byte aligned only
    assert np.array_equal(run(off1, out_pin), want[1:])
    monkeypatch.setenv('TH_NO_ZERO_COPY', '1')
    assert np.array_equal(run(pin_np, out_pin), want)
    snap.close()


def test_config2_full_size_bit_exact(th, synth, brute):
    """1M reservations x 4096 uuids, 1M candidates: the whole bitmap against the brute-force oracle."""
    cal = synth.make_calendar(1_000_000, 4096, seed=20260102)
    q = synth.make_queries(cal, 1_000_000, seed=20260103)
    snap = _load(th, cal)
    got = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
    want = brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'])
    bad = np.nonzero(got != want)[0]
    if bad.size:
        again = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
        raise AssertionError('%d mismatches, first at %s (tile %s, lane %s); second run differs from first at %d places'
                             % (bad.size, bad[:8], bad[:8] // 32, bad[:8] % 32, int((again != got).sum())))
    for _ in range(5):      # the kernel is a pure function of its inputs
        assert np.array_equal(snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id']), want)
    # size-independent properties: a row always interferes with a copy of itself unless it is the row itself
    # or cancelled; shrinking a non-conflicting candidate never creates a conflict
    live = np.nonzero(cal['cancelled'] == 0)[0][:200_000]
    same = snap.conflict_bitmap(cal['uuid_id'][live], cal['start'][live], cal['end'][live], np.full(live.size, -1, np.int64))
    assert same.all()
    free = np.nonzero(~got)[0]
    shr = snap.conflict_bitmap(q['uuid_id'][free], q['start'][free] + 1, q['end'][free] - 1, q['self_id'][free])
    assert not shr.any()
    snap.close()


# ---- fused K2 + exchange, two ranks emulated on one GPU --------------------------------------------------------------
def test_fused_exchange_two_ranks_emulated_on_one_gpu(th, synth, brute):
    """Both ranks live in this process on cuda:0; the kernels are enqueued first (wait = 0) and the wait halves
    afterwards, on ONE stream, so no kernel ever waits for one that has not been launched."""
    import torch
    from tensorhive_fixed_b200 import sharding
    from tensorhive_fixed_b200.api import unpack_bits
    dev = torch.device('cuda:0')
    cal = synth.make_calendar(120_000, 300, seed=55, skew=0.5)
    bounds = sharding.partition_uuids(np.bincount(cal['uuid_id'], minlength=300), 2)
    shards = [sharding.shard_rows(cal, bounds, r) for r in range(2)]
    snaps = []
    for s in shards:
        snap = th.Snapshot(0)
        snap.load(s['id'], s['uuid_id'], s['start'], s['end'], s['owner'], s['cancelled'], n_uuids=s['n_uuids'])
        snaps.append(snap)
    px = [sharding.PeerExchange(0, 2, r, max_queries_per_rank=50_000, connect=False) for r in range(2)]
    bases = [p.local_base() for p in px]
    for p in px:
        p.connect_local(bases)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    for step in range(3):            # three steps: both buffer generations are reused
        q = synth.make_queries(cal, 60_000 + 1000 * step, seed=900 + step)
        router = sharding.Router(q['uuid_id'], bounds)
        views, counts = [], []
        for r in range(2):
            idx = router.local_indices(r)
            counts.append(idx.size)
            views.append(px[r].conflict_words(snaps[r], t(router.local_uuid(q['uuid_id'], r)), t(q['start'][idx]),
                                              t(q['end'][idx]), t(q['self_id'][idx]), wait=False,
                                              counts=router.counts))
        for r in range(2):
            px[r].wait()
        torch.cuda.synchronize()
        for r in range(2):
            px[r].check()
        want = brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'])
        for r in range(2):           # every rank ends up with everybody's words
            got = np.zeros(len(want), bool)
            g = views[r].cpu().numpy().view(np.uint32)
            for src in range(2):
                got[router.local_indices(src)] = unpack_bits(g[src], counts[src])
            assert np.array_equal(got, want), (step, r)
    for p in px:
        p.close()
    for s in snaps:
        s.close()


def test_fused_exchange_with_more_tiles_per_cta_than_the_word_buffer(th, synth):
    """40 M queries in one step: every CTA owns more bitmap words than it buffers in shared memory, so the words leave
    as one slot store per tile.  Checked against the plain kernel on the same (device-resident) batch."""
    import torch
    from tensorhive_fixed_b200 import sharding
    dev = torch.device('cuda:0')
    cal = synth.make_calendar(200_000, 128, seed=61)
    snap = th.Snapshot(0)
    snap.load(*[cal[k] for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')], n_uuids=128)
    q = synth.make_queries(cal, 1_000_000, seed=62)
    reps = 40
    qd = [torch.from_numpy(np.ascontiguousarray(q[k])).to(dev).repeat(reps) for k in ('uuid_id', 'start', 'end', 'self_id')]
    m = int(qd[0].numel())
    px = sharding.PeerExchange(0, 1, 0, max_queries_per_rank=m, connect=False)
    px.connect_local([px.local_base()])
    for step in range(2):
        got = px.conflict_words(snap, *qd, wait=False)
        px.wait()
        torch.cuda.synchronize()
        px.check()
        want = snap.conflict_words(*qd)
        assert torch.equal(got.reshape(-1)[:want.numel()].cpu(), want.cpu()), step
    one = th.api.unpack_bits(want[:(1_000_000 + 31) // 32], 1_000_000)
    assert one.any() and not one.all()
    px.close()
    snap.close()


def test_reloading_one_handle_reuses_or_regrows_its_buffers(th, synth, brute):
    snap = th.Snapshot(0)
    for n, u, seed in [(50_000, 64, 1), (3_000, 8, 2), (0, 4, 3), (80_000, 300, 4), (80_000, 300, 5)]:
        cal = synth.make_calendar(n, u, seed=seed)
        snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
        q = synth.make_queries(cal, 5000, seed=seed + 10)
        got = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
        assert np.array_equal(got, brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'])), (n, u)
        assert snap.info()['n'] == n
    snap.close()


# ---- K4, narrow log format (uint8 utilisation) ---------------------------------------------------------------------------
@pytest.mark.parametrize('kw', [dict(n_uuids=16, samples_per_uuid=4000), dict(n_uuids=7, samples_per_uuid=1001, jitter_us=900_000),
                                dict(n_uuids=40, samples_per_uuid=300, ragged=True, cadence_us=60_000_000, jitter_us=5_000_000),
                                dict(n_uuids=3, samples_per_uuid=1), dict(n_uuids=2, samples_per_uuid=100_003)])
def test_usage_reduce_u8_matches_oracle(th, synth, brute, kw):
    s = synth.make_samples(seed=15, **kw)
    cad = kw.get('cadence_us', 1_000_000)
    w = synth.make_windows(s, 3000, seed=16, min_len_us=5 * cad, max_len_us=400 * cad)
    g8, m8 = s['gpu'].astype(np.uint8), (s['mem'] * 2 + 50).astype(np.uint8)        # values up to 250
    outs = th.usage_reduce_u8(s['offsets'], s['ts'], g8, m8, w['uuid_id'], w['start'], w['end'])
    esg, esm, ecnt = brute.usage_reduce(s['offsets'], s['ts'], g8.astype(np.int64), m8.astype(np.int64), w['uuid_id'],
                                        w['start'], w['end'])
    assert np.array_equal(outs[2], ecnt) and np.array_equal(outs[0], esg) and np.array_equal(outs[1], esm)
    assert np.array_equal(outs[3], np.array([brute.floor_avg(a, c) for a, c in zip(esg, ecnt)]))
    assert np.array_equal(outs[4], np.array([brute.floor_avg(a, c) for a, c in zip(esm, ecnt)]))
    # and it agrees with the wide-format kernel on the same data
    wide = th.usage_reduce(s['offsets'], s['ts'], g8.astype(np.int64), m8.astype(np.int64), w['uuid_id'], w['start'], w['end'])
    for a, b in zip(outs, wide):
        assert np.array_equal(a, b)


@pytest.mark.parametrize('n', [1, 15, 16, 17, 47, 48, 83])
def test_usage_reduce_u8_every_range_of_a_short_log(th, brute, n):
    """Every index range [a, b] of an n-sample log as a window: first and last 16-sample chunk the same, adjacent, or apart;
    ranges that start / end on every byte of a chunk; the column ending inside its last chunk (n not a multiple of 16)."""
    rng = np.random.default_rng(n)
    ts = (np.arange(n, dtype=np.int64) * 1_000_000 + 1_700_000_000_000_000)
    off = np.array([0, 0, n], np.int64)                               # uuid 0 has no samples
    g8 = rng.integers(0, 256, size=n, dtype=np.uint8)
    m8 = rng.integers(0, 256, size=n, dtype=np.uint8)
    ab = np.array([(a, b) for a in range(n) for b in range(a, n)], np.int64)
    wu = np.ones(len(ab), np.int64)
    ws, we = ts[ab[:, 0]], ts[ab[:, 1]] + 999_999                     # ends between two samples
    outs = th.usage_reduce_u8(off, ts, g8, m8, wu, ws, we)
    cs_g = np.concatenate([[0], np.cumsum(g8.astype(np.int64))])
    cs_m = np.concatenate([[0], np.cumsum(m8.astype(np.int64))])
    assert np.array_equal(outs[2], ab[:, 1] - ab[:, 0] + 1)
    assert np.array_equal(outs[0], cs_g[ab[:, 1] + 1] - cs_g[ab[:, 0]])
    assert np.array_equal(outs[1], cs_m[ab[:, 1] + 1] - cs_m[ab[:, 0]])
    esg, esm, ecnt = brute.usage_reduce(off, ts, g8.astype(np.int64), m8.astype(np.int64), wu, ws, we)
    assert np.array_equal(outs[0], esg) and np.array_equal(outs[1], esm) and np.array_equal(outs[2], ecnt)


def test_usage_reduce_u8_unaligned_columns(th, brute):
    import torch
    rng = np.random.default_rng(18)
    n = 70_001
    base_g = rng.integers(0, 256, size=n + 3, dtype=np.uint8)
    base_m = rng.integers(0, 101, size=n + 3, dtype=np.uint8)
    ts = np.sort(rng.integers(0, 1_000_000, size=n, dtype=np.int64))
    off = np.array([0, 30_000, 30_000, n], np.int64)
    ts[:30_000].sort(); ts[30_000:].sort()
    wu = rng.integers(0, 3, size=2000, dtype=np.int64)
    ws = rng.integers(-1000, 1_000_000, size=2000, dtype=np.int64)
    we = ws + rng.integers(-5, 60_000, size=2000, dtype=np.int64)
    dev = torch.device('cuda:0')
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    for shift in (0, 3):       # shift 3: the byte columns start 3 bytes off a 16-byte boundary
        g = t(base_g)[shift:shift + n]
        mm = t(base_m)[shift:shift + n]
        outs = th.usage_reduce_u8(t(off), t(ts), g, mm, t(wu), t(ws), t(we))
        esg, esm, ecnt = brute.usage_reduce(off, ts, base_g[shift:shift + n].astype(np.int64),
                                            base_m[shift:shift + n].astype(np.int64), wu, ws, we)
        assert np.array_equal(outs[0].cpu().numpy(), esg) and np.array_equal(outs[1].cpu().numpy(), esm)
        assert np.array_equal(outs[2].cpu().numpy(), ecnt)


@pytest.mark.parametrize('m', [1, 31, 32, 33, 4097, 100_003])
def test_packed_queries_give_identical_bitmaps(th, synth, brute, m):
    """th_conflict_bitmap_packed (uint32 uuid / self columns, 20-24 B per query) against th_conflict_bitmap and the oracle:
    host arrays (pageable and pinned, read in place), device arrays, misaligned columns, no self column, both modes."""
    import torch
    cal = synth.make_calendar(30_000, 96, seed=4242)
    q = synth.make_queries(cal, m, seed=4243)
    snap = _load(th, cal)
    want = snap.conflict_words(q['uuid_id'], q['start'], q['end'], q['self_id'])
    assert np.array_equal(th.api.unpack_bits(want, m), brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id']))
    pu, pf = th.Snapshot.pack_queries(q['uuid_id'], q['self_id'])
    assert np.array_equal(snap.conflict_words_packed(pu, q['start'], q['end'], pf), want)
    # pinned host columns: the kernel reads them over PCIe itself
    pin = [torch.from_numpy(x.copy()).pin_memory() for x in (pu.view(np.int32), q['start'], q['end'], pf.view(np.int32))]
    got = snap.conflict_words_packed(pin[0].numpy().view(np.uint32), pin[1].numpy(), pin[2].numpy(), pin[3].numpy().view(np.uint32))
    assert np.array_equal(got, want)
    # device columns, and the same shifted by one element (4-byte aligned only: the direct-load path)
    dev = [torch.from_numpy(np.concatenate([x[:1], x])).cuda() for x in (pu.view(np.int32), q['start'], q['end'], pf.view(np.int32))]
    got = snap.conflict_words_packed(*[t[1:] for t in dev])
    assert np.array_equal(got.cpu().numpy().view(np.uint32), want)
    # no self column at all = every candidate is new
    no_self = snap.conflict_words(q['uuid_id'], q['start'], q['end'], np.full(m, -1, np.int64))
    assert np.array_equal(snap.conflict_words_packed(pu, q['start'], q['end'], None), no_self)
    incl = snap.conflict_words(q['uuid_id'], q['start'], q['end'], None, th.MODE_INCLUSIVE)
    assert np.array_equal(snap.conflict_words_packed(pu, q['start'], q['end'], None, th.MODE_INCLUSIVE), incl)
    snap.close()


@pytest.mark.parametrize('plan', ['shared', 'per_uuid'])
@pytest.mark.parametrize('skew', [0.0, 1.5])
def test_both_grid_plans_give_the_same_answers(th, synth, brute, monkeypatch, plan, skew):
    """The time grid has two plans -- one bucket width for all uuids (balanced calendars) or one per uuid (hot GPUs, uuids
    with very different spans); the load picks by itself, TH_GRID_PLAN forces one.  Every search goes through it."""
    monkeypatch.setenv('TH_GRID_PLAN', plan)
    cal = synth.make_calendar(60_000, 48, seed=77, skew=skew)
    cal['start'][:50] -= 400 * synth.DAY_US            # a few outliers stretch one uuid's span far into the past
    cal['end'][:50] -= 400 * synth.DAY_US
    q = synth.make_queries(cal, 40_000, seed=78)
    snap = _load(th, cal)
    assert np.array_equal(snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id']),
                          brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id']))
    assert np.array_equal(snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], None, th.MODE_INCLUSIVE),
                          brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'], brute.INCLUSIVE))
    some = slice(0, 3000)
    off, ids, _c = snap.conflict_sets(q['uuid_id'][some], q['start'][some], q['end'][some], q['self_id'][some])
    assert csr_to_lists(off, ids) == brute.conflict_sets(cal, q['uuid_id'][some], q['start'][some], q['end'][some],
                                                         q['self_id'][some], brute.STRICT)
    now = int(np.median(cal['start']))
    res, own = snap.current_owner(now)
    want_res, want_own = brute.current_owner(cal, 48, now)
    assert np.array_equal(res, want_res) and np.array_equal(own, want_own)
    got = snap.upcoming(now, 30 * synth.MIN_US, np.arange(48, dtype=np.int64))
    want = brute.upcoming_summary(cal, np.arange(48, dtype=np.int64), now, 30 * synth.MIN_US)
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    snap.close()


def test_usage_reduce_with_stamps_at_both_ends_of_int64(th, brute):
    """Sample logs whose timestamps run from INT64_MIN + 1 to INT64_MAX (the interpolation guess works in doubles, the
    answer must still be exact), windows anywhere including the extremes: int64 columns, uint8 columns and the ring."""
    import ctypes as C
    from tensorhive_fixed_b200 import native
    I = np.iinfo(np.int64)
    rng = np.random.default_rng(44)
    segs = []
    for kind in range(4):
        lo = I.min + 1 + np.sort(rng.integers(0, 5_000, size=300))
        mid = np.sort(rng.integers(-3_000, 3_000, size=300))
        hi = I.max - 6_000 + np.sort(rng.integers(0, 5_000, size=300))
        segs.append([np.concatenate([lo, mid, hi]), np.concatenate([lo, hi]), mid, np.concatenate([lo[:1], hi[-1:]])][kind])
    off = np.concatenate([[0], np.cumsum([len(s) for s in segs])]).astype(np.int64)
    ts = np.concatenate(segs).astype(np.int64)
    s_total = int(off[-1])
    g8 = rng.integers(0, 256, size=s_total, dtype=np.uint8)
    m8 = rng.integers(0, 256, size=s_total, dtype=np.uint8)
    w = 3000
    wu = rng.integers(-1, 5, size=w).astype(np.int64)
    a = rng.choice(ts, size=w).astype(object) + rng.integers(-3, 4, size=w)
    b = rng.choice(ts, size=w).astype(object) + rng.integers(-3, 4, size=w)
    clip = lambda x: min(max(int(x), int(I.min)), int(I.max))
    ws = np.array([clip(min(x, y)) for x, y in zip(a, b)], np.int64)
    we = np.array([clip(max(x, y)) for x, y in zip(a, b)], np.int64)
    ws[:3], we[:3] = [I.min, I.max, I.min], [I.max, I.max, I.min]
    esg, esm, ecnt = brute.usage_reduce(off, ts, g8.astype(np.int64), m8.astype(np.int64), wu, ws, we)
    assert (ecnt > 300).any() and (ecnt == 0).any()
    wide = th.usage_reduce(off, ts, g8.astype(np.int64), m8.astype(np.int64), wu, ws, we)
    narrow = th.usage_reduce_u8(off, ts, g8, m8, wu, ws, we)
    for outs in (wide, narrow):
        assert np.array_equal(outs[2], ecnt) and np.array_equal(outs[0], esg) and np.array_equal(outs[1], esm)
    # the ring: appended sample by sample per uuid (stamps of a uuid must not decrease), reduced in place
    lib = native.load()
    h = C.c_void_p(0)
    native.check(lib.th_ring_create(0, 16, C.byref(h)))
    uu = np.repeat(np.arange(4, dtype=np.int64), np.diff(off))
    native.check(lib.th_ring_append(h, s_total, uu.ctypes.data, ts.ctypes.data, g8.ctypes.data, m8.ctypes.data, None))
    outs = [np.zeros(w, np.int64) for _ in range(5)]
    native.check(lib.th_ring_reduce(h, w, wu.ctypes.data, ws.ctypes.data, we.ctypes.data, *[o.ctypes.data for o in outs], 0, None))
    assert np.array_equal(outs[2], ecnt) and np.array_equal(outs[0], esg) and np.array_equal(outs[1], esm)
    lib.th_ring_destroy(h)


def test_usage_random_logs_of_six_kinds(th):
    """tools/fuzz_usage.py on 40 seeds: exact cadence, jitter, missing stretches, repeated stamps, bursts, one break in a
    run; int64 columns (also negative values), uint8 columns and the ring appended in random batches, against brute force."""
    import importlib.util
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools', 'fuzz_usage.py')
    spec = importlib.util.spec_from_file_location('fuzz_usage', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.run(40, 2000) == 0
