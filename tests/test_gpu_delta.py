"""Snapshot maintenance from deltas (th_snapshot_apply_delta): after any sequence of inserts, updates, cancels and
deletes the snapshot must be indistinguishable from one loaded from scratch with the resulting table."""
import datetime as dt
from datetime import timedelta

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

COLS = ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')


def _fresh(th, table, u):
    snap = th.Snapshot(0)
    snap.load(*[table[k] for k in COLS], n_uuids=u)
    return snap


def _same(a, b):
    ea, eb = a.export(), b.export()
    return all(np.array_equal(ea[k], eb[k]) for k in ea)


@pytest.mark.parametrize('in_place', [False, True])
def test_random_delta_sequences_equal_a_fresh_load(lib_built, in_place):
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle import brute
    rng = np.random.default_rng(7 if in_place else 8)
    u = 40
    cal = synth.make_calendar(20_000, u, seed=5)
    table = {k: cal[k].copy() for k in COLS}
    cur = _fresh(th, table, u)
    spare = th.Snapshot(0)
    next_id = int(table['id'].max()) + 1
    for step in range(6):
        n = len(table['id'])
        # deletes, updates (moved / cancelled rows keep their id), inserts (some on brand-new uuids)
        del_pos = rng.choice(n, size=int(rng.integers(0, 300)), replace=False)
        rest = np.setdiff1d(np.arange(n), del_pos)
        upd_pos = rng.choice(rest, size=int(rng.integers(0, 300)), replace=False)
        n_ins = int(rng.integers(0, 400))
        if step == 3:
            u += 5
        ins = synth.make_calendar(max(n_ins, 1), u, seed=100 + step)
        ins = {k: ins[k][:n_ins] for k in COLS}
        ins['id'] = np.arange(next_id, next_id + n_ins, dtype=np.int64)
        next_id += n_ins
        upd = {k: table[k][upd_pos].copy() for k in COLS}
        upd['start'] = upd['start'] + rng.integers(-10**9, 10**9, size=len(upd_pos))
        upd['end'] = np.maximum(upd['end'] + rng.integers(-10**9, 10**9, size=len(upd_pos)), upd['start'])
        upd['cancelled'] = (rng.random(len(upd_pos)) < 0.3).astype(np.uint8)
        ups = {k: np.concatenate([upd[k], ins[k]]) for k in COLS}
        order = rng.permutation(len(ups['id']))
        ups = {k: v[order] for k, v in ups.items()}
        del_ids = table['id'][del_pos].copy()
        if step == 2:
            del_ids = np.concatenate([del_ids, [10**12]])          # deleting an id that does not exist is a no-op
        # the truth
        keep = np.ones(n, bool)
        keep[del_pos] = False
        keep[upd_pos] = False
        table = {k: np.concatenate([table[k][keep], ups[k]]) for k in COLS}
        # the delta
        dst = cur if in_place else spare
        dst.apply_delta(cur, u, *[ups[k] for k in COLS], del_ids)
        if not in_place:
            cur, spare = dst, cur
        ref = _fresh(th, table, u)
        assert cur.info()['n'] == len(table['id'])
        assert _same(cur, ref), step
        q = synth.make_queries(dict(table, n_uuids=u), 3000, seed=200 + step)
        got = cur.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
        assert np.array_equal(got, brute.conflict_bits(dict(table, n_uuids=u), q['uuid_id'], q['start'], q['end'], q['self_id']))
        ref.close()
    cur.close(); spare.close()


def test_large_delta_with_equal_keys_merges_in_id_order(lib_built):
    """A delta above the single-CTA sort limit whose rows share (uuid, start) with surviving rows and with each other:
    the merge must interleave them by id, exactly as a sort from scratch would."""
    import tensorhive_fixed_b200 as th
    rng = np.random.default_rng(11)
    u, n = 6, 30_000
    starts = rng.integers(0, 50, size=n) * 1_000_000                # 50 distinct starts: ties everywhere
    table = {'id': rng.permutation(np.arange(1, 3 * n, 3))[:n].astype(np.int64),      # ids with gaps, shuffled
             'uuid_id': rng.integers(0, u, size=n).astype(np.int64), 'start': starts.astype(np.int64),
             'end': (starts + rng.integers(0, 10**7, size=n)).astype(np.int64),
             'owner': rng.integers(1, 9, size=n).astype(np.int64), 'cancelled': (rng.random(n) < 0.2).astype(np.uint8)}
    snap = _fresh(th, table, u)
    k = 5000
    new_ids = rng.permutation(np.setdiff1d(np.arange(1, 3 * n), table['id']))[:k].astype(np.int64)   # ids BETWEEN old ones
    ns = rng.integers(0, 50, size=k) * 1_000_000
    ups = {'id': new_ids, 'uuid_id': rng.integers(0, u, size=k).astype(np.int64), 'start': ns.astype(np.int64),
           'end': (ns + rng.integers(0, 10**7, size=k)).astype(np.int64), 'owner': rng.integers(1, 9, size=k).astype(np.int64),
           'cancelled': (rng.random(k) < 0.2).astype(np.uint8)}
    gone = rng.choice(n, size=3000, replace=False)
    snap.apply_delta(snap, u, *[ups[c] for c in COLS], table['id'][gone])
    keep = np.ones(n, bool)
    keep[gone] = False
    truth = {c: np.concatenate([table[c][keep], ups[c]]) for c in COLS}
    ref = _fresh(th, truth, u)
    assert _same(snap, ref)
    ex = snap.export()
    order = np.lexsort((truth['id'], truth['start'], truth['uuid_id']))
    assert np.array_equal(ex['id'], truth['id'][order])
    snap.close(); ref.close()


def test_delta_with_device_pointers_and_empty_cases(lib_built):
    import torch
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    dev = torch.device('cuda:0')
    cal = synth.make_calendar(5000, 8, seed=9)
    table = {k: cal[k].copy() for k in COLS}
    snap = _fresh(th, table, 8)
    z = np.zeros(0, np.int64)
    snap.apply_delta(snap, 8, z, z, z, z, z, np.zeros(0, np.uint8), z)           # empty delta: nothing changes
    assert _same(snap, _fresh(th, table, 8))
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    snap.apply_delta(snap, 8, t(z), t(z), t(z), t(z), t(z), t(np.zeros(0, np.uint8)), t(table['id']))   # delete everything
    assert snap.info()['n'] == 0
    assert not snap.conflict_bitmap(np.zeros(4, np.int64), np.zeros(4, np.int64), np.full(4, 2**62, np.int64),
                                    np.full(4, -1, np.int64)).any()
    snap.apply_delta(snap, 8, *[t(table[k]) for k in COLS], t(z))                # and insert it all back
    assert _same(snap, _fresh(th, table, 8))
    with pytest.raises(th.NativeError):
        snap.apply_delta(snap, 4, z, z, z, z, z, np.zeros(0, np.uint8), z)       # uuid dictionary cannot shrink
    snap.close()


def test_calendar_apply_delta_through_the_dropin(lib_built):
    from tensorhive_fixed_b200 import dropin
    from tensorhive_fixed_b200.calendar import Calendar
    gpu = 'GPU-d38d4de3-85ee-e837-3d87-e8e2faeb6a63'
    gpu2 = 'GPU-d38d4de3-85ee-e837-3d87-e8e2faeb6a64'
    now = dt.datetime(2026, 3, 1, 12, 0, 0)
    mk = lambda i, g, s, e, c=None: dropin.Reservation(id=i, user_id=1, title='T', description='D', resource_id=g,
                                                       start=now + timedelta(minutes=s), end=now + timedelta(minutes=e),
                                                       is_cancelled=c)
    cal = Calendar([mk(1, gpu, 0, 60)], clock=lambda: now)
    dropin.bind(cal)
    cand = mk(None, gpu, 30, 90)
    assert cand.would_interfere() is True
    v0 = cal.version
    cal.apply_delta(upserts=[mk(1, gpu, 0, 60, True)])                 # cancel row 1
    assert cal.version == v0 + 1 and cand.would_interfere() is False
    cal.apply_delta(upserts=[mk(2, gpu, 45, 120), mk(3, gpu2, 0, 600)])   # insert two rows, one on a new GPU
    assert cand.would_interfere() is True
    assert [r.id for r in dropin.Reservation.current_events(gpu2)] == [3]
    cal.apply_delta(upserts=[mk(2, gpu, 90, 150)])                     # move row 2: now it only touches the candidate
    assert cand.would_interfere() is False
    cal.apply_delta(deleted_ids=[3])
    assert dropin.Reservation.current_events(gpu2) == []
    for _ in range(4):                                                 # handle rotation
        cal.apply_delta(upserts=[mk(1, gpu, 0, 60, None)])
    assert cand.would_interfere() is True
    cal.close()


def test_duplicate_ids_are_refused_and_leave_the_snapshot_as_it_was(lib_built):
    """The primary key is unique: a load or a delta that breaks that is refused (TH_EINVAL) BEFORE the handle is touched,
    so the previous snapshot stays queryable; Calendar.apply_delta collapses a batch first (last change wins)."""
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from tensorhive_fixed_b200.calendar import Calendar
    cal = synth.make_calendar(5000, 8, seed=11)
    table = {k: cal[k].copy() for k in COLS}
    snap = _fresh(th, table, 8)
    before = snap.export()
    two = {k: np.concatenate([table[k][:1], table[k][:1]]) for k in COLS}
    two['end'][1] += 5
    with pytest.raises(th.NativeError) as e:                       # one id upserted twice
        snap.apply_delta(snap, 8, *[two[k] for k in COLS], np.zeros(0, np.int64))
    assert e.value.code == -1 and 'more than once' in str(e.value)
    with pytest.raises(th.NativeError):                            # one id upserted and deleted
        snap.apply_delta(snap, 8, *[table[k][:1] for k in COLS], table['id'][:1].copy())
    bad = {k: table[k].copy() for k in COLS}
    bad['id'][7] = bad['id'][4000]                                  # duplicate id in a full load (ids not ascending)
    with pytest.raises(th.NativeError) as e:
        snap.load(*[bad[k] for k in COLS], n_uuids=8)
    assert 'not unique' in str(e.value)
    bad = {k: table[k].copy() for k in COLS}
    bad['uuid_id'][3] = 8                                           # uuid code out of range
    with pytest.raises(th.NativeError):
        snap.load(*[bad[k] for k in COLS], n_uuids=8)
    after = snap.export()                                           # every refusal left the old table in place
    assert all(np.array_equal(before[k], after[k]) for k in before)
    q = synth.make_queries(cal, 2000, seed=12)
    from oracle import brute
    assert np.array_equal(snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id']),
                          brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id']))
    # a negative id in the delete list matches nothing
    snap.apply_delta(snap, 8, *[table[k][:0] for k in COLS], np.array([-5], np.int64))
    assert all(np.array_equal(before[k], snap.export()[k]) for k in before)
    snap.close()

    # the calendar collapses a batch: update + cancel of the same reservation -> the cancel; changed + destroyed -> gone
    rows = [{'id': 1, 'user_id': 1, 'resource_id': 'G' * 40, 'start': 0, 'end': 100, 'is_cancelled': None},
            {'id': 2, 'user_id': 1, 'resource_id': 'G' * 40, 'start': 200, 'end': 300, 'is_cancelled': None}]
    c = Calendar(rows)
    c.apply_delta(upserts=[dict(rows[0], end=150), dict(rows[0], end=150, is_cancelled=True), dict(rows[1], end=400)],
                  deleted_ids=[2])
    ex = c.state()[0].export()
    assert ex['id'].tolist() == [1] and ex['end'].tolist() == [150] and ex['cancelled'].tolist() == [1]
    with pytest.raises(ValueError):
        c.apply_delta(upserts=[dict(rows[0], id=-3)])
    c.close()


def test_delta_that_stretches_the_calendar_to_both_ends_of_int64(lib_built):
    """A delta whose rows start near INT64_MIN and INT64_MAX: the span of the calendar (and of two uuids) jumps past
    2^63 us, the grid is re-planned, and the result equals a fresh load; a second delta removes them again."""
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle import brute
    I = np.iinfo(np.int64)
    u = 12
    cal = synth.make_calendar(5_000, u, seed=31)
    table = {k: cal[k].copy() for k in COLS}
    cur = _fresh(th, table, u)
    base = int(table['id'].max()) + 1
    ext = {'id': np.arange(base, base + 6, dtype=np.int64), 'uuid_id': np.array([0, 0, 3, 3, 11, 11], np.int64),
           'start': np.array([I.min + 1, I.max - 5_000, I.min + 77, I.max - 1, I.min + 1, 0], np.int64),
           'end': np.array([I.min + 900, I.max - 10, I.min + 78, I.max, I.min + 1, I.max], np.int64),
           'owner': np.arange(1, 7, dtype=np.int64), 'cancelled': np.array([0, 0, 1, 0, 0, 0], np.uint8)}
    cur.apply_delta(cur, u, *[ext[k] for k in COLS], np.zeros(0, np.int64))
    table2 = {k: np.concatenate([table[k], ext[k]]) for k in COLS}
    ref = _fresh(th, table2, u)
    assert _same(cur, ref)
    q = synth.make_queries(dict(table, n_uuids=u), 2000, seed=32)
    qs, qe = q['start'].copy(), q['end'].copy()
    qs[:6], qe[:6] = [I.min, I.min, I.max - 3, 0, I.min + 1, I.max], [I.max, I.min + 50, I.max, I.max, I.min + 2, I.max]
    qu = q['uuid_id'].copy()
    qu[:6] = [0, 3, 3, 11, 11, 0]
    for mode, bmode in ((th.MODE_STRICT, brute.STRICT), (th.MODE_INCLUSIVE, brute.INCLUSIVE)):
        f = q['self_id'] if mode == th.MODE_STRICT else None
        fb = q['self_id'] if mode == th.MODE_STRICT else np.full(len(qu), -1, np.int64)
        got = cur.conflict_bitmap(qu, qs, qe, f, mode)
        assert np.array_equal(np.asarray(got, bool), np.asarray(brute.conflict_bits(dict(table2, n_uuids=u), qu, qs, qe, fb, bmode), bool))
    cur.apply_delta(cur, u, *[np.zeros(0, np.int64) for _ in range(5)], np.zeros(0, np.uint8), ext['id'])
    back = _fresh(th, table, u)
    assert _same(cur, back)
    for s in (cur, ref, back):
        s.close()
