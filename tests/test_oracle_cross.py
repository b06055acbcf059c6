"""The two oracle forms (SQL text through sqlite3, brute-force numpy) must agree with each other."""
import numpy as np
import pytest

from oracle import brute
from oracle.sql_ref import SqlCalendar
from tensorhive_fixed_b200 import synth


def _sql_calendar(cal):
    rows = []
    for i in range(len(cal['id'])):
        raw = int(cal['cancelled_raw'][i])
        rows.append((int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]),
                     int(cal['end'][i]), int(cal['owner'][i]), None if raw < 0 else raw))
    return SqlCalendar(rows)


@pytest.fixture(scope='module')
def world():
    cal = synth.make_calendar(n=3000, n_uuids=16, seed=20260101)
    q = synth.make_queries(cal, m=1500, seed=20260102)
    return cal, q, _sql_calendar(cal)


def test_would_interfere_sql_vs_brute(world):
    cal, q, sql = world
    bits = brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'])
    sets = brute.conflict_sets(cal, q['uuid_id'], q['start'], q['end'], q['self_id'], brute.STRICT)
    assert 0.2 < bits.mean() < 0.95          # the generator exercises both verdicts
    for i in range(len(bits)):
        sid = None if q['self_id'][i] < 0 else int(q['self_id'][i])
        u = synth.uuid_string(int(q['uuid_id'][i]))
        rows = sql.would_interfere_rows(u, int(q['start'][i]), int(q['end'][i]), sid)
        assert [r[0] for r in rows] == sets[i]
        assert any(not c for _, c in rows) == bool(bits[i])


def test_current_and_upcoming_sql_vs_brute(world):
    cal, _q, sql = world
    n_uuids = cal['n_uuids']
    mid = int(np.median(cal['start']))
    for now in (mid, int(cal['start'][5]), int(cal['end'][9]), int(cal['start'].min()) - 1):
        res, own = brute.current_owner(cal, n_uuids, now)
        for u in range(n_uuids):
            ids = sql.current_events(now, synth.uuid_string(u))
            assert (ids[0] if ids else -1) == res[u]
        for period in (0, 30 * synth.MIN_US, 3 * synth.DAY_US):
            uu = np.arange(n_uuids)
            sets = brute.conflict_sets(cal, uu, np.full(n_uuids, now), np.full(n_uuids, now + period),
                                       np.full(n_uuids, -1), brute.UPCOMING)
            for u in range(n_uuids):
                assert sql.upcoming_events_for_resource(synth.uuid_string(u), now, period) == sets[u]


def test_range_filter_sql_vs_brute(world):
    import datetime as dt
    from oracle.sql_ref import EPOCH
    cal, _q, sql = world
    rng = np.random.default_rng(5)
    for _ in range(20):
        k = int(rng.integers(1, 6))
        uu = rng.choice(cal['n_uuids'], size=k, replace=False)
        s = int(rng.integers(cal['start'].min(), cal['end'].max()))
        e = s + int(rng.integers(0, 20 * synth.DAY_US))
        sets = brute.conflict_sets(cal, uu, np.full(k, s), np.full(k, e), np.full(k, -1), brute.INCLUSIVE)
        want = sorted(i for x in sets for i in x)
        got = sql.filter_by_uuids_and_time_range([synth.uuid_string(int(u)) for u in uu],
                                                 EPOCH + dt.timedelta(microseconds=s), EPOCH + dt.timedelta(microseconds=e))
        assert sorted(got) == want


def test_equal_start_ties_follow_id_order():
    # ORDER BY _start leaves ties to SQLite: observed = ascending id (unpinned in the reference)
    base = synth.T0_US
    rows = [(7, 'g', base + 10, base + 50, 1, None), (3, 'g', base + 10, base + 60, 1, None),
            (5, 'g', base + 10, base + 70, 1, 0), (4, 'g', base + 5, base + 20, 1, None)]
    sql = SqlCalendar(rows)
    assert sql.upcoming_events_for_resource('g', base, 100) == [4, 3, 5, 7]
