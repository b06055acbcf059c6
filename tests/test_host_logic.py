"""Host-side logic of the drop-in layer on a CPU-only box: the same reference-surface tests as tests/test_gpu_dropin.py,
with the device snapshot replaced by a stand-in that answers from the brute-force oracle.  This exercises everything
above the C ABI -- uuid dictionary, datetime conversion, owner encoding, flattening and re-grouping of the process
table, the greedy pass, error behaviour -- without a GPU; the kernels themselves are covered by the `-m gpu` tests."""
import numpy as np
import pytest

import test_gpu_dropin as G
import test_gpu_model_vectors as M
import test_gpu_exec_vectors as E
from oracle import brute


class FakeSnapshot:
    """Same methods as tensorhive_fixed_b200.api.Snapshot, answered by oracle/brute.py (test double, CPU only)."""

    def __init__(self, device=0):
        self.cal = None
        self.n = self.n_uuids = 0

    def load(self, id, uuid_id, start, end, owner, cancelled, n_uuids):
        self.cal = {'id': np.asarray(id, np.int64), 'uuid_id': np.asarray(uuid_id, np.int64),
                    'start': np.asarray(start, np.int64), 'end': np.asarray(end, np.int64),
                    'owner': np.asarray(owner, np.int64), 'cancelled': np.asarray(cancelled, np.uint8), 'n_uuids': n_uuids}
        self.n, self.n_uuids = len(self.cal['id']), n_uuids
        return self

    def apply_delta(self, src, n_uuids, id, uuid_id, start, end, owner, cancelled, delete_ids):
        old = src.cal
        gone = np.isin(old['id'], np.concatenate([np.asarray(id, np.int64), np.asarray(delete_ids, np.int64)]))
        cols = {}
        for k, new in (('id', id), ('uuid_id', uuid_id), ('start', start), ('end', end), ('owner', owner), ('cancelled', cancelled)):
            cols[k] = np.concatenate([old[k][~gone], np.asarray(new, old[k].dtype)])
        return self.load(n_uuids=n_uuids, **cols)

    def info(self):
        return {'n': self.n, 'n_uuids': self.n_uuids, 'device_bytes': 0, 'radix_passes': 0}

    def close(self):
        pass

    def conflict_bitmap(self, qu, qs, qe, qf=None, mode=0):
        qf = np.full(len(qu), -1, np.int64) if qf is None else qf
        return brute.conflict_bits(self.cal, qu, qs, qe, qf, mode)

    def conflict_sets(self, qu, qs, qe, qf=None, mode=0, capacity=None):
        qf = np.full(len(qu), -1, np.int64) if qf is None else qf
        lists = brute.conflict_sets(self.cal, qu, qs, qe, qf, mode)
        off = np.zeros(len(lists) + 1, np.int64)
        off[1:] = np.cumsum([len(x) for x in lists])
        ids = np.array([i for x in lists for i in x], np.int64)
        flag = {int(i): int(c) for i, c in zip(self.cal['id'], self.cal['cancelled'])}
        return off, ids, np.array([flag[int(i)] for i in ids], np.uint8)

    def current_owner(self, now_us, device_out=False):
        return brute.current_owner(self.cal, self.n_uuids, now_us)

    def upcoming(self, now_us, period_us, uuids):
        return brute.upcoming_summary(self.cal, uuids, now_us, period_us)

    def upcoming_owner_flags(self, now_us, period_us, task_uuid, task_owner, test):
        k = len(task_uuid)
        lists = brute.conflict_sets(self.cal, task_uuid, np.full(k, now_us), np.full(k, now_us + period_us),
                                    np.full(k, -1), brute.UPCOMING)
        own = {int(i): int(o) for i, o in zip(self.cal['id'], self.cal['owner'])}
        fn = {0: lambda o, w: len(w) > 0, 1: lambda o, w: any(own[x] != o for x in w),
              2: lambda o, w: any(own[x] == o for x in w)}[test]
        return np.array([fn(int(task_owner[j]), lists[j]) for j in range(k)], np.uint8)

    def violation_join(self, now_us, strict, proc_uuid, proc_owner):
        res, own = brute.current_owner(self.cal, self.n_uuids, now_us)
        pu = np.asarray(proc_uuid, np.int64)
        ok = (pu >= 0) & (pu < self.n_uuids)
        r = np.where(ok, res[np.clip(pu, 0, max(self.n_uuids - 1, 0))], -1)
        o = np.where(ok, own[np.clip(pu, 0, max(self.n_uuids - 1, 0))], -1)
        flag = np.where(r >= 0, (o >= 0) & (np.asarray(proc_owner) != o), bool(strict))
        return flag.astype(np.uint8), np.where((r >= 0) & flag, r, -1)


def _fake_usage_reduce(off, ts, gpu, mem, wu, ws, we, device=0):
    sg, sm, cnt = brute.usage_reduce(off, ts, gpu, mem, wu, ws, we)
    ag = np.array([brute.floor_avg(a, c) for a, c in zip(sg, cnt)], np.int64)
    am = np.array([brute.floor_avg(a, c) for a, c in zip(sm, cnt)], np.int64)
    return sg, sm, cnt, ag, am


@pytest.fixture()
def dropin(monkeypatch):
    import time
    import tensorhive_fixed_b200.calendar as calmod
    from tensorhive_fixed_b200 import dropin as d
    monkeypatch.setenv('TZ', 'UTC')          # the executed-reference vectors were made with TZ=UTC (utc2local)
    time.tzset()
    monkeypatch.setattr(calmod, 'Snapshot', FakeSnapshot)
    monkeypatch.setattr(d.api, 'usage_reduce', _fake_usage_reduce)
    monkeypatch.setattr(d.api, 'usage_reduce_u8', _fake_usage_reduce)
    return d


# the reference-surface tests, now collected without the gpu marker and run against the stand-in
test_known_answers = G.test_reference_known_answers_through_the_dropin
test_invalid_time_range = G.test_invalid_reservation_time_range
test_avg = G.test_avg_matches_reference_semantics
test_protection_service = G.test_protection_service_violations_match_the_reference_loops
test_job_scheduling = G.test_job_scheduling_checks_match_the_reference_loops
test_bulk_revalidation = G.test_bulk_revalidation_batches_would_interfere
test_usage_log = G.test_usage_log_averages
test_save_destroy = G.test_save_move_save_and_destroy
# (the sample ring lives on the device: its tests are GPU tests)
test_calendar_view = G.test_calendar_view_controller
# the host logic of the drop-in against the executed reference's model vectors (queries and the write path)
test_model_methods = M.test_model_methods_match_the_executed_reference
test_write_path = M.test_write_path_matches_the_executed_reference
test_bulk_revalidation_vectors = M.test_bulk_revalidation_matches_the_executed_reference
test_calendar_view_vectors = M.test_calendar_view_matches_the_executed_reference
# ... and the service loops (V, J1-J3) against the vectors of the executed reference services
test_services = E.test_services_match_the_executed_reference


def test_calendar_delta_bookkeeping(dropin):
    import datetime as dt
    from datetime import timedelta
    from tensorhive_fixed_b200.calendar import Calendar
    now = dt.datetime(2026, 3, 1, 12)
    mk = lambda i, g, s, e, c=None: dropin.Reservation(id=i, user_id=1, title='T', description='D', resource_id=g,
                                                       start=now + timedelta(minutes=s), end=now + timedelta(minutes=e),
                                                       is_cancelled=c)
    g1, g2 = 'GPU-' + 'a' * 36, 'GPU-' + 'b' * 36
    cal = Calendar([mk(1, g1, 0, 60)], clock=lambda: now)
    dropin.bind(cal)
    cand = mk(None, g1, 30, 90)
    assert cand.would_interfere() is True
    cal.apply_delta(upserts=[mk(1, g1, 0, 60, True), mk(2, g2, 0, 600)])
    assert cand.would_interfere() is False
    assert [r.id for r in dropin.Reservation.current_events()] == [2]
    assert cal.all_resource_ids() == [g1, g2] and cal.version == 2
    cal.apply_delta(deleted_ids=[2])
    assert dropin.Reservation.current_events() == []
    with pytest.raises(RuntimeError):
        dropin._calendar = None
        dropin.Reservation.current_events()


def test_reservation_times_are_parsed_like_the_reference(dropin):
    """Reservation.py:61-79 with utils/DateUtils.py:47-60: the REST layer hands over strings in the API's input format."""
    import datetime as dt
    from tensorhive_fixed_b200.calendar import Calendar
    now = dt.datetime(2026, 3, 1, 12)
    gpu = 'GPU-' + 'c' * 36
    cal = Calendar([dropin.Reservation(id=1, user_id=1, title='T', description='D', resource_id=gpu,
                                       start='2026-03-01T12:00:00.000Z', end='2026-03-01T13:00:00.000Z')], clock=lambda: now)
    dropin.bind(cal)
    r = dropin.Reservation(user_id=1, title='T', description='D', resource_id=gpu, start='2026-03-01T12:30:00.000Z',
                           end='2026-03-01T13:30:00.500Z')
    assert r.start == dt.datetime(2026, 3, 1, 12, 30) and r.end == dt.datetime(2026, 3, 1, 13, 30, 0, 500000)
    assert r.duration == dt.timedelta(hours=1, microseconds=500000)
    assert r.would_interfere() is True
    with pytest.raises(AssertionError, match='would interfere'):
        r.check_assertions()
    r.start, r.end = '2026-03-01T13:00:00.000Z', '2026-03-01T14:00:00.000Z'       # touching: no conflict
    r.check_assertions()
    with pytest.raises(ValueError):
        r.start = '2026-03-01 13:00:00'                      # not the API format: DateUtils.parse_string re-raises
    r.end = 1_700_000_000                                    # neither str nor datetime -> None
    assert r.end is None
    with pytest.raises(AssertionError, match='end time is invalid'):
        r.check_assertions()


def test_bulk_load_from_columns_equals_the_row_wise_load(dropin):
    """Calendar.refresh_from_columns (what `SELECT id, resource_id, _start, _end, user_id, is_cancelled` yields, with
    SQLite's text datetimes and NULLs) against the row-wise constructor; rows come back lazily, by id."""
    import datetime as dt
    from datetime import timedelta
    from tensorhive_fixed_b200.calendar import Calendar
    now = dt.datetime(2026, 3, 1, 12)
    g1, g2 = 'GPU-' + 'a' * 36, 'GPU-' + 'b' * 36
    rows = [(7, g2, now - timedelta(minutes=10), now + timedelta(minutes=50), 1, None),
            (3, g1, now + timedelta(minutes=60), now + timedelta(minutes=120), 2, 1),
            (5, g1, now - timedelta(hours=1), now + timedelta(hours=1, microseconds=250), None, 0),
            (9, g1, now + timedelta(minutes=20), now + timedelta(minutes=80), 2, None)]
    users = {1: object(), 2: object()}
    mk = lambda r: dropin.Reservation(id=r[0], resource_id=r[1], start=r[2], end=r[3], user_id=r[4], is_cancelled=r[5],
                                      title='T', description='D')
    a = Calendar([mk(r) for r in rows], users=users, clock=lambda: now)
    b = Calendar([], users=users, clock=lambda: now)
    text = lambda d: d.strftime('%Y-%m-%d %H:%M:%S.%f')                 # the way SQLAlchemy stores DateTime in SQLite
    b.refresh_from_columns([r[0] for r in rows], [r[1] for r in rows], [text(r[2]) for r in rows], [text(r[3]) for r in rows],
                           [r[4] for r in rows], [r[5] for r in rows],
                           row_factory=lambda **kw: dropin.Reservation(title='T', description='D', **kw))
    sa, sb = a.state()[0].cal, b.state()[0].cal
    name_a, name_b = a.state()[1].names, b.state()[1].names
    key = lambda c, names: sorted(zip(c['id'], [names[k] for k in c['uuid_id']], c['start'], c['end'], c['owner'], c['cancelled']))
    assert key(sa, name_a) == key(sb, name_b)
    for cal in (a, b):
        dropin.bind(cal)
        assert [r.id for r in dropin.Reservation.current_events()] == [5, 7]
        assert [r.id for r in dropin.Reservation.upcoming_events_for_resource(g1, timedelta(minutes=30))] == [5, 9]
        got = dropin.Reservation.filter_by_uuids_and_time_range([g1], now, now + timedelta(hours=1))
        assert [(r.id, r.is_cancelled, r.user_id, r.start, r.end) for r in got] == \
            [(3, True, 2, rows[1][2], rows[1][3]), (5, False, None, rows[2][2], rows[2][3]), (9, False, 2, rows[3][2], rows[3][3])]
    # deltas on top of a bulk load: overlay of upserts / deletes over the lazy rows
    b.apply_delta(upserts=[mk((9, g1, now + timedelta(minutes=200), now + timedelta(minutes=260), 2, None)),
                           mk((11, g2, now, now + timedelta(minutes=30), 1, None))], deleted_ids=[5])
    dropin.bind(b)
    assert [r.id for r in dropin.Reservation.current_events()] == [7, 11]
    assert [r.id for r in dropin.Reservation.filter_by_uuids_and_time_range([g1], now, now + timedelta(hours=9))] == [3, 9]
    with pytest.raises(KeyError):
        b.rows_for([5])
    assert b.rows_for([3])[0].resource_id == g1


def test_global_calendar_is_defined_chunk_by_chunk():
    """synth.GlobalCalendar (what bench.py's config-5 share and its all-ranks parity check are built on): any process can
    materialise any chunk, chunks are deterministic, ids unique across chunks, uuid codes inside the chunk's range, and the
    partition by estimated rows per uuid balances the ranks."""
    from tensorhive_fixed_b200 import synth, sharding
    g = synth.GlobalCalendar(200_000, 2048, seed=5, chunk_uuids=256)
    assert g.n_chunks == 8 and g.n_rows == 200_000 and int(g.chunk_rows.sum()) == 200_000
    assert g.chunk_rows.min() > 0.6 * 25_000 and g.chunk_rows.max() < 1.4 * 25_000
    a, b = g.chunk(3), g.chunk(3)
    assert all(np.array_equal(a[k], b[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled'))
    ids = np.concatenate([g.chunk(c)['id'] for c in range(g.n_chunks)])
    assert len(np.unique(ids)) == 200_000 and ids.min() == 1 and ids.max() == 200_000
    for c in (0, 5, 7):
        u = g.chunk(c)['uuid_id']
        assert u.min() >= c * 256 and u.max() < (c + 1) * 256
    bounds = sharding.partition_uuids(g.rows_per_uuid_estimate(), 4)
    assert bounds[0] == 0 and bounds[-1] == 2048 and np.all(np.diff(bounds) > 0)
    rows = [sum(int((sharding.owner_rank(g.chunk(c)['uuid_id'], bounds) == r).sum()) for c in range(g.n_chunks)) for r in range(4)]
    assert sum(rows) == 200_000 and max(rows) < 1.05 * 50_000
    assert list(g.chunks_overlapping(int(bounds[1]), int(bounds[2]))) == list(range(int(bounds[1]) // 256, (int(bounds[2]) - 1) // 256 + 1))
    q = g.chunk_queries(2, 5000, seed=9, unknown_frac=0.02)
    known = q['uuid_id'] < 2048
    assert (~known).any() and q['uuid_id'][known].min() >= 512 and q['uuid_id'][known].max() < 768
    # queries of a chunk are routed to the ranks that own its uuids, unknown resources to rank 0
    assert set(np.unique(sharding.owner_rank(q['uuid_id'][~known], bounds))) == {0}
