"""The reference-facing surface (tensorhive_fixed_b200.dropin) against the reference's known answers and
against the line-by-line service oracle.  These read like the reference's own tests on purpose
(tests/unit/models/test_reservation_model.py)."""
import datetime as dt
from datetime import timedelta

import numpy as np
import pytest

from helpers import load_kat, kat_rows, EPOCH

pytestmark = pytest.mark.gpu

NOW_US, CASES = load_kat()
NOW = EPOCH + timedelta(microseconds=NOW_US)
GPU = 'GPU-d38d4de3-85ee-e837-3d87-e8e2faeb6a63'


class User:
    def __init__(self, uid, username):
        self.id, self.username, self.email = uid, username, username + '@example.org'


@pytest.fixture()
def dropin(lib_built):
    from tensorhive_fixed_b200 import dropin
    return dropin


def _calendar(dropin, rows, now_us=NOW_US, users=None):
    from tensorhive_fixed_b200.calendar import Calendar
    recs = [dropin.Reservation(id=r[0], resource_id=r[1], start=EPOCH + timedelta(microseconds=r[2]),
                               end=EPOCH + timedelta(microseconds=r[3]), user_id=r[4], is_cancelled=r[5],
                               title='TEST TITLE', description='TEST_DESCRIPTION') for r in rows]
    clock = {'now': EPOCH + timedelta(microseconds=now_us)}
    cal = Calendar(recs, users=users, clock=lambda: clock['now'])
    dropin.bind(cal)
    return cal, clock


@pytest.mark.parametrize('case', CASES, ids=[c['name'] for c in CASES])
def test_reference_known_answers_through_the_dropin(dropin, case):
    q = case['query']
    cal, clock = _calendar(dropin, kat_rows(case, NOW_US))
    sec = lambda s: NOW + timedelta(seconds=s)
    if case['op'] == 'would_interfere':
        r = dropin.Reservation(user_id=3, title='TEST TITLE', description='TEST_DESCRIPTION', resource_id=q['uuid'],
                               start=sec(q['start']), end=sec(q['end']), id=q['self_id'])
        assert r.would_interfere() is case['expect']
        if case['expect']:
            with pytest.raises(AssertionError):
                r.check_assertions()            # what save() raises in the reference
        else:
            r.check_assertions()
    elif case['op'] == 'current_events':
        clock['now'] = sec(q['now'])
        got = [r.id for r in dropin.Reservation.current_events(q['uuid'])]
        if 'expect' in case:
            assert got == case['expect']
        for i in case.get('expect_contains', []):
            assert i in got
        for i in case.get('expect_excludes', []):
            assert i not in got
    elif case['op'] == 'upcoming':
        clock['now'] = sec(q['now'])
        got = [r.id for r in dropin.Reservation.upcoming_events_for_resource(q['uuid'], timedelta(seconds=q['period']))]
        assert got == case['expect']
    elif case['op'] == 'range':
        got = [r.id for r in dropin.Reservation.filter_by_uuids_and_time_range(q['uuids'], sec(q['start']), sec(q['end']))]
        assert sorted(got) == case['expect']
    cal.close()


def test_invalid_reservation_time_range(dropin):
    # tests/unit/models/test_reservation_model.py:93-110
    cal, _ = _calendar(dropin, [])
    mk = lambda s, e: dropin.Reservation(user_id=1, title='T', description='D', resource_id=GPU, start=s, end=e)
    with pytest.raises(AssertionError):
        mk(NOW + timedelta(hours=3), NOW - timedelta(hours=3)).check_assertions()
    with pytest.raises(AssertionError):
        mk(NOW, NOW + timedelta(minutes=29, seconds=59)).check_assertions()
    with pytest.raises(AssertionError):
        mk(NOW, NOW + timedelta(days=8, seconds=1)).check_assertions()
    mk(NOW, NOW + timedelta(minutes=30)).check_assertions()
    with pytest.raises(AssertionError):
        dropin.Reservation.filter_by_uuids_and_time_range([GPU], '2026-01-01', NOW)
    cal.close()


def test_avg_matches_reference_semantics(dropin):
    cal, _ = _calendar(dropin, [])
    assert dropin.avg([1, 2, 4]) == 2 and dropin.avg([100] * 7 + [0]) == 87
    r = dropin.avg([])
    assert r == -1 and isinstance(r, float)
    cal.close()


def _world(seed, n=600, u=12, hosts=3):
    from tensorhive_fixed_b200 import synth
    from oracle.sql_ref import SqlCalendar
    from oracle import services_ref as ref
    rng = np.random.default_rng(seed)
    cal = synth.make_calendar(n, u, seed=seed)
    users = {i: User(i, 'user%d' % i) for i in range(1, 1001) if i % 7}     # some owners have no user row
    rows = [(int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]), int(cal['end'][i]),
             int(cal['owner'][i]), None if cal['cancelled_raw'][i] < 0 else int(cal['cancelled_raw'][i]))
            for i in range(n)]
    now_us = int(np.median(cal['start']))
    infra = {}
    for h in range(hosts):
        gpus = {}
        for g in range(u // hosts + 1):      # the last GPU of each host is unknown to the calendar
            code = h * (u // hosts) + g
            uuid = synth.uuid_string(code if g < u // hosts else 9000 + h)
            procs = []
            for _ in range(int(rng.integers(0, 4))):
                c = rng.random()
                owner = ('user%d' % rng.integers(1, 1001)) if c < 0.6 else ('ghost%d' % rng.integers(0, 3) if c < 0.8 else None)
                procs.append({'pid': int(rng.integers(100, 99999)), 'owner': owner,
                              'command': ['python train.py', 'Xorg', 'tensorhive_task run', '-'][int(rng.integers(0, 4))]})
            # make some processes belong to the current owner
            gpus[uuid] = {'name': 'B200', 'index': g, 'processes': procs if rng.random() > 0.1 else None}
        infra['host%d' % h] = {'GPU': gpus}
    infra['unreachable'] = {}
    sql = SqlCalendar(rows)
    world = ref.World(sql, {k: ref.UserRow(k, v.username, v.email) for k, v in users.items()})
    # plant owners: for GPUs with a current reservation, make the first process the owner's
    for h in infra.values():
        for uuid, data in (h.get('GPU') or {}).items():
            cur = world.current_events(now_us, uuid)
            if cur and cur[0].user is not None and data['processes']:
                data['processes'][0]['owner'] = cur[0].user.username
    return cal, rows, users, now_us, infra, world


class _IM:
    def __init__(self, infrastructure, dropin):
        self.infrastructure, self._d = infrastructure, dropin

    def all_nodes_with_gpu_processes(self):
        return self._d.all_nodes_with_gpu_processes(self.infrastructure)


@pytest.mark.parametrize('strict', [False, True])
@pytest.mark.parametrize('seed', [1, 2, 3])
def test_protection_service_violations_match_the_reference_loops(dropin, strict, seed):
    from oracle import services_ref as ref
    cal, rows, users, now_us, infra, world = _world(seed)
    c, _clock = _calendar(dropin, rows, now_us=now_us, users=users)
    want = ref.protection_violations(world, infra, now_us, strict)
    seen = []

    class Handler:
        def trigger_action(self, violation_data):
            seen.append(violation_data)

    svc = dropin.ProtectionService([Handler()], strict_reservations=strict, infrastructure_manager=_IM(infra, dropin))
    got = svc.collect_violations(svc.infrastructure_manager.all_nodes_with_gpu_processes())
    assert list(got.keys()) == list(want.keys())
    total = 0
    for host in want:
        assert list(got[host].keys()) == list(want[host].keys()), host      # intruders, in first-seen order
        for intruder in want[host]:
            g, w = got[host][intruder], want[host][intruder]
            assert g['VIOLATION_PIDS'] == w['VIOLATION_PIDS']
            assert len(g['RESERVATIONS']) == len(w['RESERVATIONS'])
            for a, b in zip(g['RESERVATIONS'], w['RESERVATIONS']):
                for key in ('OWNER_USERNAME', 'OWNER_EMAIL', 'GPU_UUID', 'HOSTNAME', 'RESERVATION_ID'):
                    assert a[key] == b[key], key
                assert (a['END'] is None) == (b['END'] is None)
                total += 1
    assert total > 0
    svc.do_run()
    assert len(seen) == sum(len(v) for v in want.values())
    assert all('GPUS' in v and 'OWNERS' in v for v in seen)
    c.close()


@pytest.mark.parametrize('seed', [4, 5])
def test_job_scheduling_checks_match_the_reference_loops(dropin, seed):
    from oracle import services_ref as ref
    cal, rows, users, now_us, infra, world = _world(seed, n=900, u=12)
    c, _clock = _calendar(dropin, rows, now_us=now_us, users=users)
    hosts = dropin.all_nodes_with_gpu_processes(infra)
    svc = dropin.JobSchedulingService()
    slots = svc.check_current_gpu_slots(hosts)
    want_slots = ref.check_current_gpu_slots(world, hosts, now_us)
    assert slots == want_slots                       # includes the float minutes, same expression
    kinds = {type(v) for h in slots.values() for v in h.values()}
    assert type(None) in kinds and (float in kinds or int in kinds)

    rng = np.random.default_rng(seed)
    host_names = [h for h in hosts if hosts[h]]
    ulist = list(users.values())
    jobs, ref_jobs = [], []
    for j in range(120):
        user = ulist[int(rng.integers(0, len(ulist)))]
        tasks = [ref.TaskRow(host_names[int(rng.integers(0, len(host_names)))],
                             None if rng.random() < 0.05 else int(rng.integers(0, 6))) for _ in range(int(rng.integers(1, 4)))]
        jobs.append(ref.JobRow(j, user, tasks))
        ref_jobs.append(ref.JobRow(j, world.users[user.id], tasks))
    for allow_own in (True, False):
        for period in (timedelta(0), timedelta(minutes=30), timedelta(days=2)):
            pus = (period.days * 86400 + period.seconds) * 1_000_000
            got = [dropin.JobSchedulingService.interferes_with_reservations(j, hosts, period, allow_own) for j in jobs[:25]]
            got += dropin.JobSchedulingService.interferes_with_reservations_many(jobs[25:], hosts, period, allow_own)
            want = [ref.interferes_with_reservations(world, j, hosts, now_us, pus, allow_own) for j in ref_jobs]
            assert got == want
    got = dropin.GreedyScheduler().schedule_jobs(jobs, slots)
    want = ref.greedy_schedule(world, ref_jobs, want_slots, now_us)
    assert [j.id for j in got] == [j.id for j in want]
    assert 0 < len(want) < len(jobs)
    c.close()


def test_bulk_revalidation_batches_would_interfere(dropin):
    # ReservationVerifier.update_user_reservations_statuses (ReservationVerifier.py:91-109), widening branch
    base = [(1, GPU, NOW_US, NOW_US + 3600_000_000, 1, None)]
    cal, _ = _calendar(dropin, base)
    mk = lambda i, s, e, c: dropin.Reservation(id=i, user_id=2, title='T', description='D', resource_id=GPU,
                                               start=NOW + timedelta(minutes=s), end=NOW + timedelta(minutes=e), is_cancelled=c)
    mine = [mk(10, 30, 90, True),      # overlaps the live row -> stays cancelled
            mk(11, 60, 120, True),     # touches it -> may be restored
            mk(12, 200, 260, False),   # not cancelled -> not a candidate
            mk(13, -500, -400, True),  # already over -> skipped
            mk(14, 100, 160, True),    # overlaps 11, which the loop has restored by then -> stays cancelled
            mk(15, 120, 180, True),    # touches 11, overlaps 14 (still cancelled) -> restored
            mk(16, 300, 400, True)]    # free, but the restriction walk says no
    back = dropin.reservations_that_may_be_restored(mine, is_allowed=lambda r: r.id != 16)
    assert [r.id for r in back] == [11, 15]
    cal.close()


def test_usage_log_averages(dropin):
    from tensorhive_fixed_b200 import synth
    from oracle import services_ref as ref
    cal, _ = _calendar(dropin, [])
    s = synth.make_samples(n_uuids=4, samples_per_uuid=5000, seed=3, cadence_us=60_000_000, jitter_us=2_000_000)
    codes = {synth.uuid_string(k): k for k in range(4)}
    log = dropin.UsageLog(s['offsets'], s['ts'], s['gpu'], s['mem'], codes)
    t0 = EPOCH + timedelta(microseconds=synth.T0_US)
    res = [dropin.Reservation(id=k, user_id=1, resource_id=synth.uuid_string(k % 4), start=t0 + timedelta(hours=3 * k),
                              end=t0 + timedelta(hours=3 * k + 2)) for k in range(20)]
    res.append(dropin.Reservation(id=99, user_id=1, resource_id=synth.uuid_string(1), start=t0 - timedelta(days=3),
                                  end=t0 - timedelta(days=2)))
    got = log.averages(res)
    for r, (ag, am) in zip(res, got):
        k = codes[r.resource_id]
        a, b = s['offsets'][k], s['offsets'][k + 1]
        lo = (r.start - EPOCH) // timedelta(microseconds=1)
        hi = (r.end - EPOCH) // timedelta(microseconds=1)
        sel = (s['ts'][a:b] >= lo) & (s['ts'][a:b] <= hi)
        want_g = ref.avg([int(x) for x in s['gpu'][a:b][sel]])
        want_m = ref.avg([int(x) for x in s['mem'][a:b][sel]])
        assert ag == want_g and am == want_m and type(ag) is type(want_g)
    assert got[-1] == (-1.0, -1.0)
    cal.close()


def test_usage_ring_follows_the_logger_rules(dropin):
    """UsageRing against a literal replay of the reference logger: per reservation a list that grows by one entry per
    tick while the reservation is current and both metrics are reported (Log.updated_log :92-99, log_current_usage
    :159-171), averaged with avg() (:25-30) when it expires."""
    from oracle import services_ref as ref
    cal, _ = _calendar(dropin, [])
    rng = np.random.default_rng(12)
    gpus = ['GPU-%036d' % k for k in range(3)]
    t0 = dt.datetime(2026, 4, 1, 8, 0, 0)
    res = []
    for k in range(12):
        s = t0 + timedelta(minutes=int(rng.integers(0, 600)))
        res.append(dropin.Reservation(id=k + 1, user_id=1, title='T', description='D', resource_id=gpus[k % 3], start=s,
                                      end=s + timedelta(minutes=int(rng.integers(30, 240)))))
    res.append(dropin.Reservation(id=50, user_id=1, title='T', description='D', resource_id='GPU-' + 'f' * 36,
                                  start=t0, end=t0 + timedelta(hours=1)))              # a GPU that never reports
    ring = dropin.UsageRing(initial_capacity=4)                                        # forces the buffers to grow
    lists = {r.id: ([], []) for r in res}
    dropped = 0
    for tick in range(0, 900 * 60, 47):                                                # one sample every 47 s
        now = t0 + timedelta(seconds=tick, microseconds=int(rng.integers(0, 1000)))
        for g in gpus:
            gu = None if rng.random() < 0.05 else int(rng.integers(0, 101))            # unsupported metric now and then
            mu = None if rng.random() < 0.05 else int(rng.integers(0, 101))
            kept = ring.append(g, now, gu, mu)
            assert kept == (gu is not None and mu is not None)
            dropped += 0 if kept else 1
            if kept:
                for r in res:                                                          # the reference's per-reservation log
                    if r.resource_id == g and r.start <= now <= r.end:
                        lists[r.id][0].append(gu)
                        lists[r.id][1].append(mu)
    got = ring.averages(res)
    for r, (ag, am) in zip(res, got):
        want = (ref.avg(lists[r.id][0]), ref.avg(lists[r.id][1]))
        assert (ag, am) == want and type(ag) is type(want[0]), r.id
    assert got[-1] == (-1.0, -1.0) and any(a >= 0 for a, _ in got)
    info = ring.info()
    assert len(ring) == info['samples'] == sum(1 for _ in range(0, 900 * 60, 47)) * 3 - dropped
    assert info['n_uuids'] == 3 and info['relayouts'] >= 5 and info['regular_uuids'] == 0      # jittered stamps: probed
    assert dropin.UsageRing().averages(res[:2]) == [(-1.0, -1.0)] * 2
    with pytest.raises(AssertionError):
        ring.append(gpus[0], t0, 1, 1)                                                 # time must not run backwards
    cal.close()


def test_calendar_view_controller(dropin):
    """controllers/reservation.py:31-55 over the snapshot: argument rules, status codes, id order, serialisation."""
    from oracle.sql_ref import SqlCalendar
    g1, g2 = GPU, GPU[:-1] + '4'
    rows = [(5, g1, NOW_US - 3_600_000_000, NOW_US + 3_600_000_000, 1, None),
            (2, g1, NOW_US + 3_600_000_000, NOW_US + 7_200_000_000, 2, True),        # cancelled rows are returned too
            (9, g2, NOW_US + 1_800_000_000, NOW_US + 9_000_000_000, 1, False),
            (3, g1, NOW_US + 86_400_000_000, NOW_US + 90_000_000_000, 2, None)]      # outside the view
    cal, _ = _calendar(dropin, rows, users={1: User(1, 'alice'), 2: User(2, 'bob')})
    fmt = lambda us: (EPOCH + timedelta(microseconds=us)).strftime('%Y-%m-%dT%H:%M:%S.%fZ')
    lo, hi = NOW_US, NOW_US + 3_600_000_000                                          # ends inclusive: row 2 touches hi
    body, status = dropin.get_selected([g1, g2, g1], fmt(lo), fmt(hi))
    assert status == 200
    sql = SqlCalendar([(r[0], r[1], r[2], r[3], r[4], None if r[5] is None else int(r[5])) for r in rows])
    want = sql.filter_by_uuids_and_time_range([g1, g2], EPOCH + timedelta(microseconds=lo), EPOCH + timedelta(microseconds=hi))
    assert [d['id'] for d in body] == want == [2, 5, 9]
    first = body[0]
    assert set(first) == {'id', 'title', 'description', 'resourceId', 'userId', 'gpuUtilAvg', 'memUtilAvg', 'start', 'end',
                          'createdAt', 'isCancelled', 'userName'}
    assert first['userName'] == 'bob' and first['isCancelled'] is True and first['resourceId'] == g1
    assert first['start'] == (EPOCH + timedelta(microseconds=rows[1][2])).strftime('%Y-%m-%dT%H:%M:%S') + '+00:00'
    assert dropin.get_selected([g1], fmt(lo), None) == ({'msg': 'Bad Request'}, 400)     # all three or nothing
    assert dropin.get_selected([], fmt(lo), fmt(hi))[1] == 400
    body, status = dropin.get_selected([g1], '2026-01-01 00:00', fmt(hi))             # not the API's date format
    assert status == 400 and body['msg'].startswith('Bad Request. ')
    assert dropin.get_selected(['GPU-unknown'], fmt(lo), fmt(hi)) == ([], 200)
    cal.close()



def test_bulk_column_load_on_the_device(dropin):
    """Calendar.refresh_from_columns builds the same snapshot as the row-wise path (device export compared)."""
    from tensorhive_fixed_b200 import synth
    from tensorhive_fixed_b200.calendar import Calendar
    cal = synth.make_calendar(3000, 9, seed=31)
    n = len(cal['id'])
    names = [synth.uuid_string(int(k)) for k in cal['uuid_id']]
    raw = [None if c < 0 else int(c) for c in cal['cancelled_raw']]
    rows = [dropin.Reservation(id=int(cal['id'][i]), resource_id=names[i], start=EPOCH + timedelta(microseconds=int(cal['start'][i])),
                               end=EPOCH + timedelta(microseconds=int(cal['end'][i])), user_id=int(cal['owner'][i]),
                               is_cancelled=raw[i], title='T', description='D') for i in range(n)]
    a = Calendar(rows)
    b = Calendar([])
    b.refresh_from_columns(cal['id'], names, cal['start'].astype('datetime64[us]'), cal['end'], cal['owner'], raw)
    ea, eb = a.state()[0].export(), b.state()[0].export()
    ua, ub = a.state()[1].names, b.state()[1].names
    assert [ua[k] for k in ea['uuid_id']] != [] and len(ea['id']) == n
    key = lambda e, names_: sorted(zip(e['id'].tolist(), [names_[k] for k in e['uuid_id']], e['start'].tolist(),
                                       e['end'].tolist(), e['owner'].tolist(), e['cancelled'].tolist()))
    assert key(ea, ua) == key(eb, ub)
    dropin.bind(b)
    probe = rows[17]
    got = dropin.Reservation.filter_by_uuids_and_time_range([probe.resource_id], probe.start, probe.end)
    assert probe.id in [r.id for r in got] and all(r.resource_id == probe.resource_id for r in got)
    a.close(); b.close()


def test_save_move_save_and_destroy(dropin):
    """The reference's own sequence (tests/unit/models/test_reservation_model.py:10-64) through save(): a colliding
    candidate raises on save; against a cancelled base it saves, gets the next id, and can then be moved over its own old
    slot (self-exclusion by id); destroy() removes it.  CRUDModel.py:21-50."""
    base = dropin.Reservation(id=1, user_id=1, title='TEST TITLE', description='TEST_DESCRIPTION', resource_id=GPU,
                              start=NOW, end=NOW + timedelta(minutes=60))
    cal, _ = _calendar(dropin, [])
    cal.refresh([base])
    mk = lambda s, e: dropin.Reservation(user_id=3, title='TEST TITLE', description='TEST_DESCRIPTION', resource_id=GPU,
                                         start=NOW + timedelta(minutes=s), end=NOW + timedelta(minutes=e))
    shapes = [(5, 55), (-5, 55), (5, 65), (-5, 65)]
    for s, e in shapes:                                              # :10-37 every shape collides with the live row
        with pytest.raises(AssertionError):
            mk(s, e).save()
    assert [r.id for r in dropin.Reservation.current_events(GPU)] == [1]
    base.is_cancelled = True
    base.save()                                                      # the cancel itself is a save of row 1
    assert dropin.Reservation.current_events(GPU) == []
    cand = mk(*shapes[0])
    cand.save()                                                      # :40-64 against the cancelled base: no collision
    assert cand.id == 2 and dropin.Reservation.get(2) is cand
    for s, e in shapes[1:]:
        cand.start, cand.end = NOW + timedelta(minutes=s), NOW + timedelta(minutes=e)
        cand.save()                                                  # overlaps only its own previous slot
        assert cand.id == 2
    other = mk(30, 90)
    with pytest.raises(AssertionError):
        other.save()                                                 # but another user's row now collides with it
    assert other.id is None
    cand.destroy()
    with pytest.raises(dropin.NoResultFound):
        dropin.Reservation.get(2)
    other.save()
    assert other.id == 3                                             # AUTOINCREMENT never reuses an id
    assert [r.id for r in dropin.Reservation.filter_by_uuids_and_time_range([GPU], NOW, NOW + timedelta(hours=2))] == [1, 3]
    cal.close()


def test_usage_ring_stays_on_the_device_and_knows_regular_runs(dropin):
    """The device-resident ring (th_ring_*): a tick costs one small copy whatever the log's length; uuids whose stamps
    form an exact arithmetic run are reduced without reading the ts column, and a single off-grid stamp turns a uuid back
    into a probed one -- all with the answers of th_usage_reduce_u8 over the same samples."""
    import time as _time
    import tensorhive_fixed_b200 as th
    from oracle import brute
    cal, _ = _calendar(dropin, [])
    rng = np.random.default_rng(21)
    n_gpus, ticks = 40, 3000
    names = ['GPU-%036d' % k for k in range(n_gpus)]
    ring = dropin.UsageRing(initial_capacity=16)
    t0 = 1_800_000_000_000_000
    ts_all = np.zeros((n_gpus, ticks), np.int64)
    g_all = rng.integers(0, 101, size=(n_gpus, ticks)).astype(np.uint8)
    m_all = rng.integers(0, 101, size=(n_gpus, ticks)).astype(np.uint8)
    per_tick = []
    for k in range(ticks):
        stamp = t0 + k * 1_000_000
        batch = []
        for u in range(n_gpus):
            t = stamp + (int(rng.integers(0, 900_000)) if u >= 30 else 0)       # GPUs 30.. are stamped off the grid
            if u == 7 and k == 2000:
                t += 1                                                             # one late stamp breaks GPU 7's run
            ts_all[u, k] = t
            batch.append((names[u], t, int(g_all[u, k]), int(m_all[u, k])))
        a = _time.perf_counter()
        assert ring.append_tick(batch) == n_gpus
        per_tick.append(_time.perf_counter() - a)
    info = ring.info()
    assert info['samples'] == n_gpus * ticks and info['n_uuids'] == n_gpus
    assert info['regular_uuids'] == 29                                             # 0..29 on the grid, minus GPU 7
    # the cost of a tick does not grow with the log (regions double: a handful of re-layouts in 3000 ticks)
    early, late = np.median(per_tick[200:700]), np.median(per_tick[-500:])
    assert late < 3 * early + 2e-4, (early, late)
    assert info['relayouts'] <= 12
    w = 4000
    wu = rng.integers(-1, n_gpus + 2, size=w).astype(np.int64)
    ws = t0 + rng.integers(-50, ticks, size=w).astype(np.int64) * 1_000_000 + rng.integers(0, 2, size=w) * rng.integers(0, 10**6, size=w)
    we = ws + rng.integers(0, 400, size=w).astype(np.int64) * 1_000_000 + rng.integers(0, 2, size=w) * rng.integers(0, 10**6, size=w)
    we[::50] = ws[::50] - 5                                                        # end before start: empty
    got = ring.reduce(wu, ws, we)
    # the same samples as plain CSR columns through the existing entry point, and through the oracle
    off = np.arange(n_gpus + 1, dtype=np.int64) * ticks
    code = np.array([ring.code(n) for n in names])
    assert np.array_equal(code, np.arange(n_gpus))
    want = th.usage_reduce_u8(off, ts_all.reshape(-1), g_all.reshape(-1), m_all.reshape(-1), wu, ws, we)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    sg, sm, cnt = brute.usage_reduce(off, ts_all.reshape(-1), g_all.reshape(-1).astype(np.int64),
                                     m_all.reshape(-1).astype(np.int64), wu[:600], ws[:600], we[:600])
    assert np.array_equal(got[0][:600], sg) and np.array_equal(got[1][:600], sm) and np.array_equal(got[2][:600], cnt)
    # CUDA tensors in, CUDA tensors out: nothing but the windows moves
    import torch
    dgot = ring.reduce(*[torch.from_numpy(x).cuda() for x in (wu, ws, we)])
    assert all(np.array_equal(a.cpu().numpy(), b) for a, b in zip(dgot, got))
    ring.close()
    cal.close()


@pytest.mark.parametrize('n', [1, 17, 48, 83])
def test_usage_ring_every_range_of_a_short_log(dropin, n):
    """Every index range of an n-sample ring log as a window, for a uuid on an exact cadence (both ends from divisions,
    ts never read) and one stamped off the grid (probed): the edge-chunk split of the uint8 sums in every position."""
    cal, _ = _calendar(dropin, [])
    rng = np.random.default_rng(300 + n)
    ring = dropin.UsageRing(initial_capacity=16)
    t0 = 1_800_000_000_000_000
    g = rng.integers(0, 256, size=(2, n)).astype(np.uint8)
    m = rng.integers(0, 256, size=(2, n)).astype(np.uint8)
    ts = np.zeros((2, n), np.int64)
    for k in range(n):
        ts[0, k] = t0 + k * 1_000_000
        ts[1, k] = t0 + k * 1_000_000 + int(rng.integers(0, 900_000))
        ring.append_tick([('GPU-regular', int(ts[0, k]), int(g[0, k]), int(m[0, k])),
                          ('GPU-jitter', int(ts[1, k]), int(g[1, k]), int(m[1, k]))])
    assert ring.info()['regular_uuids'] == (1 if n > 1 else ring.info()['regular_uuids'])
    ab = np.array([(u, a, b) for u in range(2) for a in range(n) for b in range(a, n)], np.int64)
    wu = np.array([ring.code('GPU-regular'), ring.code('GPU-jitter')], np.int64)[ab[:, 0]]
    ws = ts[ab[:, 0], ab[:, 1]]
    we = ts[ab[:, 0], ab[:, 2]]
    # windows at the int64 extremes, for both uuids: everything, nothing after the log, nothing before it
    I = np.iinfo(np.int64)
    xs = np.array([I.min, I.max - 5, I.min, t0 + 1, I.min], np.int64)
    xe = np.array([I.max, I.max, I.min + 5, I.max, t0 - 1], np.int64)
    for u in range(2):
        code = ring.code(['GPU-regular', 'GPU-jitter'][u])
        xg, xm, xc = ring.reduce(np.full(5, code, np.int64), xs, xe)[:3]
        first = int(ts[u, 0] >= t0 + 1)
        assert list(xc) == [n, 0, 0, n - 1 + first, 0]
        assert xg[0] == g[u].astype(np.int64).sum() and xm[3] == m[u, 1 - first:].astype(np.int64).sum()
    sg, sm, cnt = ring.reduce(wu, ws, we)[:3]
    for u in range(2):
        sel = ab[:, 0] == u
        cg = np.concatenate([[0], np.cumsum(g[u].astype(np.int64))])
        cm = np.concatenate([[0], np.cumsum(m[u].astype(np.int64))])
        a, b = ab[sel, 1], ab[sel, 2]
        assert np.array_equal(cnt[sel], b - a + 1)
        assert np.array_equal(sg[sel], cg[b + 1] - cg[a]) and np.array_equal(sm[sel], cm[b + 1] - cm[a])
    ring.close()
    cal.close()
