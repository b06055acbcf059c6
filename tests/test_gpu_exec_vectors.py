"""The reference-facing surface (tensorhive_fixed_b200.dropin, kernels behind the C ABI) against vectors produced by
EXECUTING the reference's own service sources (tests/golden/make_reference_exec_vectors.py): ProtectionService.do_run,
JobSchedulingService.check_current_gpu_slots / interferes_with_reservations, GreedyScheduler.schedule_jobs, avg and the
usage logger tick by tick.  The reference is not on the GPU box; the committed JSON holds inputs and outputs."""
import json
import os
import time
from datetime import timedelta

import pytest

from helpers import HERE, EPOCH

pytestmark = pytest.mark.gpu

with open(os.path.join(HERE, 'golden', 'reference_exec_vectors.json')) as f:
    VEC = json.load(f)
WORLDS = VEC['worlds']


class User:
    def __init__(self, uid, username, email):
        self.id, self.username, self.email = uid, username, email


class Task:
    def __init__(self, hostname, gpu_id):
        self.hostname, self.gpu_id = hostname, gpu_id


class Job:
    def __init__(self, jid, user, tasks):
        self.id, self.user, self.tasks = jid, user, tasks


class InfraManager:
    def __init__(self, infrastructure, dropin):
        self.infrastructure, self._d = infrastructure, dropin

    def all_nodes_with_gpu_processes(self):
        return self._d.all_nodes_with_gpu_processes(self.infrastructure)


@pytest.fixture()
def dropin(lib_built):
    os.environ['TZ'] = 'UTC'          # the vectors were made with TZ=UTC (utc2local, core/utils/time.py:5-9)
    time.tzset()
    from tensorhive_fixed_b200 import dropin
    return dropin


def bind_world(dropin, rows, now_us, users):
    from tensorhive_fixed_b200.calendar import Calendar
    recs = [dropin.Reservation(id=r[0], resource_id=r[1], start=EPOCH + timedelta(microseconds=r[2]),
                               end=EPOCH + timedelta(microseconds=r[3]), user_id=r[4], is_cancelled=r[5],
                               title='T', description='D') for r in rows]
    clock = {'now': EPOCH + timedelta(microseconds=now_us)}
    cal = Calendar(recs, users=users, clock=lambda: clock['now'])
    dropin.bind(cal)
    return cal, clock


@pytest.mark.parametrize('w', WORLDS, ids=['world%d' % w['seed'] for w in WORLDS])
def test_services_match_the_executed_reference(dropin, w):
    users = {u[0]: User(*u) for u in w['users']}
    cal, _clock = bind_world(dropin, w['rows'], w['now_us'], users)
    exp = w['expect']
    im = InfraManager(w['infrastructure'], dropin)
    hosts = im.all_nodes_with_gpu_processes()
    assert hosts == exp['all_nodes_with_gpu_processes']

    # V: ProtectionService.do_run -> what each handler call receives, in order
    for strict, key in ((False, 'lenient'), (True, 'strict')):
        seen = []

        class Handler:
            def trigger_action(self, violation_data):
                seen.append(violation_data)

        svc = dropin.ProtectionService([Handler()], strict_reservations=strict, infrastructure_manager=im)
        svc.do_run()
        want = exp['protection'][key]
        assert len(seen) == len(want)
        for g, x in zip(seen, want):
            assert g['INTRUDER_USERNAME'] == x['INTRUDER_USERNAME']
            assert {h: sorted(p) for h, p in g['VIOLATION_PIDS'].items()} == x['VIOLATION_PIDS']
            assert g['GPUS'] == x['GPUS'] and g['OWNERS'] == x['OWNERS']
            assert len(g['RESERVATIONS']) == len(x['RESERVATIONS'])
            for a, b in zip(g['RESERVATIONS'], x['RESERVATIONS']):
                for k in ('OWNER_USERNAME', 'OWNER_EMAIL', 'GPU_UUID', 'GPU_NAME', 'GPU_ID', 'HOSTNAME'):
                    assert a[k] == b[k], k
                end = a['END'].strftime('%Y-%m-%dT%H:%M:%S.%f') if a['END'] is not None else None
                assert end == b['END']

    # J1: minutes to the next reservation per GPU (floats by the reference's own expression)
    slots = dropin.JobSchedulingService().check_current_gpu_slots(hosts)
    assert slots == exp['gpu_slots']

    # J2 / J3
    jobs = [Job(j['id'], users.get(j['user_id']) if j['user_id'] is not None else None, [Task(*t) for t in j['tasks']])
            for j in w['jobs']]
    assert [[dropin.Scheduler.get_assigned_gpu_uid(t, hosts) for t in j.tasks] for j in jobs] == exp['assigned_gpu_uids']
    for case in exp['interferes']:
        period = timedelta(minutes=case['period_minutes'])
        got = [dropin.JobSchedulingService.interferes_with_reservations(j, hosts, period, case['allow_own'])
               for j in jobs[:6]]
        got += dropin.JobSchedulingService.interferes_with_reservations_many(jobs[6:], hosts, period, case['allow_own'])
        assert got == case['flags'], case
    assert [j.id for j in dropin.GreedyScheduler().schedule_jobs(jobs, slots)] == exp['scheduled_job_ids']
    cal.close()


def test_avg_matches_the_executed_reference(dropin):
    cal, _ = bind_world(dropin, [], 0, {})
    for case in VEC['avg']:
        got = dropin.avg(case['data'])
        assert got == case['expect'] and isinstance(got, float) == case['expect_is_float']
    cal.close()


@pytest.mark.parametrize('w', VEC['usage_worlds'], ids=['usage%d' % w['seed'] for w in VEC['usage_worlds']])
def test_usage_logger_matches_the_executed_reference(dropin, w):
    """log_current_usage + handle_expired_logs, tick by tick, through the sample ring, K4 and Reservation.save()."""
    cal, clock = bind_world(dropin, w['rows'], w['ticks'][0]['now_us'], {})
    im = InfraManager({}, dropin)
    svc = dropin.UsageLoggingService(infrastructure_manager=im)
    saved = []
    for tick in w['ticks'] + [{'now_us': w['end_us'], 'gpus': {}}]:
        clock['now'] = EPOCH + timedelta(microseconds=tick['now_us'])
        im.infrastructure = {'host0': {'GPU': tick['gpus']}}
        for r in svc.do_run():
            saved.append([r.id, r.gpu_util_avg, r.mem_util_avg])
    assert sorted(saved) == w['expect_saved']
    for (_rid, ag, _am), (_r2, eg, _e2) in zip(sorted(saved), w['expect_saved']):
        assert type(ag) is type(eg)                 # float(-1) for an empty log, an int otherwise
    # the summaries went through save(): the snapshot rows carry them
    for rid, ag, am in saved:
        row = dropin.Reservation.get(rid)
        assert (row.gpu_util_avg, row.mem_util_avg) == (ag, am)
    cal.close()
