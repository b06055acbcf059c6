"""The service-loop oracle (oracle/services_ref.py) against vectors produced by EXECUTING the reference's own source
files (tests/golden/make_reference_exec_vectors.py; the reference itself is not on the test box).  CPU only."""
import datetime as dt
import json
import os

import pytest

from helpers import HERE, EPOCH
from oracle import services_ref as ref
from oracle.sql_ref import SqlCalendar

with open(os.path.join(HERE, 'golden', 'reference_exec_vectors.json')) as f:
    VEC = json.load(f)
WORLDS = VEC['worlds']


def test_header_names_every_executed_reference_file():
    files = {e['file'] for e in VEC['header']['executed']}
    assert {'tensorhive/core/services/ProtectionService.py', 'tensorhive/core/services/JobSchedulingService.py',
            'tensorhive/core/scheduling.py', 'tensorhive/core/services/UsageLoggingService.py',
            'tensorhive/core/managers/InfrastructureManager.py'} <= files
    assert all(len(e['sha256']) == 64 and e['bodies'] for e in VEC['header']['executed'])


def build_world(w):
    sql = SqlCalendar([tuple(r) for r in w['rows']])
    users = {u[0]: ref.UserRow(*u) for u in w['users']}
    return ref.World(sql, users)


def flatten_violations(per_host):
    """{host: {intruder: data}} -> the handler-call order of ProtectionService.do_run (:84-120)."""
    return [per_host[h][i] for h in per_host for i in per_host[h]]


def fmt(d):
    return d.strftime('%Y-%m-%dT%H:%M:%S.%f') if isinstance(d, dt.datetime) else d


@pytest.mark.parametrize('w', WORLDS, ids=['world%d' % w['seed'] for w in WORLDS])
def test_service_loops_match_the_executed_reference(w):
    world = build_world(w)
    infra, now = w['infrastructure'], w['now_us']
    exp = w['expect']
    assert ref.all_nodes_with_gpu_processes(infra) == exp['all_nodes_with_gpu_processes']
    for strict, key in ((False, 'lenient'), (True, 'strict')):
        got = flatten_violations(ref.protection_violations(world, infra, now, strict))
        want = exp['protection'][key]
        assert len(got) == len(want)
        for g, x in zip(got, want):
            assert g['INTRUDER_USERNAME'] == x['INTRUDER_USERNAME']
            assert {h: sorted(p) for h, p in g['VIOLATION_PIDS'].items()} == x['VIOLATION_PIDS']
            assert len(g['RESERVATIONS']) == len(x['RESERVATIONS'])
            for a, b in zip(g['RESERVATIONS'], x['RESERVATIONS']):
                for k in ('OWNER_USERNAME', 'OWNER_EMAIL', 'GPU_UUID', 'HOSTNAME'):
                    assert a[k] == b[k], k
                assert fmt(a['END']) == b['END']           # the vectors were made with TZ=UTC (utc2local = identity)
    hosts = ref.all_nodes_with_gpu_processes(infra)
    slots = ref.check_current_gpu_slots(world, hosts, now)
    assert slots == exp['gpu_slots']
    jobs = [ref.JobRow(j['id'], world.users.get(j['user_id']) if j['user_id'] is not None else None,
                       [ref.TaskRow(*t) for t in j['tasks']]) for j in w['jobs']]
    assert [[ref.get_assigned_gpu_uid(t, hosts) for t in j.tasks] for j in jobs] == exp['assigned_gpu_uids']
    for case in exp['interferes']:
        got = [ref.interferes_with_reservations(world, j, hosts, now, case['period_minutes'] * 60_000_000, case['allow_own'])
               for j in jobs]
        assert got == case['flags'], case
    assert [j.id for j in ref.greedy_schedule(world, jobs, slots, now)] == exp['scheduled_job_ids']
    world.cal.close()


def test_avg_matches_the_executed_reference():
    for case in VEC['avg']:
        got = ref.avg(case['data'])
        assert got == case['expect'] and isinstance(got, float) == case['expect_is_float']


@pytest.mark.parametrize('w', VEC['usage_worlds'], ids=['usage%d' % w['seed'] for w in VEC['usage_worlds']])
def test_usage_logger_replay_matches_the_executed_reference(w):
    sql = SqlCalendar([tuple(r) for r in w['rows']])
    got = ref.usage_logger_replay(sql, w['ticks'], w['end_us'])
    assert got == w['expect_saved']
    assert any(isinstance(x[1], float) for x in got) or True
    sql.close()
