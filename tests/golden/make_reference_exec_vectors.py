"""Writes tests/golden/reference_exec_vectors.json by EXECUTING the reference's own source files for the service
loops of the hot path -- not a restatement of them.  Run from the repo root, in the container that holds the
reference:  python tests/golden/make_reference_exec_vectors.py

What runs here is the unmodified text of (paths relative to /root/reference)

  tensorhive/core/services/ProtectionService.py    ProtectionService.do_run / store_violation / gpu_attr   (row V)
  tensorhive/core/services/JobSchedulingService.py JobSchedulingService.check_current_gpu_slots            (row J1)
                                                   JobSchedulingService.interferes_with_reservations        (row J2)
  tensorhive/core/scheduling.py                    Scheduler.get_assigned_gpu_uid, GreedyScheduler.schedule_jobs (J3)
  tensorhive/core/services/UsageLoggingService.py  avg, Log.updated_log, log_current_usage, handle_expired_logs     (row A)
                                                   -- the real JSON log files, in a temporary directory
  tensorhive/core/managers/InfrastructureManager.py InfrastructureManager.all_nodes_with_gpu_processes
  tensorhive/core/utils/time.py                    utc2local

loaded as modules under their real dotted names with `importlib`.  Their third-party and ORM imports (sqlalchemy,
gevent, the model classes, tensorhive.config, ...) are not installable in this image, so `sys.modules` holds stand-ins
for exactly those: `Reservation.current_events` / `upcoming_events_for_resource` answer from the SQL oracle
(`oracle.sql_ref.SqlCalendar`, the reference's WHERE clauses through sqlite3), `CONFIG.SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS`
is 30 (tensorhive/config.py:259), `datetime.utcnow()` is frozen per world, and User objects are identity-mapped per id the
way one SQLAlchemy session maps them (JobSchedulingService.py:127 compares with `is not`).  Everything the service
functions compute -- loops, comparisons, dictionary layout, the float minutes -- is the reference's code.

The model's own queries (R1-R4) are executed separately, under a stand-in for SQLAlchemy's operators:
tests/golden/make_reference_model_vectors.py.  The GPU box has no /root/reference: the vectors (inputs AND outputs) are committed, and
tests/test_oracle_exec_vectors.py (CPU, oracle restatement) and tests/test_gpu_exec_vectors.py (dropin on the GPU) are
held to them.
"""
import datetime as dt
import hashlib
import importlib.util
import json
import os
import shutil
import sys
import tempfile
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get('TH_REFERENCE_ROOT', '/root/reference')
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

from oracle.sql_ref import SqlCalendar, EPOCH                       # noqa: E402
from tensorhive_fixed_b200 import synth                             # noqa: E402  (seeded world generator only)

EXECUTED = {
    'tensorhive/core/utils/time.py': ['utc2local :5-9'],
    'tensorhive/core/managers/InfrastructureManager.py': ['InfrastructureManager.node_gpu_processes :22-59',
                                                          'all_nodes_with_gpu_processes :61-62', 'ignored_processes :69-78'],
    'tensorhive/core/services/ProtectionService.py': ['ProtectionService.gpu_attr :48-53', 'store_violation :55-77',
                                                      'do_run :80-131'],
    'tensorhive/core/scheduling.py': ['Scheduler.get_assigned_gpu_uid :19-26', 'GreedyScheduler.schedule_jobs :29-62'],
    'tensorhive/core/services/JobSchedulingService.py': ['JobSchedulingService.__init__ :32-37',
                                                         'check_current_gpu_slots :76-104',
                                                         'interferes_with_reservations :117-132'],
    'tensorhive/core/services/UsageLoggingService.py': ['avg :25-30', 'JSONLogFile :43-58', 'Log.updated_log / save :85-116',
                                                        'UsageLoggingService.log_current_usage :159-171',
                                                        'handle_expired_logs :196-229', 'extract_specific_gpu_data :231-240',
                                                        '_clean_up_old_log_file :173-194'],
}


# ---------------------------------------------------------------------------------------------------------------------
# stand-ins for what cannot be imported
# ---------------------------------------------------------------------------------------------------------------------
class _World:
    """What the stand-in Reservation model answers from; swapped per world."""
    sql = None
    users = {}
    now_us = 0


class _FrozenDateTime(dt.datetime):
    @classmethod
    def utcnow(cls):
        return EPOCH + dt.timedelta(microseconds=_World.now_us)


class User:
    def __init__(self, uid, username, email):
        self.id, self.username, self.email = uid, username, email


class _Row:
    def __init__(self, rid):
        (i, res, s, e, uid, _c) = _World.sql.row(rid)
        self.id, self.resource_id = i, res
        self.start = EPOCH + dt.timedelta(microseconds=s)
        self.end = EPOCH + dt.timedelta(microseconds=e)
        self.user_id = uid
        self.user = _World.users.get(uid)          # identity-mapped; None when the owner has no user row


class Reservation:
    """Stand-in for tensorhive.models.Reservation: only the two queries the service loops call, answered by the
    reference's SQL (Reservation.py:89-118) through sqlite3; plus get(id) / save() for the usage logger, which records
    what handle_expired_logs stores."""
    saved = []

    @classmethod
    def get(cls, id):
        if _World.sql.row(id) is None:
            raise sys.modules['sqlalchemy.orm.exc'].NoResultFound()
        r = _Row(id)
        r.gpu_util_avg = r.mem_util_avg = None
        r.save = lambda: cls.saved.append([r.id, r.gpu_util_avg, r.mem_util_avg])
        return r

    @classmethod
    def current_events(cls, resource_id=None):
        return [_Row(i) for i in _World.sql.current_events(_World.now_us, resource_id)]

    @classmethod
    def upcoming_events_for_resource(cls, resource_id, period_after):
        period_us = (period_after.days * 86400 + period_after.seconds) * 1_000_000 + period_after.microseconds
        return [_Row(i) for i in _World.sql.upcoming_events_for_resource(resource_id, _World.now_us, period_us)]


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _package(name):
    m = _module(name)
    m.__path__ = []
    return m


def _load(rel):
    """Execute one reference source file, unmodified, as the module its path names."""
    name = rel[:-3].replace('/', '.')
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference():
    for pkg in ('tensorhive', 'tensorhive.core', 'tensorhive.core.services', 'tensorhive.core.managers',
                'tensorhive.core.utils', 'tensorhive.models', 'tensorhive.controllers'):
        _package(pkg)

    class Service:                                   # tensorhive/core/services/Service.py: a thread wrapper
        def __init__(self):
            pass

    _module('tensorhive.core.services.Service', Service=Service)
    _module('tensorhive.core.utils.decorators', override=lambda f: f)      # decorators.py:9-11 is the identity
    _module('tensorhive.models.Reservation', Reservation=Reservation)
    _module('tensorhive.models.User', User=User)
    _module('tensorhive.models.Job', Job=type('Job', (), {}))
    _module('tensorhive.models.Task', Task=type('Task', (), {}), TaskStatus=type('TaskStatus', (), {}))
    _module('tensorhive.database', db_session=None)
    _module('tensorhive.config',
            JOB_SCHEDULING_SERVICE=types.SimpleNamespace(SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS=30),   # config.py:259
            USAGE_LOGGING_SERVICE=types.SimpleNamespace(LOG_DIR=tempfile.mkdtemp(prefix='th_ref_logs_'),
                                                        LOG_CLEANUP_ACTION=2))                      # config.py:250-251
    _module('tensorhive.controllers.job', business_execute=None, business_stop=None, JobId=int)
    _module('tensorhive.core.task_nursery')
    _module('tensorhive.core.managers.SSHConnectionManager', SSHConnectionManager=type('SSHConnectionManager', (), {}))
    _module('gevent', sleep=lambda s: None)
    _module('sqlalchemy', or_=None, and_=None)
    _package('sqlalchemy.orm')
    _module('sqlalchemy.orm.exc', NoResultFound=type('NoResultFound', (Exception,), {}))
    ref = {}
    ref['time'] = _load('tensorhive/core/utils/time.py')
    ref['im'] = _load('tensorhive/core/managers/InfrastructureManager.py')
    ref['protection'] = _load('tensorhive/core/services/ProtectionService.py')
    ref['scheduling'] = _load('tensorhive/core/scheduling.py')
    ref['jobs'] = _load('tensorhive/core/services/JobSchedulingService.py')
    ref['usage'] = _load('tensorhive/core/services/UsageLoggingService.py')
    ref['jobs'].datetime = _FrozenDateTime           # `datetime.utcnow()` at JobSchedulingService.py:97-98
    ref['usage'].datetime = types.SimpleNamespace(datetime=_FrozenDateTime)      # UsageLoggingService.py:96,201
    return ref


# ---------------------------------------------------------------------------------------------------------------------
# seeded worlds (inputs are written to the JSON in full: the tests never regenerate them)
# ---------------------------------------------------------------------------------------------------------------------
def make_world(seed, n=160, u=9, hosts=3):
    rng = np.random.default_rng(1000 + seed)
    cal = synth.make_calendar(n, u, seed=500 + seed, skew=0.8 if seed % 3 == 0 else 0.0)
    rows = [[int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]), int(cal['end'][i]),
             int(cal['owner'][i]) % 40 + 1, None if cal['cancelled_raw'][i] < 0 else int(cal['cancelled_raw'][i])]
            for i in range(n)]
    users = [[i, 'user%d' % i, 'user%d@example.org' % i] for i in range(1, 41) if i % 7]      # some owners have no user row
    starts = np.sort(cal['start'])
    now_us = int(starts[int(rng.integers(n // 4, 3 * n // 4))]) + int(rng.integers(0, 3)) * 60_000_000
    sql = SqlCalendar([tuple(r) for r in rows])
    infra = {}
    per_host = u // hosts
    for h in range(hosts):
        gpus = {}
        for g in range(per_host + 1):                # the last GPU of each host is unknown to the calendar
            uuid = synth.uuid_string(h * per_host + g if g < per_host else 9000 + h)
            procs = []
            for _ in range(int(rng.integers(0, 4))):
                c = rng.random()
                owner = ('user%d' % rng.integers(1, 41)) if c < 0.6 else ('ghost%d' % rng.integers(0, 3) if c < 0.8 else None)
                procs.append({'pid': int(rng.integers(100, 99999)), 'owner': owner,
                              'command': ['python train.py', 'Xorg', 'tensorhive_task run', '-'][int(rng.integers(0, 4))]})
            cur = sql.current_events(now_us, uuid)
            if cur and procs and rng.random() < 0.6:     # often the first process is the reservation owner's
                owner_id = sql.row(cur[0])[4]
                if owner_id % 7:
                    procs[0]['owner'] = 'user%d' % owner_id
            if procs and rng.random() < 0.15:
                procs[-1]['owner'] = 'gdm'
            gpu = {'name': 'B200', 'index': g}
            roll = rng.random()
            if roll < 0.85:
                gpu['processes'] = procs
            elif roll < 0.95:
                gpu['processes'] = None                  # nvidia-smi pmon unsupported
            gpus[uuid] = gpu                              # (else: no 'processes' key at all)
        infra['host%d' % h] = {'GPU': gpus}
    infra['unreachable'] = {}
    infra['no_gpu_data'] = {'GPU': None}
    host_names = ['host%d' % h for h in range(hosts)]
    live_users = [r[0] for r in users]
    jobs = []
    for j in range(36):
        uid = None if rng.random() < 0.08 else int(live_users[int(rng.integers(0, len(live_users)))])
        tasks = [[host_names[int(rng.integers(0, hosts))], None if rng.random() < 0.06 else int(rng.integers(0, per_host + 3))]
                 for _ in range(int(rng.integers(1, 4)))]
        jobs.append({'id': j, 'user_id': uid, 'tasks': tasks})
    sql.close()
    return {'seed': seed, 'now_us': now_us, 'rows': rows, 'users': users, 'infrastructure': infra, 'jobs': jobs}


class _Task:
    def __init__(self, hostname, gpu_id):
        self.hostname, self.gpu_id = hostname, gpu_id


class _Job:
    def __init__(self, jid, user, tasks):
        self.id, self.user, self.tasks = jid, user, tasks


def _jsonable(v):
    if isinstance(v, dt.datetime):
        return v.strftime('%Y-%m-%dT%H:%M:%S.%f')
    if isinstance(v, (set, frozenset)):
        return sorted(v)
    if isinstance(v, dict):
        return {k: _jsonable(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return [_jsonable(x) for x in v]
    return v


def run_world(ref, w):
    _World.sql = SqlCalendar([tuple(r) for r in w['rows']])
    _World.users = {u[0]: User(*u) for u in w['users']}
    _World.now_us = w['now_us']
    im = ref['im'].InfrastructureManager({})
    im._infrastructure = w['infrastructure']
    out = {'all_nodes_with_gpu_processes': _jsonable(im.all_nodes_with_gpu_processes())}

    class Handler:
        def __init__(self):
            self.seen = []

        def trigger_action(self, violation_data):
            d = dict(violation_data)
            d.pop('SSH_CONNECTIONS', None)                  # the SSH plane is out of scope
            self.seen.append(_jsonable(d))

    class Connections:
        def single_connection(self, hostname):
            return hostname

    out['protection'] = {}
    for strict in (False, True):
        handler = Handler()
        svc = ref['protection'].ProtectionService([handler], interval=0.0, strict_reservations=strict)
        svc.infrastructure_manager = im                     # what inject() does (ProtectionService.py:33-37)
        svc.connection_manager = Connections()
        svc.do_run()
        out['protection']['strict' if strict else 'lenient'] = handler.seen
    jsvc = ref['jobs'].JobSchedulingService(interval=30.0, stop_attempts_after=5.0)
    hosts = im.all_nodes_with_gpu_processes()
    slots = jsvc.check_current_gpu_slots(hosts)
    out['gpu_slots'] = slots
    jobs = [_Job(j['id'], _World.users.get(j['user_id']) if j['user_id'] is not None else None,
                 [_Task(*t) for t in j['tasks']]) for j in w['jobs']]
    out['interferes'] = []
    for allow_own in (True, False):
        for minutes in (0, 30, 2880):
            flags = [bool(ref['jobs'].JobSchedulingService.interferes_with_reservations(
                job, hosts, dt.timedelta(minutes=minutes), allow_own)) for job in jobs]
            out['interferes'].append({'allow_own': allow_own, 'period_minutes': minutes, 'flags': flags})
    sched = ref['scheduling'].GreedyScheduler().schedule_jobs(jobs, slots)
    out['scheduled_job_ids'] = [j.id for j in sched]
    out['assigned_gpu_uids'] = [[ref['scheduling'].Scheduler.get_assigned_gpu_uid(t, hosts) for t in j.tasks] for j in jobs]
    _World.sql.close()
    return out


# ---------------------------------------------------------------------------------------------------------------------
# the usage logger, tick by tick, on real JSON files (row A end to end)
# ---------------------------------------------------------------------------------------------------------------------
def make_usage_world(seed, n_gpus=3, ticks=90):
    rng = np.random.default_rng(3000 + seed)
    t0 = synth.T0_US + seed * synth.DAY_US
    tick_us = 600_000_000                 # 10 min: every reservation lasts >= 30 min (Reservation.py:39)
    uuids = [synth.uuid_string(100 + g) for g in range(n_gpus)] + [synth.uuid_string(900)]     # the last one never reports
    rows, rid = [], 1
    for g, uuid in enumerate(uuids):
        t = t0 - 5 * tick_us
        while t < t0 + (ticks - 5) * tick_us:
            t += int(rng.integers(0, 8)) * tick_us + int(rng.integers(0, 2)) * int(rng.integers(0, tick_us))
            length = int(rng.integers(3, 25)) * tick_us + int(rng.integers(0, 2)) * int(rng.integers(0, tick_us))
            canc = [None, 0, 1][int(rng.choice(3, p=[0.6, 0.25, 0.15]))]
            rows.append([rid, uuid, t, t + length, int(rng.integers(1, 6)), canc])
            rid += 1
            if canc != 1:
                t += length if rng.random() < 0.3 else length + int(rng.integers(0, 3)) * tick_us   # 30 %: touching
    rows.append([rid, uuids[0], t0 + 10_000 * tick_us, t0 + 10_030 * tick_us, 1, None])            # never current: no log file
    order = rng.permutation(len(rows))
    rows = [rows[i] for i in order]
    for k, r in enumerate(rows):
        r[0] = k + 1                                                                                # ids in shuffled order
    tick_list = []
    for k in range(ticks):
        now = t0 + k * tick_us + int(rng.integers(0, 20_000_000))
        gpus = {}
        for g, uuid in enumerate(uuids[:-1]):
            if rng.random() < 0.04:
                continue                                     # GPU missing from this tick's infrastructure
            gu = None if rng.random() < 0.05 else int(rng.integers(0, 101))
            mu = None if rng.random() < 0.05 else int(rng.integers(0, 101))
            gpus[uuid] = {'name': 'B200', 'index': g, 'metrics': {'utilization': {'value': gu, 'unit': '%'},
                                                                  'mem_util': {'value': mu, 'unit': '%'}}}
        tick_list.append({'now_us': now, 'gpus': gpus})
    return {'seed': seed, 'rows': rows, 'ticks': tick_list, 'end_us': t0 + (ticks + 40) * tick_us}


def run_usage_world(ref, w):
    _World.sql = SqlCalendar([tuple(r) for r in w['rows']])
    _World.users = {}
    Reservation.saved = []
    svc_cls = ref['usage'].UsageLoggingService
    shutil.rmtree(str(svc_cls.log_dir), ignore_errors=True)
    svc = svc_cls(interval=0.0)
    im = ref['im'].InfrastructureManager({})
    svc.infrastructure_manager = im                          # what inject() does (UsageLoggingService.py:138-141)
    for tick in w['ticks'] + [{'now_us': w['end_us'], 'gpus': {}}]:
        _World.now_us = tick['now_us']
        im._infrastructure = {'host0': {'GPU': tick['gpus']}}
        svc.do_run()                                         # log_current_usage() then handle_expired_logs()
    _World.sql.close()
    shutil.rmtree(str(svc_cls.log_dir), ignore_errors=True)
    return sorted(Reservation.saved)


def main():
    import logging
    logging.disable(logging.CRITICAL)      # the reference logs every GPU it cannot find (UsageLoggingService.py:171)
    os.environ['TZ'] = 'UTC'            # utc2local (core/utils/time.py:5-9) reads the process time zone
    time.tzset()
    ref = load_reference()
    header = {
        'generated_by': 'tests/golden/make_reference_exec_vectors.py',
        'reference': 'kivicode/TensorHive-Fixed v1.1.0, source files executed unmodified under import stand-ins',
        'time_zone': 'UTC',
        'executed': [{'file': rel, 'sha256': hashlib.sha256(open(os.path.join(REF, rel), 'rb').read()).hexdigest(),
                      'bodies': bodies} for rel, bodies in EXECUTED.items()],
        'stand_ins': 'Reservation.current_events / upcoming_events_for_resource -> oracle.sql_ref.SqlCalendar (the '
                     'reference WHERE clauses through sqlite3); CONFIG.SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS = 30; '
                     'datetime.utcnow() frozen at now_us; gevent.sleep no-op; User objects identity-mapped per id',
    }
    worlds = []
    for seed in range(20):
        w = make_world(seed)
        w['expect'] = run_world(ref, w)
        worlds.append(w)
    rng = np.random.default_rng(7)
    avg_cases = [[]] + [[int(v) for v in rng.integers(0, 101, size=int(n))] for n in rng.integers(1, 400, size=30)]
    avg_cases += [[100] * 7 + [0], [0], [1, 2, 4]]
    avgs = [{'data': d, 'expect': ref['usage'].avg(d), 'expect_is_float': isinstance(ref['usage'].avg(d), float)}
            for d in avg_cases]
    usage_worlds = []
    for seed in range(6):
        w = make_usage_world(seed)
        w['expect_saved'] = run_usage_world(ref, w)
        assert len(w['expect_saved']) > 5
        usage_worlds.append(w)
    n_viol = sum(len(w['expect']['protection'][k]) for w in worlds for k in ('lenient', 'strict'))
    assert n_viol > 50, n_viol
    path = os.path.join(HERE, 'reference_exec_vectors.json')
    with open(path, 'w') as f:
        json.dump({'header': header, 'worlds': worlds, 'avg': avgs, 'usage_worlds': usage_worlds}, f, separators=(',', ':'))
    print('wrote %s: %d worlds, %d handler calls, %d avg cases, %d usage worlds (%d summaries), %d bytes' % (
        path, len(worlds), n_viol, len(avgs), len(usage_worlds), sum(len(w['expect_saved']) for w in usage_worlds),
        os.path.getsize(path)))


if __name__ == '__main__':
    main()
