"""Service-loop oracle (violation join, job availability, greedy scheduling, usage average).

TEST INFRASTRUCTURE, NOT PRODUCT CODE: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline leg may import this.

The reference's service code cannot be imported here (it needs SQLAlchemy, gevent, pssh, ...; see
SURVEY.md section 8c), so the loops are restated over plain dicts/objects with the SQL oracle
(``sql_ref.SqlCalendar``) standing in for the four ``Reservation.*`` queries.  Each function names the
reference lines it follows.  None of these paths has a reference test; they are pinned instead by vectors
obtained by EXECUTING the reference's own source files under import stand-ins
(``tests/golden/make_reference_exec_vectors.py`` -> ``tests/golden/reference_exec_vectors.json``), to which
``tests/test_oracle_exec_vectors.py`` holds every function below.
"""
from __future__ import annotations

import datetime as _dt
from typing import Dict, List, Optional

from .sql_ref import SqlCalendar, EPOCH

# tensorhive/core/managers/InfrastructureManager.py:69-78
IGNORED_COMMANDS = ('Xorg', '/usr/lib/xorg/Xorg', '/usr/bin/X', 'X', 'gnome-shell', '-')
# tensorhive/config.py:259 (fallback that always wins)
FREE_MINS = 30


class UserRow:
    def __init__(self, uid: int, username: str, email: str = None):
        self.id, self.username, self.email = uid, username, email or (username + '@example.org')


class ResRow:
    """What ORM hydration hands the service loops: a reservation with its ``user`` object."""

    def __init__(self, rid, resource_id, start_us, end_us, user: Optional[UserRow]):
        self.id, self.resource_id, self.user = rid, resource_id, user
        self.start = EPOCH + _dt.timedelta(microseconds=start_us)
        self.end = EPOCH + _dt.timedelta(microseconds=end_us)
        self.start_us, self.end_us = start_us, end_us


class World:
    """SQL calendar + users table; hydrates ids into ResRow objects."""

    def __init__(self, cal: SqlCalendar, users: Dict[int, UserRow]):
        self.cal, self.users = cal, users

    def _hydrate(self, ids: List[int]) -> List[ResRow]:
        out = []
        for rid in ids:
            (i, res, s, e, uid, _c) = self.cal.row(rid)
            out.append(ResRow(i, res, s, e, self.users.get(uid)))
        return out

    def current_events(self, now_us, resource_id=None):
        return self._hydrate(self.cal.current_events(now_us, resource_id))

    def upcoming(self, resource_id, now_us, period_us):
        return self._hydrate(self.cal.upcoming_events_for_resource(resource_id, now_us, period_us))


# -- InfrastructureManager.node_gpu_processes / all_nodes_with_gpu_processes (:22-63) ------------
def all_nodes_with_gpu_processes(infrastructure: Dict) -> Dict[str, Dict]:
    out = {}
    for host in infrastructure:
        gpus = infrastructure.get(host, {}).get('GPU')
        if gpus is None:
            out[host] = {}
            continue
        per_gpu = {}
        for uuid, data in gpus.items():
            if 'processes' in data:
                procs = data['processes']
                if procs is not None:
                    per_gpu[uuid] = [p for p in procs if p['command'] not in IGNORED_COMMANDS]
                else:
                    per_gpu[uuid] = []
        out[host] = per_gpu
    return out


# -- ProtectionService.do_run join (:84-102) + store_violation (:55-77) ---------------------------
def protection_violations(world: World, infrastructure: Dict, now_us: int, strict: bool) -> Dict[str, Dict[str, Dict]]:
    """Returns ``{hostname: violations}`` where ``violations`` is the per-host dict the reference builds
    before calling the handlers: intruder -> {INTRUDER_USERNAME, RESERVATIONS[...], VIOLATION_PIDS}.
    RESERVATIONS entries carry the fields of ProtectionService.py:59-67 that do not depend on the SSH
    plane or local time zone (OWNER_USERNAME, OWNER_EMAIL, END as naive-UTC datetime, GPU_UUID,
    HOSTNAME) plus RESERVATION_ID for the bit-exact comparison."""
    current = all_nodes_with_gpu_processes(infrastructure)
    result = {}
    for hostname in current:
        violations: Dict[str, Dict] = {}
        for gpu_id in current[hostname]:
            processes = current[hostname][gpu_id]
            if strict or (processes is not None and len(processes)):
                cur = world.current_events(now_us, gpu_id)
                reservation = None
                if len(cur):
                    reservation = cur[0]
                    if hostname is None or reservation.user is None:
                        continue
                    for process in processes:
                        if process['owner'] != reservation.user.username:
                            _store(violations, process, hostname, reservation, gpu_id)
                elif strict:
                    for process in processes:
                        _store(violations, process, hostname, reservation, gpu_id)
        result[hostname] = violations
    return result


def _store(storage, process, hostname, reservation, gpu_id):
    intruder = process['owner']
    data = {
        'OWNER_USERNAME': reservation.user.username if reservation else None,
        'OWNER_EMAIL': reservation.user.email if reservation else None,
        'END': reservation.end if reservation else None,
        'GPU_UUID': gpu_id,
        'HOSTNAME': hostname,
        'RESERVATION_ID': reservation.id if reservation else None,
    }
    if intruder not in storage:
        storage[intruder] = {
            'INTRUDER_USERNAME': intruder,
            'RESERVATIONS': [data],
            'VIOLATION_PIDS': {hostname: set([process['pid']])},
        }
    else:
        storage[intruder]['RESERVATIONS'].append(data)
        storage[intruder]['VIOLATION_PIDS'][hostname].add(process['pid'])


# -- JobSchedulingService.check_current_gpu_slots (:76-104) ---------------------------------------
def check_current_gpu_slots(world: World, hosts_with_gpu_occupation: Dict, now_us: int,
                            period_us: int = FREE_MINS * 60 * 1_000_000):
    now = EPOCH + _dt.timedelta(microseconds=now_us)
    ret = {}
    for host in hosts_with_gpu_occupation:
        ret[host] = {}
        for gpu_id in hosts_with_gpu_occupation[host]:
            running = hosts_with_gpu_occupation[host][gpu_id]
            if running and any('tensorhive_task' in p.get('command') and p.get('owner') != 'gdm' for p in running):
                ret[host][gpu_id] = 0
            else:
                near = world.upcoming(gpu_id, now_us, period_us)
                if len(near):
                    nearest = near[0]
                    if nearest.start > now:
                        ret[host][gpu_id] = (nearest.start - now).total_seconds() / 60
                    else:
                        ret[host][gpu_id] = 0
                else:
                    ret[host][gpu_id] = None
    return ret


# -- Scheduler.get_assigned_gpu_uid (scheduling.py:21-26) ------------------------------------------
def get_assigned_gpu_uid(task, hardware_map: Dict[str, Dict]):
    gpu_ids = list(hardware_map[task.hostname].keys())
    if task.gpu_id is None or task.gpu_id >= len(gpu_ids):
        return None
    return gpu_ids[task.gpu_id]


# -- JobSchedulingService.interferes_with_reservations (:117-132) ---------------------------------
def interferes_with_reservations(world: World, job, hosts: Dict[str, Dict], now_us: int,
                                 period_us: int = 0, allow_own: bool = True) -> bool:
    for task in job.tasks:
        gpu_id = get_assigned_gpu_uid(task, hosts)
        ups = world.upcoming(gpu_id, now_us, period_us)
        if allow_own:
            for r in ups:
                if r.user is not job.user:
                    return True
        elif len(ups):
            return True
    return False


# -- GreedyScheduler.schedule_jobs (scheduling.py:29-62) --------------------------------------------
def greedy_schedule(world: World, jobs_to_hardware, hardware_to_slots, now_us: int):
    period_us = FREE_MINS * 60 * 1_000_000
    scheduled, taken = [], []
    for job in jobs_to_hardware:
        ok_tasks = 0
        for task in job.tasks:
            uid = get_assigned_gpu_uid(task, hardware_to_slots)
            if (task.hostname, uid) in taken:
                break
            if not uid:
                ok_tasks += 1
                break
            slot = hardware_to_slots[task.hostname][uid]
            if slot is not None:
                own = [r for r in world.upcoming(uid, now_us, period_us) if r.user == job.user]
                if len(own):
                    slot = None
            if slot is None or slot >= FREE_MINS:
                ok_tasks += 1
        if ok_tasks == len(job.tasks):
            scheduled.append(job)
            taken.extend([(t.hostname, get_assigned_gpu_uid(t, hardware_to_slots)) for t in job.tasks])
    return scheduled


# -- avg (UsageLoggingService.py:25-30) ---------------------------------------------------------------
def avg(data):
    try:
        return sum(data) // len(data)
    except ZeroDivisionError:
        return float(-1)


# -- UsageLoggingService.log_current_usage (:159-171) + Log.updated_log (:85-105) + handle_expired_logs (:196-229) ------
def usage_logger_replay(cal: SqlCalendar, ticks, end_us: int):
    """The usage logger tick by tick over in-memory lists instead of JSON files.  ``ticks`` = [{'now_us', 'gpus':
    {uuid: {'metrics': {'utilization': {'value'}, 'mem_util': {'value'}}}}}], then one final tick at ``end_us`` with no
    GPUs.  Per tick (do_run :143-157): every reservation in current_events() whose GPU is present in the infrastructure
    (extract_specific_gpu_data :231-240 raises otherwise, caught at :170) gets a log; a sample is appended only when
    BOTH metrics are not None (:95); then every log whose reservation has ended (``end < now``, :207) is summarised with
    ``avg`` and dropped.  Returns sorted [[id, gpu_util_avg, mem_util_avg], ...] in the order ``save()`` would see them."""
    logs: Dict[int, Dict[str, list]] = {}
    saved = []
    for tick in list(ticks) + [{'now_us': end_us, 'gpus': {}}]:
        now = tick['now_us']
        for rid in cal.current_events(now):
            (_i, uuid, _s, _e, _u, _c) = cal.row(rid)
            gpu = tick['gpus'].get(uuid)
            if not gpu:
                continue
            log = logs.setdefault(rid, {'g': [], 'm': []})
            g, m = gpu['metrics']['utilization']['value'], gpu['metrics']['mem_util']['value']
            if g is not None and m is not None:
                log['g'].append(g)
                log['m'].append(m)
        for rid in sorted(logs):
            (_i, _uu, _s, e, _u, _c) = cal.row(rid)
            if e < now:
                saved.append([rid, avg(logs[rid]['g']), avg(logs[rid]['m'])])
                del logs[rid]
    return sorted(saved)


class TaskRow:
    def __init__(self, hostname, gpu_id):
        self.hostname, self.gpu_id = hostname, gpu_id


class JobRow:
    def __init__(self, jid, user: UserRow, tasks: List[TaskRow]):
        self.id, self.user, self.tasks = jid, user, tasks
