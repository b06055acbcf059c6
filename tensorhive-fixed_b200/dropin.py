"""The reference's call surface for the hot path, backed by the device snapshot.

Same names, argument meaning, return types and error behaviour as the reference:

  Reservation.would_interfere / current_events / upcoming_events_for_resource /
  filter_by_uuids_and_time_range              tensorhive/models/Reservation.py:89-148
  Reservation.check_assertions / save           Reservation.py:42-55, tensorhive/models/CRUDModel.py:21-38
  avg                                           tensorhive/core/services/UsageLoggingService.py:25-30
  ProtectionService.do_run / store_violation    tensorhive/core/services/ProtectionService.py:55-120
  JobSchedulingService.check_current_gpu_slots / interferes_with_reservations
                                                tensorhive/core/services/JobSchedulingService.py:76-132
  Scheduler.get_assigned_gpu_uid, GreedyScheduler.schedule_jobs
                                                tensorhive/core/scheduling.py:19-62
  ReservationVerifier.update_user_reservations_statuses (the would_interfere half, batched)
                                                tensorhive/core/utils/ReservationVerifier.py:91-109

The reference reaches its table through a module-global ``db_session``; here the module-global is the
bound :class:`~.calendar.Calendar` (``bind(calendar)``).  When the reference package is importable the same
functions can be assigned onto its classes; in this repo they are exercised on plain
row objects.  All predicate arithmetic happens in libth_b200.so -- this file only converts, orders and
re-materialises rows by id.
"""
from __future__ import annotations

import datetime as _dt
import logging
from datetime import timedelta
from typing import Dict, List, Optional

import numpy as np

from . import api
from .api import MODE_STRICT, MODE_INCLUSIVE, MODE_UPCOMING, OWNER_ANY, OWNER_OTHER, OWNER_SAME
from .calendar import Calendar, dt_to_us, us_to_dt

log = logging.getLogger(__name__)

_calendar: Optional[Calendar] = None

# tensorhive/config.py:259 (the INI section name never matches, so this fallback is what runs)
SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS = 30
# tensorhive/core/managers/InfrastructureManager.py:69-78
IGNORED_PROCESSES = ['Xorg', '/usr/lib/xorg/Xorg', '/usr/bin/X', 'X', 'gnome-shell', '-']


def bind(calendar: Calendar) -> None:
    """Point the module at the live calendar (the analogue of configuring ``db_session``)."""
    global _calendar
    _calendar = calendar


def _cal() -> Calendar:
    if _calendar is None:
        raise RuntimeError('dropin.bind(calendar) has not been called')
    return _calendar


def _period_us(period: timedelta) -> int:
    return (period.days * 86400 + period.seconds) * 1_000_000 + period.microseconds


# =========================================================================================================
# models/Reservation.py
# =========================================================================================================
class Reservation:
    """Row object with the reference model's attributes and hot-path methods."""

    _min_reservation_time = timedelta(minutes=30)   # Reservation.py:39
    _max_reservation_time = timedelta(days=8)       # Reservation.py:40

    def __init__(self, user_id=None, title='', description='', resource_id=None, start=None, end=None,
                 is_cancelled=None, id=None):
        self.id = id
        self.user_id = user_id
        self.title = title
        self.description = description
        self.resource_id = resource_id
        self.start = start
        self.end = end
        self._is_cancelled = is_cancelled
        self.gpu_util_avg = None
        self.mem_util_avg = None
        self.created_at = None

    # -- plain attributes mirroring the hybrid properties ------------------------------------------------
    # Reservation.py:61-79 + utils/DateUtils.py:9,47-60: a str is parsed with the API's input format (ValueError on a
    # malformed one, as in the reference), a datetime is kept, anything else becomes None (and fails check_assertions)
    @staticmethod
    def _try_parse(value):
        if isinstance(value, str):
            return _dt.datetime.strptime(value, '%Y-%m-%dT%H:%M:%S.%fZ')
        if isinstance(value, _dt.datetime):
            return value
        return None

    @property
    def start(self):
        return self._start

    @start.setter
    def start(self, value):
        self._start = self._try_parse(value)

    @property
    def end(self):
        return self._end

    @end.setter
    def end(self, value):
        self._end = self._try_parse(value)

    @property
    def is_cancelled(self):
        return self._is_cancelled is not None and self._is_cancelled

    @is_cancelled.setter
    def is_cancelled(self, value):
        self._is_cancelled = value

    @property
    def duration(self):
        return self.end - self.start

    @property
    def user(self):
        return _cal().users.get(self.user_id)

    # -- Reservation.py:17-18,165-168 + CRUDModel.py:72-94: camelCased public attributes, datetimes as strings ------
    __public__ = ['id', 'title', 'description', 'resource_id', 'user_id', 'gpu_util_avg', 'mem_util_avg', 'start', 'end',
                  'created_at', 'is_cancelled']

    def as_dict(self, include_private=False):
        def camel(name):
            head, *rest = name.split('_')
            return head + ''.join(w.capitalize() for w in rest)
        ret = {}
        for c in self.__public__:
            v = getattr(self, c, None)
            ret[camel(c)] = (v.strftime('%Y-%m-%dT%H:%M:%S') + '+00:00') if type(v) is _dt.datetime else v
        ret['userName'] = self.user.username
        return ret

    def __repr__(self):
        return '<Reservation id=%s resource=%s %s..%s cancelled=%s>' % (self.id, self.resource_id, self.start,
                                                                         self.end, self._is_cancelled)

    # -- Reservation.py:42-55 --------------------------------------------------------------------------------
    def check_assertions(self):
        assert self.user_id, 'Reservation owner must be given!'
        assert self.resource_id, 'Reservation must be related with a resource!'
        assert self.start, 'Reservation start time is invalid!'
        assert self.end, 'Reservation end time is invalid!'
        assert self.duration >= self._min_reservation_time, 'Reservation duration is too short!'
        assert self.duration <= self._max_reservation_time, 'Reservation duration is too long!'
        assert 0 < len(self.title) < 60, 'Reservation title length has incorrect length!'
        assert len(self.description) < 200, 'Reservation description has incorrect length!'
        assert len(self.resource_id) == 40, 'Protected resource UUID has incorrect length!'
        collision = self.would_interfere()
        assert not collision, 'Reservation would interfere with some other reservation!'

    # -- Reservation.py:120-130 ----------------------------------------------------------------------------------
    def would_interfere(self) -> bool:
        bits = _cal().would_interfere_batch([self.resource_id], [self.start], [self.end], [self.id])
        return bool(bits[0])

    @staticmethod
    def would_interfere_many(reservations: List['Reservation']) -> List[bool]:
        """Batch form: one kernel launch for any number of candidates."""
        if not reservations:
            return []
        bits = _cal().would_interfere_batch([r.resource_id for r in reservations], [r.start for r in reservations],
                                            [r.end for r in reservations], [r.id for r in reservations])
        return [bool(b) for b in bits]

    # -- Reservation.py:89-104 ---------------------------------------------------------------------------------------
    @classmethod
    def current_events(cls, resource_id: str = None):
        cal = _cal()
        now = cal.now_us()
        uuids = [resource_id] if resource_id is not None else cal.all_resource_ids()
        if not uuids:
            return []
        k = len(uuids)
        _off, ids, canc, rows = cal.sets(uuids, [now] * k, [now] * k, MODE_INCLUSIVE)
        keep = np.asarray(ids)[np.asarray(canc) == 0]
        if resource_id is None:
            keep = np.sort(keep)        # one scan over the whole table returns rowid = ascending id order
        return [rows[int(i)] for i in keep]

    # -- Reservation.py:106-118 -------------------------------------------------------------------------------------
    @classmethod
    def upcoming_events_for_resource(cls, resource_id: str, period_after: timedelta) -> List['Reservation']:
        cal = _cal()
        now = cal.now_us()
        _off, ids, _canc, rows = cal.sets([resource_id], [now], [now + _period_us(period_after)], MODE_UPCOMING)
        return [rows[int(i)] for i in ids]

    # -- Reservation.py:132-148 ----------------------------------------------------------------------------------------
    @classmethod
    def filter_by_uuids_and_time_range(cls, uuids: List[str], start: _dt.datetime, end: _dt.datetime):
        assertion_failed_msg = 'Argument must be of type datetime.datetime!'
        assert isinstance(start, _dt.datetime), assertion_failed_msg
        assert isinstance(end, _dt.datetime), assertion_failed_msg
        cal = _cal()
        uu = list(dict.fromkeys(uuids))    # IN (...) is a set test
        if not uu:
            return []
        k = len(uu)
        _off, ids, _canc, rows = cal.sets(uu, [dt_to_us(start)] * k, [dt_to_us(end)] * k, MODE_INCLUSIVE)
        return [rows[int(i)] for i in np.sort(np.asarray(ids))]


    # -- CRUDModel.py:21-50: the write path ----------------------------------------------------------------------------
    def save(self):
        """``check_assertions()`` (which runs would_interfere on the device), then commit: the row goes to the snapshot as
        a one-row delta (``th_snapshot_apply_delta``), a new row first getting the id AUTOINCREMENT would give it.  The
        check and the commit happen under the calendar's writer lock, so two racing saves cannot both pass the check.
        AssertionError is logged and re-raised, as in the reference."""
        cal = _cal()
        with cal._writer:
            try:
                self.check_assertions()
            except AssertionError as e:
                log.error(e)
                raise
            if self.id is None:
                self.id = cal.next_id()
            if self.created_at is None:
                self.created_at = cal.clock()
            cal.apply_delta(upserts=[self])
        log.debug('Saved {}'.format(self))
        return self

    def destroy(self):
        if self.id is not None:
            _cal().apply_delta(deleted_ids=[self.id])
        log.debug('Deleted {}'.format(self))
        return self

    @classmethod
    def get(cls, id):
        """CRUDModel.get (:52-66): the row with this id or ``NoResultFound``."""
        try:
            return _cal().rows_for([id])[0]
        except KeyError:
            raise NoResultFound('There is no record {} with id={}!'.format(cls.__name__, id))


class NoResultFound(Exception):
    """Stands in for sqlalchemy.orm.exc.NoResultFound (what CRUDModel.get raises, CRUDModel.py:61-64)."""


# =========================================================================================================
# controllers/reservation.py:31-55 -- the calendar view; FullCalendar.vue polls it per view
# =========================================================================================================
def get_selected(resources_ids=None, start=None, end=None):
    """Same contract as the reference controller: all three arguments or 400; API-format strings; the rows of
    ``filter_by_uuids_and_time_range`` as dicts; ValueError / AssertionError -> 400, anything else -> 500."""
    if not (resources_ids and start and end):
        return {'msg': 'Bad Request'}, 400
    try:
        start_as_datetime = _dt.datetime.strptime(start, '%Y-%m-%dT%H:%M:%S.%fZ')
        ends_as_datetime = _dt.datetime.strptime(end, '%Y-%m-%dT%H:%M:%S.%fZ')
        matches = [m.as_dict() for m in
                   Reservation.filter_by_uuids_and_time_range(resources_ids, start_as_datetime, ends_as_datetime)]
    except (ValueError, AssertionError) as reason:
        return {'msg': '{}. {}'.format('Bad Request', reason)}, 400
    except Exception as e:                                         # noqa: BLE001 -- the reference answers 500 here
        return {'msg': 'Internal server error' + str(e)}, 500
    return matches, 200


# =========================================================================================================
# core/utils/ReservationVerifier.py:91-109 -- the would_interfere half of the bulk re-validation
# =========================================================================================================
def reservations_that_may_be_restored(reservations: List[Reservation], is_allowed) -> List[Reservation]:
    """For a user whose permissions widened: the cancelled, still-running reservations that
    ``is_reservation_allowed`` accepts and that would not interfere.  ``is_allowed(reservation)`` is the
    host-side restriction walk (out of scope); all would_interfere checks against the calendar go out as ONE batch.

    The reference restores and saves one reservation at a time, so a reservation restored earlier in the loop is a
    live row for the ones after it: two of the user's cancelled reservations that overlap each other on one GPU are
    not both restored.  That part is a sequential pass over the (few) candidates on the host."""
    now = _cal().clock()
    cand = [r for r in reservations if r.end > now and r.is_cancelled and is_allowed(r)]
    flags = Reservation.would_interfere_many(cand)
    restored: List[Reservation] = []
    for r, f in zip(cand, flags):
        if f:
            continue
        if any(o.resource_id == r.resource_id and r.start < o.end and r.end > o.start and o.id != r.id for o in restored):
            continue                       # Reservation.py:120-130 against a row this very loop has just revived
        restored.append(r)
    return restored


# =========================================================================================================
# core/services/UsageLoggingService.py:25-30
# =========================================================================================================
def avg(data) -> float:
    """``sum(data) // len(data)``, ``float(-1)`` when empty -- computed by the K4 kernel (one window)."""
    data = list(data)
    if not data:
        return float(-1)
    if any(isinstance(x, float) for x in data):
        # The reference also accepts floats (Union[int, float]); its logs only ever hold ints
        # (NvidiaSmiParser.py:84-85).  Float lists are outside the int64 device contract.
        raise TypeError('usage samples are integers in the device path')
    v = np.asarray(data, np.int64)
    n = len(v)
    ts = np.arange(n, dtype=np.int64)
    _sg, _sm, _c, ag, _am = api.usage_reduce(np.array([0, n], np.int64), ts, v, v, np.array([0], np.int64),
                                             np.array([0], np.int64), np.array([n], np.int64), device=_cal().device
                                             if _calendar is not None else 0)
    return int(ag[0])


class UsageLog:
    """Columnar replacement for the per-reservation JSON logs (UsageLoggingService.py:62-116): samples of
    all GPUs, grouped by uuid and ts-ascending.  ``averages`` gives what handle_expired_logs (:196-229)
    writes into gpu_util_avg / mem_util_avg for each expired reservation."""

    def __init__(self, uuid_offsets, ts, gpu, mem, uuid_codes: Dict[str, int]):
        self.offsets, self.ts, self.gpu, self.mem = uuid_offsets, ts, gpu, mem
        self.uuid_codes = uuid_codes

    def averages(self, reservations: List[Reservation]):
        if not reservations:
            return []
        wu = np.array([self.uuid_codes.get(r.resource_id, -1) for r in reservations], np.int64)
        ws = np.array([dt_to_us(r.start) for r in reservations], np.int64)
        we = np.array([dt_to_us(r.end) for r in reservations], np.int64)
        _sg, _sm, cnt, ag, am = api.usage_reduce(self.offsets, self.ts, self.gpu, self.mem, wu, ws, we)
        out = []
        for j in range(len(reservations)):
            # avg() returns float(-1) for an empty log, an int otherwise
            out.append((float(-1), float(-1)) if cnt[j] == 0 else (int(ag[j]), int(am[j])))
        return out


class UsageRing:
    """Append-only columnar sample store ON THE DEVICE, one growing region per GPU uuid: what
    the usage logger writes instead of re-reading and re-writing one JSON file per reservation on every tick
    (UsageLoggingService.py:85-116).  A sample is kept under the rule of ``Log.updated_log`` (:92-99): only when BOTH
    gpu_util and mem_util are not None.  Timestamps must not decrease per uuid (the service appends ``utcnow()``).
    Utilisation is stored as uint8 (percentages), timestamps as int64 microseconds.

    ``append`` / ``append_tick`` cost one small host -> device copy and one scatter kernel whatever the log's length
    (``th_ring_append``); ``averages`` runs K4's narrow path over the ring in place (``th_ring_reduce``) -- the log is
    never concatenated, never re-uploaded -- and returns what ``handle_expired_logs`` (:196-229) stores per reservation."""

    def __init__(self, device: int = 0, initial_capacity: int = 1024):
        import ctypes as C
        from . import native
        self._C, self._native = C, native
        self._lib = native.load()
        self.device = int(device)
        self._h = C.c_void_p(0)
        native.check(self._lib.th_ring_create(self.device, int(initial_capacity), C.byref(self._h)))
        self._codes: Dict[str, int] = {}

    def close(self):
        if self._h:
            self._lib.th_ring_destroy(self._h)
            self._h = self._C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        C = self._C
        v = [C.c_int64(0) for _ in range(6)]
        self._native.check(self._lib.th_ring_info(self._h, *[C.byref(x) for x in v]))
        keys = ('n_uuids', 'samples', 'capacity', 'device_bytes', 'regular_uuids', 'relayouts')
        return {k: x.value for k, x in zip(keys, v)}

    def __len__(self):
        return self.info()['samples']

    def code(self, resource_id: str) -> int:
        k = self._codes.get(resource_id)
        if k is None:
            k = self._codes[resource_id] = len(self._codes)
        return k

    def append_tick(self, samples) -> int:
        """One monitoring tick: ``samples`` = iterable of ``(resource_id, ts, gpu_util, mem_util)``.  Entries with an
        unsupported metric (None) are dropped; returns how many were stored.  One H2D copy + one kernel for the tick."""
        uu, ts, g, m = [], [], [], []
        for resource_id, t, gpu_util, mem_util in samples:
            if gpu_util is None or mem_util is None:
                continue
            assert 0 <= int(gpu_util) <= 255 and 0 <= int(mem_util) <= 255, 'utilisation is a percentage'
            uu.append(self.code(resource_id))
            ts.append(dt_to_us(t) if isinstance(t, _dt.datetime) else int(t))
            g.append(int(gpu_util))
            m.append(int(mem_util))
        if not uu:
            return 0
        a = [np.array(uu, np.int64), np.array(ts, np.int64), np.array(g, np.uint8), np.array(m, np.uint8)]
        rc = self._lib.th_ring_append(self._h, len(uu), *[self._C.c_void_p(x.ctypes.data) for x in a], None)
        if rc == -1 and 'backwards' in self._native.last_error():
            raise AssertionError('samples of one GPU must be appended in time order')
        self._native.check(rc)
        return len(uu)

    def append(self, resource_id: str, ts, gpu_util, mem_util) -> bool:
        """One sample of one GPU; returns False (and stores nothing) when either metric is unsupported."""
        return self.append_tick([(resource_id, ts, gpu_util, mem_util)]) == 1

    def reduce(self, win_uuid, win_start, win_end):
        """(sum_gpu, sum_mem, count, avg_gpu, avg_mem) per window: numpy arrays in / out, or CUDA tensors in / out."""
        a = api._Args(win_uuid, win_start, win_end)
        w = int(len(win_uuid))
        outs = [a.out(w) for _ in range(5)]
        self._native.check(self._lib.th_ring_reduce(self._h, w, a.inp(win_uuid), a.inp(win_start), a.inp(win_end),
                                                    outs[0][1], outs[1][1], outs[2][1], outs[3][1], outs[4][1], a.mem,
                                                    a.stream()))
        return tuple(o[0] for o in outs)

    def averages(self, reservations: List[Reservation]):
        if not reservations:
            return []
        wu = np.array([self._codes.get(r.resource_id, -1) for r in reservations], np.int64)
        ws = np.array([dt_to_us(r.start) for r in reservations], np.int64)
        we = np.array([dt_to_us(r.end) for r in reservations], np.int64)
        _sg, _sm, cnt, ag, am = self.reduce(wu, ws, we)
        return [(float(-1), float(-1)) if cnt[j] == 0 else (int(ag[j]), int(am[j])) for j in range(len(reservations))]


class UsageLoggingService:
    """UsageLoggingService.do_run (tensorhive/core/services/UsageLoggingService.py:143-229) on the sample ring.

    ``log_current_usage`` (:159-171): for every reservation in ``current_events()`` whose GPU the infrastructure
    reports (``extract_specific_gpu_data`` :231-240 raises otherwise; the reference logs the error and goes on), the
    tick's utilisation pair is stored -- once per GPU and tick in the ring, instead of once per reservation in a JSON
    file that is re-read and re-written -- under the rule of ``Log.updated_log`` (:92-99: both metrics not None).
    ``handle_expired_logs`` (:196-229): every reservation that has a log and has ended (``end < now``) gets
    ``gpu_util_avg`` / ``mem_util_avg`` = ``avg`` of its samples (ONE K4 launch for all of them: the samples of its GPU
    with ``start <= ts <= end``) and is saved; exceptions are logged, not raised, as there."""

    def __init__(self, interval=0.0, infrastructure_manager=None, ring: 'UsageRing' = None):
        self.interval = interval
        self.infrastructure_manager = infrastructure_manager
        self.ring = ring if ring is not None else UsageRing(device=_cal().device if _calendar is not None else 0)
        self._open: Dict[int, object] = {}          # reservation id -> row, the analogue of "<id>.json exists"

    @staticmethod
    def extract_specific_gpu_data(uuid: str, infrastructure: Dict) -> Dict:
        assert isinstance(infrastructure, dict)
        assert isinstance(uuid, str) and len(uuid) == 40
        for hostname in infrastructure.keys():
            gpu_data = infrastructure[hostname].get('GPU').get(uuid)
            if gpu_data:
                return gpu_data
        raise KeyError(uuid + ' has not been found!')

    def log_current_usage(self):
        cal = _cal()
        now = cal.clock()
        infrastructure = self.infrastructure_manager.infrastructure
        seen = {}
        for reservation in Reservation.current_events():
            uuid = reservation.resource_id
            if uuid not in seen:
                try:
                    gpu_data = self.extract_specific_gpu_data(uuid=uuid, infrastructure=infrastructure)
                    self.ring.append(uuid, now, gpu_data['metrics']['utilization']['value'],
                                     gpu_data['metrics']['mem_util']['value'])
                    seen[uuid] = True
                except Exception as e:                                   # noqa: BLE001 -- :170-171
                    log.error(e)
                    seen[uuid] = False
            if seen[uuid]:
                self._open[reservation.id] = reservation                 # "<id>.json" exists from now on (:108-110)

    def handle_expired_logs(self):
        time_now = _cal().clock()
        expired = []
        for rid in sorted(self._open):
            try:
                reservation = Reservation.get(id=rid)
            except NoResultFound:
                del self._open[rid]                                      # log of a vanished reservation: cleaned up (:224-226)
                continue
            if reservation.end < time_now:
                expired.append(reservation)
        if not expired:
            return []
        averages = self.ring.averages(expired)
        saved = []
        for reservation, (ag, am) in zip(expired, averages):
            try:
                reservation.gpu_util_avg = ag
                reservation.mem_util_avg = am
                reservation.save()
                saved.append(reservation)
                del self._open[reservation.id]                           # the log is cleaned up only after a save (:218-223)
            except Exception as e:                                       # noqa: BLE001 -- :227-228
                log.debug(e)
        return saved

    def do_run(self):
        self.log_current_usage()
        return self.handle_expired_logs()


# =========================================================================================================
# core/managers/InfrastructureManager.py:22-63
# =========================================================================================================
def all_nodes_with_gpu_processes(infrastructure: Dict) -> Dict[str, Dict]:
    out = {}
    for node in infrastructure:
        if infrastructure.get(node, {}).get('GPU') is None:
            out[node] = {}
            continue
        node_processes = {}
        for uuid, gpu_data in infrastructure[node]['GPU'].items():
            if 'processes' in gpu_data:
                procs = gpu_data['processes']
                node_processes[uuid] = ([p for p in procs if p['command'] not in IGNORED_PROCESSES]
                                        if procs is not None else [])
        out[node] = node_processes
    return out


# =========================================================================================================
# core/services/ProtectionService.py
# =========================================================================================================
class ProtectionService:
    """Same constructor and ``do_run`` contract as the reference; the per-GPU ``current_events`` scans and the
    owner comparison are one K3 + K5 launch over the flattened process table."""

    def __init__(self, handlers, interval=0.0, strict_reservations=False, infrastructure_manager=None,
                 connection_manager=None):
        self.interval = interval
        self.violation_handlers = handlers
        self.strict_reservations = strict_reservations
        self.infrastructure_manager = infrastructure_manager
        self.connection_manager = connection_manager

    def gpu_attr(self, hostname: str, uuid: str, attribute='name') -> str:
        infrastructure = self.infrastructure_manager.infrastructure if self.infrastructure_manager else {}
        return infrastructure.get(hostname, {}).get('GPU', {}).get(uuid, {}).get(attribute, '<not available>')

    def collect_violations(self, current_infrastructure: Dict[str, Dict]) -> Dict[str, Dict[str, Dict]]:
        """{hostname: {intruder: violation_data}} exactly as the nested loops of do_run (:85-102) build it."""
        cal = _cal()
        names = cal.usernames()
        unknown: Dict[object, int] = {}
        result = {hostname: {} for hostname in current_infrastructure}
        with cal.reading() as (snap, uuids, by_id):       # table, dictionary and rows of ONE version
            # flatten in the reference's iteration order
            where, pu, po = [], [], []
            for hostname in current_infrastructure:
                for gpu_id in current_infrastructure[hostname]:
                    processes = current_infrastructure[hostname][gpu_id]
                    if self.strict_reservations or (processes is not None and len(processes)):
                        code = uuids.lookup(gpu_id)
                        for process in processes:
                            owner = process['owner']
                            if owner in names and owner is not None:
                                oid = names[owner]
                            else:
                                oid = unknown.setdefault(owner, -2 - len(unknown))
                            where.append((hostname, gpu_id, process))
                            pu.append(code)
                            po.append(oid)
            if not where:
                return result
            flags, res_ids = snap.violation_join(cal.now_us(), self.strict_reservations, np.array(pu, np.int64),
                                                 np.array(po, np.int64))
            for k, (hostname, gpu_id, process) in enumerate(where):
                if not flags[k] or hostname is None:
                    continue
                reservation = by_id[int(res_ids[k])] if res_ids[k] >= 0 else None
                self.store_violation(result[hostname], process, hostname, reservation, gpu_id)
        return result

    def store_violation(self, storage, process, hostname, reservation, gpu_id):
        intruder = process['owner']
        user = cal_user(reservation) if reservation else None
        reservation_data = {
            'OWNER_USERNAME': user.username if reservation else None,
            'OWNER_EMAIL': getattr(user, 'email', None) if reservation else None,
            'END': utc2local(_as_dt(reservation.end)) if reservation else None,
            'GPU_UUID': gpu_id,
            'GPU_NAME': self.gpu_attr(hostname, gpu_id, attribute='name'),
            'GPU_ID': self.gpu_attr(hostname, gpu_id, attribute='index'),
            'HOSTNAME': hostname,
            'RESERVATION_ID': reservation.id if reservation else None,
        }
        if intruder not in storage:
            storage[intruder] = {
                'INTRUDER_USERNAME': intruder,
                'RESERVATIONS': [reservation_data],
                'VIOLATION_PIDS': {hostname: set([process['pid']])},
            }
        else:
            storage[intruder]['RESERVATIONS'].append(reservation_data)
            storage[intruder]['VIOLATION_PIDS'][hostname].add(process['pid'])

    def do_run(self):
        current_infrastructure = self.infrastructure_manager.all_nodes_with_gpu_processes()
        per_host = self.collect_violations(current_infrastructure)
        for hostname in current_infrastructure:
            violations = per_host[hostname]
            for intruder in violations:
                violation_data = violations[intruder]
                reservations = violation_data['RESERVATIONS']
                hostnames = set([d['HOSTNAME'] for d in reservations])
                if self.connection_manager is not None:
                    violation_data['SSH_CONNECTIONS'] = {h: self.connection_manager.single_connection(h)
                                                         for h in hostnames}
                violation_data['GPUS'] = ',\n'.join(['{} - GPU{}: {}'.format(d['HOSTNAME'], d['GPU_ID'], d['GPU_NAME'])
                                                     for d in reservations])
                violation_data['OWNERS'] = ', '.join(['{} ({})'.format(d['OWNER_USERNAME'], d['OWNER_EMAIL'])
                                                      for d in reservations])
                for handler in self.violation_handlers:
                    try:
                        handler.trigger_action(violation_data)
                    except Exception as e:
                        log.warning('Error in violation handler: {}'.format(e))


def cal_user(reservation):
    return _cal().users.get(reservation.user_id)


def _as_dt(x):
    return x if isinstance(x, _dt.datetime) else us_to_dt(x)


def utc2local(utc: _dt.datetime) -> _dt.datetime:
    # tensorhive/core/utils/time.py:5-9
    import time
    epoch = time.mktime(utc.timetuple())
    # (the reference calls datetime.utcfromtimestamp, deprecated since Python 3.12: the same naive UTC value)
    offset = _dt.datetime.fromtimestamp(epoch) - _dt.datetime.fromtimestamp(epoch, _dt.timezone.utc).replace(tzinfo=None)
    return utc + offset


# =========================================================================================================
# core/scheduling.py + core/services/JobSchedulingService.py
# =========================================================================================================
class Scheduler:
    @staticmethod
    def get_assigned_gpu_uid(task, hardware_map: Dict[str, Dict]):
        gpu_ids = list(hardware_map[task.hostname].keys())
        if task.gpu_id is None or task.gpu_id >= len(gpu_ids):
            return None
        return gpu_ids[task.gpu_id]


def _owner_id(user) -> int:
    return -1 if user is None else int(getattr(user, 'id'))


class JobSchedulingService:
    def __init__(self, interval: float = 30.0, stop_attempts_after: float = 5.0):
        self.interval = interval
        self.stop_attempts_after = timedelta(minutes=stop_attempts_after)
        self.considered_future_period = timedelta(minutes=SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS)

    def check_current_gpu_slots(self, hosts_with_gpu_occupation: Dict[str, Dict]) -> Dict[str, Dict]:
        """JobSchedulingService.py:76-104: minutes until the next reservation per GPU (0 = occupied now,
        None = nothing upcoming).  One th_upcoming launch for every GPU that is not running a task."""
        cal = _cal()
        ret: Dict[str, Dict] = {}
        ask = []
        for host in hosts_with_gpu_occupation:
            ret[host] = {}
            for gpu_id in hosts_with_gpu_occupation[host]:
                running = hosts_with_gpu_occupation[host][gpu_id]
                if running and any('tensorhive_task' in proc.get('command') and proc.get('owner') != 'gdm'
                                   for proc in running):
                    ret[host][gpu_id] = 0
                else:
                    ret[host][gpu_id] = None
                    ask.append((host, gpu_id))
        if not ask:
            return ret
        now_dt = cal.clock()
        now = dt_to_us(now_dt)
        with cal.reading() as (snap, uuids, _):
            codes = np.array([uuids.lookup(g) for _h, g in ask], np.int64)
            _fid, fstart, _fown, count = snap.upcoming(now, _period_us(self.considered_future_period), codes)
        for k, (host, gpu_id) in enumerate(ask):
            if count[k] > 0:
                start = us_to_dt(int(fstart[k]))
                # the one floating-point operation of the path, done by the reference's own expression (:97-100)
                ret[host][gpu_id] = (start - now_dt).total_seconds() / 60 if start > now_dt else 0
        return ret

    @staticmethod
    def interferes_with_reservations(job, available_hosts_with_gpu_occupation: Dict[str, Dict],
                                     considered_future_period: timedelta = timedelta(0),
                                     allow_own: bool = True) -> bool:
        return JobSchedulingService.interferes_with_reservations_many(
            [job], available_hosts_with_gpu_occupation, considered_future_period, allow_own)[0]

    @staticmethod
    def interferes_with_reservations_many(jobs, hosts: Dict[str, Dict],
                                          considered_future_period: timedelta = timedelta(0),
                                          allow_own: bool = True) -> List[bool]:
        """JobSchedulingService.py:117-132 for a whole job list in one launch."""
        cal = _cal()
        out = [False] * len(jobs)
        with cal.reading() as (snap, uuids, _):
            tu, to, tj = [], [], []
            for j, job in enumerate(jobs):
                for task in job.tasks:
                    gpu_id = Scheduler.get_assigned_gpu_uid(task, hosts)
                    tu.append(uuids.lookup(gpu_id))
                    to.append(_owner_id(job.user))
                    tj.append(j)
            if not tu:
                return out
            flags = snap.upcoming_owner_flags(cal.now_us(), _period_us(considered_future_period), np.array(tu, np.int64),
                                              np.array(to, np.int64), OWNER_OTHER if allow_own else OWNER_ANY)
        for k, j in enumerate(tj):
            if flags[k]:
                out[j] = True
        return out


class GreedyScheduler(Scheduler):
    def schedule_jobs(self, jobs_to_hardware, hardware_to_slots) -> List:
        """scheduling.py:29-62.  The per-task "does the job's owner have an upcoming reservation here" test is
        one batched K5 launch; the taken-hosts pass is sequential by definition and stays on the host."""
        cal = _cal()
        jobs = list(jobs_to_hardware)
        tu, to, index = [], [], {}
        own = np.zeros(0, np.uint8)
        with cal.reading() as (snap, uuids, _):
            for j, job in enumerate(jobs):
                for t, task in enumerate(job.tasks):
                    uid = Scheduler.get_assigned_gpu_uid(task, hardware_to_slots)
                    index[(j, t)] = len(tu)
                    tu.append(uuids.lookup(uid))
                    to.append(_owner_id(job.user))
            if tu:
                own = snap.upcoming_owner_flags(cal.now_us(), SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS * 60 * 1_000_000,
                                                np.array(tu, np.int64), np.array(to, np.int64), OWNER_SAME)
        scheduled_jobs, taken_hosts = [], []
        for j, job in enumerate(jobs):
            scheduled_tasks = 0
            for t, task in enumerate(job.tasks):
                gpu_uid = Scheduler.get_assigned_gpu_uid(task, hardware_to_slots)
                if (task.hostname, gpu_uid) in taken_hosts:
                    break
                if not gpu_uid:
                    scheduled_tasks += 1
                    break
                slot = hardware_to_slots[task.hostname][gpu_uid]
                if slot is not None and own[index[(j, t)]]:
                    slot = None
                if slot is None or slot >= SCHEDULE_QUEUED_JOBS_WHEN_FREE_MINS:
                    scheduled_tasks += 1
            if scheduled_tasks == len(job.tasks):
                scheduled_jobs.append(job)
                taken_hosts.extend([(task.hostname, Scheduler.get_assigned_gpu_uid(task, hardware_to_slots))
                                    for task in job.tasks])
        return scheduled_jobs
