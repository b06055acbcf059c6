"""Writes tests/golden/reference_model_vectors.json by EXECUTING the reference's own model source for rows R1-R4 and the
write path -- not a restatement of it.  Run from the repo root, in the container that holds the reference:

    python tests/golden/make_reference_model_vectors.py

What runs here is the unmodified text of (paths relative to /root/reference)

  tensorhive/models/Reservation.py   would_interfere :120-130 (R1), filter_by_uuids_and_time_range :132-148 (R2),
                                     current_events :89-104 (R3), upcoming_events_for_resource :106-118 (R4),
                                     check_assertions :42-55, the is_cancelled / start / end / duration hybrids :57-86
  tensorhive/models/CRUDModel.py     save :21-39, destroy :41-50, get :52-65
  tensorhive/utils/DateUtils.py      try_parse_string (the start / end setters go through it)
  tensorhive/core/utils/ReservationVerifier.py   update_user_reservations_statuses :90-109, widening case (section 8f.2),
                                     with is_reservation_allowed -- the restriction walk, out of scope -- replaced by a fixed predicate
  tensorhive/controllers/reservation.py   get_selected :31-55, the calendar-view endpoint (section 8f.4), with the real
                                     response strings of controllers/responses.yml and Reservation.as_dict / CRUDModel.as_dict

loaded as modules under their real dotted names with `importlib`.  SQLAlchemy 1.3.0 is not installable in this image,
so `sys.modules['sqlalchemy']` is a stand-in for the handful of operators those methods use.  It evaluates the
expression objects the reference BUILDS, row by row, with the semantics SQLAlchemy + SQLite give them:

  * `Column` attributes on the class compare into expression nodes, also reflected (`current_time <= cls.end` calls
    `Column.__ge__`), `==` / `!=` against None mean IS NULL / IS NOT NULL (so `Reservation.id != self.id` keeps every row
    for a reservation that has no id yet), `.in_(list)`;
  * `and_` / `or_` / `not_` are SQL's three-valued AND / OR / NOT, `filter(a, b)` and chained `.filter()` are AND, a row is
    returned when the whole condition is TRUE;
  * `hybrid_property` calls the getter with the class (expression level) or the instance, as SQLAlchemy does;
  * `.all()` without ORDER BY yields rows in rowid (= primary key) order, the order of SQLite's table scan;
    `.order_by(col)` is a stable ascending sort on top of that, NULLs first;
  * DateTime values compare chronologically (SQLite stores them as '%Y-%m-%d %H:%M:%S.%f' text, which sorts the same way);
  * the session stand-in assigns the largest id ever used + 1 on add (sqlite_autoincrement) and keeps the table in memory.

`datetime.datetime.utcnow()` is frozen per probe.  Everything else -- which comparisons are made, strict or not, in which
order, what is filtered in Python afterwards, which assertion fires first -- is the reference's code.  Before anything is
written, the stand-in itself is checked: the reference's OWN tests for this path (all six functions of
tests/unit/models/test_reservation_model.py, loaded and called unmodified, `selfcheck`) must pass on it, and every answer
must equal the SQL oracle's (oracle/sql_ref.py, the WHERE clauses as text through sqlite3).  The GPU box has no /root/reference: inputs AND outputs are committed, and
tests/test_oracle_model_vectors.py (CPU: both oracles) and tests/test_gpu_model_vectors.py (the CUDA path and the drop-in)
are held to them.
"""
import datetime as dt
import hashlib
import importlib.util
import json
import logging
import operator
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get('TH_REFERENCE_ROOT', '/root/reference')
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

from oracle.sql_ref import SqlCalendar                               # noqa: E402
from tensorhive_fixed_b200 import synth                             # noqa: E402  (seeded world generator only)

EXECUTED = {
    'tensorhive/utils/DateUtils.py': ['DateUtils.try_parse_string :47-59', 'parse_string :13-26'],
    'tensorhive/models/CRUDModel.py': ['CRUDModel.save :21-39', 'destroy :41-50', 'get :52-65'],
    'tensorhive/core/utils/ReservationVerifier.py': ['ReservationVerifier.update_user_reservations_statuses :90-109'],
    'tensorhive/controllers/reservation.py': ['get_selected :31-55 (+ Reservation.as_dict :165-168, CRUDModel.as_dict :72-94)'],
    'tensorhive/models/Reservation.py': ['columns :14-37', 'check_assertions :42-55', 'duration / start / end hybrids :57-78',
                                         'is_cancelled :80-86', 'current_events :88-104',
                                         'upcoming_events_for_resource :106-118', 'would_interfere :120-130',
                                         'filter_by_uuids_and_time_range :132-148'],
}


# ---------------------------------------------------------------------------------------------------------------------
# stand-in for the SQLAlchemy operators the model uses
# ---------------------------------------------------------------------------------------------------------------------
class Expr:
    """A condition over one row; evaluates to True, False or None (SQL NULL)."""

    def __init__(self, fn):
        self.fn = fn

    def __call__(self, row):
        return self.fn(row)


def _value(x, row):
    return x.value_of(row) if isinstance(x, Column) else x


def _compare(op, a, b):
    def fn(row):
        x, y = _value(a, row), _value(b, row)
        if x is None or y is None:
            return None                                   # any comparison with NULL is NULL
        return bool(op(x, y))
    return Expr(fn)


class Column:
    """Class level: an expression node.  Instance level: the stored value (None when unset)."""

    def __init__(self, *args, **kwargs):
        self.attr = None

    def __set_name__(self, owner, name):
        self.attr = name

    def __get__(self, inst, owner):
        return self if inst is None else inst.__dict__.get(self.attr)

    def __set__(self, inst, value):
        inst.__dict__[self.attr] = value

    def value_of(self, row):
        return row.__dict__.get(self.attr)

    def __lt__(self, o): return _compare(operator.lt, self, o)
    def __le__(self, o): return _compare(operator.le, self, o)
    def __gt__(self, o): return _compare(operator.gt, self, o)
    def __ge__(self, o): return _compare(operator.ge, self, o)

    def __eq__(self, o):
        if o is None:
            return Expr(lambda row: self.value_of(row) is None)           # IS NULL
        return _compare(operator.eq, self, o)

    def __ne__(self, o):
        if o is None:
            return Expr(lambda row: self.value_of(row) is not None)       # IS NOT NULL
        return _compare(operator.ne, self, o)

    __hash__ = object.__hash__

    def in_(self, seq):
        vals = list(seq)
        return Expr(lambda row: None if self.value_of(row) is None else self.value_of(row) in vals)


def and_(*conds):
    def fn(row):
        out = True
        for c in conds:
            v = c(row)
            if v is False:
                return False
            if v is None:
                out = None
        return out
    return Expr(fn)


def or_(*conds):
    def fn(row):
        out = False
        for c in conds:
            v = c(row)
            if v is True:
                return True
            if v is None:
                out = None
        return out
    return Expr(fn)


def not_(cond):
    return Expr(lambda row: None if cond(row) is None else not cond(row))


class hybrid_property:
    def __init__(self, fget, fset=None):
        self.fget, self.fset = fget, fset

    def __get__(self, inst, owner):
        return self.fget(owner) if inst is None else self.fget(inst)

    def __set__(self, inst, value):
        if self.fset is None:
            raise AttributeError("can't set attribute")
        self.fset(inst, value)

    def setter(self, fset):
        return hybrid_property(self.fget, fset)


class NoResultFound(Exception):
    pass


class MultipleResultsFound(Exception):
    pass


class SQLAlchemyError(Exception):
    pass


class Table:
    """The in-memory `reservations` table of the current world; rows are instances of the reference's class."""
    rows = []
    seq = 0                        # sqlite_sequence: the largest id ever used (AUTOINCREMENT never hands one out twice)
    last_query_ids = None          # ids of the most recent .all(), in the order returned (for would_interfere's list)


class Query:
    def __init__(self, cls, conds=(), order=None):
        self.cls, self.conds, self.order = cls, tuple(conds), order

    def filter(self, *conds):
        return Query(self.cls, self.conds + tuple(conds), self.order)

    def filter_by(self, **kw):
        return self.filter(*[getattr(self.cls, k) == v for k, v in kw.items()])

    def order_by(self, col):
        return Query(self.cls, self.conds, col)

    def all(self):
        rows = sorted(Table.rows, key=lambda r: r.id)                      # rowid order
        rows = [r for r in rows if all(c(r) is True for c in self.conds)]
        if self.order is not None:
            rows.sort(key=lambda r: (self.order.value_of(r) is not None, self.order.value_of(r)))      # stable
        Table.last_query_ids = [r.id for r in rows]
        return rows

    def one(self):
        rows = self.all()
        if not rows:
            raise NoResultFound()
        if len(rows) > 1:
            raise MultipleResultsFound()
        return rows[0]


class _Session:
    def query(self, cls):
        return Query(cls)

    def add(self, obj):
        if obj.id is None:
            obj.id = max([Table.seq] + [r.id for r in Table.rows]) + 1     # Reservation.py:16 sqlite_autoincrement
        Table.seq = max(Table.seq, obj.id)
        if not any(r is obj for r in Table.rows):
            Table.rows.append(obj)

    def delete(self, obj):
        Table.rows[:] = [r for r in Table.rows if r is not obj]

    def commit(self):
        pass

    def rollback(self):
        pass


class _QueryProperty:
    def __get__(self, inst, owner):
        return Query(owner)


class Base:
    query = _QueryProperty()

    def __init__(self, **kwargs):                       # the declarative constructor: keyword -> attribute
        for k, v in kwargs.items():
            setattr(self, k, v)


class _FrozenMeta(type):
    def __instancecheck__(cls, inst):                   # the reference's `isinstance(x, datetime.datetime)` checks
        return isinstance(inst, dt.datetime)


class _FrozenDateTime(dt.datetime, metaclass=_FrozenMeta):
    now_us = 0

    @classmethod
    def utcnow(cls):
        return at(cls.now_us)


_EPOCH = _FrozenDateTime(1970, 1, 1)


def at(us):
    """A datetime `us` microseconds after the epoch (an instance of the frozen class, so isinstance checks hold)."""
    return _EPOCH + dt.timedelta(microseconds=int(us))


def us_of(d):
    delta = d - _EPOCH
    return (delta.days * 86400 + delta.seconds) * 1_000_000 + delta.microseconds


# ---------------------------------------------------------------------------------------------------------------------
# loading the reference's model under those stand-ins
# ---------------------------------------------------------------------------------------------------------------------
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


def _camelcase(string):
    """stringcase 1.2.0 `camelcase` (requirements: stringcase), restated: drop a leading separator, lower the first letter,
    upper-case the letter after each of - _ . and whitespace."""
    import re
    string = re.sub(r"^[\-_\.]", '', str(string))
    if not string:
        return string
    return string[0].lower() + re.sub(r"[\-_\.\s]([a-z])", lambda m: m.group(1).upper(), string[1:])


def load_reference():
    for pkg in ('tensorhive', 'tensorhive.models', 'tensorhive.utils'):
        _package(pkg)
    marker = lambda *a, **k: None                                  # column types / ForeignKey: declarations only
    _module('sqlalchemy', Column=Column, Boolean=marker, Integer=marker, String=marker, DateTime=marker,
            ForeignKey=marker, and_=and_, or_=or_, not_=not_, event=None)
    _package('sqlalchemy.ext')
    _module('sqlalchemy.ext.hybrid', hybrid_property=hybrid_property)
    _module('sqlalchemy.orm', relationship=lambda *a, **k: None, backref=lambda *a, **k: None)
    _module('sqlalchemy.orm.exc', NoResultFound=NoResultFound, MultipleResultsFound=MultipleResultsFound)
    _module('sqlalchemy.exc', SQLAlchemyError=SQLAlchemyError, IntegrityError=type('IntegrityError', (SQLAlchemyError,), {}))
    _module('stringcase', camelcase=_camelcase, snakecase=lambda s: s)
    _module('tensorhive.database', db_session=_Session(), Base=Base)
    ref = {}
    ref['dateutils'] = _load('tensorhive/utils/DateUtils.py')
    ref['crud'] = _load('tensorhive/models/CRUDModel.py')
    ref['model'] = _load('tensorhive/models/Reservation.py')
    _package('tensorhive.core')
    _package('tensorhive.core.utils')
    _module('tensorhive.models.Resource', Resource=type('Resource', (), {}))
    ref['verifier'] = _load('tensorhive/core/utils/ReservationVerifier.py')
    ref['verifier'].datetime = _FrozenDateTime                     # `datetime.utcnow()` at ReservationVerifier.py:100
    # `datetime.datetime.utcnow()` at Reservation.py:93,108; the isinstance checks at :141-142 see the same class
    ref['model'].datetime = types.SimpleNamespace(datetime=_FrozenDateTime, timedelta=dt.timedelta)
    import yaml
    with open(os.path.join(REF, 'tensorhive/controllers/responses.yml')) as f:
        responses = yaml.safe_load(f)                               # what tensorhive/config.py:177-180 loads into API.RESPONSES
    ident = lambda fn: fn
    _module('flask_jwt_extended', get_jwt_claims=None, get_jwt_identity=None, jwt_required=ident)
    _module('tensorhive.config', API=types.SimpleNamespace(RESPONSES=responses))
    _module('tensorhive.models.User', User=type('User', (), {}))
    _package('tensorhive.exceptions')
    _module('tensorhive.exceptions.ForbiddenException', ForbiddenException=type('ForbiddenException', (Exception,), {}))
    _package('tensorhive.controllers')
    ref['controller'] = _load('tensorhive/controllers/reservation.py')
    logging.getLogger('tensorhive').setLevel(logging.CRITICAL + 1)          # save() logs every refused row
    return ref


def fill_table(R, rows):
    """rows: [id, resource_id, start_us, end_us, user_id, is_cancelled (None / 0 / 1)] -> instances of the reference class."""
    Table.rows, Table.seq = [], max([r[0] for r in rows], default=0)
    for (i, res, s, e, uid, c) in rows:
        r = R.__new__(R)
        r.id, r.resource_id, r.user_id = i, res, uid
        r._start, r._end = at(s), at(e)
        r._is_cancelled = None if c is None else bool(c)
        r.title, r.description = 'T', 'D'
        Table.rows.append(r)


# ---------------------------------------------------------------------------------------------------------------------
# the reference's own tests for this path, run against the stand-in before it is trusted
# ---------------------------------------------------------------------------------------------------------------------
class _Faker:
    """The two faker calls of tests/unit/models/test_reservation_model.py, deterministic."""

    def future_datetime(self, end_date='+30d'):
        return _FrozenDateTime.utcnow() + dt.timedelta(hours=7, minutes=13)

    def past_datetime(self, start_date='-30d'):
        return _FrozenDateTime.utcnow() - dt.timedelta(hours=9, minutes=1)


def selfcheck(ref):
    """tests/unit/models/test_reservation_model.py, unmodified, with the fixtures of tests/fixtures/models.py:37-67
    rebuilt by hand (user ids 1 and 2, resource1's id).  Returns the names of the tests that ran."""
    R = ref['model'].Reservation
    tests = _load('tests/unit/models/test_reservation_model.py')
    tests.datetime = _FrozenDateTime                              # `datetime.utcnow()` in the last test
    ran = []
    for name in [n for n in dir(tests) if n.startswith('test_')]:
        Table.rows, Table.seq = [], 0
        _FrozenDateTime.now_us = 1_900_000_000_000_000
        now = _FrozenDateTime.utcnow()

        def fixture(uid):
            return R(user_id=uid, title='TEST TITLE', description='TEST_DESCRIPTION',
                     resource_id='GPU-d38d4de3-85ee-e837-3d87-e8e2faeb6a63', start=now, end=now + dt.timedelta(minutes=60))

        fn = getattr(tests, name)
        fn = getattr(fn, '__wrapped__', fn)
        args = fn.__code__.co_varnames[:fn.__code__.co_argcount]
        have = {'tables': None, 'new_reservation': fixture(1), 'new_reservation_2': fixture(2), 'faker': _Faker()}
        fn(*[have[a] for a in args])
        ran.append(name)
    assert len(ran) == 6, ran
    return ran


# ---------------------------------------------------------------------------------------------------------------------
# seeded worlds (inputs are written to the JSON in full: the tests never regenerate them)
# ---------------------------------------------------------------------------------------------------------------------
def make_world(seed):
    """Even seeds: a tiny time grid (ties and touching boundaries everywhere, zero-length rows, nested and duplicate
    intervals, ids with gaps in random order).  Odd seeds: a synth calendar with microsecond timestamps."""
    rng = np.random.default_rng(7000 + seed)
    if seed % 2 == 0:
        n = int(rng.integers(0, 90))
        u = int(rng.integers(1, 6))
        grid = synth.T0_US + np.arange(12) * 1000
        ids = rng.permutation(np.arange(1, 4 * n + 1))[:n]
        a, b = rng.choice(grid, size=n), rng.choice(grid, size=n)
        rows = [[int(ids[i]), synth.uuid_string(int(rng.integers(0, u))), int(min(a[i], b[i])), int(max(a[i], b[i])),
                 int(rng.integers(1, 5)), [None, 0, 1][int(rng.choice(3, p=[0.4, 0.3, 0.3]))]] for i in range(n)]
        times = [int(t) for t in grid] + [int(grid[0]) - 5, int(grid[-1]) + 7]
        jitter = [-1, 0, 1]
        periods = [0, 1, 1000, 2500, 100_000]
    else:
        n, u = 120, 7
        cal = synth.make_calendar(n, u, seed=900 + seed, skew=0.8 if seed % 3 == 0 else 0.0)
        rows = [[int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]), int(cal['end'][i]),
                 int(cal['owner'][i]) % 9 + 1, None if cal['cancelled_raw'][i] < 0 else int(cal['cancelled_raw'][i])]
                for i in range(n)]
        times = sorted({int(x) for x in rng.choice(np.concatenate([cal['start'], cal['end']]), size=10)})
        jitter = [-1, 0, 1, 1_800_000_000]
        periods = [0, 60_000_000, 1_800_000_000, 86_400_000_000]
    uuids = [synth.uuid_string(k) for k in range(u)] + [synth.uuid_string(9000)]       # the last one has no rows
    ids_pool = [r[0] for r in rows] + [10_000_000]
    probes = {'current': [], 'upcoming': [], 'range': [], 'interfere': []}
    for t in times[:8]:
        now = t + int(rng.choice(jitter))
        probes['current'].append({'now_us': now, 'resource_id': None})
        for res in uuids:
            probes['current'].append({'now_us': now, 'resource_id': res})
            probes['upcoming'].append({'now_us': now, 'resource_id': res, 'period_us': int(rng.choice(periods))})
    for _ in range(40):
        a, b = int(rng.choice(times)) + int(rng.choice(jitter)), int(rng.choice(times)) + int(rng.choice(jitter))
        k = int(rng.integers(0, len(uuids) + 1))
        probes['range'].append({'uuids': [uuids[int(x)] for x in rng.permutation(len(uuids))[:k]],
                                'start_us': min(a, b), 'end_us': max(a, b)})
    for _ in range(80):
        a, b = int(rng.choice(times)) + int(rng.choice(jitter)), int(rng.choice(times)) + int(rng.choice(jitter))
        c = rng.random()
        probes['interfere'].append({'resource_id': uuids[int(rng.integers(0, len(uuids)))], 'start_us': min(a, b),
                                    'end_us': max(a, b),
                                    'id': None if c < 0.4 else int(ids_pool[int(rng.integers(0, len(ids_pool)))])})
    return {'seed': seed, 'rows': rows, 'uuids': uuids, 'probes': probes}


def make_write_script(seed):
    """A sequence of save / move / cancel / destroy steps on a small table: the reference's own validation order."""
    rng = np.random.default_rng(8000 + seed)
    base = synth.T0_US
    H = 3_600_000_000
    res = [synth.uuid_string(k) for k in range(2)]
    steps = []
    for k in range(60):
        c = rng.random()
        start = base + int(rng.integers(0, 14)) * (H // 2)
        dur = int(rng.choice([H // 2 - 1, H // 2, H, 5 * H, 8 * 24 * H, 8 * 24 * H + 1_000_000, -H],
                             p=[0.06, 0.3, 0.3, 0.2, 0.04, 0.05, 0.05]))
        if c < 0.55:
            steps.append({'op': 'save_new', 'user_id': int(rng.choice([0, 1, 2, 3], p=[0.06, 0.4, 0.3, 0.24])),
                          'resource_id': res[int(rng.integers(0, 2))] if rng.random() < 0.95 else 'GPU-short',
                          'title': 'T' * int(rng.choice([0, 1, 10, 59, 60], p=[0.04, 0.3, 0.32, 0.3, 0.04])),
                          'description': 'D' * int(rng.choice([0, 5, 199, 200], p=[0.3, 0.36, 0.3, 0.04])),
                          'start_us': start, 'end_us': start + dur,
                          'is_cancelled': [None, False, True][int(rng.choice(3, p=[0.7, 0.15, 0.15]))]})
        elif c < 0.75:
            steps.append({'op': 'move', 'pick': int(rng.integers(0, 1000)), 'start_us': start, 'end_us': start + max(dur, H // 2)})
        elif c < 0.9:
            steps.append({'op': 'cancel', 'pick': int(rng.integers(0, 1000)), 'value': bool(rng.random() < 0.7)})
        else:
            steps.append({'op': 'destroy', 'pick': int(rng.integers(0, 1000))})
    return {'seed': seed, 'steps': steps}


# ---------------------------------------------------------------------------------------------------------------------
# running the reference on a world
# ---------------------------------------------------------------------------------------------------------------------
def run_world(ref, w):
    R = ref['model'].Reservation
    fill_table(R, w['rows'])
    sql = SqlCalendar([tuple(r) for r in w['rows']])
    out = {'current': [], 'upcoming': [], 'range': [], 'interfere': []}
    for p in w['probes']['current']:
        _FrozenDateTime.now_us = p['now_us']
        got = [r.id for r in R.current_events(p['resource_id'])] if p['resource_id'] is not None else [r.id for r in R.current_events()]
        assert got == sql.current_events(p['now_us'], p['resource_id']), ('current', w['seed'], p)
        out['current'].append(got)
    for p in w['probes']['upcoming']:
        _FrozenDateTime.now_us = p['now_us']
        got = [r.id for r in R.upcoming_events_for_resource(p['resource_id'], dt.timedelta(microseconds=p['period_us']))]
        assert got == sql.upcoming_events_for_resource(p['resource_id'], p['now_us'], p['period_us']), ('upcoming', w['seed'], p)
        out['upcoming'].append(got)
    for p in w['probes']['range']:
        got = [r.id for r in R.filter_by_uuids_and_time_range(p['uuids'], at(p['start_us']), at(p['end_us']))]
        want = sql.filter_by_uuids_and_time_range(p['uuids'], at(p['start_us']), at(p['end_us']))
        assert got == sorted(want), ('range', w['seed'], p)
        out['range'].append(got)
    for p in w['probes']['interfere']:
        cand = R(user_id=1, title='T', description='D', resource_id=p['resource_id'], start=at(p['start_us']), end=at(p['end_us']))
        cand.id = p['id']
        verdict = cand.would_interfere()
        rows = list(Table.last_query_ids)                 # what the reference's query returned, cancelled rows included
        srows = sql.would_interfere_rows(p['resource_id'], p['start_us'], p['end_us'], p['id'])
        assert rows == [r[0] for r in srows] and verdict == any(not c for _, c in srows), ('interfere', w['seed'], p)
        out['interfere'].append({'verdict': bool(verdict), 'rows': rows})
    sql.close()
    return out


def run_write_script(ref, script):
    """Executes the steps through the reference's save() / destroy(); records per step what happened and the table after it."""
    R = ref['model'].Reservation
    Table.rows, Table.seq = [], 0
    _FrozenDateTime.now_us = synth.T0_US
    log = []
    for st in script['steps']:
        rec = {}
        live = sorted(Table.rows, key=lambda r: r.id)
        try:
            if st['op'] == 'save_new':
                r = R(user_id=st['user_id'], title=st['title'], description=st['description'], resource_id=st['resource_id'],
                      start=at(st['start_us']), end=at(st['end_us']))
                if st['is_cancelled'] is not None:
                    r.is_cancelled = st['is_cancelled']
                r.save()
                rec = {'result': 'ok', 'id': r.id}
            elif not live:
                rec = {'result': 'skipped'}
            else:
                r = live[st['pick'] % len(live)]
                rec['target'] = r.id
                if st['op'] == 'move':
                    old = (r._start, r._end)
                    r.start, r.end = at(st['start_us']), at(st['end_us'])
                    try:
                        r.save()
                    except AssertionError:
                        r._start, r._end = old          # the controller's session would roll the object back
                        raise
                elif st['op'] == 'cancel':
                    old = r._is_cancelled
                    r.is_cancelled = st['value']
                    try:
                        r.save()                         # un-cancelling runs would_interfere again
                    except AssertionError:
                        r._is_cancelled = old
                        raise
                else:
                    r.destroy()
                rec['result'] = 'ok'
        except AssertionError as e:
            rec['result'] = 'AssertionError'
            rec['message'] = str(e)
        rec['ids'] = sorted(r.id for r in Table.rows)
        log.append(rec)
    table = [[r.id, r.resource_id, us_of(r._start), us_of(r._end), r.user_id,
              None if r._is_cancelled is None else int(r._is_cancelled)] for r in sorted(Table.rows, key=lambda r: r.id)]
    return {'steps': log, 'final_table': table}


def make_verifier_world(seed):
    """Valid reservations (30 min slots, 30 min ... 3 h long, two GPUs, two users), many of them cancelled, built through
    the reference's own save() so that live rows never overlap; then `now` somewhere inside the calendar."""
    rng = np.random.default_rng(9000 + seed)
    H = 1_800_000_000
    rows = []
    for _ in range(int(rng.integers(30, 90))):
        start = synth.T0_US + int(rng.integers(0, 40)) * H
        rows.append({'user_id': int(rng.integers(1, 3)), 'resource_id': synth.uuid_string(int(rng.integers(0, 2))),
                     'start_us': start, 'end_us': start + int(rng.integers(1, 7)) * H,
                     'is_cancelled': bool(rng.random() < 0.7)})
    return {'seed': seed, 'candidates': rows, 'now_us': synth.T0_US + int(rng.integers(0, 12)) * H + int(rng.integers(0, 2)),
            'user_id': int(rng.integers(1, 3)), 'denied_mod': int(rng.integers(2, 6))}


def run_verifier(ref, w):
    """Builds the table with save() (refused candidates are dropped), then runs the reference's
    update_user_reservations_statuses for the widening case with `is_reservation_allowed` replaced by a deterministic
    predicate (id % denied_mod != 0: the restriction walk itself is out of scope) and records what it restored, in order."""
    R = ref['model'].Reservation
    V = ref['verifier'].ReservationVerifier
    Table.rows, Table.seq = [], 0
    _FrozenDateTime.now_us = w['now_us']
    for c in w['candidates']:
        r = R(user_id=c['user_id'], title='T', description='D', resource_id=c['resource_id'], start=at(c['start_us']),
              end=at(c['end_us']))
        r.is_cancelled = c['is_cancelled']
        try:
            r.save()
        except AssertionError:
            pass
    table = [[r.id, r.resource_id, us_of(r._start), us_of(r._end), r.user_id, None if r._is_cancelled is None else int(r._is_cancelled)]
             for r in sorted(Table.rows, key=lambda r: r.id)]
    user = types.SimpleNamespace(id=w['user_id'])
    mine = [r for r in sorted(Table.rows, key=lambda r: r.id) if r.user_id == w['user_id']]    # the relationship loads by primary key
    user.get_reservations = lambda include_cancelled=False: mine if include_cancelled else [r for r in mine if not r.is_cancelled]
    before = {r.id: bool(r.is_cancelled) for r in mine}
    V.is_reservation_allowed = classmethod(lambda cls, u, r: r.id % w['denied_mod'] != 0)
    V.update_user_reservations_statuses(user, True)
    restored = [r.id for r in mine if before[r.id] and not r.is_cancelled]
    return {'table': table, 'restored': restored}


API_FMT = '%Y-%m-%dT%H:%M:%S.%fZ'


def make_view_world(seed):
    """A calendar for the range endpoint: rows with owners (one owner in the odd worlds has no user row), titles, results of
    the usage logger, created_at; probes in the API's string format, some malformed, some with arguments missing."""
    rng = np.random.default_rng(9500 + seed)
    H = 1_800_000_000
    uuids = [synth.uuid_string(k) for k in range(4)]
    rows = []
    for i in range(int(rng.integers(5, 40))):
        start = synth.T0_US + int(rng.integers(0, 60)) * H + int(rng.integers(0, 3)) * 250_000
        rows.append({'id': i + 1, 'resource_id': uuids[int(rng.integers(0, 3))], 'start_us': start,
                     'end_us': start + int(rng.integers(1, 9)) * H, 'user_id': int(rng.integers(1, 5)),
                     'is_cancelled': [None, False, True][int(rng.choice(3, p=[0.6, 0.2, 0.2]))], 'title': 'res %d' % i,
                     'description': 'd' * int(rng.integers(0, 5)), 'gpu_util_avg': [None, int(rng.integers(0, 101))][int(rng.integers(0, 2))],
                     'mem_util_avg': [None, int(rng.integers(0, 101))][int(rng.integers(0, 2))],
                     'created_at_us': synth.T0_US - int(rng.integers(0, 10)) * H})
    users = [[u, 'user%d' % u] for u in range(1, 5) if not (seed % 2 == 1 and u == 4)]
    probes = []
    for _ in range(14):
        a = synth.T0_US + int(rng.integers(-5, 70)) * H + int(rng.integers(0, 2)) * 250_000
        b = a + int(rng.integers(0, 30)) * H
        k = int(rng.integers(1, 5))
        probes.append({'resources_ids': [uuids[int(x)] for x in rng.permutation(4)[:k]], 'start': at(a).strftime(API_FMT),
                       'end': at(b).strftime(API_FMT)})
    probes += [{'resources_ids': uuids[:2], 'start': at(synth.T0_US).strftime('%Y_%m_%dT%H:%M:%S.%fZ'), 'end': at(synth.T0_US + H).strftime(API_FMT)},
               {'resources_ids': uuids[:2], 'start': at(synth.T0_US).strftime(API_FMT), 'end': '2026-13-45T00:00:00.000Z'},
               {'resources_ids': None, 'start': at(synth.T0_US).strftime(API_FMT), 'end': at(synth.T0_US + H).strftime(API_FMT)},
               {'resources_ids': [], 'start': at(synth.T0_US).strftime(API_FMT), 'end': at(synth.T0_US + H).strftime(API_FMT)},
               {'resources_ids': uuids[:1], 'start': None, 'end': at(synth.T0_US + H).strftime(API_FMT)},
               {'resources_ids': uuids, 'start': at(synth.T0_US - 100 * H).strftime(API_FMT), 'end': at(synth.T0_US + 200 * H).strftime(API_FMT)}]
    return {'seed': seed, 'rows': rows, 'users': users, 'probes': probes}


def run_view(ref, w):
    R = ref['model'].Reservation
    plain = lambda us: dt.datetime(1970, 1, 1) + dt.timedelta(microseconds=us)      # the ORM hands out plain datetimes
    users = {u: types.SimpleNamespace(id=u, username=name) for u, name in w['users']}
    Table.rows, Table.seq = [], len(w['rows'])
    for c in w['rows']:
        r = R.__new__(R)
        r.id, r.resource_id, r.user_id = c['id'], c['resource_id'], c['user_id']
        r._start, r._end, r._is_cancelled = plain(c['start_us']), plain(c['end_us']), c['is_cancelled']
        r.title, r.description = c['title'], c['description']
        r.gpu_util_avg, r.mem_util_avg, r.created_at = c['gpu_util_avg'], c['mem_util_avg'], plain(c['created_at_us'])
        r.user = users.get(c['user_id'])
        Table.rows.append(r)
    out = []
    for p in w['probes']:
        content, status = ref['controller'].get_selected(p['resources_ids'], p['start'], p['end'])
        out.append({'status': status, 'content': content})
    return out


def main():
    ref = load_reference()
    ran = selfcheck(ref)
    worlds = []
    for seed in range(16):
        w = make_world(seed)
        w['expect'] = run_world(ref, w)
        worlds.append(w)
    scripts = []
    for seed in range(6):
        s = make_write_script(seed)
        s['expect'] = run_write_script(ref, s)
        scripts.append(s)
    verifier_worlds = []
    for seed in range(12):
        w = make_verifier_world(seed)
        w['expect'] = run_verifier(ref, w)
        verifier_worlds.append(w)
    view_worlds = []
    for seed in range(8):
        w = make_view_world(seed)
        w['expect'] = run_view(ref, w)
        view_worlds.append(w)
    header = {
        'what': 'inputs and outputs of the reference\'s own Reservation model methods, executed under an expression-'
                'evaluating stand-in for SQLAlchemy (see tests/golden/make_reference_model_vectors.py)',
        'reference': 'kivicode/TensorHive-Fixed v1.1.0',
        'executed': [{'file': rel, 'sha256': hashlib.sha256(open(os.path.join(REF, rel), 'rb').read()).hexdigest(),
                      'bodies': bodies} for rel, bodies in EXECUTED.items()],
        'stand_in_checked_by': {'file': 'tests/unit/models/test_reservation_model.py',
                                'sha256': hashlib.sha256(open(os.path.join(REF, 'tests/unit/models/test_reservation_model.py'), 'rb').read()).hexdigest(),
                                'tests_passed': ran,
                                'and': 'every recorded answer equals oracle/sql_ref.py (sqlite3) on the same world'},
        'row_layout': ['id', 'resource_id', 'start_us', 'end_us', 'user_id', 'is_cancelled (null / 0 / 1)'],
    }
    path = os.path.join(HERE, 'reference_model_vectors.json')
    with open(path, 'w') as f:
        json.dump({'header': header, 'worlds': worlds, 'write_scripts': scripts, 'verifier_worlds': verifier_worlds,
                   'view_worlds': view_worlds}, f,
                  separators=(',', ':'))
    n_probe = sum(len(v) for w in worlds for v in w['probes'].values())
    n_true = sum(x['verdict'] for w in worlds for x in w['expect']['interfere'])
    print('wrote %s: %d worlds, %d probes (%d interfering), %d write scripts, %.0f KB' % (
        path, len(worlds), n_probe, n_true, len(scripts), os.path.getsize(path) / 1e3))
    print(' view worlds: statuses', sorted({x['status'] for w in view_worlds for x in w['expect']}),
          'rows returned', sum(len(x['content']) for w in view_worlds for x in w['expect'] if x['status'] == 200))
    print(' verifier worlds:', [(len(w['expect']['table']), len(w['expect']['restored'])) for w in verifier_worlds])
    for s in scripts:
        res = [x['result'] + (':' + x.get('message', '')[:28] if 'message' in x else '') for x in s['expect']['steps']]
        print(' script %d:' % s['seed'], {k: res.count(k) for k in sorted(set(res))})


if __name__ == '__main__':
    main()
