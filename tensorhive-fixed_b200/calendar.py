"""Columnar snapshot builder: ORM-ish rows -> int64 columns -> device snapshot.

The reference keeps reservations as SQLAlchemy rows with text datetimes and 40-character resource ids
(tensorhive/models/Reservation.py:14-37).  The snapshot is what SURVEY.md section 8a row R0 describes:
``id, uuid_id, start_us, end_us, owner`` as int64 plus a cancelled bit, where
  * ``uuid_id`` is the dense dictionary code of ``resource_id``,
  * datetimes become int64 microseconds since the epoch (order-isomorphic to the fixed-width text SQLite
    compares; the grain of DateUtils.input_date_format, tensorhive/utils/DateUtils.py:9),
  * ``cancelled`` is ``is_cancelled is not None and is_cancelled`` (Reservation.py:81-83),
  * ``owner`` is ``user_id``; rows whose user no longer exists (``reservation.user is None``) get -1.
"""
from __future__ import annotations

import contextlib
import datetime as _dt
import threading
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import numpy as np

from .api import Snapshot, MODE_STRICT, MODE_INCLUSIVE, MODE_UPCOMING

EPOCH = _dt.datetime(1970, 1, 1)


def dt_to_us(d: _dt.datetime) -> int:
    delta = d - EPOCH
    return (delta.days * 86400 + delta.seconds) * 1_000_000 + delta.microseconds


def us_to_dt(us: int) -> _dt.datetime:
    return EPOCH + _dt.timedelta(microseconds=int(us))


class UuidDictionary:
    """resource_id string <-> dense int code.  Unknown strings map to -1 (= "no reservations")."""

    def __init__(self):
        self.code: Dict[str, int] = {}
        self.names: List[str] = []

    def add(self, name: str) -> int:
        c = self.code.get(name)
        if c is None:
            c = len(self.names)
            self.code[name] = c
            self.names.append(name)
        return c

    def lookup(self, name: Optional[str]) -> int:
        if name is None:
            return -1
        return self.code.get(name, -1)

    def __len__(self):
        return len(self.names)


class ColumnRow:
    """Row object materialised on demand from the columns of a bulk load (attribute names of the reference model)."""
    __slots__ = ('id', 'user_id', 'resource_id', 'start', 'end', '_is_cancelled')

    def __init__(self, id, user_id, resource_id, start, end, is_cancelled):
        self.id, self.user_id, self.resource_id, self.start, self.end = id, user_id, resource_id, start, end
        self._is_cancelled = is_cancelled

    @property
    def is_cancelled(self):
        return self._is_cancelled is not None and bool(self._is_cancelled)


class _LazyRows:
    """id -> row for a columnar load: rows are built only for the ids a query returns (SURVEY.md section 8b:
    "re-materialises ORM objects by id only for the rows returned").  Upserts and deletes of later deltas are kept in
    an overlay, so the mapping behaves like the plain dict the row-wise path uses."""

    def __init__(self, ids, cols, factory):
        self._order = np.argsort(ids, kind='stable')
        self._sorted = np.asarray(ids, np.int64)[self._order]
        self._cols, self._factory = cols, factory
        self._over, self._gone = {}, set()

    def copy(self):
        c = _LazyRows.__new__(_LazyRows)
        c._order, c._sorted, c._cols, c._factory = self._order, self._sorted, self._cols, self._factory
        c._over, c._gone = dict(self._over), set(self._gone)
        return c

    def __setitem__(self, key, row):
        self._over[int(key)] = row
        self._gone.discard(int(key))

    def pop(self, key, default=None):
        key = int(key)
        self._gone.add(key)
        return self._over.pop(key, default)

    def __getitem__(self, key):
        key = int(key)
        if key in self._over:
            return self._over[key]
        if key in self._gone:
            raise KeyError(key)
        j = int(np.searchsorted(self._sorted, key))
        if j >= self._sorted.size or self._sorted[j] != key:
            raise KeyError(key)
        i = int(self._order[j])
        ids, users, names, codes, st, en, canc = self._cols
        uid = users[i]
        return self._factory(id=key, user_id=None if uid < 0 else int(uid), resource_id=names[int(codes[i])],
                             start=us_to_dt(st[i]), end=us_to_dt(en[i]),
                             is_cancelled=None if canc[i] < 0 else bool(canc[i]))


def _times_to_us(x) -> np.ndarray:
    """int64 microseconds from int arrays, numpy datetime64 arrays, datetime objects or SQLite's text datetimes."""
    a = np.asarray(x)
    if a.dtype.kind in 'iu':
        return a.astype(np.int64)
    if a.dtype.kind == 'M':
        return a.astype('datetime64[us]').astype(np.int64)
    return np.array(list(x), dtype='datetime64[us]').astype(np.int64)     # datetime objects or 'YYYY-MM-DD HH:MM:SS.ffffff'


class _Version:
    """One published state of the calendar and the number of readers that hold a lease on it."""
    __slots__ = ('snap', 'uuids', 'by_id', 'readers', 'retired')

    def __init__(self, snap, uuids, by_id):
        self.snap, self.uuids, self.by_id = snap, uuids, by_id
        self.readers, self.retired = 0, False


class Calendar:
    """The reservation table as the GPU path sees it.

    ``rows`` are objects (or dicts) exposing ``id, user_id, resource_id, start, end, is_cancelled`` --
    the attribute names of the reference model -- with ``start``/``end`` as naive-UTC ``datetime`` (or int
    microseconds).  ``users`` maps ``user_id`` to whatever object the callers want back as
    ``reservation.user`` (needs ``.username``; ``.email`` for the violation report).

    Thread model: service threads and API greenlets read concurrently while any of
    them may commit a write (tensorhive/core/managers/ServiceManager.py:23-25, database.py:19-23).

    * Writers (``refresh``, ``refresh_from_columns``, ``apply_delta``) are serialised by ONE writer lock held across the
      whole build, so two commits can never derive from the same source and lose one of the two.
    * Every version lives in its own device snapshot.  Readers take a lease (``with calendar.reading() as (snap, uuids,
      rows): ...``): the version they see is pinned until they leave the block, however many commits happen meanwhile.
    * A retired version's device buffers are recycled for a later build only once its last reader has left; if every
      spare is still being read, the writer allocates a fresh snapshot instead of waiting.
    """

    MAX_SPARES = 2

    def __init__(self, rows: Iterable = (), users: Optional[Dict[int, object]] = None, device: int = 0,
                 clock: Callable[[], _dt.datetime] = _dt.datetime.utcnow):
        self.device = device
        self.clock = clock
        self.users = dict(users or {})
        self._lock = threading.Lock()            # guards _state, reader counts, _spares
        self._writer = threading.RLock()         # serialises writers for the whole of a build
        self.version = 0
        self._state = None                       # _Version
        self._spares = []                        # snapshots no reader can still hold
        self._max_id = 0                         # AUTOINCREMENT counter: ids are never reused (sqlite_sequence)
        self.refresh(rows)

    # -- building -------------------------------------------------------------------------------------------
    @staticmethod
    def _get(row, name):
        if isinstance(row, dict):
            return row[name]
        return getattr(row, name)

    def refresh(self, rows: Iterable, users: Optional[Dict[int, object]] = None) -> None:
        rows = list(rows)
        if users is not None:
            self.users = dict(users)
        uuids = UuidDictionary()
        n = len(rows)
        cid = np.empty(n, np.int64)
        cuu = np.empty(n, np.int64)
        cst = np.empty(n, np.int64)
        cen = np.empty(n, np.int64)
        cow = np.empty(n, np.int64)
        cca = np.empty(n, np.uint8)
        by_id = {}
        for k, r in enumerate(rows):
            rid = int(self._get(r, 'id'))
            cid[k] = rid
            cuu[k] = uuids.add(self._get(r, 'resource_id'))
            s, e = self._get(r, 'start'), self._get(r, 'end')
            cst[k] = dt_to_us(s) if isinstance(s, _dt.datetime) else int(s)
            cen[k] = dt_to_us(e) if isinstance(e, _dt.datetime) else int(e)
            uid = self._get(r, 'user_id')
            cow[k] = int(uid) if (uid is not None and (not self.users or uid in self.users)) else -1
            raw = self._get(r, 'is_cancelled')
            cca[k] = 1 if (raw is not None and raw) else 0
            by_id[rid] = r
        with self._writer:
            snap = self._take_spare()
            snap.load(cid, cuu, cst, cen, cow, cca, n_uuids=max(len(uuids), 1))
            self._publish(snap, uuids, by_id, int(cid.max()) if n else 0)

    def refresh_from_columns(self, ids, resource_ids, starts, ends, user_ids, is_cancelled,
                             users: Optional[Dict[int, object]] = None, row_factory=ColumnRow) -> None:
        """Bulk form of ``refresh`` for large tables: the six columns as they come out of
        ``SELECT id, resource_id, _start, _end, user_id, is_cancelled FROM reservations`` (no ORM hydration, no Python
        loop over rows).  ``starts`` / ``ends``: int microseconds, ``numpy.datetime64``, ``datetime`` objects or SQLite's
        text datetimes; ``user_ids`` / ``is_cancelled`` may hold None (NULL).  Row objects are built on demand by
        ``row_factory(id=, user_id=, resource_id=, start=, end=, is_cancelled=)`` for the ids a query returns."""
        if users is not None:
            self.users = dict(users)
        cid = np.asarray(ids, np.int64)
        try:                                   # hash-based factorisation: ~10x faster than sorting 40-character strings
            import pandas as pd
            codes, names = pd.factorize(np.asarray(resource_ids, dtype=object))
        except ImportError:                    # pragma: no cover
            names, codes = np.unique(np.asarray(resource_ids, dtype=object).astype(str), return_inverse=True)
        cst, cen = _times_to_us(starts), _times_to_us(ends)

        def nullable_ints(col, as_flag):       # None (SQL NULL) -> -1
            a = np.asarray(col)
            if a.dtype != object:
                return a.astype(np.int64)
            null = np.equal(a, None)
            v = np.where(null, 0, a).astype(np.int64)
            return np.where(null, -1, (v != 0).astype(np.int64) if as_flag else v)
        uid = nullable_ints(user_ids, False)
        craw = nullable_ints(is_cancelled, True)
        cow = uid.copy()
        if self.users:
            cow[~np.isin(uid, np.fromiter(self.users.keys(), np.int64, len(self.users)))] = -1
        uuids = UuidDictionary()
        uuids.names = [str(x) for x in names]
        uuids.code = {n: k for k, n in enumerate(uuids.names)}
        by_id = _LazyRows(cid, (cid, uid, uuids.names, codes, cst, cen, craw), row_factory)
        with self._writer:
            snap = self._take_spare()
            snap.load(cid, codes.astype(np.int64), cst, cen, cow, (craw == 1).astype(np.uint8), n_uuids=max(len(uuids), 1))
            self._publish(snap, uuids, by_id, int(cid.max()) if cid.size else 0)

    # -- versions, leases, spares ------------------------------------------------------------------------------
    def _take_spare(self) -> Snapshot:
        """A snapshot handle nobody reads (writer lock held): a recycled one, else a new one."""
        with self._lock:
            if self._spares:
                return self._spares.pop()
        return Snapshot(self.device)

    def _retire(self, ver: '_Version') -> None:
        """(lock held) the version is no longer current and has no reader: its buffers may serve a later build."""
        if len(self._spares) < self.MAX_SPARES:
            self._spares.append(ver.snap)
        else:
            ver.snap.close()

    def _publish(self, snap, uuids, by_id, max_id: int) -> None:
        with self._lock:
            old = self._state
            self._state = _Version(snap, uuids, by_id)
            self.version += 1
            snap.version = self.version          # which calendar version this snapshot holds (readers may log it)
            self._max_id = max(self._max_id, int(max_id))
            if old is not None:
                old.retired = True
                if old.readers == 0:
                    self._retire(old)

    @contextlib.contextmanager
    def reading(self):
        """Lease on the current version: ``(snapshot, uuid dictionary, id -> row)``, pinned until the block is left."""
        with self._lock:
            ver = self._state
            ver.readers += 1
        try:
            yield ver.snap, ver.uuids, ver.by_id
        finally:
            with self._lock:
                ver.readers -= 1
                if ver.retired and ver.readers == 0:
                    self._retire(ver)

    def state(self):
        """The current ``(snapshot, uuids, rows)`` WITHOUT a lease: only for callers that know no writer runs
        concurrently (single-threaded tools, tests); everything else uses :meth:`reading`."""
        with self._lock:
            ver = self._state
            return ver.snap, ver.uuids, ver.by_id

    def next_id(self) -> int:
        """The id SQLite's AUTOINCREMENT would hand out next (never reuses an id, Reservation.py:20)."""
        with self._lock:
            self._max_id += 1
            return self._max_id

    def apply_delta(self, upserts: Iterable = (), deleted_ids: Iterable[int] = ()) -> None:
        """Incremental maintenance after a commit (CRUDModel.save / destroy, tensorhive/models/CRUDModel.py:21-50):
        ``upserts`` are new or changed rows (a cancel is a changed row), ``deleted_ids`` rows that were destroyed.
        Only the delta is sent to the GPU; the new snapshot is built in a spare handle from the current one and then
        swapped in, so concurrent readers are never blocked.  Several changes of one id inside a batch collapse to the
        last one; an id that is both changed and destroyed is destroyed."""
        deleted = list(dict.fromkeys(int(i) for i in deleted_ids))
        gone = set(deleted)
        last = {}
        for r in upserts:
            rid = int(self._get(r, 'id'))
            if rid < 0:
                raise ValueError('reservation ids are non-negative (got %d)' % rid)
            last.pop(rid, None)            # keep the position of the LAST occurrence
            last[rid] = r
        upserts = [r for rid, r in last.items() if rid not in gone]
        with self._writer:
            with self._lock:
                cur = self._state          # only writers replace it, and we hold the writer lock: it stays current
            snap, uuids, by_id = cur.snap, cur.uuids, cur.by_id
            new_uuids = UuidDictionary()
            new_uuids.code = dict(uuids.code)
            new_uuids.names = list(uuids.names)
            new_by_id = by_id.copy()
            n = len(upserts)
            cid = np.empty(n, np.int64); cuu = np.empty(n, np.int64); cst = np.empty(n, np.int64)
            cen = np.empty(n, np.int64); cow = np.empty(n, np.int64); cca = np.empty(n, np.uint8)
            for k, r in enumerate(upserts):
                rid = int(self._get(r, 'id'))
                cid[k] = rid
                cuu[k] = new_uuids.add(self._get(r, 'resource_id'))
                s, e = self._get(r, 'start'), self._get(r, 'end')
                cst[k] = dt_to_us(s) if isinstance(s, _dt.datetime) else int(s)
                cen[k] = dt_to_us(e) if isinstance(e, _dt.datetime) else int(e)
                uid = self._get(r, 'user_id')
                cow[k] = int(uid) if (uid is not None and (not self.users or uid in self.users)) else -1
                raw = self._get(r, 'is_cancelled')
                cca[k] = 1 if (raw is not None and raw) else 0
                new_by_id[rid] = r
            for i in deleted:
                new_by_id.pop(i, None)
            spare = self._take_spare()
            try:
                spare.apply_delta(snap, max(len(new_uuids), 1), cid, cuu, cst, cen, cow, cca, np.array(deleted, np.int64))
            except Exception:
                with self._lock:
                    self._retire(_Version(spare, None, None))
                raise
            self._publish(spare, new_uuids, new_by_id, int(cid.max()) if n else 0)

    # -- helpers ----------------------------------------------------------------------------------------------
    def now_us(self) -> int:
        return dt_to_us(self.clock())

    def usernames(self) -> Dict[str, int]:
        return {getattr(u, 'username', None): uid for uid, u in self.users.items()}

    def rows_for(self, ids: Sequence[int]) -> List:
        with self.reading() as (_snap, _uu, by_id):
            return [by_id[int(i)] for i in ids]

    # -- the four predicates, batch form ----------------------------------------------------------------------
    def would_interfere_batch(self, resource_ids: Sequence[Optional[str]], starts, ends,
                              self_ids: Sequence[Optional[int]]) -> np.ndarray:
        with self.reading() as (snap, uuids, _):
            qu = np.array([uuids.lookup(r) for r in resource_ids], np.int64)
            qs = np.array([dt_to_us(s) if isinstance(s, _dt.datetime) else int(s) for s in starts], np.int64)
            qe = np.array([dt_to_us(e) if isinstance(e, _dt.datetime) else int(e) for e in ends], np.int64)
            qf = np.array([-1 if i is None else int(i) for i in self_ids], np.int64)
            return snap.conflict_bitmap(qu, qs, qe, qf, MODE_STRICT)

    def sets(self, resource_ids: Sequence[Optional[str]], starts_us, ends_us, mode: int, self_ids=None):
        """CSR row lists AND the rows themselves, from ONE version: ``(offsets, ids, cancelled, rows)``."""
        with self.reading() as (snap, uuids, by_id):
            qu = np.array([uuids.lookup(r) for r in resource_ids], np.int64)
            qs = np.asarray(starts_us, np.int64)
            qe = np.asarray(ends_us, np.int64)
            qf = None if self_ids is None else np.array([-1 if i is None else int(i) for i in self_ids], np.int64)
            off, ids, canc = snap.conflict_sets(qu, qs, qe, qf, mode)
            return off, ids, canc, by_id

    def all_resource_ids(self) -> List[str]:
        return list(self.state()[1].names)

    def close(self) -> None:
        with self._writer, self._lock:
            if self._state is not None:
                self._state.snap.close()
            for s in self._spares:
                s.close()
            self._spares = []
