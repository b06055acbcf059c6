"""SQL oracle for the reservation predicates -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import this module.  The product path (``tensorhive_fixed_b200``) never does.

What this restates
------------------
The reference evaluates its four interval predicates by letting SQLAlchemy 1.3.0 (pinned in the
reference's ``setup.py:69``; not vendored, not installable here) compile a WHERE clause that SQLite
executes against the ``reservations`` table.  SQLAlchemy only *builds text*; the arithmetic is
SQLite's.  This module therefore creates the table the way the reference's model declares it
(``tensorhive/models/Reservation.py:14-37``: ``id INTEGER PRIMARY KEY AUTOINCREMENT``,
``_start``/``_end`` DATETIME stored by SQLAlchemy's SQLite dialect as ``'%Y-%m-%d %H:%M:%S.%f'``
text, ``is_cancelled`` BOOLEAN NULL/0/1, ``resource_id`` VARCHAR(60), **no index**) in stdlib
``sqlite3`` and runs the same WHERE clauses, followed by the same Python post-filters:

=====  ===============================================  ================================
row    reference                                        here
=====  ===============================================  ================================
R1     ``Reservation.would_interfere``     :120-130     :func:`SqlCalendar.would_interfere`
R2     ``filter_by_uuids_and_time_range``  :132-148     :func:`SqlCalendar.filter_by_uuids_and_time_range`
R3     ``Reservation.current_events``      :89-104      :func:`SqlCalendar.current_events`
R4     ``upcoming_events_for_resource``    :106-118     :func:`SqlCalendar.upcoming_events_for_resource`
=====  ===============================================  ================================

Parity pinning: the reference's own tests for this path
(``tests/unit/models/test_reservation_model.py:10-64,113-125``) are transcribed as known-answer
vectors in ``tests/golden/reference_kat.json`` and checked against this oracle by
``tests/test_oracle_kat.py``.  R2, R4 have no reference test; all four rows are additionally held to
``tests/golden/reference_model_vectors.json`` -- answers of the reference's own model methods,
executed under an expression-evaluating stand-in for SQLAlchemy
(``tests/golden/make_reference_model_vectors.py``) -- by ``tests/test_oracle_model_vectors.py``.
"""
from __future__ import annotations

import datetime as _dt
import sqlite3
from typing import Iterable, List, Optional, Sequence, Tuple

EPOCH = _dt.datetime(1970, 1, 1)
_SA_FMT = '%Y-%m-%d %H:%M:%S.%f'  # SQLAlchemy's SQLite DateTime storage format


def us_to_text(us: int) -> str:
    """int64 microseconds since epoch -> the text SQLAlchemy would store in a DATETIME column."""
    return (EPOCH + _dt.timedelta(microseconds=int(us))).strftime(_SA_FMT)


def text_to_us(text: str) -> int:
    d = _dt.datetime.strptime(text, _SA_FMT) - EPOCH
    return (d.days * 86400 + d.seconds) * 1_000_000 + d.microseconds


def dt_to_us(d: _dt.datetime) -> int:
    delta = d - EPOCH
    return (delta.days * 86400 + delta.seconds) * 1_000_000 + delta.microseconds


def _is_cancelled(raw) -> bool:
    # Reservation.is_cancelled hybrid property, Reservation.py:81-83: NULL counts as False.
    return raw is not None and bool(raw)


class SqlCalendar:
    """In-memory SQLite ``reservations`` table with the reference's columns and no index.

    Rows are ``(id, resource_id, start_us, end_us, user_id, is_cancelled)`` where ``is_cancelled`` is
    ``None``/0/1 (tri-state, as the nullable Boolean column stores it) and ``id`` may be ``None`` for
    autoincrement.
    """

    def __init__(self, rows: Iterable[Tuple] = ()):  # noqa: D401
        self.db = sqlite3.connect(':memory:', check_same_thread=False)
        self.db.execute(
            'CREATE TABLE reservations ('
            ' id INTEGER PRIMARY KEY AUTOINCREMENT,'
            ' user_id INTEGER NOT NULL,'
            ' title VARCHAR(60) NOT NULL,'
            ' description VARCHAR(200),'
            ' resource_id VARCHAR(60) NOT NULL,'
            ' is_cancelled BOOLEAN,'
            ' gpu_util_avg INTEGER,'
            ' mem_util_avg INTEGER,'
            ' _start DATETIME NOT NULL,'
            ' _end DATETIME NOT NULL,'
            ' created_at DATETIME)')
        self.insert_many(rows)

    # -- table maintenance -------------------------------------------------------------------
    def insert_many(self, rows: Iterable[Tuple]) -> None:
        self.db.executemany(
            'INSERT INTO reservations (id, user_id, title, description, resource_id, is_cancelled, _start, _end)'
            ' VALUES (?, ?, ?, ?, ?, ?, ?, ?)',
            ((rid, int(owner), 't', 'd', res, canc, us_to_text(s), us_to_text(e))
             for (rid, res, s, e, owner, canc) in rows))
        self.db.commit()

    def insert(self, resource_id: str, start_us: int, end_us: int, user_id: int, is_cancelled=None,
               rid: Optional[int] = None) -> int:
        cur = self.db.execute(
            'INSERT INTO reservations (id, user_id, title, description, resource_id, is_cancelled, _start, _end)'
            ' VALUES (?, ?, ?, ?, ?, ?, ?, ?)',
            (rid, int(user_id), 't', 'd', resource_id, is_cancelled, us_to_text(start_us), us_to_text(end_us)))
        self.db.commit()
        return int(cur.lastrowid)

    def update(self, rid: int, start_us: int = None, end_us: int = None, is_cancelled='keep') -> None:
        if start_us is not None:
            self.db.execute('UPDATE reservations SET _start = ? WHERE id = ?', (us_to_text(start_us), rid))
        if end_us is not None:
            self.db.execute('UPDATE reservations SET _end = ? WHERE id = ?', (us_to_text(end_us), rid))
        if is_cancelled != 'keep':
            self.db.execute('UPDATE reservations SET is_cancelled = ? WHERE id = ?', (is_cancelled, rid))
        self.db.commit()

    # -- R1 ------------------------------------------------------------------------------------
    def would_interfere_rows(self, resource_id: str, start_us: int, end_us: int,
                             self_id: Optional[int]) -> List[Tuple[int, bool]]:
        """The ``.all()`` of Reservation.py:121-129 as ``[(id, is_cancelled), ...]`` in scan (rowid) order.

        ``self.start < Reservation.end`` compiles to ``_end > :start``; ``self.end > Reservation.start``
        to ``_start < :end``; ``Reservation.id != None`` compiles to ``id IS NOT NULL``.
        """
        qs, qe = us_to_text(start_us), us_to_text(end_us)
        if self_id is None:
            cur = self.db.execute(
                'SELECT id, is_cancelled FROM reservations'
                ' WHERE _end > ? AND _start < ? AND id IS NOT NULL AND resource_id = ?', (qs, qe, resource_id))
        else:
            cur = self.db.execute(
                'SELECT id, is_cancelled FROM reservations'
                ' WHERE _end > ? AND _start < ? AND id != ? AND resource_id = ?', (qs, qe, int(self_id), resource_id))
        return [(int(r[0]), _is_cancelled(r[1])) for r in cur.fetchall()]

    def would_interfere(self, resource_id: str, start_us: int, end_us: int, self_id: Optional[int]) -> bool:
        # Reservation.py:130  any(not r.is_cancelled for r in conflicting_reservations)
        return any(not canc for (_id, canc) in self.would_interfere_rows(resource_id, start_us, end_us, self_id))

    # -- R2 ------------------------------------------------------------------------------------
    def filter_by_uuids_and_time_range(self, uuids: Sequence[str], start, end) -> List[int]:
        """Reservation.py:132-148.  ``start``/``end`` must be ``datetime`` (AssertionError otherwise);
        cancelled rows ARE returned; no ORDER BY."""
        msg = 'Argument must be of type datetime.datetime!'
        assert isinstance(start, _dt.datetime), msg
        assert isinstance(end, _dt.datetime), msg
        uuids = list(uuids)
        if not uuids:
            return []
        out: List[int] = []
        s, e = start.strftime(_SA_FMT), end.strftime(_SA_FMT)
        # SQLite's default variable limit is large in 3.45 (32766); chunk anyway.
        rows: List[int] = []
        if len(uuids) <= 30000:
            marks = ','.join('?' * len(uuids))
            cur = self.db.execute(
                'SELECT id FROM reservations WHERE resource_id IN (%s) AND _start <= ? AND ? <= _end' % marks,
                (*uuids, e, s))
            rows = [int(r[0]) for r in cur.fetchall()]
        else:  # pragma: no cover - not reached by the suite
            wanted = set(uuids)
            cur = self.db.execute('SELECT id, resource_id FROM reservations WHERE _start <= ? AND ? <= _end', (e, s))
            rows = [int(r[0]) for r in cur.fetchall() if r[1] in wanted]
        out.extend(rows)
        return out

    # -- R3 ------------------------------------------------------------------------------------
    def current_events(self, now_us: int, resource_id: Optional[str] = None) -> List[int]:
        """Reservation.py:89-104 with ``utcnow()`` replaced by an explicit instant."""
        now = us_to_text(now_us)
        if resource_id is None:
            cur = self.db.execute(
                'SELECT id, is_cancelled FROM reservations WHERE _start <= ? AND ? <= _end', (now, now))
        else:
            cur = self.db.execute(
                'SELECT id, is_cancelled FROM reservations WHERE _start <= ? AND ? <= _end AND resource_id = ?',
                (now, now, resource_id))
        return [int(r[0]) for r in cur.fetchall() if not _is_cancelled(r[1])]

    # -- R4 ------------------------------------------------------------------------------------
    def upcoming_events_for_resource(self, resource_id: Optional[str], now_us: int, period_us: int) -> List[int]:
        """Reservation.py:106-118 with an explicit instant.  ``resource_id=None`` compiles to
        ``resource_id IS NULL`` in SQLAlchemy, which matches nothing (the column is NOT NULL)."""
        if resource_id is None:
            return []
        now, later = us_to_text(now_us), us_to_text(now_us + period_us)
        cur = self.db.execute(
            'SELECT id, is_cancelled FROM reservations WHERE resource_id = ? AND'
            ' ((_start < ? AND _end > ?) OR (_start >= ? AND _start <= ?)) ORDER BY _start',
            (resource_id, now, now, now, later))
        return [int(r[0]) for r in cur.fetchall() if not _is_cancelled(r[1])]

    # -- row access (what ORM hydration would give the callers) -----------------------------------
    def row(self, rid: int):
        r = self.db.execute(
            'SELECT id, resource_id, _start, _end, user_id, is_cancelled FROM reservations WHERE id = ?',
            (rid,)).fetchone()
        if r is None:
            return None
        return (int(r[0]), r[1], text_to_us(r[2]), text_to_us(r[3]), int(r[4]), r[5])

    def close(self) -> None:
        self.db.close()
