"""Brute-force numpy oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline leg may import this.

An O(rows-in-uuid x queries-in-uuid) restatement of the reference's four predicates over int64
microsecond columns (order-isomorphic to the text datetimes SQLite compares; see ``sql_ref.py``).
Nothing here is clever on purpose: every predicate is the literal comparison the reference writes,
evaluated on every (query, row) pair that shares a ``resource_id``.  It is cross-checked against the
SQL oracle in ``tests/test_oracle_cross.py`` and used where SQLite's full scan per query is too slow.

Reference lines followed:
  R1 ``tensorhive/models/Reservation.py:120-130``   R2 ``:132-148``
  R3 ``:89-104``                                     R4 ``:106-118``
  A  ``tensorhive/core/services/UsageLoggingService.py:25-30`` (floor average, -1 when empty)
"""
from __future__ import annotations

import numpy as np

STRICT = 0      # R1: _end > qs AND _start < qe, id != self, verdict ignores cancelled rows
INCLUSIVE = 1   # R2/R3: _start <= qe AND qs <= _end, cancelled rows included
UPCOMING = 2    # R4: (_start < qs AND _end > qs) OR (_start >= qs AND _start <= qe)


def _group(uuid: np.ndarray):
    order = np.argsort(uuid, kind='stable')
    su = uuid[order]
    keys, first = np.unique(su, return_index=True)
    bounds = np.append(first, su.size)
    return order, {int(k): (int(bounds[i]), int(bounds[i + 1])) for i, k in enumerate(keys)}


def _match(mode, rs, re_, qs, qe):
    """Predicate matrix [queries, rows] for one uuid."""
    if mode == STRICT:
        return (re_[None, :] > qs[:, None]) & (rs[None, :] < qe[:, None])
    if mode == INCLUSIVE:
        return (rs[None, :] <= qe[:, None]) & (qs[:, None] <= re_[None, :])
    if mode == UPCOMING:
        a = (rs[None, :] < qs[:, None]) & (re_[None, :] > qs[:, None])
        b = (rs[None, :] >= qs[:, None]) & (rs[None, :] <= qe[:, None])
        return a | b
    raise ValueError(mode)


def conflict_sets(cal, q_uuid, q_start, q_end, q_self, mode, chunk=2048):
    """Per query, the ids of the rows the SQL would return, as python lists.

    ``cal`` = dict of equal-length int64 arrays ``id, uuid_id, start, end, owner`` and uint8
    ``cancelled``.  STRICT/INCLUSIVE results are ascending id (SQLite rowid scan order);
    UPCOMING results are ordered by (start, id) and exclude cancelled rows (the Python filter at
    Reservation.py:118).  STRICT applies ``id != self`` when ``q_self >= 0``.
    """
    q_uuid = np.asarray(q_uuid, np.int64)
    q_start = np.asarray(q_start, np.int64)
    q_end = np.asarray(q_end, np.int64)
    q_self = np.asarray(q_self, np.int64)
    m = q_uuid.size
    out = [[] for _ in range(m)]
    rorder, rgroups = _group(np.asarray(cal['uuid_id'], np.int64))
    qorder, qgroups = _group(q_uuid)
    rid_all = np.asarray(cal['id'], np.int64)
    rs_all = np.asarray(cal['start'], np.int64)
    re_all = np.asarray(cal['end'], np.int64)
    rc_all = np.asarray(cal['cancelled'], np.uint8)
    for u, (qa, qb) in qgroups.items():
        if u not in rgroups:
            continue
        ra, rb = rgroups[u]
        ridx = rorder[ra:rb]
        rid, rs, re_, rc = rid_all[ridx], rs_all[ridx], re_all[ridx], rc_all[ridx]
        for c0 in range(qa, qb, chunk):
            qi = qorder[c0:min(c0 + chunk, qb)]
            hit = _match(mode, rs, re_, q_start[qi], q_end[qi])
            if mode == STRICT:
                hit &= rid[None, :] != q_self[qi][:, None]   # q_self = -1 never equals an id
            if mode == UPCOMING:
                hit &= (rc == 0)[None, :]
            for k, qidx in enumerate(qi):
                sel = np.nonzero(hit[k])[0]
                if sel.size == 0:
                    continue
                if mode == UPCOMING:
                    o = np.lexsort((rid[sel], rs[sel]))
                else:
                    o = np.argsort(rid[sel], kind='stable')
                out[int(qidx)] = [int(x) for x in rid[sel][o]]
    return out


def conflict_bits(cal, q_uuid, q_start, q_end, q_self, mode=STRICT, chunk=4096):
    """Boolean verdict per query.  STRICT: ``any(not r.is_cancelled ...)`` (Reservation.py:130);
    INCLUSIVE: any row at all."""
    q_uuid = np.asarray(q_uuid, np.int64)
    q_start = np.asarray(q_start, np.int64)
    q_end = np.asarray(q_end, np.int64)
    q_self = np.asarray(q_self, np.int64)
    res = np.zeros(q_uuid.size, dtype=bool)
    rorder, rgroups = _group(np.asarray(cal['uuid_id'], np.int64))
    qorder, qgroups = _group(q_uuid)
    rid_all = np.asarray(cal['id'], np.int64)
    rs_all = np.asarray(cal['start'], np.int64)
    re_all = np.asarray(cal['end'], np.int64)
    rc_all = np.asarray(cal['cancelled'], np.uint8)
    for u, (qa, qb) in qgroups.items():
        if u not in rgroups:
            continue
        ra, rb = rgroups[u]
        ridx = rorder[ra:rb]
        rid, rs, re_, rc = rid_all[ridx], rs_all[ridx], re_all[ridx], rc_all[ridx]
        for c0 in range(qa, qb, chunk):
            qi = qorder[c0:min(c0 + chunk, qb)]
            hit = _match(mode, rs, re_, q_start[qi], q_end[qi])
            if mode == STRICT:
                hit &= rid[None, :] != q_self[qi][:, None]
                hit &= (rc == 0)[None, :]
            res[qi] = hit.any(axis=1)
    return res


def current_owner(cal, n_uuids, now_us):
    """Per uuid: the reservation ProtectionService would pick, ``current_events(uuid)[0]``
    (ProtectionService.py:90-93): the lowest-id non-cancelled row with start <= now <= end.
    Returns (res_id[U], owner[U]) with -1 where there is none."""
    rid = np.asarray(cal['id'], np.int64)
    ok = (np.asarray(cal['start']) <= now_us) & (now_us <= np.asarray(cal['end'])) & (np.asarray(cal['cancelled']) == 0)
    res_id = np.full(n_uuids, -1, np.int64)
    owner = np.full(n_uuids, -1, np.int64)
    idx = np.nonzero(ok)[0]
    if idx.size:
        u = np.asarray(cal['uuid_id'], np.int64)[idx]
        o = np.lexsort((rid[idx], u))
        idx, u = idx[o], u[o]
        firsts = np.ones(idx.size, bool)
        firsts[1:] = u[1:] != u[:-1]
        res_id[u[firsts]] = rid[idx[firsts]]
        owner[u[firsts]] = np.asarray(cal['owner'], np.int64)[idx[firsts]]
    return res_id, owner


def upcoming_summary(cal, uuids, now_us, period_us):
    """Per requested uuid: (first_id, first_start, first_owner, count) of
    ``upcoming_events_for_resource`` (Reservation.py:106-118); -1/-1/-1/0 when empty."""
    uuids = np.asarray(uuids, np.int64)
    k = uuids.size
    sets = conflict_sets(cal, uuids, np.full(k, now_us, np.int64), np.full(k, now_us + period_us, np.int64),
                         np.full(k, -1, np.int64), UPCOMING)
    by_id = {int(i): j for j, i in enumerate(np.asarray(cal['id']))}
    first_id = np.full(k, -1, np.int64)
    first_start = np.full(k, -1, np.int64)
    first_owner = np.full(k, -1, np.int64)
    count = np.zeros(k, np.int64)
    for j, ids in enumerate(sets):
        count[j] = len(ids)
        if ids:
            r = by_id[ids[0]]
            first_id[j] = ids[0]
            first_start[j] = int(cal['start'][r])
            first_owner[j] = int(cal['owner'][r])
    return first_id, first_start, first_owner, count


def floor_avg(total: int, count: int):
    """``avg`` of UsageLoggingService.py:25-30 given sum and length: ``sum // len`` or -1."""
    if count == 0:
        return -1
    return int(total) // int(count)


def usage_reduce(uuid_offsets, ts, gpu, mem, w_uuid, w_start, w_end):
    """Sums of the samples a reservation's log would hold: samples of the window's uuid with
    ``start <= ts <= end`` (appended only while the reservation is in ``current_events()``,
    UsageLoggingService.py:159-171 + Reservation.py:96-97).  Plain mask-and-sum per window."""
    uuid_offsets = np.asarray(uuid_offsets, np.int64)
    ts, gpu, mem = np.asarray(ts, np.int64), np.asarray(gpu, np.int64), np.asarray(mem, np.int64)
    w = len(w_uuid)
    sg = np.zeros(w, np.int64)
    sm = np.zeros(w, np.int64)
    cnt = np.zeros(w, np.int64)
    n_u = uuid_offsets.size - 1
    for j in range(w):
        u = int(w_uuid[j])
        if u < 0 or u >= n_u:
            continue
        a, b = int(uuid_offsets[u]), int(uuid_offsets[u + 1])
        t = ts[a:b]
        sel = (t >= int(w_start[j])) & (t <= int(w_end[j]))
        sg[j] = gpu[a:b][sel].sum()
        sm[j] = mem[a:b][sel].sum()
        cnt[j] = int(sel.sum())
    return sg, sm, cnt
