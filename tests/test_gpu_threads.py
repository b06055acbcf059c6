"""Service threads reading while one thread commits (the reference runs 3-4 service threads plus
greenlets against one table, core/managers/ServiceManager.py:23-25).  Three reader threads query in loops -- verdict
bitmap, row lists, violation join -- while a writer applies deltas; every answer is checked against the brute-force
oracle over the table AS IT WAS at the version the reader's lease pinned."""
import threading
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class Row:
    def __init__(self, rid, resource_id, start, end, user_id, is_cancelled):
        self.id, self.resource_id, self.start, self.end = rid, resource_id, start, end
        self.user_id, self.is_cancelled = user_id, is_cancelled


def test_readers_see_consistent_versions_while_a_writer_commits(lib_built):
    from tensorhive_fixed_b200 import synth
    from tensorhive_fixed_b200.calendar import Calendar
    from oracle import brute
    n, u = 3000, 16
    base = synth.make_calendar(n, u, seed=91)
    names = [synth.uuid_string(k) for k in range(u + 4)]
    table = {int(base['id'][i]): Row(int(base['id'][i]), names[int(base['uuid_id'][i])], int(base['start'][i]),
                                     int(base['end'][i]), int(base['owner'][i]), bool(base['cancelled'][i]))
             for i in range(n)}
    cal = Calendar(list(table.values()))
    truth = {cal.version: dict(table)}
    q = synth.make_queries(base, 400, seed=92)
    q_names = [names[int(k)] for k in q['uuid_id']]     # the calendar assigns its own codes: queries go by name
    now_us = int(np.median(base['start']))
    rng = np.random.default_rng(93)
    pu_names = [names[int(k)] for k in rng.integers(0, u + 2, size=300)]
    po = rng.integers(1, 1001, size=300).astype(np.int64)
    stop = threading.Event()
    records, errors = [], []

    def reader(kind):
        try:
            while not stop.is_set():
                with cal.reading() as (snap, uuids, rows):
                    version = snap.version
                    if kind == 'bitmap':
                        qu = np.array([uuids.lookup(x) for x in q_names], np.int64)
                        got = snap.conflict_bitmap(qu, q['start'], q['end'], q['self_id'])
                    elif kind == 'sets':
                        k = len(uuids)
                        codes = np.arange(k, dtype=np.int64)
                        off, ids, canc = snap.conflict_sets(codes, np.full(k, now_us, np.int64),
                                                            np.full(k, now_us + 30 * synth.DAY_US, np.int64), None, 1)
                        got = (np.asarray(off).copy(), np.asarray(ids).copy(), k)
                        for i in ids[:50]:
                            rows[int(i)]                   # every returned id has a row IN THIS VERSION
                    else:
                        pu = np.array([uuids.lookup(x) for x in pu_names], np.int64)
                        got = tuple(np.asarray(a).copy() for a in snap.violation_join(now_us, False, pu, po))
                    records.append((kind, version, got))
        except Exception as e:                          # noqa: BLE001
            errors.append(repr(e))

    def writer():
        try:
            wr = np.random.default_rng(94)
            next_id = n + 1
            while not stop.is_set():
                ups, dels = [], []
                live = list(table.keys())
                for rid in wr.choice(live, size=4, replace=False):
                    r = table[int(rid)]
                    shift = int(wr.integers(-3, 4)) * synth.HOUR_US
                    ups.append(Row(r.id, r.resource_id, r.start + shift, r.end + shift, r.user_id, bool(wr.random() < 0.2)))
                for _ in range(2):
                    s = synth.T0_US + int(wr.integers(0, 600)) * synth.DAY_US
                    ups.append(Row(next_id, names[int(wr.integers(0, u + 4))], s, s + 2 * synth.HOUR_US, int(wr.integers(1, 50)), False))
                    next_id += 1
                taken = {x.id for x in ups}
                dels = [int(x) for x in wr.choice(live, size=2, replace=False) if int(x) not in taken]
                for x in ups:
                    table[x.id] = x
                for d in dels:
                    table.pop(d, None)
                truth[cal.version + 1] = dict(table)     # single writer: the next version is the current one + 1
                cal.apply_delta(ups, dels)
        except Exception as e:                          # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=reader, args=(k,)) for k in ('bitmap', 'sets', 'join')]
    threads.append(threading.Thread(target=writer))
    for t in threads:
        t.start()
    time.sleep(2.0)
    stop.set()
    for t in threads:
        t.join(timeout=60)
    assert not errors, errors[:3]
    versions = {v for _k, v, _g in records}
    assert len(versions) >= 5 and cal.version >= 10, (len(versions), cal.version)
    assert {k for k, _v, _g in records} == {'bitmap', 'sets', 'join'}

    _snap, final_uuids, _rows = cal.state()
    code = final_uuids.code                             # the dictionary only grows: codes are stable across versions

    def columns(rows):
        rs = list(rows.values())
        return {'id': np.array([r.id for r in rs], np.int64), 'uuid_id': np.array([code[r.resource_id] for r in rs], np.int64),
                'start': np.array([r.start for r in rs], np.int64), 'end': np.array([r.end for r in rs], np.int64),
                'owner': np.array([r.user_id for r in rs], np.int64),
                'cancelled': np.array([1 if r.is_cancelled else 0 for r in rs], np.uint8)}

    checked = 0
    pick = np.random.default_rng(95).permutation(len(records))[:120]
    cols_cache = {}
    for j in pick:
        kind, version, got = records[int(j)]
        cols = cols_cache.setdefault(version, columns(truth[version]))
        if kind == 'bitmap':
            qu = np.array([code.get(x, -1) for x in q_names], np.int64)
            want = brute.conflict_bits(cols, qu, q['start'], q['end'], q['self_id'])
            assert np.array_equal(got, want), ('bitmap', version)
        elif kind == 'sets':
            off, ids, k = got
            codes = np.arange(k, dtype=np.int64)
            want = brute.conflict_sets(cols, codes, np.full(k, now_us, np.int64), np.full(k, now_us + 30 * synth.DAY_US, np.int64),
                                       np.full(k, -1, np.int64), brute.INCLUSIVE)
            assert [[int(x) for x in ids[off[i]:off[i + 1]]] for i in range(k)] == want, ('sets', version)
        else:
            flags, res = got
            want_res, want_own = brute.current_owner(cols, len(code), now_us)
            pu = np.array([code.get(x, -1) for x in pu_names], np.int64)
            for i in range(len(pu)):
                r, o = (want_res[pu[i]], want_own[pu[i]]) if pu[i] >= 0 else (-1, -1)
                viol = r >= 0 and o >= 0 and po[i] != o
                assert bool(flags[i]) == viol and res[i] == (r if viol else -1), ('join', version, i)
        checked += 1
    assert checked == len(pick)
    cal.close()
