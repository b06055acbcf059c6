"""Both oracles (the reference's WHERE clauses as SQL text through sqlite3, oracle/sql_ref.py, and the numpy brute force,
oracle/brute.py) against vectors produced by EXECUTING the reference's own model methods -- would_interfere,
filter_by_uuids_and_time_range, current_events, upcoming_events_for_resource, check_assertions / save / destroy
(tests/golden/make_reference_model_vectors.py; the reference is not needed to run this test)."""
import hashlib
import json
import os

import numpy as np
import pytest

from helpers import HERE, columns_from_rows

with open(os.path.join(HERE, 'golden', 'reference_model_vectors.json')) as f:
    VEC = json.load(f)
WORLDS = VEC['worlds']


def test_header_names_what_was_executed():
    h = VEC['header']
    files = {e['file'] for e in h['executed']}
    assert {'tensorhive/models/Reservation.py', 'tensorhive/models/CRUDModel.py', 'tensorhive/utils/DateUtils.py',
            'tensorhive/core/utils/ReservationVerifier.py', 'tensorhive/controllers/reservation.py'} <= files
    assert all(len(e['sha256']) == 64 and e['bodies'] for e in h['executed'])
    assert len(h['stand_in_checked_by']['tests_passed']) == 6           # the reference's own six model tests
    ref = os.environ.get('TH_REFERENCE_ROOT', '/root/reference')
    if os.path.isdir(ref):                                                # in the builder's container: the hashes are current
        for e in h['executed']:
            assert hashlib.sha256(open(os.path.join(ref, e['file']), 'rb').read()).hexdigest() == e['sha256']


@pytest.mark.parametrize('w', WORLDS, ids=['world%d' % w['seed'] for w in WORLDS])
def test_sql_oracle_matches_the_executed_reference(w):
    from oracle.sql_ref import SqlCalendar, EPOCH
    import datetime as dt
    sql = SqlCalendar([tuple(r) for r in w['rows']])
    at = lambda us: EPOCH + dt.timedelta(microseconds=us)
    P, X = w['probes'], w['expect']
    for p, want in zip(P['current'], X['current']):
        assert sql.current_events(p['now_us'], p['resource_id']) == want
    for p, want in zip(P['upcoming'], X['upcoming']):
        assert sql.upcoming_events_for_resource(p['resource_id'], p['now_us'], p['period_us']) == want
    for p, want in zip(P['range'], X['range']):
        assert sorted(sql.filter_by_uuids_and_time_range(p['uuids'], at(p['start_us']), at(p['end_us']))) == want
    for p, want in zip(P['interfere'], X['interfere']):
        rows = sql.would_interfere_rows(p['resource_id'], p['start_us'], p['end_us'], p['id'])
        assert [r[0] for r in rows] == want['rows']
        assert sql.would_interfere(p['resource_id'], p['start_us'], p['end_us'], p['id']) == want['verdict']
    sql.close()


@pytest.mark.parametrize('w', WORLDS, ids=['world%d' % w['seed'] for w in WORLDS])
def test_brute_oracle_matches_the_executed_reference(w):
    from oracle import brute
    cal, code = columns_from_rows([tuple(r) for r in w['rows']])
    lookup = lambda name: code.get(name, -1)
    P, X = w['probes'], w['expect']
    # R1: verdicts and row lists
    q = P['interfere']
    qu = np.array([lookup(p['resource_id']) for p in q], np.int64)
    qs = np.array([p['start_us'] for p in q], np.int64)
    qe = np.array([p['end_us'] for p in q], np.int64)
    qf = np.array([-1 if p['id'] is None else p['id'] for p in q], np.int64)
    assert brute.conflict_sets(cal, qu, qs, qe, qf, brute.STRICT) == [x['rows'] for x in X['interfere']]
    assert [bool(b) for b in brute.conflict_bits(cal, qu, qs, qe, qf)] == [x['verdict'] for x in X['interfere']]
    # R3: current events = inclusive range at one instant, cancelled rows dropped afterwards (Reservation.py:104)
    canc = {int(i): int(c) for i, c in zip(cal['id'], cal['cancelled'])}
    names = sorted(code, key=code.get)
    for p, want in zip(P['current'], X['current']):
        res = [p['resource_id']] if p['resource_id'] is not None else names
        k = len(res)
        lists = brute.conflict_sets(cal, np.array([lookup(r) for r in res], np.int64), np.full(k, p['now_us']),
                                    np.full(k, p['now_us']), np.full(k, -1), brute.INCLUSIVE) if k else []
        assert sorted(i for lst in lists for i in lst if not canc[i]) == want
    # R4
    for p, want in zip(P['upcoming'], X['upcoming']):
        got = brute.conflict_sets(cal, np.array([lookup(p['resource_id'])], np.int64), np.array([p['now_us']]),
                                  np.array([p['now_us'] + p['period_us']]), np.array([-1]), brute.UPCOMING)[0]
        assert got == want
    # R2
    for p, want in zip(P['range'], X['range']):
        uu = list(dict.fromkeys(p['uuids']))
        k = len(uu)
        lists = brute.conflict_sets(cal, np.array([lookup(r) for r in uu], np.int64), np.full(k, p['start_us']),
                                    np.full(k, p['end_us']), np.full(k, -1), brute.INCLUSIVE) if k else []
        assert sorted(i for lst in lists for i in lst) == want


def test_vectors_are_not_trivial():
    n_true = sum(x['verdict'] for w in WORLDS for x in w['expect']['interfere'])
    n_all = sum(len(w['expect']['interfere']) for w in WORLDS)
    assert 0.1 < n_true / n_all < 0.9
    assert any(x for w in WORLDS for x in w['expect']['upcoming'] if len(x) > 2)
    assert any(x for w in WORLDS for x in w['expect']['range'] if len(x) > 5)
    assert sum(len(w['expect']['restored']) for w in VEC['verifier_worlds']) >= 20
    assert {x['status'] for w in VEC['view_worlds'] for x in w['expect']} == {200, 400, 500}
    msgs = {s.get('message') for sc in VEC['write_scripts'] for s in sc['expect']['steps']}
    assert {'Reservation would interfere with some other reservation!', 'Reservation duration is too short!',
            'Reservation duration is too long!', 'Reservation owner must be given!',
            'Protected resource UUID has incorrect length!'} <= msgs
