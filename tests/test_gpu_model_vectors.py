"""The CUDA path and the drop-in against vectors produced by EXECUTING the reference's own model methods
(tests/golden/make_reference_model_vectors.py): would_interfere (R1), filter_by_uuids_and_time_range (R2), current_events
(R3), upcoming_events_for_resource (R4) on 16 small worlds full of ties, touching boundaries, zero-length rows and
NULL / 0 / 1 cancel flags, and check_assertions / save / destroy step by step.  The reference is not on the GPU box; the
committed JSON holds inputs and outputs."""
import json
import os
from datetime import timedelta

import numpy as np
import pytest

from helpers import HERE, EPOCH, columns_from_rows, csr_to_lists

pytestmark = pytest.mark.gpu

with open(os.path.join(HERE, 'golden', 'reference_model_vectors.json')) as f:
    VEC = json.load(f)
WORLDS = VEC['worlds']
at = lambda us: EPOCH + timedelta(microseconds=us)


@pytest.fixture()
def dropin(lib_built):
    from tensorhive_fixed_b200 import dropin
    return dropin


def _bind(dropin, rows, now_us=0):
    from tensorhive_fixed_b200.calendar import Calendar
    recs = [dropin.Reservation(id=r[0], resource_id=r[1], start=at(r[2]), end=at(r[3]), user_id=r[4],
                               is_cancelled=None if r[5] is None else bool(r[5]), title='T', description='D') for r in rows]
    clock = {'now': at(now_us)}
    cal = Calendar(recs, clock=lambda: clock['now'])
    dropin.bind(cal)
    return cal, clock


@pytest.mark.parametrize('w', WORLDS, ids=['world%d' % w['seed'] for w in WORLDS])
def test_model_methods_match_the_executed_reference(dropin, w):
    import tensorhive_fixed_b200 as th
    cal, clock = _bind(dropin, w['rows'])
    P, X = w['probes'], w['expect']
    R = dropin.Reservation
    for p, want in zip(P['current'], X['current']):                                     # R3, Reservation.py:89-104
        clock['now'] = at(p['now_us'])
        got = R.current_events(p['resource_id']) if p['resource_id'] is not None else R.current_events()
        assert [r.id for r in got] == want, p
    for p, want in zip(P['upcoming'], X['upcoming']):                                   # R4, :106-118 (ORDER BY start)
        clock['now'] = at(p['now_us'])
        got = R.upcoming_events_for_resource(p['resource_id'], timedelta(microseconds=p['period_us']))
        assert [r.id for r in got] == want, p
    for p, want in zip(P['range'], X['range']):                                         # R2, :132-148
        got = R.filter_by_uuids_and_time_range(p['uuids'], at(p['start_us']), at(p['end_us']))
        assert [r.id for r in got] == want, p
    cands = []
    for p, want in zip(P['interfere'], X['interfere']):                                 # R1, :120-130
        c = R(user_id=1, title='T', description='D', resource_id=p['resource_id'], start=at(p['start_us']),
              end=at(p['end_us']), id=p['id'])
        assert c.would_interfere() is want['verdict'], p
        cands.append(c)
    assert R.would_interfere_many(cands) == [x['verdict'] for x in X['interfere']]
    # the rows the reference's query returned (cancelled ones included), through the list-returning entry point
    q = P['interfere']
    _off, ids, _canc, _rows = cal.sets([p['resource_id'] for p in q], [p['start_us'] for p in q], [p['end_us'] for p in q],
                                       th.MODE_STRICT, self_ids=[p['id'] for p in q])
    assert csr_to_lists(_off, ids) == [x['rows'] for x in X['interfere']]
    cal.close()


@pytest.mark.parametrize('w', WORLDS, ids=['world%d' % w['seed'] for w in WORLDS])
def test_kernels_match_the_executed_reference(lib_built, w):
    """The same vectors through the array API (one batch per predicate), without the drop-in in between."""
    import tensorhive_fixed_b200 as th
    cal, code = columns_from_rows([tuple(r) for r in w['rows']])
    lookup = lambda name: code.get(name, -1)
    u = cal['n_uuids']
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    P, X = w['probes'], w['expect']
    q = P['interfere']
    qu = np.array([lookup(p['resource_id']) for p in q], np.int64)
    qs = np.array([p['start_us'] for p in q], np.int64)
    qe = np.array([p['end_us'] for p in q], np.int64)
    qf = np.array([-1 if p['id'] is None else p['id'] for p in q], np.int64)
    bits = snap.conflict_bitmap(qu, qs, qe, qf)
    assert [bool(b) for b in bits] == [x['verdict'] for x in X['interfere']]
    off, ids, canc = snap.conflict_sets(qu, qs, qe, qf, th.MODE_STRICT)
    assert csr_to_lists(off, ids) == [x['rows'] for x in X['interfere']]
    # packed wire format: the same verdicts
    pu, pf = snap.pack_queries(qu, qf)
    words = snap.conflict_words_packed(pu, qs, qe, pf)
    assert np.array_equal(th.api.unpack_bits(words, len(qu)), bits)
    # R4 in one batch
    q = P['upcoming']
    qu = np.array([lookup(p['resource_id']) for p in q], np.int64)
    now = np.array([p['now_us'] for p in q], np.int64)
    end = np.array([p['now_us'] + p['period_us'] for p in q], np.int64)
    off, ids, canc = snap.conflict_sets(qu, now, end, None, th.MODE_UPCOMING)
    assert csr_to_lists(off, ids) == X['upcoming']
    # R3 per uuid in one batch: inclusive at one instant, cancelled rows dropped
    q = [(p, x) for p, x in zip(P['current'], X['current']) if p['resource_id'] is not None]
    qu = np.array([lookup(p['resource_id']) for p, _ in q], np.int64)
    now = np.array([p['now_us'] for p, _ in q], np.int64)
    off, ids, canc = snap.conflict_sets(qu, now, now, None, th.MODE_INCLUSIVE)
    lists = csr_to_lists(off, ids)
    flags = csr_to_lists(off, canc)
    assert [[i for i, c in zip(l, f) if not c] for l, f in zip(lists, flags)] == [x for _, x in q]
    # ... and the owner table (the current reservation of a uuid = the smallest id among them)
    for now_us in sorted({p['now_us'] for p, _ in q})[:6]:
        res, _own = snap.current_owner(now_us)
        for p, x in q:
            if p['now_us'] == now_us and lookup(p['resource_id']) >= 0:
                assert res[lookup(p['resource_id'])] == (x[0] if x else -1)
    snap.close()


@pytest.mark.parametrize('sc', VEC['write_scripts'], ids=['script%d' % s['seed'] for s in VEC['write_scripts']])
def test_write_path_matches_the_executed_reference(dropin, sc):
    """save_new / move / cancel / destroy, step by step: the same rows accepted, the same assertion refusing the others
    (check_assertions' order, Reservation.py:42-55), the same ids handed out (AUTOINCREMENT), the same table at the end."""
    from tensorhive_fixed_b200 import synth
    cal, _clock = _bind(dropin, [], synth.T0_US)
    R = dropin.Reservation
    live = {}
    for st, want in zip(sc['steps'], sc['expect']['steps']):
        got = {}
        order = sorted(live)
        try:
            if st['op'] == 'save_new':
                r = R(user_id=st['user_id'], title=st['title'], description=st['description'], resource_id=st['resource_id'],
                      start=at(st['start_us']), end=at(st['end_us']))
                if st['is_cancelled'] is not None:
                    r.is_cancelled = st['is_cancelled']
                r.save()
                live[r.id] = r
                got = {'result': 'ok', 'id': r.id}
            elif not order:
                got = {'result': 'skipped'}
            else:
                r = live[order[st['pick'] % len(order)]]
                got['target'] = r.id
                if st['op'] == 'move':
                    old = (r.start, r.end)
                    r.start, r.end = at(st['start_us']), at(st['end_us'])
                    try:
                        r.save()
                    except AssertionError:
                        r.start, r.end = old
                        raise
                elif st['op'] == 'cancel':
                    old = r._is_cancelled
                    r.is_cancelled = st['value']
                    try:
                        r.save()
                    except AssertionError:
                        r._is_cancelled = old
                        raise
                else:
                    r.destroy()
                    del live[r.id]
                got['result'] = 'ok'
        except AssertionError as e:
            got['result'] = 'AssertionError'
            got['message'] = str(e)
        got['ids'] = sorted(live)
        assert got == want, (st, got, want)
        # the snapshot agrees with the bookkeeping: every live row can be fetched, nothing else is current
        for i in got['ids'][-3:]:
            assert R.get(i).id == i
    final = [[r.id, r.resource_id, int((r.start - EPOCH) / timedelta(microseconds=1)), int((r.end - EPOCH) / timedelta(microseconds=1)),
              r.user_id, None if r._is_cancelled is None else int(r._is_cancelled)] for r in (R.get(i) for i in sorted(live))]
    assert final == sc['expect']['final_table']
    cal.close()


@pytest.mark.parametrize('w', VEC['verifier_worlds'], ids=['verifier%d' % w['seed'] for w in VEC['verifier_worlds']])
def test_bulk_revalidation_matches_the_executed_reference(dropin, w):
    """ReservationVerifier.update_user_reservations_statuses, widening case (ReservationVerifier.py:90-109), executed from
    the reference's source with `is_reservation_allowed` replaced by a fixed predicate: the same reservations come back
    from the drop-in's batched form, in the same order (one restored earlier is a live row for the ones after it)."""
    cal, _clock = _bind(dropin, w['expect']['table'], w['now_us'])
    mine = sorted((dropin.Reservation.get(r[0]) for r in w['expect']['table'] if r[4] == w['user_id']), key=lambda r: r.id)
    back = dropin.reservations_that_may_be_restored(mine, is_allowed=lambda r: r.id % w['denied_mod'] != 0)
    assert [r.id for r in back] == w['expect']['restored']
    cal.close()


class _User:
    def __init__(self, uid, username):
        self.id, self.username = uid, username


@pytest.mark.parametrize('w', VEC['view_worlds'], ids=['view%d' % w['seed'] for w in VEC['view_worlds']])
def test_calendar_view_matches_the_executed_reference(dropin, w):
    """controllers/reservation.py:31-55 `get_selected` executed from the reference's source (with Reservation.as_dict and
    CRUDModel.as_dict): the same content and status from the drop-in -- row dicts key by key, 400 for malformed or missing
    arguments with the same message, 500 when a row's owner has no user."""
    from tensorhive_fixed_b200.calendar import Calendar
    recs = []
    for c in w['rows']:
        r = dropin.Reservation(id=c['id'], resource_id=c['resource_id'], start=at(c['start_us']), end=at(c['end_us']),
                               user_id=c['user_id'], is_cancelled=c['is_cancelled'], title=c['title'], description=c['description'])
        r.gpu_util_avg, r.mem_util_avg, r.created_at = c['gpu_util_avg'], c['mem_util_avg'], at(c['created_at_us'])
        recs.append(r)
    cal = Calendar(recs, users={u: _User(u, name) for u, name in w['users']})
    dropin.bind(cal)
    for p, want in zip(w['probes'], w['expect']):
        content, status = dropin.get_selected(p['resources_ids'], p['start'], p['end'])
        assert status == want['status'], p
        assert content == want['content'], p
    cal.close()
