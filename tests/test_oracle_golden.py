"""The committed oracle vectors (tests/golden/oracle_vectors.json, written by the SQL oracle) against the second,
independent CPU restatement (oracle/brute.py).  Neither form is pinned by a test of the reference for these rows
(SURVEY.md section 8c); the fixture freezes what the reference's source text says."""
import json
import os

import numpy as np
import pytest

from oracle import brute

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, 'golden', 'oracle_vectors.json')) as f:
    DOC = json.load(f)


def cal_of(w):
    r = w['rows']
    return {'id': np.array([x[0] for x in r], np.int64), 'uuid_id': np.array([x[1] for x in r], np.int64),
            'start': np.array([x[2] for x in r], np.int64), 'end': np.array([x[3] for x in r], np.int64),
            'owner': np.array([x[4] for x in r], np.int64),
            'cancelled': np.array([1 if x[5] else 0 for x in r], np.uint8), 'n_uuids': w['n_uuids']}


def queries_of(w):
    q = np.array(w['queries'], np.int64).reshape(-1, 4)
    return q[:, 0].copy(), q[:, 1].copy(), q[:, 2].copy(), q[:, 3].copy()


@pytest.mark.parametrize('w', DOC['worlds'], ids=lambda w: 'seed%d' % w['seed'])
def test_brute_force_oracle_reproduces_the_sql_vectors(w):
    cal, a = cal_of(w), w['answers']
    qu, qs, qe, qf = queries_of(w)
    assert [bool(b) for b in brute.conflict_bits(cal, qu, qs, qe, qf)] == a['would_interfere']
    assert brute.conflict_sets(cal, qu, qs, qe, qf, brute.STRICT) == a['strict_lists']
    assert brute.conflict_sets(cal, qu, qs, qe, qf, brute.INCLUSIVE) == a['range_lists']
    uu = np.arange(w['n_uuids'], dtype=np.int64)
    for now in w['nows']:
        res, _own = brute.current_owner(cal, w['n_uuids'], now)
        assert list(res) == a['current'][str(now)]
        for p in w['periods']:
            k = len(uu)
            got = brute.conflict_sets(cal, uu, np.full(k, now), np.full(k, now + p), np.full(k, -1), brute.UPCOMING)
            assert got == a['upcoming']['%d+%d' % (now, p)]
    assert any(a['would_interfere']) and not all(a['would_interfere'])


@pytest.mark.parametrize('c', DOC['usage'], ids=['usage0', 'usage1'])
def test_usage_vectors(c):
    win = np.array(c['windows'], np.int64)
    sg, sm, cnt = brute.usage_reduce(c['offsets'], c['ts'], c['gpu'], c['mem'], win[:, 0], win[:, 1], win[:, 2])
    assert list(cnt) == c['count']
    assert [brute.floor_avg(s, n) for s, n in zip(sg, cnt)] == c['avg_gpu']
    assert [brute.floor_avg(s, n) for s, n in zip(sm, cnt)] == c['avg_mem']
    assert -1 in c['avg_gpu'] and max(c['avg_gpu']) > 0
