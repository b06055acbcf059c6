"""The kernels against the committed oracle vectors (tests/golden/oracle_vectors.json), through the C ABI."""
import numpy as np
import pytest

from helpers import csr_to_lists
from test_oracle_golden import DOC, cal_of, queries_of

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('w', DOC['worlds'], ids=lambda w: 'seed%d' % w['seed'])
def test_kernels_reproduce_the_committed_vectors(lib_built, w):
    import tensorhive_fixed_b200 as th
    cal, a = cal_of(w), w['answers']
    u = w['n_uuids']
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    qu, qs, qe, qf = queries_of(w)
    assert [bool(b) for b in snap.conflict_bitmap(qu, qs, qe, qf)] == a['would_interfere']
    off, ids, _ = snap.conflict_sets(qu, qs, qe, qf, th.MODE_STRICT)
    assert csr_to_lists(off, ids) == a['strict_lists']
    off, ids, _ = snap.conflict_sets(qu, qs, qe, None, th.MODE_INCLUSIVE)
    assert csr_to_lists(off, ids) == a['range_lists']
    assert [bool(b) for b in snap.conflict_bitmap(qu, qs, qe, None, th.MODE_INCLUSIVE)] == [bool(x) for x in a['range_lists']]
    uu = np.arange(u, dtype=np.int64)
    for now in w['nows']:
        res, _own = snap.current_owner(now)
        assert list(res) == a['current'][str(now)]
        for p in w['periods']:
            want = a['upcoming']['%d+%d' % (now, p)]
            off, ids, _ = snap.conflict_sets(uu, np.full(u, now), np.full(u, now + p), None, th.MODE_UPCOMING)
            assert csr_to_lists(off, ids) == want
            fid, _fst, _fow, cnt = snap.upcoming(now, p, uu)
            assert list(cnt) == [len(x) for x in want] and list(fid) == [(x[0] if x else -1) for x in want]
    snap.close()


@pytest.mark.parametrize('c', DOC['usage'], ids=['usage0', 'usage1'])
def test_usage_kernels_reproduce_the_committed_vectors(lib_built, c):
    import tensorhive_fixed_b200 as th
    win = np.array(c['windows'], np.int64)
    off, ts = np.array(c['offsets'], np.int64), np.array(c['ts'], np.int64)
    gpu, mem = np.array(c['gpu'], np.int64), np.array(c['mem'], np.int64)
    for outs in (th.usage_reduce(off, ts, gpu, mem, win[:, 0].copy(), win[:, 1].copy(), win[:, 2].copy()),
                 th.usage_reduce_u8(off, ts, gpu.astype(np.uint8), mem.astype(np.uint8), win[:, 0].copy(),
                                    win[:, 1].copy(), win[:, 2].copy())):
        assert list(outs[2]) == c['count'] and list(outs[3]) == c['avg_gpu'] and list(outs[4]) == c['avg_mem']
