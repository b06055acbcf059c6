#!/usr/bin/env python
"""Short launch sequence for ncu: one rank's share of BASELINE config 5 (12.5 M rows x 8192 uuids, 12.5 M candidates),
four flushed K2 launches.  TH_PROF_N / TH_PROF_U / TH_PROF_M change the size.  Run plain first, then under ncu."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import native, synth         # noqa: E402

N = int(os.environ.get('TH_PROF_N', '12500000'))
U = int(os.environ.get('TH_PROF_U', '8192'))
M = int(os.environ.get('TH_PROF_M', '12500000'))
dev = torch.device('cuda:0')
lib = native.load()
KEYS = ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')
if N <= 16_000_000:
    cal = synth.make_calendar(N, U, seed=20260102)
    q = synth.make_queries(cal, M, seed=20260103)
else:                                                   # too many rows for one walk in int64 us: chunk by chunk
    g = synth.GlobalCalendar(N, U, seed=20260102)
    cparts, qparts = [], []
    for c in range(g.n_chunks):
        cc = g.chunk(c)
        cparts.append(cc)
        qparts.append(g.chunk_queries(c, int(round(M * g.chunk_rows[c] / g.n_rows)), seed=20260103, cal=cc))
    cal = synth.concat_columns(cparts, KEYS)
    q = synth.concat_columns(qparts, ('uuid_id', 'start', 'end', 'self_id'))
    perm = np.random.default_rng(1).permutation(len(q['uuid_id']))        # candidates arrive in no particular order
    q = {k: v[perm] for k, v in q.items()}
    N, M = len(cal['id']), len(q['uuid_id'])
    del cparts, qparts
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
snap = th.Snapshot(0)
snap.load(*[t(cal[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')], n_uuids=U)
qd = [t(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
words = torch.empty((M + 31) // 32, dtype=torch.int32, device=dev)
flush = torch.zeros(64 << 20, dtype=torch.int64, device=dev)
stream = torch.cuda.current_stream()
sp = C.c_void_p(stream.cuda_stream)
ev = []
for _ in range(4):
    flush.sum()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    native.check(lib.th_conflict_bitmap(snap._h, M, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(), qd[3].data_ptr(),
                                        0, words.data_ptr(), 1, sp))
    b.record(stream)
    ev.append((a, b))
torch.cuda.synchronize()
print('k2 us per launch:', ['%.1f' % (1e3 * a.elapsed_time(b)) for a, b in ev], 'conflicts:',
      int(th.api.unpack_bits(words, M).sum()), 'lib:', native.LIB_PATH)

# TH_PROF_CHECK=1: the verdicts of every candidate of 24 sampled uuids against the brute-force oracle over those uuids' rows
if os.environ.get('TH_PROF_CHECK'):
    from oracle import brute
    rng = np.random.default_rng(5)
    pick = np.sort(rng.choice(U, size=min(24, U), replace=False))
    rsel = np.isin(cal['uuid_id'], pick)
    qsel = np.flatnonzero(np.isin(q['uuid_id'], pick))
    sub = {k: cal[k][rsel] for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')}
    sub['n_uuids'] = U
    want = brute.conflict_bits(sub, q['uuid_id'][qsel], q['start'][qsel], q['end'][qsel], q['self_id'][qsel])
    got = th.api.unpack_bits(words, M)
    got = got.cpu().numpy() if hasattr(got, 'cpu') else np.asarray(got)
    bad = int((got[qsel].astype(bool) != np.asarray(want, bool)).sum())
    print('parity on %d uuids: %d rows, %d candidates, %d conflicts expected, %d mismatches' % (
        len(pick), int(rsel.sum()), len(qsel), int(np.asarray(want, bool).sum()), bad))
    info = snap.info()
    print('snapshot:', info)

# TH_PROF_ORDERS=group8,group32,group128,sorted: the same candidates presented in another order (reordered outside the
# timed region).  groupG = bucketed into G uuid ranges, random inside a bucket: what K2 would see AFTER a partition pass
# that blocks the index for L2 -- the upper bound of that lever, without the pass.  sorted = by (uuid, start): the
# order a whole-calendar revalidation presents.
for order in [o for o in os.environ.get('TH_PROF_ORDERS', '').split(',') if o]:
    if order == 'sorted':
        key = qd[0] * (1 << 42) + (qd[1] - qd[1].min())
    else:
        g = int(order[5:])
        key = qd[0] // ((U + g - 1) // g)
    perm = torch.sort(key, stable=True).indices
    qo = [c[perm].contiguous() for c in qd]
    del key, perm
    ev = []
    for _ in range(4):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        native.check(lib.th_conflict_bitmap(snap._h, M, qo[0].data_ptr(), qo[1].data_ptr(), qo[2].data_ptr(),
                                            qo[3].data_ptr(), 0, words.data_ptr(), 1, sp))
        b.record(stream)
        ev.append((a, b))
    torch.cuda.synchronize()
    print('order', order, 'k2 us per launch:', ['%.1f' % (1e3 * a.elapsed_time(b)) for a, b in ev], 'conflicts:',
          int(th.api.unpack_bits(words, M).sum()))
    del qo
