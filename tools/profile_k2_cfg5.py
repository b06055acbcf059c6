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
cal = synth.make_calendar(N, U, seed=20260102)
q = synth.make_queries(cal, M, seed=20260103)
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
