#!/usr/bin/env python
"""Three th_conflict_sets(TH_MODE_STRICT) calls on BASELINE config 2 with device arrays, for an ncu launch list."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import native, synth         # noqa: E402

dev = torch.device('cuda:0')
lib = native.load()
N, U, M = 1_000_000, 4096, int(os.environ.get('TH_SETS_M', '1000000'))
cal = synth.make_calendar(N, U, seed=20260102)
q = synth.make_queries(cal, M, seed=20260103)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
snap = th.Snapshot(0)
snap.load(*[t(cal[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')], n_uuids=U)
qd = [t(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
cap = 8 * M
off = torch.empty(M + 1, dtype=torch.int64, device=dev)
ids = torch.empty(cap, dtype=torch.int64, device=dev)
canc = torch.empty(cap, dtype=torch.uint8, device=dev)
needed = C.c_int64(0)
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
mode = int(os.environ.get('TH_SETS_MODE', '0'))
for _ in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    native.check(lib.th_conflict_sets(snap._h, M, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(), qd[3].data_ptr(),
                                      mode, off.data_ptr(), ids.data_ptr(), canc.data_ptr(), cap, C.byref(needed), 1, sp))
    b.record()
    torch.cuda.synchronize()
    print('conflict_sets mode %d: %.3f ms, %d rows' % (mode, a.elapsed_time(b), needed.value), flush=True)
