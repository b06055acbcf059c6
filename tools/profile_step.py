#!/usr/bin/env python
"""Short, fixed launch sequence for ncu: snapshot load (K1 + index build) on BASELINE config 2, three
flushed K2 launches, then three K4 launches on config 3 (TH_PROF_SCALE, default 1.0 = 1 B samples) and three of the
uint8 format.  Run plain first, then under ncu
(see /opt/skills/guides/B200_PROFILING.md); a number printed under ncu is never a bench value."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import native, synth         # noqa: E402

SCALE = float(os.environ.get('TH_PROF_SCALE', '1.0'))
dev = torch.device('cuda:0')
lib = native.load()
cal = synth.make_calendar(1_000_000, 4096, seed=20260102)
q = synth.make_queries(cal, 1_000_000, seed=20260103)
snap = th.Snapshot(0)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
snap.load(*[t(cal[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')], n_uuids=4096)
qd = [t(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
words = torch.empty((1_000_000 + 31) // 32, dtype=torch.int32, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    if not os.environ.get('TH_PROF_NOFLUSH'):     # warm-L2 captures: pair with ncu --cache-control none
        flush.fill_(1)
    native.check(lib.th_conflict_bitmap(snap._h, 1_000_000, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(),
                                        qd[3].data_ptr(), 0, words.data_ptr(), 1, sp))
torch.cuda.synchronize()
print('k2 conflicts:', int(th.api.unpack_bits(words, 1_000_000).sum()))

U, SPU = 65_536, int(15_259 * SCALE)
W = int(10_000_000 * SCALE)
S = U * SPU
i64 = lambda n: torch.empty(n, dtype=torch.int64, device=dev)
off, ts, gpu, mem = i64(U + 1), i64(S), i64(S), i64(S)
wu, ws, we = i64(W), i64(W), i64(W)
outs = [i64(W) for _ in range(5)]
native.check(lib.th_synth_samples(0, U, SPU, synth.T0_US, 5, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(), mem.data_ptr(), sp))
native.check(lib.th_synth_windows(0, U, SPU, synth.T0_US, 6, W, wu.data_ptr(), ws.data_ptr(), we.data_ptr(), sp))
for _ in range(3):
    native.check(lib.th_usage_reduce(0, S, U, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(), mem.data_ptr(), W,
                                     wu.data_ptr(), ws.data_ptr(), we.data_ptr(), *[o.data_ptr() for o in outs], 1, sp))
torch.cuda.synchronize()
print('k4 samples inside windows:', int(outs[2].sum().item()))
g8, m8 = gpu.to(torch.uint8), mem.to(torch.uint8)
del gpu, mem
for _ in range(3):
    native.check(lib.th_usage_reduce_u8(0, S, U, off.data_ptr(), ts.data_ptr(), g8.data_ptr(), m8.data_ptr(), W,
                                        wu.data_ptr(), ws.data_ptr(), we.data_ptr(), *[o.data_ptr() for o in outs], 1, sp))
torch.cuda.synchronize()
print('k4 u8 samples inside windows:', int(outs[2].sum().item()))
