#!/usr/bin/env python
"""Cold-L2 time of th_conflict_bitmap in both modes (STRICT = would_interfere, INCLUSIVE = the range filter) on config 2."""
import ctypes as C, os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tensorhive_fixed_b200 as th
from tensorhive_fixed_b200 import native, synth
dev = torch.device('cuda:0'); lib = native.load()
cal = synth.make_calendar(1_000_000, 4096, seed=20260102); q = synth.make_queries(cal, 1_000_000, seed=20260103)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
snap = th.Snapshot(0); snap.load(*[t(cal[k]) for k in ('id','uuid_id','start','end','owner','cancelled')], n_uuids=4096)
qd = [t(q[k]) for k in ('uuid_id','start','end','self_id')]
words = torch.empty((1_000_000+31)//32, dtype=torch.int32, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for mode in (0, 1):
    ts = []
    for i in range(8):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); native.check(lib.th_conflict_bitmap(snap._h, 1_000_000, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(), qd[3].data_ptr() if mode == 0 else None, mode, words.data_ptr(), 1, sp)); b.record()
        torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    print('mode', mode, 'cold us min %.1f median %.1f' % (min(ts[2:]), float(np.median(ts[2:]))), 'hits', int(th.api.unpack_bits(words, 1_000_000).sum()))
