#!/usr/bin/env python
"""Times K4 (th_usage_reduce and th_usage_reduce_u8) on BASELINE config 3 and checks that both sample formats give the
same sums.  TH_SWEEP_SCALE scales the workload; TH_SWEEP_GAPPY=1 makes the log irregular (a fifth of the samples
missing, 0.9 s of jitter)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tensorhive_fixed_b200 import native, synth         # noqa: E402

SCALE = float(os.environ.get('TH_SWEEP_SCALE', '1.0'))
dev = torch.device('cuda:0')
lib = native.load()
U, SPU = 65_536, int(15_259 * SCALE)
W = int(10_000_000 * SCALE)
S = U * SPU
i64 = lambda n: torch.empty(n, dtype=torch.int64, device=dev)
off, ts, gpu, mem = i64(U + 1), i64(S), i64(S), i64(S)
wu, ws, we = i64(W), i64(W), i64(W)
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
native.check(lib.th_synth_samples(0, U, SPU, synth.T0_US, 20260104, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(),
                                  mem.data_ptr(), sp))
native.check(lib.th_synth_windows(0, U, SPU, synth.T0_US, 20260105, W, wu.data_ptr(), ws.data_ptr(), we.data_ptr(), sp))
torch.cuda.synchronize()


if os.environ.get('TH_SWEEP_GAPPY'):      # an irregular log: a fifth of the samples missing at random, 0.9 s of jitter
    keep = torch.rand(S, device=dev) >= 0.2
    ts = (ts + torch.randint(0, 900_000, (S,), device=dev, dtype=torch.int64))[keep].contiguous()
    gpu, mem = gpu[keep].contiguous(), mem[keep].contiguous()
    off = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), keep.view(U, SPU).sum(1).cumsum(0)]).contiguous()
    S = int(ts.numel())
    del keep
g8, m8 = gpu.to(torch.uint8), mem.to(torch.uint8)


def run(outs, narrow=False):
    if narrow:
        native.check(lib.th_usage_reduce_u8(0, S, U, off.data_ptr(), ts.data_ptr(), g8.data_ptr(), m8.data_ptr(), W,
                                            wu.data_ptr(), ws.data_ptr(), we.data_ptr(), outs[0].data_ptr(),
                                            outs[1].data_ptr(), outs[2].data_ptr(), outs[3].data_ptr(), outs[4].data_ptr(),
                                            1, sp))
        return
    native.check(lib.th_usage_reduce(0, S, U, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(), mem.data_ptr(), W,
                                     wu.data_ptr(), ws.data_ptr(), we.data_ptr(), outs[0].data_ptr(), outs[1].data_ptr(),
                                     outs[2].data_ptr(), outs[3].data_ptr(), outs[4].data_ptr(), 1, sp))


ref = None
for narrow in (False, True):
    outs = [i64(W).fill_(-7) for _ in range(5)]
    for _ in range(2):
        run(outs, narrow)
    times = []
    for _ in range(6):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(outs, narrow); b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    if ref is None:
        ref = outs
        same = True
    else:
        same = all(bool(torch.equal(x, y)) for x, y in zip(ref, outs))
    print('K4 %-5s columns: ms min %.3f median %.3f  identical=%s' % (
        'u8' if narrow else 'int64', min(times), float(np.median(times)), same), flush=True)
