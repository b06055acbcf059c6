#!/usr/bin/env python
"""The device-resident sample ring at the FULL size of BASELINE config 3 (65 536 uuids x 15 259 samples = 1.0 B samples,
10 M windows), stamps on an exact 1 s grid: th_ring_reduce locates both window ends by division and never reads the ts
column.  (bench.py runs this leg at 1/8 scale, because building a 1 B-sample ring through th_ring_append -- host arrays,
as a monitor would deliver them -- takes about half a minute.)  A sample of the windows is checked against numpy sums.
TH_RING_U / TH_RING_SPU / TH_RING_W change the size."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tensorhive_fixed_b200 import native, synth         # noqa: E402

U = int(os.environ.get('TH_RING_U', '65536'))
SPU = int(os.environ.get('TH_RING_SPU', '15259'))
W = int(os.environ.get('TH_RING_W', '10000000'))
BATCH = 256                                              # ticks per th_ring_append call
dev = torch.device('cuda:0')
lib = native.load()
stream = torch.cuda.current_stream()
sp = C.c_void_p(stream.cuda_stream)
rng = np.random.default_rng(77)

h = C.c_void_p(0)
native.check(lib.th_ring_create(0, SPU, C.byref(h)))
g_all = np.empty((U, SPU), np.uint8)                     # uuid-major copy for the check
m_all = np.empty((U, SPU), np.uint8)
t_build = time.perf_counter()
for t0 in range(0, SPU, BATCH):
    nt = min(BATCH, SPU - t0)
    uu = np.tile(np.arange(U, dtype=np.int64), nt)       # tick-major, as a monitor reports
    ts = np.repeat(synth.T0_US + (t0 + np.arange(nt, dtype=np.int64)) * 1_000_000, U)
    g = rng.integers(0, 101, size=uu.size, dtype=np.uint8)
    m = rng.integers(0, 101, size=uu.size, dtype=np.uint8)
    native.check(lib.th_ring_append(h, uu.size, uu.ctypes.data, ts.ctypes.data, g.ctypes.data, m.ctypes.data, None))
    g_all[:, t0:t0 + nt] = g.reshape(nt, U).T
    m_all[:, t0:t0 + nt] = m.reshape(nt, U).T
t_build = time.perf_counter() - t_build
info = [C.c_int64(0) for _ in range(6)]
native.check(lib.th_ring_info(h, *[C.byref(x) for x in info]))
print('ring: uuids %d samples %d capacity %d device_bytes %d regular_uuids %d relayouts %d (built in %.1f s)' % (
    *[x.value for x in info], t_build))

i64 = lambda n: torch.empty(n, dtype=torch.int64, device=dev)
rw = [i64(W) for _ in range(3)]
native.check(lib.th_synth_windows(0, U, SPU, synth.T0_US, 20260107, W, rw[0].data_ptr(), rw[1].data_ptr(), rw[2].data_ptr(), sp))
outs = [i64(W) for _ in range(5)]
flush = torch.zeros(64 << 20, dtype=torch.int64, device=dev)


def step():
    native.check(lib.th_ring_reduce(h, W, rw[0].data_ptr(), rw[1].data_ptr(), rw[2].data_ptr(),
                                    *[o.data_ptr() for o in outs], 1, sp))


for _ in range(3):
    step()
ev = []
for _ in range(10):
    flush.sum()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream); step(); b.record(stream)
    ev.append((a, b))
torch.cuda.synchronize()
ms = [a.elapsed_time(b) for a, b in ev]
alg = 2 * int(outs[2].sum().item()) + 48 * W
print('th_ring_reduce, %d windows over %d samples, exact cadence: median %.3f ms (min %.3f)  -> %.0f GB/s on 2 B/sample + '
      '48 B/window = %.2f GB' % (W, U * SPU, float(np.median(ms)), min(ms), alg / np.median(ms) / 1e6, alg / 1e9))

# check a sample of the windows against numpy
pick = rng.integers(0, W, size=4000)
wu, ws, we = (x[torch.from_numpy(pick).to(dev)].cpu().numpy() for x in rw)
got = [o[torch.from_numpy(pick).to(dev)].cpu().numpy() for o in outs]
bad = 0
for j in range(len(pick)):
    u = int(wu[j])
    if not 0 <= u < U:
        bad += int(got[2][j] != 0)
        continue
    a = max(0, -(-(int(ws[j]) - synth.T0_US) // 1_000_000))
    b = min(SPU, (int(we[j]) - synth.T0_US) // 1_000_000 + 1) if int(we[j]) >= synth.T0_US else 0
    b = max(a, b)
    sg, sm, c = int(g_all[u, a:b].sum(dtype=np.int64)), int(m_all[u, a:b].sum(dtype=np.int64)), b - a
    ok = (got[0][j], got[1][j], got[2][j]) == (sg, sm, c) and got[3][j] == (sg // c if c else -1) and got[4][j] == (sm // c if c else -1)
    bad += 0 if ok else 1
print('checked %d windows against numpy: %d mismatches' % (len(pick), bad))
lib.th_ring_destroy(h)
sys.exit(1 if bad else 0)
