#!/usr/bin/env python
"""Times K2 (cold L2 / warm L2) on BASELINE config 2 for several time-grid densities."""
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
N = int(os.environ.get('TH_SWEEP_N', '1000000'))
U = int(os.environ.get('TH_SWEEP_U', '4096'))
M = int(os.environ.get('TH_SWEEP_M', '1000000'))
SKEW = float(os.environ.get('TH_SWEEP_SKEW', '0'))      # > 0: Zipf-like uuid popularity (hot GPUs hold most rows)
cal = synth.make_calendar(N, U, seed=20260102, skew=SKEW)
q = synth.make_queries(cal, M, seed=20260103)
if os.environ.get('TH_SWEEP_NOSELF'):
    q['self_id'][:] = -1
if os.environ.get('TH_SWEEP_SORTED'):      # queries grouped by uuid: what a caller iterating a calendar would send
    o = np.lexsort((q['end'], q['uuid_id']))
    q = {k: v[o] for k, v in q.items()}
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cols = [t(cal[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')]
qd = [t(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
words = torch.empty((M + 31) // 32, dtype=torch.int32, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
stream = torch.cuda.current_stream()
sp = C.c_void_p(stream.cuda_stream)
ref = None
VARIANTS = [v.split(':') for v in os.environ.get('TH_SWEEP_VARIANTS', '0:tma').split(',')]
for x4, (dbg, path) in [(int(v), var) for v in os.environ.get('TH_SWEEP_X4', '6').split(',') for var in VARIANTS]:
    os.environ['TH_GRID_CELLS_PER_ROW_X4'] = str(x4)
    os.environ['TH_K2_EXPERIMENT'] = dbg
    if path == 'direct':
        os.environ['TH_K2_DIRECT'] = '1'
    else:
        os.environ.pop('TH_K2_DIRECT', None)
    snap = th.Snapshot(0)
    snap.load(*cols, n_uuids=U)

    def run():
        cur = C.c_void_p(torch.cuda.current_stream().cuda_stream)      # the capture stream while a graph is recorded
        native.check(lib.th_conflict_bitmap(snap._h, M, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(),
                                            qd[3].data_ptr(), 0, words.data_ptr(), 1, cur))
    for _ in range(3):
        run()
    flush64 = flush.view(torch.int64)
    cold, cold_w = [], []
    for _ in range(10):      # cold L2, flushed by READING 512 MiB (clean lines: evicting them costs nothing)
        flush64.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); run(); b.record(stream)
        cold.append((a, b))
    torch.cuda.synchronize()
    cold = [a.elapsed_time(b) for a, b in cold]
    for _ in range(10):      # cold L2, flushed by WRITING (dirty lines must be written back during the kernel)
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); run(); b.record(stream)
        cold_w.append((a, b))
    torch.cuda.synchronize()
    cold_w = [a.elapsed_time(b) for a, b in cold_w]
    # warm L2: 20 launches in one CUDA graph so that no CPU launch latency sits between the events
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(20):
            run()
    gr.replay()
    warm = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); gr.replay(); b.record()
        torch.cuda.synchronize()
        warm.append(a.elapsed_time(b) / 20)
    w = words.cpu().numpy().copy()
    if ref is None:
        ref = w
    print('dbg=%s path=%s ' % (dbg, path), end='')
    print('x4=%2d bytes=%d cold_read_flush_us=%.1f (min %.1f) cold_write_flush_us=%.1f warm_graph_us=%.1f (min %.1f) same=%s' % (
        x4, snap.info()['device_bytes'], 1e3 * np.mean(cold), 1e3 * min(cold), 1e3 * np.mean(cold_w), 1e3 * np.mean(warm),
        1e3 * min(warm), np.array_equal(w, ref)), flush=True)
    snap.close()
