#!/usr/bin/env python
"""Where does a fused step spend its time?  Per-CTA globaltimer stamps (TH_PEER_TRACE) of the fused K2 kernel on every
rank: entry, end of the CTA's tiles, words sent, words of all peers received.  One rank per GPU (torchrun)."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                                  # noqa: E402
from tensorhive_fixed_b200 import sharding, synth, native           # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    cal = synth.make_calendar(1_000_000, 4096, seed=100 + rank)
    q = synth.make_queries(cal, 1_000_000, seed=200 + rank)
    snap = th.Snapshot(local)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=4096)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    qd = [t(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
    px = sharding.PeerExchange(local, world, rank, 1_000_000)
    flush = torch.zeros(64 << 20, dtype=torch.int64, device=dev)
    trace = torch.zeros(4 * 2048, dtype=torch.int64, device=dev)
    os.environ['TH_PEER_TRACE'] = str(trace.data_ptr())
    for m in (32768, 1_000_000):
        sub = [x[:m] for x in qd]
        rows = []
        for step in range(12):
            flush.sum()
            px.barrier()
            px.conflict_words(snap, *sub, check=False)
            torch.cuda.synchronize()
            tr = trace.cpu().numpy().reshape(-1, 4)
            tr = tr[tr[:, 0] > 0]
            trace.zero_()
            if step < 4:
                continue
            t0 = tr[:, 0].min()
            rows.append([tr[:, 0].max() - t0, np.median(tr[:, 1] - t0), (tr[:, 1] - t0).max(), (tr[:, 2] - t0).max(),
                         np.median(tr[:, 3] - tr[:, 2]), (tr[:, 3] - tr[:, 2]).max(), (tr[:, 3] - t0).max(), len(tr)])
        r = np.median(np.array(rows, float), axis=0)
        print('rank %d m=%7d ctas=%d | entry spread %.1f us | tiles done: median %.1f max %.1f | sent by %.1f | wait for peers: '
              'median %.1f max %.1f | all received by %.1f us' % (rank, m, int(r[7]), r[0] / 1e3, r[1] / 1e3, r[2] / 1e3, r[3] / 1e3,
                                                                  r[4] / 1e3, r[5] / 1e3, r[6] / 1e3), flush=True)
        dist.barrier()
    px.check()
    px.close()
    snap.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
