#!/usr/bin/env python
"""How much of a fused step is the exchange itself?  One rank per GPU (torchrun): the fused K2 + all-gather step timed
for batches from 1 K to 1 M queries per rank, beside the plain kernel on the same batch.  The difference at 1 K queries is
the protocol's fixed cost (arrival skew after the device barrier + one NVLink flight + the poll)."""
import ctypes as C
import json
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
    lib = native.load()
    cal = synth.make_calendar(1_000_000, 4096, seed=100 + rank)
    q = synth.make_queries(cal, 1_000_000, seed=200 + rank)
    snap = th.Snapshot(local)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=4096)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    qd = [t(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
    px = sharding.PeerExchange(local, world, rank, 1_000_000)
    flush = torch.zeros(64 << 20, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)
    words = torch.empty(31250, dtype=torch.int32, device=dev)
    out = {}
    for m in (1024, 32768, 262144, 1_000_000):
        sub = [x[:m] for x in qd]

        def fused():
            px.conflict_words(snap, *sub, check=False)

        def plain():
            native.check(lib.th_conflict_bitmap(snap._h, m, sub[0].data_ptr(), sub[1].data_ptr(), sub[2].data_ptr(),
                                                sub[3].data_ptr(), 0, words.data_ptr(), 1, sp))

        def timed(fn, meet, cold):
            for _ in range(5):
                fn()
            ev = []
            for _ in range(30):
                if cold:
                    flush.sum()
                if meet:
                    px.barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream); fn(); b.record(stream)
                ev.append((a, b))
            torch.cuda.synchronize()
            return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3

        dist.barrier()
        r = {'fused_cold_us': timed(fused, True, True), 'plain_cold_us': timed(plain, False, True),
             'fused_warm_us': timed(fused, True, False), 'plain_warm_us': timed(plain, False, False)}
        px.check()
        tt = torch.tensor([r[k] for k in sorted(r)], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        out[m] = dict(zip(sorted(r), [round(float(x), 2) for x in tt.tolist()]))
    if rank == 0:
        print(json.dumps({'world': world, 'max_over_ranks': out}), flush=True)
    dist.barrier()
    px.close()
    snap.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
