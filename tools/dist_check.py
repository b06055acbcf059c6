#!/usr/bin/env python
"""Multi-GPU parity check, one rank per GPU (launch with torchrun): the calendar is sharded by uuid, a replicated
query batch is routed, every rank runs K2 on its shard, the bitmap words are all-gathered over NCCL and the decoded
verdict vector must equal the brute-force oracle over the WHOLE calendar on every rank.  The per-uuid owner table (K3)
and the window averages (K4) are gathered and checked the same way."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                                  # noqa: E402
from tensorhive_fixed_b200 import sharding, synth                   # noqa: E402
from oracle import brute                                            # noqa: E402  (checker only)


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    n, u, m = 300_000, 1000, 200_000
    cal = synth.make_calendar(n, u, seed=77, skew=0.7)
    q = synth.make_queries(cal, m, seed=78)
    bounds = sharding.partition_uuids(np.bincount(cal['uuid_id'], minlength=u), world)
    mine = sharding.shard_rows(cal, bounds, rank)
    snap = th.Snapshot(local)
    snap.load(mine['id'], mine['uuid_id'], mine['start'], mine['end'], mine['owner'], mine['cancelled'],
              n_uuids=mine['n_uuids'])

    def local_words(uu, s, e, f):
        return snap.conflict_words(uu, s, e, f)

    got = sharding.sharded_conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'], bounds, rank, local_words,
                                           lambda w: sharding.torch_all_gather(w, 'cuda:%d' % local))
    want = brute.conflict_bits(cal, q['uuid_id'], q['start'], q['end'], q['self_id'])
    ok = bool(np.array_equal(got, want))
    # the same batch through the fused kernel: K2 stores its words into every rank's buffer over NVLink
    router = sharding.Router(q['uuid_id'], bounds)
    px = sharding.PeerExchange(local, world, rank, int(router.counts.max()))
    dev = 'cuda:%d' % local
    idx = router.local_indices(rank)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    qd = [t(router.local_uuid(q['uuid_id'], rank)), t(q['start'][idx]), t(q['end'][idx]), t(q['self_id'][idx])]
    for step in range(4):
        g = px.conflict_words(snap, *qd, counts=router.counts)
        torch.cuda.synchronize()
        px.check()
        fused = router.decode(g.cpu().numpy())
        ok = ok and bool(np.array_equal(fused, want))
    # ragged steps through the same peer group: odd batch sizes, ranks with very different shares, ranks with NO query at
    # all (their kernel only receives), back to back
    hot = np.flatnonzero((q['uuid_id'] >= bounds[0]) & (q['uuid_id'] < bounds[1]))          # candidates of rank 0 only
    last = np.flatnonzero(q['uuid_id'] >= bounds[world - 1])
    subsets = [np.arange(k) for k in (1, 31, 33, 1000, 4097, 77_777)] + [hot[:5000], last[:37], hot[:1], np.arange(0)]
    for sub in subsets:
        r2 = sharding.Router(q['uuid_id'][sub], bounds)
        i2 = sub[r2.local_indices(rank)]
        q2 = [t(r2.local_uuid(q['uuid_id'][sub], rank)), t(q['start'][i2]), t(q['end'][i2]), t(q['self_id'][i2])]
        g = px.conflict_words(snap, *q2, counts=r2.counts)
        torch.cuda.synchronize()
        px.check()
        ok = ok and bool(np.array_equal(r2.decode(g.cpu().numpy()), want[sub]))
    # the other two gathers of section 8e: per-uuid (cur_res_id, cur_owner) from K3 and window averages from K4
    now = int(np.median(cal['start']))
    gather64 = lambda v: sharding.torch_all_gather_i64(v, dev)
    res, own = sharding.sharded_current_owner(bounds, rank, lambda: snap.current_owner(now), gather64)
    want_res, want_own = brute.current_owner(cal, u, now)
    ok = ok and bool(np.array_equal(res, want_res)) and bool(np.array_equal(own, want_own))
    smp = synth.make_samples(u, 400, seed=79, ragged=True)
    win = synth.make_windows(smp, 3000, seed=80)
    ms = sharding.shard_samples(smp, bounds, rank)
    ag, am = sharding.sharded_window_averages(
        win['uuid_id'], win['start'], win['end'], bounds, rank,
        lambda wu, ws, we: th.usage_reduce(ms['offsets'], ms['ts'], ms['gpu'], ms['mem'], wu, ws, we, device=local),
        gather64)
    sg, sm, cnt = brute.usage_reduce(smp['offsets'], smp['ts'], smp['gpu'], smp['mem'], win['uuid_id'], win['start'],
                                     win['end'])
    ok = ok and bool(np.array_equal(ag, [brute.floor_avg(a, c) for a, c in zip(sg, cnt)]))
    ok = ok and bool(np.array_equal(am, [brute.floor_avg(a, c) for a, c in zip(sm, cnt)]))
    dist.barrier()
    px.close()
    flag = torch.tensor([1 if ok else 0], device='cuda:%d' % local)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print('DIST_CHECK world=%d rows_per_rank=%s verdicts=%d conflicts=%d ok=%s' % (
            world, [int(((cal['uuid_id'] >= bounds[r]) & (cal['uuid_id'] < bounds[r + 1])).sum()) for r in range(world)],
            m, int(want.sum()), bool(flag.item())), flush=True)
    snap.close()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == '__main__':
    main()
