#!/usr/bin/env python
"""Multi-GPU parity check, one rank per GPU (launch with torchrun): the calendar is sharded by uuid, a replicated
query batch is routed, every rank runs K2 on its shard, the bitmap words are all-gathered over NCCL and the decoded
verdict vector must equal the brute-force oracle over the WHOLE calendar on every rank."""
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
