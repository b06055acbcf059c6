"""Sharding the calendar by GPU uuid over the ranks of one node.

Every predicate of the path is keyed by ``resource_id``: rows of different uuids never interact, so rank r
owns a contiguous range of uuid codes (balanced by row count, not by uuid count) together with the
reservations, samples and processes of those uuids, and no data-path collective is needed to *compute*.
The one real exchange is the result: every service thread wants the verdicts of the whole batch, so the
per-rank bitmap words are all-gathered (torch.distributed: NCCL over NVLink on the GPU box, gloo in the CPU
tests).  Payloads are tiny (1 M queries = 125 KB of bits), i.e. latency- not bandwidth-bound.

The batch is replicated on every rank (it arrives from the same REST/service caller), so each rank can
derive the full routing table locally and decode what the others send without any metadata exchange.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np


def partition_uuids(rows_per_uuid: Sequence[int], world: int) -> np.ndarray:
    """Contiguous uuid ranges with (nearly) equal row counts.  Returns boundaries[world + 1] with
    boundaries[0] = 0 and boundaries[world] = n_uuids; rank r owns [boundaries[r], boundaries[r + 1])."""
    counts = np.asarray(rows_per_uuid, np.int64)
    n_uuids = counts.size
    csum = np.concatenate([[0], np.cumsum(counts)])
    total = int(csum[-1])
    bounds = np.zeros(world + 1, np.int64)
    bounds[world] = n_uuids
    for r in range(1, world):
        target = total * r / world
        # first uuid boundary whose prefix reaches the target, never moving backwards
        b = int(np.searchsorted(csum, target, side='left'))
        if b > 0 and abs(csum[b - 1] - target) <= abs(csum[min(b, n_uuids)] - target):
            b -= 1
        bounds[r] = min(max(b, bounds[r - 1]), n_uuids)
    return bounds


def owner_rank(uuid_codes, bounds: np.ndarray) -> np.ndarray:
    """Rank that owns each uuid code; codes outside [0, n_uuids) (unknown resources) go to rank 0, which
    answers "no reservations" for them like any other rank would."""
    u = np.asarray(uuid_codes, np.int64)
    r = np.searchsorted(bounds, u, side='right') - 1
    world = bounds.size - 1
    return np.where((u < 0) | (u >= bounds[-1]), 0, np.clip(r, 0, world - 1)).astype(np.int64)


class Router:
    """Routing of one replicated query batch: which rank evaluates which query, and where each gathered
    bit goes back to."""

    def __init__(self, q_uuid, bounds: np.ndarray):
        self.bounds = np.asarray(bounds, np.int64)
        self.world = self.bounds.size - 1
        self.m = int(len(q_uuid))
        self.owner = owner_rank(q_uuid, self.bounds)
        self.perm = np.argsort(self.owner, kind='stable')            # queries grouped by rank, input order kept
        self.counts = np.bincount(self.owner, minlength=self.world).astype(np.int64)
        self.offsets = np.concatenate([[0], np.cumsum(self.counts)])
        self.words_per_rank = int((self.counts.max() + 31) // 32) if self.m else 0

    def local_indices(self, rank: int) -> np.ndarray:
        return self.perm[self.offsets[rank]:self.offsets[rank + 1]]

    def local_uuid(self, q_uuid, rank: int) -> np.ndarray:
        """uuid codes of this rank's queries, rebased to the shard-local dictionary (unknown stay unknown)."""
        u = np.asarray(q_uuid, np.int64)[self.local_indices(rank)]
        lo, hi = self.bounds[rank], self.bounds[rank + 1]
        return np.where((u >= lo) & (u < hi), u - lo, -1)

    def decode(self, gathered_words: np.ndarray) -> np.ndarray:
        """gathered_words[world, words_per_rank] (uint32) -> bool[m] in the batch's original order."""
        out = np.zeros(self.m, bool)
        g = np.ascontiguousarray(gathered_words).view(np.uint32).reshape(self.world, -1)
        for r in range(self.world):
            c = int(self.counts[r])
            if c == 0:
                continue
            bits = np.unpackbits(g[r].view(np.uint8), bitorder='little')[:c].astype(bool)
            out[self.local_indices(r)] = bits
        return out


def sharded_conflict_bitmap(q_uuid, q_start, q_end, q_self, bounds, rank: int,
                            local_words: Callable[[np.ndarray, np.ndarray, np.ndarray, np.ndarray], np.ndarray],
                            all_gather: Callable[[np.ndarray], np.ndarray]) -> np.ndarray:
    """One replicated batch -> verdicts for the whole batch on every rank.

    ``local_words(uuid_local, start, end, self_id)`` runs K2 on this rank's snapshot and returns bitmap
    words; ``all_gather(words_padded)`` returns the [world, words] stack (see :func:`torch_all_gather`)."""
    router = Router(q_uuid, bounds)
    idx = router.local_indices(rank)
    words = np.zeros(router.words_per_rank, np.uint32)
    if idx.size:
        w = local_words(router.local_uuid(q_uuid, rank), np.asarray(q_start, np.int64)[idx],
                        np.asarray(q_end, np.int64)[idx], np.asarray(q_self, np.int64)[idx])
        w = np.asarray(w).view(np.uint32)
        words[:w.size] = w
    return router.decode(all_gather(words))


def torch_all_gather(words: np.ndarray, device: Optional[str] = None) -> np.ndarray:
    """All-gather equal-length uint32 word vectors over the default process group."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    t = torch.from_numpy(words.view(np.int32).copy())
    if device is not None:
        t = t.to(device)
    out = torch.empty(world * t.numel(), dtype=torch.int32, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return out.cpu().numpy().view(np.uint32).reshape(world, -1)


def shard_rows(cal: dict, bounds: np.ndarray, rank: int) -> dict:
    """The rows of ``cal`` this rank owns, with uuid codes rebased to the shard."""
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    sel = (cal['uuid_id'] >= lo) & (cal['uuid_id'] < hi)
    out = {k: (v[sel] if isinstance(v, np.ndarray) and v.shape[:1] == sel.shape else v) for k, v in cal.items()}
    out['uuid_id'] = out['uuid_id'] - lo
    out['n_uuids'] = max(hi - lo, 1)
    return out
