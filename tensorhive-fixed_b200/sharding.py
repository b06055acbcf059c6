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
    world = bounds.size - 1
    # a handful of comparisons against the inner boundaries (world <= 16) beats a binary search per element; counted in
    # one byte per query while the ranks fit one (1 M queries, 8 ranks: 7 ms against 19 ms in int64)
    r = np.zeros(u.shape, np.uint8 if world <= 255 else np.int64)
    for b in bounds[1:world]:
        r += (u >= b).view(np.uint8) if r.dtype == np.uint8 else (u >= b)
    r[(u < 0) | (u >= bounds[-1])] = 0
    return r


class Router:
    """Routing of one replicated query batch: which rank evaluates which query, and where each gathered
    bit goes back to."""

    def __init__(self, q_uuid, bounds: np.ndarray):
        self.bounds = np.asarray(bounds, np.int64)
        self.world = self.bounds.size - 1
        self.m = int(len(q_uuid))
        self.owner = owner_rank(q_uuid, self.bounds)
        # queries grouped by rank, input order kept: a stable argsort of `owner` (numpy sorts one-byte keys by counting:
        # 7 ms for 1 M queries; per-rank flatnonzero 18 ms; a comparison sort of int64 ranks 60 ms)
        if self.m == 0:
            self.perm = np.zeros(0, np.int64)
        elif self.owner.dtype == np.uint8:
            self.perm = np.argsort(self.owner, kind='stable')
        else:
            self.perm = np.concatenate([np.flatnonzero(self.owner == r) for r in range(self.world)])
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


def torch_all_gather_i64(values: np.ndarray, device: Optional[str] = None) -> np.ndarray:
    """All-gather equal-length int64 vectors over the default process group -> [world, len]."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    t = torch.from_numpy(np.ascontiguousarray(values, np.int64).copy())
    if device is not None:
        t = t.to(device)
    out = torch.empty(world * t.numel(), dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return out.cpu().numpy().reshape(world, -1)


def sharded_window_averages(win_uuid, win_start, win_end, bounds, rank: int,
                            local_reduce: Callable[[np.ndarray, np.ndarray, np.ndarray], Tuple],
                            all_gather_i64: Callable[[np.ndarray], np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """Replicated window list (one per reservation) -> (avg_gpu, avg_mem) of every window on every rank.

    The samples of a uuid live on the rank that owns the uuid, so rank r reduces only its windows:
    ``local_reduce(uuid_local, start, end)`` runs K4 on this rank's sample shard and returns
    ``(sum_gpu, sum_mem, count)``; the floor division of UsageLoggingService.avg (``sum // count``, -1 when the
    window holds no sample) is done here and the W x 2 int64 averages are all-gathered (section 8e)."""
    router = Router(win_uuid, bounds)
    idx = router.local_indices(rank)
    cap = int(router.counts.max()) if router.m else 0
    mine = np.full((cap, 2), -1, np.int64)
    if idx.size:
        sg, sm, cnt = (np.asarray(a, np.int64) for a in local_reduce(
            router.local_uuid(win_uuid, rank), np.asarray(win_start, np.int64)[idx], np.asarray(win_end, np.int64)[idx])[:3])
        some = cnt > 0
        safe = np.where(some, cnt, 1)
        mine[:idx.size, 0] = np.where(some, sg // safe, -1)
        mine[:idx.size, 1] = np.where(some, sm // safe, -1)
    g = all_gather_i64(mine.reshape(-1)).reshape(router.world, cap, 2)
    avg_gpu = np.full(router.m, -1, np.int64)
    avg_mem = np.full(router.m, -1, np.int64)
    for r in range(router.world):
        c = int(router.counts[r])
        if c:
            avg_gpu[router.local_indices(r)] = g[r, :c, 0]
            avg_mem[router.local_indices(r)] = g[r, :c, 1]
    return avg_gpu, avg_mem


def sharded_current_owner(bounds, rank: int, local_current: Callable[[], Tuple[np.ndarray, np.ndarray]],
                          all_gather_i64: Callable[[np.ndarray], np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """Per-uuid ``(cur_res_id, cur_owner)`` of the WHOLE calendar on every rank (the table the violation join
    and the slot check read): ``local_current()`` runs K3 on this rank's shard, the per-rank tables are
    all-gathered (padded to the widest shard) and laid back out in global uuid-code order."""
    bounds = np.asarray(bounds, np.int64)
    world = bounds.size - 1
    widths = np.diff(bounds)
    cap = int(widths.max()) if world else 0
    mine = np.full((2, cap), -1, np.int64)
    w = int(widths[rank])
    if w:
        res_id, owner = local_current()
        mine[0, :w] = np.asarray(res_id, np.int64)[:w]
        mine[1, :w] = np.asarray(owner, np.int64)[:w]
    g = all_gather_i64(mine.reshape(-1)).reshape(world, 2, cap)
    res_all = np.full(int(bounds[-1]), -1, np.int64)
    own_all = np.full(int(bounds[-1]), -1, np.int64)
    for r in range(world):
        lo, hi = int(bounds[r]), int(bounds[r + 1])
        res_all[lo:hi] = g[r, 0, :hi - lo]
        own_all[lo:hi] = g[r, 1, :hi - lo]
    return res_all, own_all


def shard_samples(samples: dict, bounds: np.ndarray, rank: int) -> dict:
    """The usage samples (CSR by uuid, ``synth.make_samples`` layout) this rank owns, offsets rebased to the shard."""
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    off = np.asarray(samples['offsets'], np.int64)
    a, b = int(off[lo]), int(off[hi])
    out = {k: samples[k][a:b] for k in ('ts', 'gpu', 'mem')}
    out['offsets'] = (off[lo:hi + 1] - a) if hi > lo else np.zeros(2, np.int64)
    out['n_uuids'] = max(hi - lo, 1)
    return out


def shard_rows(cal: dict, bounds: np.ndarray, rank: int) -> dict:
    """The rows of ``cal`` this rank owns, with uuid codes rebased to the shard."""
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    sel = (cal['uuid_id'] >= lo) & (cal['uuid_id'] < hi)
    out = {k: (v[sel] if isinstance(v, np.ndarray) and v.shape[:1] == sel.shape else v) for k, v in cal.items()}
    out['uuid_id'] = out['uuid_id'] - lo
    out['n_uuids'] = max(hi - lo, 1)
    return out


class _DevWords:
    """Library-owned device memory exposed through __cuda_array_interface__ (so torch/cupy can view it)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {'shape': (int(n),), 'typestr': '<i4', 'data': (int(ptr), False), 'version': 2}


class PeerExchange:
    """K2 fused with the all-gather of its bitmap words over peer memory (NVLink / NVSwitch), one process per GPU.

    Every rank creates its slot buffer inside the library, the 64-byte CUDA IPC handles are exchanged once through
    ``torch.distributed.all_gather_object`` (plumbing), and from then on a step is ONE call:
    ``words = px.conflict_words(snapshot, q_uuid, q_start, q_end, q_self)`` -- the kernel writes each rank's words,
    tagged with the step number, into every rank's slots and an unpack kernel hands back the compact words once all
    of them have arrived.  No NCCL call, no fence and no flag round trip per step.
    """

    def __init__(self, device: int, world: int, rank: int, max_queries_per_rank: int, connect: bool = True):
        import ctypes as C
        from . import native
        self._C, self._native = C, native
        self._lib = native.load()
        self.device, self.world, self.rank = int(device), int(world), int(rank)
        self.words_per_rank = (int(max_queries_per_rank) + 31) // 32
        self._h = C.c_void_p(0)
        self._views = {}
        self.handle = C.create_string_buffer(64)
        native.check(self._lib.th_peer_create(self.device, self.world, self.rank, self.words_per_rank,
                                              C.byref(self._h), self.handle))
        if connect:
            self.connect_over_process_group()

    def connect_over_process_group(self):
        import torch.distributed as dist
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(self.handle.raw))
        blob = self._C.create_string_buffer(b''.join(handles), 64 * self.world)
        self._native.check(self._lib.th_peer_connect(self._h, blob, None))
        dist.barrier()

    def local_base(self) -> int:
        p = self._C.c_void_p(0)
        self._native.check(self._lib.th_peer_local_base(self._h, self._C.byref(p)))
        return int(p.value)

    def connect_local(self, bases):
        """Single-process emulation: ``bases[r]`` = local_base() of the PeerExchange playing rank r."""
        arr = (self._C.c_void_p * self.world)(*[self._C.c_void_p(int(b)) for b in bases])
        self._native.check(self._lib.th_peer_connect(self._h, None, arr))

    def conflict_words(self, snapshot, q_uuid, q_start, q_end, q_self, mode: int = 0, wait: bool = True, counts=None,
                       check: bool = True, packed: bool = False):
        """int64 tensors in -- CUDA tensors, or page-locked host tensors (``.pin_memory()``), which the kernel streams
        over PCIe itself; returns an int32 CUDA tensor view [world, words_per_rank] of the gathered words (rank r's
        words for its own queries sit in row r).  ``counts[r]`` = number of queries rank r contributes this step
        (``Router.counts``); omit it when every rank has the same number.

        ``check=True`` (default) synchronises the stream and raises ``NativeError(TH_EINTERNAL)`` when a peer's words
        did not arrive in time (the words of such a step read as 0 and must not be used).  A caller that pipelines
        steps passes ``check=False`` and calls :meth:`check` before it consumes the words.  ``packed=True``: ``q_uuid`` and
        ``q_self`` are int32 tensors holding the uint32 columns of ``Snapshot.pack_queries`` (``q_self`` may be None)."""
        import torch
        C = self._C
        m = int(q_uuid.numel())
        out = C.c_void_p(0)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        cnt = None
        if counts is not None:
            cnt = (C.c_int64 * self.world)(*[int(c) for c in counts])
        fn = self._lib.th_conflict_bitmap_exchange_packed if packed else self._lib.th_conflict_bitmap_exchange
        self._native.check(fn(
            snapshot._h, self._h, m, C.c_void_p(q_uuid.data_ptr()), C.c_void_p(q_start.data_ptr()),
            C.c_void_p(q_end.data_ptr()), C.c_void_p(q_self.data_ptr()) if q_self is not None else None, int(mode),
            cnt, 1 if wait else 0, C.byref(out), stream))
        if check and wait:
            self.check()
        view = self._views.get(out.value)          # two generations -> two cached views
        if view is None:
            view = torch.as_tensor(_DevWords(out.value, self.world * self.words_per_rank),
                                   device='cuda:%d' % self.device).view(self.world, self.words_per_rank)
            self._views[out.value] = view
        return view

    def wait(self):
        import torch
        stream = self._C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self._native.check(self._lib.th_peer_wait(self._h, stream))

    def barrier(self):
        """All ranks meet on the device (enqueued on the current stream, no host synchronisation)."""
        import torch
        stream = self._C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self._native.check(self._lib.th_peer_barrier(self._h, stream))

    def check(self):
        import torch
        stream = self._C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self._native.check(self._lib.th_peer_check(self._h, stream))

    def close(self):
        if self._h:
            self._lib.th_peer_destroy(self._h)
            self._h = self._C.c_void_p(0)
