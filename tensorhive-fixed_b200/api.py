"""Array-level API over the C ABI: numpy arrays (host) or torch CUDA tensors (device) in, same kind out.

PyTorch is only a carrier here: it owns device buffers and the current stream; every computation is a
kernel of libth_b200.so.  Host arrays take the TH_MEM_HOST path (the library stages them itself), CUDA
tensors the TH_MEM_DEVICE path (no copies, work is enqueued on torch's current stream).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import native
from .native import NativeError, MEM_HOST, MEM_DEVICE  # noqa: F401

MODE_STRICT, MODE_INCLUSIVE, MODE_UPCOMING = 0, 1, 2
OWNER_ANY, OWNER_OTHER, OWNER_SAME = 0, 1, 2
TH_ECAPACITY = -5


def _is_torch_cuda(x) -> bool:
    return hasattr(x, 'is_cuda') and bool(x.is_cuda)


class _Args:
    """Collects the arrays of one call, checks they live in one place and keeps them alive."""

    def __init__(self, *probe):
        self.device_mode = any(_is_torch_cuda(p) for p in probe if p is not None)
        self.keep = []
        self.torch = None
        self.device_index = None
        if self.device_mode:
            import torch
            self.torch = torch
            for p in probe:
                if p is not None and not _is_torch_cuda(p):
                    raise TypeError('all arrays of a call must be CUDA tensors, or none of them')
            self.device_index = next(p for p in probe if p is not None).device.index

    @property
    def mem(self) -> int:
        return MEM_DEVICE if self.device_mode else MEM_HOST

    def stream(self):
        if self.device_mode:
            return C.c_void_p(self.torch.cuda.current_stream(self.device_index).cuda_stream)
        return C.c_void_p(0)

    def inp(self, x, dtype=np.int64):
        if x is None:
            return C.c_void_p(0)
        if self.device_mode:
            want = {np.int64: self.torch.int64, np.uint8: self.torch.uint8, np.uint32: self.torch.int32}[dtype]
            t = x if (x.dtype == want and x.is_contiguous()) else x.to(want).contiguous()
            self.keep.append(t)
            return C.c_void_p(t.data_ptr())
        a = np.ascontiguousarray(x, dtype=dtype)
        self.keep.append(a)
        return C.c_void_p(a.ctypes.data)

    def out(self, n: int, dtype=np.int64):
        if self.device_mode:
            want = {np.int64: self.torch.int64, np.uint8: self.torch.uint8, np.uint32: self.torch.int32}[dtype]
            t = self.torch.empty(int(n), dtype=want, device='cuda:%d' % self.device_index)
            self.keep.append(t)
            return t, C.c_void_p(t.data_ptr())
        a = np.empty(int(n), dtype=dtype)
        self.keep.append(a)
        return a, C.c_void_p(a.ctypes.data)


def unpack_bits(words, m: int):
    """bitmap words (uint32, bit i%32 of word i/32) -> bool[m] (host numpy)."""
    if _is_torch_cuda(words):
        words = words.cpu().numpy()
    w = np.ascontiguousarray(words).view(np.uint32)
    bits = np.unpackbits(w.view(np.uint8), bitorder='little')
    return bits[:m].astype(bool)


class Snapshot:
    """Device-resident, indexed copy of the ``reservations`` table (tensorhive/models/Reservation.py:14-37)."""

    def __init__(self, device: int = 0):
        self._lib = native.load()
        self._h = C.c_void_p(0)
        native.check(self._lib.th_snapshot_create(int(device), C.byref(self._h)))
        self.device = int(device)
        self.n = 0
        self.n_uuids = 0
        self._cap_hints = {}

    # -- lifecycle ---------------------------------------------------------------------------------
    def close(self) -> None:
        if self._h:
            self._lib.th_snapshot_destroy(self._h)
            self._h = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def load(self, id, uuid_id, start, end, owner, cancelled, n_uuids: int) -> 'Snapshot':
        a = _Args(id, uuid_id, start, end, owner, cancelled)
        n = int(len(id))
        for name, col in (('uuid_id', uuid_id), ('start', start), ('end', end), ('owner', owner),
                          ('cancelled', cancelled)):
            if len(col) != n:
                raise ValueError('column %s has %d rows, id has %d' % (name, len(col), n))
        native.check(self._lib.th_snapshot_load(self._h, n, int(n_uuids), a.inp(id), a.inp(uuid_id), a.inp(start),
                                                a.inp(end), a.inp(owner), a.inp(cancelled, np.uint8), a.mem, a.stream()))
        self.n, self.n_uuids = n, int(n_uuids)
        return self

    def apply_delta(self, src: 'Snapshot', n_uuids: int, id, uuid_id, start, end, owner, cancelled, delete_ids) -> 'Snapshot':
        """Make this snapshot = ``src`` minus upserted/deleted ids plus the upserted rows (``self`` may be ``src``).
        Only the delta is uploaded; the survivors are gathered and re-indexed on the device."""
        a = _Args(id, uuid_id, start, end, owner, cancelled, delete_ids)
        n_up, n_del = int(len(id)), int(len(delete_ids))
        native.check(self._lib.th_snapshot_apply_delta(self._h, src._h, int(n_uuids), n_up, a.inp(id), a.inp(uuid_id),
                                                       a.inp(start), a.inp(end), a.inp(owner), a.inp(cancelled, np.uint8),
                                                       n_del, a.inp(delete_ids), a.mem, a.stream()))
        info = self.info()
        self.n, self.n_uuids = info['n'], info['n_uuids']
        return self

    def info(self) -> dict:
        v = [C.c_int64(0) for _ in range(4)]
        native.check(self._lib.th_snapshot_info(self._h, *[C.byref(x) for x in v]))
        return {'n': v[0].value, 'n_uuids': v[1].value, 'device_bytes': v[2].value, 'radix_passes': v[3].value}

    def export(self) -> dict:
        """Indexed table back on the host, in snapshot order (uuid, start, id)."""
        a = _Args()
        cols = {}
        ptrs = []
        for name, dt, size in (('id', np.int64, self.n), ('uuid_id', np.int64, self.n), ('start', np.int64, self.n),
                               ('end', np.int64, self.n), ('owner', np.int64, self.n), ('cancelled', np.uint8, self.n),
                               ('seg_offsets', np.int64, self.n_uuids + 1)):
            arr, p = a.out(size, dt)
            cols[name] = arr
            ptrs.append(p)
        native.check(self._lib.th_snapshot_export(self._h, *ptrs, MEM_HOST, C.c_void_p(0)))
        return cols

    # -- R1 / R2 verdicts -----------------------------------------------------------------------------
    def conflict_words(self, q_uuid, q_start, q_end, q_self=None, mode: int = MODE_STRICT):
        """Raw bitmap words (uint32 numpy / int32 CUDA tensor)."""
        a = _Args(q_uuid, q_start, q_end, q_self)
        m = int(len(q_uuid))
        if q_self is None and mode == MODE_STRICT:
            q_self = np.full(m, -1, np.int64) if not a.device_mode else a.torch.full(
                (m,), -1, dtype=a.torch.int64, device=q_uuid.device)
        words, wp = a.out((m + 31) // 32, np.uint32)
        native.check(self._lib.th_conflict_bitmap(self._h, m, a.inp(q_uuid), a.inp(q_start), a.inp(q_end),
                                                  a.inp(q_self), int(mode), wp, a.mem, a.stream()))
        return words

    def conflict_words_packed(self, q_uuid_u32, q_start, q_end, q_self_u32=None, mode: int = MODE_STRICT):
        """``conflict_words`` from the narrow wire format: uuid codes and self ids as uint32 (0xFFFFFFFF = no self id,
        ``q_self_u32=None`` = no candidate is a self-update): 20-24 bytes per query cross PCIe instead of 32."""
        a = _Args(q_uuid_u32, q_start, q_end, q_self_u32)
        m = int(len(q_uuid_u32))
        words, wp = a.out((m + 31) // 32, np.uint32)
        native.check(self._lib.th_conflict_bitmap_packed(self._h, m, a.inp(q_uuid_u32, np.uint32), a.inp(q_start),
                                                         a.inp(q_end), a.inp(q_self_u32, np.uint32), int(mode), wp, a.mem,
                                                         a.stream()))
        return words

    @staticmethod
    def pack_queries(q_uuid, q_self=None):
        """int64 uuid codes / self ids (-1 = none) -> the uint32 columns of the packed format.  Codes outside
        [0, 2^32 - 2] (unknown resources) become 0xFFFFFFFE, which no snapshot has; ids must be below 2^32 - 1."""
        u = np.asarray(q_uuid, np.int64)
        pu = np.where((u < 0) | (u > 0xfffffffe), 0xfffffffe, u).astype(np.uint32)
        if q_self is None:
            return pu, None
        f = np.asarray(q_self, np.int64)
        if f.size and int(f.max()) >= 0xffffffff:
            raise ValueError('the packed format carries ids below 2^32 - 1')
        return pu, np.where(f < 0, 0xffffffff, f).astype(np.uint32)

    def conflict_bitmap(self, q_uuid, q_start, q_end, q_self=None, mode: int = MODE_STRICT) -> np.ndarray:
        """bool[m]: the value ``Reservation.would_interfere`` returns for each candidate."""
        return unpack_bits(self.conflict_words(q_uuid, q_start, q_end, q_self, mode), int(len(q_uuid)))

    # -- R1..R4 row lists -------------------------------------------------------------------------------
    def conflict_sets(self, q_uuid, q_start, q_end, q_self=None, mode: int = MODE_STRICT,
                      capacity: Optional[int] = None):
        """CSR (offsets[m+1], ids, cancelled).  Retries once with the exact capacity if the guess was short."""
        a = _Args(q_uuid, q_start, q_end, q_self)
        m = int(len(q_uuid))
        # first guess: two rows per query, or what a batch of this size and mode needed last time (a calendar view polls
        # the same ranges again and again: without the memory every poll ran the count pass twice)
        hint = self._cap_hints.get((m, int(mode)), 0)
        cap = int(capacity) if capacity is not None else max(64, 2 * m, hint + hint // 8)
        needed = C.c_int64(0)
        for _attempt in range(2):
            offsets, op = a.out(m + 1, np.int64)
            ids, ip = a.out(cap, np.int64)
            canc, cp = a.out(cap, np.uint8)
            rc = self._lib.th_conflict_sets(self._h, m, a.inp(q_uuid), a.inp(q_start), a.inp(q_end), a.inp(q_self),
                                            int(mode), op, ip, cp, cap, C.byref(needed), a.mem, a.stream())
            if rc == TH_ECAPACITY:
                cap = int(needed.value)
                continue
            native.check(rc)
            total = int(needed.value)
            if total > 2 * m:
                self._cap_hints[(m, int(mode))] = total
            return offsets, ids[:total], canc[:total]
        raise NativeError(TH_ECAPACITY, native.last_error())

    # -- R3 / R4 per-uuid lookups -------------------------------------------------------------------------
    def current_owner(self, now_us: int, device_out: bool = False):
        a = _Args()
        if device_out:
            import torch
            a.device_mode, a.torch, a.device_index = True, torch, self.device
        res, rp = a.out(self.n_uuids)
        own, op = a.out(self.n_uuids)
        native.check(self._lib.th_current_owner(self._h, int(now_us), rp, op, a.mem, a.stream()))
        return res, own

    def upcoming(self, now_us: int, period_us: int, uuids):
        a = _Args(uuids)
        k = int(len(uuids))
        outs = [a.out(k) for _ in range(4)]
        native.check(self._lib.th_upcoming(self._h, int(now_us), int(period_us), k, a.inp(uuids),
                                           outs[0][1], outs[1][1], outs[2][1], outs[3][1], a.mem, a.stream()))
        return tuple(o[0] for o in outs)   # first_id, first_start, first_owner, count

    def upcoming_owner_flags(self, now_us: int, period_us: int, task_uuid, task_owner, test: int):
        a = _Args(task_uuid, task_owner)
        t = int(len(task_uuid))
        flags, fp = a.out(t, np.uint8)
        native.check(self._lib.th_upcoming_owner_flags(self._h, int(now_us), int(period_us), t, a.inp(task_uuid),
                                                       a.inp(task_owner), int(test), fp, a.mem, a.stream()))
        return flags

    # -- V ----------------------------------------------------------------------------------------------------
    def violation_join(self, now_us: int, strict: bool, proc_uuid, proc_owner):
        a = _Args(proc_uuid, proc_owner)
        p = int(len(proc_uuid))
        flags, fp = a.out(p, np.uint8)
        res, rp = a.out(p)
        native.check(self._lib.th_violation_join(self._h, int(now_us), 1 if strict else 0, p, a.inp(proc_uuid),
                                                 a.inp(proc_owner), fp, rp, a.mem, a.stream()))
        return flags, res


    STATIC_JOIN_MAX_UUIDS = 65536

    def violation_join_static(self, now_us: int, strict: bool, proc_uuid, proc_owner, out_flags, out_res_id) -> None:
        """The violation join without a single device allocation inside the library: for callers that cannot
        allocate (the GPU's memory is already full of a reserved user's job, or the process runs with a fixed
        device footprint).  Every array is a CUDA tensor the caller already owns -- ``out_flags`` uint8[p] and
        ``out_res_id`` int64[p] included -- and the work is only enqueued on torch's current stream.  Snapshots
        with more than ``STATIC_JOIN_MAX_UUIDS`` uuids are refused.  Calls on one GPU share one built-in table:
        do not let two of them run concurrently."""
        import torch
        for name, t, dt in (('proc_uuid', proc_uuid, torch.int64), ('proc_owner', proc_owner, torch.int64),
                            ('out_flags', out_flags, torch.uint8), ('out_res_id', out_res_id, torch.int64)):
            if not (_is_torch_cuda(t) and t.dtype == dt and t.is_contiguous()):
                raise TypeError('%s must be a contiguous CUDA tensor of %s' % (name, dt))
            if t.device.index != self.device:          # the C entry point cannot tell: a tensor of another GPU would fault
                raise ValueError('%s lives on cuda:%s, the snapshot on cuda:%d' % (name, t.device.index, self.device))
        p = int(proc_uuid.numel())
        if not (proc_owner.numel() == p and out_flags.numel() == p and out_res_id.numel() == p):
            raise ValueError('all four arrays must have the same length')
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        native.check(self._lib.th_violation_join_static(self._h, int(now_us), 1 if strict else 0, p,
                                                        C.c_void_p(proc_uuid.data_ptr()), C.c_void_p(proc_owner.data_ptr()),
                                                        C.c_void_p(out_flags.data_ptr()), C.c_void_p(out_res_id.data_ptr()),
                                                        stream))


def usage_reduce(uuid_offsets, ts, gpu, mem, win_uuid, win_start, win_end, device: int = 0):
    """(sum_gpu, sum_mem, count, avg_gpu, avg_mem) per window; avg = sum // count or -1
    (tensorhive/core/services/UsageLoggingService.py:25-30)."""
    lib = native.load()
    a = _Args(uuid_offsets, ts, gpu, mem, win_uuid, win_start, win_end)
    s, w = int(len(ts)), int(len(win_uuid))
    n_uuids = int(len(uuid_offsets)) - 1
    outs = [a.out(w) for _ in range(5)]
    dev = a.device_index if a.device_mode else int(device)
    native.check(lib.th_usage_reduce(dev, s, n_uuids, a.inp(uuid_offsets), a.inp(ts), a.inp(gpu), a.inp(mem), w,
                                     a.inp(win_uuid), a.inp(win_start), a.inp(win_end),
                                     outs[0][1], outs[1][1], outs[2][1], outs[3][1], outs[4][1], a.mem, a.stream()))
    return tuple(o[0] for o in outs)


def usage_reduce_u8(uuid_offsets, ts, gpu_u8, mem_u8, win_uuid, win_start, win_end, device: int = 0):
    """usage_reduce over the narrow log format: utilisation as uint8 columns (2 bytes read per sample, not 16)."""
    lib = native.load()
    a = _Args(uuid_offsets, ts, gpu_u8, mem_u8, win_uuid, win_start, win_end)
    s, w = int(len(ts)), int(len(win_uuid))
    n_uuids = int(len(uuid_offsets)) - 1
    outs = [a.out(w) for _ in range(5)]
    dev = a.device_index if a.device_mode else int(device)
    native.check(lib.th_usage_reduce_u8(dev, s, n_uuids, a.inp(uuid_offsets), a.inp(ts), a.inp(gpu_u8, np.uint8),
                                        a.inp(mem_u8, np.uint8), w, a.inp(win_uuid), a.inp(win_start), a.inp(win_end),
                                        outs[0][1], outs[1][1], outs[2][1], outs[3][1], outs[4][1], a.mem, a.stream()))
    return tuple(o[0] for o in outs)


def sort_pairs(k_hi, k_lo, device: int = 0):
    """Stable argsort of (k_hi, k_lo) pairs by K1 (the onesweep radix sort)."""
    lib = native.load()
    a = _Args(k_hi, k_lo)
    n = int(len(k_hi))
    perm, pp = a.out(n)
    dev = a.device_index if a.device_mode else int(device)
    native.check(lib.th_sort_pairs(dev, n, a.inp(k_hi), a.inp(k_lo), pp, a.mem, a.stream()))
    return perm
