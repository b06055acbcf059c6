"""th_violation_join_static: the violation join for callers that cannot allocate device memory.

Some users of this library may not call cudaMalloc at all while they work -- a protection tick on a GPU that a
reserved user's job has filled to the brim, or a daemon that was granted a fixed device footprint.  For them the
library carries one entry point whose only working memory is a fixed array inside the library image.  This test
calls it through the C ABI twice in a row (the second call finds the first call's table still in the array) and
expects exactly what th_violation_join returns; it also checks that an oversized snapshot is refused.

The file name sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _snapshot(th, synth, n, u, seed):
    cal = synth.make_calendar(n, u, seed=seed)
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    return cal, snap


def test_static_join_matches_the_allocating_join_on_two_calls_in_a_row(lib_built):
    import torch
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import native, synth
    lib = native.load()
    dev = torch.device('cuda:0')
    cal, snap = _snapshot(th, synth, 60_000, 512, seed=314)
    t_mid = int(np.median(cal['start']))
    t_late = int(np.percentile(cal['start'], 90))
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # the caller owns every buffer, outputs included, before the calls start
    calls = []
    for k, (now, strict, p) in enumerate([(t_mid, False, 30_000), (t_late, True, 17_001)]):
        procs = synth.make_processes(cal, p, now, seed=100 + k)
        pu = torch.from_numpy(procs['uuid_id']).to(dev)
        po = torch.from_numpy(procs['owner']).to(dev)
        pu[::97] = 512 + 5            # processes on GPUs the calendar has never seen
        calls.append((now, strict, p, pu, po, torch.full((p,), 7, dtype=torch.uint8, device=dev),
                      torch.full((p,), -99, dtype=torch.int64, device=dev)))
    # First use on this device: the library looks up the array's address and the driver loads the two kernels' code
    # (lazily, on first launch) -- set-up that a caller does once, before it enters its no-allocation regime.
    now, strict, p, pu, po, flags, res = calls[0]
    rc = lib.th_violation_join_static(snap._h, now, 0, 1, pu.data_ptr(), po.data_ptr(), flags.data_ptr(), res.data_ptr(),
                                      stream)
    assert rc == 0, native.last_error()
    flags.fill_(7); res.fill_(-99)
    torch.cuda.synchronize()
    # The property in the name: nothing is allocated while the two calls run.  Watched from three sides -- the driver's
    # free-memory figure, the library's own staging pool (high-water mark reset just before) and torch's allocator.
    reserved0, used0, high = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    native.check(lib.th_pool_stats(0, C.byref(reserved0), C.byref(used0), C.byref(high), 1))
    free0, _total = torch.cuda.mem_get_info(dev)
    torch_allocs0 = torch.cuda.memory_stats(dev).get('allocation.all.allocated', 0)
    for now, strict, p, pu, po, flags, res in calls:      # two calls in a row, same built-in table
        rc = lib.th_violation_join_static(snap._h, now, 1 if strict else 0, p, pu.data_ptr(), po.data_ptr(),
                                          flags.data_ptr(), res.data_ptr(), stream)
        assert rc == 0, native.last_error()
    torch.cuda.synchronize()
    reserved1, used1 = C.c_int64(0), C.c_int64(0)
    native.check(lib.th_pool_stats(0, C.byref(reserved1), C.byref(used1), C.byref(high), 0))
    free1, _total = torch.cuda.mem_get_info(dev)
    assert free1 == free0, 'device free memory changed by %d bytes across the two static calls' % (free0 - free1)
    assert reserved1.value == reserved0.value and used1.value == used0.value
    assert high.value == used0.value, 'the staging pool served %d bytes during the static calls' % high.value
    assert torch.cuda.memory_stats(dev).get('allocation.all.allocated', 0) == torch_allocs0
    seen = []
    for now, strict, p, pu, po, flags, res in calls:
        want_flags, want_res = snap.violation_join(now, strict, pu, po)       # the existing entry point
        torch.cuda.synchronize()
        assert torch.equal(flags, want_flags)           # integer work: exact, no tolerance
        assert torch.equal(res, want_res)
        seen.append(float(flags.float().mean()))
    assert 0.0 < seen[0] < 1.0 and 0.0 < seen[1] < 1.0
    assert seen[0] != seen[1]                           # the two calls really asked different questions
    # the Python wrapper takes the same route
    now, strict, p, pu, po, flags, res = calls[0]
    flags.fill_(9); res.fill_(-5)
    snap.violation_join_static(now, strict, pu, po, flags, res)
    torch.cuda.synchronize()
    want_flags, want_res = snap.violation_join(now, strict, pu, po)
    assert torch.equal(flags, want_flags) and torch.equal(res, want_res)
    snap.close()


def test_static_join_refuses_snapshots_beyond_its_table(lib_built):
    import torch
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import native, synth
    lib = native.load()
    dev = torch.device('cuda:0')
    limit = th.Snapshot.STATIC_JOIN_MAX_UUIDS
    cal, snap = _snapshot(th, synth, 5000, limit + 1, seed=2718)
    pu = torch.zeros(8, dtype=torch.int64, device=dev)
    po = torch.ones(8, dtype=torch.int64, device=dev)
    flags = torch.full((8,), 3, dtype=torch.uint8, device=dev)
    res = torch.full((8,), -3, dtype=torch.int64, device=dev)
    rc = lib.th_violation_join_static(snap._h, int(cal['start'][0]), 0, 8, pu.data_ptr(), po.data_ptr(),
                                      flags.data_ptr(), res.data_ptr(), None)
    assert rc == -5
    msg = native.last_error()
    assert str(limit) in msg and str(limit + 1) in msg
    torch.cuda.synchronize()
    assert int(flags[0]) == 3 and int(res[0]) == -3     # refused before touching anything
    snap.close()
    # exactly at the limit it works
    cal, snap = _snapshot(th, synth, 5000, limit, seed=1618)
    now = int(np.median(cal['start']))
    rc = lib.th_violation_join_static(snap._h, now, 1, 8, pu.data_ptr(), po.data_ptr(), flags.data_ptr(),
                                      res.data_ptr(), None)
    assert rc == 0, native.last_error()
    want_flags, want_res = snap.violation_join(now, True, pu, po)
    torch.cuda.synchronize()
    assert torch.equal(flags, want_flags) and torch.equal(res, want_res)
    snap.close()
