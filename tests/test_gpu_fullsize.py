"""BASELINE-size runs checked through size-independent properties and sampled brute force (the oracle cannot
finish the full sizes): config 3 (1 B samples -> 10 M windows) and one rank's share of config 5 (12.5 M reservations,
8192 uuids, 12.5 M candidates = the 100 M-reservation calendar on 8 GPUs)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_config3_full_size_usage_reduce(lib_built):
    import torch
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import native, synth
    lib = native.load()
    free, _total = torch.cuda.mem_get_info(0)
    if free < 40 << 30:
        pytest.skip('needs ~26 GB of free device memory')
    dev = torch.device('cuda:0')
    U, SPU, W = 65_536, 15_259, 10_000_000
    S = U * SPU
    i64 = lambda n: torch.empty(n, dtype=torch.int64, device=dev)
    off, ts, gpu, mem = i64(U + 1), i64(S), i64(S), i64(S)
    wu, ws, we = i64(W), i64(W), i64(W)
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    T0 = synth.T0_US
    native.check(lib.th_synth_samples(0, U, SPU, T0, 11, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(), mem.data_ptr(), sp))
    native.check(lib.th_synth_windows(0, U, SPU, T0, 12, W, wu.data_ptr(), ws.data_ptr(), we.data_ptr(), sp))
    sg, sm, cnt, ag, am = th.usage_reduce(off, ts, gpu, mem, wu, ws, we)
    torch.cuda.synchronize()
    # (1) counts of ALL windows from first principles: samples sit at T0 + k s, k in [0, SPU)
    sec = 1_000_000
    k_lo = torch.clamp(torch.div(ws - T0 + sec - 1, sec, rounding_mode='floor'), min=0)
    k_hi = torch.clamp(torch.div(we - T0, sec, rounding_mode='floor'), max=SPU - 1)
    want_cnt = torch.clamp(k_hi - k_lo + 1, min=0)
    assert torch.equal(cnt, want_cnt)
    assert int((cnt == 0).sum()) > 50_000 and int(cnt.max()) <= 200          # empty windows exist; U[20 s, 180 s]
    # (2) averages of ALL windows: floor division, -1 when empty (UsageLoggingService.py:25-30)
    safe = torch.clamp(cnt, min=1)
    assert torch.equal(ag, torch.where(cnt > 0, torch.div(sg, safe, rounding_mode='floor'), torch.full_like(sg, -1)))
    assert torch.equal(am, torch.where(cnt > 0, torch.div(sm, safe, rounding_mode='floor'), torch.full_like(sm, -1)))
    # (3) sums: exact, for every window of 256 uuids spread over the table, via prefix sums of those uuids' samples
    for u0 in (0, 20_000, U - 256):
        sel = torch.nonzero((wu >= u0) & (wu < u0 + 256)).flatten()
        assert sel.numel() > 10_000
        a0 = u0 * SPU
        cg = torch.cumsum(gpu[a0:a0 + 256 * SPU], 0)
        cm = torch.cumsum(mem[a0:a0 + 256 * SPU], 0)
        cg = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), cg])
        cm = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), cm])
        lo = (wu[sel] - u0) * SPU + k_lo[sel]
        hi = (wu[sel] - u0) * SPU + k_hi[sel] + 1
        hi = torch.maximum(hi, lo)
        assert torch.equal(sg[sel], cg[hi] - cg[lo])
        assert torch.equal(sm[sel], cm[hi] - cm[lo])
    # (4) additivity: splitting a window at an arbitrary instant splits its sums and counts exactly
    pick = torch.arange(0, W, 997, device=dev)
    cut = ws[pick] + (we[pick] - ws[pick]) // 3
    l = th.usage_reduce(off, ts, gpu, mem, wu[pick], ws[pick], cut)
    r = th.usage_reduce(off, ts, gpu, mem, wu[pick], cut + 1, we[pick])
    torch.cuda.synchronize()
    assert torch.equal(l[0] + r[0], sg[pick]) and torch.equal(l[1] + r[1], sm[pick]) and torch.equal(l[2] + r[2], cnt[pick])


def test_config5_one_rank_share(lib_built):
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle import brute
    n, u, m = 12_500_000, 8192, 12_500_000
    cal = synth.make_calendar(n, u, seed=20260105, skew=0.3)
    q = synth.make_queries(cal, m, seed=20260106)
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    got = snap.conflict_bitmap(q['uuid_id'], q['start'], q['end'], q['self_id'])
    # sampled brute force: every query and row of 48 uuids (hot and cold ones)
    counts = np.bincount(cal['uuid_id'], minlength=u)
    order = np.argsort(-counts)
    pick = np.concatenate([order[:8], order[u // 2:u // 2 + 32], order[-8:]])
    rsel = np.isin(cal['uuid_id'], pick)
    qsel = np.isin(q['uuid_id'], pick)
    sub = {k: (v[rsel] if isinstance(v, np.ndarray) and len(v) == n else v) for k, v in cal.items()}
    want = brute.conflict_bits(sub, q['uuid_id'][qsel], q['start'][qsel], q['end'][qsel], q['self_id'][qsel])
    assert qsel.sum() > 50_000
    assert np.array_equal(got[qsel], want)
    # properties over the whole batch
    live = np.nonzero(cal['cancelled'] == 0)[0][::5]
    assert snap.conflict_bitmap(cal['uuid_id'][live], cal['start'][live], cal['end'][live],
                                np.full(live.size, -1, np.int64)).all()              # a live row conflicts with its copy
    free = np.nonzero(~got)[0]
    assert not snap.conflict_bitmap(q['uuid_id'][free], q['start'][free] + 1, q['end'][free] - 1,
                                    q['self_id'][free]).any()                         # shrinking never creates a conflict
    unknown = q['uuid_id'] >= u
    assert unknown.any() and not got[unknown].any()
    # the same verdicts from the row lists: a conflict <=> some non-cancelled row in the STRICT list
    some = np.nonzero(qsel)[0][:50_000]          # >= 32768: the 4-lanes-per-query list kernels
    off, ids, canc = snap.conflict_sets(q['uuid_id'][some], q['start'][some], q['end'][some], q['self_id'][some])
    has_live = np.add.reduceat(np.concatenate([(np.asarray(canc) == 0).astype(np.int64), [0]]), np.asarray(off)[:-1]) > 0
    has_live &= np.diff(np.asarray(off)) > 0
    assert np.array_equal(has_live, got[some])
    snap.close()


def test_config4_full_size_service_sweep(lib_built):
    """BASELINE configs[3] at full size: 1 M live GPU processes and 100 k queued jobs (~250 k tasks) against the 1 M-row
    calendar.  The violation join is checked for EVERY process (the oracle's per-uuid owner table is vectorised); the
    job-task flags against the brute-force row lists on a sampled uuid set plus whole-batch identities.
    Reference: ProtectionService.py:84-102, JobSchedulingService.py:117-132, scheduling.py:48-53."""
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth
    from oracle import brute
    n, u = 1_000_000, 4096
    cal = synth.make_calendar(n, u, seed=20260102)
    now = int(np.median(cal['start']))
    period = 30 * synth.MIN_US
    procs = synth.make_processes(cal, 1_000_000, now, seed=20260108)
    tasks = synth.make_job_tasks(cal, 100_000, seed=20260109)
    snap = th.Snapshot(0)
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=u)
    want_res, want_own = brute.current_owner(cal, u, now)
    pu, po = procs['uuid_id'], procs['owner']
    r, o = want_res[pu], want_own[pu]
    for strict in (False, True):
        flags, res = snap.violation_join(now, strict, pu, po)
        viol = (r >= 0) & (o >= 0) & (po != o)
        want_flag = viol | ((r < 0) if strict else False)
        assert np.array_equal(np.asarray(flags) != 0, want_flag), strict
        assert np.array_equal(np.asarray(res), np.where(viol, r, -1)), strict
    assert 0.05 < (np.asarray(flags) != 0).mean() < 0.95
    tu, to = tasks['uuid_id'], tasks['owner']
    assert len(tu) > 200_000
    any_f = np.asarray(snap.upcoming_owner_flags(now, period, tu, to, th.OWNER_ANY)) != 0
    oth_f = np.asarray(snap.upcoming_owner_flags(now, period, tu, to, th.OWNER_OTHER)) != 0
    same_f = np.asarray(snap.upcoming_owner_flags(now, period, tu, to, th.OWNER_SAME)) != 0
    assert np.array_equal(any_f, oth_f | same_f)                       # every upcoming row is the owner's or somebody else's
    _fid, _fs, _fo, count = snap.upcoming(now, period, np.arange(u, dtype=np.int64))
    has_up = np.asarray(count) > 0
    assert np.array_equal(any_f, np.where(tu >= 0, has_up[np.clip(tu, 0, u - 1)], False))
    assert not any_f[tu < 0].any()                                     # tasks without an assigned GPU never interfere
    # sampled uuids: the row lists of upcoming_events_for_resource from the brute-force oracle, task by task
    pick = np.random.default_rng(5).choice(u, size=96, replace=False)
    sets = brute.conflict_sets(cal, pick, np.full(96, now, np.int64), np.full(96, now + period, np.int64),
                               np.full(96, -1, np.int64), brute.UPCOMING)
    owner_of = dict(zip(cal['id'].tolist(), cal['owner'].tolist()))
    owners = {int(g): [owner_of[i] for i in ids] for g, ids in zip(pick, sets)}
    sel = np.nonzero(np.isin(tu, pick))[0]
    assert sel.size > 3000
    for k in sel:
        ow = owners[int(tu[k])]
        assert any_f[k] == (len(ow) > 0)
        assert oth_f[k] == any(x != to[k] for x in ow)
        assert same_f[k] == any(x == to[k] for x in ow)
    fid, fs, fo, cnt = (np.asarray(a) for a in snap.upcoming(now, period, pick))
    wid, wst, wow, wc = brute.upcoming_summary(cal, pick, now, period)
    assert np.array_equal(fid, wid) and np.array_equal(fs, wst) and np.array_equal(fo, wow) and np.array_equal(cnt, wc)
    snap.close()
