#!/usr/bin/env python
"""Benchmark of the reservation hot path (BASELINE.json metric: conflict queries/s + usage-reduce GB/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one pass of the hot path over one batch: K2 (`th_conflict_bitmap`) over 1 M candidate
reservations against the resident 1 M-reservation / 4096-uuid snapshot of BASELINE config 2
(`configs[1]`), and -- when N > 1 -- the all-gather of the bitmap words, which is the path's one exchange.
With N ranks every rank owns its own 4096-uuid shard and 1 M routed queries (weak scaling).
Every line also carries "cfg5_share": one rank's share of BASELINE configs[4] -- ONE global calendar of N x 12.5 M
reservations over N x 8192 uuids (100 M / 65 536 at N = 8), partitioned by `sharding.partition_uuids`, each rank holding
only its uuid range, 12.5 M routed candidates per GPU through the fused exchange -- with an all-ranks parity check of a
replicated, `Router`-routed batch against the brute-force oracle, and the two other gathers of SURVEY 8(e) timed.
The same run also measures K4 (`th_usage_reduce`) on BASELINE config 3 (1 B samples -> 10 M windows,
generated on the device) and reports it under "usage_reduce".

Timing: CUDA events on the launching stream around every step, L2 flushed between steps and outside the
events by READING a 512 MiB buffer (a write-flush leaves L2 full of dirty lines whose write-back is then
billed to the kernel: +6 us on a 25 us kernel, measured), W >= 3 warm-up steps, barrier + synchronize on both
sides, max over ranks.
`--impl reference` times the reference's own CPU path for the same metric: the SQL text of
Reservation.would_interfere through sqlite3 (SQLAlchemy is not installable here) on every
host core, each step a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_ROWS, N_UUIDS, N_QUERIES = 1_000_000, 4096, 1_000_000          # BASELINE config 2
S_UUIDS, S_PER_UUID, S_WINDOWS = 65_536, 15_259, 10_000_000       # BASELINE config 3 (65536 * 15259 ~ 1.0e9)
C5_ROWS, C5_UUIDS = 12_500_000, 8192                               # one rank's share of BASELINE config 5
C5_SAMPLES_PER_UUID, C5_WINDOWS = 2000, 1_000_000                  # usage log behind the K4 gather of the same block
SEED = 20260101


def _peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs, burst copy)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


def _traffic(kernel: str):
    """DRAM bytes per launch from the committed ncu capture, if one has been summarised."""
    p = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(kernel)
    return None


def _bind_to_gpu_numa_node(torch, index: int):
    """Pin this process to the CPUs next to its GPU (what NCCL launchers do): pinned host buffers are then allocated on
    the GPU's NUMA node and the PCIe copies of the e2e path do not cross the socket interconnect.  Best effort."""
    try:
        p = torch.cuda.get_device_properties(index)
        bdf = '%04x:%02x:%02x.0' % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open('/sys/bus/pci/devices/%s/local_cpulist' % bdf) as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(','):
            a, _, b = part.partition('-')
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return {'pci': bdf, 'cpus': spec}
    except Exception as e:                                    # no sysfs entry, no permission, old torch: leave it
        return {'unbound': str(e)[:80]}
    return {'unbound': 'empty cpu list'}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    FIELDS = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
              'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.FIELDS,
                                          '--format=csv,noheader,nounits', '-lms', '200'], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(',')])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        return {'sm_mhz': float(np.median(sm)), 'sm_max_mhz': float(max(mx)), 'reasons': sorted(reasons),
                'samples': len(sm)}


# =============================================================================================================
# reference arm / CPU baseline: the reference's SQL through sqlite3
# =============================================================================================================
_W = {}


def _ref_init(seed, n_rows, n_uuids):
    from oracle.sql_ref import SqlCalendar
    from tensorhive_fixed_b200 import synth
    cal = synth.make_calendar(n_rows, n_uuids, seed=seed)
    rows = ((int(cal['id'][i]), synth.uuid_string(int(cal['uuid_id'][i])), int(cal['start'][i]), int(cal['end'][i]),
             int(cal['owner'][i]), None if cal['cancelled_raw'][i] < 0 else int(cal['cancelled_raw'][i]))
            for i in range(n_rows))
    _W['sql'] = SqlCalendar(rows)
    _W['q'] = synth.make_queries(cal, N_QUERIES, seed=seed + 1)
    _W['uuid'] = [synth.uuid_string(k) for k in range(n_uuids + 8)]
    return True


def _ref_ready(_):
    return 'sql' in _W


def _ref_run(span):
    lo, hi = span
    sql, q, names = _W['sql'], _W['q'], _W['uuid']
    hits = 0
    for i in range(lo, hi):
        sid = None if q['self_id'][i] < 0 else int(q['self_id'][i])
        hits += sql.would_interfere(names[int(q['uuid_id'][i])], int(q['start'][i]), int(q['end'][i]), sid)
    return hits


def cpu_baseline_single_core(budget_s: float = 12.0):
    """The oracle's SQL path on ONE core (the reference is one Python thread under the GIL with one SQLite
    connection): as many config-2 candidates as fit in ~budget_s."""
    t0 = time.perf_counter()
    _ref_init(SEED + 1, N_ROWS, N_UUIDS)
    build_s = time.perf_counter() - t0
    done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        _ref_run((done, done + 8))
        done += 8
    dt = time.perf_counter() - t0
    return {'value': done / dt, 'unit': 'queries/s', 'cores': 1, 'kind': 'port',
            'sample': '%d of the 1M config-2 candidates against the full 1M-row in-memory SQLite table '
                      '(no index, as in the reference); table build %.1f s not counted' % (done, build_s)}


def run_reference(args):
    """--impl reference: every host core runs the SQL path on its own in-memory copy of the config-2 table."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 32))
    per_step_each = 6            # queries per worker per step: ~0.1 s per full-scan query
    ctx = mp.get_context('fork')
    t_build = time.perf_counter()
    with ctx.Pool(workers, initializer=_ref_init, initargs=(SEED + 1, N_ROWS, N_UUIDS)) as pool:
        assert all(pool.map(_ref_ready, range(workers * 4)))
        build_s = time.perf_counter() - t_build
        cursor = 0

        def step():
            nonlocal cursor
            spans = [(cursor + w * per_step_each, cursor + (w + 1) * per_step_each) for w in range(workers)]
            cursor += workers * per_step_each
            t0 = time.perf_counter()
            pool.map(_ref_run, spans, chunksize=1)
            return time.perf_counter() - t0

        for _ in range(args.warmup):
            step()
        times = [step() for _ in range(args.steps)]
    per_step = workers * per_step_each
    total = float(sum(times))
    value = per_step * args.steps / total
    sample = ('%d config-2 candidates per step (%d per worker) against a private 1M-row in-memory SQLite table per '
              'worker; table build %.1f s not counted' % (per_step, per_step_each, build_s))
    line = {'impl': 'reference', 'metric': 'conflict_queries_per_s', 'value': value, 'unit': 'queries/s', 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * total / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'int64', 'data': 'synthetic',
            'config': _config(args.gpus, ref=True),
            'cpu_baseline': {'value': value, 'unit': 'queries/s', 'cores': workers, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': 'queries/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0,
            'note': 'reference SQL text (Reservation.py:121-130) through stdlib sqlite3; SQLAlchemy 1.3.0 is not '
                    'installable on this image, so this is an upper bound on the reference ORM path'}
    print(json.dumps(line), flush=True)


def _config(n_gpus, ref=False):
    name = ('BASELINE configs[1]: 1M reservations x 4096 GPU UUIDs, 1M candidate conflict queries per GPU'
            if N_ROWS == 1_000_000 else
            "one rank's share of BASELINE configs[4]: %d reservations x %d GPU UUIDs, %d candidate conflict queries per GPU"
            % (N_ROWS, N_UUIDS, N_QUERIES))
    return {'workload': name,
            'reservations_per_gpu': N_ROWS, 'uuids_per_gpu': N_UUIDS, 'queries_per_step_per_gpu': N_QUERIES,
            'parallelism': ('uuid-sharded x%d, K2 fused with the all-gather of bitmap words over NVLink peer memory' % n_gpus
                            if n_gpus > 1 else 'single GPU'),
            'l2': 'n/a (CPU)' if ref else 'flushed between steps (512 MiB read sweep, outside the timed events)',
            'seed': SEED}


# =============================================================================================================
# our arm
# =============================================================================================================
def run_ours(args):
    import torch
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import synth, native
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    affinity = _bind_to_gpu_numa_node(torch, local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    lib = native.load()
    steps, warmup = args.steps, max(args.warmup, 3)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- workload: this rank's shard of the calendar and its routed queries -------------------------------
    cal = synth.make_calendar(N_ROWS, N_UUIDS, seed=SEED + 1 + 100 * rank)
    q = synth.make_queries(cal, N_QUERIES, seed=SEED + 2 + 100 * rank)
    snap = th.Snapshot(local)
    t0 = time.perf_counter()
    snap.load(cal['id'], cal['uuid_id'], cal['start'], cal['end'], cal['owner'], cal['cancelled'], n_uuids=N_UUIDS)
    load_ms_host = 1e3 * (time.perf_counter() - t0)
    tdev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    d_cal = [tdev(cal[k]) for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')]
    qd = [tdev(q[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]
    words = torch.empty((N_QUERIES + 31) // 32, dtype=torch.int32, device=dev)
    px = None
    if world > 1:       # K2 fused with the all-gather of its words over peer memory (no NCCL call per step)
        from tensorhive_fixed_b200 import sharding
        px = sharding.PeerExchange(local, world, rank, N_QUERIES)
    flush = torch.zeros(64 << 20, dtype=torch.int64, device=dev)     # 512 MiB, read to evict L2 with clean lines
    stream = torch.cuda.current_stream(dev)
    import ctypes as C
    sp = C.c_void_p(stream.cuda_stream)

    def k2_step():
        if px is None:
            native.check(lib.th_conflict_bitmap(snap._h, N_QUERIES, qd[0].data_ptr(), qd[1].data_ptr(),
                                                qd[2].data_ptr(), qd[3].data_ptr(), 0, words.data_ptr(), 1, sp))
            return words
        return px.conflict_words(snap, qd[0], qd[1], qd[2], qd[3], check=False)      # [world, words] on every rank

    def timed(step_fn, n):
        """n steps, each bracketed by its own CUDA events; L2 flushed before every step, outside the events."""
        evs = []
        for _ in range(n):
            flush.sum()
            if px is not None:
                px.barrier()      # ranks meet on the device: the step's time is the exchange, not arrival skew
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            step_fn()
            b.record(stream)
            evs.append((a, b))
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    # snapshot (re)load from device-resident columns: K1 + index build, reported separately
    ld = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        snap.load(*d_cal, n_uuids=N_UUIDS)
        ld.append(1e3 * (time.perf_counter() - t0))
    info = snap.info()
    # maintenance from a delta (1000 changed rows): survivors gathered on the device, no table re-upload
    up = {k: cal[k][:1000].copy() for k in ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')}
    up['end'] = up['end'] + 1
    dl = []
    for _ in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        snap.apply_delta(snap, N_UUIDS, up['id'], up['uuid_id'], up['start'], up['end'], up['owner'], up['cancelled'],
                         np.zeros(0, np.int64))
        dl.append(1e3 * (time.perf_counter() - t0))
    snap.load(*d_cal, n_uuids=N_UUIDS)      # back to the benchmark table

    for _ in range(warmup):
        flush.sum()
        k2_step()
    barrier()
    with ClockSampler(local) as clocks:
        barrier()
        times = timed(k2_step, steps)
        barrier()
        # keep the GPU busy long enough for at least a few clock samples (kernel only: ranks run this loop a
        # different number of times, so it must not contain a collective)
        t_end = time.perf_counter() + 1.0
        while time.perf_counter() < t_end:
            for _ in range(50):
                native.check(lib.th_conflict_bitmap(snap._h, N_QUERIES, qd[0].data_ptr(), qd[1].data_ptr(),
                                                    qd[2].data_ptr(), qd[3].data_ptr(), 0, words.data_ptr(), 1, sp))
            torch.cuda.synchronize()
    total_ms = max_over_ranks(float(sum(times)))
    ms_per_step = total_ms / steps
    value = world * N_QUERIES * steps / (total_ms * 1e-3)

    # kernel-only duration (no collective) for the roofline
    def k2_only():
        native.check(lib.th_conflict_bitmap(snap._h, N_QUERIES, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(),
                                            qd[3].data_ptr(), 0, words.data_ptr(), 1, sp))
    ktimes = timed(k2_only, max(steps, 10))
    k_ms = float(np.mean(ktimes))
    # warm-L2 figure for context (snapshot resident in L2, as between service ticks): 20 launches inside one
    # CUDA graph so that no CPU launch latency sits between the two events
    k_ms_warm = None
    if dist is None:      # (no graph capture next to a live NCCL communicator: its watchdog polls events)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            cap = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            for _ in range(20):
                native.check(lib.th_conflict_bitmap(snap._h, N_QUERIES, qd[0].data_ptr(), qd[1].data_ptr(),
                                                    qd[2].data_ptr(), qd[3].data_ptr(), 0, words.data_ptr(), 1, cap))
        graph.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); graph.replay(); b.record()
        torch.cuda.synchronize()
        k_ms_warm = a.elapsed_time(b) / 20.0
    # service steady state for context: the snapshot stays resident while 8 DIFFERENT query batches (8 x 32 MB > L2)
    # stream through, no flush -- queries come from HBM every time, the index mostly from L2
    k_ms_stream = None
    if dist is None and N_QUERIES <= 2_000_000:
        batches = [[torch.roll(t, shifts=1000 * (j + 1)) for t in qd] for j in range(8)]
        outs_b = [torch.empty_like(words) for _ in range(8)]
        g8 = torch.cuda.CUDAGraph()        # one graph = the 8 batches back to back: no CPU launch gaps inside
        with torch.cuda.graph(g8):
            cap = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            for j in range(8):
                native.check(lib.th_conflict_bitmap(snap._h, N_QUERIES, batches[j][0].data_ptr(), batches[j][1].data_ptr(),
                                                    batches[j][2].data_ptr(), batches[j][3].data_ptr(), 0,
                                                    outs_b[j].data_ptr(), 1, cap))
        flush.sum()
        g8.replay()
        reps = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); g8.replay(); b.record()
            torch.cuda.synchronize()
            reps.append(a.elapsed_time(b) / 8.0)
        k_ms_stream = float(np.median(reps))
        del batches, outs_b
    # K2 on a calendar with Zipf-skewed uuid popularity (a quarter of all rows on the hottest GPU): the load gives every
    # uuid its own time grid there, so hot uuids do not fall back to long bisections
    k_ms_skew = None
    if dist is None and N_ROWS == 1_000_000:
        cal_s = synth.make_calendar(N_ROWS, N_UUIDS, seed=SEED + 11, skew=1.2)
        q_s = synth.make_queries(cal_s, N_QUERIES, seed=SEED + 12)
        snap_s = th.Snapshot(local)
        snap_s.load(cal_s['id'], cal_s['uuid_id'], cal_s['start'], cal_s['end'], cal_s['owner'], cal_s['cancelled'],
                    n_uuids=N_UUIDS)
        qs_d = [tdev(q_s[k]) for k in ('uuid_id', 'start', 'end', 'self_id')]

        def k2_skew():
            native.check(lib.th_conflict_bitmap(snap_s._h, N_QUERIES, qs_d[0].data_ptr(), qs_d[1].data_ptr(),
                                                qs_d[2].data_ptr(), qs_d[3].data_ptr(), 0, words.data_ptr(), 1, sp))
        k_ms_skew = float(np.mean(timed(k2_skew, 10)))
        snap_s.close()
        del qs_d, cal_s, q_s
    peak, peak_src = _peaks()
    alg_bytes = N_QUERIES * 32 + N_ROWS * 25 + 8 * (N_UUIDS + 1) + N_QUERIES // 8
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9

    # ---- e2e: the public API with pinned HOST buffers, copies inside the timed region ------------------------
    pin = [torch.from_numpy(q[k]).pin_memory() for k in ('uuid_id', 'start', 'end', 'self_id')]
    pin_np = [p.numpy() for p in pin]
    out_pin = torch.empty((N_QUERIES + 31) // 32, dtype=torch.int32).pin_memory()

    out_all = torch.empty(world * ((N_QUERIES + 31) // 32), dtype=torch.int32).pin_memory() if px is not None else None

    # the packed wire format (uint32 uuid / self columns: 24 B per query instead of 32) in pinned buffers as well
    pu, pf = th.Snapshot.pack_queries(q['uuid_id'], q['self_id'])
    pin_u32 = torch.from_numpy(pu.view(np.int32)).pin_memory()
    pin_f32 = torch.from_numpy(pf.view(np.int32)).pin_memory()
    out_pin_p = torch.empty((N_QUERIES + 31) // 32, dtype=torch.int32).pin_memory()

    def e2e_step(packed=True):
        if px is None:
            if packed:
                native.check(lib.th_conflict_bitmap_packed(snap._h, N_QUERIES, pin_u32.data_ptr(), pin[1].data_ptr(),
                                                           pin[2].data_ptr(), pin_f32.data_ptr(), 0, out_pin_p.data_ptr(), 0, sp))
            else:
                native.check(lib.th_conflict_bitmap(snap._h, N_QUERIES, pin_np[0].ctypes.data, pin_np[1].ctypes.data,
                                                    pin_np[2].ctypes.data, pin_np[3].ctypes.data, 0, out_pin.data_ptr(), 0,
                                                    sp))
            return
        # the fused kernel reads the pinned columns in place over PCIe (no staging copies); everybody's verdicts
        # come back to the host
        if packed:
            g = px.conflict_words(snap, pin_u32, pin[1], pin[2], pin_f32, check=False, packed=True)
        else:
            g = px.conflict_words(snap, pin[0], pin[1], pin[2], pin[3], check=False)
        out_all.copy_(g.reshape(-1), non_blocking=True)
        torch.cuda.synchronize()

    def e2e_timed(packed):
        for _ in range(3):
            e2e_step(packed)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step(packed)
        torch.cuda.synchronize()
        dt_ = max_over_ranks(time.perf_counter() - t0)
        barrier()
        return dt_

    # for context: the same 32 MB as bare pinned-host -> device copies on this box (what the e2e step cannot beat)
    bare = [torch.empty_like(t) for t in qd]
    copy_ms = []
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for d, h in zip(bare, pin):
            d.copy_(h, non_blocking=True)
        torch.cuda.synchronize()
        copy_ms.append(1e3 * (time.perf_counter() - t0))
    del bare
    e2e_wide_s = e2e_timed(False)           # int64 columns, 32 B per query
    out_wide = out_all.clone() if px is not None else None
    e2e_s = e2e_timed(True)                 # packed columns, 24 B per query: the headline
    e2e_value = world * N_QUERIES * steps / e2e_s
    # check the last e2e result against the device-resident run (same queries)
    k2_only(); torch.cuda.synchronize()
    exchange_check = None
    if px is None:
        assert np.array_equal(out_pin.numpy(), words.cpu().numpy()), 'host-path and device-path bitmaps differ'
        assert np.array_equal(out_pin_p.numpy(), words.cpu().numpy()), 'packed host path and device path bitmaps differ'
    else:
        # EVERY rank's words as delivered by the fused exchange (device-resident step and e2e step) against the same
        # words moved by NCCL: all_gather of each rank's plain-kernel bitmap
        px.check()
        nccl_all = torch.empty(world * words.numel(), dtype=torch.int32, device=dev)
        dist.all_gather_into_tensor(nccl_all, words)
        fused = k2_step()
        torch.cuda.synchronize()
        px.check()
        want_all = nccl_all.view(world, -1).cpu().numpy()
        bad = int((fused.cpu().numpy()[:, :words.numel()] != want_all).sum())
        bad += int((out_all.numpy().reshape(world, -1) != want_all).sum())          # e2e, packed columns
        bad += int((out_wide.numpy().reshape(world, -1) != want_all).sum())         # e2e, int64 columns
        t_bad = torch.tensor([bad], dtype=torch.int64, device=dev)
        dist.all_reduce(t_bad)
        exchange_check = {'words_compared_per_rank': 3 * int(want_all.size), 'ranks': world,
                          'mismatching_words_all_ranks': int(t_bad.item()),
                          'what': 'fused-exchange words of ALL ranks (device step and e2e step) vs an NCCL all_gather of the '
                                  'plain-kernel words, on every rank'}
        assert int(t_bad.item()) == 0, 'fused-exchange words differ from the NCCL all-gather of the plain kernel'

    line = {'metric': 'conflict_queries_per_s', 'value': value, 'unit': 'queries/s', 'n_gpus': world, 'steps': steps,
            'warmup': warmup, 'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'int64', 'data': 'synthetic', 'config': _config(world),
            'clocks': clocks.summary(),
            'e2e': {'value': e2e_value, 'unit': 'queries/s', 'h2d_bytes_per_step': 24 * N_QUERIES,
                    'int64_columns': {'value': world * N_QUERIES * steps / e2e_wide_s, 'ms_per_step': 1e3 * e2e_wide_s / steps,
                                      'h2d_bytes_per_step': 32 * N_QUERIES,
                                      'api': 'th_conflict_bitmap / th_conflict_bitmap_exchange on four int64 columns'},
                    'd2h_bytes_per_step': 4 * ((N_QUERIES + 31) // 32) * (world if px is not None else 1),
                    'ms_per_step': 1e3 * e2e_s / steps, 'host_affinity': affinity,
                    'bare_h2d_copy_ms': min(copy_ms), 'bare_h2d_GBs': 32 * N_QUERIES / (min(copy_ms) * 1e-3) / 1e9,
                    'api': ('th_conflict_bitmap_packed(TH_MEM_HOST) on pinned buffers (uuid / self id as uint32, start / end int64: '
                            '24 B per query): the kernel reads the columns in place over PCIe and writes the bitmap into the '
                            'pinned output, no staging copies' if px is None else
                            'th_conflict_bitmap_exchange_packed on pinned buffers (fused K2 + peer all-gather; the kernel reads '
                            'the packed query columns in place over PCIe) + D2H of all ranks\' words')},
            'gpu_launches': steps,      # one K2 launch per step (send and receive halves of the exchange are inside it)
            'roofline': {'kernel': 'k2_bitmap_kernel', 'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak, 'traffic': _traffic('k2_bitmap_kernel') if N_ROWS == 1_000_000 else None,
                         'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': alg_bytes, 'kernel_ms': k_ms, 'kernel_ms_warm_l2': k_ms_warm,
                         'kernel_ms_streaming_8_batches_no_flush': k_ms_stream,
                         'kernel_ms_zipf_skewed_calendar': k_ms_skew,
                         'frac_of_nominal_8TBs': achieved / 8000.0,
                         'what_bounds_it': 'the number of scattered L2 fills, not their bytes: 1.215 probe fills + 0.25 stream line '
                                           'fills per query against ~46 G fills/s out of HBM on B200 (profiles/r2_random_access_rate.md); '
                                           'at one fill per query the byte fraction could reach 0.36 at most'},
            'snapshot_load': {'ms_device_columns': float(np.median(ld)), 'ms_first_from_host': load_ms_host,
                              'ms_apply_delta_1000_rows_host_arrays': float(np.median(dl[1:])),
                              'radix_passes': info['radix_passes'], 'device_bytes': info['device_bytes']}}

    if exchange_check is not None:
        line['exchange_check'] = exchange_check
    # ---- BASELINE config 5: this rank's share of ONE global calendar, cross-rank parity, the other two gathers ------
    if not args.skip_cfg5:
        c5 = bench_cfg5(torch, lib, native, dist, dev, local, rank, world, sp, stream, flush, steps, warmup, peak, peak_src,
                        barrier, max_over_ranks)
        line['cfg5_share'] = c5
        line['parity'] = c5.pop('parity')
        line['gpu_launches'] += steps
    # ---- BASELINE config 4: ProtectionService + job-queue sweep over the same calendar ---------------------------
    if not args.skip_sweep:
        line['service_sweep'] = bench_sweep(torch, lib, native, snap, cal, dev, sp, stream, flush, peak, qd)

    # ---- K4 on BASELINE config 3, generated on the device -------------------------------------------------------
    if not args.skip_usage:
        line['usage_reduce'] = bench_usage(torch, lib, native, dev, local, sp, stream, flush, steps, warmup, barrier,
                                           max_over_ranks, world, peak, peak_src)
        line['gpu_launches'] += steps
    if rank == 0 and world == 1 and not args.skip_cpu:
        line['cpu_baseline'] = cpu_baseline_single_core()
    elif rank == 0:
        line['cpu_baseline'] = None
    if rank == 0:
        print(json.dumps(line), flush=True)
    if px is not None:
        dist.barrier()
        px.close()
    snap.close()
    if dist is not None:
        dist.destroy_process_group()


def bench_sweep(torch, lib, native, snap, cal, dev, sp, stream, flush, peak, qd=None):
    """BASELINE configs[3]: 1M live GPU processes joined against the current reservation owners (K3 + K5,
    ProtectionService.py:84-102) and 100k queued jobs' tasks tested against upcoming reservations (K5,
    JobSchedulingService.py:117-132), device-resident inputs, cold L2, median of 10."""
    from tensorhive_fixed_b200 import synth
    now = int(np.median(cal['start']))
    period = 30 * synth.MIN_US
    procs = synth.make_processes(cal, 1_000_000, now, seed=SEED + 7)
    tasks = synth.make_job_tasks(cal, 100_000, seed=SEED + 8)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    pu, po = t(procs['uuid_id']), t(procs['owner'])
    tu, to = t(tasks['uuid_id']), t(tasks['owner'])
    P, T, U, N = int(pu.numel()), int(tu.numel()), N_UUIDS, N_ROWS
    flags = torch.empty(P, dtype=torch.uint8, device=dev)
    res = torch.empty(P, dtype=torch.int64, device=dev)
    tflags = torch.empty(T, dtype=torch.uint8, device=dev)
    uu = torch.arange(U, dtype=torch.int64, device=dev)
    up = [torch.empty(U, dtype=torch.int64, device=dev) for _ in range(4)]

    def med(fn):
        for _ in range(3):
            fn()
        ev = []
        for _ in range(10):
            flush.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream); fn(); b.record(stream)
            ev.append((a, b))
        torch.cuda.synchronize()
        return float(np.median([a.elapsed_time(b) for a, b in ev]))

    ms_join = med(lambda: native.check(lib.th_violation_join(snap._h, now, 0, P, pu.data_ptr(), po.data_ptr(),
                                                             flags.data_ptr(), res.data_ptr(), 1, sp)))
    ms_task = med(lambda: native.check(lib.th_upcoming_owner_flags(snap._h, now, period, T, tu.data_ptr(), to.data_ptr(),
                                                                   1, tflags.data_ptr(), 1, sp)))
    ms_up = med(lambda: native.check(lib.th_upcoming(snap._h, now, period, U, uu.data_ptr(), up[0].data_ptr(),
                                                     up[1].data_ptr(), up[2].data_ptr(), up[3].data_ptr(), 1, sp)))
    # calendar view (controllers/reservation.py:31-55 -> filter_by_uuids_and_time_range): one 30-day range per uuid,
    # row lists through th_conflict_sets(TH_MODE_INCLUSIVE) on the host path (two passes + D2H of the CSR)
    r_s = np.full(U, now, np.int64)
    r_e = r_s + 30 * synth.DAY_US
    snapshot = snap
    r_u = np.arange(U, dtype=np.int64)
    snapshot.conflict_sets(r_u, r_s, r_e, None, 1)          # first host-path call grows the library's staging pool
    reps = []
    for _ in range(5):
        t0 = time.perf_counter()
        off, ids, canc = snapshot.conflict_sets(r_u, r_s, r_e, None, 1)
        reps.append(1e3 * (time.perf_counter() - t0))
    ms_range = float(np.median(reps))
    torch.cuda.synchronize()
    # conflict SETS of the whole candidate batch (R1's list form: ids ascending, cancelled rows included with their flag),
    # device-resident; the call counts, scans, synchronises once for the capacity check, fills and ranks by id
    # one candidate at a time, the way Reservation.would_interfere() is called from a controller: host arrays of length
    # 1, wall clock per call (staging, launch, copy back, synchronisation)
    one = [np.ascontiguousarray(cal[k][:1]) for k in ('uuid_id', 'start', 'end')] + [np.full(1, -1, np.int64)]
    one_out = np.zeros(1, np.uint32)
    for _ in range(20):
        native.check(lib.th_conflict_bitmap(snap._h, 1, one[0].ctypes.data, one[1].ctypes.data, one[2].ctypes.data,
                                            one[3].ctypes.data, 0, one_out.ctypes.data, 0, sp))
    t0 = time.perf_counter()
    for _ in range(200):
        native.check(lib.th_conflict_bitmap(snap._h, 1, one[0].ctypes.data, one[1].ctypes.data, one[2].ctypes.data,
                                            one[3].ctypes.data, 0, one_out.ctypes.data, 0, sp))
    us_single = 1e6 * (time.perf_counter() - t0) / 200
    sets = None
    if qd is not None:
        import ctypes as C
        M = int(qd[0].numel())
        cap = 8 * M
        s_off = torch.empty(M + 1, dtype=torch.int64, device=dev)
        s_ids = torch.empty(cap, dtype=torch.int64, device=dev)
        s_canc = torch.empty(cap, dtype=torch.uint8, device=dev)
        needed = C.c_int64(0)

        def sets_step():
            native.check(lib.th_conflict_sets(snap._h, M, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(),
                                              qd[3].data_ptr(), 0, s_off.data_ptr(), s_ids.data_ptr(), s_canc.data_ptr(),
                                              cap, C.byref(needed), 1, sp))
        ms_sets = med(sets_step)
        sets = {'ms': ms_sets, 'queries_per_s': M / (ms_sets * 1e-3), 'rows_returned': int(needed.value),
                'note': 'th_conflict_sets(TH_MODE_STRICT) on the config-2 batch, device arrays; four kernels and one '
                        'host synchronisation (capacity check) inside the call'}
    join_bytes = P * 16 + P * 9 + U * 16          # the K5 join proper
    return {'workload': 'BASELINE configs[3]: 1M GPU processes + 100k queued jobs (%d tasks) vs the config-2 calendar' % T,
            'violation_join': {'ms': ms_join, 'processes_per_s': P / (ms_join * 1e-3), 'violations': int(flags.sum().item()),
                               'algorithmic_bytes': join_bytes, 'achieved_GBs': join_bytes / (ms_join * 1e-3) / 1e9,
                               'frac_of_peak': join_bytes / (ms_join * 1e-3) / 1e9 / peak,
                               'note': 'two launches: K3 current-owner table over %d uuids, then the K5 join' % U},
            'job_tasks': {'ms': ms_task, 'tasks_per_s': T / (ms_task * 1e-3), 'interfering': int(tflags.sum().item())},
            'gpu_slots': {'ms': ms_up, 'uuids_per_s': U / (ms_up * 1e-3)},
            'conflict_sets': sets,
            'single_would_interfere_call': {'us_per_call': us_single, 'note': 'th_conflict_bitmap(m = 1, TH_MEM_HOST), wall clock'},
            'calendar_range': {'ms_host_call': ms_range, 'ranges': U, 'rows_returned': int(len(ids)),
                               'note': 'th_conflict_sets(TH_MODE_INCLUSIVE), host arrays, wall clock incl. copies'}}


def bench_usage(torch, lib, native, dev, local, sp, stream, flush, steps, warmup, barrier, max_over_ranks, world, peak,
                peak_src):
    s_total = S_UUIDS * S_PER_UUID
    i64 = lambda n: torch.empty(n, dtype=torch.int64, device=dev)
    off, ts, gpu, mem = i64(S_UUIDS + 1), i64(s_total), i64(s_total), i64(s_total)
    wu, ws, we = i64(S_WINDOWS), i64(S_WINDOWS), i64(S_WINDOWS)
    outs = [i64(S_WINDOWS) for _ in range(5)]
    from tensorhive_fixed_b200 import synth
    native.check(lib.th_synth_samples(local, S_UUIDS, S_PER_UUID, synth.T0_US, SEED + 3, off.data_ptr(), ts.data_ptr(),
                                      gpu.data_ptr(), mem.data_ptr(), sp))
    native.check(lib.th_synth_windows(local, S_UUIDS, S_PER_UUID, synth.T0_US, SEED + 4, S_WINDOWS, wu.data_ptr(),
                                      ws.data_ptr(), we.data_ptr(), sp))
    torch.cuda.synchronize()

    def k4_step():
        native.check(lib.th_usage_reduce(local, s_total, S_UUIDS, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(),
                                         mem.data_ptr(), S_WINDOWS, wu.data_ptr(), ws.data_ptr(), we.data_ptr(),
                                         outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(), outs[3].data_ptr(),
                                         outs[4].data_ptr(), 1, sp))

    for _ in range(warmup):
        k4_step()
    barrier()
    evs = []
    for _ in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); k4_step(); b.record(stream)
        evs.append((a, b))
    torch.cuda.synchronize()
    ms = max_over_ranks(float(np.mean([a.elapsed_time(b) for a, b in evs])))
    barrier()
    # the same 10 M windows in RANDOM order (a reservation table in id order is not sorted by (uuid, start); the generator
    # above deals windows to uuids in runs with increasing starts, i.e. a stream): results must be the same set
    perm = torch.randperm(S_WINDOWS, device=dev)
    wu_s, ws_s, we_s = wu[perm].contiguous(), ws[perm].contiguous(), we[perm].contiguous()
    outs_s = [i64(S_WINDOWS) for _ in range(5)]

    def k4_shuffled_step():
        native.check(lib.th_usage_reduce(local, s_total, S_UUIDS, off.data_ptr(), ts.data_ptr(), gpu.data_ptr(),
                                         mem.data_ptr(), S_WINDOWS, wu_s.data_ptr(), ws_s.data_ptr(), we_s.data_ptr(),
                                         outs_s[0].data_ptr(), outs_s[1].data_ptr(), outs_s[2].data_ptr(),
                                         outs_s[3].data_ptr(), outs_s[4].data_ptr(), 1, sp))

    for _ in range(2):
        k4_shuffled_step()
    evs_s = []
    for _ in range(max(3, steps // 4)):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); k4_shuffled_step(); b.record(stream)
        evs_s.append((a, b))
    torch.cuda.synchronize()
    ms_shuffled = max_over_ranks(float(np.mean([a.elapsed_time(b) for a, b in evs_s])))
    shuffled_same = all(bool(torch.equal(x[perm], y)) for x, y in zip(outs, outs_s))
    del wu_s, ws_s, we_s, outs_s, perm
    # SURVEY section 8d counts 24 B per sample (ts, gpu, mem) and defines the reported "usage-reduce GB/s" as that figure
    # over time.  The kernel never streams the ts column (it is probed: 2 x 128 B per window), so against the roofline
    # only what it must move is counted: 16 B per sample, the window columns in, five int64 results out, the uuid table.
    alg_survey = s_total * 24 + S_WINDOWS * 24 + S_WINDOWS * 24 + 8 * (S_UUIDS + 1)
    alg = s_total * 16 + S_WINDOWS * 24 + S_WINDOWS * 40 + 8 * (S_UUIDS + 1)
    gbs_survey = alg_survey / (ms * 1e-3) / 1e9
    gbs = alg / (ms * 1e-3) / 1e9
    # the narrow log format (uint8 utilisation): same windows, 2 bytes per sample instead of 16 behind the same search
    g8, m8 = gpu.to(torch.uint8), mem.to(torch.uint8)
    outs8 = [i64(S_WINDOWS) for _ in range(5)]

    def k4_u8_step():
        native.check(lib.th_usage_reduce_u8(local, s_total, S_UUIDS, off.data_ptr(), ts.data_ptr(), g8.data_ptr(),
                                            m8.data_ptr(), S_WINDOWS, wu.data_ptr(), ws.data_ptr(), we.data_ptr(),
                                            outs8[0].data_ptr(), outs8[1].data_ptr(), outs8[2].data_ptr(),
                                            outs8[3].data_ptr(), outs8[4].data_ptr(), 1, sp))

    for _ in range(warmup):
        k4_u8_step()
    evs8 = []
    for _ in range(steps):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); k4_u8_step(); b.record(stream)
        evs8.append((a, b))
    torch.cuda.synchronize()
    ms8 = max_over_ranks(float(np.mean([a.elapsed_time(b) for a, b in evs8])))
    same = all(bool(torch.equal(x, y)) for x, y in zip(outs, outs8))
    alg8 = s_total * 2 + S_WINDOWS * 48 + 8 * (S_UUIDS + 1)
    gbs8 = alg8 / (ms8 * 1e-3) / 1e9
    cnt = outs[2]
    covered = int(cnt.sum().item())
    cpu = None
    if world == 1:
        # the reference's averager on the host: avg(values) = sum(values) // len(values) over Python lists
        # (UsageLoggingService.py:25-30), on windows of the same length distribution; one core, bounded sample
        from oracle.services_ref import avg as ref_avg
        rng = np.random.default_rng(SEED)
        lists = [[int(v) for v in rng.integers(0, 101, size=int(n))] for n in rng.integers(20, 181, size=100_000)]
        t0 = time.perf_counter()
        for v in lists:
            ref_avg(v); ref_avg(v)          # gpu_util and mem_util
        dt = time.perf_counter() - t0
        n_samples = sum(len(v) for v in lists)
        cpu = {'value': n_samples / dt, 'unit': 'samples/s', 'cores': 1, 'kind': 'port',
               'sample': '%d windows (%d samples) of in-memory Python lists; the reference additionally re-reads and '
                         're-writes each JSON log on every tick' % (len(lists), n_samples)}
    # ---- the device-resident sample ring (th_ring_*) at 1/8 of config 3: ticks appended as a monitor would, then the
    # same windows reduced from a ring whose stamps lie on an exact grid (run-length timestamps: the ts column is never
    # read) and from one with jittered stamps (probed), against th_usage_reduce_u8 over plain columns
    ring = None
    if world == 1:
        import ctypes as C
        RU, RSPU, RW, BATCH = 8192, S_PER_UUID, S_WINDOWS // 8, 1024
        rng = np.random.default_rng(SEED + 5)
        rw = [i64(RW) for _ in range(3)]
        native.check(lib.th_synth_windows(local, RU, RSPU, synth.T0_US, SEED + 6, RW, rw[0].data_ptr(), rw[1].data_ptr(),
                                          rw[2].data_ptr(), sp))
        rings, tick_ms, cols = {}, [], {}
        for kind in ('regular', 'jittered'):
            h = C.c_void_p(0)
            native.check(lib.th_ring_create(local, RSPU, C.byref(h)))
            parts = {'ts': [], 'g': [], 'm': []}
            for t0_ in range(0, RSPU, BATCH):
                nt_ = min(BATCH, RSPU - t0_)
                uu = np.tile(np.arange(RU, dtype=np.int64), nt_)                      # tick-major, as a monitor reports
                ts_ = np.repeat(synth.T0_US + (t0_ + np.arange(nt_, dtype=np.int64)) * 1_000_000, RU)
                if kind == 'jittered':
                    ts_ = ts_ + rng.integers(0, 900_000, size=ts_.size)
                g_ = rng.integers(0, 101, size=ts_.size).astype(np.uint8)
                m_ = rng.integers(0, 101, size=ts_.size).astype(np.uint8)
                a0 = time.perf_counter()
                native.check(lib.th_ring_append(h, uu.size, uu.ctypes.data, ts_.ctypes.data, g_.ctypes.data, m_.ctypes.data, None))
                tick_ms.append(1e3 * (time.perf_counter() - a0) / nt_)
                for k_, v_ in (('ts', ts_), ('g', g_), ('m', m_)):
                    parts[k_].append(v_.reshape(nt_, RU))
            rings[kind] = h
            cols[kind] = [torch.from_numpy(np.ascontiguousarray(np.concatenate(parts[k_]).T).reshape(-1)).to(dev)
                          for k_ in ('ts', 'g', 'm')]                               # uuid-major CSR columns of the same log
            del parts
        r_off = torch.arange(RU + 1, dtype=torch.int64, device=dev) * RSPU
        r_out = {k_: [i64(RW) for _ in range(5)] for k_ in ('regular', 'jittered', 'plain')}

        def ring_step(kind):
            native.check(lib.th_ring_reduce(rings[kind], RW, rw[0].data_ptr(), rw[1].data_ptr(), rw[2].data_ptr(),
                                            *[o.data_ptr() for o in r_out[kind]], 1, sp))

        def plain_step(kind):
            c_ = cols[kind]
            native.check(lib.th_usage_reduce_u8(local, RU * RSPU, RU, r_off.data_ptr(), c_[0].data_ptr(), c_[1].data_ptr(),
                                                c_[2].data_ptr(), RW, rw[0].data_ptr(), rw[1].data_ptr(), rw[2].data_ptr(),
                                                *[o.data_ptr() for o in r_out['plain']], 1, sp))

        def med_ms(fn):
            for _ in range(3):
                fn()
            ev = []
            for _ in range(10):
                flush.sum()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream); fn(); b.record(stream)
                ev.append((a, b))
            torch.cuda.synchronize()
            return float(np.median([a.elapsed_time(b) for a, b in ev]))

        res = {}
        for kind in ('regular', 'jittered'):
            res[kind] = {'ring_ms': med_ms(lambda: ring_step(kind)), 'plain_columns_ms': med_ms(lambda: plain_step(kind))}
            torch.cuda.synchronize()
            res[kind]['identical_to_th_usage_reduce_u8'] = all(bool(torch.equal(x, y)) for x, y in
                                                               zip(r_out[kind], r_out['plain']))
        info = [C.c_int64(0) for _ in range(6)]
        native.check(lib.th_ring_info(rings['regular'], *[C.byref(x) for x in info]))
        ring = {'samples': RU * RSPU, 'uuids': RU, 'windows': RW, 'append_ms_per_tick_of_8192_samples': float(np.median(tick_ms)),
                'relayouts': info[5].value, 'regular_uuids': info[4].value, 'device_bytes': info[3].value,
                'regular_stamps': res['regular'], 'jittered_stamps': res['jittered'],
                'note': '1/8 of config 3, appended through th_ring_append in batches of 1024 ticks; th_ring_reduce reads the ring in '
                        'place (device window arrays in, device results out); with stamps on an exact grid both window ends '
                        'come from divisions and the ts column is not read'}
        for h in rings.values():
            lib.th_ring_destroy(h)
        del cols, r_out
    traffic = _traffic('k4_usage_kernel')
    return {'metric': 'usage_reduce_GB_per_s', 'value': world * gbs_survey, 'unit': 'GB/s', 'ms_per_step': ms,
            'value_note': 'SURVEY 8d definition: 24 B per logged sample (ts, gpu_util, mem_util) / time; above the HBM peak '
                          'is possible because the ts column is probed, not streamed -- see roofline for the physical rate',
            'samples_per_s': world * s_total / (ms * 1e-3),
            'config': {'workload': 'BASELINE configs[2]: 1B samples (1 s cadence) -> 10M reservation windows per GPU',
                       'samples': s_total, 'windows': S_WINDOWS, 'uuids': S_UUIDS,
                       'l2': 'inputs (24 GB) exceed L2; no flush needed', 'samples_inside_windows': covered,
                       'empty_windows': int((cnt == 0).sum().item())},
            'roofline': {'kernel': 'k4_usage_kernel', 'bound': 'hbm', 'achieved': gbs, 'peak': peak, 'unit': 'GB/s',
                         'frac': gbs / peak, 'traffic': traffic, 'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': alg, 'frac_of_nominal_8TBs': gbs / 8000.0,
                         'algorithmic_bytes_note': '16 B/sample (gpu_util + mem_util int64) + 64 B/window + 8 B/uuid; the ts '
                                                   'column (8 of SURVEY 8d\'s 24 B/sample) is only probed',
                         'dram_GBs_from_ncu_traffic': (traffic / (ms * 1e-3) / 1e9) if traffic else None,
                         'frac_survey_8d_bytes': gbs_survey / peak},
            'cpu_baseline': cpu,
            'windows_shuffled': {'ms_per_step': ms_shuffled, 'roofline_frac': alg / (ms_shuffled * 1e-3) / 1e9 / peak,
                                 'slowdown_vs_sorted_windows': ms_shuffled / ms, 'results_identical': shuffled_same,
                                 'note': 'the same 10M windows in random order (as a reservation table in id order would '
                                         'present them) instead of grouped by uuid with increasing starts'},
            'sample_ring': ring,
            'narrow_u8': {'ms_per_step': ms8, 'samples_per_s': world * s_total / (ms8 * 1e-3),
                          'speedup_vs_int64_columns': ms / ms8, 'results_identical_to_int64_path': same,
                          'roofline': {'kernel': 'k4_usage_u8_kernel', 'bound': 'hbm', 'achieved': gbs8, 'peak': peak,
                                       'unit': 'GB/s', 'frac': gbs8 / peak, 'algorithmic_bytes_per_launch': alg8,
                                       'traffic': _traffic('k4_usage_u8_kernel'),
                                       'note': 'algorithmic bytes = 2 B/sample (uint8 gpu_util + mem_util) + 48 B/window; '
                                               'ts is probed, not streamed'}}}


def bench_cfg5(torch, lib, native, dist, dev, local, rank, world, sp, stream, flush, steps, warmup, peak, peak_src, barrier,
               max_over_ranks):
    """BASELINE configs[4] at this job's size: ONE calendar of world x 12.5 M reservations over world x 8192 uuids (100 M /
    65 536 at 8 GPUs), defined chunk by chunk so that no process ever holds all of it.  `sharding.partition_uuids` cuts it
    by rows; each rank materialises and loads only its uuid range and the candidates `sharding.owner_rank` routes to it
    (12.5 M per GPU).  Before timing, a REPLICATED batch over sampled chunks (one inside every rank's range, some
    straddling a range boundary, 1 % unknown resources) goes through `sharding.Router` and the fused exchange, and every
    rank decodes ALL ranks' words and compares them with the brute-force oracle."""
    import ctypes as C
    import tensorhive_fixed_b200 as th
    from tensorhive_fixed_b200 import sharding, synth
    from oracle import brute                                      # the checker; nothing timed goes through it
    t_start = time.perf_counter()
    U, N = C5_UUIDS * world, C5_ROWS * world
    gcal = synth.GlobalCalendar(N, U, seed=SEED + 50)
    bounds = sharding.partition_uuids(gcal.rows_per_uuid_estimate(), world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    cols = ('id', 'uuid_id', 'start', 'end', 'owner', 'cancelled')
    qcols = ('uuid_id', 'start', 'end', 'self_id')
    parts, qparts = [], []
    for c in gcal.chunks_overlapping(lo, hi):
        cal_c = gcal.chunk(c)
        q_c = gcal.chunk_queries(c, len(cal_c['id']), seed=SEED + 60, cal=cal_c)
        sel = sharding.owner_rank(cal_c['uuid_id'], bounds) == rank
        parts.append({k: cal_c[k][sel] for k in cols})
        qsel = sharding.owner_rank(q_c['uuid_id'], bounds) == rank          # the Router's rule, chunk by chunk
        qparts.append({k: q_c[k][qsel] for k in qcols})
    mine = synth.concat_columns(parts, cols)
    q = synth.concat_columns(qparts, qcols)
    del parts, qparts
    mine['uuid_id'] = mine['uuid_id'] - lo
    q['uuid_id'] = q['uuid_id'] - lo
    n_r, m_r, u_r = int(len(mine['id'])), int(len(q['uuid_id'])), max(hi - lo, 1)
    gen_s = time.perf_counter() - t_start
    snap = th.Snapshot(local)
    t0 = time.perf_counter()
    snap.load(*[mine[k] for k in cols], n_uuids=u_r)
    load_s = time.perf_counter() - t0
    del mine
    tdev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    qd = [tdev(q[k]) for k in qcols]
    words = torch.empty((m_r + 31) // 32, dtype=torch.int32, device=dev)
    m_all = [m_r]
    n_all = [n_r]
    if dist is not None:
        t = torch.tensor([m_r, n_r], dtype=torch.int64, device=dev)
        g = torch.empty(2 * world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(g, t)
        g = g.view(world, 2).cpu().numpy()
        m_all, n_all = [int(x) for x in g[:, 0]], [int(x) for x in g[:, 1]]

    # ---- the parity batch: replicated on every rank, routed by Router, all ranks' words decoded everywhere ---------
    cpu_ = gcal.chunk_uuids
    pick = set()
    for r in range(world):
        cs = list(gcal.chunks_overlapping(int(bounds[r]), int(bounds[r + 1])))
        if cs:
            pick.add(cs[len(cs) // 2])
    for r in sorted({1, world // 2, world - 1}):
        if 0 < r < world:
            pick.add(int(bounds[r]) // cpu_)                 # the chunk a range boundary falls into (or starts)
    if world == 1:
        pick.update({0, gcal.n_chunks - 1})
    pick = sorted(pick)
    pcal_parts, pq_parts = [], []
    for c in pick:
        cal_c = gcal.chunk(c)
        pcal_parts.append(cal_c)
        pq_parts.append(gcal.chunk_queries(c, 8000, seed=SEED + 70, unknown_frac=0.01, cal=cal_c))
    pcal = synth.concat_columns(pcal_parts, cols)
    pq = synth.concat_columns(pq_parts, qcols)
    del pcal_parts, pq_parts
    router = sharding.Router(pq['uuid_id'], bounds)
    idx = router.local_indices(rank)
    pqd = [tdev(router.local_uuid(pq['uuid_id'], rank)), tdev(pq['start'][idx]), tdev(pq['end'][idx]),
           tdev(pq['self_id'][idx])]
    px = None
    if world > 1:
        px = sharding.PeerExchange(local, world, rank, max(max(m_all), int(router.counts.max())))
        gw = px.conflict_words(snap, *pqd, counts=router.counts)          # check=True: raises if a peer was late
        gathered = gw.cpu().numpy()
    else:
        gathered = snap.conflict_words(*pqd).cpu().numpy().reshape(1, -1)
    got = router.decode(gathered)
    t0 = time.perf_counter()
    want = brute.conflict_bits(pcal, pq['uuid_id'], pq['start'], pq['end'], pq['self_id'])
    oracle_s = time.perf_counter() - t0
    mism = int((got != want).sum())

    def total_over_ranks(x: int) -> int:
        if dist is None:
            return int(x)
        t = torch.tensor([int(x)], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        return int(t.item())

    # ---- K3 + all-gather: the per-uuid (current reservation, owner) table of the whole calendar on every rank ----------
    now5 = synth.T0_US + 1000 * synth.DAY_US
    dname = 'cuda:%d' % local
    if dist is not None:
        gather64 = lambda v: sharding.torch_all_gather_i64(v, dname)
    else:
        gather64 = lambda v: np.asarray(v, np.int64).reshape(1, -1)
    owner_ms = []
    for _ in range(4):
        barrier()
        t0 = time.perf_counter()
        res_all, own_all = sharding.sharded_current_owner(bounds, rank, lambda: snap.current_owner(now5), gather64)
        owner_ms.append(1e3 * (time.perf_counter() - t0))
    want_res, want_own = brute.current_owner(pcal, U, now5)
    covered = np.zeros(U, bool)
    for c in pick:
        covered[c * cpu_:(c + 1) * cpu_] = True
    mism_owner = int((res_all[covered] != want_res[covered]).sum() + (own_all[covered] != want_own[covered]).sum())

    # ---- K4 + all-gather: window averages of a replicated window list, samples sharded by the same bounds ---------------
    spu = C5_SAMPLES_PER_UUID

    def chunk_samples(c):
        return synth.make_samples(cpu_, spu, seed=SEED + 80 + c)

    s_parts = []
    for c in gcal.chunks_overlapping(lo, hi):
        sc_ = chunk_samples(c)
        a, b = max(lo, c * cpu_) - c * cpu_, min(hi, (c + 1) * cpu_) - c * cpu_
        s_parts.append({k: sc_[k][a * spu:b * spu] for k in ('ts', 'gpu', 'mem')})
    smp = synth.concat_columns(s_parts, ('ts', 'gpu', 'mem'))
    del s_parts
    d_off = tdev(np.arange(u_r + 1, dtype=np.int64) * spu if hi > lo else np.zeros(2, np.int64))
    d_ts, d_g, d_m = tdev(smp['ts']), tdev(smp['gpu']), tdev(smp['mem'])
    win = synth.make_window_batch(U, spu, C5_WINDOWS, seed=SEED + 90)

    def local_reduce(wu, ws, we):
        return tuple(x.cpu().numpy() for x in th.usage_reduce(d_off, d_ts, d_g, d_m, tdev(wu), tdev(ws), tdev(we)))

    win_ms = []
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        ag, am = sharding.sharded_window_averages(win['uuid_id'], win['start'], win['end'], bounds, rank, local_reduce,
                                                  gather64)
        win_ms.append(1e3 * (time.perf_counter() - t0))
    wsel = np.nonzero(covered[win['uuid_id']])[0][:20000]
    # oracle over the sampled chunks only: their samples laid out as a CSR over the global uuid space
    off_g = np.zeros(U + 1, np.int64)
    cnt_g = np.zeros(U, np.int64)
    cnt_g[covered] = spu
    off_g[1:] = np.cumsum(cnt_g)
    ps = synth.concat_columns([chunk_samples(c) for c in pick], ('ts', 'gpu', 'mem'))
    sg, sm, cn = brute.usage_reduce(off_g, ps['ts'], ps['gpu'], ps['mem'], win['uuid_id'][wsel], win['start'][wsel],
                                    win['end'][wsel])
    want_ag = np.array([brute.floor_avg(a, c) for a, c in zip(sg, cn)], np.int64)
    want_am = np.array([brute.floor_avg(a, c) for a, c in zip(sm, cn)], np.int64)
    mism_win = int((ag[wsel] != want_ag).sum() + (am[wsel] != want_am).sum())
    del d_ts, d_g, d_m, smp, ps

    # ---- timing: the fused step over this rank's 12.5 M routed candidates ------------------------------------------------
    def step():
        if px is None:
            native.check(lib.th_conflict_bitmap(snap._h, m_r, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(),
                                                qd[3].data_ptr(), 0, words.data_ptr(), 1, sp))
        else:
            px.conflict_words(snap, qd[0], qd[1], qd[2], qd[3], counts=m_all, check=False)

    def plain():
        native.check(lib.th_conflict_bitmap(snap._h, m_r, qd[0].data_ptr(), qd[1].data_ptr(), qd[2].data_ptr(),
                                            qd[3].data_ptr(), 0, words.data_ptr(), 1, sp))

    def timed(fn, n, meet):
        evs = []
        for _ in range(n):
            flush.sum()
            if meet and px is not None:
                px.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream); fn(); b.record(stream)
            evs.append((a, b))
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    for _ in range(max(3, min(warmup, 5))):
        flush.sum()
        step()
    barrier()
    times = timed(step, steps, True)
    barrier()
    if px is not None:
        px.check()
    total_ms = max_over_ranks(float(sum(times)))
    k_ms = float(np.mean(timed(plain, 10, False)))
    k_ms_max = max_over_ranks(k_ms)
    # the words the fused step left on this rank against the plain kernel: own row here, every row via the parity batch
    if px is not None:
        gw = px.conflict_words(snap, qd[0], qd[1], qd[2], qd[3], counts=m_all)
        plain(); torch.cuda.synchronize()
        assert bool(torch.equal(gw[rank, :words.numel()], words)), 'config-5 share: fused words differ from the plain kernel'
    value = float(sum(m_all)) * steps / (total_ms * 1e-3)
    single = (float(sum(m_all)) / world) / (k_ms_max * 1e-3)          # one GPU on one share, no exchange
    alg = m_r * 32 + n_r * 25 + 8 * (u_r + 1) + m_r // 8
    achieved = alg / (k_ms * 1e-3) / 1e9
    traffic = _traffic('k2_bitmap_kernel_cfg5_share')
    out = {'workload': "BASELINE configs[4]: ONE calendar of %d reservations x %d GPU UUIDs sharded by uuid range over %d GPU(s); "
                       '%d routed candidate conflict queries per step in total' % (gcal.n_rows, U, world, sum(m_all)),
           'value': value, 'unit': 'queries/s', 'ms_per_step': total_ms / steps, 'steps': steps,
           'speedup_vs_n1': value / single,
           'speedup_note': 'value / (queries per GPU / plain-kernel time of one share on one GPU, slowest rank, this run); '
                           "per_n.1.cfg5_share.value is the same quantity measured by the N = 1 run",
           'rows_per_rank': n_all, 'queries_per_rank': m_all, 'uuid_bounds': [int(b) for b in bounds],
           'routing': 'sharding.partition_uuids on rows per uuid; candidates routed by sharding.owner_rank chunk by chunk '
                      '(what sharding.Router does for a whole batch)',
           'l2': 'flushed between steps (512 MiB read sweep, outside the timed events)',
           'roofline': {'kernel': 'k2_bitmap_kernel', 'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                        'frac': achieved / peak, 'traffic': traffic, 'peak_source': peak_src,
                        'algorithmic_bytes_per_launch': alg, 'kernel_ms': k_ms, 'kernel_ms_slowest_rank': k_ms_max,
                        'frac_of_nominal_8TBs': achieved / 8000.0,
                        'dram_over_algorithmic': (traffic / alg) if traffic else None},
           'gathers': {'current_owner_table': {'ms_wall': float(np.median(owner_ms)), 'uuids': U, 'bytes_gathered': 16 * U,
                                               'what': 'sharding.sharded_current_owner: K3 on the shard + all-gather of '
                                                       '(cur_res_id, cur_owner) + host layout, wall clock'},
                       'window_averages': {'ms_wall': float(np.median(win_ms)), 'windows': C5_WINDOWS,
                                           'samples_per_gpu': int(u_r * spu), 'bytes_gathered': 16 * C5_WINDOWS,
                                           'what': 'sharding.sharded_window_averages: Router over the replicated windows, H2D, '
                                                   'K4 on the sample shard, floor averages, all-gather, host layout, '
                                                   'wall clock'}},
           'host_s': {'generate': gen_s, 'first_load_from_host_arrays': load_s, 'parity_oracle': oracle_s},
           'snapshot_device_bytes': snap.info()['device_bytes'],
           'parity': {'checked': True, 'ranks': world, 'queries': int(router.m),
                      'rows_checked_against': int(len(pcal['id'])), 'chunks': [int(c) for c in pick],
                      'conflicts_expected': int(want.sum()),
                      'mismatches': total_over_ranks(mism),
                      'owner_table_mismatches': total_over_ranks(mism_owner),
                      'window_average_mismatches': total_over_ranks(mism_win), 'windows_checked': int(wsel.size),
                      'what': 'a replicated batch routed by sharding.Router, evaluated by every rank on its shard of the '
                              'global calendar, exchanged by the fused kernel; EVERY rank decodes ALL ranks\' words and '
                              'compares with the brute-force oracle over the sampled chunks (mismatches are summed over '
                              'ranks); the K3 owner table and the K4 window averages after their all-gathers likewise'}}
    if px is not None:
        dist.barrier()
        px.close()
    snap.close()
    return out


def main():
    global N_ROWS, N_UUIDS, N_QUERIES
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--skip-usage', action='store_true', help='skip the config-3 usage-reduce measurement')
    ap.add_argument('--skip-sweep', action='store_true', help='skip the config-4 service sweep measurement')
    ap.add_argument('--skip-cpu', action='store_true', help='skip the single-core CPU baseline leg')
    ap.add_argument('--skip-cfg5', action='store_true', help="skip the config-5 share (global calendar, parity, gathers)")
    ap.add_argument('--workload', default='cfg2', choices=['cfg2', 'cfg5'],
                    help="cfg2 = BASELINE configs[1] (default, the metric's configuration); cfg5 = one rank's share of "
                         'configs[4]: 12.5M reservations x 8192 uuids, 12.5M queries per GPU (100M calendar on 8 GPUs)')
    args = ap.parse_args()
    if args.workload == 'cfg5':
        N_ROWS, N_UUIDS, N_QUERIES = 12_500_000, 8192, 12_500_000
        args.skip_sweep = True
        args.skip_cfg5 = True
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
