"""Synthetic calendars, candidate batches, usage logs and process tables.

Pure numpy, seeded: the same arrays feed the CUDA path and the oracle.  Sizes above ~10^7 rows are
generated on the device instead (``th_synth_*`` in the C ABI) because host generation + upload would
dominate the run.
"""
from __future__ import annotations

import numpy as np

T0_US = 1767225600 * 1_000_000          # 2026-01-01T00:00:00Z
MIN_US = 60 * 1_000_000
HOUR_US = 60 * MIN_US
DAY_US = 24 * HOUR_US
MIN_DURATION_US = 30 * MIN_US           # tensorhive/models/Reservation.py:39
MAX_DURATION_US = 8 * DAY_US            # tensorhive/models/Reservation.py:40


def uuid_string(k: int) -> str:
    """40-character resource id in the reference's format (Reservation.py:52)."""
    h = '%032x' % (0xd38d4de385eee8373d87e8e2faeb0000 + int(k))
    return 'GPU-%s-%s-%s-%s-%s' % (h[:8], h[8:12], h[12:16], h[16:20], h[20:])


def make_calendar(n: int, n_uuids: int, seed: int, skew: float = 0.0, shuffle_ids: bool = True) -> dict:
    """``n`` reservations over ``n_uuids`` GPUs.

    Per uuid a walk from T0: gap ~ Exp(mean 2 h) rounded to us, exactly 0 for 10 % (touching neighbours);
    duration ~ U[30 min, 8 d] with us jitter.  10 % extra rows are cancelled and may overlap anything.
    ``is_cancelled`` raw values: NULL (-1 here) 60 %, 0 30 %, 1 10 %.  owner ~ U[1, 1000].
    ``skew`` > 0 makes uuid popularity Zipf-like.  ids are a random permutation of 1..n when
    ``shuffle_ids`` (insertion order != calendar order), else ascending.
    """
    rng = np.random.default_rng(seed)
    if n == 0:
        z = np.zeros(0, np.int64)
        return {'id': z, 'uuid_id': z.copy(), 'start': z.copy(), 'end': z.copy(), 'owner': z.copy(),
                'cancelled': np.zeros(0, np.uint8), 'cancelled_raw': z.copy(), 'n_uuids': n_uuids}
    if skew > 0:
        p = 1.0 / np.arange(1, n_uuids + 1) ** skew
        p /= p.sum()
        uuid = rng.choice(n_uuids, size=n, p=p).astype(np.int64)
    else:
        uuid = rng.integers(0, n_uuids, size=n, dtype=np.int64)
    uuid.sort()
    is_canc = rng.random(n) < 0.10
    gap = np.round(rng.exponential(2 * HOUR_US, size=n)).astype(np.int64)
    gap[rng.random(n) < 0.10] = 0
    dur = rng.integers(MIN_DURATION_US, MAX_DURATION_US + 1, size=n, dtype=np.int64)
    # live rows advance the walk; cancelled rows are placed afterwards
    step = np.where(is_canc, 0, gap + dur)
    if float(step.sum(dtype=np.float64)) > 9.0e18:
        raise ValueError('make_calendar(%d rows): the walk would overflow int64 microseconds -- build calendars of this '
                         'size chunk by chunk (GlobalCalendar)' % n)
    csum = np.cumsum(step)
    first = np.ones(n, bool)
    first[1:] = uuid[1:] != uuid[:-1]
    seg_base = np.maximum.accumulate(np.where(first, csum - step, 0))
    end = T0_US + (csum - seg_base)
    start = end - dur
    # cancelled rows: anywhere inside the span their uuid has reached so far, any legal duration
    span = np.maximum(end - T0_US, DAY_US)
    cs = T0_US + (rng.random(n) * span).astype(np.int64)
    start = np.where(is_canc, cs, start)
    end = np.where(is_canc, cs + dur, end)
    raw = np.where(is_canc, 1, np.where(rng.random(n) < 2.0 / 3.0, -1, 0)).astype(np.int64)
    owner = rng.integers(1, 1001, size=n, dtype=np.int64)
    ids = np.arange(1, n + 1, dtype=np.int64)
    if shuffle_ids:
        ids = rng.permutation(ids)
    # present rows in insertion (id) order when shuffled so that input order is NOT calendar order
    order = np.argsort(ids, kind='stable') if shuffle_ids else np.arange(n)
    return {'id': ids[order], 'uuid_id': uuid[order], 'start': start[order], 'end': end[order],
            'owner': owner[order], 'cancelled': (raw[order] == 1).astype(np.uint8), 'cancelled_raw': raw[order],
            'n_uuids': n_uuids}


def make_queries(cal: dict, m: int, seed: int, unknown_frac: float = 0.01) -> dict:
    """Candidate reservations: 50 % perturbations of an existing row in the four shapes of the reference's
    test (inside / left / right / enclosing, tests/unit/models/test_reservation_model.py:10-37), 25 %
    placed at random (mostly gaps), 15 % exactly touching a row, 10 % self-updates (self_id = row id)."""
    rng = np.random.default_rng(seed)
    n = len(cal['id'])
    n_uuids = int(cal['n_uuids'])
    if n == 0:
        u = rng.integers(0, max(n_uuids, 1), size=m, dtype=np.int64)
        s = T0_US + rng.integers(0, 30 * DAY_US, size=m, dtype=np.int64)
        return {'uuid_id': u, 'start': s, 'end': s + HOUR_US, 'self_id': np.full(m, -1, np.int64)}
    pick = rng.integers(0, n, size=m)
    rs, re_, ru, rid = cal['start'][pick], cal['end'][pick], cal['uuid_id'][pick], cal['id'][pick]
    cls = rng.random(m)
    off = rng.integers(1, 5 * MIN_US + 1, size=m, dtype=np.int64)
    shape = rng.integers(0, 4, size=m)
    qs = np.where((shape == 0) | (shape == 2), rs + off, rs - off)
    qe = np.where((shape == 0) | (shape == 1), re_ - off, re_ + off)
    self_id = np.full(m, -1, np.int64)
    # random placement
    r = (cls >= 0.50) & (cls < 0.75)
    span = int(cal['end'].max() - T0_US) + DAY_US
    rnd_s = T0_US + (rng.random(m) * span).astype(np.int64)
    rnd_d = rng.integers(MIN_DURATION_US, 2 * DAY_US, size=m, dtype=np.int64)
    qs = np.where(r, rnd_s, qs)
    qe = np.where(r, rnd_s + rnd_d, qe)
    # touching: starts exactly at a row's end, or ends exactly at a row's start
    t = (cls >= 0.75) & (cls < 0.90)
    side = rng.random(m) < 0.5
    t_d = rng.integers(MIN_DURATION_US, DAY_US, size=m, dtype=np.int64)
    qs = np.where(t, np.where(side, re_, rs - t_d), qs)
    qe = np.where(t, np.where(side, re_ + t_d, rs), qe)
    # self-update: move the row a little, pass its own id
    s = cls >= 0.90
    qs = np.where(s, rs + off - 2 * MIN_US, qs)
    qe = np.where(s, re_ + off, qe)
    self_id = np.where(s, rid, self_id)
    # a few queries for resources the calendar has never seen
    unknown = rng.random(m) < unknown_frac
    ru = np.where(unknown, n_uuids + 3, ru)
    return {'uuid_id': ru.astype(np.int64), 'start': qs.astype(np.int64), 'end': qe.astype(np.int64),
            'self_id': self_id.astype(np.int64)}


def make_samples(n_uuids: int, samples_per_uuid: int, seed: int, cadence_us: int = 1_000_000, jitter_us: int = 0,
                 ragged: bool = False) -> dict:
    """Usage log: per uuid ``samples_per_uuid`` samples at ``cadence_us`` (+- jitter), ts ascending.
    gpu_util: 0 w.p. 0.3 else U{0..100}; mem_util: U{0..100} (ints, NvidiaSmiParser.py:84-85)."""
    rng = np.random.default_rng(seed)
    if ragged:
        counts = rng.integers(0, 2 * samples_per_uuid + 1, size=n_uuids)
        counts[rng.random(n_uuids) < 0.2] = 0
    else:
        counts = np.full(n_uuids, samples_per_uuid)
    offsets = np.zeros(n_uuids + 1, np.int64)
    offsets[1:] = np.cumsum(counts)
    s = int(offsets[-1])
    k = np.arange(s, dtype=np.int64) - np.repeat(offsets[:-1], counts)
    ts = T0_US + k * cadence_us
    if jitter_us:
        assert jitter_us < cadence_us, 'jitter must stay below the cadence so that ts remains ascending'
        ts = ts + rng.integers(0, jitter_us + 1, size=s, dtype=np.int64)
    gpu = np.where(rng.random(s) < 0.3, 0, rng.integers(0, 101, size=s)).astype(np.int64)
    mem = rng.integers(0, 101, size=s, dtype=np.int64)
    return {'offsets': offsets, 'ts': ts.astype(np.int64), 'gpu': gpu, 'mem': mem, 'n_uuids': n_uuids}


def make_windows(samples: dict, w: int, seed: int, min_len_us: int = 20_000_000, max_len_us: int = 180_000_000) -> dict:
    """Reservation windows over a usage log: mostly inside the logged span, 10 % touching the previous
    window's end exactly, a few empty (before the first sample), a few covering everything, a few on
    uuids with no samples, ends landing exactly on sample timestamps half of the time."""
    rng = np.random.default_rng(seed)
    n_uuids = samples['n_uuids']
    off, ts = samples['offsets'], samples['ts']
    wu = rng.integers(0, n_uuids, size=w, dtype=np.int64)
    ws = np.zeros(w, np.int64)
    we = np.zeros(w, np.int64)
    for j in range(w):
        a, b = int(off[wu[j]]), int(off[wu[j] + 1])
        lo, hi = (int(ts[a]), int(ts[b - 1])) if b > a else (T0_US, T0_US + HOUR_US)
        c = rng.random()
        length = int(rng.integers(min_len_us, max_len_us + 1))
        if c < 0.03:
            ws[j], we[j] = lo - 10 * HOUR_US, lo - 9 * HOUR_US          # nothing logged -> -1
        elif c < 0.06:
            ws[j], we[j] = lo - HOUR_US, hi + HOUR_US                    # everything
        elif c < 0.16 and j > 0 and wu[j - 1] == wu[j]:
            ws[j], we[j] = we[j - 1], we[j - 1] + length                 # touches the previous window
        else:
            s0 = int(rng.integers(lo - length, hi + 1))
            if b > a and rng.random() < 0.5:
                s0 = int(ts[int(rng.integers(a, b))])                    # exactly on a sample
            e0 = s0 + length
            if b > a and rng.random() < 0.5:
                e0 = max(s0, int(ts[int(rng.integers(a, b))]))
            ws[j], we[j] = s0, e0
    return {'uuid_id': wu, 'start': ws, 'end': we}


def make_processes(cal: dict, p: int, now_us: int, seed: int, n_users: int = 1000) -> dict:
    """GPU process table for the ProtectionService join: 20 % of GPUs hold 80 % of the processes, 30 % of
    GPUs have none; owner = the current reservation's owner w.p. 0.7, another known user 0.25,
    unknown / None (negative id) 0.05."""
    rng = np.random.default_rng(seed)
    n_uuids = int(cal['n_uuids'])
    active = np.nonzero(rng.random(n_uuids) >= 0.30)[0]
    if active.size == 0:
        active = np.arange(n_uuids)
    hot = active[: max(1, active.size // 5)]
    pick_hot = rng.random(p) < 0.8
    uuid = np.where(pick_hot, hot[rng.integers(0, hot.size, size=p)], active[rng.integers(0, active.size, size=p)])
    ok = (cal['start'] <= now_us) & (now_us <= cal['end']) & (cal['cancelled'] == 0)
    cur_owner = np.full(n_uuids, -1, np.int64)
    idx = np.nonzero(ok)[0]
    order = np.argsort(-cal['id'][idx], kind='stable')       # lowest id written last wins
    cur_owner[cal['uuid_id'][idx][order]] = cal['owner'][idx][order]
    c = rng.random(p)
    other = rng.integers(1, n_users + 1, size=p, dtype=np.int64)
    owner = np.where(c < 0.70, np.where(cur_owner[uuid] > 0, cur_owner[uuid], other),
                     np.where(c < 0.95, other, -2 - rng.integers(0, 5, size=p)))
    return {'uuid_id': uuid.astype(np.int64), 'owner': owner.astype(np.int64),
            'pid': rng.integers(100, 4_000_000, size=p, dtype=np.int64),
            'host_id': (uuid // 8).astype(np.int64)}


def make_job_tasks(cal: dict, jobs: int, seed: int, n_users: int = 1000) -> dict:
    """Queued jobs with U{1..4} tasks each; a task names a GPU uuid (or -1: no assigned GPU)."""
    rng = np.random.default_rng(seed)
    n_uuids = int(cal['n_uuids'])
    per = rng.integers(1, 5, size=jobs)
    t = int(per.sum())
    job_id = np.repeat(np.arange(jobs, dtype=np.int64), per)
    job_owner = rng.integers(1, n_users + 1, size=jobs, dtype=np.int64)
    uuid = rng.integers(0, n_uuids, size=t, dtype=np.int64)
    uuid[rng.random(t) < 0.02] = -1
    return {'job_id': job_id, 'uuid_id': uuid, 'owner': job_owner[job_id], 'tasks_per_job': per.astype(np.int64)}


# ---- one calendar too large for one host process: generated chunk by chunk (BASELINE configs[4]) -----------------------
class GlobalCalendar:
    """A calendar of ``n_uuids`` GPUs and about ``n_rows`` reservations defined CHUNK BY CHUNK (``chunk_uuids`` uuids per
    chunk, every chunk from its own seed), so that each rank of a multi-GPU job can materialise exactly the uuid range it
    owns -- and any rank can materialise any other chunk to check the answers that arrive from its peers -- without a
    process ever holding all rows.  Chunk sizes vary by +-25 % (mild popularity skew), so
    the uuid partition is balanced by rows, not by uuid count.  Reservation ids are globally unique."""

    def __init__(self, n_rows: int, n_uuids: int, seed: int, chunk_uuids: int = 256):
        assert n_uuids % chunk_uuids == 0
        self.n_uuids, self.chunk_uuids, self.seed = int(n_uuids), int(chunk_uuids), int(seed)
        self.n_chunks = n_uuids // chunk_uuids
        rng = np.random.default_rng(seed)
        weight = rng.uniform(0.75, 1.25, size=self.n_chunks)
        rows = np.floor(weight / weight.sum() * n_rows).astype(np.int64)
        rows[0] += n_rows - int(rows.sum())
        self.chunk_rows = rows
        self.id_base = np.concatenate([[0], np.cumsum(rows)])
        self.n_rows = int(rows.sum())

    def rows_per_uuid_estimate(self) -> np.ndarray:
        """What the partitioner balances on: a chunk's rows spread evenly over its uuids."""
        return np.repeat(self.chunk_rows / self.chunk_uuids, self.chunk_uuids)

    def chunk(self, c: int) -> dict:
        """Rows of chunk ``c`` with GLOBAL uuid codes and ids."""
        cal = make_calendar(int(self.chunk_rows[c]), self.chunk_uuids, seed=self.seed + 1000 + c)
        cal['uuid_id'] = cal['uuid_id'] + c * self.chunk_uuids
        cal['id'] = cal['id'] + int(self.id_base[c])
        cal['n_uuids'] = self.n_uuids
        return cal

    def chunk_queries(self, c: int, m: int, seed: int, unknown_frac: float = 0.0, cal: dict = None) -> dict:
        """``m`` candidates against chunk ``c`` (global uuid codes; unknown resources get code n_uuids + 3)."""
        cal = cal if cal is not None else self.chunk(c)
        local = dict(cal)
        local['n_uuids'] = self.n_uuids
        return make_queries(local, m, seed=seed + c, unknown_frac=unknown_frac)

    def chunks_overlapping(self, lo: int, hi: int):
        return range(lo // self.chunk_uuids, (max(hi, lo + 1) - 1) // self.chunk_uuids + 1) if hi > lo else range(0)


def concat_columns(parts, keys) -> dict:
    return {k: (np.concatenate([p[k] for p in parts]) if parts else np.zeros(0, np.int64)) for k in keys}


def make_window_batch(n_uuids: int, samples_per_uuid: int, w: int, seed: int, cadence_us: int = 1_000_000) -> dict:
    """Vectorised window list over a regular-cadence log (every uuid: ``samples_per_uuid`` samples from T0): lengths
    U[20 s, 180 s], starts anywhere in the logged span (some windows run past either end), 2 % before the log."""
    rng = np.random.default_rng(seed)
    wu = rng.integers(0, n_uuids, size=w, dtype=np.int64)
    span = samples_per_uuid * cadence_us
    ws = T0_US + rng.integers(-100_000_000, span, size=w, dtype=np.int64)
    ws = np.where(rng.random(w) < 0.5, ws - ws % cadence_us, ws)         # half of the starts exactly on a sample
    we = ws + rng.integers(20_000_000, 180_000_001, size=w, dtype=np.int64)
    empty = rng.random(w) < 0.02
    ws = np.where(empty, T0_US - 10 * HOUR_US, ws)
    we = np.where(empty, T0_US - 9 * HOUR_US, we)
    return {'uuid_id': wu, 'start': ws, 'end': we}
