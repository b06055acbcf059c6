#!/usr/bin/env python
"""Differential fuzzer for K4: random usage logs (exact cadence, jitter, missing stretches, repeated stamps, bursts, empty
uuids), random windows (on samples, beside them, far outside, reversed), through th_usage_reduce (int64 columns),
th_usage_reduce_u8 and the device ring (appended in random batches), against the brute-force oracle.
usage: python tools/fuzz_usage.py [n_logs] [first_seed]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensorhive_fixed_b200 as th                      # noqa: E402
from tensorhive_fixed_b200 import native                # noqa: E402
from oracle import brute                                # noqa: E402

T0 = 1_700_000_000_000_000


def make_log(seed):
    rng = np.random.default_rng(seed)
    u = int(rng.integers(1, 30))
    segs = []
    for k in range(u):
        n = int(rng.choice([0, 1, 2, 15, 16, 17, 33, int(rng.integers(0, 3000))]))
        kind = int(rng.integers(0, 6))
        cad = int(rng.choice([1, 7, 1000, 1_000_000, 60_000_000]))
        t = T0 + int(rng.integers(0, 10**9)) + np.arange(n, dtype=np.int64) * cad            # 0: exact cadence
        if kind == 1:
            t = t + rng.integers(0, max(cad - 1, 1), size=n)                                 # jitter below one step
        elif kind == 2:
            t = t[rng.random(n) < 0.7]                                                       # missing samples
        elif kind == 3:
            t = np.sort(rng.choice(t, size=n)) if n else t                                    # repeated stamps
        elif kind == 4:
            t = T0 + np.sort(rng.integers(0, 50, size=n) + rng.choice([0, 10**9, 10**12], size=n))   # bursts
        elif kind == 5 and n > 4:
            t[n // 2:] += int(rng.choice([1, cad, 10**10]))                                   # one break in the run
        segs.append(np.sort(t.astype(np.int64)))
    off = np.concatenate([[0], np.cumsum([len(s) for s in segs])]).astype(np.int64)
    ts = np.concatenate(segs) if segs else np.zeros(0, np.int64)
    return u, segs, off, ts, rng


def make_windows(u, segs, ts, rng, w):
    wu = rng.integers(-1, u + 1, size=w).astype(np.int64)
    ws, we = np.empty(w, np.int64), np.empty(w, np.int64)
    for j in range(w):
        s = segs[wu[j]] if 0 <= wu[j] < u else np.zeros(0, np.int64)
        c = rng.random()
        if len(s) and c < 0.75:
            i, k = sorted(rng.integers(0, len(s), size=2))
            if rng.random() < 0.5:
                k = min(len(s) - 1, i + int(rng.integers(0, 40)))
            ws[j], we[j] = s[i] + rng.integers(-1, 2), s[k] + rng.integers(-1, 2)
        elif c < 0.85:
            ws[j], we[j] = T0 - 10**6, T0 + 10**14
        elif c < 0.95:
            a = T0 + int(rng.integers(-10**9, 10**13))
            ws[j], we[j] = a, a + int(rng.integers(0, 10**9))
        else:
            ws[j], we[j] = T0 + 500, T0 + 100                                                 # end before start
    return wu, ws, we


def check_log(seed, verbose=True):
    u, segs, off, ts, rng = make_log(seed)
    s_total = int(off[-1])
    g8 = rng.integers(0, 256, size=s_total, dtype=np.uint8)
    m8 = rng.integers(0, 256, size=s_total, dtype=np.uint8)
    w = 800
    wu, ws, we = make_windows(u, segs, ts, rng, w)
    esg, esm, ecnt = brute.usage_reduce(off, ts, g8.astype(np.int64), m8.astype(np.int64), wu, ws, we)
    eag = np.array([brute.floor_avg(a, c) for a, c in zip(esg, ecnt)], np.int64)
    bad = []
    neg = rng.random() < 0.3                                       # int64 columns may hold negative values
    gi = g8.astype(np.int64) - (300 if neg else 0)
    wide = th.usage_reduce(off, ts, gi, m8.astype(np.int64), wu, ws, we)
    wsg = esg - (300 * ecnt if neg else 0)
    if not (np.array_equal(wide[2], ecnt) and np.array_equal(wide[0], wsg) and np.array_equal(wide[1], esm)):
        bad.append('int64 sums')
    if not np.array_equal(wide[3], np.array([brute.floor_avg(a, c) for a, c in zip(wsg, ecnt)], np.int64)):
        bad.append('int64 floor averages')
    narrow = th.usage_reduce_u8(off, ts, g8, m8, wu, ws, we)
    if not (np.array_equal(narrow[2], ecnt) and np.array_equal(narrow[0], esg) and np.array_equal(narrow[1], esm)
            and np.array_equal(narrow[3], eag)):
        bad.append('uint8')
    # the ring: samples arrive in time order over all uuids, in batches of random size
    lib = native.load()
    h = C.c_void_p(0)
    native.check(lib.th_ring_create(0, int(rng.choice([1, 16, 4096])), C.byref(h)))
    if s_total:
        uu_all = np.repeat(np.arange(u, dtype=np.int64), np.diff(off))
        order = np.argsort(ts, kind='stable')
        pos = 0
        while pos < s_total:
            k = int(rng.integers(1, max(2, s_total // 3)))
            sl = order[pos:pos + k]
            a_u, a_t = np.ascontiguousarray(uu_all[sl]), np.ascontiguousarray(ts[sl])
            a_g, a_m = np.ascontiguousarray(g8[sl]), np.ascontiguousarray(m8[sl])
            native.check(lib.th_ring_append(h, len(sl), a_u.ctypes.data, a_t.ctypes.data, a_g.ctypes.data, a_m.ctypes.data, None))
            pos += k
    outs = [np.zeros(w, np.int64) for _ in range(5)]
    # uuids beyond the last one that ever logged a sample are unknown to the ring: same answer (no samples)
    native.check(lib.th_ring_reduce(h, w, wu.ctypes.data, ws.ctypes.data, we.ctypes.data, *[o.ctypes.data for o in outs], 0, None))
    if not (np.array_equal(outs[2], ecnt) and np.array_equal(outs[0], esg) and np.array_equal(outs[1], esm)
            and np.array_equal(outs[3], eag)):
        bad.append('ring')
    lib.th_ring_destroy(h)
    if bad and verbose:
        print('seed', seed, 'uuids', u, 'samples', s_total, '->', bad)
    return len(bad)


def run(n_logs=40, first=0, verbose=True):
    return sum(check_log(s, verbose) for s in range(first, first + n_logs))


if __name__ == '__main__':
    k = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    f0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    n_bad = run(k, f0)
    print('logs', k, 'from seed', f0, 'mismatching checks:', n_bad)
    sys.exit(1 if n_bad else 0)
