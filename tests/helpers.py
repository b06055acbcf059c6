"""Shared test helpers: KAT loading and tiny conversions."""
import datetime as dt
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
EPOCH = dt.datetime(1970, 1, 1)


def load_kat():
    with open(os.path.join(HERE, 'golden', 'reference_kat.json')) as f:
        data = json.load(f)
    now = dt.datetime.strptime(data['now'], '%Y-%m-%dT%H:%M:%S.%f')
    d = now - EPOCH
    now_us = (d.days * 86400 + d.seconds) * 1_000_000 + d.microseconds
    return now_us, data['cases']


def kat_rows(case, now_us):
    """rows of a KAT case as (id, resource_id, start_us, end_us, user_id, is_cancelled)."""
    return [(r['id'], r['uuid'], now_us + r['start'] * 1_000_000, now_us + r['end'] * 1_000_000, r['user'],
             r['cancelled']) for r in case['rows']]


def columns_from_rows(rows):
    """-> (cal dict of numpy columns with dictionary-encoded uuids, uuid -> code map)."""
    uu = {}
    for r in rows:
        uu.setdefault(r[1], len(uu))
    cal = {
        'id': np.array([r[0] for r in rows], np.int64),
        'uuid_id': np.array([uu[r[1]] for r in rows], np.int64),
        'start': np.array([r[2] for r in rows], np.int64),
        'end': np.array([r[3] for r in rows], np.int64),
        'owner': np.array([r[4] for r in rows], np.int64),
        'cancelled': np.array([1 if (r[5] is not None and r[5]) else 0 for r in rows], np.uint8),
        'n_uuids': max(len(uu), 1),
    }
    return cal, uu


def csr_to_lists(offsets, ids):
    offsets = np.asarray(offsets)
    ids = np.asarray(ids)
    return [[int(x) for x in ids[offsets[i]:offsets[i + 1]]] for i in range(len(offsets) - 1)]
