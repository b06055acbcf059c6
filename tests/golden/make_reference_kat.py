"""Writes tests/golden/reference_kat.json: the reference's own tests for the hot path, transcribed as
known-answer vectors.  Run from the repo root:  python tests/golden/make_reference_kat.py

The reference tests build their rows relative to ``datetime.utcnow()``; here NOW is fixed so the vectors
are reproducible.  Offsets are seconds relative to NOW.  ``expect`` comes from what the reference test
ASSERTS (``pytest.raises(AssertionError)`` on save == would_interfere() is True; a save that succeeds ==
False; membership asserts for current_events) -- never from running this repo's oracle.  Cases marked
``"pinned": false`` have no reference test; their expectation is read off the source text.
"""
import json
import os

NOW = '2026-03-01T12:00:00.000000'
GPU = 'GPU-d38d4de3-85ee-e837-3d87-e8e2faeb6a63'      # tests/fixtures/models.py:133
GPU2 = 'GPU-d38d4de3-85ee-e837-3d87-e8e2faeb6a64'     # tests/fixtures/models.py:141
M = 60
H = 3600

cases = []

# tests/unit/models/test_reservation_model.py:10-37 -- the four overlap shapes all raise
base = {'id': 1, 'uuid': GPU, 'start': 0, 'end': 60 * M, 'user': 1, 'cancelled': None}
shapes = {
    'inside   | A    A |': (+5 * M, 60 * M - 5 * M),
    'left     A |     A |': (-5 * M, 60 * M - 5 * M),
    'right    | A     | A': (+5 * M, 60 * M + 5 * M),
    'enclose  A |      | A': (-5 * M, 60 * M + 5 * M),
}
for name, (s, e) in shapes.items():
    cases.append({'name': 'interfering/' + name, 'source': 'tests/unit/models/test_reservation_model.py:10-37',
                  'pinned': True, 'rows': [base], 'op': 'would_interfere',
                  'query': {'uuid': GPU, 'start': s, 'end': e, 'self_id': None}, 'expect': True})

# :40-64 -- same shapes against a CANCELLED base row never raise; the candidate is saved (id 2) after the
# first shape and then moved, so from the second shape on it excludes itself by id.
cbase = dict(base, cancelled=True)
prev = None
for k, (name, (s, e)) in enumerate(shapes.items()):
    rows = [cbase] if prev is None else [cbase, {'id': 2, 'uuid': GPU, 'start': prev[0], 'end': prev[1], 'user': 3,
                                                'cancelled': None}]
    cases.append({'name': 'cancelled-base/' + name, 'source': 'tests/unit/models/test_reservation_model.py:40-64',
                  'pinned': True, 'rows': rows, 'op': 'would_interfere',
                  'query': {'uuid': GPU, 'start': s, 'end': e, 'self_id': None if prev is None else 2},
                  'expect': False})
    prev = (s, e)

# :113-125 -- current_events returns only non-cancelled rows
r1 = {'id': 1, 'uuid': GPU, 'start': -10 * M, 'end': 60 * M, 'user': 1, 'cancelled': None}
cases.append({'name': 'current/active row is listed', 'source': 'tests/unit/models/test_reservation_model.py:113-117',
              'pinned': True, 'rows': [r1], 'op': 'current_events', 'query': {'uuid': None, 'now': 1},
              'expect_contains': [1], 'expect_excludes': []})
r1c = dict(r1, cancelled=True)
r2 = {'id': 2, 'uuid': GPU, 'start': 0, 'end': 60 * M, 'user': 3, 'cancelled': None}
cases.append({'name': 'current/cancelled row is dropped', 'source': 'tests/unit/models/test_reservation_model.py:119-125',
              'pinned': True, 'rows': [r1c, r2], 'op': 'current_events', 'query': {'uuid': None, 'now': 1},
              'expect_contains': [2], 'expect_excludes': [1]})
# the save of r2 at :121 succeeds although it overlaps the cancelled r1
cases.append({'name': 'current/second row may be saved over cancelled one',
              'source': 'tests/unit/models/test_reservation_model.py:119-121', 'pinned': True, 'rows': [r1c],
              'op': 'would_interfere', 'query': {'uuid': GPU, 'start': 0, 'end': 60 * M, 'self_id': None},
              'expect': False})

# tests/unit/models/test_user_model.py:48-57 pins the tri-state flag: an unset (NULL) flag is "not cancelled"
cases.append({'name': 'flag/NULL is not cancelled', 'source': 'tests/unit/models/test_user_model.py:43-45',
              'pinned': True, 'rows': [base], 'op': 'current_events', 'query': {'uuid': GPU, 'now': 30 * M},
              'expect_contains': [1], 'expect_excludes': []})

# fixtures past / active / future (tests/fixtures/models.py:70-115): classes for R3 / R4.  No reference test
# calls current_events / upcoming_events_for_resource on them: expectations follow Reservation.py:96-97,112-117.
past = {'id': 1, 'uuid': GPU, 'start': -5 * H, 'end': -4 * H, 'user': 1, 'cancelled': None}
active = {'id': 2, 'uuid': GPU, 'start': -5 * H, 'end': 5 * H, 'user': 1, 'cancelled': None}
future = {'id': 3, 'uuid': GPU, 'start': 5 * H, 'end': 15 * H, 'user': 1, 'cancelled': None}
other = {'id': 4, 'uuid': GPU2, 'start': -1 * H, 'end': 1 * H, 'user': 2, 'cancelled': None}
trio = [past, active, future, other]
cases.append({'name': 'classes/current_events(gpu)', 'source': 'tests/fixtures/models.py:70-115 + Reservation.py:96-100',
              'pinned': False, 'rows': trio, 'op': 'current_events', 'query': {'uuid': GPU, 'now': 0}, 'expect': [2]})
cases.append({'name': 'classes/current_events()', 'source': 'tests/fixtures/models.py:70-115 + Reservation.py:96-97',
              'pinned': False, 'rows': trio, 'op': 'current_events', 'query': {'uuid': None, 'now': 0}, 'expect': [2, 4]})
cases.append({'name': 'classes/upcoming 30 min', 'source': 'Reservation.py:106-118', 'pinned': False, 'rows': trio,
              'op': 'upcoming', 'query': {'uuid': GPU, 'now': 0, 'period': 30 * M}, 'expect': [2]})
cases.append({'name': 'classes/upcoming 6 h', 'source': 'Reservation.py:106-118', 'pinned': False, 'rows': trio,
              'op': 'upcoming', 'query': {'uuid': GPU, 'now': 0, 'period': 6 * H}, 'expect': [2, 3]})
cases.append({'name': 'classes/upcoming exactly reaches start', 'source': 'Reservation.py:114-115 (start <= now + period)',
              'pinned': False, 'rows': trio, 'op': 'upcoming', 'query': {'uuid': GPU, 'now': 0, 'period': 5 * H},
              'expect': [2, 3]})
cases.append({'name': 'classes/range filter is inclusive and keeps cancelled',
              'source': 'Reservation.py:132-148', 'pinned': False,
              'rows': [past, dict(active, cancelled=True), future, other], 'op': 'range',
              'query': {'uuids': [GPU], 'start': -4 * H, 'end': 5 * H}, 'expect': [1, 2, 3]})
# touching boundaries: strict for would_interfere (Reservation.py:124-125)
cases.append({'name': 'touching/starts at end', 'source': 'Reservation.py:124', 'pinned': False, 'rows': [base],
              'op': 'would_interfere', 'query': {'uuid': GPU, 'start': 60 * M, 'end': 120 * M, 'self_id': None},
              'expect': False})
cases.append({'name': 'touching/ends at start', 'source': 'Reservation.py:125', 'pinned': False, 'rows': [base],
              'op': 'would_interfere', 'query': {'uuid': GPU, 'start': -60 * M, 'end': 0, 'self_id': None},
              'expect': False})
cases.append({'name': 'other gpu never interferes', 'source': 'Reservation.py:129', 'pinned': False, 'rows': [base],
              'op': 'would_interfere', 'query': {'uuid': GPU2, 'start': 5 * M, 'end': 50 * M, 'self_id': None},
              'expect': False})

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'reference_kat.json')
with open(out, 'w') as f:
    json.dump({'now': NOW, 'cases': cases}, f, indent=1)
print('wrote %d cases to %s' % (len(cases), out))
