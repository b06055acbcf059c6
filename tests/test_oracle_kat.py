"""The oracle against the reference's own known-answer tests (tests/golden/reference_kat.json)."""
import datetime as dt

import numpy as np
import pytest

from helpers import load_kat, kat_rows, columns_from_rows, EPOCH
from oracle import brute
from oracle.sql_ref import SqlCalendar

NOW_US, CASES = load_kat()


def _sec(now_us, s):
    return now_us + s * 1_000_000


@pytest.mark.parametrize('case', CASES, ids=[c['name'] for c in CASES])
def test_sql_oracle_matches_reference_kat(case):
    cal = SqlCalendar(kat_rows(case, NOW_US))
    q = case['query']
    if case['op'] == 'would_interfere':
        got = cal.would_interfere(q['uuid'], _sec(NOW_US, q['start']), _sec(NOW_US, q['end']), q['self_id'])
        assert got is case['expect']
    elif case['op'] == 'current_events':
        got = cal.current_events(_sec(NOW_US, q['now']), q['uuid'])
        _check_list(case, got)
    elif case['op'] == 'upcoming':
        got = cal.upcoming_events_for_resource(q['uuid'], _sec(NOW_US, q['now']), q['period'] * 1_000_000)
        assert got == case['expect']
    elif case['op'] == 'range':
        s = EPOCH + dt.timedelta(microseconds=_sec(NOW_US, q['start']))
        e = EPOCH + dt.timedelta(microseconds=_sec(NOW_US, q['end']))
        assert sorted(cal.filter_by_uuids_and_time_range(q['uuids'], s, e)) == case['expect']
    else:
        raise AssertionError(case['op'])


def _check_list(case, got):
    if 'expect' in case:
        assert got == case['expect']
    for i in case.get('expect_contains', []):
        assert i in got
    for i in case.get('expect_excludes', []):
        assert i not in got


@pytest.mark.parametrize('case', CASES, ids=[c['name'] for c in CASES])
def test_brute_oracle_matches_reference_kat(case):
    rows = kat_rows(case, NOW_US)
    cal, uu = columns_from_rows(rows)
    q = case['query']

    def code(u):
        return uu.get(u, len(uu) + 7)

    if case['op'] == 'would_interfere':
        sid = -1 if q['self_id'] is None else q['self_id']
        got = brute.conflict_bits(cal, [code(q['uuid'])], [_sec(NOW_US, q['start'])], [_sec(NOW_US, q['end'])], [sid])
        assert bool(got[0]) is case['expect']
    elif case['op'] == 'current_events':
        now = _sec(NOW_US, q['now'])
        codes = list(uu.values()) if q['uuid'] is None else [code(q['uuid'])]
        k = len(codes)
        sets = brute.conflict_sets(cal, codes, [now] * k, [now] * k, [-1] * k, brute.INCLUSIVE)
        canc = {int(i): int(c) for i, c in zip(cal['id'], cal['cancelled'])}
        got = sorted(i for s in sets for i in s if not canc[i])
        _check_list(case, got)
    elif case['op'] == 'upcoming':
        now = _sec(NOW_US, q['now'])
        sets = brute.conflict_sets(cal, [code(q['uuid'])], [now], [now + q['period'] * 1_000_000], [-1], brute.UPCOMING)
        assert sets[0] == case['expect']
    elif case['op'] == 'range':
        codes = [code(u) for u in q['uuids']]
        k = len(codes)
        sets = brute.conflict_sets(cal, codes, [_sec(NOW_US, q['start'])] * k, [_sec(NOW_US, q['end'])] * k, [-1] * k,
                                   brute.INCLUSIVE)
        assert sorted(i for s in sets for i in s) == case['expect']


def test_non_datetime_range_arguments_raise():
    # Reservation.py:140-142 -> AssertionError (HTTP 400 at controllers/reservation.py:43-45)
    cal = SqlCalendar()
    with pytest.raises(AssertionError):
        cal.filter_by_uuids_and_time_range(['x'], '2026-01-01', dt.datetime(2026, 1, 2))


def test_floor_average():
    # UsageLoggingService.py:25-30
    from oracle.services_ref import avg
    assert avg([1, 2, 4]) == 2
    assert avg([]) == -1.0
    assert isinstance(avg([]), float)
    assert brute.floor_avg(7, 3) == 2 and brute.floor_avg(0, 0) == -1
