"""rtc_portable_math.h (the transcendentals both the oracle's PORTABLE mode and the GPU's exact kernel use) against
double precision: <= 2 ulp on the ranges the render path feeds them."""
from __future__ import annotations

import math

import numpy as np
import pytest

SIN, COS, LOG, POW5, ATAN2, ACOS = range(6)


def ulps(got, want):
    want32 = np.float32(want)
    spacing = np.spacing(np.abs(want32)) if want32 != 0 else np.float32(1e-45)
    return abs(float(got) - float(want)) / float(spacing)


@pytest.mark.parametrize("which,fn,lo,hi", [
    (SIN, math.sin, 0.0, 2 * math.pi), (COS, math.cos, 0.0, 2 * math.pi), (SIN, math.sin, -50.0, 50.0),
    (LOG, math.log, 1e-7, 1.0), (ACOS, math.acos, -1.0, 1.0), (POW5, lambda x: x ** 5, 0.0, 1.0),
])
def test_one_argument_functions(oracle, which, fn, lo, hi):
    rs = np.random.RandomState(which + 1)
    xs = rs.uniform(lo, hi, 4000).astype(np.float32)
    worst = 0.0
    for x in xs:
        got = oracle.oracle().oracle_pm(which, float(x), 0.0)
        want = fn(float(x))
        if which in (SIN, COS) and abs(want) < 1e-3:
            assert abs(got - want) < 2e-7  # absolute near the zeros, as for any reduced-argument sinf/cosf
            continue
        worst = max(worst, ulps(got, want))
    # x^5 by three roundings of repeated multiplication can be 3 ulp from the exact power; the rest stay within 2
    assert worst <= (3.0 if which == POW5 else 2.0), worst


def test_atan2(oracle):
    rs = np.random.RandomState(7)
    pts = rs.normal(size=(4000, 2)).astype(np.float32)
    worst = 0.0
    for y, x in pts:
        got = oracle.oracle().oracle_pm(ATAN2, float(x), float(y))
        want = math.atan2(float(y), float(x))
        if abs(want) < 1e-3:
            assert abs(got - want) < 2e-7
            continue
        worst = max(worst, ulps(got, want))
    assert worst <= 2.5, worst


def test_special_values(oracle):
    pm = oracle.oracle().oracle_pm
    assert pm(LOG, 0.0, 0.0) == -math.inf
    assert math.isnan(pm(LOG, -1.0, 0.0))
    assert pm(LOG, 1.0, 0.0) == 0.0
    assert pm(ACOS, 1.0, 0.0) == 0.0
    assert abs(pm(ACOS, -1.0, 0.0) - math.pi) < 1e-6
    assert math.isnan(pm(ACOS, 1.5, 0.0))
    assert pm(ATAN2, 1.0, 0.0) == 0.0
    assert abs(pm(ATAN2, -1.0, 0.0) - math.pi) < 1e-6
    assert pm(SIN, 0.0, 0.0) == 0.0 and pm(COS, 0.0, 0.0) == 1.0
