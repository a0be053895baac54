"""The argument behind the register-resident emission kernel (fricc_escan.cu: escan_emit_kernel),
replayed on the CPU in IEEE doubles: locating samples PER SAMPLE inside a thread's run of consecutive elements
(sample i lands on the first element whose inclusive running sum exceeds uk_i = fl(i + rn)) gives exactly the
assignment of the reference's per-element interval test  intervals[j] <= uk_i < intervals[j+1]
(/root/reference/src/fri_utils.cpp:147-157), including elements that own no sample, elements whose interval is empty
(inclusive == exclusive sum) and running sums that sit exactly on sample points."""
import math

import numpy as np


def first_sample_at_or_above(X, rn, nsamp):
    """smallest i in [0, nsamp] with fl(i + rn) >= X  (fricc_psrc.cuh)"""
    g = X - rn
    if not (g > 0.0):
        i = 0
    elif g >= float(nsamp):
        i = nsamp
    else:
        i = int(math.ceil(g))
    while i > 0 and float(i - 1) + rn >= X:
        i -= 1
    while i < nsamp and float(i) + rn < X:
        i += 1
    return i


def per_element(excl0, incl, rn, nsamp):
    """the general path / the reference: element q owns samples [first(excl_q), first(incl_q)) if its interval is non-empty"""
    owner = {}
    excl = excl0
    for q, inc in enumerate(incl):
        if inc > excl:
            for i in range(first_sample_at_or_above(excl, rn, nsamp), first_sample_at_or_above(inc, rn, nsamp)):
                assert i not in owner
                owner[i] = q
        excl = inc
    return owner


def per_sample(excl0, incl, rn, nsamp):
    """the register path: the run covers [excl0, incl[-1]); each sample inside goes to the first element whose inclusive
    sum exceeds uk_i"""
    owner = {}
    if incl[-1] > excl0:
        for i in range(first_sample_at_or_above(excl0, rn, nsamp), first_sample_at_or_above(incl[-1], rn, nsamp)):
            uk = float(i) + rn
            owner[i] = sum(1 for t in incl[:-1] if t <= uk)
    return owner


def test_per_sample_location_equals_the_interval_test():
    rng = np.random.default_rng(2024)
    nsamp = 50
    for case in range(4000):
        rn = [0.0, 1.0, 0.5, float(rng.random()), float(np.nextafter(1.0, 0.0))][case % 5]
        start = float(rng.random() * 40.0)
        kind = case % 4
        if kind == 0:       # sparse: most elements own nothing
            steps = rng.random(8) * 0.08
        elif kind == 1:     # dense: several samples per element
            steps = rng.random(8) * 3.0
        elif kind == 2:     # empty intervals and tiny increments mixed in
            steps = np.where(rng.random(8) < 0.4, 0.0, rng.random(8) * 0.7)
        else:               # running sums exactly on sample points fl(i + rn)
            steps = None
        if steps is None:
            pts = sorted(float(i) + rn for i in rng.integers(0, nsamp, size=8))
            excl0, incl = pts[0], pts
        else:
            excl0 = start
            incl = list(np.add.accumulate(np.concatenate(([start], steps)))[1:])
            incl = [float(x) for x in incl]
        a, b = per_element(excl0, incl, rn, nsamp), per_sample(excl0, incl, rn, nsamp)
        assert a == b, (case, rn, excl0, incl)


def test_first_sample_is_the_threshold_of_a_monotone_map():
    """i -> fl(i + rn) is non-decreasing, and first_sample_at_or_above is its generalised inverse -- what both paths rest on."""
    rng = np.random.default_rng(7)
    nsamp = 200
    for _ in range(2000):
        rn = float(rng.random())
        X = float(rng.random() * (nsamp + 2) - 1.0)
        i = first_sample_at_or_above(X, rn, nsamp)
        assert 0 <= i <= nsamp
        assert i == nsamp or float(i) + rn >= X
        assert i == 0 or float(i - 1) + rn < X
