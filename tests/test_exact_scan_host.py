"""CPU check of the exact-scan algebra (fricc_b200/csrc/fricc_exact.cuh): the map monoid + binade restarts reproduce
a left-to-right fp64 running sum bit for bit, including round-half-even ties and binade crossings."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "exact_scan_host.cpp")
SO = os.path.join(HERE, "host", "libexact_scan_host.so")


@pytest.fixture(scope="module")
def H():
    if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(SRC):
        subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", SRC, "-o", SO], check=True)
    L = ctypes.CDLL(SO)
    L.exact_running_sum.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                    ctypes.POINTER(ctypes.c_int64)]
    L.block_map_apply.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_double, ctypes.POINTER(ctypes.c_double),
                                  ctypes.POINTER(ctypes.c_int)]
    return L


def run(H, p, chunk=256, group=4):
    p = np.ascontiguousarray(p, dtype=np.float64)
    out = np.empty_like(p)
    nc = ctypes.c_int64(0)
    H.exact_running_sum(p.ctypes.data, p.size, chunk, group, out.ctypes.data, ctypes.byref(nc))
    return out, nc.value


def cases():
    rng = np.random.default_rng(123)
    yield "uniform", rng.random(20000)
    yield "probs", rng.random(50000) * 0.08                      # like p = |x|/r*n at density 4%
    yield "with_zeros", np.where(rng.random(30000) < 0.3, 0.0, rng.random(30000))
    yield "leading_zeros", np.concatenate([np.zeros(700), rng.random(3000)])
    yield "wide_range", rng.random(40000) * 10.0 ** rng.uniform(-20, 0, 40000)
    yield "tiny_then_big", np.concatenate([rng.random(1000) * 1e-300, rng.random(1000) * 1e-5, rng.random(1000)])
    yield "subnormal", np.concatenate([np.full(50, 5e-324), rng.random(500) * 1e-310, rng.random(500)])
    # adversarial ties: addends that are exact odd multiples of half an ulp of the running sum
    yield "ties_small_ints", rng.integers(0, 4, 20000).astype(np.float64) * 2.0 ** -53 + 0.0
    yield "ties_after_one", np.concatenate([[1.0], rng.integers(0, 8, 30000) * 2.0 ** -53])
    yield "ties_dyadic", np.concatenate([[0.75], rng.integers(1, 64, 30000) * 2.0 ** -54])
    yield "ones", np.ones(5000)
    yield "powers", 2.0 ** rng.integers(-60, 3, 20000).astype(np.float64)
    yield "all_zero", np.zeros(1000)
    yield "single", np.array([0.3])


@pytest.mark.parametrize("name,p", list(cases()), ids=[c[0] for c in cases()])
@pytest.mark.parametrize("chunk,group", [(256, 4), (64, 1), (1024, 8), (7, 3)])
def test_exact_running_sum(H, name, p, chunk, group):
    want = np.cumsum(p)  # numpy's cumsum is the plain left-to-right recurrence (= std::partial_sum)
    got, ncross = run(H, p, chunk, group)
    assert np.array_equal(got, want), name
    # sanity: the number of restarts is O(#binades), not O(n)
    assert ncross <= 2200


def test_block_map(H):
    """A whole block's composed map applied to an exact start equals the sequential sum whenever it reports valid."""
    rng = np.random.default_rng(5)
    nvalid = 0
    for _ in range(300):
        c = float(rng.random() * 10.0 ** rng.uniform(-3, 6))
        p = rng.random(int(rng.integers(1, 400))) * 10.0 ** rng.uniform(-8, 0)
        if rng.random() < 0.3:
            p = np.round(p * 2.0 ** 60) * 2.0 ** -60 * 2.0 ** int(np.floor(np.log2(c)) - 45)  # provoke ties
        out, valid = ctypes.c_double(0), ctypes.c_int(0)
        H.block_map_apply(p.ctypes.data, p.size, c, ctypes.byref(out), ctypes.byref(valid))
        seq = c
        for v in p:
            seq = seq + v
        inside = np.floor(np.log2(seq)) == np.floor(np.log2(c)) and seq < 2.0 ** (np.floor(np.log2(c)) + 1)
        if valid.value:
            nvalid += 1
            assert out.value == seq
        else:
            assert not inside
    assert nvalid > 50
