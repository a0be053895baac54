"""Pins the oracle (oracle/fricc_oracle.c): against the golden vectors recorded from the compiled reference, against
the compiled reference itself where oracle/_ref exists, and against numpy.einsum for the contraction formulas."""
import ctypes

import numpy as np
import pytest
import os

from conftest import GOLDEN, TERMS, create_sparse_ndarray, synth_t2


def test_glibc_rand_restatement(port):
    libc = ctypes.CDLL(None)
    for seed in (1, 20, 99999):
        libc.srand(seed)
        want = np.array([libc.rand() for _ in range(1500)], dtype=np.int32)
        assert (port.glibc_rand(seed, 1500) == want).all()
        assert (port.glibc_rand(seed, 500, skip=1000) == want[1000:]).all()
    assert port.glibc_rand(1, 1)[0] == 1804289383  # unseeded libc rand()


def test_port_matches_golden_compression(port, golden_compression):
    g = golden_compression
    for name in g["names"]:
        x, m, draws = g[f"{name}/x"], int(g[f"{name}/m"]), g[f"{name}/draws"]
        assert port.one_norm(x) == float(g[f"{name}/one_norm"])
        # argsort: the canonical tie order coincides with the reference wherever magnitudes are distinct
        if len(np.unique(np.abs(x))) == len(x):
            assert (port.argsort(x, m) == g[f"{name}/argsort"]).all()
        D, r = port.get_d_largest(x, m)
        assert (D == g[f"{name}/D"]).all() and r == float(g[f"{name}/remaining_norm"])
        p = port.make_probability_vector(x, m - len(D), D, r)
        assert np.array_equal(p, g[f"{name}/p"])
        for method in ("pivotal", "systematic"):
            S, used = getattr(port, f"sample_{method}")(m - len(D), p, draws)
            assert (S == g[f"{name}/{method}/S"]).all()
            assert used == int(g[f"{name}/{method}/consumed"])
            c = port.fri_compression(x, m, method, draws)
            assert (c["idx"] == g[f"{name}/{method}/idx"]).all()
            assert np.array_equal(c["vals"], g[f"{name}/{method}/vals"])


def test_port_matches_golden_contractions(port, golden_contractions):
    g = golden_contractions
    for key in g["names"]:
        t2 = g[f"{key}/t2"]
        flat, vals = g[f"{key}/flat"], g[f"{key}/vals"]
        for term in TERMS:
            W, o0, o1 = g[f"{key}/{term}/W"], g[f"{key}/{term}/out0"], g[f"{key}/{term}/out1"]
            out = o0.copy()
            port.contract(term, W, flat, vals, t2.shape, out)
            assert np.array_equal(out, o1), term


@pytest.mark.parametrize("no,nv,frac", [(4, 8, 0.1), (5, 10, 0.01), (10, 20, 0.001)])
def test_port_contractions_match_einsum(port, no, nv, frac):
    """The reference's own test for every term (fricc/tests/test_sparse_tensors.py:48-269), run on the restatement."""
    rng = np.random.RandomState(20)
    ext = {"o": no, "v": nv}
    shape = (no, no, nv, nv)
    t2 = create_sparse_ndarray(shape, int(np.prod(shape) * frac), rng)
    m = int(t2.size * frac)
    flat, vals = port.largest(t2, m)
    for term, (wax, oax, scale, es) in TERMS.items():
        W = rng.rand(*[ext[c] for c in wax])
        want = scale * np.einsum(es, W, t2)
        out = np.zeros([ext[c] for c in oax])
        port.contract(term, W, flat, vals, shape, out)
        np.testing.assert_almost_equal(out, want)


def test_port_matches_compiled_reference_random(port, ref):
    rng = np.random.default_rng(7)
    for trial in range(24):
        n = int(rng.integers(50, 6000))
        m = int(rng.integers(1, max(2, n // 2)))
        x = synth_t2(rng, n, ("heavy", "heavier", "uniform", "diff")[trial % 4])
        draws = rng.integers(0, 2**31 - 1, size=n + 5).astype(np.int32)
        assert port.one_norm(x) == ref.one_norm(x)
        assert (port.argsort(x, m) == ref.argsort(x, m)).all()
        for method in ("pivotal", "systematic"):
            a, b = port.fri_compression(x, m, method, draws), ref.fri_compression(x, m, method, draws)
            assert (a["idx"] == b["idx"]).all() and np.array_equal(a["vals"], b["vals"])
            assert a["consumed"] == b["consumed"]


def test_systematic_fallthrough_and_pivotal_tail(port, ref):
    """Edge cases of the samplers: a uk beyond the last interval leaves S[i] = 0 (fri_utils.cpp:147-157); the pivotal
    tail rule pushes the current candidate when the loop ends one sample short (:106-109)."""
    p = np.array([0.25, 0.5, 0.25, 0.5, 0.4999])       # sums to 1.9999
    draws = np.array([2**31 - 2] * 8, dtype=np.int32)   # rn ~ 1: uk = [~1, ~2): the second lands past the end
    S, _ = port.sample_systematic(2, p, draws)
    S2, _ = ref.sample_systematic(2, p, draws)
    assert (S == S2).all() and S[1] == 0
    p = np.array([0.25, 0.5, 0.25, 0.5, 0.4999999])    # sums to 2 - 1e-7: one crossing, then the tail rule
    draws = np.array([5, 2**30, 7, 2**29, 11, 13], dtype=np.int32)
    S, used = port.sample_pivotal(2, p, draws)
    S2, used2 = ref.sample_pivotal(2, p, draws)
    assert (S == S2).all() and used == used2


def _golden_parallel_cases():
    g = np.load(os.path.join(GOLDEN, "parallel_sample.npz"))
    for name in g["names"]:
        name = str(name)
        ns, T = (int(v) for v in g[name + "/meta"])
        q = g[name + "/q"]
        d = g[name + "/draws"]
        draws = d if T == 1 else np.full(q.size + 8, int(d[0]), dtype=np.int32)
        yield name, q, ns, T, draws, g[name + "/S"]


def test_port_parallel_sample_matches_golden(port):
    """parallel_sample (fri_utils.cpp:176-362): the restatement (draws consumed in chunk order) against outputs of the
    compiled reference recorded by tests/golden/make_golden_parallel_sample.py -- one thread with random draws, and
    2/3/8/16 threads on inputs where the reference's rand() race and reduction order cannot matter."""
    n = 0
    for name, q, ns, T, draws, want in _golden_parallel_cases():
        S, used = port.parallel_sample(ns, q, T, draws)
        assert np.array_equal(S, want), name
        assert len(np.unique(S)) == ns and S.min() >= 0 and S.max() < q.size     # a sample without replacement
        n += 1
    assert n >= 50


def test_port_parallel_sample_matches_compiled_reference(port, ref):
    rng = np.random.default_rng(5)
    compared = 0
    for trial in range(12):          # one thread: any probabilities, random draws, draw count included
        n = int(rng.integers(40, 5000))
        ns = int(rng.integers(1, max(2, n // 6)))
        x = np.abs(rng.standard_normal(n)) * np.exp(-3 * rng.random(n))
        p = x / x.sum() * ns
        while p.max() > 1:
            p = np.minimum(p, 1.0)
            p = p / p.sum() * ns
        draws = rng.integers(0, 2**31 - 1, size=n + 8).astype(np.int32)
        # the restatement goes FIRST: where it reports a sample-count error the reference calls exit() (with one chunk,
        # n_sample * |q| / |q| can round to just below n_sample, which leaves one sample for a one-element pivotal draw
        # that cannot deliver it, fri_utils.cpp:121-124) -- and would take this test process with it
        try:
            a, ua = port.parallel_sample(ns, p, 1, draws)
        except Exception as e:
            assert getattr(e, "code", None) == -5, e
            continue
        compared += 1
        b, ub = ref.parallel_sample(ns, p, 1, draws)
        assert np.array_equal(a, b) and ua == ub
    assert compared >= 6
