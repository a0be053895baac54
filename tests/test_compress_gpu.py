"""GPU parity of the compression stages (P1-P5 of SURVEY.md 8c): bit-exact against the oracle and against the golden
vectors recorded from the compiled reference, with the same injected uniform draws."""
import numpy as np
import pytest
from conftest import synth_t2

pytestmark = pytest.mark.gpu


def test_golden_stages(rccsd, golden_compression):
    g = golden_compression
    for name in g["names"]:
        x, m, draws = g[f"{name}/x"], int(g[f"{name}/m"]), g[f"{name}/draws"]
        assert rccsd.one_norm(x) == float(g[f"{name}/one_norm"]), name                      # P2 (norm)
        got = rccsd.argsort(x, m)
        want = g[f"{name}/argsort"]
        assert np.array_equal(np.abs(x[got]), np.abs(x[want])), name                         # P1
        if len(np.unique(np.abs(x))) == len(x):
            assert (got == want).all(), name
        D, r = rccsd.get_d_largest(x, m)
        assert (D == g[f"{name}/D"]).all() and r == float(g[f"{name}/remaining_norm"]), name  # P2
        p = rccsd.make_probability_vector(x, m - len(D), D, r)
        assert np.array_equal(p, g[f"{name}/p"]), name                                       # P3
        for method in ("pivotal", "systematic"):
            S = getattr(rccsd, f"sample_{method}")(m - len(D), p, draws=draws)
            assert (np.array(S, dtype=np.int64) == g[f"{name}/{method}/S"]).all(), (name, method)   # P4
            idx, vals = rccsd.fri_compression(x, m, method, draws=draws)
            assert (idx == g[f"{name}/{method}/idx"]).all(), (name, method)
            assert np.array_equal(vals, g[f"{name}/{method}/vals"]), (name, method)          # P5


@pytest.mark.parametrize("kind", ["heavy", "heavier", "uniform", "diff"])
@pytest.mark.parametrize("n,m", [(1, 1), (2, 1), (33, 33), (5000, 1), (20000, 3000), (300000, 40000)])
def test_fri_vs_oracle(rccsd, port, kind, n, m):
    rng = np.random.default_rng(n * 7 + m)
    x = synth_t2(rng, n, kind)
    draws = rng.integers(0, 2**31 - 1, size=n + 4).astype(np.int32)
    for method in ("systematic", "pivotal"):
        try:
            want = port.fri_compression(x, m, method, draws)
        except Exception as e:  # the reference would exit here; the GPU path must raise the matching error
            with pytest.raises(Exception):
                rccsd.fri_compression(x, m, method, draws=draws)
            continue
        idx, vals = rccsd.fri_compression(x, m, method, draws=draws)
        assert (idx == want["idx"]).all(), method
        assert np.array_equal(vals, want["vals"]), method


def test_sparse_tensor_fri_ctor_and_consumed_draws(rccsd, port):
    rng = np.random.default_rng(11)
    no, nv = 6, 14
    t2 = synth_t2(rng, no * no * nv * nv, "heavier").reshape(no, no, nv, nv)
    m = 700
    draws = rng.integers(0, 2**31 - 1, size=t2.size).astype(np.int32)
    for method in ("systematic", "pivotal"):
        T = rccsd.SparseTensor4d(t2, t2.shape, m, "fri", method, False, draws=draws)
        want = port.fri_compression(t2, m, method, draws)
        flat, vals = T.export()
        assert (flat == want["idx"]).all() and np.array_equal(vals, want["vals"])
        assert T.consumed_draws == want["consumed"]
        idx, v = T.get_element(m - 1)
        assert np.ravel_multi_index(tuple(idx), t2.shape) == want["idx"][-1] and v == want["vals"][-1]


def test_library_rand_stream_is_glibc(rccsd, port):
    """Without injected draws the module consumes glibc's rand() stream exactly like a stock reference build:
    srand(seed), then one draw per systematic compression / one per visited element for pivotal."""
    rng = np.random.default_rng(5)
    x = synth_t2(rng, 4000, "heavy")
    stream = port.glibc_rand(20, 3 * x.size)
    rccsd.srand(20)
    pos = 0
    for method in ("systematic", "pivotal", "systematic", "pivotal"):
        want = port.fri_compression(x, 300, method, stream[pos:])
        idx, vals = rccsd.fri_compression(x, 300, method)
        assert (idx == want["idx"]).all() and np.array_equal(vals, want["vals"])
        pos += want["consumed"]
        assert rccsd.rand_consumed() == pos


def test_largest_with_ties_and_zeros(rccsd, port):
    """The reference's contraction tests feed mostly-zero tensors with m larger than the number of distinct nonzeros,
    so the cut lands inside a huge group of exact ties (zeros); pair-symmetric amplitudes tie in pairs."""
    rng = np.random.default_rng(2)
    x = np.zeros(50000)
    nz = rng.choice(x.size, 400, replace=False)
    x[nz] = rng.random(400)
    x[rng.choice(nz, 50)] = 0.5  # exact ties among the kept values
    for m in (100, 400, 450, 5000):
        got = rccsd.argsort(x, m)
        want = port.argsort(x, m)
        assert (got == want).all()  # canonical order: |x| descending, index ascending
    # big tie group straddling the cut (more ties than the small-bin threshold is not needed to hit the path)
    y = np.full(3_000_000, 0.25)
    y[::1000] = 1.0
    got = rccsd.argsort(y, 3000 + 1234)
    assert (got[:3000] == np.arange(0, 3_000_000, 1000)).all()
    rest = got[3000:]
    assert (rest == np.array([i for i in range(2000) if i % 1000][:1234])).all()


def test_error_conditions(rccsd):
    """Conditions on which the reference prints to stderr and exits become exceptions."""
    from fricc_b200._lib import FriccError

    with pytest.raises(FriccError):  # p > 1: fri_utils.cpp:383-395
        rccsd.make_probability_vector(np.array([1.0, 1.0, 5.0]), 2, np.array([], dtype=np.int64), 3.0)
    with pytest.raises(FriccError):  # wrong number of samples: fri_utils.cpp:110-124
        rccsd.sample_pivotal(3, np.array([0.2, 0.2, 0.2, 0.2]), draws=np.arange(10, dtype=np.int32))
    with pytest.raises(ValueError):  # m > N: the Python driver's own check (fricc/friccsd.py:69-72)
        rccsd.SparseTensor4d(np.zeros((1, 1, 2, 2)), (1, 1, 2, 2), 5, "largest")
    with pytest.raises(ValueError):
        rccsd.SparseTensor4d(np.zeros((1, 1, 2, 2)), (1, 1, 2, 2), 2, "bogus")
    with pytest.raises(TypeError):  # F-ordered tensor_arr is rejected (array_t<double, c_style>, noconvert)
        rccsd.SparseTensor4d(np.asfortranarray(np.zeros((2, 2, 3, 3))), (2, 2, 3, 3), 2, "largest")
    t = rccsd.SparseTensor4d([2, 2, 2, 2], 3)
    assert t.get_element(0) == ([0, 0, 0, 0], 0.0)


@pytest.mark.parametrize("kind", ["heavy", "heavier", "uniform"])
def test_parallel_exact_scan_equals_chain(rccsd, kind):
    """The parallel binade-segmented scan (fricc_escan.cu) and the one-thread chain kernels are bit-identical at a size
    where a tree-ordered sum is known to differ from the sequential one."""
    rng = np.random.default_rng(77)
    n, m = 6_000_000, 400_000
    x = synth_t2(rng, n, kind)
    x[: 5000] = 0.0                      # leading zeros: the zero-state path
    draws = np.array([123456789], dtype=np.int32)
    try:
        rccsd.set_option("escan", 0)
        norm0 = rccsd.one_norm(x)
        idx0, val0 = rccsd.fri_compression(x, m, "systematic", draws=draws)
    finally:
        rccsd.set_option("escan", 1)
    norm1 = rccsd.one_norm(x)
    idx1, val1 = rccsd.fri_compression(x, m, "systematic", draws=draws)
    assert norm0 == norm1
    assert norm1 == float(np.cumsum(np.abs(x))[-1])   # numpy cumsum is the left-to-right recurrence
    assert (idx0 == idx1).all() and np.array_equal(val0, val1)


def test_exact_norm_adversarial(rccsd):
    """Ties (addends that are odd multiples of half an ulp of the running sum), wide dynamic range, subnormals."""
    rng = np.random.default_rng(3)
    cases = [
        np.concatenate([[1.0], rng.integers(0, 8, 300000) * 2.0 ** -53]),
        np.concatenate([[0.75], rng.integers(1, 64, 300000) * 2.0 ** -54]),
        rng.random(500000) * 10.0 ** rng.uniform(-20, 0, 500000),
        np.concatenate([np.full(50, 5e-324), rng.random(5000) * 1e-310, rng.random(100000)]),
        np.zeros(10000),
        2.0 ** rng.integers(-60, 3, 200000).astype(np.float64),
    ]
    for x in cases:
        assert rccsd.one_norm(x) == float(np.cumsum(np.abs(x))[-1])


@pytest.mark.parametrize("kind,n,m", [("heavier", 400000, 300000), ("uniform", 200000, 150000), ("heavy", 300000, 1000)])
def test_adaptive_d_candidate_list(rccsd, port, kind, n, m):
    """compress_fri orders only a candidate prefix when it looks for D (sized by a histogram estimate, grown x16 when the
    greedy test does not stop inside it).  Every sizing policy must give the oracle's result, including large |D|."""
    rng = np.random.default_rng(n + m)
    x = synth_t2(rng, n, kind)
    draws = np.array([987654321], dtype=np.int32)
    want = port.fri_compression(x, m, "systematic", draws)
    try:
        for cand, est in ((4, 0), (4, 1), (1 << 16, 1), (1 << 30, 0)):
            rccsd.set_option("d_candidates", cand)
            rccsd.set_option("d_estimate", est)
            idx, vals = rccsd.fri_compression(x, m, "systematic", draws=draws)
            assert (idx == want["idx"]).all(), (cand, est)
            assert np.array_equal(vals, want["vals"]), (cand, est)
    finally:
        rccsd.set_option("d_candidates", 1 << 16)
        rccsd.set_option("d_estimate", 1)


@pytest.mark.parametrize("kind", ["heavy", "heavier", "uniform", "diff"])
@pytest.mark.parametrize("n,m", [(1, 1), (2, 1), (33, 33), (5000, 1), (20000, 3000), (300000, 40000), (2_000_000, 150_000)])
def test_parallel_pivotal_equals_oracle(rccsd, port, kind, n, m):
    """The parallel exact pivotal sampler (fricc_pivotal.cu) forced on at every size: bit-identical sample lists, values
    and consumed-draw counts to the sequential walk of the oracle (and to the one-thread chain kernel)."""
    rng = np.random.default_rng(n * 13 + m)
    x = synth_t2(rng, n, kind)
    draws = rng.integers(0, 2**31 - 1, size=n + 4).astype(np.int32)
    try:
        want = port.fri_compression(x, m, "pivotal", draws)
    except Exception:
        want = None
    try:
        rccsd.set_option("pivotal_parallel", 1)
        if want is None:
            with pytest.raises(Exception):
                rccsd.fri_compression(x, m, "pivotal", draws=draws)
            return
        idx, vals = rccsd.fri_compression(x, m, "pivotal", draws=draws)
        restarts = rccsd.get_stat("pivotal_restarts")
    finally:
        rccsd.set_option("pivotal_parallel", -1)
    assert (idx == want["idx"]).all()
    assert np.array_equal(vals, want["vals"])
    assert restarts <= 2, restarts   # a restart costs time, never correctness; more than a couple means the model is off


def test_parallel_pivotal_probability_vectors(rccsd, port):
    """Stage-wise sampler on explicit probability vectors with adversarial structure: exact ties at the wrap step (all
    probabilities multiples of 2^-k), long zero stretches, probabilities equal to 1, tiny probabilities."""
    rng = np.random.default_rng(5)
    cases = []
    p = rng.integers(0, 2**20, 200000) / 2.0**22            # dyadic: many exact ties
    cases.append(p)
    p = rng.random(300000) * (rng.random(300000) < 0.05)    # 95 % zeros
    cases.append(p)
    p = rng.random(100000) ** 8
    p[::1000] = 1.0                                         # probabilities of exactly one
    cases.append(p)
    p = np.concatenate([np.full(70000, 2.0**-30), rng.random(70000), np.full(70000, 1e-300)])
    cases.append(p)
    p = rng.integers(1, 4, 150000) * 2.0**-53 + 0.25        # half-ulp offsets around a coarse value
    cases.append(p)
    try:
        rccsd.set_option("pivotal_parallel", 1)
        for k, p in enumerate(cases):
            ns = int(np.floor(p.sum()))
            p = p * (ns / p.sum())                          # sum ~ ns so that the walk ends with a ~ 1 or fills S
            p = np.minimum(p, 1.0)
            draws = rng.integers(0, 2**31 - 1, size=p.size).astype(np.int32)
            try:
                want, _ = port.sample_pivotal(ns, p, draws)
            except Exception:
                with pytest.raises(Exception):
                    rccsd.sample_pivotal(ns, p, draws=draws)
                continue
            got = rccsd.sample_pivotal(ns, p, draws=draws)
            assert (np.array(got, dtype=np.int64) == np.array(want, dtype=np.int64)).all(), k
    finally:
        rccsd.set_option("pivotal_parallel", -1)


def test_parallel_d_selection(rccsd, port):
    """get_d_largest by the parallel rounding-map scan (fricc_escan.cu: d_greedy_parallel) vs the one-warp walk vs the
    oracle: identical D and bit-identical remaining norm, including lists where r drops through several binades, exact
    ties (dyadic magnitudes), zeros, and D == everything."""
    rng = np.random.default_rng(17)
    cases = []
    x = synth_t2(rng, 400000, "heavier"); cases.append((x, 300000))
    x = synth_t2(rng, 200000, "uniform"); cases.append((x, 150000))
    x = np.concatenate([rng.random(50) * 1e6, rng.random(300000) * 1e-3]); cases.append((x, 200000))     # r falls ~30 binades
    x = rng.integers(1, 2**12, 300000) * 2.0**-20; cases.append((x, 250000))                             # dyadic: ties
    x = np.concatenate([2.0 ** rng.integers(-40, 10, 100000).astype(np.float64), np.zeros(1000)]); cases.append((x, 100500))
    x = np.full(70000, 0.125); cases.append((x, 70000))                                                  # D == everything
    x = np.concatenate([[1.0], rng.integers(0, 8, 100000) * 2.0**-53]); cases.append((x, 60000))
    for k, (x, m) in enumerate(cases):
        x = x * np.where(rng.random(x.size) < 0.5, -1.0, 1.0)
        wantD, wantr = port.get_d_largest(x, m)
        try:
            res = {}
            for mode in (0, 1):
                rccsd.set_option("d_parallel", mode)
                res[mode] = rccsd.get_d_largest(x, m)
        finally:
            rccsd.set_option("d_parallel", -1)
        for mode in (0, 1):
            D, r = res[mode]
            assert len(D) == len(wantD), (k, mode, len(D), len(wantD))
            assert r == wantr, (k, mode)
            assert np.array_equal(np.abs(x[D]), np.abs(x[wantD])), (k, mode)


def test_device_rand_stream_equals_host_stream(rccsd):
    """The device generator (per-thread jump-ahead, fricc_rand.cu) produces exactly the host/libc stream, and a pivotal
    compression that draws from the library stream consumes it exactly like one fed the same draws explicitly."""
    import ctypes

    import torch
    from fricc_b200 import _lib

    L = _lib.lib()
    for seed, n in ((1, 5), (4, 992), (4, 993), (11, 3_000_001)):
        _lib.check(L.fricc_srand(seed))
        d = torch.empty(n, dtype=torch.int32, device="cuda")
        _lib.check(L.fricc_rand_fill_device(ctypes.c_void_p(d.data_ptr()), n))
        assert L.fricc_rand_consumed() == 0
        want = np.empty(n, dtype=np.int32)
        _lib.check(L.fricc_rand_fill(want.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), n))
        assert (d.cpu().numpy() == want).all(), (seed, n)
    rng = np.random.default_rng(2)
    x = synth_t2(rng, 500000, "heavier")
    m = 30000
    _lib.check(L.fricc_srand(77))
    draws = np.empty(x.size, dtype=np.int32)
    _lib.check(L.fricc_rand_fill(draws.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), x.size))
    idx0, val0 = rccsd.fri_compression(x, m, "pivotal", draws=draws)
    _lib.check(L.fricc_srand(77))
    idx1, val1 = rccsd.fri_compression(x, m, "pivotal")
    used = L.fricc_rand_consumed()
    assert (idx0 == idx1).all() and np.array_equal(val0, val1)
    # the stream stands where the reference's rand() would: the next value is draws[used]
    nxt = np.empty(1, dtype=np.int32)
    _lib.check(L.fricc_rand_fill(nxt.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), 1))
    assert 0 < used <= x.size - 1 and nxt[0] == draws[used]
