"""Parity mode of the contractions (option "spmm_exact_order") and the library's own rand() stream.

* every contraction term BIT-IDENTICAL to the compiled reference (oracle/_ref) when the entries of an owner pair are
  added in storage order with the reference's roundings (src/sparse_tensor.cpp:66-73; a plain x86-64 build, no FMA);
* >= 10 consecutive pivotal compressions drawn from the library's glibc-compatible stream against the reference drawing
  from libc, same seed: same entries, same number of draws consumed after every call (early stops included).

These are the two ingredients of the energy parity P7 (tests/test_rccsd_energy_gpu.py): an FRI-CCSD trajectory feeds
each residual into the next compression, and the pivotal sampler's last window is rounding-sensitive
(src/fri_utils.cpp:81-109), so equal energies need equal bits."""
import ctypes

import numpy as np
import pytest
from conftest import TERMS, create_sparse_ndarray, synth_t2

pytestmark = pytest.mark.gpu


@pytest.fixture()
def exact_order(rccsd):
    rccsd.set_option("spmm_exact_order", 1)
    yield rccsd
    rccsd.set_option("spmm_exact_order", 0)


def _pair(rccsd, ref, t2, m, compression, sampling, draws):
    """The same compression on both sides (same injected draws) -> (T_ref, T_ours)."""
    ref.set_draws(draws)
    T_ref = ref.mod.SparseTensor4d(t2, list(t2.shape), m, compression, sampling, False)
    ref.clear_draws()
    T_our = rccsd.SparseTensor4d(t2, t2.shape, m, compression, sampling, False, draws=draws)
    flat, vals = T_our.export()
    for mi in (0, m // 2, m - 1):
        idx, v = T_ref.get_element(mi)
        assert np.ravel_multi_index(idx, t2.shape) == flat[mi] and v == vals[mi]
    return T_ref, T_our


@pytest.mark.parametrize("compression,sampling", [("fri", "systematic"), ("fri", "pivotal"), ("largest", "pivotal")])
@pytest.mark.parametrize("no,nv,frac", [(4, 8, 0.2), (7, 13, 0.05)])
def test_every_term_bit_identical_to_reference(exact_order, ref, compression, sampling, no, nv, frac):
    rccsd = exact_order
    rng = np.random.default_rng(100 * no + nv)
    shape = (no, no, nv, nv)
    n = int(np.prod(shape))
    t2 = synth_t2(rng, n, "heavier").reshape(shape)
    m = max(3, int(frac * n))
    draws = rng.integers(0, 2**31 - 1, size=n + 8).astype(np.int32)
    T_ref, T_our = _pair(rccsd, ref, t2, m, compression, sampling, draws)
    ext = {"o": no, "v": nv}
    for term, (wax, oax, _, _) in sorted(TERMS.items()):
        W = rng.random([ext[c] for c in wax]) - 0.4          # both signs
        o0 = rng.standard_normal([ext[c] for c in oax])
        a, b = o0.copy(), o0.copy()
        for _ in range(2):                                    # accumulate twice onto a non-zero output
            ref.mod.contract_DTSpT(W, T_ref, a, term)
            rccsd.contract_DTSpT(W, T_our, b, term)
        assert np.array_equal(a, b), (term, np.abs(a - b).max())


def test_bit_identical_with_strided_operands_and_duplicates(exact_order, ref):
    """Transposed views of W and of the output (fricc/frilccsd.py:171-174) and a tensor whose kept entries include
    zeros (indices drawn with replacement, fricc/tests/test_sparse_tensors.py:8-15)."""
    rccsd = exact_order
    rng = np.random.RandomState(11)
    no, nv = 5, 9
    shape = (no, no, nv, nv)
    t2 = create_sparse_ndarray(shape, 400, rng)
    T_ref = ref.mod.SparseTensor4d(t2, list(shape), 350, "largest", "pivotal", False)
    T_our = rccsd.SparseTensor4d(t2, shape, 350, "largest")
    # ties (the zeros) may be ordered differently: compare as sets of non-zero entries
    ours = {(int(f), v) for f, v in zip(*T_our.export()) if v != 0.0}
    theirs = set()
    for mi in range(350):
        idx, v = T_ref.get_element(mi)
        if v != 0.0:
            theirs.add((int(np.ravel_multi_index(idx, shape)), v))
    assert ours == theirs
    for term, wshape, perm_w, perm_o in (("0101", (no, no, no, no), (0, 2, 1, 3), (1, 0, 3, 2)),
                                         ("1302", (nv, no, no, nv), (0, 1, 2, 3), (1, 0, 3, 2)),
                                         ("2323", (nv, nv, nv, nv), (1, 0, 3, 2), (0, 1, 2, 3))):
        W = (rng.rand(*wshape) - 0.5).transpose(perm_w)
        ext = {"o": no, "v": nv}
        oshape = [ext[c] for c in TERMS[term][1]]
        base = rng.rand(*oshape)
        a = base.copy().transpose(perm_o)
        b = base.copy().transpose(perm_o)
        if a.shape != tuple(oshape):
            a, b = base.copy(), base.copy()
        ref.mod.contract_DTSpT(W, T_ref, a, term)
        rccsd.contract_DTSpT(W, T_our, b, term)
        assert np.array_equal(a, b), term


def test_fast_and_exact_forms_agree_to_1e10(rccsd, ref):
    rng = np.random.default_rng(5)
    no, nv = 6, 11
    shape = (no, no, nv, nv)
    t2 = synth_t2(rng, int(np.prod(shape)), "heavy").reshape(shape)
    T = rccsd.SparseTensor4d(t2, shape, 500, "fri", "systematic", False, draws=np.array([12345], dtype=np.int32))
    W = rng.random((nv, nv, nv, nv))
    fast, exact = np.zeros(shape), np.zeros(shape)
    rccsd.contract_DTSpT(W, T, fast, "2323")
    rccsd.set_option("spmm_exact_order", 1)
    try:
        rccsd.contract_DTSpT(W, T, exact, "2323")
    finally:
        rccsd.set_option("spmm_exact_order", 0)
    assert np.abs(fast - exact).max() <= 1e-10 * np.abs(exact).max()


@pytest.mark.parametrize("seed,n,m", [(4321, 1600, 160), (7, 1999, 400), (99, 600, 37)])
def test_consecutive_pivotal_compressions_from_the_library_stream(rccsd, ref, seed, n, m):
    """12 compressions in a row, the library drawing from ITS glibc-compatible stream, the reference from libc rand()
    after srand(seed).  The tensors change from call to call; every third one ends in a block of zeros so that the
    pivotal walk collects its n' samples before the end of the vector and stops early (fri_utils.cpp:103-105) --
    the number of draws consumed then differs from N-1 and must still match."""
    libc = ctypes.CDLL(None)
    rng = np.random.default_rng(seed)
    shape = (1, 1, 1, n)
    rccsd.srand(seed)
    libc.srand(seed)
    ref.clear_draws()
    saw_early_stop = False
    for it in range(12):
        x = synth_t2(rng, n, "heavier" if it % 2 else "heavy")
        if it % 3 == 2:
            x[n - n // 5:] = 0.0
        t2 = np.ascontiguousarray(x.reshape(shape))
        ref.reset_rand_calls()
        T_ref = ref.mod.SparseTensor4d(t2, list(shape), m, "fri", "pivotal", False)
        used_ref = ref.rand_calls()
        before = rccsd.rand_consumed()
        T_our = rccsd.SparseTensor4d(t2, shape, m, "fri", "pivotal", False)
        used_our = rccsd.rand_consumed() - before
        assert used_our == used_ref, (it, used_our, used_ref)
        saw_early_stop |= used_ref < n - 1
        flat, vals = T_our.export()
        for mi in range(m):
            idx, v = T_ref.get_element(mi)
            assert idx[3] == flat[mi] and v == vals[mi], (it, mi)
    assert saw_early_stop
    # the two streams are still aligned afterwards
    nxt = np.empty(4, dtype=np.int32)
    rccsd_lib = __import__("fricc_b200._lib", fromlist=["lib"]).lib()
    rccsd_lib.fricc_rand_fill(nxt.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), 4)
    assert [int(v) for v in nxt] == [libc.rand() for _ in range(4)]
