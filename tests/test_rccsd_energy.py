"""Energy-level checks around the hot path (parity item P7) on synthetic closed-shell Hamiltonians
(PySCF is not available in this image, so BASELINE.json's molecules cannot be run): tests/rccsd_driver.py restates the
reference's `update_amps` (fricc/friccsd.py:41-297) with the native module as a parameter.

CPU part (this file): (1) the restated RCCSD equations are exact -- CCSD equals exact diagonalisation for two electrons,
is size consistent and invariant under occupied-occupied / virtual-virtual rotations; (2) with the reference's own
compiled module as the backend and m = N ("largest" keeps everything) every compressed contraction class reproduces the
dense equations, i.e. the 11 sparse terms are wired with the reference's signs and prefactors; (3) an FRI run (m < N)
with the reference backend is reproducible from the rand() seed.  The GPU part (test_rccsd_energy_gpu.py) runs the same
driver with fricc_b200.py_rccsd as the backend against the reference backend at equal m and seed."""
import ctypes

import numpy as np
import pytest

import rccsd_driver as D


def _rot(rng, n, s):
    a = rng.standard_normal((n, n)) * s
    a = a - a.T
    w, v = np.linalg.eigh(1j * a)
    return (v @ np.diag(np.exp(-1j * w)) @ v.conj().T).real


def _system(seed, norb, nocc, strength=0.35):
    rng = np.random.default_rng(seed)
    h, g = D.random_hamiltonian(rng, norb, strength=strength)
    c, _, ehf = D.rhf(h, g, nocc, iters=2000, tol=1e-15)
    return h, g, c, ehf


def test_ccsd_is_exact_for_two_electrons():
    h, g, c, ehf = _system(0, 6, 1)
    er = D.Eris(h, g, c, 1)
    exact = D.fci_two_electrons(h, g) - ehf
    es, _, _ = D.run(er, 200, as_reference=False)
    assert abs(es[-1] - exact) < 1e-12
    # the reference driver's t1*t1 ladder term as written (fricc/friccsd.py:225) is not the RCCSD term: small but real
    es_ref, _, _ = D.run(er, 200, as_reference=True)
    assert 1e-12 < abs(es_ref[-1] - exact) < 1e-7


def test_size_consistency_and_orbital_invariance():
    rng = np.random.default_rng(1)
    hA, gA = D.random_hamiltonian(rng, 4)
    hB, gB = D.random_hamiltonian(rng, 5)
    _, _, eA = D.rhf(hA, gA, 1, iters=2000, tol=1e-15)
    _, _, eB = D.rhf(hB, gB, 1, iters=2000, tol=1e-15)
    exact = D.fci_two_electrons(hA, gA) - eA + D.fci_two_electrons(hB, gB) - eB
    h, g = D.block_diagonal([(hA, gA), (hB, gB)])
    c, _, ehf = D.rhf(h, g, 2, iters=2000, tol=1e-15)
    assert abs(ehf - (eA + eB)) < 1e-12
    es, _, _ = D.run(D.Eris(h, g, c, 2), 300, as_reference=False)
    assert abs(es[-1] - exact) < 1e-12                      # two non-interacting pairs: CCSD = FCI_A + FCI_B
    c2 = c.copy()                                           # delocalise: mixes the fragments' occupied / virtual orbitals
    c2[:, :2] = c[:, :2] @ _rot(rng, 2, 0.7)
    c2[:, 2:] = c[:, 2:] @ _rot(rng, 7, 0.3)
    es2, _, _ = D.run(D.Eris(h, g, c2, 2), 600, as_reference=False)
    assert abs(es2[-1] - exact) < 1e-11


class _RefBackend:
    """The reference's own compiled module (oracle/_ref), called exactly as fricc/friccsd.py calls it."""

    def __init__(self, ref):
        self.SparseTensor4d = ref.mod.SparseTensor4d
        self.contract_DTSpT = ref.mod.contract_DTSpT

    @staticmethod
    def srand(seed):
        ctypes.CDLL(None).srand(int(seed))      # the libc state the reference's rand() reads


@pytest.mark.parametrize("compressed", [("O^2V^4",), ("O^3V^3",), ("O^4V^2",), D.ALL_COMPRESSED])
def test_sparse_terms_are_wired_like_the_dense_equations(ref, compressed):
    """m = N with "largest" keeps every amplitude: each compressed class must then reproduce the dense update."""
    h, g, c, _ = _system(2, 9, 3)
    er = D.Eris(h, g, c, 3)
    B = _RefBackend(ref)
    t1, t2 = D.init_amps(er)
    for _ in range(3):
        d1, d2 = D.update_amps(t1, t2, er)
        s1, s2 = D.update_amps(t1, t2, er, B, t2.size, compressed, "largest")
        assert np.abs(s1 - d1).max() < 1e-13 and np.abs(s2 - d2).max() < 1e-12
        t1, t2 = d1, d2


@pytest.mark.parametrize("sampling", ["systematic", "pivotal"])
def test_fri_run_with_the_reference_backend_is_reproducible(ref, sampling):
    h, g, c, _ = _system(4, 9, 3, strength=0.3)      # a well-gapped system: the plain iteration converges
    er = D.Eris(h, g, c, 3)
    B = _RefBackend(ref)
    m = er.ovov.size // 4
    runs = []
    for _ in range(2):
        B.srand(1234)
        es, _, _ = D.run(er, 6, B, m, D.ALL_COMPRESSED, "fri", sampling)
        runs.append(es)
    assert runs[0] == runs[1]
    dense, _, _ = D.run(er, 6)
    assert np.isfinite(runs[0]).all() and abs(runs[0][-1] - dense[-1]) < 0.2 * abs(dense[-1])
