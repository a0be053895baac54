"""TEST INFRASTRUCTURE: a small RCCSD driver that calls the `py_rccsd` surface exactly where the reference's driver does.

Purpose: parity item P7 -- "the CCSD correlation energy must match within 1e-8 Eh at equal m and seed" --
needs a driver around the hot path.  The reference's driver (fricc/friccsd.py) subclasses PySCF's CCSD, and PySCF is not
installed in this image, so the molecules of BASELINE.json cannot be run here.  This file restates the reference's
`update_amps` (fricc/friccsd.py:41-297) and its sparse intermediates (`sparse_cc_Woooo` :306-320, `cc_Wvvvv` :323-329,
`sparse_cc_Wvoov` :332-360, `sparse_cc_Wvovo` :363-378) on plain numpy arrays, with the native module as a parameter:

    backend.SparseTensor4d(t2, t2.shape, m, compression, sampling_method, verbose)
    backend.contract_DTSpT(W, T, out, term)

`backend` is either the reference's own compiled module (oracle/_ref: fricc/py_rccsd*.so) or fricc_b200.py_rccsd, so the
two can be run side by side on the same synthetic closed-shell Hamiltonian with the same rand() seed.  The dense
intermediates the reference takes from pyscf.cc.rintermediates (cc_Foo, cc_Fvv, cc_Fov, Loo, Lvv, cc_Woooo, cc_Wvoov,
cc_Wvovo) are restated from the RCCSD equations; tests/test_rccsd_energy.py checks the restatement against exact
diagonalisation (two-electron systems, for which CCSD is exact) and size consistency / orbital invariance.
Nothing under fricc_b200/ imports this file."""
import numpy as np

ALL_COMPRESSED = ("O^2V^4", "O^3V^3", "O^4V^2")


# ----------------------------------------------------------------------------------------------------------------------
# a synthetic closed-shell "molecule": one-electron matrix h, two-electron integrals (pq|rs), RHF orbitals
# ----------------------------------------------------------------------------------------------------------------------
def random_hamiltonian(rng, norb, strength=0.35, naux=None):
    """Symmetric h with well separated levels and positive semi-definite (pq|rs) = sum_P B_P[pq] B_P[rs] (8-fold symmetric)."""
    naux = naux or 3 * norb
    h = np.diag(np.linspace(-2.0, 1.5, norb)) + 0.05 * rng.standard_normal((norb, norb))
    h = 0.5 * (h + h.T)
    B = strength * rng.standard_normal((naux, norb, norb)) / np.sqrt(naux)
    B = 0.5 * (B + B.transpose(0, 2, 1))
    eri = np.einsum("Ppq,Prs->pqrs", B, B)
    return h, eri


def block_diagonal(systems):
    """Non-interacting union of fragments: h and (pq|rs) are zero between fragments."""
    n = sum(h.shape[0] for h, _ in systems)
    H = np.zeros((n, n))
    E = np.zeros((n, n, n, n))
    o = 0
    for h, e in systems:
        k = h.shape[0]
        H[o:o + k, o:o + k] = h
        E[o:o + k, o:o + k, o:o + k, o:o + k] = e
        o += k
    return H, E


def rhf(h, eri, nocc, iters=200, tol=1e-13):
    """Restricted Hartree-Fock in an orthonormal basis -> (mo_coeff, mo_energy, e_hf)."""
    norb = h.shape[0]
    e, c = np.linalg.eigh(h)
    for _ in range(iters):
        d = c[:, :nocc] @ c[:, :nocc].T
        f = h + 2 * np.einsum("pqrs,rs->pq", eri, d) - np.einsum("prqs,rs->pq", eri, d)
        e_new, c_new = np.linalg.eigh(f)
        d_new = c_new[:, :nocc] @ c_new[:, :nocc].T
        done = np.abs(d_new - d).max() < tol
        e, c = e_new, c_new
        if done:
            break
    d = c[:, :nocc] @ c[:, :nocc].T
    f = h + 2 * np.einsum("pqrs,rs->pq", eri, d) - np.einsum("prqs,rs->pq", eri, d)
    e_hf = np.einsum("pq,pq->", h + f, d)
    assert norb >= nocc
    return c, e, float(e_hf)


class Eris:
    """The integral blocks the reference reads from PySCF's `_ChemistsERIs` (chemists' notation, MO basis):
    fock, mo_energy, oooo, ovoo, ovov, oovv, ovvo, ovvv (get_ovvv), vvvv (imd._get_vvvv)."""

    def __init__(self, h_ao, eri_ao, mo_coeff, nocc):
        c = mo_coeff
        h = c.T @ h_ao @ c
        g = np.einsum("pqrs,pi,qj,rk,sl->ijkl", eri_ao, c, c, c, c, optimize=True)
        o, v = slice(0, nocc), slice(nocc, None)
        self.nocc = nocc
        self.fock = h + 2 * np.einsum("pqkk->pq", g[:, :, o, o]) - np.einsum("pkkq->pq", g[:, o, o, :])
        self.mo_energy = np.diag(self.fock).copy()
        self.oooo = g[o, o, o, o].copy()
        self.ovoo = g[o, v, o, o].copy()
        self.ovov = g[o, v, o, v].copy()
        self.oovv = g[o, o, v, v].copy()
        self.ovvo = g[o, v, v, o].copy()
        self.ovvv = g[o, v, v, v].copy()
        self.vvvv = g[v, v, v, v].copy()
        self.h_mo, self.g_mo = h, g

    def get_ovvv(self):
        return self.ovvv


# ----------------------------------------------------------------------------------------------------------------------
# dense intermediates (the functions fricc/friccsd.py takes from pyscf.cc.rintermediates: called at :144-146,190-191,
# 196,204-205 of the reference driver).  Closed-shell RCCSD, chemists' integrals.
# ----------------------------------------------------------------------------------------------------------------------
def cc_Foo(t1, t2, eris):
    nocc = t1.shape[0]
    ovov = eris.ovov
    F = 2 * np.einsum("kcld,ilcd->ki", ovov, t2) - np.einsum("kdlc,ilcd->ki", ovov, t2)
    F += 2 * np.einsum("kcld,ic,ld->ki", ovov, t1, t1) - np.einsum("kdlc,ic,ld->ki", ovov, t1, t1)
    return F + eris.fock[:nocc, :nocc]


def cc_Fvv(t1, t2, eris):
    nocc = t1.shape[0]
    ovov = eris.ovov
    F = -2 * np.einsum("kcld,klad->ac", ovov, t2) + np.einsum("kdlc,klad->ac", ovov, t2)
    F += -2 * np.einsum("kcld,ka,ld->ac", ovov, t1, t1) + np.einsum("kdlc,ka,ld->ac", ovov, t1, t1)
    return F + eris.fock[nocc:, nocc:]


def cc_Fov(t1, t2, eris):
    nocc = t1.shape[0]
    ovov = eris.ovov
    F = 2 * np.einsum("kcld,ld->kc", ovov, t1) - np.einsum("kdlc,ld->kc", ovov, t1)
    return F + eris.fock[:nocc, nocc:]


def Loo(t1, t2, eris):
    nocc = t1.shape[0]
    fov = eris.fock[:nocc, nocc:]
    L = cc_Foo(t1, t2, eris) + np.einsum("kc,ic->ki", fov, t1)
    L += 2 * np.einsum("lcki,lc->ki", eris.ovoo, t1) - np.einsum("kcli,lc->ki", eris.ovoo, t1)
    return L


def Lvv(t1, t2, eris):
    nocc = t1.shape[0]
    fov = eris.fock[:nocc, nocc:]
    ovvv = eris.get_ovvv()
    L = cc_Fvv(t1, t2, eris) - np.einsum("kc,ka->ac", fov, t1)
    L += 2 * np.einsum("kdac,kd->ac", ovvv, t1) - np.einsum("kcad,kd->ac", ovvv, t1)
    return L


def cc_Woooo(t1, t2, eris):
    W = np.einsum("lcki,jc->klij", eris.ovoo, t1) + np.einsum("kclj,ic->klij", eris.ovoo, t1)
    W += np.einsum("kcld,ijcd->klij", eris.ovov, t2) + np.einsum("kcld,ic,jd->klij", eris.ovov, t1, t1)
    return W + eris.oooo.transpose(0, 2, 1, 3)


def cc_Wvoov(t1, t2, eris):
    ovvv, ovoo, ovov = eris.get_ovvv(), eris.ovoo, eris.ovov
    W = np.einsum("kcad,id->akic", ovvv, t1) - np.einsum("kcli,la->akic", ovoo, t1)
    W += eris.ovvo.transpose(2, 0, 3, 1)
    W -= 0.5 * np.einsum("ldkc,ilda->akic", ovov, t2)
    W -= 0.5 * np.einsum("lckd,ilad->akic", ovov, t2)
    W -= np.einsum("ldkc,id,la->akic", ovov, t1, t1)
    W += np.einsum("ldkc,ilad->akic", ovov, t2)
    return W


def cc_Wvovo(t1, t2, eris):
    ovvv, ovoo, ovov = eris.get_ovvv(), eris.ovoo, eris.ovov
    W = np.einsum("kdac,id->akci", ovvv, t1) - np.einsum("lcki,la->akci", ovoo, t1)
    W += eris.oovv.transpose(2, 0, 3, 1)
    W -= 0.5 * np.einsum("lckd,ilda->akci", ovov, t2)
    W -= np.einsum("lckd,id,la->akci", ovov, t1, t1)
    return W


def cc_Wvvvv(t1, t2, eris):
    """fricc/friccsd.py:323-329 (t2 is not used)."""
    ovvv = eris.get_ovvv()
    W = np.einsum("kdac,kb->abcd", ovvv, -t1) - np.einsum("kcbd,ka->abcd", ovvv, t1)
    return W + eris.vvvv.transpose(0, 2, 1, 3)


# ---- the reference's sparse forms of the same intermediates -----------------------------------------------------------
def sparse_cc_Woooo(t1, T, eris, backend):
    """fricc/friccsd.py:306-320: the t2 term is contract_DTSpT(ovov, T, W, "1323")."""
    W = np.einsum("lcki,jc->klij", eris.ovoo, t1) + np.einsum("kclj,ic->klij", eris.ovoo, t1)
    W = np.ascontiguousarray(W)
    backend.contract_DTSpT(eris.ovov, T, W, "1323")
    W += np.einsum("kcld,ic,jd->klij", eris.ovov, t1, t1)
    W += eris.oooo.transpose(0, 2, 1, 3)
    return W


def sparse_cc_Wvoov(t1, T, eris, backend):
    """fricc/friccsd.py:332-360: the three t2 terms are the contractions 0112, 0313, 0113."""
    W = np.einsum("kcad,id->akic", eris.get_ovvv(), t1) - np.einsum("kcli,la->akic", eris.ovoo, t1)
    W = np.ascontiguousarray(W + eris.ovvo.transpose(2, 0, 3, 1))
    for term in ("0112", "0313", "0113"):
        backend.contract_DTSpT(eris.ovov, T, W, term)
    W -= np.einsum("ldkc,id,la->akic", eris.ovov, t1, t1)
    return W


def sparse_cc_Wvovo(t1, T, eris, backend):
    """fricc/friccsd.py:363-378: the t2 term is the contraction 0312."""
    W = np.einsum("kdac,id->akci", eris.get_ovvv(), t1) - np.einsum("lcki,la->akci", eris.ovoo, t1)
    W = np.ascontiguousarray(W + eris.oovv.transpose(2, 0, 3, 1))
    backend.contract_DTSpT(eris.ovov, T, W, "0312")
    W -= np.einsum("lckd,id,la->akci", eris.ovov, t1, t1)
    return W


# ----------------------------------------------------------------------------------------------------------------------
# update_amps: fricc/friccsd.py:41-297
# ----------------------------------------------------------------------------------------------------------------------
def update_amps(t1, t2, eris, backend=None, m_keep=None, compressed=(), compression="fri", sampling="systematic",
                level_shift=0.0, as_reference=True):
    """One RCCSD amplitude update.  compressed: subset of ALL_COMPRESSED -- the contraction classes evaluated through
    backend.contract_DTSpT on compressed amplitudes (the reference's fri_settings["compressed_contractions"]); empty =
    the dense equations.  The two compressions per update (t2_sparse, t2_sparse_alt, :100-136) are made whenever a class
    is compressed, in the reference's order, so that both backends consume the same rand() draws.
    as_reference: fricc/friccsd.py:225 writes the t1*t1 part of the particle-particle ladder as
    einsum("abcd,ia,jb->ijab", Wvvvv, t1, t1) -- c and d are summed without being contracted with t1; the RCCSD equation
    (tau_ijcd = t2_ijcd + t1_ic t1_jd) needs "abcd,ic,jd->ijab".  True (default) mirrors the reference as it is, which is
    what energy parity between two native backends needs; False uses the RCCSD term (exact for two electrons, invariant
    under virtual-virtual rotations -- tests/test_rccsd_energy.py shows both)."""
    nocc, nvir = t1.shape
    fock = eris.fock
    mo_e_o = eris.mo_energy[:nocc]
    mo_e_v = eris.mo_energy[nocc:] + level_shift
    fov = fock[:nocc, nocc:].copy()
    t2 = np.ascontiguousarray(t2)
    T = T_alt = None
    if compressed:
        if not (0 < m_keep <= t2.size):
            raise ValueError("Bad m_keep value! The following condition wasn't met 0 < m_keep <= t2.size")   # :69-72
        T = backend.SparseTensor4d(t2, t2.shape, int(m_keep), compression, sampling, False)                 # :100-107
        if "O^3V^3" in compressed or "O^4V^2" in compressed:
            T_alt = backend.SparseTensor4d(t2, t2.shape, int(m_keep), compression, sampling, False)         # :108-117

    Foo, Fvv, Fov = cc_Foo(t1, t2, eris), cc_Fvv(t1, t2, eris), cc_Fov(t1, t2, eris)
    Foo[np.diag_indices(nocc)] -= mo_e_o
    Fvv[np.diag_indices(nvir)] -= mo_e_v

    # ---- T1 equation (:152-176)
    ovvv, ovoo = eris.get_ovvv(), eris.ovoo
    t1new = -2 * np.einsum("kc,ka,ic->ia", fov, t1, t1)
    t1new += np.einsum("ac,ic->ia", Fvv, t1) - np.einsum("ki,ka->ia", Foo, t1)
    t1new += 2 * np.einsum("kc,kica->ia", Fov, t2) - np.einsum("kc,ikca->ia", Fov, t2)
    t1new += np.einsum("kc,ic,ka->ia", Fov, t1, t1)
    t1new += fov
    t1new += 2 * np.einsum("kcai,kc->ia", eris.ovvo, t1) - np.einsum("kiac,kc->ia", eris.oovv, t1)
    t1new += 2 * np.einsum("kdac,ikcd->ia", ovvv, t2) - np.einsum("kcad,ikcd->ia", ovvv, t2)
    t1new += 2 * np.einsum("kdac,kd,ic->ia", ovvv, t1, t1) - np.einsum("kcad,kd,ic->ia", ovvv, t1, t1)
    t1new += -2 * np.einsum("lcki,klac->ia", ovoo, t2) + np.einsum("kcli,klac->ia", ovoo, t2)
    t1new += -2 * np.einsum("lcki,lc,ka->ia", ovoo, t1, t1) + np.einsum("kcli,lc,ka->ia", ovoo, t1, t1)

    # ---- T2 equation (:178-282)
    tmp2 = np.einsum("kibc,ka->abic", eris.oovv, -t1) + ovvv.transpose(1, 3, 0, 2)
    tmp = np.einsum("abic,jc->ijab", tmp2, t1)
    t2new = tmp + tmp.transpose(1, 0, 3, 2)
    tmp2 = np.einsum("kcai,jc->akij", eris.ovvo, t1) + ovoo.transpose(1, 3, 0, 2)
    tmp = np.einsum("akij,kb->ijab", tmp2, t1)
    t2new -= tmp + tmp.transpose(1, 0, 3, 2)
    t2new += eris.ovov.transpose(0, 2, 1, 3)
    t2new = np.ascontiguousarray(t2new)

    Lo, Lv = Loo(t1, t2, eris), Lvv(t1, t2, eris)
    Lo[np.diag_indices(nocc)] -= mo_e_o
    Lv[np.diag_indices(nvir)] -= mo_e_v

    Woooo = sparse_cc_Woooo(t1, T_alt, eris, backend) if "O^4V^2" in compressed else cc_Woooo(t1, t2, eris)
    if "O^3V^3" in compressed:
        Wvoov = sparse_cc_Wvoov(t1, T_alt, eris, backend)
        Wvovo = sparse_cc_Wvovo(t1, T_alt, eris, backend)
    else:
        Wvoov, Wvovo = cc_Wvoov(t1, t2, eris), cc_Wvovo(t1, t2, eris)
    Wvvvv = np.ascontiguousarray(cc_Wvvvv(t1, t2, eris))

    if "O^4V^2" in compressed:                                                                   # :212-218
        t2new += np.einsum("klij,ka,lb->ijab", Woooo, t1, t1)
        backend.contract_DTSpT(np.ascontiguousarray(Woooo), T, t2new, "0101")
    else:
        tau = t2 + np.einsum("ia,jb->ijab", t1, t1)
        t2new += np.einsum("klij,klab->ijab", Woooo, tau)

    if as_reference:
        t2new += np.einsum("abcd,ia,jb->ijab", Wvvvv, t1, t1)                                    # :225, as written
    else:
        t2new += np.einsum("abcd,ic,jd->ijab", Wvvvv, t1, t1)
    if "O^2V^4" in compressed:
        backend.contract_DTSpT(Wvvvv, T, t2new, "2323")                                          # :229-231
    else:
        t2new += np.einsum("abcd,ijcd->ijab", Wvvvv, t2)

    tmp = np.einsum("ac,ijcb->ijab", Lv, t2)                                                     # :242-245
    t2new += tmp + tmp.transpose(1, 0, 3, 2)
    tmp = np.einsum("ki,kjab->ijab", Lo, t2)
    t2new -= tmp + tmp.transpose(1, 0, 3, 2)

    if "O^3V^3" in compressed:                                                                   # :248-271
        Wvoov, Wvovo = np.ascontiguousarray(Wvoov), np.ascontiguousarray(Wvovo)
        tmp = np.zeros_like(t2new)
        backend.contract_DTSpT(Wvoov, T, tmp, "1302")
        backend.contract_DTSpT(Wvovo, T, tmp, "1202")
        t2new += tmp + tmp.transpose(1, 0, 3, 2)
        tmp = np.zeros_like(t2new)
        backend.contract_DTSpT(Wvoov, T, tmp, "1303")
        t2new -= tmp + tmp.transpose(1, 0, 3, 2)
        tmp = np.zeros_like(t2new)
        backend.contract_DTSpT(Wvovo, T, tmp, "1203")
        t2new -= tmp + tmp.transpose(1, 0, 3, 2)
    else:                                                                                        # :273-282
        tmp = 2 * np.einsum("akic,kjcb->ijab", Wvoov, t2) - np.einsum("akci,kjcb->ijab", Wvovo, t2)
        t2new += tmp + tmp.transpose(1, 0, 3, 2)
        tmp = np.einsum("akic,kjbc->ijab", Wvoov, t2)
        t2new -= tmp + tmp.transpose(1, 0, 3, 2)
        tmp = np.einsum("bkci,kjac->ijab", Wvovo, t2)
        t2new -= tmp + tmp.transpose(1, 0, 3, 2)

    eia = mo_e_o[:, None] - mo_e_v                                                                # :284-287
    eijab = eia[:, None, :, None] + eia[None, :, None, :]
    return t1new / eia, t2new / eijab


def energy(t1, t2, eris):
    """RCCSD correlation energy (what PySCF's `energy` returns for the reference driver)."""
    nocc = t1.shape[0]
    fov = eris.fock[:nocc, nocc:]
    tau = t2 + np.einsum("ia,jb->ijab", t1, t1)
    e = 2 * np.einsum("ia,ia", fov, t1)
    e += 2 * np.einsum("ijab,iajb", tau, eris.ovov) - np.einsum("ijab,ibja", tau, eris.ovov)
    return float(e)


def init_amps(eris):
    """MP2 guess (PySCF's get_init_guess): t1 = fov / eia, t2 = (ia|jb) / eijab."""
    nocc = eris.nocc
    mo_e = eris.mo_energy
    eia = mo_e[:nocc, None] - mo_e[None, nocc:]
    eijab = eia[:, None, :, None] + eia[None, :, None, :]
    t1 = eris.fock[:nocc, nocc:] / eia
    t2 = eris.ovov.transpose(0, 2, 1, 3) / eijab
    return t1, np.ascontiguousarray(t2)


def run(eris, n_iter, backend=None, m_keep=None, compressed=(), compression="fri", sampling="systematic",
        as_reference=True):
    """Plain fixed-point iteration (no DIIS, as the reference configures its runs) -> list of correlation energies."""
    t1, t2 = init_amps(eris)
    energies = [energy(t1, t2, eris)]
    for _ in range(n_iter):
        t1, t2 = update_amps(t1, t2, eris, backend, m_keep, compressed, compression, sampling,
                             as_reference=as_reference)
        energies.append(energy(t1, t2, eris))
    return energies, t1, t2


def fci_two_electrons(h, eri):
    """Exact singlet ground-state energy of two electrons in the orbitals of (h, eri): the spatial wavefunction
    sum_pq C_pq phi_p(1) phi_q(2) with symmetric C; H C = h C + C h + sum_rs (pr|qs) C_rs."""
    n = h.shape[0]
    H = np.einsum("pr,qs->pqrs", h, np.eye(n)) + np.einsum("pr,qs->pqrs", np.eye(n), h) + eri.transpose(0, 2, 1, 3)
    H = H.reshape(n * n, n * n)
    # restrict to the symmetric (singlet) subspace
    basis = []
    for p in range(n):
        for q in range(p, n):
            v = np.zeros((n, n))
            v[p, q] += 1
            v[q, p] += 1
            basis.append(v.ravel() / np.linalg.norm(v))
    Bm = np.array(basis).T
    return float(np.linalg.eigvalsh(Bm.T @ H @ Bm)[0])
