"""Device-resident FRI-CCSD iteration: the caller of the hot path.

The reference's driver (fricc/friccsd.py) keeps t1, t2, the integrals and every intermediate in numpy arrays on the host
and calls into the native module for the two compressions and the eleven sparse contractions of an iteration; everything
else of `update_amps` -- the T1 equation, the Foo/Fvv/Fov/Loo/Lvv intermediates, the t1-dressed parts of Woooo, Wvoov,
Wvovo and Wvvvv, the `tmp + tmp.transpose(1,0,3,2)` symmetrisations and the division by the orbital-energy denominators
(fricc/friccsd.py:144-192,238-241,281-287) -- is numpy einsum.  Once the sparse path runs on the GPU those host einsums
and the 2 x 2 GB of T2 / residual crossing PCIe every iteration are what is left, so this module restates the same
`update_amps` on torch tensors that stay on the device: dense pieces are torch.einsum (cuBLAS DGEMM) and elementwise
kernels, the compressed pieces are the SAME plugin calls (`py_rccsd.SparseTensor4d`, `py_rccsd.contract_DTSpT`) on device
tensors, and nothing but the scalar energy leaves the GPU.

    eris = DeviceEris.from_host(host_eris, device)        # fock, mo_energy, oooo, ovoo, ovov, oovv, ovvo, ovvv, vvvv
    t1, t2 = init_amps(eris)                              # MP2 guess
    t1, t2 = update_amps(t1, t2, eris, m_keep=m, compressed=("O^2V^4", "O^3V^3", "O^4V^2"))
    e_corr  = energy(t1, t2, eris)
    conv, e_corr, t1, t2 = kernel(eris, max_cycle=50, ...)   # the loop of fricc/friccsd.py:379-456 (no DIIS)
    run(eris, n_iter, ...)                                # n_iter updates, no convergence test -> (energies, t1, t2)
    mcmc_std(energies)                                    # fricc/__init__.py:8-11

Line references are to fricc/friccsd.py.  Two notes on fidelity:
  * :225 writes the t1*t1 part of the particle-particle ladder as einsum("abcd,ia,jb->ijab", Wvvvv, t1, t1); the RCCSD
    equation needs "abcd,ic,jd->ijab".  `as_reference=True` (default) mirrors the reference as it is, `False` is the
    RCCSD term (see tests/rccsd_driver.py and tests/test_rccsd_energy.py, which show the difference on exact cases).
  * the intermediates the reference imports from pyscf.cc.rintermediates (cc_Foo, cc_Fvv, cc_Fov, Loo, Lvv, cc_Woooo,
    cc_Wvoov, cc_Wvovo) are restated from the closed-shell RCCSD equations (PySCF is not in this image); the restatement
    is checked against exact diagonalisation in tests/test_rccsd_energy.py through the numpy twin of this file.

torch on a CPU device works for the dense equations (compressed=()) -- that is how the CPU tests compare this file with
its numpy twin; any compressed class needs the CUDA library (no fallback)."""
import torch

ALL_COMPRESSED = ("O^2V^4", "O^3V^3", "O^4V^2")
_BLOCKS = ("fock", "mo_energy", "oooo", "ovoo", "ovov", "oovv", "ovvo", "ovvv", "vvvv")


def _es(expr, *ops):
    return torch.einsum(expr, *ops)


class DeviceEris:
    """The integral blocks of PySCF's `_ChemistsERIs` (chemists' notation, MO basis) as tensors on one device."""

    def __init__(self, nocc, **blocks):
        self.nocc = int(nocc)
        for k in _BLOCKS:
            setattr(self, k, blocks[k])

    @classmethod
    def from_host(cls, eris, device="cuda", dtype=torch.float64):
        """`eris`: any object with numpy attributes fock, mo_energy, oooo, ovoo, ovov, oovv, ovvo, vvvv and either an
        attribute `ovvv` or PySCF's `get_ovvv()`."""
        import numpy as np

        get = lambda k: eris.get_ovvv() if k == "ovvv" and not hasattr(eris, "ovvv") else getattr(eris, k)
        blocks = {k: torch.as_tensor(np.ascontiguousarray(get(k)), dtype=dtype).to(device) for k in _BLOCKS}
        nocc = getattr(eris, "nocc", blocks["oooo"].shape[0])
        return cls(nocc, **blocks)

    def get_ovvv(self):
        return self.ovvv


# ----------------------------------------------------------------------------------------------------------------------
# dense intermediates (pyscf.cc.rintermediates, called at :144-146,190-191,196,204-205)
# ----------------------------------------------------------------------------------------------------------------------
def cc_Foo(t1, t2, eris):
    nocc = t1.shape[0]
    ovov = eris.ovov
    F = 2 * _es("kcld,ilcd->ki", ovov, t2) - _es("kdlc,ilcd->ki", ovov, t2)
    F += 2 * _es("kcld,ic,ld->ki", ovov, t1, t1) - _es("kdlc,ic,ld->ki", ovov, t1, t1)
    return F + eris.fock[:nocc, :nocc]


def cc_Fvv(t1, t2, eris):
    nocc = t1.shape[0]
    ovov = eris.ovov
    F = -2 * _es("kcld,klad->ac", ovov, t2) + _es("kdlc,klad->ac", ovov, t2)
    F += -2 * _es("kcld,ka,ld->ac", ovov, t1, t1) + _es("kdlc,ka,ld->ac", ovov, t1, t1)
    return F + eris.fock[nocc:, nocc:]


def cc_Fov(t1, t2, eris):
    nocc = t1.shape[0]
    ovov = eris.ovov
    F = 2 * _es("kcld,ld->kc", ovov, t1) - _es("kdlc,ld->kc", ovov, t1)
    return F + eris.fock[:nocc, nocc:]


def Loo(t1, t2, eris):
    nocc = t1.shape[0]
    fov = eris.fock[:nocc, nocc:]
    L = cc_Foo(t1, t2, eris) + _es("kc,ic->ki", fov, t1)
    L += 2 * _es("lcki,lc->ki", eris.ovoo, t1) - _es("kcli,lc->ki", eris.ovoo, t1)
    return L


def Lvv(t1, t2, eris):
    nocc = t1.shape[0]
    fov = eris.fock[:nocc, nocc:]
    ovvv = eris.get_ovvv()
    L = cc_Fvv(t1, t2, eris) - _es("kc,ka->ac", fov, t1)
    L += 2 * _es("kdac,kd->ac", ovvv, t1) - _es("kcad,kd->ac", ovvv, t1)
    return L


def cc_Woooo(t1, t2, eris):
    W = _es("lcki,jc->klij", eris.ovoo, t1) + _es("kclj,ic->klij", eris.ovoo, t1)
    W += _es("kcld,ijcd->klij", eris.ovov, t2) + _es("kcld,ic,jd->klij", eris.ovov, t1, t1)
    return W + eris.oooo.permute(0, 2, 1, 3)


def cc_Wvoov(t1, t2, eris):
    ovvv, ovoo, ovov = eris.get_ovvv(), eris.ovoo, eris.ovov
    W = _es("kcad,id->akic", ovvv, t1) - _es("kcli,la->akic", ovoo, t1)
    W += eris.ovvo.permute(2, 0, 3, 1)
    W -= 0.5 * _es("ldkc,ilda->akic", ovov, t2)
    W -= 0.5 * _es("lckd,ilad->akic", ovov, t2)
    W -= _es("ldkc,id,la->akic", ovov, t1, t1)
    W += _es("ldkc,ilad->akic", ovov, t2)
    return W


def cc_Wvovo(t1, t2, eris):
    ovvv, ovoo, ovov = eris.get_ovvv(), eris.ovoo, eris.ovov
    W = _es("kdac,id->akci", ovvv, t1) - _es("lcki,la->akci", ovoo, t1)
    W += eris.oovv.permute(2, 0, 3, 1)
    W -= 0.5 * _es("lckd,ilda->akci", ovov, t2)
    W -= _es("lckd,id,la->akci", ovov, t1, t1)
    return W


def cc_Wvvvv(t1, t2, eris, out=None):
    """:323-329 (t2 is not used).  Built straight into its native [a,b,c,d] layout on the device -- the layout the ladder's
    plane builder reads (each column (a,b) is one contiguous run of v^2 doubles), so no [cd][ab] copy and no host
    hand-off of the v^4 block.  `out`: a preallocated [v,v,v,v] buffer that is rewritten every iteration;
    coming back with the same buffer lets the library re-use the previous iteration's column exponents."""
    ovvv = eris.get_ovvv()
    W = out if out is not None else torch.empty_like(eris.vvvv)
    W.copy_(eris.vvvv.permute(0, 2, 1, 3))
    W -= _es("kdac,kb->abcd", ovvv, t1)
    W -= _es("kcbd,ka->abcd", ovvv, t1)
    return W


# ---- the reference's sparse forms of the same intermediates ---------------------------------------------------------------
def sparse_cc_Woooo(t1, T, eris, backend):
    """:306-320: the t2 term is contract_DTSpT(ovov, T, W, "1323")."""
    W = (_es("lcki,jc->klij", eris.ovoo, t1) + _es("kclj,ic->klij", eris.ovoo, t1)).contiguous()
    backend.contract_DTSpT(eris.ovov, T, W, "1323")
    W += _es("kcld,ic,jd->klij", eris.ovov, t1, t1)
    W += eris.oooo.permute(0, 2, 1, 3)
    return W


def sparse_cc_Wvoov(t1, T, eris, backend):
    """:332-360: the three t2 terms are the contractions 0112, 0313, 0113."""
    W = _es("kcad,id->akic", eris.get_ovvv(), t1) - _es("kcli,la->akic", eris.ovoo, t1)
    W = (W + eris.ovvo.permute(2, 0, 3, 1)).contiguous()
    for term in ("0112", "0313", "0113"):
        backend.contract_DTSpT(eris.ovov, T, W, term)
    W -= _es("ldkc,id,la->akic", eris.ovov, t1, t1)
    return W


def sparse_cc_Wvovo(t1, T, eris, backend):
    """:363-378: the t2 term is the contraction 0312."""
    W = _es("kdac,id->akci", eris.get_ovvv(), t1) - _es("lcki,la->akci", eris.ovoo, t1)
    W = (W + eris.oovv.permute(2, 0, 3, 1)).contiguous()
    backend.contract_DTSpT(eris.ovov, T, W, "0312")
    W -= _es("lckd,id,la->akci", eris.ovov, t1, t1)
    return W


def _backend():
    from . import py_rccsd

    return py_rccsd


def _sym(tmp):
    """tmp + tmp.transpose(1, 0, 3, 2)  (:181,186,245,...)"""
    return tmp + tmp.permute(1, 0, 3, 2)


# ----------------------------------------------------------------------------------------------------------------------
# update_amps (:41-297)
# ----------------------------------------------------------------------------------------------------------------------
def update_amps(t1, t2, eris, m_keep=None, compressed=(), compression="fri", sampling="systematic", level_shift=0.0,
                as_reference=True, backend=None, wvvvv_buffer=None):
    """One RCCSD amplitude update on the device.  Same arguments and behaviour as the numpy twin
    (tests/rccsd_driver.py:update_amps); `backend` defaults to fricc_b200.py_rccsd."""
    nocc, nvir = t1.shape
    fock = eris.fock
    mo_e_o = eris.mo_energy[:nocc]
    mo_e_v = eris.mo_energy[nocc:] + level_shift
    fov = fock[:nocc, nocc:].clone()
    t2 = t2.contiguous()
    T = T_alt = None
    if compressed:
        backend = backend or _backend()
        if m_keep is None or not (0 < m_keep <= t2.numel()):
            raise ValueError("Bad m_keep value! The following condition wasn't met 0 < m_keep <= t2.size")   # :69-72
        T = backend.SparseTensor4d(t2, tuple(t2.shape), int(m_keep), compression, sampling, False)         # :100-107
        if "O^3V^3" in compressed or "O^4V^2" in compressed:
            T_alt = backend.SparseTensor4d(t2, tuple(t2.shape), int(m_keep), compression, sampling, False)  # :108-117

    Foo, Fvv, Fov = cc_Foo(t1, t2, eris), cc_Fvv(t1, t2, eris), cc_Fov(t1, t2, eris)
    Foo.diagonal().sub_(mo_e_o)
    Fvv.diagonal().sub_(mo_e_v)

    # ---- T1 equation (:152-176)
    ovvv, ovoo = eris.get_ovvv(), eris.ovoo
    t1new = -2 * _es("kc,ka,ic->ia", fov, t1, t1)
    t1new += _es("ac,ic->ia", Fvv, t1) - _es("ki,ka->ia", Foo, t1)
    t1new += 2 * _es("kc,kica->ia", Fov, t2) - _es("kc,ikca->ia", Fov, t2)
    t1new += _es("kc,ic,ka->ia", Fov, t1, t1)
    t1new += fov
    t1new += 2 * _es("kcai,kc->ia", eris.ovvo, t1) - _es("kiac,kc->ia", eris.oovv, t1)
    t1new += 2 * _es("kdac,ikcd->ia", ovvv, t2) - _es("kcad,ikcd->ia", ovvv, t2)
    t1new += 2 * _es("kdac,kd,ic->ia", ovvv, t1, t1) - _es("kcad,kd,ic->ia", ovvv, t1, t1)
    t1new += -2 * _es("lcki,klac->ia", ovoo, t2) + _es("kcli,klac->ia", ovoo, t2)
    t1new += -2 * _es("lcki,lc,ka->ia", ovoo, t1, t1) + _es("kcli,lc,ka->ia", ovoo, t1, t1)

    # ---- T2 equation (:178-282)
    tmp2 = _es("kibc,ka->abic", eris.oovv, -t1) + ovvv.permute(1, 3, 0, 2)
    t2new = _sym(_es("abic,jc->ijab", tmp2, t1))
    tmp2 = _es("kcai,jc->akij", eris.ovvo, t1) + ovoo.permute(1, 3, 0, 2)
    t2new -= _sym(_es("akij,kb->ijab", tmp2, t1))
    t2new += eris.ovov.permute(0, 2, 1, 3)
    t2new = t2new.contiguous()

    Lo, Lv = Loo(t1, t2, eris), Lvv(t1, t2, eris)
    Lo.diagonal().sub_(mo_e_o)
    Lv.diagonal().sub_(mo_e_v)

    Woooo = sparse_cc_Woooo(t1, T_alt, eris, backend) if "O^4V^2" in compressed else cc_Woooo(t1, t2, eris)
    if "O^3V^3" in compressed:
        Wvoov = sparse_cc_Wvoov(t1, T_alt, eris, backend)
        Wvovo = sparse_cc_Wvovo(t1, T_alt, eris, backend)
    else:
        Wvoov, Wvovo = cc_Wvoov(t1, t2, eris), cc_Wvovo(t1, t2, eris)
    Wvvvv = cc_Wvvvv(t1, t2, eris, out=wvvvv_buffer)

    if "O^4V^2" in compressed:                                                                   # :212-218
        t2new += _es("klij,ka,lb->ijab", Woooo, t1, t1)
        backend.contract_DTSpT(Woooo.contiguous(), T, t2new, "0101")
    else:
        tau = t2 + _es("ia,jb->ijab", t1, t1)
        t2new += _es("klij,klab->ijab", Woooo, tau)

    if as_reference:
        t2new += _es("abcd,ia,jb->ijab", Wvvvv, t1, t1)                                          # :225, as written
    else:
        t2new += _es("abcd,ic,jd->ijab", Wvvvv, t1, t1)
    if "O^2V^4" in compressed:
        backend.contract_DTSpT(Wvvvv, T, t2new, "2323")                                          # :229-231
    else:
        t2new += _es("abcd,ijcd->ijab", Wvvvv, t2)

    t2new += _sym(_es("ac,ijcb->ijab", Lv, t2))                                                  # :242-245
    t2new -= _sym(_es("ki,kjab->ijab", Lo, t2))

    if "O^3V^3" in compressed:                                                                   # :248-271
        Wvoov, Wvovo = Wvoov.contiguous(), Wvovo.contiguous()
        tmp = torch.zeros_like(t2new)
        backend.contract_DTSpT(Wvoov, T, tmp, "1302")
        backend.contract_DTSpT(Wvovo, T, tmp, "1202")
        t2new += _sym(tmp)
        tmp.zero_()
        backend.contract_DTSpT(Wvoov, T, tmp, "1303")
        t2new -= _sym(tmp)
        tmp.zero_()
        backend.contract_DTSpT(Wvovo, T, tmp, "1203")
        t2new -= _sym(tmp)
    else:                                                                                        # :273-282
        t2new += _sym(2 * _es("akic,kjcb->ijab", Wvoov, t2) - _es("akci,kjcb->ijab", Wvovo, t2))
        t2new -= _sym(_es("akic,kjbc->ijab", Wvoov, t2))
        t2new -= _sym(_es("bkci,kjac->ijab", Wvovo, t2))

    eia = mo_e_o[:, None] - mo_e_v                                                                # :284-287
    eijab = eia[:, None, :, None] + eia[None, :, None, :]
    return t1new / eia, t2new / eijab


def energy(t1, t2, eris):
    """RCCSD correlation energy (fricc/frilccsd.py:504-522; what PySCF's `energy` returns for the reference driver).
    The only number that leaves the device per iteration."""
    nocc = t1.shape[0]
    fov = eris.fock[:nocc, nocc:]
    tau = t2 + _es("ia,jb->ijab", t1, t1)
    e = 2 * _es("ia,ia", fov, t1)
    e += 2 * _es("ijab,iajb", tau, eris.ovov) - _es("ijab,ibja", tau, eris.ovov)
    return float(e)


def init_amps(eris):
    """MP2 guess (PySCF's get_init_guess): t1 = fov / eia, t2 = (ia|jb) / eijab."""
    nocc = eris.nocc
    mo_e = eris.mo_energy
    eia = mo_e[:nocc, None] - mo_e[None, nocc:]
    eijab = eia[:, None, :, None] + eia[None, :, None, :]
    t1 = eris.fock[:nocc, nocc:] / eia
    t2 = eris.ovov.permute(0, 2, 1, 3) / eijab
    return t1, t2.contiguous()


def amplitude_norm(t1, t2, t1_old, t2_old):
    """|| amplitudes_to_vector(new) - amplitudes_to_vector(old) ||  (the convergence measure of kernel(), :431-434)."""
    return float(torch.sqrt((t1 - t1_old).pow(2).sum() + (t2 - t2_old).pow(2).sum()))


def run(eris, n_iter, m_keep=None, compressed=(), compression="fri", sampling="systematic", as_reference=True,
        backend=None):
    """n_iter plain fixed-point updates from the MP2 guess -> (energies, t1, t2); the twin of tests/rccsd_driver.py:run."""
    res = kernel(eris, max_cycle=n_iter, tol=0.0, tolnormt=0.0, m_keep=m_keep, compressed=compressed, compression=compression,
                 sampling=sampling, as_reference=as_reference, backend=backend)
    return res.energies, res.t1, res.t2


class KernelResult:
    """What kernel() returns: conv, e_corr, t1, t2 (the reference's tuple, :456) + the energy trace (`mycc.energies`, :411)
    with the initial energy in front, and the running average of the amplitudes when it was asked for."""

    def __init__(self, conv, e_corr, t1, t2, energies, average=None):
        self.conv, self.e_corr, self.t1, self.t2, self.energies, self.average = conv, e_corr, t1, t2, energies, average

    def __iter__(self):                      # conv, eccsd, t1, t2 = kernel(...)
        return iter((self.conv, self.e_corr, self.t1, self.t2))


def kernel(eris, t1=None, t2=None, max_cycle=50, tol=1e-8, tolnormt=1e-6, m_keep=None, compressed=(), compression="fri",
           sampling="systematic", as_reference=True, backend=None, average_from=None, callback=None):
    """The iteration loop of fricc/friccsd.py:379-456 with the amplitudes resident on the device: update_amps, the norm of
    the amplitude change (:431-434), no DIIS (:438-439), the energy (:440), and the three exits -- converged (|dE| < tol
    and norm < tolnormt, :455), NaN (:458-461), |dE| > 1e5 (:462-467).  Only the energy and the norm -- two scalars --
    leave the GPU per cycle.

    average_from: cycle from which the amplitudes are also averaged -- the "unbiased" estimator the reference sketches at
    :119-127 (its branch is dead code there); the result's `.average` is then (t1_avg, t2_avg, energy of the average)."""
    if t1 is None and t2 is None:
        t1, t2 = init_amps(eris)
    elif t2 is None:
        t2 = init_amps(eris)[1]
    eold = 0.0
    eccsd = energy(t1, t2, eris)
    energies = [eccsd]
    wbuf = torch.empty_like(eris.vvvv)       # Wvvvv is rebuilt into the same buffer every cycle
    # the integrals never change during the loop: with the B200 backend the digit planes of eris.ovov (five of the
    # eleven terms read it, :347-378) are then built in the first cycle and kept
    declared = None
    if compressed and eris.ovov.is_cuda and hasattr(backend or _backend(), "constant_operand"):
        declared = backend or _backend()
        declared.constant_operand(eris.ovov, True)
    t1_sum = t2_sum = None
    n_avg = 0
    conv = False
    try:
        for istep in range(max_cycle):
            t1new, t2new = update_amps(t1, t2, eris, m_keep, compressed, compression, sampling, as_reference=as_reference,
                                       backend=backend, wvvvv_buffer=wbuf)
            normt = amplitude_norm(t1new, t2new, t1, t2)
            t1, t2 = t1new, t2new
            eold, eccsd = eccsd, energy(t1, t2, eris)
            energies.append(eccsd)
            if callback is not None:
                callback(istep, eccsd, eccsd - eold, normt)
            if average_from is not None and istep + 1 >= average_from:
                t1_sum = t1.clone() if t1_sum is None else t1_sum.add_(t1)
                t2_sum = t2.clone() if t2_sum is None else t2_sum.add_(t2)
                n_avg += 1
            if abs(eccsd - eold) < tol and normt < tolnormt:
                conv = True
                break
            if eccsd != eccsd or eold != eold:
                break                                           # "FAILED THERE ARE NANS"
            if abs(eccsd - eold) > 1e5:
                break                                           # "FAILED BECAUSE dE = ..."
    finally:
        if declared is not None:
            declared.constant_operand(eris.ovov, False)
    average = None
    if average_from is not None and n_avg > 0:
        t1a, t2a = t1_sum / n_avg, t2_sum / n_avg
        average = (t1a, t2a, energy(t1a, t2a, eris))
    return KernelResult(conv, eccsd, t1, t2, energies, average)


def integrated_time(x, c=5.0):
    """Integrated autocorrelation time of a 1-D trace: normalised autocorrelation by FFT, tau(M) = 2 sum_{t<=M} rho(t) - 1,
    window M = the first lag with M >= c tau(M) (Sokal's automatic windowing; emcee.autocorr.integrated_time, which the
    reference calls at fricc/__init__.py:9, without its chain-length check)."""
    import numpy as np

    x = np.asarray(x, dtype=np.float64).ravel()
    n = x.size
    if n < 2:
        raise ValueError("need at least two samples")
    nfft = 1 << (2 * n - 1).bit_length()
    f = np.fft.rfft(x - x.mean(), n=nfft)
    acf = np.fft.irfft(f * np.conjugate(f))[:n]
    if acf[0] == 0.0:
        return 1.0
    acf = acf / acf[0]
    taus = 2.0 * np.cumsum(acf) - 1.0
    ok = np.arange(n) >= c * taus
    window = int(np.argmax(ok)) if ok.any() else n - 1
    return float(taus[window])


def mcmc_std(vals):
    """fricc/__init__.py:8-11: (sqrt(tau_f / n * var(vals)), tau_f) -- the standard error of the mean of a correlated energy
    trace."""
    import numpy as np

    vals = np.asarray(vals, dtype=np.float64)
    tau_f = integrated_time(vals)
    return float(np.sqrt(tau_f / vals.size * np.var(vals))), tau_f
