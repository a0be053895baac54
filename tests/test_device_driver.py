"""fricc_b200/friccsd.py (the device-resident update_amps / kernel loop) against its numpy twin
tests/rccsd_driver.py (= the reference's update_amps, fricc/friccsd.py:41-297).

CPU part (torch on the CPU, dense equations -- no native module involved): one update and a short run agree to 1e-12;
the two-electron case reproduces exact diagonalisation; mcmc_std against a direct evaluation of its definition.
GPU part: the same run with every contraction class compressed, our plugin as the native module on both sides --
numpy driver with host arrays vs the device driver with device tensors -- at equal m and seed."""
import numpy as np
import pytest
import torch

import rccsd_driver as D

from fricc_b200 import friccsd as F


def _eris(seed, norb, nocc, strength=0.3):
    rng = np.random.default_rng(seed)
    h, g = D.random_hamiltonian(rng, norb, strength=strength)
    c, _, _ = D.rhf(h, g, nocc, iters=2000, tol=1e-15)
    return D.Eris(h, g, c, nocc), h, g


@pytest.mark.parametrize("as_reference", [True, False])
def test_dense_update_matches_the_numpy_twin(as_reference):
    er, _, _ = _eris(3, 10, 3)
    de = F.DeviceEris.from_host(er, "cpu")
    t1, t2 = D.init_amps(er)
    d1, d2 = F.init_amps(de)
    assert np.allclose(d1.numpy(), t1, atol=1e-15) and np.allclose(d2.numpy(), t2, atol=1e-15)
    for _ in range(3):
        t1, t2 = D.update_amps(t1, t2, er, as_reference=as_reference)
        d1, d2 = F.update_amps(d1, d2, de, as_reference=as_reference)
        assert np.abs(d1.numpy() - t1).max() <= 1e-12 and np.abs(d2.numpy() - t2).max() <= 1e-12
        assert abs(F.energy(d1, d2, de) - D.energy(t1, t2, er)) <= 1e-12


def test_two_electrons_are_exact_on_the_device_driver():
    er, h, g = _eris(8, 7, 1, strength=0.4)
    de = F.DeviceEris.from_host(er, "cpu")
    res = F.kernel(de, max_cycle=300, tol=1e-13, tolnormt=1e-11, as_reference=False)
    _, _, e_hf = D.rhf(h, g, 1, iters=2000, tol=1e-15)
    assert res.conv
    assert abs(e_hf + res.e_corr - D.fci_two_electrons(h, g)) <= 1e-9
    conv, e_corr, t1, t2 = res                       # the reference's return tuple
    assert conv and e_corr == res.energies[-1] and t1.shape == (1, 6)


def test_kernel_exits_and_average():
    er, _, _ = _eris(5, 9, 3)
    de = F.DeviceEris.from_host(er, "cpu")
    res = F.kernel(de, max_cycle=6, tol=0.0, tolnormt=0.0, average_from=4)
    assert not res.conv and len(res.energies) == 7
    t1a, t2a, ea = res.average
    assert np.isfinite(ea) and t2a.shape == res.t2.shape
    es, t1, t2 = F.run(de, 6)
    assert es == res.energies and torch.equal(t2, res.t2)
    er.fock = er.fock * np.nan                         # NaN exit (:458-461)
    bad = F.kernel(F.DeviceEris.from_host(er, "cpu"), max_cycle=5)
    assert not bad.conv and len(bad.energies) == 2


def test_mcmc_std_follows_its_definition():
    rng = np.random.default_rng(0)
    x = np.zeros(4000)
    for i in range(1, x.size):                         # AR(1), rho = 0.8: tau = (1 + rho) / (1 - rho) = 9
        x[i] = 0.8 * x[i - 1] + rng.standard_normal()
    err, tau = F.mcmc_std(x)
    assert 6.0 < tau < 13.0
    assert abs(err - np.sqrt(tau / x.size * np.var(x))) <= 1e-15
    err_w, tau_w = F.mcmc_std(rng.standard_normal(4000))
    assert 0.7 < tau_w < 1.4                           # white noise: tau = 1


@pytest.mark.gpu
@pytest.mark.parametrize("sampling", ["systematic", "pivotal"])
def test_device_driver_matches_the_host_driver_at_equal_m_and_seed(rccsd, sampling):
    """Every class compressed.  The dense pieces are cuBLAS here and numpy there (different summation orders, ~1e-15), so
    what is promised is what the fast kernels promise (tests/test_rccsd_energy_gpu.py): within 1e-8 for the systematic
    sampler; for the pivotal sampler the trajectory is another draw of the same estimator and is compared loosely."""

    class Ours:
        SparseTensor4d = rccsd.SparseTensor4d
        contract_DTSpT = staticmethod(rccsd.contract_DTSpT)

    er, _, _ = _eris(11, 14, 4)
    m = max(1, int(0.1 * er.ovov.size))
    rccsd.srand(99)
    e_host, _, t2_host = D.run(er, 4, Ours, m, D.ALL_COMPRESSED, "fri", sampling)
    de = F.DeviceEris.from_host(er, "cuda")
    rccsd.srand(99)
    e_dev, t1_dev, t2_dev = F.run(de, 4, m, F.ALL_COMPRESSED, "fri", sampling)
    assert t2_dev.is_cuda
    tol = 1e-8 if sampling == "systematic" else 5e-3
    assert np.abs(np.array(e_dev) - np.array(e_host)).max() <= tol
    if sampling == "systematic":
        assert np.abs(t2_dev.cpu().numpy() - t2_host).max() <= 1e-8


@pytest.mark.gpu
def test_device_kernel_keeps_everything_on_the_gpu(rccsd):
    er, _, _ = _eris(4, 12, 4)
    de = F.DeviceEris.from_host(er, "cuda")
    m = er.ovov.size // 5
    before = rccsd.get_stat("tc_contract_calls") + rccsd.get_stat("spmm_contract_calls")
    res = F.kernel(de, max_cycle=3, tol=0.0, tolnormt=0.0, m_keep=m, compressed=F.ALL_COMPRESSED, average_from=2)
    assert rccsd.get_stat("tc_contract_calls") + rccsd.get_stat("spmm_contract_calls") - before == 3 * 11
    assert res.t1.is_cuda and res.t2.is_cuda and res.average[1].is_cuda and np.isfinite(res.energies).all()
