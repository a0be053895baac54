import os
import sys

# the compiled reference's one_norm is an OpenMP reduction: one thread makes it bit-deterministic
os.environ.setdefault("OMP_NUM_THREADS", "1")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import pytest  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TERMS = {  # term -> (W axes, out axes, einsum with prefactor)   reference: fricc/tests/test_sparse_tensors.py:48-269
    "0101": ("oooo", "oovv", 1.0, "klij,klab->ijab"),
    "2323": ("vvvv", "oovv", 1.0, "abcd,ijcd->ijab"),
    "1302": ("voov", "oovv", 2.0, "akic,kjcb->ijab"),
    "1202": ("vovo", "oovv", -1.0, "akci,kjcb->ijab"),
    "1303": ("voov", "oovv", 1.0, "akic,kjbc->ijab"),
    "1203": ("vovo", "oovv", 1.0, "bkci,kjac->ijab"),
    "1323": ("ovov", "oooo", 1.0, "kcld,ijcd->klij"),
    "0112": ("ovov", "voov", -0.5, "ldkc,ilda->akic"),
    "0313": ("ovov", "voov", -0.5, "lckd,ilad->akic"),
    "0113": ("ovov", "voov", 1.0, "ldkc,ilad->akic"),
    "0312": ("ovov", "vovo", -0.5, "lckd,ilda->akci"),
}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle import oracle

    return oracle.port()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref); tests that need it are skipped where it was not built."""
    from oracle import oracle

    r = oracle.ref()
    if r is None:
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return r


@pytest.fixture(scope="session")
def golden_compression():
    return np.load(os.path.join(GOLDEN, "compression.npz"))


@pytest.fixture(scope="session")
def golden_contractions():
    return np.load(os.path.join(GOLDEN, "contractions.npz"))


@pytest.fixture(scope="session")
def rccsd():
    """The product module, on a GPU.  Fails (does not skip) if the CUDA library is missing: no silent fallback."""
    from fricc_b200 import _lib, py_rccsd

    _lib.require_cuda()
    return py_rccsd


def create_sparse_ndarray(t_shape, nnz, rng=np.random):
    """fricc/tests/test_sparse_tensors.py:8-15 (indices drawn with replacement)."""
    size = int(np.prod(t_shape))
    tensor = np.zeros(size, order="C")
    idx = rng.choice(np.arange(size), size=int(nnz))
    tensor[idx] = rng.rand(int(nnz))
    return tensor.reshape(t_shape)


def synth_t2(rng, n, kind):
    if kind == "heavy":
        return rng.standard_normal(n) * np.exp(-4.0 * rng.random(n))
    if kind == "heavier":
        return rng.standard_normal(n) * 10.0 ** (-6.0 * rng.random(n) ** 2)
    if kind == "uniform":
        return rng.random(n)
    if kind == "diff":
        return rng.random(n) - rng.random(n)
    raise ValueError(kind)
