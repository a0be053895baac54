"""GPU parity of the 11 contraction terms -- the reference's own tests (fricc/tests/test_sparse_tensors.py,
test_sparse_tensors_large.py) re-run against fricc_b200.py_rccsd, plus the oracle at 1e-10 relative (P6)."""
import numpy as np
import pytest
from conftest import TERMS, create_sparse_ndarray

pytestmark = pytest.mark.gpu
npt = np.testing


def rel_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("no,nv,frac", [(4, 8, 0.1), (10, 15, 0.001), (100, 2, 0.001)])
def test_sparse_init(rccsd, no, nv, frac):
    """fricc/tests/test_sparse_tensors.py:18-45"""
    rng = np.random.RandomState(20)
    a = np.ascontiguousarray(rng.rand(no, no, nv, nv))
    m = int(a.size * frac)
    t = rccsd.SparseTensor4d(a, a.shape, m, "largest")
    for mi in range(m):
        idx, value = t.get_element(mi)
        npt.assert_equal(value, a[tuple(idx)])


SHAPES = [(4, 8, 0.1), (5, 10, 0.01), (10, 20, 0.001)]


@pytest.mark.parametrize("term", sorted(TERMS))
@pytest.mark.parametrize("no,nv,frac", SHAPES)
def test_term_vs_einsum_and_oracle(rccsd, port, term, no, nv, frac):
    wax, oax, scale, es = TERMS[term]
    rng = np.random.RandomState(20 + no)
    ext = {"o": no, "v": nv}
    shape = (no, no, nv, nv)
    W = np.ascontiguousarray(rng.rand(*[ext[c] for c in wax]))
    t2 = create_sparse_ndarray(shape, int(np.prod(shape) * frac), rng)
    m = int(t2.size * frac)
    T = rccsd.SparseTensor4d(t2, t2.shape, m, "largest")
    want = scale * np.einsum(es, W, t2, order="C")
    got = np.zeros([ext[c] for c in oax], order="C")
    assert rccsd.contract_DTSpT(W, T, got, term) is None
    npt.assert_almost_equal(got, want)  # the reference's own tolerance
    # oracle, accumulating onto a non-zero output, 1e-10 relative
    flat, vals = T.export()
    o0 = rng.rand(*[ext[c] for c in oax])
    o_gpu, o_cpu = o0.copy(), o0.copy()
    rccsd.contract_DTSpT(W, T, o_gpu, term)
    port.contract(term, W, flat, vals, shape, o_cpu)
    assert rel_err(o_gpu, o_cpu) <= 1e-10


@pytest.mark.parametrize("no,nv,frac", [(20, 40, 1e-4), (50, 90, 1e-6)])
def test_2323_large(rccsd, no, nv, frac):
    """fricc/tests/test_sparse_tensors_large.py:18-42"""
    rng = np.random.RandomState(20)
    w = rng.rand(nv, nv, nv, nv)
    shape = (no, no, nv, nv)
    t2 = create_sparse_ndarray(shape, int(np.prod(shape) * frac), rng)
    m = int(t2.size * frac)
    T = rccsd.SparseTensor4d(t2, t2.shape, m, "largest")
    want = np.einsum("abcd,ijcd->ijab", w, t2, order="C")
    got = np.zeros(t2.shape, order="C")
    rccsd.contract_DTSpT(w, T, got, "2323")
    npt.assert_almost_equal(got, want)


def test_golden_contractions(rccsd, golden_contractions):
    """Outputs recorded from the compiled reference (tests/golden/make_golden.py)."""
    g = golden_contractions
    for key in g["names"]:
        t2 = g[f"{key}/t2"]
        T = rccsd.SparseTensor4d.from_coo(g[f"{key}/flat"], g[f"{key}/vals"], t2.shape)
        for term in TERMS:
            out = g[f"{key}/{term}/out0"].copy()
            rccsd.contract_DTSpT(g[f"{key}/{term}/W"], T, out, term)
            assert rel_err(out, g[f"{key}/{term}/out1"]) <= 1e-10, (key, term)


def test_strided_operands(rccsd, port):
    """W and the output may be arbitrary strided views (pybind11 unchecked accessors honour numpy strides;
    FRILCCSD passes transposed views, fricc/frilccsd.py:171-174)."""
    rng = np.random.RandomState(3)
    no, nv = 5, 9
    shape = (no, no, nv, nv)
    t2 = create_sparse_ndarray(shape, 300, rng)
    T = rccsd.SparseTensor4d(t2, shape, 250, "largest")
    flat, vals = T.export()
    # 0101 with W = oooo.transpose(0,2,1,3) (a view), out a transposed view too
    base = rng.rand(no, no, no, no)
    W = base.transpose(0, 2, 1, 3)
    obase = rng.rand(no, no, nv, nv)
    o_gpu = obase.copy().transpose(1, 0, 3, 2)
    o_cpu = obase.copy().transpose(1, 0, 3, 2)
    rccsd.contract_DTSpT(W, T, o_gpu, "0101")
    port.contract("0101", W, flat, vals, shape, o_cpu)
    assert rel_err(o_gpu, o_cpu) <= 1e-10
    # 2323 with a sliced (non-compact) W
    big = rng.rand(nv, nv, nv, 2 * nv)
    W = big[:, :, :, ::2]
    o_gpu = np.zeros(shape)
    o_cpu = np.zeros(shape)
    rccsd.contract_DTSpT(W, T, o_gpu, "2323")
    port.contract("2323", W, flat, vals, shape, o_cpu)
    assert rel_err(o_gpu, o_cpu) <= 1e-10


def test_experimental_variants_and_errors(rccsd):
    rng = np.random.RandomState(5)
    no, nv = 4, 6
    shape = (no, no, nv, nv)
    t2 = create_sparse_ndarray(shape, 60, rng)
    T = rccsd.SparseTensor4d(t2, shape, 50, "largest")
    W = rng.rand(nv, nv, nv, nv)
    a, b, c = np.zeros(shape), np.zeros(shape), np.zeros(shape)
    rccsd.contract_DTSpT(W, T, a, "2323")
    rccsd.contract_DTSpT_experimental(W, T, b, "taskloop")
    rccsd.contract_DTSpT_experimental(W, T, c, "atomic")
    assert np.array_equal(a, b) and np.array_equal(a, c)
    with pytest.raises(ValueError):
        rccsd.contract_DTSpT(W, T, a, "9999")  # reference: "CASE NOT IMPLEMENTED" + exit
    with pytest.raises(TypeError):
        rccsd.contract_DTSpT(W.astype(np.float32), T, a, "2323")  # noconvert
    with pytest.raises(ValueError):
        rccsd.contract_DTSpT(W[0], T, a, "2323")  # ndim


def test_device_resident_operands(rccsd, port):
    """torch CUDA tensors are used in place (no host copies)."""
    import torch

    rng = np.random.RandomState(9)
    no, nv = 6, 12
    shape = (no, no, nv, nv)
    t2 = create_sparse_ndarray(shape, 500, rng)
    t2_d = torch.from_numpy(t2).cuda()
    T = rccsd.SparseTensor4d(t2_d, shape, 400, "largest")
    flat, vals = T.export()
    W = rng.rand(nv, nv, nv, nv)
    out_d = torch.zeros(shape, dtype=torch.float64, device="cuda")
    rccsd.contract_DTSpT(torch.from_numpy(W).cuda(), T, out_d, "2323")
    o_cpu = np.zeros(shape)
    port.contract("2323", W, flat, vals, shape, o_cpu)
    assert rel_err(out_d.cpu().numpy(), o_cpu) <= 1e-10
