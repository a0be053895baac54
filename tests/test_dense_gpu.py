"""Tensor-core form of the contractions (fricc_b200/csrc/fricc_dense.cu), through the C-ABI:
  * the 8-bit digit-plane expansion of W is error-free up to its stated tail (fricc_wslices_reconstruct),
  * sign matrix x planes (tcgen05.mma.kind::i8) + fp64 remainder == the fp64 SpMM == the oracle, <= 1e-10 relative
    (the P6 tolerance) for every term, on FRI-compressed tensors, at ragged shapes (rows % 128, columns % 64,
    contracted pair % 128 all non-zero), onto non-zero outputs,
  * "largest" compressions (no common magnitude) fall through to the fp64 SpMM.
"""
import ctypes

import numpy as np
import pytest

from conftest import TERMS, synth_t2

pytestmark = pytest.mark.gpu

TOL = 1e-10


def _fri_tensor(rccsd, rng, no, nv, frac, method="systematic"):
    n = no * no * nv * nv
    t2 = synth_t2(rng, n, "heavier").reshape(no, no, nv, nv)
    m = max(2, int(frac * n))
    draws = rng.integers(0, 2**31 - 1, size=n).astype(np.int32)
    T = rccsd.SparseTensor4d(t2, t2.shape, m, "fri", method, False, draws=draws)
    flat, vals = T.export()
    return T, flat, vals, (no, no, nv, nv)


@pytest.mark.parametrize("n_planes", [1, 3, 6, 7])
def test_digit_planes_are_error_free(rccsd, n_planes):
    import torch

    from fricc_b200 import _lib

    L = _lib.lib()
    rng = np.random.default_rng(n_planes)
    nK, ld, lcnt = 300, 200, 150
    W = rng.standard_normal((nK, ld)) * 10.0 ** rng.uniform(-8, 3, size=(1, ld))   # wide range of column scales
    W[:, 5] = 0.0
    W[7, :] = 0.0
    Wd = torch.from_numpy(W).cuda()
    h = ctypes.c_void_p()
    _lib.check(L.fricc_wslices_create(ctypes.c_void_p(Wd.data_ptr()), ld, nK, lcnt, n_planes, ctypes.byref(h)))
    rec = torch.empty(nK, lcnt, dtype=torch.float64, device="cuda")
    _lib.check(L.fricc_wslices_reconstruct(h, ctypes.c_void_p(rec.data_ptr())))
    _lib.check(L.fricc_wslices_destroy(h))
    rec = rec.cpu().numpy()
    colmax = np.abs(W[:, :lcnt]).max(axis=0)
    e = np.where(colmax > 0, np.frexp(colmax)[1], 0)
    # one rounding to the 8S-bit grid: |tail| <= 2^(e_l - 8 S); the fp64 recombination itself rounds at 2^(e_l - 52);
    # an element within half a grid step of +2^e_l is clamped to the top of the grid (one more half step)
    bound = np.maximum(np.ldexp(1.0, e - 8 * n_planes), np.ldexp(1.0, e - 51))
    err = np.abs(rec - W[:, :lcnt])
    clamped = W[:, :lcnt] >= np.ldexp(1.0, e)[None, :] * (1.0 - 2.0 ** (-8 * n_planes))
    assert (err <= bound[None, :] * np.where(clamped, 2.0, 1.0)).all()
    assert clamped.sum() <= 2


@pytest.mark.parametrize("term", sorted(TERMS))
@pytest.mark.parametrize("no,nv,frac", [(3, 7, 0.5), (6, 21, 0.2), (13, 17, 0.3)])
def test_tc_equals_fp64_and_oracle(rccsd, port, term, no, nv, frac):
    import torch

    from fricc_b200 import dist as fd

    rng = np.random.default_rng(1000 * no + nv)
    T, flat, vals, dims = _fri_tensor(rccsd, rng, no, nv, frac)
    ext = {"o": no, "v": nv}
    waxes, oaxes, _, _ = TERMS[term]
    W = rng.random([ext[c] for c in waxes]) - 0.3
    out0 = rng.random([ext[c] for c in oaxes])
    want = out0.copy()
    port.contract(term, W, flat, vals, dims, want)

    dev = torch.device("cuda", 0)
    Wd = torch.from_numpy(W).to(dev)
    world = 2
    for n_planes in (6,):
        outd = torch.from_numpy(out0).to(dev)
        used = 0
        for g in range(world):
            st = fd.ShardedTerm(term, no, nv, fd.Comm(g, world), dev)
            st.permute_from(Wd)
            st.build_planes(n_planes)
            st.tc_min_fraction = 0.0     # always take the tensor-core form (the sign matrix may be small here)
            used += int(st.use_tc(T))
            st.local_into(T, outd)
            st.drop_planes()
        assert used == world
        got = outd.cpu().numpy()
        assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max())


def test_tc_panel_form_and_shared_panels(rccsd, port):
    """out_panel form (multi-GPU layout) + accumulate; two terms adding into one panel."""
    import torch

    from fricc_b200 import dist as fd

    rng = np.random.default_rng(3)
    no, nv, world = 5, 19, 3
    T, flat, vals, dims = _fri_tensor(rccsd, rng, no, nv, 0.3, "pivotal")
    dev = torch.device("cuda", 0)
    for term in ("2323", "1302", "0313"):
        ext = {"o": no, "v": nv}
        waxes, oaxes, _, _ = TERMS[term]
        W = rng.random([ext[c] for c in waxes])
        out0 = rng.random([ext[c] for c in oaxes])
        want = out0.copy()
        port.contract(term, W, flat, vals, dims, want)
        Wd = torch.from_numpy(W).to(dev)
        ranks = [fd.ShardedTerm(term, no, nv, fd.Comm(g, world), dev) for g in range(world)]
        nR, lc = ranks[0].nR, ranks[0].lc
        gathered = torch.zeros(world * nR * lc, dtype=torch.float64, device=dev)
        for g, st in enumerate(ranks):
            st.permute_from(Wd)
            st.build_planes(6)
            st.tc_min_fraction = 0.0
            gathered[g * nR * lc:(g + 1) * nR * lc].copy_(st.local(T))
        outd = torch.from_numpy(out0).to(dev)
        fd.check(ranks[0].L.fricc_accumulate_panels(fd._ptr(gathered), world, lc, T._h, term.encode(), fd._ptr(outd),
                                                    fd.i64x4(outd.shape), fd.i64x4(outd.stride())))
        got = outd.cpu().numpy()
        assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max()), term


def test_largest_compression_has_no_sign_matrix(rccsd, port):
    import torch

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd

    rng = np.random.default_rng(11)
    no, nv = 4, 12
    t2 = rng.standard_normal((no, no, nv, nv))
    T = rccsd.SparseTensor4d(t2, t2.shape, 500, "largest")
    f = ctypes.c_double(-1.0)
    _lib.check(_lib.lib().fricc_st_sign_fraction(T._h, b"2323", ctypes.byref(f)))
    assert f.value <= 2.0 / 500
    flat, vals = T.export()
    W = rng.random((nv, nv, nv, nv))
    want = np.zeros((no, no, nv, nv))
    port.contract("2323", W, flat, vals, t2.shape, want)
    dev = torch.device("cuda", 0)
    st = fd.ShardedTerm("2323", no, nv, fd.Comm(0, 1), dev)
    st.permute_from(torch.from_numpy(W).to(dev))
    st.build_planes(6)
    assert not st.use_tc(T)
    st.tc_min_fraction = 0.0      # forced: everything goes through the remainder, still correct
    outd = torch.zeros(no, no, nv, nv, dtype=torch.float64, device=dev)
    st.local_into(T, outd)
    assert np.abs(outd.cpu().numpy() - want).max() <= TOL * np.abs(want).max()


def test_tc_accuracy_follows_the_plane_count(rccsd, port):
    """error ~ 2^(-8 S): each extra plane buys 8 bits; S = 6 is at fp64 rounding level."""
    import torch

    from fricc_b200 import dist as fd

    rng = np.random.default_rng(5)
    no, nv = 6, 24
    T, flat, vals, dims = _fri_tensor(rccsd, rng, no, nv, 0.25)
    W = rng.random((nv, nv, nv, nv))
    want = np.zeros((no, no, nv, nv))
    port.contract("2323", W, flat, vals, dims, want)
    dev = torch.device("cuda", 0)
    st = fd.ShardedTerm("2323", no, nv, fd.Comm(0, 1), dev)
    st.permute_from(torch.from_numpy(W).to(dev))
    st.tc_min_fraction = 0.0
    errs = {}
    for S in (2, 3, 5, 6):
        st.build_planes(S)
        outd = torch.zeros(no, no, nv, nv, dtype=torch.float64, device=dev)
        st.local_into(T, outd)
        errs[S] = np.abs(outd.cpu().numpy() - want).max() / np.abs(want).max()
    assert errs[2] > errs[3] > errs[5] >= errs[6] * 0.5
    assert errs[3] < 1e-5 and errs[5] <= TOL and errs[6] <= 1e-13


def test_false_provenance_hint_poisons_instead_of_lying(rccsd, port):
    """fricc_st_set_n_exact promises "everything behind the first n_exact entries is +-c at distinct positions"; when
    that is false the result must be NaN, not a plausible wrong number."""
    import torch

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd

    rng = np.random.default_rng(2)
    no, nv = 4, 10
    n = no * no * nv * nv
    flat = rng.choice(n, 300, replace=False).astype(np.int64)
    vals = rng.standard_normal(300)                     # not a common magnitude
    T = rccsd.SparseTensor4d.from_coo(flat, vals, (no, no, nv, nv))
    dev = torch.device("cuda", 0)
    st = fd.ShardedTerm("2323", no, nv, fd.Comm(0, 1), dev)
    st.permute_from(torch.rand(nv, nv, nv, nv, dtype=torch.float64, device=dev))
    st.build_planes(6)
    st.tc_min_fraction = 0.0
    # honest (unknown provenance): everything is classified, the result is right
    outd = torch.zeros(no, no, nv, nv, dtype=torch.float64, device=dev)
    st.local_into(T, outd)
    assert torch.isfinite(outd).all()
    # false hint on a fresh handle
    T2 = rccsd.SparseTensor4d.from_coo(flat, vals, (no, no, nv, nv))
    _lib.check(_lib.lib().fricc_st_set_n_exact(T2._h, 0))
    outd2 = torch.zeros(no, no, nv, nv, dtype=torch.float64, device=dev)
    st.local_into(T2, outd2)
    assert torch.isnan(outd2).any()
    # a position sampled twice behind a true-magnitude hint is NOT an error: the sign matrix carries the multiplicity
    # (the systematic sampler's unmatched samples land on flat index 0 more than once, fri_utils.cpp:132,149-157)
    fdup = np.concatenate([flat[:50], flat[:1], flat[:1]])
    vdup = np.full(52, 0.25)
    T3 = rccsd.SparseTensor4d.from_coo(fdup, vdup, (no, no, nv, nv))
    _lib.check(_lib.lib().fricc_st_set_n_exact(T3._h, 0))
    outd3 = torch.zeros(no, no, nv, nv, dtype=torch.float64, device=dev)
    st.local_into(T3, outd3)
    want = np.zeros((no, no, nv, nv))
    port.contract("2323", st.panel.cpu().numpy().reshape(nv, nv, nv, nv).transpose(2, 3, 0, 1), fdup, vdup,
                  (no, no, nv, nv), want)
    assert np.abs(outd3.cpu().numpy() - want).max() <= TOL * np.abs(want).max()


@pytest.mark.parametrize("term,no,nv", [("2323", 13, 17), ("1302", 6, 40), ("0312", 18, 23)])
def test_pair_and_one_cta_kernels_agree_bitwise(rccsd, port, term, no, nv):
    """tc_sign_gemm_pair_kernel (cta_group::2, 256-row tiles, soft lockstep) against tc_sign_gemm_kernel: the integer
    accumulation is exact and the epilogue arithmetic is the same code, so the outputs must be the same BITS; both within
    1e-10 of the oracle.  Shapes: rows % 256 in (128, 256) and in (0, 128), several row tiles, ragged column tiles."""
    rng = np.random.default_rng(no * 100 + nv)
    T, flat, vals, dims = _fri_tensor(rccsd, rng, no, nv, 0.3)
    ext = {"o": no, "v": nv}
    waxes, oaxes, _, _ = TERMS[term]
    W = rng.random([ext[c] for c in waxes]) - 0.4
    out0 = rng.random([ext[c] for c in oaxes])
    want = out0.copy()
    port.contract(term, W, flat, vals, dims, want)
    got = {}
    rccsd.set_option("tc_mode", 2)
    try:
        for pair, lockstep in ((0, 0), (1, 0), (1, 1)):
            rccsd.set_option("tc_pair", pair)
            rccsd.set_option("tc_lockstep", lockstep)
            out = out0.copy()
            rccsd.contract_DTSpT(W, T, out, term)
            got[(pair, lockstep)] = out
    finally:
        rccsd.set_option("tc_mode", 1)
        rccsd.set_option("tc_pair", 1)
        rccsd.set_option("tc_lockstep", 1)
    for k, out in got.items():
        assert np.abs(out - want).max() <= TOL * max(1.0, np.abs(want).max()), k
    assert np.array_equal(got[(0, 0)], got[(1, 0)]) and np.array_equal(got[(1, 0)], got[(1, 1)])
