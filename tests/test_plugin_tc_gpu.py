"""The tensor-core form through the PLUGIN CALL: `py_rccsd.contract_DTSpT` (= fricc_contract, the one symbol the
reference's drivers call, src/py_rccsd.cpp:50-51) picks the kernel itself and builds the int8 digit planes straight from
the 4-D operand it is handed (any strides, host or device), so a maintainer who only swaps the module gets the kernel the
sharded engine runs.  Checked against the oracle at the P6 tolerance (1e-10 relative) with the form forced on
("tc_mode" 2) and off (0), for every term, at the molecular shapes of BASELINE.json's configs (H2O
o=5,v=19; N2 o=7,v=53; benzene o=21,v=93 with the clamped m), at a shape that crosses many K blocks and more tiles
than there are SMs, with heavy-tailed W, and for the single-read plane builder against the two-pass one."""
import ctypes

import numpy as np
import pytest
from conftest import TERMS, synth_t2

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture()
def tc(rccsd):
    """contract_DTSpT with the tensor-core form forced; restores the cost model afterwards."""
    rccsd.set_option("tc_mode", 2)
    yield rccsd
    rccsd.set_option("tc_mode", 1)
    rccsd.set_option("tc_planes", 6)
    rccsd.set_option("tc_fused_planes", 1)
    rccsd.set_option("tc_plane_cache", 0)


def _fri(rccsd, rng, no, nv, m, method="systematic", kind="heavier"):
    n = no * no * nv * nv
    t2 = synth_t2(rng, n, kind).reshape(no, no, nv, nv)
    draws = rng.integers(0, 2**31 - 1, size=n).astype(np.int32)
    T = rccsd.SparseTensor4d(t2, t2.shape, int(m), "fri", method, False, draws=draws)
    flat, vals = T.export()
    return T, flat, vals, (no, no, nv, nv)


def _close(got, want, tol=TOL):
    return np.abs(got - want).max() <= tol * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("no,nv,m", [(5, 19, 1000), (5, 19, 4000), (7, 53, 50000)])
@pytest.mark.parametrize("method", ["systematic", "pivotal"])
def test_molecular_shapes_every_term_both_forms(tc, port, no, nv, m, method):
    """H2O / cc-pVDZ (o=5, v=19, N=9025: m=1e4 of the literal config is rejected by the reference, fricc/friccsd.py:69-72)
    and N2 / cc-pVTZ (o=7, v=53, N=137641: m=1e6 rejected) at the clamped m of SURVEY.md 8d."""
    rccsd = tc
    rng = np.random.default_rng(no * 1000 + nv + m)
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, m, method)
    ext = {"o": no, "v": nv}
    before = rccsd.get_stat("tc_contract_calls")
    for term, (wax, oax, _, _) in sorted(TERMS.items()):
        W = rng.random([ext[c] for c in wax]) - 0.3
        o0 = rng.random([ext[c] for c in oax])
        want = o0.copy()
        port.contract(term, W, flat, vals, dims, want)
        got = o0.copy()
        rccsd.contract_DTSpT(W, T, got, term)                 # tensor cores
        assert _close(got, want), (term, "tc")
        rccsd.set_option("tc_mode", 0)
        got2 = o0.copy()
        rccsd.contract_DTSpT(W, T, got2, term)                # fp64 SpMM
        rccsd.set_option("tc_mode", 2)
        assert _close(got2, want), (term, "spmm")
    assert rccsd.get_stat("tc_contract_calls") - before == len(TERMS)


def test_benzene_shape_against_the_compiled_reference(tc, ref):
    """benzene / cc-pVDZ shape (o=21, v=93, N=3 814 209; the literal m=5e6 is rejected by the reference) at m=1e6:
    both samplers' compressions bit-identical to the compiled reference, then the ladder and two O^3V^3 / O^4V^2 terms
    against the reference's own kernels (all host cores; its contractions give the same bits for any thread count)."""
    import os

    rccsd = tc
    no, nv, m = 21, 93, 1_000_000
    rng = np.random.default_rng(2193)
    shape = (no, no, nv, nv)
    n = int(np.prod(shape))
    t2 = synth_t2(rng, n, "heavier").reshape(shape)
    draws = rng.integers(0, 2**31 - 1, size=n + 8).astype(np.int32)
    for method in ("systematic", "pivotal"):
        ref.set_draws(draws)
        want = ref.fri_compression(t2, m, method, draws)
        T = rccsd.SparseTensor4d(t2, shape, m, "fri", method, False, draws=draws)
        flat, vals = T.export()
        assert np.array_equal(flat, want["idx"]) and np.array_equal(vals, want["vals"]), method
    ref.set_draws(draws)
    T_ref = ref.mod.SparseTensor4d(t2, list(shape), m, "fri", "pivotal", False)
    ref.clear_draws()
    ref.set_num_threads(os.cpu_count() or 1)
    try:
        ext = {"o": no, "v": nv}
        for term in ("2323", "1302", "0313", "0101"):
            wax, oax, _, _ = TERMS[term]
            W = rng.random([ext[c] for c in wax])
            want = np.zeros([ext[c] for c in oax])
            ref.mod.contract_DTSpT(W, T_ref, want, term)
            for mode in (2, 0):
                rccsd.set_option("tc_mode", mode)
                got = np.zeros_like(want)
                rccsd.contract_DTSpT(W, T, got, term)
                assert _close(got, want), (term, mode)
            rccsd.set_option("tc_mode", 2)
    finally:
        ref.set_num_threads(1)


def test_auto_mode_picks_the_tensor_cores_for_a_dense_fri_tensor(rccsd, port):
    """Default "tc_mode" 1: the cost model sends a 20 %-dense FRI tensor on a (289 x 4096) x (4096 x 4096) product to
    the tensor cores and a "largest" compression of the same size to the SpMM.  K = 4096 is 32 pipeline blocks, the 192
    output tiles are more than the 148 persistent CTAs (every CTA loops), and row sums reach ~800."""
    rng = np.random.default_rng(17)
    no, nv = 17, 64
    n = no * no * nv * nv
    # (a flat magnitude distribution: nothing is large enough to be kept exactly, so all entries are sampled ones --
    #  with the heavy-tailed generator most of a 20 % compression is the exact set D and the SpMM is the right choice)
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, n // 5, kind="diff")
    W = rng.random((nv, nv, nv, nv)) - 0.25
    want = np.zeros(dims)
    port.contract("2323", W, flat, vals, dims, want)
    tc0, sp0 = rccsd.get_stat("tc_contract_calls"), rccsd.get_stat("spmm_contract_calls")
    got = np.zeros(dims)
    rccsd.contract_DTSpT(W, T, got, "2323")
    assert rccsd.get_stat("tc_contract_calls") == tc0 + 1
    assert _close(got, want)
    t2 = rng.standard_normal(dims)
    Tl = rccsd.SparseTensor4d(t2, dims, n // 5, "largest")
    got = np.zeros(dims)
    rccsd.contract_DTSpT(W, Tl, got, "2323")
    assert rccsd.get_stat("spmm_contract_calls") == sp0 + 1


def test_strided_and_device_operands_through_the_plugin_call(tc, port):
    import torch

    rccsd = tc
    rng = np.random.default_rng(23)
    no, nv = 6, 14
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, 1500, "pivotal")
    # FRILCCSD's operands are transposed views (fricc/frilccsd.py:171-174)
    oooo = rng.random((no, no, no, no))
    W = oooo.transpose(0, 2, 1, 3)
    want = np.zeros(dims)
    port.contract("0101", np.ascontiguousarray(W), flat, vals, dims, want)
    got = np.zeros(dims)
    rccsd.contract_DTSpT(W, T, got, "0101")
    assert _close(got, want)
    # device tensors, the output a transposed view, an L-slice of the operand (what one rank of the sharded engine passes)
    Wv = rng.random((nv, nv, nv, nv))
    want = np.zeros(dims)
    port.contract("2323", Wv, flat, vals, dims, want)
    Wd = torch.from_numpy(Wv).cuda()
    outd = torch.zeros(dims, dtype=torch.float64, device="cuda")
    for a0, a1 in ((0, 5), (5, nv)):
        rccsd.contract_DTSpT(Wd[a0:a1], T, outd[:, :, a0:a1, :], "2323")
    assert _close(outd.cpu().numpy(), want)
    outT = torch.zeros(no, no, nv, nv, dtype=torch.float64, device="cuda").permute(1, 0, 3, 2)
    rccsd.contract_DTSpT(Wd, T, outT, "2323")
    assert _close(outT.cpu().numpy(), want)


@pytest.mark.parametrize("planes", [6, 7])
def test_heavy_tailed_operand(tc, port, planes):
    """One power-of-two scale per column means an element 10^-6 of its column's largest keeps 48 - 20 = 28 significant
    bits at 6 planes.  What the contract (P6) bounds is the OUTPUT: the dropped tail is <= 2^-48 of the column maximum per
    term, i.e. <= nnz_row * 2^-48 * max|W_col| * c absolute -- checked against sum |T||W| as well as against 1e-10."""
    rccsd = tc
    rccsd.set_option("tc_planes", planes)
    rng = np.random.default_rng(31)
    no, nv = 8, 24
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, 9000)
    W = (rng.random((nv, nv, nv, nv)) - 0.5) * 10.0 ** (-6.0 * rng.random((nv, nv, nv, nv)))
    want = np.zeros(dims)
    port.contract("2323", W, flat, vals, dims, want)
    scale = np.zeros(dims)
    port.contract("2323", np.abs(W), flat, np.abs(vals), dims, scale)
    got = np.zeros(dims)
    rccsd.contract_DTSpT(W, T, got, "2323")
    err = np.abs(got - want)
    assert err.max() <= TOL * np.abs(want).max()
    assert (err <= 2.0 ** (-8 * planes + 12) * np.maximum(scale, scale.max() * 1e-3)).all()


def _planes(L, check, W, term, l0, lc, S):
    import torch

    h = ctypes.c_void_p()
    i64 = lambda v: (ctypes.c_int64 * 4)(*[int(x) for x in v])
    check(L.fricc_wslices_create_view(ctypes.c_void_p(W.data_ptr()), i64(W.shape), i64(W.stride()), term.encode(), l0, lc, S,
                                      ctypes.byref(h)))
    nK = W.shape[2] * W.shape[3]
    rec = torch.empty(nK, lc, dtype=torch.float64, device="cuda")
    check(L.fricc_wslices_reconstruct(h, ctypes.c_void_p(rec.data_ptr())))
    check(L.fricc_wslices_destroy(h))
    return rec


@pytest.mark.parametrize("shape,l0,lc", [((2, 40, 64, 64), 3, 70),        # 4096-long runs: one CTA per column
                                         ((2, 24, 250, 240), 0, 48),      # 60 000-long runs: clusters of 4
                                         ((1, 9, 370, 372), 2, 7)])       # 137 640-long runs: clusters of 8, ragged tail
def test_single_read_plane_builder_equals_the_two_pass_builder(rccsd, shape, l0, lc):
    """The ladder's native W[a,b,c,d] through the cluster kernel (each column held once in distributed shared memory)
    gives the same planes as colmax + slice, and both reconstruct W to 2^-48 of each column's largest element."""
    import torch

    from fricc_b200 import _lib

    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(shape[2])
    W = (torch.rand(shape, generator=g, device="cuda", dtype=torch.float64) - 0.5)
    W *= torch.pow(10.0, -4.0 * torch.rand(shape[0], shape[1], 1, 1, generator=g, device="cuda", dtype=torch.float64))
    W[0, 1] = 0.0
    try:
        rccsd.set_option("tc_fused_planes", 1)
        fused = _planes(L, _lib.check, W, "2323", l0, lc, 6)
        rccsd.set_option("tc_fused_planes", 0)
        twop = _planes(L, _lib.check, W, "2323", l0, lc, 6)
    finally:
        rccsd.set_option("tc_fused_planes", 1)
    assert torch.equal(fused, twop)
    ref = W.reshape(shape[0] * shape[1], -1)[l0:l0 + lc].t()
    colmax = ref.abs().max(dim=0).values
    assert ((fused - ref).abs() <= colmax[None, :] * 2.0 ** -47 + 1e-300).all()


def test_plane_cache_follows_the_operand(tc, port):
    """"tc_plane_cache" 1: same device operand twice -> built once; rewritten in place -> rebuilt (fingerprint)."""
    import torch

    rccsd = tc
    rccsd.set_option("tc_plane_cache", 1)
    rng = np.random.default_rng(41)
    no, nv = 5, 12
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, 800)
    Wd = torch.from_numpy(rng.random((nv, nv, nv, nv))).cuda()
    for round_ in range(2):
        want = np.zeros(dims)
        port.contract("2323", Wd.cpu().numpy(), flat, vals, dims, want)
        for _ in range(2):
            out = torch.zeros(dims, dtype=torch.float64, device="cuda")
            rccsd.contract_DTSpT(Wd, T, out, "2323")
            assert _close(out.cpu().numpy(), want)
        Wd.mul_(-1.5).add_(0.1)      # same address, new values
    from fricc_b200 import _lib

    _lib.check(_lib.lib().fricc_plane_cache_clear())


def test_prepared_operand_is_used_once_and_only_by_the_tensor_core_form(tc, port):
    """py_rccsd.prepare_operand (fricc_prepare_operand): planes encoded ahead of the call on a second stream are what the
    next tensor-core call on that operand uses -- once, same bits; a call that takes the fp64 form drops them, so a later
    tensor-core call on a rewritten operand cannot see stale planes."""
    import torch

    from fricc_b200 import dist as fd

    rccsd = tc
    rng = np.random.default_rng(91)
    no, nv = 6, 20
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, 2500)
    W = torch.from_numpy(rng.random((7, nv, nv, nv)) - 0.45).cuda()          # a rank's a-range of Wvvvv
    fd.use_torch_stream()

    def ladder():
        out = torch.zeros(no, no, nv, nv, dtype=torch.float64, device="cuda")
        ov = out[:, :, 3:10, :]
        rccsd.contract_DTSpT(W, T, ov, "2323")
        return out

    plain = ladder()
    side = fd.SideStream(torch.device("cuda"))
    out = torch.zeros(no, no, nv, nv, dtype=torch.float64, device="cuda")
    ov = out[:, :, 3:10, :]
    h0 = rccsd.get_stat("prepared_plane_hits")
    side.run(lambda: rccsd.prepare_operand(W, ov, "2323"))
    side.join()
    rccsd.contract_DTSpT(W, T, ov, "2323")
    assert rccsd.get_stat("prepared_plane_hits") == h0 + 1
    assert torch.equal(out, plain)
    assert torch.equal(ladder(), plain) and rccsd.get_stat("prepared_plane_hits") == h0 + 1     # used once
    # prepared, then the fp64 form runs on it: dropped
    side.run(lambda: rccsd.prepare_operand(W, ov, "2323"))
    side.join()
    rccsd.set_option("tc_mode", 0)
    spmm = ladder()
    rccsd.set_option("tc_mode", 2)
    assert _close(spmm.cpu().numpy(), plain.cpu().numpy())
    W.mul_(-0.75).add_(0.02)
    fresh = ladder()
    assert rccsd.get_stat("prepared_plane_hits") == h0 + 1
    rccsd.set_option("tc_mode", 0)
    spmm2 = ladder()
    rccsd.set_option("tc_mode", 2)
    assert _close(fresh.cpu().numpy(), spmm2.cpu().numpy()) and not _close(fresh.cpu().numpy(), plain.cpu().numpy())


def test_constant_operand_planes_are_built_once_and_follow_the_declaration(tc, port):
    """py_rccsd.constant_operand (fricc_constant_operand): the ovov integrals of a CC driver never change, so the planes
    of their views are kept -- same bits as undeclared calls for every term that reads ovov, on the whole operand and on
    a narrowed view (a rank's slice); a declaration made again after rewriting the array picks up the new content, a
    withdrawn one goes back to building per call."""
    import torch

    rccsd = tc
    rng = np.random.default_rng(77)
    no, nv = 6, 13
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, 1500)
    ovov = torch.from_numpy(rng.random((no, nv, no, nv)) - 0.4).cuda()
    ext = {"o": no, "v": nv}
    terms = ("0112", "0113", "0313", "0312", "1323")

    def run(W):
        outs = {}
        for term in terms:
            out = torch.zeros([ext[c] for c in TERMS[term][1]], dtype=torch.float64, device="cuda")
            rccsd.contract_DTSpT(W, T, out, term)
            outs[term] = out
        return outs

    plain = run(ovov)
    rccsd.constant_operand(ovov, True)
    try:
        h0 = rccsd.get_stat("constant_plane_hits")
        first = run(ovov)                                   # builds (terms that read the same view share one build)
        h1 = rccsd.get_stat("constant_plane_hits")
        again = run(ovov)
        h2 = rccsd.get_stat("constant_plane_hits")
        assert h2 - h1 == len(terms) and h1 - h0 < len(terms)
        for term in terms:
            assert torch.equal(first[term], plain[term]) and torch.equal(again[term], plain[term]), term
            want = np.zeros(plain[term].shape)
            port.contract(term, ovov.cpu().numpy(), flat, vals, dims, want)
            assert _close(plain[term].cpu().numpy(), want), term
        # a rank's slice of the operand starts inside the declared range: its planes are kept too
        part = ovov.narrow(2, 2, 3)
        out_a = torch.zeros(nv, 3, nv, no, dtype=torch.float64, device="cuda")
        out_b = torch.zeros_like(out_a)
        rccsd.contract_DTSpT(part, T, out_a, "0312")
        h3 = rccsd.get_stat("constant_plane_hits")
        rccsd.contract_DTSpT(part, T, out_b, "0312")
        assert rccsd.get_stat("constant_plane_hits") == h3 + 1 and torch.equal(out_a, out_b)
        assert torch.equal(out_a, plain["0312"].narrow(1, 2, 3))
        # rewritten in place: the old planes would be wrong -- the caller declares again
        ovov.mul_(-1.25).add_(0.05)
        rccsd.constant_operand(ovov, True)
        fresh = run(ovov)
        rccsd.constant_operand(ovov, False)
        undeclared = run(ovov)
        for term in terms:
            assert torch.equal(fresh[term], undeclared[term]), term
            assert not torch.equal(fresh[term], plain[term]), term
        h4 = rccsd.get_stat("constant_plane_hits")
        run(ovov)
        assert rccsd.get_stat("constant_plane_hits") == h4     # withdrawn: built per call again
    finally:
        rccsd.constant_operand(ovov, False)


@pytest.mark.parametrize("shape,l0,lc,S", [((2, 40, 64, 64), 3, 70, 6), ((1, 9, 370, 372), 2, 7, 6), ((3, 20, 33, 34), 0, 60, 4),
                                           ((2, 30, 50, 50), 5, 50, 7)])
def test_predicted_update_equals_a_fresh_build(rccsd, shape, l0, lc, S):
    """fricc_wslices_update_view on planes that exist: digits cut with the previous exponents in one streaming pass, the
    columns whose exponent moved (here: scaled by 4, by 1/8, zeroed, made non-finite) cut again -- the same planes and
    scales as a fresh build of the new operand, and the count of corrected columns is reported."""
    import torch

    from fricc_b200 import _lib

    L, check = _lib.lib(), _lib.check
    i64 = lambda v: (ctypes.c_int64 * 4)(*[int(x) for x in v])
    g = torch.Generator(device="cuda").manual_seed(shape[2] + S)
    W = (torch.rand(shape, generator=g, device="cuda", dtype=torch.float64) - 0.5)
    W *= torch.pow(10.0, -3.0 * torch.rand(shape[0], shape[1], 1, 1, generator=g, device="cuda", dtype=torch.float64))
    nK = shape[2] * shape[3]
    h = ctypes.c_void_p()
    check(L.fricc_wslices_create_view(ctypes.c_void_p(W.data_ptr()), i64(W.shape), i64(W.stride()), b"2323", l0, lc, S, ctypes.byref(h)))
    n = ctypes.c_int64(7)
    check(L.fricc_wslices_redone(h, ctypes.byref(n)))
    assert n.value == -1                                  # a first build predicts nothing
    before = rccsd.get_stat("planes_predicted_builds")
    # new values everywhere (same magnitudes: exponents mostly stay), and a few columns that move
    W2 = W * (1.0 + 0.05 * (torch.rand(shape, generator=g, device="cuda", dtype=torch.float64) - 0.5))
    flat = W2.view(shape[0] * shape[1], -1)
    moved = {l0 + 1: 4.0, l0 + 2: 0.125, l0 + lc - 1: 0.0}
    for col, f in moved.items():
        flat[col] *= f
    flat[l0 + 4, 5] = float("inf")
    check(L.fricc_wslices_update_view(h, ctypes.c_void_p(W2.data_ptr()), i64(W2.shape), i64(W2.stride()), b"2323", l0))
    assert rccsd.get_stat("planes_predicted_builds") == before + 1
    check(L.fricc_wslices_redone(h, ctypes.byref(n)))
    assert 4 <= n.value <= max(4, lc // 2)                 # the four edited columns (+ the few whose maximum crossed a power of two)
    rec = torch.empty(nK, lc, dtype=torch.float64, device="cuda")
    check(L.fricc_wslices_reconstruct(h, ctypes.c_void_p(rec.data_ptr())))
    check(L.fricc_wslices_destroy(h))
    fresh = _planes(L, check, W2, "2323", l0, lc, S)
    assert torch.equal(torch.nan_to_num(rec, nan=-7.0), torch.nan_to_num(fresh, nan=-7.0))
    # and an update with the prediction switched off is the same thing
    try:
        rccsd.set_option("tc_predict", 0)
        h2 = ctypes.c_void_p()
        check(L.fricc_wslices_create_view(ctypes.c_void_p(W.data_ptr()), i64(W.shape), i64(W.stride()), b"2323", l0, lc, S, ctypes.byref(h2)))
        check(L.fricc_wslices_update_view(h2, ctypes.c_void_p(W2.data_ptr()), i64(W2.shape), i64(W2.stride()), b"2323", l0))
        rec2 = torch.empty(nK, lc, dtype=torch.float64, device="cuda")
        check(L.fricc_wslices_reconstruct(h2, ctypes.c_void_p(rec2.data_ptr())))
        check(L.fricc_wslices_destroy(h2))
    finally:
        rccsd.set_option("tc_predict", 1)
    assert torch.equal(torch.nan_to_num(rec2, nan=-7.0), torch.nan_to_num(fresh, nan=-7.0))


@pytest.mark.parametrize("ta,tb", [("0112", "0113"), ("0113", "0313"), ("1302", "1202")])
@pytest.mark.parametrize("no,nv,sliced", [(5, 19, False), (9, 30, True)])
def test_fused_pair_equals_the_two_calls(tc, port, ta, tb, no, nv, sliced):
    """contract_DTSpT_pair: the pairs the drivers issue back to back into one array, as ONE dense product (a weighted
    sign matrix for 0112 + 0113, planes of the combined operand for 0113 + 0313 and 1302 + 1202) -- against the oracle's
    two contractions, on whole operands and on the L-slice views a rank of the sharded step hands over."""
    import torch

    rccsd = tc
    rng = np.random.default_rng(no * 77 + nv + int(ta))
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, int(0.3 * no * no * nv * nv), kind="diff")
    ext = {"o": no, "v": nv}
    (wa_axes, o_axes, _, _), (wb_axes, ob_axes, _, _) = TERMS[ta], TERMS[tb]
    assert o_axes == ob_axes
    Wa = rng.random([ext[c] for c in wa_axes]) - 0.4
    # 0112 + 0113 and 0113 + 0313 read the same ovov block (the second pair through two different views of it)
    Wb = Wa if wa_axes == wb_axes else rng.random([ext[c] for c in wb_axes]) - 0.6
    out0 = rng.random([ext[c] for c in o_axes])
    want = out0.copy()
    port.contract(ta, Wa, flat, vals, dims, want)
    port.contract(tb, Wb, flat, vals, dims, want)
    da = torch.from_numpy(Wa).cuda()
    db = da if Wb is Wa else torch.from_numpy(Wb).cuda()
    got = torch.from_numpy(out0).cuda()
    before = rccsd.get_stat("fused_pairs")
    if not sliced:
        rccsd.contract_DTSpT_pair(da, db, T, got, ta, tb)
        calls = 1
    else:
        from fricc_b200 import dist as fd

        calls = 0
        for r in range(3):                           # three ranks' slices, one after the other on this GPU
            eng = fd.ViewEngine.__new__(fd.ViewEngine)
            eng.comm, eng.device, eng.arena, eng.side, eng.py = fd.Comm(r, 3), None, None, None, rccsd
            calls += eng.contract_pair(ta, tb, da, db, T, got) is not None
    assert rccsd.get_stat("fused_pairs") == before + calls
    assert _close(got.cpu().numpy(), want)
    # and the pair on host arrays is simply the two calls
    host = out0.copy()
    rccsd.contract_DTSpT_pair(Wa, Wb, T, host, ta, tb)
    assert _close(host, want)


def test_pair_that_cannot_fuse_falls_back_to_two_calls(tc, port):
    import torch

    rccsd = tc
    rng = np.random.default_rng(8)
    no, nv = 5, 12
    T, flat, vals, dims = _fri(rccsd, rng, no, nv, 2000, kind="diff")
    Wvoov, Wvovo = rng.random((nv, no, no, nv)), rng.random((nv, no, nv, no))
    want = np.zeros((no, no, nv, nv))
    port.contract("1303", Wvoov, flat, vals, dims, want)       # 1303 and 1203 write through different output maps
    port.contract("1203", Wvovo, flat, vals, dims, want)
    got = torch.zeros(no, no, nv, nv, dtype=torch.float64, device="cuda")
    before = rccsd.get_stat("fused_pairs")
    rccsd.contract_DTSpT_pair(torch.from_numpy(Wvoov).cuda(), torch.from_numpy(Wvovo).cuda(), T, got, "1303", "1203")
    assert rccsd.get_stat("fused_pairs") == before
    assert _close(got.cpu().numpy(), want)


def test_pair_with_a_mismatched_tensor_reports_the_plain_calls_error(tc):
    """A T whose extents do not fit term b axis by axis must get the error the two plain calls give, never a fused
    product with the wrong owner split (round-2 advice: the fused path checked term b's owner pair only by its product)."""
    import torch

    rccsd = tc
    rng = np.random.default_rng(81)
    no, nv = 6, 14
    W = torch.from_numpy(rng.random((no, nv, no, nv))).cuda()          # ovov, read by 0112 and 0113
    for dims in [(no, no, nv, nv + 1), (no, no, nv + 1, nv), (no + 1, no, nv, nv)]:
        x = rng.random(dims) - 0.5
        T = rccsd.SparseTensor4d(x, dims, int(0.6 * x.size), "fri", "systematic")
        out = torch.zeros(nv, no, no, nv, dtype=torch.float64, device="cuda")
        before = rccsd.get_stat("fused_pairs")
        with pytest.raises(ValueError):
            rccsd.contract_DTSpT_pair(W, W, T, out, "0112", "0113")
        assert rccsd.get_stat("fused_pairs") == before
        plain = torch.zeros_like(out)
        with pytest.raises(ValueError):
            rccsd.contract_DTSpT(W, T, plain, "0112")
            rccsd.contract_DTSpT(W, T, plain, "0113")


def test_pivotal_with_fast_contractions_warns_once(tc):
    """The default kernels meet 1e-10 per call but not equal-seed 1e-8 energies with the pivotal sampler:
    the mirror says so once per process, and not at all in parity mode or for the systematic sampler."""
    import warnings

    from fricc_b200 import py_rccsd

    rccsd = tc
    rng = np.random.default_rng(4)
    no, nv = 4, 9
    x = rng.random((no, no, nv, nv)) - 0.5
    W = rng.random((nv, nv, nv, nv))
    out = np.zeros((no, no, nv, nv))
    Ts = rccsd.SparseTensor4d(x, x.shape, 300, "fri", "systematic")
    Tp = rccsd.SparseTensor4d(x, x.shape, 300, "fri", "pivotal")
    py_rccsd._warned_pivotal = False
    try:
        with warnings.catch_warnings(record=True) as rec:
            warnings.simplefilter("always")
            rccsd.contract_DTSpT(W, Ts, out, "2323")
            rccsd.set_option("spmm_exact_order", 1)
            rccsd.contract_DTSpT(W, Tp, out, "2323")
            rccsd.set_option("spmm_exact_order", 0)
            assert not [w for w in rec if "pivotal" in str(w.message)]
            rccsd.contract_DTSpT(W, Tp, out, "2323")
            rccsd.contract_DTSpT(W, Tp, out, "2323")
            assert len([w for w in rec if "spmm_exact_order" in str(w.message)]) == 1
    finally:
        rccsd.set_option("spmm_exact_order", 0)
