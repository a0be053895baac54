"""The column-sharded contraction path on ONE GPU: the G ranks of a box are emulated as G column shards
processed one after another (never as concurrent processes), through the same C-ABI calls fricc_b200/dist.py makes:
fricc_permute_w -> fricc_contract_panel -> [gather] -> fricc_accumulate_panels, and the single-rank form
fricc_contract_panel_into.  Checked against fricc_contract on the same inputs and against the oracle (<= 1e-10 relative,
the P6 tolerance)."""
import numpy as np
import pytest

from conftest import TERMS

pytestmark = pytest.mark.gpu

TOL = 1e-10  # P6: contractions within 1e-10 relative (max-norm)


def _setup(rccsd, rng, no, nv, frac):
    import torch

    dims = (no, no, nv, nv)
    n = no * no * nv * nv
    m = max(1, int(frac * n))
    flat = rng.choice(n, m, replace=False).astype(np.int64)
    vals = rng.standard_normal(m)
    T = rccsd.SparseTensor4d.from_coo(flat, vals, dims)
    return torch, dims, flat, vals, T


@pytest.mark.parametrize("term", sorted(TERMS))
@pytest.mark.parametrize("no,nv,world", [(3, 5, 2), (5, 9, 3), (6, 37, 2), (7, 40, 4)])
def test_sharded_equals_direct(rccsd, port, term, no, nv, world):
    from fricc_b200 import dist as fd

    rng = np.random.default_rng(100 * no + nv + world)
    torch, dims, flat, vals, T = _setup(rccsd, rng, no, nv, 0.3)
    ext = {"o": no, "v": nv}
    waxes, oaxes, _, _ = TERMS[term]
    W = rng.random([ext[c] for c in waxes])
    out0 = rng.random([ext[c] for c in oaxes])       # accumulate onto a non-zero output, as the reference does
    want = out0.copy()
    port.contract(term, W, flat, vals, dims, want)

    dev = torch.device("cuda", 0)
    Wd = torch.from_numpy(W).to(dev)
    ranks = [fd.ShardedTerm(term, no, nv, fd.Comm(g, world), dev) for g in range(world)]
    nR, lc = ranks[0].nR, ranks[0].lc
    gathered = torch.zeros(world * nR * lc, dtype=torch.float64, device=dev)
    for g, st in enumerate(ranks):
        st.permute_from(Wd)
        gathered[g * nR * lc:(g + 1) * nR * lc].copy_(st.local(T))
    outd = torch.from_numpy(out0).to(dev)
    fd.check(ranks[0].L.fricc_accumulate_panels(fd._ptr(gathered), world, lc, T._h, term.encode(), fd._ptr(outd),
                                                fd.i64x4(outd.shape), fd.i64x4(outd.stride())))
    got = outd.cpu().numpy()
    assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max())

    # single-rank form: every shard accumulated straight into the 4-D output
    outd2 = torch.from_numpy(out0).to(dev)
    for st in ranks:
        st.local_into(T, outd2)
    got2 = outd2.cpu().numpy()
    assert np.abs(got2 - want).max() <= TOL * max(1.0, np.abs(want).max())


def test_accumulate_into_strided_output(rccsd, port):
    """out is a non-contiguous view: the tiled transposing kernels must not be chosen on the wrong stride."""
    from fricc_b200 import dist as fd

    rng = np.random.default_rng(7)
    no, nv, world = 4, 33, 2
    torch, dims, flat, vals, T = _setup(rccsd, rng, no, nv, 0.25)
    dev = torch.device("cuda", 0)
    for term in ("1302", "1203", "0312", "0101", "2323"):
        ext = {"o": no, "v": nv}
        waxes, oaxes, _, _ = TERMS[term]
        W = rng.random([ext[c] for c in waxes])
        oshape = [ext[c] for c in oaxes]
        big = rng.random([2 * s for s in oshape])
        view = big[::2, ::2, ::2, ::2]
        want = view.copy()
        port.contract(term, W, flat, vals, dims, want)
        bigd = torch.from_numpy(big).to(dev)
        viewd = bigd[::2, ::2, ::2, ::2]
        Wd = torch.from_numpy(W).to(dev)
        ranks = [fd.ShardedTerm(term, no, nv, fd.Comm(g, world), dev) for g in range(world)]
        nR, lc = ranks[0].nR, ranks[0].lc
        gathered = torch.zeros(world * nR * lc, dtype=torch.float64, device=dev)
        for g, st in enumerate(ranks):
            st.permute_from(Wd)
            gathered[g * nR * lc:(g + 1) * nR * lc].copy_(st.local(T))
        fd.check(ranks[0].L.fricc_accumulate_panels(fd._ptr(gathered), world, lc, T._h, term.encode(), fd._ptr(viewd),
                                                    fd.i64x4(viewd.shape), fd.i64x4(viewd.stride())))
        got = viewd.cpu().numpy()
        assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max()), term
        # untouched elements of the enclosing buffer stay untouched
        mask = np.ones_like(big, dtype=bool)
        mask[::2, ::2, ::2, ::2] = False
        assert np.array_equal(bigd.cpu().numpy()[mask], big[mask])


@pytest.mark.parametrize("term", sorted(TERMS))
@pytest.mark.parametrize("no,nv,world,side", [(3, 5, 2, False), (6, 21, 3, True), (5, 40, 4, False), (5, 40, 4, True)])
def test_peer_push_exchange(rccsd, port, term, no, nv, world, side):
    """The exchange over peer memory (fricc_peer.cu), ranks emulated one after another on one GPU: rank g accumulates
    its L-columns into ITS copy of the output and stores them into every other copy; afterwards all copies hold the
    full result.  The "peer" copies are ordinary buffers of this process (same kernel, same addressing)."""
    import ctypes

    from fricc_b200 import dist as fd

    rng = np.random.default_rng(7 * no + nv + world)
    torch, dims, flat, vals, T = _setup(rccsd, rng, no, nv, 0.3)
    ext = {"o": no, "v": nv}
    waxes, oaxes, _, _ = TERMS[term]
    W = rng.random([ext[c] for c in waxes])
    out0 = rng.random([ext[c] for c in oaxes])   # replicated, non-zero initial content
    want = out0.copy()
    port.contract(term, W, flat, vals, dims, want)
    dev = torch.device("cuda", 0)
    Wd = torch.from_numpy(W).to(dev)
    copies = [torch.from_numpy(out0).to(dev) for _ in range(world)]
    peers = (ctypes.c_void_p * world)(*[int(c.data_ptr()) for c in copies])

    class FakeArena:   # the copies are ordinary buffers of this process; same kernels, same addressing
        def peers_of(self, t):
            return peers

    fd.use_torch_stream()
    S = fd.SideStream(dev) if side else None   # peer stores issued on a second stream behind the compute
    for g in range(world):
        st = fd.ShardedTerm(term, no, nv, fd.Comm(g, world), dev)
        st.permute_from(Wd)
        # direct epilogue write + column push, or coalesced panel + transposing scatter (1302/1202/1303/1203)
        st.accumulate(T, copies[g], arena=FakeArena(), side=S)
        if S is not None:
            S.join()
    for c in copies:
        got = c.cpu().numpy()
        assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max())
    # single rank, no peers: the same entry point is the 1-GPU path
    one = torch.from_numpy(out0).to(dev)
    st = fd.ShardedTerm(term, no, nv, fd.Comm(0, 1), dev)
    st.permute_from(Wd)
    st.accumulate(T, one)
    assert np.abs(one.cpu().numpy() - want).max() <= TOL * max(1.0, np.abs(want).max())
