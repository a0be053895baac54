"""ViewEngine's host logic on CPU (fricc_b200/dist.py): for every term and world size the per-rank L-slice views of the
operand and of the output tile the whole contraction -- summing what each emulated rank adds equals the unsharded
result.  The plugin call is replaced by numpy's einsum of the reference's own expression (fricc/tests/test_sparse_tensors.py:
48-269) on the same views (numpy shares memory with the torch CPU tensors; the free indices of a sliced operand give the
sliced output by themselves), and the whole is compared with the oracle port on the unsliced operands.  What is tested is exactly the view arithmetic a rank would hand to fricc_contract: which axis is cut,
its range, the matching output view, and the pre-sliced (w_range) form the ladder uses."""
import numpy as np
import pytest
import torch

from conftest import TERMS

from fricc_b200 import dist as fd


class _EinsumPlugin:
    """contract_DTSpT with the library's argument meaning (out += prefactor * einsum(W, T)), in numpy."""

    def contract_DTSpT(self, W, T, out, term):
        flat, vals, dims = T
        dense = np.zeros(int(np.prod(dims)))
        np.add.at(dense, flat, vals)
        _, _, scale, expr = TERMS[term]
        o = out.numpy()
        o += scale * np.einsum(expr, W.numpy(), dense.reshape(dims))


def _engine(port, rank, world):
    eng = fd.ViewEngine.__new__(fd.ViewEngine)           # no device, no arena: host logic only
    eng.comm, eng.device, eng.arena, eng.side = fd.Comm(rank, world), None, None, None
    eng.py = _EinsumPlugin()
    return eng


def _case(rng, term, no, nv, nnz):
    ext = {"o": no, "v": nv}
    waxes, oaxes, _, _ = TERMS[term]
    dims = (no, no, nv, nv)
    n = int(np.prod(dims))
    flat = rng.choice(n, size=nnz, replace=False).astype(np.int64)
    vals = rng.standard_normal(nnz)
    W = rng.random([ext[c] for c in waxes])
    out0 = rng.random([ext[c] for c in oaxes])
    return W, out0, (flat, vals, dims)


@pytest.mark.parametrize("world", [1, 2, 3, 5])
@pytest.mark.parametrize("term", sorted(TERMS))
def test_slices_tile_the_contraction(port, term, world):
    rng = np.random.default_rng(17)
    no, nv = 3, 7
    W, out0, T = _case(rng, term, no, nv, 60)
    want = out0.copy()
    port.contract(term, W, T[0], T[1], T[2], want)
    Wt = torch.from_numpy(W)
    got = torch.from_numpy(out0.copy())
    covered = torch.zeros_like(got)
    for r in range(world):
        eng = _engine(port, r, world)
        own = eng.own_view(term, Wt.shape, got)
        ov = eng.contract(term, Wt, T, got)
        assert (own is None) == (ov is None)
        if ov is not None:
            assert own.data_ptr() == ov.data_ptr() and own.shape == ov.shape and own.stride() == ov.stride()
            assert ov.data_ptr() >= got.data_ptr()               # a view of the output, not a copy
            j = eng.slice_axis(term, Wt.shape, world)
            lo, hi = eng.my_range(Wt.shape[fd.TERM_AXES[term]["wL"][j]])
            covered.narrow(fd.TERM_AXES[term]["oL"][j], lo, hi - lo).add_(1.0)
    assert np.abs(got.numpy() - want).max() <= 1e-12 * np.abs(want).max()
    assert (covered == 1.0).all()                                # every element updated by exactly one rank


@pytest.mark.parametrize("world", [2, 4])
def test_presliced_ladder_operand(port, world):
    """w_range: the rank holds only ITS part of Wvvvv (possibly in sub-shards) and says which part it is."""
    rng = np.random.default_rng(3)
    no, nv = 3, 9
    W, out0, T = _case(rng, "2323", no, nv, 80)
    want = out0.copy()
    port.contract("2323", W, T[0], T[1], T[2], want)
    got = torch.from_numpy(out0.copy())
    for r in range(world):
        eng = _engine(port, r, world)
        lo, hi = eng.my_range(nv)
        mid = (lo + hi) // 2
        for a, b in ((lo, mid), (mid, hi)):                      # two sub-shards per rank
            part = torch.from_numpy(np.ascontiguousarray(W[a:b]))
            eng.contract("2323", part, T, got, w_range=(a, b))
    assert np.abs(got.numpy() - want).max() <= 1e-12 * np.abs(want).max()


def test_cut_axis_is_the_longer_free_axis():
    for term, ax in fd.TERM_AXES.items():
        shape = [{"o": 4, "v": 9}[c] for c in ax["w"]]
        for world in (1, 3):                                     # 4 occupied orbitals do not split evenly among 3 ranks
            j = fd.ViewEngine.slice_axis(term, shape, world)
            e = [shape[ax["wL"][0]], shape[ax["wL"][1]]]
            assert e[j] == max(e) and (e[0] != e[1] or j == 0)


def test_cut_axis_avoids_short_output_rows_when_the_other_axis_splits_evenly():
    """Headline shape on 8 ranks: cutting v would leave 400-byte rows in Wvoov (0112, 0113, 0313) and in tmp2 (1203);
    o = 40 splits evenly, so those terms are cut along o and every rank's output slice is contiguous."""
    dims = {"o": 40, "v": 400}
    switched = set()
    for term, ax in fd.TERM_AXES.items():
        shape = [dims[c] for c in ax["w"]]
        j1, j8 = fd.ViewEngine.slice_axis(term, shape, 1), fd.ViewEngine.slice_axis(term, shape, 8)
        if j1 != j8:
            switched.add(term)
            assert ax["oL"][j1] == 3 and ax["oL"][j8] != 3
        assert fd.ViewEngine.slice_axis(term, shape, 7) == j1    # 40 does not split among 7: the longer axis stays
    assert switched == {"0112", "0113", "0313", "1203"}


@pytest.mark.parametrize("parts", [1, 2, 3])
def test_contract_in_parts_is_the_same_contraction(port, parts):
    rng = np.random.default_rng(5)
    no, nv = 3, 8
    W, out0, T = _case(rng, "1203", no, nv, 70)
    want = out0.copy()
    port.contract("1203", W, T[0], T[1], T[2], want)
    got = torch.from_numpy(out0.copy())
    for r in range(2):
        _engine(port, r, 2).contract_in_parts("1203", torch.from_numpy(W), T, got, parts=parts)
    assert np.abs(got.numpy() - want).max() <= 1e-12 * np.abs(want).max()
