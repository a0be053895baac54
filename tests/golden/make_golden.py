"""Generate the golden vectors under tests/golden/ from the COMPILED, UNMODIFIED reference (oracle/_ref).

Run in the build container (needs /root/reference for `make -C oracle ref`):

    OMP_NUM_THREADS=1 python tests/golden/make_golden.py

The reference has no golden vectors or known-answer tests of its own for FRI compression (SURVEY.md section 4:
"parity unpinned by the reference's tests"), so its own outputs on seeded inputs are recorded here.  OMP_NUM_THREADS=1
makes one_norm (an OpenMP reduction) bit-deterministic.  The files are small on purpose (< 1 MB in total).
"""
import os
import sys

os.environ["OMP_NUM_THREADS"] = "1"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from oracle import oracle  # noqa: E402

TERMS = {  # term -> (W axes, out axes) in terms of o / v extents   
    "0101": ("oooo", "oovv"), "2323": ("vvvv", "oovv"), "1302": ("voov", "oovv"), "1202": ("vovo", "oovv"),
    "1303": ("voov", "oovv"), "1203": ("vovo", "oovv"), "1323": ("ovov", "oooo"), "0112": ("ovov", "voov"),
    "0313": ("ovov", "voov"), "0113": ("ovov", "voov"), "0312": ("ovov", "vovo"),
}


def synth_t2(rng, n, kind):
    """Synthetic amplitude vectors: heavy-tailed, the reference tests' own distributions, ties."""
    if kind == "heavy":
        return rng.standard_normal(n) * np.exp(-4.0 * rng.random(n))
    if kind == "heavier":  # a few dominant entries -> non-empty D
        return rng.standard_normal(n) * 10.0 ** (-6.0 * rng.random(n) ** 2)
    if kind == "uniform":  # fricc/tests/test_sparse_tensors.py:30
        return rng.random(n)
    if kind == "diff":  # fricc/tests/test_fri_utils.py:92
        return rng.random(n) - rng.random(n)
    raise ValueError(kind)


def compression_cases():
    rng = np.random.default_rng(20240501)
    cases = []
    for kind in ("heavy", "heavier", "uniform", "diff"):
        for (n, m) in ((97, 13), (1000, 100), (4096, 1500), (9025, 4000)):
            x = synth_t2(rng, n, kind)
            draws = rng.integers(0, 2**31 - 1, size=n + 8).astype(np.int32)
            cases.append((f"{kind}_{n}_{m}", x, m, draws))
    return cases


def main():
    R = oracle.ref()
    if R is None:
        raise SystemExit("oracle/_ref is not built: run `make -C oracle ref` first (needs /root/reference)")
    out = {}
    names = []
    for name, x, m, draws in compression_cases():
        names.append(name)
        out[f"{name}/x"] = x
        out[f"{name}/m"] = np.int64(m)
        out[f"{name}/draws"] = draws
        out[f"{name}/one_norm"] = np.float64(R.one_norm(x))
        out[f"{name}/argsort"] = R.argsort(x, m)
        D, r = R.get_d_largest(x, m)
        out[f"{name}/D"] = D
        out[f"{name}/remaining_norm"] = np.float64(r)
        p = R.make_probability_vector(x, m - len(D), D, r)
        out[f"{name}/p"] = p
        for method in ("pivotal", "systematic"):
            S, used = getattr(R, f"sample_{method}")(m - len(D), p, draws)
            out[f"{name}/{method}/S"] = S
            out[f"{name}/{method}/consumed"] = np.int64(used)
            c = R.fri_compression(x, m, method, draws)
            out[f"{name}/{method}/idx"] = c["idx"]
            out[f"{name}/{method}/vals"] = c["vals"]
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "compression.npz"), **out)

    # contractions: the reference's own test recipe (fricc/tests/test_sparse_tensors.py:8-15, 48-269), all 11 terms,
    # accumulating onto a NON-zero output, through the reference's pybind11 module
    np.random.seed(20)
    cout = {}
    cnames = []
    for (no, nv, frac) in ((4, 8, 0.1), (5, 10, 0.01), (3, 7, 0.3)):
        shape = (no, no, nv, nv)
        size = int(np.prod(shape))
        nnz = int(size * frac)
        t2 = np.zeros(size)
        t2[np.random.choice(np.arange(size), size=nnz)] = np.random.rand(nnz)
        t2 = t2.reshape(shape)
        m = int(t2.size * frac)
        T = R.mod.SparseTensor4d(t2, t2.shape, m, "largest")
        flat = np.array([np.ravel_multi_index(tuple(T.get_element(i)[0]), shape) for i in range(m)], dtype=np.int64)
        vals = np.array([T.get_element(i)[1] for i in range(m)])
        key = f"{no}_{nv}"
        cnames.append(key)
        cout[f"{key}/t2"] = t2
        cout[f"{key}/m"] = np.int64(m)
        cout[f"{key}/flat"] = flat
        cout[f"{key}/vals"] = vals
        ext = {"o": no, "v": nv}
        for term, (wax, oax) in TERMS.items():
            W = np.random.rand(*[ext[c] for c in wax])
            o0 = np.random.rand(*[ext[c] for c in oax])
            o1 = o0.copy()
            R.mod.contract_DTSpT(W, T, o1, term)
            cout[f"{key}/{term}/W"] = W
            cout[f"{key}/{term}/out0"] = o0
            cout[f"{key}/{term}/out1"] = o1
    cout["names"] = np.array(cnames)
    np.savez_compressed(os.path.join(HERE, "contractions.npz"), **cout)
    for f in ("compression.npz", "contractions.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
