"""parallel_sample ("Alg. A.2", src/fri_utils.cpp:176-362) on the GPU (fricc_psample.cu) through the C-ABI:
against outputs of the compiled reference (tests/golden/parallel_sample.npz) and against the pinned C restatement."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _valid_probs(rng, n, ns):
    x = np.abs(rng.standard_normal(n)) * np.exp(-3 * rng.random(n))
    p = x / x.sum() * ns
    while p.max() > 1:
        p = np.minimum(p, 1.0)
        p = p / p.sum() * ns
    return p


def test_matches_reference_golden_vectors(rccsd):
    g = np.load(os.path.join(GOLDEN, "parallel_sample.npz"))
    for name in g["names"]:
        name = str(name)
        ns, T = (int(v) for v in g[name + "/meta"])
        q = g[name + "/q"]
        d = g[name + "/draws"]
        draws = d if T == 1 else np.full(q.size + 8, int(d[0]), dtype=np.int32)
        got = rccsd.parallel_sample(ns, q, n_chunks=T, draws=draws)
        assert np.array_equal(np.asarray(got, dtype=np.int64), g[name + "/S"]), name


@pytest.mark.parametrize("T", [1, 2, 5, 16])
def test_matches_restatement_with_random_draws(rccsd, port, T):
    """Random draws, several chunks: draws are consumed in chunk order on both sides; also through the library's own
    rand() stream, which must end where the restatement's consumption says."""
    rng = np.random.default_rng(100 + T)
    done = 0
    for n, ns in ((60, 9), (700, 41), (9000, 800), (150000, 20000)):
        if n < T:
            continue
        p = _valid_probs(rng, n, ns)
        draws = rng.integers(0, 2**31 - 1, size=n + 64).astype(np.int32)
        try:
            want, used = port.parallel_sample(ns, p, T, draws)
        except Exception:                      # the reference exits on this input: the library must raise as well
            with pytest.raises(Exception):
                rccsd.parallel_sample(ns, p, n_chunks=T, draws=draws)
            continue
        got = rccsd.parallel_sample(ns, p, n_chunks=T, draws=draws)
        assert np.array_equal(np.asarray(got, dtype=np.int64), want), (n, ns)
        stream = port.glibc_rand(77, n + 64)
        want2, used2 = port.parallel_sample(ns, p, T, stream)
        rccsd.srand(77)
        got2 = rccsd.parallel_sample(ns, p, n_chunks=T)
        assert np.array_equal(np.asarray(got2, dtype=np.int64), want2) and rccsd.rand_consumed() == used2
        done += 1
    assert done >= 2


def test_rejects_bad_arguments(rccsd):
    with pytest.raises(ValueError):
        rccsd.parallel_sample(1, [0.5, 0.5], n_chunks=3)        # fewer elements than chunks
    with pytest.raises(ValueError):
        rccsd.parallel_sample(1, [0.5, 0.5], n_chunks=0)
