"""Golden vectors for parallel_sample (fri_utils.cpp:176-362) from the COMPILED, UNMODIFIED reference (oracle/_ref).

    python tests/golden/make_golden_parallel_sample.py        (needs /root/reference for `make -C oracle ref`)

The reference's threads call rand() concurrently and its one_norm is an OpenMP reduction, so for more than one thread its
output is reproducible only when (a) the draw stream is constant (the race over rand() cannot matter) and (b) the
probabilities are dyadic rationals whose sums are exact in any order.  Cases: one thread with random draws (any
probabilities), and 2/3/8/16 threads on such inputs.  File: tests/golden/parallel_sample.npz (< 200 KB)."""
import os
import sys

os.environ["OMP_NUM_THREADS"] = "1"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from oracle import oracle  # noqa: E402


def dyadic_probs(rng, n, ns, k=14):
    """n probabilities, multiples of 2^-k, each <= 1, summing to ns exactly (every partial sum is exact in fp64)."""
    W = ns << k
    w = rng.multinomial(W, np.ones(n) / n).astype(np.int64)
    for _ in range(n // 4):     # make it uneven
        i, j = rng.integers(0, n, 2)
        d = min(w[i] // 2, (1 << k) - w[j])
        w[i] -= d
        w[j] += d
    assert w.sum() == W and w.max() <= (1 << k) and w.min() >= 0
    return w.astype(np.float64) / (1 << k)


def general_probs(rng, n, ns):
    x = np.abs(rng.standard_normal(n)) * np.exp(-3 * rng.random(n))
    p = x / x.sum() * ns
    while p.max() > 1:
        p = np.minimum(p, 1.0)
        p = p / p.sum() * ns
    return p


def cases():
    rng = np.random.default_rng(20241104)
    out = []
    for n, ns in ((50, 7), (400, 37), (5000, 300)):
        q = general_probs(rng, n, ns)
        out.append((f"t1_general_{n}_{ns}", q, ns, 1, rng.integers(0, 2**31 - 1, size=n + 8).astype(np.int32)))
    for n, ns in ((64, 8), (400, 37), (5000, 300)):
        q = dyadic_probs(rng, n, ns)
        out.append((f"t1_dyadic_{n}_{ns}", q, ns, 1, rng.integers(0, 2**31 - 1, size=n + 8).astype(np.int32)))
        for T in (2, 3, 8, 16):
            for const in (0, 12345, 2**30, 2**31 - 2):
                out.append((f"t{T}_dyadic_{n}_{ns}_c{const}", q, ns, T, np.full(n + 8, const, dtype=np.int32)))
    return out


def main():
    R = oracle.ref()
    if R is None:
        raise SystemExit("oracle/_ref is not built: run `make -C oracle ref` first (needs /root/reference)")
    out, names = {}, []
    for name, q, ns, T, draws in cases():
        S, _ = R.parallel_sample(ns, q, T, draws)
        assert len(S) == ns
        names.append(name)
        out[name + "/q"] = q
        out[name + "/meta"] = np.array([ns, T], dtype=np.int64)
        out[name + "/draws"] = draws[:1] if T > 1 else draws          # constant streams are stored as one value
        out[name + "/S"] = S
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "parallel_sample.npz"), **out)
    print(f"wrote {len(names)} cases")


if __name__ == "__main__":
    main()
