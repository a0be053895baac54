"""Oracle results at the headline scale, computed once per session and cached in /tmp (never committed: GB-sized).

north_star asks for bit-exact compression of a 1e8 - 1e9 entry T2; SURVEY.md 7.2-1 measured that this is where parallel
scans start to disagree with the sequential reference.  The C port of the reference (oracle/fricc_oracle.c, pinned to
oracle/_ref and the golden vectors by tests/test_oracle.py) needs ~40 s per 1e8 elements for its full sort, so the stages
are run once -- get_d_largest, make_probability_vector, both samplers on the same probabilities -- and the compression
result [D..., S...], x[D], x[S]/p[S] (fri_utils.cpp:484-504) is assembled from them.  test_oracle.py pins that this
composition IS the port's fri_compression.  Test infrastructure: only tests/ import this."""
import os

import numpy as np

CACHE = os.environ.get("FRICC_SCALE_CACHE", "/tmp")


def make_x(n, seed):
    """heavy-tailed amplitudes (conftest.synth_t2 'heavy'), generated in blocks so every rank can rebuild any range"""
    rng = np.random.default_rng(seed)
    return rng.standard_normal(n) * np.exp(-4.0 * rng.random(n))


def make_draws(n, seed):
    return np.random.default_rng(seed + 1).integers(0, 2**31 - 1, size=n + 4, dtype=np.int64).astype(np.int32)


def at_scale(n, m, seed=2026):
    """-> dict(D, r, S_systematic, S_pivotal, used_pivotal, vals_systematic, vals_pivotal) for x = make_x(n, seed)."""
    path = os.path.join(CACHE, f"fricc_scale_oracle_{n}_{m}_{seed}.npz")
    if os.path.exists(path):
        return dict(np.load(path))
    from oracle import oracle

    P = oracle.port()
    x, draws = make_x(n, seed), make_draws(n, seed)
    D, r = P.get_d_largest(x, m)
    nsamp = m - D.size
    p = P.make_probability_vector(x, nsamp, D, r)
    Ss, us = P.sample_systematic(nsamp, p, draws)
    Sp, up = P.sample_pivotal(nsamp, p, draws)
    Ss, Sp = np.asarray(Ss, dtype=np.int64), np.asarray(Sp, dtype=np.int64)
    out = dict(D=np.asarray(D, dtype=np.int64), r=np.float64(r), S_systematic=Ss, S_pivotal=Sp,
               used_systematic=np.int64(us), used_pivotal=np.int64(up),
               vals_systematic=np.concatenate([x[D], x[Ss] / p[Ss]]), vals_pivotal=np.concatenate([x[D], x[Sp] / p[Sp]]))
    tmp = path + f".{os.getpid()}.tmp.npz"
    np.savez(tmp, **out)
    os.replace(tmp, path)
    return out
