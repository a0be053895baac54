"""Bit-exact FRI compression at the scale north_star names (N = 1e8 and the headline T2 of 40 x 40 x 400 x 400 = 2.56e8
entries, m = 1e7), both samplers, against the C port of the reference (fri_utils.cpp:439-522 with :63-127, :129-159,
:400-437) fed the same draws; and the same through the range-sharded path at world 2 and 5.  At these sizes the running
sum of p reaches 1e7 (about 23 binade crossings), the D search takes the candidate-growth path and the parallel pivotal
sampler runs over 1e8+ elements -- none of which the smaller cases exercise."""
import os
import subprocess
import sys

import numpy as np
import pytest

import scale_oracle

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
SEED = 2026


@pytest.mark.parametrize("n,m", [(100_000_000, 10_000_000), (40 * 40 * 400 * 400, 10_000_000)])
def test_compression_equals_the_oracle_at_scale(rccsd, n, m):
    import torch

    want = scale_oracle.at_scale(n, m, SEED)
    x, draws = scale_oracle.make_x(n, SEED), scale_oracle.make_draws(n, SEED)
    xd = torch.from_numpy(x).cuda()
    nD = want["D"].size
    assert 0 < nD < m
    for method in ("systematic", "pivotal"):
        idx, vals = rccsd.fri_compression(xd, m, method, draws=draws)
        assert np.array_equal(idx[:nD], want["D"]), method
        assert np.array_equal(idx[nD:], want[f"S_{method}"]), method
        assert np.array_equal(vals.view(np.int64), want[f"vals_{method}"].view(np.int64)), method
    assert rccsd.one_norm(xd) == float(np.cumsum(np.abs(x))[-1])


@pytest.mark.parametrize("world", [2, 5])
def test_range_sharded_compression_equals_the_oracle_at_scale(rccsd, world):
    n, m = 100_000_000, 10_000_000
    scale_oracle.at_scale(n, m, SEED)          # computed (or found in the cache) before the workers start
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29760 + world), os.path.join(HERE, "dist_scale_worker.py"), str(n), str(m),
           str(SEED)]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert "SCALE_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
