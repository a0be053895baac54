"""Range-sharded FRI compression (fricc_b200/ranges.py, csrc/fricc_shard.cu): bit-identical to the single-process
compression (reference: fri_compression, src/fri_utils.cpp:439-522) and to the oracle."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import synth_t2

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 3, 5])
def test_compress_ranges_equals_single_process(rccsd, world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29720 + world), os.path.join(HERE, "dist_ranges_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert "RANGES_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


@pytest.mark.parametrize("kind", ["heavy", "heavier", "uniform", "diff"])
@pytest.mark.parametrize("dims,frac", [((4, 4, 9, 9), 0.2), ((7, 7, 40, 40), 0.05), ((10, 10, 100, 100), 0.03)])
def test_single_rank_stages_vs_oracle(rccsd, port, kind, dims, frac):
    """world = 1 exercises every staged kernel (carry-in scan, threshold compaction, candidate sort, greedy, emit,
    values) without a second process; checked against the C oracle's fri_compression, bit for bit."""
    import torch

    from fricc_b200 import dist as fd
    from fricc_b200 import ranges

    rng = np.random.default_rng(sum(kind.encode()) * 1000 + sum(dims))
    n = int(np.prod(dims))
    x = synth_t2(rng, n, kind)
    m = max(1, int(frac * n))
    draw = rng.integers(0, 2**31 - 1, size=1).astype(np.int32)
    T = ranges.compress_ranges(torch.from_numpy(x).cuda(), 0, dims, m, fd.Comm(), "fri", "systematic", draws=draw)
    flat, vals = T.export()
    want = port.fri_compression(x, m, "systematic", draw)
    assert np.array_equal(flat, want["idx"])
    assert np.array_equal(vals.view(np.int64), want["vals"].view(np.int64))


@pytest.mark.parametrize("kind", ["heavy", "uniform"])
@pytest.mark.parametrize("dims,frac", [((4, 4, 9, 9), 0.2), ((7, 7, 40, 40), 0.05)])
def test_single_chunk_parallel_sampler_vs_oracle(rccsd, port, kind, dims, frac):
    """sampling_method="parallel" with one rank = the reference's parallel_sample with one thread behind the same D and
    probabilities (the multi-rank form is in dist_ranges_worker.py)."""
    import torch

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd
    from fricc_b200 import ranges

    rng = np.random.default_rng(sum(kind.encode()) * 7 + sum(dims))
    n = int(np.prod(dims))
    x = synth_t2(rng, n, kind)
    m = max(1, int(frac * n))
    draws = rng.integers(0, 2**31 - 1, size=n).astype(np.int32)
    D, r = port.get_d_largest(x, m)
    nsamp = m - D.size
    p = port.make_probability_vector(x, nsamp, D, r)
    try:
        S, used = port.parallel_sample(nsamp, p, 1, draws)
    except Exception:
        # n' * |q| / |q| came out a hair off an integer: the reference's one-element budget walk exits (fri_utils.cpp:121)
        with pytest.raises(_lib.FriccError):
            ranges.compress_ranges(torch.from_numpy(x).cuda(), 0, dims, m, fd.Comm(), "fri", "parallel", draws=(draws[:0], draws))
        return
    T = ranges.compress_ranges(torch.from_numpy(x).cuda(), 0, dims, m, fd.Comm(), "fri", "parallel", draws=(draws[:0], draws))
    flat, vals = T.export()
    assert np.array_equal(flat, np.concatenate([D, S]))
    assert np.array_equal(vals.view(np.int64), np.concatenate([x[D], x[S] / p[S]]).view(np.int64))
    assert T.consumed_draws_parts == (0, used) and T.consumed_draws == used


def test_carry_in_scan_is_the_sequential_sum(rccsd):
    """The staged scan with a non-zero exact carry: sum(x[:k]) carried into the shard x[k:] must give the plain
    left-to-right fp64 sum of all of x (one_norm, fri_utils.cpp:164-171) for every split point."""
    import torch

    from fricc_b200 import _lib
    from fricc_b200.ranges import _Shard, _ptr

    L = _lib.lib()
    rng = np.random.default_rng(11)
    n = 50_000
    x = rng.standard_normal(n) * np.exp(rng.uniform(-12, 6, n))
    seq = np.add.accumulate(np.abs(x))          # numpy accumulates strictly left to right
    xd = torch.from_numpy(x).cuda()
    for k in (0, 1, 777, 2048, 30_001, n - 1, n):
        sh = _Shard(xd[k:], k, n)
        L.fricc_shard_use_abs(sh.h)
        approx = torch.zeros(1, dtype=torch.float64, device="cuda")
        _lib.check(L.fricc_shard_scan_local(sh.h, _ptr(approx)))
        carry_val = float(seq[k - 1]) if k > 0 else 0.0
        rough = torch.tensor([carry_val * (1 + 1e-9)], dtype=torch.float64, device="cuda")  # prediction may be off
        _lib.check(L.fricc_shard_scan_maps(sh.h, _ptr(rough)))
        carry = torch.tensor([carry_val], dtype=torch.float64, device="cuda")
        _lib.check(L.fricc_shard_scan_resolve(sh.h, _ptr(carry), _ptr(carry)))
        assert float(carry.item()) == float(seq[-1]), k
        sh.close()


def test_rejects_what_it_does_not_do(rccsd):
    import torch

    from fricc_b200 import dist as fd
    from fricc_b200 import ranges

    x = torch.rand(2 * 2 * 3 * 3, dtype=torch.float64, device="cuda")
    with pytest.raises(NotImplementedError):
        ranges.compress_ranges(x, 0, (2, 2, 3, 3), 5, fd.Comm(), "fri", "pivotal")
    with pytest.raises(ValueError):
        ranges.compress_ranges(x, 0, (2, 2, 3, 3), 37, fd.Comm(), "fri", "systematic")   # m > N, friccsd.py:69-72
    with pytest.raises(TypeError):
        ranges.compress_ranges(x.float(), 0, (2, 2, 3, 3), 5, fd.Comm())
