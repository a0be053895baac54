"""Host logic of the range-sharded compression (fricc_b200/ranges.py) on CPU: index ranges, and the digit walk over
ALL-REDUCED key histograms (world_size 2 and 3, gloo).  The histograms are computed with numpy here -- the CUDA kernel
that produces them on the GPU (fricc_key_histogram) is covered by tests/test_ranges_gpu.py."""
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def np_hist_fn(shards):
    keys = [np.abs(s).view(np.uint64) for s in shards]

    def fn(shift, prefix, prefix_bits):
        nbits = 3 if shift == 0 else 12
        h = np.zeros(4096, dtype=np.int64)
        for k in keys:
            sel = k if prefix_bits == 0 else k[(k >> np.uint64(shift + nbits)) == np.uint64(prefix)]
            d = ((sel >> np.uint64(shift)) & np.uint64((1 << nbits) - 1)).astype(np.int64)
            h += np.bincount(d, minlength=4096)
        return h

    return fn


def test_index_range_properties():
    from fricc_b200.ranges import index_range

    for n in (0, 1, 7, 120, 10**6 + 3):
        for world in (1, 2, 3, 5, 8):
            nxt = 0
            for r in range(world):
                j0, cnt = index_range(n, r, world)
                assert j0 == nxt and cnt >= 0
                nxt = j0 + cnt
            assert nxt == n


@pytest.mark.parametrize("kind", ["normal", "ties", "tiny"])
def test_locate_threshold_against_sort(kind):
    from fricc_b200.ranges import locate_threshold

    rng = np.random.default_rng(3)
    n = 20000
    if kind == "normal":
        x = rng.standard_normal(n) * np.exp(-5 * rng.random(n))
    elif kind == "ties":
        x = rng.integers(-4, 5, n).astype(np.float64)
    else:
        x = rng.standard_normal(n) * 1e-300
    shards = np.array_split(x, 3)
    keys = np.sort(np.abs(x).view(np.uint64))[::-1]
    for want in (1, 17, 4096, 12345, n):
        for small in (1, 64, 1 << 16):
            key_min, count = locate_threshold(np_hist_fn(shards), want, small_bin=small)
            assert count == int((keys >= np.uint64(key_min)).sum())
            assert count >= want
            assert key_min <= int(keys[want - 1])          # the want-th largest key is at or above the threshold
    assert locate_threshold(np_hist_fn(shards), 0)[1] == 0


@pytest.mark.parametrize("world", [2, 3])
def test_threshold_with_allreduced_histograms(world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29640 + world), os.path.join(HERE, "dist_threshold_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "THRESHOLD_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


# ---- Alg. A.2 with one chunk per rank (sampling_method="parallel"): the host side -----------------------------------
def test_chunk_range_is_the_references_chunking():
    from fricc_b200.ranges import chunk_range

    for n in (1, 7, 120, 10**6 + 3):
        for world in (1, 2, 3, 5, 8):
            if n < world:
                continue
            step = n // world
            for r in range(world):
                j0, cnt = chunk_range(n, r, world)
                assert j0 == r * step                                   # fri_utils.cpp:216-219
                assert cnt == step + (n % world if r == world - 1 else 0)


def test_pivotal_host_is_the_oracles_sample_pivotal(port):
    from fricc_b200 import _lib
    from fricc_b200.ranges import pivotal_host

    rng = np.random.default_rng(77)
    for trial in range(300):
        T = int(rng.integers(1, 9))
        t = rng.random(T)
        c = int(round(t.sum()))
        t *= c / t.sum() if t.sum() > 0 else 0.0                       # fractional parts that add up to an integer
        t = np.minimum(t, 1.0)
        draws = rng.integers(0, 2**31 - 1, size=T).astype(np.int32)
        try:
            want, used_w = port.sample_pivotal(c, t, draws)
        except Exception:
            with pytest.raises(_lib.FriccError):
                pivotal_host(c, t, draws)
            continue
        got, used = pivotal_host(c, t, draws)
        assert got == want.tolist() and used == used_w


def _rescale(q, t, a, g, s):
    """fri_utils.cpp:303-322 on one chunk (numpy float64 scalars: the reference's operations in its order)."""
    if g > a:
        for i in range(q.size):
            qi = q[i]
            y = min(1.0, qi / t)
            s += y - qi
            q[i] = y
            if s >= g:
                q[i] = y + g - s
                break
    elif g < a:
        for i in range(q.size):
            qi = q[i]
            y = max(0.0, (qi - t) / (1.0 - t))
            s += y - qi
            q[i] = y
            if s <= g:
                q[i] = y + g - s
                break
    return q


@pytest.mark.parametrize("world", [1, 2, 3, 5, 8])
def test_chunks_sampled_apart_reassemble_to_parallel_sample(port, world):
    """What the ranks do (budgets on the host from exact norms, every chunk rescaled and sampled on its own with its own
    draws) equals the oracle's parallel_sample with `world` chunks when the draws each part actually consumed are put
    one after the other -- the reference's behaviour whenever its threads happen to run in chunk order."""
    from fricc_b200.ranges import chunk_budgets, chunk_range

    rng = np.random.default_rng(1000 + world)
    done = 0
    for trial in range(40):
        n = int(rng.integers(world * 4, 3000))
        nsamp = int(rng.integers(1, max(2, n // 8)))
        w = rng.random(n) * np.exp(-3.0 * rng.random(n))
        p = w / w.sum() * nsamp
        if p.max() > 1.0:
            continue
        budget_draws = rng.integers(0, 2**31 - 1, size=world).astype(np.int32)
        chunk_draws = [rng.integers(0, 2**31 - 1, size=n).astype(np.int32) for _ in range(world)]
        q_norm = float(np.add.accumulate(p)[-1])                        # left to right, like one_norm
        rngs = [chunk_range(n, r, world) for r in range(world)]
        norms = [float(np.add.accumulate(p[j0:j0 + c])[-1]) for j0, c in rngs]
        try:
            a, t, g, used0 = chunk_budgets(nsamp, norms, q_norm, budget_draws)
        except Exception:
            continue                                                    # the reference exits too (short budget walk)
        S, seq = [], [budget_draws[:used0]]
        for r, (j0, c) in enumerate(rngs):
            q = _rescale(p[j0:j0 + c].copy(), t[r], a[r], float(g[r]), norms[r])
            Sr, u = port.sample_pivotal(g[r], q, chunk_draws[r])
            S.append(Sr + j0)
            seq.append(chunk_draws[r][:u])
        want, used = port.parallel_sample(nsamp, p, world, np.concatenate(seq))
        assert used == sum(s.size for s in seq)
        assert np.array_equal(np.concatenate(S), want)
        done += 1
    assert done >= 20
