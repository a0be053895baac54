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
