"""The N>1 path's host logic on CPU: world_size 2 and 3, gloo backend (no GPU, no NCCL)."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 3])
def test_column_sharding_and_gather(world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    port = 29600 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "dist_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "DIST_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


def test_column_range_properties():
    from fricc_b200.dist import column_range

    for nL in (1, 7, 16, 1600, 160000, 225):
        for world in (1, 2, 3, 4, 8):
            seen = 0
            for r in range(world):
                l0, cnt, lc = column_range(nL, r, world)
                assert l0 == min(r * lc, nL) and 0 <= cnt <= lc
                seen += cnt
            assert seen == nL
