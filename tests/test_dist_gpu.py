"""Multi-rank paths that need the CUDA library, run as several processes on ONE GPU with gloo for the exchange."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 3])
def test_compress_pair_equals_two_compressions(rccsd, world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29700 + world), os.path.join(HERE, "dist_pair_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    assert "PAIR_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
