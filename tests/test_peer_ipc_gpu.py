"""The N>1 data path as separate processes (tests/dist_peer_worker.py): SymmArena IPC mappings, fricc_push_view stores
into the peers' copies, and -- where every process has its own GPU -- the device-side epoch barrier.  On a one-GPU box
the processes share the device and the barrier is a host one (see the worker's header)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_step_in_separate_processes(rccsd, world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29750 + world), os.path.join(HERE, "dist_peer_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert "PEER_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
