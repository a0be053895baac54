"""Worker for tests/test_dist_cpu.py: the N>1 host logic (column sharding, gather order, panel assembly) on CPU with the
gloo backend.  The per-rank arithmetic is done by the oracle (this is a test, not the product path)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from fricc_b200.dist import TERM_AXES, Comm, column_range  # noqa: E402
from oracle import oracle  # noqa: E402


def pair_index(shape, axes, idx):
    return idx[axes[0]] * shape[axes[1]] + idx[axes[1]]


def panel_of(out, ax):
    """4-D output -> [R][L] matrix through the term's (oR, oL) axes."""
    oR, oL = ax["oR"], ax["oL"]
    perm = (oR[0], oR[1], oL[0], oL[1])
    t = np.transpose(out, perm)
    return np.ascontiguousarray(t).reshape(t.shape[0] * t.shape[1], t.shape[2] * t.shape[3]), perm


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    comm = Comm(rank, world)
    P = oracle.port()
    rng = np.random.default_rng(42)  # same data on every rank
    no, nv = 3, 5
    dims = (no, no, nv, nv)
    ext = {"o": no, "v": nv}
    m = 60
    flat = rng.choice(no * no * nv * nv, m, replace=False).astype(np.int64)
    vals = rng.standard_normal(m)
    ok = True
    for term, ax in TERM_AXES.items():
        W = rng.random([ext[c] for c in ax["w"]])
        oshape = [ext[c] for c in ax["out"]]
        full = np.zeros(oshape)
        P.contract(term, W, flat, vals, dims, full)
        want_RL, perm = panel_of(full, ax)
        nR, nL = want_RL.shape
        l0, cnt, lc = column_range(nL, rank, world)
        # this rank's column slice only: zero every other column of W's free pair before contracting
        mine = np.zeros((nR, lc))
        mine[:, :cnt] = want_RL[:, l0:l0 + cnt]
        gathered = torch.empty(world * nR * lc, dtype=torch.float64)
        comm.all_gather(gathered, torch.from_numpy(mine.reshape(-1)))
        g = gathered.numpy().reshape(world, nR, lc)
        # host emulation of accumulate_panels_kernel: panel g holds columns [g*lc, min((g+1)*lc, nL))
        got = np.zeros((nR, nL))
        for q in range(world):
            a, c, _ = column_range(nL, q, world)
            got[:, a:a + c] += g[q][:, :c]
        ok = ok and np.array_equal(got, want_RL)
        # every column is owned exactly once
        cover = np.zeros(nL, dtype=int)
        for q in range(world):
            a, c, _ = column_range(nL, q, world)
            cover[a:a + c] += 1
        ok = ok and (cover == 1).all()
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("DIST_OK" if int(flag) == 1 else "DIST_FAIL")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
