"""Worker for tests/test_scale_parity_gpu.py: ranges.compress_ranges at N = 1e8 with several processes sharing cuda:0
(gloo moves the bytes).  Every rank puts only ITS flat-index range of x on the GPU; the assembled compression must equal
the oracle's (cached by the test before the launch), indices and values bit for bit."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import scale_oracle  # noqa: E402
from fricc_b200 import _lib  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402
from fricc_b200 import ranges  # noqa: E402


def main():
    n, m, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    comm = fd.Comm(rank, world)
    _lib.require_cuda()
    torch.cuda.set_device(0)
    want = scale_oracle.at_scale(n, m, seed)
    x, draws = scale_oracle.make_x(n, seed), scale_oracle.make_draws(n, seed)
    dims = (10, 10, 1000, 1000)
    assert int(np.prod(dims)) == n
    j0, cnt = ranges.index_range(n, rank, world)
    mine = torch.from_numpy(x[j0:j0 + cnt].copy()).cuda()
    del x
    T = ranges.compress_ranges(mine, j0, dims, m, comm, "fri", "systematic", draws=draws[:1])
    flat, vals = T.export()
    nD = want["D"].size
    ok = np.array_equal(flat[:nD], want["D"]) and np.array_equal(flat[nD:], want["S_systematic"]) and \
        np.array_equal(vals.view(np.int64), want["vals_systematic"].view(np.int64))
    flag = torch.tensor([int(ok)])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("SCALE_OK" if int(flag.item()) else "SCALE_MISMATCH", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
