"""Worker for tests/test_dist_gpu.py: dist.compress_pair with several processes sharing cuda:0 (gloo moves the entries;
nothing on the GPU waits for another process).  Every rank checks the pair against two consecutive compressions."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from fricc_b200 import _lib, py_rccsd  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402


def same(A, B):
    fa, va = A.export()
    fb, vb = B.export()
    return np.array_equal(fa, fb) and np.array_equal(va, vb)


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    comm = fd.Comm(rank, world)
    _lib.require_cuda()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(9)
    no, nv = 6, 30
    n = no * no * nv * nv
    x = rng.standard_normal(n) * 10.0 ** (-6.0 * rng.random(n) ** 2)
    t2 = torch.from_numpy(x.reshape(no, no, nv, nv)).cuda()
    m = n // 10
    ok = True
    L = _lib.lib()
    for method in ("systematic", "pivotal"):
        # (a) explicit draws
        d0 = rng.integers(0, 2**31 - 1, size=n).astype(np.int32)
        d1 = rng.integers(0, 2**31 - 1, size=n).astype(np.int32)
        A, B = fd.compress_pair(t2, m, comm, "fri", method, draws=(d0, d1))
        ok &= same(A, fd.compress(t2, m, "fri", method, draws=d0)) and same(B, fd.compress(t2, m, "fri", method, draws=d1))
        for T in (A, B):
            ne = fd.ctypes.c_int64(-5)
            fd.check(L.fricc_st_get_n_exact(T._h, fd.ctypes.byref(ne)))
            ok &= 0 <= ne.value <= m
        # (b) the library's own rand() stream: the pair must equal two consecutive constructor calls after srand(seed),
        #     and leave the stream where those leave it
        py_rccsd.srand(77)
        A, B = fd.compress_pair(t2, m, comm, "fri", method, draws=None)
        pos_pair = py_rccsd.rand_consumed()
        py_rccsd.srand(77)
        A2 = fd.compress(t2, m, "fri", method)
        B2 = fd.compress(t2, m, "fri", method)
        ok &= same(A, A2) and same(B, B2) and pos_pair == py_rccsd.rand_consumed()
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("PAIR_OK" if int(flag) == 1 else "PAIR_FAIL")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
