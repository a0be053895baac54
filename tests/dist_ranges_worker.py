"""Worker for tests/test_ranges_gpu.py: dist.compress_ranges with several processes sharing cuda:0 (gloo moves the bytes).
Every rank holds only ITS contiguous flat-index range of x and checks the assembled sparse tensor against the
single-process constructor on the whole tensor (same draw): indices and values must be bit-identical."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from fricc_b200 import _lib, py_rccsd  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402
from fricc_b200 import ranges  # noqa: E402
from oracle import oracle  # noqa: E402


def same(A, B):
    fa, va = A.export()
    fb, vb = B.export()
    return np.array_equal(fa, fb) and np.array_equal(va.view(np.int64), vb.view(np.int64))


def make(rng, n, kind):
    if kind == "heavy":
        return rng.standard_normal(n) * np.exp(-4.0 * rng.random(n))
    if kind == "heavier":
        return rng.standard_normal(n) * 10.0 ** (-6.0 * rng.random(n) ** 2)
    if kind == "uniform":
        return rng.random(n)
    if kind == "ties":                      # many exactly equal magnitudes straddling every cut
        return rng.integers(-3, 4, n).astype(np.float64) * 0.25
    if kind == "growing":                   # the running sums cross many binades inside every shard
        return rng.random(n) * np.exp(np.linspace(-20, 8, n))
    raise ValueError(kind)


def parallel_cases(rng, rank, world, comm):
    """sampling_method="parallel" (Alg. A.2, one chunk per rank) against the oracle: D and the probabilities from the
    oracle's stages, the samples from its parallel_sample with `world` chunks fed the draws every part consumed, one
    after the other; values x[D] and x[S] / p[S] (fri_utils.cpp:484-494)."""
    P = oracle.port()
    ok = True
    cases = [((6, 6, 30, 30), 0.10, "heavier"), ((5, 5, 41, 41), 0.05, "heavy"), ((4, 4, 50, 50), 0.30, "uniform"),
             ((8, 8, 64, 64), 0.02, "growing"), ((3, 3, 7, 11), 0.2, "heavy")]
    for dims, frac, kind in cases:
        n = int(np.prod(dims))
        x = make(rng, n, kind)
        m = max(1, int(frac * n))
        budget = rng.integers(0, 2**31 - 1, size=max(world, 1)).astype(np.int32)
        chunk_draws = [rng.integers(0, 2**31 - 1, size=n).astype(np.int32) for _ in range(world)]
        j0, cnt = ranges.chunk_range(n, rank, world)
        mine = torch.from_numpy(x[j0:j0 + cnt].copy()).cuda()
        T = ranges.compress_ranges(mine, j0, dims, m, comm, "fri", "parallel", draws=(budget, chunk_draws[rank]))
        used = [None] * world
        dist.all_gather_object(used, T.consumed_draws_parts)
        D, r = P.get_d_largest(x, m)
        nsamp = m - D.size
        flat, vals = T.export()
        good = np.array_equal(flat[:D.size], D)
        if nsamp > 0:
            p = P.make_probability_vector(x, nsamp, D, r)
            seq = np.concatenate([budget[:used[0][0]]] + [chunk_draws[g][:used[g][1]] for g in range(world)])
            S, n_used = P.parallel_sample(nsamp, p, world, seq)
            good &= n_used == seq.size and np.array_equal(flat[D.size:], S)
            want_vals = np.concatenate([x[D], x[S] / p[S]])
        else:
            want_vals = x[D]
        good &= np.array_equal(vals.view(np.int64), want_vals.view(np.int64))
        if not good:
            print(f"rank {rank}: PARALLEL MISMATCH dims={dims} kind={kind}", flush=True)
        ok &= bool(good)
        # the library's own stream: same result on every rank, every rank's copy advanced by (world - 1) + N
        py_rccsd.srand(31)
        A = ranges.compress_ranges(mine, j0, dims, m, comm, "fri", "parallel")
        pos = py_rccsd.rand_consumed()
        fa, va = A.export()
        mine_sum = torch.tensor([float(fa.sum()), float(va.sum()), float(pos)], dtype=torch.float64)
        lo, hi = mine_sum.clone(), mine_sum.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        good = bool((lo == hi).all()) and (nsamp == 0 or pos == world - 1 + n) and fa.size == m
        good &= np.unique(fa).size == m                                  # a sample without replacement
        if not good:
            print(f"rank {rank}: PARALLEL (stream) MISMATCH dims={dims} kind={kind} pos={pos}", flush=True)
        ok &= bool(good)
    return ok


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    comm = fd.Comm(rank, world)
    _lib.require_cuda()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(123)  # same data on every rank
    ok = True
    cases = [((6, 6, 30, 30), 0.10, "heavier"), ((5, 5, 41, 41), 0.05, "heavy"), ((4, 4, 50, 50), 0.30, "uniform"),
             ((3, 3, 20, 20), 0.50, "ties"), ((8, 8, 64, 64), 0.02, "growing"), ((2, 2, 3, 3), 0.5, "heavy")]
    for dims, frac, kind in cases:
        n = int(np.prod(dims))
        x = make(rng, n, kind)
        m = max(1, int(frac * n))
        draw = rng.integers(0, 2**31 - 1, size=1).astype(np.int32)
        full = torch.from_numpy(x.reshape(dims)).cuda()
        j0, cnt = ranges.index_range(n, rank, world)
        mine = torch.from_numpy(x[j0:j0 + cnt].copy()).cuda()       # this rank never touches the rest
        for compression in ("fri", "largest"):
            T = ranges.compress_ranges(mine, j0, dims, m, comm, compression, "systematic", draws=draw)
            want = py_rccsd.SparseTensor4d(full, dims, m, compression, "systematic", False, draws=draw)
            good = same(T, want)
            if not good:
                print(f"rank {rank}: MISMATCH dims={dims} kind={kind} {compression}", flush=True)
            ok &= good
        # the two compressions of one iteration, sharing norm / D / running sums: equal to two constructor calls
        draw2 = rng.integers(0, 2**31 - 1, size=1).astype(np.int32)
        A, B = ranges.compress_ranges_pair(mine, j0, dims, m, comm, "fri", "systematic", draws=(draw, draw2))
        ok &= same(A, py_rccsd.SparseTensor4d(full, dims, m, "fri", "systematic", False, draws=draw))
        ok &= same(B, py_rccsd.SparseTensor4d(full, dims, m, "fri", "systematic", False, draws=draw2))
        py_rccsd.srand(9)
        A, B = ranges.compress_ranges_pair(mine, j0, dims, m, comm, "fri", "systematic")
        pos2 = py_rccsd.rand_consumed()
        py_rccsd.srand(9)
        ok &= same(A, py_rccsd.SparseTensor4d(full, dims, m, "fri", "systematic"))
        ok &= same(B, py_rccsd.SparseTensor4d(full, dims, m, "fri", "systematic")) and pos2 == 2
        # the library's own rand() stream: one draw consumed, as the reference's lin_space_add_const does
        py_rccsd.srand(5)
        T = ranges.compress_ranges(mine, j0, dims, m, comm, "fri", "systematic")
        pos = py_rccsd.rand_consumed()
        py_rccsd.srand(5)
        want = py_rccsd.SparseTensor4d(full, dims, m, "fri", "systematic")
        ok &= same(T, want) and pos == py_rccsd.rand_consumed() == 1
    ok &= parallel_cases(rng, rank, world, comm)
    # m = 0 and m = N
    dims = (2, 3, 4, 5)
    n = 120
    x = make(rng, n, "heavy")
    j0, cnt = ranges.index_range(n, rank, world)
    mine = torch.from_numpy(x[j0:j0 + cnt].copy()).cuda()
    draw = np.array([12345], dtype=np.int32)
    for m in (0, n):
        T = ranges.compress_ranges(mine, j0, dims, m, comm, "fri", "systematic", draws=draw)
        want = py_rccsd.SparseTensor4d(torch.from_numpy(x.reshape(dims)).cuda(), dims, m, "fri", "systematic", False,
                                       draws=draw)
        ok &= same(T, want)
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("RANGES_OK" if int(flag) == 1 else "RANGES_FAIL")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
