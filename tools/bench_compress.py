#!/usr/bin/env python
"""Compression-only sweep (BASELINE.json config 5; SURVEY.md 8d "Algorithmic bytes -- compression").

One FRI compression (top-m select + D greedy + sampler) of a synthetic heavy-tailed T2 of N entries resident in HBM, for
N in --sizes and both samplers.  Reports ms per compression (CUDA events on the launching stream, median of --reps after
--warmup), elements/s, and the algorithmic-bytes roofline fraction  B_comp = 16 N + 16 m  against MEASURED_PEAKS.json.
Prints one JSON line per (N, sampler).  Default: single process / one GPU (SparseTensor4d constructor).

--ranges: the range-sharded compression (fricc_b200/ranges.py: x split by contiguous flat-index ranges over the ranks,
all-reduced key histograms, exact sums carried from rank to rank).  Launch with
  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tools/bench_compress.py --ranges
Every rank generates only its own N/G elements; time = max over ranks (barrier + synchronize on both sides, CUDA
events); rank 0 prints.  systematic sampler only."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1e8,2.56e8,1e9")
    ap.add_argument("--m", type=float, default=1e7)
    ap.add_argument("--samplers", default="systematic,pivotal")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--ranges", action="store_true")
    args = ap.parse_args()
    if args.ranges:
        return main_ranges(args)
    import numpy as np
    import torch

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd
    from fricc_b200.py_rccsd import SparseTensor4d

    _lib.require_cuda()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    fd.use_torch_stream()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    m = int(args.m)
    for size in args.sizes.split(","):
        n = int(float(size))
        v = int(round((n / 1600) ** 0.5))          # shape (40, 40, v, v) with 1600 v^2 ~ N
        n = 1600 * v * v
        g = torch.Generator(device=dev)
        g.manual_seed(20)
        u = torch.rand(n, generator=g, device=dev, dtype=torch.float64)
        z = torch.randn(n, generator=g, device=dev, dtype=torch.float64)
        t2 = (z * torch.pow(10.0, -6.0 * u * u)).reshape(40, 40, v, v).contiguous()
        del u, z
        for sampler in args.samplers.split(","):
            draws = None
            if sampler == "systematic":
                draws = np.array([123456789], dtype=np.int32)
            times = []
            for it in range(args.warmup + args.reps):
                L0 = _lib.lib().fricc_kernel_launches()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                _lib.lib().fricc_srand(1)
                torch.cuda.synchronize()
                e0.record()
                T = SparseTensor4d(t2, tuple(t2.shape), m, "fri", sampler, False, draws=draws)
                e1.record()
                torch.cuda.synchronize()
                if it >= args.warmup:
                    times.append(e0.elapsed_time(e1))
                launches = _lib.lib().fricc_kernel_launches() - L0
                del T
            times.sort()
            ms = times[len(times) // 2]
            bcomp = 16.0 * n + 16.0 * m
            print(json.dumps({"workload": "FRI compression only", "N": n, "m": m, "sampler": sampler,
                              "ms_per_compression": ms, "elements_per_s": n / (ms * 1e-3),
                              "algorithmic_bytes": bcomp, "achieved_gbs": bcomp / (ms * 1e-3) / 1e9,
                              "hbm_peak_gbs": peak, "frac_of_measured_hbm": bcomp / (ms * 1e-3) / 1e9 / peak,
                              "gpu_launches": int(launches), "reps": args.reps}), flush=True)
        del t2
        torch.cuda.empty_cache()


def main_ranges(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd
    from fricc_b200 import ranges

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    _lib.require_cuda()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    comm = fd.Comm(rank, world)
    fd.use_torch_stream()
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
    except Exception:
        pass
    m = int(args.m)
    draw = np.array([123456789], dtype=np.int32)
    for size in args.sizes.split(","):
        n = int(float(size))
        v = int(round((n / 1600) ** 0.5))
        n = 1600 * v * v
        dims = (40, 40, v, v)
        j0, cnt = ranges.index_range(n, rank, world)
        g = torch.Generator(device=dev)
        g.manual_seed(20 + rank)
        u = torch.rand(cnt, generator=g, device=dev, dtype=torch.float64)
        z = torch.randn(cnt, generator=g, device=dev, dtype=torch.float64)
        x = (z * torch.pow(10.0, -6.0 * u * u)).contiguous()
        del u, z
        times = []
        for it in range(args.warmup + args.reps):
            L0 = _lib.lib().fricc_kernel_launches()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            comm.barrier()
            torch.cuda.synchronize()
            e0.record()
            T = ranges.compress_ranges(x, j0, dims, m, comm, "fri", "systematic", draws=draw)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if it >= args.warmup:
                times.append(float(t.item()))
            launches = _lib.lib().fricc_kernel_launches() - L0
            nnz = T.size()
            del T
        times.sort()
        ms = times[len(times) // 2]
        bcomp = 16.0 * n + 16.0 * m
        if rank == 0:
            print(json.dumps({"workload": "FRI compression only, x sharded by contiguous flat-index ranges", "N": n,
                              "m": m, "nnz": int(nnz), "sampler": "systematic", "n_gpus": world,
                              "ms_per_compression": ms, "elements_per_s": n / (ms * 1e-3), "algorithmic_bytes": bcomp,
                              "achieved_gbs": bcomp / (ms * 1e-3) / 1e9, "hbm_peak_gbs_per_gpu": peak,
                              "frac_of_measured_hbm": bcomp / (ms * 1e-3) / 1e9 / (peak * world),
                              "gpu_launches_rank0": int(launches), "reps": args.reps}), flush=True)
        del x
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
