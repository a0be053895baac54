#!/usr/bin/env python
"""Diagnostic: time the peer-memory pushes (fricc_peer.cu) per output map on N GPUs, next to an NCCL all-gather of the
same volume.   torchrun --nproc-per-node N tools/bench_peer.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from fricc_b200 import _lib
from fricc_b200 import dist as fd


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    _lib.check(_lib.lib().fricc_set_device(lr))
    dist.init_process_group("nccl", device_id=dev)
    comm = fd.Comm(rank, world)
    fd.use_torch_stream()
    o, v = 40, 400
    arena = fd.SymmArena(8 * (3 * o * o * v * v) + 65536, comm, dev)
    shapes = {"2323": (o, o, v, v), "0112": (v, o, o, v), "0312": (v, o, v, o), "0101": (o, o, v, v), "1302": (o, o, v, v),
              "1203": (o, o, v, v)}
    bufs = {}
    for shp in set(shapes.values()):
        bufs[shp] = arena.tensor(shp).zero_()
    flag = torch.zeros(1, device=dev)
    res = {}
    sweep = "--sweep" in sys.argv   # push throughput against the grid cap of the peer-store kernels
    caps = (1, 2, 4, 8, 16) if sweep else (None,)
    if sweep:
        shapes = {k: shapes[k] for k in ("2323", "1302")}
    for cap in caps:
      if cap is not None:
        _lib.check(_lib.lib().fricc_set_option(b"push_ctas_per_sm", cap))
      for term, shp in shapes.items():
        out = bufs[shp]
        st = fd.ShardedTerm(term, o, v, comm, dev)
        st.out_panel = torch.rand(st.nR * st.lc, dtype=torch.float64, device=dev)
        for mode in ("push", "scatter"):
            ts = []
            for rep in range(3):
                dist.all_reduce(flag)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                if mode == "push":
                    fd.push_columns(st, (o, o, v, v), out, arena)
                else:
                    fd.check(st.L.fricc_scatter_panel(fd._ptr(st.out_panel), st.lc, st.l0, st.lcnt, term.encode(),
                                                      fd.i64x4((o, o, v, v)), fd._ptr(out), arena.peers_of(out), world, rank,
                                                      fd.i64x4(out.shape), fd.i64x4(out.stride())))
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            res[f"{term}_{mode}" + (f"_cap{cap}" if cap is not None else "")] = round(min(ts), 3)
    # NCCL all-gather of one output's worth of panels
    n = o * o * v * v
    src = torch.rand(n // world, dtype=torch.float64, device=dev)
    dst = torch.empty(n, dtype=torch.float64, device=dev)
    ts = []
    for rep in range(3):
        dist.all_reduce(flag)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dist.all_gather_into_tensor(dst, src)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res["nccl_all_gather_2GB"] = round(min(ts), 3)
    t = torch.tensor(list(res.values()), device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"world": world, "ms_max_over_ranks": dict(zip(res.keys(), [round(x, 3) for x in t.tolist()])),
                          "bytes_pushed_per_rank": 8 * n // world * (world - 1)}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
