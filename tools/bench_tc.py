#!/usr/bin/env python
"""A/B of the two contraction kernels on one GPU at the headline shapes: fp64 bucketed SpMM vs the tensor-core form
(int8 sign matrix x digit planes).  Prints one JSON line per (term, kernel).

  python tools/bench_tc.py [--o 40 --v 400 --m 1e7 --cols 20000 --planes 7 --terms 2323,1302]
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--o", type=int, default=40)
    ap.add_argument("--v", type=int, default=400)
    ap.add_argument("--m", type=float, default=1e7)
    ap.add_argument("--cols", type=int, default=20000, help="columns of the ladder slice held by this GPU")
    ap.add_argument("--planes", type=int, default=6)
    ap.add_argument("--terms", default="2323,1302")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--check", action="store_true", help="compare the two kernels' outputs")
    a = ap.parse_args()
    import torch

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd

    _lib.require_cuda()
    dev = torch.device("cuda", 0)
    _lib.check(_lib.lib().fricc_set_device(0))
    fd.use_torch_stream()
    o, v, m = a.o, a.v, int(a.m)
    N = o * o * v * v
    g = torch.Generator(device=dev)
    g.manual_seed(20)
    u = torch.rand(N, generator=g, device=dev, dtype=torch.float64)
    z = torch.randn(N, generator=g, device=dev, dtype=torch.float64)
    t2 = (z * torch.pow(10.0, -6.0 * u * u)).reshape(o, o, v, v).contiguous()
    del u, z
    T = fd.compress(t2, m, "fri", "systematic", draws=None)
    del t2
    for term in a.terms.split(","):
        world = max(1, (v * v) // a.cols) if term == "2323" else 8
        st = fd.ShardedTerm(term, o, v, fd.Comm(0, world), dev)
        st.panel = torch.empty(st.nK * st.lcnt, dtype=torch.float64, device=dev)
        step = 1 << 28
        for s in range(0, st.panel.numel(), step):
            st.panel[s:s + step].uniform_(0.0, 1.0, generator=g)
        out = torch.zeros(st.nR * st.lc, dtype=torch.float64, device=dev)
        res = {}
        for kind in ("fp64_spmm", "tensor_core", "tensor_core_one_cta_per_tile"):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            _lib.check(_lib.lib().fricc_set_option(b"tc_persistent", 0 if kind.endswith("per_tile") else 1))
            if kind.startswith("tensor_core"):
                ev[0].record()
                st.build_planes(a.planes)
                ev[1].record()
                st.tc_min_fraction = 0.5
                assert st.use_tc(T), "T has no common magnitude?"
            else:
                st.drop_planes()
            def run():
                st.out_panel = out
                out.zero_()
                return st_local(st, T, out)
            run()
            torch.cuda.synchronize()
            ev[2].record()
            for _ in range(a.reps):
                run()
            ev[3].record()
            torch.cuda.synchronize()
            ms = ev[2].elapsed_time(ev[3]) / a.reps
            line = {"term": term, "kernel": kind, "o": o, "v": v, "m": m, "rows": st.nR, "cols": st.lcnt, "nK": st.nK,
                    "ms": ms, "fma_per_s": m * st.lcnt / (ms * 1e-3)}
            if kind.startswith("tensor_core"):
                macs = (-(-st.nR // 128) * 128) * (-(-st.lcnt // 64) * 64) * (-(-st.nK // 128) * 128) * a.planes
                b = ctypes.c_int64(0)
                _lib.check(_lib.lib().fricc_wslices_bytes(st.planes, ctypes.byref(b)))
                line.update(planes=a.planes, plane_build_ms=ev[0].elapsed_time(ev[1]), int8_macs=macs,
                            int8_tops=2 * macs / (ms * 1e-3) / 1e12, plane_bytes=b.value)
            if a.check:
                res[kind] = out.clone()
            print(json.dumps(line), flush=True)
        if a.check:
            _lib.check(_lib.lib().fricc_set_option(b"tc_persistent", 1))
            d = (res["fp64_spmm"] - res["tensor_core"]).abs().max().item()
            print(json.dumps({"term": term, "max_abs_diff": d, "max_abs": res["fp64_spmm"].abs().max().item()}), flush=True)
        st.drop_planes()
        del st, out, res
        torch.cuda.empty_cache()


def st_local(st, T, out):
    from fricc_b200.dist import _ptr, check

    if st.planes is not None:
        check(st.L.fricc_contract_panel_tc(st.planes, _ptr(st.panel), st.lcnt, st.l0, T._h, _ptr(out), st.lc,
                                           st.term.encode(), 0))
    else:
        check(st.L.fricc_contract_panel(_ptr(st.panel), st.lcnt, st.l0, st.lcnt, T._h, _ptr(out), st.lc, st.term.encode()))


if __name__ == "__main__":
    main()
