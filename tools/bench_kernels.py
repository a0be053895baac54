#!/usr/bin/env python
"""Kernel-level timings on one GPU at the headline ladder slice (o=40, v=400, m=1e7, 20 000 of the 160 000 columns):

  * plane building from the ladder's NATIVE operand W[a,b,c,d] (no [K][L] fp64 copy): the single-read cluster kernel
    against the two-pass (column maximum + slicing) kernels, as ms and as GB/s of the algorithmic bytes
    (8 bytes read + S bytes written per element);
  * the tensor-core kernel for several plane counts S (the ratio of bytes staged per k-block to MMA work changes with
    S: 16 KB of sign tile + 8*S KB of planes against S*64 accumulator columns), which tells whether it is bound by the
    tensor pipe or by the shared-memory fill;
  * the fp64 remainder on the native layout.
One JSON line each.  python tools/bench_kernels.py [--cols 20000] [--planes 6,3,2] [--reps 3]
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
    ap.add_argument("--cols", type=int, default=20000)
    ap.add_argument("--planes", default="6,3,2")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--skip-two-pass", action="store_true")
    ap.add_argument("--debug", action="store_true", help="also time the kernels with every CTA loading tile 0 (L2-hit ceiling)")
    ap.add_argument("--bands", default="0,1000,4", help="tc_band values for the O^3V^3 term (0 = automatic, 1000 = one band)")
    ap.add_argument("--sync", default="4:2", help="lockstep variants shift:lag,...")
    a = ap.parse_args()
    a.sync = [tuple(int(x) for x in v.split(":")) for v in a.sync.split(",") if v]
    import torch

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd

    _lib.require_cuda()
    L = _lib.lib()
    check = _lib.check
    dev = torch.device("cuda", 0)
    check(L.fricc_set_device(0))
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
    na = a.cols // v
    cols = na * v
    W = torch.empty(na, v, v, v, device=dev, dtype=torch.float64)
    flat = W.view(-1)
    for s in range(0, flat.numel(), 1 << 28):
        flat[s:s + (1 << 28)].uniform_(0.0, 1.0, generator=g)
    out = torch.zeros(o, o, v, v, device=dev, dtype=torch.float64)
    outv = out[:, :, :na, :]
    i64 = _lib.i64x4
    ptr = lambda t: ctypes.c_void_p(int(t.data_ptr()))
    nK = v * v

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    for S in [int(x) for x in a.planes.split(",")]:
        handles = {}
        for fused in ((1,) if a.skip_two_pass else (1, 0)):
            check(L.fricc_set_option(b"tc_fused_planes", fused))
            h = ctypes.c_void_p()

            def build():
                nonlocal h
                if h.value:
                    check(L.fricc_wslices_destroy(h))
                    h = ctypes.c_void_p()
                check(L.fricc_wslices_create_view(ptr(W), i64(W.shape), i64(W.stride()), b"2323", 0, cols, S, ctypes.byref(h)))

            ms = timed(build, a.reps)
            alg = nK * cols * (8.0 + S)
            print(json.dumps({"kernel": "planes_fused_kernel (single read)" if fused else "colmax_view + slice_view (two passes)",
                              "planes": S, "nK": nK, "cols": cols, "ms": round(ms, 3),
                              "algorithmic_GB": round(alg / 1e9, 2), "algorithmic_GBps": round(alg / (ms * 1e-3) / 1e9, 1)}),
                  flush=True)
            handles[fused] = h
        check(L.fricc_set_option(b"tc_fused_planes", 1))
        h = handles[1]
        upd = lambda: check(L.fricc_wslices_update_view(h, ptr(W), i64(W.shape), i64(W.stride()), b"2323", 0))
        ms = timed(upd, a.reps)
        alg = nK * cols * (8.0 + S)
        print(json.dumps({"kernel": "planes_predict_kernel x2 + check (update of existing planes)", "planes": S, "nK": nK,
                          "cols": cols, "ms": round(ms, 3), "algorithmic_GBps": round(alg / (ms * 1e-3) / 1e9, 1)}), flush=True)
        if 0 in handles:
            check(L.fricc_wslices_destroy(handles[0]))

        def gemm():
            check(L.fricc_contract_planes(h, ptr(W), i64(W.shape), i64(W.stride()), 0, T._h, ptr(outv), i64(outv.shape),
                                          i64(outv.stride()), b"2323"))

        rows = o * o
        ops = 2.0 * rows * nK * cols * S
        variants = [(0, 0, 0, 4, 2), (1, 0, 0, 4, 2)] + [(1, 1, 0, sh, lag) for sh, lag in a.sync]
        if a.debug:
            variants += [(0, 0, 1, 4, 2), (1, 0, 1, 4, 2)]
        for pair, lock, dbg, sh, lag in variants:
            check(L.fricc_set_option(b"tc_pair", pair))
            check(L.fricc_set_option(b"tc_lockstep", lock))
            check(L.fricc_set_option(b"tc_debug", dbg))
            check(L.fricc_set_option(b"tc_sync_shift", sh))
            check(L.fricc_set_option(b"tc_sync_lag", lag))
            ms = timed(gemm, a.reps)
            print(json.dumps({"kernel": ("tc_sign_gemm_pair_kernel" if pair else "tc_sign_gemm_kernel") + " + remainder",
                              "lockstep": [lock, 1 << sh, lag], "all_ctas_load_tile_0": dbg, "planes": S, "rows": rows, "nK": nK, "cols": cols,
                              "ms": round(ms, 3), "int8_tops_unpadded": round(ops / (ms * 1e-3) / 1e12, 1)}), flush=True)
        check(L.fricc_set_option(b"tc_pair", 1))
        check(L.fricc_set_option(b"tc_lockstep", 1))
        check(L.fricc_set_option(b"tc_debug", 0))
        check(L.fricc_set_option(b"tc_sync_shift", 4))
        check(L.fricc_set_option(b"tc_sync_lag", 2))
        check(L.fricc_wslices_destroy(h))
    del W, out
    torch.cuda.empty_cache()
    # an O^3V^3 term through the plugin call (planes + sign matrix + tensor-core kernel + remainder), per raster band
    from fricc_b200 import py_rccsd
    Wv = torch.rand(v, o, o, v, device=dev, dtype=torch.float64, generator=g)
    tmp = torch.zeros(o, o, v, v, device=dev, dtype=torch.float64)
    py_rccsd.set_option("tc_mode", 2)
    for band in [int(x) for x in a.bands.split(",") if x]:
        check(L.fricc_set_option(b"tc_band", band))
        ms = timed(lambda: py_rccsd.contract_DTSpT(Wv, T, tmp, "1302"), a.reps)
        ops = 2.0 * (o * v) ** 3 * 6
        print(json.dumps({"term": "1302 via contract_DTSpT (planes + sign matrix + pair kernel + remainder)", "band": band,
                          "ms": round(ms, 3), "int8_tops_unpadded": round(ops / (ms * 1e-3) / 1e12, 1)}), flush=True)
    check(L.fricc_set_option(b"tc_band", 0))
    py_rccsd.set_option("tc_mode", 1)


if __name__ == "__main__":
    main()
