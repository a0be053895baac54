#!/usr/bin/env python
"""One process that launches each kernel worth a `--set full` capture once, at the headline shapes:
the range-sharded systematic compression (N = 2.56e8, m = 1e7), the first build and the predicted update of the planes
of a 20 000-column ladder slice, the pair tensor-core kernel on it, and an O^3V^3 term through the plugin call.

  ncu --set full --clock-control none --import-source on -k regex:"tc_sign_gemm_pair|planes_|escan_|key_hist|shard_" \
      -c 24 -f -o gpurun_out/prof_r02_targets python tools/profile_targets.py
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch

    from fricc_b200 import _lib, py_rccsd, ranges
    from fricc_b200 import dist as fd

    _lib.require_cuda()
    L, check = _lib.lib(), _lib.check
    dev = torch.device("cuda", 0)
    check(L.fricc_set_device(0))
    fd.use_torch_stream()
    o, v, m = 40, 400, 10_000_000
    N = o * o * v * v
    g = torch.Generator(device=dev)
    g.manual_seed(20)
    u = torch.rand(N, generator=g, device=dev, dtype=torch.float64)
    z = torch.randn(N, generator=g, device=dev, dtype=torch.float64)
    t2 = (z * torch.pow(10.0, -6.0 * u * u)).reshape(o, o, v, v).contiguous()
    del u, z
    draws = (np.array([123456789], dtype=np.int32), np.array([987654321], dtype=np.int32))
    T, T_alt = ranges.compress_ranges_pair(t2.view(-1), 0, (o, o, v, v), m, fd.Comm(0, 1), "fri", "systematic", draws=draws)
    del t2
    na = 50
    W = torch.empty(na, v, v, v, device=dev, dtype=torch.float64)
    check(L.fricc_fill_uniform(ctypes.c_void_p(W.data_ptr()), W.numel(), 0, 1000))
    i64 = _lib.i64x4
    ptr = lambda t: ctypes.c_void_p(int(t.data_ptr()))
    h = ctypes.c_void_p()
    check(L.fricc_wslices_create_view(ptr(W), i64(W.shape), i64(W.stride()), b"2323", 0, na * v, 6, ctypes.byref(h)))   # planes_fused_kernel
    check(L.fricc_wslices_update_view(h, ptr(W), i64(W.shape), i64(W.stride()), b"2323", 0))                            # planes_predict_kernel x2
    out = torch.zeros(o, o, v, v, device=dev, dtype=torch.float64)
    ov = out[:, :, :na, :]
    check(L.fricc_contract_planes(h, ptr(W), i64(W.shape), i64(W.stride()), 0, T._h, ptr(ov), i64(ov.shape), i64(ov.stride()), b"2323"))
    check(L.fricc_wslices_destroy(h))
    del W
    torch.cuda.empty_cache()
    Wv = torch.rand(v, o, o, v, device=dev, dtype=torch.float64, generator=g)
    py_rccsd.set_option("tc_mode", 2)
    py_rccsd.contract_DTSpT(Wv, T, out, "1302")
    torch.cuda.synchronize()
    print("profile targets done")


if __name__ == "__main__":
    main()
