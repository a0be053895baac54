#!/usr/bin/env python
"""Diagnostic (ONE GPU): can the peer-store kernels (fricc_peer.cu) run UNDER the tensor-core contraction kernel?

Emulates rank 0 of 8 at the headline shape: the 0312-class term (16 000 x 16 000 sign matrix x 2 000 columns, 6 digit
planes) runs on the compute stream while `fricc_push_columns` stores a 2 GB output's columns into 7 "peer" copies on a
side stream (the copies are ordinary buffers of this device: same kernel, same addressing, HBM instead of NVLink).
Reports, for each (push_ctas_per_sm, push_carveout): the contraction alone, the push alone, both serialised on one stream
and both on two streams.  overlapped ~ max(alone) => the kernels share the SMs; overlapped ~ sum => they take turns."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from fricc_b200 import _lib  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402
from fricc_b200.py_rccsd import SparseTensor4d  # noqa: E402


def main():
    _lib.require_cuda()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    fd.use_torch_stream()
    L = _lib.lib()
    o, v, m, world = 40, 400, 10_000_000, 8
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    n = o * o * v * v
    t2 = (torch.randn(n, generator=g, device=dev, dtype=torch.float64) *
          torch.pow(10.0, -6.0 * torch.rand(n, generator=g, device=dev, dtype=torch.float64) ** 2)).reshape(o, o, v, v)
    T = SparseTensor4d(t2, (o, o, v, v), m, "fri", "systematic", False, draws=np.array([99], dtype=np.int32))
    del t2
    st = fd.ShardedTerm("0312", o, v, fd.Comm(0, world), dev)
    W = torch.rand(o, v, o, v, generator=g, device=dev, dtype=torch.float64)
    st.auto_planes = 6
    st.permute_from(W)
    out = torch.zeros(v, o, v, o, dtype=torch.float64, device=dev)
    pushed = torch.rand(o, o, v, v, generator=g, device=dev, dtype=torch.float64)       # a finished 2 GB output
    copies = [pushed] + [torch.empty_like(pushed) for _ in range(world - 1)]
    peers = (ctypes.c_void_p * world)(*[int(c.data_ptr()) for c in copies])
    side = fd.SideStream(dev)

    def contraction():
        st.local_into(T, out)

    def push():
        fd.check(L.fricc_push_columns(fd._ptr(pushed), peers, world, 0, b"2323", fd.i64x4((o, o, v, v)),
                                      fd.i64x4(pushed.shape), fd.i64x4(pushed.stride()), 0, v * v))   # ALL columns: 14 GB of stores, ~ the contraction's time

    def timed(fn, reps=5):
        ts = []
        for _ in range(reps + 1):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return round(min(ts[1:]), 3)

    def both_one_stream():
        push()
        contraction()

    def both_two_streams():
        side.run(push)
        contraction()
        side.join()

    contraction()
    for cap, carve in ((64, -1), (3, -1), (3, 100), (2, 100), (4, 100)):
        fd.check(L.fricc_set_option(b"push_ctas_per_sm", cap))
        fd.check(L.fricc_set_option(b"push_carveout", carve))
        print(json.dumps({"push_ctas_per_sm": cap, "push_carveout": carve, "contraction_ms": timed(contraction),
                          "push_ms": timed(push), "serial_ms": timed(both_one_stream),
                          "two_streams_ms": timed(both_two_streams)}), flush=True)


if __name__ == "__main__":
    main()
