"""Experiment: does a column-blocked panel layout [tile][K][512] beat the row-major panel [K][cols] for the ladder?"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from fricc_b200 import _lib, py_rccsd  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402

o, v, m = 40, 400, 10_000_000
dev = torch.device("cuda", 0)
fd.use_torch_stream()
L = _lib.lib()
g = torch.Generator(device=dev)
g.manual_seed(1)
N = o * o * v * v
t2 = (torch.randn(N, generator=g, device=dev, dtype=torch.float64) *
      torch.pow(10.0, -6.0 * torch.rand(N, generator=g, device=dev, dtype=torch.float64) ** 2)).reshape(o, o, v, v)
T = py_rccsd.SparseTensor4d(t2, (o, o, v, v), m, "fri", "systematic", False, draws=[12345])
del t2
nK, nR = v * v, o * o
py_rccsd.set_option("spmm_nv2", 0)


def ptr(t, off=0):
    import ctypes
    return ctypes.c_void_p(int(t.data_ptr()) + 8 * off)


def timeit(fn, reps=3):
    ts = []
    for r in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if r:
            ts.append(e0.elapsed_time(e1))
    return min(ts)


for cols in (4096, 8192, 16384, 24576):
    panel = torch.rand(nK * cols, generator=g, device=dev, dtype=torch.float64)
    out = torch.zeros(nR * cols, device=dev, dtype=torch.float64)
    fm = m * cols
    ms = timeit(lambda: _lib.check(L.fricc_contract_panel(ptr(panel), cols, 0, cols, T._h, ptr(out), cols, b"2323")))
    print(f"cols={cols:6d} row-major [K][cols]      : {ms:8.2f} ms  {fm/ms/1e9:5.2f} TFMA/s")
    for tc in (512, 2048):
        def blocked():
            for t in range(cols // tc):
                _lib.check(L.fricc_contract_panel(ptr(panel, t * nK * tc), tc, t * tc, tc, T._h, ptr(out, t * tc), cols, b"2323"))
        ms = timeit(blocked)
        print(f"cols={cols:6d} blocked  [tile][K][{tc:4d}] : {ms:8.2f} ms  {fm/ms/1e9:5.2f} TFMA/s  ({cols//tc} launches)")
    del panel, out
