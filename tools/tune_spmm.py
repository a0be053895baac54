"""GPU tuning harness for the ladder kernel: times fricc_contract_panel for the SpMM launch variants at a mid-size
problem with the headline density (3.9 %), and checks the variants against each other.
   python tools/tune_spmm.py [o v m cols]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from fricc_b200 import _lib, py_rccsd  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402

o, v, m, cols = 40, 400, 10_000_000, 8192
if len(sys.argv) >= 5:
    o, v, m, cols = (int(float(a)) for a in sys.argv[1:5])
dev = torch.device("cuda", 0)
fd.use_torch_stream()
L = _lib.lib()
g = torch.Generator(device=dev)
g.manual_seed(1)
N = o * o * v * v
t2 = (torch.randn(N, generator=g, device=dev, dtype=torch.float64) *
      torch.pow(10.0, -6.0 * torch.rand(N, generator=g, device=dev, dtype=torch.float64) ** 2)).reshape(o, o, v, v)
T = py_rccsd.SparseTensor4d(t2, (o, o, v, v), m, "fri", "systematic", False, draws=[12345])
nK, nR = v * v, o * o
panel = torch.rand(nK * cols, generator=g, device=dev, dtype=torch.float64)
out = torch.zeros(nR * cols, device=dev, dtype=torch.float64)
alg = 8.0 * m * cols + 8.0 * nR * cols + 12.0 * m
print(f"o={o} v={v} m={m} cols={cols}: panel {panel.numel()*8/1e9:.2f} GB, algorithmic {alg/1e9:.1f} GB, FMAs {m*cols:.3e}")


def run(nv2, threads, reps=5):
    py_rccsd.set_option("spmm_nv2", nv2)
    py_rccsd.set_option("spmm_threads", threads)
    out.zero_()
    ts = []
    for r in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(L.fricc_contract_panel(fd._ptr(panel), cols, 0, cols, T._h, fd._ptr(out), cols, b"2323"))
        e1.record()
        torch.cuda.synchronize()
        if r:
            ts.append(e0.elapsed_time(e1))
    return min(ts), out.clone()


base = None
for nv2, threads in ((0, 64), (0, 32), (0, 16), (2, 64), (1, 64), (2, 32), (4, 32), (1, 32), (2, 64), (0, 64)):
    ms, res = run(nv2, threads)
    if base is None:
        base = res
    err = float((res - base).abs().max() / base.abs().max())
    print(f"nv2={nv2} threads={threads:3d}: {ms:8.3f} ms  {alg/ms/1e6:8.1f} GB/s algorithmic  "
          f"{m*cols/ms/1e9:6.2f} TFMA/s  rel.diff vs base {err:.1e}")
