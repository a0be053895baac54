"""Host-side pieces of bench.py that the driver's contract rests on (no GPU): defaults of the command line, the clock
sampler's window filter, the operation-count models behind the reference arm's extrapolation."""
import importlib.util
import os
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _args(bench, argv):
    old = sys.argv
    sys.argv = ["bench.py"] + argv
    try:
        return bench.parse_args()
    finally:
        sys.argv = old


def test_defaults_are_the_contract(bench):
    a = _args(bench, [])
    assert a.gpus == 1 and a.steps >= 1 and a.warmup >= 3 and a.impl == "b200"
    assert (a.o, a.v, int(a.m), a.sampling) == (40, 400, 10_000_000, "systematic")          # BASELINE.json's config
    assert a.w_mode == "refreshed" and a.overlap_planes and not a.no_constant_eris
    assert not _args(bench, ["--no-overlap-planes"]).overlap_planes
    d = _args(bench, ["--gpus", "8", "--steps", "20", "--warmup", "5", "--impl", "reference"])
    assert (d.gpus, d.steps, d.warmup, d.impl) == (8, 20, 5, "reference")
    assert bench.metric_name(40, 400, 1e7) == "FRI-CCSD s/iter at o=40,v=400,m=1e7" and bench.UNIT == "s/iter"


def test_clock_samples_are_cut_to_the_timed_window(bench, tmp_path):
    """nvidia-smi lines: sm, max sm, power, four throttle reasons, timestamp.  Only the samples inside the window count;
    unreadable timestamps fall back to all samples."""
    import datetime
    import subprocess

    now = time.time()
    fmt = lambda t: datetime.datetime.fromtimestamp(t).strftime("%Y/%m/%d %H:%M:%S.%f")[:-3]
    rows = [(now - 10.0, 345, "Not Active"), (now - 9.5, 400, "Not Active"),             # warm-up: outside
            (now - 5.0, 1700, "Active"), (now - 4.8, 1650, "Active"), (now - 4.6, 1800, "Not Active"),
            (now + 30.0, 210, "Not Active")]                                            # after the region: outside
    path = tmp_path / "clocks.csv"
    path.write_text("".join(f"{sm}, 1965, 900.1, Not Active, Not Active, Not Active, {cap}, {fmt(t)}\n" for t, sm, cap in rows))
    proc = subprocess.Popen([sys.executable, "-c", "import time; time.sleep(30)"])
    got = bench.stop_clock_sampler(proc, str(path), (now - 5.1, now - 4.5))
    assert got["samples"] == 3 and got["sm_max_mhz"] == 1965.0 and got["reasons"] == ["sw_power_cap"]
    assert got["sm_mhz"] in (1700.0, 1800.0)            # median of the upper half of the in-window samples
    assert bench._smi_time("2026/11/05 01:23:45.678") is not None and bench._smi_time("garbage") is None
    path2 = tmp_path / "clocks2.csv"
    path2.write_text("1500, 1965, 800.0, Not Active, Not Active, Not Active, Not Active, n/a\n")
    proc = subprocess.Popen([sys.executable, "-c", "import time; time.sleep(30)"])
    got = bench.stop_clock_sampler(proc, str(path2), (now - 1.0, now))
    assert got["samples"] == 1 and got["sm_mhz"] == 1500.0 and got["reasons"] == []
    assert bench.stop_clock_sampler(None, "", None)["reasons"] == ["nvidia-smi unavailable"]


def test_operation_count_models(bench):
    """FMAs per iteration: ladder m v^2, six O^3V^3 ... terms m o v each, two O^4V^2 terms m o^2."""
    f = bench.flops_model(40, 400, 10_000_000)
    total = f if isinstance(f, (int, float)) else sum(f.values()) if isinstance(f, dict) else sum(f)
    m, o, v = 1e7, 40, 400
    assert total == pytest.approx(m * v * v + 8 * m * o * v + 2 * m * o * o, rel=1e-12)


def test_spot_check_recomputes_every_output_from_the_term_definitions():
    """bench.spot_check_outputs on a tiny CPU stand-in of the bench engine: outputs built with np.einsum from the dense
    form of T / T_alt pass at rounding level, and a single wrong element anywhere in an output is flagged when hit."""
    import types

    import numpy as np
    import torch

    import bench

    o, v = 3, 5
    rng = np.random.default_rng(3)

    class FakeT:
        def __init__(self, n_entries):
            self.flat = rng.choice(o * o * v * v, size=n_entries, replace=False).astype(np.int64)
            self.vals = rng.standard_normal(n_entries)

        def export(self):
            return self.flat, self.vals

        def dense(self):
            d = np.zeros(o * o * v * v)
            d[self.flat] = self.vals
            return d.reshape(o, o, v, v)

    T, T_alt = FakeT(90), FakeT(70)
    W = bench.fill_uniform_host(np.arange(v ** 4, dtype=np.uint64), 1000).reshape(v, v, v, v)
    assert 0.0 <= W.min() and W.max() < 1.0 and len(np.unique(W)) == v ** 4
    ovov = rng.random((o, v, o, v))
    d, da = T.dense(), T_alt.dense()

    def ein(term, Wop, Td):
        pref, w_idx, t_idx, o_idx = bench.SPOT_TERMS[term]
        return pref * np.einsum(f"{w_idx},{t_idx}->{o_idx}", Wop, Td)

    Wklij = ein("1323", ovov, da)
    Wakic = ein("0112", ovov, da) + ein("0113", ovov, da) + ein("0313", ovov, da)
    Wakci = ein("0312", ovov, da)
    t2new = ein("2323", W, d) + ein("0101", Wklij, d)
    tmp = [ein("1302", Wakic, d) + ein("1202", Wakci, d), ein("1303", Wakic, d), ein("1203", Wakci, d)]
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    it = types.SimpleNamespace(o=o, v=v, last_pair=(T, T_alt), ovov=tt(ovov), Wklij=tt(Wklij), Wakic=tt(Wakic),
                               Wakci=tt(Wakci), t2new=tt(t2new), tmp=[tt(x) for x in tmp])
    res = bench.spot_check_outputs(it, torch, n_t2new=40, n_other=20)
    assert res["points"] == 40 + 6 * 20
    assert res["max_rel_to_abs_sum"] < 1e-14
    it.tmp[1] = it.tmp[1] + 1e-6                      # every element off: must be seen
    bad = bench.spot_check_outputs(it, torch, n_t2new=8, n_other=8)
    assert bad["by_output"]["tmp1"]["max_rel_to_abs_sum"] > 1e-9
    assert bad["by_output"]["tmp0"]["max_rel_to_abs_sum"] < 1e-14
