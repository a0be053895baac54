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
