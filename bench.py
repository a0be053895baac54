#!/usr/bin/env python
"""bench.py -- one FRI-CCSD inner-loop iteration (the hot path of jamesETsmith/fri-cc) on synthetic tensors.

Metric (BASELINE.json): FRI-CCSD s/iter at o=40, v=400, m=1e7 on 1/2/4/8 GPUs; ladder roofline.

A "step" is one pass of the hot path as fricc/friccsd.py:update_amps drives it with all three contraction classes
compressed: 2 FRI compressions of T2 (N = o^2 v^2 entries -> m) and the 11 sparse contractions
  1323 (ovov -> Woooo), 0112 0313 0113 (ovov -> Wvoov), 0312 (ovov -> Wvovo),
  0101 (Woooo), 2323 (Wvvvv: the particle-particle ladder), 1302 1303 (Wvoov), 1202 1203 (Wvovo)   -> t2new / tmp.
Two pairs that go into one array back to back run as one dense product each (0112 + 0113, 1302 + 1202:
py_rccsd.contract_DTSpT_pair, an extension of the reference's surface; --no-fuse issues the eleven plain calls).

Operand upkeep -- `--w-mode` (both values are in the line, `value` is the one selected):
  refreshed (default) the FRICCSD regime: Wvvvv = cc_Wvvvv(t1, t2, eris) is rebuilt EVERY iteration
            (fricc/friccsd.py:207,227,325-331), so the re-encoding of the ladder operand into int8 digit planes is part
            of every step and is timed.  The operand is read in its native W[a,b,c,d] layout (no [cd][ab] copy exists
            on this path any more).  Producing W itself is the caller's dense work and is untimed.
  invariant the FRILCCSD regime: W is the bare integral block, the same every iteration (fricc/frilccsd.py:171-179);
            planes are built once, outside the timed region.
  In both modes the ladder's first sub-shard is encoded on a second stream BESIDE the compression of T2 (W does not
  depend on T2; --no-overlap-planes for the A/B).  Woooo / Wvoov / Wvovo are outputs of this very step and are
  re-encoded inside it in both modes.  The ovov integrals never change between iterations in either driver
  (fricc/friccsd.py:316,348-356,372 pass eris.ovov to five of the eleven terms): they are declared constant
  (py_rccsd.constant_operand), so their planes are built in the first warm-up step and kept (--no-constant-eris for
  the A/B).

Sharding (N > 1, strong scaling): every contraction is cut along its free index pair; a rank hands the plugin call
`py_rccsd.contract_DTSpT` (= fricc_contract) its slice of the operand and of the output as plain 4-D views
(fricc_b200/dist.py: ViewEngine) -- the library picks the kernel -- and stores the finished slice into every peer's copy
over NVLink (symmetric allocation, stores on a side stream under the next term's kernel, two device-side barriers
per step, no NCCL on the data path).  The step's two compressions come from one range-sharded pass.

Timing: W >= 3 untimed warm-up steps, then K steps; barrier + torch.cuda.synchronize on both sides; CUDA events on the
launching stream; max over ranks.  Inputs are far larger than L2 (>= 25 GB of W per GPU per step).  When the ladder's W
slice does not fit in HBM next to its planes (1 GPU: 204.8 GB; 2 GPUs: 102.4 GB + 76.8 GB of planes) it is processed in
sub-shards that are re-generated in HBM between timed segments; the step time is the sum of the timed segments.

`e2e`: the same step through the plugin calls with T2 arriving from pinned host memory and the residual returning to
it inside the timed region.  1 GPU: one upload of T2, `SparseTensor4d(t2, ...)` twice and `contract_DTSpT` eleven
times -- the calls fricc/friccsd.py makes (W and the intermediates are device tensors: Wvvvv is 204.8 GB, more than
the host has) --, t2new copied back while the last four contractions run.  N > 1: every rank uploads its 1/N of T2,
the range-sharded compression needs nothing else, the contractions are the same plugin calls on views, and every rank
reads back the 1/N of the residual it owns (the i-range its own 0101 call completes) on a copy stream under the
remaining contractions.

Measurement hygiene: the nvidia-smi clock sampler starts before the warm-up and its samples are cut to the timed
window; `stage_ms_slowest_step` / `stage_host_issue_ms_of_that_step` show the slowest instance of every stage next to
the host time spent issuing it; `scratch_allocations_in_timed_region` counts the library's driver allocation calls
(0 in steady state: freed scratch blocks are cached, see fricc_runtime.cu -- a cudaMallocAsync that has to map memory
queues behind nvidia-smi's queries and used to stall single steps by up to 2.7 s).

`--impl reference`: the reference's own CPU implementation (oracle/_ref, the unmodified sources compiled by
oracle/Makefile) on the box's host cores: the full workload when it fits the host (e.g. --o 21 --v 93 --m 1e6),
otherwise a set of sample points and a per-term fit  t = alpha * (FMA count) + beta * (index tests)  extrapolated to
(o, v, m), with the measured points and the fit residuals in the line.
"""
import argparse
import ctypes
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

UNIT = "s/iter"
CONTRACTION_ORDER = ["1323", "0112", "0313", "0113", "0312", "0101", "2323", "1302", "1202", "1303", "1203"]
# checksum of t2new after one step of the default workload (o=40, v=400, m=1e7, systematic, seed 20): the same on
# every rank and for every GPU count (each output element is computed by one rank in a fixed order)
EXPECTED_T2NEW_CHECKSUM = 3286248761.738817


def metric_name(o, v, m):
    return f"FRI-CCSD s/iter at o={o},v={v},m={m:.0e}".replace("e+0", "e").replace("e+", "e")


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--o", "--nocc", dest="o", type=int, default=40)
    p.add_argument("--v", "--nvir", dest="v", type=int, default=400)
    p.add_argument("--m", "--nkeep", dest="m", type=float, default=1e7)
    p.add_argument("--sampling", default="systematic", choices=["systematic", "pivotal"])
    p.add_argument("--w-mode", default="refreshed", choices=["refreshed", "invariant"],
                   help="refreshed: the ladder operand is re-encoded inside every step (FRICCSD rebuilds Wvvvv each "
                        "iteration); invariant: planes built once (FRILCCSD)")
    p.add_argument("--max-slice-gb", type=float, default=26.0, help="largest resident ladder W sub-shard (fp64 GB)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-pairs", action="store_true",
                   help="skip the small configurations both arms run in full (measured_configs / same_config_pairs)")
    p.add_argument("--no-spot-check", action="store_true", help="skip the host fp64 recomputation of random output elements")
    p.add_argument("--no-clocks", action="store_true", help="diagnostic: do not run the nvidia-smi clock sampler")
    p.add_argument("--clock-period-ms", type=int, default=200, help="sampling period of the nvidia-smi clock sampler")
    p.add_argument("--stall-trace", type=float, default=0.0,
                   help="diagnostic: dump the main thread's Python stack to stderr whenever it has sat in one place for "
                        "this many milliseconds outside the compression (finds a host-side stall)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-sweep", action="store_true", help="skip the compression-only sweep (config 5)")
    p.add_argument("--planes", type=int, default=6,
                   help="int8 digit planes of W for the tensor-core form of the contractions (0 = fp64 SpMM only)")
    p.add_argument("--last-term-parts", type=int, default=2,
                   help="N>1: the last term before the closing barrier is issued in this many parts, each part's peer stores "
                        "running under the next part's kernel")
    p.add_argument("--push-copy-engine", type=int, default=0,
                   help="N>1: peer stores of long-row slices through cudaMemcpy(2D)Async instead of the store kernel (A/B): "
                        "1 = on the side stream, 2 = one private stream per peer (see fricc_peer.cu)")
    p.add_argument("--uniform-carveout", type=int, default=-1,
                   help="N>1: every library kernel asks for this L1/shared split (percent shared; 100 = the split of the "
                        "tensor-core and peer-store kernels), so that the short kernels of a term can share SMs with the "
                        "peer stores of the previous one (A/B; -1 = the driver's choice per kernel)")
    p.add_argument("--overlap-planes", dest="overlap_planes", action="store_true", default=True,
                   help="(default) encode the ladder's first sub-shard on a second, lower-priority stream beside the "
                        "compression instead of after it -- W does not depend on T2; the encoder runs one 64 KB unit per "
                        "block so that the compression's kernels get in between (8 GPUs: 0.0680 -> 0.0639 s/iter)")
    p.add_argument("--no-overlap-planes", dest="overlap_planes", action="store_false", help="A/B: encode after the compression")
    p.add_argument("--push-inline", action="store_true",
                   help="N>1: issue the peer stores on the compute stream (A/B of the side-stream push)")
    p.add_argument("--no-constant-eris", action="store_true",
                   help="A/B: do not declare the ovov integrals constant (their digit planes are then rebuilt in every step)")
    p.add_argument("--no-fuse", action="store_true",
                   help="issue 0112/0113 and 1302/1202 as the eleven separate plugin calls of the reference's driver instead of "
                        "the two fused pairs (py_rccsd.contract_DTSpT_pair)")
    p.add_argument("--as-rank-of", type=int, default=0,
                   help="profiling aid (1 GPU): do rank 0's share of the contractions of an N-rank job, without the exchange; "
                        "the line is marked and is not a benchmark value")
    p.add_argument("--nccl-barrier", action="store_true", help="N>1: 4-byte NCCL all-reduce barriers instead of the "
                                                               "device-side flag barrier (A/B)")
    return p.parse_args()


def flops_model(o, v, m):
    """FMAs per term at (o, v, m)."""
    f = {"2323": m * v * v, "0101": m * o * o, "1323": m * o * o}
    for t in ("1302", "1202", "1303", "1203", "0112", "0313", "0113", "0312"):
        f[t] = m * o * v
    return f


def scan_model(o, v, m):
    """Index tests per term: every owner pair scans all m stored entries (src/sparse_tensor.cpp:62-68)."""
    s = {"2323": o * o * m, "1323": o * o * m, "0101": v * v * m}
    for t in ("1302", "1202", "1303", "1203", "0112", "0313", "0113", "0312"):
        s[t] = o * v * m
    return s


def workload_config(args, world):
    o, v, m = args.o, args.v, int(args.m)
    regime = ("Wvvvv re-encoded every step (FRICCSD, fricc/friccsd.py:207,325-331)" if args.w_mode == "refreshed" else
              "Wvvvv iteration-invariant, planes built once (FRILCCSD, fricc/frilccsd.py:171-179)")
    return {
        "workload": f"synthetic FRI-CCSD iteration o={o},v={v},m={m}: 2 FRI compressions ({args.sampling}) of T2 "
                    f"(N={o*o*v*v}) + 11 sparse contractions; {regime}; vvvv operand sharded by (a,b) columns over "
                    f"{world} GPU(s)",
        "o": o, "v": v, "m": m, "sampling": args.sampling, "w_mode": args.w_mode,
        "parallelism": f"free-pair column shards x{world}", "l2": "inputs_larger_than_L2",
    }


# ======================================================================================================================
# reference arm / CPU baseline: the reference's own OpenMP implementation on the host cores
# ======================================================================================================================
def _ref_modules(threads):
    os.environ["OMP_NUM_THREADS"] = str(threads)
    from oracle import oracle

    R = oracle.ref()
    if R is not None:
        R.set_num_threads(threads)
    return R, oracle.port()


def _time_point(R, P, o, v, m_comp, m_con, sampling, terms, rng, n_comp=2):
    """Seconds for `n_comp` compressions at (o, v, m_comp) and each contraction term in `terms` at (o, v, m_con), on an
    FRI-compressed tensor (the layout the drivers hand over)."""
    import numpy as np

    from fricc_b200.dist import TERM_AXES

    n = o * o * v * v
    t2 = (rng.standard_normal(n) * 10.0 ** (-6.0 * rng.random(n) ** 2)).reshape(o, o, v, v)
    out = {}
    if n_comp:
        t0 = time.perf_counter()
        for _ in range(n_comp):
            if R is not None:
                T = R.mod.SparseTensor4d(t2, list(t2.shape), int(m_comp), "fri", sampling, False)
            else:
                P.fri_compression(t2, int(m_comp), sampling, rng.integers(0, 2**31 - 1, size=n).astype(np.int32))
        out["compress"] = (time.perf_counter() - t0) / n_comp
    if terms:
        if R is not None:
            T = R.mod.SparseTensor4d(t2, list(t2.shape), int(m_con), "fri", "systematic", False)
        else:
            cmp_ = P.fri_compression(t2, int(m_con), "systematic", np.array([12345], dtype=np.int32))
            flat, vals = cmp_["idx"], cmp_["vals"]
        ext = {"o": o, "v": v}
        for term in terms:
            ax = TERM_AXES[term]
            W = rng.random([ext[c] for c in ax["w"]])
            dst = np.zeros([ext[c] for c in ax["out"]])
            t0 = time.perf_counter()
            if R is not None:
                R.mod.contract_DTSpT(W, T, dst, term)
            else:
                P.contract(term, W, flat, vals, t2.shape, dst)
            out[term] = time.perf_counter() - t0
    return out


# host-feasible configurations that BOTH arms run in full: a measured, same-config pair beside the extrapolated headline
# (BASELINE.json config 2's benzene shape, and the largest o = 40 shape whose 11 terms the host finishes in about a minute)
SAME_CONFIG_PAIRS = [(21, 93, 1_000_000), (40, 120, 1_000_000)]


def measure_full_config(R, P, o, v, m, sampling):
    """The reference's CPU path RUN IN FULL at (o, v, m): 2 x SparseTensor4d(fri) + the 11 contract_DTSpT terms, no model."""
    import numpy as np

    pt = _time_point(R, P, o, v, m, m, sampling, CONTRACTION_ORDER, np.random.default_rng(2), n_comp=2)
    total = 2 * pt["compress"] + sum(pt[t] for t in CONTRACTION_ORDER)
    return {"o": o, "v": v, "m": int(m), "sampling": sampling, "cpu_s": total, "same_config": True,
            "parts_s": {k: round(x, 4) for k, x in pt.items()}}


def cpu_reference_sample(o, v, m, sampling, threads, scale=1.0):
    """One bounded sample of the reference's CPU path -> seconds per FRI-CCSD iteration at (o, v, m).

    If the workload itself is small enough for the host (W = 8 v^4 bytes <= 4 GB and <= ~3e10 FMAs) it is RUN as it is:
    2 compressions + the 11 terms at (o, v, m) -- `same_config` true, no model.  Otherwise sample points at o' = min(o, 40),
    v' in {60, 80, 100, 120} and two m' are timed and each part is fitted and extrapolated:
      ladder 2323      t = a m v^2 + b o^2 m + c  (FMAs + the o^2 m index tests of the owner scan, sparse_tensor.cpp:62-73)
      0101             t = a m o^2 + b v^2 m + c
      1323, O^3V^3     t = a * (FMAs + index tests)   (both scale alike: one coefficient)
      compression      t = g N log2 N               (the full merge sort of N pairs is 77 % of it)
    `scale` < 1 shrinks the sample (warm-up steps)."""
    import numpy as np

    R, P = _ref_modules(threads)
    kind = "reference" if R is not None else "port"
    rng = np.random.default_rng(1)
    N = o * o * v * v
    fm = flops_model(o, v, m)
    total_fma = sum(fm.values())
    if 8.0 * v ** 4 <= 4e9 and total_fma <= 3e10 and scale >= 1.0:
        pt = _time_point(R, P, o, v, m, m, sampling, CONTRACTION_ORDER, rng, n_comp=2)
        total = 2 * pt["compress"] + sum(pt[t] for t in CONTRACTION_ORDER)
        sample = (f"oracle/{'_ref' if R is not None else 'port'}: the workload itself -- 2x SparseTensor4d(fri,{sampling}) + "
                  f"the 11 contract_DTSpT terms at o={o},v={v},m={int(m)}; {total:.2f} s")
        return dict(kind=kind, cores=threads, sample=sample, measured_s=total, extrapolated_s=total, same_config=True,
                    parts={k: round(x, 4) for k, x in pt.items()}, fit=None)
    so = min(o, 40)
    vs = [vv for vv in (60, 80, 100, 120) if vv <= max(v, 60)]
    if scale < 1.0:
        vs, m_lo, m_hi = [40, 60], 2e4, 4e4
    else:
        m_lo, m_hi = 5e4, 2e5
    points = []
    t_start = time.perf_counter()
    # ladder: every v' at m_lo, one more at m_hi (separates the v^2 term from the o^2 scan term)
    lad = [(vv, m_lo) for vv in vs] + [(vs[min(1, len(vs) - 1)], m_hi)]
    for vv, mm in lad:
        pt = _time_point(R, P, so, vv, mm, mm, sampling, ["2323", "0101", "1323"], rng, n_comp=0)
        points.append(dict(o=so, v=vv, m=int(mm), **{k: round(x, 5) for k, x in pt.items()}))
    # O^3V^3 terms: one coefficient each, two points
    o3 = ["1302", "1202", "1303", "1203", "0112", "0313", "0113", "0312"]
    for vv, mm in ((vs[min(1, len(vs) - 1)], m_lo), (vs[-1], m_hi)):
        pt = _time_point(R, P, so, vv, mm, mm, sampling, o3, rng, n_comp=0)
        points.append(dict(o=so, v=vv, m=int(mm), **{k: round(x, 5) for k, x in pt.items()}))
    # compression at two sizes, the target's density
    for vv in (vs[0], vs[-2] if len(vs) > 2 else vs[-1]):
        n = so * so * vv * vv
        mc = max(1, min(int(round(n * m / N)), n))
        pt = _time_point(R, P, so, vv, mc, 0, sampling, [], rng, n_comp=1)
        points.append(dict(o=so, v=vv, m_comp=mc, N=n, compress=round(pt["compress"], 5)))
    measured = time.perf_counter() - t_start

    def lstsq_nonneg(rows, ys):
        A, y = np.array(rows, dtype=float), np.array(ys, dtype=float)
        coef, *_ = np.linalg.lstsq(A, y, rcond=None)
        if (coef < 0).any():            # drop the negative feature and refit the other
            keep = [i for i in range(A.shape[1]) if coef[i] >= 0] or [0]
            c2, *_ = np.linalg.lstsq(A[:, keep], y, rcond=None)
            coef = np.zeros(A.shape[1])
            coef[keep] = np.maximum(c2, 0.0)
        pred = A @ coef
        resid = float(np.max(np.abs(pred - y) / np.maximum(y, 1e-12)))
        return coef, resid

    fit, extrap = {}, 0.0
    for term in CONTRACTION_ORDER:
        pts = [p_ for p_ in points if term in p_]
        if term in ("2323", "0101"):
            # + a constant per call (thread start-up, first touch of the output): keeps it out of the per-FMA rate
            rows = [[flops_model(p_["o"], p_["v"], p_["m"])[term], scan_model(p_["o"], p_["v"], p_["m"])[term], 1.0] for p_ in pts]
            target = [fm[term], scan_model(o, v, m)[term], 1.0]
        else:
            rows = [[flops_model(p_["o"], p_["v"], p_["m"])[term] + scan_model(p_["o"], p_["v"], p_["m"])[term]] for p_ in pts]
            target = [fm[term] + scan_model(o, v, m)[term]]
        coef, resid = lstsq_nonneg(rows, [p_[term] for p_ in pts])
        t_term = float(np.dot(coef, target))
        fit[term] = dict(ns_per_unit=[round(float(c) * 1e9, 4) for c in coef[:2]], max_rel_residual=round(resid, 3),
                         extrapolated_s=round(t_term, 3))
        extrap += t_term
    cp = [p_ for p_ in points if "compress" in p_]
    coef, resid = lstsq_nonneg([[p_["N"] * math.log2(p_["N"])] for p_ in cp], [p_["compress"] for p_ in cp])
    t_comp = 2.0 * float(coef[0]) * N * math.log2(N)
    fit["compress_x2"] = dict(ns_per_NlogN=round(float(coef[0]) * 1e9, 5), max_rel_residual=round(resid, 3),
                              extrapolated_s=round(t_comp, 3))
    extrap += t_comp
    sample = (f"oracle/{'_ref' if R is not None else 'port'}: {len(points)} sample points at o={so}, v in {vs}, "
              f"m in {{{int(m_lo)}, {int(m_hi)}}} on FRI-compressed tensors ({measured:.1f} s measured); per-term fit "
              f"t = a*FMAs + b*index_tests (compression: g*N*log2 N) extrapolated to o={o},v={v},m={int(m)}")
    return dict(kind=kind, cores=threads, sample=sample, measured_s=measured, extrapolated_s=extrap, same_config=False,
                parts=None, fit=fit, points=points)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    o, v, m = args.o, args.v, int(args.m)
    for _ in range(max(args.warmup, 0)):
        cpu_reference_sample(o, v, m, args.sampling, threads, scale=0.25)   # warm-up: page cache, thread pool
    vals, meas, r = [], [], None
    for _ in range(max(args.steps, 1)):
        r = cpu_reference_sample(o, v, m, args.sampling, threads)
        vals.append(r["extrapolated_s"])
        meas.append(r["measured_s"])
    value = sum(vals) / len(vals)
    cfg = workload_config(args, args.gpus)
    cfg["same_config"] = bool(r["same_config"])
    line = {
        "impl": "reference", "metric": metric_name(o, v, m), "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(meas) / len(meas), "higher_is_better": False,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "fit": r["fit"], "sample_points": r.get("points"), "parts_s": r["parts"],
        "per_step_values": [round(x, 3) for x in vals],
    }
    if not args.no_pairs:
        R, P = _ref_modules(threads)
        line["measured_configs"] = [measure_full_config(R, P, o2, v2, m2, args.sampling) for o2, v2, m2 in SAME_CONFIG_PAIRS]
        for c in line["measured_configs"]:
            c["cores"], c["kind"] = threads, ("reference" if R is not None else "port")
    print(json.dumps(line))


# ======================================================================================================================
# the B200 arm
# ======================================================================================================================
def _ptr(t, elem_offset=0):
    return ctypes.c_void_p(int(t.data_ptr()) + 8 * int(elem_offset))


class SegmentTimer:
    """CUDA-event stopwatch on the current stream that can be paused around untimed input regeneration."""

    def __init__(self, torch):
        self.torch = torch
        self.segments = []
        self.open = None

    def resume(self):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        self.open = e

    def pause(self):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        self.segments.append((self.open, e))
        self.open = None

    def total_ms(self):
        self.torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in self.segments)


class Iteration:
    """Device-resident synthetic FRI-CCSD iteration, one instance per rank (ViewEngine: plugin calls on views)."""

    def __init__(self, args, comm, device, view_comm=None):
        import torch

        from fricc_b200 import _lib
        from fricc_b200 import dist as fd
        from fricc_b200 import py_rccsd

        self.torch, self.fd, self.py, self.lib = torch, fd, py_rccsd, _lib.lib()
        self.check = _lib.check
        self.args, self.comm, self.device = args, comm, device
        o, v = args.o, args.v
        self.o, self.v, self.m = o, v, int(args.m)
        self.N = o * o * v * v
        g = torch.Generator(device=device)
        g.manual_seed(20)
        # heavy-tailed amplitudes, identical on every rank
        u = torch.rand(self.N, generator=g, device=device, dtype=torch.float64)
        z = torch.randn(self.N, generator=g, device=device, dtype=torch.float64)
        self.t2 = (z * torch.pow(10.0, -6.0 * u * u)).reshape(o, o, v, v).contiguous()
        del u, z
        # U(0,1) operands like the reference's tests (fricc/tests/test_sparse_tensors.py:78)
        self.ovov = torch.rand(o, v, o, v, generator=g, device=device, dtype=torch.float64)
        self.tc = args.planes > 0
        self.fuse = self.tc and not args.no_fuse
        py_rccsd.set_option("tc_mode", 1 if self.tc else 0)
        if self.tc:
            py_rccsd.set_option("tc_planes", args.planes)
            if not args.no_constant_eris:
                # the integrals (eris.ovov, read by 0112/0113/0313/0312/1323) never change between iterations in either
                # driver: declared constant, their digit planes are built in the first (warm-up) step and kept
                py_rccsd.constant_operand(self.ovov, True)
        self.multi = comm.world > 1
        self.side = None
        if self.multi:
            nb = 8 * (o ** 4 + 2 * v * o * o * v + 4 * o * o * v * v) + 8 * 4096
            self.arena = fd.SymmArena(nb, comm, device)
            mk = lambda *shape: self.arena.tensor(shape).zero_()
            if not args.push_inline:
                # (same priority as the step's own stream, which --overlap-planes raises)
                self.side = fd.SideStream(device, priority=-1 if args.overlap_planes else 0)
            py_rccsd.set_option("push_copy_engine", args.push_copy_engine)
            py_rccsd.set_option("uniform_carveout", args.uniform_carveout)
        else:
            self.arena = None
            mk = lambda *shape: torch.zeros(*shape, device=device, dtype=torch.float64)
        self.Wklij = mk(o, o, o, o)
        self.Wakic = mk(v, o, o, v)
        self.Wakci = mk(v, o, v, o)
        self.t2new = mk(o, o, v, v)
        self.tmp = [mk(o, o, v, v) for _ in range(3)]
        self.eng = fd.ViewEngine(view_comm or comm, device, None if args.nccl_barrier else self.arena, self.side)
        if args.nccl_barrier:
            self.eng.arena = self.arena          # peer stores still go through the arena; only the barrier is NCCL
        # ---- the ladder operand: this rank's a-range of the NATIVE W[a,b,c,d], possibly in sub-shards
        self.a_lo, self.a_hi = self.eng.my_range(v)
        na = max(0, self.a_hi - self.a_lo)
        per_a = 8.0 * v ** 3
        # resident needs W (8 B) and its planes (S B) per element next to everything else
        na_sub = max(1, int(args.max_slice_gb * 1e9 / per_a))
        self.sub = []
        a = self.a_lo
        while a < self.a_hi:
            w = min(na_sub, self.a_hi - a)
            self.sub.append((a, a + w))
            a += w
        self.resident = len(self.sub) <= 1
        width = max((b - a for a, b in self.sub), default=1)
        self.Wlad = torch.empty(width, v, v, v, device=device, dtype=torch.float64)
        self.planes = None                      # fricc_wslices_t of the sub-shard currently in Wlad
        self.plane_ms, self.ladder_ms = [], []
        self.cur_sub = -1
        self.prebuilt = -1
        self.prepared_plugin = -1
        self.plane_side = fd.SideStream(device) if self.tc and args.overlap_planes else None
        if self.sub:
            self.fill_sub(0)
            if self.tc and args.w_mode == "invariant":
                self.build_planes(0)
        self.draws = torch.randint(0, 2**31 - 1, (64,), generator=torch.Generator().manual_seed(5)).numpy().astype("int32")
        self.draw_pos = 0
        self.stage_events = []

    # -- the ladder operand ---------------------------------------------------------------------------------------
    def wsub(self, s):
        a, b = self.sub[s]
        return self.Wlad[: b - a]

    def fill_sub(self, s):
        """(Re)generate sub-shard s of W in HBM -- an INPUT (the caller's dense work), not part of the timed step.
        Values depend on the GLOBAL element index only, so every sharding holds the same tensor."""
        a, b = self.sub[s]
        n = (b - a) * self.v ** 3
        self.check(self.lib.fricc_fill_uniform(_ptr(self.Wlad), n, a * self.v ** 3, 1000))
        self.cur_sub = s

    def build_planes(self, s, timed=False, stream=None):
        """int8 digit planes of sub-shard s, straight from the native layout (single-read cluster kernel).  `stream`:
        the stream the library is currently issuing on when that is not torch's current one (events go there)."""
        torch, fd = self.torch, self.fd
        W = self.wsub(s)
        cols = W.shape[0] * self.v
        stream = stream or torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        if self.planes is not None and self._planes_cols == cols:
            self.check(self.lib.fricc_wslices_update_view(self.planes, _ptr(W), fd.i64x4(W.shape), fd.i64x4(W.stride()),
                                                          b"2323", 0))
        else:
            if self.planes is not None:
                self.lib.fricc_wslices_destroy(self.planes)
            h = ctypes.c_void_p()
            self.check(self.lib.fricc_wslices_create_view(_ptr(W), fd.i64x4(W.shape), fd.i64x4(W.stride()), b"2323", 0, cols,
                                                          self.args.planes, ctypes.byref(h)))
            self.planes, self._planes_cols = h, cols
        e1.record(stream)
        if timed:
            self.plane_ms.append((e0, e1, cols, stream != torch.cuda.current_stream()))

    def ladder(self, T, timer=None, plugin=False):
        """2323 on this rank's a-range.  `plugin`: through contract_DTSpT (planes built inside the call); otherwise the
        same two kernels as two explicit steps, so that each can be timed on its own."""
        torch, fd = self.torch, self.fd
        refreshed = self.args.w_mode == "refreshed"
        for s, (a, b) in enumerate(self.sub):
            if self.cur_sub != s:
                if timer:
                    timer.pause()
                self.fill_sub(s)
                if self.tc and not refreshed:
                    self.build_planes(s)
                if timer:
                    timer.resume()
            W = self.wsub(s)
            ov = self.t2new[:, :, a:b, :]
            if plugin or not self.tc:
                if self.prepared_plugin == s:
                    self.plane_side.join()       # py_rccsd.prepare_operand ran on the side stream during upload + compression
                    self.prepared_plugin = -1
                if self.tc and not refreshed:
                    self.check(self.lib.fricc_contract_planes(self.planes, _ptr(W), fd.i64x4(W.shape), fd.i64x4(W.stride()), 0,
                                                              T._h, _ptr(ov), fd.i64x4(ov.shape), fd.i64x4(ov.stride()), b"2323"))
                else:
                    self.py.contract_DTSpT(W, T, ov, "2323")
                continue
            if refreshed and self.prebuilt == s:
                self.plane_side.join()           # encoded on the side stream while T2 was being compressed
                self.prebuilt = -1
            elif refreshed:
                self.build_planes(s, timed=True)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            self.check(self.lib.fricc_contract_planes(self.planes, _ptr(W), fd.i64x4(W.shape), fd.i64x4(W.stride()), 0, T._h,
                                                      _ptr(ov), fd.i64x4(ov.shape), fd.i64x4(ov.stride()), b"2323"))
            e1.record()
            self.ladder_ms.append((e0, e1, (b - a) * self.v))
        if self.sub:
            return self.t2new[:, :, self.a_lo:self.a_hi, :]
        return None

    # -- the step ---------------------------------------------------------------------------------------------------
    def mark(self, name):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        self.stage_events.append((name, e, time.perf_counter()))

    def compress_mode(self):
        if self.args.sampling == "systematic":
            return "ranges"
        return "pair" if self.multi else "replicated"

    def compress_two(self, t2_slice, j0, full=None):
        """The step's two compressions.  t2_slice = this rank's contiguous flat range [j0, j0+len) of T2 (systematic:
        all a rank needs); full = the whole T2 (pivotal)."""
        fd = self.fd
        if self.compress_mode() == "ranges":
            d = (self.draws[self.draw_pos % 32:], self.draws[(self.draw_pos + 1) % 32:])
            self.draw_pos += 2
            return fd.compress_ranges_pair(t2_slice, j0, (self.o, self.o, self.v, self.v), self.m, self.comm, "fri",
                                           "systematic", draws=d)
        if self.multi:
            return fd.compress_pair(full, self.m, self.comm, "fri", self.args.sampling, draws=None)
        return (fd.compress(full, self.m, "fri", self.args.sampling), fd.compress(full, self.m, "fri", self.args.sampling))

    def contractions(self, T, T_alt, timer=None, plugin=False, after_t2new=None):
        eng, c = self.eng, self.eng.contract
        # every output accumulates from zero; a rank zeroes only the part it computes -- the rest of its copy is overwritten
        # by its peers' stores, and nobody stores into a slice its owner still works on, so no barrier is needed here
        # (the previous step's B2 has made sure that nobody still reads last step's outputs)
        for view in (eng.own_view("2323", None, self.t2new, w_range=(self.a_lo, self.a_hi)),
                     eng.own_view("0113", self.ovov.shape, self.Wakic), eng.own_view("0312", self.ovov.shape, self.Wakci),
                     eng.own_view("1323", self.ovov.shape, self.Wklij), eng.own_view("1302", self.Wakic.shape, self.tmp[0]),
                     eng.own_view("1303", self.Wakic.shape, self.tmp[1]), eng.own_view("1203", self.Wakci.shape, self.tmp[2])):
            if view is not None:
                view.zero_()
        self.mark("zero_own_slices")
        # the ladder first: its 2 GB of peer stores then hide under the ovov terms
        eng.publish(self.ladder(T, timer, plugin))
        self.mark("2323" if self.resident else "2323_incl_regen")
        if self.fuse:      # 0112 + 0113 read the same operand view: one weighted sign matrix, one dense product
            eng.contract_pair("0112", "0113", self.ovov, self.ovov, T_alt, self.Wakic)
        else:
            c("0112", self.ovov, T_alt, self.Wakic)
            c("0113", self.ovov, T_alt, self.Wakic)
        eng.publish(c("0313", self.ovov, T_alt, self.Wakic))
        self.mark("0112_0313_0113")
        eng.publish(c("0312", self.ovov, T_alt, self.Wakci))
        self.mark("0312")
        eng.publish(c("1323", self.ovov, T_alt, self.Wklij))      # the smallest output last: its stores are all B1 waits for
        self.mark("1323")
        eng.barrier()                                            # B1: Woooo / Wvoov / Wvovo and the ladder's t2new are complete everywhere
        self.mark("B1")
        eng.publish(c("0101", self.Wklij, T, self.t2new))        # t2new is final on this rank from here on
        self.mark("0101")
        if after_t2new is not None:
            after_t2new()
        if self.fuse:      # 1302 + 1202 share the sign matrix: planes of 2 Wvoov - Wvovo, one dense product
            eng.publish(eng.contract_pair("1302", "1202", self.Wakic, self.Wakci, T, self.tmp[0]))
        else:
            c("1302", self.Wakic, T, self.tmp[0])
            eng.publish(c("1202", self.Wakci, T, self.tmp[0]))
        self.mark("1302_1202")
        eng.publish(c("1303", self.Wakic, T, self.tmp[1]))
        self.mark("1303")
        if self.multi:
            eng.contract_in_parts("1203", self.Wakci, T, self.tmp[2], parts=self.args.last_term_parts)   # last before B2
        else:
            c("1203", self.Wakci, T, self.tmp[2])
        self.mark("1203")
        eng.barrier()                                            # B2: the step's outputs are complete on every rank
        self.mark("B2")

    def prepare_ladder_operand(self, timer=None):
        """e2e (plugin calls only): encode the ladder's first sub-shard on the second stream NOW -- W does not depend on
        T2, which is still on its way from the host -- through py_rccsd.prepare_operand; the coming
        contract_DTSpT(W, T, t2new[..], "2323") then finds its planes ready."""
        if self.plane_side is None or not self.sub or self.args.w_mode != "refreshed":
            return
        if self.cur_sub != 0:
            if timer:
                timer.pause()
            self.fill_sub(0)
            if timer:
                timer.resume()
        a, b = self.sub[0]
        W, ov = self.wsub(0), self.t2new[:, :, a:b, :]
        self.plane_side.run(lambda: self.py.prepare_operand(W, ov, "2323"))
        self.prepared_plugin = 0

    def step(self, t2, timer=None):
        self.mark("begin")
        if self.plane_side is not None and self.sub and self.args.w_mode == "refreshed":
            # the ladder operand does not depend on T2: its first sub-shard is encoded on a second stream while the
            # compression (short kernels, host round trips and, N>1, small collectives) leaves most of the GPU idle
            if self.cur_sub != 0:
                if timer:
                    timer.pause()
                self.fill_sub(0)
                if timer:
                    timer.resume()
            self.plane_side.run(lambda: self.build_planes(0, timed=True, stream=self.plane_side.stream))
            self.prebuilt = 0
        flat = t2.view(-1)
        j0, cnt = self.fd.index_range(flat.numel(), self.comm.rank, self.comm.world)
        T, T_alt = self.compress_two(flat[j0:j0 + cnt], j0, full=t2)
        self.mark("compress_x2")
        self.contractions(T, T_alt, timer, plugin=False)
        self.last_pair = (T, T_alt)
        return T

    def stage_breakdown(self, steps):
        self.torch.cuda.synchronize()
        out, worst, host = {}, {}, {}
        for (n0, e0, h0), (n1, e1, h1) in zip(self.stage_events, self.stage_events[1:]):
            if n1 == "begin":
                continue
            dt = e0.elapsed_time(e1)
            out[n1] = out.get(n1, 0.0) + dt
            if dt >= worst.get(n1, 0.0):
                # host_ms: what the HOST spent issuing the stage (it runs ahead of the device): a device time far above
                # both the stage's average and its host time is a stall on the device side, not late launches
                worst[n1], host[n1] = dt, (h1 - h0) * 1e3
        self.stage_events.clear()
        self.stage_worst = {k: round(x, 3) for k, x in worst.items()}    # the slowest single step, per stage
        self.stage_worst_host = {k: round(x, 3) for k, x in host.items()}
        return {k: round(x / steps, 3) for k, x in out.items()}


# ---- host spot check of a step's outputs (outside the timed region) ---------------------------------------------------
# term -> (prefactor, W indices, T indices, output indices): the einsum strings of the reference's own tests
# (fricc/tests/test_sparse_tensors.py:48-269), i.e. what src/sparse_tensor.cpp:12-346 computes term by term
SPOT_TERMS = {"2323": (1.0, "abcd", "ijcd", "ijab"), "0101": (1.0, "klij", "klab", "ijab"),
              "1302": (2.0, "akic", "kjcb", "ijab"), "1202": (-1.0, "akci", "kjcb", "ijab"),
              "1303": (1.0, "akic", "kjbc", "ijab"), "1203": (1.0, "bkci", "kjac", "ijab"),
              "1323": (1.0, "kcld", "ijcd", "klij"), "0112": (-0.5, "ldkc", "ilda", "akic"),
              "0313": (-0.5, "lckd", "ilad", "akic"), "0113": (1.0, "ldkc", "ilad", "akic"),
              "0312": (-0.5, "lckd", "ilda", "akci")}


def fill_uniform_host(idx, seed):
    """fricc_fill_uniform on the host: the value at GLOBAL element index idx (uint64 array) -- a function of (seed, idx) only."""
    import numpy as np

    with np.errstate(over="ignore"):
        z = idx.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * 1.1102230246251565e-16


def spot_check_outputs(it, torch, n_t2new=256, n_other=64, seed=77):
    """Recompute randomly chosen elements of every output of the last step ON THE HOST in fp64, straight from the term
    definitions: the exported entries of T / T_alt (fricc_st_export), the ladder operand from the index-defined generator
    behind fricc_fill_uniform, the other operands read element-wise from the arrays the step was given (ovov) or produced
    one stage earlier (Woooo, Wvoov, Wvovo -- each of which is itself spot-checked against ovov and T_alt, so the chain
    is checked stage by stage).  Independent of the library's kernels, layouts and plane expansion.
    -> {"points", "max_rel" (|got - want| / |want|), "max_rel_to_abs_sum" (/ sum |terms|: the error bound's own scale)}"""
    import numpy as np

    o, v = it.o, it.v
    T, T_alt = it.last_pair
    ent = {}
    for name, S in (("T", T), ("T_alt", T_alt)):
        flat, vals = S.export()
        flat = np.asarray(flat, dtype=np.int64)
        r, b = np.divmod(flat, v)
        r, a = np.divmod(r, v)
        i, j = np.divmod(r, o)
        ent[name] = ([i.astype(np.int32), j.astype(np.int32), a.astype(np.int32), b.astype(np.int32)], np.asarray(vals))

    def ladder_w(ix):
        a, b, c, d = (np.asarray(q, dtype=np.uint64) for q in ix)
        vv = np.uint64(v)
        return fill_uniform_host(((a * vv + b) * vv + c) * vv + d, 1000)

    def from_device(t):
        def get(ix):
            if len(ix[0]) == 0:
                return np.zeros(0)
            dev = tuple(torch.as_tensor(np.asarray(q, dtype=np.int64), device=t.device) for q in ix)
            return t[dev].cpu().numpy()
        return get

    ovov, Wklij, Wakic, Wakci = (from_device(t) for t in (it.ovov, it.Wklij, it.Wakic, it.Wakci))
    outputs = [("t2new", it.t2new, n_t2new, [("2323", ladder_w, "T"), ("0101", Wklij, "T")]),
               ("Woooo", it.Wklij, n_other, [("1323", ovov, "T_alt")]),
               ("Wvoov", it.Wakic, n_other, [("0112", ovov, "T_alt"), ("0113", ovov, "T_alt"), ("0313", ovov, "T_alt")]),
               ("Wvovo", it.Wakci, n_other, [("0312", ovov, "T_alt")]),
               ("tmp0", it.tmp[0], n_other, [("1302", Wakic, "T"), ("1202", Wakci, "T")]),
               ("tmp1", it.tmp[1], n_other, [("1303", Wakic, "T")]),
               ("tmp2", it.tmp[2], n_other, [("1203", Wakci, "T")])]
    rng = np.random.default_rng(seed)
    worst_rel, worst_abs, n_points, per_out = 0.0, 0.0, 0, {}
    for name, out, count, terms in outputs:
        pts = np.stack([rng.integers(0, n, size=count) for n in out.shape], axis=1)
        got = out[tuple(torch.as_tensor(pts[:, q], device=out.device) for q in range(4))].cpu().numpy()
        w_rel = w_abs = 0.0
        for pt, g in zip(pts, got):
            want = scale_sum = 0.0
            for term, get_w, which in terms:
                pref, w_idx, t_idx, o_idx = SPOT_TERMS[term]
                fixed = dict(zip(o_idx, (int(q) for q in pt)))
                tix, tv = ent[which]
                mask = None
                for q, c in enumerate(t_idx):
                    if c in fixed:
                        mq = tix[q] == fixed[c]
                        mask = mq if mask is None else (mask & mq)
                sel = np.nonzero(mask)[0]
                val = dict((c, np.full(sel.size, x, dtype=np.int64)) for c, x in fixed.items())
                for q, c in enumerate(t_idx):
                    if c not in fixed:
                        val[c] = tix[q][sel].astype(np.int64)
                prod = get_w([val[c] for c in w_idx]) * tv[sel]
                want += pref * float(np.sum(prod))
                scale_sum += abs(pref) * float(np.sum(np.abs(prod)))
            err = abs(float(g) - want)
            if want != 0.0:
                w_rel = max(w_rel, err / abs(want))
            if scale_sum != 0.0:
                w_abs = max(w_abs, err / scale_sum)
            elif err != 0.0:
                w_rel = w_abs = float("inf")
            n_points += 1
        per_out[name] = {"points": int(count), "max_rel": w_rel, "max_rel_to_abs_sum": w_abs}
        worst_rel, worst_abs = max(worst_rel, w_rel), max(worst_abs, w_abs)
    return {"points": n_points, "max_rel": worst_rel, "max_rel_to_abs_sum": worst_abs, "by_output": per_out,
            "how": "host fp64 recomputation of random output elements from the exported T / T_alt and the term definitions; "
                   "ladder operand from the index-defined generator of fricc_fill_uniform (see spot_check_outputs)"}


def compress_parity_at_scale(it):
    """The two compressions of the verification step against the oracle AT THE WORKLOAD'S OWN SIZE (part of the
    cpu_baseline leg: rank 0, N = 1, outside every timed region).  T2 goes to the host, the C port of the reference runs
    get_d_largest / make_probability_vector once and sample_systematic once per draw (fri_utils.cpp:400-437, :367-398,
    :129-159; the composition is fri_compression, :439-522), and the exported index lists and values of T and T_alt must
    equal it bit for bit."""
    import numpy as np

    from oracle import oracle

    t0 = time.time()
    P = oracle.port()
    x = it.t2.reshape(-1).cpu().numpy()
    D, r = P.get_d_largest(x, it.m)
    D = np.asarray(D, dtype=np.int64)
    nsamp = it.m - D.size
    p = P.make_probability_vector(x, nsamp, D, r)
    idx_equal = vals_equal = True
    for S_gpu, draw in zip(it.last_pair, (it.draws[0:1], it.draws[1:2])):
        S, _ = P.sample_systematic(nsamp, p, draw)
        S = np.asarray(S, dtype=np.int64)
        flat, vals = S_gpu.export()
        idx_equal &= bool(np.array_equal(np.asarray(flat, dtype=np.int64), np.concatenate([D, S])))
        want = np.concatenate([x[D], x[S] / p[S]])
        vals_equal &= bool(np.array_equal(np.asarray(vals).view(np.int64), want.view(np.int64)))
    return {"N": int(x.size), "m": int(it.m), "n_exact": int(D.size), "compressions": 2, "sampler": "systematic",
            "idx_equal": idx_equal, "vals_equal": vals_equal, "oracle": "oracle/fricc_oracle.c (C port, 1 thread)",
            "oracle_seconds": time.time() - t0}


def compression_sweep(comm, device, torch, fd, sizes=(1.0e8, 2.56e8, 1.0e9), m=10_000_000, reps=3):
    """BASELINE config 5: compression only (top-m select + D + sampler) of a synthetic heavy-tailed T2 of N entries, at
    THIS run's GPU count.  systematic: range-sharded (every rank holds and scans N/G entries); pivotal: the walk is not
    range-sharded -- one rank's time for one compression of the full tensor.  ms = median of `reps`, max over ranks."""
    import numpy as np
    import torch.distributed as dist

    from fricc_b200 import ranges
    from fricc_b200.py_rccsd import SparseTensor4d

    out = []
    world, rank = comm.world, comm.rank
    draw = np.array([123456789], dtype=np.int32)
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
    except Exception:
        pass

    def timed(fn):
        ts = []
        for it in range(1 + reps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if it > 0:
                ts.append(float(t.item()))
        ts.sort()
        return ts[len(ts) // 2]

    for size in sizes:
        v = int(round((size / 1600) ** 0.5))
        n = 1600 * v * v
        dims = (40, 40, v, v)
        mm = min(m, n)
        j0, cnt = ranges.index_range(n, rank, world)
        g = torch.Generator(device=device)
        g.manual_seed(20 + rank)
        u = torch.rand(cnt, generator=g, device=device, dtype=torch.float64)
        z = torch.randn(cnt, generator=g, device=device, dtype=torch.float64)
        x = (z * torch.pow(10.0, -6.0 * u * u)).contiguous()
        del u, z
        ms = timed(lambda: ranges.compress_ranges(x, j0, dims, mm, comm, "fri", "systematic", draws=draw))
        bcomp = 16.0 * n + 16.0 * mm
        out.append({"N": n, "m": mm, "sampler": "systematic", "how": f"range-sharded over {world} GPU(s)",
                    "ms_per_compression": round(ms, 3), "elements_per_s": n / (ms * 1e-3),
                    "frac_of_measured_hbm": bcomp / (ms * 1e-3) / 1e9 / (peak * world)})
        # the chunked pivotal sampler (Alg. A.2, fri_utils.cpp:176-362) with one chunk per rank: the pivotal family's
        # range-sharded form (no state crosses a chunk boundary); library rand() stream, every rank's copy in step
        try:
            ms = timed(lambda: ranges.compress_ranges(x, j0, dims, mm, comm, "fri", "parallel"))
            out.append({"N": n, "m": mm, "sampler": "parallel (chunked pivotal, Alg. A.2)",
                        "how": f"range-sharded over {world} GPU(s), one chunk per rank",
                        "ms_per_compression": round(ms, 3), "elements_per_s": n / (ms * 1e-3),
                        "frac_of_measured_hbm": bcomp / (ms * 1e-3) / 1e9 / (peak * world)})
        except Exception as e:       # e.g. the reference's own exit when a one-chunk budget is a hair off an integer
            out.append({"N": n, "m": mm, "sampler": "parallel (chunked pivotal, Alg. A.2)", "error": str(e)[:200]})
        del x
        if rank == 0 and n <= 3.0e8:
            g.manual_seed(20)
            u = torch.rand(n, generator=g, device=device, dtype=torch.float64)
            z = torch.randn(n, generator=g, device=device, dtype=torch.float64)
            xf = (z * torch.pow(10.0, -6.0 * u * u)).reshape(dims).contiguous()
            del u, z
        else:
            xf = None
        # pivotal: one rank, the others wait in the barrier (its walk is sequential in index order and not sharded)
        def piv():
            if xf is not None:
                SparseTensor4d(xf, dims, mm, "fri", "pivotal", False)
        if n <= 3.0e8:
            ms = timed(piv)
            out.append({"N": n, "m": mm, "sampler": "pivotal", "how": "one GPU (the pivotal walk is not range-sharded)",
                        "ms_per_compression": round(ms, 3), "elements_per_s": n / (ms * 1e-3),
                        "frac_of_measured_hbm": bcomp / (ms * 1e-3) / 1e9 / peak})
        del xf
        torch.cuda.empty_cache()
    return out


def start_stall_trace(it, limit_ms):
    """Diagnostic watchdog thread: when the main thread has not passed a stage mark for `limit_ms` and is not inside the
    compression (whose read-backs legitimately wait for the device), print where it is."""
    import threading
    import traceback

    main_id = threading.main_thread().ident

    def watch():
        last, since, reported = None, time.perf_counter(), False
        while True:
            time.sleep(0.02)
            key = len(it.stage_events)
            now = time.perf_counter()
            if key != last:
                last, since, reported = key, now, False
                continue
            name = it.stage_events[-1][0] if it.stage_events else ""
            if not reported and (now - since) * 1e3 > limit_ms and name not in ("begin", "B2", ""):
                frame = sys._current_frames().get(main_id)
                sys.stderr.write(f"[stall-trace] {1e3 * (now - since):.0f} ms behind mark '{name}':\n" +
                                 "".join(traceback.format_stack(frame)[-6:]))
                sys.stderr.flush()
                reported = True

    threading.Thread(target=watch, daemon=True).start()


def start_clock_sampler(gpu_index, period_ms=200):
    path = tempfile.NamedTemporaryFile(prefix="clocks_", suffix=".csv", delete=False).name
    # started BEFORE the warm-up steps: nvidia-smi's start-up (NVML attaches every GPU of the box) then falls into the
    # warm-up, not into the timed region; the samples are cut to the timed window by their timestamps afterwards
    q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,timestamp")
    try:
        proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                 "-lms", str(int(period_ms))], stdout=open(path, "w"), stderr=subprocess.DEVNULL)
    except Exception:
        return None, path
    return proc, path


def _smi_time(text):
    import datetime

    try:
        return datetime.datetime.strptime(text.strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
    except Exception:
        return None


def stop_clock_sampler(proc, path, window=None):
    """window = (t0, t1) in time.time() seconds: only samples taken inside it count (all of them if the timestamps
    cannot be read or none falls inside)."""
    if proc is None:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    lines = [ln for ln in open(path)]
    if window is not None:
        inside = []
        for ln in lines:
            ts = _smi_time(ln.split(",")[-1]) if ln.count(",") >= 7 else None
            if ts is not None and window[0] - 0.2 <= ts <= window[1] + 0.2:
                inside.append(ln)
        lines = inside or lines
    for line in lines:
        f = [x.strip() for x in line.split(",")]
        if len(f) < 7:
            continue
        try:
            sm.append(float(f[0])); mx.append(float(f[1]))
        except ValueError:
            continue
        for nme, val in zip(names, f[3:7]):
            if val.lower().startswith("active"):
                reasons.add(nme)
    os.unlink(path)
    sm.sort()
    load = sm[len(sm) // 2:] if sm else []      # upper half of the samples = under load (the sampler also sees idle gaps)
    return {"sm_mhz": (load[len(load) // 2] if load else None), "sm_max_mhz": (max(mx) if mx else None),
            "reasons": sorted(reasons), "samples": len(sm)}


NCU_CAPTURE_CONFIG = {"o": 40, "v": 400, "m": 10_000_000, "planes": 6}
_run_matches_ncu_capture = False   # set by run_b200_arm: the committed captures were taken at the default workload on one GPU


def ncu_traffic_per_launch(which):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch from the newest committed `ncu --set full` summary
    profiles/ncu_full_r*_<which>.csv; None if absent -- and None whenever this run's (o, v, m, planes, GPUs) is not the
    configuration those captures were taken at (a per-launch byte count of another shape says nothing about this one)."""
    import csv
    import glob

    if not _run_matches_ncu_capture:
        return None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", f"ncu_full_r*_{which}.csv")))
    if not files:
        return None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tot = 0.0
    try:
        for row in csv.DictReader(open(files[-1])):
            if row["metric"] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(row["value"]) * scale[row["unit"]]
    except Exception:
        return None
    return tot or None


def run_b200_arm(args):
    global _run_matches_ncu_capture
    _run_matches_ncu_capture = ((args.o, args.v, int(args.m), args.planes) ==
                                (NCU_CAPTURE_CONFIG["o"], NCU_CAPTURE_CONFIG["v"], NCU_CAPTURE_CONFIG["m"], NCU_CAPTURE_CONFIG["planes"])
                                and int(os.environ.get("WORLD_SIZE", "1")) == 1 and not getattr(args, "as_rank_of", 0))
    if args.overlap_planes and not getattr(args, "_on_priority_stream", False):
        # the step's own stream gets the higher priority, so that the short kernels of the compression are scheduled
        # ahead of the pending blocks of the plane encode that runs beside them on the (default-priority) second stream
        import torch

        args._on_priority_stream = True
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if int(os.environ.get("WORLD_SIZE", "1")) > 1 or args.gpus <= 1:
            torch.cuda.set_device(local_rank)
            with torch.cuda.stream(torch.cuda.Stream(device=local_rank, priority=-1)):
                return run_b200_arm(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    import numpy as np
    import torch
    import torch.distributed as dist

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd
    from fricc_b200 import py_rccsd

    _lib.require_cuda()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    _lib.check(_lib.lib().fricc_set_device(local_rank))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    comm = fd.Comm(rank, world)
    fd.use_torch_stream()
    L = _lib.lib()
    view_comm = fd.Comm(0, args.as_rank_of) if (args.as_rank_of > 1 and world == 1) else None
    it = Iteration(args, comm, device, view_comm)
    o, v, m = it.o, it.v, it.m

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(step_fn, steps):
        sync_all()
        timer = SegmentTimer(torch)
        timer.resume()
        for _ in range(steps):
            step_fn(timer)
        timer.pause()
        sync_all()
        return timer.total_ms()

    # ---- warm-up (the clock sampler starts here; only its samples from the timed window are used)
    proc, cpath = start_clock_sampler(local_rank, args.clock_period_ms) if (rank == 0 and not args.no_clocks) else (None, "")
    if args.stall_trace > 0:
        start_stall_trace(it, args.stall_trace)
    for _ in range(max(args.warmup, 0)):
        it.step(it.t2)
    sync_all()
    it.ladder_ms.clear(); it.plane_ms.clear(); it.stage_events.clear()

    # ---- timed: device-resident inputs
    launches0 = L.fricc_kernel_launches()
    tc0 = py_rccsd.get_stat("tc_contract_calls")
    alloc0, hits0 = py_rccsd.get_stat("alloc_driver_calls"), py_rccsd.get_stat("alloc_cache_hits")
    wall0 = time.time()
    ms = timed_steps(lambda timer: it.step(it.t2, timer), args.steps)
    wall1 = time.time()
    launches = L.fricc_kernel_launches() - launches0
    tc_calls = py_rccsd.get_stat("tc_contract_calls") - tc0
    alloc_calls, alloc_hits = py_rccsd.get_stat("alloc_driver_calls") - alloc0, py_rccsd.get_stat("alloc_cache_hits") - hits0
    clocks = stop_clock_sampler(proc, cpath, (wall0, wall1)) if rank == 0 else None
    stage_ms = it.stage_breakdown(args.steps)
    stage_worst, stage_worst_host = it.stage_worst, it.stage_worst_host
    torch.cuda.synchronize()
    waits_by_rank = None
    if world > 1 and not args.as_rank_of:
        # what each rank waits at the two barriers: tells rank imbalance (one rank waits little) from stores still in flight
        t = torch.tensor([stage_ms.get("B1", 0.0), stage_ms.get("B2", 0.0)], device=device, dtype=torch.float64)
        allw = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allw, t)
        waits_by_rank = {"B1": [round(float(x[0]), 3) for x in allw], "B2": [round(float(x[1]), 3) for x in allw]}
    lad_ms = [a.elapsed_time(b) for a, b, _ in it.ladder_ms]
    lad_cols = [c for _, _, c in it.ladder_ms]
    pl_over_ms = [a.elapsed_time(b) for a, b, _, ov in it.plane_ms if ov]
    solo = [(a, b, c) for a, b, c, ov in it.plane_ms if not ov]
    if it.tc and it.sub and args.w_mode == "refreshed" and not solo:
        # every timed encode ran beside the compression: time the kernel ALONE for its roofline line (untimed extra)
        it.plane_ms.clear()
        for _ in range(3):
            it.build_planes(it.cur_sub, timed=True)
        torch.cuda.synchronize()
        solo = [(a, b, c) for a, b, c, _ in it.plane_ms]
    pl_ms = [a.elapsed_time(b) for a, b, _ in solo]
    pl_cols = [c for _, _, c in solo]
    it.ladder_ms.clear(); it.plane_ms.clear()

    # ---- one more (untimed) step from a fixed position of the draw sequence: its outputs do not depend on --steps /
    #      --warmup, on the GPU count or on the sharding, and their sums are compared with stored values below
    it.draw_pos = 0
    it.step(it.t2)
    it.ladder_ms.clear(); it.plane_ms.clear(); it.stage_events.clear()
    checks = {k: float(x.double().sum().item()) for k, x in
              (("t2new", it.t2new), ("tmp0", it.tmp[0]), ("tmp1", it.tmp[1]), ("tmp2", it.tmp[2]),
               ("Wvoov", it.Wakic), ("Wvovo", it.Wakci), ("Woooo", it.Wklij))}

    spot = parity = None
    if rank == 0 and view_comm is None and not args.no_spot_check:
        spot = spot_check_outputs(it, torch)
    if (rank == 0 and world == 1 and view_comm is None and not args.no_cpu_baseline and args.sampling == "systematic"
            and it.N <= 300_000_000):
        parity = compress_parity_at_scale(it)
    # ---- the other operand regime, for the record (same kernels, planes in or out of the step)
    other_ms = None
    if it.tc and it.sub:
        other = "invariant" if args.w_mode == "refreshed" else "refreshed"
        args.w_mode = other
        if other == "invariant":
            it.build_planes(it.cur_sub)
        it.step(it.t2)
        other_ms = timed_steps(lambda timer: it.step(it.t2, timer), max(1, min(args.steps, 2))) / max(1, min(args.steps, 2))
        args.w_mode = "refreshed" if other == "invariant" else "invariant"
        it.ladder_ms.clear(); it.plane_ms.clear(); it.stage_events.clear()

    # ---- e2e: T2 from pinned host memory through the plugin calls, residual back to pinned host memory
    e2e_ms = 0.0
    if not args.no_e2e:
        nel = it.t2.numel()
        if world == 1:
            # the calls fricc/friccsd.py makes -- SparseTensor4d x2, contract_DTSpT x11 -- on a T2 that arrives from pinned
            # host memory ONCE per step (the caller's upload; both constructors then read the device copy) and with the
            # residual t2new going back to pinned host memory as soon as its last term (0101) is done, under the
            # remaining contractions
            h_t2 = torch.empty(it.t2.shape, dtype=torch.float64, pin_memory=True)
            h_t2.copy_(it.t2)
            d_t2 = torch.empty_like(it.t2)
            h_out = torch.empty(it.t2.shape, dtype=torch.float64, pin_memory=True)
            draws = it.draws
            copy_stream = torch.cuda.Stream(device=device)

            def send_back():
                copy_stream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(copy_stream):
                    h_out.copy_(it.t2new, non_blocking=True)

            def e2e_step(timer):
                it.mark("begin")
                it.prepare_ladder_operand(timer)
                d_t2.copy_(h_t2, non_blocking=True)
                d0, d1 = draws[it.draw_pos % 32:], draws[(it.draw_pos + 1) % 32:]
                it.draw_pos += 2
                sys_ = args.sampling == "systematic"
                T = py_rccsd.SparseTensor4d(d_t2, d_t2.shape, m, "fri", args.sampling, False, draws=d0 if sys_ else None)
                T_alt = py_rccsd.SparseTensor4d(d_t2, d_t2.shape, m, "fri", args.sampling, False, draws=d1 if sys_ else None)
                it.mark("compress_x2")
                it.contractions(T, T_alt, timer, plugin=True, after_t2new=send_back)
                torch.cuda.current_stream().wait_stream(copy_stream)
            h2d = nel * 8
            d2h = nel * 8
        else:
            j0, cnt = fd.index_range(nel, rank, world)
            h_t2 = torch.empty(cnt, dtype=torch.float64, pin_memory=True)
            h_t2.copy_(it.t2.view(-1)[j0:j0 + cnt])
            d_slice = torch.empty(cnt, dtype=torch.float64, device=device)
            pivotal_full = torch.empty(nel, dtype=torch.float64, device=device) if args.sampling != "systematic" else None
            # the residual goes back in the pieces the ranks OWN: a rank's slice of t2new for 0101 (a contiguous i-range)
            # holds the whole ladder contribution after B1 and is final on this rank once its own 0101 is done, so its
            # read-back runs on a copy stream under the remaining contractions (the slices of all ranks tile t2new)
            own = it.eng.own_view("0101", it.Wklij.shape, it.t2new)
            early = own is None or own.is_contiguous()
            if early:
                h_out = torch.empty(own.numel() if own is not None else 0, dtype=torch.float64, pin_memory=True)
            else:
                h_out = torch.empty(cnt, dtype=torch.float64, pin_memory=True)
            copy_stream = torch.cuda.Stream(device=device)

            def send_back():
                if own is None:
                    return
                copy_stream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(copy_stream):
                    h_out.copy_(own.view(-1), non_blocking=True)

            def e2e_step(timer):
                it.mark("begin")
                it.prepare_ladder_operand(timer)
                d_slice.copy_(h_t2, non_blocking=True)
                full = None
                if pivotal_full is not None:          # the pivotal walk needs the whole tensor: assemble it over NVLink
                    sizes = [fd.index_range(nel, r, world)[1] for r in range(world)]
                    dist.all_gather(list(pivotal_full.split(sizes)), d_slice)
                    full = pivotal_full.view(it.t2.shape)
                T, T_alt = it.compress_two(d_slice, j0, full=full)
                it.mark("compress_x2")
                it.contractions(T, T_alt, timer, plugin=True, after_t2new=send_back if early else None)
                if early:
                    torch.cuda.current_stream().wait_stream(copy_stream)
                else:
                    h_out.copy_(it.t2new.view(-1)[j0:j0 + cnt], non_blocking=True)
            h2d = nel * 8
            d2h = nel * 8
        e2e_step(None)
        torch.cuda.synchronize()
        it.stage_events.clear()
        e2e_ms = timed_steps(e2e_step, args.steps)
        e2e_stage_ms = it.stage_breakdown(args.steps)
        e2e_stage_worst = it.stage_worst
        it.ladder_ms.clear(); it.plane_ms.clear()

    # ---- compression-only sweep (config 5) at this GPU count, outside the timed regions
    sweep = None
    if not args.no_sweep:
        it.Wlad = None
        if it.planes is not None:
            L.fricc_wslices_destroy(it.planes)
            it.planes = None
        torch.cuda.empty_cache()
        sweep = compression_sweep(comm, device, torch, fd)

    # ---- measured int8 peak on this device (rank 0, nothing else running)
    peak_burst = peak_sust = None
    if rank == 0 and it.tc:
        p = ctypes.c_double(0.0)
        _lib.check(L.fricc_int8_peak(0.02, ctypes.byref(p)))
        peak_burst = p.value
        _lib.check(L.fricc_int8_peak(1.5, ctypes.byref(p)))
        peak_sust = p.value

    # ---- reduce over ranks (max of times, sum of launches)
    stats = torch.tensor([ms, e2e_ms, float(launches), other_ms or 0.0], dtype=torch.float64, device=device)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm_ = stats.clone()
        dist.all_reduce(sm_, op=dist.ReduceOp.SUM)
        ms, e2e_ms, launches, other_ms = float(mx[0]), float(mx[1]), int(sm_[2]), (float(mx[3]) if other_ms else None)
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        hbm_kind = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        nK = v * v
        lad_rows = o * o
        if it.tc and lad_ms:
            # dominant kernel: tc_sign_gemm_kernel (+ the fp64 remainder launch) on the ladder.  Work per launch: the dense
            # sign-matrix x digit-plane product, 2 * rows * nK * cols * planes int8 operations (unpadded).  Peak: measured on
            # this device in this run by fricc_int8_peak (back-to-back tcgen05.mma.kind::i8 on resident operands) -- the
            # sustained figure, the kernel sits inside a long step.
            alg = [2.0 * lad_rows * nK * c * args.planes for c in lad_cols]
            ach = sum(alg) / (sum(lad_ms) * 1e-3) / 1e12
            useful = sum(2.0 * m * c for c in lad_cols) / (sum(lad_ms) * 1e-3) / 1e12
            roof = {"bound": "tensor", "achieved": ach, "peak": peak_sust, "unit": "TFLOP/s", "frac": ach / peak_sust,
                    "traffic": ncu_traffic_per_launch("tc_ladder"),
                    "kernel": f"tc_sign_gemm_kernel (ladder 2323, tcgen05.mma.kind::i8, {args.planes} digit planes)",
                    "peak_source": "measured in this run: fricc_int8_peak, 1.5 s of back-to-back tcgen05.mma.kind::i8 "
                                   "(M=128,N=256,K=32) on operands resident in shared memory, one CTA per SM (under the power cap)",
                    "peak_burst": peak_burst, "peak_nominal_note": "2 x measured bf16_tflops_sustained = "
                    f"{2.0 * float(peaks.get('bf16_tflops_sustained', 1400.0)):.0f}",
                    "unit_note": "int8 TOP/s (2 x MAC) of the dense product = tensor-pipe utilisation",
                    "useful_tflops_fp64_equivalent": useful,
                    "useful_note": "2*m*cols fp64 FMA-equivalents per launch: the work the reference's loop nest does",
                    "fp64_equivalent_algorithmic_gbs": sum(8.0 * m * c for c in lad_cols) / (sum(lad_ms) * 1e-3) / 1e9,
                    "launches_timed": len(lad_ms), "avg_launch_ms": sum(lad_ms) / len(lad_ms),
                    "algorithmic_per_launch": sum(alg) / len(alg)}
        elif lad_ms:
            alg = [8.0 * m * c + 8.0 * o * o * c + 12.0 * m for c in lad_cols]
            ach = sum(alg) / (sum(lad_ms) * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
                    "traffic": ncu_traffic_per_launch("spmm_ladder"), "kernel": "spmm_rows_wide_kernel<2> (ladder 2323)",
                    "peak_source": hbm_kind, "launches_timed": len(lad_ms), "avg_launch_ms": sum(lad_ms) / len(lad_ms),
                    "algorithmic_per_launch": sum(alg) / len(alg)}
        else:
            roof = None
        s_per_iter = ms * 1e-3 / args.steps
        line = {
            **({"NOT_A_BENCHMARK": f"rank 0's share of a {args.as_rank_of}-rank job, no exchange (--as-rank-of)"} if view_comm else {}),
            "metric": metric_name(o, v, m), "value": s_per_iter, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, world),
            "roofline": roof, "clocks": clocks, "gpu_launches": int(launches), "peer_copies": int(py_rccsd.get_stat("push_copies")),
            "ladder_ms_per_step": sum(lad_ms) / args.steps if lad_ms else None, "stage_ms_per_step": stage_ms,
            "stage_ms_slowest_step": stage_worst, "stage_host_issue_ms_of_that_step": stage_worst_host,
            "scratch_allocations_in_timed_region": {"cudaMallocAsync_calls": int(alloc_calls), "served_from_cached_blocks": int(alloc_hits),
                                                    "note": "rank 0; the library keeps freed scratch blocks per stream, so a steady-state "
                                                            "step does not enter the driver (a cudaMallocAsync that has to map memory "
                                                            "queues behind nvidia-smi's queries and can stall the host for 0.03-2.7 s)"},
            "output_checksums": checks, "barrier_wait_ms_by_rank": waits_by_rank,
            "schedule": ("single rank: plugin calls (py_rccsd.contract_DTSpT) on the full operands" if world == 1 else
                         "plugin calls on L-slice views; each rank stores its finished slice into every peer's copy over "
                         "NVLink" + (" (copy engines where the slice is made of rows >= 4 KB, a store kernel otherwise)" if args.push_copy_engine else "") +
                         (" on the compute stream" if args.push_inline else " on a side stream under the next term's kernel") +
                         "; 2 " + ("NCCL 4-byte all-reduce" if args.nccl_barrier else "device-side flag") + " barriers per step, no NCCL on the data path"),
            "timed_segments": "whole steps" if it.resident else
                              f"sum of segments; the ladder's W is regenerated in HBM between {len(it.sub)} sub-shards (an input)",
        }
        if it.tc:
            # achieved int8 rate of the O^3V^3 stages (whole stage: planes of the operand, sign matrix, tensor-core kernel,
            # fp64 remainder), rank 0's share: 2 * (o v)^3 / N * planes operations per dense product
            prod = {"0112_0313_0113": 2 if it.fuse else 3, "0312": 1, "1302_1202": 1 if it.fuse else 2, "1303": 1, "1203": 1}
            per = 2.0 * float(o * v) ** 3 / world * args.planes
            line["o3v3_stage_int8_tops"] = {
                k: round(n * per / (stage_ms[k] * 1e-3) / 1e12, 1) for k, n in prod.items() if stage_ms.get(k, 0.0) > 0.0}
            line["o3v3_stage_int8_tops_note"] = ("dense products per stage " + json.dumps(prod) + "; unpadded int8 operations of the "
                                                 "whole stage on rank 0 over its CUDA-event time; the ladder's kernel-only figure is roofline.achieved")
        if other_ms is not None:
            other = "invariant" if args.w_mode == "refreshed" else "refreshed"
            line["value_by_w_mode"] = {args.w_mode: s_per_iter, other: other_ms * 1e-3}
        if pl_ms:
            palg = sum((8.0 + args.planes) * nK * c for c in pl_cols)
            line["roofline_planes"] = {
                "bound": "hbm", "kernel": "planes_predict_kernel (ladder operand -> int8 digit planes: one streaming pass with the previous "
                                         "build's column exponents + verification; the first build of a geometry is planes_fused_kernel)",
                "achieved": palg / (sum(pl_ms) * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                "frac": palg / (sum(pl_ms) * 1e-3) / 1e9 / hbm, "peak_source": hbm_kind,
                "avg_launch_ms": sum(pl_ms) / len(pl_ms), "launches_timed": len(pl_ms),
                "avg_launch_ms_beside_the_compression": (sum(pl_over_ms) / len(pl_over_ms)) if pl_over_ms else None,
                "algorithmic_per_launch": palg / len(pl_ms), "traffic": ncu_traffic_per_launch("planes_predict")}
        line["config"]["ladder_subshards"] = len(it.sub)
        line["config"]["ladder_planes"] = (
            "encoded inside the step" + (": the first sub-shard on a second stream beside the compression of T2 (W does not depend "
                                         "on T2), the others in line" if it.plane_side is not None else ", in line")
            if args.w_mode == "refreshed" and it.tc else "built once outside the step" if it.tc else None)
        line["config"]["compression"] = {
            "ranges": "range-sharded: each rank scans its contiguous 1/N of T2, all-reduced key histograms, exact sums "
                      "carried rank to rank; norm, D and running sums shared by the step's two draws",
            "pair": "task-parallel: even ranks compute the first compression, odd ranks the second, entries broadcast",
            "replicated": "both compressions on every rank"}[it.compress_mode()]
        line["config"]["precision"] = (
            f"fp64 results via {args.planes}x8-bit block fixed point: each column of W (one power-of-two scale per column) is "
            f"expanded into {args.planes} int8 digit planes, tail <= 2^-{8 * args.planes} of the column's largest element per "
            "term, exact int32 accumulation, fp64 recombination; contraction parity <= 1e-10 relative (tests/test_plugin_tc_gpu.py, "
            "incl. heavy-tailed W)") if it.tc else "fp64 FMA"
        line["config"]["constant_operands"] = (
            "eris.ovov declared constant (py_rccsd.constant_operand): its digit planes are built once, "
            f"{py_rccsd.get_stat('constant_plane_hits')} plugin calls reused them" if (it.tc and not args.no_constant_eris)
            else "none declared")
        line["config"]["fused_pairs"] = ("0112+0113 (one weighted sign matrix) and 1302+1202 (planes of 2 Wvoov - Wvovo) through "
                                         "py_rccsd.contract_DTSpT_pair: 9 dense products for the 11 terms" if it.fuse else
                                         "none: eleven plugin calls")
        line["config"]["contraction_kernel"] = (f"int8 tensor cores, {args.planes} digit planes + fp64 remainder, chosen by "
                                                f"fricc_contract's cost model ({tc_calls} of the timed plugin calls)"
                                                if it.tc else "fp64 bucketed SpMM")
        if not args.no_e2e:
            line["e2e"] = {"value": e2e_ms * 1e-3 / args.steps, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "stage_ms_per_step": e2e_stage_ms, "stage_ms_slowest_step": e2e_stage_worst,
                           "through": ("T2 from pinned host memory (one upload), py_rccsd.SparseTensor4d x2 + contract_DTSpT x11 (W "
                                       "and the intermediates are device tensors), residual t2new to pinned host memory under "
                                       "the last four contractions; the ladder's first sub-shard is encoded ahead of its call "
                                       "(py_rccsd.prepare_operand, second stream) while T2 is uploaded and compressed" if world == 1 else
                                       "each rank: its 1/N of T2 from pinned host memory, range-sharded compression, "
                                       "py_rccsd.contract_DTSpT on L-slice views, its own 1/N of the residual (the i-range "
                                       "its 0101 call completes) back to pinned host memory under the remaining contractions; "
                                       "the ladder operand is encoded ahead of its call (py_rccsd.prepare_operand, second stream) "
                                       "while T2 is uploaded and compressed"),
                           "outputs_read_back": "t2new only (a device-resident caller: tmp0..2, which the reference driver "
                                                "symmetrises on the host, fricc/friccsd.py:252-284, stay in HBM; a numpy caller that "
                                                "wants all four O2V2 arrays on the host pays 4x these d2h bytes)",
                           "prepared_plane_hits": int(py_rccsd.get_stat("prepared_plane_hits"))}
        if sweep is not None:
            line["compress_sweep"] = sweep
        if spot is not None:
            # the evidence at this scale that does not rest on a stored value: random elements of all seven outputs of the
            # verification step against a host recomputation (tolerance of north_star: 1e-10 of the terms' own scale)
            line["spot_check"] = spot
            line["spot_check_max_rel"] = spot["max_rel"]
            assert spot["max_rel_to_abs_sum"] <= 1e-10, f"spot check failed: {spot}"
        if parity is not None:
            line["compress_parity_at_scale"] = parity
            assert parity["idx_equal"] and parity["vals_equal"], f"compression differs from the oracle: {parity}"
        default_cfg = (o, v, m, args.sampling) == (40, 400, 10_000_000, "systematic") and args.planes == 6
        if default_cfg and world == 1 and view_comm is None and not args.no_pairs:
            # the same small configurations the reference arm runs in full (its "measured_configs"), each in a fresh process
            # of this script on this GPU: a measured same-config pair beside the extrapolated headline
            pairs = []
            for o2, v2, m2 in SAME_CONFIG_PAIRS:
                cmd = [sys.executable, os.path.abspath(__file__), "--o", str(o2), "--v", str(v2), "--m", str(m2), "--steps", "10",
                       "--warmup", "3", "--sampling", args.sampling, "--no-cpu-baseline", "--no-pairs", "--no-sweep", "--no-clocks"]
                try:
                    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
                    sub = json.loads(res.stdout.strip().splitlines()[-1])
                    pairs.append({"o": o2, "v": v2, "m": m2, "gpu_s": sub["value"], "gpu_e2e_s": sub.get("e2e", {}).get("value"),
                                  "spot_check_max_rel": sub.get("spot_check_max_rel"), "n_gpus": 1})
                except Exception as e:   # reported, never fatal for the headline line
                    pairs.append({"o": o2, "v": v2, "m": m2, "error": repr(e)[:300]})
            line["same_config_pairs"] = pairs
        if default_cfg and EXPECTED_T2NEW_CHECKSUM is not None and view_comm is None:
            rel = abs(checks["t2new"] - EXPECTED_T2NEW_CHECKSUM) / abs(EXPECTED_T2NEW_CHECKSUM)
            line["t2new_checksum_vs_stored"] = {"rel_diff": rel, "ok": rel <= 1e-12}
            assert rel <= 1e-9, f"t2new checksum {checks['t2new']!r} differs from the stored {EXPECTED_T2NEW_CHECKSUM!r}"
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference_sample(o, v, m, args.sampling, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": r["extrapolated_s"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                                    "sample": r["sample"], "same_config": r["same_config"]}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if os.environ.get("FRICC_BENCH_NOGC"):      # diagnostic
        import gc

        gc.disable()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
