#!/usr/bin/env python
"""bench.py -- one FRI-CCSD inner-loop iteration (the hot path of jamesETsmith/fri-cc) on synthetic tensors.

Metric (BASELINE.json): FRI-CCSD s/iter at o=40, v=400, m=1e7 on 1/2/4/8 GPUs; ladder HBM GB/s vs peak.

A "step" is one pass of the hot path exactly as fricc/friccsd.py:update_amps drives it with all three contraction
classes compressed: 2 FRI compressions of T2 (N = o^2 v^2 entries -> m) and the 11 sparse contractions
  1323 (ovov -> Woooo), 0112 0313 0113 (ovov -> Wvoov), 0312 (ovov -> Wvovo),
  0101 (Woooo), 2323 (Wvvvv: the particle-particle ladder), 1302 1303 (Wvoov), 1202 1203 (Wvovo)   -> t2new / tmp.
Multi-GPU: every contraction is sharded along its free index pair (column slices of the [K][L] panel of W); each rank
accumulates its columns into its own copy of the 4-D output and stores them into every peer's copy over NVLink
(symmetric allocation, stores issued on a side stream under the next term's kernel; `--exchange nccl` = all-gather of
panels instead), so every rank holds the full result; strong scaling (fixed problem).  The step's two compressions come
from one range-sharded pass (every rank scans its contiguous 1/N of T2; `--compress`).  For FRI-compressed tensors the
contractions run in their tensor-core form (`--planes`, fricc_dense.cu; `--planes 0` = fp64 SpMM).

Timing: W >= 3 untimed warm-up steps, then K steps; barrier + torch.cuda.synchronize on both sides; CUDA events on the
launching stream; max over ranks.  Inputs are far larger than L2 (the ladder streams >= 25 GB of W per GPU per step).
When the ladder's W slice does not fit in HBM (1 GPU: 204.8 GB > 180 GB) it is processed in column sub-shards that are
re-generated in HBM between timed segments (inputs resident when each timed segment starts); the step time is then
the sum of the timed segments.  `e2e` repeats the measurement with T2 arriving from pinned host memory and the
residual going back to pinned host memory inside the timed region (W stays resident: it is 204.8 GB, more than the
host has, and is iteration-invariant in the linearised driver, fricc/frilccsd.py:171-174).

`--impl reference` times the reference's own CPU implementation (oracle/_ref, the unmodified sources compiled by
oracle/Makefile) on a bounded sample of the same workload and extrapolates by operation count.
"""
import argparse
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

METRIC = "FRI-CCSD s/iter at o=40,v=400,m=1e7"
UNIT = "s/iter"

CONTRACTION_ORDER = ["1323", "0112", "0313", "0113", "0312", "0101", "2323", "1302", "1202", "1303", "1203"]


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
    p.add_argument("--max-panel-gb", type=float, default=26.0, help="largest resident ladder W slice per sub-shard")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--only-ladder", action="store_true", help="debug: time only the ladder launches")
    p.add_argument("--planes", type=int, default=6,
                   help="int8 digit planes of W for the tensor-core form of the contractions (0 = fp64 SpMM only)")
    p.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                   help="N>1: how column slices reach the other ranks: stores into peer memory over NVLink (default) or "
                        "NCCL all-gather of panels + accumulate")
    p.add_argument("--compress", default="auto", choices=["auto", "ranges", "pair", "replicated"],
                   help="how the two compressions of a step are computed: 'ranges' = every rank scans its contiguous 1/N "
                        "of T2, all-reduced key histograms, norm / D / running sums shared by both draws (systematic "
                        "sampler; default for it); 'pair' = whole compressions task-parallel on even/odd ranks; "
                        "'replicated' = both on every rank")
    p.add_argument("--push-inline", action="store_true",
                   help="N>1, peer exchange: issue the stores into the peers on the compute stream (A/B of the side-stream "
                        "push that runs them under the next term's kernel)")
    p.add_argument("--push-ctas", type=int, default=0, help="peer-store kernels: CTAs per SM (0 = library default)")
    p.add_argument("--no-overlap", action="store_true", help="N>1: one blocking exchange per term (A/B of the pipelined step)")
    return p.parse_args()


# ======================================================================================================================
# reference arm / CPU baseline: the reference's own OpenMP implementation on the host cores
# ======================================================================================================================
def flops_model(o, v, m):
    """FMAs per term at (o, v, m)."""
    f = {"2323": m * v * v, "0101": m * o * o, "1323": m * o * o}
    for t in ("1302", "1202", "1303", "1203", "0112", "0313", "0113", "0312"):
        f[t] = m * o * v
    return f


def cpu_reference_sample(o, v, m, sampling, threads):
    """Time oracle/_ref (kind 'reference') -- or the C restatement (kind 'port') if the compiled reference is not
    there -- on a bounded sample of the workload and extrapolate each part to (o, v, m) by its operation count."""
    os.environ["OMP_NUM_THREADS"] = str(threads)
    import numpy as np

    from oracle import oracle

    R = oracle.ref()
    kind = "reference" if R is not None else "port"
    P = oracle.port()
    so, sv = min(o, 40), min(v, 100)
    sN = so * so * sv * sv
    sm_comp = max(1, min(int(round(sN * m / (o * o * v * v))), sN))   # same density as the target
    sm_con = max(1, min(100_000, sN // 10))
    rng = np.random.default_rng(1)
    t2 = (rng.standard_normal(sN) * 10.0 ** (-6.0 * rng.random(sN) ** 2)).reshape(so, so, sv, sv)
    ext = {"o": so, "v": sv}
    from fricc_b200.dist import TERM_AXES

    times = {}
    t0 = time.perf_counter()
    if R is not None:
        T1 = R.mod.SparseTensor4d(t2, t2.shape, sm_comp, "fri", sampling, False)
        T2_ = R.mod.SparseTensor4d(t2, t2.shape, sm_comp, "fri", sampling, False)
        del T1, T2_
    else:
        draws = rng.integers(0, 2**31 - 1, size=sN).astype(np.int32)
        P.fri_compression(t2, sm_comp, sampling, draws)
        P.fri_compression(t2, sm_comp, sampling, draws)
    times["compress_x2"] = time.perf_counter() - t0
    if R is not None:
        T = R.mod.SparseTensor4d(t2, t2.shape, sm_con, "largest")
    else:
        flat, vals = P.largest(t2, sm_con)
    for term in CONTRACTION_ORDER:
        ax = TERM_AXES[term]
        W = rng.random([ext[c] for c in ax["w"]])
        out = np.zeros([ext[c] for c in ax["out"]])
        t0 = time.perf_counter()
        if R is not None:
            R.mod.contract_DTSpT(W, T, out, term)
        else:
            P.contract(term, W, flat, vals, t2.shape, out)
        times[term] = time.perf_counter() - t0
    fs, ft = flops_model(so, sv, sm_con), flops_model(o, v, m)
    extrap = times["compress_x2"] * (o * o * v * v) / sN   # N log N treated as linear: favours the reference
    for term in CONTRACTION_ORDER:
        extrap += times[term] * ft[term] / fs[term]
    sample = (f"oracle/{'_ref' if R is not None else 'port'}: 2x SparseTensor4d(fri,{sampling}) at o={so},v={sv},"
              f"m={sm_comp} + the 11 contract_DTSpT terms at o={so},v={sv},m={sm_con}; measured {sum(times.values()):.2f} s;"
              f" extrapolated to o={o},v={v},m={int(m)} per part by FMA count (contractions) and N (compression)")
    return dict(kind=kind, cores=threads, sample=sample, measured_s=sum(times.values()), extrapolated_s=extrap,
                parts={k: round(x, 4) for k, x in times.items()})


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    o, v, m = args.o, args.v, int(args.m)
    vals, meas = [], []
    for it in range(args.warmup + args.steps):
        r = cpu_reference_sample(o, v, m, args.sampling, threads)
        if it >= args.warmup:
            vals.append(r["extrapolated_s"])
            meas.append(r["measured_s"])
    value = sum(vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(meas) / len(meas), "higher_is_better": False,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, world):
    o, v, m = args.o, args.v, int(args.m)
    return {
        "workload": f"synthetic FRI-CCSD iteration o={o},v={v},m={m}: 2 FRI compressions ({args.sampling}) of T2 "
                    f"(N={o*o*v*v}) + 11 sparse contractions; vvvv panel sharded by (a,b) columns over {world} GPU(s)",
        "o": o, "v": v, "m": m, "sampling": args.sampling, "parallelism": f"free-pair column shards x{world}",
        "l2": "inputs_larger_than_L2",
    }


# ======================================================================================================================
# the B200 arm
# ======================================================================================================================
class Iteration:
    """Device-resident synthetic FRI-CCSD iteration, one instance per rank."""

    def __init__(self, args, comm, device):
        import torch

        from fricc_b200 import dist as fd

        self.torch, self.fd = torch, fd
        self.args, self.comm, self.device = args, comm, device
        o, v = args.o, args.v
        self.o, self.v, self.m = o, v, int(args.m)
        self.N = o * o * v * v
        g = torch.Generator(device=device)
        g.manual_seed(20)
        # heavy-tailed amplitudes (a few dominant entries -> non-empty D), identical on every rank
        u = torch.rand(self.N, generator=g, device=device, dtype=torch.float64)
        z = torch.randn(self.N, generator=g, device=device, dtype=torch.float64)
        self.t2 = (z * torch.pow(10.0, -6.0 * u * u)).reshape(o, o, v, v).contiguous()
        del u, z
        self.terms = {t: fd.ShardedTerm(t, o, v, comm, device) for t in CONTRACTION_ORDER}
        # tensor-core form when T is dense enough for it to win (fricc_dense.cu: break-even ~1.3 % at 6 planes)
        self.planes = args.planes if (args.planes > 0 and self.m / self.N >= 0.015) else 0
        for t in self.terms.values():
            t.auto_planes = self.planes
        # ---- dense operands (U(0,1) like the reference's tests, fricc/tests/test_sparse_tensors.py:78)
        self.ovov = torch.rand(o, v, o, v, generator=g, device=device, dtype=torch.float64)
        # iteration-invariant operand -> its panels are permuted once, outside the timed region
        for t in ("1323", "0112", "0313"):
            self.terms[t].permute_from(self.ovov)
        # terms that read the same operand through the same (K, L) axes share one panel
        self.terms["0113"].set_panel(self.terms["0112"].panel, planes_from=self.terms["0112"])
        self.terms["0312"].set_panel(self.terms["0313"].panel, planes_from=self.terms["0313"])
        self.peer = comm.world > 1 and args.exchange == "peer" and not args.no_overlap and not args.only_ladder
        self.side = None
        if self.peer:
            # the outputs live in one symmetric allocation per rank: every rank stores its columns into all copies
            nb = 8 * (o ** 4 + 2 * v * o * o * v + 4 * o * o * v * v) + 8 * 4096
            self.arena = fd.SymmArena(nb, comm, device)
            mk = lambda *shape: self.arena.tensor(shape).zero_()
        else:
            self.arena = None
            mk = lambda *shape: torch.zeros(*shape, device=device, dtype=torch.float64)
        self.Wklij = mk(o, o, o, o)
        self.Wakic = mk(v, o, o, v)
        self.Wakci = mk(v, o, v, o)
        self.t2new = mk(o, o, v, v)
        self.tmp = [mk(o, o, v, v) for _ in range(3)]
        self.flag = torch.zeros(1, device=device, dtype=torch.float32)
        # ---- the ladder operand: this rank's (a,b)-column slice of Wp[(c,d)][(a,b)], possibly in sub-shards
        lad = self.terms["2323"]
        max_cols = max(1, int(args.max_panel_gb * 1e9 / (8 * lad.nK)))
        self.sub = []
        c = 0
        while c < lad.lcnt:
            w = min(max_cols, lad.lcnt - c)
            self.sub.append((c, w))
            c += w
        self.lad_width = max((w for _, w in self.sub), default=1)
        self.lad_panel = torch.empty(lad.nK * self.lad_width, device=device, dtype=torch.float64)
        self.resident = len(self.sub) <= 1
        self.plane_build_ms = []
        if self.resident and self.sub:
            self.fill_ladder_panel(0)
        self.draws = torch.randint(0, 2**31 - 1, (64,), generator=torch.Generator().manual_seed(5)).numpy().astype("int32")
        self.draw_pos = 0
        self.ladder_events = []
        self.stage_events = []

    def fill_ladder_panel(self, s):
        """(Re)generate sub-shard s of the ladder panel in HBM -- an INPUT, not part of the timed work."""
        torch = self.torch
        c, w = self.sub[s]
        lad = self.terms["2323"]
        g = torch.Generator(device=self.device)
        g.manual_seed(1000 + lad.l0 + c)
        view = self.lad_panel[: lad.nK * w]
        # chunked fill keeps the generator's temporary small
        step = 1 << 28
        for a in range(0, view.numel(), step):
            view[a: a + step].uniform_(0.0, 1.0, generator=g)
        if self.planes > 0:
            # the digit planes are a re-encoding of this INPUT (like its [K][L] permutation): built with it, untimed,
            # and reported separately (`ladder_plane_build_ms`)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if lad.planes is not None and lad._planes_lcnt == w:
                lad.refresh_planes(self.lad_panel)
            else:
                lad.build_planes(self.planes, panel=self.lad_panel, lcnt=w)
            e1.record()
            self.plane_build_ms.append((e0, e1))

    def compress(self, t2):
        d = self.draws[self.draw_pos % 32:]
        self.draw_pos += 1
        return self.fd.compress(t2, self.m, "fri", self.args.sampling, draws=d if self.args.sampling == "systematic" else None)

    def ladder_local(self, T, timer=None):
        """Particle-particle ladder on this rank's column slice (into t2new directly when there is one rank)."""
        torch = self.torch
        lad = self.terms["2323"]
        L = lad.L
        single = self.comm.world == 1   # one rank: accumulate straight into t2new (no panel, no gather)
        if not single:
            if lad.out_panel is None:
                lad.out_panel = torch.empty(lad.nR * lad.lc, dtype=torch.float64, device=self.device)
            lad.out_panel.zero_()
        for s, (c, w) in enumerate(self.sub):
            if not self.resident:
                if timer:
                    timer.pause()
                self.fill_ladder_panel(s)
                if timer:
                    timer.resume()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if single:
                lad.local_into(T, self.t2new, panel=self.lad_panel, l0=lad.l0 + c, lcnt=w)
            else:
                if lad.planes is not None and lad._planes_lcnt == w and lad.use_tc(T):
                    self.fd.check(L.fricc_contract_panel_tc(lad.planes, self.fd._ptr(self.lad_panel), w, lad.l0 + c, T._h,
                                                            ctypes_ptr(lad.out_panel, c), lad.lc, b"2323", 0))
                else:
                    self.fd.check(L.fricc_contract_panel(self.fd._ptr(self.lad_panel), w, lad.l0 + c, w, T._h,
                                                         ctypes_ptr(lad.out_panel, c), lad.lc, b"2323"))
            e1.record()
            self.ladder_events.append((e0, e1, w))

    def ladder(self, T, timer=None):
        self.ladder_local(T, timer)
        if self.comm.world > 1:
            self.terms["2323"].exchange_and_accumulate(T, self.t2new)

    def mark(self, name):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        self.stage_events.append((name, e))

    def compress_mode(self):
        c = self.args.compress
        if c == "auto":
            if self.args.sampling == "systematic":
                return "ranges"
            return "pair" if (self.comm.world > 1 and not self.args.no_overlap) else "replicated"
        return c

    def step(self, t2, timer=None):
        self.mark("begin")
        mode = self.compress_mode()
        if mode == "ranges":
            # every rank scans only its contiguous 1/N of T2 (a view: T2 is replicated); the two compressions differ
            # only in the draw, so norm, D and the running sums of p are computed once (fricc_b200/ranges.py)
            d = (self.draws[self.draw_pos % 32:], self.draws[(self.draw_pos + 1) % 32:])
            self.draw_pos += 2
            flat = t2.view(-1)
            j0, cnt = self.fd.index_range(flat.numel(), self.comm.rank, self.comm.world)
            T, T_alt = self.fd.compress_ranges_pair(flat[j0:j0 + cnt], j0, tuple(t2.shape), self.m, self.comm, "fri",
                                                    "systematic", draws=d)
        elif mode == "pair" and self.comm.world > 1:
            # task-parallel: even ranks compress with the first draws, odd ranks with the second, entries broadcast
            d = None
            if self.args.sampling == "systematic":
                d = (self.draws[self.draw_pos % 32:], self.draws[(self.draw_pos + 1) % 32:])
            self.draw_pos += 2
            T, T_alt = self.fd.compress_pair(t2, self.m, self.comm, "fri", self.args.sampling, draws=d)
        else:
            T = self.compress(t2)
            T_alt = self.compress(t2)
        self.mark("compress_x2")
        if self.peer:
            return self.step_peer(T, T_alt, timer)
        if self.comm.world > 1 and not self.args.only_ladder and not self.args.no_overlap:
            return self.step_pipelined(T, T_alt, timer)
        tm = self.terms
        if not self.args.only_ladder:
            self.Wklij.zero_(); self.Wakic.zero_(); self.Wakci.zero_()
            self.t2new.zero_()
            for t in self.tmp:
                t.zero_()
            self.mark("zero_outputs")
            tm["1323"](T_alt, self.Wklij)
            self.mark("1323")
            tm["0112"](T_alt, self.Wakic)
            self.mark("0112")
            tm["0313"](T_alt, self.Wakic)
            self.mark("0313")
            tm["0113"](T_alt, self.Wakic)
            self.mark("0113")
            tm["0312"](T_alt, self.Wakci)
            self.mark("0312")
            tm["0101"](T, self.t2new, W=self.Wklij)
            self.mark("0101")
        self.ladder(T, timer)
        self.mark("2323_incl_regen" if not self.resident else "2323")
        if not self.args.only_ladder:
            tm["1302"](T, self.tmp[0], W=self.Wakic)
            self.mark("1302")
            tm["1202"](T, self.tmp[0], W=self.Wakci)
            self.mark("1202")
            tm["1303"].set_panel(tm["1302"].panel, planes_from=tm["1302"])   # same Wvoov panel (K=(k,c), L=(a,i))
            tm["1203"].set_panel(tm["1202"].panel, planes_from=tm["1202"])   # same Wvovo panel
            tm["1303"](T, self.tmp[1])
            self.mark("1303")
            tm["1203"](T, self.tmp[2])
            self.mark("1203")
        return T

    def rank_barrier(self):
        """Stream-ordered barrier (a 4-byte NCCL all-reduce): everything every rank queued before it -- including its
        stores into this rank's memory -- has completed when the work queued behind it starts.  Does not block the host."""
        import torch.distributed as dist

        dist.all_reduce(self.flag, group=self.comm.group)

    def step_peer(self, T, T_alt, timer=None):
        """N > 1, exchange over peer memory: every term accumulates its L-columns straight into this rank's copy of the
        4-D output and the finished columns are stored into every peer's copy over NVLink (fricc_peer.cu) -- no panels,
        no all-gather, no re-accumulation pass.  Three barriers order consumers behind writers:
          B0 after the outputs are zeroed (a peer must not zero what was already pushed to it),
          B1 before Woooo / Wvoov / Wvovo are read as operands and before 0101 (column axis (i,j)) accumulates onto the
             ladder's t2new (column axis (a,b)): an output must be complete everywhere before a term with a different
             column axis adds to it,
          B2 at the end of the step."""
        tm, dims, push = self.terms, (self.o, self.o, self.v, self.v), self.fd.push_columns
        if self.side is None and not self.args.push_inline:
            self.side = self.fd.SideStream(self.device)
        S = None if self.args.push_inline else self.side   # peer stores run on it, under the next term's kernel
        self.Wklij.zero_(); self.Wakic.zero_(); self.Wakci.zero_()
        self.t2new.zero_()
        for t in self.tmp:
            t.zero_()
        self.rank_barrier()
        self.mark("zero_outputs_B0")
        A = self.arena
        # the ladder first: its 2 GB push then hides under the ovov terms; the o^4-sized output goes last, so that
        # only its (tiny) push is exposed in front of the barrier
        lad = tm["2323"]
        for s_, (c, w) in enumerate(self.sub):
            if not self.resident:
                if timer:
                    timer.pause()
                self.fill_ladder_panel(s_)
                if timer:
                    timer.resume()
            e0, e1 = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
            e0.record()
            lad.local_into(T, self.t2new, panel=self.lad_panel, l0=lad.l0 + c, lcnt=w)
            e1.record()
            self.ladder_events.append((e0, e1, w))
        push(lad, dims, self.t2new, A, side=S)
        self.mark("2323" if self.resident else "2323_incl_regen")
        tm["0112"].accumulate(T_alt, self.Wakic, also=(tm["0313"], tm["0113"]), arena=A, side=S)
        self.mark("0112_0313_0113")
        tm["0312"].accumulate(T_alt, self.Wakci, arena=A, side=S)
        self.mark("0312")
        tm["1323"].accumulate(T_alt, self.Wklij, arena=A, side=S)
        self.mark("1323")
        if S is not None:
            S.join()
        self.rank_barrier()
        self.mark("B1")
        tm["1302"].permute_from(self.Wakic); tm["1202"].permute_from(self.Wakci)
        tm["1302"].accumulate(T, self.tmp[0], also=(tm["1202"],), arena=A, side=S)
        self.mark("1302_1202")
        tm["1303"].set_panel(tm["1302"].panel, planes_from=tm["1302"]); tm["1303"].accumulate(T, self.tmp[1], arena=A, side=S)
        self.mark("1303")
        tm["1203"].set_panel(tm["1202"].panel, planes_from=tm["1202"]); tm["1203"].accumulate(T, self.tmp[2], arena=A, side=S)
        self.mark("1203")
        tm["0101"].permute_from(self.Wklij); tm["0101"].accumulate(T, self.t2new, arena=A, side=S)
        self.mark("0101")
        if S is not None:
            S.join()
        self.rank_barrier()
        self.mark("B2")
        return T

    def step_pipelined(self, T, T_alt, timer=None):
        """N > 1: the same 11 contractions, ordered so that every all-gather (NCCL, its own stream) runs under the next
        term's SpMM, and terms that feed the same output through the same (R, L) map share one panel and one exchange
        (0112+0313+0113 -> Wvoov, 1302+1202 -> tmp).  Data dependences: 0101 reads Woooo (1323), 1302/1303 read Wvoov,
        1202/1203 read Wvovo (0312)."""
        tm = self.terms
        self.Wklij.zero_(); self.Wakic.zero_(); self.Wakci.zero_()
        self.t2new.zero_()
        for t in self.tmp:
            t.zero_()
        self.mark("zero_outputs")
        tm["1323"].local(T_alt); tm["1323"].start_exchange()
        self.mark("1323_local")
        tm["0112"].local(T_alt)
        tm["0313"].local(T_alt, into=tm["0112"])
        tm["0113"].local(T_alt, into=tm["0112"])
        tm["0112"].start_exchange()
        self.mark("0112_0313_0113_local")
        tm["0312"].local(T_alt); tm["0312"].start_exchange()
        self.mark("0312_local")
        self.ladder_local(T, timer)                       # the long one: hides the three gathers above
        tm["2323"].start_exchange()
        self.mark("2323_local" if self.resident else "2323_local_incl_regen")
        tm["1323"].finish(T_alt, self.Wklij)
        tm["0112"].finish(T_alt, self.Wakic)
        tm["0312"].finish(T_alt, self.Wakci)
        self.mark("finish_Woooo_Wvoov_Wvovo")
        tm["0101"].permute_from(self.Wklij); tm["0101"].local(T); tm["0101"].start_exchange()
        self.mark("0101_local")
        tm["1302"].permute_from(self.Wakic); tm["1302"].local(T)
        tm["1202"].permute_from(self.Wakci); tm["1202"].local(T, into=tm["1302"])
        tm["1302"].start_exchange()
        self.mark("1302_1202_local")
        tm["2323"].finish(T, self.t2new)
        self.mark("finish_2323")
        tm["1303"].set_panel(tm["1302"].panel, planes_from=tm["1302"]); tm["1303"].local(T); tm["1303"].start_exchange()
        self.mark("1303_local")
        tm["0101"].finish(T, self.t2new)
        tm["1203"].set_panel(tm["1202"].panel, planes_from=tm["1202"]); tm["1203"].local(T); tm["1203"].start_exchange()
        self.mark("finish_0101_1203_local")
        tm["1302"].finish(T, self.tmp[0])
        tm["1303"].finish(T, self.tmp[1])
        tm["1203"].finish(T, self.tmp[2])
        self.mark("finish_tmp0_tmp1_tmp2")
        return T

    def stage_breakdown(self, steps):
        """ms per step spent between consecutive marks (rank-local; includes any exchange inside the stage)."""
        self.torch.cuda.synchronize()
        out = {}
        for (n0, e0), (n1, e1) in zip(self.stage_events, self.stage_events[1:]):
            if n1 == "begin":
                continue
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        self.stage_events.clear()
        return {k: round(x / steps, 3) for k, x in out.items()}


def ncu_traffic_per_launch(which="spmm_ladder"):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE ladder launch (1600 x 20000-column slice, m=1e7), from the
    committed `ncu --set full` capture (profiles/ncu_full_*_spmm_ladder.csv, newest round); None if absent."""
    import csv
    import glob

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


def ctypes_ptr(t, elem_offset=0):
    import ctypes

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


def start_clock_sampler(gpu_index):
    path = tempfile.NamedTemporaryFile(prefix="clocks_", suffix=".csv", delete=False).name
    q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                 "-lms", "200"], stdout=open(path, "w"), stderr=subprocess.DEVNULL)
    except Exception:
        return None, path
    return proc, path


def stop_clock_sampler(proc, path):
    if proc is None:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for line in open(path):
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
    # median over the upper half of the samples = under load (the sampler also sees the idle gaps)
    load = sm[len(sm) // 2:] if sm else []
    return {"sm_mhz": (load[len(load) // 2] if load else None), "sm_max_mhz": (max(mx) if mx else None),
            "reasons": sorted(reasons), "samples": len(sm)}


def run_b200_arm(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    import torch
    import torch.distributed as dist

    from fricc_b200 import _lib
    from fricc_b200 import dist as fd

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

    if args.push_ctas > 0:
        _lib.check(L.fricc_set_option(b"push_ctas_per_sm", args.push_ctas))
    it = Iteration(args, comm, device)
    o, v, m = it.o, it.v, it.m
    lad = it.terms["2323"]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up
    for _ in range(max(args.warmup, 0)):
        it.step(it.t2)
    sync_all()
    it.ladder_events.clear()
    it.stage_events.clear()

    # ---- timed: device-resident inputs
    proc, cpath = start_clock_sampler(local_rank) if rank == 0 else (None, "")
    launches0 = L.fricc_kernel_launches()
    sync_all()
    timer = SegmentTimer(torch)
    timer.resume()
    for _ in range(args.steps):
        T = it.step(it.t2, timer)
    timer.pause()
    sync_all()
    ms = timer.total_ms()
    launches = L.fricc_kernel_launches() - launches0
    clocks = stop_clock_sampler(proc, cpath) if rank == 0 else None
    stage_ms = it.stage_breakdown(args.steps)
    # checksums of the step's outputs (after the last timed step): identical between the pipelined and the blocking
    # schedule, between ranks, and between GPU counts up to summation order
    checks = {k: float(v.double().sum().item()) for k, v in
              (("t2new", it.t2new), ("tmp0", it.tmp[0]), ("tmp1", it.tmp[1]), ("tmp2", it.tmp[2]),
               ("Wvoov", it.Wakic), ("Wvovo", it.Wakci), ("Woooo", it.Wklij))}
    lad_ms = [a.elapsed_time(b) for a, b, _ in it.ladder_events]
    lad_cols = [w for _, _, w in it.ladder_events]
    it.ladder_events.clear()
    nD = None

    # ---- e2e: T2 from pinned host memory, residual back to pinned host memory, inside the timed region
    e2e = None
    if not args.no_e2e:
        # every rank moves only ITS 1/world slice over PCIe; the full T2 is assembled over NVLink (NCCL all-gather) and
        # each rank returns its slice of the residual -- h2d/d2h bytes below are the totals over all ranks
        nel = it.t2.numel()
        per = -(-nel // world)
        lo, hi = min(rank * per, nel), min((rank + 1) * per, nel)
        h_t2 = torch.empty(per, dtype=torch.float64, pin_memory=True)
        h_t2[: hi - lo].copy_(it.t2.view(-1)[lo:hi])
        h_out = torch.empty(per, dtype=torch.float64, pin_memory=True)
        d_slice = torch.zeros(per, dtype=torch.float64, device=device)
        d_full = torch.empty(per * world, dtype=torch.float64, device=device)
        sync_all()
        t2timer = SegmentTimer(torch)
        t2timer.resume()
        for _ in range(args.steps):
            d_slice.copy_(h_t2, non_blocking=True)
            if world > 1:
                dist.all_gather_into_tensor(d_full, d_slice)
            else:
                d_full = d_slice
            it.step(d_full[:nel].view(it.t2.shape), t2timer)
            h_out[: hi - lo].copy_(it.t2new.view(-1)[lo:hi], non_blocking=True)
        t2timer.pause()
        sync_all()
        e2e_ms = t2timer.total_ms()
        it.ladder_events.clear()
        it.stage_events.clear()
    # ---- reduce over ranks (max of times, sum of launches)
    stats = torch.tensor([ms, e2e_ms if not args.no_e2e else 0.0, float(launches)], dtype=torch.float64, device=device)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm_ = stats.clone()
        dist.all_reduce(sm_, op=dist.ReduceOp.SUM)
        ms, e2e_ms, launches = float(mx[0]), float(mx[1]), int(sm_[2])
    if rank == 0:
        import json as _json

        peaks = {}
        try:
            peaks = _json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_kind = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        tc = it.planes > 0
        if tc:
            # dominant kernel: tc_sign_gemm_kernel on the ladder (int8 tensor cores).  Algorithmic work per launch:
            # 2 * rows * cols * nK * planes int8 operations (the dense sign-matrix x digit-plane product, unpadded).
            # Peak: int8 runs at twice the bf16 rate on B200 (4.5 vs 2.25 POP/s nominal), so 2 x the MEASURED cuBLAS bf16
            # figure; the ladder launches sit inside a long step -> the sustained (power-capped) figure.
            alg = [2.0 * lad.nR * c * lad.nK * it.planes for c in lad_cols]
            ach = sum(alg) / (sum(lad_ms) * 1e-3) / 1e12 if lad_ms else 0.0
            bf16 = float(peaks.get("bf16_tflops_sustained", 1400.0))
            peak = 2.0 * bf16
            peak_kind = ("2 x measured bf16_tflops_sustained (MEASURED_PEAKS.json); int8 = 2 x bf16 rate" if
                         "bf16_tflops_sustained" in peaks else "2 x fallback 1400 TFLOP/s sustained bf16")
            roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                    "traffic": ncu_traffic_per_launch("tc_ladder"),
                    "kernel": f"tc_sign_gemm_kernel (ladder 2323, int8 tcgen05.mma, {it.planes} digit planes)",
                    "peak_source": peak_kind, "peak_burst": 2.0 * float(peaks.get("bf16_tflops", 1590.0)),
                    "unit_note": "int8 TOP/s (2 x MAC)",
                    "fp64_equivalent_algorithmic_gbs": sum(8.0 * m * c for c in lad_cols) / (sum(lad_ms) * 1e-3) / 1e9
                    if lad_ms else None}
        else:
            # dominant kernel: spmm_rows_kernel on the ladder; algorithmic bytes per launch:
            # 8*m*cols (one W row slice per nonzero) + 8*o^2*cols (one R2 write) + 12*m (packed entries)
            alg = [8.0 * m * c + 8.0 * o * o * c + 12.0 * m for c in lad_cols]
            ach = sum(alg) / (sum(lad_ms) * 1e-3) / 1e9 if lad_ms else 0.0
            roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": ncu_traffic_per_launch("spmm_ladder"), "kernel": "spmm_rows_wide_kernel<2> (ladder 2323)",
                    "peak_source": peak_kind}
        roof.update(launches_timed=len(lad_ms), avg_launch_ms=(sum(lad_ms) / len(lad_ms)) if lad_ms else None,
                    algorithmic_per_launch=(sum(alg) / len(alg)) if alg else None)
        s_per_iter = ms * 1e-3 / args.steps
        line = {
            "metric": METRIC, "value": s_per_iter, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, world),
            "roofline": roof,
            "clocks": clocks, "gpu_launches": int(launches),
            "ladder_ms_per_step": sum(lad_ms) / args.steps, "stage_ms_per_step": stage_ms, "output_checksums": checks,
            "schedule": ("single rank" if world == 1 else "blocking NCCL exchange per term" if args.no_overlap else
                         "peer memory: each rank accumulates its columns in place and stores them into every peer's copy "
                         "over NVLink" + (" on the compute stream" if args.push_inline else " on a side stream under the next term's kernel") +
                         "; 3 stream-ordered barriers per step, no NCCL on the data path" if it.peer else
                         "pipelined: NCCL all-gathers overlap the next term's kernel; 7 exchanges for 11 terms"),
            "timed_segments": "whole steps" if it.resident else
                              f"sum of segments; ladder W regenerated in HBM between {len(it.sub)} column sub-shards",
        }
        line["config"]["ladder_subshards"] = len(it.sub)
        line["config"]["compression"] = {
            "ranges": "range-sharded: each rank scans its contiguous 1/N of T2, all-reduced key histograms, exact sums "
                      "carried rank to rank; norm, D and running sums shared by the step's two draws",
            "pair": "task-parallel: even ranks compute the first compression, odd ranks the second, entries broadcast",
            "replicated": "both compressions on every rank"}[it.compress_mode()]
        line["config"]["precision"] = ("fp64 results: W expanded error-free into int8 digit planes (tail 2^-48 of each column's "
                                       "largest element), exact int32 accumulation, fp64 recombination; measured max "
                                       "deviation from the fp64 SpMM 7e-15 relative (profiles/tc_vs_fp64_r01_H_1gpu.jsonl)"
                                       ) if tc else "fp64 FMA"
        line["config"]["contraction_kernel"] = (f"int8 tensor cores, {it.planes} digit planes + fp64 remainder" if tc
                                                else "fp64 bucketed SpMM")
        if tc and it.plane_build_ms:
            pb = [a.elapsed_time(b) for a, b in it.plane_build_ms]
            line["ladder_plane_build_ms"] = {"per_build": sum(pb) / len(pb), "builds": len(pb),
                                             "note": "untimed: re-encoding of the iteration-invariant ladder operand"}
        if not args.no_e2e:
            line["e2e"] = {"value": e2e_ms * 1e-3 / args.steps, "unit": UNIT,
                           "h2d_bytes_per_step": int(it.t2.numel() * 8), "d2h_bytes_per_step": int(it.t2new.numel() * 8)}
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference_sample(o, v, m, args.sampling, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": r["extrapolated_s"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                                    "sample": r["sample"]}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
