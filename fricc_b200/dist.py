"""One-process-per-GPU orchestration of the FRI-CCSD inner loop.

Sharding: every contraction  out[R, L] += scale * sum_K T[R, K] * W[K, L]  is split along its FREE pair L (the
non-contracted pair of W): rank g owns the contiguous L-columns [g*lc, (g+1)*lc) of the [K][L] panel of W (and of its
int8 digit planes, fricc_dense.cu) and produces the same columns of out.  The compressed tensor T (<= 16*m bytes) is
replicated.  There is no data-path collective inside a contraction; the one exchange step per output is either
  * stores into NVLink peer memory (SymmArena + ShardedTerm.accumulate(arena=...) / push_columns, fricc_peer.cu),
    optionally issued on a SideStream so that they run under the next term's kernel -- the default of bench.py --, or
  * an NCCL all-gather of column panels + fricc_accumulate_panels (ShardedTerm.local / start_exchange / finish),
after which every rank holds the full output -- which is what the next compression needs.

Compression: `compress` (replicated), `compress_pair` (the two compressions of an iteration task-parallel on even / odd
ranks; any sampler), `compress_ranges[_pair]` (fricc_b200/ranges.py: every rank scans its contiguous 1/G of T2,
all-reduced key histograms, exact sums carried rank to rank; systematic sampler).

torch is used for device buffers, streams and torch.distributed only; all arithmetic is in libfricc_b200.so.
"""
import ctypes

import torch

from . import _lib
from ._lib import check, i64x4
from .py_rccsd import SparseTensor4d

# term -> (W axes, out axes) in o/v extents, and which W axes form (K, L) / which out axes form (R, L)
# (mirror of the C++ term table, fricc_contract.cu; used for sizes only)
TERM_AXES = {
    "2323": dict(w="vvvv", out="oovv", wK=(2, 3), wL=(0, 1), oR=(0, 1), oL=(2, 3)),
    "0101": dict(w="oooo", out="oovv", wK=(0, 1), wL=(2, 3), oR=(2, 3), oL=(0, 1)),
    "1323": dict(w="ovov", out="oooo", wK=(1, 3), wL=(0, 2), oR=(2, 3), oL=(0, 1)),
    "1302": dict(w="voov", out="oovv", wK=(1, 3), wL=(0, 2), oR=(3, 1), oL=(2, 0)),
    "1202": dict(w="vovo", out="oovv", wK=(1, 2), wL=(0, 3), oR=(3, 1), oL=(2, 0)),
    "1303": dict(w="voov", out="oovv", wK=(1, 3), wL=(0, 2), oR=(3, 1), oL=(2, 0)),
    "1203": dict(w="vovo", out="oovv", wK=(1, 2), wL=(0, 3), oR=(2, 1), oL=(3, 0)),
    "0112": dict(w="ovov", out="voov", wK=(0, 1), wL=(2, 3), oR=(0, 2), oL=(1, 3)),
    "0313": dict(w="ovov", out="voov", wK=(0, 3), wL=(2, 1), oR=(0, 2), oL=(1, 3)),
    "0113": dict(w="ovov", out="voov", wK=(0, 1), wL=(2, 3), oR=(0, 2), oL=(1, 3)),
    "0312": dict(w="ovov", out="vovo", wK=(0, 3), wL=(2, 1), oR=(0, 3), oL=(1, 2)),
}


def _ptr(t):
    return ctypes.c_void_p(int(t.data_ptr()))


def column_range(nL, rank, world):
    """Columns of the free pair owned by `rank`: uniform width lc = ceil(nL/world) (the last shard may be short)."""
    lc = -(-nL // world)
    l0 = min(rank * lc, nL)
    return l0, min(lc, nL - l0), lc


class Comm:
    """Thin wrapper: torch.distributed when world > 1, no-ops otherwise."""

    def __init__(self, rank=0, world=1, group=None):
        self.rank, self.world, self.group = rank, world, group

    def all_gather(self, out, inp, async_op=False):
        if self.world == 1:
            if out.data_ptr() != inp.data_ptr():
                out.view(-1)[: inp.numel()].copy_(inp.view(-1))
            return None
        import torch.distributed as dist

        return dist.all_gather_into_tensor(out, inp, group=self.group, async_op=async_op)

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)


class ShardedTerm:
    """One contraction term sharded by L-columns.  The W panel shard is either handed over already permuted
    (`set_panel`, for operands that do not change between iterations) or permuted from a 4-D W on every call."""

    def __init__(self, term, o, v, comm, device):
        self.term, self.comm, self.device = term, comm, device
        ax = TERM_AXES[term]
        ext = {"o": o, "v": v}
        self.w_shape = tuple(ext[c] for c in ax["w"])
        self.out_shape = tuple(ext[c] for c in ax["out"])
        self.nK = self.w_shape[ax["wK"][0]] * self.w_shape[ax["wK"][1]]
        self.nL = self.w_shape[ax["wL"][0]] * self.w_shape[ax["wL"][1]]
        self.nR = self.out_shape[ax["oR"][0]] * self.out_shape[ax["oR"][1]]
        self.t_dims = (o, o, v, v)
        self.l0, self.lcnt, self.lc = column_range(self.nL, comm.rank, comm.world)
        self.panel = None        # [nK][lcnt] W panel shard
        self.out_panel = None    # [nR][lc]
        self.gathered = None     # [world][nR][lc]
        self.L = _lib.lib()
        self.planes = None       # int8 digit planes of `panel` (tensor-core form, fricc_dense.cu); None = fp64 SpMM only
        self.n_planes = 0
        self.tc_min_fraction = 0.5   # use the tensor-core form when at least this share of T is in the sign matrix
        self.auto_planes = 0         # > 0: permute_from() also (re)builds that many digit planes

    # -- operands ---------------------------------------------------------------------------------------------
    def panel_bytes(self):
        return self.nK * max(self.lcnt, 1) * 8

    def set_panel(self, panel, planes_from=None):
        """Use an already permuted panel (optionally together with the digit planes another term built from it)."""
        assert panel.dtype == torch.float64 and panel.numel() >= self.nK * self.lcnt
        self.panel = panel
        if planes_from is not None and planes_from.planes is not None:
            self.share_planes(planes_from)

    def permute_from(self, W):
        """W: 4-D float64 CUDA tensor (any strides) -> this rank's [nK][lcnt] panel (hand-written transpose kernel)."""
        if self.panel is None:
            self.panel = torch.empty(self.nK * max(self.lcnt, 1), dtype=torch.float64, device=self.device)
        if self.lcnt > 0:
            check(self.L.fricc_permute_w(_ptr(W), _lib.FRICC_DEVICE, i64x4(W.shape), i64x4(W.stride()),
                                         self.term.encode(), self.l0, self.lcnt, _ptr(self.panel)))
            if self.auto_planes > 0:     # keep the tensor-core operand in step with the panel
                if self.planes is None:
                    self.build_planes(self.auto_planes)
                else:
                    self.refresh_planes()

    # -- tensor-core form ------------------------------------------------------------------------------------
    def build_planes(self, n_planes=6, panel=None, lcnt=None):
        """Expand this rank's W panel into int8 digit planes (error-free up to 2^(-8*n_planes) of each column's largest
        element).  Call again (or `refresh_planes`) whenever the panel's values change."""
        panel = self.panel if panel is None else panel
        lcnt = self.lcnt if lcnt is None else lcnt
        self.drop_planes()
        if lcnt <= 0:
            return
        h = ctypes.c_void_p()
        check(self.L.fricc_wslices_create(_ptr(panel), lcnt, self.nK, lcnt, int(n_planes), ctypes.byref(h)))
        self.planes, self.n_planes, self._planes_lcnt = h, int(n_planes), lcnt

    def refresh_planes(self, panel=None):
        panel = self.panel if panel is None else panel
        if self.planes is not None:
            check(self.L.fricc_wslices_update(self.planes, _ptr(panel), self._planes_lcnt))

    def share_planes(self, other):
        """Terms that read the same panel share its planes (not owned: only `other` frees them)."""
        self.planes, self.n_planes, self._planes_lcnt, self._planes_borrowed = other.planes, other.n_planes, other._planes_lcnt, True

    def drop_planes(self):
        if self.planes is not None and not getattr(self, "_planes_borrowed", False):
            self.L.fricc_wslices_destroy(self.planes)
        self.planes, self._planes_borrowed = None, False

    def __del__(self):
        try:
            self.drop_planes()
        except Exception:
            pass

    def use_tc(self, T):
        """Tensor-core form for this (term, T)?  Needs planes and a T whose entries mostly share one magnitude (an FRI
        compression); density gating is the caller's choice of whether to build planes at all."""
        if self.planes is None:
            return False
        f = ctypes.c_double(0.0)
        check(self.L.fricc_st_sign_fraction(T._h, self.term.encode(), ctypes.byref(f)))
        return f.value >= self.tc_min_fraction

    # -- compute ----------------------------------------------------------------------------------------------
    def local(self, T, into=None):
        """This rank's column slice of the output: out_panel[R][c] (+)= scale * sum_K T[R,K] * Wp[K][c].
        `into`: another ShardedTerm with the same (R, L) output mapping whose panel this term is ADDED to (terms that
        feed the same output share one exchange, e.g. 0112/0313/0113 -> Wvoov)."""
        fresh = into is None
        if into is not None:
            assert (into.nR, into.lc, into.l0) == (self.nR, self.lc, self.l0)
            target = into.out_panel
        else:
            if self.out_panel is None:
                self.out_panel = torch.empty(self.nR * self.lc, dtype=torch.float64, device=self.device)
            target = self.out_panel
        tc = self.lcnt > 0 and self.use_tc(T)
        if fresh and not (tc and self.lcnt == self.lc):
            target.zero_()       # the tensor-core kernel overwrites its columns; everything else accumulates
        if self.lcnt > 0:
            if tc:
                check(self.L.fricc_contract_panel_tc(self.planes, _ptr(self.panel), self.lcnt, self.l0, T._h,
                                                     _ptr(target), self.lc, self.term.encode(), 1 if fresh else 0))
            else:
                check(self.L.fricc_contract_panel(_ptr(self.panel), self.lcnt, self.l0, self.lcnt, T._h,
                                                  _ptr(target), self.lc, self.term.encode()))
        return target

    def start_exchange(self):
        """Launch the all-gather of the column slices without waiting for it: NCCL runs on its own stream behind the work
        already queued on the current one, so the next term's SpMM overlaps the transfer."""
        self._work = None
        if self.comm.world > 1:
            if self.gathered is None:
                self.gathered = torch.empty(self.comm.world * self.nR * self.lc, dtype=torch.float64, device=self.device)
            self._work = self.comm.all_gather(self.gathered, self.out_panel, async_op=True)

    def finish(self, T, out):
        """Wait for the gather started by start_exchange and add the panels into the full 4-D output."""
        w = self.comm.world
        if w == 1:
            g = self.out_panel
        else:
            if self._work is not None:
                self._work.wait()
                self._work = None
            g = self.gathered
        check(self.L.fricc_accumulate_panels(_ptr(g), w, self.lc, T._h, self.term.encode(), _ptr(out),
                                             i64x4(out.shape), i64x4(out.stride())))

    def exchange_and_accumulate(self, T, out):
        """all-gather the column slices and add them into the full 4-D output on every rank."""
        self.start_exchange()
        self.finish(T, out)

    def local_into(self, T, out, panel=None, l0=None, lcnt=None):
        """Single-rank form: accumulate this panel's columns straight into the 4-D output (no intermediate panel)."""
        panel = self.panel if panel is None else panel
        l0 = self.l0 if l0 is None else l0
        lcnt = self.lcnt if lcnt is None else lcnt
        if lcnt > 0:
            if self.planes is not None and lcnt == self._planes_lcnt and self.use_tc(T):
                check(self.L.fricc_contract_panel_into_tc(self.planes, _ptr(panel), lcnt, l0, T._h, _ptr(out),
                                                          i64x4(out.shape), i64x4(out.stride()), self.term.encode()))
            else:
                check(self.L.fricc_contract_panel_into(_ptr(panel), lcnt, l0, lcnt, T._h, _ptr(out), i64x4(out.shape),
                                                       i64x4(out.stride()), self.term.encode()))

    def direct_ok(self, out):
        """Is a kernel epilogue writing straight into `out` coalesced?  (threads run over R-minor, each thread over 8
        consecutive L-minor indices: one of the two must be out's unit-stride axis)"""
        ax = TERM_AXES[self.term]
        st = out.stride()
        return st[ax["oL"][1]] == 1 or st[ax["oR"][1]] == 1

    def accumulate(self, T, out, also=(), arena=None, side=None):
        """out[.., this rank's L-columns, ..] += this term (+ the terms in `also`, which must map (R, L) onto `out` the
        same way and use the same T).  With an arena (SymmArena holding `out` on every rank) the finished columns are
        also stored into every peer's copy -- the exchange step.  Terms whose output's unit-stride axis is R-major /
        L-major (1302, 1202, 1303, 1203 on a C-contiguous t2new) go through a coalesced [R][lc] panel and one
        transposing scatter pass; the others write straight from the kernel's epilogue.
        side: a SideStream -- the stores into the peers (NVLink-bound, a few SMs' worth of work) are issued on it behind
        the compute, so they run under the NEXT term's kernel; the caller joins it before the next barrier."""
        terms = (self,) + tuple(also)
        for t in terms:
            assert (t.nR, t.lc, t.l0) == (self.nR, self.lc, self.l0)
        peers = arena.peers_of(out) if (arena is not None and self.comm.world > 1) else None
        if peers is None:
            side = None
        if self.direct_ok(out):
            for t in terms:
                t.local_into(T, out)
            if peers is not None and self.lcnt > 0:
                def push():
                    check(self.L.fricc_push_columns(_ptr(out), peers, self.comm.world, self.comm.rank,
                                                    self.term.encode(), i64x4(self.t_dims), i64x4(out.shape),
                                                    i64x4(out.stride()), self.l0, self.lcnt))
                side.run(push) if side is not None else push()
            return
        self.local(T)
        for t in also:
            t.local(T, into=self)
        if self.lcnt > 0:
            def scatter():
                check(self.L.fricc_scatter_panel(_ptr(self.out_panel), self.lc, self.l0, self.lcnt, self.term.encode(),
                                                 i64x4(self.t_dims), _ptr(out), peers, self.comm.world if peers else 1,
                                                 self.comm.rank if peers else 0, i64x4(out.shape),
                                                 i64x4(out.stride())))
            side.run(scatter) if side is not None else scatter()

    def __call__(self, T, out, W=None):
        if W is not None:
            self.permute_from(W)
        if self.comm.world == 1:
            self.accumulate(T, out)
            return
        self.local(T)
        self.exchange_and_accumulate(T, out)


def use_torch_stream():
    """Issue the library's work on torch's current stream (so torch.cuda.Event timing and NCCL ordering see it)."""
    s = torch.cuda.current_stream().cuda_stream
    check(_lib.lib().fricc_set_stream(ctypes.c_void_p(s)))


def compress(t2, m, compression="fri", sampling_method="systematic", draws=None):
    """Replicated compression of the (replicated) dense T2: every rank runs the same deterministic kernels on the
    same data with the same draws, so the sparse tensors are identical without any exchange."""
    return SparseTensor4d(t2, tuple(t2.shape), int(m), compression, sampling_method, False, draws=draws)


def _assumed_consumption(n, sampling_method):
    """rand() draws one compression of an n-element tensor takes from the global stream: systematic draws once
    (fri_utils.cpp:47); pivotal draws once per visited element, n - 1 unless the walk stops early (fri_utils.cpp:74,103)."""
    return 1 if sampling_method == "systematic" else max(n - 1, 0)


def compress_pair(t2, m, comm, compression="fri", sampling_method="systematic", draws=None):
    """The TWO compressions of one FRICCSD iteration (fricc/friccsd.py:100-136: the same t2 compressed twice with fresh
    draws -> t2_sparse, t2_sparse_alt), computed task-parallel: even ranks compute the first, odd ranks the second, and
    the entries (16*m bytes each) are broadcast over NVLink, so no rank pays for both.

    draws: None -> both compressions draw from the library's single glibc-compatible stream, in the reference's order
    (the second starts where the first stopped); (d_first, d_second) -> explicit draw buffers (parity tests, bench).
    Results are identical to two consecutive `compress` calls on every rank."""
    world = comm.world
    L = _lib.lib()
    if world == 1:
        if draws is None:
            return (compress(t2, m, compression, sampling_method), compress(t2, m, compression, sampling_method))
        return (compress(t2, m, compression, sampling_method, draws=draws[0]),
                compress(t2, m, compression, sampling_method, draws=draws[1]))
    import torch.distributed as dist

    n = t2.numel()
    mine = comm.rank % 2
    state = None
    if draws is None:
        state = ctypes.create_string_buffer(L.fricc_rand_state_size())
        check(L.fricc_rand_get_state(state))
        assumed = _assumed_consumption(n, sampling_method)
        if mine == 1:
            check(L.fricc_rand_skip(assumed))
        T_mine = compress(t2, m, compression, sampling_method)
    else:
        T_mine = compress(t2, m, compression, sampling_method, draws=draws[mine])
    nnz = T_mine.size()
    # payload per compression: [flat (nnz) | n_exact | consumed draws] and the values
    idx = [torch.empty(nnz + 2, dtype=torch.int64, device=t2.device) for _ in range(2)]
    val = [torch.empty(max(nnz, 1), dtype=torch.float64, device=t2.device) for _ in range(2)]
    ne = ctypes.c_int64(0)
    check(L.fricc_st_get_n_exact(T_mine._h, ctypes.byref(ne)))
    check(L.fricc_st_export(T_mine._h, _ptr(idx[mine]), _ptr(val[mine]), _lib.FRICC_DEVICE))
    idx[mine][nnz:] = torch.tensor([ne.value, int(T_mine.consumed_draws)], dtype=torch.int64, device=t2.device)
    works = []
    for which in (0, 1):   # rank 0 holds the first compression, rank 1 the second
        works.append(dist.broadcast(idx[which], src=which, group=comm.group, async_op=True))
        works.append(dist.broadcast(val[which], src=which, group=comm.group, async_op=True))
    for w in works:
        w.wait()
    tail = torch.stack([idx[0][nnz:], idx[1][nnz:]]).cpu().tolist()   # [[n_exact, consumed], [n_exact, consumed]]
    other = 1 - mine
    T_other = SparseTensor4d.from_coo(idx[other][:nnz], val[other][:nnz], tuple(t2.shape))
    check(L.fricc_st_set_n_exact(T_other._h, int(tail[other][0])))
    T_other.consumed_draws = int(tail[other][1])
    pair = [None, None]
    pair[mine], pair[other] = T_mine, T_other
    if draws is None:
        c1 = int(tail[0][1])
        check(L.fricc_rand_set_state(state))
        check(L.fricc_rand_skip(c1))
        if c1 != assumed:
            # the first walk stopped early: the second compression started from the wrong stream position on the odd
            # ranks -- redo it everywhere from the right one (rare; keeps the result exact)
            pair[1] = compress(t2, m, compression, sampling_method)
        else:
            check(L.fricc_rand_skip(int(tail[1][1])))
    return pair[0], pair[1]


class _CudaArray:
    """Zero-copy view of a raw device pointer for torch.as_tensor (the __cuda_array_interface__ protocol)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class SymmArena:
    """One symmetric device allocation per rank for the contraction outputs (fricc_peer.cu): `tensor(shape)` carves the
    same region out of every rank's allocation (every rank must request the same shapes in the same order), and
    `peers_of(t)` gives the addresses of that region in all ranks' copies, so a rank can store its L-columns of an output
    straight into its peers' HBM over NVLink.  world == 1: a plain allocation, no handles."""

    def __init__(self, nbytes, comm, device):
        self.comm, self.device, self.L = comm, device, _lib.lib()
        self.nbytes = int(nbytes)
        self.base = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        check(self.L.fricc_symm_alloc(self.nbytes, ctypes.byref(self.base), handle))
        self.peer_base = [None] * comm.world
        self.peer_base[comm.rank] = self.base.value
        self._opened = []
        if comm.world > 1:
            import torch.distributed as dist

            handles = [None] * comm.world
            dist.all_gather_object(handles, bytes(handle.raw), group=comm.group)
            for q, h in enumerate(handles):
                if q == comm.rank:
                    continue
                p = ctypes.c_void_p()
                check(self.L.fricc_symm_open(ctypes.create_string_buffer(h, 64), ctypes.byref(p)))
                self.peer_base[q] = p.value
                self._opened.append(p)
        self._bytes = torch.as_tensor(_CudaArray(self.base.value, self.nbytes), device=device)
        self._used = 0

    def tensor(self, shape, dtype=torch.float64):
        """A tensor carved from the arena (uninitialised)."""
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        off = (self._used + 255) // 256 * 256
        if off + nbytes > self.nbytes:
            raise MemoryError("SymmArena exhausted")
        self._used = off + nbytes
        return self._bytes[off: off + nbytes].view(dtype).view(*shape)

    def peers_of(self, t):
        """Array of the world's addresses of `t` (a tensor carved from this arena)."""
        off = int(t.data_ptr()) - self.base.value
        assert 0 <= off < self.nbytes
        return (ctypes.c_void_p * self.comm.world)(*[b + off for b in self.peer_base])

    def close(self):
        self._bytes = None
        for p in self._opened:
            self.L.fricc_symm_close(p)
        self._opened = []
        if self.base.value:
            self.L.fricc_symm_free(self.base)
            self.base = ctypes.c_void_p()


def push_columns(term_obj, t_dims, out, arena, side=None):
    """Store this rank's L-columns of `out` (already accumulated locally) into every peer's copy."""
    if term_obj.comm.world == 1 or term_obj.lcnt <= 0:
        return

    def push():
        check(term_obj.L.fricc_push_columns(_ptr(out), arena.peers_of(out), term_obj.comm.world, term_obj.comm.rank,
                                            term_obj.term.encode(), i64x4(t_dims), i64x4(out.shape),
                                            i64x4(out.stride()), term_obj.l0, term_obj.lcnt))
    side.run(push) if side is not None else push()


class SideStream:
    """A second CUDA stream for the peer stores of the exchange step.  run(fn): fn's library calls are issued on the
    side stream, ordered behind everything queued on the current stream so far; join(): the current stream waits for
    everything issued on the side stream (call it before a barrier, or before the pushed output is written again)."""

    def __init__(self, device, priority=0):
        self.stream = torch.cuda.Stream(device=device, priority=priority)
        self.L = _lib.lib()

    def run(self, fn):
        main = torch.cuda.current_stream()
        self.stream.wait_stream(main)
        check(self.L.fricc_set_stream(ctypes.c_void_p(self.stream.cuda_stream)))
        try:
            fn()
        finally:
            check(self.L.fricc_set_stream(ctypes.c_void_p(main.cuda_stream)))

    def join(self):
        torch.cuda.current_stream().wait_stream(self.stream)


# range-sharded compression (x split by contiguous flat-index ranges; all-reduced key histograms): fricc_b200/ranges.py
from .ranges import compress_ranges, compress_ranges_pair, index_range, chunk_range  # noqa: E402,F401


# ======================================================================================================================
# the sharded step on top of the PLUGIN CALL: every rank hands `py_rccsd.contract_DTSpT` its L-slice of the operand and
# of the output as plain 4-D views -- the call the reference's drivers make (src/py_rccsd.cpp:50-51) -- and the library
# picks the kernel (tensor cores for FRI-compressed tensors, planes built straight from the view).  No panels, no
# term-specific entry points: what a single-GPU user of the module runs is exactly what each rank runs here.
# ======================================================================================================================
class ViewEngine:
    """Views + plugin calls + peer stores.  One instance per rank.

    slice_axis(term, shape, world): which of the two free-pair axes is cut across the ranks (see there).
    contract(term, W, T, out, w_range=None): this rank's slice of `term`.  `W` is the FULL operand (replicated small
        tensors: ovov, Woooo, Wvoov, Wvovo) unless w_range = (lo, hi) says that `W` already IS the slice [lo, hi) of the
        cut axis (the ladder's vvvv operand, of which a rank only ever holds its part, possibly in sub-shards).
    publish(term, out, w_range=None): stores the slice of `out` this rank owns into every peer's copy (SymmArena).
    barrier(): stream-ordered barrier of the ranks through the arena (fricc_peer_barrier), NCCL-free."""

    def __init__(self, comm, device, arena=None, side=None):
        from . import py_rccsd

        self.comm, self.device, self.arena, self.side = comm, device, arena, side
        self.py = py_rccsd
        self.L = _lib.lib()
        self._epoch = 0
        self._flags = None
        if arena is not None and comm.world > 1:
            self._flags = arena.tensor((16,), dtype=torch.int64).zero_()
            self._flag_peers = arena.peers_of(self._flags)
            torch.cuda.synchronize()
            comm.barrier()          # every rank's counters are zero before anyone stores an epoch

    MIN_ROW_BYTES = 4096          # rows a copy engine moves at link speed (fricc_push_view's own threshold)

    @staticmethod
    def slice_axis(term, w_shape, world=1):
        """0 or 1: which axis of the free pair is cut.  The longer one (ties: the major one) -- except when that would
        cut the OUTPUT along its innermost axis into rows shorter than MIN_ROW_BYTES while the other axis splits evenly
        among the ranks: then the other one, so that a rank's slice of the output is made of long contiguous rows (peer
        stores through the copy engines, contiguous runs for the kernels' own loads and stores)."""
        ax = TERM_AXES[term]
        e = (w_shape[ax["wL"][0]], w_shape[ax["wL"][1]])
        j = 0 if e[0] >= e[1] else 1
        if world > 1 and ax["oL"][j] == 3 and -(-e[j] // world) * 8 < ViewEngine.MIN_ROW_BYTES and e[1 - j] % world == 0:
            j = 1 - j
        return j

    def my_range(self, extent):
        lo, cnt, _ = column_range(extent, self.comm.rank, self.comm.world)
        return lo, lo + cnt

    def out_view(self, term, out, j, lo, hi):
        return out.narrow(TERM_AXES[term]["oL"][j], lo, hi - lo)

    def contract(self, term, W, T, out, w_range=None):
        ax = TERM_AXES[term]
        if w_range is None:
            j = self.slice_axis(term, W.shape, self.comm.world)
            lo, hi = self.my_range(W.shape[ax["wL"][j]])
            if hi <= lo:
                return None
            Wv = W.narrow(ax["wL"][j], lo, hi - lo)
        else:
            j, (lo, hi), Wv = 0, w_range, W
            if hi <= lo:
                return None
        ov = self.out_view(term, out, j, lo, hi)
        self.py.contract_DTSpT(Wv, T, ov, term)
        return ov

    def own_view(self, term, w_shape, out, w_range=None):
        """The part of `out` this rank computes for `term` (None if its range is empty): what has to be zero before the
        term accumulates -- the rest of a rank's copy is overwritten by its peers' stores."""
        ax = TERM_AXES[term]
        if w_range is None:
            j = self.slice_axis(term, w_shape, self.comm.world)
            lo, hi = self.my_range(w_shape[ax["wL"][j]])
        else:
            j, (lo, hi) = 0, w_range
        return self.out_view(term, out, j, lo, hi) if hi > lo else None

    def contract_pair(self, term_a, term_b, W_a, W_b, T, out):
        """This rank's slice of two terms that go into the same output through the same slicing, as one fused call
        (py_rccsd.contract_DTSpT_pair)."""
        axa, axb = TERM_AXES[term_a], TERM_AXES[term_b]
        j = self.slice_axis(term_a, W_a.shape, self.comm.world)
        if j != self.slice_axis(term_b, W_b.shape, self.comm.world) or axa["oL"] != axb["oL"]:
            raise ValueError(f"terms {term_a} and {term_b} are not sliced alike")
        lo, hi = self.my_range(W_a.shape[axa["wL"][j]])
        if hi <= lo:
            return None
        ov = self.out_view(term_a, out, j, lo, hi)
        self.py.contract_DTSpT_pair(W_a.narrow(axa["wL"][j], lo, hi - lo), W_b.narrow(axb["wL"][j], lo, hi - lo), T, ov,
                                    term_a, term_b)
        return ov

    def contract_in_parts(self, term, W, T, out, parts=2):
        """`contract` + `publish` with this rank's range cut into `parts`: the stores of one part run under the kernel of
        the next, so that only the last part's stores are left when the term is the last one before a barrier."""
        ax = TERM_AXES[term]
        j = self.slice_axis(term, W.shape, self.comm.world)
        lo, hi = self.my_range(W.shape[ax["wL"][j]])
        parts = max(1, min(int(parts), hi - lo))
        for q in range(parts):
            a = lo + (hi - lo) * q // parts
            b = lo + (hi - lo) * (q + 1) // parts
            if b <= a:
                continue
            ov = self.out_view(term, out, j, a, b)
            self.py.contract_DTSpT(W.narrow(ax["wL"][j], a, b - a), T, ov, term)
            self.publish(ov)

    def publish(self, view):
        """Store a 4-D view of an arena tensor into the same view of every peer's copy (on the side stream if any)."""
        if view is None or self.arena is None or self.comm.world == 1:
            return
        peers = self.arena.peers_of(view)

        def push():
            check(self.L.fricc_push_view(_ptr(view), peers, self.comm.world, self.comm.rank, i64x4(view.shape),
                                         i64x4(view.stride())))
        self.side.run(push) if self.side is not None else push()

    def join(self):
        if self.side is not None:
            self.side.join()

    def barrier(self):
        if self.comm.world == 1:
            return
        self.join()
        if self._flags is None:
            import torch.distributed as dist

            if dist.is_available() and dist.is_initialized():
                dist.barrier(group=self.comm.group)
            return
        self._epoch += 1
        check(self.L.fricc_peer_barrier(self._flag_peers, self.comm.world, self.comm.rank, self._epoch))
