"""FRI compression of a tensor sharded by contiguous flat-index ranges over the ranks ("distributed
top-m uses an all-reduced histogram"; BASELINE config 5, the compression-only sweep at 1/2/4/8 GPUs).

Rank g holds x[j0_g : j0_g + n_g] of the flattened (o,o,v,v) tensor, rank order = index order.  The result is the SAME
SparseTensor4d, bit for bit, that the single-process compression produces (reference: fri_compression,
/root/reference/src/fri_utils.cpp:439-522, behind SparseTensor4d's constructor, include/sparse_tensor.hpp:58-67), and
every rank ends up holding it (the contractions replicate T).

What crosses NVLink per compression (m entries kept, G ranks):
  * 2 x (G doubles all-gather + G one-double broadcasts): the exact left-to-right sums (norm; running sum of p)
  * <= 6 all-reduces of a 4096-bin key histogram per candidate round (the top-|D| select)
  * an all-gather of the candidates at or above the cut (~65 536 (key, index) pairs in the common case)
  * all-reduce(SUM) of the m-entry index list and of the m-entry value list (each entry is written by exactly one rank)
All arithmetic is in libfricc_b200.so (fricc_shard.cu, fricc_escan.cu); torch moves bytes.
"""
import ctypes
import math

import numpy as np
import torch

from . import _lib
from ._lib import check
from .py_rccsd import SparseTensor4d

KEY_SHIFTS = (51, 39, 27, 15, 3, 0)       # digit schedule over the 63 magnitude bits (fricc_select.cu)
KEY_DIGIT_BITS = (12, 12, 12, 12, 12, 3)
FIRST_CANDIDATES = 65536                  # same first candidate-list size as the single-GPU compress_fri
SMALL_BIN = 1 << 16


def _ptr(t):
    return ctypes.c_void_p(int(t.data_ptr()))


def index_range(n_total, rank, world):
    """Contiguous flat-index range [j0, j0 + n) owned by `rank` (equal split, the last ranks may be one short)."""
    base, extra = divmod(int(n_total), int(world))
    j0 = rank * base + min(rank, extra)
    return j0, base + (1 if rank < extra else 0)


def chunk_range(n_total, rank, world):
    """parallel_sample's own chunking (fri_utils.cpp:183-184,215-218): `world` chunks of n // world elements, the last
    one takes the remainder.  The ranges compress_ranges(..., sampling_method="parallel") has to be given for its result
    to be the reference's parallel_sample with `world` threads."""
    step, rem = divmod(int(n_total), int(world))
    return rank * step, step + (rem if rank == world - 1 else 0)


class _Solo:
    """A communicator of one: exact_sum(shard, _SOLO) is the exact left-to-right sum of THIS shard alone."""
    rank, world, group = 0, 1, None


_SOLO = _Solo()


def pivotal_host(n_sample, probs, draws):
    """sample_pivotal (fri_utils.cpp:63-127) on a handful of numbers -- the fractional budgets of parallel_sample's
    chunks (:268), one per rank -- in host arithmetic (IEEE double, the reference's operations in its order).
    draws: rand() values, consumed one per element visited.  Returns (S, draws consumed); the reference's two exits
    (acceptance probability outside [0,1]; wrong number of samples) are raised as FriccError."""
    probs = [float(p) for p in probs]
    S, i, used = [], 0, 0
    if not probs:
        raise _lib.FriccError(_lib.EINVAL, "pivotal_host: empty probability vector")
    a = probs[0]
    last = len(probs) - 1
    for j in range(1, len(probs)):
        a += probs[j]
        if used >= len(draws):
            raise _lib.FriccError(_lib.EDRAWS, "parallel sampling: not enough budget draws (one per rank after the first)")
        rn = int(draws[used]) / 2147483647.0
        used += 1
        if a < 1.0:
            if a != 0.0 and rn < probs[j] / a:       # 0/0 is NaN in the reference: the comparison is false
                i = j
        else:
            a -= 1.0
            pa = (1.0 - probs[j]) / (1.0 - a)
            if pa > 1 or pa < 0:
                raise _lib.FriccError(_lib.EPIVOTAL, "ERROR IN PIVOTAL SAMPLING: acceptance probability outside [0,1] "
                                                     "(reference: fri_utils.cpp:84-94)")
            if rn < pa:
                S.append(i)
                i = j
            else:
                S.append(j)
        if len(S) == n_sample:
            break
        if j == last:
            if abs(a - 1.0) < 1e-5:
                S.append(i)
            else:
                raise _lib.FriccError(_lib.ECOUNT, "ERROR IN PIVOTAL SAMPLING: the walk ended short of n_sample "
                                                   "(reference: fri_utils.cpp:106-118)")
    if len(S) != n_sample:
        raise _lib.FriccError(_lib.ECOUNT, "ERROR IN PIVOTAL SAMPLING: wrong number of samples (reference: fri_utils.cpp:121-124)")
    return S, used


def chunk_budgets(n_sample, norms, q_norm, budget_draws):
    """parallel_sample's budget arithmetic (fri_utils.cpp:214-275) for chunks with exact 1-norms `norms`, |q|_1 = q_norm:
    a_j = n_sample * |q_j| / |q|, t_j = a_j - floor(a_j), g_j = floor(a_j), and the c = n_sample - sum(g) spare samples go
    to the chunks picked by sample_pivotal(c, t).  Returns (a, t, g, draws consumed).  Pure host logic."""
    a = [n_sample * float(nj) / q_norm for nj in norms]
    t = [aj - math.floor(aj) for aj in a]
    g = [int(math.floor(aj)) for aj in a]
    c = int(n_sample) - sum(g)
    if c < 0:
        raise _lib.FriccError(_lib.ECOUNT, "parallel_sample: the integer budgets exceed n_sample")
    picked, used = pivotal_host(c, t, budget_draws)
    for j in picked:
        g[j] += 1
    return a, t, g, used


def locate_threshold(hist_fn, want, small_bin=SMALL_BIN):
    """Walk the key digits from the top using GLOBAL (already all-reduced) histograms until the bin that straddles the
    `want`-th largest key holds at most `small_bin` elements (or the bits are exhausted).

    hist_fn(shift, prefix, prefix_bits) -> sequence of counts of the digit at `shift` among keys whose higher bits equal
    `prefix`.  Returns (key_min, count): `count` >= min(want, N) elements have key >= key_min, and key_min is the
    largest threshold of that digit depth with this property.  Pure host logic (tested on CPU with gloo)."""
    prefix, prefix_bits, above, rem = 0, 0, 0, int(want)
    if rem <= 0:
        return (1 << 63), 0          # above every key: nothing selected
    shift = 63
    bin_count = 0
    for shift, nbits in zip(KEY_SHIFTS, KEY_DIGIT_BITS):
        h = hist_fn(shift, prefix, prefix_bits)
        nb = 1 << nbits
        acc, b = 0, nb - 1
        while b >= 0:
            c = int(h[b])
            if acc + c >= rem:
                break
            acc += c
            b -= 1
        if b < 0:                    # fewer than `want` elements exist below this prefix: take them all
            return prefix << (shift + nbits), above + acc
        above += acc
        rem -= acc
        bin_count = int(h[b])
        prefix = (prefix << nbits) | b
        prefix_bits += nbits
        if bin_count == rem or bin_count <= small_bin:
            break
    return prefix << shift, above + bin_count


class _Shard:
    """RAII wrapper of fricc_shard_t."""

    def __init__(self, x_local, j0, n_total):
        self.L = _lib.lib()
        self.h = ctypes.c_void_p(None)
        self.x = x_local
        self.n = int(x_local.numel())
        self.j0, self.n_total = int(j0), int(n_total)
        check(self.L.fricc_shard_create(_ptr(x_local), self.n, self.j0, self.n_total, ctypes.byref(self.h)))

    def close(self):
        if self.h is not None and self.h.value:
            self.L.fricc_shard_destroy(self.h)
            self.h = ctypes.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _src(comm, g):
    if comm.group is None:
        return g
    import torch.distributed as dist

    return dist.get_global_rank(comm.group, g)


def _all_reduce(comm, t, op="sum"):
    if comm.world > 1:
        import torch.distributed as dist

        dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX, group=comm.group)
    return t


def exact_sum(shard, comm):
    """Exact left-to-right sum of the current addends (|x| or p) over ALL shards in index order; the per-chunk exact
    running sums stay in the shard for fricc_shard_emit_systematic.  Returns a 1-element device tensor (same on all
    ranks).  Stages: approximate shard sums (all-gather) -> chunk maps (parallel on every rank) -> one small resolve
    kernel per rank, the exact carry handed from rank to rank by a one-double broadcast (stream-ordered, no host sync)."""
    L, dev = shard.L, shard.x.device
    approx = torch.zeros(1, dtype=torch.float64, device=dev)
    check(L.fricc_shard_scan_local(shard.h, _ptr(approx)))
    if comm.world > 1:
        allg = torch.empty(comm.world, dtype=torch.float64, device=dev)
        comm.all_gather(allg, approx)
        carry_approx = allg[: comm.rank].sum().reshape(1) if comm.rank > 0 else torch.zeros_like(approx)
    else:
        carry_approx = torch.zeros_like(approx)
    check(L.fricc_shard_scan_maps(shard.h, _ptr(carry_approx)))
    carry = torch.zeros(1, dtype=torch.float64, device=dev)
    for g in range(comm.world):
        if g == comm.rank:
            check(L.fricc_shard_scan_resolve(shard.h, _ptr(carry), _ptr(carry)))
        if comm.world > 1:
            import torch.distributed as dist

            dist.broadcast(carry, src=_src(comm, g), group=comm.group)
    return carry


def global_candidates(shard, comm, want):
    """(keys, idx) of the `want` largest |x| over all shards, ordered by (|x| descending, flat index ascending), as
    device tensors (identical on every rank).  The cut is located with all-reduced key histograms; only the elements at
    or above it are gathered and sorted."""
    L, dev = shard.L, shard.x.device
    want = min(int(want), shard.n_total)
    hist = torch.zeros(4096, dtype=torch.int64, device=dev)

    def hist_fn(shift, prefix, prefix_bits):
        hist.zero_()
        check(L.fricc_key_histogram(_ptr(shard.x), shard.n, shift, ctypes.c_uint64(prefix), prefix_bits, _ptr(hist)))
        _all_reduce(comm, hist)
        return hist.cpu().numpy()

    key_min, total = locate_threshold(hist_fn, want)
    cap = max(min(total, shard.n), 1)
    keys = torch.empty(cap, dtype=torch.int64, device=dev)      # uint64 bit patterns of |x| (sign bit clear)
    idx = torch.empty(cap, dtype=torch.int64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    check(L.fricc_shard_candidates(shard.h, ctypes.c_uint64(key_min), _ptr(keys), _ptr(idx), cap, _ptr(cnt)))
    if comm.world > 1:
        counts = torch.empty(comm.world, dtype=torch.int64, device=dev)
        comm.all_gather(counts, cnt)
        counts = counts.cpu().tolist()
        if sum(counts) != total or counts[comm.rank] > cap:
            raise RuntimeError("distributed select: candidate counts disagree with the all-reduced histograms")
        width = max(max(counts), 1)
        pad_k = torch.zeros(width, dtype=torch.int64, device=dev)
        pad_i = torch.zeros(width, dtype=torch.int64, device=dev)
        pad_k[: counts[comm.rank]] = keys[: counts[comm.rank]]
        pad_i[: counts[comm.rank]] = idx[: counts[comm.rank]]
        all_k = torch.empty(comm.world * width, dtype=torch.int64, device=dev)
        all_i = torch.empty(comm.world * width, dtype=torch.int64, device=dev)
        comm.all_gather(all_k, pad_k)
        comm.all_gather(all_i, pad_i)
        keys = torch.cat([all_k[g * width: g * width + c] for g, c in enumerate(counts)])
        idx = torch.cat([all_i[g * width: g * width + c] for g, c in enumerate(counts)])
    else:
        c = int(cnt.item())
        if c != total:
            raise RuntimeError("select: candidate count disagrees with the histograms")
        keys, idx = keys[:c].clone(), idx[:c].clone()
    if total > 0:
        check(L.fricc_sort_candidates(_ptr(keys), _ptr(idx), total, shard.n_total))
    return keys[:want], idx[:want]


def _sample_systematic(shard, comm, nsamp, nD, idx, flat, mm, count, draws):
    """sample_systematic (fri_utils.cpp:129-159) over p = |x| / r * n', p[D] = 0 (never materialised): the running sums of
    p are computed once and shared by the `count` draws.  Fills flat[k*mm + ...] with this rank's part; returns the draws
    one compression consumes (1)."""
    L = shard.L
    the_draws = []
    for k in range(count):
        if draws is None:
            d = (ctypes.c_int32 * 1)()
            check(L.fricc_rand_fill(d, 1))
            the_draws.append(int(d[0]))
        else:
            d = np.asarray(draws[k]).reshape(-1)
            if d.size < 1:
                raise _lib.FriccError(_lib.EDRAWS, "systematic sampling needs one draw")
            the_draws.append(int(d[0]))
    if nsamp > 0:
        exact_sum(shard, comm)          # exact running sums of p at the chunk starts: shared by all draws
    for k in range(count):
        if nD > 0 and comm.rank == 0:   # summed over ranks below: rank 0 contributes D
            flat[k * mm: k * mm + nD] = idx[:nD]
    if nsamp > 0:                       # all draws from one pass over the shard
        d = (ctypes.c_int32 * count)(*the_draws)
        check(L.fricc_shard_emit_systematic_multi(shard.h, nsamp, d, count, _ptr(flat[nD:]), mm))
    return 1


_PEER_ERRORS = {_lib.EPROB: "ERROR MAKING PROBABILITY VECTOR", _lib.EPIVOTAL: "ERROR IN PIVOTAL SAMPLING",
                _lib.ECOUNT: "sampler produced the wrong number of samples", _lib.EDRAWS: "draw buffer too short",
                _lib.EINVAL: "bad argument", _lib.ECUDA: "CUDA error"}


def _sample_parallel(shard, comm, nsamp, nD, idx, flat, mm, count, draws):
    """parallel_sample (fri_utils.cpp:176-362, "Alg. A.2") with ONE CHUNK PER RANK over the same probabilities: exact
    |q|_1 over all ranks and |q_j|_1 per chunk (the scan stages), budgets on the host (chunk_budgets), then every rank
    rescales and samples its own chunk with the parallel pivotal sampler -- no state crosses a chunk boundary, so the
    ranks run at the same time.

    draws: None -> the library's rand() stream.  Every rank's copy advances by (world - 1) + N per compression: the
    budget draw reads the first world - 1 values, chunk j reads from position (world - 1) + j0 on (it consumes fewer than
    its length), so the chunks' draws never overlap and the result does not depend on which rank finishes first (the
    reference's threads share rand() in a race, :345).  Or, per compression, a pair (budget_draws, chunk_draws) of int32
    arrays: the values for the budget draw (same on every rank) and for THIS rank's chunk.
    Returns ((budget draws consumed, this chunk's draws consumed) of the last compression, first error or None)."""
    L, dev = shard.L, shard.x.device
    rank, world = comm.rank, comm.world
    consumed, failure = (0, 0), None
    norms = q_norm = None
    if nsamp > 0:
        q_norm = float(exact_sum(shard, comm).item())                 # one_norm(q), :191
        own = exact_sum(shard, _SOLO)                                  # one_norm(qj[j]), :221
        if world > 1:
            allg = torch.empty(world, dtype=torch.float64, device=dev)
            comm.all_gather(allg, own)
            norms = allg.cpu().tolist()
        else:
            norms = [float(own.item())]
    for k in range(count):
        if nD > 0 and rank == 0:
            flat[k * mm: k * mm + nD] = idx[:nD]
        if nsamp == 0:
            continue
        if draws is None:
            nb = world - 1
            # every rank's copy of the stream ends this compression exactly (world - 1) + N draws further on, whatever
            # happens in between (a failing rank that stopped short would leave the copies at different positions)
            stream_end = int(L.fricc_rand_consumed()) + nb + shard.n_total
            bd = (ctypes.c_int32 * max(nb, 1))()
            if nb > 0:
                check(L.fricc_rand_fill(bd, nb))
            budget_draws, chunk_draws = [int(v) for v in bd[:nb]], None
        else:
            budget_draws = np.asarray(draws[k][0], dtype=np.int32).reshape(-1)
            chunk_draws = np.ascontiguousarray(np.asarray(draws[k][1], dtype=np.int32).reshape(-1))
        used = ctypes.c_int64(0)
        try:
            a, t, g, used0 = chunk_budgets(nsamp, norms, q_norm, budget_draws)
            lb = sum(g[:rank])
            if chunk_draws is None:
                check(L.fricc_rand_skip(shard.j0))
                dptr, dn = None, 0
            else:
                dptr, dn = ctypes.c_void_p(chunk_draws.ctypes.data), chunk_draws.size
            check(L.fricc_shard_sample_chunk(shard.h, g[rank], t[rank], a[rank], norms[rank], dptr, dn, ctypes.byref(used),
                                             _ptr(flat[k * mm + nD + lb:])))
            consumed = (used0, int(used.value))
        except _lib.FriccError as e:          # raised on every rank after the collectives below (compress_ranges_multi)
            failure = failure or e
        if chunk_draws is None:
            check(L.fricc_rand_skip(max(0, stream_end - int(L.fricc_rand_consumed()))))
    return consumed, failure


def compress_ranges(x_local, j0, dims, m, comm, compression="fri", sampling_method="systematic", draws=None):
    """SparseTensor4d(tensor, dims, m, compression, sampling_method) for a tensor whose flattened elements
    [j0, j0 + x_local.numel()) live on this rank (CUDA, float64, contiguous).  Every rank returns the full, identical
    sparse tensor -- bit-identical to the single-process constructor on the concatenated tensor with the same draws.

    draws: None -> the library's glibc-compatible rand() stream (every rank advances its copy by the one draw the
    reference's systematic sampler consumes); or an int32 array whose first element is that draw.
    sampling_method "parallel": the chunked pivotal sampler (parallel_sample, "Alg. A.2", fri_utils.cpp:176-362) with one
    chunk per rank -- see _sample_parallel for the draws; with ranges from chunk_range() the samples are the reference's
    parallel_sample with `world` threads running in chunk order.
    sampling_method "pivotal" (ONE ordered walk over all N elements) is not range-sharded: use dist.compress_pair, which
    runs whole compressions task-parallel; here it raises NotImplementedError."""
    return compress_ranges_multi(x_local, j0, dims, m, comm, compression, sampling_method,
                                 draws=None if draws is None else [draws], count=1)[0]


def compress_ranges_pair(x_local, j0, dims, m, comm, compression="fri", sampling_method="systematic", draws=None):
    """The TWO compressions of one FRICCSD iteration (fricc/friccsd.py:100-136: the same t2 compressed twice, the
    second with the next draw -> t2_sparse, t2_sparse_alt).  With the systematic sampler the two differ ONLY in the one
    uniform draw: the norm, the exactly preserved set D, the remaining norm and the running sums of p are the same, so
    they are computed once and only the sample emission and the values are done per draw.  Results are bit-identical to
    two consecutive constructor calls.  draws: None (library stream, two draws consumed) or (d_first, d_second)."""
    T = compress_ranges_multi(x_local, j0, dims, m, comm, compression, sampling_method, draws=draws, count=2)
    return T[0], T[1]


def compress_ranges_multi(x_local, j0, dims, m, comm, compression="fri", sampling_method="systematic", draws=None,
                          count=1):
    """`count` consecutive compressions of the same tensor (the k-th uses the k-th draw); see compress_ranges[_pair]."""
    from .dist import use_torch_stream

    dims = tuple(int(d) for d in dims)
    n_total = dims[0] * dims[1] * dims[2] * dims[3]
    m, count = int(m), int(count)
    if not (0 <= m <= n_total):
        raise ValueError(f"need 0 <= m <= N (m={m}, N={n_total})")   # the reference's driver check, friccsd.py:69-72
    if compression not in ("fri", "largest"):
        raise ValueError(f"unknown compression '{compression}'")
    if compression == "fri" and sampling_method not in ("systematic", "parallel"):
        raise NotImplementedError("range-sharded compression implements the systematic sampler and the chunked pivotal "
                                  "sampler ('parallel', Alg. A.2 with one chunk per rank); the single ordered pivotal walk "
                                  "is not range-sharded -- use dist.compress_pair for it")
    if x_local.dtype != torch.float64 or not x_local.is_cuda or not x_local.is_contiguous():
        raise TypeError("x_local must be a contiguous CUDA float64 tensor")
    if draws is not None and len(draws) < count:
        raise _lib.FriccError(_lib.EDRAWS, "one draw buffer per compression is needed")
    use_torch_stream()
    L, dev = _lib.lib(), x_local.device
    shard = _Shard(x_local.view(-1), j0, n_total)
    mm = max(m, 1)
    try:
        consumed, failure = 0, None
        # the `count` index / value lists side by side: ONE all-reduce each
        flat = torch.zeros(count * mm, dtype=torch.int64, device=dev)
        vals = torch.zeros(count * mm, dtype=torch.float64, device=dev)
        if compression == "largest":
            keys, idx = global_candidates(shard, comm, m)
            for k in range(count):
                flat[k * mm: k * mm + m] = idx
            n_exact = m
            check(L.fricc_shard_use_probabilities(shard.h, 1.0, 0, ctypes.c_uint64(2**64 - 1), ctypes.c_uint64(0)))
        else:
            # ---- get_d_largest (fri_utils.cpp:400-437): norm, ordered candidates, greedy test
            check(L.fricc_shard_use_abs(shard.h))
            norm = float(exact_sum(shard, comm).item())
            nD, r = ctypes.c_int64(0), ctypes.c_double(norm)
            cand = min(m, FIRST_CANDIDATES)
            keys = idx = None
            while m > 0:
                keys, idx = global_candidates(shard, comm, cand)
                check(L.fricc_d_greedy(_ptr(keys), cand, m, norm, ctypes.byref(nD), ctypes.byref(r)))
                if nD.value < cand or cand == m:
                    break
                cand = m if cand > m // 16 else cand * 16
            nD, r = nD.value, r.value
            key_cut, idx_cut = 2**64 - 1, 0
            if nD > 0:
                key_cut, idx_cut = (int(v) for v in torch.stack([keys[nD - 1], idx[nD - 1]]).cpu().tolist())
            nsamp = m - nD
            check(L.fricc_shard_use_probabilities(shard.h, r, nsamp, ctypes.c_uint64(key_cut), ctypes.c_uint64(idx_cut)))
            if sampling_method == "parallel":
                consumed, failure = _sample_parallel(shard, comm, nsamp, nD, idx, flat, mm, count, draws)
            else:
                consumed = _sample_systematic(shard, comm, nsamp, nD, idx, flat, mm, count, draws)
            _all_reduce(comm, flat)
            n_exact = nD
        for k in range(count):
            check(L.fricc_shard_values(shard.h, _ptr(flat[k * mm:]), m, n_exact, _ptr(vals[k * mm:])))
        _all_reduce(comm, vals)
        # a kernel error on any rank (some p > 1: fri_utils.cpp:383-395) is raised on every rank
        # (the status code travels, negated, through a max-reduction: peers raise the same kind of error, not a generic one)
        msg = failure
        if compression == "fri" and failure is None:
            try:
                check(L.fricc_shard_check(shard.h))
            except _lib.FriccError as e:
                msg = e
        err = torch.full((1,), 0 if msg is None else -int(msg.code), dtype=torch.int64, device=dev)
        _all_reduce(comm, err, "max")
        code = -int(err.item())
        if code:
            raise msg if msg is not None else _lib.FriccError(code, f"{_PEER_ERRORS.get(code, 'error')} on another rank")
        out = []
        for k in range(count):
            T = SparseTensor4d.from_coo(flat[k * mm: k * mm + m], vals[k * mm: k * mm + m], dims)
            check(L.fricc_st_set_n_exact(T._h, int(n_exact)))
            # consumed_draws: always an int -- the draws THIS rank's samplers read in the last compression (1 for the
            # systematic sampler; budget draws + this chunk's draws for "parallel", whose split is consumed_draws_parts)
            if isinstance(consumed, tuple):
                T.consumed_draws, T.consumed_draws_parts = int(consumed[0]) + int(consumed[1]), consumed
            else:
                T.consumed_draws, T.consumed_draws_parts = int(consumed), None
            out.append(T)
        return out
    finally:
        shard.close()
