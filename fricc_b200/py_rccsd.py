"""Drop-in mirror of the reference's pybind11 module ``fricc/py_rccsd`` on top of libfricc_b200.so.

Reference surface: /root/reference/src/py_rccsd.cpp:16-55.  Same names, same positional/keyword arguments (the output
kwarg of ``contract_DTSpT`` really is spelled ``ouput`` there), same argument checks (``noconvert``: arrays must already
be float64; the dense tensor handed to ``SparseTensor4d`` must be C-contiguous; W and the output may have any
strides), results accumulated in place, ``None`` returned.  Differences, all deliberate:

* every array argument may also be a CUDA tensor (anything with ``__cuda_array_interface__``/``data_ptr``, e.g. a
  ``torch.float64`` tensor on the GPU): it is then used in place, without host<->device copies;
* algorithmic failures raise (``ValueError`` / ``fricc_b200._lib.FriccError``) instead of ``exit(EXIT_FAILURE)``;
* the five utilities that the reference binds with ``std::span`` arguments and therefore cannot be called from Python
  at HEAD (``argsort``, ``get_d_largest``, ``make_probability_vector``, ``fri_compression``, ``parallel_sample``) work
  here with ndarray arguments;
* optional keyword ``draws=`` (int32 values in [0, 2**31-1]) injects the uniform draws; without it the module keeps a
  glibc-compatible ``rand()`` stream (``srand(seed)`` below), i.e. the same sequence a stock reference build sees.

There is no CPU implementation behind this module: without the built library and a CUDA device every call raises.
"""
import ctypes
from ctypes import byref, c_double, c_int64, c_void_p

import os

import numpy as np

from . import _lib
from ._lib import FRICC_DEVICE, FRICC_HOST, check, i64x4

__all__ = [
    "SparseTensor4d", "contract_DTSpT", "contract_DTSpT_experimental", "argsort", "sample_pivotal",
    "sample_systematic", "parallel_sample", "make_probability_vector", "get_d_largest", "fri_compression", "srand",
    "rand_consumed",
]


# ----------------------------------------------------------------------------------------------------------------
# argument plumbing
# ----------------------------------------------------------------------------------------------------------------
def _is_cuda_tensor(a):
    return hasattr(a, "data_ptr") and getattr(a, "is_cuda", False)


def _is_cpu_tensor(a):
    return hasattr(a, "data_ptr") and hasattr(a, "numpy") and not getattr(a, "is_cuda", False)


class _Arr:
    """A float64 array argument resolved to (pointer, memory space, shape, element strides)."""

    __slots__ = ("ptr", "mem", "shape", "strides", "keep", "size")

    def __init__(self, a, name, ndim=None, c_contiguous=False, writeable=False):
        if _is_cuda_tensor(a):
            if "float64" not in str(a.dtype):
                raise TypeError(f"{name}: expected a float64 array (noconvert), got {a.dtype}")
            if c_contiguous and not a.is_contiguous():
                raise TypeError(f"{name}: expected a C-contiguous array")
            self.ptr, self.mem = int(a.data_ptr()), FRICC_DEVICE
            self.shape, self.strides = tuple(a.shape), tuple(a.stride())
            self.keep = a
        else:
            if _is_cpu_tensor(a):
                a = a.numpy()
            if not isinstance(a, np.ndarray):
                raise TypeError(f"{name}: expected a numpy.ndarray[float64] (noconvert), got {type(a).__name__}")
            if a.dtype != np.float64:
                raise TypeError(f"{name}: expected a float64 array (noconvert), got {a.dtype}")
            if c_contiguous and not a.flags.c_contiguous:
                raise TypeError(f"{name}: expected a C-contiguous array (array_t<double, c_style>, noconvert)")
            if writeable and not a.flags.writeable:
                raise ValueError(f"{name}: array is not writeable")
            if any(s % a.itemsize for s in a.strides):
                raise TypeError(f"{name}: strides must be multiples of the item size")
            self.ptr, self.mem = a.ctypes.data, FRICC_HOST
            self.shape, self.strides = tuple(a.shape), tuple(s // a.itemsize for s in a.strides)
            self.keep = a
        if ndim is not None and len(self.shape) != ndim:
            raise ValueError(f"{name}: array has incorrect number of dimensions: {len(self.shape)}; expected {ndim}")
        self.size = int(np.prod(self.shape)) if len(self.shape) else 1


def _strided4(a, name, writeable=False):
    """W / output argument of a contraction: 4-D float64, any strides (negative strides are normalised by a copy
    for read-only operands and rejected for the in-place output)."""
    arr = _Arr(a, name, ndim=4, writeable=writeable)
    if any(s < 0 for s in arr.strides):
        if writeable:
            raise ValueError(f"{name}: negative strides are not supported for the in-place output")
        if arr.mem == FRICC_HOST:
            arr = _Arr(np.ascontiguousarray(arr.keep), name, ndim=4)
        else:
            arr = _Arr(arr.keep.contiguous(), name, ndim=4)
    if arr.mem == FRICC_HOST and not writeable:
        # a host view that spans much more memory than it addresses would be staged wastefully: compact it first
        span = 1 + sum((n - 1) * s for n, s in zip(arr.shape, arr.strides)) if arr.size else 0
        if span > 2 * arr.size:
            arr = _Arr(np.ascontiguousarray(arr.keep), name, ndim=4)
    return arr


def _draws_arg(draws):
    if draws is None:
        return None, 0, None
    d = np.ascontiguousarray(draws, dtype=np.int32)
    return d.ctypes.data, int(d.size), d


def _flat_f64(x, name):
    arr = _Arr(x, name)
    if arr.mem == FRICC_HOST:
        if not arr.keep.flags.c_contiguous:
            raise TypeError(f"{name}: expected a contiguous array")
    elif not arr.keep.is_contiguous():
        raise TypeError(f"{name}: expected a contiguous array")
    return arr


def srand(seed):
    """Seed the module's rand() stream exactly like libc ``srand(seed)`` does for the reference."""
    check(_lib.lib().fricc_srand(int(seed) & 0xFFFFFFFF))


_options = {}            # what set_option was last given, per key (the library holds the authoritative values)
_warned_pivotal = False


def _warn_pivotal_fast_mode(T):
    """Once per process: a tensor sampled with the ORDERED pivotal walk is being contracted by the fast kernels.  Those
    are within 1e-10 of the reference per call, but they add an owner pair's entries in another order, and the last
    pivotal window (fri_utils.cpp:81-109) turns such last-bit differences of t2 into a different sample now and then: CC
    energies then agree with the reference's at equal seed only statistically, not to 1e-8.
    set_option("spmm_exact_order", 1) restores seed parity (bit-identical contractions, 3-5x slower)."""
    global _warned_pivotal
    if _warned_pivotal or getattr(T, "sampling_method", None) != "pivotal" or _options.get("spmm_exact_order", 0):
        return
    _warned_pivotal = True
    import warnings

    warnings.warn('fricc_b200: SparseTensor4d sampled with "pivotal" contracted by the fast kernels -- equal-seed energies '
                  'match the reference only statistically; set_option("spmm_exact_order", 1) for seed parity '
                  "(see DESIGN.md section 2)", UserWarning, stacklevel=3)


def set_option(key, value):
    """Library switches (include/fricc_b200.h: fricc_set_option), e.g. ``set_option("escan", 0)``."""
    check(_lib.lib().fricc_set_option(str(key).encode(), int(value)))
    _options[str(key)] = int(value)


def get_stat(key):
    """Counters of the last library call (include/fricc_b200.h: fricc_get_stat), e.g. ``get_stat("pivotal_restarts")``."""
    v = c_int64(0)
    check(_lib.lib().fricc_get_stat(str(key).encode(), byref(v)))
    return v.value


def rand_consumed():
    return int(_lib.lib().fricc_rand_consumed())


# ----------------------------------------------------------------------------------------------------------------
# SparseTensor4d   (reference: include/sparse_tensor.hpp:13-100, bound at src/py_rccsd.cpp:32-47)
# ----------------------------------------------------------------------------------------------------------------
class SparseTensor4d:
    """COO sparse 4-tensor living on the GPU.

    ``SparseTensor4d(dims, nnz)``                         -- nnz zero entries (py_rccsd.cpp:33)
    ``SparseTensor4d(tensor_arr, dims, m, compression="fri", sampling_method="pivotal", verbose=False)``
                                                          -- compress a dense tensor (py_rccsd.cpp:34-39)
    """

    def __init__(self, *args, **kwargs):
        self._h = c_void_p(None)
        self._L = _lib.lib()
        draws = kwargs.pop("draws", None)
        names = ("tensor_arr", "dims", "m", "compression", "sampling_method", "verbose")
        if len(args) == 2 and not kwargs and not hasattr(args[0], "dtype"):
            dims, nnz = args
            dims = self._dims(dims)
            check(self._L.fricc_st_create_empty(i64x4(dims), int(nnz), byref(self._h)))
            self.consumed_draws = 0
            return
        params = dict(compression="fri", sampling_method="pivotal", verbose=False)
        if len(args) > len(names):
            raise TypeError("__init__(): incompatible constructor arguments")
        for k, v in zip(names, args):
            params[k] = v
        for k, v in kwargs.items():
            if k not in names or k in names[: len(args)]:
                raise TypeError(f"__init__(): incompatible constructor arguments (unexpected '{k}')")
            params[k] = v
        for k in ("tensor_arr", "dims", "m"):
            if k not in params:
                raise TypeError(f"__init__(): incompatible constructor arguments (missing '{k}')")
        if not isinstance(params["compression"], str) or not isinstance(params["sampling_method"], str):
            raise TypeError("__init__(): compression and sampling_method must be str")
        t = _Arr(params["tensor_arr"], "tensor_arr", c_contiguous=True)
        dims = self._dims(params["dims"])
        if int(np.prod(dims)) != t.size:
            raise ValueError(f"dims {dims} do not match tensor_arr of {t.size} elements")
        m = params["m"]
        if isinstance(m, (float, np.floating)) and not float(m).is_integer():
            raise TypeError("__init__(): m must be an integer")
        dptr, dn, keep = _draws_arg(draws)
        consumed = c_int64(0)
        if params["compression"] == "fri":
            self.sampling_method = params["sampling_method"]      # remembered for _warn_pivotal_fast_mode
        if dptr is None:
            before = rand_consumed()
            check(self._L.fricc_st_create(t.ptr, t.mem, i64x4(dims), int(m), params["compression"].encode(),
                                          params["sampling_method"].encode(), int(bool(params["verbose"])),
                                          byref(self._h)))
            self.consumed_draws = rand_consumed() - before
        else:
            check(self._L.fricc_st_create_with_draws(t.ptr, t.mem, i64x4(dims), int(m),
                                                     params["compression"].encode(),
                                                     params["sampling_method"].encode(), int(bool(params["verbose"])),
                                                     dptr, dn, byref(consumed), byref(self._h)))
            self.consumed_draws = consumed.value

    @staticmethod
    def _dims(dims):
        dims = [int(d) for d in dims]
        if len(dims) != 4:
            raise TypeError("dims must have exactly 4 entries (std::array<size_t, 4>)")
        return dims

    @classmethod
    def from_coo(cls, flat_idx, vals, dims):
        """Build from flat row-major indices and values (host ndarrays or CUDA tensors) -- multi-GPU gather path."""
        self = cls.__new__(cls)
        self._h = c_void_p(None)
        self._L = _lib.lib()
        self.consumed_draws = 0
        dims = cls._dims(dims)
        if _is_cuda_tensor(flat_idx):
            m = int(flat_idx.numel())
            check(self._L.fricc_st_create_from_coo(int(flat_idx.data_ptr()), int(vals.data_ptr()), FRICC_DEVICE,
                                                   i64x4(dims), m, byref(self._h)))
        else:
            fi = np.ascontiguousarray(flat_idx, dtype=np.int64)
            va = np.ascontiguousarray(vals, dtype=np.float64)
            check(self._L.fricc_st_create_from_coo(fi.ctypes.data, va.ctypes.data, FRICC_HOST, i64x4(dims),
                                                   int(fi.size), byref(self._h)))
        return self

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                self._L.fricc_st_destroy(h)
            except Exception:
                pass
            self._h = None  # (at interpreter shutdown module globals such as c_void_p may already be gone)

    # -- accessors -------------------------------------------------------------------------------------------
    def size(self):
        n = c_int64(0)
        check(self._L.fricc_st_size(self._h, byref(n)))
        return n.value

    def dimensions(self):
        d = (c_int64 * 4)()
        check(self._L.fricc_st_dims(self._h, d))
        return [int(v) for v in d]

    def get_element(self, mi):
        """-> ([i, j, a, b], value)   (py_rccsd.cpp:40-46)"""
        idx = (c_int64 * 4)()
        val = c_double(0.0)
        check(self._L.fricc_st_get_element(self._h, int(mi), idx, byref(val)))
        return [int(v) for v in idx], val.value

    def set_element(self, mi, flat_idx, value):
        check(self._L.fricc_st_set_element(self._h, int(mi), int(flat_idx), float(value)))

    def export(self):
        """All entries in storage order: (flat row-major indices int64[m], values float64[m]) as host arrays."""
        m = self.size()
        idx = np.empty(m, dtype=np.int64)
        val = np.empty(m, dtype=np.float64)
        check(self._L.fricc_st_export(self._h, idx.ctypes.data, val.ctypes.data, FRICC_HOST))
        return idx, val

    def print(self):
        """Same dump as operator<< (sparse_tensor.hpp:90-99)."""
        idx, val = self.export()
        d = self.dimensions()
        print("i   j   k   l        val")
        for f, v in zip(idx.tolist(), val.tolist()):
            i0 = f // (d[1] * d[2] * d[3])
            i1 = f % (d[1] * d[2] * d[3]) // (d[2] * d[3])
            i2 = f % (d[2] * d[3]) // d[3]
            i3 = f % d[3]
            print(f"{i0}   {i1}   {i2}   {i3}        {v:g}")
        print()


# ----------------------------------------------------------------------------------------------------------------
# contractions   (reference: contract_SparseTensor4d_wrapper, src/sparse_tensor.cpp:349-382)
# ----------------------------------------------------------------------------------------------------------------
def _contract(W, T, ouput, term):
    if not isinstance(T, SparseTensor4d):
        raise TypeError("contract_DTSpT(): T must be a SparseTensor4d")
    if not isinstance(term, str):
        raise TypeError("contract_DTSpT(): term must be a str")
    _warn_pivotal_fast_mode(T)
    w = _strided4(W, "W")
    o = _strided4(ouput, "ouput", writeable=True)
    check(_lib.lib().fricc_contract(w.ptr, w.mem, i64x4(w.shape), i64x4(w.strides), T._h, o.ptr, o.mem,
                                    i64x4(o.shape), i64x4(o.strides), term.encode()))
    return None


def contract_DTSpT(W, T, ouput, term):
    """``ouput += scale * contraction(W, T)`` for term in 0101 2323 1302 1202 1303 1203 1323 0112 0313 0113 0312."""
    return _contract(W, T, ouput, term)


def contract_DTSpT_pair(W_a, W_b, T, ouput, term_a, term_b):
    """``contract_DTSpT(W_a, T, ouput, term_a); contract_DTSpT(W_b, T, ouput, term_b)`` -- as ONE dense product when the two
    terms allow it (include/fricc_b200.h: fricc_contract_pair): the pairs the reference's drivers issue back to back into
    one array, 0112 + 0113 (same operand, fricc/friccsd.py:347-356), 0113 + 0313, 1302 + 1202 (:252-258).  An extension of
    the reference's surface.  Device tensors take the fused path; anything else is the two plain calls."""
    if not isinstance(T, SparseTensor4d):
        raise TypeError("contract_DTSpT_pair(): T must be a SparseTensor4d")
    if not (_is_cuda_tensor(W_a) and _is_cuda_tensor(W_b) and _is_cuda_tensor(ouput)):
        _contract(W_a, T, ouput, term_a)
        return _contract(W_b, T, ouput, term_b)
    a, b = _strided4(W_a, "W_a"), _strided4(W_b, "W_b")
    o = _strided4(ouput, "ouput", writeable=True)
    check(_lib.lib().fricc_contract_pair(a.ptr, i64x4(a.shape), i64x4(a.strides), b.ptr, i64x4(b.shape), i64x4(b.strides),
                                         T._h, o.ptr, i64x4(o.shape), i64x4(o.strides), str(term_a).encode(),
                                         str(term_b).encode()))
    return None


def constant_operand(W, on=True):
    """Declare the CUDA tensor `W` (its whole storage span) constant between plugin calls, or withdraw that (`on=False`).
    The drivers hand ``eris.ovov`` -- integrals that never change -- to five of the eleven terms in every iteration
    (fricc/friccsd.py:347-378); the digit planes of a declared operand's views are built on first use and kept
    (include/fricc_b200.h: fricc_constant_operand).  Call it again after rewriting the array.  An extension of the
    reference's surface; results are bit-identical with and without it."""
    if not _is_cuda_tensor(W):
        raise TypeError("constant_operand(): a CUDA tensor is needed (host arrays are staged per call)")
    span = 1 + sum((int(n) - 1) * int(st) for n, st in zip(W.shape, W.stride())) if W.numel() else 0
    if span <= 0:
        return None
    check(_lib.lib().fricc_constant_operand(c_void_p(int(W.data_ptr())), int(span), 1 if on else 0))
    return None


def prepare_operand(W, ouput, term):
    """Encode the digit planes of the CUDA tensor `W` for the coming ``contract_DTSpT(W, T, ouput, term)`` now, on the
    library's current stream (include/fricc_b200.h: fricc_prepare_operand) -- e.g. on a second stream while T2 is still
    on its way to the GPU: Wvvvv does not depend on it (fricc/friccsd.py:207).  `ouput` is only looked at for its
    strides (which of the two free axes is the minor one decides the orientation of the planes).  The next call that
    takes the tensor-core form on the same operand uses these planes, once; order the streams yourself and do not write
    `W` in between.  An extension of the reference's surface; results are bit-identical with and without it."""
    if not (_is_cuda_tensor(W) and _is_cuda_tensor(ouput)):
        raise TypeError("prepare_operand(): CUDA tensors are needed (host arrays are staged per call)")
    w = _strided4(W, "W")
    o = _strided4(ouput, "ouput", writeable=True)
    check(_lib.lib().fricc_prepare_operand(w.ptr, i64x4(w.shape), i64x4(w.strides), i64x4(o.strides), str(term).encode()))
    return None


def contract_DTSpT_experimental(W, T, ouput, term):
    """The reference's alternative 2323 kernels (``"taskloop"``, ``"atomic"``; src/sparse_tensor.cpp:445-459) differ only
    in CPU threading; both names run the same GPU ladder kernel here."""
    if term not in ("taskloop", "atomic"):
        raise ValueError("CASE NOT IMPLEMENTED")
    return _contract(W, T, ouput, term)


# ----------------------------------------------------------------------------------------------------------------
# FRI utilities   (reference: include/fri_utils.hpp:8-32, bound at src/py_rccsd.cpp:17-30)
# ----------------------------------------------------------------------------------------------------------------
def argsort(array, m):
    """Indices of the m largest-magnitude entries, |x| descending (ties: index ascending)  (fri_utils.cpp:11-36)."""
    x = _flat_f64(array, "array")
    out = np.empty(int(m), dtype=np.int64)
    check(_lib.lib().fricc_argsort(x.ptr, x.mem, x.size, int(m), out.ctypes.data))
    return out


def one_norm(x):
    xa = _flat_f64(x, "x")
    r = c_double(0.0)
    check(_lib.lib().fricc_one_norm(xa.ptr, xa.mem, xa.size, byref(r)))
    return r.value


def one_norm_static(x, out=None):
    """one_norm for callers that must not allocate device memory while they run (an extension; the reference has no
    counterpart).  A driver whose own arena fills the GPU, a thread that cannot afford to block in the kernel driver,
    or a process under a fixed device-memory budget uses this instead of one_norm: the working memory of the exact
    scan is a fixed 1.3 MB array inside the library, not a pooled buffer.  x: a CUDA float64 array of at most
    FRICC_ONE_NORM_STATIC_MAX_N = 2**26 elements (more -> ValueError).  out: None -> the sum is returned as a float (the
    stream is synchronised); a 1-element CUDA float64 tensor -> written there asynchronously, returned as is.
    All calls on a GPU share the one array: keep them on one stream (overlapping calls race), and make the first call on
    each device during set-up, before any stream capture.  Bit-identical to one_norm."""
    xa = _Arr(x, "x", c_contiguous=True)
    if xa.mem != FRICC_DEVICE:
        raise TypeError("one_norm_static: x must be a CUDA array (staging a host array would allocate)")
    if out is None:
        r = c_double(0.0)
        check(_lib.lib().fricc_one_norm_static(xa.ptr, xa.size, ctypes.addressof(r), FRICC_HOST))
        return r.value
    oa = _Arr(out, "out", c_contiguous=True)
    if oa.mem != FRICC_DEVICE or oa.size != 1:
        raise TypeError("one_norm_static: out must be a 1-element CUDA float64 tensor")
    check(_lib.lib().fricc_one_norm_static(xa.ptr, xa.size, oa.ptr, FRICC_DEVICE))
    return out


def get_d_largest(x, n_sample):
    """-> (D, remaining_norm)   (fri_utils.cpp:400-437).  Positional order is (x, n_sample) as in the C++ signature."""
    xa = _flat_f64(x, "x")
    D = np.empty(int(n_sample), dtype=np.int64)
    nD = c_int64(0)
    r = c_double(0.0)
    check(_lib.lib().fricc_get_d_largest(xa.ptr, xa.mem, xa.size, int(n_sample), D.ctypes.data, byref(nD), byref(r)))
    return D[: nD.value].copy(), r.value


def make_probability_vector(x, n_sample, D, remaining_norm):
    """p_i = |x_i| / remaining_norm * n_sample, p[D] = 0; raises if any p_i > 1   (fri_utils.cpp:367-398)."""
    xa = _flat_f64(x, "x")
    Dh = np.ascontiguousarray(D, dtype=np.int64)
    p = np.empty(xa.size, dtype=np.float64)
    check(_lib.lib().fricc_make_probability_vector(xa.ptr, xa.mem, xa.size, int(n_sample), Dh.ctypes.data,
                                                   int(Dh.size), float(remaining_norm), p.ctypes.data, FRICC_HOST))
    return p


def _probs(probs):
    if isinstance(probs, (list, tuple)):
        probs = np.asarray(probs, dtype=np.float64)
    return _flat_f64(probs, "probs")


def sample_pivotal(n_sample, probs, draws=None):
    """Ordered pivotal sampling (fri_utils.cpp:63-127) -> list[int]."""
    p = _probs(probs)
    S = np.empty(int(n_sample), dtype=np.int64)
    dptr, dn, keep = _draws_arg(draws)
    consumed = c_int64(0)
    check(_lib.lib().fricc_sample_pivotal(int(n_sample), p.ptr, p.mem, p.size, dptr, dn, byref(consumed),
                                          S.ctypes.data))
    return S.tolist()


def sample_systematic(n_sample, probs, draws=None):
    """Systematic sampling (fri_utils.cpp:129-159) -> list[int]."""
    p = _probs(probs)
    S = np.empty(int(n_sample), dtype=np.int64)
    dptr, dn, keep = _draws_arg(draws)
    consumed = c_int64(0)
    check(_lib.lib().fricc_sample_systematic(int(n_sample), p.ptr, p.mem, p.size, dptr, dn, byref(consumed),
                                             S.ctypes.data))
    return S.tolist()


def parallel_sample(n_sample, probs, n_chunks=None, draws=None):
    """Thread-partitioned pivotal sampling, "Alg. A.2" (fri_utils.cpp:176-362) -> list[int].

    The reference cuts `probs` into omp_get_max_threads() chunks; `n_chunks` is that number here (default: the value
    omp_get_max_threads() would have in this process -- OMP_NUM_THREADS if set, else the CPU count).  The reference's
    threads call rand() concurrently, so with more than one chunk the order of its draws is a race; this
    implementation consumes them in chunk order (chunk j after chunk j-1), which is what the reference does when its
    threads happen to run one after another and exactly what it does with one thread."""
    if n_chunks is None:
        env = os.environ.get("OMP_NUM_THREADS", "")
        n_chunks = int(env.split(",")[0]) if env.split(",")[0].strip().isdigit() else (os.cpu_count() or 1)
    p = _probs(probs)
    S = np.empty(int(n_sample), dtype=np.int64)
    dptr, dn, keep = _draws_arg(draws)
    consumed = c_int64(0)
    check(_lib.lib().fricc_parallel_sample(int(n_sample), p.ptr, p.mem, p.size, int(n_chunks), dptr, dn,
                                           byref(consumed), S.ctypes.data))
    return S.tolist()


def fri_compression(x, n_sample, sampling_method="pivotal", verbose=False, draws=None):
    """-> (compressed_idx int64[n_sample], compressed_vals float64[n_sample]), [D..., S...]   (fri_utils.cpp:439-522)."""
    xa = _flat_f64(x, "x")
    idx = np.empty(int(n_sample), dtype=np.int64)
    val = np.empty(int(n_sample), dtype=np.float64)
    dptr, dn, keep = _draws_arg(draws)
    consumed = c_int64(0)
    check(_lib.lib().fricc_fri_compression(xa.ptr, xa.mem, xa.size, int(n_sample), str(sampling_method).encode(),
                                           int(bool(verbose)), dptr, dn, byref(consumed), idx.ctypes.data,
                                           val.ctypes.data))
    return idx, val
