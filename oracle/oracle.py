"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY: Python access to the two CPU checkers.

* ``port``  : oracle/fricc_oracle.c, our plain-C restatement of the reference algorithm (always available; built by
              ``make -C oracle port`` / ``__graft_entry__.build()``).
* ``ref``   : the UNMODIFIED reference sources compiled by oracle/Makefile into oracle/_ref/ (available when it was
              built in a container that has /root/reference; the .so travels to the GPU box, the sources do not).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The product (fricc_b200/) never does.
"""
import ctypes
import glob
import importlib.util
import os
import subprocess
from ctypes import POINTER, c_char_p, c_double, c_int, c_int32, c_int64, c_long, c_uint32, c_void_p

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "_build", "libfricc_oracle.so")

ORC_OK, ORC_EINVAL, ORC_EPROB, ORC_EPIVOTAL, ORC_ECOUNT, ORC_EDRAWS = 0, -1, -3, -4, -5, -6


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__(f"oracle status {code}")
        self.code = code


def build_port():
    subprocess.run(["make", "-C", HERE, "port"], check=True, capture_output=True)
    return PORT_SO


def build_ref():
    """Needs /root/reference (this container only)."""
    subprocess.run(["make", "-C", HERE, "ref"], check=True, capture_output=True)


def _p(a):
    return a.ctypes.data_as(c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


# ------------------------------------------------------------------------------------------------------------------
# the C restatement
# ------------------------------------------------------------------------------------------------------------------
class Port:
    def __init__(self):
        if not os.path.exists(PORT_SO):
            build_port()
        L = ctypes.CDLL(PORT_SO)
        L.orc_one_norm.restype = c_double
        L.orc_one_norm.argtypes = [c_void_p, c_int64]
        L.orc_glibc_rand_fill.restype = None
        L.orc_glibc_rand_fill.argtypes = [c_uint32, c_int64, c_int64, c_void_p]
        L.orc_argsort.argtypes = [c_void_p, c_int64, c_int64, c_void_p]
        L.orc_get_d_largest.argtypes = [c_void_p, c_int64, c_int64, c_void_p, POINTER(c_int64), POINTER(c_double)]
        L.orc_make_probability_vector.argtypes = [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_double, c_void_p]
        L.orc_sample_pivotal.argtypes = [c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, POINTER(c_int64)]
        L.orc_sample_systematic.argtypes = [c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, POINTER(c_int64)]
        L.orc_fri_compression.argtypes = [c_void_p, c_int64, c_int64, c_int, c_void_p, c_int64, c_void_p, c_void_p,
                                          POINTER(c_int64), POINTER(c_double), POINTER(c_int64)]
        L.orc_largest.argtypes = [c_void_p, c_int64, c_int64, c_void_p, c_void_p]
        L.orc_parallel_sample.argtypes = [c_int64, c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p,
                                          POINTER(c_int64), c_void_p]
        L.orc_contract.argtypes = [c_char_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p,
                                   c_void_p]
        self.L = L

    @staticmethod
    def _chk(rc):
        if rc != ORC_OK:
            raise OracleError(rc)

    def glibc_rand(self, seed, n, skip=0):
        out = np.empty(int(n), dtype=np.int32)
        self.L.orc_glibc_rand_fill(int(seed) & 0xFFFFFFFF, int(skip), int(n), _p(out))
        return out

    def one_norm(self, x):
        x = _f64(x).ravel()
        return float(self.L.orc_one_norm(_p(x), x.size))

    def argsort(self, x, m):
        x = _f64(x).ravel()
        out = np.empty(int(m), dtype=np.int64)
        self._chk(self.L.orc_argsort(_p(x), x.size, int(m), _p(out)))
        return out

    def get_d_largest(self, x, n_sample):
        x = _f64(x).ravel()
        D = np.empty(max(int(n_sample), 1), dtype=np.int64)
        nD, r = c_int64(0), c_double(0)
        self._chk(self.L.orc_get_d_largest(_p(x), x.size, int(n_sample), _p(D), ctypes.byref(nD), ctypes.byref(r)))
        return D[: nD.value].copy(), r.value

    def make_probability_vector(self, x, n_sample, D, remaining_norm):
        x = _f64(x).ravel()
        D = _i64(D)
        p = np.empty(x.size, dtype=np.float64)
        self._chk(self.L.orc_make_probability_vector(_p(x), x.size, int(n_sample), _p(D), D.size, float(remaining_norm),
                                                     _p(p)))
        return p

    def sample_pivotal(self, n_sample, probs, draws):
        p, d = _f64(probs).ravel(), _i32(draws)
        S = np.zeros(max(int(n_sample), 1), dtype=np.int64)
        used = c_int64(0)
        self._chk(self.L.orc_sample_pivotal(int(n_sample), _p(p), p.size, _p(d), d.size, _p(S), ctypes.byref(used)))
        return S[: int(n_sample)].copy(), used.value

    def sample_systematic(self, n_sample, probs, draws):
        p, d = _f64(probs).ravel(), _i32(draws)
        S = np.zeros(max(int(n_sample), 1), dtype=np.int64)
        used = c_int64(0)
        self._chk(self.L.orc_sample_systematic(int(n_sample), _p(p), p.size, _p(d), d.size, _p(S), ctypes.byref(used)))
        return S[: int(n_sample)].copy(), used.value

    def parallel_sample(self, n_sample, probs, n_chunks, draws):
        """fri_utils.cpp:176-362 with `n_chunks` chunks, draws consumed in chunk order -> (S, consumed)"""
        p, d = _f64(probs).ravel(), _i32(draws)
        S = np.zeros(max(int(n_sample), 1), dtype=np.int64)
        work = np.empty(max(p.size, 1), dtype=np.float64)
        used = c_int64(0)
        self._chk(self.L.orc_parallel_sample(int(n_sample), _p(p), p.size, int(n_chunks), _p(d), d.size, _p(S),
                                             ctypes.byref(used), _p(work)))
        return S[: int(n_sample)].copy(), used.value

    def fri_compression(self, x, n_sample, method, draws):
        """-> dict(idx, vals, nD, remaining_norm, consumed)"""
        x, d = _f64(x).ravel(), _i32(draws)
        m = int(n_sample)
        idx = np.zeros(max(m, 1), dtype=np.int64)
        vals = np.zeros(max(m, 1), dtype=np.float64)
        nD, r, used = c_int64(0), c_double(0), c_int64(0)
        code = {"pivotal": 0, "systematic": 1}[method]
        self._chk(self.L.orc_fri_compression(_p(x), x.size, m, code, _p(d), d.size, _p(idx), _p(vals),
                                             ctypes.byref(nD), ctypes.byref(r), ctypes.byref(used)))
        return dict(idx=idx[:m].copy(), vals=vals[:m].copy(), nD=nD.value, remaining_norm=r.value, consumed=used.value)

    def largest(self, x, m):
        x = _f64(x).ravel()
        idx = np.zeros(max(int(m), 1), dtype=np.int64)
        vals = np.zeros(max(int(m), 1), dtype=np.float64)
        self._chk(self.L.orc_largest(_p(x), x.size, int(m), _p(idx), _p(vals)))
        return idx[: int(m)].copy(), vals[: int(m)].copy()

    def contract(self, term, W, flat, vals, dims, out):
        """out (float64 ndarray, any strides, modified in place) += contraction; W any strides."""
        assert W.dtype == np.float64 and out.dtype == np.float64 and W.ndim == 4 and out.ndim == 4
        flat, vals = _i64(flat), _f64(vals)
        ws = _i64([s // 8 for s in W.strides])
        os_ = _i64([s // 8 for s in out.strides])
        dm = _i64(dims)
        self._chk(self.L.orc_contract(term.encode(), c_void_p(W.ctypes.data), _p(ws), _p(flat), _p(vals), flat.size,
                                      _p(dm), c_void_p(out.ctypes.data), _p(os_)))


_port = None


def port():
    global _port
    if _port is None:
        _port = Port()
    return _port


# ------------------------------------------------------------------------------------------------------------------
# the compiled reference
# ------------------------------------------------------------------------------------------------------------------
class Ref:
    """oracle/_ref/py_rccsd*.so: the reference's own pybind11 module (``.mod``) plus the C-ABI adapter
    (oracle/ref_adapter.cpp) for the span functions and draw injection."""

    def __init__(self, path):
        spec = importlib.util.spec_from_file_location("py_rccsd", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        self.mod = mod
        L = ctypes.CDLL(path)
        L.ref_one_norm.restype = c_double
        L.ref_one_norm.argtypes = [c_void_p, c_long]
        L.ref_set_draws.argtypes = [c_void_p, c_long]
        L.ref_rand_calls.restype = c_long
        L.ref_argsort.argtypes = [c_void_p, c_long, c_long, c_void_p]
        L.ref_get_d_largest.restype = c_long
        L.ref_get_d_largest.argtypes = [c_void_p, c_long, c_long, c_void_p, POINTER(c_double)]
        L.ref_make_probability_vector.argtypes = [c_void_p, c_long, c_long, c_void_p, c_long, c_double, c_void_p]
        L.ref_sample_pivotal.argtypes = [c_long, c_void_p, c_long, c_void_p]
        L.ref_sample_systematic.argtypes = [c_long, c_void_p, c_long, c_void_p]
        L.ref_fri_compression.argtypes = [c_void_p, c_long, c_long, c_char_p, c_int, c_void_p, c_void_p]
        if hasattr(L, "ref_parallel_sample"):
            L.ref_parallel_sample.restype = c_long
            L.ref_parallel_sample.argtypes = [c_long, c_void_p, c_long, c_int, c_void_p]
        self.L = L

    def set_num_threads(self, n):
        """OpenMP threads of the reference from here on (contractions are bit-identical for any count)."""
        if hasattr(self.L, "ref_set_num_threads"):
            self.L.ref_set_num_threads(int(n))

    def reset_rand_calls(self):
        self.L.ref_reset_rand_calls()

    def set_draws(self, draws):
        d = _i32(draws)
        self.L.ref_set_draws(_p(d), d.size)

    def clear_draws(self):
        self.L.ref_clear_draws()

    def rand_calls(self):
        return int(self.L.ref_rand_calls())

    def one_norm(self, x):
        x = _f64(x).ravel()
        return float(self.L.ref_one_norm(_p(x), x.size))

    def argsort(self, x, m):
        x = _f64(x).ravel()
        out = np.empty(int(m), dtype=np.int64)
        self.L.ref_argsort(_p(x), x.size, int(m), _p(out))
        return out

    def get_d_largest(self, x, n_sample):
        x = _f64(x).ravel()
        D = np.empty(max(int(n_sample), 1), dtype=np.int64)
        r = c_double(0)
        nD = self.L.ref_get_d_largest(_p(x), x.size, int(n_sample), _p(D), ctypes.byref(r))
        return D[:nD].copy(), r.value

    def make_probability_vector(self, x, n_sample, D, remaining_norm):
        x, D = _f64(x).ravel(), _i64(D)
        p = np.empty(x.size, dtype=np.float64)
        self.L.ref_make_probability_vector(_p(x), x.size, int(n_sample), _p(D), D.size, float(remaining_norm), _p(p))
        return p

    def sample_pivotal(self, n_sample, probs, draws):
        p = _f64(probs).ravel()
        self.set_draws(draws)
        S = np.zeros(max(int(n_sample), 1), dtype=np.int64)
        self.L.ref_sample_pivotal(int(n_sample), _p(p), p.size, _p(S))
        used = self.rand_calls()
        self.clear_draws()
        return S[: int(n_sample)].copy(), used

    def sample_systematic(self, n_sample, probs, draws):
        p = _f64(probs).ravel()
        self.set_draws(draws)
        S = np.zeros(max(int(n_sample), 1), dtype=np.int64)
        self.L.ref_sample_systematic(int(n_sample), _p(p), p.size, _p(S))
        used = self.rand_calls()
        self.clear_draws()
        return S[: int(n_sample)].copy(), used

    def parallel_sample(self, n_sample, probs, n_threads, draws):
        """The reference's parallel_sample with omp_set_num_threads(n_threads).  Reproducible only for n_threads == 1
        or a constant draw stream (its threads call rand() concurrently).  -> (S, rand() calls)"""
        p = _f64(probs).ravel()
        self.set_draws(draws)
        S = np.zeros(max(int(n_sample), 1), dtype=np.int64)
        n = self.L.ref_parallel_sample(int(n_sample), _p(p), p.size, int(n_threads), _p(S))
        used = self.rand_calls()
        self.clear_draws()
        return S[: int(n)].copy(), used

    def fri_compression(self, x, n_sample, method, draws):
        x = _f64(x).ravel()
        m = int(n_sample)
        self.set_draws(draws)
        idx = np.zeros(max(m, 1), dtype=np.int64)
        vals = np.zeros(max(m, 1), dtype=np.float64)
        self.L.ref_fri_compression(_p(x), x.size, m, method.encode(), 0, _p(idx), _p(vals))
        used = self.rand_calls()
        self.clear_draws()
        return dict(idx=idx[:m].copy(), vals=vals[:m].copy(), consumed=used)


_ref = None


def ref_path():
    c = sorted(glob.glob(os.path.join(HERE, "_ref", "py_rccsd*.so")))
    return c[0] if c else None


def ref():
    """The compiled reference, or None when oracle/_ref has not been built."""
    global _ref
    if _ref is None:
        path = ref_path()
        if path is None:
            return None
        _ref = Ref(path)
    return _ref
