"""ctypes binding of libfricc_b200.so (the C-ABI declared in include/fricc_b200.h)."""
import ctypes
import os
import subprocess
from ctypes import POINTER, c_char_p, c_double, c_int, c_int32, c_int64, c_uint32, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfricc_b200.so")
CSRC = os.path.join(_HERE, "csrc")

FRICC_HOST = 0
FRICC_DEVICE = 1

OK = 0
EINVAL, ECUDA, EPROB, EPIVOTAL, ECOUNT, EDRAWS, ERANGE = -1, -2, -3, -4, -5, -6, -7


class FriccError(RuntimeError):
    """An algorithmic error the reference reports with std::cerr + exit(EXIT_FAILURE)."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def build(verbose=False):
    """Compile every CUDA source for sm_100a into fricc_b200/libfricc_b200.so (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j8"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise RuntimeError("building libfricc_b200.so failed")
    return LIB_PATH


_lib = None

_i64p = POINTER(c_int64)
_f64p = POINTER(c_double)

# name -> (restype, argtypes); every symbol include/fricc_b200.h declares
SIGNATURES = {
    "fricc_last_error": (c_char_p, []),
    "fricc_version": (c_char_p, []),
    "fricc_device_count": (c_int, [POINTER(c_int)]),
    "fricc_set_device": (c_int, [c_int]),
    "fricc_set_stream": (c_int, [c_void_p]),
    "fricc_stream_synchronize": (c_int, []),
    "fricc_set_option": (c_int, [c_char_p, c_int]),
    "fricc_kernel_launches": (c_int64, []),
    "fricc_plane_cache_clear": (c_int, []),
    "fricc_get_stat": (c_int, [c_char_p, _i64p]),
    "fricc_srand": (c_int, [c_uint32]),
    "fricc_rand_fill": (c_int, [POINTER(c_int32), c_int64]),
    "fricc_rand_consumed": (c_int64, []),
    "fricc_rand_skip": (c_int, [c_int64]),
    "fricc_rand_fill_device": (c_int, [c_void_p, c_int64]),
    "fricc_st_create_empty": (c_int, [_i64p, c_int64, POINTER(c_void_p)]),
    "fricc_st_create": (c_int, [c_void_p, c_int, _i64p, c_int64, c_char_p, c_char_p, c_int, POINTER(c_void_p)]),
    "fricc_st_create_with_draws": (c_int, [c_void_p, c_int, _i64p, c_int64, c_char_p, c_char_p, c_int, c_void_p,
                                           c_int64, _i64p, POINTER(c_void_p)]),
    "fricc_st_create_from_coo": (c_int, [c_void_p, c_void_p, c_int, _i64p, c_int64, POINTER(c_void_p)]),
    "fricc_st_get_n_exact": (c_int, [c_void_p, _i64p]),
    "fricc_st_set_n_exact": (c_int, [c_void_p, c_int64]),
    "fricc_rand_state_size": (c_int, []),
    "fricc_rand_get_state": (c_int, [c_void_p]),
    "fricc_rand_set_state": (c_int, [c_void_p]),
    "fricc_st_destroy": (c_int, [c_void_p]),
    "fricc_st_size": (c_int, [c_void_p, _i64p]),
    "fricc_st_dims": (c_int, [c_void_p, _i64p]),
    "fricc_st_get_element": (c_int, [c_void_p, c_int64, _i64p, _f64p]),
    "fricc_st_set_element": (c_int, [c_void_p, c_int64, c_int64, c_double]),
    "fricc_st_export": (c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    "fricc_contract": (c_int, [c_void_p, c_int, _i64p, _i64p, c_void_p, c_void_p, c_int, _i64p, _i64p, c_char_p]),
    "fricc_contract_panel": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_char_p]),
    "fricc_contract_panel_into": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, _i64p, _i64p,
                                          c_char_p]),
    "fricc_accumulate_panels": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_char_p, c_void_p, _i64p, _i64p]),
    "fricc_wslices_create": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, POINTER(c_void_p)]),
    "fricc_wslices_create_view": (c_int, [c_void_p, _i64p, _i64p, c_char_p, c_int64, c_int64, c_int,
                                          POINTER(c_void_p)]),
    "fricc_contract_planes": (c_int, [c_void_p, c_void_p, _i64p, _i64p, c_int64, c_void_p, c_void_p, _i64p, _i64p,
                                      c_char_p]),
    "fricc_wslices_update_view": (c_int, [c_void_p, c_void_p, _i64p, _i64p, c_char_p, c_int64]),
    "fricc_wslices_redone": (c_int, [c_void_p, POINTER(c_int64)]),
    "fricc_contract_pair": (c_int, [c_void_p, _i64p, _i64p, c_void_p, _i64p, _i64p, c_void_p, c_void_p, _i64p, _i64p, c_char_p,
                                    c_char_p]),
    "fricc_int8_peak": (c_int, [c_double, _f64p]),
    "fricc_wslices_update": (c_int, [c_void_p, c_void_p, c_int64]),
    "fricc_wslices_destroy": (c_int, [c_void_p]),
    "fricc_wslices_bytes": (c_int, [c_void_p, _i64p]),
    "fricc_wslices_reconstruct": (c_int, [c_void_p, c_void_p]),
    "fricc_st_sign_fraction": (c_int, [c_void_p, c_char_p, _f64p]),
    "fricc_contract_panel_tc": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_char_p,
                                        c_int]),
    "fricc_contract_panel_into_tc": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, _i64p, _i64p,
                                             c_char_p]),
    "fricc_symm_alloc": (c_int, [c_int64, POINTER(c_void_p), c_void_p]),
    "fricc_symm_open": (c_int, [c_void_p, POINTER(c_void_p)]),
    "fricc_symm_close": (c_int, [c_void_p]),
    "fricc_symm_free": (c_int, [c_void_p]),
    "fricc_push_columns": (c_int, [c_void_p, POINTER(c_void_p), c_int, c_int, c_char_p, _i64p, _i64p, _i64p, c_int64,
                                   c_int64]),
    "fricc_push_view": (c_int, [c_void_p, POINTER(c_void_p), c_int, c_int, _i64p, _i64p]),
    "fricc_peer_barrier": (c_int, [POINTER(c_void_p), c_int, c_int, c_uint64]),
    "fricc_fill_uniform": (c_int, [c_void_p, c_int64, c_uint64, c_uint64]),
    "fricc_scatter_panel": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_char_p, _i64p, c_void_p, POINTER(c_void_p),
                                    c_int, c_int, _i64p, _i64p]),
    "fricc_permute_w": (c_int, [c_void_p, c_int, _i64p, _i64p, c_char_p, c_int64, c_int64, c_void_p]),
    "fricc_one_norm": (c_int, [c_void_p, c_int, c_int64, _f64p]),
    "fricc_one_norm_static": (c_int, [c_void_p, c_int64, c_void_p, c_int]),
    "fricc_argsort": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p]),
    "fricc_get_d_largest": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, _i64p, _f64p]),
    "fricc_make_probability_vector": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_int64, c_double,
                                              c_void_p, c_int]),
    "fricc_sample_pivotal": (c_int, [c_int64, c_void_p, c_int, c_int64, c_void_p, c_int64, _i64p, c_void_p]),
    "fricc_sample_systematic": (c_int, [c_int64, c_void_p, c_int, c_int64, c_void_p, c_int64, _i64p, c_void_p]),
    "fricc_parallel_sample": (c_int, [c_int64, c_void_p, c_int, c_int64, c_int, c_void_p, c_int64, _i64p, c_void_p]),
    "fricc_fri_compression": (c_int, [c_void_p, c_int, c_int64, c_int64, c_char_p, c_int, c_void_p, c_int64, _i64p,
                                      c_void_p, c_void_p]),
    "fricc_key_histogram": (c_int, [c_void_p, c_int64, c_int, c_uint64, c_int, c_void_p]),
    "fricc_shard_create": (c_int, [c_void_p, c_int64, c_int64, c_int64, POINTER(c_void_p)]),
    "fricc_shard_destroy": (c_int, [c_void_p]),
    "fricc_shard_use_abs": (c_int, [c_void_p]),
    "fricc_shard_use_probabilities": (c_int, [c_void_p, c_double, c_int64, c_uint64, c_uint64]),
    "fricc_shard_scan_local": (c_int, [c_void_p, c_void_p]),
    "fricc_shard_scan_maps": (c_int, [c_void_p, c_void_p]),
    "fricc_shard_scan_resolve": (c_int, [c_void_p, c_void_p, c_void_p]),
    "fricc_shard_emit_systematic": (c_int, [c_void_p, c_int64, c_int32, c_void_p]),
    "fricc_constant_operand": (c_int, [c_void_p, c_int64, c_int]),
    "fricc_prepare_operand": (c_int, [c_void_p, _i64p, _i64p, _i64p, c_char_p]),
    "fricc_shard_emit_systematic_multi": (c_int, [c_void_p, c_int64, c_void_p, c_int, c_void_p, c_int64]),
    "fricc_shard_sample_chunk": (c_int, [c_void_p, c_int64, c_double, c_double, c_double, c_void_p, c_int64, _i64p, c_void_p]),
    "fricc_shard_check": (c_int, [c_void_p]),
    "fricc_shard_candidates": (c_int, [c_void_p, c_uint64, c_void_p, c_void_p, c_int64, c_void_p]),
    "fricc_sort_candidates": (c_int, [c_void_p, c_void_p, c_int64, c_int64]),
    "fricc_d_greedy": (c_int, [c_void_p, c_int64, c_int64, c_double, _i64p, _f64p]),
    "fricc_shard_values": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p]),
}


def lib():
    """The loaded library.  Fails loudly if it has not been built -- there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                f"or `make -C {CSRC}` (nvcc, sm_100a).  fricc_b200 has no CPU fallback."
            )
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status):
    """Turn a C-ABI status into the exception the reference surface would have produced."""
    if status == OK:
        return
    msg = lib().fricc_last_error().decode("utf-8", "replace")
    if status == EINVAL:
        raise ValueError(msg)
    if status == ERANGE:
        raise IndexError(msg)
    if status == ECUDA:
        raise RuntimeError(msg)
    raise FriccError(status, msg)


def i64x4(seq):
    arr = (c_int64 * 4)(*[int(v) for v in seq])
    return arr


def require_cuda():
    n = c_int(0)
    check(lib().fricc_device_count(ctypes.byref(n)))
    if n.value <= 0:
        raise RuntimeError("fricc_b200: no CUDA device available (there is no CPU fallback)")
    return n.value
