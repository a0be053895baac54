"""The C-ABI library loads and exports every symbol include/fricc_b200.h declares (no compute calls: CPU-only)."""
import ctypes
import os
import re

import pytest

from fricc_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "fricc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fricc_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built_in_tree():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"


def test_every_declared_symbol_is_exported():
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/fricc_b200.h but not exported"


def test_binding_table_covers_header():
    assert sorted(_lib.SIGNATURES) == declared_symbols()


def test_version_and_error_string():
    L = _lib.lib()
    assert b"sm_100a" in L.fricc_version()
    assert isinstance(L.fricc_last_error(), bytes)


def test_no_cpu_fallback_without_device():
    """Without a CUDA device a compute entry point must fail loudly, not fall back."""
    import numpy as np

    n = ctypes.c_int(0)
    _lib.check(_lib.lib().fricc_device_count(ctypes.byref(n)))
    if n.value > 0:
        pytest.skip("a CUDA device is present")
    from fricc_b200 import py_rccsd

    x = np.arange(16, dtype=np.float64).reshape(1, 1, 4, 4)
    with pytest.raises(RuntimeError):
        py_rccsd.SparseTensor4d(x, x.shape, 3, "largest")


def test_glibc_rand_stream_matches_libc():
    """The library's rand() stream is glibc's (the reference draws with libc rand(), fri_utils.cpp:47,74)."""
    import numpy as np

    libc = ctypes.CDLL(None)
    L = _lib.lib()
    for seed in (1, 7, 20, 123456):
        libc.srand(seed)
        want = np.array([libc.rand() for _ in range(2000)], dtype=np.int32)
        _lib.check(L.fricc_srand(seed))
        got = np.empty(2000, dtype=np.int32)
        _lib.check(L.fricc_rand_fill(got.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), 1200))
        _lib.check(L.fricc_rand_fill(got[1200:].ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), 800))
        assert (got == want).all()
        assert L.fricc_rand_consumed() == 2000


def test_rand_skip_is_a_jump_ahead_of_the_same_stream():
    """fricc_rand_skip(n) (O(log n) polynomial jump over the linear recurrence) == drawing n values and dropping them."""
    import numpy as np

    libc = ctypes.CDLL(None)
    L = _lib.lib()
    p32 = ctypes.POINTER(ctypes.c_int32)
    for seed, skip in ((1, 1), (3, 30), (3, 31), (9, 992), (20, 12345), (5, 1_000_003)):
        libc.srand(seed)
        for _ in range(skip):
            libc.rand()
        want = np.array([libc.rand() for _ in range(200)], dtype=np.int32)
        _lib.check(L.fricc_srand(seed))
        _lib.check(L.fricc_rand_skip(skip))
        got = np.empty(200, dtype=np.int32)
        _lib.check(L.fricc_rand_fill(got.ctypes.data_as(p32), 200))
        assert (got == want).all(), (seed, skip)
        assert L.fricc_rand_consumed() == skip + 200


def test_plane_extensions_reject_host_arrays_and_bad_arguments():
    """constant_operand / prepare_operand are about DEVICE operands (host arrays are staged per call, there is nothing to
    keep or to prepare): the mirror raises TypeError, the C-ABI reports bad arguments by status code."""
    import numpy as np

    from fricc_b200 import py_rccsd

    w = np.zeros((2, 3, 2, 3))
    with pytest.raises(TypeError):
        py_rccsd.constant_operand(w)
    with pytest.raises(TypeError):
        py_rccsd.prepare_operand(w, np.zeros((3, 3, 2, 2)), "0112")
    L = _lib.lib()
    assert L.fricc_constant_operand(None, 10, 1) == _lib.EINVAL
    assert L.fricc_constant_operand(ctypes.c_void_p(4096), 0, 1) == _lib.EINVAL
    i4 = _lib.i64x4((2, 2, 2, 2))
    assert L.fricc_prepare_operand(None, i4, i4, i4, b"2323") == _lib.EINVAL
    assert L.fricc_prepare_operand(ctypes.c_void_p(4096), i4, i4, i4, b"9999") == _lib.EINVAL
    assert b"unknown contraction term" in L.fricc_last_error()
    # declaring and withdrawing a range touches no memory: allowed without a device
    assert L.fricc_constant_operand(ctypes.c_void_p(4096), 16, 1) == 0
    assert L.fricc_constant_operand(ctypes.c_void_p(4096), 16, 0) == 0
