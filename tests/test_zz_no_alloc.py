"""fricc_one_norm_static: the exact one_norm for callers that must not allocate device memory.

Some programs that drive libfricc_b200 cannot let it allocate while they run: a coupled-cluster driver whose own arena
already fills the GPU (nothing is left for the library's pool, and a failed allocation in the middle of an iteration
cannot be recovered from), a host thread that must not block inside the kernel driver (an allocation that has to map
memory can wait there for seconds), or a process held to a fixed device-memory budget by its launcher.  For them the
library offers one entry point whose working memory -- the chunk sums, chunk maps and carries of the exact scan -- is a
fixed 1.3 MB array inside the library image instead of buffers from the block cache.  Every call on a GPU shares that
array and finds in it whatever the previous call left, so what this file checks is exactly that: the result equals
fricc_one_norm's bit for bit (integer-exact scan: the tolerance of the compression tests is equality), on two calls in
a row whose inputs differ in size and content, for a host result and for a device result, and inputs beyond the
array's capacity are refused with a message instead of overrunning it.

The file sorts last on purpose (pytest -x stops at the first failure: a failure here must not hide other tests) and
does not depend on anything the earlier tests did: it runs alone in a fresh process as well as after the whole suite.
Everything goes through the C-ABI with ctypes; torch only provides the device arrays.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FRICC_HOST, FRICC_DEVICE = 0, 1
FRICC_EINVAL = -1
MAX_N = 1 << 26  # FRICC_ONE_NORM_STATIC_MAX_N (include/fricc_b200.h)


@pytest.fixture(scope="module")
def L():
    from fricc_b200 import _lib

    _lib.require_cuda()  # fails, does not skip, when the CUDA library or the device is missing
    lib = _lib.lib()
    _lib.check(lib.fricc_set_stream(None))  # the default stream, whatever an earlier test left selected
    return lib


def _existing(L, x_dev):
    """fricc_one_norm on a device array (this one takes its scratch from the library's allocator)."""
    from fricc_b200 import _lib

    r = ctypes.c_double(-1.0)
    _lib.check(L.fricc_one_norm(x_dev.data_ptr(), FRICC_DEVICE, x_dev.numel(), ctypes.byref(r)))
    return r.value


def _static_host(L, x_dev, n=None):
    from fricc_b200 import _lib

    r = ctypes.c_double(-1.0)
    n = x_dev.numel() if n is None else n
    _lib.check(L.fricc_one_norm_static(x_dev.data_ptr(), n, ctypes.addressof(r), FRICC_HOST))
    return r.value


def _static_device(L, x_dev, out_dev):
    from fricc_b200 import _lib

    _lib.check(L.fricc_one_norm_static(x_dev.data_ptr(), x_dev.numel(), out_dev.data_ptr(), FRICC_DEVICE))


def _inputs(torch):
    """Two inputs that differ in length and content; the first is the longer one, so the second call finds chunk sums,
    maps and carries of the first one behind its own in the shared array.  Both cross many binades of the running sum
    and carry a few entries that dwarf their neighbours (the case the exact scan has to redo chunk by chunk)."""
    rng = np.random.default_rng(20261106)
    a = rng.standard_normal(3_000_017) * 10.0 ** (-6.0 * rng.random(3_000_017) ** 2)
    a[[5, 70_001, 2_999_999]] = [3.0e7, -8.0e9, 1.0e-300]
    b = rng.random(1_234_567) - rng.random(1_234_567)
    b[123_456] = -4.0e12
    return [torch.from_numpy(v).cuda() for v in (a, b)], (a, b)


def test_two_calls_in_a_row_equal_the_allocating_entry_point(L):
    import torch

    (xa, xb), (ha, hb) = _inputs(torch)
    want_a, want_b = _existing(L, xa), _existing(L, xb)
    # the allocating entry point is itself pinned to the sequential sum elsewhere; restate it here so that this file
    # stands alone (np.cumsum adds left to right in IEEE doubles, like fri_utils.cpp:164-171 on one thread)
    assert want_a == float(np.cumsum(np.abs(ha))[-1])
    assert want_b == float(np.cumsum(np.abs(hb))[-1])

    # two calls in a row, host result: long input first, short input second
    got_a = _static_host(L, xa)
    got_b = _static_host(L, xb)
    assert got_a == want_a
    assert got_b == want_b

    # and again the other way round with the result left on the device (nothing synchronises between the two calls:
    # stream order alone separates their use of the shared array)
    out = torch.full((2,), -1.0, dtype=torch.float64, device="cuda")
    _static_device(L, xb, out[0:1])
    _static_device(L, xa, out[1:2])
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [want_b, want_a]

    # a prefix of the same array: the call after a longer one must not pick up its chunks
    n = 4097
    assert _static_host(L, xa, n) == float(np.cumsum(np.abs(ha[:n]))[-1])


def test_static_entry_point_makes_no_allocation_call(L):
    """The library counts its own cudaMallocAsync calls and block-cache hits: neither moves across the static calls."""
    import torch
    from fricc_b200 import _lib

    x = torch.rand(500_000, dtype=torch.float64, device="cuda")
    _static_host(L, x)  # this device's address look-up is behind us now

    def counters():
        a, h = ctypes.c_int64(0), ctypes.c_int64(0)
        _lib.check(L.fricc_get_stat(b"alloc_driver_calls", ctypes.byref(a)))
        _lib.check(L.fricc_get_stat(b"alloc_cache_hits", ctypes.byref(h)))
        return a.value, h.value

    before = counters()
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        _static_host(L, x)
    assert counters() == before
    assert torch.cuda.mem_get_info()[0] >= free_before  # no direct cudaMalloc either
    assert _static_host(L, x) == _existing(L, x)
    assert counters() != before  # the allocating entry point does show up in these counters


def test_limits_and_edge_cases(L):
    import torch

    x = torch.rand(1000, dtype=torch.float64, device="cuda")
    r = ctypes.c_double(-1.0)
    # beyond the capacity of the fixed array: refused before anything is launched (the pointer is never touched)
    rc = L.fricc_one_norm_static(x.data_ptr(), MAX_N + 1, ctypes.addressof(r), FRICC_HOST)
    assert rc == FRICC_EINVAL
    msg = L.fricc_last_error().decode()
    assert "FRICC_ONE_NORM_STATIC_MAX_N" in msg and str(MAX_N) in msg
    assert r.value == -1.0
    # bad arguments
    assert L.fricc_one_norm_static(x.data_ptr(), -1, ctypes.addressof(r), FRICC_HOST) == FRICC_EINVAL
    assert L.fricc_one_norm_static(x.data_ptr(), 10, None, FRICC_HOST) == FRICC_EINVAL
    assert L.fricc_one_norm_static(None, 10, ctypes.addressof(r), FRICC_HOST) == FRICC_EINVAL
    assert L.fricc_one_norm_static(x.data_ptr(), 10, ctypes.addressof(r), 7) == FRICC_EINVAL
    # the empty sum, and a still usable entry point after the refusals
    assert _static_host(L, x, 0) == 0.0
    assert _static_host(L, x) == _existing(L, x)
    # all zeros / subnormals: no binade to predict, every chunk goes through the fallback
    z = torch.zeros(10_000, dtype=torch.float64, device="cuda")
    z[777] = 5e-324
    assert _static_host(L, z) == 5e-324


def test_largest_supported_size(L):
    """n = FRICC_ONE_NORM_STATIC_MAX_N exactly (512 MB of doubles, 32768 chunks: every slot of the array in use)."""
    import torch

    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.rand(MAX_N, dtype=torch.float64, device="cuda", generator=g) - 0.5
    assert _static_host(L, x) == _existing(L, x)
    # followed at once by a tiny input
    assert _static_host(L, x, 3) == float(np.cumsum(np.abs(x[:3].cpu().numpy()))[-1])


def test_python_mirror(L):
    """py_rccsd.one_norm_static: the same call for a driver that holds torch tensors."""
    import torch
    from fricc_b200 import py_rccsd

    x = torch.rand(70_000, dtype=torch.float64, device="cuda") - 0.25
    want = py_rccsd.one_norm(x)
    assert py_rccsd.one_norm_static(x) == want
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    assert py_rccsd.one_norm_static(x, out) is out
    assert out.item() == want
    with pytest.raises(TypeError):
        py_rccsd.one_norm_static(x.cpu().numpy())  # a host array would have to be staged, i.e. allocated for
