"""The block cache in front of cudaMallocAsync (fricc_runtime.cu): blocks are cached under the device that owns them,
carry an event recorded at their free, and the stream that takes one out again waits for that event -- so a block freed
while the library issued on one stream can be reused from another without an ordering hole (round-2 advice: the cache
used to be keyed by the raw stream handle, which the driver can hand out again after the stream is destroyed)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _stat(L, key):
    from fricc_b200 import _lib

    v = ctypes.c_int64(0)
    _lib.check(L.fricc_get_stat(key, ctypes.byref(v)))
    return v.value


def test_blocks_are_reused_across_streams_and_results_stay_exact(rccsd):
    import torch
    from fricc_b200 import _lib

    L = _lib.lib()
    rng = np.random.default_rng(5)
    xs = [rng.standard_normal(n) * np.exp(-3.0 * rng.random(n)) for n in (400_003, 400_003, 400_003)]
    want = [float(np.cumsum(np.abs(x))[-1]) for x in xs]
    dev = [torch.from_numpy(x).cuda() for x in xs]
    torch.cuda.synchronize()
    try:
        _lib.check(L.fricc_set_option(b"release_cached", 1))
        assert rccsd.one_norm(dev[0]) == want[0]          # default stream: allocates, then leaves its blocks cached
        calls0, hits0 = _stat(L, b"alloc_driver_calls"), _stat(L, b"alloc_cache_hits")
        for rep in range(3):                               # a fresh stream each time (handles may repeat)
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            _lib.check(L.fricc_set_stream(ctypes.c_void_p(s.cuda_stream)))
            for x, w in zip(dev, want):
                assert rccsd.one_norm(x) == w
            _lib.check(L.fricc_set_stream(None))
            assert rccsd.one_norm(dev[rep]) == want[rep]
            del s
        # same-size requests on other streams were served from the blocks the default stream freed
        assert _stat(L, b"alloc_cache_hits") > hits0
        assert _stat(L, b"alloc_driver_calls") == calls0
        _lib.check(L.fricc_set_option(b"release_cached", 1))
        assert rccsd.one_norm(dev[2]) == want[2]
        assert _stat(L, b"alloc_driver_calls") > calls0   # the cache was really emptied
    finally:
        _lib.check(L.fricc_set_stream(None))
