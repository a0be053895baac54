"""fricc_peer_barrier when a rank never arrives: the wait gives up after "peer_barrier_timeout_s" WITHOUT trapping (the
CUDA context of the healthy rank survives -- round-2 advice), and the next fricc_stream_synchronize reports it."""
import ctypes

import pytest

pytestmark = pytest.mark.gpu


def test_missing_peer_is_reported_and_the_context_survives(rccsd):
    import torch
    from fricc_b200 import _lib

    L = _lib.lib()
    _lib.check(L.fricc_set_stream(None))
    mine = torch.zeros(16, dtype=torch.int64, device="cuda")     # this rank's epoch counters
    theirs = torch.zeros(16, dtype=torch.int64, device="cuda")   # stands in for the absent peer's
    ptrs = (ctypes.c_void_p * 2)(mine.data_ptr(), theirs.data_ptr())
    x = torch.rand(10_000, dtype=torch.float64, device="cuda")
    want = rccsd.one_norm(x)
    try:
        rccsd.set_option("peer_barrier_timeout_s", 1)
        _lib.check(L.fricc_peer_barrier(ptrs, 2, 0, 1))           # the peer never stores epoch 1 into mine[1]
        rc = L.fricc_stream_synchronize()
        assert rc == _lib.ECUDA
        assert "gave up waiting" in L.fricc_last_error().decode()
        assert theirs[0].item() == 1                              # our own arrival was published
        assert L.fricc_stream_synchronize() == 0                  # reported once
        assert rccsd.one_norm(x) == want                          # no trap: the context is intact
        mine[1] = 2                                               # the peer "arrives" at barrier 2
        _lib.check(L.fricc_peer_barrier(ptrs, 2, 0, 2))
        assert L.fricc_stream_synchronize() == 0
        assert rccsd.get_stat("peer_barrier_timeouts") == 0
    finally:
        rccsd.set_option("peer_barrier_timeout_s", 300)
