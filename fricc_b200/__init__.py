"""fricc_b200 -- B200 (sm_100a) implementation of fri-cc's FRI-CCSD inner loop.

The package holds only what the hot path needs:

* ``csrc/``          hand-written CUDA kernels + the C-ABI (``include/fricc_b200.h``) -> ``libfricc_b200.so``
* ``py_rccsd``       host-side mirror of the reference's pybind11 module ``fricc/py_rccsd`` (same names, argument
                     meaning and error behaviour; /root/reference/src/py_rccsd.cpp:16-55)
* ``dist``           one-process-per-GPU orchestration of the sharded ladder / distributed compression (torch.distributed)

Importing the package does not need a GPU; any compute call without the built CUDA library or without a CUDA device
raises -- there is no CPU fallback.
"""
from . import _lib  # noqa: F401

__all__ = ["py_rccsd", "_lib"]
