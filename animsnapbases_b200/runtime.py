"""Process-wide device context of the basis engine: one CudaOps (libasb.so on the current CUDA device) and one Comm
(torch.distributed when initialised, else single rank).  There is no CPU backend in the product: `get()` raises AsbError
without a GPU or without the built library.  `install()` exists so that tests can inject the numpy statement of the
kernels (tests/numpy_ops.py) to exercise the host logic on machines without a GPU."""
from __future__ import annotations

_ops = None
_comm = None


def install(ops, comm=None):
    global _ops, _comm
    _ops, _comm = ops, comm


def get():
    global _ops, _comm
    if _ops is None:
        from .ops import CudaOps
        _ops = CudaOps()
    if _comm is None:
        from .engine import Comm
        _comm = Comm()
    return _ops, _comm
