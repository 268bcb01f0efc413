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


def _join_torchrun():
    """`torchrun --nproc-per-node N main.py`: the reference's main.py knows nothing about ranks, so the first
    use of the engine binds this process to its GPU (LOCAL_RANK) and joins the NCCL group torchrun described in the
    environment.  A process group that the host program initialised itself is left alone."""
    import os
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return
    import torch
    import torch.distributed as dist
    if not dist.is_available() or dist.is_initialized():
        return
    local = int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def get():
    global _ops, _comm
    if _ops is None:
        _join_torchrun()
        from .ops import CudaOps
        _ops = CudaOps()
    if _comm is None:
        from .engine import Comm
        _comm = Comm()
    return _ops, _comm
