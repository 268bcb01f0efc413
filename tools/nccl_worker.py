"""torchrun worker: the golden pipeline checks (tests/pipeline_checks.py) on N GPUs over NCCL, one rank per GPU, rows
sharded across ranks -- the GPU counterpart of tests/gloo_worker.py.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 tools/nccl_worker.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import pipeline_checks  # noqa: E402
from animsnapbases_b200 import engine  # noqa: E402
from animsnapbases_b200.ops import CudaOps  # noqa: E402

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
comm = engine.Comm()
ops = CudaOps(f"cuda:{local}")
pipeline_checks.check_positions(ops, comm, "pos_grid")
pipeline_checks.check_positions(ops, comm, "pos_sphere")
pipeline_checks.check_bar(ops, comm)
pipeline_checks.check_bar_geom(ops, comm)
pipeline_checks.check_bar_reconstruction(ops, comm)
print(f"RANK{comm.rank}/{comm.world} r01-validated checks OK", flush=True)
# added after the last multi-GPU session of r01 (green on two and five gloo ranks, not yet run over NCCL)
pipeline_checks.check_bar_reduced(ops, comm)
pipeline_checks.check_verts_geom(ops, comm)
pipeline_checks.check_zero_residual(ops, comm)
pipeline_checks.check_pd_local(ops, comm)
if "--mirror-classes" in sys.argv:           # main.py's two protocols through the drop-in classes, rows sharded over the GPUs
    import pathlib
    import tempfile
    import api_checks
    base = [tempfile.mkdtemp(prefix="asb_nccl_") if comm.rank == 0 else None]
    dist.broadcast_object_list(base, src=0)
    for name, check in (("pos", api_checks.check_pos_protocol), ("con", api_checks.check_constraints_protocol)):
        tmp = pathlib.Path(base[0]) / name
        if comm.rank == 0:
            tmp.mkdir(parents=True, exist_ok=True)
        dist.barrier()
        assert check(ops, comm, tmp)
        dist.barrier()
print(f"RANK{comm.rank}/{comm.world} OK ({ops.launch_count()} launches)", flush=True)
dist.destroy_process_group()
