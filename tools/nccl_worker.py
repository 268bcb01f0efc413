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
print(f"RANK{comm.rank}/{comm.world} OK ({ops.launch_count()} launches)", flush=True)
dist.destroy_process_group()
