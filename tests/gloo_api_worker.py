"""Worker for tests/test_host_logic.py::test_two_rank_gloo_mirror_classes (torchrun, gloo, CPU): the drop-in classes
(posComponents / constraintsComponents protocols of main.py) with their rows sharded over the ranks; rank 0 writes the files."""
import os
import sys
import pathlib

import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

from numpy_ops import NumpyOps  # noqa: E402
import api_checks  # noqa: E402
from animsnapbases_b200 import engine  # noqa: E402

dist.init_process_group("gloo")
comm = engine.Comm()
assert comm.world >= 2
base = pathlib.Path(sys.argv[1])
for name, check in (("pos", api_checks.check_pos_protocol), ("con", api_checks.check_constraints_protocol)):
    tmp = base / name
    if comm.rank == 0:
        tmp.mkdir(parents=True, exist_ok=True)
    dist.barrier()
    assert check(NumpyOps(), comm, tmp)
    dist.barrier()
print(f"RANK{comm.rank} OK", flush=True)
dist.destroy_process_group()
