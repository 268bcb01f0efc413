"""Worker for tests/test_host_logic.py::test_two_rank_gloo_row_sharding (torchrun, gloo, CPU)."""
import os
import sys

import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

from numpy_ops import NumpyOps  # noqa: E402
import pipeline_checks  # noqa: E402
from animsnapbases_b200 import engine  # noqa: E402

dist.init_process_group("gloo")
comm = engine.Comm()
assert comm.world >= 2
ops = NumpyOps()
pipeline_checks.check_positions(ops, comm, "pos_grid")
pipeline_checks.check_bar(ops, comm)
pipeline_checks.check_bar_geom(ops, comm)
pipeline_checks.check_bar_reconstruction(ops, comm)
pipeline_checks.check_bar_reduced(ops, comm)
pipeline_checks.check_verts_geom(ops, comm)
pipeline_checks.check_zero_residual(ops, comm)
pipeline_checks.check_pd_local(ops, comm)
print(f"RANK{comm.rank} OK", flush=True)
dist.destroy_process_group()
