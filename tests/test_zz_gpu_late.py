"""`-m gpu` parity checks added at the very end of r01, kept in a file that sorts last: the consumer-side
operators, the '_verts' geometric DEIM and the real-mesh runs against the live oracle.  They go through
kernels whose own parity tests run earlier in the suite (asb_spmv_csr, asb_gemm_nn, asb_ingest / asb_export_aos, ...)."""
import pytest

import pipeline_checks
from animsnapbases_b200 import engine

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mesh,F,K,splocs,seed", [("bunny", 40, 12, (2, 5, 2.0, 10.0), 2),
                                                  ("armadillo", 60, 24, (2, 10, 2.0, 10.0), 4)])
def test_real_mesh_against_live_oracle(cuda_ops, mesh, F, K, splocs, seed):
    """BASELINE.json configs 1/2/4 on their real meshes, frame / component counts the oracle finishes in seconds."""
    assert pipeline_checks.check_positions_live(cuda_ops, engine.Comm(enabled=False), mesh, F, K, splocs, seed)


def test_bar_reduced_group(cuda_ops):
    assert pipeline_checks.check_bar_reduced(cuda_ops, engine.Comm(enabled=False))


def test_verts_geometric_deim(cuda_ops):
    """'_verts' flavour of the geometric DEIM (p = 1: single-row asb_gemm_nn / asb_spmv_csr calls)."""
    assert pipeline_checks.check_verts_geom(cuda_ops, engine.Comm(enabled=False))


def test_constraints_protocol_with_consumer(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_constraints_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path, extended=True)


def test_verts_geometric_deim_protocol(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_verts_geom_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path)


def test_pos_protocol_with_figure_tables(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_pos_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path, extended=True)


def test_zero_residual(cuda_ops):
    """Vanishing DEIM residual: print + no result, like the reference, in all three selections."""
    assert pipeline_checks.check_zero_residual(cuda_ops, engine.Comm(enabled=False))


def test_pd_local_snapshot_producer(cuda_ops, tmp_path):
    assert pipeline_checks.check_pd_local(cuda_ops, engine.Comm(enabled=False), tmp_path)
