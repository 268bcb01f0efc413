"""`-m gpu` parity of the consumer-side operators through libasb.so.  Kept in a file that sorts last: these
checks were added after the round's last GPU session, on kernels (asb_spmv_csr, asb_gemm_nn, asb_ingest / asb_export_aos)
whose own parity tests run earlier in the suite."""
import pytest

import pipeline_checks
from animsnapbases_b200 import engine

pytestmark = pytest.mark.gpu


def test_bar_reduced_group(cuda_ops):
    assert pipeline_checks.check_bar_reduced(cuda_ops, engine.Comm(enabled=False))


def test_verts_geometric_deim(cuda_ops):
    """'_verts' flavour of the geometric DEIM (p = 1: single-row asb_gemm_nn / asb_spmv_csr calls)."""
    assert pipeline_checks.check_verts_geom(cuda_ops, engine.Comm(enabled=False))


def test_constraints_protocol_with_consumer(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_constraints_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path, consumer=True)


def test_verts_geometric_deim_protocol(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_verts_geom_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path)
