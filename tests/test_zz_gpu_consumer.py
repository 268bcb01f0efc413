"""`-m gpu` parity of the consumer-side operators through libasb.so.  Kept in a file that sorts last: these
checks were added after the round's last GPU session, on kernels (asb_spmv_csr, asb_gemm_nn, asb_ingest / asb_export_aos)
whose own parity tests run earlier in the suite."""
import pytest

import pipeline_checks
from animsnapbases_b200 import engine

pytestmark = pytest.mark.gpu


def test_bar_reduced_group(cuda_ops):
    assert pipeline_checks.check_bar_reduced(cuda_ops, engine.Comm(enabled=False))


def test_constraints_protocol_with_consumer(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_constraints_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path, consumer=True)
