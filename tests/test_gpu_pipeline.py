"""-m gpu: the full hot path through the C-ABI kernels against the golden outputs of the reference."""
import pytest

import pipeline_checks
from animsnapbases_b200 import engine

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["pos_grid", "pos_sphere"])
def test_position_pipeline(cuda_ops, name):
    assert pipeline_checks.check_positions(cuda_ops, engine.Comm(enabled=False), name)


def test_bar_pipeline(cuda_ops):
    assert pipeline_checks.check_bar(cuda_ops, engine.Comm(enabled=False))


def test_bar_blocks_pipeline(cuda_ops):
    assert pipeline_checks.check_bar_blocks(cuda_ops, engine.Comm(enabled=False))


def test_bar_pod_per_row(cuda_ops):
    assert pipeline_checks.check_bar_pod_per_row(cuda_ops, engine.Comm(enabled=False))


def test_bar_reconstruction(cuda_ops):
    assert pipeline_checks.check_bar_reconstruction(cuda_ops, engine.Comm(enabled=False))


def test_bar_geometric_deim(cuda_ops):
    assert pipeline_checks.check_bar_geom(cuda_ops, engine.Comm(enabled=False))


def test_pos_components_protocol(cuda_ops, tmp_path):
    """The reference's main.py:71-100 protocol through libasb.so, against the reference's golden outputs."""
    import api_checks
    from animsnapbases_b200 import engine
    assert api_checks.check_pos_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path)


@pytest.mark.gpu
def test_constraints_components_protocol(cuda_ops, tmp_path):
    import api_checks
    assert api_checks.check_constraints_protocol(cuda_ops, engine.Comm(enabled=False), tmp_path)


def test_bar_blocks_with_St(cuda_ops):
    """pca_blocks_with_St (constraintsComponents.py:156-271) against the unmodified reference's output."""
    assert pipeline_checks.check_bar_blocks_st(cuda_ops, engine.Comm(enabled=False))


@pytest.mark.parametrize("name", ["pos_grid", "pos_sphere"])
def test_pos_orth_branch(cuda_ops, name):
    assert pipeline_checks.check_pos_orth(cuda_ops, engine.Comm(enabled=False), name)


def test_admm_rejects_more_than_256_components(cuda_ops):
    """asb_admm_solve / asb_admm_solve_ws refuse K > 256 with ASB_ERR_ARG (include/asb.h) instead of launching."""
    import torch
    from animsnapbases_b200._lib import AsbError
    K, N, ld = 264, 40, 64
    Hr = cuda_ops.zeros((272, 272))
    B0, Z, U = cuda_ops.zeros((K, 3, ld)), cuda_ops.zeros((K, 3, ld)), cuda_ops.zeros((K, 3, ld))
    Lam = cuda_ops.zeros((K, ld))
    with pytest.raises(AsbError, match="K <= 256"):
        cuda_ops.admm_solve(Hr, B0, Z, U, Lam, K, N, 0.1, 2)
    torch.cuda.synchronize()
