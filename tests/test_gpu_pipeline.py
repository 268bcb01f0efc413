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
