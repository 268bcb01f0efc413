"""-m gpu: parity AT BASELINE.json's OWN SIZES against results of the unmodified reference run in full on bit-identical inputs
(tests/golden/scale_*.npz, oracle/make_golden_scale.py): bunny and armadillo F = 200 / K = 200 / 20 x 10 SPLOCS, the full bar
(3 120 tets, F = 70, K = 70, pod_vectorized -> post-process -> deim end to end), and the N = 1e5 / F = 500 / K = 200 sample of
the synthetic headline config.  The per-test error report is written to gpurun_out/ when that directory exists."""
import json
import os

import pytest

import pipeline_checks
from animsnapbases_b200 import engine

pytestmark = pytest.mark.gpu


def _save(name, rep):
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, f"parity_scale_{name}.json"), "w") as fh:
            json.dump(rep, fh, indent=1, default=float)


@pytest.mark.parametrize("name", ["bunny_splocs", "armadillo_splocs", "armadillo_splocs_f1000", "synthetic_sample"])
def test_position_configs_at_scale(cuda_ops, name):
    if not os.path.exists(os.path.join(os.path.dirname(__file__), "golden", f"scale_{name}.npz")):
        pytest.skip(f"fixture scale_{name}.npz not generated")
    rep = {}
    try:
        assert pipeline_checks.check_scale_pos(cuda_ops, engine.Comm(enabled=False), name, report=rep)
    finally:
        _save(name, rep)


def test_bar_config_at_scale(cuda_ops):
    rep = {}
    try:
        assert pipeline_checks.check_scale_bar(cuda_ops, engine.Comm(enabled=False), report=rep)
    finally:
        _save("bar_deim", rep)
