"""TEST INFRASTRUCTURE ONLY -- at-scale parity fixtures: runs the UNMODIFIED reference (through oracle/ref_shim.py) IN FULL on
BASELINE.json's configs at their own sizes (animsnapbases_b200/configs.py presets) and commits COMPACT results under
tests/golden/scale_<preset>.npz: index sets, singular values, weights, the SPLOCS history, and for every (K, N, 3) field its
per-component norms, a seeded K x 3 x 8 sketch and a few full rows (the fields themselves are 70 MB and more).  The inputs
are regenerated bit-identically on the GPU box (configs.build_inputs; `digest` pins them).  The same run is the measured
CPU time of the reference on this container's cores -> profiles/r02_reference_cpu_<preset>.json.

    python oracle/make_golden_scale.py bunny_splocs bar_deim armadillo_splocs synthetic_sample
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from animsnapbases_b200 import configs  # noqa: E402
from oracle import ref_runner  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
PROF = os.path.join(ROOT, "profiles")
FULL_ROWS = 3          # components committed in full: first, middle, last


def field_summary(C, G, prefix):
    """Compact form of a (K, N, 3) field: norms, per-dimension sketch, three full rows."""
    K = C.shape[0]
    rows = np.array(sorted({0, K // 2, K - 1}))
    return {prefix + "_norms": np.sqrt((C ** 2).sum(axis=(1, 2))),
            prefix + "_sketch": np.einsum("knd,ndj->kdj", C, G),
            prefix + "_rows": rows, prefix + "_full": C[rows].copy()}


def pos_golden(name):
    inputs = configs.build_inputs(name)
    N = inputs["anim"].shape[1]
    t0 = time.time()
    r = ref_runner.run_pos(inputs, capture_hist=True)
    wall = time.time() - t0
    assert r["kind"] == "reference", "at-scale fixtures come from the unmodified reference only"
    G = configs.sketch_matrix(N)
    out = dict(digest=np.array(inputs["digest"]), picked=r["picked"], meas=r["meas"], weigs=r["weigs"], scale=np.array([r["scale"]]),
               mean_sketch=np.einsum("nd,ndj->dj", r["mean"], G), X_norm2=np.array([float((r["X"] ** 2).sum())]))
    out.update(field_summary(r["comps"], G, "comps"))
    ids = np.array([0, N // 3, N - 1])
    sample = np.random.default_rng(7).choice(N, size=min(N, 2048), replace=False)
    out["geo_ids"], out["geo_sample"] = ids, sample
    out["geo_phi"] = np.array([r["geodesic"](int(i))[sample] for i in ids])
    if "splocs_C" in r:
        out.update(centres=r["centres"], hist=r["hist"], r2=r["r2"], splocs_W=r["splocs_W"])
        out.update(field_summary(r["splocs_C"], G, "splocs_C"))
    np.savez_compressed(os.path.join(GOLD, f"scale_{name}.npz"), **out)
    rec = {"preset": name, "kind": r["kind"], "config": {k: v for k, v in inputs["preset"].items()}, "vertices": N,
           "t_prepare_s": r["t_prepare"], "t_deflation_s": r["t_deflation"], "t_splocs_s": r["t_splocs"],
           "build_s": r["t_deflation"] + r["t_splocs"], "wall_s": wall, "host": ref_runner.host_info(),
           "where": "build container (no GPU); reference's own posComponents.compute_components_store_singvalues via oracle/ref_shim.py"}
    json.dump(rec, open(os.path.join(PROF, f"r02_reference_cpu_{name}.json"), "w"), indent=1)
    print(name, json.dumps(rec))


def bar_golden(name):
    inputs = configs.build_inputs(name)
    ep = inputs["snaps"].shape[1]
    t0 = time.time()
    r = ref_runner.run_bar(inputs)
    wall = time.time() - t0
    assert r["kind"] == "reference"
    G = configs.sketch_matrix(ep)
    out = dict(digest=np.array(inputs["digest"]), sing=r["sing"], Pt=r["Pt"], alpha=r["alpha"], scale=np.array([r["scale"]]),
               mean_sketch=np.einsum("nd,ndj->dj", r["mean"], G), X_norm2=np.array([float((r["X"] ** 2).sum())]))
    out.update(field_summary(r["pod_comps"], G, "pod"))
    out.update(field_summary(r["post_comps"], G, "post"))
    np.savez_compressed(os.path.join(GOLD, f"scale_{name}.npz"), **out)
    rec = {"preset": name, "kind": r["kind"], "config": dict(inputs["preset"]), "rows": ep, "t_prepare_s": r["t_prepare"],
           "t_pod_s": r["t_pod"], "t_post_s": r["t_post"], "t_deim_s": r["t_deim"], "build_s": r["t_pod"] + r["t_post"] + r["t_deim"],
           "wall_s": wall, "host": ref_runner.host_info(),
           "where": "build container (no GPU); reference's own constraintsComponents methods via oracle/ref_shim.py"}
    json.dump(rec, open(os.path.join(PROF, f"r02_reference_cpu_{name}.json"), "w"), indent=1)
    print(name, json.dumps(rec))


if __name__ == "__main__":
    for nm in sys.argv[1:]:
        (bar_golden if configs.PRESETS[nm]["kind"] == "bar" else pos_golden)(nm)
