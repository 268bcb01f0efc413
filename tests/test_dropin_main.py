"""Drop-in check of the boundary: the reference's UNMODIFIED `main.main(param)` and its own
`config.config.Config_parameters` (json-driven) run over this package's posComponents / constraintsComponents, which are
put in place of `snapbases.posComponents` / `snapbases.constraintsComponents` the way INTEGRATION.md describes.  Kernels are
the numpy statement (tests/numpy_ops.py): this is a host-protocol test, it needs the reference mounted (skipped on the GPU
box) and no GPU.  Outputs are compared with the reference's golden results."""
import json
import os
import sys

import numpy as np
import pytest

from conftest import golden, rel
from numpy_ops import NumpyOps
from animsnapbases_b200 import engine, hostio, runtime
from oracle import port, ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.reference_available(), reason="reference not mounted (GPU box)")


@pytest.fixture
def reference_main(monkeypatch, tmp_path):
    """`main` of the reference, importable, with this package's classes under the reference's module names; CWD = scratch
    (config.py and log_time write relative paths)."""
    ref_shim.load_reference()
    import animsnapbases_b200.snapbases.constraintsComponents as ours_con
    import animsnapbases_b200.snapbases.posComponents as ours_pos
    monkeypatch.setitem(sys.modules, "snapbases.posComponents", ours_pos)
    monkeypatch.setitem(sys.modules, "snapbases.constraintsComponents", ours_con)
    monkeypatch.chdir(tmp_path)
    runtime.install(NumpyOps(), engine.Comm(enabled=False))
    monkeypatch.delitem(sys.modules, "main", raising=False)
    import main as ref_main
    assert os.path.samefile(ref_main.__file__, os.path.join(ref_shim.REFERENCE_ROOT, "main.py"))
    yield ref_main
    runtime.install(None, None)
    sys.modules.pop("main", None)


def _params(cfg):
    from config.config import Config_parameters
    with open("cfg.json", "w") as fh:
        json.dump(cfg, fh)
    param = Config_parameters()
    param.reset("cfg.json")
    return param


NO_POS = {"computeState": {"compute": False}}
NO_CON = {"computeState": {"compute": False, "run_main": True}}


def test_reference_main_drives_pos_components(reference_main):
    """main.py:19-100 with a config.json shaped like config/examples/bunny_gFall_posSubspace.json (repaired: valid JSON,
    `volumetric` and `run_main` keys present), SPLOCS on, local support, run_tests on."""
    g = golden("pos_grid")
    F, N, _ = g["anim"].shape
    K = int(g["K"])
    its, admm, lam, rho = g["splocs"]
    cfg = {"object": {"mesh": "grid", "volumetric": False, "experiment": "_bumps", "experiment_dir": "data_in/", "snap_format": ".off"},
           "vertexPos_bases": {
               "computeState": {"compute": True, "testingComputations": "_Release"},
               "snapshots": {"max_numFrames": F, "frame_increment": 1, "numFrames": F, "format": ".off", "snaps_folder": "fom",
                             "read_all_from_first": True, "anims_folder": "FOM_animations_h5", "preAlignement": "_centered",
                             "anim_folder_ready": False, "visualize_aligned_animations": False, "reduced_snaps_available": False},
               "dim": 3, "rest_shape": "first", "massWeighted": "_Volkwein", "standarized": "_Standarized",
               "orthogonalized": "_nonOrthogonalized", "support": {"min_dist": 0.1, "max_dist": 0.25},
               "pca": {"compute": True, "numComponents": K, "supported": "_Local", "store_sing_val": True},
               "splocs": {"compute": True, "max_itrs": int(its), "admm_num_itrs": int(admm), "lambda": float(lam), "rho": float(rho)},
               "store": True, "run_tests": True, "visualize": False},
           "constraintProj_bases": NO_CON}
    param = _params(cfg)
    os.makedirs(param.aligned_snapshots_directory, exist_ok=True)
    for name in (param.train_aligned_snapshots_animation_file, param.test_aligned_snapshots_animation_file):
        path = os.path.join(param.aligned_snapshots_directory, name)
        twin = hostio.write_animation(path, g["anim"], g["tris"])
        if twin != path:                     # no h5py here: the data lives in the .npz twin; main.py:29 only tests existence
            open(path, "wb").close()
    os.makedirs(os.path.dirname(param.vertPos_masses_file), exist_ok=True)
    hostio.write_mass_bin(param.vertPos_masses_file, g["mass"])

    reference_main.main(param)

    out = param.vertPos_output_directory
    assert os.path.exists(os.path.join(out, "time_logs.txt")), "main.py:85 moves function_timings.txt there"
    rows = open(os.path.join(out, "grid_posBases_pcaExtraction_singValues_errorNorm.csv")).read().splitlines()
    assert rows[0] == "component,singVal,norm_R" and len(rows) == K + 1
    meas = np.array([[float(v) for v in r.split(",")] for r in rows[1:]])
    assert np.allclose(meas, g["meas_local"], rtol=1e-7)
    # main.py:96-100: one .bin per k = 1..K; the last one holds all post-processed components
    prep = port.prepare_snapshots(g["anim"], g["mass"], "first")
    post = port.post_process_pos(g["comps_local"].copy(), prep["mean"], prep["pre_scale_factor"], prep["invMassL"], True, False, True)
    for k in (1, K):
        raw = open(os.path.join(out, "q_pos_F%dK%d.bin" % (F, k)), "rb").read()
        n, cols = np.frombuffer(raw[:8], "<i4")
        assert (n, cols) == (N, 3 * k) and len(raw) == 8 + 8 * N * 3 * k
        B = np.frombuffer(raw[8:], "<f8").reshape(3, k, N).transpose(1, 2, 0)
        assert rel(B, post[:k]) < 1e-7
    assert os.path.exists(os.path.join(out, "components.npz")) or os.path.exists(os.path.join(out, "components.h5"))
    # run_tests = true: the reference's own figure script (generate_figures/pos_reduction_tests.py, pyplot stubbed) consumed the
    # object: its singular-value table is there and carries what reduction_tests.pos_pca_tests (headless) writes
    tab = open(os.path.join(out, "grid_posBases_xyz_fullBasesRange_K_singVals.csv")).read().splitlines()
    assert tab[0] == "row,x,y,z" and len(tab) == K + 1
    ref_s = port.bases_sing_vals(post)
    got_s = np.array([[float(v) for v in r.split(",")[1:]] for r in tab[1:]])
    assert np.allclose(got_s, ref_s, atol=1e-7)


def test_reference_main_drives_constraints_components(reference_main):
    """main.py:102-142 with a config.json shaped like config/examples/bar_automated_deim_tetDeformationGradientSubspace.json:
    pod_vectorized -> post-process -> deim -> one .npz for the simulator."""
    b = golden("bar")
    F, ep, _ = b["snaps"].shape
    K = 10
    folder = "/constraint_projection/FOM/tets_deformation_gradient"
    cfg = {"object": {"experiment_dir": "pd_out/", "mesh": "bar", "volumetric": True, "experiment": "bar_defgrad/", "snap_format": ".off"},
           "vertexPos_bases": NO_POS,
           "constraintProj_bases": {
               "computeState": {"compute": True, "run_main": True, "testingComputations": "_Release"},
               "constraintType": {"name": "tets_deformation_gradient", "elements": "_tets", "p_snaps_folder": folder,
                                  "pos_snaps_folder": "/position_snapshots/FOM", "geom_pos_snaps_folder": "/position_snapshots/deim/",
                                  "assembly_file_name": "assembly_ST.npz", "assembly_key": "tets_deformation_gradient",
                                  "snaps_pattern_full_p": "/tets_deformation_gradient_p.npz", "constrained_elements": "", "rowSize": 3},
               "snapshots": {"max_numFrames": F, "numFrames": F, "frame_increment": 1, "read_all_from_first": True,
                             "processed_snapshots_file": "/processed_snapshots/", "preAlignement": "_noAlignement",
                             "processed_snapshots_ready": False, "reduced_snaps_available": False},
               "basis_type": "pod_vectorized", "interpolation_type": "deim", "desired_num_components": K, "numComponents_verts": 5,
               "bases_res_tol": 98, "dim": 3, "constraintsDim": 1, "max_element_per_geom_vert": 100000000, "rest_shape": "first",
               "massWeighted": "_Volkwein", "standarized": "_Standarized", "supported": "_Global", "orthogonalized": "_Orthogonalized",
               "store_sing_val": True, "store_to_files": True, "run_tests": True, "visualize_elements_at_bases_num": 30,
               "visualize_geom_elements": 528}}
    param = _params(cfg)
    os.makedirs(os.path.dirname(param.constProj_input_snapshots_pattern), exist_ok=True)
    np.savez(param.constProj_input_snapshots_pattern, **{str(i): b["snaps"][i] for i in range(F)})
    os.makedirs(os.path.dirname(param.constProj_masses_file), exist_ok=True)
    hostio.write_mass_bin(param.constProj_masses_file, b["mass"])
    os.makedirs(param.vertPos_output_directory or ".", exist_ok=True)

    reference_main.main(param)

    z = np.load(os.path.join(param.constProj_output_directory, "components_interpol_alphas_interpol_verts_interpol_alpha_ranges.npz"),
                allow_pickle=True)
    assert set(z.files) == {"components", "interpol_alphas", "Pt", "interpol_verts", "interpol_alpha_ranges"}
    assert z["components"].shape == (K, ep, 3)
    rows = open(os.path.join(param.constProj_output_directory,
                             "bar_tets_deformation_gradient_constrprojBases_pcaExtraction_singValues.csv")).read().splitlines()
    sig = np.array([float(r.split(",")[1]) for r in rows[1:]])
    assert np.all(np.abs(sig[:K] / b["pod_sing"][:K] - 1) < 1e-9)
    # SVD signs are LAPACK's choice and post-processing adds the mean to every mode, so the emitted vectors themselves are
    # compared sign-pinned in tests/api_checks.py::check_constraints_protocol; here: what any sign choice must satisfy
    for d in range(3):
        U = z["components"][:, :, d]
        assert np.allclose((U * b["mass"][None]) @ U.T, np.eye(K), atol=1e-8), "orthogonalised + un-weighted: U^T M U = I"
    assert z["Pt"].shape == (K,) and len(np.unique(z["Pt"])) == K and np.array_equal(z["interpol_alphas"], z["Pt"] // 3)
    assert np.array_equal(z["interpol_alpha_ranges"], np.arange(1, K + 1))
    # run_tests = true: the reference's nl_reduction_tests figure script (pyplot / polyscope stubbed) drove geom_constructed
    # and the error statics over the object and wrote its convergence tables
    stem = os.path.join(param.constProj_output_directory, "bar_tets_deformation_gradient_deim_pod_vectorized")
    tab = open(stem + "_train_convergence_tests.csv").read().splitlines()
    assert tab[0].startswith("numPoints,fro_error,max_err") and len(tab) == K + 1
    ref_rows = np.array([[float(v) for v in r.split(",")] for r in tab[1:]])
    assert os.path.exists(stem + "_test_convergence_tests.csv") and os.path.exists(stem + "_num_interpol_elemnets.csv")
    # (not monotone: the reference reconstructs from element ids used as row ids, constraintsComponents.py:512)
    assert np.isfinite(ref_rows).all() and ref_rows[0, 1] > ref_rows[-1, 1] > 0 and np.array_equal(ref_rows[:, 0], np.arange(1, K + 1))


def _write_medit(path, verts, tets):
    with open(path, "w") as fh:
        fh.write("MeshVersionFormatted 1\nDimension 3\nVertices\n%d\n" % len(verts))
        fh.writelines("%.17g %.17g %.17g 0\n" % tuple(v) for v in verts)
        fh.write("Tetrahedra\n%d\n" % len(tets))
        fh.writelines("%d %d %d %d 0\n" % tuple(t + 1) for t in tets)
        fh.write("End\n")


def test_reference_main_drives_geometric_interpolation(reference_main):
    """main.py:102-142 with `basis_type: pca_blocks`, `interpolation_type: geom` (shaped like
    config/examples/bar_automated_geom_tetDeformationGradientSubspace.json): block deflation -> post-process -> geometric DEIM
    through the pickled `assembly_ST.npz` -> the simulator's .npz.  Block signs are free (and post-processing adds the mean),
    so the picks are checked structurally here; the sign-pinned comparison with the reference's picks is check_bar_geom."""
    from scipy import sparse
    b, g, m = golden("bar"), golden("bar_geom"), golden("meshes")
    F, ep, _ = b["snaps"].shape
    tets = b["tets"].astype(np.int64)
    nb, per_vert = 4, 2
    folder = "/constraint_projection/FOM/tets_deformation_gradient"
    cfg = {"object": {"experiment_dir": "pd_out/", "mesh": "bar", "volumetric": True, "experiment": "bar_defgrad/", "snap_format": ".off"},
           "vertexPos_bases": NO_POS,
           "constraintProj_bases": {
               "computeState": {"compute": True, "run_main": True, "testingComputations": "_Release"},
               "constraintType": {"name": "tets_deformation_gradient", "elements": "_tets", "p_snaps_folder": folder,
                                  "pos_snaps_folder": "/position_snapshots/FOM", "geom_pos_snaps_folder": "/position_snapshots/geom/",
                                  "assembly_file_name": "assembly_ST.npz", "assembly_key": "tets_deformation_gradient",
                                  "snaps_pattern_full_p": "/tets_deformation_gradient_p.npz", "constrained_elements": "", "rowSize": 3},
               "snapshots": {"max_numFrames": F, "numFrames": F, "frame_increment": 1, "read_all_from_first": True,
                             "processed_snapshots_file": "/processed_snapshots/", "preAlignement": "_noAlignement",
                             "processed_snapshots_ready": False, "reduced_snaps_available": False},
               "basis_type": "pca_blocks", "interpolation_type": "geom", "desired_num_components": nb, "numComponents_verts": 5,
               "bases_res_tol": 1e-12, "dim": 3, "constraintsDim": 1, "max_element_per_geom_vert": per_vert, "rest_shape": "first",
               "massWeighted": "_Volkwein", "standarized": "_Standarized", "supported": "_Global", "orthogonalized": "_Orthogonalized",
               "store_sing_val": True, "store_to_files": True, "run_tests": False, "visualize_elements_at_bases_num": 30,
               "visualize_geom_elements": 528}}
    param = _params(cfg)
    os.makedirs(os.path.dirname(param.constProj_input_snapshots_pattern), exist_ok=True)
    np.savez(param.constProj_input_snapshots_pattern, **{str(i): b["snaps"][i] for i in range(F)})
    St = sparse.csr_matrix((g["St_vals"], g["St_colidx"], g["St_rowptr"]), shape=tuple(g["St_shape"]))
    np.savez(param.constProj_weightedSt, tets_deformation_gradient=St)          # pickled scipy sparse, as the simulator stores it
    os.makedirs(os.path.dirname(param.constProj_masses_file), exist_ok=True)
    hostio.write_mass_bin(param.constProj_masses_file, b["mass"])
    _write_medit(param.tet_mesh_file, m["bar_verts"], tets)

    reference_main.main(param)

    z = np.load(os.path.join(param.constProj_output_directory, "components_interpol_alphas_interpol_verts_interpol_alpha_ranges.npz"),
                allow_pickle=True)
    comps, alphas, Pt, verts, ranges = (z[k] for k in ("components", "interpol_alphas", "Pt", "interpol_verts", "interpol_alpha_ranges"))
    assert comps.shape == (3 * nb, ep, 3) and len(verts) == nb and len(ranges) == nb
    assert np.array_equal(Pt, (alphas[:, None] * 3 + np.arange(3)[None]).ravel()) and len(np.unique(alphas)) == len(alphas)
    steps = np.diff(np.concatenate([[0], ranges]))
    assert ranges[-1] == len(alphas) and np.all(steps >= 0) and np.all(steps <= per_vert)
    start = 0
    for v, stop in zip(verts, ranges):
        assert all(v in tets[a] for a in alphas[start:stop]), "every new element is incident to the vertex picked in that step"
        start = stop
    # and the emitted file is what the simulator-side precompute consumes (Simulators.py:157-218)
    pm = port.prepare_reduced_group(z, nb, 3, St, "geom_pca_blocks_withSt")[2]
    assert pm.shape == (St.shape[0], 3 * nb, 3) and np.isfinite(pm).all()


def test_reference_main_imports_off_sequence(reference_main, monkeypatch):
    """main.py:28-63, the branch that runs when the aligned animation files do not exist yet: OFF sequence -> animation file
    -> Procrustes alignment -> bases.  `convert_sequence_to_hdf5` / `load_off` / `align` are the names main.py imports from
    utils/process.py (h5py-bound); the maintainer's import swap of INTEGRATION.md points them at hostio.  Result against
    the oracle pipeline run on the same raw frames."""
    from animsnapbases_b200 import synth
    monkeypatch.setattr(reference_main, "convert_sequence_to_hdf5", hostio.convert_sequence_to_hdf5)
    monkeypatch.setattr(reference_main, "load_off", hostio.load_off)
    monkeypatch.setattr(reference_main, "align", hostio.align)
    m = synth.golden_mesh("sphere")               # a genuinely 3-D shape: rigid Procrustes of a near-planar one is ill-posed
    verts0, tris = m["verts"], m["tris"]
    F, K = 12, 5
    rng = np.random.default_rng(21)
    raw = (2.5 * synth.bump_animation(verts0, 2 * F, n_bumps=4, sigma=0.2, seed=6) + np.array([1.0, 2.0, 3.0])
           + 0.05 * rng.standard_normal((2 * F, 1, 3))).astype(np.float32)          # unnormalised, drifting frames
    cfg = {"object": {"mesh": "grid", "volumetric": False, "experiment": "_bumps", "experiment_dir": "data_in/", "snap_format": ".off"},
           "vertexPos_bases": {
               "computeState": {"compute": True, "testingComputations": "_Release"},
               "snapshots": {"max_numFrames": 2 * F, "frame_increment": 1, "numFrames": F, "format": ".off", "snaps_folder": "fom",
                             "read_all_from_first": True, "anims_folder": "FOM_animations_h5", "preAlignement": "_alignedRigid",
                             "anim_folder_ready": False, "visualize_aligned_animations": False, "reduced_snaps_available": False},
               "dim": 3, "rest_shape": "first", "massWeighted": "_Volkwein", "standarized": "_Standarized",
               "orthogonalized": "_nonOrthogonalized", "support": {"min_dist": 0.1, "max_dist": 0.25},
               "pca": {"compute": True, "numComponents": K, "supported": "_Global", "store_sing_val": True},
               "splocs": {"compute": False, "max_itrs": 2, "admm_num_itrs": 2, "lambda": 2, "rho": 10.0},
               "store": True, "run_tests": False, "visualize": False},
           "constraintProj_bases": NO_CON}
    param = _params(cfg)
    os.makedirs(os.path.dirname(param.input_snapshots_pattern), exist_ok=True)
    for f, fr in enumerate(raw):
        with open(param.input_snapshots_files_name + "%d.off" % f, "w") as fh:
            fh.write("OFF\n%d %d 0\n" % (len(fr), len(tris)))
            fh.writelines("%.9g %.9g %.9g\n" % tuple(p) for p in fr)
            fh.writelines("3 %d %d %d\n" % tuple(t) for t in tris)
    # oracle: the same frames through the restated ingest (first F files, increment 1), rigid alignment, masses of frame 0
    pre, _, _ = port.preprocess_unit_cube(raw[:F])
    aligned = port.align_frames(pre, rigid=True)
    mass = synth.lumped_vertex_masses(aligned[0].astype(float), tris)
    os.makedirs(os.path.dirname(param.vertPos_masses_file), exist_ok=True)
    hostio.write_mass_bin(param.vertPos_masses_file, mass)

    reference_main.main(param)

    got, got_tris = hostio.read_animation(os.path.join(param.aligned_snapshots_directory, param.train_aligned_snapshots_animation_file))
    assert got.dtype == np.float32 and got.shape == aligned.shape and np.array_equal(got_tris, tris)
    assert np.abs(got - aligned).max() < 2e-6, "float32 Procrustes: device reductions vs LAPACK"
    test_anim, _ = hostio.read_animation(os.path.join(param.aligned_snapshots_directory, param.test_aligned_snapshots_animation_file))
    assert test_anim.shape == (F, len(verts0), 3), "testing frames: every (increment + train_test_jump)-th file (main.py:49-51)"
    prep = port.prepare_snapshots(got, mass, "first")
    C, W, meas, picked, _ = port.extract_k_components(prep["snapTensor"], K, "global")
    post = port.post_process_pos(C, prep["mean"], prep["pre_scale_factor"], prep["invMassL"], True, False, True)
    raw_bin = open(os.path.join(param.vertPos_output_directory, "q_pos_F%dK%d.bin" % (F, K)), "rb").read()
    B = np.frombuffer(raw_bin[8:], "<f8").reshape(3, K, len(verts0)).transpose(1, 2, 0)
    # global support: the sign of each (w, c) pair is free and un-standardising adds the mean, so compare per component
    # against both signs of the oracle's c_k
    flipped = port.post_process_pos(-C, prep["mean"], prep["pre_scale_factor"], prep["invMassL"], True, False, True)
    for k in range(K):
        assert min(rel(B[k], post[k]), rel(B[k], flipped[k])) < 1e-7, k
