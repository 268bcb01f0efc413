"""Drop-in protocol checks: drive animsnapbases_b200.snapbases.* exactly like the reference's main.py:71-100 / :106-142
drives snapbases.* and compare attributes and written files with the golden outputs of the reference."""
import os
import types

import numpy as np

from conftest import golden, rel
from animsnapbases_b200 import hostio, runtime
from oracle import port


def pos_param(tmp, g, bases_type="PCA", support="local", orthogonal=False):
    d = str(tmp)
    hostio.write_animation(os.path.join(d, "train.h5"), g["anim"], g["tris"])
    hostio.write_animation(os.path.join(d, "test.h5"), g["anim"][::2], g["tris"])
    hostio.write_mass_bin(os.path.join(d, "mass.bin"), g["mass"])
    it, admm, lam, rho = g["splocs"]
    return types.SimpleNamespace(
        vertPos_bases_type=bases_type, aligned_snapshots_directory=d, train_aligned_snapshots_animation_file="train.h5",
        test_aligned_snapshots_animation_file="test.h5", vertPos_rest_shape="first",
        vertPos_masses_file=os.path.join(d, "mass.bin"), tet_mesh_file="", q_standarize=True, q_massWeight=True,
        vertPos_numComponents=int(g["K"]), q_support=support, store_vertPos_PCA_sing_val=True,
        vertPos_smooth_min_dist=0.1, vertPos_smooth_max_dist=0.25, splocs_max_itrs=int(it), splocs_admm_num_itrs=int(admm),
        splocs_lambda=float(lam), splocs_rho=float(rho), q_orthogonal=orthogonal, vertPos_output_directory=d, name="fixture")


def check_pos_protocol(ops, comm, tmp, name="pos_grid"):
    from animsnapbases_b200.snapbases.posComponents import posComponents
    runtime.install(ops, comm)
    cwd = os.getcwd()
    os.chdir(str(tmp))                     # log_time("") writes function_timings.txt into the CWD, like the reference
    try:
        g = golden(name)
        K = int(g["K"])
        # ---- 'SPLOCS' mode, local support
        bases = posComponents(pos_param(tmp, g, "SPLOCS", "local"))
        snaps = bases.pos_snapshots
        assert snaps.frs == g["anim"].shape[0] and snaps.nVerts == g["anim"].shape[1]
        assert rel(snaps.snapTensor, g["X"]) < 1e-14 and rel(snaps.mean, g["mean"]) < 1e-15
        assert abs(snaps.pre_scale_factor - float(g["scale"])) <= 1e-13 * float(g["scale"])
        assert rel(snaps.compute_geodesic_distance(int(g["geo_ids"][1])), g["geo_phi"][1]) < 1e-8
        bases.compute_components_store_singvalues()
        assert np.array_equal(bases.picked_vertices, g["picked_local"])
        assert bases.comps.shape == (K, snaps.nVerts, 3) and bases.weigs.shape == (snaps.frs, K)
        # reference quirk: the SPLOCS result is not written back -- comps stay the deflation result
        assert rel(bases.comps, g["comps_local"]) < 1e-7 and rel(bases.weigs, g["weigs_local"]) < 1e-9
        assert rel(bases.splocs_comps, g["splocs_C"]) < 1e-6 and rel(bases.splocs_weigs, g["splocs_W"]) < 1e-6
        assert np.allclose(bases.measures_at_largeDeforVerts, g["meas_local"], rtol=1e-7)
        rows = open(os.path.join(str(tmp), "fixture_posBases_pcaExtraction_singValues_errorNorm.csv")).read().splitlines()
        assert rows[0] == "component,singVal,norm_R" and len(rows) == K + 1 and rows[1].split(",")[0] == "0"
        assert os.path.exists("function_timings.txt")
        # ---- 'PCA' mode, global support, full post-processing + writers
        bases = posComponents(pos_param(tmp, g, "PCA", "global"))
        bases.compute_components_store_singvalues()
        W = bases.weigs
        sg = np.sign((W * g["weigs_global"]).sum(axis=0))
        assert rel(bases.comps * sg[:, None, None], g["comps_global"]) < 1e-9
        bases.post_process_components()
        keep = sg > 0
        assert rel(bases.comps[keep], g["post_global"][keep]) < 1e-9
        bases.store_components_to_files(1, K, 3, ".bin")
        for k in range(1, K + 1, 3):
            got = open(os.path.join(str(tmp), "q_pos_F%dK%d.bin" % (snaps.frs, k)), "rb").read()
            assert got == port.components_bin_bytes(bases.comps[:k]), "bin layout: <i N><i 3K> + d,k,i order"
        bases.store_components_to_files(K, K, 1, ".npy")
        assert np.array_equal(np.load(os.path.join(str(tmp), "q_pos_%dK%d.npy" % (snaps.frs, K))), bases.comps)
        bases.store_animations(str(tmp))
        s = bases.test_basesSingVals()
        ref_s = port.bases_sing_vals(bases.comps)
        assert s.shape == (K, 3) and np.allclose(s, ref_s, atol=1e-7)
        fro, mx, rx, ry, rz = bases.test_convergence(1, K, 4)
        assert len(fro) == len(range(1, K + 1, 4)) and all(np.isfinite(fro))
        # ---- orthogonalised variant: U^T M U = I is asserted inside post_process_components
        bases = posComponents(pos_param(tmp, g, "PCA", "global", orthogonal=True))
        bases.compute_components_store_singvalues()
        bases.post_process_components()
        Mu = bases.comps[:, :, 0] * g["mass"][None]
        assert np.allclose(bases.comps[:, :, 0] @ Mu.T, np.eye(K), atol=1e-8)
    finally:
        os.chdir(cwd)
        runtime.install(None, None)
    return True
