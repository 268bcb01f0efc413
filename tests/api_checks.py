"""Drop-in protocol checks: drive animsnapbases_b200.snapbases.* exactly like the reference's main.py:71-100 / :106-142
drives snapbases.* and compare attributes and written files with the golden outputs of the reference."""
import os
import types

import numpy as np

from conftest import golden, rel
from animsnapbases_b200 import hostio, runtime
from oracle import port


def _inputs_once(tmp, write, tag):
    """Input fixtures of a protocol check: written by rank 0 only (several ranks share the directory under torchrun), once."""
    _, comm = runtime.get()
    marker = os.path.join(str(tmp), ".inputs_written_" + tag)
    if comm.rank == 0 and not os.path.exists(marker):
        write()
        open(marker, "w").close()
    comm.barrier()


def pos_param(tmp, g, bases_type="PCA", support="local", orthogonal=False):
    d = str(tmp)

    def write():
        hostio.write_animation(os.path.join(d, "train.h5"), g["anim"], g["tris"])
        hostio.write_animation(os.path.join(d, "test.h5"), g["anim"][::2], g["tris"])
        hostio.write_mass_bin(os.path.join(d, "mass.bin"), g["mass"])
    _inputs_once(tmp, write, "pos")
    it, admm, lam, rho = g["splocs"]
    return types.SimpleNamespace(
        vertPos_bases_type=bases_type, aligned_snapshots_directory=d, train_aligned_snapshots_animation_file="train.h5",
        test_aligned_snapshots_animation_file="test.h5", vertPos_rest_shape="first",
        vertPos_masses_file=os.path.join(d, "mass.bin"), tet_mesh_file="", q_standarize=True, q_massWeight=True,
        vertPos_numComponents=int(g["K"]), q_support=support, store_vertPos_PCA_sing_val=True,
        vertPos_smooth_min_dist=0.1, vertPos_smooth_max_dist=0.25, splocs_max_itrs=int(it), splocs_admm_num_itrs=int(admm),
        splocs_lambda=float(lam), splocs_rho=float(rho), q_orthogonal=orthogonal, vertPos_output_directory=d, name="fixture")


def check_pos_protocol(ops, comm, tmp, name="pos_grid", extended=False):
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
        if extended:        # headless figure-script table (generate_figures/pos_reduction_tests.py:62-92)
            from animsnapbases_b200 import reduction_tests
            sv = reduction_tests.pos_pca_tests(bases, bases.param)
            tab = open(os.path.join(str(tmp), "fixture_posBases_xyz_fullBasesRange_K_singVals.csv")).read().splitlines()
            assert tab[0] == "row,x,y,z" and len(tab) == K + 1 and tab[1].split(",")[0] == "1"
            assert np.allclose([float(v) for v in tab[K].split(",")[1:]], sv[K - 1])
            assert np.allclose(sv, port.bases_sing_vals(bases.comps), atol=1e-7)
    finally:
        os.chdir(cwd)
        runtime.install(None, None)
    return True


def constraints_param(tmp, b, k_desired):
    d = str(tmp)
    snaps = b["snaps"]

    def write():
        np.savez(os.path.join(d, "bar_p.npz"), **{str(i): snaps[i] for i in range(snaps.shape[0])})
        hostio.write_mass_bin(os.path.join(d, "elem_mass.bin"), b["mass"])
    _inputs_once(tmp, write, "constraints")
    return types.SimpleNamespace(
        constProj_input_snapshots_pattern=os.path.join(d, "bar_p.npz"), constProj_rest_shape="first", constProj_dim=3,
        constProj_masses_file=os.path.join(d, "elem_mass.bin"), constProj_numFrames=snaps.shape[0], constProj_p_size=3,
        constProj_element_type="_tets", tet_mesh_file="", tri_mesh_file="", constProj_output_directory=d,
        constProj_frame_increment=1, constProj_train_test_jump=1, constProj_massWeight=True, constProj_standarize=True,
        constProj_orthogonal=True, constProj_bases_interpolation_type="deim", constProj_support="global",
        constProj_store_sing_val=True, constProj_weightedSt=os.path.join(d, "missing_ST.npz"), costProj_St_key="St",
        constProj_basis_type="pod_vectorized", deim_desired_num_components=k_desired, name="fixture", constProj_name="tets")


def check_consumer(tmp, K, ep):
    """the emitted `.npz` goes through the consumer-side precompute (projective_dynamics/Simulators.py:157-218)
    and the reduced term (:366-399) -- mirror on the installed ops against the CPU restatement reading the same file."""
    import shutil
    from scipy import sparse
    from animsnapbases_b200 import reduced_group
    g = golden("bar_geom")
    St = sparse.csr_matrix((g["St_vals"], g["St_colidx"], g["St_rowptr"]), shape=tuple(g["St_shape"]))
    group, name = "tets_deformation_gradient", "components_interpol_alphas_interpol_verts_interpol_alpha_ranges.npz"
    os.makedirs(os.path.join(str(tmp), group), exist_ok=True)
    shutil.copy(os.path.join(str(tmp), name), os.path.join(str(tmp), group, name))
    m = K - 2
    grp = reduced_group.prepare_reduced_group(True, True, group, m, 3, St, str(tmp), name, reduction_type="deim_pod_vectorized")
    alpha, Pt, pm, solvers = grp
    local = np.load(os.path.join(str(tmp), group, name))
    r_alpha, r_Pt, r_pm, r_normal, r_PtVT = port.prepare_reduced_group(local, m, 3, St, "deim_pod_vectorized")
    assert np.array_equal(alpha, r_alpha) and np.array_equal(Pt, r_Pt) and pm.shape == (St.shape[0], m, 3)
    assert rel(pm, r_pm) < 1e-13
    P = np.random.default_rng(5).standard_normal((ep, 3))
    from scipy.linalg import lu_solve
    for d in range(3):
        assert rel(solvers[d][1], r_PtVT[:, :, d]) < 1e-15
        want = np.linalg.solve(r_normal[d], r_PtVT[:, :, d] @ P[r_Pt, d])
        assert rel(lu_solve(solvers[d][0], solvers[d][1] @ P[Pt, d]), want) < 1e-9     # the reference's own use of solver_list
    assert rel(grp.reduced_term(P[Pt]), port.group_reduced_term(P[r_Pt], r_pm, r_normal, r_PtVT)) < 1e-9
    assert reduced_group.prepare_reduced_group(False, True, group, m, 3, St, str(tmp), name) == (None, [], None, [])
    try:
        reduced_group.prepare_reduced_group(True, True, group, m, 3, St, str(tmp), name, reduction_type="nope")
        raise RuntimeError("expected ValueError")
    except ValueError:
        pass
    return True


def check_constraints_protocol(ops, comm, tmp, extended=False):
    """main.py:106-142 on the bar fixture: pod_vectorized -> post-process (QR, un-weight) -> deim -> writers."""
    from animsnapbases_b200.snapbases.constraintsComponents import constraintsComponents
    runtime.install(ops, comm)
    cwd = os.getcwd()
    os.chdir(str(tmp))
    try:
        b = golden("bar")
        K = 10
        bases = constraintsComponents(constraints_param(tmp, b, K))
        bases.nonlinearSnapshots.config()
        bases.config(fileNameBases="p_nl_", fileName_geom_points="p_nl_interpol_points_")
        bases.nonlinearSnapshots.snapshots_prepare()
        snaps = bases.nonlinearSnapshots
        F, ep, _ = b["snaps"].shape
        assert snaps.num_constained_elements == ep // 3 and snaps.frs == F
        assert snaps.test_snapTensor.shape == (F - 1, ep, 3)
        assert rel(snaps.snapTensor, b["X"]) < 1e-14 and rel(snaps.mean, b["mean"]) < 1e-15
        assert abs(snaps.pre_scale_factor - float(b["scale"])) <= 1e-13 * float(b["scale"])
        bases.compute_components_store_singvalues()
        assert bases.numComp == K and bases.comps.shape == (K, ep, 3)
        ref = b["pod_sing"]
        assert np.all(np.abs(bases.singular_values[:K] / ref[:K] - 1) < 1e-9)
        rows = open(os.path.join(str(tmp), "fixture_tets_constrprojBases_pcaExtraction_singValues.csv")).read().splitlines()
        assert rows[0] == "component,singVal" and len(rows) == F + 1 and rows[1].split(",")[0] == "1"
        sg = np.sign((bases.comps * b["pod_comps"][:K]).sum(axis=(1, 2)))
        assert rel(bases.comps * sg[:, None, None], b["pod_comps"][:K]) < 1e-7
        # SVD modes are defined up to sign (north_star: "bases equal up to per-column sign"), but post-processing adds the mean
        # to every mode (:422), so everything downstream depends on LAPACK's sign choice: adopt the reference's signs here.
        import torch
        bases._C *= torch.as_tensor(sg, dtype=bases._C.dtype, device=bases._C.device)[:, None, None]
        bases._host = {}
        bases.post_process_components()
        post = b["post_comps"][:K]                  # economic QR: the first K columns do not depend on the later ones
        for d in range(3):
            s = np.sign((bases.comps[:, :, d] * post[:, :, d]).sum(axis=1))
            assert rel(bases.comps[:, :, d] * s[:, None], post[:, :, d]) < 1e-7
        assert rel(snaps.snapTensor, b["post_snap"]) < 1e-12, "post-processing also un-standardises the snapshots (:426-443)"
        assert bases.is_utmu_orthogonal() == [True, True, True], "QR then M^-1/2 scaling: U^T M U = I per dimension (:450-459)"
        Rt = np.swapaxes(b["X"], 0, 1)                                       # (ep, F, 3) as indxLargestDeformation expects
        assert bases.indxLargestDeformation(Rt, 3, ep // 3) == port.index_largest_deformation(Rt, 3, ep // 3)
        bases.deim()
        assert np.array_equal(bases.geom_Pt, b["deim_Pt"][:K]), "DEIM indices must be bit-exact"
        assert np.array_equal(bases.geom_alpha, b["deim_alpha"][:K])
        assert np.array_equal(bases.geom_alpha_ranges, np.arange(1, K + 1))
        bases.store_components_n_interpol_points()
        z = np.load(os.path.join(str(tmp), "components_interpol_alphas_interpol_verts_interpol_alpha_ranges.npz"),
                    allow_pickle=True)
        assert set(z.files) == {"components", "interpol_alphas", "Pt", "interpol_verts", "interpol_alpha_ranges"}
        assert np.array_equal(z["components"], bases.comps) and np.array_equal(z["Pt"], bases.geom_Pt)
        if extended:
            assert check_consumer(tmp, K, ep)
        # plain DEIM leaves geom_interpol_verts empty, so the reference's third writer asserts (utils/utils.py:86) after the
        # bases and interpolation-point files of the first k were written; same behaviour here.  With vertices present all
        # three files are written for every k.
        try:
            bases.store_components_gradually_to_files(2, 3, 1, ".bin")
            raise RuntimeError("expected the reference's AssertionError")
        except AssertionError:
            pass
        assert os.path.exists(os.path.join(str(tmp), "p_nl_F%dKp%d.bin" % (F, 6)))
        bases.geom_interpol_verts = np.arange(K)
        bases.store_components_gradually_to_files(2, 3, 1, ".bin")
        for k in (2, 3):
            got = open(os.path.join(str(tmp), "p_nl_F%dKp%d.bin" % (F, 3 * k)), "rb").read()
            assert got == port.components_bin_bytes(bases.comps[:3 * k])
            pts = hostio.load_vector_values(os.path.join(str(tmp), "p_nl_interpol_points_F%dK%d_points%d.bin" % (F, k, k)))
            assert np.array_equal(pts, bases.geom_alpha[:k].astype(float))
        s = bases.test_basesSingVals()
        assert s.shape == (K, 3) and np.allclose(s, port.bases_sing_vals(bases.comps), atol=1e-7)
        # reconstruction from r = K interpolation elements (generate_figures/nl_reduction_tests.py:117-150 -> :489-521)
        rec = bases.geom_constructed(K, "train")
        assert rec.shape == (F, ep, 3)
        err = bases.frobenius_error(snaps.snapTensor, rec)
        assert abs(err - float(b["deim_rec10_err"][0])) < 1e-6 * float(b["deim_rec10_err"][0])
        if extended:
            # headless figure-script numbers (generate_figures/nl_reduction_tests.py:120-225) against the host statics
            from animsnapbases_b200 import reduction_tests
            train, test = reduction_tests.nl_geom_tests(bases, steps=4)
            assert [r[0] for r in train] == list(range(1, K + 1, 4)) and len(test) == len(train)
            for rows, case, frames in ((train, "train", snaps.snapTensor), (test, "test", snaps.test_snapTensor)):
                for r, fro, mx, rx, ry, rz, r3 in rows:
                    got = bases.geom_constructed(r, case)
                    want = [port.frobenius_error(frames, got), port.max_pointwise_error(frames, got)] + \
                        list(port.relative_error_per_component(frames, got))
                    assert np.allclose([fro, mx, rx, ry, rz], want, rtol=1e-9), (case, r)
                    assert abs(r3 - (rx + ry + rz) / 3) < 1e-15
            stem = os.path.join(str(tmp), "fixture_tets_deim_pod_vectorized")
            lines = open(stem + "_train_convergence_tests.csv").read().splitlines()
            assert lines[0] == "numPoints,fro_error,max_err,relative_errors_x,relative_errors_y,relative_errors_z,relative3d"
            assert len(lines) == 1 + len(train) and os.path.exists(stem + "_test_convergence_tests.csv")
            assert open(stem + "_num_interpol_elemnets.csv").read().splitlines()[:2] == ["numPoints,num_elements", "1,1"]
            sv = reduction_tests.nl_pca_tests(bases)
            tab = open(os.path.join(str(tmp), "fixture_tets_constrprojBases_xyz_fullBasesRange_Kp_singVals.csv")).read().splitlines()
            assert tab[0] == "i,x,y,z" and len(tab) == 1 + 2 * K and tab[1].split(",")[0] == "1" and sv.shape == (K, 3)
            assert reduction_tests.nl_postprocess_tests(bases) == [True, True, True]
        # geometric DEIM through the differential operator (main.py:132-134): picks are a prefix of the reference's
        from scipy import sparse
        g = golden("bar_geom")
        bases.St = sparse.csr_matrix((g["St_vals"], g["St_colidx"], g["St_rowptr"]), shape=tuple(g["St_shape"]))
        snaps.tets, snaps.ele_type = b["tets"].astype(np.int64), "_tets"
        bases.param.geom_ele_per_vert = 1
        bases.numComp = K // 3                       # blocks of p = 3 rows
        bases.geom_block_form_utilizing_differential_operator(True)
        nb = K // 3
        assert np.array_equal(bases.geom_alpha, g["pod_pos1_alpha"][:nb]) and np.array_equal(bases.geom_Pt, g["pod_pos1_Pt"][:3 * nb])
        assert np.array_equal(bases.geom_interpol_verts, g["pod_pos1_verts"][:nb])
        bases.geom_block_form_utilizing_differential_operator(False)
        assert np.array_equal(bases.geom_alpha, g["pod_con_alpha"][:nb])
    finally:
        os.chdir(cwd)
        runtime.install(None, None)
    return True


def check_verts_geom_protocol(ops, comm, tmp):
    """constraintsComponents.geom_block_form_utilizing_differential_operator(True) of the mirror class for '_verts'
    constraints of type verts_bending: reads the constrained-vertex file the reference reads (:636-638) and returns the
    reference's picks (tests/golden/verts_geom.npz)."""
    import types
    from scipy import sparse
    from animsnapbases_b200 import engine
    from animsnapbases_b200.snapbases.constraintsComponents import constraintsComponents
    from animsnapbases_b200.snapbases.nonlinear_snapshots import nonlinearSnapshots
    runtime.install(ops, comm)
    cwd = os.getcwd()
    os.chdir(str(tmp))                     # log_time("") writes function_timings.txt into the CWD
    try:
        g = golden("verts_geom")
        comps, idx = g["bend_comps"], g["bend_idx"]
        f = os.path.join(str(tmp), "verts_bending_constrained_indices.npz")
        np.savez(f, indices=idx)
        ns = object.__new__(nonlinearSnapshots)
        ns.constraintsSize, ns.num_constained_elements, ns.dim = 1, len(idx), 3
        ns.ele_type, ns.tris, ns.tets, ns.edges = "_verts", g["tris"], None, None
        ns.shard = engine.Shard.make(len(idx), comm.world, comm.rank)
        cc = object.__new__(constraintsComponents)
        cc.param = types.SimpleNamespace(geom_ele_per_vert=3, constProj_input_snaps_constrained_elements=f,
                                         constProj_snapshots_type="verts_bending")
        cc.nonlinearSnapshots, cc.numComp, cc.geom_interpol_verts = ns, comps.shape[0], []
        cc.St = sparse.csr_matrix((g["bend_St_vals"], g["bend_St_colidx"], g["bend_St_rowptr"]), shape=tuple(g["bend_St_shape"]))
        rows = slice(ns.shard.row0, ns.shard.row0 + ns.shard.n)
        cc._C = ops.ingest(ops.to_device(np.ascontiguousarray(comps[:, rows])), None, ns.shard.ld)
        cc.geom_block_form_utilizing_differential_operator(True)
        assert np.array_equal(cc.geom_Pt, g["bend3_Pt"]) and np.array_equal(cc.geom_alpha, g["bend3_alpha"])
        assert np.array_equal(cc.geom_alpha_ranges, g["bend3_ranges"]) and np.array_equal(cc.geom_interpol_verts, g["bend3_verts"])
        assert np.array_equal(cc.constrianed_verts, idx)
    finally:
        os.chdir(cwd)
        runtime.install(None, None)
    return True
