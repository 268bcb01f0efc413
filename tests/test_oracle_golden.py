"""CPU suite: the oracle restatement (oracle/port.py) against the golden vectors produced by the reference itself
(tests/golden/*.npz, oracle/make_golden.py).  When /root/reference is mounted the port is also checked against the
live reference on a fresh seed.  LAPACK results may differ in the last bits between hosts, hence rtol 1e-9 (not ==)."""
import numpy as np
import pytest

from conftest import golden, rel
from oracle import port

RT = 1e-9


def _pos(name):
    g = golden(name)
    prep = port.prepare_snapshots(g["anim"], g["mass"], "first")
    geo = port.GeodesicHeat(g["anim"][0].astype(float), g["tris"].astype(np.int64))
    return g, prep, geo


@pytest.mark.parametrize("name", ["pos_grid", "pos_sphere"])
def test_port_positions(name):
    g, prep, geo = _pos(name)
    K = int(g["K"])
    assert np.array_equal(prep["snapTensor"], g["X"]) and prep["pre_scale_factor"] == float(g["scale"])
    for i, gid in enumerate(g["geo_ids"]):
        assert rel(geo(int(gid)), g["geo_phi"][i]) < 1e-8
    C, W, meas, picked, R = port.extract_k_components(prep["snapTensor"], K, "local", geo)
    assert np.array_equal(picked, g["picked_local"])
    assert rel(C, g["comps_local"]) < RT and rel(W, g["weigs_local"]) < RT and rel(meas, g["meas_local"]) < RT
    it, admm, lam, rho = g["splocs"]
    Wf, Cf, hist, cent = port.splocs_glob_optimization(prep["snapTensor"], C, W, R, geo, int(it), int(admm), lam, rho)
    assert rel(Wf, g["splocs_W"]) < RT and rel(Cf, g["splocs_C"]) < RT and np.array_equal(cent, g["splocs_centres"])
    Cg, Wg, mg, pg, _ = port.extract_k_components(prep["snapTensor"], K, "global")
    assert np.array_equal(pg, g["picked_global"])
    sg = np.sign((Wg * g["weigs_global"]).sum(axis=0))
    assert rel(Cg * sg[:, None, None], g["comps_global"]) < RT
    post = port.post_process_pos(Cg, prep["mean"], prep["pre_scale_factor"], prep["invMassL"], True, False, True)
    assert rel(post[sg > 0], g["post_global"][sg > 0]) < RT


def test_port_bar():
    b = golden("bar")
    prep = port.prepare_snapshots(b["snaps"], b["mass"], "first")
    assert np.array_equal(prep["snapTensor"], b["X"])
    C, S = port.pod_vectorized(prep["snapTensor"], 200)
    assert np.allclose(S[:-1], b["pod_sing"][:-1], rtol=1e-9, atol=1e-12 * S[0])
    d = port.deim(b["post_comps"], 3)
    assert np.array_equal(d["Pt"], b["deim_Pt"]) and np.array_equal(d["alpha"], b["deim_alpha"])
    rec = port.deim_reconstruct(b["post_comps"], b["post_snap"], d["alpha"][:10], 10)
    assert abs(port.frobenius_error(b["post_snap"], rec) - b["deim_rec10_err"][0]) < 1e-7 * b["deim_rec10_err"][0]
    Cb, Wb, mb, sidx, sp_ = port.nonlinearity_bases_blocks(b["X"], len(b["tets"]), 3, 5)
    assert np.array_equal(sidx, b["blk_points"]) and np.array_equal(sp_, b["blk_rows"])
    sg = np.sign((Wb * b["blk_weigs"]).sum(axis=0))
    assert rel(Cb * sg[:, None, None], b["blk_comps"]) < RT
    db = port.deim_blocks(b["blk_post"], 3)
    assert np.array_equal(db["Pt"], b["blk_deim_Pt"])


def test_port_align_and_writers():
    a = golden("align")
    for rigid in (True, False):
        out = port.align_frames(a["frames"], rigid)
        assert np.abs(out - a[f"aligned_rigid{int(rigid)}"]).max() <= 1e-6      # float32 LAPACK may differ across hosts
    pre, mean, scale = port.preprocess_unit_cube(a["frames"])
    assert np.allclose(pre, a["pre"]) and np.isclose(scale, a["pre_scale"][0])
    w = golden("writers")
    assert port.components_bin_bytes(w["B"]) == w["bin"].tobytes()
    assert port.points_bin_bytes(w["pts"]) == w["pts_bin"].tobytes()


def test_port_geometric_deim_golden():
    """oracle/port.py::deim_geom against the picks of the unmodified reference (constraintsComponents.py:619-731)."""
    from scipy import sparse
    b, g = golden("bar"), golden("bar_geom")
    tets = b["tets"].astype(np.int64)
    St = sparse.csr_matrix((g["St_vals"], g["St_colidx"], g["St_rowptr"]), shape=tuple(g["St_shape"]))
    for name, comps, pos, per_vert in (("blk_pos2", b["blk_post"], True, 2), ("pod_pos1", b["post_comps"][:24], True, 1),
                                       ("pod_pos3", b["post_comps"][:24], True, 3), ("pod_con", b["post_comps"][:24], False, 1)):
        out = port.deim_geom(comps, 3, len(tets), tets, St if pos else None, per_vert)
        assert np.array_equal(out["Pt"], g[name + "_Pt"]) and np.array_equal(out["alpha"], g[name + "_alpha"])
        assert np.array_equal(out["alpha_ranges"], g[name + "_ranges"]) and np.array_equal(out["verts"], g[name + "_verts"])


def test_port_verts_geometric_deim_golden():
    """oracle/port.py::deim_geom, '_verts' flavour, against the picks of the unmodified reference (:684-702)."""
    from scipy import sparse
    g = golden("verts_geom")
    for gname in ("bend", "all"):
        St = sparse.csr_matrix((g[gname + "_St_vals"], g[gname + "_St_colidx"], g[gname + "_St_rowptr"]),
                               shape=tuple(g[gname + "_St_shape"]))
        for per_vert in (1, 3):
            out = port.deim_geom(g[gname + "_comps"], 1, len(g[gname + "_idx"]), None, St, per_vert, star_faces=g["tris"],
                                 constrained_verts=g[gname + "_idx"] if gname == "bend" else None)
            key = f"{gname}{per_vert}"
            assert np.array_equal(out["Pt"], g[key + "_Pt"]) and np.array_equal(out["alpha"], g[key + "_alpha"])
            assert np.array_equal(out["alpha_ranges"], g[key + "_ranges"]) and np.array_equal(out["verts"], g[key + "_verts"])


def test_port_reduced_group_golden():
    """oracle/port.py::prepare_reduced_group / group_reduced_term against the outputs of the unmodified consumer code
    (projective_dynamics/Simulators.py:157-218, :366-399) stored in tests/golden/bar_reduced.npz."""
    from scipy import sparse
    b, g, r = golden("bar"), golden("bar_geom"), golden("bar_reduced")
    St = sparse.csr_matrix((g["St_vals"], g["St_colidx"], g["St_rowptr"]), shape=tuple(g["St_shape"]))
    fx = {**{k: b[k] for k in b.files}, **{k: g[k] for k in g.files}}
    cases = {"pod": ("deim_pod_vectorized", "post_comps", ("deim_Pt", "deim_alpha", "deim_ranges"), 10),
             "blk": ("geom_pca_blocks_withSt", "blk_post", ("blk_pos2_Pt", "blk_pos2_alpha", "blk_pos2_ranges"), 4)}
    for name, (rtype, ckey, (kpt, kal, krg), ncomp) in cases.items():
        local = dict(components=fx[ckey], interpol_alphas=fx[kal], Pt=fx[kpt], interpol_alpha_ranges=fx[krg])
        alpha, Pt, pm, normal, PtVT = port.prepare_reduced_group(local, ncomp, 3, St, rtype)
        assert np.array_equal(alpha, r[name + "_alpha"]) and np.array_equal(Pt, r[name + "_Pt"])
        assert rel(pm, r[name + "_StV"]) < 1e-14 and np.array_equal(PtVT, r[name + "_PtVT"])
        coef = np.stack([np.linalg.solve(normal[d], PtVT[:, :, d] @ r["P"][Pt, d]) for d in range(3)], axis=1)
        assert rel(coef, r[name + "_coef"]) < 1e-9
        assert rel(port.group_reduced_term(r["P"][Pt], pm, normal, PtVT), r[name + "_term"]) < 1e-9


def test_pd_local_step_golden():
    """oracle/pd_local.py (the local step that produces the constraint-projection snapshots) against get_pi /
    S_i^T of the unmodified constraint classes (projective_dynamics/Constraint_projections.py) stored in pd_local.npz; and
    the synthetic bar generator of the fixtures (animsnapbases_b200.synth) against the same statement."""
    from oracle import pd_local
    from animsnapbases_b200 import synth
    g = golden("pd_local")
    wi, (smin, smax) = float(g["wi"]), g["sigma"]
    cases = {"tet_dg": (pd_local.TetDeformationGradient(g["bar_verts"], g["bar_tets"], wi), g["bar_q"]),
             "tet_strain": (pd_local.TetStrain(g["bar_verts"], g["bar_tets"], wi, smin, smax), g["bar_q"]),
             "tri_strain": (pd_local.TriStrain(g["grid_verts"], g["grid_tris"], wi, smin, smax), g["grid_q"]),
             "edge_spring": (pd_local.EdgeSpring(g["grid_verts"], g["grid_edges"], wi), g["grid_q"])}
    cases["vert_bending"] = (pd_local.VertBending(g["sphere_verts"], g["sphere_tris"], wi), g["sphere_q"])
    for name, (group, q) in cases.items():
        p = group.project(q)
        assert p.shape == g[name + "_p"].shape == (group.rows * len(p) // group.rows, 3)
        assert rel(p, g[name + "_p"]) < 1e-12, name
        assert rel(group.assembly() @ p, g[name + "_StP"]) < 1e-12, name
    # the generator behind tests/golden/bar*.npz: same rotation rows, and its gradient operator is S (unit weights)
    dg = pd_local.TetDeformationGradient(g["bar_verts"], g["bar_tets"], 1.0)
    S = synth.tet_gradient_operator(g["bar_verts"], g["bar_tets"], weighted=False)
    unit = dg.assembly() @ sparse_diag(1.0 / np.repeat(np.abs(dg.V0), 3))
    assert abs(S.T - unit).max() < 1e-12 * abs(unit).max()
    Ft = (S @ g["bar_q"]).reshape(-1, 3, 3)                                    # rows of F^T per tet
    U, _, Vt = np.linalg.svd(np.swapaxes(Ft, 1, 2))
    R = U @ Vt
    R[np.linalg.det(R) < 0, :, 2] *= -1
    assert rel(np.swapaxes(R, 1, 2).reshape(-1, 3), g["tet_dg_p"]) < 1e-12


def sparse_diag(v):
    from scipy import sparse
    return sparse.diags(v)


def test_port_against_live_reference():
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference not mounted (GPU box)")
    import contextlib, io
    from animsnapbases_b200 import synth
    import oracle.make_golden as mg
    ref_shim.load_reference()
    with ref_shim.scratch_cwd():
        verts0, tris = synth.grid_mesh(15, 11)
        anim = synth.bump_animation(verts0, 14, n_bumps=4, sigma=0.15, seed=7)
        mass = synth.lumped_vertex_masses(anim[0].astype(float), tris)
        with contextlib.redirect_stdout(io.StringIO()):
            ps, pc = mg.ref_pos_objects(anim, tris, mass, "average", "local", 6, "SPLOCS", dict(max_itrs=2, admm=3, lam=1.5, rho=8.0))
            cap, _ = mg.run_ref_pos(pc, True)
    prep = port.prepare_snapshots(anim, mass, "average")
    assert np.array_equal(prep["snapTensor"], ps.snapTensor)
    geo = port.GeodesicHeat(np.mean(anim.astype(float), axis=0), tris)
    C, W, meas, picked, R = port.extract_k_components(prep["snapTensor"], 6, "local", geo)
    assert np.array_equal(C, pc.comps) and np.array_equal(W, pc.weigs)
    Wf, Cf, _, _ = port.splocs_glob_optimization(prep["snapTensor"], C, W, R, geo, 2, 3, 1.5, 8.0)
    assert rel(Wf, cap["W"]) < 1e-12 and rel(Cf, cap["C"]) < 1e-12
