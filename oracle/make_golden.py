"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
(imported from /root/reference through oracle/ref_shim.py) on seeded synthetic inputs, and checks the
CPU restatement (oracle/port.py) against it on the way.  Run in the build container:

    python oracle/make_golden.py            # writes tests/golden/{meshes,pos_grid,pos_sphere,bar,align,writers}.npz

The fixtures are the pin for the oracle (the reference has no golden vectors of its own).
Nothing at test/bench time on the GPU box reads /root/reference.
"""
from __future__ import annotations

import io
import os
import sys
import types
import contextlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import port  # noqa: E402
from oracle.ref_shim import REFERENCE_ROOT, load_reference, scratch_cwd  # noqa: E402
from animsnapbases_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


# ------------------------------------------------------------------------------------------------ meshes
def make_meshes():
    data = os.path.join(REFERENCE_ROOT, "data")
    out = {}
    for name in ("bunny", "armadillo", "sphere"):
        v, t = synth.read_obj_mesh(os.path.join(data, f"{name}.obj"))
        out[f"{name}_verts"] = v
        out[f"{name}_tris"] = t.astype(np.int32)
    v, tets, tris = synth.read_medit_mesh(os.path.join(data, "bar.mesh"))
    out["bar_verts"], out["bar_tets"], out["bar_tris"] = v, tets.astype(np.int32), tris.astype(np.int32)
    np.savez_compressed(os.path.join(GOLD, "meshes.npz"), **out)
    print("meshes:", {k: v.shape for k, v in out.items()})


# ------------------------------------------------------------------------------------------------ reference drivers
def ref_pos_objects(verts_f32, tris, mass, rest_shape, support, K, bases_type, splocs=None,
                    min_dist=0.1, max_dist=0.25):
    """Build reference posSnapshots/posComponents with object.__new__ (bypasses h5py)."""
    from snapbases.posComponents import posComponents
    from snapbases.posSnapshots import posSnapshots
    from utils.support import GeodesicDistanceComputation

    ps = object.__new__(posSnapshots)
    ps.verts = np.asarray(verts_f32).astype(float)
    ps.tris = tris
    ps.frs, ps.nVerts, _ = ps.verts.shape
    ps.rest_shape = rest_shape
    ps.snapTensor = ps.verts.copy()
    ps.mass = mass.copy()
    ps.massL = np.sqrt(mass)            # == diagonal of cholesky(diag(m)), posSnapshots.py:156-159
    ps.invMassL = 1.0 / ps.massL
    ps.snapTensor *= ps.massL[:, None]  # posSnapshots.py:82
    ps.mean = ps.snapTensor[0].copy() if rest_shape == "first" else np.mean(ps.snapTensor, axis=0)
    rest = ps.verts[0] if rest_shape == "first" else np.mean(ps.verts, axis=0)
    ps.compute_geodesic_distance = GeodesicDistanceComputation(rest, tris)
    ps.pre_scale_factor = 1
    ps.standarize()

    pc = object.__new__(posComponents)
    pc.basesType = bases_type
    pc.pos_snapshots = ps
    pc.numComp = K
    pc.support = support
    pc.storeSingVal = False
    pc.comps = pc.weigs = None
    pc.smooth_min_dist, pc.smooth_max_dist = min_dist, max_dist
    pc.fileNameBases = "q_pos_"
    sp = splocs or dict(max_itrs=1, admm=1, lam=2.0, rho=10.0)
    pc.param = types.SimpleNamespace(splocs_max_itrs=sp["max_itrs"], splocs_admm_num_itrs=sp["admm"],
                                     splocs_lambda=sp["lam"], splocs_rho=sp["rho"], q_standarize=True,
                                     q_orthogonal=False, q_massWeight=True)
    return ps, pc


def run_ref_pos(pc, capture_splocs):
    """Runs extract_k_components; captures SPLOCS' discarded (W, C) by wrapping the module-level
    `tensordot` (last call with axes == (1, 0) is posComponents.py:183)."""
    import snapbases.posComponents as mod
    captured = {}
    orig = mod.tensordot

    def spy(a, b, axes=2):
        if axes == (1, 0):
            captured["W"], captured["C"] = np.array(a, copy=True), np.array(b, copy=True)
        return orig(a, b, axes)

    log = io.StringIO()
    mod.tensordot = spy
    try:
        with contextlib.redirect_stdout(log):
            pc.extract_k_components(None)
    finally:
        mod.tensordot = orig
    return captured, log.getvalue()


def pos_case(name, verts0, tris, F, K, n_bumps, sigma, splocs, seed):
    verts0, _, _ = synth.normalize_unit_cube(verts0)
    anim = synth.bump_animation(verts0, F, n_bumps=n_bumps, sigma=sigma, seed=seed)
    anim = port.align_frames(anim.astype(np.float64), rigid=False)          # "_centered" pre-alignment
    mass = synth.lumped_vertex_masses(anim[0].astype(float), tris)
    out = dict(anim=anim, tris=tris.astype(np.int32), mass=mass, K=K, splocs=np.array(
        [splocs["max_itrs"], splocs["admm"], splocs["lam"], splocs["rho"]], float))

    # ---- reference: local support + SPLOCS
    ps, pc = ref_pos_objects(anim, tris, mass, "first", "local", K, "SPLOCS", splocs)
    cap, log = run_ref_pos(pc, True)
    out.update(X=ps.snapTensor.copy(), mean=ps.mean.copy(), scale=ps.pre_scale_factor,
               comps_local=pc.comps.copy(), weigs_local=pc.weigs.copy(), meas_local=pc.measures_at_largeDeforVerts,
               splocs_W=cap["W"], splocs_C=cap["C"], splocs_log=np.array(log))
    geo_ids = np.array([0, len(verts0) // 3, len(verts0) - 1])
    out["geo_ids"] = geo_ids
    out["geo_phi"] = np.array([ps.compute_geodesic_distance(int(i)) for i in geo_ids])
    # ---- reference: global support PCA
    ps2, pc2 = ref_pos_objects(anim, tris, mass, "first", "global", K, "PCA")
    run_ref_pos(pc2, False)
    out.update(comps_global=pc2.comps.copy(), weigs_global=pc2.weigs.copy(), meas_global=pc2.measures_at_largeDeforVerts)
    # post-process (standardised, mass-weighted, not orthogonalised / orthogonalised)
    with contextlib.redirect_stdout(io.StringIO()):
        pc2.param.q_orthogonal = False
        pc2.post_process_components()
    out["post_global"] = pc2.comps.copy()
    # ---- q_orthogonal branch (posComponents.py:285-287): scipy orth per dimension, on a fresh copy of the global components
    ps3, pc3 = ref_pos_objects(anim, tris, mass, "first", "global", K, "PCA")
    run_ref_pos(pc3, False)
    pre = pc3.comps.copy()
    pre /= ps3.pre_scale_factor
    pre += ps3.mean[np.newaxis]
    out["pre_orth"] = pre.copy()                      # what `orth` is applied to (:279-282)
    with contextlib.redirect_stdout(io.StringIO()):
        pc3.param.q_orthogonal = True
        pc3.post_process_components()
    out["post_global_orth"] = pc3.comps.copy()
    out["post_global_orth_raw"] = pc3.comps / ps3.invMassL[None, :, None]      # undo the final `*= invMassL` (:289-292)

    # ---- port vs reference
    prep = port.prepare_snapshots(anim, mass, "first")
    assert np.array_equal(prep["snapTensor"], ps.snapTensor), "standardise mismatch"
    geo = port.GeodesicHeat(anim[0].astype(float), tris)
    for i, gid in enumerate(geo_ids):
        assert rel(geo(int(gid)), out["geo_phi"][i]) < 1e-9, "geodesic mismatch"
    C, W, meas, picked, R = port.extract_k_components(prep["snapTensor"], K, "local", geo)
    assert np.array_equal(C, out["comps_local"]) and np.array_equal(W, out["weigs_local"]), "deflation(local) mismatch"
    assert np.array_equal(meas, out["meas_local"])
    out["picked_local"] = picked
    Wf, Cf, hist, centres = port.splocs_glob_optimization(prep["snapTensor"], C, W, R, geo, splocs["max_itrs"],
                                                          splocs["admm"], splocs["lam"], splocs["rho"])
    assert rel(Wf, cap["W"]) < 1e-12 and rel(Cf, cap["C"]) < 1e-12, ("splocs mismatch", rel(Wf, cap["W"]), rel(Cf, cap["C"]))
    out["splocs_hist"], out["splocs_centres"] = hist, centres
    Cg, Wg, measg, pickedg, _ = port.extract_k_components(prep["snapTensor"], K, "global")
    assert np.array_equal(Cg, out["comps_global"]) and np.array_equal(Wg, out["weigs_global"])
    out["picked_global"] = pickedg
    post = port.post_process_pos(Cg, prep["mean"], prep["pre_scale_factor"], prep["invMassL"], True, False, True)
    assert np.array_equal(post, out["post_global"])
    np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), **out)
    print(f"{name}: N={len(verts0)} F={F} K={K}  port == reference (deflation bit-equal, splocs rel "
          f"{rel(Wf, cap['W']):.1e}/{rel(Cf, cap['C']):.1e})")


# ------------------------------------------------------------------------------------------------ bar / constraints
def bar_case():
    from snapbases.constraintsComponents import constraintsComponents
    from snapbases.nonlinear_snapshots import nonlinearSnapshots

    m = np.load(os.path.join(GOLD, "meshes.npz"))
    verts, tets = m["bar_verts"], m["bar_tets"].astype(np.int64)
    tets = tets[:600]
    e, p, F = len(tets), 3, 24
    snaps = synth.tet_rotation_snapshots(verts, tets, F, seed=0)           # (F, 3e, 3)
    mass = synth.tet_element_masses(verts, tets, p)

    param = types.SimpleNamespace(constProj_massWeight=True, constProj_standarize=True, constProj_orthogonal=True,
                                  deim_desired_num_components=200, constProj_bases_interpolation_type="deim",
                                  constProj_snapshots_type="tets_deformation_gradient", constProj_output_directory="")
    ns = object.__new__(nonlinearSnapshots)
    ns.param = param
    ns.constraintsSize, ns.num_constained_elements, ns.dim, ns.frs = p, e, 3, F
    ns.rest_shape = "first"
    ns.snapTensor = snaps.copy()
    ns.mass = mass.copy()
    ns.massL = np.sqrt(mass)
    ns.invMassL = 1.0 / ns.massL
    ns.snapTensor *= ns.massL[:, None]
    ns.standarize()
    cc = object.__new__(constraintsComponents)
    cc.param, cc.nonlinearSnapshots = param, ns
    cc.storeSingVal, cc.file_name_sing, cc.support = False, "", "global"
    cc.geom_interpol_verts = []
    cc.comps = None

    out = dict(snaps=snaps, mass=mass, tets=tets.astype(np.int32), X=ns.snapTensor.copy(), mean=ns.mean,
               scale=ns.pre_scale_factor)
    rows = []
    writer = types.SimpleNamespace(writerow=lambda r: rows.append(list(r)))
    with contextlib.redirect_stdout(io.StringIO()):
        cc.compute_pod_for_vectorized_nonlinear_snapshots_tensor(writer)
    out["pod_comps"] = cc.comps.copy()
    out["pod_sing"] = np.array([r[1] for r in rows])
    with contextlib.redirect_stdout(io.StringIO()):
        cc.post_process_components()
    out["post_comps"] = cc.comps.copy()
    out["post_snap"] = ns.snapTensor.copy()
    with contextlib.redirect_stdout(io.StringIO()):
        cc.deim()
    out["deim_Pt"], out["deim_alpha"], out["deim_ranges"] = cc.geom_Pt, cc.geom_alpha, cc.geom_alpha_ranges
    with contextlib.redirect_stdout(io.StringIO()):
        rec = cc.geom_constructed(10, "train")
    out["deim_rec10_err"] = np.array([cc.frobenius_error(ns.snapTensor, rec)])

    # port checks
    prep = port.prepare_snapshots(snaps, mass, "first")
    assert np.array_equal(prep["snapTensor"], out["X"])
    C, S = port.pod_vectorized(prep["snapTensor"], 200)
    assert np.array_equal(C, out["pod_comps"]) and np.array_equal(S, out["pod_sing"])
    pc, psn = port.post_process_constraints(C, prep["snapTensor"], prep["mean"], prep["pre_scale_factor"],
                                            prep["invMassL"], True, True, True)
    assert np.array_equal(pc, out["post_comps"]) and np.array_equal(psn, out["post_snap"])
    d = port.deim(pc, p)
    assert np.array_equal(d["Pt"], out["deim_Pt"]) and np.array_equal(d["alpha"], out["deim_alpha"])
    rec_p = port.deim_reconstruct(pc, psn, d["alpha"][:10], 10)      # reference quirk: alpha used as rows (H6)
    assert abs(port.frobenius_error(psn, rec_p) - out["deim_rec10_err"][0]) < 1e-9 * out["deim_rec10_err"][0]

    # ---- block deflation + block DEIM (reference)
    ns2 = object.__new__(nonlinearSnapshots)
    ns2.__dict__.update(ns.__dict__)
    ns2.snapTensor = out["X"].copy()
    cc2 = object.__new__(constraintsComponents)
    cc2.param = types.SimpleNamespace(**{**param.__dict__, "deim_desired_num_components": 5,
                                         "constProj_bases_interpolation_type": "deim_block_form"})
    cc2.nonlinearSnapshots, cc2.storeSingVal, cc2.support, cc2.geom_interpol_verts = ns2, False, "global", []
    with contextlib.redirect_stdout(io.StringIO()):
        cc2.compute_nonlinearity_bases_blocks(None)
    out.update(blk_comps=cc2.comps.copy(), blk_weigs=cc2.weigs.copy(), blk_meas=cc2.measures_at_largeDeforVerts,
               blk_points=cc2.largeDeforPoints, blk_rows=cc2.largeDeforBlocks)
    with contextlib.redirect_stdout(io.StringIO()):
        cc2.post_process_components()
        cc2.deim_blocksForm()
    out.update(blk_post=cc2.comps.copy(), blk_deim_Pt=cc2.geom_Pt, blk_deim_alpha=cc2.geom_alpha)
    Cb, Wb, mb, sidx, sp_ = port.nonlinearity_bases_blocks(out["X"], e, p, 5)
    assert np.array_equal(Cb, out["blk_comps"]) and np.array_equal(Wb, out["blk_weigs"]) and np.array_equal(mb, out["blk_meas"])
    pcb, _ = port.post_process_constraints(Cb, out["X"], prep["mean"], prep["pre_scale_factor"], prep["invMassL"], True, True, True)
    assert np.array_equal(pcb, out["blk_post"])
    db = port.deim_blocks(pcb, p)
    assert np.array_equal(db["Pt"], out["blk_deim_Pt"]) and np.array_equal(db["alpha"], out["blk_deim_alpha"])
    np.savez_compressed(os.path.join(GOLD, "bar.npz"), **out)
    print(f"bar: e={e} p={p} F={F}: pod/post/deim/blocks port == reference; Pt[:8]={out['deim_Pt'][:8]}")


# ------------------------------------------------------------------------------------------------ alignment + writers
def align_case():
    from utils.process import find_rbm_procrustes, transform

    m = np.load(os.path.join(GOLD, "meshes.npz"))
    v0, _, _ = synth.normalize_unit_cube(m["sphere_verts"])
    rng = np.random.default_rng(3)
    F = 6
    frames = []
    for f in range(F):
        R = synth._rot(rng.standard_normal(3), 0.4 * rng.standard_normal())
        if f == 4:
            R = R @ np.diag([1.0, 1.0, -1.0])      # reflection: exercises the det(R) < 0 branch
        frames.append((v0 + 0.01 * rng.standard_normal(v0.shape)) @ R.T + 0.1 * rng.standard_normal(3))
    frames[0] = v0
    frames = np.array(frames, np.float32)
    out = dict(frames=frames)
    for rigid in (True, False):
        ref = []
        for v in frames:
            M = find_rbm_procrustes(v, frames[0], rigid)
            ref.append(transform(v, M))
        ref = np.array(ref, np.float32)
        mine = port.align_frames(frames, rigid)
        assert np.array_equal(mine, ref), "align mismatch"
        out[f"aligned_rigid{int(rigid)}"] = ref
    pre, mean, scale = port.preprocess_unit_cube(frames)
    out.update(pre=pre, pre_mean=mean, pre_scale=np.array([scale]))
    np.savez_compressed(os.path.join(GOLD, "align.npz"), **out)
    print("align: port == reference (rigid and translation-only, incl. reflection branch)")


def writers_case():
    import utils.utils as ru
    rng = np.random.default_rng(5)
    B = rng.standard_normal((4, 7, 3))
    with scratch_cwd() as d:
        ru.store_components(os.path.join(d, "q_pos_"), 9, 4, 7, 3, B, ".bin", "K")
        ref = open(os.path.join(d, "q_pos_F9K4.bin"), "rb").read()
        pts = np.array([5, 3, 11, 2])
        with contextlib.redirect_stdout(io.StringIO()):
            ru.store_interpol_points_vector(os.path.join(d, "pts_"), 9, 4, pts, ".bin")
        refp = open(os.path.join(d, "pts_F9K4_points4.bin"), "rb").read()
    assert port.components_bin_bytes(B) == ref and port.points_bin_bytes(pts) == refp
    np.savez_compressed(os.path.join(GOLD, "writers.npz"), B=B, bin=np.frombuffer(ref, np.uint8), pts=pts,
                        pts_bin=np.frombuffer(refp, np.uint8))
    print("writers: .bin layouts pinned")


# ------------------------------------------------------------------------------------------------ geometric DEIM
def bar_geom_case():
    """geom_block_form_utilizing_differential_operator (constraintsComponents.py:619-731) of the unmodified reference on the
    bar fixture's post-processed bases, in position space (St = transpose of the tet gradient operator) and in constraint
    space -> tests/golden/bar_geom.npz.  Reads tests/golden/bar.npz (does not regenerate it)."""
    from snapbases.constraintsComponents import constraintsComponents
    from snapbases.nonlinear_snapshots import nonlinearSnapshots

    m = np.load(os.path.join(GOLD, "meshes.npz"))
    bar = np.load(os.path.join(GOLD, "bar.npz"))
    verts, tets = m["bar_verts"], bar["tets"].astype(np.int64)
    e, p = len(tets), 3
    St = synth.tet_gradient_operator(verts, tets).T.tocsr()                 # |V| x ep
    out = dict(St_rowptr=St.indptr.astype(np.int32), St_colidx=St.indices.astype(np.int32), St_vals=St.data,
               St_shape=np.array(St.shape))
    cases = {"blk_pos2": (bar["blk_post"], True, 2), "pod_pos1": (bar["post_comps"][:24], True, 1),
             "pod_pos3": (bar["post_comps"][:24], True, 3), "pod_con": (bar["post_comps"][:24], False, 1)}
    for name, (comps, pos, per_vert) in cases.items():
        ns = object.__new__(nonlinearSnapshots)
        ns.constraintsSize, ns.num_constained_elements, ns.dim = p, e, 3
        ns.ele_type, ns.tets = "_tets", tets
        cc = object.__new__(constraintsComponents)
        cc.param = types.SimpleNamespace(geom_ele_per_vert=per_vert, constProj_snapshots_type="tets_deformation_gradient")
        cc.nonlinearSnapshots, cc.comps, cc.numComp, cc.St, cc.geom_interpol_verts = ns, comps.copy(), comps.shape[0] // p, St, []
        with contextlib.redirect_stdout(io.StringIO()):
            cc.geom_block_form_utilizing_differential_operator(pos)
        g = port.deim_geom(comps, p, e, tets, St if pos else None, per_vert)
        assert np.array_equal(g["Pt"], cc.geom_Pt) and np.array_equal(g["alpha"], cc.geom_alpha)
        assert np.array_equal(g["alpha_ranges"], cc.geom_alpha_ranges) and np.array_equal(g["verts"], cc.geom_interpol_verts)
        out.update({name + "_Pt": cc.geom_Pt, name + "_alpha": cc.geom_alpha, name + "_ranges": cc.geom_alpha_ranges,
                    name + "_verts": np.asarray(cc.geom_interpol_verts, dtype=np.int64)})
        print(f"bar_geom {name}: port == reference; alpha={cc.geom_alpha[:8]} ranges={cc.geom_alpha_ranges}")
    np.savez_compressed(os.path.join(GOLD, "bar_geom.npz"), **out)


# ------------------------------------------------------------------------------------------------ pca_blocks_with_St (a11)
def bar_blocks_st_case():
    """compute_nonlinearity_bases_blocks_utilizing_diffirential_operator (constraintsComponents.py:156-271) of the unmodified
    reference on the bar fixture's standardised snapshots -> tests/golden/bar_blocks_st.npz.  The function is well defined on
    these inputs: every rank-1 step removes one frame-space direction from all rows, so the residual norm passes bases_R_tol
    after about rank(X) steps (8 blocks for tol = 1e-8, 11 for 1e-13), before the incident-element loop of the first
    vertex runs out.  Also pins the f32 per-row POD (`pod`, :274-294, torch CPU) for oracle/port.py::pod_per_row_f32."""
    from snapbases.constraintsComponents import constraintsComponents
    from snapbases.nonlinear_snapshots import nonlinearSnapshots
    m = np.load(os.path.join(GOLD, "meshes.npz"))
    bar = np.load(os.path.join(GOLD, "bar.npz"))
    verts, tets = m["bar_verts"], bar["tets"].astype(np.int64)
    e, p = len(tets), 3
    St = synth.tet_gradient_operator(verts, tets).T.tocsr()
    out = {}
    for tag, tol in (("tol8", 1e-8), ("tol13", 1e-13)):
        ns = object.__new__(nonlinearSnapshots)
        ns.constraintsSize, ns.num_constained_elements, ns.dim = p, e, 3
        ns.ele_type, ns.tets = "_tets", tets
        ns.snapTensor = bar["X"].copy()
        cc = object.__new__(constraintsComponents)
        cc.param = types.SimpleNamespace(bases_R_tol=tol)
        cc.nonlinearSnapshots, cc.St, cc.support, cc.storeSingVal = ns, St, "global", False
        log = io.StringIO()
        with contextlib.redirect_stdout(log):
            cc.compute_nonlinearity_bases_blocks_utilizing_diffirential_operator(None)
        assert not np.isnan(cc.comps).any()
        vs = np.array([int(l.split()[1]) for l in log.getvalue().split("\n") if l.startswith("vert")])
        C, W, M, V = port.nonlinearity_bases_blocks_with_St(bar["X"], e, p, St, tets, tol)
        assert np.array_equal(C, cc.comps) and np.array_equal(W, cc.weigs) and np.array_equal(M, cc.measures_at_largeDeforVerts)
        assert np.array_equal(V, vs)
        out.update({tag + "_comps": cc.comps, tag + "_weigs": cc.weigs, tag + "_meas": cc.measures_at_largeDeforVerts,
                    tag + "_verts": vs, tag + "_tol": np.array([tol])})
        print(f"bar_blocks_st {tag}: {cc.comps.shape[0] // p} blocks, verts {vs}, port == reference (bit-equal)")
    # ---- `pod` basis type: p * 3 float32 SVDs through torch on the CPU (:274-294)
    ns = object.__new__(nonlinearSnapshots)
    ns.constraintsSize, ns.num_constained_elements, ns.dim = p, e, 3
    ns.snapTensor = bar["X"].copy()
    cc = object.__new__(constraintsComponents)
    cc.param = types.SimpleNamespace(deim_desired_num_components=8)
    cc.nonlinearSnapshots = ns
    with contextlib.redirect_stdout(io.StringIO()):
        cc.compute_pod_for_nonlinear_snapshots_tensor(None)
    ref = np.asarray(cc.comps)
    mine = port.pod_per_row_f32(bar["X"], e, p, 8)
    worst = 0.0
    for k in range(ref.shape[0]):
        for i in range(p):
            for d in range(3):
                a, b = ref[k, i::p, d], mine[k, i::p, d]
                sg = np.sign(float(a @ b)) or 1.0
                worst = max(worst, float(np.linalg.norm(a - sg * b)))
    print(f"pod (f32, per row): port (LAPACK sgesdd) vs reference (torch.linalg.svd f32): max mode difference up to sign {worst:.2e}")
    assert worst < 5e-3, worst          # float32 SVDs from two LAPACK drivers: close modes rotate at 1e-4 .. 1e-3
    out["pod_f32_comps"] = ref
    np.savez_compressed(os.path.join(GOLD, "bar_blocks_st.npz"), **out)


# ------------------------------------------------------------------------------------------------ geometric DEIM, '_verts'
def verts_geom_inputs():
    """Vertex-based constraint snapshots on an open triangle grid: rows = cotangent-Laplacian vectors (L x_f) at the
    constrained vertices, St = L[constrained, :]^T (|V| x e), bases = POD of the centred snapshots.  Two groups:
    'bend' = interior vertices only (verts_bending: row index != vertex id), 'all' = every vertex."""
    verts0, tris = synth.grid_mesh(14, 11)
    anim = synth.bump_animation(verts0, 18, n_bumps=5, sigma=0.18, seed=3).astype(np.float64)
    L, _ = port.mesh_laplacian(verts0.astype(np.float64), tris)
    nv = len(verts0)
    edge_count = {}
    for t in tris:
        for a, b in ((t[0], t[1]), (t[1], t[2]), (t[2], t[0])):
            key = (min(a, b), max(a, b))
            edge_count[key] = edge_count.get(key, 0) + 1
    boundary = sorted({v for (a, b), c in edge_count.items() if c == 1 for v in (a, b)})
    interior = np.setdiff1d(np.arange(nv), boundary)
    groups = {}
    for name, idx in (("bend", interior), ("all", np.arange(nv))):
        Lc = L.tocsr()[idx]
        X = np.stack([Lc @ f for f in anim])                      # (F, e, 3)
        X = X - X.mean(axis=0)
        comps, _ = port.pod_vectorized(X, 8)
        groups[name] = dict(idx=idx.astype(np.int64), St=Lc.T.tocsr(), comps=comps)
    return tris.astype(np.int64), groups


def verts_geom_case():
    """'_verts' flavour of geom_block_form_utilizing_differential_operator (constraintsComponents.py:684-702) of the
    unmodified reference -> tests/golden/verts_geom.npz: vertex stars as candidate sets, with (verts_bending) and without
    the constrained-vertex index mapping, 1 and 3 elements per picked vertex."""
    import tempfile
    from snapbases.constraintsComponents import constraintsComponents
    from snapbases.nonlinear_snapshots import nonlinearSnapshots
    tris, groups = verts_geom_inputs()
    out = dict(tris=tris)
    for gname, g in groups.items():
        St, comps, idx = g["St"], g["comps"], g["idx"]
        out.update({f"{gname}_idx": idx, f"{gname}_comps": comps, f"{gname}_St_rowptr": St.indptr.astype(np.int32),
                    f"{gname}_St_colidx": St.indices.astype(np.int32), f"{gname}_St_vals": St.data,
                    f"{gname}_St_shape": np.array(St.shape)})
        for per_vert in (1, 3):
            bending = gname == "bend"
            ns = object.__new__(nonlinearSnapshots)
            ns.constraintsSize, ns.num_constained_elements, ns.dim = 1, len(idx), 3
            ns.ele_type, ns.tris = "_verts", tris
            cc = object.__new__(constraintsComponents)
            with tempfile.TemporaryDirectory() as tmp:
                f = os.path.join(tmp, "verts_bending_constrained_indices.npz")
                np.savez(f, indices=idx)
                cc.param = types.SimpleNamespace(geom_ele_per_vert=per_vert, constProj_input_snaps_constrained_elements=f,
                                                 constProj_snapshots_type="verts_bending" if bending else "verts_other")
                cc.nonlinearSnapshots, cc.comps, cc.numComp, cc.St, cc.geom_interpol_verts = ns, comps.copy(), comps.shape[0], St, []
                with contextlib.redirect_stdout(io.StringIO()):
                    cc.geom_block_form_utilizing_differential_operator(True)
            gp = port.deim_geom(comps, 1, len(idx), None, St, per_vert, star_faces=tris, constrained_verts=idx if bending else None)
            assert np.array_equal(gp["Pt"], cc.geom_Pt) and np.array_equal(gp["alpha"], cc.geom_alpha)
            assert np.array_equal(gp["alpha_ranges"], cc.geom_alpha_ranges) and np.array_equal(gp["verts"], cc.geom_interpol_verts)
            key = f"{gname}{per_vert}"
            out.update({key + "_Pt": np.asarray(cc.geom_Pt, dtype=np.int64), key + "_alpha": np.asarray(cc.geom_alpha, dtype=np.int64),
                        key + "_ranges": cc.geom_alpha_ranges, key + "_verts": np.asarray(cc.geom_interpol_verts, dtype=np.int64)})
            print(f"verts_geom {key}: port == reference; verts={cc.geom_interpol_verts} alpha={cc.geom_alpha[:8]} "
                  f"Pt={cc.geom_Pt[:8]} ranges={cc.geom_alpha_ranges}")
    np.savez_compressed(os.path.join(GOLD, "verts_geom.npz"), **out)


# ------------------------------------------------------------------------------------------------ consumer side (n2)
REDUCED_CASES = {   # name: (reduction type, components key, (Pt, alpha, ranges) keys, num_components)
    "pod": ("deim_pod_vectorized", "post_comps", ("deim_Pt", "deim_alpha", "deim_ranges"), 10),
    "blk": ("geom_pca_blocks_withSt", "blk_post", ("blk_pos2_Pt", "blk_pos2_alpha", "blk_pos2_ranges"), 4),
}


def bar_reduced_case():
    """The reference's consumer of the emitted `.npz` (projective_dynamics/Simulators.py:157-218 prepare_reduced_group and
    :366-399 get_group_reduced_term, unmodified, called unbound on a stand-in `self`) on the bar fixture's bases and
    interpolation sets -> tests/golden/bar_reduced.npz.  Reads bar.npz / bar_geom.npz (does not regenerate them)."""
    import tempfile
    from scipy import sparse
    from scipy.linalg import lu_solve
    from oracle.ref_shim import load_pd_simulators
    Solver = load_pd_simulators().animSnapBasesSolver

    bar = np.load(os.path.join(GOLD, "bar.npz"))
    geo = np.load(os.path.join(GOLD, "bar_geom.npz"))
    St = sparse.csr_matrix((geo["St_vals"], geo["St_colidx"], geo["St_rowptr"]), shape=tuple(geo["St_shape"]))
    p, group = 3, "tets_deformation_gradient"
    fixtures = {**{k: bar[k] for k in bar.files}, **{k: geo[k] for k in geo.files}}
    P = np.random.default_rng(11).standard_normal((St.shape[1], 3))          # stand-in constraint projections (ep, 3)

    class Projection:                                                        # what get_pi returns for element alpha
        def __init__(self, alpha):
            self.alpha = alpha

        def get_pi(self, q_t):
            return P[self.alpha * p:(self.alpha + 1) * p]

    constraints = [Projection(a) for a in range(St.shape[1] // p)]
    out = dict(P=P)
    for name, (rtype, ckey, (kpt, kal, krg), ncomp) in REDUCED_CASES.items():
        local = dict(components=fixtures[ckey], interpol_alphas=fixtures[kal], Pt=fixtures[kpt],
                     interpol_verts=np.zeros(0, dtype=np.int64), interpol_alpha_ranges=fixtures[krg])
        me = types.SimpleNamespace(constraint_projection_reduction_type=rtype, reduced_position=False)
        with tempfile.TemporaryDirectory() as tmp:
            os.makedirs(os.path.join(tmp, group))
            np.savez(os.path.join(tmp, group, "basis.npz"), **local)
            with contextlib.redirect_stdout(io.StringIO()):
                alpha, Pt, pm, solvers = Solver.prepare_reduced_group(me, True, True, group, ncomp, p, St, tmp, "basis.npz")
        term = Solver.get_group_reduced_term(me, None, constraints, p, alpha, Pt, pm, solvers)
        g_alpha, g_Pt, g_pm, g_normal, g_PtVT = port.prepare_reduced_group(local, ncomp, p, St, rtype)
        assert np.array_equal(g_alpha, alpha) and np.array_equal(g_Pt, np.asarray(Pt))
        assert rel(g_pm, pm) < 1e-14, rel(g_pm, pm)
        g_term = port.group_reduced_term(P[np.asarray(Pt)], g_pm, g_normal, g_PtVT)
        coef = np.stack([lu_solve(solvers[d][0], solvers[d][1] @ P[np.asarray(Pt), d]) for d in range(3)], axis=1)
        g_coef = np.stack([np.linalg.solve(g_normal[d], g_PtVT[:, :, d] @ P[np.asarray(Pt), d]) for d in range(3)], axis=1)
        print(f"bar_reduced {name}: |Pt|={len(Pt)} mp={pm.shape[1]}  port vs reference: S^T V {rel(g_pm, pm):.1e}, "
              f"coefficients {rel(g_coef, coef):.1e}, term {rel(g_term, term):.1e}")
        assert rel(g_term, term) < 1e-9
        out.update({name + "_alpha": np.asarray(alpha), name + "_Pt": np.asarray(Pt, dtype=np.int64), name + "_StV": pm,
                    name + "_PtVT": np.stack([solvers[d][1] for d in range(3)], axis=2), name + "_coef": coef,
                    name + "_term": term})
    np.savez_compressed(os.path.join(GOLD, "bar_reduced.npz"), **out)


# ------------------------------------------------------------------------------------------------ PD local step (n4)
def pd_local_inputs():
    """A few tets / triangles / edges of the bar and of a grid with seeded deformed positions (some tets inverted so that the
    reflection branches run)."""
    m = np.load(os.path.join(GOLD, "meshes.npz"))
    rng = np.random.default_rng(17)
    bv, bt = m["bar_verts"].astype(np.float64), m["bar_tets"].astype(np.int64)[::130][:24]
    gv, gt = synth.grid_mesh(6, 5)
    gv = gv.astype(np.float64) + 0.02 * rng.standard_normal(gv.shape)
    gt = gt.astype(np.int64)[:30]
    ge = np.unique(np.sort(np.concatenate([gt[:, [0, 1]], gt[:, [1, 2]], gt[:, [2, 0]]]), axis=1), axis=0)[:40]
    qb = bv + 0.15 * rng.standard_normal(bv.shape)
    flip = np.unique(bt[::5].ravel())
    qb[flip, 0] = -qb[flip, 0]                                              # inverts some tets
    qg = gv + 0.1 * rng.standard_normal(gv.shape)
    return dict(bar_verts=bv, bar_tets=bt, bar_q=qb, grid_verts=gv, grid_tris=gt, grid_edges=ge, grid_q=qg)


def m_sphere():
    m = np.load(os.path.join(GOLD, "meshes.npz"))
    return m["sphere_verts"].astype(np.float64), m["sphere_tris"].astype(np.int64)


def pd_local_case():
    """get_pi and S_i^T of the reference's constraint classes (projective_dynamics/Constraint_projections.py, unmodified, loaded
    by ref_shim.load_pd_module) per element -> tests/golden/pd_local.npz, and oracle/pd_local.py checked against them."""
    from scipy import sparse
    from oracle import pd_local
    from oracle.ref_shim import load_pd_module
    cp = load_pd_module("Constraint_projections")
    d = pd_local_inputs()
    wi, smin, smax = 3.5, 0.9, 1.1
    groups = {
        "tet_dg": (pd_local.TetDeformationGradient(d["bar_verts"], d["bar_tets"], wi),
                   [cp.TetDeformationGradientConstraint(list(t), wi, d["bar_verts"]) for t in d["bar_tets"]], d["bar_q"]),
        "tet_strain": (pd_local.TetStrain(d["bar_verts"], d["bar_tets"], wi, smin, smax),
                       [cp.TetStrainConstraint(list(t), wi, d["bar_verts"], smin, smax) for t in d["bar_tets"]], d["bar_q"]),
        "tri_strain": (pd_local.TriStrain(d["grid_verts"], d["grid_tris"], wi, smin, smax),
                       [cp.TriStrainConstraint(list(t), wi, d["grid_verts"], smin, smax) for t in d["grid_tris"]], d["grid_q"]),
        "edge_spring": (pd_local.EdgeSpring(d["grid_verts"], d["grid_edges"], wi),
                        [cp.EdgeSpringConstraint(list(t), wi, d["grid_verts"]) for t in d["grid_edges"]], d["grid_q"]),
    }
    # vertex bending on a curved, closed surface: every vertex is constrained; the builder of the reference's model class
    # (vertex_star / vertex_masses / add_vertex_bending_constraint, :1086-1223) runs unbound on a stand-in `self`
    sv, sf = m_sphere()
    rng = np.random.default_rng(23)
    sq = sv * (1.0 + 0.3 * np.sin(3.0 * sv[:, :1])) + 0.03 * rng.standard_normal(sv.shape)
    sq[::7] = sv[::7] * 0.6                                                  # dents: some stars flip to the other side
    Model = [getattr(cp, n) for n in dir(cp) if hasattr(getattr(cp, n), "add_vertex_bending_constraint")][0]
    model = types.SimpleNamespace(positions=sv, faces=sf, constraints=[], verts_bending_indicies=[], verts_bending_constraints=[],
                                  triangle_area=lambda tri, pos: Model.triangle_area(None, tri, pos))
    model.vertex_masses = lambda tris, pos: Model.vertex_masses(model, tris, pos)
    model.vertex_star = lambda: Model.vertex_star(model)
    Model.add_vertex_bending_constraint(model, wi)
    d.update(sphere_verts=sv, sphere_tris=sf, sphere_q=sq)
    groups["vert_bending"] = (pd_local.VertBending(sv, sf, wi), model.verts_bending_constraints, sq)
    out = dict(d, wi=wi, sigma=np.array([smin, smax]))
    for name, (mine, theirs, q) in groups.items():
        p_ref = np.concatenate([c.get_pi(q.reshape(-1)) for c in theirs], axis=0)
        St_ref = sparse.hstack([sparse.csr_matrix(c.selection_matrix) for c in theirs]).tocsr()
        p_mine, St_mine = mine.project(q), mine.assembly()
        assert p_mine.shape == p_ref.shape and rel(p_mine, p_ref) < 1e-13, (name, rel(p_mine, p_ref))
        assert St_mine.shape == St_ref.shape and abs(St_mine - St_ref).max() < 1e-12 * abs(St_ref).max(), name
        rhs_ref = np.zeros_like(q)
        for c in theirs:
            c.project_wi_SiT_pi(q.reshape(-1), rhs_ref)                   # the assembled local-step term S^T p
        assert rel(St_mine @ p_mine, rhs_ref) < 1e-12, name
        if name == "vert_bending":
            assert np.array_equal(mine.verts, model.verts_bending_indicies) and len(mine.verts) == len(sv)
            free = pd_local.VertBending(sv, sf, wi, prevent_bending_flips=False).project(q)
            flips = int((np.abs(free - p_mine).max(axis=1) > 0).sum())
            assert 0 < flips < len(sv), "the fixture must exercise both sides of the flip test (:211-214)"
            print(f"pd_local vert_bending: {flips} of {len(sv)} corrections flipped back")
        out[name + "_p"], out[name + "_StP"] = p_ref, rhs_ref
        print(f"pd_local {name}: {len(theirs)} elements, port vs reference p {rel(p_mine, p_ref):.1e}, S^T p {rel(St_mine @ p_mine, rhs_ref):.1e}")
    np.savez_compressed(os.path.join(GOLD, "pd_local.npz"), **out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    load_reference()
    if "--pd-only" in sys.argv:                # writes pd_local.npz only
        with scratch_cwd():
            return pd_local_case()
    if "--pos-only" in sys.argv:               # rewrites pos_grid.npz / pos_sphere.npz only
        with scratch_cwd():
            m = np.load(os.path.join(GOLD, "meshes.npz"))
            gv, gt = synth.grid_mesh(24, 20)
            pos_case("pos_grid", gv, gt, F=24, K=10, n_bumps=5, sigma=0.12, splocs=dict(max_itrs=3, admm=5, lam=2.0, rho=10.0), seed=0)
            pos_case("pos_sphere", m["sphere_verts"], m["sphere_tris"].astype(np.int64), F=30, K=12, n_bumps=6, sigma=0.2,
                     splocs=dict(max_itrs=2, admm=10, lam=2.0, rho=10.0), seed=1)
            return
    if "--blocks-st-only" in sys.argv:         # reads bar.npz, writes bar_blocks_st.npz only
        with scratch_cwd():
            return bar_blocks_st_case()
    if "--verts-only" in sys.argv:             # writes verts_geom.npz only
        with scratch_cwd():
            return verts_geom_case()
    if "--reduced-only" in sys.argv:           # reads bar.npz / bar_geom.npz, writes bar_reduced.npz only
        with scratch_cwd():
            return bar_reduced_case()
    if "--geom-only" not in sys.argv:
        make_meshes()
    with scratch_cwd():
        m = np.load(os.path.join(GOLD, "meshes.npz"))
        gv, gt = synth.grid_mesh(24, 20)
        pos_case("pos_grid", gv, gt, F=24, K=10, n_bumps=5, sigma=0.12,
                 splocs=dict(max_itrs=3, admm=5, lam=2.0, rho=10.0), seed=0)
        pos_case("pos_sphere", m["sphere_verts"], m["sphere_tris"].astype(np.int64), F=30, K=12, n_bumps=6, sigma=0.2,
                 splocs=dict(max_itrs=2, admm=10, lam=2.0, rho=10.0), seed=1)
        if "--geom-only" in sys.argv:          # the other fixtures are committed; only (re)write bar_geom.npz
            return bar_geom_case()
        bar_case()
        bar_geom_case()
        bar_blocks_st_case()
        bar_reduced_case()
        verts_geom_case()
        pd_local_case()
        align_case()
        writers_case()


if __name__ == "__main__":
    main()
