"""TEST INFRASTRUCTURE ONLY -- runs the reference's CPU implementation of the basis path on the inputs of a preset
(animsnapbases_b200/configs.py), timed.  Two back ends:

  * kind "reference": the UNMODIFIED reference (/root/reference, imported through oracle/ref_shim.py): its own
    posComponents.compute_components_store_singvalues (snapbases/posComponents.py:259 -> :67-189) and
    constraintsComponents.compute_components_store_singvalues / post_process_components / deim
    (snapbases/constraintsComponents.py:108, :415-446, :797-860).  Only where /root/reference is mounted (build container).
  * kind "port": oracle/port.py, the restatement pinned bit-equal to it (GPU box: the reference tree does not travel).

Callers: oracle/make_golden_scale.py (fixtures), bench.py's `--impl reference` and `cpu_baseline` legs, tests/.
"""
from __future__ import annotations

import contextlib
import io
import os
import re
import time
import types

import numpy as np

from . import port
from .ref_shim import load_reference, reference_available, scratch_cwd


def backend_kind(prefer_reference=True) -> str:
    return "reference" if prefer_reference and reference_available() else "port"


def host_info():
    import platform
    model = ""
    try:
        with open("/proc/cpuinfo") as fh:
            for line in fh:
                if line.startswith("model name"):
                    model = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    import scipy
    return {"cpu": model or platform.processor(), "cores": os.cpu_count() or 1, "numpy": np.__version__, "scipy": scipy.__version__}


# ------------------------------------------------------------------------------------------------ position path
def _ref_pos_objects(anim_f32, tris, mass, K, bases_type, splocs, min_dist, max_dist):
    """Reference posSnapshots / posComponents built with object.__new__ (bypasses h5py); every attribute
    below is what read() / read_factorize_masses() / do_snapshots_precomputations() would have set (posSnapshots.py:64-160)."""
    from snapbases.posComponents import posComponents
    from snapbases.posSnapshots import posSnapshots
    from utils.support import GeodesicDistanceComputation

    ps = object.__new__(posSnapshots)
    ps.verts = np.asarray(anim_f32).astype(float)
    ps.tris = tris
    ps.frs, ps.nVerts, _ = ps.verts.shape
    ps.rest_shape = "first"
    ps.snapTensor = ps.verts.copy()
    ps.mass = mass.copy()
    ps.massL = np.sqrt(mass)
    ps.invMassL = 1.0 / ps.massL
    ps.snapTensor *= ps.massL[:, None]
    ps.mean = ps.snapTensor[0].copy()
    ps.compute_geodesic_distance = GeodesicDistanceComputation(ps.verts[0], tris)
    ps.pre_scale_factor = 1
    ps.standarize()
    pc = object.__new__(posComponents)
    pc.basesType = bases_type
    pc.pos_snapshots = ps
    pc.numComp = K
    pc.support = "local"
    pc.storeSingVal = False
    pc.comps = pc.weigs = None
    pc.smooth_min_dist, pc.smooth_max_dist = min_dist, max_dist
    pc.fileNameBases = "q_pos_"
    its, admm, lam, rho = splocs or (1, 1, 2.0, 10.0)
    pc.param = types.SimpleNamespace(splocs_max_itrs=int(its), splocs_admm_num_itrs=int(admm), splocs_lambda=lam, splocs_rho=rho,
                                     q_standarize=True, q_orthogonal=False, q_massWeight=True, name="scale",
                                     vertPos_output_directory=".")
    return ps, pc


def run_pos(inputs, K=None, splocs="preset", prefer_reference=True, min_dist=0.1, max_dist=0.25, capture_hist=False):
    """Deflation 'PCA' with local support (+ SPLOCS) of a 'pos' preset on the CPU.
    Returns dict(kind, t_prepare, t_deflation, t_splocs, comps, weigs, meas, picked (port only, else derived), splocs_W,
    splocs_C, hist, centres, X, mean, scale, geo_phi0)."""
    cfg = inputs["preset"]
    K = int(cfg["comps"] if K is None else K)
    sp = cfg["splocs"] if splocs == "preset" else splocs
    anim, tris, mass = inputs["anim"], inputs["tris"], inputs["mass"]
    kind = backend_kind(prefer_reference)
    out = {"kind": kind}
    if kind == "reference":
        load_reference()
        import snapbases.posComponents as mod
        with scratch_cwd():
            t0 = time.perf_counter()
            ps, pc = _ref_pos_objects(anim, tris, mass, K, "SPLOCS" if sp else "PCA", sp, min_dist, max_dist)
            out["t_prepare"] = time.perf_counter() - t0
            cap, tsp, tspy = {"r2": []}, [0.0], [0.0]
            orig_td, orig_sp, orig_geo = mod.tensordot, pc.splocs_glob_optimization, ps.compute_geodesic_distance
            queried = []

            def geo_logged(idx):            # every support map goes through here (posComponents.py:62): the picked vertices
                queried.append(int(idx))
                return orig_geo(idx)

            def spy(a, b, axes=2):          # each call with axes == (1, 0) is posComponents.py:183; the last carries the final (W, C)
                res = orig_td(a, b, axes)
                if axes == (1, 0):
                    t = time.perf_counter()
                    cap["W"], cap["C"] = a, b
                    if capture_hist:
                        cap["r2"].append(float(((ps.snapTensor - res) ** 2).sum()))
                    tspy[0] += time.perf_counter() - t
                return res

            def timed_splocs(*a, **k):
                t = time.perf_counter()
                r = orig_sp(*a, **k)
                tsp[0] = time.perf_counter() - t - tspy[0]
                return r

            ps.compute_geodesic_distance = geo_logged
            pc.splocs_glob_optimization = timed_splocs
            mod.tensordot = spy
            log = io.StringIO()
            try:
                t0 = time.perf_counter()
                with contextlib.redirect_stdout(log):
                    pc.compute_components_store_singvalues()          # posComponents.py:259
                total = time.perf_counter() - t0 - tspy[0]
            finally:
                mod.tensordot = orig_td
                ps.compute_geodesic_distance = orig_geo
        out.update(t_deflation=total - tsp[0], t_splocs=tsp[0], comps=pc.comps, weigs=pc.weigs,
                   meas=np.asarray(pc.measures_at_largeDeforVerts), X=ps.snapTensor, mean=ps.mean, scale=ps.pre_scale_factor)
        out["picked"] = np.array(queried[:K], dtype=np.int64)
        if sp:
            out.update(splocs_W=np.array(cap["W"]), splocs_C=np.array(cap["C"]))
            out["centres"] = np.array(queried[K:], dtype=np.int64).reshape(int(sp[0]), K)
            rows = [re.findall(r"=\s*([-+0-9.eE]+|nan|inf)", ln) for ln in log.getvalue().split("\n") if ln.startswith("itr ")]
            out["hist"] = np.array([[float(r[0]), float(r[1])] for r in rows])          # the reference's own printed %f values
            out["r2"] = np.array(cap["r2"])                                              # |X - W C|_F^2 per iteration, full precision
        out["geodesic"] = ps.compute_geodesic_distance
        return out
    t0 = time.perf_counter()
    prep = port.prepare_snapshots(anim, mass, "first")
    geo = port.GeodesicHeat(anim[0].astype(float), tris)
    out["t_prepare"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    C, W, meas, picked, R = port.extract_k_components(prep["snapTensor"], K, "local", geo, min_dist, max_dist)
    out["t_deflation"] = time.perf_counter() - t0
    out.update(comps=C, weigs=W, meas=meas, picked=picked, X=prep["snapTensor"], mean=prep["mean"], scale=prep["pre_scale_factor"],
               geodesic=geo, t_splocs=0.0)
    if sp:
        its, admm, lam, rho = sp
        t0 = time.perf_counter()
        Wf, Cf, hist, centres = port.splocs_glob_optimization(prep["snapTensor"], C, W, R, geo, int(its), int(admm), lam, rho,
                                                              min_dist, max_dist)
        out["t_splocs"] = time.perf_counter() - t0
        out.update(splocs_W=Wf, splocs_C=Cf, hist=hist, centres=centres)
    return out


# ------------------------------------------------------------------------------------------------ constraint path
def run_bar(inputs, prefer_reference=True):
    """pod_vectorized -> post_process_components -> deim of a 'bar' preset on the CPU (config 3).
    Returns dict(kind, t_prepare, t_pod, t_post, t_deim, sing, pod_comps, post_comps, post_snap, Pt, alpha, X, mean, scale)."""
    cfg = inputs["preset"]
    snaps, mass, p = inputs["snaps"], inputs["mass"], inputs["p"]
    F, ep, _ = snaps.shape
    e = ep // p
    kind = backend_kind(prefer_reference)
    out = {"kind": kind}
    if kind == "reference":
        load_reference()
        from snapbases.constraintsComponents import constraintsComponents
        from snapbases.nonlinear_snapshots import nonlinearSnapshots
        with scratch_cwd():
            param = types.SimpleNamespace(constProj_massWeight=True, constProj_standarize=True, constProj_orthogonal=True,
                                          deim_desired_num_components=int(cfg["comps"]), constProj_bases_interpolation_type="deim",
                                          constProj_snapshots_type="tets_deformation_gradient", constProj_output_directory="")
            t0 = time.perf_counter()
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
            out["t_prepare"] = time.perf_counter() - t0
            cc = object.__new__(constraintsComponents)
            cc.param, cc.nonlinearSnapshots = param, ns
            cc.storeSingVal, cc.file_name_sing, cc.support = False, "", "global"
            cc.geom_interpol_verts = []
            cc.comps = None
            out.update(X=ns.snapTensor.copy(), mean=ns.mean, scale=ns.pre_scale_factor)
            rows = []
            writer = types.SimpleNamespace(writerow=lambda r: rows.append(list(r)))
            sink = io.StringIO()
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(sink):
                cc.compute_pod_for_vectorized_nonlinear_snapshots_tensor(writer)      # constraintsComponents.py:298-320
            out["t_pod"] = time.perf_counter() - t0
            out["pod_comps"] = cc.comps.copy()
            out["sing"] = np.array([r[1] for r in rows])
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(sink):
                cc.post_process_components()                                          # :415-446
            out["t_post"] = time.perf_counter() - t0
            out["post_comps"], out["post_snap"] = cc.comps.copy(), ns.snapTensor.copy()
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(sink):
                cc.deim()                                                             # :797-860
            out["t_deim"] = time.perf_counter() - t0
            out["Pt"], out["alpha"] = np.asarray(cc.geom_Pt), np.asarray(cc.geom_alpha)
        return out
    t0 = time.perf_counter()
    prep = port.prepare_snapshots(snaps, mass, "first")
    out["t_prepare"] = time.perf_counter() - t0
    out.update(X=prep["snapTensor"], mean=prep["mean"], scale=prep["pre_scale_factor"])
    t0 = time.perf_counter()
    C, S = port.pod_vectorized(prep["snapTensor"], int(cfg["comps"]))
    out["t_pod"] = time.perf_counter() - t0
    out["pod_comps"], out["sing"] = C, S
    t0 = time.perf_counter()
    pc, psn = port.post_process_constraints(C, prep["snapTensor"], prep["mean"], prep["pre_scale_factor"], prep["invMassL"],
                                            True, True, True)
    out["t_post"] = time.perf_counter() - t0
    out["post_comps"], out["post_snap"] = pc, psn
    t0 = time.perf_counter()
    d = port.deim(pc, p)
    out["t_deim"] = time.perf_counter() - t0
    out["Pt"], out["alpha"] = np.asarray(d["Pt"]), np.asarray(d["alpha"])
    return out
