"""CPU suite (-m "not gpu"): host-side orchestration with the numpy stand-in ops against the golden fixtures, the
C-ABI surface (library loads, exports every symbol include/asb.h declares -- no compute without a GPU), row sharding."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT
from numpy_ops import NumpyOps
import pipeline_checks
from animsnapbases_b200 import _lib, engine


def test_library_exports_every_declared_symbol():
    from animsnapbases_b200.build import build_library
    path = build_library()
    lib = ctypes.CDLL(path)
    protos = _lib.parse_header()
    assert len(protos) >= 50
    for name in protos:
        assert hasattr(lib, name), f"{name} declared in include/asb.h but not exported by libasb.so"
    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T asb_" in l}
    assert exported == set(protos), f"header/library mismatch: {exported ^ set(protos)}"
    assert _lib.load().asb_version() == 100


def test_product_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from animsnapbases_b200.ops import CudaOps
    with pytest.raises(_lib.AsbError):
        CudaOps()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "animsnapbases_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "numpy_ops" not in txt.replace(
                    "tests/numpy_ops.py", ""), f"{f} references test infrastructure"


def test_shard_partition():
    for n, world in ((1000, 1), (1000, 3), (14290, 8), (33, 2), (1_000_000, 8)):
        shards = [engine.Shard.make(n, world, r) for r in range(world)]
        assert sum(s.n for s in shards) == n
        assert all(s.row0 % 32 == 0 and s.ld % 32 == 0 and s.ld >= s.n for s in shards)
        assert all(shards[i].row0 + shards[i].n == shards[i + 1].row0 for i in range(world - 1))


@pytest.mark.parametrize("name", ["pos_grid", "pos_sphere"])
def test_position_pipeline_host_logic(name):
    assert pipeline_checks.check_positions(NumpyOps(), engine.Comm(enabled=False), name)


def test_position_pipeline_with_pick_prediction(monkeypatch):
    """engine.deflate with the next pick predicted one component ahead (default from 4 ranks on, forced here): the picks, the
    components and everything downstream must be what they are without it."""
    monkeypatch.setenv("ASB_DEFLATE_PREDICT", "1")
    assert pipeline_checks.check_positions(NumpyOps(), engine.Comm(enabled=False), "pos_grid")


def test_bunny_live_oracle_host_logic():
    """BASELINE config 1/2 mesh (bunny, N = 14 290) at an oracle-friendly frame count: engine over the numpy statement of the
    kernels == oracle/port.py run live (deflation with geodesic support + SPLOCS)."""
    assert pipeline_checks.check_positions_live(NumpyOps(), engine.Comm(enabled=False), "bunny")


def test_deflation_block_size_invariance():
    """Blocking the rank-1 updates must not change which vertices are picked nor the components."""
    import torch
    from conftest import golden, planar_to_aos, rel
    g = golden("pos_sphere")
    ops, comm = NumpyOps(), engine.Comm(enabled=False)
    N = g["anim"].shape[1]
    sh = engine.Shard.make(N)
    st = engine.prepare_snapshots(ops, comm, torch.from_numpy(g["anim"]), g["mass"], "first", sh)
    outs = [engine.deflate(ops, comm, st.X, sh, int(g["K"]), False, block=b) for b in (1, 5, 16)]
    for C, W, meas, picked, R in outs[1:]:
        assert np.array_equal(picked.numpy(), outs[0][3].numpy())
        assert rel(planar_to_aos(C, N), planar_to_aos(outs[0][0], N)) < 1e-12
        assert rel(R.numpy(), outs[0][4].numpy()) < 1e-10


def test_bar_pipeline_host_logic():
    assert pipeline_checks.check_bar(NumpyOps(), engine.Comm(enabled=False))


def test_bar_blocks_host_logic():
    assert pipeline_checks.check_bar_blocks(NumpyOps(), engine.Comm(enabled=False))


def test_bar_pod_per_row_host_logic():
    assert pipeline_checks.check_bar_pod_per_row(NumpyOps(), engine.Comm(enabled=False))


def test_bar_reconstruction_host_logic():
    assert pipeline_checks.check_bar_reconstruction(NumpyOps(), engine.Comm(enabled=False))


def test_verts_geometric_deim_protocol(tmp_path):
    import api_checks
    assert api_checks.check_verts_geom_protocol(NumpyOps(), engine.Comm(enabled=False), tmp_path)


def test_verts_geometric_deim_host_logic():
    """'_verts' geometric DEIM (vertex stars, verts_bending index mapping) == the reference's picks (verts_geom.npz)."""
    assert pipeline_checks.check_verts_geom(NumpyOps(), engine.Comm(enabled=False))


def test_pd_local_snapshot_producer_host_logic(tmp_path):
    """pd_local.ConstraintGroup over the numpy statement of asb_pd_project == the reference's constraint classes."""
    assert pipeline_checks.check_pd_local(NumpyOps(), engine.Comm(enabled=False), tmp_path)


def test_pd_math_on_host(tmp_path):
    """The arithmetic of the asb_pd_project kernel (csrc/pd_elements.cuh + pd_math.cuh: Jacobi eigen-pairs of F^T F instead of
    LAPACK's SVD, host + device inline functions) compiled for the CPU and compared with get_pi of the reference's
    constraint classes (tests/golden/pd_local.npz, some tets inverted) -- a check of the kernel's math that needs no GPU."""
    import ctypes
    import shutil
    from conftest import golden, rel
    from oracle import pd_local as oracle_pd
    if shutil.which("g++") is None:
        pytest.skip("no host compiler")
    so = str(tmp_path / "libpd_host.so")
    r = subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-shared", "-fPIC", "-I",
                        os.path.join(ROOT, "animsnapbases_b200", "csrc"), os.path.join(ROOT, "tools", "host", "pd_host.cpp"), "-o", so],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lib = ctypes.CDLL(so)
    g = golden("pd_local")
    wi, (smin, smax) = float(g["wi"]), g["sigma"]

    def run(kind, nv, elems, rest, q, rows):
        ld = (nv + 31) // 32 * 32
        qp = np.zeros((1, 3, ld))
        qp[0, :, :nv] = q.T
        e = len(elems)
        ldo = (rows * e + 31) // 32 * 32
        out = np.zeros((1, 3, ldo))
        el = np.ascontiguousarray(elems, dtype=np.int32)
        rs = np.ascontiguousarray(rest, dtype=np.float64)
        vp = ctypes.c_void_p
        lib.pd_host(ctypes.c_int(kind), qp.ctypes.data_as(vp), ctypes.c_int64(ld), ctypes.c_int64(1), el.ctypes.data_as(vp),
                    rs.ctypes.data_as(vp), ctypes.c_int64(e), ctypes.c_double(smin), ctypes.c_double(smax), out.ctypes.data_as(vp),
                    ctypes.c_int64(ldo))
        return out[0, :, :rows * e].T

    dg = oracle_pd.TetDeformationGradient(g["bar_verts"], g["bar_tets"], wi)
    ts = oracle_pd.TriStrain(g["grid_verts"], g["grid_tris"], wi, smin, smax)
    es = oracle_pd.EdgeSpring(g["grid_verts"], g["grid_edges"], wi)
    nb, ng = len(g["bar_verts"]), len(g["grid_verts"])
    assert np.any(np.linalg.det(np.stack([g["bar_q"][g["bar_tets"][:, c]] - g["bar_q"][g["bar_tets"][:, 3]] for c in range(3)], axis=2)
                                @ dg.DmInv) < 0), "the fixture exercises the inverted-element branches"
    assert rel(run(0, nb, g["bar_tets"], dg.DmInv.reshape(-1, 9), g["bar_q"], 3), g["tet_dg_p"]) < 1e-11
    assert rel(run(1, nb, g["bar_tets"], dg.DmInv.reshape(-1, 9), g["bar_q"], 3), g["tet_strain_p"]) < 1e-11
    assert rel(run(2, ng, g["grid_tris"], np.concatenate([ts.P.reshape(-1, 6), ts.DmInv.reshape(-1, 4)], axis=1), g["grid_q"], 2),
               g["tri_strain_p"]) < 1e-11
    assert rel(run(3, ng, g["grid_edges"], es.d.reshape(-1, 1), g["grid_q"], 1), g["edge_spring_p"]) < 1e-13
    # vertex bending: rest data from the product's vectorised set-up, projection by the kernel's inline code
    from animsnapbases_b200.pd_local import bending_stars
    sv, sq = g["sphere_verts"], g["sphere_q"]
    verts, ptr, nbr, w, rest, _ = bending_stars(sv, g["sphere_tris"])
    ld = (len(sv) + 31) // 32 * 32
    qp = np.zeros((1, 3, ld))
    qp[0, :, :len(sv)] = sq.T
    out = np.zeros((1, 3, ld))
    arrs = [np.ascontiguousarray(a, dtype=t) for a, t in ((verts, np.int32), (ptr, np.int32), (nbr, np.int32), (w, np.float64),
                                                          (rest, np.float64))]
    vp = ctypes.c_void_p
    lib.pd_host_vertex_bending(qp.ctypes.data_as(vp), ctypes.c_int64(ld), ctypes.c_int64(1), *[a.ctypes.data_as(vp) for a in arrs],
                               ctypes.c_int64(len(verts)), ctypes.c_int(1), out.ctypes.data_as(vp), ctypes.c_int64(ld))
    assert rel(out[0, :, :len(verts)].T, g["vert_bending_p"]) < 1e-12


def test_zero_residual_host_logic():
    assert pipeline_checks.check_zero_residual(NumpyOps(), engine.Comm(enabled=False))


def test_bar_reduced_group_host_logic():
    """engine.reduced_group_operators / reduced_group_term over the numpy statement of the kernels == the reference's own
    prepare_reduced_group / get_group_reduced_term outputs (tests/golden/bar_reduced.npz)."""
    assert pipeline_checks.check_bar_reduced(NumpyOps(), engine.Comm(enabled=False))


def test_bar_geometric_deim_host_logic():
    """engine.deim_geom over the numpy statement of the kernels == the reference's picks (tests/golden/bar_geom.npz)."""
    assert pipeline_checks.check_bar_geom(NumpyOps(), engine.Comm(enabled=False))


@pytest.mark.parametrize("world", [2, 5])
def test_two_rank_gloo_row_sharding(world):
    """world_size 2 (and 5: uneven shards, the last ranks short or empty) over gloo: sharded deflation + SPLOCS (with the
    geodesic all-to-all) + POD + DEIM + geometric DEIM + reconstruction reproduce the single-process golden results."""
    port_no = str(29613 + world)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=port_no, PYTHONPATH=ROOT + os.pathsep + os.path.join(ROOT, "tests"))
    if world == 2:
        env["ASB_DEFLATE_PREDICT"] = "1"            # 5 ranks predict by default (>= 4); cover the forced path on 2 as well
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
                        "127.0.0.1", "--master-port", port_no, os.path.join(ROOT, "tests", "gloo_worker.py")],
                       capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert all(f"RANK{i} OK" in r.stdout for i in range(world))


def test_two_rank_gloo_mirror_classes(tmp_path):
    """The drop-in classes under torchrun (gloo, 2 ranks): main.py's two protocols with the rows sharded, host attribute views
    complete on every rank, files written by rank 0 and visible to every rank when the writer returns."""
    port_no = "29671"
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=port_no, PYTHONPATH=ROOT + os.pathsep + os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", port_no, os.path.join(ROOT, "tests", "gloo_api_worker.py"), str(tmp_path)],
                       capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert all(f"RANK{i} OK" in r.stdout for i in range(2))


@pytest.mark.parametrize("nx,leaf,ground", [(40, 16, None), (40, 16, 3), (7, 64, None)])
def test_multifrontal_symbolic_and_numeric(nx, leaf, ground):
    """Host side of animsnapbases_b200/mfsolve.py (ordering, symbolic maps, batched numeric factorisation with torch on the
    CPU) driven through the numpy statement of the two solve kernels, against SuperLU."""
    from scipy import sparse
    from scipy.sparse.linalg import splu
    from animsnapbases_b200 import synth
    from animsnapbases_b200.geodesic import cotan_laplacian
    from animsnapbases_b200.mfsolve import MultifrontalSolver
    from numpy_ops import NumpyOps
    verts, tris = synth.grid_mesh(nx, nx + 3)
    n = len(verts)
    Lc, va = cotan_laplacian(verts, tris)
    A = (sparse.diags(va) - 10.0 * float(np.mean(va)) * Lc).tocsc() if ground is None else (-Lc).tocsc()
    mf = MultifrontalSolver(NumpyOps(), A, verts, leaf_size=leaf, ground=ground)
    rng = np.random.default_rng(0)
    rhs = np.zeros((2, mf.ldx))
    if ground is None:
        rhs[[0, 1], [n // 3, n - 2]] = 1.0
        x = mf.solve(torch.from_numpy(rhs)).numpy()[:, :n]
        lu = splu(A)
        ref = np.stack([lu.solve(rhs[b, :n]) for b in range(2)])
        assert np.max(np.abs(x - ref) / ref) < 1e-12
    else:
        d = rng.standard_normal((2, n))
        rhs[:, :n] = d - d.mean(axis=1, keepdims=True)
        x = mf.solve(torch.from_numpy(rhs)).numpy()[:, :n]
        assert np.linalg.norm((A @ x.T).T - rhs[:, :n]) / np.linalg.norm(rhs) < 1e-12


def test_pos_components_protocol_host_logic(tmp_path):
    """posSnapshots / posComponents drop-in protocol with the numpy statement of the kernels (host logic only)."""
    import api_checks
    assert api_checks.check_pos_protocol(NumpyOps(), engine.Comm(enabled=False), tmp_path, extended=True)


def test_constraints_components_protocol_host_logic(tmp_path):
    """nonlinearSnapshots / constraintsComponents drop-in protocol (main.py:106-142) with the numpy statement of the kernels."""
    import api_checks
    assert api_checks.check_constraints_protocol(NumpyOps(), engine.Comm(enabled=False), tmp_path, extended=True)


def test_preprocess_and_off_sequence_ingest(tmp_path):
    """utils/process.py:46-148: OFF sequence -> preprocessed animation file.  The normalisation is checked
    against the oracle restatement, the filtering (degenerate triangle, small disconnected component) structurally."""
    from animsnapbases_b200 import hostio, synth
    from oracle import port
    v0, tris = synth.grid_mesh(6, 5)
    n = len(v0)
    island = np.array([[5.0, 5.0, 0.0], [5.1, 5.0, 0.0], [5.0, 5.1, 0.0]])          # second, smaller component
    verts0 = np.concatenate([v0, island])
    tris_all = np.concatenate([tris, [[n, n + 1, n + 2]], [[0, 0, 1]]])               # + one zero-area triangle
    rng = np.random.default_rng(5)
    frames = np.stack([verts0 + 0.01 * f * rng.standard_normal(verts0.shape) for f in range(4)]).astype(np.float32)
    for f, fr in enumerate(frames):
        with open(tmp_path / f"mesh_{f}.off", "w") as fh:
            fh.write("OFF\n# comment\n%d %d 0\n" % (len(fr), len(tris_all)))
            fh.writelines("%.9g %.9g %.9g\n" % tuple(p) for p in fr)
            fh.writelines("3 %d %d %d\n" % tuple(t) for t in tris_all)
    vv, ff = hostio.load_off(str(tmp_path / "mesh_2.off"), no_colors=True)
    assert np.allclose(vv, frames[2], rtol=1e-6) and np.array_equal(ff, tris_all)
    out = hostio.convert_sequence_to_hdf5(str(tmp_path / "mesh_*.off"), lambda p: hostio.load_off(p, no_colors=True),
                                          str(tmp_path / "anim.h5"), max_frames=10, icreament=1)
    verts, tr = hostio.read_animation(out)
    assert verts.shape == (4, n, 3) and verts.dtype == np.float32, "the 3-vertex island is dropped"
    assert np.array_equal(tr, tris), "degenerate and island triangles are dropped, indices unchanged for the kept block"
    ref, mean, scale = port.preprocess_unit_cube(np.array([hostio.load_off(str(tmp_path / f"mesh_{f}.off"), True)[0][:n]
                                                           for f in range(4)], np.float32))
    assert np.array_equal(verts, ref)
    ext = verts.max(axis=1) - verts.min(axis=1)
    assert abs(ext.max() - 1.0) < 1e-6 and np.abs(verts.mean(axis=(0, 1))).max() < 1e-6


def test_sparse_operator_files(tmp_path):
    """utils/utils.py:289-323: the `.bin` triplet container of the differential operator St and the pickled
    `.npz` variant; checked against the unmodified reference reader when /root/reference is mounted."""
    from scipy import sparse
    from animsnapbases_b200 import hostio
    rng = np.random.default_rng(8)
    A = sparse.random(37, 53, density=0.08, random_state=3, format="csr")
    path = str(tmp_path / "St.bin")
    hostio.write_sparse_matrix_to_bin(path, A)
    raw = open(path, "rb").read()
    assert len(raw) == 12 + 16 * A.nnz and np.frombuffer(raw[:12], "<i4").tolist() == [37, 53, A.nnz]
    B = hostio.read_sparse_matrix_from_bin(path)
    assert B.shape == A.shape and (abs(A - B)).nnz == 0
    assert (abs(hostio.read_sparse_matrix(path, ".bin") - A)).nnz == 0
    # duplicate triplets add up (COO -> CSR), as in the reference's csr_matrix((values, (rows, cols)))
    with open(path, "wb") as f:
        f.write(np.array([2, 2, 3], "<i4").tobytes())
        for r, c, v in ((0, 1, 1.5), (0, 1, 2.0), (1, 0, -1.0)):
            f.write(np.array([r, c], "<i4").tobytes() + np.array([v], "<f8").tobytes())
    assert np.array_equal(hostio.read_sparse_matrix_from_bin(path).toarray(), [[0.0, 3.5], [-1.0, 0.0]])
    np.savez(str(tmp_path / "assembly_ST.npz"), tets_deformation_gradient=A)
    got = hostio.read_sparse_matrix(str(tmp_path / "assembly_ST.npz"), ".npz", key="tets_deformation_gradient")
    assert (abs(got.item() - A)).nnz == 0
    with pytest.raises(ValueError):
        hostio.read_sparse_matrix(path, ".npz")
    with pytest.raises(ValueError):
        hostio.read_sparse_matrix(path, ".txt")
    from oracle import ref_shim
    if ref_shim.reference_available():
        ref_shim.load_reference()
        import utils.utils as ref_utils
        hostio.write_sparse_matrix_to_bin(path, A)
        assert (abs(ref_utils.read_sparse_matrix_from_bin(path) - A)).nnz == 0
    del rng


def test_header_is_plain_c_and_links(tmp_path):
    """include/asb.h is the boundary a non-Python host binds: it must compile as strict C99 and as C++17, and a program
    linked against libasb.so must resolve the entry points (no compute call: there is no GPU here)."""
    import shutil
    if shutil.which("gcc") is None or shutil.which("g++") is None:
        pytest.skip("no host compiler")
    lib_dir = os.path.join(ROOT, "animsnapbases_b200", "lib")
    if not os.path.exists(os.path.join(lib_dir, "libasb.so")):
        pytest.skip("libasb.so not built")
    src = tmp_path / "host.c"
    src.write_text('#include "asb.h"\n#include <stdio.h>\n'
                   'int main(void) {\n'
                   '    int (*solve)(const double*, int64_t, const double*, double*, double*, const double*, int64_t, int64_t, int64_t,\n'
                   '                 double, int64_t, asb_stream_t) = asb_admm_solve;   /* signature check against the header */\n'
                   '    const char* e = asb_last_error();\n'
                   '    printf("%d %d %s\\n", asb_version(), solve != 0, e ? "str" : "null");\n'
                   '    return 0;\n}\n')
    for cc, std, extra in (("gcc", "-std=c99", ["-pedantic"]), ("g++", "-std=c++17", ["-x", "c++"])):
        exe = tmp_path / ("host_" + cc.replace("+", "p"))
        r = subprocess.run([cc, std, "-Wall", "-Wextra", "-Werror", *extra, "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                            "-L", lib_dir, "-lasb", "-Wl,-rpath," + lib_dir], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
        assert out.returncode == 0 and out.stdout.split() == ["100", "1", "str"], out.stdout + out.stderr


def test_mirror_error_behaviour(tmp_path):
    """SURVEY 8(b) 'Errors': the drop-in classes keep the reference's exception types -- AssertionError for an unknown position
    basis type (posComponents.py:22), SystemExit(1) for an unknown rest shape (posSnapshots.py:93, nonlinear_snapshots.py:278),
    ValueError for an unknown constraint basis type (constraintsComponents.py:141,153), for local support on constraint
    bases (:376) and for unknown frames to reconstruct (:503); a missing mass file is a FileNotFoundError that names the absent
    third-party fallback (libigl) instead of a CPU path."""
    import api_checks
    from conftest import golden
    from animsnapbases_b200 import runtime
    from animsnapbases_b200.snapbases.constraintsComponents import constraintsComponents
    from animsnapbases_b200.snapbases.posComponents import posComponents
    runtime.install(NumpyOps(), engine.Comm(enabled=False))
    cwd = os.getcwd()
    os.chdir(str(tmp_path))
    try:
        g, b = golden("pos_grid"), golden("bar")
        with pytest.raises(AssertionError):
            posComponents(api_checks.pos_param(tmp_path, g, bases_type="NMF"))
        p = api_checks.pos_param(tmp_path, g)
        p.vertPos_rest_shape = "median"
        with pytest.raises(SystemExit) as ex:
            posComponents(p)
        assert ex.value.code == 1
        p = api_checks.pos_param(tmp_path, g)
        p.vertPos_masses_file = os.path.join(str(tmp_path), "absent.bin")
        with pytest.raises(FileNotFoundError, match="libigl"):
            posComponents(p)

        def constraints(**over):
            cp = api_checks.constraints_param(tmp_path, b, 6)
            for k, v in over.items():
                setattr(cp, k, v)
            cc = constraintsComponents(cp)
            cc.nonlinearSnapshots.config()
            cc.config(fileNameBases="p_nl_", fileName_geom_points="p_nl_interpol_points_")
            return cc
        cc = constraints(constProj_rest_shape="median")
        with pytest.raises(SystemExit) as ex:
            cc.nonlinearSnapshots.snapshots_prepare()
        assert ex.value.code == 1
        cc = constraints(constProj_basis_type="nmf")
        cc.nonlinearSnapshots.snapshots_prepare()
        with pytest.raises(ValueError):
            cc.compute_components_store_singvalues()
        cc = constraints(constProj_basis_type="pca_blocks", constProj_support="local")
        cc.nonlinearSnapshots.snapshots_prepare()
        with pytest.raises(ValueError, match="Local support"):
            cc.compute_components_store_singvalues()
        cc = constraints(constProj_basis_type="pca_blocks_with_St")
        cc.nonlinearSnapshots.snapshots_prepare()
        with pytest.raises(ValueError, match="differential operator St"):   # this fixture's config names no St file
            cc.compute_components_store_singvalues()
        cc = constraints()
        cc.nonlinearSnapshots.snapshots_prepare()
        cc.compute_components_store_singvalues()
        cc.post_process_components()
        cc.deim()
        with pytest.raises(ValueError, match="unknown frames"):
            cc.geom_constructed(3, "validation")
        cc = constraints(constProj_masses_file=os.path.join(str(tmp_path), "absent.bin"))
        with pytest.raises(FileNotFoundError, match="libigl"):
            cc.nonlinearSnapshots.snapshots_prepare()
    finally:
        os.chdir(cwd)
        runtime.install(None, None)


def test_ply_reader(tmp_path):
    """hostio.load_ply (main.py:52-58 `.ply` snapshots; the reference needs TVTK for it): ascii and binary files, extra vertex
    properties skipped, non-triangle faces refused."""
    from animsnapbases_b200 import hostio, synth
    verts, tris = synth.grid_mesh(5, 4)
    verts = verts.astype(np.float32)
    head = "ply\nformat %s 1.0\ncomment made by a test\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n" \
           "property uchar red\nelement face %d\nproperty list uchar int vertex_indices\nend_header\n"
    with open(tmp_path / "a.ply", "w") as f:
        f.write(head % ("ascii", len(verts), len(tris)))
        f.writelines("%.9g %.9g %.9g 7\n" % tuple(v) for v in verts)
        f.writelines("3 %d %d %d\n" % tuple(t) for t in tris)
    for fmt, end in (("binary_little_endian", "<"), ("binary_big_endian", ">")):
        with open(tmp_path / (fmt + ".ply"), "wb") as f:
            f.write((head % (fmt, len(verts), len(tris))).encode())
            vt = np.empty(len(verts), dtype=[("x", end + "f4"), ("y", end + "f4"), ("z", end + "f4"), ("red", "u1")])
            vt["x"], vt["y"], vt["z"], vt["red"] = verts[:, 0], verts[:, 1], verts[:, 2], 7
            f.write(vt.tobytes())
            ft = np.empty(len(tris), dtype=[("n", "u1"), ("v", end + "i4", (3,))])
            ft["n"], ft["v"] = 3, tris
            f.write(ft.tobytes())
    for name in ("a.ply", "binary_little_endian.ply", "binary_big_endian.ply"):
        v, t = hostio.load_ply(str(tmp_path / name))
        assert np.array_equal(v.astype(np.float32), verts) and np.array_equal(t, tris), name
    with open(tmp_path / "quad.ply", "w") as f:
        f.write(head % ("ascii", 4, 1))
        f.writelines("%d %d 0 7\n" % (i % 2, i // 2) for i in range(4))
        f.write("4 0 1 3 2\n")
    with pytest.raises(AssertionError):
        hostio.load_ply(str(tmp_path / "quad.ply"))


def test_tet_grid_generator():
    """synth.tet_grid (the block the snapshot-producer timing runs on): conforming, positively oriented, fills the box."""
    from animsnapbases_b200 import synth
    v, t = synth.tet_grid(3, 2, 4)
    assert v.shape == (4 * 3 * 5, 3) and t.shape == (6 * 24, 4) and len(np.unique(t)) == len(v)
    vol = np.einsum("ij,ij->i", np.cross(v[t[:, 1]] - v[t[:, 0]], v[t[:, 2]] - v[t[:, 0]]), v[t[:, 3]] - v[t[:, 0]]) / 6
    assert vol.min() > 0 and abs(vol.sum() - 24 / 4 ** 3) < 1e-12
    faces = np.sort(np.concatenate([t[:, [0, 1, 2]], t[:, [0, 1, 3]], t[:, [0, 2, 3]], t[:, [1, 2, 3]]]), axis=1)
    _, counts = np.unique(faces, axis=0, return_counts=True)
    assert set(counts) == {1, 2}, "every interior face is shared by exactly two tets"


# ---- BASELINE configs at their own sizes (tests/golden/scale_*.npz): host logic of the checks over the numpy statement
def test_scale_inputs_are_reproducible():
    """configs.build_inputs regenerates, bit for bit, the inputs the at-scale fixtures were computed on (IEEE-only arithmetic)."""
    from animsnapbases_b200 import configs
    for name in ("bunny_splocs", "bar_deim", "armadillo_splocs", "armadillo_splocs_f1000"):
        f = os.path.join(GOLDEN, f"scale_{name}.npz")
        if os.path.exists(f):
            assert configs.build_inputs(name)["digest"] == str(np.load(f)["digest"]), name


def test_scale_bar_end_to_end_numpy():
    """Full bar (3 120 tets, F = 70, K = 70): multi-pass Gram POD resolves sigma_1 / sigma_69 = 1e8, DEIM end to end."""
    import pipeline_checks
    rep = {}
    assert pipeline_checks.check_scale_bar(NumpyOps(), engine.Comm(enabled=False), rep)
    assert rep["sigma"] < 1e-10 and rep["modes"] < 1e-8


def test_scale_bunny_prefix_numpy():
    """First components of the bunny F = 200 fixture (the deflation prefix does not depend on K)."""
    import pipeline_checks
    assert pipeline_checks.check_scale_pos(NumpyOps(), engine.Comm(enabled=False), "bunny_splocs", K=6)


def test_bar_blocks_with_St_numpy():
    """pca_blocks_with_St against the unmodified reference's output (loop-counter row indexing included)."""
    import pipeline_checks
    assert pipeline_checks.check_bar_blocks_st(NumpyOps(), engine.Comm(enabled=False))


@pytest.mark.parametrize("name", ["pos_grid", "pos_sphere"])
def test_pos_orth_branch_numpy(name):
    """q_orthogonal branch of posComponents.post_process_components against the reference's scipy.linalg.orth output."""
    import pipeline_checks
    assert pipeline_checks.check_pos_orth(NumpyOps(), engine.Comm(enabled=False), name)


# ---- HDF5 containers without h5py: animsnapbases_b200/h5mini.py
def test_h5mini_roundtrip_and_layout(tmp_path):
    from animsnapbases_b200 import h5mini, hostio
    rng = np.random.default_rng(0)
    data = {"default": rng.standard_normal((7, 3)), "tris": rng.integers(0, 7, (9, 3)).astype(np.int32)}
    data.update({"comp%03d" % i: rng.standard_normal((7, 3)) for i in range(203)})        # more links than one default SNOD holds
    p = h5mini.write_datasets(str(tmp_path / "components.h5"), data, attrs={"scale": np.float64(2.5), "mean": np.arange(3.0)})
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0 and raw[13] == 8 and raw[14] == 8
    assert struct_u64(raw, 40) == len(raw), "end-of-file address of the superblock"
    out = h5mini.read_datasets(p)
    assert sorted(out) == sorted(data) and all(np.array_equal(out[k], v) and out[k].dtype == v.dtype for k, v in data.items())
    # the writers of the drop-in classes emit real HDF5 now
    q = hostio.write_components_animation(str(tmp_path / "c2.h5"), data["default"], data["tris"], [data["comp%03d" % i] for i in range(5)])
    assert h5mini.is_hdf5(q) and sorted(h5mini.read_datasets(q)) == ["comp000", "comp001", "comp002", "comp003", "comp004", "default", "tris"]
    with pytest.raises(h5mini.H5Error):
        open(tmp_path / "bad.h5", "wb").write(b"not hdf5 at all")
        h5mini.read_datasets(str(tmp_path / "bad.h5"))


def struct_u64(raw, off):
    import struct
    return struct.unpack_from("<Q", raw, off)[0]


@pytest.mark.parametrize("shuffle", [False, True])
def test_h5mini_reads_gzip_chunked_animation(tmp_path, shuffle):
    """An aligned-animation file laid out the way h5py writes `create_dataset('verts', ..., compression='gzip')`
    (utils/process.py:88-91): chunked storage, v1 B-tree chunk index, deflate (+ shuffle) filter pipeline, edge chunks."""
    import h5_fixture
    from animsnapbases_b200 import hostio
    rng = np.random.default_rng(1)
    verts = rng.standard_normal((5, 11, 3)).astype(np.float32)
    tris = rng.integers(0, 11, (13, 3))
    p = h5_fixture.write_chunked(str(tmp_path / "anim.h5"), verts, tris, chunk=(2, 4, 3), shuffle=shuffle)
    v, t = hostio.read_animation(p)
    assert v.dtype == np.float32 and np.array_equal(v, verts) and np.array_equal(t, tris)
