"""constraintsComponents -- B200 mirror of snapbases/constraintsComponents.py:26-860 (reference), hot-path part.

Same protocol as main.py:106-142 drives:  `.nonlinearSnapshots.config()`, `.config(...)`,
`.nonlinearSnapshots.snapshots_prepare()`, `.compute_components_store_singvalues()`, `.post_process_components()`,
`.deim()`, `.store_components_n_interpol_points()` / `.store_components_gradually_to_files(...)`.
Device work goes through animsnapbases_b200.engine -> libasb.so:

  pod_vectorized  (:298-320)  Gram X X^T on the DMMA pipe -> F x F eigensolve -> back-projection GEMM      engine.pod
  post-process    (:415-446)  unscale + mean, per-dimension economic QR (CGS2), M^-1/2 row scaling         engine.orthogonalize_rows
  deim            (:797-860)  greedy point selection, one index set for x/y/z, bit-exact first-max argmax   engine.deim

  pca_blocks      (:324-412)  block deflation: element energy argmax + p rank-1 steps per block               engine.deflate_blocks
  deim_blocksForm (:733-795)  block DEIM, p rows per pick                                                   engine.deim_blocks

  pod             (:274-294)  p * 3 independent PODs of the (e x F) blocks (f32 torch-CPU SVDs in the reference)    engine.pod_per_row

  pca_blocks_with_St (:156-271)  block deflation picked in position space through St, until |R| <= bases_R_tol; keeps the
                              reference's loop-counter row indexing                              engine.deflate_blocks_with_St
The plotting / polyscope methods of the reference (:865-1156) are out of scope (graphics: unavailable on this box).
"""
from __future__ import annotations

import csv
import os

import numpy as np

from .. import engine, hostio, runtime
from ..hostio import log_time
from .nonlinear_snapshots import nonlinearSnapshots


class constraintsComponents:  # Components == bases
    def __init__(self, param):
        self.basesType = ""
        self.numComp = 0
        self.support = ""
        self.storeSingVal = False
        self.nonlinearSnapshots = nonlinearSnapshots(param)
        self.param = param
        self._W = None
        self.ortho_comps = None
        self.largeDeforPoints = None
        self.largeDeforBlocks = None
        self.measures_at_largeDeforVerts = None
        self.fileNameBases = ""
        self.fileName_geom_points = ""
        self.file_name_sing = ""
        self.geom_interpol_verts = []
        self.geom_alpha = None
        self.geom_alpha_ranges = None
        self.geom_Pt = None
        self.geom_ep_idxs = None
        self.St = None
        self.singular_values = None
        self._C = None               # device (K, 3, ld) planar components of this rank's row shard
        self._host = {}

    # ------------------------------------------------------------------ host view (reference attribute name)
    @property
    def comps(self):
        if self._C is None:
            return None
        if "comps" not in self._host:
            ops, comm = runtime.get()
            self._host["comps"] = engine.host_rows(ops, comm, self._C, self.nonlinearSnapshots.shard)
        return self._host["comps"]

    @property
    def weigs(self):
        return None if self._W is None else self._W.cpu().numpy()[:, :self._C.shape[0]]

    @property
    def _rows(self):
        """Number of component rows: numComp for POD, numComp * p for block bases."""
        return 0 if self._C is None else self._C.shape[0]

    def config(self, fileNameBases="p_nl_", fileName_geom_points="p_nl_interpol_points_",
               file_name_sing="_constrprojBases_pcaExtraction_singValues"):
        p = self.param
        self.basesType = p.constProj_bases_interpolation_type
        self.support = p.constProj_support
        self.storeSingVal = p.constProj_store_sing_val
        self.fileNameBases = fileNameBases
        self.fileName_geom_points = fileName_geom_points
        self.file_name_sing = file_name_sing
        # the differential operator St is only used by the geometric / St variants (a11, a14): load it when present
        st_file = getattr(p, "constProj_weightedSt", None)
        if st_file and os.path.exists(st_file):
            data = np.load(st_file, allow_pickle=True)
            St = data[getattr(p, "costProj_St_key", data.files[0])]
            if isinstance(St, np.ndarray) and St.dtype == object:
                St = St.item()
            self.St = St

    @staticmethod
    def project_weight(x):
        x = np.maximum(0., x)
        m = x.max()
        return x if m == 0 else x / m

    @staticmethod
    def indxLargestDeformation(R, p, e):
        """Host statement of the element pick (:86-96; R (ep, F, d)): kept for callers of the static; the build itself picks
        on the device (asb_vertex_energy + asb_group_sum + asb_argmax_first, same first-max rule)."""
        magnitude = (R ** 2).sum(axis=2).reshape(e, p, -1)
        return np.argmax(magnitude.sum(axis=(1, 2)))

    @staticmethod
    def indxLargestDeformation_e(r, e, p, d):
        assert r.shape == (e, p * d)
        return np.argmax((r ** 2).sum(axis=1))

    # ------------------------------------------------------------------ basis build
    @log_time("")
    def compute_components_store_singvalues(self):
        p = self.param
        kind = p.constProj_basis_type
        if kind in ("pca_blocks", "pca_blocks_with_St"):
            headerSing = ['component', 'idx', 'residual_matrix_norm'] + \
                         ['singVal' + str(i) for i in range(self.nonlinearSnapshots.constraintsSize)]
        else:
            headerSing = ['component', 'singVal']
        if kind not in ("pod", "pod_vectorized", "pca_blocks", "pca_blocks_with_St"):
            raise ValueError("Uknown basis type: ", kind)
        compute = {"pod_vectorized": self.compute_pod_for_vectorized_nonlinear_snapshots_tensor,
                   "pod": self.compute_pod_for_nonlinear_snapshots_tensor,
                   "pca_blocks": self.compute_nonlinearity_bases_blocks,
                   "pca_blocks_with_St": self.compute_nonlinearity_bases_blocks_utilizing_diffirential_operator}[kind]
        _, comm = runtime.get()
        file_name = os.path.join(p.constProj_output_directory, p.name + "_" + p.constProj_name + self.file_name_sing)
        rows = []                                      # collected on every rank (collectives inside), written by rank 0

        class _Rows:
            writerow = staticmethod(rows.append)

        compute(_Rows if self.storeSingVal else None)
        if self.storeSingVal:
            def write():
                with open(file_name + '.csv', 'w', encoding='UTF8') as singFile:
                    writer = csv.writer(singFile)
                    writer.writerow(headerSing)
                    writer.writerows(rows)
            comm.on_rank0(write)

    @log_time("")
    def compute_nonlinearity_bases_blocks(self, writer=None):
        """Block deflation (:324-412): K blocks of p rank-1 components; numComp counts BLOCKS, comps has K p rows."""
        if self.support == 'local':
            raise ValueError(' Local support is not yet available for nonlinearity')
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        p = snaps.constraintsSize
        C, W, table, elems, rows, _ = engine.deflate_blocks(ops, comm, snaps.state.X, snaps.shard, p,
                                                            int(self.param.deim_desired_num_components),
                                                            block=int(getattr(self.param, "engine_deflation_block", 8)))
        self._C, self._W, self._host = C, W, {}
        self.measures_at_largeDeforVerts = table
        self.largeDeforPoints, self.largeDeforBlocks = elems, rows
        self.numComp = C.shape[0] // p
        if writer is not None:
            for r in table:
                writer.writerow([int(r[0]), int(r[1])] + [float(v) for v in r[2:]])
        print("bases shape", (C.shape[0], snaps.shard.n_global, 3), "number of components", self.numComp)

    @log_time("")
    def compute_nonlinearity_bases_blocks_utilizing_diffirential_operator(self, writer=None):
        """`pca_blocks_with_St` (:156-271): block deflation whose picks are made in position space through St, until the
        residual norm drops under param.bases_R_tol (engine.deflate_blocks_with_St keeps the loop-counter row indexing)."""
        if self.support == 'local':
            raise ValueError("Local support maps are not yet available for nonlinear-term components")
        if self.St is None:
            raise ValueError("basis type 'pca_blocks_with_St' needs the differential operator St (constProj_weightedSt)")
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        p = snaps.constraintsSize
        elems = {"_tets": snaps.tets, "_tris": snaps.tris, "_edges": snaps.edges}.get(snaps.ele_type)
        if elems is None:
            print("ERROR! unknown constraints projection type")
            return
        C, W, table, verts = engine.deflate_blocks_with_St(ops, comm, snaps.state.X, snaps.shard, p, self.St, elems,
                                                           float(self.param.bases_R_tol))
        self._C, self._W, self._host = C, W, {}
        self.measures_at_largeDeforVerts = table
        self.largeDeforPoints = verts
        self.numComp = C.shape[0] // p
        if writer is not None:
            for r in table:
                writer.writerow([int(r[0]), int(r[1])] + [float(x) for x in r[2:]])
        print("bases shape", (C.shape[0], snaps.shard.n_global, 3), "number of components", self.numComp)

    @log_time("")
    def compute_pod_for_nonlinear_snapshots_tensor(self, writer=None):
        """`pod` (:274-294): p * 3 independent PODs of the (e x F) blocks (float32 torch-CPU SVDs in the reference)."""
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        C, sig = engine.pod_per_row(ops, comm, snaps.state.X, snaps.shard, snaps.constraintsSize,
                                    int(self.param.deim_desired_num_components))
        self.singular_values = sig
        self._C, self._host = C, {}
        self.numComp = C.shape[0]

    @log_time("")
    def compute_pod_for_vectorized_nonlinear_snapshots_tensor(self, writer=None):
        """Thin SVD of the (3 e p) x F snapshot matrix (:298-320) through its F x F Gram."""
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        F = snaps.state.X.shape[0]
        k_desired = min(int(self.param.deim_desired_num_components), F)
        C, sig = engine.pod(ops, comm, snaps.state.X, snaps.shard, k_desired)
        self.singular_values = sig
        if writer is not None:
            for ai, bi in zip(range(1, sig.shape[0] + 1), sig):
                writer.writerow([ai, bi])
        self._C, self._host = C, {}
        self.numComp = C.shape[0]

    # ------------------------------------------------------------------ post-processing
    @log_time("")
    def post_process_components(self):
        ops, comm = runtime.get()
        p, snaps, sh = self.param, self.nonlinearSnapshots, self.nonlinearSnapshots.shard
        C, X = self._C, snaps.state.X
        if p.constProj_standarize:
            ops.unstandardize(C, snaps.state.mean, snaps.pre_scale_factor, sh.n)   # comps /= scale; comps += mean
            ops.unstandardize(X, snaps.state.mean, snaps.pre_scale_factor, sh.n)   # the snapshots too (:426-430)
        if p.constProj_orthogonal:
            engine.orthogonalize_rows(ops, comm, C, sh, self._rows)                # economic QR per dimension
        if p.constProj_massWeight:
            assert snaps.invMassL.shape[0] == sh.n_global
            self._C = ops.ingest(ops.export_aos(C, snaps.state.invMassL, sh.n), None, sh.ld)
            snaps.state.X = ops.ingest(ops.export_aos(X, snaps.state.invMassL, sh.n), None, sh.ld)
        self._host = {}
        snaps.invalidate_host_views()
        print("Post-processing, Undo standardization: ", p.constProj_standarize, ". Orthogonal-ized", p.constProj_orthogonal,
              ". Mass weighting", p.constProj_massWeight, ", and bases shape",
              (self._rows, sh.n_global, 3))

    # ------------------------------------------------------------------ DEIM
    @log_time("")
    def deim(self):
        """Classical DEIM (:797-860): geom_Pt row ids, geom_alpha = Pt // p, geom_alpha_ranges = 1..K."""
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        p, K = snaps.constraintsSize, self.numComp
        Pt = engine.deim(ops, comm, self._C, snaps.shard, K)
        if Pt is None:                       # zero residual: the reference prints and returns (:832-834)
            return
        self.geom_Pt = np.array(Pt)
        self.geom_alpha = self.geom_Pt // p
        self.geom_ep_idxs = np.array(set(int(i) for i in Pt))       # 0-d object array, as the reference stores it (:856)
        self.geom_alpha_ranges = np.arange(1, K + 1)
        self.geom_interpol_verts = np.array(self.geom_interpol_verts)
        for k in range(K):
            print(k, int(self.geom_alpha[k]))
        print("Regular Deim interpolation, used", self.geom_alpha.shape[0], "constrained elements")

    @log_time("")
    def deim_blocksForm(self):
        """Block DEIM (:733-795): p rows per pick; geom_Pt has numComp * p row ids."""
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        out = engine.deim_blocks(ops, comm, self._C, snaps.shard, snaps.constraintsSize)
        if out is None:
            return
        self.geom_Pt, self.geom_alpha, self.geom_alpha_ranges = out["Pt"], out["alpha"], out["alpha_ranges"]
        self.geom_ep_idxs = np.array(set(int(i) for i in out["idxs"]))
        self.geom_interpol_verts = np.array(self.geom_interpol_verts)
        for k in range(self.geom_alpha.shape[0]):
            print(k, int(self.geom_alpha[k]))
        print("Regular Deim interpolation, used", self.geom_alpha.shape[0], "constrained elements")

    @log_time("")
    def geom_block_form_utilizing_differential_operator(self, error_in_pos_space=False):
        """Geometric block DEIM (:619-731): residual measured in position space through the differential operator St (worst
        vertex -> up to geom_ele_per_vert of its incident elements) or in constraint space (worst element)."""
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        elems = {"_tets": snaps.tets, "_tris": snaps.tris, "_edges": snaps.edges}.get(snaps.ele_type)
        St = star_faces = constrained = None
        if error_in_pos_space:
            if snaps.ele_type not in ["_tets", "_tris", "_edges", "_verts"]:
                print("ERROR! Unknown constained elements nonliner snapshots type!")
                return
            if self.St is None:
                raise ValueError("geometric DEIM in position space needs the differential operator St (constProj_weightedSt)")
            St = self.St
            if snaps.ele_type == "_verts":                      # vertex stars (:687-692), index mapping for verts_bending
                star_faces = snaps.tris
                if self.param.constProj_snapshots_type == "verts_bending":
                    self.constrianed_verts = np.load(self.param.constProj_input_snaps_constrained_elements)["indices"]
                    constrained = self.constrianed_verts
        out = engine.deim_geom(ops, comm, self._C, snaps.shard, snaps.constraintsSize, elems, St,
                               int(getattr(self.param, "geom_ele_per_vert", 1)), star_faces, constrained)
        if out is None:
            return
        self.geom_Pt, self.geom_alpha, self.geom_alpha_ranges = out["Pt"], out["alpha"], out["alpha_ranges"]
        self.geom_interpol_verts = out["verts"]
        print("Computing interpolation elements utilizing differential operator, used", self.geom_alpha.shape[0],
              "constrained elements")

    def _geom_constructed_device(self, r, case):
        """(frames, reconstruction) planar on the device for geom_constructed / geom_reconstruction_errors."""
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        p = snaps.constraintsSize if self.param.constProj_bases_interpolation_type in ("geom", "deim_block_form") else 1
        sh = snaps.shard
        if case == "train":
            X = snaps.state.X
        elif case == "test":
            rows = slice(sh.row0, sh.row0 + sh.n)
            X = ops.ingest(ops.to_device(np.ascontiguousarray(snaps.test_snapTensor[:, rows])), None, sh.ld)
        else:
            raise ValueError("unknown frames to reconstruct.")
        if getattr(self.param, "constProj_snapshots_type", "") == "verts_bending":
            Pt = self.geom_Pt[:self.geom_alpha_ranges[r - 1]]
        else:
            Pt = self.geom_alpha[:self.geom_alpha_ranges[r - 1]]
        return X, engine.deim_reconstruct(ops, comm, self._C, X, sh, Pt, r * p)

    @log_time("")
    def geom_constructed(self, r, case, interpol="geom"):
        """DEIM reconstruction of the train / test frames from r interpolation elements (:489-521); returns the host
        (F, ep, 3) array the figure scripts feed to the error statics.  Keeps the reference's row choice (geom_alpha, i.e.
        element ids used as row ids unless constProj_snapshots_type == 'verts_bending', :507-512)."""
        ops, comm = runtime.get()
        sh = self.nonlinearSnapshots.shard
        _, rec = self._geom_constructed_device(r, case)
        out = ops.export_aos(rec, None, sh.n).cpu().numpy()
        if comm.world > 1:
            out = engine.gather_rows_host(comm, out, sh, axis=1)
        return out

    def geom_reconstruction_errors(self, r, case):
        """(frobenius_error, max_pointwise_error, relative_error_per_component) of geom_constructed(r, case) against the
        frames (generate_figures/nl_reduction_tests.py:137-140) with both tensors left on the device."""
        _, comm = runtime.get()
        X, rec = self._geom_constructed_device(r, case)
        return engine.tensor_error_measures(comm, X, rec, self.nonlinearSnapshots.shard)

    # ------------------------------------------------------------------ checks
    @log_time("")
    def test_basesSingVals(self):
        """(Kp, 3): singular values of comps[:, :, d] normalised by the largest, per dimension (:555-569)."""
        ops, comm = runtime.get()
        s = np.empty((self._rows, 3))
        for d, sig in enumerate(engine.row_singular_values(ops, comm, self._C, self.nonlinearSnapshots.shard, self._rows)):
            s[:, d] = sig / sig.max()
        return s

    def is_utmu_orthogonal(self):
        """Prints 'True..' / 'False..' per dimension for U^T M U = I (:450-459; prints, does not assert): mass-weighted
        Kp x Kp Grams on the DMMA pipe."""
        print('... testing M orthogonality, U^T M U = I (Kp x Kp) ...', end='', flush=True)
        ops, comm = runtime.get()
        snaps, sh = self.nonlinearSnapshots, self.nonlinearSnapshots.shard
        rows = slice(sh.row0, sh.row0 + sh.n)
        verdicts = []
        for G in engine.weighted_row_grams(ops, comm, self._C, sh, self._rows, ops.to_device(np.sqrt(snaps.mass[rows]))):
            verdicts.append(bool(np.allclose(G, np.eye(self._rows))))
            print('True..' if verdicts[-1] else "False..")
        return verdicts

    @staticmethod
    def frobenius_error(f, f_reconstructed):
        return np.linalg.norm(f - f_reconstructed)

    @staticmethod
    def relative_error_per_component(f, f_reconstructed):
        return [np.linalg.norm(f[:, :, i] - f_reconstructed[:, :, i]) / np.linalg.norm(f[:, :, i]) for i in range(3)]

    @staticmethod
    def max_pointwise_error(f, f_reconstructed):
        return np.max(np.abs(f - f_reconstructed)) / np.max(f)

    # ------------------------------------------------------------------ outputs
    def _gathered_comps(self):
        return self.comps                    # complete (all row shards) on every rank

    @log_time("")
    def store_components_gradually_to_files(self, start, end, step, fileType):
        """'.bin' or '.npy', one file per k (:572-593): bases (k*p components), interpolation elements, vertices."""
        print('Storing bases ...', end='', flush=True)
        ops, comm = runtime.get()
        snaps = self.nonlinearSnapshots
        p = snaps.constraintsSize
        out = self.param.constProj_output_directory
        basesFile = os.path.join(out, self.fileNameBases)
        pointsFile = os.path.join(out, self.fileName_geom_points)
        vertsFile = os.path.join(out, "corrVerts")
        numverts = snaps.num_constained_elements * p
        packed = ops.pack_bin(self._C, self._rows, snaps.shard.n).cpu().numpy()        # (3, Kp, ep): the .bin body order
        if comm.world > 1:
            packed = engine.gather_rows_host(comm, packed, snaps.shard, axis=2)
        def write():
            for k in range(start, end + 1, step):
                hostio.store_components_packed(basesFile, snaps.frs, k * p, numverts, 3, packed, fileType, 'Kp')
                hostio.store_interpol_points_vector(pointsFile, snaps.frs, k,
                                                    self.geom_alpha[:self.geom_alpha_ranges[k - 1]], fileType)
                hostio.store_interpol_points_vector(vertsFile, snaps.frs, k, self.geom_interpol_verts[:k], fileType)
        comm.on_rank0(write)
        print('done.')

    @log_time("")
    def store_components_n_interpol_points(self):
        """One .npz with the keys projective_dynamics/Simulators.py:179-188 reads (:596-611)."""
        print('Storing bases and interpolation points to one file ...', end='', flush=True)
        _, comm = runtime.get()
        comps = self._gathered_comps()
        comm.on_rank0(lambda: np.savez(
            os.path.join(self.param.constProj_output_directory, "components_interpol_alphas_interpol_verts_interpol_alpha_ranges.npz"),
            components=comps, interpol_alphas=self.geom_alpha, Pt=self.geom_Pt,
            interpol_verts=self.geom_interpol_verts, interpol_alpha_ranges=self.geom_alpha_ranges))
        print("done!")
