"""posComponents -- B200 mirror of snapbases/posComponents.py:18-356 (reference): same constructor argument (the
reference's `Config_parameters` object or anything carrying the same attributes), same method names, same attributes,
same files written.  Device work goes through animsnapbases_b200.engine -> libasb.so.

Reference behaviours kept on purpose:
  * in 'SPLOCS' mode the reference optimises local copies and never writes them back (posComponents.py:124-127,:181):
    `comps` / `weigs` stay the deflation result.  The optimised pair is exposed as `splocs_comps` / `splocs_weigs`
    (and written back only if `param.engine_splocs_writeback` is true).
  * post-processing adds the mean shape to every component (:282).
"""
from __future__ import annotations

import csv
import os

import numpy as np

from .. import engine, hostio, runtime
from ..hostio import log_time
from .posSnapshots import posSnapshots


class posComponents:  # Components == bases
    def __init__(self, param):
        self.basesType = param.vertPos_bases_type
        assert self.basesType == 'PCA' or self.basesType == 'SPLOCS'
        train = os.path.join(param.aligned_snapshots_directory, param.train_aligned_snapshots_animation_file)
        test = os.path.join(param.aligned_snapshots_directory, param.test_aligned_snapshots_animation_file)
        self.pos_snapshots = posSnapshots(train, test, param.vertPos_rest_shape, param.vertPos_masses_file,
                                          param.tet_mesh_file, param.q_standarize, param.q_massWeight)
        self.numComp = param.vertPos_numComponents
        self.support = param.q_support
        self.storeSingVal = param.store_vertPos_PCA_sing_val
        self.ortho_comps = None
        self.smooth_min_dist = param.vertPos_smooth_min_dist
        self.smooth_max_dist = param.vertPos_smooth_max_dist
        self.output_components_file = "components.h5"
        self.measures_at_largeDeforVerts = None
        self.fileNameBases = "q_pos_"
        self.param = param
        self.picked_vertices = None
        self.splocs_history = None
        self._C = self._W = None              # device: (K, 3, ld) planar components, (F, Kp) weights
        self._Cs = self._Ws = None            # device: SPLOCS-optimised pair
        self._host = {}

    # ------------------------------------------------------------------ host views (reference attribute names)
    def _aos(self, T):
        ops, comm = runtime.get()
        return engine.host_rows(ops, comm, T, self.pos_snapshots.shard)     # complete (all row shards) on every rank

    @property
    def comps(self):
        if self._C is None:
            return None
        if "comps" not in self._host:
            self._host["comps"] = self._aos(self._C)
        return self._host["comps"]

    @property
    def weigs(self):
        return None if self._W is None else self._W.cpu().numpy()[:, :self.numComp]

    @property
    def splocs_comps(self):
        return None if self._Cs is None else self._aos(self._Cs)

    @property
    def splocs_weigs(self):
        return None if self._Ws is None else self._Ws.cpu().numpy()[:, :self.numComp]

    def local_components_host(self, splocs=False):
        """This rank's row shard of `comps` (or of the SPLOCS-optimised components) as a host array (K, n_local, 3), through
        pinned memory.  `comps` itself is complete on every rank (an all-gather of host objects when the rows are sharded),
        which is a convenience for scripts written against the reference, not something a sharded pipeline wants to pay."""
        ops, _ = runtime.get()
        T = self._Cs if splocs else self._C
        return None if T is None else ops.to_host(ops.export_aos(T, None, self.pos_snapshots.shard.n))

    # ------------------------------------------------------------------ small host statics (API parity)
    @staticmethod
    def project_weight(x):
        x = np.maximum(0., x)
        m = x.max()
        return x if m == 0 else x / m

    @staticmethod
    def compute_support_map(idx, geodesics, min_dist, max_dist):
        phi = geodesics(idx)
        return (np.clip(phi, min_dist, max_dist) - min_dist) / (max_dist - min_dist)

    @staticmethod
    def prox_l1l2(Lambda, x, beta):
        xlen = np.sqrt((x ** 2).sum(axis=-1))
        with np.errstate(divide='ignore'):
            shrinkage = np.maximum(0.0, 1 - beta * Lambda / xlen)
        return x * shrinkage[..., np.newaxis]

    @staticmethod
    def frobenius_error(f, f_reconstructed):
        return np.linalg.norm((f - f_reconstructed).ravel())

    @staticmethod
    def relative_error_per_component(f, f_reconstructed):
        return [np.linalg.norm(f[:, :, i] - f_reconstructed[:, :, i]) / np.linalg.norm(f[:, :, i]) for i in range(3)]

    @staticmethod
    def max_pointwise_error(f, f_reconstructed):
        return np.max(np.abs(f - f_reconstructed)) / np.max(f)

    # ------------------------------------------------------------------ basis build
    @log_time("")
    def extract_k_components(self, writer, num_iters_max=20, num_admm_iterations=10):
        ops, comm = runtime.get()
        snaps = self.pos_snapshots
        snaps._geo.reset()                    # memoised support maps belong to one build (bounded: geodesic.cache_rows)
        local = self.support == 'local'
        block = int(getattr(self.param, "engine_deflation_block", 8))
        keep_r = False
        C, W, meas, picked, _ = engine.deflate(ops, comm, snaps.state.X, snaps.shard, self.numComp, local, snaps._geo,
                                               self.smooth_min_dist, self.smooth_max_dist, block=block, keep_residual=keep_r)
        self._C, self._W, self._host = C, W, {}
        m = meas.cpu().numpy()
        self.picked_vertices = picked.cpu().numpy()
        self.measures_at_largeDeforVerts = np.stack([np.arange(self.numComp, dtype=float), m[:, 0], m[:, 1]], axis=1)
        if self.storeSingVal and writer is not None:
            for k in range(self.numComp):
                writer.writerow([k, m[k, 0], m[k, 1]])
        if self.basesType == 'SPLOCS':
            self.splocs_glob_optimization(self.param.splocs_max_itrs, self.param.splocs_admm_num_itrs, None,
                                          snaps.compute_geodesic_distance)
        print("Computed '", self.basesType, "' bases size ", (self.numComp, snaps.nVerts, 3))

    @log_time("")
    def splocs_glob_optimization(self, num_iters_max, num_admm_iterations, R, compute_geodesic_distance):
        """posComponents.py:132-189 in Gram form (engine.splocs); prints the reference's per-iteration line."""
        ops, comm = runtime.get()
        snaps = self.pos_snapshots
        Cs, Ws = self._C.clone(), self._W.clone()
        hist, centres = engine.splocs(ops, comm, snaps.state.X, snaps.shard, Cs, Ws, self.numComp, snaps._geo,
                                      num_iters_max, num_admm_iterations, self.param.splocs_lambda, self.param.splocs_rho,
                                      self.smooth_min_dist, self.smooth_max_dist)
        self._Cs, self._Ws, self.splocs_history = Cs, Ws, hist
        if bool(getattr(self.param, "engine_splocs_writeback", False)):
            self._C, self._W, self._host = Cs, Ws, {}

    @log_time("")
    def compute_components_store_singvalues(self):
        headerSing = ['component', 'singVal', 'norm_R']
        file_name = os.path.join(self.param.vertPos_output_directory,
                                 self.param.name + "_posBases_pcaExtraction_singValues_errorNorm")
        _, comm = runtime.get()
        rows = []                                      # collected on every rank (collectives inside), written by rank 0

        class _Rows:
            writerow = staticmethod(rows.append)

        self.extract_k_components(_Rows if self.storeSingVal else None)
        if self.storeSingVal:
            def write():
                with open(file_name + '.csv', 'w', encoding='UTF8') as singFile:
                    writer = csv.writer(singFile)
                    writer.writerow(headerSing)
                    writer.writerows(rows)
            comm.on_rank0(write)

    # ------------------------------------------------------------------ post-processing
    @log_time("")
    def post_process_components(self):
        print("Post-processing pos components ...")
        ops, comm = runtime.get()
        snaps, sh = self.pos_snapshots, self.pos_snapshots.shard
        C = self._C
        if self.param.q_standarize:
            ops.unstandardize(C, snaps.state.mean, snaps.pre_scale_factor, sh.n)       # comps /= scale; comps += mean
        if self.param.q_orthogonal:
            rank = engine.orth_rows(ops, comm, C, sh, self.numComp)                    # scipy.linalg.orth per dimension
            if min(rank) < self.numComp:
                raise ValueError("could not broadcast: orth() dropped rank-deficient directions (rank %s < K = %d)"
                                 % (rank, self.numComp))
        if self.param.q_massWeight:
            assert snaps.nVerts == snaps.invMassL.shape[0]
            aos = ops.export_aos(C, snaps.state.invMassL, sh.n)                        # comps *= invMassL[:, None]
            self._C = C = ops.ingest(aos, None, sh.ld)
        self._host = {}
        sparsity, ranks = engine.sparsity_and_rank(ops, comm, C, sh, self.numComp)
        if min(sparsity) > 0.5:
            print("sparse, min %" + str(100 * min(sparsity)) + " zero entries.")
        else:
            print("... not sparse.")
        for r in ranks:
            if r == self.numComp:
                print(".. linear independent.")
            else:
                print("... not linear independent, with rank: " + str(r) + " != " + str(self.numComp) + ".")
        if self.param.q_orthogonal:
            self.is_utmu_orthogonal()
        print("... Volkwein (" + str(self.param.q_massWeight) + ")... standerized (" + str(self.param.q_standarize) +
              ")... support (" + str(self.support) + "), orthogonalized (" + str(self.param.q_orthogonal) + ").")

    @log_time("")
    def is_utmu_orthogonal(self):
        print('... testing M orthogonality, U^T M U = I (K x K) ...', end='', flush=True)
        ops, comm = runtime.get()
        snaps, sh = self.pos_snapshots, self.pos_snapshots.shard
        rows = slice(sh.row0, sh.row0 + sh.n)
        grams = engine.weighted_row_grams(ops, comm, self._C, sh, self.numComp, ops.to_device(np.sqrt(snaps.mass[rows])))
        for G in grams:
            assert np.allclose(G, np.eye(self.numComp))
        print('(True).')

    # ------------------------------------------------------------------ outputs
    @log_time("")
    def store_components_to_files(self, start, end, step, fileType):
        """fileType '.bin' or '.npy'; one file per k in range(start, end + 1, step), as main.py:96-100 asks."""
        print('Storing bases ...', end='', flush=True)
        ops, comm = runtime.get()
        snaps = self.pos_snapshots
        basesFile = os.path.join(self.param.vertPos_output_directory, self.fileNameBases)
        packed = ops.pack_bin(self._C, self.numComp, snaps.shard.n).cpu().numpy()     # (3, K, N): the .bin body order
        if comm.world > 1:
            packed = engine.gather_rows_host(comm, packed, snaps.shard, axis=2)
        def write():
            for k in range(start, end + 1, step):
                hostio.store_components_packed(basesFile, snaps.frs, k, snaps.nVerts, 3, packed, fileType, 'K')
        comm.on_rank0(write)
        print('done.')

    @log_time("")
    def store_animations(self, output_bases_dir):
        _, comm = runtime.get()
        out = os.path.join(output_bases_dir, self.output_components_file)
        comps = self.comps
        comm.on_rank0(lambda: hostio.write_components_animation(out, self.pos_snapshots.verts[0], self.pos_snapshots.tris, comps))

    @log_time("")
    def test_basesSingVals(self):
        """(K, 3): singular values of comps[:, :, d] normalised by the largest, per dimension (posComponents.py:344-356)."""
        ops, comm = runtime.get()
        s = np.empty((self.numComp, 3))
        for d, sig in enumerate(engine.row_singular_values(ops, comm, self._C, self.pos_snapshots.shard, self.numComp)):
            s[:, d] = sig / sig.max()
        return s

    @log_time("")
    def test_convergence(self, start, end, step, writer=None):
        """Reconstruction errors of the first k components, k in range(start, end + 1, step) (posComponents.py:192-214)."""
        ops, comm = runtime.get()
        snaps = self.pos_snapshots
        return engine.reconstruction_errors(ops, comm, snaps.state.X, snaps.shard, self._C, self._W, range(start, end + 1, step))
