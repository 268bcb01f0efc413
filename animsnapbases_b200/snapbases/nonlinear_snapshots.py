"""nonlinearSnapshots -- device-resident mirror of snapbases/nonlinear_snapshots.py:17-288 (reference).

Same constructor argument (the reference's `Config_parameters` or any object with the same attributes), same protocol
(`config()` then `snapshots_prepare()`), same attributes.  The (F, ep, 3) constraint-projection tensor is read on the host
(one vectorised pass instead of the reference's O(F^2) `np.concatenate` growth, :109-123), its row shard is handed to the
device once, and mass weighting / rest shape / population-std standardisation run in libasb.so kernels
(engine.prepare_snapshots).  Host numpy views (`snapTensor`, `mean`) are materialised lazily.
"""
from __future__ import annotations

import os
import struct
import sys

import numpy as np

from .. import engine, runtime
from ..hostio import log_time, read_mass_bin


def element_masses_from_vertex_masses(vertex_mass, elements, p):
    """compute_tetMasses / compute_triMasses / compute_edgeMasses (utils/support.py:12-38, :62-74): the mass of a
    constrained element is the sum of its vertex masses, replicated over its p constraint rows."""
    w = np.asarray(vertex_mass, dtype=np.float64)[np.asarray(elements)].sum(axis=1)
    return np.repeat(w, p)


class nonlinearSnapshots:
    def __init__(self, param):
        self.snapshots_file = ""
        self.rest_shape = ""
        self.dim = 0
        self.mass_file = ""
        self.frs = 0
        self.constraintsSize = 0
        self.num_constained_elements = 0
        self.pre_scale_factor = 1
        self.mass = None
        self.massL = None
        self.invMassL = None
        self.test_snapTensor = None
        self.verts = None
        self.tris = None
        self.tets = None
        self.edges = None
        self.tet_mesh = None
        self.ele_type = ""
        self.param = param
        self.state = None            # engine.SnapshotState: this rank's row shard on the device
        self.shard = None
        self._raw = None             # host (F, ep, 3) as read
        self._host = {}

    # ---- lazily materialised host views (reference attribute names)
    @property
    def snapTensor(self):
        if self.state is None:
            return self._raw
        if "snapTensor" not in self._host:
            ops, comm = runtime.get()
            self._host["snapTensor"] = engine.host_rows(ops, comm, self.state.X, self.shard)
        return self._host["snapTensor"]

    @property
    def mean(self):
        if self.state is None or self.state.mean is None:
            return None
        ops, comm = runtime.get()
        return engine.host_rows(ops, comm, self.state.mean[None], self.shard)[0]

    def invalidate_host_views(self):
        self._host = {}

    def config(self):
        p = self.param
        self.snapshots_file = p.constProj_input_snapshots_pattern
        self.rest_shape = p.constProj_rest_shape
        self.dim = p.constProj_dim
        self.mass_file = p.constProj_masses_file
        self.frs = p.constProj_numFrames
        self.constraintsSize = p.constProj_p_size
        self.ele_type = p.constProj_element_type
        self.tet_mesh = p.tet_mesh_file
        self.tri_mesh = getattr(p, "tri_mesh_file", None)

    @log_time("")
    def snapshots_prepare(self):
        ops, comm = runtime.get()
        self.read()
        p = self.param
        if p.constProj_massWeight:
            self.load_factorize_masses()
            assert self._raw.shape[1] == self.massL.shape[0]
        if p.constProj_standarize and self.rest_shape not in ("first", "average"):
            print('Error! unknown rest shape: ', self.rest_shape)
            sys.exit(1)
        ep = self._raw.shape[1]
        mult = int(np.lcm(32, max(int(self.constraintsSize), 1)))       # whole elements per shard (block deflation / DEIM)
        self.shard = engine.Shard.make(ep, comm.world, comm.rank, multiple=mult)
        rows = slice(self.shard.row0, self.shard.row0 + self.shard.n)
        src = ops.to_device(np.ascontiguousarray(self._raw[:, rows]))
        rest = self.rest_shape if p.constProj_standarize else "first"
        self.state = engine.prepare_snapshots(ops, comm, src, self.mass[rows] if p.constProj_massWeight else None, rest,
                                              self.shard, standarize=bool(p.constProj_standarize),
                                              mass_weight=bool(p.constProj_massWeight))
        if not p.constProj_standarize:
            self.state.mean = None
        self.pre_scale_factor = self.state.scale
        self._host = {}
        print('nonlinearSnapshots ready ... Volkwein (' + str(p.constProj_massWeight) + '), standarized (' +
              str(p.constProj_standarize) + ').')

    # ------------------------------------------------------------------ host I/O
    @staticmethod
    def _read_bin_matrix(path):
        """`<i ni><i mi>` then ni*mi little-endian doubles, column-major (nonlinear_snapshots.py:128-139)."""
        with open(path, "rb") as f:
            ni, mi = struct.unpack('<ii', f.read(8))
            body = np.frombuffer(f.read(8 * ni * mi), dtype='<f8')
        return body.reshape(mi, ni).T.astype(np.float64)

    @log_time("")
    def read(self, file_type=".npz"):
        print("reading the nonlinear snapshots tensor ...")
        p = self.param
        inc, jump = p.constProj_frame_increment, p.constProj_train_test_jump
        train_ids = range(0, self.frs * inc, inc)
        test_ids = range(jump, self.frs * inc, inc)
        if file_type == ".npz":
            data = np.load(self.snapshots_file, allow_pickle=True)
            get = lambda i: np.asarray(data[str(i)], dtype=np.float64)  # noqa: E731
        elif file_type == ".bin":
            get = lambda i: self._read_bin_matrix(self.snapshots_file + str(i) + ".bin")  # noqa: E731
        else:
            raise ValueError("unknown snapshots file type: " + str(file_type))
        X = np.stack([get(i) for i in train_ids], axis=0)
        Xt = np.stack([get(j) for j in test_ids], axis=0) if len(test_ids) else np.empty((0,) + X.shape[1:])
        self.num_constained_elements = X.shape[1] // self.constraintsSize
        self._raw = X
        self.test_snapTensor = Xt
        print("loaded snapshots size", X.shape)
        print("No. constrained verts: ", self.num_constained_elements)
        print("loaded test snapshots size", Xt.shape)

    @log_time("")
    def load_factorize_masses(self):
        n = self.num_constained_elements * self.constraintsSize
        if os.path.exists(self.mass_file):
            self.mass = read_mass_bin(self.mass_file, None)[:n].copy()
        else:
            raise FileNotFoundError(
                f"{self.mass_file}: the reference derives element masses from libigl's mass matrix here "
                "(snapbases/nonlinear_snapshots.py:193-241); libigl is a third-party dependency that is not available. "
                "Supply the mass .bin (hostio.write_mass_bin of element_masses_from_vertex_masses(...))")
        self._read_connectivity()
        massL = np.sqrt(self.mass)
        assert np.allclose(massL * massL - self.mass, 0.0)
        invMassL = np.zeros(n)
        nz = massL != 0
        invMassL[nz] = 1.0 / massL[nz]
        assert np.allclose(invMassL * massL, np.ones(n))
        self.massL, self.invMassL = massL, invMassL

    def _read_connectivity(self):
        """verts / tets / tris of the rest mesh when its file exists.  The reference reads them only on its libigl branch
        (nonlinear_snapshots.py:197-241, i.e. when no mass file is given), yet the geometric DEIM needs them (:684-690 of
        constraintsComponents.py); reading them alongside the mass file lets `interpolation_type: geom` run under main.py."""
        from ..synth import read_medit_mesh, read_obj_mesh
        if self.tets is None and self.tet_mesh and os.path.exists(self.tet_mesh):
            self.verts, self.tets, tris = read_medit_mesh(self.tet_mesh)
            if len(tris):
                self.tris = tris
        if self.tris is None and self.tri_mesh and os.path.exists(self.tri_mesh):
            self.verts, self.tris = read_obj_mesh(self.tri_mesh)

    @log_time("")
    def standarize(self):
        """Kept for API parity; the work is part of snapshots_prepare (engine.prepare_snapshots)."""
        return None
