"""posSnapshots -- device-resident mirror of snapbases/posSnapshots.py:26-172 (reference).

Same constructor, attributes and method names; the arithmetic (f32 -> f64, sqrt(m) row scaling, rest shape, population
std over all entries, centring/scaling) runs in libasb.so kernels on the planar device tensor `X`.  Host numpy views
(`snapTensor`, `mean`) are materialised lazily on attribute access so that the reference's figure scripts keep working.
"""
from __future__ import annotations

import os
import sys

import numpy as np

from .. import engine, hostio, runtime
from ..geodesic import HeatGeodesics
from ..hostio import log_time


class posSnapshots:
    def __init__(self, input_train_animation_file, input_test_animation_file, rest_shape, masses_file, tet_mesh_file,
                 standarize=True, massWeight=True):
        self.input_animation_file = input_train_animation_file
        self.input_test_animation_file = input_test_animation_file
        self.rest_shape = rest_shape
        self._stored_verts = None
        self._stored_test = None
        self._verts64 = None
        self._test64 = None
        self.tris = None
        self.test_tris = None
        self.frs = 0
        self.nVerts = 0
        self.pre_scale_factor = 1
        self.massesFile = masses_file
        self.mass = None
        self.massL = None
        self.invMassL = None
        self.compute_geodesic_distance = None
        self.tet_mesh = tet_mesh_file
        self.state = None            # engine.SnapshotState (device tensors of this rank's row shard)
        self.shard = None
        self._geo = None
        self.do_snapshots_precomputations(standarize, massWeight)

    # ---- the reference keeps float64 host copies of both animations (`verts`, `test_verts`, posSnapshots.py:110-118); here
    # they are made on first access only: the build reads the stored (float32) frames straight into HBM
    @staticmethod
    def _as_float(stored):
        if isinstance(stored, hostio.MemoryAnimation):
            stored = stored.complete()
        return np.asarray(stored).astype(float)

    @property
    def verts(self):
        if self._verts64 is None and self._stored_verts is not None:
            self._verts64 = self._as_float(self._stored_verts)
        return self._verts64

    @property
    def test_verts(self):
        if self._test64 is None and self._stored_test is not None:
            self._test64 = self._as_float(self._stored_test)
        return self._test64

    # ---- lazily materialised host views of device tensors (reference attribute names)
    @property
    def snapTensor(self):
        ops, comm = runtime.get()
        return engine.host_rows(ops, comm, self.state.X, self.shard)

    @property
    def mean(self):
        ops, comm = runtime.get()
        return engine.host_rows(ops, comm, self.state.mean[None], self.shard)[0]

    @log_time("")
    def do_snapshots_precomputations(self, standarize, massWeight):
        ops, comm = runtime.get()
        self.read()
        if massWeight:
            self.read_factorize_masses()
            assert self.nVerts == self.massL.shape[0]
        if self.rest_shape not in ("first", "average"):
            print('Error! unknown rest shape: ', self.rest_shape)
            sys.exit(1)
        self.shard = engine.Shard.make(self.nVerts, comm.world, comm.rank)
        rows = slice(self.shard.row0, self.shard.row0 + self.shard.n)
        # geodesics on the un-weighted rest shape (posSnapshots.py:96-99); the mesh operator is global
        stored = self._stored_verts
        if isinstance(stored, hostio.MemoryAnimation):
            assert self.rest_shape == "first", "an animation handed over in memory carries frame 0 as its rest shape"
            rest = stored.rest
            local = stored.rows(self.shard.row0, self.shard.n)
            src = local if hasattr(local, "is_pinned") else np.ascontiguousarray(local)
        else:
            rest = stored[0] if self.rest_shape == "first" else np.mean(self.verts, axis=0)
            src = np.ascontiguousarray(stored[:, rows])
        self._geo = HeatGeodesics.cached(ops, np.asarray(rest, dtype=np.float64), self.tris)
        self.compute_geodesic_distance = self._distance_query
        src = ops.to_device(src)
        self.state = engine.prepare_snapshots(ops, comm, src, self.mass[rows] if massWeight else None, self.rest_shape,
                                              self.shard, standarize=standarize, mass_weight=massWeight)
        self.pre_scale_factor = self.state.scale
        print('Snapshots ready... Volkwein (' + str(massWeight) + '), standarized (' + str(standarize) + ').')

    def _distance_query(self, idx):
        """GeodesicDistanceComputation.__call__ (utils/support.py:173-208): phi (N,) from one source vertex."""
        ops, _ = runtime.get()
        out = ops.zeros((1, self._geo.ldx))
        phi = self._geo.distances(ops.to_device(np.asarray([int(idx)], dtype=np.int64)))
        ops.support_map(phi, self._geo.n, 0.0, 0.0, 2, 0.0, 0, self._geo.n, out)
        return out.cpu().numpy()[0, :self._geo.n]

    @log_time("")
    def read(self):
        verts, self.tris = hostio.read_animation(self.input_animation_file)
        self._stored_verts = verts                       # float32 on disk (utils/process.py:85,:246)
        self.frs, self.nVerts, _ = verts.shape
        print("Vertices: ", self.nVerts)
        print("Faces: ", self.tris.shape[0])
        print("Frames: ", self.frs)
        self._stored_test, self.test_tris = hostio.read_animation(self.input_test_animation_file)

    @log_time("")
    def read_factorize_masses(self, mass_on_tet_mesh=False):
        if not os.path.exists(self.massesFile):
            raise FileNotFoundError(
                f"{self.massesFile}: the reference falls back to libigl's Voronoi mass matrix here "
                "(snapbases/posSnapshots.py:131-139); libigl is a third-party dependency that is not available, supply the "
                "mass .bin (animsnapbases_b200.hostio.write_mass_bin)")
        self.mass = hostio.read_mass_bin(self.massesFile, self.nVerts)
        # cholesky / inverse of a DIAGONAL matrix (posSnapshots.py:156-160) == sqrt / reciprocal, bit-identical
        self.massL = np.sqrt(self.mass)
        self.invMassL = 1.0 / self.massL

    @log_time("")
    def standarize(self):
        """Kept for API parity; the work is part of do_snapshots_precomputations (engine.prepare_snapshots)."""
        return None
