"""Consumer side of the emitted constraint-projection bases: the reduced operator precompute of
projective_dynamics/Simulators.py:157-218 (`prepare_reduced_group`) and the per-step reduced term of :366-399
(`get_group_reduced_term`), for one constraint group, on the device.

Only these two functions of the reference's solver are mirrored -- the step right after the basis build, which proves the
emitted `components / Pt / interpol_alphas / interpol_alpha_ranges` `.npz` is consumable.  The projective-dynamics solver
around them (global step, constraint classes, polyscope loop) is out of scope.

    grp = prepare_reduced_group(True, True, "tets_deformation_gradient", num_components, row_dim, assembly_ST, dir, file,
                                reduction_type="deim_pod_vectorized")
    interpolation_alpha, Pt, projecting_mat, solver_list = grp          # the reference's 4-tuple (host numpy, lazily)
    rhs = grp.reduced_term(p_rows)                                      # (|V|, 3): S^T V (A^T A + la I)^-1 A^T p

Device work goes through animsnapbases_b200.engine (reduced_group_operators / reduced_group_term) -> libasb.so; there is no
CPU path.
"""
from __future__ import annotations

import os

import numpy as np

from . import engine, runtime

ROW_SELECTING_TYPES = {"deim_pod", "deim_pod_vectorized"}                       # Simulators.py:173, :187, :382
KNOWN_TYPES = ROW_SELECTING_TYPES | {"deim_pca_blocks", "geom_pca_blocks_withSt"}   # Simulators.py:283


class ReducedGroup:
    """Device-resident operators of one reduced constraint group.  Iterating yields the reference's return tuple
    `(interpolation_alpha, Pt, projecting_mat, solver_list)` (Simulators.py:218)."""

    def __init__(self, ops, comm, interpolation_alpha, Pt, state):
        self._ops, self._comm = ops, comm
        self.interpolation_alpha = interpolation_alpha
        self.Pt = Pt
        self._state = state
        self._projecting_mat = None

    @property
    def projecting_mat(self):
        """S^T V as the reference holds it: host (|V|, mp, 3)."""
        if self._projecting_mat is None:
            s = self._state
            aos = self._ops.export_aos(s["StV"], None, s["nv"])                # (mp, |V|, 3)
            self._projecting_mat = np.ascontiguousarray(np.swapaxes(aos.cpu().numpy(), 0, 1))
        return self._projecting_mat

    @property
    def solver_list(self):
        """[[(lu, piv), A_d^T] for d in x, y, z] usable with scipy.linalg.lu_solve (Simulators.py:211-214)."""
        s = self._state
        lu, piv, A = s["LU"].cpu().numpy(), s["piv"].cpu().numpy(), s["A"].cpu().numpy()
        return [[(lu[d], piv[d].astype(np.int32) - 1), np.ascontiguousarray(A[d].T)] for d in range(3)]

    def __iter__(self):
        return iter((self.interpolation_alpha, self.Pt, self.projecting_mat, self.solver_list))

    def reduced_term_device(self, p_rows):
        """(B, 3, L) planar on the device for a batch of projections (B, |Pt|, 3) at the interpolation rows."""
        return engine.reduced_group_term(self._ops, self._state, p_rows)

    def reduced_term(self, p_rows):
        """Simulators.py:394-399 for one set of projections p_rows (|Pt|, 3) -> host (|V|, 3)."""
        out = self.reduced_term_device(np.asarray(p_rows, dtype=np.float64)[None])
        return np.ascontiguousarray(out[0, :, :self._state["nv"]].T.cpu().numpy())


def prepare_reduced_group(has_group_constraints, reduced_group, group_name, num_components, row_dim, assembly_ST, dir, file,
                          reduction_type="deim_pod_vectorized"):
    """Argument meaning as projective_dynamics/Simulators.py:157 (`reduction_type` stands for
    `self.constraint_projection_reduction_type`).  Reads `<dir>/<group_name>/<file>` (the `.npz` written by
    constraintsComponents.store_components_n_interpol_points), returns a ReducedGroup, or `(None, [], None, [])` when the
    group has no reduced constraints (:219-220).  Unknown reduction types raise ValueError (:287)."""
    if reduction_type not in KNOWN_TYPES:
        raise ValueError("Unknown reduction type for constraint projections")
    if not (has_group_constraints and reduced_group):
        return None, [], None, []
    if reduction_type in ROW_SELECTING_TYPES:
        row_dim = 1
    ops, comm = runtime.get()
    local_data = np.load(os.path.join(dir, group_name, file))
    components = local_data["components"]                                        # (K, ep, 3)
    mp = num_components * row_dim
    assert components.ndim == 3 and components.shape[2] == 3 and mp <= components.shape[0]
    alpha_range = int(local_data["interpol_alpha_ranges"][num_components - 1])
    interpolation_alpha = local_data["interpol_alphas"][:alpha_range]
    if reduction_type in ROW_SELECTING_TYPES:
        Pt = local_data["Pt"][:alpha_range]
    else:
        Pt = [int(alpha) * row_dim + l for alpha in interpolation_alpha for l in range(row_dim)]
    ep = components.shape[1]
    assert assembly_ST.shape[1] == ep, "assembly_ST must be |V| x ep"
    shard = engine.Shard.make(ep, comm.world, comm.rank)
    rows = slice(shard.row0, shard.row0 + shard.n)
    C = ops.ingest(ops.to_device(np.ascontiguousarray(components[:mp, rows], dtype=np.float64)), None, shard.ld)
    state = engine.reduced_group_operators(ops, comm, C, shard, Pt, assembly_ST)
    print(group_name + f" basis file loaded with: \n Basis shape {(ep, mp, 3)} and {interpolation_alpha.shape} interpolation points.")
    return ReducedGroup(ops, comm, interpolation_alpha, Pt, state)
