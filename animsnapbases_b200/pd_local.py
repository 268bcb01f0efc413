"""Headless snapshot producer: the projective-dynamics LOCAL step of one constraint group over a batch of
position frames, on the device, decoupled from the reference's polyscope callback loop
(projective_dynamics/Simulators.py:342-364 get_group_ST_p calls `c.get_pi(q_t)` element by element in Python and stores
`p` per frame in `<group>_p.npz`; projective_dynamics/Constraint_projections.py holds the projections).

    group = ConstraintGroup("tets_deformation_gradient", rest_positions, tets)
    X = group.project(frames)             # (F, 3, ld) planar: exactly the stacked `p` of every frame = the snapshot tensor
    group.save_p_npz(path, X)             # the `<group>_p.npz` file nonlinearSnapshots.read consumes (keys '0', '1', ...)

Rest data (inverse rest matrices, rest frames, rest lengths) is an O(e) numpy pass at set-up; the per-frame work is
asb_pd_project (one thread per element and frame).  The global step of the solver is not part of this: frames come from
wherever positions come from (a simulator, an animation file, synth.bar_deformation).
"""
from __future__ import annotations

import numpy as np

from . import runtime
from .ops import pad_ld

KINDS = {"tets_deformation_gradient": (0, 4, 3), "tets_strain": (1, 4, 3), "tris_strain": (2, 3, 2), "edge_spring": (3, 2, 1)}


class ConstraintGroup:
    def __init__(self, name, rest_positions, elems, sigma_min=0.95, sigma_max=1.05):
        if name not in KINDS:
            raise ValueError("unknown constraint group: " + str(name))
        self.name = name
        self.kind, nodes, self.rows = KINDS[name]
        x = np.asarray(rest_positions, dtype=np.float64)
        el = np.asarray(elems, dtype=np.int64)
        assert x.ndim == 2 and x.shape[1] == 3 and el.ndim == 2 and el.shape[1] == nodes
        self.n_verts, self.num_elements = len(x), len(el)
        self.sigma_min, self.sigma_max = float(sigma_min), float(sigma_max)
        if self.kind <= 1:          # Constraint_projections.py:501-504 / :636-639: Dm = [p1-p4, p2-p4, p3-p4], DmInv
            Dm = np.stack([x[el[:, c]] - x[el[:, 3]] for c in range(3)], axis=2)
            rest = np.linalg.inv(Dm).reshape(-1, 9)
        elif self.kind == 2:        # :366-381: orthonormal rest frame P (3 x 2) of the triangle and the inverse 2 x 2 rest matrix
            e1, e2 = x[el[:, 1]] - x[el[:, 0]], x[el[:, 2]] - x[el[:, 0]]
            P0 = e1 / np.linalg.norm(e1, axis=1, keepdims=True)
            P1 = e2 - (e2 * P0).sum(axis=1, keepdims=True) * P0
            P1 /= np.linalg.norm(P1, axis=1, keepdims=True)
            P = np.stack([P0, P1], axis=2)
            Di = np.linalg.inv(np.swapaxes(P, 1, 2) @ np.stack([e1, e2], axis=2))
            rest = np.concatenate([P.reshape(-1, 6), Di.reshape(-1, 4)], axis=1)
        else:                       # :280: rest length
            rest = np.linalg.norm(x[el[:, 0]] - x[el[:, 1]], axis=1)[:, None]
        ops, _ = runtime.get()
        self._elems = ops.to_device(np.ascontiguousarray(el, dtype=np.int32))
        self._rest = ops.to_device(np.ascontiguousarray(rest, dtype=np.float64))

    def project(self, frames):
        """frames: host (F, |V|, 3) positions, or a planar device tensor (F, 3, ldq) -> planar snapshot tensor (F, 3, ld) with
        ld = padded rows * e; row rows * t + j of frame f is row j of get_pi of element t."""
        ops, _ = runtime.get()
        if isinstance(frames, np.ndarray):
            assert frames.ndim == 3 and frames.shape[1:] == (self.n_verts, 3)
            q = ops.ingest(ops.to_device(np.ascontiguousarray(frames, dtype=np.float64)), None, pad_ld(self.n_verts))
        else:
            q = frames
        out = ops.zeros((q.shape[0], 3, pad_ld(self.rows * self.num_elements)))
        return ops.pd_project(self.kind, q, self._elems, self._rest, self.sigma_min, self.sigma_max, out)

    def project_host(self, frames):
        """(F, rows * e, 3) host array: the reference's stacked `p` per frame."""
        ops, _ = runtime.get()
        return ops.export_aos(self.project(frames), None, self.rows * self.num_elements).cpu().numpy()

    def save_p_npz(self, path, X=None, frames=None):
        """`<group>_p.npz` with one (rows * e, 3) array per frame under the keys '0', '1', ... (Simulators.py:357-360)."""
        ops, _ = runtime.get()
        if X is None:
            X = self.project(frames)
        p = ops.export_aos(X, None, self.rows * self.num_elements).cpu().numpy()
        np.savez(path, **{str(i): p[i] for i in range(p.shape[0])})
        return path
