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


def bending_stars(rest_positions, faces, flat_bending=False):
    """Rest data of the verts_bending group (Constraint_projections.py:132-187, :1086-1104, :1129-1161, :1195-1223), vectorised
    over the half-edges instead of the reference's per-triangle Python loops: the constrained vertices (every star edge has two
    triangles), their stars in CSR form with cotangent weights / Voronoi area, and per vertex [rest mean curvature, rest
    normal, rest normal . rest mean-curvature vector].  Returns (verts, ptr, nbr, w, rest, voronoi_area)."""
    x, f = np.asarray(rest_positions, dtype=np.float64), np.asarray(faces, dtype=np.int64)
    n, T = len(x), len(f)
    nrm = np.cross(x[f[:, 1]] - x[f[:, 0]], x[f[:, 2]] - x[f[:, 0]])
    nlen = np.linalg.norm(nrm, axis=1)
    area = np.bincount(f.ravel(), np.repeat(0.5 * nlen / 3.0, 3), minlength=n)
    area[area < 1e-7] = 1e-7
    v, nb, o = f[:, [0, 0, 1, 1, 2, 2]].ravel(), f[:, [1, 2, 0, 2, 0, 1]].ravel(), f[:, [2, 1, 2, 0, 1, 0]].ravel()
    u, w = x[v] - x[o], x[nb] - x[o]
    cosang = np.clip((u * w).sum(axis=1) / (np.linalg.norm(u, axis=1) * np.linalg.norm(w, axis=1)), -1, 1)
    half_cot = 0.5 / np.tan(np.arccos(cosang))
    key, inv, cnt = np.unique(v * n + nb, return_inverse=True, return_counts=True)      # one entry per directed star edge
    ev, enb = key // n, key % n
    ecot = np.bincount(inv, half_cot, minlength=len(key))
    open_edge = np.bincount(ev, cnt != 2, minlength=n) > 0
    has_star = np.bincount(ev, minlength=n) > 0
    verts = np.nonzero(has_star & ~open_edge)[0]
    keep = np.isin(ev, verts)
    ev, enb, ew = ev[keep], enb[keep], ecot[keep] / area[ev[keep]]
    ptr = np.concatenate([[0], np.cumsum(np.bincount(ev, minlength=n)[verts])])
    mc = np.zeros((n, 3))
    np.add.at(mc, ev, (x[ev] - x[enb]) * ew[:, None])
    valid = nlen > 1e-10
    unit = np.where(valid[:, None], nrm / np.where(valid, nlen, 1.0)[:, None], 0.0)
    nsum, ncnt = np.zeros((n, 3)), np.zeros(n)
    for c in range(3):
        np.add.at(nsum, f[:, c], unit)
        np.add.at(ncnt, f[:, c], valid.astype(np.float64))
    normal = np.where(ncnt[:, None] > 0, nsum / np.where(ncnt > 0, ncnt, 1.0)[:, None], np.array([0.0, 0.0, 1.0]))
    rest = np.empty((len(verts), 5))
    rest[:, 0] = 0.0 if flat_bending else np.linalg.norm(mc[verts], axis=1)
    rest[:, 1:4] = normal[verts]
    rest[:, 4] = (normal[verts] * mc[verts]).sum(axis=1)
    return verts, ptr, enb, ew, rest, area


class VertexBendingGroup:
    """verts_bending: one row per constrained vertex (`verts`, the reference's verts_bending_indicies)."""
    name, rows = "verts_bending", 1

    def __init__(self, rest_positions, faces, prevent_bending_flips=True, flat_bending=False):
        self.verts, ptr, nbr, w, rest, self.voronoi_area = bending_stars(rest_positions, faces, flat_bending)
        self.n_verts, self.num_elements, self.prevent = len(rest_positions), len(self.verts), bool(prevent_bending_flips)
        ops, _ = runtime.get()
        self._dev = [ops.to_device(np.ascontiguousarray(a, dtype=t)) for a, t in
                     ((self.verts, np.int32), (ptr, np.int32), (nbr, np.int32), (w, np.float64), (rest, np.float64))]

    def project(self, frames):
        ops, _ = runtime.get()
        if isinstance(frames, np.ndarray):
            assert frames.ndim == 3 and frames.shape[1:] == (self.n_verts, 3)
            q = ops.ingest(ops.to_device(np.ascontiguousarray(frames, dtype=np.float64)), None, pad_ld(self.n_verts))
        else:
            q = frames
        out = ops.zeros((q.shape[0], 3, pad_ld(self.num_elements)))
        return ops.pd_vertex_bending(q, *self._dev, self.prevent, out)

    def project_host(self, frames):
        ops, _ = runtime.get()
        return ops.export_aos(self.project(frames), None, self.num_elements).cpu().numpy()


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
