"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the projective-dynamics LOCAL STEP that produces the constraint-projection
snapshots: per element, the projection `get_pi` onto the constraint manifold and the weighted selection
operator `S_i^T`, for the element types whose snapshots the basis path consumes.  Vectorised numpy over all elements; each
function cites the reference lines it follows (projective_dynamics/Constraint_projections.py).

Pinned against the reference's own constraint classes (loaded unmodified through oracle/ref_shim.load_pd_module) by
oracle/make_golden.py --pd-only -> tests/golden/pd_local.npz and by tests/test_oracle_golden.py.  This is the oracle for
the snapshot-producer kernels of a later round; nothing in animsnapbases_b200/ imports it.

Layout of the results: `project` returns the stacked rows `p (e * rows, 3)` exactly as Simulators.get_group_ST_p stacks the
per-element `get_pi` blocks (projective_dynamics/Simulators.py:354-356); `assembly` returns S^T (|V| x e * rows, CSR).
"""
from __future__ import annotations

import numpy as np
from scipy import sparse


def _edges3(x, elems):
    """Columns [x1 - x4, x2 - x4, x3 - x4] per tet -> (e, 3, 3)."""
    return np.stack([x[elems[:, 0]] - x[elems[:, 3]], x[elems[:, 1]] - x[elems[:, 3]], x[elems[:, 2]] - x[elems[:, 3]]], axis=2)


def _assemble(n_verts, elems, G, weight):
    """S^T from per-element gradient blocks G (e, rows, nodes): entry (elems[t, c], rows * t + j) = weight[t] * G[t, j, c]
    (the build_SiT methods: :388-405 triangles, :510-532 / :645-667 tets, :285-289 edges)."""
    e, rows, nodes = G.shape
    r = np.repeat(elems[:, None, :], rows, axis=1)                                   # vertex id of (t, j, c)
    c = np.broadcast_to((rows * np.arange(e))[:, None, None] + np.arange(rows)[None, :, None], G.shape)
    v = G * weight[:, None, None]
    return sparse.csr_matrix((v.ravel(), (r.ravel(), c.ravel())), shape=(n_verts, rows * e))


# ------------------------------------------------------------------------------------------ tets: deformation gradient
class TetDeformationGradient:
    """TetDeformationGradientConstraint (:627-687): rows = R^T of the polar rotation of F = Ds Dm^-1 (3 rows per tet)."""
    rows = 3

    def __init__(self, positions, tets, wi=1.0):
        self.tets = np.asarray(tets)
        Dm = _edges3(positions, self.tets)                                           # :638
        self.DmInv = np.linalg.inv(Dm)                                               # :639
        self.V0 = np.linalg.det(Dm) / 6.0                                            # :640
        self.wi, self.n_verts = wi, len(positions)

    def assembly(self):
        G = np.concatenate([np.swapaxes(self.DmInv, 1, 2), -self.DmInv.sum(axis=1)[:, :, None]], axis=2)   # :651-654 (3 x 4)
        return _assemble(self.n_verts, self.tets, G, self.wi * np.abs(self.V0))      # :667

    def project(self, q):
        F = _edges3(q, self.tets) @ self.DmInv                                       # :679-680
        U, _, Vt = np.linalg.svd(F)
        R = U @ Vt                                                                   # :682-683
        R[np.linalg.det(R) < 0, :, 2] *= -1                                          # :684-685
        return np.swapaxes(R, 1, 2).reshape(-1, 3)                                   # :687 (R^T)


# ------------------------------------------------------------------------------------------ tets: strain limiting
class TetStrain(TetDeformationGradient):
    """TetStrainConstraint (:483-554): rows = Fhat = U clip(s) V^T with the last singular value negated for inverted tets."""

    def __init__(self, positions, tets, wi=1.0, sigma_min=0.95, sigma_max=1.05):
        super().__init__(positions, tets, wi)
        self.sigma_min, self.sigma_max = sigma_min, sigma_max

    def project(self, q):
        F = _edges3(q, self.tets) @ self.DmInv                                       # :544-545
        U, s, Vt = np.linalg.svd(F)
        s = np.clip(s, self.sigma_min, self.sigma_max)                               # :548
        s[np.linalg.det(F) < 0.0, 2] *= -1                                           # :550-551
        return ((U * s[:, None, :]) @ Vt).reshape(-1, 3)                             # :553-554 (Fhat, not transposed)


# ------------------------------------------------------------------------------------------ triangles: strain limiting
class TriStrain:
    """TriStrainConstraint (:353-426): 2 rows per triangle, (P Fhat)^T with Fhat the clipped 2 x 2 strain in the rest frame."""
    rows = 2

    def __init__(self, positions, tris, wi=1.0, sigma_min=0.95, sigma_max=1.05):
        self.tris = np.asarray(tris)
        p1, p2, p3 = (positions[self.tris[:, i]] for i in range(3))
        e1, e2 = p2 - p1, p3 - p1
        P0 = e1 / np.linalg.norm(e1, axis=1, keepdims=True)                          # :372
        P1 = e2 - (e2 * P0).sum(axis=1, keepdims=True) * P0                          # :373
        P1 /= np.linalg.norm(P1, axis=1, keepdims=True)                              # :374
        self.P = np.stack([P0, P1], axis=2)                                          # (e, 3, 2)
        rest2d = np.swapaxes(self.P, 1, 2) @ np.stack([e1, e2], axis=2)              # :379-380
        self.DmInv = np.linalg.inv(rest2d)                                           # :381
        self.A0 = 0.5 * np.linalg.det(rest2d)                                        # :383
        self.sigma_min, self.sigma_max, self.wi, self.n_verts = sigma_min, sigma_max, wi, len(positions)

    def assembly(self):
        DT = np.swapaxes(self.DmInv, 1, 2)
        G = np.concatenate([DT, -DT.sum(axis=2)[:, :, None]], axis=2)                # :395 (2 x 3)
        return _assemble(self.n_verts, self.tris, G, self.wi * np.abs(self.A0))      # :405

    def project(self, q):
        q1, q2, q3 = (q[self.tris[:, i]] for i in range(3))
        Ds2d = np.swapaxes(self.P, 1, 2) @ np.stack([q2 - q1, q3 - q1], axis=2)      # :416-417
        U, s, Vt = np.linalg.svd(Ds2d @ self.DmInv)                                  # :419
        s = np.clip(s, self.sigma_min, self.sigma_max)                               # :420
        Fhat = (U * s[:, None, :]) @ Vt                                              # :421
        return np.swapaxes(self.P @ Fhat, 1, 2).reshape(-1, 3)                       # :425


# ------------------------------------------------------------------------------------------ edges: springs
class EdgeSpring:
    """EdgeSpringConstraint (:274-312): one row per edge, half the edge vector rescaled to the mean of its current and rest
    length (zero-length edges: the reference returns None, :303-304; here a zero row)."""
    rows = 1

    def __init__(self, positions, edges, wi=1.0):
        self.edges = np.asarray(edges)
        self.d = np.linalg.norm(positions[self.edges[:, 0]] - positions[self.edges[:, 1]], axis=1)   # :280
        self.wi, self.n_verts = wi, len(positions)

    def assembly(self):
        G = np.broadcast_to(np.array([-1.0, 1.0])[None, None, :], (len(self.edges), 1, 2))           # :288-289
        return _assemble(self.n_verts, self.edges, np.array(G), np.full(len(self.edges), self.wi))

    def project(self, q):
        p1, p2 = q[self.edges[:, 0]], q[self.edges[:, 1]]
        spring = p2 - p1
        length = np.linalg.norm(spring, axis=1)
        safe = np.where(length == 0, 1.0, length)
        n = spring / safe[:, None]
        delta = 0.5 * (length - self.d)                                              # :307
        pi1 = p1 + delta[:, None] * n
        pi2 = p2 - delta[:, None] * n
        out = 0.5 * (pi2 - pi1)                                                      # :311
        out[length == 0] = 0.0
        return out
