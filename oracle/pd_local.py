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


# ------------------------------------------------------------------------------------------ vertices: bending
class VertBending:
    """VertBendingConstraint (:116-216) for every vertex whose star edges all have two triangles (add_vertex_bending_constraint,
    :1195-1223): one row per constrained vertex, the cotangent-weighted star sum rescaled to the rest mean-curvature length
    (optionally flipped back to the rest side of the surface)."""
    rows = 1

    def __init__(self, positions, faces, wi=1.0, prevent_bending_flips=True, flat_bending=False):
        x, faces = np.asarray(positions, dtype=np.float64), np.asarray(faces)
        n = len(x)
        area = np.zeros(n)                                                            # vertex_masses :1086-1104
        for tri in faces:
            a = 0.5 * np.linalg.norm(np.cross(x[tri[1]] - x[tri[0]], x[tri[2]] - x[tri[0]])) / 3.0
            for vi in tri:
                area[vi] += a
        area[area < 1e-7] = 1e-7
        stars = [[] for _ in range(n)]                                                # vertex_star :1129-1161: [v2, vOtherT1, t1, vOtherT2, t2]
        for t, tri in enumerate(faces):
            for v in range(3):
                for ov in range(3):
                    if v == ov:
                        continue
                    nb, third = tri[ov], tri[3 - (v + ov)]
                    for edge in stars[tri[v]]:
                        if edge[0] == nb:
                            edge[4], edge[3] = t, third
                            break
                    else:
                        stars[tri[v]].append([nb, third, t, -1, -1])

        def angle(a, b, c):                                                           # :138-143, angle at b
            u, w = a - b, c - b
            return np.arccos(np.clip(np.dot(u, w) / (np.linalg.norm(u) * np.linalg.norm(w)), -1, 1))

        self.verts, self.nbrs, self.weights, self.rest_mc, self.normal, self.dot_n, self.scale = [], [], [], [], [], [], []
        for v in range(n):
            star = stars[v]
            if not star or any(e[4] < 0 for e in star):                               # :1207-1212
                continue
            w, tris_seen, tris = [], set(), []
            for v2, o1, t1, o2, t2 in star:                                           # :156-175
                cot = 0.5 / np.tan(angle(x[v], x[o1], x[v2]))
                if t2 >= 0:
                    cot += 0.5 / np.tan(angle(x[v], x[o2], x[v2]))
                w.append(cot / area[v])
                for t in (t1, t2):
                    if t >= 0 and t not in tris_seen:
                        tris.append(faces[t])
                        tris_seen.add(t)
            w = np.array(w)
            mc = np.zeros(3)
            for e, wi_e in zip(star, w):                                              # :180-182
                mc += (x[v] - x[e[0]]) * wi_e
            normals = []
            for tri in tris:                                                          # :145-154
                nrm = np.cross(x[tri[1]] - x[tri[0]], x[tri[2]] - x[tri[0]])
                ln = np.linalg.norm(nrm)
                if ln > 1e-10:
                    normals.append(nrm / ln)
            nm = np.mean(normals, axis=0) if normals else np.array([0.0, 0.0, 1.0])
            self.verts.append(v)
            self.nbrs.append(np.array([e[0] for e in star]))
            self.weights.append(w)
            self.rest_mc.append(0.0 if flat_bending else np.linalg.norm(mc))          # :184
            self.normal.append(nm)
            self.dot_n.append(nm @ mc)                                                # :187
            self.scale.append(wi * area[v])                                           # super().__init__([v], wi * voronoi_area) :119
        self.verts = np.array(self.verts, dtype=np.int64)
        self.prevent, self.n_verts = prevent_bending_flips, n

    def assembly(self):
        """|V| x (constrained vertices): column i = wi_i (sum_j w_ij at v_i, -w_ij at its neighbours) (:190-197, :1220-1222)."""
        r, c, val = [], [], []
        for i, v in enumerate(self.verts):
            r += [v] + list(self.nbrs[i])
            c += [i] * (1 + len(self.nbrs[i]))
            val += [self.scale[i] * self.weights[i].sum()] + list(-self.scale[i] * self.weights[i])
        return sparse.csr_matrix((val, (r, c)), shape=(self.n_verts, len(self.verts)))

    def project(self, q):
        out = np.zeros((len(self.verts), 3))
        for i, v in enumerate(self.verts):                                            # :197-216
            star_sum = np.zeros(3)
            for nb, w in zip(self.nbrs[i], self.weights[i]):
                star_sum += (q[v] - q[nb]) * w
            norm = np.linalg.norm(star_sum)
            if norm < 1e-10:
                corr = self.normal[i] * self.rest_mc[i]
            else:
                corr = star_sum * (self.rest_mc[i] / norm)
            if self.prevent:
                if norm > 1e-5 and (self.normal[i] @ corr) * self.dot_n[i] < 0:
                    corr = -corr
            out[i] = corr
        return out
