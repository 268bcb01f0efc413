"""TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's basis-computation hot path.

Plain numpy/scipy, one function per reference function, each citing the reference file:line it
follows (paths relative to the reference root, ShMonem/animSnapBases).  It is the *checker* for the
CUDA path: only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it.  `animsnapbases_b200/` never does; the product fails loudly without its CUDA library.

Pinning: the reference holds no golden vectors or known-answer tests for this path (SURVEY.md 0.9,
section 4), so the port is pinned against *outputs of the reference itself*: `oracle/make_golden.py`
imports the unmodified reference through `oracle/ref_shim.py` in the build container, runs both on the
same seeded inputs, asserts agreement and commits the reference's outputs under `tests/golden/`.
`tests/test_oracle_golden.py` re-checks the port against those fixtures everywhere (and against the
live reference when /root/reference is mounted).

The port keeps the reference's arithmetic *and* its cost structure (S-sized temporaries, three passes
per Gauss-Seidel step, ...) so that timing it is a fair CPU baseline ("kind": "port").
"""
from __future__ import annotations

import struct

import numpy as np
import scipy.linalg as sla
from scipy import sparse
from scipy.sparse.linalg import splu


# ======================================================================================= helpers
def veclen(v):
    """utils/process.py:150-152"""
    return np.sqrt(np.sum(v ** 2, axis=-1))


def normalized(v):
    """utils/process.py:154-156"""
    return v / veclen(v)[..., np.newaxis]


def project_weight(x):
    """snapbases/posComponents.py:52-58 (same in constraintsComponents.py:76-83):
    clamp to >= 0 and divide by the maximum (unless that is 0)."""
    x = np.maximum(0.0, x)
    m = x.max()
    return x if m == 0 else x / m


def support_map(phi, min_dist, max_dist):
    """snapbases/posComponents.py:61-64, eq. (6) before lambda."""
    return (np.clip(phi, min_dist, max_dist) - min_dist) / (max_dist - min_dist)


def prox_l1l2(Lambda, x, beta):
    """snapbases/posComponents.py:252-256: group soft-threshold of the 3-vectors."""
    xlen = np.sqrt((x ** 2).sum(axis=-1))
    with np.errstate(divide="ignore"):
        shrink = np.maximum(0.0, 1 - beta * Lambda / xlen)
    return x * shrink[..., np.newaxis]


# ======================================================================================= alignment (a1, a2)
def preprocess_unit_cube(verts):
    """utils/process.py:140-144 with the numpy-2-broken `ndarray.ptp` line restated as max-min.
    verts: (F, N, 3).  Returns (verts', mean(3), scale)."""
    verts = np.array(verts, copy=True)
    mean = verts.mean(axis=0).mean(axis=0)
    verts -= mean
    scale = np.abs(verts.max(axis=1) - verts.min(axis=1)).max()
    verts /= scale
    return verts, mean, scale


def find_rbm_procrustes(frompts, topts, rigid):
    """utils/process.py:210-234.  NB the translation column uses R even when rigid is False (H6)."""
    t0 = frompts.mean(0)
    t1 = topts.mean(0)
    M = np.dot((topts - t1).T, frompts - t0)
    U, _, Vt = np.linalg.svd(M)
    R = np.dot(U, Vt)
    if np.linalg.det(R) < 0:
        R *= -1
    T = np.eye(4)
    if rigid:
        T[:3, :3] = R
    T[:3, 3] = t1 - np.dot(R, t0)
    return T


def transform_points(v, T):
    """utils/process.py:198-208 for a 4x4 homogeneous matrix (w = 1)."""
    h = np.concatenate([v, np.ones((len(v), 1))], axis=1) @ T.T
    return h[:, :3] / h[:, 3:4]


def align_frames(verts, rigid):
    """utils/process.py:235-250 without the h5 wrapper: every frame is fitted to frame 0, result is
    stored as float32."""
    v0 = verts[0]
    out = [transform_points(v, find_rbm_procrustes(v, v0, rigid)) for v in verts]
    return np.array(out, np.float32)


# ======================================================================================= snapshots (a3, a8)
def factorize_diag_masses(mass):
    """snapbases/posSnapshots.py:155-160: cholesky/inv of diag(m) -- for a diagonal matrix that is
    sqrt and reciprocal (bit-identical; the dense N x N route is O(N^3)).
    snapbases/nonlinear_snapshots.py:246-262 does the same with an explicit zero guard."""
    massL = np.sqrt(mass)
    inv = np.zeros_like(massL)
    nz = massL != 0
    inv[nz] = 1.0 / massL[nz]
    return massL, inv


def read_mass_bin(path):
    """snapbases/posSnapshots.py:141-149: `<i N><i 1>` then N little-endian doubles."""
    with open(path, "rb") as fh:
        n = struct.unpack("<i", fh.read(4))[0]
        struct.unpack("<i", fh.read(4))
        return np.frombuffer(fh.read(8 * n), dtype="<f8").copy()


def prepare_snapshots(verts_f32, mass, rest_shape, standarize=True, mass_weight=True):
    """posSnapshots.do_snapshots_precomputations + standarize (snapbases/posSnapshots.py:64-105,
    :163-172); identical arithmetic in nonlinear_snapshots.py:74-96, :268-288.
    Returns dict(snapTensor, mean, pre_scale_factor, massL, invMassL)."""
    X = np.asarray(verts_f32).astype(float).copy()       # posSnapshots.py:110
    massL = invMassL = None
    if mass_weight:
        massL, invMassL = factorize_diag_masses(mass)
        X *= massL[:, None]                               # :82
    if rest_shape == "first":
        mean = X[0].copy()                                # :86
    elif rest_shape == "average":
        mean = np.mean(X, axis=0)                         # :89
    else:
        raise SystemExit(1)                               # :91-93
    scale = 1
    if standarize:
        X -= mean[np.newaxis]                             # :168
        scale = 1 / np.std(X)                             # :171 population std over all entries
        X *= scale                                        # :172
    return dict(snapTensor=X, mean=mean, pre_scale_factor=scale, massL=massL, invMassL=invMassL)


# ======================================================================================= geodesics (a4)
def mesh_laplacian(verts, tris):
    """utils/support.py:81-136: cotangent Laplacian Lc (row sums 0) and lumped vertex areas."""
    n = len(verts)
    W, I, J = [], [], []
    for i1, i2, i3 in [(0, 1, 2), (1, 2, 0), (2, 0, 1)]:
        vi1, vi2, vi3 = tris[:, i1], tris[:, i2], tris[:, i3]
        u = verts[vi2] - verts[vi1]
        v = verts[vi3] - verts[vi1]
        cotan = (u * v).sum(axis=1) / veclen(np.cross(u, v))
        W += [0.5 * cotan, 0.5 * cotan]
        I += [vi2, vi3]
        J += [vi3, vi2]
    L = sparse.csr_matrix((np.concatenate(W), (np.concatenate(I), np.concatenate(J))), shape=(n, n))
    L = (L - sparse.spdiags(L * np.ones(n), 0, n, n)).tocsr()
    e1 = verts[tris[:, 1]] - verts[tris[:, 0]]
    e2 = verts[tris[:, 2]] - verts[tris[:, 0]]
    ta3 = 0.5 * veclen(np.cross(e1, e2)) / 3
    area = np.zeros(n)
    for i in range(3):
        bc = np.bincount(tris[:, i].astype(int), ta3)
        area[:len(bc)] += bc
    return L, sparse.spdiags(area, 0, n, n)


class GeodesicHeat:
    """utils/support.py:139-208 (heat method, m = 10): two SuperLU factorisations at construction,
    per query: heat solve -> normalised negative gradient per triangle -> cotan divergence -> Poisson
    solve -> shift so that min = 0."""

    def __init__(self, verts, tris, m=10.0):
        self.verts, self.tris = verts, tris
        e01 = verts[tris[:, 1]] - verts[tris[:, 0]]
        e12 = verts[tris[:, 2]] - verts[tris[:, 1]]
        e20 = verts[tris[:, 0]] - verts[tris[:, 2]]
        self.area = 0.5 * veclen(np.cross(e01, e12))
        un = normalized(np.cross(normalized(e01), normalized(e12)))
        self.n_e01, self.n_e12, self.n_e20 = np.cross(un, e01), np.cross(un, e12), np.cross(un, e20)
        h = np.mean(list(map(veclen, [e01, e12, e20])))
        self.t = m * h ** 2
        Lc, A = mesh_laplacian(verts, tris)
        self.Lc, self.A = Lc, A
        self._heat = splu((A - self.t * Lc).tocsc()).solve
        self._poisson = splu(Lc.tocsc()).solve
        self.calls = 0

    def divergence_of_unit_gradient(self, u):
        """utils/support.py:183-204."""
        tris, verts = self.tris, self.verts
        g = 1 / (2 * self.area)[:, None] * (self.n_e01 * u[tris[:, 2]][:, None]
                                            + self.n_e12 * u[tris[:, 0]][:, None]
                                            + self.n_e20 * u[tris[:, 1]][:, None])
        X = -g / veclen(g)[:, None]
        div = np.zeros(len(verts))
        for i1, i2, i3 in [(0, 1, 2), (1, 2, 0), (2, 0, 1)]:
            vi1, vi2, vi3 = tris[:, i1], tris[:, i2], tris[:, i3]
            e1 = verts[vi2] - verts[vi1]
            e2 = verts[vi3] - verts[vi1]
            eo = verts[vi3] - verts[vi2]
            cot1 = 1 / np.tan(np.arccos((normalized(-e2) * normalized(-eo)).sum(axis=1)))
            cot2 = 1 / np.tan(np.arccos((normalized(-e1) * normalized(eo)).sum(axis=1)))
            div += np.bincount(vi1.astype(int), 0.5 * (cot1 * (e1 * X).sum(axis=1) + cot2 * (e2 * X).sum(axis=1)),
                               minlength=len(verts))
        return div

    def __call__(self, idx):
        self.calls += 1
        u0 = np.zeros(len(self.verts))
        u0[idx] = 1.0
        u = self._heat(u0).ravel()
        phi = self._poisson(self.divergence_of_unit_gradient(u)).ravel()
        phi -= phi.min()
        return phi


# ======================================================================================= deflation "PCA" (a5)
def extract_k_components(X, K, support, geodesic=None, min_dist=0.1, max_dist=0.25, on_row=None):
    """snapbases/posComponents.py:67-122.  X: standardised snapTensor (F, N, 3).
    Returns comps (K,N,3), weigs (F,K), measures (K,3) = [k, sigma0, ||R||_F], picked vertex ids (K,), R."""
    R = X.copy()
    C, W, meas, picked = [], [], [], []
    s = None
    for k in range(K):
        magnitude = (R ** 2).sum(axis=2)                                   # :78
        idx = int(np.argmax(magnitude.sum(axis=0)))                        # :79
        _, sing, Vt = sla.svd(R[:, idx, :].reshape(R.shape[0], -1).T, full_matrices=False)  # :83
        wk = sing[0] * Vt[0, :]                                            # :85
        if support == "local":
            a, b = project_weight(wk), project_weight(-wk)                 # :90-94
            wk = a if sla.norm(a) > sla.norm(b) else b
            s = 1 - support_map(geodesic(idx), min_dist, max_dist)         # :95-96
            ck = (np.tensordot(wk, R, (0, 0)) * s[:, None]) / np.inner(wk, wk)   # :102-103
        else:
            ck = np.tensordot(wk, R, (0, 0)) / np.inner(wk, wk)            # :105
        C.append(ck)
        W.append(wk)
        R -= np.outer(wk, ck).reshape(R.shape)                             # :111
        row = [k, sing[0], sla.norm(R)]                                    # :113
        meas.append(row)
        picked.append(idx)
        if on_row is not None:
            on_row(row)
    return np.array(C), np.array(W).T, np.array(meas), np.array(picked), R


# ======================================================================================= SPLOCS (a6)
def splocs_glob_optimization(X, comps, weigs, R, geodesic, num_iters_max, num_admm_iterations, lam, rho,
                             min_dist=0.1, max_dist=0.25, log=None):
    """snapbases/posComponents.py:132-189.  Returns the loop's internal final (W, C) -- which the
    reference discards -- plus per-iteration (energy, E_rms) and the centre ids."""
    K, N = comps.shape[0], comps.shape[1]
    F = X.shape[0]
    Lambda = np.empty((K, N))
    U = np.zeros((K, N, 3))
    C = comps.copy()
    W = weigs.copy()
    X = X.copy()
    R = R.copy()
    hist, centres = [], []
    for it in range(num_iters_max):
        Rflat = R.reshape(F, N * 3)
        for k in range(K):                                                 # :144-156
            Ck = C[k].ravel()
            Ck_norm = np.inner(Ck, Ck)
            if Ck_norm <= 1.e-8:
                W[:, k] = 0
                continue
            Rflat += np.outer(W[:, k], Ck)
            opt = np.dot(Rflat, Ck) / Ck_norm
            W[:, k] = project_weight(opt)
            Rflat -= np.outer(W[:, k], Ck)
        ids = []
        for k in range(K):                                                 # :158-165
            idx = int((C[k] ** 2).sum(axis=1).argmax())
            ids.append(idx)
            Lambda[k] = lam * support_map(geodesic(idx), min_dist, max_dist)
        centres.append(ids)
        Z = C.copy()                                                       # :168
        G = np.dot(W.T, W)
        c = np.dot(W.T, X.reshape(F, -1))
        fac = sla.cho_factor(G + rho * np.eye(K))                          # :172
        for _ in range(num_admm_iterations):                               # :175-178
            C = sla.cho_solve(fac, c + rho * (Z - U).reshape(c.shape)).reshape(C.shape)
            Z = prox_l1l2(Lambda, C + U, 1. / rho)
            U = U + C - Z
        C = Z                                                              # :181
        R = X - np.tensordot(W, C, (1, 0))                                 # :183
        sparsity = np.sum(Lambda * np.sqrt((C ** 2).sum(axis=2)))
        e_rms = sla.norm(R) / np.sqrt(3 * N * F)
        energy = (R ** 2).sum() + sparsity
        hist.append((energy, e_rms))
        if log is not None:
            log("itr %03d, Energy =%f, Error =%f" % (it, energy, e_rms))   # :189
    return W, C, np.array(hist), np.array(centres)


def post_process_pos(comps, mean, pre_scale_factor, invMassL, standarize, orthogonal, mass_weight):
    """snapbases/posComponents.py:275-293."""
    comps = comps.copy()
    if standarize:
        comps /= pre_scale_factor
        comps += mean[np.newaxis]
    if orthogonal:
        for l in range(comps.shape[2]):
            comps[:, :, l] = sla.orth(comps[:, :, l].T).T
    if mass_weight:
        comps *= invMassL[:, None]
    return comps


# ======================================================================================= POD (a9, a10)
def pod_vectorized(X, k_desired):
    """snapbases/constraintsComponents.py:298-320: thin SVD of the (3ep x F) matrix.
    Returns comps (min(k_desired, F), ep, 3) and all F singular values."""
    F = X.shape[0]
    ep = X.shape[1]
    U, S, _ = sla.svd(X.reshape(F, -1).T, full_matrices=False)
    C = U.T.reshape((F, ep, -1))
    if k_desired < C.shape[0]:
        C = C[:k_desired]
    return np.array(C), S


def pod_per_row_f32(X, e, p, k_desired):
    """snapbases/constraintsComponents.py:274-294 (`pod`): p*3 independent float32 SVDs of (e x F)
    blocks (torch CPU in the reference; LAPACK sgesdd here -- same algorithm family, float32)."""
    F = X.shape[0]
    R = X.reshape(F, e, p, 3)
    R = np.moveaxis(R, [0, 1, 2, 3], [3, 2, 0, 1]).astype(np.float32)      # (p, d, e, F)
    Cs = np.empty((p, 3, e, min(e, F)), dtype=np.float32)
    for i in range(p):
        for d in range(3):
            Cs[i, d] = sla.svd(R[i, d], full_matrices=False)[0]
    C = np.moveaxis(Cs, [0, 1, 2, 3], [2, 3, 1, 0])
    C = C.reshape(C.shape[0], e * p, C.shape[3])
    if k_desired < C.shape[0]:
        C = C[:k_desired]
    return np.array(C)


# ======================================================================================= block deflation (a11)
def index_largest_deformation(R_epFd, p, e):
    """snapbases/constraintsComponents.py:86-96: element with the largest energy over its p rows."""
    magnitude = (R_epFd ** 2).sum(axis=2).reshape(e, p, -1)
    return int(np.argmax(magnitude.sum(axis=(1, 2))))


def nonlinearity_bases_blocks(X, e, p, k_desired, on_row=None):
    """snapbases/constraintsComponents.py:324-412 (global support only; local raises ValueError :376)."""
    R = X.copy()
    C, W, S_idx, S_p, meas = [], [], [], [], []
    for k in range(k_desired):
        idx = index_largest_deformation(np.swapaxes(R, 0, 1), p, e)
        S_idx.append(idx)
        sigma = []
        for i in range(p):
            _, sing, Vt = sla.svd(R[:, idx * p + i, :].reshape(R.shape[0], -1).T, full_matrices=False)
            wk = sing[0] * Vt[0, :]
            ck = np.tensordot(wk, R, (0, 0)) / np.inner(wk, wk)
            R -= np.outer(wk, ck).reshape(R.shape)
            S_p.append(idx * p + i)
            C.append(ck)
            W.append(wk)
            sigma.append(sing[0])
        row = [k, idx, sla.norm(R)] + sigma
        meas.append(row)
        if on_row is not None:
            on_row(row)
    return np.array(C), np.array(W).T, np.array(meas), np.array(S_idx), np.array(S_p)


def nonlinearity_bases_blocks_with_St(X, e, p, St, elems, tol, on_row=None):
    """snapbases/constraintsComponents.py:156-271 (`pca_blocks_with_St`, global support, '_tets' / '_tris' / '_edges').
    While |R| > tol: the vertex with the largest residual pushed through St (:178) fixes HOW MANY blocks follow (its incident
    elements); the blocks themselves deflate rows  idx * p + i  with idx the LOOP COUNTER, not the element id (:203-209, SURVEY
    H6) -- every rank-1 step projects one frame-space direction out of ALL rows, so the loop ends once rank(X) directions are
    gone.  Returns comps (nb p, ep, 3), weigs (F, nb p), measures rows [count, idx, |R|, sigma_0..], picked vertices."""
    R = X.copy()
    C, W, meas, verts = [], [], [], []
    count = 0
    elems = np.asarray(elems)
    while sla.norm(R) > tol:
        v = int(np.argmax(((St @ np.swapaxes(R, 0, 1).reshape(e * p, -1)) ** 2).sum(axis=1)))       # :178
        nel = int((elems == v).any(axis=1).sum())                                                    # utils/support.py:210-257
        verts.append(v)
        for idx in range(nel):
            sigma = []
            for i in range(p):
                _, sing, Vt = sla.svd(R[:, idx * p + i, :].reshape(R.shape[0], -1).T, full_matrices=False)
                wk = sing[0] * Vt[0, :]
                ck = np.tensordot(wk, R, (0, 0)) / np.inner(wk, wk)
                R -= np.outer(wk, ck).reshape(R.shape)
                C.append(ck)
                W.append(wk)
                sigma.append(sing[0])
            count += 1
            row = [count, idx, sla.norm(R)] + sigma
            meas.append(row)
            if on_row is not None:
                on_row(row)
            if sla.norm(R) < tol:
                break
    return np.array(C), np.array(W).T, np.array(meas), np.array(verts, dtype=np.int64)


def post_process_constraints(comps, snapTensor, mean, pre_scale_factor, invMassL, standarize, orthogonal, mass_weight):
    """snapbases/constraintsComponents.py:415-446.  Returns (comps', snapTensor')."""
    comps, snap = comps.copy(), snapTensor.copy()
    if standarize:
        comps /= pre_scale_factor
        comps += mean[np.newaxis]
        snap /= pre_scale_factor
        snap += mean[np.newaxis]
    if orthogonal:
        for l in range(comps.shape[2]):
            comps[:, :, l] = sla.qr(comps[:, :, l].T, mode="economic")[0].T
    if mass_weight:
        comps *= invMassL[:, None]
        snap *= invMassL[:, None]
    return comps, snap


# ======================================================================================= DEIM (a13, a14)
def deim(comps, p):
    """snapbases/constraintsComponents.py:797-860.  comps: (K, ep, 3).
    Returns dict(Pt, alpha, alpha_ranges, residual_norms) or None on a zero residual (:832-834)."""
    bases = comps.swapaxes(0, 1)                       # (ep, K, 3)
    K = comps.shape[0]
    Pt, alphas, ranges = [], [], []
    V = None
    for k in range(K):
        if k == 0:
            r = bases[:, 0, :]
            vk = r
        else:
            vk = bases[:, k, :]
            c = np.empty(vk.shape)
            for i in range(3):
                c[:, i] = V[:, :, i] @ np.linalg.lstsq(V[Pt, :, i], vk[Pt, i], rcond=None)[0]
            r = c - vk
            if np.allclose(r, np.zeros(r.shape)):
                return None
        idx = int(np.argmax((r ** 2).sum(axis=1)))
        Pt.append(idx)
        alphas.append(idx // p)
        ranges.append(k + 1)
        V = vk[:, None, :] if k == 0 else np.concatenate((V, vk[:, None, :]), axis=1)
    return dict(Pt=np.array(Pt), alpha=np.array(alphas), alpha_ranges=np.array(ranges))


def deim_blocks(comps, p):
    """snapbases/constraintsComponents.py:733-795: p rows per pick, comps (K*p, ep, 3)."""
    bases = comps.swapaxes(0, 1)                       # (ep, Kp, 3)
    K = comps.shape[0] // p
    Pt, alphas, ranges, idxs = [], [], [], []
    V = None
    for k in range(K):
        vk = bases[:, k * p:(k + 1) * p, :]
        if k == 0:
            r = vk
        else:
            c = np.empty(vk.shape)
            for i in range(3):
                c[:, :, i] = V[:, :, i] @ np.linalg.lstsq(V[Pt, :, i], vk[Pt, :, i], rcond=None)[0]
            r = c - vk
            if np.allclose(r, np.zeros(r.shape)):
                return None
        idx = int(np.argmax((r ** 2).sum(axis=(1, 2))))
        alpha = idx // p
        idxs.append(idx)
        alphas.append(alpha)
        Pt += [alpha * p + m for m in range(p)]
        ranges.append(k + 1)
        V = vk if k == 0 else np.concatenate((V, vk), axis=1)
    return dict(Pt=np.array(Pt), alpha=np.array(alphas), alpha_ranges=np.array(ranges), idxs=np.array(idxs))


def elements_per_vert(v, elems):
    """utils/support.py:210-257 (get_triangles/tetrahedrons/edges_per_vert): ascending ids of the elements containing v."""
    return [int(i) for i in np.nonzero((np.asarray(elems) == v).any(axis=1))[0]]


def vert_star_per_vert(v, faces):
    """utils/support.py:239-246 (get_vert_star_per_vert as constraintsComponents.py:688 calls it): the vertices of all
    faces containing v, v included, in the iteration order of a CPython set filled face by face (`list(neighbors)`)."""
    neighbors = set()
    for face in np.asarray(faces):
        if v in face:
            neighbors.update(face)
    return list(neighbors)


def deim_geom(comps, p, e, elems, St=None, ele_per_vert=1, star_faces=None, constrained_verts=None):
    """snapbases/constraintsComponents.py:619-731 (geom_block_form_utilizing_differential_operator).  Element types
    '_tets' / '_tris' / '_edges': incident elements by get_*_per_vert on `elems`.  '_verts' (:687-692): pass the triangles as
    `star_faces`, the candidates are the vertex star; with `constrained_verts` (constProj_snapshots_type 'verts_bending')
    the star is intersected with the constrained vertices (np.intersect1d: ascending) and the selected ROW is the position
    in that list (:700-702, p == 1) while geom_alpha keeps the vertex id.  comps (K p, ep, 3).  St (|V| x ep, sparse or dense) given: error measured in
    position space (error_in_pos_space=True: worst VERTEX, up to ele_per_vert new incident elements per step, rectangular
    least squares on V[Pt]); St None: error in constraint space (worst ELEMENT by block energy, :86-96).
    Returns dict(Pt, alpha, alpha_ranges, verts) or None on a zero residual (:676-678)."""
    bases = comps.swapaxes(0, 1)                       # (ep, Kp, 3)
    K = comps.shape[0] // p
    Pt, e_points, e_jump, e_range, verts = [], [], [], [], []
    V = None
    for k in range(K):
        vk = bases[:, k * p:(k + 1) * p, :]
        if k == 0:
            r = vk
        else:
            c = np.empty(vk.shape)
            for i in range(3):
                c[:, :, i] = V[:, :, i] @ np.linalg.lstsq(V[Pt, :, i], vk[Pt, :, i], rcond=None)[0]
            r = c - vk
        if St is not None:
            r = St @ r.reshape(r.shape[0], -1)          # (|V|, p * 3)
        if k > 0 and np.allclose(r, np.zeros(r.shape)):
            return None
        if St is not None:
            v = int(np.argmax((r ** 2).sum(axis=1)))
            verts.append(v)
            jump = 0
            mapped = None
            if star_faces is not None:
                alpha_list = vert_star_per_vert(v, star_faces)
                if constrained_verts is not None:
                    alpha_list, mapped, _ = np.intersect1d(constrained_verts, alpha_list, return_indices=True)
            else:
                alpha_list = elements_per_vert(v, elems)
            for al, alpha in enumerate(alpha_list):
                if alpha not in e_points and jump < ele_per_vert:
                    jump += 1
                    e_points.append(alpha)
                    Pt += [mapped[al]] if mapped is not None else [alpha * p + m for m in range(p)]
            e_jump.append(jump)
        else:
            alpha = index_largest_deformation(r, p, e)
            assert alpha not in e_points
            e_points.append(alpha)
            Pt += [alpha * p + m for m in range(p)]
            e_jump.append(1)
        e_range.append(int(sum(e_jump)))
        V = vk if k == 0 else np.concatenate((V, vk), axis=1)
    return dict(Pt=np.array(Pt), alpha=np.array(e_points), alpha_ranges=np.array(e_range), verts=np.array(verts))


def deim_reconstruct(comps, frames, Pt_rows, r, p_block=1):
    """snapbases/constraintsComponents.py:489-521 (`geom_constructed`) for an explicit row set:
    per-dimension normal-equation solve on the selected rows, then expansion with the full basis."""
    V_r = comps.swapaxes(0, 1)[:, :r * p_block, :]
    F, ep, _ = frames.shape
    out = np.zeros((F, ep, 3))
    for l in range(3):
        A = V_r[Pt_rows, :, l]
        lu = sla.lu_factor(A.T @ A)
        for f in range(F):
            out[f, :, l] = V_r[:, :, l] @ sla.lu_solve(lu, A.T @ frames[f, Pt_rows, l])
    return out


# ======================================================================================= consumer side
def prepare_reduced_group(local_data, num_components, row_dim, assembly_ST, reduction_type):
    """projective_dynamics/Simulators.py:157-218 (reduced_position False).  local_data: mapping with the `.npz` keys
    `components (K, ep, 3)`, `interpol_alphas`, `Pt`, `interpol_alpha_ranges` written by
    constraintsComponents.store_components_n_interpol_points (:596-611).  Returns (interpolation_alpha, Pt, S^T V (N, mp, 3),
    [regularised normal matrices (mp, mp) per dimension], PtV_T (mp, |Pt|, 3))."""
    pod_rows = reduction_type in {"deim_pod", "deim_pod_vectorized"}
    if pod_rows:
        row_dim = 1                                                                          # :173-174
    Vj = np.swapaxes(np.asarray(local_data["components"]), 0, 1)[:, :num_components * row_dim, :]   # :180-181
    alpha_range = int(np.asarray(local_data["interpol_alpha_ranges"])[num_components - 1])  # :182
    interpolation_alpha = np.asarray(local_data["interpol_alphas"])[:alpha_range]           # :184
    if pod_rows:
        Pt = np.asarray(local_data["Pt"])[:alpha_range]                                      # :187-188
    else:
        Pt = np.array([a * row_dim + l for a in interpolation_alpha for l in range(row_dim)], dtype=np.int64)   # :190-193
    St = sparse.csr_matrix(assembly_ST)
    projecting_mat = np.stack([St @ Vj[:, :, d] for d in range(3)], axis=2)                  # :197 (einsum on the dense St)
    PtV = Vj[Pt, :, :]                                                                       # :205
    PtV_T = np.swapaxes(PtV, 0, 1)                                                           # :206
    AtA = np.einsum("nai,ami->nmi", PtV_T, PtV)                                              # :207
    la = 1e-8 * np.trace(AtA) / AtA.shape[0]                                                 # :209 (trace over axes 0,1 -> (3,))
    normal = [AtA[:, :, d] + la[d] * np.eye(AtA.shape[0]) for d in range(3)]                 # :214
    return interpolation_alpha, Pt, projecting_mat, normal, PtV_T


def group_reduced_term(p_rows, projecting_mat, normal, PtV_T):
    """projective_dynamics/Simulators.py:394-399: S^T V (A^T A + la I)^-1 A^T p per dimension (lu_factor / lu_solve);
    p_rows (|Pt|, 3) = the constraint projections at the interpolation rows (:382-392)."""
    cols = []
    for d in range(3):
        coef = sla.lu_solve(sla.lu_factor(normal[d]), PtV_T[:, :, d] @ p_rows[:, d])
        cols.append(projecting_mat[:, :, d] @ coef)
    return np.column_stack(cols)


# ======================================================================================= error measures / writers
def frobenius_error(f, g):
    """snapbases/posComponents.py:221-227."""
    return sla.norm(f - g)


def relative_error_per_component(f, g):
    """snapbases/posComponents.py:229-241."""
    return [sla.norm(f[:, :, i] - g[:, :, i]) / sla.norm(f[:, :, i]) for i in range(3)]


def max_pointwise_error(f, g):
    """snapbases/posComponents.py:243-250."""
    return np.max(np.abs(f - g)) / np.max(f)


def bases_sing_vals(comps):
    """snapbases/posComponents.py:344-356: per-dimension normalised singular values of (K, N)."""
    s = np.empty((comps.shape[0], 3))
    for i in range(3):
        sv = sla.svd(comps[:, :, i], compute_uv=False)
        s[:, i] = sv / sv.max()
    return s


def components_bin_bytes(B):
    """utils/utils.py:14-35 as bytes: `<i N><i 3K>` then doubles in the order d, k, i."""
    K, N, dim = B.shape
    out = bytearray(struct.pack("<i", N) + struct.pack("<i", dim * K))
    for d in range(dim):
        for k in range(K):
            for i in range(N):
                out += struct.pack("<d", B[k, i, d])
    return bytes(out)


def points_bin_bytes(points):
    """utils/utils.py:77-95: `<i n><i 1>` then the indices as little-endian doubles."""
    out = bytearray(struct.pack("<i", len(points)) + struct.pack("<i", 1))
    for v in points:
        out += struct.pack("<d", v)
    return bytes(out)
