"""Synthetic snapshot inputs.

The reference ships no snapshot data (input_data/ and results/ are git-ignored, *.h5/*.npy are
missing LFS blobs), so every config runs on synthetic animations built on rest meshes:

  * position snapshots: `x_f = x_0 + sum_b a_b(f) g_b(x_0)` with Gaussian bumps `g_b`, smooth random
    amplitudes and a small per-frame translation, rounded to float32 exactly like the reference's
    `.h5` pipeline does (utils/process.py:85,246) before posSnapshots.read casts to float64
    (snapbases/posSnapshots.py:110);
  * constraint-projection snapshots for tetrahedral meshes: per-tet polar rotation of the
    deformation gradient, rows = R^T (projective_dynamics/Constraint_projections.py:669-687).

Host generators are plain numpy (they define the *inputs*; no hot-path arithmetic lives here).
`device_bump_animation` builds the large 1M x 2000 workload directly in HBM, shard by shard.
"""
from __future__ import annotations

import os

import numpy as np

_GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


# --------------------------------------------------------------------------- meshes
def grid_mesh(nx: int, ny: int):
    """Planar nx x ny vertex grid in [-0.5, 0.5]^2 (z = 0), two triangles per cell."""
    xs = np.linspace(-0.5, 0.5, nx)
    ys = np.linspace(-0.5, 0.5, ny)
    gx, gy = np.meshgrid(xs, ys, indexing="ij")
    verts = np.stack([gx.ravel(), gy.ravel(), np.zeros(nx * ny)], axis=1)
    idx = np.arange(nx * ny, dtype=np.int64).reshape(nx, ny)
    a = idx[:-1, :-1].ravel()
    b = idx[1:, :-1].ravel()
    c = idx[1:, 1:].ravel()
    d = idx[:-1, 1:].ravel()
    tris = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], axis=0)
    return verts, tris


def tet_grid(nx: int, ny: int, nz: int):
    """Structured tetrahedral block: (nx+1)(ny+1)(nz+1) vertices in the unit cube scaled by the cell counts, every cell
    split into 6 positively oriented tets around its main diagonal.  Returns (verts (V, 3) float64, tets (6 nx ny nz, 4))."""
    gx, gy, gz = np.meshgrid(np.arange(nx + 1), np.arange(ny + 1), np.arange(nz + 1), indexing="ij")
    verts = np.stack([gx.ravel() / max(nx, ny, nz), gy.ravel() / max(nx, ny, nz), gz.ravel() / max(nx, ny, nz)], axis=1).astype(np.float64)
    vid = lambda i, j, k: (i * (ny + 1) + j) * (nz + 1) + k          # noqa: E731
    ci, cj, ck = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    ci, cj, ck = ci.ravel(), cj.ravel(), ck.ravel()
    c = {(a, b, d): vid(ci + a, cj + b, ck + d) for a in (0, 1) for b in (0, 1) for d in (0, 1)}
    paths = [((1, 0, 0), (1, 1, 0)), ((1, 0, 0), (1, 0, 1)), ((0, 1, 0), (1, 1, 0)), ((0, 1, 0), (0, 1, 1)),
             ((0, 0, 1), (1, 0, 1)), ((0, 0, 1), (0, 1, 1))]                 # the 6 monotone paths from corner 000 to 111
    tets = np.concatenate([np.stack([c[(0, 0, 0)], c[p1], c[p2], c[(1, 1, 1)]], axis=1) for p1, p2 in paths])
    x = verts
    vol = np.einsum("ij,ij->i", np.cross(x[tets[:, 1]] - x[tets[:, 0]], x[tets[:, 2]] - x[tets[:, 0]]), x[tets[:, 3]] - x[tets[:, 0]])
    flip = vol < 0
    tets[flip] = tets[flip][:, [0, 2, 1, 3]]
    return verts, tets.astype(np.int64)


def read_obj_mesh(path: str):
    """Minimal Wavefront OBJ reader: `v x y z` and triangular `f a b c` (1-based, a/b/c forms)."""
    vs, fs = [], []
    with open(path, "r", errors="replace") as fh:
        for line in fh:
            if line.startswith("v "):
                vs.append([float(t) for t in line.split()[1:4]])
            elif line.startswith("f "):
                fs.append([int(t.split("/")[0]) - 1 for t in line.split()[1:4]])
    return np.asarray(vs, dtype=np.float64), np.asarray(fs, dtype=np.int64)


def read_medit_mesh(path: str):
    """Minimal Medit `.mesh` reader -> (verts, tets, tris), zero-based indices."""
    verts, tets, tris = [], [], []
    with open(path, "r", errors="replace") as fh:
        toks = fh.read().split("\n")
    i = 0
    while i < len(toks):
        key = toks[i].strip()
        if key in ("Vertices", "Tetrahedra", "Triangles"):
            n = int(toks[i + 1].strip())
            rows = [toks[i + 2 + j].split() for j in range(n)]
            if key == "Vertices":
                verts = [[float(r[0]), float(r[1]), float(r[2])] for r in rows]
            elif key == "Tetrahedra":
                tets = [[int(t) - 1 for t in r[:4]] for r in rows]
            else:
                tris = [[int(t) - 1 for t in r[:3]] for r in rows]
            i += 2 + n
        else:
            i += 1
    return (np.asarray(verts, dtype=np.float64), np.asarray(tets, dtype=np.int64),
            np.asarray(tris, dtype=np.int64))


def golden_mesh(name: str):
    """Rest meshes committed as fixtures under tests/golden/meshes.npz (bunny, armadillo, bar).
    Returns a dict with verts/tris (and tets for the bar)."""
    data = np.load(os.path.join(_GOLDEN_DIR, "meshes.npz"))
    out = {"verts": data[f"{name}_verts"].astype(np.float64), "tris": data[f"{name}_tris"].astype(np.int64)}
    if f"{name}_tets" in data.files:
        out["tets"] = data[f"{name}_tets"].astype(np.int64)
    return out


def normalize_unit_cube(verts: np.ndarray):
    """Centre on the global mean and divide by the largest per-frame extent: the scale
    convention of preprocess_mesh_animation (utils/process.py:140-144) that d_min/d_max assume."""
    v = np.asarray(verts, dtype=np.float64)
    anim = v if v.ndim == 3 else v[None]
    mean = anim.mean(axis=0).mean(axis=0)
    anim = anim - mean
    scale = np.abs(anim.max(axis=1) - anim.min(axis=1)).max()
    anim = anim / scale
    return (anim if v.ndim == 3 else anim[0]), mean, scale


def lumped_vertex_masses(verts: np.ndarray, tris: np.ndarray):
    """Barycentric lumped vertex areas, normalised `m / sum(m) * 2` like the reference's igl branch
    (snapbases/posSnapshots.py:137-139) -- used to *generate* the mass file the configs need."""
    e1 = verts[tris[:, 1]] - verts[tris[:, 0]]
    e2 = verts[tris[:, 2]] - verts[tris[:, 0]]
    area = 0.5 * np.linalg.norm(np.cross(e1, e2), axis=1)
    m = np.zeros(len(verts))
    for c in range(3):
        m += np.bincount(tris[:, c], area / 3.0, minlength=len(verts))
    return m / m.sum() * 2.0


# --------------------------------------------------------------------------- animations
def bump_parameters(verts0: np.ndarray, frames: int, n_bumps: int, seed: int = 0, amp: float = 0.05,
                    translate: float = 0.02):
    """Random centres / smooth amplitudes / per-frame translations (seeded, host-side, tiny)."""
    rng = np.random.default_rng(seed)
    centres = rng.choice(len(verts0), size=n_bumps, replace=False)
    steps = rng.standard_normal((frames, n_bumps, 3))
    amps = amp * np.cumsum(steps, axis=0) / np.sqrt(frames)
    amps[0] = 0.0
    trans = translate * rng.standard_normal((frames, 3))
    trans[0] = 0.0
    return centres, amps, trans


def bump_animation(verts0: np.ndarray, frames: int, n_bumps: int = 12, sigma: float = 0.15, seed: int = 0,
                   amp: float = 0.05, translate: float = 0.02) -> np.ndarray:
    """(F, N, 3) float32 animation: Gaussian-bump displacement fields with smooth random amplitudes."""
    centres, amps, trans = bump_parameters(verts0, frames, n_bumps, seed, amp, translate)
    d2 = ((verts0[:, None, :] - verts0[centres][None, :, :]) ** 2).sum(axis=2)  # (N, B)
    g = np.exp(-d2 / (2.0 * sigma * sigma))
    disp = np.einsum("nb,fbc->fnc", g, amps)
    anim = verts0[None] + disp + trans[:, None, :]
    return anim.astype(np.float32)


def device_bump_animation_from_params(verts0_dev, centres_xyz, amps, sigma: float, frame_chunk: int = 25):
    """verts0_dev: torch CUDA (n,3) f64; centres_xyz: (B,3) numpy; amps: (F,B,3) numpy.
    Returns torch CUDA float32 (F, n, 3) = rest + bump displacement (no translation: the big
    workload is fed to the basis build directly, alignment has its own bench line)."""
    import torch

    dev = verts0_dev.device
    c = torch.as_tensor(centres_xyz, dtype=torch.float64, device=dev)
    a = torch.as_tensor(amps, dtype=torch.float64, device=dev)
    n = verts0_dev.shape[0]
    out = torch.empty((a.shape[0], n, 3), dtype=torch.float32, device=dev)
    row_chunk = 1 << 17
    for r0 in range(0, n, row_chunk):
        v = verts0_dev[r0:r0 + row_chunk]
        d2 = ((v[:, None, :] - c[None, :, :]) ** 2).sum(dim=2)
        g = torch.exp(-d2 / (2.0 * sigma * sigma))  # (rows, B)
        for f0 in range(0, a.shape[0], frame_chunk):
            disp = torch.einsum("nb,fbc->fnc", g, a[f0:f0 + frame_chunk])
            out[f0:f0 + frame_chunk, r0:r0 + row_chunk] = (v[None] + disp).to(torch.float32)
    return out


# --------------------------------------------------------------------------- constraint projections
def _rot(axis, ang):
    axis = np.asarray(axis, dtype=np.float64)
    axis = axis / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)


def bar_deformation(verts0: np.ndarray, t: float, seed_vec: np.ndarray) -> np.ndarray:
    """Smooth bend + twist + sag of a bar whose long axis is x in [-0.5, 0.5]."""
    x = verts0[:, 0] + 0.5
    out = np.empty_like(verts0)
    bend = seed_vec[0] * t
    twist = seed_vec[1] * t
    sag = seed_vec[2] * t
    for i, (p, s) in enumerate(zip(verts0, x)):
        q = _rot([1, 0, 0], twist * s) @ np.array([0.0, p[1], p[2]])
        q = q + np.array([p[0], 0.0, 0.0])
        q = _rot([0, 0, 1], bend * s) @ q
        q[1] -= sag * s * s
        out[i] = q
    return out


def bar_motion_frames(verts0: np.ndarray, frames: int, seed: int = 0) -> np.ndarray:
    """(F, |V|, 3) positions of the synthetic bar motion: bend / twist / sag along a smooth, non-periodic parameter path, so
    that the snapshot matrix built from it has a decaying spectrum."""
    rng = np.random.default_rng(seed)
    path = np.cumsum(rng.standard_normal((frames, 3)), axis=0) / np.sqrt(frames)
    path += np.linspace(0, 1, frames)[:, None] * np.array([1.2, 2.0, 0.3])
    return np.stack([bar_deformation(verts0, 1.0, path[f]) for f in range(frames)])


def tet_rotation_snapshots(verts0: np.ndarray, tets: np.ndarray, frames: int, seed: int = 0) -> np.ndarray:
    """(F, 3e, 3) float64: for each frame and tet, `R^T` of the polar decomposition of the deformation
    gradient `Ds Dm^{-1}` (restates Constraint_projections.py:669-687; det-fix flips the last column)."""
    e = len(tets)
    v = verts0
    Dm = np.stack([v[tets[:, 0]] - v[tets[:, 3]], v[tets[:, 1]] - v[tets[:, 3]], v[tets[:, 2]] - v[tets[:, 3]]], axis=2)
    DmInv = np.linalg.inv(Dm)
    positions = bar_motion_frames(verts0, frames, seed)
    out = np.empty((frames, 3 * e, 3))
    for f in range(frames):
        q = positions[f]
        Ds = np.stack([q[tets[:, 0]] - q[tets[:, 3]], q[tets[:, 1]] - q[tets[:, 3]], q[tets[:, 2]] - q[tets[:, 3]]], axis=2)
        Fg = Ds @ DmInv
        U, _, Vt = np.linalg.svd(Fg)
        R = U @ Vt
        neg = np.linalg.det(R) < 0
        R[neg, :, 2] *= -1
        out[f] = np.swapaxes(R, 1, 2).reshape(3 * e, 3)
    return out


def tet_element_masses(verts: np.ndarray, tets: np.ndarray, p: int = 3) -> np.ndarray:
    """Element masses replicated over the p rows of each element: sum of the 4 lumped vertex masses
    (utils/support.py:12-23 on top of a volume-lumped vertex mass normalised to total 1, :41-59)."""
    a, b, c, d = (verts[tets[:, i]] for i in range(4))
    vol = np.abs(np.einsum("ij,ij->i", np.cross(b - a, c - a), d - a)) / 6.0
    mv = np.zeros(len(verts))
    for i in range(4):
        mv += np.bincount(tets[:, i], vol / 4.0, minlength=len(verts))
    mv /= mv.sum()
    me = mv[tets].sum(axis=1)
    return np.repeat(me, p)


def tet_gradient_operator(verts: np.ndarray, tets: np.ndarray, weighted: bool = True):
    """The differential operator of tets_deformation_gradient constraints as a scipy CSR matrix S (3e x |V|): row 3t+i
    applied to the positions gives row i of F_t^T, F_t = Ds Dm^{-1} (same vertex order as `tet_rotation_snapshots`).
    `weighted` scales the rows of tet t by sqrt(volume_t) (the PD assembly weight).  The `St` the reference loads from
    assembly_ST.npz (snapbases/constraintsComponents.py:61-84) is the transpose, |V| x 3e."""
    from scipy import sparse
    e = len(tets)
    v = verts
    Dm = np.stack([v[tets[:, 0]] - v[tets[:, 3]], v[tets[:, 1]] - v[tets[:, 3]], v[tets[:, 2]] - v[tets[:, 3]]], axis=2)
    DmInv = np.linalg.inv(Dm)                                       # (e, 3, 3)
    w = np.sqrt(np.abs(np.linalg.det(Dm)) / 6.0) if weighted else np.ones(e)
    rows, cols, vals = [], [], []
    for i in range(3):
        r = 3 * np.arange(e) + i
        for j in range(3):
            rows.append(r); cols.append(tets[:, j]); vals.append(w * DmInv[:, j, i])
        rows.append(r); cols.append(tets[:, 3]); vals.append(-w * DmInv[:, :, i].sum(axis=1))
    S = sparse.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(3 * e, len(verts)))
    S.sum_duplicates()
    return S
