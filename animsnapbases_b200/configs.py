"""Workload presets for BASELINE.json's five configs and their INPUT generators.

The reference ships no snapshot data, so every config runs on a synthetic animation of the config's own rest mesh
(tests/golden/meshes.npz: bunny, armadillo, bar; planar grids for the synthetic cases).  The inputs of the at-scale parity
fixtures (tests/golden/scale_*.npz) are far too large to commit, so they are REGENERATED wherever a test or a bench leg runs
-- in the build container next to the unmodified reference (oracle/make_golden_scale.py) and on the GPU box.  That only
works if generation is bit-reproducible across CPUs, so the generators here use nothing but correctly rounded IEEE
arithmetic in a fixed order: + - * / sqrt on numpy arrays (one ufunc per operation: no fused multiply-add, no reduction
whose association depends on the SIMD width), `math.sin / math.cos` on Python floats (libm of the image), and
numpy.random.Generator streams.  No np.exp, no BLAS (einsum / matmul / LAPACK).  `input_digest` is stored in every golden
and re-checked before any comparison.

No hot-path arithmetic lives here: these functions define inputs only.
"""
from __future__ import annotations

import hashlib
import math

import numpy as np

from . import synth

# name -> parameters.  `kind`: "pos" = posComponents path (deflation 'PCA' with local support [+ SPLOCS]);
# "bar" = constraintsComponents path (pod_vectorized -> post-process -> deim).
PRESETS = {
    # config 1: config/examples/bunny_gFall_posSubspace.json:16-46 (K = 200, local 0.1 / 0.25, rest 'first', standardised, Volkwein)
    "bunny_pca": dict(kind="pos", mesh="bunny", frames=200, comps=200, bumps=12, sigma=0.15, seed=0, splocs=None),
    # config 2: same + splocs 20 x 10, lambda 2, rho 10 (:47-54)
    "bunny_splocs": dict(kind="pos", mesh="bunny", frames=200, comps=200, bumps=12, sigma=0.15, seed=0,
                         splocs=(20, 10, 2.0, 10.0)),
    # config 3: config/examples/bar_automated_deim_tetDeformationGradientSubspace.json:40-61 (3 120 tets, every 2nd of 140 frames)
    "bar_deim": dict(kind="bar", mesh="bar", frames=70, comps=200, seed=0),
    # config 4: armadillo, F = 200 (config/examples/armadillo_gFall_tetstrainSubspace.json:36-38), as config 2
    "armadillo_splocs": dict(kind="pos", mesh="armadillo", frames=200, comps=200, bumps=12, sigma=0.15, seed=4,
                             splocs=(20, 10, 2.0, 10.0)),
    # config 4 at the longer of the two frame counts SURVEY 8(d) names for "full range of motion" (F = 200 and F = 1 000)
    "armadillo_splocs_f1000": dict(kind="pos", mesh="armadillo", frames=1000, comps=200, bumps=12, sigma=0.15, seed=4,
                                   splocs=(20, 10, 2.0, 10.0)),
    # config 5's CPU-runnable sample: N = 1e5, F = 500, K = 200, 2 x 10 SPLOCS
    "synthetic_sample": dict(kind="pos", mesh="grid", grid=(400, 250), frames=500, comps=200, bumps=32, sigma=0.08, seed=0,
                             splocs=(2, 10, 2.0, 10.0)),
    # config 5 (the headline): generated on the device per shard by bench.py, listed here for its parameters only
    "synthetic": dict(kind="pos", mesh="grid", grid=(1000, 1000), frames=2000, comps=200, bumps=64, sigma=0.05, seed=0,
                      splocs=(50, 10, 2.0, 10.0)),
}


def _digest(*arrays) -> str:
    h = hashlib.sha256()
    for a in arrays:
        a = np.ascontiguousarray(a)
        h.update(str(a.dtype).encode() + str(a.shape).encode())
        h.update(a.tobytes())
    return h.hexdigest()


def _seq_sum(a):
    """Sum in index order (np.cumsum is sequential): independent of the reduction blocking of np.sum."""
    return float(np.cumsum(np.asarray(a, dtype=np.float64).ravel())[-1])


def exact_bump_animation(verts0, frames, n_bumps, sigma, seed, amp=0.05):
    """(F, N, 3) float32: x_f = x_0 + sum_b a_b(f) g_b(x_0) with rational bumps g_b = 1 / (1 + |x - c_b|^2 / (2 sigma^2))^2
    and random-walk amplitudes (the generator of synth.bump_animation without np.exp / einsum / the per-frame translation:
    the frames are already 'centered', so no alignment pass is needed before the build)."""
    v = np.asarray(verts0, dtype=np.float64)
    centres, amps, _ = synth.bump_parameters(v, frames, n_bumps, seed=seed, amp=amp, translate=0.0)
    inv = 1.0 / (2.0 * sigma * sigma)
    out = np.empty((frames, len(v), 3), dtype=np.float32)
    g = np.empty((n_bumps, len(v)))
    for b in range(n_bumps):
        d = v - v[centres[b]]
        d2 = d[:, 0] * d[:, 0]
        d2 = d2 + d[:, 1] * d[:, 1]
        d2 = d2 + d[:, 2] * d[:, 2]
        t = 1.0 + d2 * inv
        g[b] = 1.0 / (t * t)
    for f in range(frames):
        x = v.copy()
        for b in range(n_bumps):
            for c in range(3):
                x[:, c] = x[:, c] + g[b] * amps[f, b, c]
        out[f] = x.astype(np.float32)
    return out


def exact_vertex_masses(verts, tris):
    """Barycentric lumped vertex areas normalised to m / sum(m) * 2 (the mass .bin the configs supply so that the reference's
    libigl branch is not taken, snapbases/posSnapshots.py:131-139), in reproducible arithmetic."""
    v = np.asarray(verts, dtype=np.float64)
    a, b, c = v[tris[:, 0]], v[tris[:, 1]], v[tris[:, 2]]
    e1, e2 = b - a, c - a
    cx = e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1]
    cy = e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2]
    cz = e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]
    area = np.sqrt(cx * cx + cy * cy + cz * cz) * 0.5
    m = np.zeros(len(v))
    for k in range(3):
        m = m + np.bincount(tris[:, k], area / 3.0, minlength=len(v))
    return m / _seq_sum(m) * 2.0


def exact_unit_cube(verts):
    """Rest mesh centred and scaled to unit extent (the scale convention of utils/process.py:140-144 that d_min / d_max assume)."""
    v = np.asarray(verts, dtype=np.float64)
    mean = np.array([_seq_sum(v[:, c]) / len(v) for c in range(3)])
    v = v - mean
    return v / float((v.max(axis=0) - v.min(axis=0)).max())


def _bar_frame(verts0, bend, twist, sag):
    """Twist about x, bend about z and a quadratic sag of a bar whose long axis is x in [-0.5, 0.5] (math.sin / cos per vertex).
    NB: on a structured mesh under a smooth motion neighbouring tets can have deformation gradients that agree to round-off
    (two tets between the same two cross-sections): DEIM then meets exact ties, which the parity checks treat as such."""
    out = np.empty_like(verts0)
    for i in range(len(verts0)):
        px, py, pz = (float(t) for t in verts0[i])
        s = px + 0.5
        ct, st = math.cos(twist * s), math.sin(twist * s)
        qx, qy, qz = px, ct * py - st * pz, st * py + ct * pz
        cb, sb = math.cos(bend * s), math.sin(bend * s)
        out[i, 0] = cb * qx - sb * qy
        out[i, 1] = sb * qx + cb * qy - sag * s * s
        out[i, 2] = qz
    return out


def _inv3(A):
    """Inverse of a stack of 3 x 3 matrices by cofactors (elementwise arithmetic only)."""
    a, b, c = A[:, 0, 0], A[:, 0, 1], A[:, 0, 2]
    d, e, f = A[:, 1, 0], A[:, 1, 1], A[:, 1, 2]
    g, h, i = A[:, 2, 0], A[:, 2, 1], A[:, 2, 2]
    c00, c01, c02 = e * i - f * h, f * g - d * i, d * h - e * g
    det = a * c00 + b * c01 + c * c02
    out = np.empty_like(A)
    out[:, 0, 0], out[:, 0, 1], out[:, 0, 2] = c00, c * h - b * i, b * f - c * e
    out[:, 1, 0], out[:, 1, 1], out[:, 1, 2] = c01, a * i - c * g, c * d - a * f
    out[:, 2, 0], out[:, 2, 1], out[:, 2, 2] = c02, b * g - a * h, a * e - b * d
    return out / det[:, None, None], det


def _mul3(A, B):
    out = np.empty_like(A)
    for r in range(3):
        for c in range(3):
            out[:, r, c] = A[:, r, 0] * B[:, 0, c] + A[:, r, 1] * B[:, 1, c] + A[:, r, 2] * B[:, 2, c]
    return out


def exact_tet_rotation_snapshots(verts0, tets, frames, seed):
    """(F, 3e, 3) float64 constraint-projection snapshots of the tets_deformation_gradient group: rows = R^T of the polar
    rotation of F_t = Ds Dm^-1 (what projective_dynamics/Constraint_projections.py:669-687 returns for an un-inverted tet),
    the rotation by 30 Newton steps R <- (R + R^-T) / 2 instead of a LAPACK SVD (reproducible; converged to machine
    precision for these deformations)."""
    v = np.asarray(verts0, dtype=np.float64)
    rng = np.random.default_rng(seed)
    path = np.cumsum(rng.standard_normal((frames, 3)), axis=0) / math.sqrt(frames)
    path = path + (np.arange(frames) / float(max(frames - 1, 1)))[:, None] * np.array([1.2, 2.0, 0.3])
    e = len(tets)
    Dm = np.stack([v[tets[:, 0]] - v[tets[:, 3]], v[tets[:, 1]] - v[tets[:, 3]], v[tets[:, 2]] - v[tets[:, 3]]], axis=2)
    DmInv, _ = _inv3(Dm)
    out = np.empty((frames, 3 * e, 3))
    for f in range(frames):
        q = _bar_frame(v, path[f, 0], path[f, 1], path[f, 2])
        Ds = np.stack([q[tets[:, 0]] - q[tets[:, 3]], q[tets[:, 1]] - q[tets[:, 3]], q[tets[:, 2]] - q[tets[:, 3]]], axis=2)
        R = _mul3(Ds, DmInv)
        for _ in range(30):
            Ri, _ = _inv3(R)
            R = (R + np.swapaxes(Ri, 1, 2)) * 0.5
        out[f] = np.swapaxes(R, 1, 2).reshape(3 * e, 3)
    return out


def exact_tet_masses(verts, tets, p=3):
    """Element masses (sum of the 4 volume-lumped vertex masses, utils/support.py:12-23, total vertex mass 1) over the p rows."""
    v = np.asarray(verts, dtype=np.float64)
    a, b, c, d = (v[tets[:, i]] for i in range(4))
    e1, e2, e3 = b - a, c - a, d - a
    cx = e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1]
    cy = e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2]
    cz = e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]
    vol = np.abs(cx * e3[:, 0] + cy * e3[:, 1] + cz * e3[:, 2]) / 6.0
    mv = np.zeros(len(v))
    for i in range(4):
        mv = mv + np.bincount(tets[:, i], vol / 4.0, minlength=len(v))
    mv = mv / _seq_sum(mv)
    me = mv[tets[:, 0]] + mv[tets[:, 1]] + mv[tets[:, 2]] + mv[tets[:, 3]]
    return np.repeat(me, p)


def build_inputs(name):
    """Host inputs of a preset, bit-reproducible.  kind 'pos': dict(anim (F, N, 3) f32, tris, mass (N,), rest (N, 3) f64);
    kind 'bar': dict(snaps (F, 3e, 3) f64, mass (3e,), tets, verts, p).  Both carry `digest` and `preset`."""
    cfg = dict(PRESETS[name])
    if cfg["kind"] == "pos":
        if cfg["mesh"] == "grid":
            verts0, tris = synth.grid_mesh(*cfg["grid"])
        else:
            m = synth.golden_mesh(cfg["mesh"])
            verts0, tris = exact_unit_cube(m["verts"]), m["tris"]
        anim = exact_bump_animation(verts0, cfg["frames"], cfg["bumps"], cfg["sigma"], cfg["seed"])
        rest = anim[0].astype(np.float64)
        mass = exact_vertex_masses(rest, tris)
        return dict(preset=cfg, anim=anim, tris=np.asarray(tris, dtype=np.int64), mass=mass, rest=rest,
                    digest=_digest(anim, mass))
    m = synth.golden_mesh("bar")
    verts, tets = m["verts"], m["tets"]
    snaps = exact_tet_rotation_snapshots(verts, tets, cfg["frames"], cfg["seed"])
    mass = exact_tet_masses(verts, tets, 3)
    return dict(preset=cfg, snaps=snaps, mass=mass, tets=tets, verts=verts, p=3, digest=_digest(snaps, mass))


def sketch_matrix(n_rows, cols=8, seed=12345):
    """Seeded Gaussian test matrix (n_rows, 3, cols) for the compact fixtures: a (K, N, 3) field is committed as its K x cols
    sketch sum_{n,d} C[k, n, d] G[n, d, j] (plus norms and a few full rows), and compared through the same sketch."""
    return np.random.default_rng(seed).standard_normal((n_rows, 3, cols))


def sketch(C, G):
    """(K, N, 3) x (N, 3, cols) -> (K, cols), accumulated in float64 (comparison helper, host side)."""
    C = np.asarray(C, dtype=np.float64)
    return C.reshape(C.shape[0], -1) @ G.reshape(-1, G.shape[2])
