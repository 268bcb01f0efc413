"""TEST INFRASTRUCTURE ONLY -- numpy stand-in for animsnapbases_b200.ops.CudaOps.

Same method names and tensor conventions (torch CPU tensors, planar layout), each method a few lines of numpy
stating what the corresponding CUDA kernel must compute.  It lets the `-m "not gpu"` suite exercise the host-side
orchestration in animsnapbases_b200/engine.py (blocked deflation, Gram-form SPLOCS, DEIM stepping, row sharding
with gloo) against the oracle, and it is the per-kernel reference in the `-m gpu` parity tests.
It is never imported by the product.
"""
from __future__ import annotations

import numpy as np
import torch

F64 = torch.float64


def _np(t):
    return None if t is None else t.numpy()


class NumpyOps:
    name = "numpy"

    def __init__(self):
        self.device = torch.device("cpu")
        self.sm_count = 148
        self.launches = 0

    # plumbing -------------------------------------------------------------------------------------
    def fork(self, after=None, lane=0):
        import contextlib
        return contextlib.nullcontext()

    def join(self):
        pass

    def record_event(self):
        return None

    def empty(self, shape, dtype=F64):
        return torch.zeros(shape, dtype=dtype)

    def zeros(self, shape, dtype=F64):
        return torch.zeros(shape, dtype=dtype)

    def to_device(self, a, dtype=None):
        t = torch.as_tensor(np.ascontiguousarray(a))
        return t.to(dtype) if dtype is not None else t.clone()

    def to_host(self, t):
        return t.numpy()

    def launch_count(self):
        return self.launches

    def reset_launch_count(self):
        self.launches = 0

    def synchronize(self):
        pass

    # boundary -------------------------------------------------------------------------------------
    def ingest(self, src, row_scale, ld):
        R, N, _ = src.shape
        X = torch.zeros((R, 3, ld), dtype=F64)
        v = src.numpy().astype(np.float64)
        if row_scale is not None:
            v = v * row_scale.numpy()[None, :N, None]
        X.numpy()[:, :, :N] = np.swapaxes(v, 1, 2)
        return X

    def export_aos(self, T, row_scale, N):
        v = np.swapaxes(T.numpy()[:, :, :N], 1, 2).copy()
        if row_scale is not None:
            v = v * row_scale.numpy()[None, :N, None]
        return torch.from_numpy(np.ascontiguousarray(v))

    def pack_bin(self, T, K, N):
        return torch.from_numpy(np.ascontiguousarray(np.swapaxes(T.numpy()[:K, :, :N], 0, 1)))

    # centering ------------------------------------------------------------------------------------
    def col_mean(self, X):
        F = X.shape[0]
        s = np.zeros(X.shape[1:])
        for f in range(F):
            s += X.numpy()[f]
        return torch.from_numpy(s / F)

    def center_stats(self, X, mean, N):
        y = X.numpy()[:, :, :N] - mean.numpy()[None, :, :N]
        return torch.tensor([y.sum(), (y * y).sum()], dtype=F64)

    def center_apply(self, X, mean, scale, N):
        x = X.numpy()
        x -= mean.numpy()[None]
        x *= scale

    def unstandardize(self, T, mean, scale, N):
        t = T.numpy()
        t[:, :, :N] = t[:, :, :N] / scale + mean.numpy()[None, :, :N]

    # deflation ------------------------------------------------------------------------------------
    def vertex_energy(self, R, N, out=None):
        e = torch.zeros((R.shape[2],), dtype=F64) if out is None else out
        r = R.numpy()
        e.numpy()[:] = 0
        e.numpy()[:N] = ((r[:, 0, :N] ** 2 + r[:, 1, :N] ** 2) + r[:, 2, :N] ** 2).sum(axis=0)
        return e

    def argmax_outside_ball(self, v, n, index_offset, pos, centre, radius, out_pair):
        out_pair[0] = -np.inf
        out_pair[1:2].view(torch.int64)[0] = np.iinfo(np.int64).max
        if n == 0:
            return
        P = pos.numpy()
        d = P[index_offset:index_offset + n] - P[int(centre[0])]
        ok = (d * d).sum(axis=1) > float(radius) ** 2
        if not ok.any():
            return
        a = np.where(ok, v.numpy()[:n], -np.inf)
        i = int(np.argmax(a))
        out_pair[0] = float(a[i])
        out_pair[1:2].view(torch.int64)[0] = i + index_offset

    def select_pair(self, all_pairs, world, fallback, out_idx):
        A = all_pairs.reshape(-1)[:2 * world].reshape(world, 2)
        vals = A[:, 0].numpy()
        ids = A[:, 1].contiguous().view(torch.int64).numpy()
        best = vals.max()
        if best == -np.inf:
            out_idx[0] = int(fallback[0])
            return
        out_idx[0] = int(ids[vals == best].min())

    def wait_event(self, ev):
        pass

    def argmax_first(self, v, n, index_offset, out_val, out_idx):
        if n == 0:
            out_val[0] = -np.inf
            out_idx[0] = np.iinfo(np.int64).max
            return
        a = v.numpy()[:n]
        i = int(np.argmax(a))
        out_val[0] = float(a[i])
        out_idx[0] = i + index_offset

    def group_sum(self, v, n, group, out):
        out.numpy()[:n // group] = v.numpy()[:n].reshape(n // group, group).sum(axis=1)

    def gather_traj(self, R, N, idx_dev, row0, W, k0, C, npend, traj):
        i = int(idx_dev[0]) - row0
        if i < 0 or i >= N:
            return
        t = R.numpy()[:, :, i].T.copy()                      # (3, F)
        for j in range(npend):
            t -= np.outer(C.numpy()[k0 + j, :, i], W.numpy()[:, k0 + j])
        traj.numpy()[:] = t

    def gather_candidate(self, R, N, val, idx, row0, W, k0, C, npend, cand):
        F = R.shape[0]
        i = int(idx[0]) - row0
        c = cand.numpy()
        if i < 0 or i >= N:
            c[0] = -np.inf
            c[1:2].view(np.int64)[0] = np.iinfo(np.int64).max
            c[2:2 + 3 * F] = 0.0
            return
        c[0] = float(val[0])
        c[1:2].view(np.int64)[0] = int(idx[0])
        t = torch.zeros((3, F), dtype=F64)
        self.gather_traj(R, N, idx, row0, W, k0, C, npend, t)
        c[2:2 + 3 * F] = t.numpy().ravel()

    def select_candidate(self, cand_all, F, val, idx, traj):
        a = cand_all.numpy()
        best, bi, bw = -np.inf, np.iinfo(np.int64).max, 0
        for r in range(a.shape[0]):
            v, i = float(a[r, 0]), int(a[r, 1:2].view(np.int64)[0])
            if v > best or (v == best and i < bi):
                best, bi, bw = v, i, r
        val[0] = best
        idx[0] = bi
        traj.numpy()[:] = a[bw, 2:2 + 3 * F].reshape(3, F)

    def traj_rank1(self, traj, F, local_support, W, k, k0, coef, scal):
        T = traj.numpy()
        lam, V = np.linalg.eigh(T @ T.T)
        u = V[:, -1]
        if u[np.argmax(np.abs(u))] < 0:
            u = -u
        w = u @ T
        sigma0 = np.sqrt(max(lam[-1], 0.0))
        if local_support:
            def proj(x):
                x = np.maximum(0.0, x)
                m = x.max()
                return x if m == 0 else x / m
            a, b = proj(w), proj(-w)
            w = a if np.linalg.norm(a) > np.linalg.norm(b) else b
        W.numpy()[:, k] = w
        scal[0] = sigma0
        scal[1] = float(w @ w)
        for j in range(k - k0):
            coef[j] = float(w @ W.numpy()[:, k0 + j])

    def deflate_gemv(self, R, N, W, k, g):
        g.numpy()[:] = np.tensordot(W.numpy()[:, k], R.numpy(), (0, 0))

    def deflate_finish(self, g, C, k, k0, coef, s, scal, e, N, meas):
        gg = g.numpy().copy()
        for j in range(k - k0):
            gg -= float(coef[j]) * C.numpy()[k0 + j]
        ww = float(scal[1])
        sn = np.ones(gg.shape[1]) if s is None else s.numpy()
        c = gg * sn[None, :] / ww
        c[:, N:] = 0
        C.numpy()[k] = c
        g2 = (gg ** 2).sum(axis=0)
        en = e.numpy()
        en[:N] = np.maximum(en[:N] - ((2 * sn - sn * sn) * g2 / ww)[:N], 0.0)
        tot = float(en[:N].sum())
        meas[0] = float(scal[0])
        meas[1] = np.sqrt(tot)
        meas[2] = tot

    def rank_update(self, R, N, W, C, k0, nb, e):
        r = R.numpy()
        for j in range(nb):
            r -= W.numpy()[:, k0 + j][:, None, None] * C.numpy()[k0 + j][None]
        self.vertex_energy(R, N, out=e)

    # dense contractions ---------------------------------------------------------------------------
    def gemm_nt(self, A, ra, B, rb, M, symmetric=False, out=None):
        D = torch.zeros((ra, rb), dtype=F64) if out is None else out
        D.numpy()[:ra, :rb] = A.numpy()[:ra, :M] @ B.numpy()[:rb, :M].T
        return D

    def gemm_nn(self, A, r, q, B, M, alpha, out, row_scale=None):
        res = alpha * (A.numpy()[:r, :q] @ B.numpy()[:q, :M])
        if row_scale is not None:
            res = res * row_scale.numpy()[:r, None]
        out.numpy()[:r, :M] = res
        return out

    def admm_iter(self, Hr, B0, Z, U, Lambda, K, N, inv_rho):
        ld = Z.shape[2]
        z, u = Z.numpy(), U.numpy()
        Cc = B0.numpy().reshape(K, -1) + Hr.numpy()[:K, :K] @ (z - u).reshape(K, -1)
        V = Cc.reshape(K, 3, ld) + u
        xlen = np.sqrt((V[:, 0] ** 2 + V[:, 1] ** 2) + V[:, 2] ** 2)
        with np.errstate(divide="ignore", invalid="ignore"):
            shrink = np.maximum(0.0, 1 - inv_rho * Lambda.numpy() / xlen)
        Zn = V * shrink[:, None, :]
        Un = V - Zn
        Zn[:, :, N:] = 0
        Un[:, :, N:] = 0
        z[:] = Zn
        u[:] = Un

    def admm_solve(self, Hr, B0, Z, U, Lambda, K, N, inv_rho, iters):
        for _ in range(int(iters)):
            self.admm_iter(Hr, B0, Z, U, Lambda, K, N, inv_rho)

    # small dense ----------------------------------------------------------------------------------
    def gram_small(self, W, F, K, diag_add, G):
        w = W.numpy()[:F, :K]
        G.numpy()[:K, :K] = w.T @ w + diag_add * np.eye(K)

    def spd_inverse(self, H, K, Hinv, info):
        h = H.numpy()[:K, :K]
        try:
            L = np.linalg.cholesky(h)
            Li = np.linalg.inv(L)
            Hinv.numpy()[:K, :K] = Li.T @ Li
            info[0] = 0
        except np.linalg.LinAlgError:
            info[0] = 1

    def small_gemm(self, transA, transB, m, n, k, alpha, A, B, D):
        a = A.numpy().T[:m, :k] if transA else A.numpy()[:m, :k]
        b = B.numpy().T[:k, :n] if transB else B.numpy()[:k, :n]
        D.numpy()[:m, :n] = alpha * (a @ b)

    def weights_sweep(self, P, Q, W, F, K):
        p, q, w = P.numpy(), Q.numpy(), W.numpy()
        weff = w[:F, :K].copy()
        for k in range(K):
            qkk = q[k, k]
            if qkk <= 1.e-8:
                w[:F, k] = 0
                continue
            opt = (p[:F, k] - weff @ q[k, :K] + weff[:, k] * qkk) / qkk
            x = np.maximum(0.0, opt)
            m = x.max()
            if m != 0:
                x = x / m
            weff[:, k] = x
            w[:F, k] = x

    def component_centers(self, C, Lambda, K, N, index_offset, val, idx, sp):
        if N == 0:
            val.numpy()[:K] = -np.inf
            idx.numpy()[:K] = np.iinfo(np.int64).max
            if Lambda is not None:
                sp.numpy()[:K] = 0.0
            return
        c = C.numpy()[:K, :, :N]
        v = (c[:, 0] ** 2 + c[:, 1] ** 2) + c[:, 2] ** 2
        ii = np.argmax(v, axis=1)
        val.numpy()[:K] = v[np.arange(K), ii]
        idx.numpy()[:K] = ii + index_offset
        if Lambda is not None:
            sp.numpy()[:K] = (Lambda.numpy()[:K, :N] * np.sqrt(v)).sum(axis=1)

    def syevd(self, A, n):
        a = A.numpy()[:n, :n]
        w, V = np.linalg.eigh(a)
        A.numpy()[:n, :n] = V.T
        return torch.from_numpy(w.copy()), torch.zeros((1,), dtype=torch.int32)

    def cgs_dots(self, T, k, d, N, h):
        t = T.numpy()
        h.numpy()[:k] = t[:k, d, :N] @ t[k, d, :N]

    def cgs_project(self, T, k, d, N, h, nrm2):
        t = T.numpy()
        if k > 0:
            t[k, d, :N] -= h.numpy()[:k] @ t[:k, d, :N]
        nrm2[0] = float(t[k, d, :N] @ t[k, d, :N])

    def row_scale(self, T, k, d, nrm2):
        T.numpy()[k, d] /= np.sqrt(float(nrm2[0]))

    # DEIM -----------------------------------------------------------------------------------------
    def deim_gather(self, C, N, row0, Pt, k, Vp):
        c, vp = C.numpy(), Vp.numpy()
        for j in range(k):
            row = int(Pt[j]) - row0
            if 0 <= row < N:
                vp[:, j, :k + 1] = c[:k + 1, :, row].T

    def deim_gather_rhs(self, C, N, row0, Pt, k, rhs, Vp):
        c, vp = C.numpy(), Vp.numpy()
        for j in range(k):
            row = int(Pt[j]) - row0
            if 0 <= row < N:
                vp[:, j, :k] = c[:k, :, row].T
                vp[:, j, k] = c[rhs, :, row]

    def deim_residual_block(self, C, N, k, target, accumulate, a, e, maxabs):
        c = C.numpy()
        if k == 0:
            r = c[target, :, :N].copy()
        else:
            r = np.einsum("dl,ldn->dn", a.numpy()[:, :k], c[:k, :, :N]) - c[target, :, :N]
        en = (r[0] ** 2 + r[1] ** 2) + r[2] ** 2
        if accumulate:
            e.numpy()[:N] += en
            maxabs[0] = max(float(maxabs[0]), float(np.abs(r).max()))
        else:
            e.numpy()[:] = 0
            e.numpy()[:N] = en
            maxabs[0] = float(np.abs(r).max())

    def deim_solve(self, Vp, k, a, info):
        vp = Vp.numpy()
        for d in range(3):
            try:
                a.numpy()[d, :k] = np.linalg.solve(vp[d, :k, :k], vp[d, :k, k])
            except np.linalg.LinAlgError:          # the kernel flags a zero pivot and leaves a_d as it was
                info[0] = max(int(info[0]), 1)

    def deim_residual(self, C, N, k, a, e, maxabs):
        c = C.numpy()
        if k == 0:
            r = c[0, :, :N].copy()
        else:
            r = np.einsum("dl,ldn->dn", a.numpy()[:, :k], c[:k, :, :N]) - c[k, :, :N]
        e.numpy()[:] = 0
        e.numpy()[:N] = (r[0] ** 2 + r[1] ** 2) + r[2] ** 2
        maxabs[0] = float(np.abs(r).max())

    # geodesics ------------------------------------------------------------------------------------
    def spmv(self, csr, x, y):
        from scipy import sparse
        rowptr, colidx, vals = (t.numpy() for t in csr)
        n = len(rowptr) - 1
        A = sparse.csr_matrix((vals, colidx, rowptr), shape=(n, n))
        y.numpy()[:, :n] = (A @ x.numpy()[:, :n].T).T

    def batched_dot(self, x, y, n, out):
        out.numpy()[:] = (x.numpy()[:, :n] * y.numpy()[:, :n]).sum(axis=1)

    def pcg_update_xr(self, x, r, z, p, Ap, dinv, rz, pAp, n):
        with np.errstate(all="ignore"):
            alpha = np.where(pAp.numpy() > 0, rz.numpy() / pAp.numpy(), 0.0)[:, None]
        x.numpy()[:, :n] += alpha * p.numpy()[:, :n]
        r.numpy()[:, :n] -= alpha * Ap.numpy()[:, :n]
        if z is not None:
            z.numpy()[:, :n] = r.numpy()[:, :n] * (dinv.numpy()[None, :n] if dinv is not None else 1.0)

    def pcg_update_p(self, p, z, rz_new, rz_old, n):
        with np.errstate(all="ignore"):
            beta = np.where(rz_old.numpy() > 0, rz_new.numpy() / rz_old.numpy(), 0.0)[:, None]
        p.numpy()[:, :n] = z.numpy()[:, :n] + beta * p.numpy()[:, :n]

    def axpby(self, y, alpha, a, row_scale, beta, c, n):
        v = alpha * a.numpy()[:, :n]
        if row_scale is not None:
            v = v * row_scale.numpy()[None, :n]
        if c is not None:
            v = v + beta * c.numpy()[:, :n]
        y.numpy()[:, :n] = v

    def remove_mean(self, x, n):
        x.numpy()[:, :n] -= x.numpy()[:, :n].mean(axis=1, keepdims=True)

    def heat_divergence(self, u, tri, gconst, dconst, vc_ptr, vc_idx, n, Xt, div):
        uu, t, g = u.numpy(), tri.numpy(), gconst.numpy()
        T = t.shape[0]
        grad = (g[None, :, 0:3] * uu[:, t[:, 2]][:, :, None] + g[None, :, 3:6] * uu[:, t[:, 0]][:, :, None]
                + g[None, :, 6:9] * uu[:, t[:, 1]][:, :, None])
        X = -grad / np.sqrt((grad ** 2).sum(axis=2))[:, :, None]
        Xt.numpy()[:] = X
        dc = dconst.numpy().reshape(T, 3, 8)
        out = np.zeros((uu.shape[0], n))
        for c in range(3):
            contrib = 0.5 * (dc[None, :, c, 0] * (dc[None, :, c, 1:4] * X).sum(axis=2)
                             + dc[None, :, c, 4] * (dc[None, :, c, 5:8] * X).sum(axis=2))
            for b in range(uu.shape[0]):
                out[b] += np.bincount(t[:, c], contrib[b], minlength=n)
        div.numpy()[:, :n] = out

    def support_map(self, phi, n, dmin, dmax, mode, lam, row0, nloc, out):
        p = phi.numpy()[:, :n]
        p = p - p.min(axis=1, keepdims=True)
        p = p[:, row0:row0 + nloc]
        if mode == 2:
            v = p
        else:
            m = (np.clip(p, dmin, dmax) - dmin) / (dmax - dmin)
            v = 1.0 - m if mode == 0 else lam * m
        out.numpy()[:] = 0
        out.numpy()[:, :nloc] = v

    def delta_rhs(self, idx, n, rhs):
        rhs.numpy()[:] = 0
        for b in range(rhs.shape[0]):
            rhs.numpy()[b, int(idx[b])] = 1.0

    # multifrontal solve (animsnapbases_b200/mfsolve.py) ---------------------------------------------
    def gather_rows(self, src, idx, out, n):
        out.numpy()[:, :n] = src.numpy()[:, idx.numpy()[:n]]

    def mf_forward(self, desc, vals, src, slots, work, rhs, y, upd):
        D, V, S = desc.numpy(), vals.numpy(), src.numpy()
        r, yy, uu = rhs.numpy(), y.numpy(), upd.numpy()
        for t in np.unique(work.numpy()[:, 0]):
            s0, n1, n2, mo, so, uo, ldm, row2 = (int(v) for v in D[t])
            f = np.zeros((r.shape[0], n1 + n2))
            f[:, :n1] = r[:, s0:s0 + n1]
            for ci in range(2):
                sl = S[ci, so:so + n1 + n2]
                ok = sl >= 0
                f[:, ok] += uu[:, sl[ok]]
            M = V[mo: mo + (row2 + n2) * ldm].reshape(row2 + n2, ldm)
            yy[:, s0:s0 + n1] = f[:, :n1] @ M[:n1, :n1].T
            if n2:
                uu[:, uo:uo + n2] = f[:, n1:] + f[:, :n1] @ M[row2:row2 + n2, :n1].T

    def mf_backward(self, desc, vals, bnd, work, y, x):
        D, V, Bd = desc.numpy(), vals.numpy(), bnd.numpy()
        yy, xx = y.numpy(), x.numpy()
        for t in np.unique(work.numpy()[:, 0]):
            s0, n1, n2, mo, so, uo, ldm, row2 = (int(v) for v in D[t])
            M = V[mo: mo + (row2 + n2) * ldm].reshape(row2 + n2, ldm)
            acc = yy[:, s0:s0 + n1] @ M[:n1, :n1]
            if n2:
                acc = acc + xx[:, Bd[uo:uo + n2]] @ M[row2:row2 + n2, :n1]
            xx[:, s0:s0 + n1] = acc

    # alignment ------------------------------------------------------------------------------------
    def pd_project(self, kind, q, elems, rest, sigma_min, sigma_max, out):
        """numpy statement of asb_pd_project through LAPACK's SVD (the kernel uses Jacobi eigen-pairs of F^T F instead)."""
        el, rs = elems.numpy(), rest.numpy()
        e = len(el)
        for f in range(q.shape[0]):
            x = q.numpy()[f].T                                            # (|V| padded, 3)
            if kind in (0, 1):
                Ds = np.stack([x[el[:, c]] - x[el[:, 3]] for c in range(3)], axis=2)
                Fm = Ds @ rs.reshape(e, 3, 3)
                U, s, Vt = np.linalg.svd(Fm)
                if kind == 0:
                    R = U @ Vt
                    R[np.linalg.det(R) < 0, :, 2] *= -1
                    rows = np.swapaxes(R, 1, 2)
                else:
                    s = np.clip(s, sigma_min, sigma_max)
                    s[np.linalg.det(Fm) < 0, 2] *= -1
                    rows = (U * s[:, None, :]) @ Vt
                rows = rows.reshape(3 * e, 3)
            elif kind == 2:
                P, Di = rs[:, :6].reshape(e, 3, 2), rs[:, 6:].reshape(e, 2, 2)
                Ds = np.stack([x[el[:, 1]] - x[el[:, 0]], x[el[:, 2]] - x[el[:, 0]]], axis=2)
                U, s, Vt = np.linalg.svd(np.swapaxes(P, 1, 2) @ Ds @ Di)
                Fh = (U * np.clip(s, sigma_min, sigma_max)[:, None, :]) @ Vt
                rows = np.swapaxes(P @ Fh, 1, 2).reshape(2 * e, 3)
            else:
                sp = x[el[:, 1]] - x[el[:, 0]]
                ln = np.linalg.norm(sp, axis=1)
                rows = 0.5 * rs[:, 0:1] * sp / np.where(ln == 0, 1.0, ln)[:, None]
                rows[ln == 0] = 0.0
            out.numpy()[f, :, :rows.shape[0]] = rows.T
        return out

    def pd_vertex_bending(self, q, verts, ptr, nbr, w, rest, prevent_flips, out):
        vs, pt, nb, ww, rs = (t.numpy() for t in (verts, ptr, nbr, w, rest))
        seg = np.repeat(np.arange(len(vs)), np.diff(pt))
        for f in range(q.shape[0]):
            x = q.numpy()[f].T
            star = np.zeros((len(vs), 3))
            np.add.at(star, seg, (x[vs[seg]] - x[nb]) * ww[:, None])
            norm = np.linalg.norm(star, axis=1)
            small = norm < 1e-10
            c = np.where(small[:, None], rs[:, 1:4] * rs[:, 0:1], star * (rs[:, 0] / np.where(small, 1.0, norm))[:, None])
            if prevent_flips:
                flip = (norm > 1e-5) & ((rs[:, 1:4] * c).sum(axis=1) * rs[:, 4] < 0)
                c[flip] = -c[flip]
            out.numpy()[f, :, :len(vs)] = c.T
        return out

    def align(self, frames, rigid):
        v = frames.numpy().astype(np.float64)
        F = v.shape[0]
        out = np.empty_like(frames.numpy())
        Ts = np.zeros((F, 12))
        t1 = v[0].mean(0)
        for f in range(F):
            t0 = v[f].mean(0)
            M = (v[0] - t1).T @ (v[f] - t0)
            U, _, Vt = np.linalg.svd(M)
            Rm = U @ Vt
            if np.linalg.det(Rm) < 0:
                Rm *= -1
            A = Rm if rigid else np.eye(3)
            t = t1 - Rm @ t0
            out[f] = (v[f] @ A.T + t).astype(np.float32)
            Ts[f] = np.concatenate([A, t[:, None]], axis=1).ravel()
        return torch.from_numpy(out), torch.from_numpy(Ts)
