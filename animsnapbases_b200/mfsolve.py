"""Sparse direct solver for the two heat-method systems: nested-dissection multifrontal Cholesky whose *solve* phase is
a short sequence of dense, coalesced GPU passes (one launch per tree level and direction).

Why a direct solver: the reference factorises `A - t Lc` and `Lc` once with SuperLU (utils/support.py:169-171) and every
geodesic query is two pairs of triangular solves (:182, :205).  The heat solution spans hundreds of orders of magnitude
(u ~ exp(-d / sqrt(t))) and only its *gradient direction* is used (:188-190), so the far field needs componentwise
relative accuracy -- which elimination on these (near) M-matrices delivers and a Krylov method does not.

Set-up (once per mesh; geometry only, like the reference's `__init__`):
  1. nested dissection of the mesh graph by coordinate bisection -> binary tree of fronts; every front owns a set of
     unknowns (a leaf interior or a separator) and sees a boundary made of ancestor-separator unknowns   [host, index work];
  2. symbolic factorisation, one vectorised pass per tree level                                            [host, index work];
  3. numeric multifrontal Cholesky, bottom-up, fronts of one level batched by size class:
        F = [F11 F21^T; F21 F22]  ->  L11 = chol(F11),  L21 = F21 L11^-T,  update U = F22 - L21 L21^T
     with batched dense device routines (torch.linalg on the engine's device: set-up plumbing, not the hot path);
  4. per front the dense *solve operator*  Mt = [ L11^-1 ; -L21 L11^-1 ]  ((n1 + n2) x n1, row-major) is kept, so that
        forward :  [y1 ; upd] = [0 ; f2] + Mt f1        (f = rhs on own rows + children's update vectors)
        backward:  x1 = Mt^T [y1 ; x2]                   (x2 = already-solved ancestor unknowns)
     are dense mat-vecs with no sequential dependence inside a front: asb_mf_forward / asb_mf_backward, one launch per
     level.  For these matrices L11^-1 >= 0 and -L21 L11^-1 >= 0, so folding the triangular solves into explicit
     inverses introduces no cancellation.
The singular Poisson system is grounded at one vertex (the solution is defined up to a constant, which
`phi -= phi.min()`, utils/support.py:206, removes anyway).
"""
from __future__ import annotations

import sys

import numpy as np
import torch
from scipy import sparse

ROWS_PER_BLOCK = 32     # forward: rows of Mt per CTA;  backward: columns of Mt per CTA  (must match mf_solve.cu)


def _row_entries(indptr, rows):
    """(row position, flat CSR position) of every stored entry of the CSR rows `rows`."""
    cnt = indptr[rows + 1] - indptr[rows]
    tot = int(cnt.sum())
    if tot == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    rep = np.repeat(np.arange(len(rows)), cnt)
    flat = np.repeat(indptr[rows], cnt) + (np.arange(tot) - np.repeat(np.cumsum(cnt) - cnt, cnt))
    return rep, flat


def _ranges(starts, lengths):
    """Concatenation of arange(starts[i], starts[i] + lengths[i]) and the index i of every element."""
    tot = int(lengths.sum())
    who = np.repeat(np.arange(len(starts)), lengths)
    if tot == 0:
        return np.zeros(0, np.int64), who
    within = np.arange(tot) - np.repeat(np.cumsum(lengths) - lengths, lengths)
    return np.repeat(starts, lengths) + within, who


def _size_class(x):
    """Padded dimension of a front block: multiples of 8 up to 64, then four steps per octave (<= 25 % padding)."""
    x = np.asarray(x, dtype=np.int64)
    small = (x + 7) // 8 * 8
    lg = np.floor(np.log2(np.maximum(x, 1))).astype(np.int64)
    step = np.left_shift(1, np.maximum(lg - 2, 3))
    big = (x + step - 1) // step * step
    return np.where(x <= 64, small, big)


def nested_dissection(indptr, indices, coords, leaf_size=64):
    """Binary dissection tree by coordinate bisection.  Returns (own lists, children (T,2), depth (T,)) in post-order:
    children precede parents, the root is last."""
    n = len(indptr) - 1
    lab = np.zeros(n, dtype=np.int8)
    owns, children, depths = [], [], []

    def leaf(S, depth):
        owns.append(S)
        children.append((-1, -1))
        depths.append(depth)
        return len(owns) - 1

    def rec(S, depth):
        if len(S) <= leaf_size:
            return leaf(S, depth)
        c = coords[S]
        axis = int(np.argmax(c.max(axis=0) - c.min(axis=0)))
        half = len(S) // 2
        part = np.argpartition(c[:, axis], half)
        A, B = S[part[:half]], S[part[half:]]
        lab[A], lab[B] = 1, 2
        rep, flat = _row_entries(indptr, A)
        sep_mask = np.zeros(len(A), dtype=bool)
        sep_mask[rep[lab[indices[flat]] == 2]] = True
        lab[S] = 0
        if not sep_mask.any():
            sep_mask[0] = True                      # disconnected halves: any vertex is a valid (super)separator
        sep, A2 = A[sep_mask], A[~sep_mask]
        if len(A2) == 0 or len(B) == 0:
            return leaf(S, depth)
        ca = rec(A2, depth + 1)
        cb = rec(B, depth + 1)
        owns.append(sep)
        children.append((ca, cb))
        depths.append(depth)
        return len(owns) - 1

    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 10000))
    try:
        rec(np.arange(n, dtype=np.int64), 0)
    finally:
        sys.setrecursionlimit(old)
    return owns, np.asarray(children, dtype=np.int64).reshape(-1, 2), np.asarray(depths, dtype=np.int64)


class SymbolicFactor:
    """Ordering, fronts, boundaries, size-class buckets and every index map of the factorisation (host, integers only)."""

    def __init__(self, A, coords, leaf_size=64):
        A = sparse.csr_matrix(A, dtype=np.float64)
        A.sum_duplicates()
        A.sort_indices()
        n = self.n = A.shape[0]
        pattern = A.copy()
        pattern.setdiag(0)
        pattern.eliminate_zeros()
        owns, ch, depth = nested_dissection(pattern.indptr.astype(np.int64), pattern.indices.astype(np.int64),
                                            np.asarray(coords, dtype=np.float64), leaf_size)
        T = self.T = len(owns)
        n1 = np.array([len(o) for o in owns], dtype=np.int64)
        start = np.cumsum(n1) - n1
        end = start + n1
        perm = np.concatenate(owns)
        order = np.empty(n, dtype=np.int64)
        order[perm] = np.arange(n)
        Ap = A[perm][:, perm].tocsr()
        Ap.sort_indices()
        ip, ix, av = Ap.indptr.astype(np.int64), Ap.indices.astype(np.int64), Ap.data
        self.perm, self.order, self.children, self.depth = perm, order, ch, depth
        self.n1, self.start = n1, start
        n2 = np.zeros(T, dtype=np.int64)
        lvl_idx = np.zeros(T, dtype=np.int64)
        dmax = int(depth.max())
        levels = [None] * (dmax + 1)
        for d in range(dmax, -1, -1):
            ft = np.nonzero(depth == d)[0]
            m = len(ft)
            lvl_idx[ft] = np.arange(m)
            st, en = start[ft], end[ft]
            rows, fr = _ranges(st, n1[ft])
            rep, flat = _row_entries(ip, rows)
            nb, val, fe = ix[flat], av[flat], fr[rep]
            lrow = rows[rep] - st[fe]
            keep = nb >= st[fe]
            nb, val, fe, lrow = nb[keep], val[keep], fe[keep], lrow[keep]
            is_b = nb >= en[fe]
            cand = [fe[is_b] * n + nb[is_b]]
            kids = []
            for ci in range(2):
                c = ch[ft, ci]
                has = np.nonzero(c >= 0)[0]
                if len(has) == 0 or d == dmax:
                    kids.append(None)
                    continue
                lo = levels[d + 1]
                cidx = lvl_idx[c[has]]
                pos, who = _ranges(lo["bptr"][cidx], lo["bptr"][cidx + 1] - lo["bptr"][cidx])
                cb = lo["bnd"][pos]
                pi = has[who]
                up = cb >= en[pi]
                cand.append(pi[up] * n + cb[up])
                kids.append({"child": c[has], "parent": has, "cb": cb, "pi": pi})
            keys = np.unique(np.concatenate(cand))
            bf = keys // n
            bptr = np.zeros(m + 1, dtype=np.int64)
            np.cumsum(np.bincount(bf, minlength=m), out=bptr[1:])
            n2[ft] = bptr[1:] - bptr[:-1]
            e_pos = np.where(is_b, np.searchsorted(keys, fe * n + nb) - bptr[fe], -1)
            e_col = np.where(is_b, 0, nb - st[fe])
            for kd in kids:
                if kd is None:
                    continue
                cb, pi = kd.pop("cb"), kd["pi"]
                inown = cb < en[pi]
                kd["inown"] = inown
                kd["rel"] = np.where(inown, cb - st[pi], np.searchsorted(keys, pi * n + cb) - bptr[pi])
            levels[d] = {"ft": ft, "bnd": keys % n, "bptr": bptr, "e_front": fe, "e_row": lrow, "e_col": e_col,
                         "e_pos": e_pos, "e_val": val, "kids": kids}
        self.n2, self.levels, self.lvl_idx = n2, levels, lvl_idx
        nn = n1 + n2
        self.slot_off = np.cumsum(nn) - nn
        self.upd_off = np.cumsum(n2) - n2
        self.slots, self.upd_len = int(nn.sum()), int(n2.sum())
        # ---- size-class buckets per level, operator storage offsets
        self.n1p = _size_class(n1)
        self.n2p = np.where(n2 > 0, _size_class(n2), 0)
        self.mt_off = np.zeros(T, dtype=np.int64)
        self.bucket = np.zeros(T, dtype=np.int64)      # bucket id within the level
        self.bpos = np.zeros(T, dtype=np.int64)        # batch position within the bucket
        off = 0
        for d, lv in enumerate(levels):
            ft = lv["ft"]
            key = self.n1p[ft] * (1 << 32) + self.n2p[ft]
            uk, inv = np.unique(key, return_inverse=True)
            lv["buckets"] = []
            for b in range(len(uk)):
                mem = np.nonzero(inv == b)[0]
                a, c = int(self.n1p[ft[mem[0]]]), int(self.n2p[ft[mem[0]]])
                lv["buckets"].append({"members": mem, "n1p": a, "n2p": c, "base": off})
                self.bucket[ft[mem]] = b
                self.bpos[ft[mem]] = np.arange(len(mem))
                self.mt_off[ft[mem]] = off + np.arange(len(mem)) * ((a + c) * a)
                off += len(mem) * (a + c) * a
        self.vals_len = off
        # ---- compact device maps: boundary lists, children -> parent slot gather maps, per-level work lists
        bnd = np.zeros(max(self.upd_len, 1), dtype=np.int32)
        src = np.full((2, max(self.slots, 1)), -1, dtype=np.int32)
        for d, lv in enumerate(levels):
            ft = lv["ft"]
            pos, who = _ranges(self.upd_off[ft], n2[ft])
            bnd[pos] = lv["bnd"]
            for ci, kd in enumerate(lv["kids"]):
                if kd is None:
                    continue
                p = ft[kd["pi"]]
                slot = np.where(kd["inown"], kd["rel"], n1[p] + kd["rel"])
                cpos, _ = _ranges(self.upd_off[kd["child"]], n2[kd["child"]])
                src[ci, self.slot_off[p] + slot] = cpos
        self.bnd, self.src = bnd, src
        self.work_fwd, self.work_bwd = [], []
        for lv in levels:
            ft = lv["ft"]
            for arr, extent in ((self.work_fwd, nn[ft]), (self.work_bwd, n1[ft])):
                nblk = (extent + ROWS_PER_BLOCK - 1) // ROWS_PER_BLOCK
                r0, who = _ranges(np.zeros(len(ft), dtype=np.int64), nblk)
                arr.append(np.stack([ft[who], r0 * ROWS_PER_BLOCK], axis=1).astype(np.int32))
        # front descriptor rows: own_start, n1, n2, mt_off, slot_off, upd_off, ldm, row2
        self.desc = np.stack([start, n1, n2, self.mt_off, self.slot_off, self.upd_off, self.n1p, self.n1p], axis=1).copy()


def numeric_factor(sym: SymbolicFactor, device):
    """Bottom-up multifrontal Cholesky with the fronts of every (level, size class) bucket batched.  Returns the flat
    operator buffer `vals` (device, float64) addressed by SymbolicFactor.desc."""
    dev = torch.device(device)
    vals = torch.zeros(max(sym.vals_len, 1), dtype=torch.float64, device=dev)
    n1_all, n2_all = sym.n1, sym.n2

    def t(a, dtype=torch.int64):
        return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(dev)

    upd_below = None                                  # per bucket of the level below: U (Bc, n2p, n2p)
    for d in range(len(sym.levels) - 1, -1, -1):
        lv = sym.levels[d]
        ft = lv["ft"]
        bk_of, pos_of = sym.bucket[ft], sym.bpos[ft]
        upd_here = []
        for b, bk in enumerate(lv["buckets"]):
            mem, a, c = bk["members"], bk["n1p"], bk["n2p"]
            nm, Bt = a + c, len(bk["members"])
            Fm = torch.zeros((Bt, nm, nm), dtype=torch.float64, device=dev)
            flatF = Fm.view(-1)
            # ---- original matrix entries of the own rows (F11 in full, F21 and its mirror)
            sel = np.nonzero(bk_of[lv["e_front"]] == b)[0]
            fe, row = lv["e_front"][sel], lv["e_row"][sel]
            col = np.where(lv["e_pos"][sel] >= 0, a + lv["e_pos"][sel], lv["e_col"][sel])
            base = pos_of[fe] * (nm * nm)
            isb = lv["e_pos"][sel] >= 0
            idx = np.concatenate([base + row * nm + col, (base + col * nm + row)[isb]])
            v = np.concatenate([lv["e_val"][sel], lv["e_val"][sel][isb]])
            flatF[t(idx)] = t(v, torch.float64)
            # ---- identity on the padded own slots
            pr, who = _ranges(n1_all[ft[mem]], a - n1_all[ft[mem]])
            if len(pr):
                flatF[t(who * (nm * nm) + pr * (nm + 1))] = 1.0
            # ---- extend-add of the children's update matrices (first children, then second: fixed order)
            for ci, kd in enumerate(lv["kids"]):
                if kd is None:
                    continue
                child = kd["child"]
                par = kd["parent"]
                mine = np.nonzero((bk_of[par] == b) & (n2_all[child] > 0))[0]
                if len(mine) == 0:
                    continue
                lo = sym.levels[d + 1]
                cb_of, cpos_of = sym.bucket[child[mine]], sym.bpos[child[mine]]
                eptr = np.concatenate([[0], np.cumsum(n2_all[child])])
                slot_all = np.where(kd["inown"], kd["rel"], a + kd["rel"])
                for cbk in np.unique(cb_of):
                    g = mine[cb_of == cbk]
                    c2p = lo["buckets"][cbk]["n2p"]
                    rel = np.zeros((len(g), c2p), dtype=np.int64)
                    pos, who = _ranges(eptr[g], n2_all[child[g]])
                    rel[who, pos - eptr[g][who]] = slot_all[pos]
                    relt = t(rel)
                    pb = t(pos_of[par[g]])
                    flat = pb[:, None, None] * (nm * nm) + relt[:, :, None] * nm + relt[:, None, :]
                    U = upd_below[cbk][t(sym.bpos[child[g]])]
                    flatF.index_add_(0, flat.reshape(-1), U.reshape(-1))
            # ---- partial Cholesky and the solve operator
            L11, info = torch.linalg.cholesky_ex(Fm[:, :a, :a])
            if int(info.max().item()) != 0:
                raise FloatingPointError("multifrontal Cholesky: front is not positive definite")
            eye = torch.eye(a, dtype=torch.float64, device=dev).expand(Bt, a, a)
            Linv = torch.linalg.solve_triangular(L11, eye, upper=False).tril_()
            out = vals[bk["base"]: bk["base"] + Bt * nm * a].view(Bt, nm, a)
            out[:, :a] = Linv
            if c:
                L21 = Fm[:, a:, :a] @ Linv.mT
                upd_here.append(Fm[:, a:, a:] - L21 @ L21.mT)
                out[:, a:] = -(L21 @ Linv)
            else:
                upd_here.append(None)
            del Fm, flatF
        upd_below = upd_here
    return vals


class MultifrontalSolver:
    """Device-resident factorisation of one SPD matrix; `solve` runs 2 x (tree depth + 1) kernel launches."""

    def __init__(self, ops, A, coords, leaf_size=64, ground=None):
        A = sparse.csr_matrix(A, dtype=np.float64)
        n = A.shape[0]
        if ground is not None:                      # decouple the grounded unknown: row/col -> identity
            keep = np.ones(n)
            keep[ground] = 0.0
            Dk = sparse.diags(keep)
            A = (Dk @ A @ Dk + sparse.diags(1.0 - keep)).tocsr()
        self.ops, self.n, self.ground = ops, n, ground
        sym = SymbolicFactor(A, coords, leaf_size)
        self.vals = numeric_factor(sym, ops.device)
        self.ldx = (n + 31) // 32 * 32
        self.desc = ops.to_device(sym.desc)
        self.bnd = ops.to_device(sym.bnd)
        self.src = ops.to_device(sym.src)
        self.perm = ops.to_device(sym.perm.astype(np.int32))
        self.order = ops.to_device(sym.order.astype(np.int32))
        self.order_host = sym.order
        self.work_fwd = [ops.to_device(w) for w in sym.work_fwd]
        self.work_bwd = [ops.to_device(w) for w in sym.work_bwd]
        # host table for asb_mf_solve (root level first): work-list pointers and lengths of every level, and the largest n1 /
        # n1 + n2 of its fronts (the kernels pick their staging size and the small-front variants by them)
        self.levels_host = None
        if hasattr(ops, "mf_solve"):
            nn_all = sym.n1 + sym.n2
            self.levels_host = np.array([[wf.data_ptr(), wf.shape[0], wb.data_ptr(), wb.shape[0],
                                          int(sym.n1[sym.depth == d].max()), int(nn_all[sym.depth == d].max())]
                                         for d, (wf, wb) in enumerate(zip(self.work_fwd, self.work_bwd))], dtype=np.int64)
        # tree tables for asb_mf_solve_delta: parent / depth / first elimination position of every front, and the number
        # of 32-row blocks of the largest front of every level
        self.tree = None
        if hasattr(ops, "mf_solve_delta"):
            parent = np.full(sym.T, -1, dtype=np.int32)
            for c in range(2):
                has = sym.children[:, c] >= 0
                parent[sym.children[has, c]] = np.nonzero(has)[0]
            nn = sym.n1 + sym.n2
            nlev = len(sym.levels)
            self.maxblk_host = np.array([int(((nn[sym.depth == d] + ROWS_PER_BLOCK - 1) // ROWS_PER_BLOCK).max())
                                         for d in range(nlev)], dtype=np.int64)
            self.tree = (ops.to_device(sym.start.astype(np.int64)), ops.to_device(parent),
                         ops.to_device(sym.depth.astype(np.int32)))
        self.upd_len = (sym.upd_len + 31) // 32 * 32 + 32
        self.slots = sym.src.shape[1]
        self.stats = {"fronts": sym.T, "levels": len(sym.levels), "operator_bytes": int(sym.vals_len) * 8}

    def solve_permuted(self, rhs_p):
        """rhs_p: (B, ldx) in elimination order -> x in elimination order."""
        ops = self.ops
        B = rhs_p.shape[0]
        y = ops.empty((B, self.ldx))
        x = ops.empty((B, self.ldx))
        upd = ops.empty((B, self.upd_len))
        if self.levels_host is not None:
            ops.mf_solve(self.desc, self.vals, self.src, self.slots, self.bnd, self.levels_host, rhs_p, y, x, upd)
            return x
        for w in reversed(self.work_fwd):
            ops.mf_forward(self.desc, self.vals, self.src, self.slots, w, rhs_p, y, upd)
        for w in self.work_bwd:
            ops.mf_backward(self.desc, self.vals, self.bnd, w, y, x)
        return x

    def solve_delta(self, idx_dev):
        """A^-1 e_i for the B source vertices idx_dev (int64, device), in vertex order: (B, ldx).  The forward pass only
        walks the tree paths above the sources (asb_mf_solve_delta); the grounded variant is not needed (heat system)."""
        ops = self.ops
        if self.tree is None or self.ground is not None:
            rhs = ops.empty((idx_dev.numel(), self.ldx))
            ops.delta_rhs(idx_dev, self.n, rhs)
            return self.solve(rhs)
        B = idx_dev.numel()
        fstart, parent, depth = self.tree
        rp, y, xp = ops.empty((B, self.ldx)), ops.empty((B, self.ldx)), ops.empty((B, self.ldx))
        upd = ops.empty((B, self.upd_len))
        pathfront = torch.empty((B, len(self.work_fwd)), dtype=torch.int32, device=rp.device)
        ops.mf_solve_delta(self.desc, self.vals, self.src, self.slots, self.bnd, self.levels_host, self.maxblk_host, fstart, parent,
                           depth, self.order, idx_dev, pathfront, rp, y, xp, upd)
        out = ops.zeros((B, self.ldx))
        ops.gather_rows(xp, self.order, out, self.n)
        return out

    def solve(self, rhs):
        """rhs: (B, ldx) in vertex order -> A^-1 rhs in vertex order (grounded unknown = 0)."""
        ops = self.ops
        rp = ops.empty(rhs.shape)
        ops.gather_rows(rhs, self.perm, rp, self.n)
        if self.ground is not None:
            rp[:, int(self.order_host[self.ground])] = 0.0
        xp = self.solve_permuted(rp)
        out = ops.zeros(rhs.shape)
        ops.gather_rows(xp, self.order, out, self.n)
        return out
