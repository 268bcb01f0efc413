"""Device operator set of the basis engine: one Python method per C-ABI entry point of libasb.so.

PyTorch is used for device memory, streams and (in engine.py) torch.distributed only; every arithmetic
step of the hot path is a hand-written sm_100a kernel behind include/asb.h.  The class refuses to
construct without a CUDA device and the built library -- there is no CPU path in the product.
(tests/numpy_ops.py holds a numpy stand-in with the same interface, used only to exercise the host-side
orchestration and the world_size-2 gloo path on machines without a GPU.)
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

F64 = torch.float64
I64 = torch.int64
I32 = torch.int32


def pad_ld(n: int) -> int:
    """Plane stride of the planar layout: next multiple of 32."""
    return (int(n) + 31) // 32 * 32


class CudaOps:
    name = "cuda"

    def __init__(self, device=None):
        if not torch.cuda.is_available():
            raise _lib.AsbError("animsnapbases_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        torch.cuda.set_device(self.device)
        sms, major, minor = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.asb_device_info(ctypes.byref(sms), ctypes.byref(major), ctypes.byref(minor)), "device_info")
        self.sm_count, self.cc = sms.value, (major.value, minor.value)
        if self.cc[0] < 10:
            raise _lib.AsbError(f"libasb.so is built for sm_100a only; found compute capability {self.cc}")
        self._ws = {}          # scratch per stream
        self._side = None      # side stream of fork() / join()
        self._ws2 = None
        self._admm_ws = None
        self.timers = None

    # ------------------------------------------------------------------ plumbing
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @staticmethod
    def _p(t):
        if t is None:
            return None
        return ctypes.c_void_p(t.data_ptr())

    def _call(self, name, *args):
        ev = self.timers.get(name) if self.timers is not None else None
        if ev is not None:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            _lib.check(getattr(self.lib, name)(*args, self._stream()), name)
            b.record()
            ev.append((a, b))
            return
        _lib.check(getattr(self.lib, name)(*args, self._stream()), name)

    def time_kernels(self, names):
        """Record CUDA events (on torch's current stream, the one every kernel is launched on) around each launch
        of the named C-ABI entry points; `kernel_times()` returns {name: (launches, total_ms)} after a sync."""
        self.timers = {n: [] for n in names} if names else None

    def kernel_times(self):
        self.synchronize()
        out = {}
        for n, evs in (self.timers or {}).items():
            out[n] = (len(evs), float(sum(a.elapsed_time(b) for a, b in evs)))
        return out

    def empty(self, shape, dtype=F64):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def zeros(self, shape, dtype=F64):
        return torch.zeros(shape, dtype=dtype, device=self.device)

    def to_device(self, a, dtype=None):
        t = torch.as_tensor(a)
        if dtype is not None:
            t = t.to(dtype)
        return t.to(self.device, non_blocking=True).contiguous()

    def to_host(self, t):
        """Device tensor -> host numpy array through pinned memory (asynchronous copy + stream synchronise)."""
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return h.numpy()

    def workspace(self, nbytes: int):
        """Scratch for the reductions of the CURRENT stream (one buffer per stream: the geodesic solves run on a side
        stream next to the streaming / DMMA kernels of the main one)."""
        key = torch.cuda.current_stream(self.device).cuda_stream
        ws = self._ws.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = self._ws[key] = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=self.device)
        return ws

    def fork(self, after=None, lane=0):
        """Context manager: run the enclosed launches on a side stream (`lane` 0 or 1: two independent side streams), ordered
        after everything issued so far on the current stream -- or, with `after` (a recorded CUDA event), only after that
        event; `join()` makes the current stream wait for them."""
        if self._side is None:
            self._side = {}
        st = self._side.get(lane)
        if st is None:
            import os
            prio = int(os.environ.get("ASB_SIDE_PRIORITY", "0"))     # -1: the side stream's CTAs are dispatched first
            st = self._side[lane] = torch.cuda.Stream(self.device, priority=prio)
        if after is not None:
            st.wait_event(after)
        else:
            st.wait_stream(torch.cuda.current_stream(self.device))
        return torch.cuda.stream(st)

    def record_event(self):
        """A CUDA event recorded on the current stream (ordering handle for `fork(after=...)`)."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        return ev

    def join(self):
        if self._side is not None:
            for st in self._side.values():
                torch.cuda.current_stream(self.device).wait_stream(st)

    def ws_for(self, rows: int, cols: int):
        return self.workspace(self.lib.asb_workspace_bytes(int(rows), int(cols)))

    def launch_count(self) -> int:
        return int(self.lib.asb_launch_count())

    def reset_launch_count(self):
        self.lib.asb_reset_launch_count()

    def add_launch_count(self, n: int):
        """Kernel nodes of a replayed CUDA graph (captured launches are counted once, at capture time)."""
        self.lib.asb_add_launch_count(int(n))

    def synchronize(self):
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ boundary
    def ingest(self, src, row_scale, ld):
        R, N, _ = src.shape
        if N == 0:                                        # empty row shard (more ranks than 32-row blocks)
            return self.zeros((R, 3, ld))
        if src.dtype not in (torch.float32, torch.float64):
            src = src.to(torch.float64)                   # the reference converts everything with .astype(float)
        src = src.contiguous()
        X = self.empty((R, 3, ld))
        fn = "asb_ingest_f32" if src.dtype == torch.float32 else "asb_ingest_f64"
        self._call(fn, self._p(src), self._p(row_scale), self._p(X), R, N, ld)
        return X

    def export_aos(self, T, row_scale, N):
        R, _, ld = T.shape
        out = self.empty((R, N, 3))
        if N == 0:
            return out
        self._call("asb_export_aos", self._p(T), self._p(row_scale), self._p(out), R, N, ld)
        return out

    def pack_bin(self, T, K, N):
        out = self.empty((3, K, N))
        if N == 0:
            return out
        self._call("asb_pack_bin", self._p(T), self._p(out), K, N, T.shape[2])
        return out

    # ------------------------------------------------------------------ centering
    def col_mean(self, X):
        F, _, ld = X.shape
        mean = self.empty((3, ld))
        self._call("asb_col_mean", self._p(X), self._p(mean), F, ld)
        return mean

    def center_stats(self, X, mean, N):
        F, _, ld = X.shape
        out = self.empty((2,))
        if N == 0:
            return out.zero_()
        self._call("asb_center_stats", self._p(X), self._p(mean), F, N, ld, self._p(out), self._p(self.ws_for(F, 3 * ld)))
        return out

    def center_apply(self, X, mean, scale, N):
        F, _, ld = X.shape
        if N == 0:
            return
        self._call("asb_center_apply", self._p(X), self._p(mean), float(scale), F, N, ld)

    def unstandardize(self, T, mean, scale, N):
        R, _, ld = T.shape
        if N == 0:
            return
        self._call("asb_unstandardize", self._p(T), self._p(mean), float(scale), R, N, ld)

    # ------------------------------------------------------------------ deflation
    def vertex_energy(self, R, N, out=None):
        F, _, ld = R.shape
        e = self.empty((ld,)) if out is None else out
        if N == 0:
            return e.zero_()
        self._call("asb_vertex_energy", self._p(R), self._p(e), F, N, ld, self._p(self.ws_for(F, 3 * ld)))
        return e

    def argmax_first(self, v, n, index_offset, out_val, out_idx):
        if n == 0:                                        # empty shard: never wins a pair reduction
            out_val.fill_(-float("inf"))
            out_idx.fill_(torch.iinfo(torch.int64).max)
            return
        self._call("asb_argmax_first", self._p(v), n, index_offset, self._p(out_val), self._p(out_idx),
                   self._p(self.ws_for(1, n)))

    def argmax_outside_ball(self, v, n, index_offset, pos, centre, radius, out_pair):
        """out_pair (2,) f64 = {largest v[i] whose vertex i + index_offset lies outside the ball, bits of its global index}."""
        if n == 0:                                        # empty shard: never wins
            out_pair[0] = -float("inf")
            out_pair[1:2].view(I64).fill_(torch.iinfo(torch.int64).max)
            return
        self._call("asb_argmax_outside_ball", self._p(v), n, index_offset, self._p(pos), self._p(centre), float(radius),
                   self._p(out_pair), self._p(self.ws_for(1, n)))

    def select_pair(self, all_pairs, world, fallback, out_idx):
        self._call("asb_select_pair", self._p(all_pairs), int(world), self._p(fallback), self._p(out_idx))

    def wait_event(self, ev):
        """Make the current stream wait for a recorded event (finer than join(): the side stream may keep running)."""
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)

    def group_sum(self, v, n, group, out):
        if n == 0:
            return
        self._call("asb_group_sum", self._p(v), n, group, self._p(out))

    def gather_traj(self, R, N, idx_dev, row0, W, k0, C, npend, traj):
        """W: (F, Kmax) weights, C: (Kmax, 3, ld) components; pending block = columns/rows [k0, k0+npend)."""
        F, _, ld = R.shape
        if N == 0:
            return
        Wp = ctypes.c_void_p(W.data_ptr() + 8 * k0) if npend else None
        Cp = ctypes.c_void_p(C.data_ptr() + 8 * k0 * 3 * ld) if npend else None
        self._call("asb_gather_traj", self._p(R), F, N, ld, self._p(idx_dev), row0, Wp, W.shape[1], Cp, npend,
                   self._p(traj))

    def gather_candidate(self, R, N, val, idx, row0, W, k0, C, npend, cand):
        """This rank's pick packed for one all-gather: cand = [value, index bits, trajectory (3 F)] (asb_gather_candidate)."""
        F, _, ld = R.shape
        Wp = ctypes.c_void_p(W.data_ptr() + 8 * k0) if npend else None
        Cp = ctypes.c_void_p(C.data_ptr() + 8 * k0 * 3 * ld) if npend else None
        self._call("asb_gather_candidate", self._p(R), F, N, ld, self._p(val), self._p(idx), row0, Wp, W.shape[1], Cp, npend,
                   self._p(cand))

    def select_candidate(self, cand_all, F, val, idx, traj):
        """Winner of the gathered candidates (largest value, ties to the lowest global index) and its trajectory."""
        world, stride = cand_all.shape
        self._call("asb_select_candidate", self._p(cand_all), world, stride, F, self._p(val), self._p(idx), self._p(traj))

    def traj_rank1(self, traj, F, local_support, W, k, k0, coef, scal):
        npend = k - k0
        ldw = W.shape[1]
        wout = ctypes.c_void_p(W.data_ptr() + 8 * k)
        Wp = ctypes.c_void_p(W.data_ptr() + 8 * k0) if npend else None
        self._call("asb_traj_rank1", self._p(traj), F, int(bool(local_support)), wout, ldw, Wp, npend,
                   self._p(coef) if npend else None, self._p(scal))

    def deflate_gemv(self, R, N, W, k, g):
        F, _, ld = R.shape
        if N == 0:
            return
        w = ctypes.c_void_p(W.data_ptr() + 8 * k)
        self._call("asb_deflate_gemv", self._p(R), F, N, ld, w, W.shape[1], self._p(g), self._p(self.ws_for(F, 3 * ld)))

    def deflate_finish(self, g, C, k, k0, coef, s, scal, e, N, meas):
        ld = C.shape[2]
        npend = k - k0
        if N == 0:                                        # C[k] stays zero; sigma0 is replicated, the norm partial is 0
            meas.zero_()
            meas[0:1].copy_(scal[0:1])
            return
        Cp = ctypes.c_void_p(C.data_ptr() + 8 * k0 * 3 * ld) if npend else None
        cout = ctypes.c_void_p(C.data_ptr() + 8 * k * 3 * ld)
        self._call("asb_deflate_finish", self._p(g), Cp, npend, self._p(coef) if npend else None, self._p(s), self._p(scal),
                   cout, self._p(e), N, ld, self._p(meas), self._p(self.ws_for(1, 3 * ld)))

    def rank_update(self, R, N, W, C, k0, nb, e):
        F, _, ld = R.shape
        if N == 0:
            return
        Wb = ctypes.c_void_p(W.data_ptr() + 8 * k0)
        Cb = ctypes.c_void_p(C.data_ptr() + 8 * k0 * 3 * ld)
        self._call("asb_rank_update", self._p(R), F, N, ld, Wb, W.shape[1], Cb, nb, self._p(e),
                   self._p(self.ws_for(F, 3 * ld)))

    # ------------------------------------------------------------------ dense contractions
    def gemm_nt(self, A, ra, B, rb, M, symmetric=False, out=None):
        """D = A[:ra, :M] @ B[:rb, :M].T for row-major 2-D views A, B (stride(0) = leading dimension)."""
        D = self.empty((ra, rb)) if out is None else out
        nbytes = self.lib.asb_gemm_nt_workspace_bytes(ra, rb, M, int(symmetric))
        if self._ws2 is None or self._ws2.numel() < nbytes:
            self._ws2 = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=self.device)
        self._call("asb_gemm_nt", self._p(A), A.stride(0), ra, self._p(B), B.stride(0), rb, M, self._p(D), D.stride(0),
                   int(symmetric), self._p(self._ws2))
        return D

    def gemm_nn(self, A, r, q, B, M, alpha, out, row_scale=None):
        """out[:r, :M] = alpha * diag(row_scale) * A[:r, :q] @ B[:q, :M]."""
        self._call("asb_gemm_nn", self._p(A), A.stride(0), r, q, self._p(B), B.stride(0), M, float(alpha),
                   self._p(row_scale), self._p(out), out.stride(0))
        return out

    def admm_iter(self, Hr, B0, Z, U, Lambda, K, N, inv_rho):
        """One ADMM iteration (asb_admm_iter == asb_admm_solve with iters = 1)."""
        self.admm_solve(Hr, B0, Z, U, Lambda, K, N, inv_rho, 1)

    def admm_solve(self, Hr, B0, Z, U, Lambda, K, N, inv_rho, iters):
        """`iters` fused ADMM iterations in one launch (Z, U in place).  The kernel wants Hr zero-padded to
        ceil16(K) x ceil16(K); a tensor that is not (tests, ad-hoc callers) is copied into a padded one here."""
        ld = Z.shape[2]
        if N == 0:
            return
        k8 = (K + 15) // 16 * 16
        if Hr.shape[0] < k8 or Hr.stride(0) < k8:
            Hp = self.zeros((k8, k8))
            Hp[:K, :K] = Hr[:K, :K]
            Hr = Hp
        nb = int(self.lib.asb_admm_workspace_bytes(int(K)))
        if self._admm_ws is None or self._admm_ws.numel() < nb:
            self._admm_ws = torch.empty(nb + 256, dtype=torch.uint8, device=self.device)
        self._call("asb_admm_solve_ws", self._p(Hr), Hr.stride(0), self._p(B0), self._p(Z), self._p(U), self._p(Lambda), K, N, ld,
                   float(inv_rho), int(iters), self._p(self._admm_ws))

    # ------------------------------------------------------------------ small dense
    def gram_small(self, W, F, K, diag_add, G):
        self._call("asb_gram_small", self._p(W), W.stride(0), F, K, float(diag_add), self._p(G), G.stride(0))

    def spd_inverse(self, H, K, Hinv, info):
        scratch = self.empty((K, Hinv.stride(0)))
        self._call("asb_spd_inverse", self._p(H), H.stride(0), K, self._p(Hinv), Hinv.stride(0), self._p(scratch), self._p(info))

    def small_gemm(self, transA, transB, m, n, k, alpha, A, B, D):
        self._call("asb_small_gemm", int(transA), int(transB), m, n, k, float(alpha), self._p(A), A.stride(0), self._p(B),
                   B.stride(0), self._p(D), D.stride(0))

    def weights_sweep(self, P, Q, W, F, K):
        self._call("asb_weights_sweep", self._p(P), P.stride(0), self._p(Q), Q.stride(0), self._p(W), W.stride(0), F, K,
                   self._p(self.ws_for(1, 4096)))

    def component_centers(self, C, Lambda, K, N, index_offset, val, idx, sp):
        ld = C.shape[2]
        if N == 0:
            val.fill_(-float("inf"))
            idx.fill_(torch.iinfo(torch.int64).max)
            if Lambda is not None:
                sp.zero_()
            return
        self._call("asb_component_centers", self._p(C), self._p(Lambda), K, N, ld, index_offset, self._p(val), self._p(idx),
                   self._p(sp), self._p(self.workspace(8 * (64 * K * 3 + 1024))))

    def syevd(self, A, n):
        """In place: rows of A become the eigenvectors (ascending eigenvalues returned)."""
        w = self.empty((n,))
        info = self.zeros((1,), I32)
        nbytes = self.lib.asb_syevd_workspace_bytes(n)
        if nbytes < 0:
            raise _lib.AsbError("cuSOLVER workspace query failed")
        ws = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=self.device)
        self._call("asb_syevd", self._p(A), A.stride(0), n, self._p(w), self._p(ws), int(nbytes), self._p(info))
        return w, info

    def cgs_dots(self, T, k, d, N, h):
        if N == 0:
            h[:k].zero_()
            return
        self._call("asb_cgs_dots", self._p(T), k, d, N, T.shape[2], self._p(h), self._p(self.workspace(8 * (2 * self.sm_count * max(k, 1) + 1024))))

    def cgs_project(self, T, k, d, N, h, nrm2):
        if N == 0:
            nrm2.zero_()
            return
        self._call("asb_cgs_project", self._p(T), k, d, N, T.shape[2], self._p(h), self._p(nrm2), self._p(self.workspace(8 * (T.shape[2] // 256 + 1024))))

    def row_scale(self, T, k, d, nrm2):
        self._call("asb_row_scale", self._p(T), k, d, T.shape[2], self._p(nrm2))

    # ------------------------------------------------------------------ DEIM
    def deim_gather(self, C, N, row0, Pt, k, Vp):
        self._call("asb_deim_gather", self._p(C), C.shape[2], N, row0, self._p(Pt), k, self._p(Vp), Vp.shape[1])

    def deim_gather_rhs(self, C, N, row0, Pt, k, rhs, Vp):
        self._call("asb_deim_gather_rhs", self._p(C), C.shape[2], N, row0, self._p(Pt), k, rhs, self._p(Vp), Vp.shape[1])

    def deim_residual_block(self, C, N, k, target, accumulate, a, e, maxabs):
        """The per-CTA max|r| partials live in `ws` and, with `accumulate`, are carried from one target of a block to the
        next: they get a buffer of their own -- the shared scratch is rewritten by asb_deim_solve between the targets (on a
        B200 that turned max|r| of a block into garbage, so a vanishing block residual went unnoticed)."""
        if N == 0:
            if not accumulate:
                e.zero_()
                maxabs.zero_()
            return
        need = (int(C.shape[2]) + 255) // 256 + 8
        buf = getattr(self, "_deim_pmax", None)
        if buf is None or buf.numel() < need:
            buf = self._deim_pmax = torch.zeros(need, dtype=F64, device=self.device)
        self._call("asb_deim_residual_block", self._p(C), C.shape[2], N, k, target, int(accumulate), self._p(a), a.stride(0),
                   self._p(e), self._p(maxabs), self._p(buf))

    def deim_solve(self, Vp, k, a, info):
        ws = self.workspace(3 * k * (k + 1) * 8 + 1024)
        self._call("asb_deim_solve", self._p(Vp), Vp.shape[1], k, self._p(a), a.stride(0), self._p(ws), self._p(info))

    def deim_residual(self, C, N, k, a, e, maxabs):
        if N == 0:
            e.zero_()
            maxabs.zero_()
            return
        self._call("asb_deim_residual", self._p(C), C.shape[2], N, k, self._p(a), a.stride(0), self._p(e), self._p(maxabs),
                   self._p(self.ws_for(1, C.shape[2])))

    # ------------------------------------------------------------------ geodesics
    def spmv(self, csr, x, y):
        rowptr, colidx, vals = csr
        nrhs, ldx = x.shape
        self._call("asb_spmv_csr", self._p(rowptr), self._p(colidx), self._p(vals), rowptr.numel() - 1, self._p(x), self._p(y),
                   nrhs, ldx)

    def batched_dot(self, x, y, n, out):
        nrhs, ldx = x.shape
        self._call("asb_batched_dot", self._p(x), self._p(y), n, nrhs, ldx, self._p(out), self._p(self.workspace(8 * (65 * nrhs + 1024))))

    def pcg_update_xr(self, x, r, z, p, Ap, dinv, rz, pAp, n):
        nrhs, ldx = x.shape
        self._call("asb_pcg_update_xr", self._p(x), self._p(r), self._p(z), self._p(p), self._p(Ap), self._p(dinv), self._p(rz),
                   self._p(pAp), None, n, nrhs, ldx)

    def pcg_update_p(self, p, z, rz_new, rz_old, n):
        nrhs, ldx = p.shape
        self._call("asb_pcg_update_p", self._p(p), self._p(z), self._p(rz_new), self._p(rz_old), None, n, nrhs, ldx)

    def axpby(self, y, alpha, a, row_scale, beta, c, n):
        nrhs, ldx = y.shape
        self._call("asb_axpby", self._p(y), float(alpha), self._p(a), self._p(row_scale), float(beta), self._p(c), n, nrhs, ldx)

    def remove_mean(self, x, n):
        nrhs, ldx = x.shape
        self._call("asb_remove_mean", self._p(x), n, nrhs, ldx, self._p(self.workspace(8 * (66 * nrhs + 1024))))

    def gather_rows(self, src, idx, out, n):
        self._call("asb_gather_rows", self._p(src), self._p(idx), self._p(out), n, src.shape[0], src.stride(0), out.stride(0))

    def mf_forward(self, desc, vals, src, slots, work, rhs, y, upd):
        self._call("asb_mf_forward", self._p(desc), self._p(vals), self._p(src), slots, self._p(work), work.shape[0], self._p(rhs),
                   self._p(y), self._p(upd), rhs.shape[0], rhs.stride(0), upd.stride(0))

    def mf_backward(self, desc, vals, bnd, work, y, x):
        self._call("asb_mf_backward", self._p(desc), self._p(vals), self._p(bnd), self._p(work), work.shape[0], self._p(y), self._p(x),
                   y.shape[0], y.stride(0))

    def mf_solve(self, desc, vals, src, slots, bnd, levels_host, rhs, y, x, upd):
        """All levels of one multifrontal solve from a single C call; levels_host: (nlevels, 6) int64 numpy (root first):
        forward work list and length, backward work list and length, max n1, max n1 + n2 of the level."""
        self._call("asb_mf_solve", self._p(desc), self._p(vals), self._p(src), slots, self._p(bnd),
                   ctypes.c_void_p(levels_host.ctypes.data), levels_host.shape[0], self._p(rhs), self._p(y), self._p(x), self._p(upd),
                   rhs.shape[0], rhs.stride(0), upd.stride(0))

    def mf_solve_delta(self, desc, vals, src, slots, bnd, levels_host, maxblk_host, fstart, parent, depth, order, idx, pathfront,
                       rp, y, x, upd):
        """asb_mf_solve for unit right-hand sides at the source vertices `idx` (forward pass along the tree paths only)."""
        self._call("asb_mf_solve_delta", self._p(desc), self._p(vals), self._p(src), slots, self._p(bnd),
                   ctypes.c_void_p(levels_host.ctypes.data), levels_host.shape[0], ctypes.c_void_p(maxblk_host.ctypes.data),
                   self._p(fstart), self._p(parent), self._p(depth), fstart.numel(), self._p(order), self._p(idx),
                   self._p(pathfront), self._p(rp), self._p(y), self._p(x), self._p(upd), idx.numel(), rp.stride(0), upd.stride(0))

    def heat_divergence(self, u, tri, gconst, dconst, vc_ptr, vc_idx, n, Xt, div):
        nrhs, ldx = u.shape
        self._call("asb_heat_divergence", self._p(u), nrhs, ldx, self._p(tri), self._p(gconst), self._p(dconst), tri.shape[0],
                   self._p(vc_ptr), self._p(vc_idx), n, self._p(Xt), self._p(div))

    def support_map(self, phi, n, dmin, dmax, mode, lam, row0, nloc, out):
        nrhs, ldx = phi.shape
        if nloc == 0:
            return
        self._call("asb_support_map", self._p(phi), n, nrhs, ldx, float(dmin), float(dmax), int(mode), float(lam), row0, nloc,
                   self._p(out), out.stride(0), self._p(self.workspace(8 * (66 * nrhs + 1024))))

    def delta_rhs(self, idx, n, rhs):
        nrhs, ldx = rhs.shape
        self._call("asb_delta_rhs", self._p(idx), n, nrhs, ldx, self._p(rhs))

    # ------------------------------------------------------------------ snapshot producer (PD local step)
    def pd_project(self, kind, q, elems, rest, sigma_min, sigma_max, out):
        """out[f][d][rows * t + j] = get_pi of element t in frame f (asb_pd_project); q (F, 3, ldq), out (F, 3, ldo)."""
        F, _, ldq = q.shape
        self._call("asb_pd_project", int(kind), self._p(q), ldq, F, self._p(elems), self._p(rest), elems.shape[0],
                   float(sigma_min), float(sigma_max), self._p(out), out.shape[2])
        return out

    def pd_vertex_bending(self, q, verts, ptr, nbr, w, rest, prevent_flips, out):
        F, _, ldq = q.shape
        self._call("asb_pd_vertex_bending", self._p(q), ldq, F, self._p(verts), self._p(ptr), self._p(nbr), self._p(w), self._p(rest),
                   verts.shape[0], int(bool(prevent_flips)), self._p(out), out.shape[2])
        return out

    # ------------------------------------------------------------------ alignment
    def align(self, frames, rigid):
        """frames: (F, N, 3) float32 on the device -> aligned float32 (F, N, 3) and the (F, 12) transforms."""
        F, N, _ = frames.shape
        cent = self.empty((F, 3))
        M = self.empty((F, 9))
        T = self.empty((F, 12))
        out = torch.empty_like(frames)
        ws = self._p(self.workspace(8 * (32 * 9 * F + 1024)))
        self._call("asb_align_centroids", self._p(frames), F, N, self._p(cent), ws)
        self._call("asb_align_cov", self._p(frames), F, N, self._p(cent), self._p(M), ws)
        self._call("asb_align_solve", self._p(M), self._p(cent), F, int(bool(rigid)), self._p(T))
        self._call("asb_align_apply", self._p(frames), F, N, self._p(T), self._p(out))
        return out, T
