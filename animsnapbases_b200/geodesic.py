"""Heat-method geodesic distances on the GPU (reference: utils/support.py:139-208, GeodesicDistanceComputation).

Set-up (once per mesh, like the reference's `__init__` which pre-factorises with SuperLU, utils/support.py:153-171)
assembles the cotangent Laplacian `Lc`, the lumped areas `A`, the time step `t = 10 h^2` and the per-triangle /
per-corner geometric constants on the host with numpy/scipy.sparse -- geometry only, no snapshot data.  Every
*query* then runs on the device through libasb kernels:

    u   = (A - t Lc)^-1 delta_idx          multifrontal Cholesky solve (mfsolve.py; asb_mf_forward / asb_mf_backward)
    X   = -grad u / |grad u|               per triangle
    div = integrated divergence            per-vertex gather over incident corners (deterministic)
    phi = Lc^+ div                          multifrontal solve of -Lc grounded at one vertex (defined up to a constant)
    phi -= min(phi); support map            fused in asb_support_map

A direct solver, like the reference's SuperLU factors (utils/support.py:169-171): u decays like exp(-d / sqrt(t)) over
hundreds of orders of magnitude and only its gradient *direction* is used, so the far field needs componentwise
relative accuracy, which a Krylov solver cannot give (it also under-reaches: k iterations see k hops).

Several sources are solved at once (multi right-hand side); maps are memoised by vertex id because phi depends on
the source vertex only (SPLOCS re-queries the same centres every outer iteration, posComponents.py:158-165).
The mesh operator is global: with row-sharded snapshots every rank solves on the full mesh and keeps its row slice.
"""
from __future__ import annotations

import numpy as np
from scipy import sparse


def _veclen(v):
    return np.sqrt(np.sum(v ** 2, axis=-1))


def _normalized(v):
    return v / _veclen(v)[..., None]


def cotan_laplacian(verts, tris):
    """Cotangent Laplacian (row sums zero, off-diagonals 0.5*(cot a + cot b)) and lumped vertex areas
    (utils/support.py:81-136)."""
    n = len(verts)
    I, J, W = [], [], []
    for i1, i2, i3 in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
        a, b, c = tris[:, i1], tris[:, i2], tris[:, i3]
        u = verts[b] - verts[a]
        v = verts[c] - verts[a]
        cot = (u * v).sum(axis=1) / _veclen(np.cross(u, v))
        I += [b, c]
        J += [c, b]
        W += [0.5 * cot, 0.5 * cot]
    L = sparse.csr_matrix((np.concatenate(W), (np.concatenate(I), np.concatenate(J))), shape=(n, n))
    L = (L - sparse.diags(np.asarray(L.sum(axis=1)).ravel())).tocsr()
    e1 = verts[tris[:, 1]] - verts[tris[:, 0]]
    e2 = verts[tris[:, 2]] - verts[tris[:, 0]]
    ta3 = 0.5 * _veclen(np.cross(e1, e2)) / 3.0
    area = np.zeros(n)
    for c in range(3):
        area += np.bincount(tris[:, c], ta3, minlength=n)
    return L, area


class HeatGeodesics:
    def __init__(self, ops, verts, tris, m: float = 10.0, leaf_size: int = 64):
        from .mfsolve import MultifrontalSolver
        self.ops = ops
        verts = np.asarray(verts, dtype=np.float64)
        tris = np.asarray(tris, dtype=np.int64)
        self.n = n = len(verts)
        self.T = len(tris)
        self.ldx = (n + 31) // 32 * 32
        e01 = verts[tris[:, 1]] - verts[tris[:, 0]]
        e12 = verts[tris[:, 2]] - verts[tris[:, 1]]
        e20 = verts[tris[:, 0]] - verts[tris[:, 2]]
        area = 0.5 * _veclen(np.cross(e01, e12))
        un = _normalized(np.cross(_normalized(e01), _normalized(e12)))
        inv2a = (1.0 / (2.0 * area))[:, None]
        # gradient constants in the order the kernel consumes them: (n x e01) u2 + (n x e12) u0 + (n x e20) u1
        gconst = np.concatenate([np.cross(un, e01) * inv2a, np.cross(un, e12) * inv2a, np.cross(un, e20) * inv2a], axis=1)
        # per-corner divergence constants (cot1, e1, cot2, e2) with the reference's 1/tan(arccos(.)) cotangents
        dconst = np.empty((self.T, 3, 8))
        for c, (i1, i2, i3) in enumerate(((0, 1, 2), (1, 2, 0), (2, 0, 1))):
            v1, v2, v3 = tris[:, i1], tris[:, i2], tris[:, i3]
            e1 = verts[v2] - verts[v1]
            e2 = verts[v3] - verts[v1]
            eo = verts[v3] - verts[v2]
            with np.errstate(all="ignore"):
                cot1 = 1 / np.tan(np.arccos((_normalized(-e2) * _normalized(-eo)).sum(axis=1)))
                cot2 = 1 / np.tan(np.arccos((_normalized(-e1) * _normalized(eo)).sum(axis=1)))
            dconst[:, c, 0] = cot1
            dconst[:, c, 1:4] = e1
            dconst[:, c, 4] = cot2
            dconst[:, c, 5:8] = e2
        # vertex -> incident corner CSR (corner id = 3 t + c), stable order = (c, t) like the reference's three passes
        corner_vertex = tris.T.ravel()                              # c-major: corner (c, t) -> vertex
        corner_id = (np.arange(self.T)[None, :] * 3 + np.arange(3)[:, None]).ravel()
        order = np.argsort(corner_vertex, kind="stable")
        vc_idx = corner_id[order].astype(np.int32)
        vc_ptr = np.zeros(n + 1, dtype=np.int32)
        np.cumsum(np.bincount(corner_vertex, minlength=n), out=vc_ptr[1:])

        h = np.mean([_veclen(e01), _veclen(e12), _veclen(e20)])
        self.t = m * h ** 2
        Lc, va = cotan_laplacian(verts, tris)
        self.vpos = ops.to_device(np.ascontiguousarray(verts))        # rest positions (n, 3): pick prediction of engine.deflate
        self.heat = MultifrontalSolver(ops, (sparse.diags(va) - self.t * Lc).tocsr(), verts, leaf_size)
        self.pois = MultifrontalSolver(ops, (-Lc).tocsr(), verts, leaf_size, ground=0)
        self.tri = ops.to_device(tris.astype(np.int32))
        self.gconst = ops.to_device(np.ascontiguousarray(gconst))
        self.dconst = ops.to_device(np.ascontiguousarray(dconst.reshape(self.T * 3, 8)))
        self.vc_ptr = ops.to_device(vc_ptr)
        self.vc_idx = ops.to_device(vc_idx)
        self.stats = {"queries": 0, "cache_hits": 0, "solves": 0, "fronts": self.heat.stats["fronts"],
                      "levels": self.heat.stats["levels"],
                      "operator_bytes": self.heat.stats["operator_bytes"] + self.pois.stats["operator_bytes"]}
        self._cache = {}
        self._graphs = {}            # captured single-source queries (support_single); None = capture failed, stay eager
        import os
        self.cache_rows = int(os.environ.get("ASB_GEO_CACHE_ROWS", "4"))   # memoised maps kept: at most cache_rows * B rows per key (LRU)

    _instances = {}              # (device, digest of mesh, leaf size) -> HeatGeodesics: the factorisation depends on the mesh only

    @classmethod
    def cached(cls, ops, verts, tris, leaf_size: int = 64, keep: int = 2):
        """The geodesic set-up of a mesh (ordering, symbolic and numeric multifrontal factorisation: ~10 s and 1.9 GB at 1M
        vertices) built once and shared by every snapshot object on the same rest mesh -- the reference redoes its two
        `splu` factorisations in every posSnapshots constructor (snapbases/posSnapshots.py:97-99)."""
        import hashlib
        v = np.ascontiguousarray(verts, dtype=np.float64)
        t = np.ascontiguousarray(tris, dtype=np.int64)
        key = (str(getattr(ops, "device", "")), id(ops), hashlib.sha1(v.tobytes() + t.tobytes()).hexdigest(), int(leaf_size))
        geo = cls._instances.get(key)
        if geo is None:
            while len(cls._instances) >= keep:
                cls._instances.pop(next(iter(cls._instances)))
            geo = cls._instances[key] = cls(ops, v, t, leaf_size=leaf_size)
        else:
            geo.reset()
        return geo

    def reset(self):
        """Drop the memoised support maps (a new basis build starts cold)."""
        self._cache = {}

    def _remember(self, cache, ids_now, new_rows):
        """Insert {id: row} (each row a tensor of its own, not a view into a chunk buffer) and evict the least recently used
        ids that the current request does not need once the table exceeds cache_rows * len(request) rows."""
        for i, row in new_rows:
            cache[i] = row
        for i in ids_now:
            cache.move_to_end(i)
        cap = max(self.cache_rows * len(ids_now), 64)
        while len(cache) > cap:
            cache.popitem(last=False)

    # ------------------------------------------------------------------ queries
    def support_single(self, idx_dev, mode, lam, dmin, dmax, row0, nloc, out):
        """`support` for ONE source vertex, replayed from a CUDA graph: the ~75 launches of a single-source query (memsets,
        tree-path forward pass, one launch per level and direction of two triangular solves, divergence, support map) are
        captured once per (mode, shard) with buffers the graph owns and cost one launch on the host afterwards.  The 200
        deflation queries are sequential and latency bound; on 8 GPUs the host time to issue them (2 ms each) was longer than
        the sharded HBM pass they hide behind (0.9 ms).  out: (1, ld) device tensor; idx_dev: (1,) int64 device tensor."""
        ops = self.ops
        if getattr(ops, "name", "") != "cuda" or self._graphs is None:
            return self.support(idx_dev, mode, lam, dmin, dmax, row0, nloc, out)
        import torch
        key = (int(mode), float(lam), float(dmin), float(dmax), int(row0), int(nloc), int(out.shape[1]))
        ent = self._graphs.get(key)
        if ent is None:
            try:
                sidx = ops.zeros((1,), torch.int64)
                sout = ops.zeros((1, out.shape[1]))
                sidx.copy_(idx_dev)
                self.support(sidx, mode, lam, dmin, dmax, row0, nloc, sout)        # eager warm-up (workspaces, attributes)
                torch.cuda.current_stream(ops.device).synchronize()
                n0 = ops.launch_count()
                graph = torch.cuda.CUDAGraph()
                timers, ops.timers = ops.timers, None                               # no per-launch timing events inside a capture
                try:
                    import os
                    prio = int(os.environ.get("ASB_SIDE_PRIORITY", "0"))   # kernel nodes inherit the capture stream's priority
                    cap = torch.cuda.Stream(ops.device, priority=prio)
                    with torch.cuda.graph(graph, stream=cap):
                        self.support(sidx, mode, lam, dmin, dmax, row0, nloc, sout)
                finally:
                    ops.timers = timers
                ent = self._graphs[key] = (graph, sidx, sout, ops.launch_count() - n0)
                self.stats["queries"] -= 2                                          # warm-up + capture are not queries
                self.stats["solves"] -= 2
            except Exception as e:                                                  # noqa: BLE001 -- eager launches still work
                print("[animsnapbases_b200] CUDA-graph capture of the single-source geodesic query failed (%s); "
                      "queries stay eager" % (e,))
                self._graphs = None
                return self.support(idx_dev, mode, lam, dmin, dmax, row0, nloc, out)
        graph, sidx, sout, nodes = ent
        sidx.copy_(idx_dev)
        graph.replay()
        out.copy_(sout)
        ops.add_launch_count(nodes)
        self.stats["queries"] += 1
        self.stats["solves"] += 1
        return out

    def distances(self, idx_dev):
        """phi (B, ldx) for the B source vertices in idx_dev (int64 device tensor); phi is NOT yet shifted by its
        minimum (asb_support_map does that, utils/support.py:206)."""
        ops, n = self.ops, self.n
        B = idx_dev.numel()
        u = self.heat.solve_delta(idx_dev)            # heat step from a delta source: forward pass along the tree path only
        Xt = ops.empty((B, self.T, 3))
        div = ops.zeros((B, self.ldx))
        ops.heat_divergence(u, self.tri, self.gconst, self.dconst, self.vc_ptr, self.vc_idx, n, Xt, div)
        # reference solves Lc phi = div; we solve (-Lc) psi = div  =>  phi = -psi
        psi = self.pois.solve(div)
        phi = ops.empty((B, self.ldx))
        ops.axpby(phi, -1.0, psi, None, 0.0, None, n)
        self.stats["solves"] += B
        return phi

    def support(self, idx_dev, mode, lam, dmin, dmax, row0, nloc, out, ids_host=None, chunk=64, comm=None):
        """out[b, :nloc] = support map of source idx_dev[b] restricted to rows [row0, row0+nloc).
        mode 0: 1 - map (deflation mask), 1: lam * map (SPLOCS Lambda), 2: phi - min(phi).
        With `ids_host` (the same indices as a host list) rows are memoised per vertex id.  With `comm` (world > 1) the
        new sources of a batch are dealt round-robin to the ranks: every rank solves its share on the full mesh (the heat
        method is global), cuts the support map of each field into the ranks' row slices and one all-to-all hands every
        rank its rows of every field -- the geodesic work of a SPLOCS iteration is divided by the world size instead of
        being repeated on every rank."""
        ops = self.ops
        B = idx_dev.numel()
        self.stats["queries"] += B
        if ids_host is None:
            for b0 in range(0, B, chunk):
                sl = slice(b0, min(B, b0 + chunk))
                phi = self.distances(idx_dev[sl])
                ops.support_map(phi, self.n, dmin, dmax, mode, lam, row0, nloc, out[sl])
            return out
        key = (mode, float(lam), float(dmin), float(dmax), row0, nloc)
        from collections import OrderedDict
        cache = self._cache.setdefault(key, OrderedDict())
        ids_now = [int(i) for i in ids_host]
        missing = sorted({i for i in ids_now if i not in cache})
        world = comm.world if comm is not None else 1
        if world > 1 and len(missing) > 1:
            new_rows = self._support_sharded(comm, missing, mode, lam, dmin, dmax, row0, nloc, out.shape[1], chunk)
            missing_local = []
        else:
            missing_local, new_rows = missing, []
        for b0 in range(0, len(missing_local), chunk):
            ids = missing_local[b0:b0 + chunk]
            idt = ops.to_device(np.asarray(ids, dtype=np.int64))
            phi = self.distances(idt)
            rows = ops.zeros((len(ids), out.shape[1]))
            ops.support_map(phi, self.n, dmin, dmax, mode, lam, row0, nloc, rows)
            new_rows += [(i, rows[j].clone()) for j, i in enumerate(ids)]
        self._remember(cache, ids_now, new_rows)
        self.stats["cache_hits"] += B - len(missing)
        for b, i in enumerate(ids_now):
            out[b].copy_(cache[i])
        return out

    def _support_sharded(self, comm, missing, mode, lam, dmin, dmax, row0, nloc, ld, chunk):
        """Round-robin the sources over the ranks, solve locally, exchange row slices (see `support`)."""
        ops = self.ops
        world, rank = comm.world, comm.rank
        skey = (row0, nloc, ld)
        if getattr(self, "_shards_key", None) != skey:
            self._shards = comm.allgather_ints([row0, nloc, ld])      # every rank's (row0, nloc, ld), once per sharding
            self._shards_key = skey
        ldmax = max(s[2] for s in self._shards)
        bmax = (len(missing) + world - 1) // world
        mine = missing[rank::world]
        send = [ops.zeros((bmax, ldmax)) for _ in range(world)]
        for b0 in range(0, len(mine), chunk):
            ids = mine[b0:b0 + chunk]
            phi = self.distances(ops.to_device(np.asarray(ids, dtype=np.int64)))
            for j, (r0j, nj, ldj) in enumerate(self._shards):
                if nj > 0:
                    ops.support_map(phi, self.n, dmin, dmax, mode, lam, r0j, nj, send[j][b0:b0 + len(ids)])
        recv = comm.all_to_all(send)
        return [(i, recv[j][b, :ld].clone()) for j in range(world) for b, i in enumerate(missing[j::world])]
