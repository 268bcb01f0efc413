#!/usr/bin/env python
"""Per-kernel timing of the hot path at a given snapshot shape (CUDA events on the launching stream, inputs >> L2).
Prints one JSON line per C-ABI entry point with its algorithmic bytes / flops and the achieved rate, next to a plain
device copy (HBM ceiling) and cuBLAS DGEMM (fp64 tensor ceiling).  Used to pick the kernel to optimise next."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from animsnapbases_b200.ops import CudaOps, pad_ld  # noqa: E402


def timed(fn, reps=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best * 1e-3


def pd_bench(cells, frames):
    """asb_pd_project / asb_pd_vertex_bending on a structured block: HBM-bound gather/scatter, one thread per
    (element, frame).  Algorithmic bytes per element and frame: rest data + node ids + rows * 24 B written +
    the element's share of the vertex planes."""
    import numpy as np
    from animsnapbases_b200 import pd_local, runtime, synth
    from animsnapbases_b200.engine import Comm
    ops = CudaOps("cuda:0")
    runtime.install(ops, Comm(enabled=False))
    verts, tets = synth.tet_grid(cells, cells, cells)
    nv, e = len(verts), len(tets)
    q0 = ops.ingest(ops.to_device(np.ascontiguousarray(verts[None])), None, pad_ld(nv))
    q = q0.repeat(frames, 1, 1)
    q += 0.05 * torch.sin(7.0 * q.roll(1, dims=1) + torch.arange(frames, device=q.device, dtype=q.dtype)[:, None, None])
    for name, rest_b, ids_b in (("tets_deformation_gradient", 72, 16), ("tets_strain", 72, 16)):
        grp = pd_local.ConstraintGroup(name, verts, tets)
        out = ops.zeros((frames, 3, pad_ld(3 * e)))
        t = timed(lambda: ops.pd_project(grp.kind, q, grp._elems, grp._rest, grp.sigma_min, grp.sigma_max, out))
        byts = frames * (e * (rest_b + ids_b + 3 * 24.0) + nv * 24.0)
        print(json.dumps({"kernel": "asb_pd_project/" + name, "elements": e, "verts": nv, "frames": frames, "ms": t * 1e3,
                          "GB/s": byts / t / 1e9, "Melem_frames/s": e * frames / t / 1e6}), flush=True)
    tris = np.concatenate([tets[:, [0, 1, 2]], tets[:, [0, 1, 3]]])[: 2 * e]
    es = pd_local.ConstraintGroup("edge_spring", verts, np.unique(np.sort(tris[:, :2], axis=1), axis=0))
    out = ops.zeros((frames, 3, pad_ld(es.num_elements)))
    t = timed(lambda: ops.pd_project(es.kind, q, es._elems, es._rest, 0.0, 0.0, out))
    print(json.dumps({"kernel": "asb_pd_project/edge_spring", "elements": es.num_elements, "frames": frames, "ms": t * 1e3,
                      "GB/s": frames * (es.num_elements * (8 + 8 + 24.0) + nv * 24.0) / t / 1e9}), flush=True)


def geo_bench(nx, ny, leaf_size=64):
    """Heat-method geodesics on an nx x ny grid: factorisation set-up, single-source queries (the 200 sequential deflation
    queries: latency bound, replicated on every rank) and a batch of 40 (the new centres of a SPLOCS iteration)."""
    import time
    import numpy as np
    from animsnapbases_b200 import synth
    from animsnapbases_b200.geodesic import HeatGeodesics
    ops = CudaOps("cuda:0")
    verts, tris = synth.grid_mesh(nx, ny)
    t0 = time.perf_counter()
    geo = HeatGeodesics(ops, verts, tris, leaf_size=leaf_size)
    torch.cuda.synchronize()
    setup = time.perf_counter() - t0
    n = nx * ny
    rng = np.random.default_rng(0)
    out = ops.zeros((1, geo.ldx))
    srcs = [ops.to_device(np.array([int(i)], dtype=np.int64)) for i in rng.integers(0, n, size=16)]
    state = {"i": 0}

    def one():
        geo.support(srcs[state["i"] % len(srcs)], 0, 1.0, 0.1, 0.25, 0, n, out)
        state["i"] += 1
    t1 = timed(one, reps=12, warm=3)

    def one_graph():
        geo.support_single(srcs[state["i"] % len(srcs)], 0, 1.0, 0.1, 0.25, 0, n, out)
        state["i"] += 1
    tg = timed(one_graph, reps=24, warm=4)
    ids = ops.to_device(rng.integers(0, n, size=40).astype(np.int64))
    outb = ops.zeros((40, geo.ldx))
    tb = timed(lambda: geo.support(ids, 1, 2.0, 0.1, 0.25, 0, n, outb), reps=3, warm=1)
    ob = geo.stats["operator_bytes"]
    print(json.dumps({"kernel": "geodesic_query", "grid": [nx, ny], "setup_s": setup, "operator_bytes": ob, "levels": geo.stats["levels"],
                      "leaf_size": leaf_size, "single_source_graph_ms": tg * 1e3,
                      "single_source_ms": t1 * 1e3, "single_source_GB/s": 1.5 * ob / t1 / 1e9, "batch40_ms": tb * 1e3,
                      "batch40_ms_per_query": tb * 1e3 / 40, "mf_tall": os.environ.get("ASB_MF_TALL", "1"),
                      "note": "single source: 3 of the 4 triangular passes stream every operator once (1.5 x operator bytes)"}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geo", type=int, nargs=2, metavar=("NX", "NY"), default=None, help="time heat-method geodesic queries on a grid")
    ap.add_argument("--leaf-size", type=int, default=64)
    ap.add_argument("--verts", type=int, default=1000000)
    ap.add_argument("--frames", type=int, default=2000)
    ap.add_argument("--comps", type=int, default=200)
    ap.add_argument("--only", default="")
    ap.add_argument("--once", action="store_true", help="one launch per kernel, no warm-up (for ncu --set full captures)")
    ap.add_argument("--pd", type=int, nargs=2, metavar=("CELLS", "FRAMES"), default=None,
                    help="time the snapshot-producer kernels instead: CELLS^3 grid cells x 6 tets, FRAMES position frames")
    args = ap.parse_args()
    if args.pd:
        return pd_bench(args.pd[0], args.pd[1])
    if args.geo:
        return geo_bench(args.geo[0], args.geo[1], args.leaf_size)
    if args.once:
        global timed
        _t = timed
        timed = lambda fn, reps=1, warm=0: _t(fn, 1, 0)  # noqa: E731
    ops = CudaOps("cuda:0")
    N, F, K = args.verts, args.frames, args.comps
    ld = pad_ld(N)
    M = 3 * ld
    S = 8.0 * F * M
    only = set(args.only.split(",")) if args.only else None

    def want(n):
        return only is None or n in only

    X = ops.empty((F, 3, ld))
    for f0 in range(0, F, 50):
        X[f0:f0 + 50].normal_()
    Kp = (K + 3) // 4 * 4
    W = ops.empty((F, Kp)).normal_()
    C = ops.empty((K, 3, ld)).normal_()
    out = []

    def rec(name, secs, nbytes=None, flops=None, note=""):
        r = {"kernel": name, "ms": secs * 1e3}
        if nbytes:
            r["GB/s"] = nbytes / secs / 1e9
        if flops:
            r["TFLOP/s"] = flops / secs / 1e12
        if note:
            r["note"] = note
        print(json.dumps(r), flush=True)
        out.append(r)

    # ceilings
    if want("copy"):
        a = torch.empty(1 << 30, dtype=torch.float64, device=ops.device)
        b = torch.empty_like(a)
        rec("torch_copy_8GiB", timed(lambda: b.copy_(a)), nbytes=2.0 * a.numel() * 8)
        rec("torch_sum_read_8GiB", timed(lambda: a.sum()), nbytes=1.0 * a.numel() * 8)
        del a, b
    if want("dgemm"):
        a = torch.randn(8192, 8192, dtype=torch.float64, device=ops.device)
        b = torch.randn(8192, 8192, dtype=torch.float64, device=ops.device)
        rec("cublas_dgemm_8192", timed(lambda: torch.matmul(a, b), reps=5), flops=2.0 * 8192 ** 3)
        del a, b
        xs = X.view(F, M)[:, : min(M, 300000)]
        rec("cublas_dgemm_WtX_K%dxF%dx%d" % (K, F, xs.shape[1]), timed(lambda: torch.matmul(W[:, :K].t(), xs)),
            flops=2.0 * K * F * xs.shape[1])
        rec("cublas_gram_XXt_F%dx%d" % (F, xs.shape[1]), timed(lambda: torch.matmul(xs, xs.t())), flops=2.0 * F * F * xs.shape[1],
            note="full (non-symmetric) flop count")
    # streaming
    e = ops.empty((ld,))
    g = ops.empty((3, ld))
    if want("stream"):
        rec("asb_vertex_energy", timed(lambda: ops.vertex_energy(X, N, out=e)), nbytes=S)
        rec("asb_deflate_gemv", timed(lambda: ops.deflate_gemv(X, N, W, 0, g)), nbytes=S)
        Wz = ops.zeros((F, Kp))
        for nb in ((8,) if args.once else (4, 8, 16)):
            rec("asb_rank_update_nb%d" % nb, timed(lambda: ops.rank_update(X, N, Wz, C, 0, nb, e), reps=2), nbytes=2 * S)
        mean = ops.zeros((3, ld))
        rec("asb_center_stats", timed(lambda: ops.center_stats(X, mean, N)), nbytes=S)
        rec("asb_center_apply", timed(lambda: ops.center_apply(X, mean, 1.0, N), reps=2), nbytes=2 * S)
        val = ops.zeros((K,)); idx = ops.zeros((K,), torch.int64); sp = ops.zeros((K,))
        rec("asb_component_centers", timed(lambda: ops.component_centers(C, None, K, N, 0, val, idx, sp)), nbytes=8.0 * K * M)
    Xm, Cm = X.view(F, M), C.view(K, M)
    if want("gemm"):
        rec("asb_gemm_nt_XCt", timed(lambda: ops.gemm_nt(Xm, F, Cm, K, M)), flops=2.0 * F * K * M, nbytes=S + 8.0 * K * M)
        rec("asb_gemm_nt_CCt_sym", timed(lambda: ops.gemm_nt(Cm, K, Cm, K, M, symmetric=True)), flops=1.0 * K * (K + 1) * M,
            nbytes=8.0 * K * M)
        Fp = (F + 1) // 2 * 2
        A = ops.empty((K, Fp)).normal_()
        B0 = ops.empty((K, 3, ld))
        rec("asb_gemm_nn_B0", timed(lambda: ops.gemm_nn(A, K, F, Xm, M, 1.0, B0.view(K, M))), flops=2.0 * K * F * M,
            nbytes=S + 8.0 * K * M)
        del B0
    if want("gram"):
        G = ops.zeros((F, (F + 1) // 2 * 2))
        rec("asb_gemm_nt_gram_sym", timed(lambda: ops.gemm_nt(Xm, F, Xm, F, M, symmetric=True, out=G), reps=1, warm=1),
            flops=1.0 * F * (F + 1) * M, nbytes=S, note="symmetric-half flops")
    if want("admm"):
        B0 = ops.empty((K, 3, ld)).normal_()
        U = ops.zeros((K, 3, ld))
        Lam = ops.empty((K, ld)).uniform_(0, 2)
        K8 = (K + 15) // 16 * 16
        Hr = ops.zeros((K8, K8))
        Hr[:K, :K] = torch.eye(K, dtype=torch.float64, device=ops.device) * 0.5 + 0.01 * torch.randn(K, K, dtype=torch.float64, device=ops.device)
        if not args.once:
            rec("asb_admm_iter", timed(lambda: ops.admm_iter(Hr, B0, C, U, Lam, K, N, 0.1)), flops=2.0 * K * K * M,
                nbytes=8.0 * K * M * 5 + 8.0 * K * ld)
        rec("asb_admm_solve_x10", timed(lambda: ops.admm_solve(Hr, B0, C, U, Lam, K, N, 0.1, 10), reps=2), flops=20.0 * K * K * M,
            nbytes=8.0 * K * M * 5 + 8.0 * K * ld)
    json.dump(out, open(os.path.join("gpurun_out", "kbench_%dx%d.json" % (N, F)), "w"), indent=1)


if __name__ == "__main__":
    main()
