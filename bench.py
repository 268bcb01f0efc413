#!/usr/bin/env python
"""bench.py -- PCA (greedy deflation) + SPLOCS basis build time on synthetic snapshot matrices (BASELINE.json).

One step = one full basis build (`compute_components_store_singvalues()` with splocs.compute = true):
K deflation components (local geodesic support) + I_s SPLOCS outer iterations of I_a fused ADMM iterations each, on a
standardised snapshot tensor already resident in HBM.  `e2e` times the same build through the public API from pinned
HOST buffers (H2D of the float32 animation, ingest, standardisation, build, D2H of components and weights).

    python bench.py                              # 1 GPU, default workload
    torchrun --nproc-per-node N bench.py --gpus N # row-sharded over N GPUs (strong scaling, same total workload)
    python bench.py --impl reference             # CPU oracle port of the reference path on a bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pca_splocs_basis_build_time"
UNIT = "s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--grid", type=int, nargs=2, default=[1000, 1000], help="synthetic planar grid (nx ny) -> N = nx*ny vertices")
    ap.add_argument("--frames", type=int, default=2000)
    ap.add_argument("--comps", type=int, default=200)
    ap.add_argument("--splocs-iters", type=int, default=50)
    ap.add_argument("--admm-iters", type=int, default=10)
    ap.add_argument("--bumps", type=int, default=64)
    ap.add_argument("--sigma", type=float, default=0.05)
    ap.add_argument("--block", type=int, default=8, help="deflation rank-update block size")
    ap.add_argument("--leaf-size", type=int, default=64, help="nested-dissection leaf size of the geodesic factorisation")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", type=int, nargs=5, default=[160, 120, 140, 28, 2],
                    help="CPU baseline sample: nx ny frames comps splocs_iters")
    return ap.parse_args()


def work_units(N, F, K, Is, Ia):
    """Algorithmic work of one build: deflation bytes (K+1)*S and SPLOCS flops."""
    S = 24.0 * F * N
    defl_bytes = (K + 1) * S
    splocs_flops = Is * (4.0 * F * K * 3 * N + K * K * 3.0 * N + Ia * 2.0 * K * K * 3 * N)
    return defl_bytes, splocs_flops


# ====================================================================================== CPU oracle leg
def cpu_port_build(nx, ny, F, K, Is, Ia, bumps, sigma):
    """The reference path restated on the CPU (oracle/port.py): returns seconds for one build on this sample."""
    from animsnapbases_b200 import synth
    from oracle import port
    verts0, tris = synth.grid_mesh(nx, ny)
    anim = synth.bump_animation(verts0, F, n_bumps=min(bumps, 16), sigma=max(sigma, 0.1), seed=0, translate=0.0)
    mass = synth.lumped_vertex_masses(verts0, tris)
    prep = port.prepare_snapshots(anim, mass, "first")
    geo = port.GeodesicHeat(anim[0].astype(float), tris)
    t0 = time.perf_counter()
    C, W, meas, picked, R = port.extract_k_components(prep["snapTensor"], K, "local", geo)
    port.splocs_glob_optimization(prep["snapTensor"], C, W, R, geo, Is, Ia, 2.0, 10.0)
    return time.perf_counter() - t0


def cpu_baseline(args, N, F, K, Is, Ia):
    nx, ny, f, k, its = args.cpu_sample
    secs = cpu_port_build(nx, ny, f, k, its, Ia, args.bumps, args.sigma)
    # the reference's cost is ~ proportional to N*F per component / per (iteration, component): extrapolate linearly
    n = nx * ny
    full = secs * (N * F * (K + Is * K)) / (n * f * (k + its * k))
    cores = os.cpu_count() or 1
    return {"value": full, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle/port.py (numpy/scipy restatement of the reference) timed {secs:.2f} s on a {nx}x{ny} grid "
                      f"(N={n}), F={f}, K={k}, {its} SPLOCS x {Ia} ADMM iterations; value = linear extrapolation in "
                      f"N*F*K*(1+I_s) to the full workload (the reference cannot hold the full tensor: 4-5 S-sized copies)",
            "measured_sample_s": secs}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N = args.grid[0] * args.grid[1]
    F, K, Is, Ia = args.frames, args.comps, args.splocs_iters, args.admm_iters
    nx, ny, f, k, its = args.cpu_sample
    for _ in range(max(args.warmup, 0) and 1):
        cpu_port_build(max(nx // 4, 8), max(ny // 4, 8), max(f // 4, 8), min(k, 4), 1, 2, args.bumps, args.sigma)
    times = [cpu_port_build(nx, ny, f, k, its, Ia, args.bumps, args.sigma) for _ in range(max(1, args.steps))]
    secs = float(np.mean(times))
    n = nx * ny
    full = secs * (N * F * (K + Is * K)) / (n * f * (k + its * k))
    cores = os.cpu_count() or 1
    line = {"impl": "reference", "metric": METRIC, "value": full, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": full * 1e3, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, N, F, K, Is, Ia),
            "cpu_baseline": {"value": full, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"each step = oracle/port.py build on a {nx}x{ny} grid, F={f}, K={k}, {its}x{Ia} SPLOCS "
                                       f"({secs:.2f} s measured), linearly extrapolated in N*F*K*(1+I_s) to the config"},
            "e2e": {"value": full, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args, N, F, K, Is, Ia):
    return {"workload": f"synthetic {args.grid[0]}x{args.grid[1]} grid (N={N} vertices) x {F} frames, K={K} deflation "
                        f"components with local geodesic support 0.1/0.25 + {Is} SPLOCS iterations x {Ia} ADMM, lambda=2, rho=10",
            "vertices": N, "frames": F, "components": K, "splocs_iters": Is, "admm_iters": Ia,
            "sharding": "vertex rows over ranks", "l2_policy": "inputs larger than L2 (snapshot tensor >> 126 MB)"
            if 24.0 * N * F > 4 * 126e6 else "snapshot tensor fits in L2; L2 flushed between timed steps"}


# ====================================================================================== clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                          str(self.index), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([t.strip() for t in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ====================================================================================== GPU arm
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from animsnapbases_b200 import engine, synth
    from animsnapbases_b200.geodesic import HeatGeodesics
    from animsnapbases_b200.ops import CudaOps

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ops = CudaOps(f"cuda:{local_rank}")
    comm = engine.Comm()
    rank = comm.rank

    nx, ny = args.grid
    N, F, K, Is, Ia = nx * ny, args.frames, args.comps, args.splocs_iters, args.admm_iters
    sh = engine.Shard.make(N, world, rank)
    verts0, tris = synth.grid_mesh(nx, ny)
    mass = synth.lumped_vertex_masses(verts0, tris)
    centres, amps, _ = synth.bump_parameters(verts0, F, args.bumps, seed=0)
    rows = slice(sh.row0, sh.row0 + sh.n)
    v_loc = ops.to_device(verts0[rows])
    anim_dev = synth.device_bump_animation_from_params(v_loc, verts0[centres], amps, args.sigma)   # (F, n_loc, 3) f32
    geo = HeatGeodesics(ops, verts0, tris, leaf_size=args.leaf_size)
    mass_loc = mass[rows]

    def prepare(src_dev):
        return engine.prepare_snapshots(ops, comm, src_dev, mass_loc, "first", sh)

    phase_ev = []        # (deflation start, deflation end = SPLOCS start, SPLOCS end) CUDA events of every build

    def build(st):
        geo.reset()      # every build starts cold: support maps are memoised by source vertex WITHIN a build only
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
        C, W, meas, picked, _ = engine.deflate(ops, comm, st.X, sh, K, True, geo, block=args.block, keep_residual=False)
        ev[1].record()
        engine.splocs(ops, comm, st.X, sh, C, W, K, geo, Is, Ia, 2.0, 10.0, log=None)
        ev[2].record()
        phase_ev.append(ev)
        return C, W, meas

    st = prepare(anim_dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=ops.device) if 24.0 * N * F <= 4 * 126e6 else None

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        build(st)
    sync_all()
    # ---- timed region (device-resident inputs)
    timed = ["asb_deflate_gemv", "asb_rank_update", "asb_admm_solve", "asb_gemm_nn", "asb_gemm_nt", "asb_weights_sweep",
             "asb_component_centers", "asb_mf_solve", "asb_mf_solve_delta", "asb_support_map", "asb_vertex_energy"]
    ops.time_kernels(timed)
    ops.reset_launch_count()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    phase_ev.clear()
    ev0.record()
    for _ in range(args.steps):
        if flush is not None:
            flush.fill_(1)
        build(st)
    ev1.record()
    sync_all()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=ops.device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    launches = ops.launch_count()
    ktimes = ops.kernel_times()
    ops.time_kernels(None)
    clk = clocks.stop() if rank == 0 else None
    ms_per_step = float(ms.item()) / args.steps
    phase_ms = {"deflation": sum(e[0].elapsed_time(e[1]) for e in phase_ev) / args.steps,
                "splocs": sum(e[1].elapsed_time(e[2]) for e in phase_ev) / args.steps}
    phase_ev.clear()
    geo_stats = dict(geo.stats)

    # ---- e2e: host float32 animation (pinned) -> H2D -> ingest -> standardise -> build -> D2H comps + weights
    e2e = None
    if not args.no_e2e:
        host_in = torch.empty(anim_dev.shape, dtype=torch.float32, pin_memory=True)
        host_in.copy_(anim_dev)
        del anim_dev, st                                   # one resident snapshot tensor at a time (48 GB at 1M x 2000)
        torch.cuda.empty_cache()
        host_C = torch.empty((K, 3, sh.ld), dtype=torch.float64, pin_memory=True)
        host_W = torch.empty((F, (K + 3) // 4 * 4), dtype=torch.float64, pin_memory=True)
        n_e2e = max(1, min(args.steps, 2))
        sync_all()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            d_in = host_in.to(ops.device, non_blocking=True)
            st2 = prepare(d_in)
            del d_in
            C, W, meas = build(st2)
            host_C.copy_(C, non_blocking=True)
            host_W.copy_(W, non_blocking=True)
            torch.cuda.synchronize()
            del st2, C, W
        sync_all()
        t = torch.tensor([(time.perf_counter() - t0) / n_e2e], dtype=torch.float64, device=ops.device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": float(t.item()), "unit": UNIT, "h2d_bytes_per_step": int(host_in.numel() * 4 * world),
               "d2h_bytes_per_step": int((host_C.numel() + host_W.numel()) * 8 * world), "steps": n_e2e}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (by device time inside the timed region)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"
    # fp64 peak is not in MEASURED_PEAKS.json: measure cuBLAS DGEMM here (the library kernel to beat)
    a = torch.randn(8192, 8192, dtype=torch.float64, device=ops.device)
    b = torch.randn(8192, 8192, dtype=torch.float64, device=ops.device)
    best = 1e30
    for i in range(6):
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        torch.matmul(a, b)
        s1.record()
        torch.cuda.synchronize()
        if i:
            best = min(best, s0.elapsed_time(s1))
    fp64_peak = 2 * 8192.0 ** 3 / (best * 1e-3) / 1e12
    del a, b
    n_loc, ld = sh.n, sh.ld
    M = 3.0 * ld
    # ALGORITHMIC work of all launches of one entry point in one step (local shard)
    per_step = {
        "asb_deflate_gemv": ("hbm", K * 8.0 * F * M),                       # K read passes over R
        "asb_rank_update": ("hbm", -(-K // args.block) * 2 * 8.0 * F * M),   # read + write pass per block of components
        "asb_admm_solve": ("tensor", Is * Ia * 2.0 * K * K * M),             # Hr (Z - U), Ia iterations fused per launch
        "asb_gemm_nn": ("tensor", Is * 2.0 * K * F * M),                     # B0 = (H^-1 W^T) X
        "asb_gemm_nt": ("tensor", (Is + 1) * (2.0 * F * K + K * (K + 1.0)) * M),   # P = X C^T and the symmetric Q = C C^T
    }
    dom = max(ktimes.items(), key=lambda kv: kv[1][1])[0] if ktimes else None
    roof = None
    if dom in per_step and ktimes[dom][0]:
        bound, work = per_step[dom]
        work *= args.steps
        tot_s = ktimes[dom][1] * 1e-3
        traffic = None      # measured DRAM bytes per launch (ncu capture of the same kernel at this shape), if we have one
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
            w = tr["workload"]
            if world == 1 and (w["vertices"], w["frames"], w["components"], w["admm_iters"]) == (N, F, K, Ia):
                traffic = tr["per_launch_bytes"].get(dom)
        except Exception:
            pass
        common = {"kernel": dom, "bound": bound, "traffic": traffic, "avg_launch_ms": ktimes[dom][1] / ktimes[dom][0],
                  "launches": ktimes[dom][0], "work_per_launch": work / ktimes[dom][0]}
        if bound == "hbm":
            ach = work / tot_s / 1e9
            roof = dict(common, achieved=ach, peak=hbm_peak, unit="GB/s", frac=ach / hbm_peak, peak_source=hbm_src)
        else:
            ach = work / tot_s / 1e12
            roof = dict(common, achieved=ach, peak=fp64_peak, unit="TFLOP/s", frac=ach / fp64_peak,
                        peak_source="cuBLAS DGEMM 8192^3 measured in this run (fp64 is not in MEASURED_PEAKS.json; "
                                    "sm_100a has no fp64 tcgen05 kind: DMMA.8x8x4 is the fp64 tensor path)")
    defl_bytes, splocs_flops = work_units(N, F, K, Is, Ia)
    ideal = defl_bytes / world / (hbm_peak * 1e9) + splocs_flops / world / (fp64_peak * 1e12)
    line = {
        "metric": METRIC, "value": ms_per_step / 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args, N, F, K, Is, Ia), "gpu_launches": int(launches),
        "clocks": clk, "e2e": e2e, "roofline": roof,
        "path_roofline": {"ideal_s": ideal, "frac": ideal / (ms_per_step / 1e3), "hbm_gbs": hbm_peak, "fp64_tflops": fp64_peak,
                          "algorithmic": {"deflation_bytes": defl_bytes, "splocs_flops": splocs_flops}},
        "phase_ms_per_step": phase_ms,
        "kernel_ms_per_step": {k: v[1] / args.steps for k, v in ktimes.items()},
        "kernel_launches_per_step": {k: v[0] / args.steps for k, v in ktimes.items()},
        "geodesic": geo_stats,
    }
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(args, N, F, K, Is, Ia)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
