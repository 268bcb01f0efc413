#!/usr/bin/env python
"""Bitwise A/B of two ADMM kernel layouts on the same inputs (needs a GPU).

    python tools/admm_ab.py N K ITERS [REPS] [--layout 5] [--variants 0]

Runs asb_admm_solve_ws with layout 4 (cp.async pair, r01's kernel) and layout 5 (bulk-copy pair, the product default; or
--layout 6 / 7 with --variants = the duet's ASB_ADMM_SKEW values) and reports, per variant and repetition, how many entries
of Z differ bit for bit, which vertex tiles they fall into (tile -> CTA slot, round of that CTA) and which
rows (-> which row-warp's m-tile).  The two kernels accumulate in the same order, so a correct layout 5 differs nowhere.
"""
from __future__ import annotations

import argparse
import collections
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from animsnapbases_b200 import ops as ops_mod  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("N", type=int)
    ap.add_argument("K", type=int)
    ap.add_argument("iters", type=int)
    ap.add_argument("reps", type=int, nargs="?", default=2)
    ap.add_argument("--variants", default="0")
    ap.add_argument("--layout", type=int, default=5)
    a = ap.parse_args()
    ops = ops_mod.CudaOps()
    N, K = a.N, a.K
    ld = ops_mod.pad_ld(N)
    g = torch.Generator(device="cpu").manual_seed(1)
    W = torch.randn((300, K), generator=g, dtype=torch.float64)
    rho = 10.0
    H = rho * torch.linalg.inv(W.T @ W / 300 + rho * torch.eye(K, dtype=torch.float64))
    k16 = (K + 15) // 16 * 16
    Hr = torch.zeros((k16, k16), dtype=torch.float64)
    Hr[:K, :K] = H
    Hr = Hr.cuda()
    B0 = torch.zeros((K, 3, ld), dtype=torch.float64)
    B0[:, :, :N] = torch.randn((K, 3, N), generator=g, dtype=torch.float64)
    Z0 = torch.zeros_like(B0)
    Z0[:, :, :N] = torch.randn((K, 3, N), generator=g, dtype=torch.float64)
    U0 = torch.zeros_like(B0)
    U0[:, :, :N] = 0.1 * torch.randn((K, 3, N), generator=g, dtype=torch.float64)
    Lam = torch.zeros((K, ld), dtype=torch.float64)
    Lam[:, :N] = torch.rand((K, N), generator=g, dtype=torch.float64) * 0.5
    B0, Z0, U0, Lam = B0.cuda(), Z0.cuda(), U0.cuda(), Lam.cuda()

    def run(layout, skew):
        ops.lib.asb_admm_select_layout(layout, skew)
        Z, U = Z0.clone(), U0.clone()
        ops.admm_solve(Hr, B0, Z, U, Lam, K, N, 1.0 / rho, a.iters)
        torch.cuda.synchronize()
        return Z, U

    Za, Ua = run(4, 0)
    Za2, _ = run(4, 0)
    print(f"N={N} K={K} iters={a.iters}: layout 4 run-to-run differing entries: {int((Za != Za2).sum())}")
    sms = ops.sm_count
    ntiles = (ld + 7) // 8
    grid = min(ntiles, 2 * sms)
    mtiles = (K + 7) // 8
    for v in [int(x) for x in a.variants.split(",")]:
        for rep in range(a.reps):
            Zb, Ub = run(a.layout, v)
            bad = (Zb != Za)
            nbad = int(bad.sum())
            line = f"layout {a.layout} variant {v:3d} rep {rep}: differing Z entries {nbad}"
            if nbad:
                badn = bad.any(dim=0).any(dim=0).cpu().numpy()            # per vertex
                tiles = np.unique(np.nonzero(badn)[0] // 8)
                slots = collections.Counter(int((t % grid) // max(1, grid // 2)) for t in tiles)
                rounds = collections.Counter(int(t // grid) for t in tiles)
                badk = bad.any(dim=1).cpu().numpy()                         # (K, ld)
                rows_per_tile = collections.Counter()
                mtile_hist = collections.Counter()
                for t in tiles[:4000]:
                    rk = np.nonzero(badk[:, t * 8:(t + 1) * 8].any(axis=1))[0]
                    rows_per_tile[len(rk)] += 1
                    if len(rk) <= 8:
                        mtile_hist[int(rk[0] // 8)] += 1
                line += (f" | tiles {len(tiles)} slots {sorted(slots.items())} rounds {sorted(rounds.items())}"
                         f" | rows/tile {sorted(rows_per_tile.items())} | single m-tile index {sorted(mtile_hist.items())}"
                         f" (of {mtiles}) | max |dZ| {float((Zb - Za).abs().max()):.3e}")
            print(line, flush=True)
    ops.lib.asb_admm_select_layout(-1, -1)


if __name__ == "__main__":
    main()
