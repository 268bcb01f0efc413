"""-m gpu: every C-ABI kernel of libasb.so against its numpy statement (tests/numpy_ops.py) on seeded inputs,
including ragged sizes (N not a multiple of the tile sizes, odd K/F), padding, and tie-breaking cases."""
import numpy as np
import pytest
import torch

from conftest import rel
from numpy_ops import NumpyOps

pytestmark = pytest.mark.gpu

REF = NumpyOps()


def dev(ops, t):
    return None if t is None else t.to(ops.device)


def planar(rng, R, N, ld, scale=1.0):
    t = torch.zeros((R, 3, ld), dtype=torch.float64)
    t[:, :, :N] = torch.from_numpy(rng.standard_normal((R, 3, N)) * scale)
    return t


def close(a, b, tol, what=""):
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    b = b.detach().cpu().numpy() if hasattr(b, "detach") else np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = rel(a, b) if np.linalg.norm(b) > 0 else float(np.abs(a).max())
    assert err <= tol, f"{what}: rel err {err:.3e} > {tol:.1e}"


# ------------------------------------------------------------------------------------------------ boundary
@pytest.mark.parametrize("R,N,dtype", [(3, 37, torch.float32), (5, 1000, torch.float64), (2, 32, torch.float32)])
def test_ingest_export_pack(cuda_ops, R, N, dtype):
    rng = np.random.default_rng(0)
    ld = (N + 31) // 32 * 32
    src = torch.from_numpy(rng.standard_normal((R, N, 3))).to(dtype)
    rs = torch.from_numpy(rng.random(N) + 0.5)
    for scale in (None, rs):
        Xr = REF.ingest(src, scale, ld)
        Xg = cuda_ops.ingest(dev(cuda_ops, src), dev(cuda_ops, scale), ld)
        assert torch.equal(Xg.cpu(), Xr), "ingest must be bit-exact (one multiply per element)"
        Er = REF.export_aos(Xr, scale, N)
        Eg = cuda_ops.export_aos(Xg, dev(cuda_ops, scale), N)
        assert torch.equal(Eg.cpu(), Er)
    assert torch.equal(cuda_ops.pack_bin(Xg, R, N).cpu(), REF.pack_bin(Xr, R, N))


@pytest.mark.parametrize("F,N", [(7, 45), (40, 3000)])
def test_center(cuda_ops, F, N):
    rng = np.random.default_rng(1)
    ld = (N + 31) // 32 * 32
    X = planar(rng, F, N, ld)
    X[:, :, :N] += 3.0
    Xg = dev(cuda_ops, X)
    mr, mg = REF.col_mean(X), cuda_ops.col_mean(Xg)
    assert torch.equal(mg.cpu(), mr), "column mean is sequential in f: bit-exact"
    close(cuda_ops.center_stats(Xg, mg, N), REF.center_stats(X, mr, N), 1e-12, "center_stats")
    REF.center_apply(X, mr, 1.7, N)
    cuda_ops.center_apply(Xg, mg, 1.7, N)
    assert torch.equal(Xg.cpu(), X), "center_apply: two roundings like numpy -> bit-exact"
    REF.unstandardize(X, mr, 1.7, N)
    cuda_ops.unstandardize(Xg, mg, 1.7, N)
    assert torch.equal(Xg.cpu(), X)


# ------------------------------------------------------------------------------------------------ deflation pieces
@pytest.mark.parametrize("F,N", [(9, 50), (33, 5000), (300, 700)])
def test_vertex_energy_argmax(cuda_ops, F, N):
    rng = np.random.default_rng(2)
    ld = (N + 31) // 32 * 32
    R = planar(rng, F, N, ld)
    Rg = dev(cuda_ops, R)
    er, eg = REF.vertex_energy(R, N), cuda_ops.vertex_energy(Rg, N)
    close(eg, er, 1e-14, "vertex_energy")
    # argmax with an exact tie: lowest index must win
    e = er.clone()
    e[N // 2] = e[N - 1] = 10 * float(e.max())
    val, idx = cuda_ops.zeros((1,)), cuda_ops.zeros((1,), torch.int64)
    cuda_ops.argmax_first(dev(cuda_ops, e), N, 1000, val, idx)
    assert int(idx.item()) == N // 2 + 1000 and float(val.item()) == float(e[N // 2])


def test_argmax_large(cuda_ops):
    rng = np.random.default_rng(3)
    n = 3_000_017
    v = torch.from_numpy(rng.random(n))
    v[2_500_000] = v[17] = 2.0
    val, idx = cuda_ops.zeros((1,)), cuda_ops.zeros((1,), torch.int64)
    cuda_ops.argmax_first(dev(cuda_ops, v), n, 0, val, idx)
    assert int(idx.item()) == 17


def test_argmax_outside_ball(cuda_ops):
    """asb_argmax_outside_ball against the numpy statement: first maximum among the vertices outside the ball, shard offset,
    and the no-candidate answer (-inf, INT64_MAX)."""
    from numpy_ops import NumpyOps
    rng = np.random.default_rng(5)
    nglob, row0, n = 50_000, 12_345, 30_011
    pos = torch.from_numpy(rng.random((nglob, 3)))
    v = torch.from_numpy(rng.random(n))
    centre_id = row0 + 777
    far = np.nonzero(((pos.numpy()[row0:row0 + n] - pos.numpy()[centre_id]) ** 2).sum(axis=1) > 0.3 ** 2)[0]
    v[777] = 5.0                                   # the global maximum sits at the centre: must be skipped
    v[int(far[10])] = v[int(far[3])] = 3.0         # tie outside the ball: the lower index wins
    centre = torch.tensor([centre_id], dtype=torch.int64)
    ref = NumpyOps()
    rp = torch.zeros(2, dtype=torch.float64)
    ref.argmax_outside_ball(v, n, row0, pos, centre, 0.3, rp)
    pair, idx = cuda_ops.zeros((2,)), cuda_ops.zeros((1,), torch.int64)
    vd, pd, cd = dev(cuda_ops, v), dev(cuda_ops, pos), dev(cuda_ops, centre)
    cuda_ops.argmax_outside_ball(vd, n, row0, pd, cd, 0.3, pair)
    got = pair.cpu()
    assert float(got[0]) == float(rp[0]) == 3.0
    assert int(got[1:2].view(torch.int64)[0]) == int(rp[1:2].view(torch.int64)[0]) == row0 + int(far[3])
    # reduction over "ranks": the record of this shard against a better one, a tie with a lower index, and nothing at all
    allp = torch.zeros((3, 2), dtype=torch.float64)
    allp[0] = got
    allp[1, 0], allp[2, 0] = 3.0, 2.0
    allp[1, 1:2].view(torch.int64)[0] = 11
    allp[2, 1:2].view(torch.int64)[0] = 5
    cuda_ops.select_pair(dev(cuda_ops, allp), 3, cd, idx)
    assert int(idx.item()) == 11
    cuda_ops.argmax_outside_ball(vd, n, row0, pd, cd, 10.0, pair)
    got = pair.cpu()
    assert float(got[0]) == -np.inf and int(got[1:2].view(torch.int64)[0]) == np.iinfo(np.int64).max
    cuda_ops.select_pair(pair, 1, cd, idx)
    assert int(idx.item()) == centre_id


@pytest.mark.parametrize("local", [False, True])
def test_traj_and_gemv_finish_update(cuda_ops, local):
    rng = np.random.default_rng(4)
    F, N, K = 37, 203, 8
    ld = (N + 31) // 32 * 32
    R = planar(rng, F, N, ld)
    W = torch.zeros((F, 8), dtype=torch.float64)
    C = torch.zeros((K, 3, ld), dtype=torch.float64)
    W[:, :3] = torch.from_numpy(rng.random((F, 3)))
    C[:3, :, :N] = torch.from_numpy(rng.standard_normal((3, 3, N)) * 0.1)
    k0, k = 1, 3                                   # two pending components (1, 2), computing component 3
    Rg, Wg, Cg = dev(cuda_ops, R), dev(cuda_ops, W), dev(cuda_ops, C)
    idx = torch.tensor([57 + 500], dtype=torch.int64)
    tr, tg = torch.zeros((3, F), dtype=torch.float64), cuda_ops.zeros((3, F))
    REF.gather_traj(R, N, idx, 500, W, k0, C, k - k0, tr)
    cuda_ops.gather_traj(Rg, N, dev(cuda_ops, idx), 500, Wg, k0, Cg, k - k0, tg)
    close(tg, tr, 1e-15, "gather_traj")
    # not the owner: untouched
    t2 = cuda_ops.zeros((3, F))
    cuda_ops.gather_traj(Rg, N, dev(cuda_ops, torch.tensor([5], dtype=torch.int64)), 500, Wg, k0, Cg, k - k0, t2)
    assert float(t2.abs().max().item()) == 0.0
    cr, sr = torch.zeros(16, dtype=torch.float64), torch.zeros(2, dtype=torch.float64)
    cg, sg = cuda_ops.zeros((16,)), cuda_ops.zeros((2,))
    REF.traj_rank1(tr, F, local, W, k, k0, cr, sr)
    cuda_ops.traj_rank1(tg, F, local, Wg, k, k0, cg, sg)
    close(Wg[:, k], W[:, k], 1e-12, "traj_rank1 w")
    close(sg, sr, 1e-12, "traj_rank1 scalars")
    close(cg[:2], cr[:2], 1e-12, "traj_rank1 coef")
    gr, gg = torch.zeros((3, ld), dtype=torch.float64), cuda_ops.zeros((3, ld))
    REF.deflate_gemv(R, N, W, k, gr)
    cuda_ops.deflate_gemv(Rg, N, Wg, k, gg)
    close(gg, gr, 1e-13, "deflate_gemv")
    s = torch.zeros(ld, dtype=torch.float64)
    s[:N] = torch.from_numpy(rng.random(N))
    er = REF.vertex_energy(R, N) * 5
    eg = dev(cuda_ops, er.clone())
    mr, mg = torch.zeros(3, dtype=torch.float64), cuda_ops.zeros((3,))
    REF.deflate_finish(gr, C, k, k0, cr, s if local else None, sr, er, N, mr)
    cuda_ops.deflate_finish(gg, Cg, k, k0, cg, dev(cuda_ops, s) if local else None, sg, eg, N, mg)
    close(Cg[k], C[k], 1e-12, "deflate_finish c")
    close(eg, er, 1e-12, "deflate_finish e")
    close(mg, mr, 1e-12, "deflate_finish meas")
    REF.rank_update(R, N, W, C, k0, 3, er)
    cuda_ops.rank_update(Rg, N, Wg, Cg, k0, 3, eg)
    close(Rg, R, 1e-14, "rank_update R")
    close(eg, er, 1e-13, "rank_update e")


@pytest.mark.parametrize("nb", [1, 5, 8, 13, 16])
def test_rank_update_sizes(cuda_ops, nb):
    rng = np.random.default_rng(5)
    F, N = 150, 1111
    ld = (N + 31) // 32 * 32
    R = planar(rng, F, N, ld)
    W = torch.from_numpy(rng.random((F, 16)))
    C = planar(rng, 16, N, ld, 0.1)
    e = torch.zeros(ld, dtype=torch.float64)
    Rg, eg = dev(cuda_ops, R), cuda_ops.zeros((ld,))
    REF.rank_update(R, N, W, C, 0, nb, e)
    cuda_ops.rank_update(Rg, N, dev(cuda_ops, W), dev(cuda_ops, C), 0, nb, eg)
    close(Rg, R, 1e-14, "rank_update R")
    close(eg, e, 1e-13, "rank_update e")


def test_gemv_large_is_one_slab(cuda_ops):
    """N large enough that the column blocks alone fill the machine (no frame split): exercises the streaming path."""
    rng = np.random.default_rng(6)
    F, N = 64, 200_000
    ld = (N + 31) // 32 * 32
    R = planar(rng, F, N, ld)
    W = torch.from_numpy(rng.standard_normal((F, 4)))
    g, gg = torch.zeros((3, ld), dtype=torch.float64), cuda_ops.zeros((3, ld))
    REF.deflate_gemv(R, N, W, 2, g)
    cuda_ops.deflate_gemv(dev(cuda_ops, R), N, dev(cuda_ops, W), 2, gg)
    close(gg, g, 1e-13, "deflate_gemv large")


# ------------------------------------------------------------------------------------------------ DMMA contractions
@pytest.mark.parametrize("ra,rb,M", [(5, 7, 96), (130, 70, 3 * 352), (200, 200, 3 * 4000), (257, 129, 6400), (300, 200, 3 * 2400),
                                     (201, 333, 3 * 1024), (1000, 256, 3 * 640), (300, 290, 3 * 800)])
def test_gemm_nt(cuda_ops, ra, rb, M):
    rng = np.random.default_rng(7)
    A = torch.from_numpy(rng.standard_normal((ra, M)))
    B = torch.from_numpy(rng.standard_normal((rb, M)))
    D = cuda_ops.gemm_nt(dev(cuda_ops, A), ra, dev(cuda_ops, B), rb, M)
    close(D, REF.gemm_nt(A, ra, B, rb, M), 1e-13, "gemm_nt")


@pytest.mark.parametrize("r,M", [(24, 3 * 320), (200, 3 * 2016), (300, 3 * 1024)])
def test_gemm_nt_symmetric(cuda_ops, r, M):
    rng = np.random.default_rng(8)
    A = torch.from_numpy(rng.standard_normal((r, M)))
    Ag = dev(cuda_ops, A)
    D = cuda_ops.gemm_nt(Ag, r, Ag, r, M, symmetric=True)
    close(D, REF.gemm_nt(A, r, A, r, M), 1e-13, "gemm_nt symmetric")
    assert torch.equal(D, D.T.contiguous()), "Gram must be exactly symmetric"


@pytest.mark.parametrize("r,q,M", [(10, 24, 3 * 64), (200, 200, 3 * 1056), (70, 71, 3 * 320), (24, 24, 3 * 1824), (200, 333, 3 * 2080),
                                   (256, 100, 3 * 352), (129, 50, 3 * 96), (300, 64, 3 * 352)])
def test_gemm_nn(cuda_ops, r, q, M):
    rng = np.random.default_rng(9)
    lda = (q + 1) // 2 * 2
    A = torch.zeros((r, lda), dtype=torch.float64)
    A[:, :q] = torch.from_numpy(rng.standard_normal((r, q)))
    B = torch.from_numpy(rng.standard_normal((q, M)))
    rs = torch.from_numpy(rng.random(r) + 0.5)
    for scale in (None, rs):
        out_r = torch.zeros((r, M), dtype=torch.float64)
        out_g = cuda_ops.zeros((r, M))
        REF.gemm_nn(A, r, q, B, M, 0.5, out_r, scale)
        cuda_ops.gemm_nn(dev(cuda_ops, A), r, q, dev(cuda_ops, B), M, 0.5, out_g, dev(cuda_ops, scale))
        close(out_g, out_r, 1e-13, "gemm_nn")


@pytest.mark.parametrize("K,N", [(10, 100), (64, 333), (100, 1000), (200, 520), (230, 77)])
def test_admm_iter(cuda_ops, K, N):
    rng = np.random.default_rng(10)
    ld = (N + 31) // 32 * 32
    Kp = (K + 3) // 4 * 4
    Hr = torch.zeros((K, Kp), dtype=torch.float64)
    H = rng.standard_normal((K, K)) * 0.05
    Hr[:, :K] = torch.from_numpy(H @ H.T + 0.5 * np.eye(K))
    B0, Z, U = planar(rng, K, N, ld), planar(rng, K, N, ld), planar(rng, K, N, ld, 0.3)
    Lam = torch.zeros((K, ld), dtype=torch.float64)
    Lam[:, :N] = torch.from_numpy(rng.random((K, N)) * 16.0)
    Lam[:, : N // 3] = 0.0                              # inside the support centre: no shrinkage
    Zg, Ug = dev(cuda_ops, Z), dev(cuda_ops, U)
    REF.admm_iter(Hr, B0, Z, U, Lam, K, N, 0.1)
    cuda_ops.admm_iter(dev(cuda_ops, Hr), dev(cuda_ops, B0), Zg, Ug, dev(cuda_ops, Lam), K, N, 0.1)
    close(Zg, Z, 1e-12, "admm Z")
    close(Ug, U, 1e-12, "admm U")
    assert float((Z[:, :, :N] == 0).double().mean()) > 0.01, "test must exercise the shrink-to-zero branch"
    assert float(Zg[:, :, N:].abs().max().item()) == 0.0 and float(Ug[:, :, N:].abs().max().item()) == 0.0


@pytest.mark.parametrize("K,N,iters", [(10, 100, 3), (64, 333, 4), (128, 200, 2), (200, 520, 10), (201, 77, 3), (256, 70, 2)])
def test_admm_solve(cuda_ops, K, N, iters):
    """All ADMM iterations of one outer step in one launch == the same number of single iterations of the numpy statement."""
    rng = np.random.default_rng(12)
    ld = (N + 31) // 32 * 32
    Kp = (K + 3) // 4 * 4
    Hr = torch.zeros((K, Kp), dtype=torch.float64)
    H = rng.standard_normal((K, K)) * 0.03
    Hr[:, :K] = torch.from_numpy(H @ H.T + 0.4 * np.eye(K))
    B0, Z, U = planar(rng, K, N, ld), planar(rng, K, N, ld), planar(rng, K, N, ld, 0.3)
    Lam = torch.zeros((K, ld), dtype=torch.float64)
    Lam[:, :N] = torch.from_numpy(rng.random((K, N)) * 16.0)
    Lam[:, : N // 3] = 0.0
    Zg, Ug = dev(cuda_ops, Z), dev(cuda_ops, U)
    REF.admm_solve(Hr, B0, Z, U, Lam, K, N, 0.1, iters)
    cuda_ops.admm_solve(dev(cuda_ops, Hr), dev(cuda_ops, B0), Zg, Ug, dev(cuda_ops, Lam), K, N, 0.1, iters)
    close(Zg, Z, 1e-11, "admm_solve Z")
    close(Ug, U, 1e-11, "admm_solve U")
    assert float((Z[:, :, :N] == 0).double().mean()) > 0.01
    assert float(Zg[:, :, N:].abs().max().item()) == 0.0 and float(Ug[:, :, N:].abs().max().item()) == 0.0


def test_admm_bulk_copy_kernel_is_bit_identical_to_cp_async_kernel(cuda_ops):
    """The default ADMM kernel (bulk copies + mbarriers, layout 5) and r01's cp.async kernel (layout 4) accumulate every entry
    in the same order, so they must agree bit for bit -- at a size where a CTA runs several vertex tiles (N = 14 290: 1 787
    tiles on 296 CTAs) and the co-resident CTAs of an SM drift apart.  Guards the hand-over of the slab ring: a stage released
    before its last fragment load has returned shows up here as whole vertex tiles that differ from run to run
    (profiles/r02_admm_layouts.md)."""
    K, N, iters = 200, 14290, 10
    ld = (N + 31) // 32 * 32
    g = torch.Generator(device="cpu").manual_seed(1)
    Wm = torch.randn((300, K), generator=g, dtype=torch.float64)
    H = 10.0 * torch.linalg.inv(Wm.T @ Wm / 300 + 10.0 * torch.eye(K, dtype=torch.float64))
    k16 = (K + 15) // 16 * 16
    Hr = torch.zeros((k16, k16), dtype=torch.float64)
    Hr[:K, :K] = H

    def field(scale):
        t = torch.zeros((K, 3, ld), dtype=torch.float64)
        t[:, :, :N] = scale * torch.randn((K, 3, N), generator=g, dtype=torch.float64)
        return dev(cuda_ops, t)
    B0, Z0, U0 = field(1.0), field(1.0), field(0.1)
    Lam = torch.zeros((K, ld), dtype=torch.float64)
    Lam[:, :N] = 0.5 * torch.rand((K, N), generator=g, dtype=torch.float64)
    Hr, Lam = dev(cuda_ops, Hr), dev(cuda_ops, Lam)

    def run(layout):
        cuda_ops.lib.asb_admm_select_layout(layout, 0)
        Z, U = Z0.clone(), U0.clone()
        cuda_ops.admm_solve(Hr, B0, Z, U, Lam, K, N, 0.1, iters)
        torch.cuda.synchronize()
        return Z, U
    try:
        Z4, U4 = run(4)
        for rep in range(4):
            Z5, U5 = run(5)
            assert int((Z5 != Z4).sum()) == 0 and int((U5 != U4).sum()) == 0, f"repetition {rep}"
    finally:
        cuda_ops.lib.asb_admm_select_layout(-1, -1)


# ------------------------------------------------------------------------------------------------ small dense
@pytest.mark.parametrize("F,K", [(24, 10), (200, 64), (333, 201)])
def test_small_dense(cuda_ops, F, K):
    rng = np.random.default_rng(11)
    Kp = (K + 3) // 4 * 4
    W = torch.zeros((F, Kp), dtype=torch.float64)
    W[:, :K] = torch.from_numpy(rng.random((F, K)))
    Wg = dev(cuda_ops, W)
    G, Gg = torch.zeros((K, Kp), dtype=torch.float64), cuda_ops.zeros((K, Kp))
    REF.gram_small(W, F, K, 10.0, G)
    cuda_ops.gram_small(Wg, F, K, 10.0, Gg)
    close(Gg, G, 1e-13, "gram_small")
    Hi, Hig = torch.zeros((K, Kp), dtype=torch.float64), cuda_ops.zeros((K, Kp))
    info = cuda_ops.zeros((1,), torch.int32)
    REF.spd_inverse(G.clone(), K, Hi, torch.zeros(1, dtype=torch.int32))
    cuda_ops.spd_inverse(Gg.clone(), K, Hig, info)
    assert int(info.item()) == 0
    close(Hig, Hi, 1e-11, "spd_inverse")
    Fp = (F + 1) // 2 * 2
    A, Ag = torch.zeros((K, Fp), dtype=torch.float64), cuda_ops.zeros((K, Fp))
    REF.small_gemm(0, 1, K, F, K, 1.0, Hi, W, A)
    cuda_ops.small_gemm(0, 1, K, F, K, 1.0, Hig, Wg, Ag)
    close(Ag, A, 1e-11, "small_gemm")
    # weight sweep with one dead component (Q_kk <= 1e-8)
    Cm = rng.standard_normal((K, 300))
    Cm[K // 2] *= 1e-6
    X = rng.standard_normal((F, 300))
    P = torch.from_numpy(X @ Cm.T)
    Q = torch.from_numpy(Cm @ Cm.T)
    W2, W2g = W.clone(), dev(cuda_ops, W.clone())
    REF.weights_sweep(P, Q, W2, F, K)
    cuda_ops.weights_sweep(dev(cuda_ops, P), dev(cuda_ops, Q), W2g, F, K)
    close(W2g, W2, 1e-10, "weights_sweep")
    assert float(W2[:, K // 2].abs().max()) == 0.0


def test_spd_inverse_flags_indefinite(cuda_ops):
    H = cuda_ops.to_device(np.array([[1.0, 2.0, 0, 0], [2.0, 1.0, 0, 0]]))
    info = cuda_ops.zeros((1,), torch.int32)
    cuda_ops.spd_inverse(H, 2, cuda_ops.zeros((2, 4)), info)
    assert int(info.item()) != 0


@pytest.mark.parametrize("K,N", [(5, 100), (200, 1500), (3, 70_001)])
def test_component_centers(cuda_ops, K, N):
    rng = np.random.default_rng(12)
    ld = (N + 31) // 32 * 32
    C = planar(rng, K, N, ld)
    C[1, :, 3] = C[1, :, N - 2] = 50.0                       # tie -> lowest index
    Lam = torch.zeros((K, ld), dtype=torch.float64)
    Lam[:, :N] = torch.from_numpy(rng.random((K, N)))
    vr, ir, sr = torch.zeros(K, dtype=torch.float64), torch.zeros(K, dtype=torch.int64), torch.zeros(K, dtype=torch.float64)
    vg, ig, sg = cuda_ops.zeros((K,)), cuda_ops.zeros((K,), torch.int64), cuda_ops.zeros((K,))
    REF.component_centers(C, Lam, K, N, 7, vr, ir, sr)
    cuda_ops.component_centers(dev(cuda_ops, C), dev(cuda_ops, Lam), K, N, 7, vg, ig, sg)
    assert torch.equal(ig.cpu(), ir) and int(ir[1]) == 3 + 7
    assert torch.equal(vg.cpu(), vr), "centre energies are (x^2+y^2)+z^2: bit-exact"
    close(sg, sr, 1e-13, "sparsity")


@pytest.mark.parametrize("n", [6, 70, 301])
def test_syevd(cuda_ops, n):
    rng = np.random.default_rng(13)
    A = rng.standard_normal((n, 3 * n))
    G = A @ A.T
    lda = (n + 1) // 2 * 2
    Gp = torch.zeros((n, lda), dtype=torch.float64)
    Gp[:, :n] = torch.from_numpy(G)
    Gg = dev(cuda_ops, Gp)
    w, info = cuda_ops.syevd(Gg, n)
    assert int(info.item()) == 0
    wr = np.linalg.eigvalsh(G)
    close(w, torch.from_numpy(wr), 1e-12, "eigenvalues")
    V = Gg.cpu().numpy()[:, :n]                               # rows = eigenvectors
    assert np.abs(V @ V.T - np.eye(n)).max() < 1e-12
    assert np.abs(V @ G @ V.T - np.diag(wr)).max() < 1e-10 * wr[-1]


def test_cgs(cuda_ops):
    rng = np.random.default_rng(14)
    K, N = 12, 1000
    ld = (N + 31) // 32 * 32
    T = planar(rng, K, N, ld)
    T[:, :, :N] += 5.0                                        # nearly parallel rows (mean added to every component)
    Tg = dev(cuda_ops, T)
    h, nrm2 = cuda_ops.zeros((K,)), cuda_ops.zeros((1,))
    for d in range(3):
        for k in range(K):
            for _ in range(2):
                if k:
                    cuda_ops.cgs_dots(Tg, k, d, N, h)
                cuda_ops.cgs_project(Tg, k, d, N, h if k else None, nrm2)
            cuda_ops.row_scale(Tg, k, d, nrm2)
    Q = Tg.cpu().numpy()
    for d in range(3):
        q = Q[:, d, :N]
        assert np.abs(q @ q.T - np.eye(K)).max() < 1e-13
        qr = np.linalg.qr(T.numpy()[:, d, :N].T)[0].T
        sg = np.sign((q * qr).sum(axis=1))
        assert rel(q * sg[:, None], qr) < 1e-10


# ------------------------------------------------------------------------------------------------ DEIM pieces
def test_deim_pieces(cuda_ops):
    rng = np.random.default_rng(15)
    K, N = 9, 400
    ld = (N + 31) // 32 * 32
    C = planar(rng, K, N, ld)
    Cg = dev(cuda_ops, C)
    k = 6
    Pt = torch.from_numpy(rng.choice(N, size=K, replace=False).astype(np.int64) + 100)
    Vr, Vg = torch.zeros((3, K + 1, K + 1), dtype=torch.float64), cuda_ops.zeros((3, K + 1, K + 1))
    REF.deim_gather(C, N, 100, Pt, k, Vr)
    cuda_ops.deim_gather(Cg, N, 100, dev(cuda_ops, Pt), k, Vg)
    assert torch.equal(Vg.cpu(), Vr)
    ar, ag = torch.zeros((3, 12), dtype=torch.float64), cuda_ops.zeros((3, 12))
    info = cuda_ops.zeros((1,), torch.int32)
    REF.deim_solve(Vr, k, ar, None)
    cuda_ops.deim_solve(Vg, k, ag, info)
    assert int(info.item()) == 0
    close(ag, ar, 1e-11, "deim_solve")
    for kk, a_r, a_g in ((k, ar, ag), (0, None, None)):
        er, eg = torch.zeros(ld, dtype=torch.float64), cuda_ops.zeros((ld,))
        mr, mg = torch.zeros(1, dtype=torch.float64), cuda_ops.zeros((1,))
        REF.deim_residual(C, N, kk, a_r if a_r is not None else ar, er, mr)
        cuda_ops.deim_residual(Cg, N, kk, a_g if a_g is not None else ag, eg, mg)
        close(eg, er, 1e-11, "deim_residual e")
        close(mg, mr, 1e-11, "deim_residual max")


# ------------------------------------------------------------------------------------------------ geodesic kernels
def test_geodesic_kernels(cuda_ops):
    from animsnapbases_b200 import synth
    from animsnapbases_b200.geodesic import HeatGeodesics
    verts, tris = synth.grid_mesh(17, 13)
    verts = verts + 0.01 * np.random.default_rng(16).standard_normal(verts.shape)
    gr, gg = HeatGeodesics(REF, verts, tris), HeatGeodesics(cuda_ops, verts, tris)
    n, ldx = gr.n, gr.ldx
    rng = np.random.default_rng(17)
    x = torch.zeros((3, ldx), dtype=torch.float64)
    x[:, :n] = torch.from_numpy(rng.standard_normal((3, n)))
    xg = dev(cuda_ops, x)
    yr, yg = torch.zeros_like(x), cuda_ops.zeros((3, ldx))
    from animsnapbases_b200.geodesic import cotan_laplacian
    Lc = cotan_laplacian(verts, tris)[0].tocsr()
    Lc.sort_indices()
    csr = (Lc.indptr.astype(np.int32), Lc.indices.astype(np.int32), Lc.data)
    REF.spmv(tuple(REF.to_device(a) for a in csr), x, yr)
    cuda_ops.spmv(tuple(cuda_ops.to_device(a) for a in csr), xg, yg)
    close(yg, yr, 1e-14, "spmv")
    dr, dg = torch.zeros(3, dtype=torch.float64), cuda_ops.zeros((3,))
    REF.batched_dot(x, yr, n, dr)
    cuda_ops.batched_dot(xg, yg, n, dg)
    close(dg, dr, 1e-13, "batched_dot")
    Xr, Xg = torch.zeros((3, gr.T, 3), dtype=torch.float64), cuda_ops.zeros((3, gr.T, 3))
    vr, vg = torch.zeros_like(x), cuda_ops.zeros((3, ldx))
    REF.heat_divergence(x, gr.tri, gr.gconst, gr.dconst, gr.vc_ptr, gr.vc_idx, n, Xr, vr)
    cuda_ops.heat_divergence(xg, gg.tri, gg.gconst, gg.dconst, gg.vc_ptr, gg.vc_idx, n, Xg, vg)
    close(Xg, Xr, 1e-13, "unit gradient")
    close(vg, vr, 1e-12, "divergence")
    for mode in (0, 1, 2):
        orr, og = torch.zeros((3, 64), dtype=torch.float64), cuda_ops.zeros((3, 64))
        REF.support_map(x, n, 0.1, 0.9, mode, 2.0, 32, 50, orr)
        cuda_ops.support_map(xg, n, 0.1, 0.9, mode, 2.0, 32, 50, og)
        close(og, orr, 1e-14, f"support_map mode {mode}")
    REF.remove_mean(x, n)
    cuda_ops.remove_mean(xg, n)
    close(xg, x, 1e-13, "remove_mean")
    ids = torch.tensor([0, n // 2, n - 1], dtype=torch.int64)
    pr, pg = gr.distances(ids), gg.distances(dev(cuda_ops, ids))
    orr, og = torch.zeros((3, ldx), dtype=torch.float64), cuda_ops.zeros((3, ldx))
    REF.support_map(pr, n, 0, 0, 2, 0, 0, n, orr)
    cuda_ops.support_map(pg, n, 0, 0, 2, 0, 0, n, og)
    close(og, orr, 1e-11, "heat-method distances (multifrontal solve on device vs numpy statement)")


@pytest.mark.parametrize("B", [1, 5, 19])    # 1: single-RHS kernels; 5: one partial group of 8; 19: 2 full groups + 3
@pytest.mark.parametrize("nx,ny,leaf,ground", [(9, 7, 64, None), (60, 45, 16, None), (150, 120, 64, None), (60, 45, 16, 5),
                                               (150, 120, 32, 0)])
def test_multifrontal_solve(cuda_ops, nx, ny, leaf, ground, B):
    """asb_mf_forward / asb_mf_backward (+ batched device factorisation) against SuperLU, the reference's solver
    (utils/support.py:169-171): componentwise relative accuracy for the heat system, residual for the grounded Poisson."""
    from scipy import sparse
    from scipy.sparse.linalg import splu
    from animsnapbases_b200 import synth
    from animsnapbases_b200.geodesic import cotan_laplacian
    from animsnapbases_b200.mfsolve import MultifrontalSolver
    verts, tris = synth.grid_mesh(nx, ny)
    verts = verts + (0.2 / max(nx, ny)) * np.random.default_rng(3).standard_normal(verts.shape)
    n = len(verts)
    Lc, va = cotan_laplacian(verts, tris)
    h2 = float(np.mean(va))
    A = (sparse.diags(va) - 10.0 * h2 * Lc).tocsc() if ground is None else (-Lc).tocsc()
    mf = MultifrontalSolver(cuda_ops, A, verts, leaf_size=leaf, ground=ground)
    ldx = mf.ldx
    rng = np.random.default_rng(4)
    rhs = np.zeros((B, ldx))
    if ground is None:
        rhs[np.arange(B), rng.integers(0, n, B)] = 1.0              # heat sources
        x = mf.solve(cuda_ops.to_device(rhs)).cpu().numpy()[:, :n]
        lu = splu(A)
        ref = np.stack([lu.solve(rhs[b, :n]) for b in range(B)])
        assert ref.min() > 0
        err = np.max(np.abs(x - ref) / ref)
        assert err < 1e-11, f"componentwise relative error {err:.2e} (smallest value {ref.min():.1e})"
        # unit right-hand sides through the tree-path forward pass (asb_mf_solve_delta)
        src_idx = np.argmax(rhs[:, :n], axis=1).astype(np.int64)
        xd = mf.solve_delta(cuda_ops.to_device(src_idx)).cpu().numpy()[:, :n]
        errd = np.max(np.abs(xd - ref) / ref)
        assert errd < 1e-11, f"delta solve: componentwise relative error {errd:.2e}"
    else:
        d = rng.standard_normal((B, n))
        rhs[:, :n] = d - d.mean(axis=1, keepdims=True)
        x = mf.solve(cuda_ops.to_device(rhs)).cpu().numpy()[:, :n]
        assert np.all(x[:, ground] == 0.0)
        res = np.linalg.norm((A @ x.T).T - rhs[:, :n]) / np.linalg.norm(rhs)
        assert res < 1e-11, res


# ------------------------------------------------------------------------------------------------ alignment
@pytest.mark.parametrize("rigid", [True, False])
def test_align(cuda_ops, rigid):
    from conftest import golden
    g = golden("align")
    frames = torch.from_numpy(g["frames"])
    out, T = cuda_ops.align(dev(cuda_ops, frames), rigid)
    ref64, Tr = REF.align(frames, rigid)                      # fp64 statement of the same arithmetic
    assert np.abs(out.cpu().numpy() - ref64.numpy()).max() <= 2.5e-7, "vs fp64 numpy statement (<= 2 float32 ulp at |x|~1)"
    close(T, Tr, 1e-9, "transforms")
    # the reference estimates R in float32 (utils/process.py:219-233 on float32 .h5 data): parity within float32 accuracy
    assert np.abs(out.cpu().numpy() - g[f"aligned_rigid{int(rigid)}"]).max() <= 5e-6
