// SPLOCS ADMM inner loop (snapbases/posComponents.py:175-178, prox :252-256), all iterations of one outer step fused.
//
//   repeat iters times:   C = B0 + Hr (Z - U);   V = C + U;   Z = V * max(0, 1 - Lambda / (rho |V|_3));   U = V - Z
//
// The iteration couples the K components of ONE vertex only (Hr is K x K, the prox acts on the 3-vector of a (k, n)
// pair), so a CTA owns a tile of NV = 8*WN vertices (x 3 planes = 24*WN columns) for the whole inner loop:
// D = Z - U and U stay in shared memory, the K x 24*WN accumulators in registers, and Z, U travel HBM <-> SM exactly once
// per outer step instead of once per ADMM iteration.  What is left per iteration is the K x K x 24*WN DMMA product, whose
// left operand Hr (320 KB at K = 200, does not fit beside the tile) streams L2 -> shared through a 3-stage cp.async
// ring that keeps running across iterations.  B0 and Lambda of the tile are re-read in the epilogue (L2 hits: the
// 148 resident tiles hold 15 MB).  Persistent grid: one CTA per SM looping over vertex tiles.
//
// Warp layout: 8 warps = WM row-warps x WN vertex groups, warp id = wm * WN + wn.  The ceil(K/8) m-tiles are dealt to the
// row-warps as evenly as possible at run time (K = 200: 7,6,6,6), and with this warp numbering the four SM sub-partitions
// (warp id mod 4) get 13,13,12,12 tiles x 3 planes: 96 % DMMA balance without padding K to a multiple of 32.
#include "dmma.cuh"

namespace asb {

constexpr int ADMM_BK = 8;             // k-slab of Hr per pipeline stage
constexpr int ADMM_ALD = ADMM_BK + 4;  // 12 doubles: conflict-free 8x4 A-fragment reads
constexpr int ADMM_STAGES = 3;

template <int WM, int WN, int MT>
__global__ void __launch_bounds__(WM * WN * 32, 1)
admm_solve_kernel(const double* __restrict__ Hr, int64_t ldh, const double* __restrict__ B0, double* __restrict__ Z,
                  double* __restrict__ U, const double* __restrict__ Lambda, int K, int64_t N, int64_t ld,
                  double inv_rho, int iters, int64_t ntiles) {
    constexpr int THREADS = WM * WN * 32;
    constexpr int NV = 8 * WN;        // vertices per tile
    constexpr int NC = 24 * WN;       // columns per tile: [vertex group][plane][8 vertices]
    constexpr int DLD = NC + 4;       // 28 / 52: stride mod 16 in {4, 12} -> conflict-free 4x8 B-fragment reads
    extern __shared__ __align__(16) double smem[];
    const int rows8 = (K + 7) & ~7;   // rows held in shared memory (rows >= K are zero)
    double* Ds = smem;                          // [rows8][DLD]  Z - U
    double* Us = Ds + (size_t)rows8 * DLD;      // [rows8][DLD]  U
    double* As = Us + (size_t)rows8 * DLD;      // ADMM_STAGES x [rows8][ADMM_ALD]
    const int astage = rows8 * ADMM_ALD;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int wm = warp / WN, wn = warp % WN;
    int mt0, mtn;
    mtile_range(rows8 >> 3, WM, wm, mt0, mtn);
    const int64_t C3 = 3 * ld;
    const int nslab = rows8 / ADMM_BK;
    const int total = iters * nslab;

    auto load_a = [&](int stage, int k0) {  // Hr[:, k0 : k0+8] -> As[stage]; rows / columns >= K zero-filled
        for (int q = threadIdx.x; q < rows8 * 4; q += THREADS) {
            const int row = q >> 2, cc = (q & 3) * 2;
            const bool ok = row < K && k0 + cc < K;
            const double* src = ok ? Hr + (int64_t)row * ldh + k0 + cc : Hr;
            cp_async16(As + stage * astage + row * ADMM_ALD + cc, src, ok ? 16 : 0);
        }
    };

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t n0 = tile * NV;
        // ---- Hr ring prologue (overlaps the tile load below)
#pragma unroll
        for (int s = 0; s < ADMM_STAGES - 1; ++s) {
            if (s < total) load_a(s, (s % nslab) * ADMM_BK);
            cp_async_commit();
        }
        // ---- tile load: D = Z - U and U, rows8 x 3 planes x NV vertices, 16-byte chunks (128 B runs per row for WN = 2)
        for (int t = threadIdx.x; t < rows8 * 3 * (NV / 2); t += THREADS) {
            const int cc = t % (NV / 2), kd = t / (NV / 2), d = kd % 3, k = kd / 3;
            const int v = cc * 2, g = v >> 3, c = v & 7;
            double2 dz = make_double2(0.0, 0.0), du = dz;
            if (k < K) {
                const int64_t off = (int64_t)k * C3 + (int64_t)d * ld + n0 + v;
                const double2 z = ld2(Z + off);
                du = ld2(U + off);
                dz.x = z.x - du.x;
                dz.y = z.y - du.y;
            }
            st2(Ds + k * DLD + g * 24 + d * 8 + c, dz);
            st2(Us + k * DLD + g * 24 + d * 8 + c, du);
        }

        int g = 0;  // running slab counter of this tile (ring position)
        for (int it = 0; it < iters; ++it) {
            double acc[MT][3][2];
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
                for (int d = 0; d < 3; ++d) acc[i][d][0] = acc[i][d][1] = 0.0;

            for (int s = 0; s < nslab; ++s, ++g) {
                cp_async_wait<ADMM_STAGES - 2>();
                __syncthreads();  // slab g landed; stage (g-1)%3 free; Ds/Us writes of the previous epilogue visible
                const int nx = g + ADMM_STAGES - 1;
                if (nx < total) load_a(nx % ADMM_STAGES, (nx % nslab) * ADMM_BK);
                cp_async_commit();
                const double* as = As + (g % ADMM_STAGES) * astage + (mt0 * 8 + lr) * ADMM_ALD + lc;
                const double* ds = Ds + (s * ADMM_BK + lc) * DLD + wn * 24 + lr;
#pragma unroll
                for (int kk = 0; kk < ADMM_BK / 4; ++kk) {
                    double b[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d) b[d] = ds[kk * 4 * DLD + d * 8];
#pragma unroll
                    for (int i = 0; i < MT; ++i) {
                        if (i < mtn) {  // warp-uniform
                            const double a = as[i * 8 * ADMM_ALD + kk * 4];
#pragma unroll
                            for (int d = 0; d < 3; ++d) dmma(acc[i][d][0], acc[i][d][1], a, b[d]);
                        }
                    }
                }
            }
            __syncthreads();  // every warp has finished reading Ds for this iteration

            // ---- epilogue: prox + dual update on the thread's own (k, 2 vertices, xyz) entries
            const bool last = it == iters - 1;
            const int64_t n = n0 + wn * 8 + 2 * lc;
            const bool ok0 = n < N, ok1 = n + 1 < N;
#pragma unroll
            for (int i = 0; i < MT; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                if (i >= mtn || k >= K) continue;
                const int64_t base = (int64_t)k * C3 + n;
                double* us = Us + k * DLD + wn * 24 + 2 * lc;
                double* dsw = Ds + k * DLD + wn * 24 + 2 * lc;
                double2 v[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    const double2 b0 = ld2(B0 + base + (int64_t)d * ld);
                    const double2 u = ld2(us + d * 8);
                    v[d].x = (b0.x + acc[i][d][0]) + u.x;  // V = C + U
                    v[d].y = (b0.y + acc[i][d][1]) + u.y;
                }
                const double2 lam = ld2(Lambda + (int64_t)k * ld + n);
                double sh[2];
                {
                    const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(v[0].x, v[0].x), __dmul_rn(v[1].x, v[1].x)),
                                                      __dmul_rn(v[2].x, v[2].x)));
                    const double t = 1.0 - __ddiv_rn(__dmul_rn(inv_rho, lam.x), len);
                    sh[0] = (t != t) ? t : fmax(0.0, t);  // np.maximum propagates the 0/0 NaN (posComponents.py:253-256)
                }
                {
                    const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(v[0].y, v[0].y), __dmul_rn(v[1].y, v[1].y)),
                                                      __dmul_rn(v[2].y, v[2].y)));
                    const double t = 1.0 - __ddiv_rn(__dmul_rn(inv_rho, lam.y), len);
                    sh[1] = (t != t) ? t : fmax(0.0, t);
                }
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    double2 z, u2;
                    z.x = ok0 ? v[d].x * sh[0] : 0.0;
                    z.y = ok1 ? v[d].y * sh[1] : 0.0;
                    u2.x = ok0 ? v[d].x - z.x : 0.0;
                    u2.y = ok1 ? v[d].y - z.y : 0.0;
                    if (last) {
                        st2(Z + base + (int64_t)d * ld, z);
                        st2(U + base + (int64_t)d * ld, u2);
                    } else {
                        st2(us + d * 8, u2);
                        st2(dsw + d * 8, make_double2(z.x - u2.x, z.y - u2.y));
                    }
                }
            }
        }
        cp_async_wait<0>();
        __syncthreads();  // the next tile overwrites Ds / Us / As
    }
}

template <int WM, int WN, int MT>
static int launch_admm(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                       int64_t K, int64_t N, int64_t ld, double inv_rho, int iters, cudaStream_t st) {
    const int rows8 = (int)((K + 7) & ~7LL);
    const size_t smem = (size_t)rows8 * (2 * (24 * WN + 4) + ADMM_STAGES * ADMM_ALD) * sizeof(double);
    ASB_CUDA(cudaFuncSetAttribute(admm_solve_kernel<WM, WN, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = ld / (8 * WN);
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, sm_count());
    admm_solve_kernel<WM, WN, MT><<<grid, WM * WN * 32, smem, st>>>(Hr, ldh, B0, Z, U, Lambda, (int)K, N, ld, inv_rho, iters,
                                                                   ntiles);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_admm_solve(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                   int64_t N, int64_t ld, double inv_rho, int64_t iters, asb_stream_t stream) {
    ASB_REQUIRE(Hr && B0 && Z && U && Lambda && K > 0 && K <= 256 && N > 0 && ld >= N && ld % 32 == 0,
                "admm_solve: bad arguments (K <= 256)");
    ASB_REQUIRE(ldh % 2 == 0 && ((uintptr_t)Hr % 16 == 0), "admm_solve: Hr must be 16-byte aligned with even ldh");
    ASB_REQUIRE(iters >= 0 && iters < (1 << 20), "admm_solve: bad iteration count");
    if (iters == 0) return ASB_OK;
    const cudaStream_t st = S(stream);
    const int it = (int)iters;
    // 16-vertex tiles while two K x 52 tiles + the Hr ring fit in 227 KB of shared memory (K <= 200), else 8-vertex tiles
    if (K <= 64) return launch_admm<4, 2, 2>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, it, st);
    if (K <= 128) return launch_admm<4, 2, 4>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, it, st);
    if (K <= 200) return launch_admm<4, 2, 7>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, it, st);
    return launch_admm<8, 1, 4>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, it, st);
}

int asb_admm_iter(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                  int64_t N, int64_t ld, double inv_rho, asb_stream_t stream) {
    return asb_admm_solve(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, 1, stream);
}

}  // extern "C"
