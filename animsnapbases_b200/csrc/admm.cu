// SPLOCS ADMM inner loop (snapbases/posComponents.py:175-178, prox :252-256), all iterations of one outer step fused.
//
//   repeat iters times:   C = B0 + Hr (Z - U);   V = C + U;   Z = V * max(0, 1 - Lambda / (rho |V|_3));   U = V - Z
//
// The iteration couples the K components of ONE vertex only (Hr is K x K, the prox acts on the 3-vector of a (k, n)
// pair), so a CTA owns a tile of 16 vertices (x 3 planes = 48 columns) for the whole inner loop: D = Z - U lives in
// shared memory (the DMMA right operand), the K x 48 accumulators AND the dual U in registers, and Z, U travel
// HBM <-> SM exactly once per outer step instead of once per ADMM iteration.  What is left per iteration is the
// K x K x 48 DMMA product, whose left operand Hr (320 KB at K = 200, does not fit beside the tile) streams L2 -> shared
// through a 3-stage cp.async ring (16-column slabs) that keeps running across iterations.  B0 and Lambda of the tile are
// re-read in the epilogue (L2 hits: the 148 resident tiles hold 15 MB).  Persistent grid: one CTA per SM.
//
// Warp layout: 8 warps = 4 row-warps x 2 vertex groups, warp id = wm * 2 + wn.  The ceil(K/8) m-tiles are dealt to the
// row-warps MT or MT-1 each (K = 200: 7,6,6,6 -> SM sub-partition loads 13,13,12,12); the body is compiled for both
// counts and chosen by a warp-uniform branch (a predicated-off DMMA still occupies the tensor pipe).
#include "dmma.cuh"

namespace asb {

// Two small CTAs per SM (4 row-warps x 1 vertex group each) instead of one 8-warp CTA: the slab barriers, the FP64
// prox epilogue and the tile load/store of one CTA overlap with the DMMA stream of the other (ncu on the 8-warp version:
// 14 % of the warp samples sat behind BAR.SYNC, DMMA pipe 61 % active).
constexpr int ADMM_WM = 4, ADMM_WN = 1, ADMM_THREADS = ADMM_WM * ADMM_WN * 32, ADMM_CTAS_PER_SM = 2;
constexpr int ADMM_BK = 8;             // k-slab of Hr per pipeline stage
constexpr int ADMM_ALD = ADMM_BK + 4;  // 12 doubles: conflict-free 8x4 A-fragment reads
constexpr int ADMM_STAGES = 3;
constexpr int ADMM_CPR = ADMM_BK / 2;              // 16-byte chunks per staged row
constexpr int ADMM_RPP = ADMM_THREADS / ADMM_CPR;  // rows staged per pass
constexpr int ADMM_NV = 8 * ADMM_WN;   // vertices per tile
constexpr int ADMM_NC = 24 * ADMM_WN;  // columns per tile: [vertex group][plane][8 vertices]
constexpr int ADMM_DLD = ADMM_NC + 4;  // 28 / 52: stride mod 16 in {12, 4} -> conflict-free 4x8 B-fragment reads

template <int MT, int MTN>
__device__ __forceinline__ void admm_body(double* smem, const double* __restrict__ Hr, int64_t ldh,
                                          const double* __restrict__ B0, double* __restrict__ Z, double* __restrict__ U,
                                          const double* __restrict__ Lambda, int K, int64_t N, int64_t ld, double inv_rho,
                                          int iters, int64_t ntiles, int mt0, bool live) {
    const int rows8 = (K + 7) & ~7;   // rows held in shared memory (rows >= K are zero)
    double* Ds = smem;                                  // [rows8][ADMM_DLD]  Z - U
    double* As = Ds + (size_t)rows8 * ADMM_DLD;         // ADMM_STAGES x [rows8][ADMM_ALD]
    const int astage = rows8 * ADMM_ALD;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int wn = warp % ADMM_WN;
    const int64_t C3 = 3 * ld;
    const int nslab = (K + ADMM_BK - 1) / ADMM_BK;
    const int total = iters * nslab;

    // Hr slab stager: chunk t of a thread = row tid / CPR + RPP t, doubles (tid % CPR) * 2 .. +1 of the slab
    constexpr int NCH = (32 * MT + ADMM_RPP - 1) / ADMM_RPP;
    const int srow = threadIdx.x / ADMM_CPR, scc = (threadIdx.x % ADMM_CPR) * 2;
    unsigned okmask = 0;
#pragma unroll
    for (int t = 0; t < NCH; ++t)
        if (srow + ADMM_RPP * t < K) okmask |= 1u << t;
    const double* hp = Hr + (int64_t)min(srow, K - 1) * ldh + scc;
    const int64_t hstep = ADMM_RPP * ldh;
    const unsigned sdst = srow * ADMM_ALD + scc;
    auto load_a = [&](int stage, int slab) {  // Hr[:, BK slab : BK slab + BK] -> As[stage]; rows / columns >= K zero-filled
        const int k0 = slab * ADMM_BK;
        const bool kin = k0 + scc < K;
        const double* q = hp + k0;
        double* dst = As + stage * astage + sdst;
#pragma unroll
        for (int t = 0; t < NCH; ++t) {
            if (srow + ADMM_RPP * t < rows8) {
                const bool ok = kin && ((okmask >> t) & 1u);
                cp_async16(dst + t * ADMM_RPP * ADMM_ALD, ok ? q : Hr, ok ? 16 : 0);
            }
            q += hstep;
        }
    };
    const double* afrag = As + (mt0 * 8 + lr) * ADMM_ALD + lc;
    const double* bfrag = Ds + lc * ADMM_DLD + wn * 24 + lr;
    double* down = Ds + wn * 24 + 2 * lc;   // + k * ADMM_DLD + d * 8: the thread's own entries of D

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t n = tile * ADMM_NV + wn * 8 + 2 * lc;   // the thread's two vertices
        const bool ok0 = n < N, ok1 = n + 1 < N;
        // ---- Hr ring prologue (overlaps the tile load below)
        int gl = 0, ls = 0, lslab = 0;   // slabs issued, load stage, slab index of the next load
#pragma unroll
        for (int s = 0; s < ADMM_STAGES - 1; ++s) {
            if (gl < total) {
                load_a(ls, lslab);
                ++gl;
                ls = ls + 1 == ADMM_STAGES ? 0 : ls + 1;
                lslab = lslab + 1 == nslab ? 0 : lslab + 1;
            }
            cp_async_commit();
        }
        // ---- tile load: every thread fetches its own (k, 2 vertices, xyz) entries; U stays in registers
        double u[MTN][3][2];
#pragma unroll
        for (int i = 0; i < MTN; ++i) {
            const int k = (mt0 + i) * 8 + lr;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                double2 z = make_double2(0.0, 0.0), uu = z;
                if (live && k < K) {
                    const int64_t off = (int64_t)k * C3 + (int64_t)d * ld + n;
                    z = ld2(Z + off);
                    uu = ld2(U + off);
                }
                u[i][d][0] = uu.x;
                u[i][d][1] = uu.y;
                if (live && k < rows8) st2(down + k * ADMM_DLD + d * 8, make_double2(z.x - uu.x, z.y - uu.y));
            }
        }

        int cs = 0;
        for (int it = 0; it < iters; ++it) {
            // accumulators start at B0 (C = B0 + Hr D): the L2 loads are all in flight before the first slab is consumed
            double acc[MTN][3][2];
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    double2 b0 = make_double2(0.0, 0.0);
                    if (live && k < K) b0 = ld2(B0 + (int64_t)k * C3 + (int64_t)d * ld + n);
                    acc[i][d][0] = b0.x;
                    acc[i][d][1] = b0.y;
                }
            }

            for (int s = 0; s < nslab; ++s) {
                cp_async_wait<ADMM_STAGES - 2>();
                __syncthreads();  // slab landed; previous stage free; D writes of the previous epilogue / tile load visible
                if (gl < total) {
                    load_a(ls, lslab);
                    ++gl;
                    ls = ls + 1 == ADMM_STAGES ? 0 : ls + 1;
                    lslab = lslab + 1 == nslab ? 0 : lslab + 1;
                }
                cp_async_commit();
                const double* as = afrag + cs * astage;
                const double* ds = bfrag + s * ADMM_BK * ADMM_DLD;
                cs = cs + 1 == ADMM_STAGES ? 0 : cs + 1;
#pragma unroll
                for (int kk = 0; kk < ADMM_BK / 4; ++kk) {
                    if (s * ADMM_BK + kk * 4 < rows8) {   // the last slab of K = 200 has 8 live columns
                        double b[3];
#pragma unroll
                        for (int d = 0; d < 3; ++d) b[d] = ds[kk * 4 * ADMM_DLD + d * 8];
#pragma unroll
                        for (int i = 0; i < MTN; ++i) {
                            const double a = as[i * 8 * ADMM_ALD + kk * 4];
#pragma unroll
                            for (int d = 0; d < 3; ++d) dmma(acc[i][d][0], acc[i][d][1], a, b[d]);
                        }
                    }
                }
            }
            __syncthreads();  // every warp has finished reading D for this iteration

            // ---- epilogue: prox + dual update on the thread's own entries
            const bool last = it == iters - 1;
            double2 lam[MTN];
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                lam[i] = (live && k < K) ? ld2(Lambda + (int64_t)k * ld + n) : make_double2(0.0, 0.0);
            }
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                if (!live || k >= K) continue;
                const int64_t base = (int64_t)k * C3 + n;
                double2 v[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    v[d].x = acc[i][d][0] + u[i][d][0];  // V = C + U
                    v[d].y = acc[i][d][1] + u[i][d][1];
                }
                // shrink = max(0, 1 - (lambda / rho) / |v|): NaN where lambda = 0 and v = 0 (0 * inf), like the reference's 0/0
                // (np.maximum propagates it, posComponents.py:253-256); 0 where v = 0 and lambda > 0 (1 - inf).
                double sh[2];
                {
                    const double s2 = __dadd_rn(__dadd_rn(__dmul_rn(v[0].x, v[0].x), __dmul_rn(v[1].x, v[1].x)),
                                                __dmul_rn(v[2].x, v[2].x));
                    const double t = 1.0 - __dmul_rn(__dmul_rn(inv_rho, lam[i].x), rsqrt(s2));
                    sh[0] = (t != t) ? t : fmax(0.0, t);
                }
                {
                    const double s2 = __dadd_rn(__dadd_rn(__dmul_rn(v[0].y, v[0].y), __dmul_rn(v[1].y, v[1].y)),
                                                __dmul_rn(v[2].y, v[2].y));
                    const double t = 1.0 - __dmul_rn(__dmul_rn(inv_rho, lam[i].y), rsqrt(s2));
                    sh[1] = (t != t) ? t : fmax(0.0, t);
                }
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    double2 z, u2;
                    z.x = ok0 ? v[d].x * sh[0] : 0.0;
                    z.y = ok1 ? v[d].y * sh[1] : 0.0;
                    u2.x = ok0 ? v[d].x - z.x : 0.0;
                    u2.y = ok1 ? v[d].y - z.y : 0.0;
                    u[i][d][0] = u2.x;
                    u[i][d][1] = u2.y;
                    if (last) {
                        st2(Z + base + (int64_t)d * ld, z);
                        st2(U + base + (int64_t)d * ld, u2);
                    } else {
                        st2(down + k * ADMM_DLD + d * 8, make_double2(z.x - u2.x, z.y - u2.y));
                    }
                }
            }
        }
        cp_async_wait<0>();
        __syncthreads();  // the next tile overwrites D and the Hr ring
    }
}

template <int MT>
__global__ void __launch_bounds__(ADMM_THREADS, ADMM_CTAS_PER_SM)
admm_solve_kernel(const double* __restrict__ Hr, int64_t ldh, const double* __restrict__ B0, double* __restrict__ Z,
                  double* __restrict__ U, const double* __restrict__ Lambda, int K, int64_t N, int64_t ld,
                  double inv_rho, int iters, int64_t ntiles) {
    extern __shared__ __align__(16) double smem[];
    int mt0, mtn;
    // co-resident CTAs (blockIdx.x and blockIdx.x + #SMs under round-robin placement) rotate the row-warp roles by two
    // sub-partitions, so the 7-tile warps of the two CTAs sit on different SM sub-partitions (13,12,13,12 instead of 14,12,12,12)
    const int slot = (blockIdx.x / max(1u, gridDim.x / ADMM_CTAS_PER_SM)) & 1;
    const int wm = (((threadIdx.x >> 5) / ADMM_WN) + 2 * slot) % ADMM_WM;
    mtile_range(((K + 7) & ~7) >> 3, ADMM_WM, wm, mt0, mtn);
    const bool live = mtn > 0;   // K <= 24: some row-warps own no m-tile; they idle on tile 0 and never store
    if (!live) mt0 = 0;
    if (mtn == MT) admm_body<MT, MT>(smem, Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, ntiles, mt0, live);
    else admm_body<MT, (MT > 1 ? MT - 1 : 1)>(smem, Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, ntiles, mt0, live);
}

template <int MT>
static int launch_admm(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                       int64_t K, int64_t N, int64_t ld, double inv_rho, int iters, cudaStream_t st) {
    const int rows8 = (int)((K + 7) & ~7LL);
    const size_t smem = (size_t)rows8 * (ADMM_DLD + ADMM_STAGES * ADMM_ALD) * sizeof(double);
    ASB_CUDA(cudaFuncSetAttribute(admm_solve_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = ld / ADMM_NV;
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)ADMM_CTAS_PER_SM * sm_count());
    admm_solve_kernel<MT><<<grid, ADMM_THREADS, smem, st>>>(Hr, ldh, B0, Z, U, Lambda, (int)K, N, ld, inv_rho, iters, ntiles);
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
    switch ((int)cdiv(cdiv(K, 8), ADMM_WM)) {  // m-tiles per row-warp
#define ASB_ADMM_CASE(MT) case MT: return launch_admm<MT>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, it, st);
        ASB_ADMM_CASE(1) ASB_ADMM_CASE(2) ASB_ADMM_CASE(3) ASB_ADMM_CASE(4) ASB_ADMM_CASE(5) ASB_ADMM_CASE(6) ASB_ADMM_CASE(7)
        default: return launch_admm<8>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, it, st);
#undef ASB_ADMM_CASE
    }
}

int asb_admm_iter(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                  int64_t N, int64_t ld, double inv_rho, asb_stream_t stream) {
    return asb_admm_solve(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, 1, stream);
}

}  // extern "C"
