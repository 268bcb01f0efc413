// SPLOCS ADMM inner loop (snapbases/posComponents.py:175-178, prox :252-256), all iterations of one outer step fused.
//
//   repeat iters times:   C = B0 + Hr (Z - U);   V = C + U;   Z = V * max(0, 1 - Lambda / (rho |V|_3));   U = V - Z
//
// The iteration couples the K components of ONE vertex only (Hr is K x K, the prox acts on the 3-vector of a (k, n)
// pair), so a CTA owns a tile of vertices (x 3 planes) for the whole inner loop: D = Z - U lives in shared memory (the
// DMMA right operand), the K x columns accumulators in registers, and Z travels HBM <-> SM exactly once per outer step
// instead of once per ADMM iteration.  What is left per iteration is the K x K x columns DMMA product, whose left operand
// Hr (320 KB at K = 200, does not fit beside the tile) streams L2 -> shared through a 3-stage cp.async ring (8-column
// slabs) that keeps running across iterations.  Persistent grid.
//
// Layouts of the same body (template AdmmCfg), chosen by ASB_ADMM_LAYOUT; measured on B200 at 1M x K = 200, 10 iterations
// per launch (profiles/r01_admm_layouts.md), cuBLAS DGEMM peak 35.5 TFLOP/s:
//   4    Pair    two CTAs per SM, 4 row-warps x 1 vertex group each (8 vertices, 24 columns), U in registers (254 regs),
//                8-column slabs, 3 stages                                           100.2 ms  24.0 TFLOP/s   <- default
//   416  Pair16  same, 16-column slabs (half the barriers), 2 stages                 109.4 ms
//   12   Wide    one CTA per SM, 12 warps = 4 row-warps x 3 vertex groups (24 vertices, 72 columns); no room for U in
//                registers (168 per thread): U goes through L2 every iteration       108.9 ms
//   1216 Wide16  same, 16-column slabs                                               112.5 ms
//   16   Tall    one CTA per SM, 16 warps = 8 row-warps x 2 vertex groups (<= 4 m-tiles per warp, 128 registers),
//                U through L2                                                        105.2 ms
//   8/81 Duo     two CTAs per SM of 8 row-warps x 1 vertex group (U through L2 / in registers)  115.7 / 106.8 ms
//   5    TmaPair same tile as Pair; Hr streamed from a slab-major copy with ONE cp.async.bulk (TMA, SASS UBLKCP) per slab,
//                handed over through mbarriers (no CTA barrier per slab), ring running on across tiles    97.0 ms  24.7 TFLOP/s
//                <- default (r02); bit-identical to layout 4 on the same inputs (tools/admm_ab.py, profiles/r02_admm_layouts.md)
//   6    TmaDuet one CTA per SM, two groups with an optional phase offset (ASB_ADMM_SKEW 0 / 1 / 2)       98.7 / 98.8 / 132.5 ms
//   7    TmaTrio one CTA per SM, three groups (12 warps), U through L2, 4-column slabs                     111.1 ms
//        (r02 measurements and what they rule out: profiles/r02_admm_layouts.md)
// ASB_ADMM_LOCK=1 adds a per-SM epilogue mutex to the Pair layouts (keeps the two CTAs of an SM half a phase apart so one
// runs the FP64 prox while the other owns the tensor pipe): 103.5 ms (4) / 104.6 ms (416) -- the two global atomics per
// iteration cost what the overlap wins, so it is off by default.
//
// Warp layout: warp id = wm * WN + wn.  The ceil(K/8) m-tiles are dealt to the 4 row-warps MT or MT-1 each (K = 200:
// 7,6,6,6); the body is compiled for both counts and chosen by a warp-uniform branch (a predicated-off DMMA still
// occupies the tensor pipe).  Hr must be zero-padded by the caller to ceil8(K) x ceil8(K): the slab stager and the k-loop
// run without bounds tests (ncu r01 before: 0.66 ISETP and 0.4 LDC per DMMA in the inner loop).
#include "dmma.cuh"
#include <cstdlib>
#ifndef ASB_ADMM_LAYOUT_DEFAULT
#define ASB_ADMM_LAYOUT_DEFAULT 5
#endif

namespace asb {

template <int WN_, bool UREG_, int CTAS_, int BK_, int STAGES_, int WM_ = 4>
struct AdmmCfg {
    static constexpr int WM = WM_;                    // row-warps: the ceil(K/8) m-tiles are dealt to them
    static constexpr int WN = WN_;                    // vertex groups (8 vertices each) per CTA
    static constexpr bool UREG = UREG_;               // dual U in registers (else through L2 every iteration)
    static constexpr int CTAS = CTAS_;                // CTAs per SM
    static constexpr int BK = BK_;                    // k-slab of Hr per pipeline stage (one barrier per slab)
    static constexpr int STAGES = STAGES_;            // cp.async ring depth
    static constexpr int ALD = BK + 4;                // 12 / 20 doubles: conflict-free 8x4 A-fragment reads
    static constexpr int CPR = BK / 2;                // 16-byte chunks per staged row
    static constexpr int THREADS = WM * WN * 32;
    static constexpr int RPP = THREADS / CPR;         // Hr rows staged per pass of the CTA
    static constexpr int NV = 8 * WN;                 // vertices per tile
    static constexpr int NC = 24 * WN;                // columns per tile: [vertex group][plane][8 vertices]
    static constexpr int DLD = NC + 4;                // 28 / 76: stride mod 16 == 12 -> conflict-free 4x8 B-fragment reads
};
using AdmmPair = AdmmCfg<1, true, 2, 8, 3>;      // 102 KB per CTA at K = 200
using AdmmPair16 = AdmmCfg<1, true, 2, 16, 2>;   // 109 KB: half the barriers, one stage less
using AdmmWide = AdmmCfg<3, false, 1, 8, 3>;     // 179 KB
using AdmmWide16 = AdmmCfg<3, false, 1, 16, 3>;  // 218 KB
using AdmmTall = AdmmCfg<2, false, 1, 8, 3, 8>;  // 141 KB: 8 row-warps x 2 vertex groups = 16 warps, <= 4 m-tiles per warp
using AdmmDuo = AdmmCfg<1, false, 2, 8, 3, 8>;   // 102 KB: two CTAs per SM of 8 row-warps x 1 vertex group (16 warps per SM in
                                                 // two independent CTAs: one keeps 2 warps per sub-partition on the tensor pipe
                                                 // while the other is in its epilogue), U through L2, 128 registers
using AdmmDuoU = AdmmCfg<1, true, 2, 8, 3, 8>;   // same with U in registers

// Epilogue mutex, one per SM (Pair layout).  Two identical co-resident CTAs that share the DMMA pipe fall into lock-step
// (a CTA that is ahead runs its GEMM phase at half rate while the one behind catches up during the leader's epilogue:
// the skew halves every iteration), so both sit in the FP64 prox epilogue at the same time and the DMMA pipe idles.
// Letting only one CTA of an SM into the epilogue at a time keeps them half a phase apart.  Performance-only: the spin is
// bounded, a CTA that gives up simply proceeds; the host clears the array before every launch.
__device__ int g_admm_epilogue_lock[1024];

__device__ __forceinline__ unsigned smid() {
    unsigned r;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(r));
    return r;
}
__device__ __forceinline__ bool epilogue_lock_acquire(int* lock) {
    for (int spin = 0; spin < (1 << 14); ++spin) {
        if (atomicCAS(lock, 0, 1) == 0) return true;
        __nanosleep(64);
    }
    return false;
}

// shrink = max(0, 1 - (lambda / rho) / |v|): NaN where lambda = 0 and v = 0 (0 * inf), like the reference's 0/0
// (np.maximum propagates it, posComponents.py:253-256); 0 where v = 0 and lambda > 0 (1 - inf).
__device__ __forceinline__ double prox_shrink(double vx, double vy, double vz, double lam, double inv_rho) {
    const double s2 = __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
    const double t = 1.0 - __dmul_rn(__dmul_rn(inv_rho, lam), rsqrt(s2));
    return (t != t) ? t : fmax(0.0, t);
}

template <class Cfg, int MT, int MTN>
__device__ __forceinline__ void admm_body(double* smem, const double* __restrict__ Hr, int64_t ldh,
                                          const double* __restrict__ B0, double* __restrict__ Z, double* __restrict__ U,
                                          const double* __restrict__ Lambda, int K, int64_t N, int64_t ld, double inv_rho,
                                          int iters, int64_t ntiles, int mt0, bool live, int* lock) {
    constexpr int WN = Cfg::WN, DLD = Cfg::DLD, RPP = Cfg::RPP;
    constexpr int ADMM_BK = Cfg::BK, ADMM_ALD = Cfg::ALD, ADMM_STAGES = Cfg::STAGES, ADMM_CPR = Cfg::CPR;
    constexpr bool UREG = Cfg::UREG;
    constexpr int UR = UREG ? MTN : 1;   // register copy of U only in the Pair layout
    const int rows8 = (K + 7) & ~7;      // rows held in shared memory (rows >= K are zero)
    const int rowsk = (K + ADMM_BK - 1) / ADMM_BK * ADMM_BK;   // contraction length: whole slabs (Hr is padded to it)
    bool held = false;                   // thread 0: this CTA owns the SM's epilogue mutex
    double* Ds = smem;                                  // [rowsk][DLD]  Z - U (rows >= K zero)
    double* As = Ds + (size_t)rowsk * DLD;              // ADMM_STAGES x [rows8][ADMM_ALD]
    const int astage = rows8 * ADMM_ALD;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int wn = warp % WN;
    const int64_t C3 = 3 * ld;
    const int nslab = rowsk / ADMM_BK;
    const int total = iters * nslab;

    // Hr slab stager: chunk t of a thread = row tid / CPR + RPP t, doubles (tid % CPR) * 2 .. +1 of the slab.  Hr is padded
    // by the caller (zeros) to whole slabs, so no chunk needs a column test, and the NFULL passes that end below
    // 8 WM (MT - 1) < rows8 need no row test either; the remaining one or two passes use a mask computed once.
    constexpr int NCH = (8 * Cfg::WM * MT + RPP - 1) / RPP;
    constexpr int NFULL = (8 * Cfg::WM * (MT - 1)) / RPP;
    static_assert(NCH - NFULL <= 4, "tail passes");
    const int srow = threadIdx.x / ADMM_CPR, scc = (threadIdx.x % ADMM_CPR) * 2;
    unsigned tailmask = 0;
#pragma unroll
    for (int t = NFULL; t < NCH; ++t)
        if (srow + RPP * t < rows8) tailmask |= 1u << (t - NFULL);
    const double* hp = Hr + (int64_t)min(srow, rows8 - 1) * ldh + scc;
    const int64_t hstep = RPP * ldh;
    const unsigned sdst = srow * ADMM_ALD + scc;
    auto load_a = [&](int stage, int slab, bool valid) {  // Hr[:, BK slab : BK slab + BK] -> As[stage] (zero-fill if !valid)
        const double* q = hp + slab * ADMM_BK;
        double* dst = As + stage * astage + sdst;
        const int nbytes = valid ? 16 : 0;
#pragma unroll
        for (int t = 0; t < NFULL; ++t) {
            cp_async16(dst + t * RPP * ADMM_ALD, q, nbytes);
            q += hstep;
        }
#pragma unroll
        for (int t = NFULL; t < NCH; ++t) {
            if ((tailmask >> (t - NFULL)) & 1u) cp_async16(dst + t * RPP * ADMM_ALD, q, nbytes);
            q += hstep;
        }
    };
    // D rows [rows8, rowsk) (BK = 16, K % 16 in 1..8) are never written by a tile load or an epilogue: zero them once
    for (int e = rows8 * DLD + threadIdx.x; e < rowsk * DLD; e += Cfg::THREADS) Ds[e] = 0.0;
    const double* afrag = As + (mt0 * 8 + lr) * ADMM_ALD + lc;
    const double* bfrag = Ds + lc * DLD + wn * 24 + lr;
    double* down = Ds + wn * 24 + 2 * lc;   // + k * DLD + d * 8: the thread's own entries of D

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t n = tile * Cfg::NV + wn * 8 + 2 * lc;   // the thread's two vertices
        const bool ok0 = n < N, ok1 = n + 1 < N;
        const bool inb = live && n < ld;   // Wide: 24 does not divide ld, the last tile hangs over the padded plane
        // ---- Hr ring prologue (overlaps the tile load below)
        int gl = 0, ls = 0, lslab = 0;   // slabs issued, load stage, slab index of the next load
#pragma unroll
        for (int s = 0; s < ADMM_STAGES - 1; ++s) {
            load_a(ls, lslab, gl < total);
            ++gl;
            ls = ls + 1 == ADMM_STAGES ? 0 : ls + 1;
            lslab = lslab + 1 == nslab ? 0 : lslab + 1;
            cp_async_commit();
        }
        // ---- tile load: every thread fetches its own (k, 2 vertices, xyz) entries of Z and U
        double u[UR][3][2];
#pragma unroll
        for (int i = 0; i < MTN; ++i) {
            const int k = (mt0 + i) * 8 + lr;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                double2 z = make_double2(0.0, 0.0), uu = z;
                if (inb && k < K) {
                    const int64_t off = (int64_t)k * C3 + (int64_t)d * ld + n;
                    z = ld2(Z + off);
                    uu = ld2(U + off);
                }
                if (UREG) {
                    u[i % UR][d][0] = uu.x;
                    u[i % UR][d][1] = uu.y;
                }
                if (live && k < rows8) st2(down + k * DLD + d * 8, make_double2(z.x - uu.x, z.y - uu.y));   // k >= K: zeros
            }
        }

        int cs = 0;
        // accumulators start at B0 (C = B0 + Hr D).  First iteration: loaded here (in flight while the first slab lands);
        // later iterations: re-loaded in the epilogue as soon as an m-tile's accumulators have been consumed, so the L2
        // latency hides behind the rest of the epilogue instead of standing in front of the first DMMA.
        double acc[MTN][3][2];
        auto load_b0 = [&](int i) {
            const int k = (mt0 + i) * 8 + lr;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                double2 b0 = make_double2(0.0, 0.0);
                if (inb && k < K) b0 = ld2(B0 + (int64_t)k * C3 + (int64_t)d * ld + n);
                acc[i][d][0] = b0.x;
                acc[i][d][1] = b0.y;
            }
        };
#pragma unroll
        for (int i = 0; i < MTN; ++i) load_b0(i);
        for (int it = 0; it < iters; ++it) {
            for (int s = 0; s < nslab; ++s) {
                cp_async_wait<ADMM_STAGES - 2>();
                __syncthreads();  // slab landed; previous stage free; D writes of the previous epilogue / tile load visible
                load_a(ls, lslab, gl < total);   // past the end: a zero-fill of a stage nobody reads (no branch in the loop)
                ++gl;
                ls = ls + 1 == ADMM_STAGES ? 0 : ls + 1;
                lslab = lslab + 1 == nslab ? 0 : lslab + 1;
                cp_async_commit();
                const double* as = afrag + cs * astage;
                const double* ds = bfrag + s * ADMM_BK * DLD;
                cs = cs + 1 == ADMM_STAGES ? 0 : cs + 1;
#pragma unroll
                for (int kk = 0; kk < ADMM_BK / 4; ++kk) {   // whole slabs: Hr columns and D rows >= K are zero
                    double b[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d) b[d] = ds[kk * 4 * DLD + d * 8];
#pragma unroll
                    for (int i = 0; i < MTN; ++i) {
                        const double a = as[i * 8 * ADMM_ALD + kk * 4];
#pragma unroll
                        for (int d = 0; d < 3; ++d) dmma(acc[i][d][0], acc[i][d][1], a, b[d]);
                    }
                }
            }
            if (lock && threadIdx.x == 0) held = epilogue_lock_acquire(lock);
            __syncthreads();  // every warp has finished reading D for this iteration (and the epilogue mutex is ours)

            // ---- epilogue: prox + dual update on the thread's own entries.  Lambda of all m-tiles is fetched up front (one L2
            // latency); in the Wide layout the old dual U (L2 hits: written by this thread one iteration ago) runs UDEPTH
            // m-tiles ahead of the arithmetic.
            const bool last = it == iters - 1;
            double2 lam[MTN];
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                lam[i] = (inb && k < K) ? ld2(Lambda + (int64_t)k * ld + n) : make_double2(0.0, 0.0);
            }
            constexpr int UDMAX = Cfg::THREADS > 384 ? 2 : 3;   // 128-register cap at 512 threads
            constexpr int UDEPTH = UREG ? 1 : (MTN < UDMAX ? MTN : UDMAX);
            double2 uo[UDEPTH][3];
            auto fetch_u = [&](int i) {
                const int k = (mt0 + i) * 8 + lr;
                if (!inb || k >= K) return;
                const int64_t base = (int64_t)k * C3 + n;
#pragma unroll
                for (int d = 0; d < 3; ++d) uo[i % UDEPTH][d] = ld2(U + base + (int64_t)d * ld);
            };
            if (!UREG) {
#pragma unroll
                for (int i = 0; i < UDEPTH; ++i) fetch_u(i);
            }
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                const bool on = inb && k < K;
                const int64_t base = (int64_t)k * C3 + n;
                double2 v[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {   // V = C + U
                    v[d].x = acc[i][d][0] + (UREG ? u[i % UR][d][0] : uo[i % UDEPTH][d].x);
                    v[d].y = acc[i][d][1] + (UREG ? u[i % UR][d][1] : uo[i % UDEPTH][d].y);
                }
                if (!UREG && i + UDEPTH < MTN) fetch_u(i + UDEPTH);   // slot i % UDEPTH is free again
                if (!last) load_b0(i);                                // acc[i] is consumed: start the next iteration's C = B0 + ...
                if (on) {
                    const double sh0 = prox_shrink(v[0].x, v[1].x, v[2].x, lam[i].x, inv_rho);
                    const double sh1 = prox_shrink(v[0].y, v[1].y, v[2].y, lam[i].y, inv_rho);
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        double2 z, u2;
                        z.x = ok0 ? v[d].x * sh0 : 0.0;
                        z.y = ok1 ? v[d].y * sh1 : 0.0;
                        u2.x = ok0 ? v[d].x - z.x : 0.0;
                        u2.y = ok1 ? v[d].y - z.y : 0.0;
                        if (UREG) {
                            u[i % UR][d][0] = u2.x;
                            u[i % UR][d][1] = u2.y;
                        }
                        if (!UREG || last) st2(U + base + (int64_t)d * ld, u2);
                        if (last) st2(Z + base + (int64_t)d * ld, z);
                        else st2(down + k * DLD + d * 8, make_double2(z.x - u2.x, z.y - u2.y));
                    }
                }
            }
            if (lock) {
                __syncthreads();  // all warps are through the epilogue
                if (threadIdx.x == 0 && held) {
                    atomicExch(lock, 0);
                    held = false;
                }
            }
        }
        cp_async_wait<0>();
        __syncthreads();  // the next tile overwrites D and the Hr ring
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Bulk-copy (TMA) variant of the Pair layout.  Same tile, same fragments, same epilogue; what changes is how the left operand
// reaches shared memory and how the warps of a CTA meet:
//   * Hr is re-laid once per launch into slab-major order  Hpk[slab][row][12]  (8 columns + 4 padding doubles: the padded row
//     stride the A-fragment reads are conflict-free with), so a slab is ONE contiguous 19.2 KB block and ONE thread moves it
//     with `cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes` (SASS UBLKCP) instead of every thread issuing 7
//     16-byte cp.async per slab;
//   * slabs are handed over through mbarriers (full[stage]: armed with expect_tx by the issuing thread, completed by the copy
//     engine; empty[stage]: one arrival per warp), so there is NO CTA-wide barrier per slab: a warp waits for data, never for
//     its neighbours.  Two __syncthreads per ADMM iteration remain (D = Z - U is read by every warp and rewritten by the
//     epilogue);
//   * the ring never drains: Hr is the same for every vertex tile, so the slab stream runs on across iterations AND tiles
//     (the r01 kernel refilled its ring from empty at every tile).
// The refill of the stage that held slab g-1 is issued by thread 0 at the start of step g, when the other warps have had a
// whole slab of slack to release it.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    const uint32_t a = smem_u32(b);
    uint32_t ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}"
            : "=r"(ok)
            : "r"(a), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Shapes of the bulk-copy layouts.  G groups of 4 row-warps per CTA; BK columns of Hr per slab, staged [row][ALD]:
// ALD = 12 for BK = 8 and ALD = 4 for BK = 4 are the strides at which the 64-bit A-fragment loads of a half-warp
// (rows lr = 0..3, columns lc = 0..3) touch 16 distinct 8-byte slots.
template <int G_, bool UREG_, int BK_>
struct AdmmTmaCfg {
    static constexpr int G = G_;                       // groups (vertex tiles in flight) per CTA
    static constexpr bool UREG = UREG_;                // dual U in registers (else through L2 every iteration)
    static constexpr int BK = BK_;
    static constexpr int ALD = BK_ == 8 ? 12 : 4;
    static constexpr int STAGES = BK_ == 8 ? 3 : 4;
    static constexpr int WM = 4, NV = 8, NC = 24, DLD = 28;
    static constexpr int THREADS = 128 * G_;
    static constexpr int CTAS = G_ == 1 ? 2 : 1;       // CTAs per SM
};
using AdmmTmaPair = AdmmTmaCfg<1, true, 8>;    // 5: two CTAs per SM, 4 warps each, U in registers (255 registers)
using AdmmTmaDuet = AdmmTmaCfg<2, true, 8>;    // 6: one CTA per SM, two groups that can hold a phase offset
using AdmmTmaTrio = AdmmTmaCfg<3, false, 4>;   // 7: one CTA per SM, THREE free-running groups (12 warps, <= 168 registers, U
                                               //    through L2): a lone warp per SM sub-partition issues one DMMA per ~37
                                               //    cycles, two reach ~18 -- with three groups two are always in their GEMM
                                               //    phase while the third runs its FP64 prox epilogue
struct AdmmTma {                               // what the host side needs without a layout
    static constexpr int NV = 8;
};

// Hpk[slab][row][0..ALD): columns BK slab .. BK slab + BK - 1 of Hr row `row`, then zeros
__global__ void admm_pack_kernel(const double* __restrict__ Hr, int64_t ldh, int rows8, int nslab, int bk, int ald,
                                 double* __restrict__ Hpk) {
    const int total = nslab * rows8 * ald;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        const int c = t % ald, r = (t / ald) % rows8, sl = t / (ald * rows8);
        Hpk[t] = c < bk ? Hr[(int64_t)r * ldh + sl * bk + c] : 0.0;
    }
}

// G = 1: the CTA is one group of 4 row-warps (two CTAs per SM).  G >= 2: ONE CTA per SM holds G independent groups (warps
// 4g .. 4g + 3), each with its own vertex tile, D, slab ring and named barrier -- the same work as G co-resident CTAs.  With
// G = 2 the groups can additionally see each other's progress through two mbarriers and hold a fixed phase offset (`skew`):
//   skew 0   free running;
//   skew 1   a group may start iteration i only when its partner is half way through the slabs of its own current iteration;
//   skew 2   strict alternation of the GEMM phases (one group on the pipe at a time).
// Measured on B200 at 1M x K = 200, 10 iterations (profiles/r02_admm_layouts.md): skew changes nothing (98.7 / 98.8 ms) and
// strict alternation costs 34 %, because ONE group cannot feed the pipe (a 4-warp CTA alone on the SM: 147.7 ms, 37 cycles
// per DMMA and sub-partition): what the epilogue of one group needs beside it is TWO groups in their GEMM phase -> G = 3.
template <class Cfg, int MT, int MTN>
__device__ __forceinline__ void admm_tma_body(double* smem_all, const double* __restrict__ Hpk, const double* __restrict__ B0,
                                              double* __restrict__ Z, double* __restrict__ U, const double* __restrict__ Lambda,
                                              int K, int64_t N, int64_t ld, double inv_rho, int iters, int64_t ntiles, int mt0,
                                              bool live, int skew, int zero) {
    constexpr int G = Cfg::G, DLD = Cfg::DLD, ALD = Cfg::ALD, STAGES = Cfg::STAGES, BK = Cfg::BK;
    constexpr bool UREG = Cfg::UREG;
    constexpr int UR = UREG ? MTN : 1;
    const int rows8 = (K + 7) & ~7;
    const int nslab = rows8 / BK;                         // rows8 is a multiple of 8, BK is 4 or 8
    const int astage = rows8 * ALD;
    const uint32_t slab_bytes = (uint32_t)astage * 8u;
    const int grp = G > 1 ? (int)(threadIdx.x >> 7) : 0;
    const int gt = (int)threadIdx.x - 128 * grp;          // thread index inside the group
    const size_t group_doubles = (size_t)rows8 * DLD + (size_t)STAGES * astage;
    double* Ds = smem_all + grp * group_doubles;          // [rows8][DLD]  Z - U (rows >= K zero)
    double* As = Ds + (size_t)rows8 * DLD;                // STAGES x [rows8][ALD]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_all + G * group_doubles);
    uint64_t* full = bars + grp * 2 * STAGES;
    uint64_t* empty = full + STAGES;
    uint64_t* mid = bars + G * 2 * STAGES;                // mid[g]: progress signal of group g (G == 2)
    const int lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int64_t C3 = 3 * ld;
    auto tiles_of = [&](int g) -> int64_t {
        const int64_t first = (int64_t)blockIdx.x * G + g, step = (int64_t)gridDim.x * G;
        return first < ntiles ? (ntiles - first + step - 1) / step : 0;
    };
    const int64_t my_tiles = tiles_of(grp);
    const int64_t total_all = my_tiles * iters * nslab;   // slabs this group consumes over its whole life
    const int64_t partner_iters = G == 2 ? tiles_of(grp ^ 1) * iters : 0;
    auto group_sync = [&]() {
        if (G > 1) asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        else __syncthreads();
    };
    if (threadIdx.x == 0) {
        for (int i = 0; i < G * STAGES; ++i) {
            mbar_init(bars + (i / STAGES) * 2 * STAGES + (i % STAGES), 1);                    // full
            mbar_init(bars + (i / STAGES) * 2 * STAGES + STAGES + (i % STAGES), Cfg::WM);      // empty
        }
        if (G == 2) {
            mbar_init(mid + 0, Cfg::WM);
            mbar_init(mid + 1, Cfg::WM);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // producer state (thread 0 of the group): next slab to issue
    int64_t issued = 0;
    int pst = 0, pslab = 0;
    uint32_t pwrap = 0;
    auto issue = [&]() {   // slab `issued` -> stage pst; from the second lap on the stage must have been released
        if (issued >= STAGES) mbar_wait(empty + pst, pwrap ^ 1u);
        mbar_arrive_expect_tx(full + pst, slab_bytes);
        bulk_g2s(As + (size_t)pst * astage, Hpk + (size_t)pslab * astage, slab_bytes, full + pst);
        ++issued;
        if (++pst == STAGES) { pst = 0; pwrap ^= 1u; }
        if (++pslab == nslab) pslab = 0;
    };
    if (gt == 0)
        for (int i = 0; i < STAGES && issued < total_all; ++i) issue();
    // consumer state: next slab to use
    int cst = 0;
    uint32_t cwrap = 0;
    bool first = true;   // no refill before the very first step (the prologue filled every stage)
    int64_t git = 0;     // iterations this group has started (over all its tiles)
    const int sig_slab = skew == 2 ? nslab - 1 : (nslab - 1) / 2;   // the slab after which the group signals its progress
    const double* afrag = As + (mt0 * 8 + lr) * ALD + lc;
    const double* bfrag = Ds + lc * DLD + lr;
    double* down = Ds + 2 * lc;   // + k * DLD + d * 8: the thread's own entries of D

    for (int64_t tile = (int64_t)blockIdx.x * G + grp; tile < ntiles; tile += (int64_t)gridDim.x * G) {
        const int64_t n = tile * Cfg::NV + 2 * lc;   // the thread's two vertices
        const bool ok0 = n < N, ok1 = n + 1 < N;
        const bool inb = live && n < ld;
        // ---- tile load: every thread fetches its own (k, 2 vertices, xyz) entries of Z and U
        double u[UR][3][2];
#pragma unroll
        for (int i = 0; i < MTN; ++i) {
            const int k = (mt0 + i) * 8 + lr;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                double2 z = make_double2(0.0, 0.0), uu = z;
                if (inb && k < K) {
                    const int64_t off = (int64_t)k * C3 + (int64_t)d * ld + n;
                    z = ld2(Z + off);
                    uu = ld2(U + off);
                }
                if (UREG) {
                    u[i % UR][d][0] = uu.x;
                    u[i % UR][d][1] = uu.y;
                }
                if (live && k < rows8) st2(down + k * DLD + d * 8, make_double2(z.x - uu.x, z.y - uu.y));   // k >= K: zeros
            }
        }
        double acc[MTN][3][2];
        auto load_b0 = [&](int i) {
            const int k = (mt0 + i) * 8 + lr;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                double2 b0 = make_double2(0.0, 0.0);
                if (inb && k < K) b0 = ld2(B0 + (int64_t)k * C3 + (int64_t)d * ld + n);
                acc[i][d][0] = b0.x;
                acc[i][d][1] = b0.y;
            }
        };
#pragma unroll
        for (int i = 0; i < MTN; ++i) load_b0(i);
        group_sync();   // D of this tile is complete (and every warp is done with the previous tile's D)
        for (int it = 0; it < iters; ++it, ++git) {
            if (G == 2 && skew) {
                // group 0 starts iteration a after group 1 signalled in its iteration a - 1; group 1 starts iteration b after
                // group 0 signalled in ITS iteration b (no wait once the partner has run out of iterations)
                const int64_t need = grp == 0 ? git - 1 : git;
                if (need >= 0 && need < partner_iters) mbar_wait(mid + (grp ^ 1), (uint32_t)(need & 1));
            }
            for (int s = 0; s < nslab; ++s) {
                if (gt == 0 && !first && issued < total_all) issue();   // refill the stage released one step ago
                first = false;
                mbar_wait(full + cst, cwrap);
                const double* as = afrag + cst * astage;
                const double* ds = bfrag + s * BK * DLD;
                // `tok` collects a word of every A fragment this lane loads from the stage; the release below is issued with
                // an address that depends on it ((tok & zero) == 0 at run time, `zero` being a kernel argument the compiler cannot
                // fold).  Without that dependency ptxas schedules SYNCS.ARRIVE right behind the ISSUE of the last LDS (the DMMAs
                // that consume the fragments sink below it), the arrival can overtake the shared-memory read, the issuing thread
                // refills the stage through the async proxy and the last m-tile of the slowest warp multiplies with the NEXT
                // lap's slab (seen on a B200 at N = 14 290, K = 200: wrong entries only in m-tile 6 of the 7-tile warp, only in
                // CTAs whose issuing thread sits in a 6-tile warp, run-to-run different -- tools/admm_ab.py).
                unsigned tok = 0;
#pragma unroll
                for (int kk = 0; kk < BK / 4; ++kk) {
                    double b[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d) b[d] = ds[kk * 4 * DLD + d * 8];
#pragma unroll
                    for (int i = 0; i < MTN; ++i) {
                        const double a = as[i * 8 * ALD + kk * 4];
                        tok |= (unsigned)__double2hiint(a);
#pragma unroll
                        for (int d = 0; d < 3; ++d) dmma(acc[i][d][0], acc[i][d][1], a, b[d]);
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(empty + cst + (tok & (unsigned)zero));   // every LDS of this stage has returned its data
                    if (G == 2 && s == sig_slab) mbar_arrive(mid + grp);
                }
                if (++cst == STAGES) { cst = 0; cwrap ^= 1u; }
            }
            group_sync();  // every warp has finished reading D for this iteration

            // ---- epilogue: prox + dual update on the thread's own entries.  Lambda of all m-tiles is fetched up front (one L2
            // latency); without the register copy the old dual U (L2 hits: written by this thread one iteration ago) runs
            // UDEPTH m-tiles ahead of the arithmetic.
            const bool last = it == iters - 1;
            double2 lam[MTN];
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                lam[i] = (inb && k < K) ? ld2(Lambda + (int64_t)k * ld + n) : make_double2(0.0, 0.0);
            }
            constexpr int UDEPTH = UREG ? 1 : (MTN < 2 ? MTN : 2);
            double2 uo[UDEPTH][3];
            auto fetch_u = [&](int i) {
                const int k = (mt0 + i) * 8 + lr;
#pragma unroll
                for (int d = 0; d < 3; ++d) uo[i % UDEPTH][d] = make_double2(0.0, 0.0);
                if (!inb || k >= K) return;
                const int64_t base = (int64_t)k * C3 + n;
#pragma unroll
                for (int d = 0; d < 3; ++d) uo[i % UDEPTH][d] = ld2(U + base + (int64_t)d * ld);
            };
            if (!UREG) {
#pragma unroll
                for (int i = 0; i < UDEPTH; ++i) fetch_u(i);
            }
#pragma unroll
            for (int i = 0; i < MTN; ++i) {
                const int k = (mt0 + i) * 8 + lr;
                const bool on = inb && k < K;
                const int64_t base = (int64_t)k * C3 + n;
                double2 v[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {   // V = C + U
                    v[d].x = acc[i][d][0] + (UREG ? u[i % UR][d][0] : uo[i % UDEPTH][d].x);
                    v[d].y = acc[i][d][1] + (UREG ? u[i % UR][d][1] : uo[i % UDEPTH][d].y);
                }
                if (!UREG && i + UDEPTH < MTN) fetch_u(i + UDEPTH);   // slot i % UDEPTH is free again
                if (!last) load_b0(i);          // acc[i] is consumed: start the next iteration's C = B0 + ...
                if (on) {
                    const double sh0 = prox_shrink(v[0].x, v[1].x, v[2].x, lam[i].x, inv_rho);
                    const double sh1 = prox_shrink(v[0].y, v[1].y, v[2].y, lam[i].y, inv_rho);
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        double2 z, u2;
                        z.x = ok0 ? v[d].x * sh0 : 0.0;
                        z.y = ok1 ? v[d].y * sh1 : 0.0;
                        u2.x = ok0 ? v[d].x - z.x : 0.0;
                        u2.y = ok1 ? v[d].y - z.y : 0.0;
                        if (UREG) {
                            u[i % UR][d][0] = u2.x;
                            u[i % UR][d][1] = u2.y;
                        }
                        if (!UREG || last) st2(U + base + (int64_t)d * ld, u2);
                        if (last) st2(Z + base + (int64_t)d * ld, z);
                        else st2(down + k * DLD + d * 8, make_double2(z.x - u2.x, z.y - u2.y));
                    }
                }
            }
            if (!last) group_sync();  // the new D is complete before any warp multiplies with it
        }
    }
}

template <class Cfg, int MT>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::CTAS)
admm_tma_kernel(const double* __restrict__ Hpk, const double* __restrict__ B0, double* __restrict__ Z, double* __restrict__ U,
                const double* __restrict__ Lambda, int K, int64_t N, int64_t ld, double inv_rho, int iters, int64_t ntiles,
                int skew, int zero) {
    extern __shared__ __align__(128) double smem[];
    int mt0, mtn;
    int wm = (threadIdx.x >> 5) & 3;
    // co-resident groups rotate the row-warp roles: the 7-tile warps (K = 200: 7,6,6,6) sit on different SM sub-partitions
    const int slot = Cfg::G > 1 ? (int)(threadIdx.x >> 7) : (int)((blockIdx.x / max(1u, gridDim.x / 2)) & 1);
    wm = (wm + (Cfg::G == 3 ? slot : 2 * slot)) % Cfg::WM;
    mtile_range(((K + 7) & ~7) >> 3, Cfg::WM, wm, mt0, mtn);
    const bool live = mtn > 0;
    if (!live) mt0 = 0;
    if (mtn == MT) admm_tma_body<Cfg, MT, MT>(smem, Hpk, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, ntiles, mt0, live, skew, zero);
    else admm_tma_body<Cfg, MT, (MT > 1 ? MT - 1 : 1)>(smem, Hpk, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, ntiles, mt0, live, skew,
                                                        zero);
}

template <class Cfg>
static size_t admm_tma_smem_bytes(int64_t K) {
    const int64_t rows8 = (K + 7) & ~7LL;
    return (size_t)Cfg::G * ((size_t)rows8 * Cfg::DLD + (size_t)rows8 * Cfg::STAGES * Cfg::ALD) * sizeof(double) +
           (size_t)(Cfg::G * 2 * Cfg::STAGES + 2) * 8;
}


template <class Cfg, int MT>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::CTAS)
admm_solve_kernel(const double* __restrict__ Hr, int64_t ldh, const double* __restrict__ B0, double* __restrict__ Z,
                  double* __restrict__ U, const double* __restrict__ Lambda, int K, int64_t N, int64_t ld,
                  double inv_rho, int iters, int64_t ntiles, int use_lock) {
    extern __shared__ __align__(16) double smem[];
    int* lock = (use_lock && Cfg::CTAS > 1) ? g_admm_epilogue_lock + (smid() & 1023) : nullptr;
    int mt0, mtn;
    int wm = (threadIdx.x >> 5) / Cfg::WN;
    if (Cfg::CTAS == 2) {
        // co-resident CTAs (blockIdx.x and blockIdx.x + #SMs under round-robin placement) rotate the row-warp roles by two
        // sub-partitions, so the 7-tile warps of the two CTAs sit on different SM sub-partitions (13,12,13,12, not 14,12,12,12)
        const int slot = (blockIdx.x / max(1u, gridDim.x / 2)) & 1;
        wm = (wm + 2 * slot) % Cfg::WM;
    }
    mtile_range(((K + 7) & ~7) >> 3, Cfg::WM, wm, mt0, mtn);
    const bool live = mtn > 0;   // K <= 24: some row-warps own no m-tile; they idle on tile 0 and never store
    if (!live) mt0 = 0;
    if (mtn == MT) admm_body<Cfg, MT, MT>(smem, Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, ntiles, mt0, live, lock);
    else admm_body<Cfg, MT, (MT > 1 ? MT - 1 : 1)>(smem, Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, ntiles, mt0, live,
                                                  lock);
}

template <class Cfg>
static size_t admm_smem_bytes(int64_t K) {
    const int64_t rows8 = (K + 7) & ~7LL, rowsk = cdiv(K, Cfg::BK) * Cfg::BK;
    return ((size_t)rowsk * Cfg::DLD + (size_t)rows8 * Cfg::STAGES * Cfg::ALD) * sizeof(double);
}

static int g_admm_layout = -1, g_admm_skew = -1;   // -1: from the environment (asb_admm_select_layout overrides both)
static int env_int(const char* name, int dflt) {
    const char* e = std::getenv(name);
    return e ? std::atoi(e) : dflt;
}

template <class Cfg, int MT>
static int launch_admm_cfg(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                           int64_t K, int64_t N, int64_t ld, double inv_rho, int iters, cudaStream_t st) {
    const size_t smem = admm_smem_bytes<Cfg>(K);
    ASB_CUDA(cudaFuncSetAttribute(admm_solve_kernel<Cfg, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = cdiv(ld, Cfg::NV);
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)Cfg::CTAS * sm_count());
    static const int use_lock = env_int("ASB_ADMM_LOCK", 0);
    if (use_lock && Cfg::CTAS > 1) {   // a previous launch that died mid-kernel must not leave a mutex taken
        void* lk = nullptr;
        ASB_CUDA(cudaGetSymbolAddress(&lk, g_admm_epilogue_lock));
        ASB_CUDA(cudaMemsetAsync(lk, 0, sizeof(int) * 1024, st));
    }
    admm_solve_kernel<Cfg, MT><<<grid, Cfg::THREADS, smem, st>>>(Hr, ldh, B0, Z, U, Lambda, (int)K, N, ld, inv_rho, iters, ntiles,
                                                                use_lock);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

// m-tiles per row-warp decide the instantiation (accumulator registers): 1..8 with 4 row-warps, 1..4 with 8
template <class Cfg>
static int launch_admm_layout(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                              int64_t K, int64_t N, int64_t ld, double inv_rho, int iters, cudaStream_t st) {
    constexpr int MTMAX = 32 / Cfg::WM;   // K <= 256
    switch ((int)cdiv(cdiv(K, 8), Cfg::WM)) {
#define ASB_ADMM_CASE(MT) case MT: return launch_admm_cfg<Cfg, (MT <= MTMAX ? MT : MTMAX)>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
        ASB_ADMM_CASE(1) ASB_ADMM_CASE(2) ASB_ADMM_CASE(3) ASB_ADMM_CASE(4) ASB_ADMM_CASE(5) ASB_ADMM_CASE(6) ASB_ADMM_CASE(7)
        default: return launch_admm_cfg<Cfg, MTMAX>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
#undef ASB_ADMM_CASE
    }
}

template <class Cfg, int MT>
static int launch_admm_tma_mt(const double* Hpk, const double* B0, double* Z, double* U, const double* Lambda, int64_t K, int64_t N,
                              int64_t ld, double inv_rho, int iters, int skew, cudaStream_t st) {
    static const int solo = env_int("ASB_ADMM_TMA_SOLO", 0);   // experiment: one 4-warp CTA per SM (extra shared memory)
    const size_t smem = admm_tma_smem_bytes<Cfg>(K) + ((Cfg::G == 1 && solo) ? 24 * 1024 : 0);
    ASB_CUDA(cudaFuncSetAttribute(admm_tma_kernel<Cfg, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = cdiv(ld, Cfg::NV);
    const int64_t slots = Cfg::G > 1 ? sm_count() : (solo ? sm_count() : (int64_t)Cfg::CTAS * sm_count());
    const unsigned grid = (unsigned)std::min<int64_t>(cdiv(ntiles, Cfg::G), slots);
    static volatile int runtime_zero = 0;   // see admm_tma_body: a zero the compiler cannot see through
    admm_tma_kernel<Cfg, MT><<<grid, Cfg::THREADS, smem, st>>>(Hpk, B0, Z, U, Lambda, (int)K, N, ld, inv_rho, iters, ntiles, skew,
                                                              runtime_zero);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

// bytes of the slab-major copy of Hr the bulk-copy layouts stream from (the larger of the two slab formats)
static size_t admm_pack_bytes(int64_t K) {
    const int64_t rows8 = (K + 7) & ~7LL;
    return (size_t)(rows8 / 8) * rows8 * 12 * sizeof(double);
}

template <class Cfg>
static int launch_admm_tma(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                           int64_t N, int64_t ld, double inv_rho, int iters, double* pack, cudaStream_t st) {
    const int rows8 = (int)((K + 7) & ~7LL), nslab = rows8 / Cfg::BK;
    static const int env_skew = env_int("ASB_ADMM_SKEW", 0);
    const int skew = g_admm_skew >= 0 ? g_admm_skew : env_skew;
    admm_pack_kernel<<<64, 256, 0, st>>>(Hr, ldh, rows8, nslab, Cfg::BK, Cfg::ALD, pack);
    ASB_LAUNCH_CHECK();
    switch ((int)cdiv(cdiv(K, 8), Cfg::WM)) {
#define ASB_ADMM_CASE(MT) case MT: return launch_admm_tma_mt<Cfg, MT>(pack, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, skew, st);
        ASB_ADMM_CASE(1) ASB_ADMM_CASE(2) ASB_ADMM_CASE(3) ASB_ADMM_CASE(4) ASB_ADMM_CASE(5) ASB_ADMM_CASE(6) ASB_ADMM_CASE(7)
        default: return launch_admm_tma_mt<Cfg, 8>(pack, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, skew, st);
#undef ASB_ADMM_CASE
    }
}


static int admm_layout() {
    static const int layout = env_int("ASB_ADMM_LAYOUT", ASB_ADMM_LAYOUT_DEFAULT);   // 5: TMA Pair, 4 / 416: Pair, 12 / 1216: Wide, 16: Tall
    return g_admm_layout >= 0 ? g_admm_layout : layout;
}

static int launch_admm(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                       int64_t K, int64_t N, int64_t ld, double inv_rho, int iters, double* pack, cudaStream_t st) {
    const int layout = admm_layout();
    constexpr size_t SMEM_MAX = 227 * 1024;   // a layout whose tile does not fit at this K falls through to Pair (139 KB at K = 256)
    if (pack) {   // bulk-copy layouts: 7 trio (default), 6 duet, 5 pair; a layout whose tiles do not fit at this K falls through
        if (layout == 7 && admm_tma_smem_bytes<AdmmTmaTrio>(K) <= SMEM_MAX)
            return launch_admm_tma<AdmmTmaTrio>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, pack, st);
        if ((layout == 6 || layout == 7) && admm_tma_smem_bytes<AdmmTmaDuet>(K) <= SMEM_MAX)
            return launch_admm_tma<AdmmTmaDuet>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, pack, st);
        if (layout == 5 || layout == 6 || layout == 7)
            return launch_admm_tma<AdmmTmaPair>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, pack, st);
    }
    if (layout == 16 && admm_smem_bytes<AdmmTall>(K) <= SMEM_MAX)
        return launch_admm_layout<AdmmTall>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
    if (layout == 8) return launch_admm_layout<AdmmDuo>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
    if (layout == 81) return launch_admm_layout<AdmmDuoU>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
    if (layout == 12 && admm_smem_bytes<AdmmWide>(K) <= SMEM_MAX)
        return launch_admm_layout<AdmmWide>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
    if (layout == 1216 && admm_smem_bytes<AdmmWide16>(K) <= SMEM_MAX)
        return launch_admm_layout<AdmmWide16>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
    if (layout == 416) return launch_admm_layout<AdmmPair16>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
    return launch_admm_layout<AdmmPair>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, st);
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_admm_select_layout(int layout, int skew) {
    g_admm_layout = layout;
    g_admm_skew = skew;
    return ASB_OK;
}

int64_t asb_admm_workspace_bytes(int64_t K) { return K > 0 && K <= 256 ? (int64_t)admm_pack_bytes(K) : 0; }

int asb_admm_solve_ws(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                      int64_t N, int64_t ld, double inv_rho, int64_t iters, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(Hr && B0 && Z && U && Lambda && K > 0 && K <= 256 && N > 0 && ld >= N && ld % 32 == 0,
                "admm_solve: bad arguments (K <= 256)");
    ASB_REQUIRE(ldh % 2 == 0 && ((uintptr_t)Hr % 16 == 0), "admm_solve: Hr must be 16-byte aligned with even ldh");
    ASB_REQUIRE(ldh >= ((K + 15) & ~15LL), "admm_solve: Hr must be zero-padded to ceil8(K) rows x ceil16(K) columns");
    ASB_REQUIRE(iters >= 0 && iters < (1 << 20), "admm_solve: bad iteration count");
    ASB_REQUIRE(!ws || (uintptr_t)ws % 16 == 0, "admm_solve: workspace must be 16-byte aligned");
    if (iters == 0) return ASB_OK;
    return launch_admm(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, (int)iters, reinterpret_cast<double*>(ws), S(stream));
}

int asb_admm_solve(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                   int64_t N, int64_t ld, double inv_rho, int64_t iters, asb_stream_t stream) {
    // workspace-free form: the slab-major copy of Hr lives in a per-process device buffer (786 KB at K = 256), so calls of
    // this form must not overlap on different streams; asb_admm_solve_ws has no such restriction
    static void* pack = nullptr;
    if (!pack && admm_layout() >= 5 && admm_layout() <= 7) ASB_CUDA(cudaMalloc(&pack, admm_pack_bytes(256)));
    return asb_admm_solve_ws(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, iters, pack, stream);
}

int asb_admm_iter(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                  int64_t N, int64_t ld, double inv_rho, asb_stream_t stream) {
    return asb_admm_solve(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, 1, stream);
}

}  // extern "C"
