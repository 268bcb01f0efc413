// Skinny DMMA contractions: one operand has at most 256 rows (the K components / the K x F weight operator).
//
// That operand becomes the DMMA m-side and is covered COMPLETELY by one CTA (no 128-row tile quantisation: K = 200 costs
// 25 m-tiles, not 32).  Its ceil(rows/8) m-tiles are dealt to WM = 4 row-warps, MT or MT-1 each (K = 200: 7,6,6,6); with
// warp id = wm * 2 + wn the four SM sub-partitions (warp id mod 4) carry 13,13,12,12 tiles.  The kernel body is compiled
// for both tile counts and selected by a warp-uniform BRANCH: a predicated-off DMMA still occupies the tensor pipe
// (measured: 80 % pipe-active for 70 % useful work with predication).  The long operand is the n-side, 64 rows/columns
// per CTA (2 column-warps x 4 n-tiles).  4-stage cp.async ring, BK = 16; every thread keeps running source pointers so
// the in-loop address arithmetic is one 64-bit add per 16-byte chunk.
#pragma once
#include "dmma.cuh"

namespace asb {

constexpr int SK_BK = 16, SK_ALD = SK_BK + 4;  // 20 doubles per staged row (conflict-free fragments)
constexpr int SK_STAGES = 4, SK_WM = 4, SK_NT = 4;
// WN column-warps (2 or 3): threads = 128 WN, n-side tile = 32 WN, B stored [k][n] with stride 32 WN + 4 (mod 16 == 4)
#define SK_THREADS (SK_WM * WN * 32)
#define SK_BN (WN * SK_NT * 8)
#define SK_BLD_NN (SK_BN + 4)
#define SK_RPP (SK_THREADS / 8)   /* rows staged per pass of the CTA */

// Rows-by-BK staging of a row-major operand: chunk t of a thread is row (tid>>3) + RPP t, doubles (tid&7)*2 .. +1.
template <int NCH, int RPP>
struct RowStager {
    const double* p;      // source of chunk 0 at the current k position
    const double* safe;   // always-valid address for zero-filled chunks
    int64_t step;         // 32 rows further
    unsigned dst;         // shared offset (doubles) of chunk 0 inside a stage
    unsigned okmask;      // chunk t is inside the operand
    int cc;
    __device__ __forceinline__ void init(const double* g, int64_t ld, int64_t row0, int64_t nrows, int rows_staged,
                                         int64_t k0) {
        const int r = threadIdx.x >> 3;
        cc = (threadIdx.x & 7) * 2;
        okmask = 0;
#pragma unroll
        for (int t = 0; t < NCH; ++t)
            if (r + RPP * t < rows_staged && row0 + r + RPP * t < nrows) okmask |= 1u << t;
        const int64_t rr = min(row0 + r, nrows - 1);
        p = g + rr * ld + k0 + cc;
        safe = g;
        step = RPP * ld;
        dst = r * SK_ALD + cc;
    }
    // stage the BK slice starting at the current position (kleft = doubles left on the contraction axis), then advance
    __device__ __forceinline__ void issue(double* stage, int rows_staged, int64_t kleft) {
        const bool kin = cc < kleft;
        const double* q = p;
        const int r = threadIdx.x >> 3;
#pragma unroll
        for (int t = 0; t < NCH; ++t) {
            if (r + RPP * t < rows_staged) {
                const bool ok = kin && ((okmask >> t) & 1u);
                cp_async16(stage + dst + t * RPP * SK_ALD, ok ? q : safe, ok ? 16 : 0);
            }
            q += step;
        }
        p += SK_BK;
    }
};

template <int MTN, bool NT_LAYOUT, int WN>
__device__ __forceinline__ void sk_mma_stage(const double* __restrict__ as, const double* __restrict__ bs,
                                             double (&acc)[MTN][SK_NT][2]) {
#pragma unroll
    for (int kk = 0; kk < SK_BK / 4; ++kk) {
        double b[SK_NT];
#pragma unroll
        for (int j = 0; j < SK_NT; ++j) b[j] = NT_LAYOUT ? bs[j * 8 * SK_ALD + kk * 4] : bs[kk * 4 * SK_BLD_NN + j * 8];
#pragma unroll
        for (int i = 0; i < MTN; ++i) {
            const double a = as[i * 8 * SK_ALD + kk * 4];
#pragma unroll
            for (int j = 0; j < SK_NT; ++j) dmma(acc[i][j][0], acc[i][j][1], a, b[j]);
        }
    }
}

// ------------------------------------------------------------------------------------------ gemm_nn, r <= 256
// D (r x M) = alpha * diag(row_scale) * A (r x q) * B (q x M).  grid.x = column tile of 64.
template <int MT, int MTN, int WN>
__device__ __forceinline__ void nn_skinny_body(double* smem, const double* __restrict__ A, int64_t lda, int r, int64_t q,
                                               const double* __restrict__ B, int64_t ldb, int64_t M, double alpha,
                                               const double* __restrict__ row_scale, double* __restrict__ D, int64_t ldd,
                                               int mt0, bool live) {
    const int rows8 = (r + 7) & ~7;
    const int astage = rows8 * SK_ALD, bstage = SK_BK * SK_BLD_NN;
    double* As = smem;
    double* Bs = smem + SK_STAGES * astage;
    const int64_t col0 = (int64_t)blockIdx.x * SK_BN;
    const int64_t nk = (q + SK_BK - 1) / SK_BK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wn = warp % WN, lr = lane >> 2, lc = lane & 3;
    double acc[MTN][SK_NT][2];
#pragma unroll
    for (int i = 0; i < MTN; ++i)
#pragma unroll
        for (int j = 0; j < SK_NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    RowStager<(32 * MT + SK_RPP - 1) / SK_RPP, SK_RPP> sa;
    sa.init(A, lda, 0, r, rows8, 0);
    // B tile: BK x 32 WN -> [k][SK_BLD_NN]; 2 chunks per thread: k rows tid / (16 WN) and + 8, doubles (tid % (16 WN)) * 2
    const int brow = threadIdx.x / (SK_BN / 2), bcc = (threadIdx.x % (SK_BN / 2)) * 2;
    const bool bcol_ok = col0 + bcc < M;
    const double* bp = B + (int64_t)brow * ldb + (bcol_ok ? col0 + bcc : 0);
    const int64_t bstep = 8 * ldb;
    const unsigned bdst = brow * SK_BLD_NN + bcc;
    auto issue_b = [&](double* stage, int64_t kleft) {
        const bool ok0 = bcol_ok && brow < kleft, ok1 = bcol_ok && brow + 8 < kleft;
        cp_async16(stage + bdst, ok0 ? bp : B, ok0 ? 16 : 0);
        cp_async16(stage + bdst + 8 * SK_BLD_NN, ok1 ? bp + bstep : B, ok1 ? 16 : 0);
        bp += 2 * bstep;
    };
#pragma unroll
    for (int s = 0; s < SK_STAGES - 1; ++s) {
        if (s < nk) {
            sa.issue(As + s * astage, rows8, q - (int64_t)s * SK_BK);
            issue_b(Bs + s * bstage, q - (int64_t)s * SK_BK);
        }
        cp_async_commit();
    }
    const double* afrag = As + (mt0 * 8 + lr) * SK_ALD + lc;
    const double* bfrag = Bs + lc * SK_BLD_NN + wn * SK_NT * 8 + lr;
    int cs = 0, ls = SK_STAGES - 1;  // compute stage, load stage
    for (int64_t kt = 0; kt < nk; ++kt) {
        cp_async_wait<SK_STAGES - 2>();
        __syncthreads();
        const int64_t nx = kt + SK_STAGES - 1;
        if (nx < nk) {
            sa.issue(As + ls * astage, rows8, q - nx * SK_BK);
            issue_b(Bs + ls * bstage, q - nx * SK_BK);
        }
        cp_async_commit();
        sk_mma_stage<MTN, false, WN>(afrag + cs * astage, bfrag + cs * bstage, acc);
        cs = cs + 1 == SK_STAGES ? 0 : cs + 1;
        ls = ls + 1 == SK_STAGES ? 0 : ls + 1;
    }
    cp_async_wait<0>();
#pragma unroll
    for (int i = 0; i < MTN; ++i) {
        const int row = (mt0 + i) * 8 + lr;
        if (!live || row >= r) continue;
        const double sc = row_scale ? alpha * row_scale[row] : alpha;
#pragma unroll
        for (int j = 0; j < SK_NT; ++j) {
            const int64_t c = col0 + wn * SK_NT * 8 + j * 8 + 2 * lc;
            if (c < M) st2(D + (int64_t)row * ldd + c, make_double2(sc * acc[i][j][0], sc * acc[i][j][1]));
        }
    }
}

template <int MT, int WN>
__global__ void __launch_bounds__(SK_THREADS, 1)
gemm_nn_skinny_kernel(const double* __restrict__ A, int64_t lda, int r, int64_t q, const double* __restrict__ B, int64_t ldb,
                      int64_t M, double alpha, const double* __restrict__ row_scale, double* __restrict__ D, int64_t ldd) {
    extern __shared__ __align__(16) double smem[];
    int mt0, mtn;
    mtile_range(((r + 7) & ~7) >> 3, SK_WM, (threadIdx.x >> 5) / WN, mt0, mtn);
    const bool live = mtn > 0;   // r <= 24: some row-warps own no m-tile; they recompute tile 0 and never store
    if (!live) mt0 = 0;
    if (mtn == MT) nn_skinny_body<MT, MT, WN>(smem, A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd, mt0, live);
    else nn_skinny_body<MT, (MT > 1 ? MT - 1 : 1), WN>(smem, A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd, mt0, live);
}

// ------------------------------------------------------------------------------------------ gemm_nt, small side <= 256
// part[slab][...] = S L^T over one slab of the long axis;  S: rs x M (the m-side), L: rl x M, 64 rows of L per CTA.
// Output element (s, l) goes to part[l * ldo + s] (transpose_out) or part[s * ldo + l].
// grid.x = L tile (fastest: CTAs sharing a slab of S run together -> S is read from HBM once), grid.y = slab.
template <int MT, int MTN, int WN>
__device__ __forceinline__ void nt_skinny_body(double* smem, const double* __restrict__ Sm, int64_t lds, int rs,
                                               const double* __restrict__ L, int64_t ldl, int64_t rl, int64_t M,
                                               int64_t slab_len, int transpose_out, double* __restrict__ out, int64_t ldo,
                                               int64_t slab_stride, int mt0, bool live) {
    const int rows8 = (rs + 7) & ~7;
    const int astage = rows8 * SK_ALD, bstage = SK_BN * SK_ALD;
    double* As = smem;
    double* Bs = smem + SK_STAGES * astage;
    const int64_t l0 = (int64_t)blockIdx.x * SK_BN;
    const int64_t m0 = (int64_t)blockIdx.y * slab_len;
    const int64_t m1 = min(M, m0 + slab_len);
    const int64_t nk = (m1 - m0 + SK_BK - 1) / SK_BK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wn = warp % WN, lr = lane >> 2, lc = lane & 3;
    double acc[MTN][SK_NT][2];
#pragma unroll
    for (int i = 0; i < MTN; ++i)
#pragma unroll
        for (int j = 0; j < SK_NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    RowStager<(32 * MT + SK_RPP - 1) / SK_RPP, SK_RPP> sa;
    sa.init(Sm, lds, 0, rs, rows8, m0);
    RowStager<2, SK_RPP> sl;  // 32 WN rows of L = 2 passes
    sl.init(L, ldl, l0, rl, SK_BN, m0);
#pragma unroll
    for (int s = 0; s < SK_STAGES - 1; ++s) {
        if (s < nk) {
            sa.issue(As + s * astage, rows8, m1 - m0 - (int64_t)s * SK_BK);
            sl.issue(Bs + s * bstage, SK_BN, m1 - m0 - (int64_t)s * SK_BK);
        }
        cp_async_commit();
    }
    const double* afrag = As + (mt0 * 8 + lr) * SK_ALD + lc;
    const double* bfrag = Bs + (wn * SK_NT * 8 + lr) * SK_ALD + lc;
    int cs = 0, ls = SK_STAGES - 1;
    for (int64_t kt = 0; kt < nk; ++kt) {
        cp_async_wait<SK_STAGES - 2>();
        __syncthreads();
        const int64_t nx = kt + SK_STAGES - 1;
        if (nx < nk) {
            sa.issue(As + ls * astage, rows8, m1 - m0 - nx * SK_BK);
            sl.issue(Bs + ls * bstage, SK_BN, m1 - m0 - nx * SK_BK);
        }
        cp_async_commit();
        sk_mma_stage<MTN, true, WN>(afrag + cs * astage, bfrag + cs * bstage, acc);
        cs = cs + 1 == SK_STAGES ? 0 : cs + 1;
        ls = ls + 1 == SK_STAGES ? 0 : ls + 1;
    }
    cp_async_wait<0>();
    double* o = out + (int64_t)blockIdx.y * slab_stride;
#pragma unroll
    for (int i = 0; i < MTN; ++i) {
        const int srow = (mt0 + i) * 8 + lr;
        if (!live || srow >= rs) continue;
#pragma unroll
        for (int j = 0; j < SK_NT; ++j) {
            const int64_t l = l0 + wn * SK_NT * 8 + j * 8 + 2 * lc;
            if (transpose_out) {
                if (l < rl) o[l * ldo + srow] = acc[i][j][0];
                if (l + 1 < rl) o[(l + 1) * ldo + srow] = acc[i][j][1];
            } else {
                if (l < rl) o[(int64_t)srow * ldo + l] = acc[i][j][0];
                if (l + 1 < rl) o[(int64_t)srow * ldo + l + 1] = acc[i][j][1];
            }
        }
    }
}

template <int MT, int WN>
__global__ void __launch_bounds__(SK_THREADS, 1)
gemm_nt_skinny_kernel(const double* __restrict__ Sm, int64_t lds, int rs, const double* __restrict__ L, int64_t ldl,
                      int64_t rl, int64_t M, int64_t slab_len, int transpose_out, double* __restrict__ out, int64_t ldo,
                      int64_t slab_stride) {
    extern __shared__ __align__(16) double smem[];
    int mt0, mtn;
    mtile_range(((rs + 7) & ~7) >> 3, SK_WM, (threadIdx.x >> 5) / WN, mt0, mtn);
    const bool live = mtn > 0;
    if (!live) mt0 = 0;
    if (mtn == MT) nt_skinny_body<MT, MT, WN>(smem, Sm, lds, rs, L, ldl, rl, M, slab_len, transpose_out, out, ldo, slab_stride, mt0,
                                          live);
    else nt_skinny_body<MT, (MT > 1 ? MT - 1 : 1), WN>(smem, Sm, lds, rs, L, ldl, rl, M, slab_len, transpose_out, out, ldo,
                                                    slab_stride, mt0, live);
}

}  // namespace asb
