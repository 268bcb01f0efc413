// FP64 tensor-core (DMMA m8n8k4) contractions of the basis build.
//
//   gemm_nt : D = A B^T, contraction over the long (vertex) axis M  -> small outputs (F x F Gram, X C^T, C C^T)
//   gemm_nn : D = A B,   small left matrix times the wide snapshot/component matrix (c = W^T X, back-projection)
//   admm    : one fused SPLOCS ADMM iteration (K x K times K x 3N GEMM + group-lasso prox + dual update)
//
// sm_100a has no fp64 kind of tcgen05.mma; fp64 tensor work is `mma.sync.aligned.m8n8k4.f64` (SASS DMMA.8x8x4).
// Operand tiles are staged global -> shared with a 3-stage cp.async (LDGSTS) pipeline; shared tiles use padded
// strides (stride mod 16 doubles == 4 or 12) so that every 8x4 / 4x8 fragment read is bank-conflict free.
#include "dmma.cuh"
#include "dmma_skinny.cuh"
#include <cstdlib>
#ifndef ASB_SK_WN_DEFAULT
#define ASB_SK_WN_DEFAULT 3
#endif

namespace asb {

constexpr int BM = 128, BN = 128, BK = 16, STAGES = 3, GEMM_THREADS = 256;
constexpr int A_LD = BK + 4;         // 20 doubles: rows 160 B apart -> conflict-free 8x4 fragment reads
constexpr int BNT_LD = BK + 4;       // B stored [n][k]
constexpr int BNN_LD = BN + 4;       // B stored [k][n], 132
constexpr int A_STAGE = BM * A_LD;   // doubles
constexpr int BNT_STAGE = BN * BNT_LD;
constexpr int BNN_STAGE = BK * BNN_LD;

// rows x BK tile (row-major source, `ld` doubles between rows) -> smem [row][A_LD]; zero-fill out of range
__device__ __forceinline__ void load_tile_rows(double* sm, const double* __restrict__ g, int64_t ld, int64_t row0,
                                               int64_t nrows, int64_t k0, int64_t klen) {
    // 128 rows x 8 chunks(16 B)
#pragma unroll
    for (int t = 0; t < (BM * BK / 2) / GEMM_THREADS; ++t) {
        const int q = threadIdx.x + t * GEMM_THREADS;
        const int row = q >> 3, cc = (q & 7) * 2;
        const int64_t gr = row0 + row, gk = k0 + cc;
        const bool ok = gr < nrows && gk < klen;
        const double* src = ok ? g + gr * ld + gk : g;
        cp_async16(sm + row * A_LD + cc, src, ok ? 16 : 0);
    }
}
// BK x BN tile of a row-major (q x M) matrix -> smem [k][BNN_LD]
__device__ __forceinline__ void load_tile_cols(double* sm, const double* __restrict__ g, int64_t ld, int64_t k0,
                                               int64_t klen, int64_t col0, int64_t ncols) {
#pragma unroll
    for (int t = 0; t < (BK * BN / 2) / GEMM_THREADS; ++t) {
        const int q = threadIdx.x + t * GEMM_THREADS;
        const int row = q >> 6, cc = (q & 63) * 2;
        const int64_t gk = k0 + row, gc = col0 + cc;
        const bool ok = gk < klen && gc < ncols;
        const double* src = ok ? g + gk * ld + gc : g;
        cp_async16(sm + row * BNN_LD + cc, src, ok ? 16 : 0);
    }
}

// one BK slice of the 128x128 tile: warp tile 64 x 32 -> acc[8][4][2]
template <bool NT>
__device__ __forceinline__ void mma_stage(const double* __restrict__ As, const double* __restrict__ Bs,
                                          double (&acc)[8][4][2], int wm, int wn, int lane) {
    const int lr = lane >> 2, lc = lane & 3;
#pragma unroll
    for (int kk = 0; kk < BK / 4; ++kk) {
        double a[8], b[4];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = As[(wm * 64 + i * 8 + lr) * A_LD + kk * 4 + lc];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            b[j] = NT ? Bs[(wn * 32 + j * 8 + lr) * BNT_LD + kk * 4 + lc] : Bs[(kk * 4 + lc) * BNN_LD + wn * 32 + j * 8 + lr];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
}

// ------------------------------------------------------------------------------------------ gemm_nt
// grid.x = output tile (fastest: CTAs that run together share the same M slab -> L2 reuse of A/B rows),
// grid.y = M slab. Partial tile -> part[slab][ra][rb] (or D directly when there is one slab).
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_nt_kernel(const double* __restrict__ A, int64_t lda, int64_t ra, const double* __restrict__ B, int64_t ldb,
               int64_t rb, int64_t M, int64_t slab_len, int symmetric, int tiles_j, double* __restrict__ out,
               int64_t ldo, int64_t slab_stride) {
    extern __shared__ __align__(16) double smem[];
    double* As = smem;
    double* Bs = smem + STAGES * A_STAGE;
    int ti, tj;
    if (symmetric) {  // enumerate upper-triangular tile pairs (ti <= tj)
        int t = blockIdx.x;
        ti = 0;
        int rowlen = tiles_j;
        while (t >= rowlen) { t -= rowlen; ++ti; --rowlen; }
        tj = ti + t;
    } else {
        ti = blockIdx.x / tiles_j;
        tj = blockIdx.x % tiles_j;
    }
    const int64_t m0 = (int64_t)blockIdx.y * slab_len;
    const int64_t m1 = min(M, m0 + slab_len);
    const int64_t nk = (m1 - m0 + BK - 1) / BK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wm = warp >> 2, wn = warp & 3;
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    const int64_t arow0 = (int64_t)ti * BM, brow0 = (int64_t)tj * BN;
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) {
            load_tile_rows(As + s * A_STAGE, A, lda, arow0, ra, m0 + (int64_t)s * BK, m1);
            load_tile_rows(Bs + s * BNT_STAGE, B, ldb, brow0, rb, m0 + (int64_t)s * BK, m1);
        }
        cp_async_commit();
    }
    for (int64_t kt = 0; kt < nk; ++kt) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        const int64_t nx = kt + STAGES - 1;
        if (nx < nk) {
            const int s = (int)(nx % STAGES);
            load_tile_rows(As + s * A_STAGE, A, lda, arow0, ra, m0 + nx * BK, m1);
            load_tile_rows(Bs + s * BNT_STAGE, B, ldb, brow0, rb, m0 + nx * BK, m1);
        }
        cp_async_commit();
        const int s = (int)(kt % STAGES);
        mma_stage<true>(As + s * A_STAGE, Bs + s * BNT_STAGE, acc, wm, wn, lane);
    }
    cp_async_wait<0>();
    double* o = out + (int64_t)blockIdx.y * slab_stride;
    const int lr = lane >> 2, lc = lane & 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int64_t r = arow0 + wm * 64 + i * 8 + lr;
        if (r >= ra) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t c = brow0 + wn * 32 + j * 8 + 2 * lc;
            if (c < rb) o[r * ldo + c] = acc[i][j][0];
            if (c + 1 < rb) o[r * ldo + c + 1] = acc[i][j][1];
        }
    }
}

// D[i][j] = sum_s part[s][i][j]; symmetric: lower triangle mirrored from the upper one
__global__ void gemm_nt_reduce_kernel(const double* __restrict__ part, int64_t slabs, int64_t slab_stride, int64_t ra,
                                      int64_t rb, int64_t ldp, int symmetric, double* __restrict__ D, int64_t ldd) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ra * rb) return;
    const int64_t i = t / rb, j = t % rb;
    int64_t si = i, sj = j;
    if (symmetric == 1 && ((j / BN) < (i / BM))) { si = j; sj = i; }  // tile (ti > tj) was not computed
    if (symmetric == 2 && j < i) { si = j; sj = i; }                   // every tile computed: mirror element-wise
    double s = 0.0;
    for (int64_t k = 0; k < slabs; ++k) s += part[k * slab_stride + si * ldp + sj];
    D[i * ldd + j] = s;
}

// ------------------------------------------------------------------------------------------ gemm_nn
// grid.x = row tile (fastest: CTAs sharing a column block of B run together), grid.y = column tile.
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_nn_kernel(const double* __restrict__ A, int64_t lda, int64_t r, int64_t q, const double* __restrict__ B,
               int64_t ldb, int64_t M, double alpha, const double* __restrict__ row_scale, double* __restrict__ D,
               int64_t ldd) {
    extern __shared__ __align__(16) double smem[];
    double* As = smem;
    double* Bs = smem + STAGES * A_STAGE;
    const int64_t arow0 = (int64_t)blockIdx.x * BM, col0 = (int64_t)blockIdx.y * BN;
    const int64_t nk = (q + BK - 1) / BK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wm = warp >> 2, wn = warp & 3;
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) {
            load_tile_rows(As + s * A_STAGE, A, lda, arow0, r, (int64_t)s * BK, q);
            load_tile_cols(Bs + s * BNN_STAGE, B, ldb, (int64_t)s * BK, q, col0, M);
        }
        cp_async_commit();
    }
    for (int64_t kt = 0; kt < nk; ++kt) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        const int64_t nx = kt + STAGES - 1;
        if (nx < nk) {
            const int s = (int)(nx % STAGES);
            load_tile_rows(As + s * A_STAGE, A, lda, arow0, r, nx * BK, q);
            load_tile_cols(Bs + s * BNN_STAGE, B, ldb, nx * BK, q, col0, M);
        }
        cp_async_commit();
        const int s = (int)(kt % STAGES);
        mma_stage<false>(As + s * A_STAGE, Bs + s * BNN_STAGE, acc, wm, wn, lane);
    }
    cp_async_wait<0>();
    const int lr = lane >> 2, lc = lane & 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int64_t row = arow0 + wm * 64 + i * 8 + lr;
        if (row >= r) continue;
        const double sc = row_scale ? alpha * row_scale[row] : alpha;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t c = col0 + wn * 32 + j * 8 + 2 * lc;
            if (c < M) st2(D + row * ldd + c, make_double2(sc * acc[i][j][0], sc * acc[i][j][1]));
        }
    }
}


}  // namespace asb

using namespace asb;

struct NtPlan {
    bool skinny;       // one operand <= 256 rows: gemm_nt_skinny_kernel
    bool swap;         // skinny: A is the small (m-side) operand -> output written untransposed
    int64_t tiles, slabs, slab_len;
    int tiles_j;
};

// Slab count: fill whole waves of the SM count (a 2.2-wave grid runs as long as a 3-wave one), prefer fewer slabs among
// equally efficient choices, keep >= 8 k-tiles per slab and <= 64 partial outputs.
static int64_t pick_slabs(int64_t tiles, int64_t M) {
    const int64_t sms = sm_count();
    const int64_t max_by_len = std::max<int64_t>(1, M / (8 * BK));
    const int64_t smax = std::max<int64_t>(1, std::min<int64_t>(64, max_by_len));
    int64_t best = 1;
    double best_eff = -1.0;
    for (int64_t s = 1; s <= smax; ++s) {
        const int64_t ctas = tiles * s;
        double eff = (double)ctas / (double)(cdiv(ctas, sms) * sms);
        if (ctas >= 2 * sms) eff += 0.004;  // >= 2 waves: the tail of uneven CTAs is amortised
        if (eff > best_eff + 0.005) { best_eff = eff; best = s; }
    }
    return best;
}

// column-warps of the skinny kernels: 2 (256 threads, 64-wide n-tile) or 3 (384 threads, 96-wide); ASB_SK_WN overrides
static int sk_wn() {
    static int wn = 0;
    if (!wn) {
        const char* e = getenv("ASB_SK_WN");
        wn = (e && e[0] == '2') ? 2 : (e && e[0] == '3') ? 3 : ASB_SK_WN_DEFAULT;
    }
    return wn;
}

// 8 m-tiles per row-warp (rows > 224) need 128 accumulator registers: no room under the 168-register cap of 384 threads
static int sk_wn_for(int64_t small_rows) { return cdiv(cdiv(small_rows, 8), SK_WM) >= 8 ? 2 : sk_wn(); }

static NtPlan nt_plan(int64_t ra, int64_t rb, int64_t M, int symmetric) {
    NtPlan p{};
    const int64_t small = std::min(ra, rb);
    p.skinny = small <= 256;
    if (p.skinny) {
        p.swap = ra <= rb;  // A small: m-side = A, n-side = B
        p.tiles = cdiv(p.swap ? rb : ra, 32 * sk_wn_for(small));
        p.tiles_j = 1;
    } else {
        const int64_t ti = cdiv(ra, BM), tj = cdiv(rb, BN);
        p.tiles_j = (int)tj;
        p.tiles = symmetric ? ti * (ti + 1) / 2 : ti * tj;
    }
    const int64_t s = pick_slabs(p.tiles, M);
    p.slab_len = cdiv(cdiv(M, s), BK) * BK;
    p.slabs = cdiv(M, p.slab_len);
    return p;
}

template <int MT, int WN>
static int launch_nt_skinny(const double* Sm, int64_t lds, int64_t rs, const double* L, int64_t ldl, int64_t rl, int64_t M,
                            const NtPlan& p, int transpose_out, double* part, int64_t ldo, int64_t slab_stride,
                            cudaStream_t st) {
    const int rows8 = (int)((rs + 7) & ~7LL);
    const size_t smem = (size_t)SK_STAGES * (rows8 * SK_ALD + SK_BN * SK_ALD) * sizeof(double);
    ASB_CUDA(cudaFuncSetAttribute(gemm_nt_skinny_kernel<MT, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gemm_nt_skinny_kernel<MT, WN><<<dim3((unsigned)p.tiles, (unsigned)p.slabs), SK_THREADS, smem, st>>>(
        Sm, lds, (int)rs, L, ldl, rl, M, p.slab_len, transpose_out, part, ldo, slab_stride);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

template <int MT, int WN>
static int launch_nn_skinny(const double* A, int64_t lda, int64_t r, int64_t q, const double* B, int64_t ldb, int64_t M,
                            double alpha, const double* row_scale, double* D, int64_t ldd, cudaStream_t st) {
    const int rows8 = (int)((r + 7) & ~7LL);
    const size_t smem = (size_t)SK_STAGES * (rows8 * SK_ALD + SK_BK * SK_BLD_NN) * sizeof(double);
    ASB_CUDA(cudaFuncSetAttribute(gemm_nn_skinny_kernel<MT, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gemm_nn_skinny_kernel<MT, WN><<<(unsigned)cdiv(M, SK_BN), SK_THREADS, smem, st>>>(A, lda, (int)r, q, B, ldb, M, alpha,
                                                                                  row_scale, D, ldd);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

extern "C" {

int64_t asb_gemm_nt_workspace_bytes(int64_t ra, int64_t rb, int64_t M, int symmetric) {
    const NtPlan p = nt_plan(ra, rb, M, symmetric);
    return p.slabs * ra * rb * (int64_t)sizeof(double);
}

int asb_gemm_nt(const double* A, int64_t lda, int64_t ra, const double* B, int64_t ldb, int64_t rb, int64_t M, double* D,
                int64_t ldd, int symmetric, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(A && B && D && ws && ra > 0 && rb > 0 && M > 0, "gemm_nt: bad arguments");
    ASB_REQUIRE(lda % 2 == 0 && ldb % 2 == 0 && M % 2 == 0, "gemm_nt: leading dimensions and M must be even");
    ASB_REQUIRE(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0), "gemm_nt: operands must be 16-byte aligned");
    ASB_REQUIRE(!symmetric || (A == B && ra == rb), "gemm_nt: symmetric needs A == B");
    const NtPlan p = nt_plan(ra, rb, M, symmetric);
    double* part = reinterpret_cast<double*>(ws);
    const int64_t slab_stride = ra * rb;
    if (p.skinny) {
        // part[slab][i][j], i over A rows, j over B rows.  m-side = the small operand.
        const double* Sm = p.swap ? A : B;
        const double* L = p.swap ? B : A;
        const int64_t lds = p.swap ? lda : ldb, ldl = p.swap ? ldb : lda, rs = p.swap ? ra : rb, rl = p.swap ? rb : ra;
        const int tr = p.swap ? 0 : 1;  // element (s, l): swap -> part[s][l], else part[l][s]
        int rc;
        const bool wn3 = sk_wn_for(rs) == 3;
        switch ((int)cdiv(cdiv(rs, 8), SK_WM)) {  // m-tiles per row-warp
#define ASB_NT_CASE(MT) case MT: rc = wn3 ? launch_nt_skinny<MT, 3>(Sm, lds, rs, L, ldl, rl, M, p, tr, part, rb, slab_stride, S(stream)) : launch_nt_skinny<MT, 2>(Sm, lds, rs, L, ldl, rl, M, p, tr, part, rb, slab_stride, S(stream)); break;
            ASB_NT_CASE(1) ASB_NT_CASE(2) ASB_NT_CASE(3) ASB_NT_CASE(4) ASB_NT_CASE(5) ASB_NT_CASE(6) ASB_NT_CASE(7)
            default: rc = wn3 ? launch_nt_skinny<8, 3>(Sm, lds, rs, L, ldl, rl, M, p, tr, part, rb, slab_stride, S(stream))
                              : launch_nt_skinny<8, 2>(Sm, lds, rs, L, ldl, rl, M, p, tr, part, rb, slab_stride, S(stream));
#undef ASB_NT_CASE
        }
        if (rc != ASB_OK) return rc;
        symmetric = symmetric ? 2 : 0;  // the skinny kernel computes every tile; the Gram is mirrored element-wise
    } else {
        const size_t smem = (size_t)STAGES * (A_STAGE + BNT_STAGE) * sizeof(double);
        static bool attr = false;
        if (!attr) {
            ASB_CUDA(cudaFuncSetAttribute(gemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr = true;
        }
        gemm_nt_kernel<<<dim3((unsigned)p.tiles, (unsigned)p.slabs), GEMM_THREADS, smem, S(stream)>>>(
            A, lda, ra, B, ldb, rb, M, p.slab_len, symmetric, p.tiles_j, part, rb, slab_stride);
        ASB_LAUNCH_CHECK();
    }
    gemm_nt_reduce_kernel<<<(unsigned)cdiv(ra * rb, 256), 256, 0, S(stream)>>>(part, p.slabs, slab_stride, ra, rb, rb,
                                                                             symmetric, D, ldd);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_gemm_nn(const double* A, int64_t lda, int64_t r, int64_t q, const double* B, int64_t ldb, int64_t M, double alpha,
                const double* row_scale, double* D, int64_t ldd, asb_stream_t stream) {
    ASB_REQUIRE(A && B && D && r > 0 && q > 0 && M > 0, "gemm_nn: bad arguments");
    ASB_REQUIRE(lda % 2 == 0 && lda >= q + (q & 1) && ldb % 2 == 0 && ldd % 2 == 0 && M % 2 == 0,
                "gemm_nn: leading dimensions and M must be even (A zero-padded to an even row length)");
    ASB_REQUIRE(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && ((uintptr_t)D % 16 == 0),
                "gemm_nn: operands must be 16-byte aligned");
    if (r <= 256) {
        const bool wn3 = sk_wn_for(r) == 3;
        switch ((int)cdiv(cdiv(r, 8), SK_WM)) {
#define ASB_NN_CASE(MT) case MT: return wn3 ? launch_nn_skinny<MT, 3>(A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd, S(stream)) : launch_nn_skinny<MT, 2>(A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd, S(stream));
            ASB_NN_CASE(1) ASB_NN_CASE(2) ASB_NN_CASE(3) ASB_NN_CASE(4) ASB_NN_CASE(5) ASB_NN_CASE(6) ASB_NN_CASE(7)
            default: return wn3 ? launch_nn_skinny<8, 3>(A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd, S(stream))
                                : launch_nn_skinny<8, 2>(A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd, S(stream));
#undef ASB_NN_CASE
        }
    }
    ASB_REQUIRE(cdiv(M, BN) < 65536, "gemm_nn: M too large for one launch");
    const size_t smem = (size_t)STAGES * (A_STAGE + BNN_STAGE) * sizeof(double);
    static bool attr = false;
    if (!attr) {
        ASB_CUDA(cudaFuncSetAttribute(gemm_nn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    gemm_nn_kernel<<<dim3((unsigned)cdiv(r, BM), (unsigned)cdiv(M, BN)), GEMM_THREADS, smem, S(stream)>>>(
        A, lda, r, q, B, ldb, M, alpha, row_scale, D, ldd);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // extern "C"
