// FP64 tensor-core (DMMA m8n8k4) contractions of the basis build.
//
//   gemm_nt : D = A B^T, contraction over the long (vertex) axis M  -> small outputs (F x F Gram, X C^T, C C^T)
//   gemm_nn : D = A B,   small left matrix times the wide snapshot/component matrix (c = W^T X, back-projection)
//   admm    : one fused SPLOCS ADMM iteration (K x K times K x 3N GEMM + group-lasso prox + dual update)
//
// sm_100a has no fp64 kind of tcgen05.mma; fp64 tensor work is `mma.sync.aligned.m8n8k4.f64` (SASS DMMA.8x8x4).
// Operand tiles are staged global -> shared with a 3-stage cp.async (LDGSTS) pipeline; shared tiles use padded
// strides (stride mod 16 doubles == 4 or 12) so that every 8x4 / 4x8 fragment read is bank-conflict free.
#include "common.cuh"

namespace asb {

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

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
    if (symmetric && ((j / BN) < (i / BM))) { si = j; sj = i; }  // tile (ti > tj) was not computed
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

// ------------------------------------------------------------------------------------------ fused ADMM iteration
// CTA = 8 vertices (x 3 planes = 24 columns) x all K rows. 4 warps = 4 row groups of MTW m-tiles; each warp owns the
// x, y, z accumulators of the same (k, n) so the group-lasso prox is thread-local.
constexpr int ADMM_BK = 8;
constexpr int ADMM_ALD = ADMM_BK + 4;  // 12
constexpr int ADMM_DLD = 24 + 4;       // 28
template <int MTW>
__global__ void __launch_bounds__(128)
admm_iter_kernel(const double* __restrict__ Hr, int64_t ldh, const double* __restrict__ B0, double* __restrict__ Z,
                 double* __restrict__ U, const double* __restrict__ Lambda, int64_t K, int64_t N, int64_t ld,
                 double inv_rho) {
    constexpr int KP = 4 * MTW * 8;  // padded rows
    constexpr int ASTAGE = KP * ADMM_ALD;
    extern __shared__ __align__(16) double smem[];
    double* Ds = smem;                   // [KP][28]   (Z - U) tile, k-major
    double* As = smem + KP * ADMM_DLD;   // STAGES x [KP][12]
    const int64_t n0 = (int64_t)blockIdx.x * 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int64_t C3 = 3 * ld;

    // stage D = Z - U (rows >= K are zero): KP rows x 3 planes x 4 double2
    for (int t = threadIdx.x; t < KP * 12; t += 128) {
        const int k = t / 12, rem = t % 12, d = rem >> 2, c2 = (rem & 3) * 2;
        double2 v = make_double2(0.0, 0.0);
        if (k < K) {
            const int64_t off = (int64_t)k * C3 + d * ld + n0 + c2;
            const double2 z = ld2(Z + off), u = ld2(U + off);
            v.x = z.x - u.x;
            v.y = z.y - u.y;
        }
        st2(Ds + k * ADMM_DLD + d * 8 + c2, v);
    }
    auto load_a = [&](int stage, int64_t k0) {
        // KP rows x 4 chunks of 16 B
        for (int q = threadIdx.x; q < KP * 4; q += 128) {
            const int row = q >> 2, cc = (q & 3) * 2;
            const bool ok = row < K && k0 + cc < K;
            const double* src = ok ? Hr + (int64_t)row * ldh + k0 + cc : Hr;
            cp_async16(As + stage * ASTAGE + row * ADMM_ALD + cc, src, ok ? 16 : 0);
        }
    };
    const int64_t nk = (K + ADMM_BK - 1) / ADMM_BK;
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) load_a(s, (int64_t)s * ADMM_BK);
        cp_async_commit();
    }
    double acc[MTW][3][2];
#pragma unroll
    for (int i = 0; i < MTW; ++i)
#pragma unroll
        for (int d = 0; d < 3; ++d) acc[i][d][0] = acc[i][d][1] = 0.0;

    for (int64_t kt = 0; kt < nk; ++kt) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();  // also orders the Ds stores before the first fragment read
        const int64_t nx = kt + STAGES - 1;
        if (nx < nk) load_a((int)(nx % STAGES), nx * ADMM_BK);
        cp_async_commit();
        const double* as = As + (int)(kt % STAGES) * ASTAGE;
#pragma unroll
        for (int kk = 0; kk < ADMM_BK / 4; ++kk) {
            const int krow = (int)kt * ADMM_BK + kk * 4 + lc;
            double b[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) b[d] = Ds[krow * ADMM_DLD + d * 8 + lr];
#pragma unroll
            for (int i = 0; i < MTW; ++i) {
                const double a = as[(warp * MTW * 8 + i * 8 + lr) * ADMM_ALD + kk * 4 + lc];
#pragma unroll
                for (int d = 0; d < 3; ++d) dmma(acc[i][d][0], acc[i][d][1], a, b[d]);
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();  // every warp has finished reading Z/U through Ds before anyone overwrites them

    const int64_t n = n0 + 2 * lc;
#pragma unroll
    for (int i = 0; i < MTW; ++i) {
        const int64_t k = (int64_t)warp * MTW * 8 + i * 8 + lr;
        if (k >= K || n >= ld) continue;
        const int64_t base = k * C3 + n;
        double2 v[3], un[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const double2 b0 = ld2(B0 + base + d * ld);
            const double2 u = ld2(U + base + d * ld);
            v[d].x = (b0.x + acc[i][d][0]) + u.x;  // V = C + U
            v[d].y = (b0.y + acc[i][d][1]) + u.y;
        }
        const double2 lam = ld2(Lambda + k * ld + n);
        double sh[2];
        {
            const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(v[0].x, v[0].x), __dmul_rn(v[1].x, v[1].x)), __dmul_rn(v[2].x, v[2].x)));
            const double t = 1.0 - __ddiv_rn(__dmul_rn(inv_rho, lam.x), len);
            sh[0] = (t != t) ? t : fmax(0.0, t);  // np.maximum propagates the 0/0 NaN (posComponents.py:253-256)
        }
        {
            const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(v[0].y, v[0].y), __dmul_rn(v[1].y, v[1].y)), __dmul_rn(v[2].y, v[2].y)));
            const double t = 1.0 - __ddiv_rn(__dmul_rn(inv_rho, lam.y), len);
            sh[1] = (t != t) ? t : fmax(0.0, t);
        }
        const bool ok0 = n < N, ok1 = n + 1 < N;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            double2 z, u2;
            z.x = ok0 ? v[d].x * sh[0] : 0.0;
            z.y = ok1 ? v[d].y * sh[1] : 0.0;
            u2.x = ok0 ? v[d].x - z.x : 0.0;
            u2.y = ok1 ? v[d].y - z.y : 0.0;
            un[d] = u2;
            st2(Z + base + d * ld, z);
        }
#pragma unroll
        for (int d = 0; d < 3; ++d) st2(U + base + d * ld, un[d]);
    }
}

}  // namespace asb

using namespace asb;

static int nt_plan(int64_t ra, int64_t rb, int64_t M, int symmetric, int64_t* tiles, int64_t* slabs, int64_t* slab_len,
                   int* tiles_j) {
    const int64_t ti = cdiv(ra, BM), tj = cdiv(rb, BN);
    *tiles_j = (int)tj;
    *tiles = symmetric ? ti * (ti + 1) / 2 : ti * tj;
    int64_t s = cdiv(2LL * sm_count(), *tiles);
    const int64_t max_by_len = std::max<int64_t>(1, M / (8 * BK));  // at least 8 k-tiles per slab
    s = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(s, 64), max_by_len));
    int64_t len = cdiv(cdiv(M, s), BK) * BK;
    *slab_len = len;
    *slabs = cdiv(M, len);
    return 0;
}

template <int MTW>
static int launch_admm(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                       int64_t K, int64_t N, int64_t ld, double inv_rho, cudaStream_t st) {
    constexpr int KP = 4 * MTW * 8;
    const size_t smem = (size_t)(KP * ADMM_DLD + STAGES * KP * ADMM_ALD) * sizeof(double);
    ASB_CUDA(cudaFuncSetAttribute(admm_iter_kernel<MTW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    admm_iter_kernel<MTW><<<(unsigned)(ld / 8), 128, smem, st>>>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

extern "C" {

int64_t asb_gemm_nt_workspace_bytes(int64_t ra, int64_t rb, int64_t M, int symmetric) {
    int64_t tiles, slabs, len;
    int tj;
    nt_plan(ra, rb, M, symmetric, &tiles, &slabs, &len, &tj);
    return slabs * ra * rb * (int64_t)sizeof(double);
}

int asb_gemm_nt(const double* A, int64_t lda, int64_t ra, const double* B, int64_t ldb, int64_t rb, int64_t M, double* D,
                int64_t ldd, int symmetric, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(A && B && D && ws && ra > 0 && rb > 0 && M > 0, "gemm_nt: bad arguments");
    ASB_REQUIRE(lda % 2 == 0 && ldb % 2 == 0 && M % 2 == 0, "gemm_nt: leading dimensions and M must be even");
    ASB_REQUIRE(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0), "gemm_nt: operands must be 16-byte aligned");
    ASB_REQUIRE(!symmetric || (A == B && ra == rb), "gemm_nt: symmetric needs A == B");
    int64_t tiles, slabs, slab_len;
    int tiles_j;
    nt_plan(ra, rb, M, symmetric, &tiles, &slabs, &slab_len, &tiles_j);
    const size_t smem = (size_t)STAGES * (A_STAGE + BNT_STAGE) * sizeof(double);
    static bool attr = false;
    if (!attr) {
        ASB_CUDA(cudaFuncSetAttribute(gemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    double* part = reinterpret_cast<double*>(ws);
    const int64_t slab_stride = ra * rb;
    gemm_nt_kernel<<<dim3((unsigned)tiles, (unsigned)slabs), GEMM_THREADS, smem, S(stream)>>>(
        A, lda, ra, B, ldb, rb, M, slab_len, symmetric, tiles_j, part, rb, slab_stride);
    ASB_LAUNCH_CHECK();
    gemm_nt_reduce_kernel<<<(unsigned)cdiv(ra * rb, 256), 256, 0, S(stream)>>>(part, slabs, slab_stride, ra, rb, rb,
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

int asb_admm_iter(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda, int64_t K,
                  int64_t N, int64_t ld, double inv_rho, asb_stream_t stream) {
    ASB_REQUIRE(Hr && B0 && Z && U && Lambda && K > 0 && K <= 256 && N > 0 && ld >= N && ld % 32 == 0,
                "admm_iter: bad arguments (K <= 256)");
    ASB_REQUIRE(ldh % 2 == 0 && ((uintptr_t)Hr % 16 == 0), "admm_iter: Hr must be 16-byte aligned with even ldh");
    if (K <= 64) return launch_admm<2>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, S(stream));
    if (K <= 128) return launch_admm<4>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, S(stream));
    if (K <= 224) return launch_admm<7>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, S(stream));
    return launch_admm<8>(Hr, ldh, B0, Z, U, Lambda, K, N, ld, inv_rho, S(stream));
}

}  // extern "C"
