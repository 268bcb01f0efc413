// Heat-method geodesics on the GPU (k6) and the support maps built from them (k7).
// The reference solves two sparse systems with SuperLU per query (utils/support.py:169-171, :182, :205); here the
// SPD heat system and the semi-definite cotan Poisson system are solved by batched preconditioned CG whose building
// blocks are the kernels below (multi right-hand-side CSR SpMV, fixed-order batched dots, fused vector updates).
// The per-triangle gradient / per-vertex divergence use precomputed geometric constants and a vertex->corner CSR so
// that the scatter of utils/support.py:201 (np.bincount) becomes a deterministic gather.
#include "common.cuh"

namespace asb {

__global__ void spmv_csr_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                                const double* __restrict__ vals, int64_t n, const double* __restrict__ x,
                                double* __restrict__ y, int64_t ldx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* xb = x + (int64_t)blockIdx.y * ldx;
    double s = 0.0;
    const int p1 = rowptr[i + 1];
    for (int p = rowptr[i]; p < p1; ++p) s = fma(vals[p], xb[colidx[p]], s);
    y[(int64_t)blockIdx.y * ldx + i] = s;
}

// partial[b][chunk] = sum_{i in chunk} x[b][i] * y[b][i]
__global__ void batched_dot_stage1(const double* __restrict__ x, const double* __restrict__ y, int64_t n, int64_t ldx,
                                   int64_t chunk, double* __restrict__ partial) {
    __shared__ double sh[33];
    const int64_t b = blockIdx.y;
    const int64_t i0 = (int64_t)blockIdx.x * chunk, i1 = min(n, i0 + chunk);
    const double* xb = x + b * ldx;
    const double* yb = y + b * ldx;
    double s = 0.0;
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) s = fma(xb[i], yb[i], s);
    s = block_sum(s, sh);
    if (threadIdx.x == 0) partial[b * gridDim.x + blockIdx.x] = s;
}
__global__ void batched_reduce_stage2(const double* __restrict__ partial, int64_t chunks, double* __restrict__ out,
                                      int op /*0 sum, 1 min*/) {
    const int64_t b = blockIdx.x;
    if (threadIdx.x != 0) return;
    double s = op == 0 ? 0.0 : INFINITY;
    for (int64_t c = 0; c < chunks; ++c) {
        const double v = partial[b * chunks + c];
        s = op == 0 ? s + v : fmin(s, v);
    }
    out[b] = s;
}
__global__ void batched_min_stage1(const double* __restrict__ x, int64_t n, int64_t ldx, int64_t chunk,
                                   double* __restrict__ partial) {
    __shared__ double sh[33];
    const int64_t b = blockIdx.y;
    const int64_t i0 = (int64_t)blockIdx.x * chunk, i1 = min(n, i0 + chunk);
    const double* xb = x + b * ldx;
    double m = -INFINITY;  // max of negated values
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) m = fmax(m, -xb[i]);
    m = block_max(m, sh);
    if (threadIdx.x == 0) partial[b * gridDim.x + blockIdx.x] = -m;
}
__global__ void batched_sum_stage1(const double* __restrict__ x, int64_t n, int64_t ldx, int64_t chunk,
                                   double* __restrict__ partial) {
    __shared__ double sh[33];
    const int64_t b = blockIdx.y;
    const int64_t i0 = (int64_t)blockIdx.x * chunk, i1 = min(n, i0 + chunk);
    const double* xb = x + b * ldx;
    double s = 0.0;
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) s += xb[i];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) partial[b * gridDim.x + blockIdx.x] = s;
}

__global__ void pcg_update_xr_kernel(double* __restrict__ x, double* __restrict__ r, double* __restrict__ z,
                                     const double* __restrict__ p, const double* __restrict__ Ap,
                                     const double* __restrict__ dinv, const double* __restrict__ rz,
                                     const double* __restrict__ pAp, const int32_t* __restrict__ active, int64_t n,
                                     int64_t ldx) {
    const int64_t b = blockIdx.y;
    if (active && !active[b]) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double alpha = pAp[b] > 0.0 ? rz[b] / pAp[b] : 0.0;
    const int64_t o = b * ldx + i;
    x[o] = fma(alpha, p[o], x[o]);
    const double rn = fma(-alpha, Ap[o], r[o]);
    r[o] = rn;
    if (z) z[o] = dinv ? dinv[i] * rn : rn;
}
__global__ void pcg_update_p_kernel(double* __restrict__ p, const double* __restrict__ z,
                                    const double* __restrict__ rz_new, const double* __restrict__ rz_old,
                                    const int32_t* __restrict__ active, int64_t n, int64_t ldx) {
    const int64_t b = blockIdx.y;
    if (active && !active[b]) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double beta = rz_old[b] > 0.0 ? rz_new[b] / rz_old[b] : 0.0;
    const int64_t o = b * ldx + i;
    p[o] = fma(beta, p[o], z[o]);
}
// y[b] = alpha * a[b] .* s + beta * c[b]   (s: per-row scale shared by all rhs, may be NULL)
__global__ void axpby_kernel(double* __restrict__ y, double alpha, const double* __restrict__ a,
                             const double* __restrict__ s, double beta, const double* __restrict__ c, int64_t n,
                             int64_t ldx) {
    const int64_t b = blockIdx.y;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t o = b * ldx + i;
    double v = alpha * a[o];
    if (s) v *= s[i];
    if (c) v = fma(beta, c[o], v);
    y[o] = v;
}
__global__ void shift_kernel(double* __restrict__ x, const double* __restrict__ sums, double inv_n, int64_t n, int64_t ldx) {
    const int64_t b = blockIdx.y;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    x[b * ldx + i] -= sums[b] * inv_n;
}

// Both heat kernels keep HEAT_NB right-hand sides per thread: the per-triangle / per-corner constants (84 B and 64 B) are read
// once per group instead of once per right-hand side (r02 ncu, 40 sources on the 1M-vertex grid: 7.3 GB + 29.3 GB of DRAM
// traffic and 6.6 of the 16.9 ms of the query before).  Per right-hand side the arithmetic and its order are unchanged.
constexpr int HEAT_NB = 8;

__global__ void __launch_bounds__(128) heat_gradient_kernel(const double* __restrict__ u, int64_t ldx, const int32_t* __restrict__ tri,
                                                           const double* __restrict__ gconst, int64_t T, int nrhs,
                                                           double* __restrict__ Xt) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    const int b0 = blockIdx.y * HEAT_NB;
    const int i0 = tri[t * 3], i1 = tri[t * 3 + 1], i2 = tri[t * 3 + 2];
    double g[9];  // (n x e01)/(2A), (n x e12)/(2A), (n x e20)/(2A)
#pragma unroll
    for (int k = 0; k < 9; ++k) g[k] = gconst[t * 9 + k];
#pragma unroll 4
    for (int bb = 0; bb < HEAT_NB; ++bb) {
        const int64_t b = b0 + bb;
        if (b >= nrhs) break;
        const double* ub = u + b * ldx;
        const double u0 = ub[i0], u1 = ub[i1], u2 = ub[i2];
        double gx = g[0] * u2 + g[3] * u0 + g[6] * u1;
        double gy = g[1] * u2 + g[4] * u0 + g[7] * u1;
        double gz = g[2] * u2 + g[5] * u0 + g[8] * u1;
        const double len = sqrt(gx * gx + gy * gy + gz * gz);
        double* o = Xt + (b * T + t) * 3;
        o[0] = -gx / len;
        o[1] = -gy / len;
        o[2] = -gz / len;
    }
}
__global__ void __launch_bounds__(128) heat_divergence_kernel(const double* __restrict__ Xt, int64_t T, const double* __restrict__ dconst,
                                                             const int32_t* __restrict__ vc_ptr, const int32_t* __restrict__ vc_idx,
                                                             int64_t n, int nrhs, double* __restrict__ div, int64_t ldx) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int b0 = blockIdx.y * HEAT_NB;
    const int nb = min(HEAT_NB, nrhs - b0);
    double s[HEAT_NB];
#pragma unroll
    for (int bb = 0; bb < HEAT_NB; ++bb) s[bb] = 0.0;
    const int p1 = vc_ptr[v + 1];
    for (int p = vc_ptr[v]; p < p1; ++p) {
        const int corner = vc_idx[p];  // t*3 + c
        double c[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) c[k] = dconst[(int64_t)corner * 8 + k];
        const double* X0 = Xt + ((int64_t)b0 * T + corner / 3) * 3;
#pragma unroll
        for (int bb = 0; bb < HEAT_NB; ++bb) {
            if (bb < nb) {
                const double* X = X0 + (int64_t)bb * T * 3;
                const double d1 = c[1] * X[0] + c[2] * X[1] + c[3] * X[2];
                const double d2 = c[5] * X[0] + c[6] * X[1] + c[7] * X[2];
                s[bb] += 0.5 * (c[0] * d1 + c[4] * d2);
            }
        }
    }
#pragma unroll
    for (int bb = 0; bb < HEAT_NB; ++bb)
        if (bb < nb) div[(int64_t)(b0 + bb) * ldx + v] = s[bb];
}

__global__ void support_map_kernel(const double* __restrict__ phi, const double* __restrict__ mins, int64_t ldx,
                                   double dmin, double dmax, int mode, double lambda, int64_t row0, int64_t nloc,
                                   double* __restrict__ out, int64_t ldo) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ldo) return;
    const int64_t b = blockIdx.y;
    double v = 0.0;
    if (i < nloc) {
        const double p = phi[b * ldx + row0 + i] - mins[b];
        if (mode == 2) {
            v = p;
        } else {
            const double m = (fmin(fmax(p, dmin), dmax) - dmin) / (dmax - dmin);
            v = mode == 0 ? 1.0 - m : lambda * m;
        }
    }
    out[b * ldo + i] = v;
}

__global__ void delta_rhs_kernel(const int64_t* __restrict__ idx, int64_t n, int64_t ldx, double* __restrict__ rhs) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ldx) return;
    const int64_t b = blockIdx.y;
    rhs[b * ldx + i] = (i < n && i == idx[b]) ? 1.0 : 0.0;
}

inline void chunks_for(int64_t n, int64_t nrhs, int64_t* chunks, int64_t* chunk) {
    int64_t c = std::max<int64_t>(1, std::min<int64_t>(64, cdiv(2LL * sm_count(), nrhs)));
    c = std::min<int64_t>(c, cdiv(n, 1024));
    c = std::max<int64_t>(c, 1);
    *chunk = cdiv(n, c);
    *chunks = cdiv(n, *chunk);
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_spmv_csr(const int32_t* rowptr, const int32_t* colidx, const double* vals, int64_t n, const double* x, double* y,
                 int64_t nrhs, int64_t ldx, asb_stream_t stream) {
    ASB_REQUIRE(rowptr && colidx && vals && x && y && n > 0 && nrhs > 0 && nrhs < 65536 && ldx >= n, "spmv: bad arguments");
    spmv_csr_kernel<<<dim3((unsigned)cdiv(n, 128), (unsigned)nrhs), 128, 0, S(stream)>>>(rowptr, colidx, vals, n, x, y, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_batched_dot(const double* x, const double* y, int64_t n, int64_t nrhs, int64_t ldx, double* dots, void* ws,
                    asb_stream_t stream) {
    ASB_REQUIRE(x && y && dots && ws && n > 0 && nrhs > 0 && nrhs < 65536, "batched_dot: bad arguments");
    int64_t chunks, chunk;
    chunks_for(n, nrhs, &chunks, &chunk);
    double* partial = reinterpret_cast<double*>(ws);
    batched_dot_stage1<<<dim3((unsigned)chunks, (unsigned)nrhs), 256, 0, S(stream)>>>(x, y, n, ldx, chunk, partial);
    ASB_LAUNCH_CHECK();
    batched_reduce_stage2<<<(unsigned)nrhs, 32, 0, S(stream)>>>(partial, chunks, dots, 0);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_pcg_update_xr(double* x, double* r, double* z, const double* p, const double* Ap, const double* dinv,
                      const double* rz, const double* pAp, const int32_t* active, int64_t n, int64_t nrhs, int64_t ldx,
                      asb_stream_t stream) {
    ASB_REQUIRE(x && r && p && Ap && rz && pAp && n > 0 && nrhs > 0 && nrhs < 65536, "pcg_update_xr: bad arguments");
    pcg_update_xr_kernel<<<dim3((unsigned)cdiv(n, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(x, r, z, p, Ap, dinv, rz, pAp,
                                                                                         active, n, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_pcg_update_p(double* p, const double* z, const double* rz_new, const double* rz_old, const int32_t* active,
                     int64_t n, int64_t nrhs, int64_t ldx, asb_stream_t stream) {
    ASB_REQUIRE(p && z && rz_new && rz_old && n > 0 && nrhs > 0 && nrhs < 65536, "pcg_update_p: bad arguments");
    pcg_update_p_kernel<<<dim3((unsigned)cdiv(n, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(p, z, rz_new, rz_old, active, n, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_axpby(double* y, double alpha, const double* a, const double* row_scale, double beta, const double* c, int64_t n,
              int64_t nrhs, int64_t ldx, asb_stream_t stream) {
    ASB_REQUIRE(y && a && n > 0 && nrhs > 0 && nrhs < 65536, "axpby: bad arguments");
    axpby_kernel<<<dim3((unsigned)cdiv(n, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(y, alpha, a, row_scale, beta, c, n, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_remove_mean(double* x, int64_t n, int64_t nrhs, int64_t ldx, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(x && ws && n > 0 && nrhs > 0 && nrhs < 65536, "remove_mean: bad arguments");
    int64_t chunks, chunk;
    chunks_for(n, nrhs, &chunks, &chunk);
    double* partial = reinterpret_cast<double*>(ws);
    double* sums = partial + chunks * nrhs;
    batched_sum_stage1<<<dim3((unsigned)chunks, (unsigned)nrhs), 256, 0, S(stream)>>>(x, n, ldx, chunk, partial);
    ASB_LAUNCH_CHECK();
    batched_reduce_stage2<<<(unsigned)nrhs, 32, 0, S(stream)>>>(partial, chunks, sums, 0);
    ASB_LAUNCH_CHECK();
    shift_kernel<<<dim3((unsigned)cdiv(n, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(x, sums, 1.0 / (double)n, n, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_heat_divergence(const double* u, int64_t nrhs, int64_t ldx, const int32_t* tri, const double* gconst,
                        const double* dconst, int64_t T, const int32_t* vc_ptr, const int32_t* vc_idx, int64_t n,
                        double* Xt, double* div, asb_stream_t stream) {
    ASB_REQUIRE(u && tri && gconst && dconst && vc_ptr && vc_idx && Xt && div && T > 0 && n > 0 && nrhs > 0 && nrhs < 65536,
                "heat_divergence: bad arguments");
    const unsigned groups = (unsigned)cdiv(nrhs, HEAT_NB);
    heat_gradient_kernel<<<dim3((unsigned)cdiv(T, 128), groups), 128, 0, S(stream)>>>(u, ldx, tri, gconst, T, (int)nrhs, Xt);
    ASB_LAUNCH_CHECK();
    heat_divergence_kernel<<<dim3((unsigned)cdiv(n, 128), groups), 128, 0, S(stream)>>>(Xt, T, dconst, vc_ptr, vc_idx, n, (int)nrhs,
                                                                                     div, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_support_map(const double* phi, int64_t n, int64_t nrhs, int64_t ldx, double dmin, double dmax, int mode,
                    double lambda, int64_t row0, int64_t nloc, double* out, int64_t ldo, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(phi && out && ws && n > 0 && nrhs > 0 && nrhs < 65536 && row0 >= 0 && row0 + nloc <= n && ldo >= nloc,
                "support_map: bad arguments");
    int64_t chunks, chunk;
    chunks_for(n, nrhs, &chunks, &chunk);
    double* partial = reinterpret_cast<double*>(ws);
    double* mins = partial + chunks * nrhs;
    batched_min_stage1<<<dim3((unsigned)chunks, (unsigned)nrhs), 256, 0, S(stream)>>>(phi, n, ldx, chunk, partial);
    ASB_LAUNCH_CHECK();
    batched_reduce_stage2<<<(unsigned)nrhs, 32, 0, S(stream)>>>(partial, chunks, mins, 1);
    ASB_LAUNCH_CHECK();
    support_map_kernel<<<dim3((unsigned)cdiv(ldo, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(phi, mins, ldx, dmin, dmax, mode,
                                                                                           lambda, row0, nloc, out, ldo);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_delta_rhs(const int64_t* idx, int64_t n, int64_t nrhs, int64_t ldx, double* rhs, asb_stream_t stream) {
    ASB_REQUIRE(idx && rhs && n > 0 && nrhs > 0 && nrhs < 65536 && ldx >= n, "delta_rhs: bad arguments");
    delta_rhs_kernel<<<dim3((unsigned)cdiv(ldx, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(idx, n, ldx, rhs);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // extern "C"
