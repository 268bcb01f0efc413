// k1: per-frame Procrustes alignment to frame 0 (utils/process.py:198-250).
// Streaming passes over the float32 (F, N, 3) animation: centroids, centred 3x3 cross-covariances, a 3x3 SVD per
// frame (one-sided Jacobi, one thread), and the affine apply that writes float32 like the reference's .h5 file.
#include "common.cuh"

namespace asb {

// part[f][chunk][0..2] = sum_n v[f][n][d]
__global__ void align_centroid_kernel(const float* __restrict__ frames, int64_t N, int64_t chunk, double* __restrict__ part) {
    __shared__ double sh[33];
    const int64_t f = blockIdx.y;
    const int64_t n0 = (int64_t)blockIdx.x * chunk, n1 = min(N, n0 + chunk);
    const float* p = frames + f * N * 3;
    double s[3] = {0, 0, 0};
    for (int64_t n = n0 + threadIdx.x; n < n1; n += blockDim.x) {
        s[0] += (double)p[n * 3];
        s[1] += (double)p[n * 3 + 1];
        s[2] += (double)p[n * 3 + 2];
    }
    for (int d = 0; d < 3; ++d) {
        const double t = block_sum(s[d], sh);
        if (threadIdx.x == 0) part[(f * gridDim.x + blockIdx.x) * 3 + d] = t;
    }
}
__global__ void align_centroid_final(const double* __restrict__ part, int64_t chunks, int64_t N, double* __restrict__ cent) {
    const int64_t f = blockIdx.x;
    if (threadIdx.x >= 3) return;
    double s = 0.0;
    for (int64_t c = 0; c < chunks; ++c) s += part[(f * chunks + c) * 3 + threadIdx.x];
    cent[f * 3 + threadIdx.x] = s / (double)N;
}

// part[f][chunk][a*3+b] = sum_n (v0[n][a] - t1[a]) (v_f[n][b] - t0[b])
__global__ void align_cov_kernel(const float* __restrict__ frames, int64_t N, int64_t chunk, const double* __restrict__ cent,
                                 double* __restrict__ part) {
    __shared__ double sh[33];
    const int64_t f = blockIdx.y;
    const int64_t n0 = (int64_t)blockIdx.x * chunk, n1 = min(N, n0 + chunk);
    const float* p0 = frames;
    const float* pf = frames + f * N * 3;
    const double t1[3] = {cent[0], cent[1], cent[2]};
    const double t0[3] = {cent[f * 3], cent[f * 3 + 1], cent[f * 3 + 2]};
    double m[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t n = n0 + threadIdx.x; n < n1; n += blockDim.x) {
        double a[3], b[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            a[d] = (double)p0[n * 3 + d] - t1[d];
            b[d] = (double)pf[n * 3 + d] - t0[d];
        }
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) m[i * 3 + j] = fma(a[i], b[j], m[i * 3 + j]);
    }
    for (int i = 0; i < 9; ++i) {
        const double t = block_sum(m[i], sh);
        if (threadIdx.x == 0) part[(f * gridDim.x + blockIdx.x) * 9 + i] = t;
    }
}
__global__ void align_cov_final(const double* __restrict__ part, int64_t chunks, double* __restrict__ M) {
    const int64_t f = blockIdx.x;
    if (threadIdx.x >= 9) return;
    double s = 0.0;
    for (int64_t c = 0; c < chunks; ++c) s += part[(f * chunks + c) * 9 + threadIdx.x];
    M[f * 9 + threadIdx.x] = s;
}

// One-sided Jacobi SVD of a 3x3 matrix: M = U diag(s) V^T; returns R = U V^T (the orthogonal polar factor).
__device__ inline void polar_rotation3(const double* M, double* R) {
    double a[3][3], v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) a[i][j] = M[i * 3 + j];
    for (int sweep = 0; sweep < 40; ++sweep) {
        double rot = 0.0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double al = 0, be = 0, ga = 0;
                for (int i = 0; i < 3; ++i) {
                    al += a[i][p] * a[i][p];
                    be += a[i][q] * a[i][q];
                    ga += a[i][p] * a[i][q];
                }
                if (fabs(ga) <= 1e-17 * sqrt(al * be) || ga == 0.0) continue;
                rot = fmax(rot, fabs(ga) / sqrt(al * be));
                const double zeta = (be - al) / (2.0 * ga);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int i = 0; i < 3; ++i) {
                    const double x = a[i][p], y = a[i][q];
                    a[i][p] = c * x - s * y;
                    a[i][q] = s * x + c * y;
                    const double vx = v[i][p], vy = v[i][q];
                    v[i][p] = c * vx - s * vy;
                    v[i][q] = s * vx + c * vy;
                }
            }
        if (rot < 1e-16) break;
    }
    // columns of a are u_j * s_j
    double sn[3], u[3][3];
    int order[3] = {0, 1, 2};
    for (int j = 0; j < 3; ++j) sn[j] = sqrt(a[0][j] * a[0][j] + a[1][j] * a[1][j] + a[2][j] * a[2][j]);
    for (int i = 0; i < 2; ++i)
        for (int j = i + 1; j < 3; ++j)
            if (sn[order[j]] > sn[order[i]]) { int t = order[i]; order[i] = order[j]; order[j] = t; }
    const double smax = sn[order[0]];
    int good = 0;
    for (int r = 0; r < 3; ++r) {
        const int j = order[r];
        if (sn[j] > 1e-14 * smax && sn[j] > 0.0) {
            for (int i = 0; i < 3; ++i) u[i][j] = a[i][j] / sn[j];
            ++good;
        }
    }
    if (good == 2) {  // rank 2 (e.g. planar mesh): complete U with the cross product, orientation from V
        const int j0 = order[0], j1 = order[1], j2 = order[2];
        double cx = u[1][j0] * u[2][j1] - u[2][j0] * u[1][j1];
        double cy = u[2][j0] * u[0][j1] - u[0][j0] * u[2][j1];
        double cz = u[0][j0] * u[1][j1] - u[1][j0] * u[0][j1];
        // sign so that det(U) has the sign of det(V) (proper rotation R)
        const double dv = v[0][j0] * (v[1][j1] * v[2][j2] - v[2][j1] * v[1][j2]) - v[1][j0] * (v[0][j1] * v[2][j2] - v[2][j1] * v[0][j2]) +
                          v[2][j0] * (v[0][j1] * v[1][j2] - v[1][j1] * v[0][j2]);
        const double sg = dv >= 0 ? 1.0 : -1.0;
        u[0][j2] = sg * cx; u[1][j2] = sg * cy; u[2][j2] = sg * cz;
    } else if (good < 2) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) u[i][j] = (i == j) ? 1.0 : 0.0, v[i][j] = (i == j) ? 1.0 : 0.0;
    }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[i * 3 + j] = u[i][0] * v[j][0] + u[i][1] * v[j][1] + u[i][2] * v[j][2];
}

__global__ void align_solve_kernel(const double* __restrict__ M, const double* __restrict__ cent, int64_t F, int rigid,
                                   double* __restrict__ T) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    double R[9];
    polar_rotation3(M + f * 9, R);
    const double det = R[0] * (R[4] * R[8] - R[5] * R[7]) - R[1] * (R[3] * R[8] - R[5] * R[6]) + R[2] * (R[3] * R[7] - R[4] * R[6]);
    if (det < 0)
        for (int i = 0; i < 9; ++i) R[i] = -R[i];  // `R *= -1`, utils/process.py:227-228
    const double* t0 = cent + f * 3;
    const double* t1 = cent;
    double* o = T + f * 12;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) o[i * 4 + j] = rigid ? R[i * 3 + j] : (i == j ? 1.0 : 0.0);
        o[i * 4 + 3] = t1[i] - (R[i * 3] * t0[0] + R[i * 3 + 1] * t0[1] + R[i * 3 + 2] * t0[2]);  // uses R even if !rigid
    }
}

__global__ void align_apply_kernel(const float* __restrict__ frames, int64_t N, const double* __restrict__ T,
                                   float* __restrict__ out) {
    const int64_t f = blockIdx.y;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const double* t = T + f * 12;
    const float* p = frames + (f * N + n) * 3;
    const double x = p[0], y = p[1], z = p[2];
    float* o = out + (f * N + n) * 3;
#pragma unroll
    for (int i = 0; i < 3; ++i) o[i] = (float)(t[i * 4] * x + t[i * 4 + 1] * y + t[i * 4 + 2] * z + t[i * 4 + 3]);
}

}  // namespace asb

using namespace asb;

extern "C" {

static void align_chunks(int64_t F, int64_t N, int64_t* chunks, int64_t* chunk) {
    int64_t c = std::max<int64_t>(1, std::min<int64_t>(32, cdiv(4LL * sm_count(), F)));
    c = std::max<int64_t>(1, std::min<int64_t>(c, cdiv(N, 2048)));
    *chunk = cdiv(N, c);
    *chunks = cdiv(N, *chunk);
}

int asb_align_centroids(const float* frames, int64_t F, int64_t N, double* cent, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(frames && cent && ws && F > 0 && F < 65536 && N > 0, "align_centroids: bad arguments");
    int64_t chunks, chunk;
    align_chunks(F, N, &chunks, &chunk);
    double* part = reinterpret_cast<double*>(ws);
    align_centroid_kernel<<<dim3((unsigned)chunks, (unsigned)F), 256, 0, S(stream)>>>(frames, N, chunk, part);
    ASB_LAUNCH_CHECK();
    align_centroid_final<<<(unsigned)F, 32, 0, S(stream)>>>(part, chunks, N, cent);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_align_cov(const float* frames, int64_t F, int64_t N, const double* cent, double* M, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(frames && cent && M && ws && F > 0 && F < 65536 && N > 0, "align_cov: bad arguments");
    int64_t chunks, chunk;
    align_chunks(F, N, &chunks, &chunk);
    double* part = reinterpret_cast<double*>(ws);
    align_cov_kernel<<<dim3((unsigned)chunks, (unsigned)F), 256, 0, S(stream)>>>(frames, N, chunk, cent, part);
    ASB_LAUNCH_CHECK();
    align_cov_final<<<(unsigned)F, 32, 0, S(stream)>>>(part, chunks, M);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_align_solve(const double* M, const double* cent, int64_t F, int rigid, double* T, asb_stream_t stream) {
    ASB_REQUIRE(M && cent && T && F > 0, "align_solve: bad arguments");
    align_solve_kernel<<<(unsigned)cdiv(F, 64), 64, 0, S(stream)>>>(M, cent, F, rigid, T);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_align_apply(const float* frames, int64_t F, int64_t N, const double* T, float* out, asb_stream_t stream) {
    ASB_REQUIRE(frames && T && out && F > 0 && F < 65536 && N > 0, "align_apply: bad arguments");
    align_apply_kernel<<<dim3((unsigned)cdiv(N, 256), (unsigned)F), 256, 0, S(stream)>>>(frames, N, T, out);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // extern "C"
