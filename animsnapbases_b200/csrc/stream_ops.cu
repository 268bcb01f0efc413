// HBM-bound streaming kernels: boundary layout conversion, mean-centering/standardisation (k2),
// greedy-deflation passes (k3, k8, k9).  All are coalesced planar passes; reductions are fixed-order.
#include <cstdlib>
#include <mutex>
#include "common.cuh"

namespace asb {

thread_local char g_err[512] = {0};
std::atomic<long long> g_launches{0};

int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

// ------------------------------------------------------------------------------------------ ingest / export
template <typename TIn>
__global__ void ingest_kernel(const TIn* __restrict__ src, const double* __restrict__ row_scale,
                              double* __restrict__ X, int64_t N, int64_t ld) {
    const int64_t r = blockIdx.y;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= ld) return;
    double x = 0, y = 0, z = 0;
    if (n < N) {
        const TIn* p = src + (r * N + n) * 3;
        const double s = row_scale ? row_scale[n] : 1.0;
        x = (double)p[0] * s;
        y = (double)p[1] * s;
        z = (double)p[2] * s;
    }
    double* o = X + r * 3 * ld + n;
    o[0] = x;
    o[ld] = y;
    o[2 * ld] = z;
}

__global__ void export_aos_kernel(const double* __restrict__ T, const double* __restrict__ row_scale,
                                  double* __restrict__ dst, int64_t N, int64_t ld) {
    const int64_t r = blockIdx.y;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const double* p = T + r * 3 * ld + n;
    const double s = row_scale ? row_scale[n] : 1.0;
    double* o = dst + (r * N + n) * 3;
    if (row_scale) {
        o[0] = p[0] * s;
        o[1] = p[ld] * s;
        o[2] = p[2 * ld] * s;
    } else {
        o[0] = p[0];
        o[1] = p[ld];
        o[2] = p[2 * ld];
    }
}

__global__ void pack_bin_kernel(const double* __restrict__ T, double* __restrict__ out, int64_t K, int64_t N,
                                int64_t ld) {
    // out[d][k][i] = T[k][d][i]
    const int64_t k = blockIdx.y, d = blockIdx.z;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    out[(d * K + k) * N + i] = T[(k * 3 + d) * ld + i];
}

// ------------------------------------------------------------------------------------------ centering
__global__ void col_mean_kernel(const double* __restrict__ X, double* __restrict__ mean, int64_t F, int64_t C) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s = 0.0;
    for (int64_t f = 0; f < F; ++f) s += X[f * C + c];  // sequential in f, like np.mean(axis=0)
    mean[c] = s / (double)F;
}

// partial[block][0..1] = sum (x-m), sum (x-m)^2 over a contiguous range of frames
__global__ void center_stats_kernel(const double* __restrict__ X, const double* __restrict__ mean, int64_t F,
                                    int64_t C, int64_t frames_per_block, double* __restrict__ partial) {
    __shared__ double sh[33];
    const int64_t f0 = (int64_t)blockIdx.y * frames_per_block;
    const int64_t f1 = min(F, f0 + frames_per_block);
    double s1 = 0.0, s2 = 0.0;
    for (int64_t c = 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x); c < C; c += 2 * (int64_t)gridDim.x * blockDim.x) {
        const double2 m = ld2(mean + c);
        for (int64_t f = f0; f < f1; ++f) {
            const double2 x = ld2_stream(X + f * C + c);
            const double a = x.x - m.x, b = x.y - m.y;
            s1 += a;
            s1 += b;
            s2 = fma(a, a, s2);
            s2 = fma(b, b, s2);
        }
    }
    const double t1 = block_sum(s1, sh);
    const double t2 = block_sum(s2, sh);
    if (threadIdx.x == 0) {
        const int64_t b = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;
        partial[2 * b] = t1;
        partial[2 * b + 1] = t2;
    }
}

// out[j] = sum_b partial[b*stride + j], j < stride, fixed order, single block
__global__ void sum_partials_kernel(const double* __restrict__ partial, int64_t nb, int stride, double* __restrict__ out) {
    __shared__ double sh[33];
    for (int j = 0; j < stride; ++j) {
        double s = 0.0;
        for (int64_t b = threadIdx.x; b < nb; b += blockDim.x) s += partial[b * stride + j];
        s = block_sum(s, sh);
        if (threadIdx.x == 0) out[j] = s;
    }
}

__global__ void center_apply_kernel(double* __restrict__ X, const double* __restrict__ mean, double scale,
                                    int64_t F, int64_t C) {
    const int64_t f = blockIdx.y;
    for (int64_t c = 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x); c < C; c += 2 * (int64_t)gridDim.x * blockDim.x) {
        const double2 m = ld2(mean + c);
        double2 x = ld2(X + f * C + c);
        x.x = __dmul_rn(__dsub_rn(x.x, m.x), scale);
        x.y = __dmul_rn(__dsub_rn(x.y, m.y), scale);
        st2(X + f * C + c, x);
    }
}

__global__ void unstandardize_kernel(double* __restrict__ T, const double* __restrict__ mean, double scale,
                                     int64_t N, int64_t ld) {
    const int64_t r = blockIdx.y;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        double* p = T + (r * 3 + d) * ld + n;
        *p = __dadd_rn(__ddiv_rn(*p, scale), mean[d * ld + n]);
    }
}

// ------------------------------------------------------------------------------------------ deflation
// e_part[fs][n] = sum_{f in slab fs} ((x^2 + y^2) + z^2)
__global__ void vertex_energy_kernel(const double* __restrict__ R, double* __restrict__ e_part, int64_t F,
                                     int64_t N, int64_t ld, int64_t frames_per_slab) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= ld) return;
    const int64_t f0 = (int64_t)blockIdx.y * frames_per_slab, f1 = min(F, f0 + frames_per_slab);
    double e = 0.0;
    if (n < N) {
        const double* p = R + f0 * 3 * ld + n;
        for (int64_t f = f0; f < f1; ++f, p += 3 * ld) {
            const double x = p[0], y = p[ld], z = p[2 * ld];
            e += __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
        }
    }
    e_part[(int64_t)blockIdx.y * ld + n] = e;
}

__global__ void sum_slabs_kernel(const double* __restrict__ part, double* __restrict__ out, int64_t n, int64_t slabs) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double s = part[i];
    for (int64_t k = 1; k < slabs; ++k) s += part[k * n + i];
    out[i] = s;
}

// out[j] = sum of `group` consecutive entries (fixed ascending order)
__global__ void group_sum_kernel(const double* __restrict__ v, int64_t ng, int64_t group, double* __restrict__ out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ng) return;
    double s = 0.0;
    for (int64_t i = 0; i < group; ++i) s = __dadd_rn(s, v[j * group + i]);
    out[j] = s;
}

__global__ void argmax_stage1_kernel(const double* __restrict__ v, int64_t n, int64_t chunk, double* __restrict__ pv,
                                     long long* __restrict__ pi) {
    __shared__ double shv[33];
    __shared__ long long shi[33];
    const int64_t i0 = (int64_t)blockIdx.x * chunk, i1 = min(n, i0 + chunk);
    ArgMax a;
    a.v = -INFINITY;
    a.i = 0x7fffffffffffffffLL;
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        const double x = v[i];
        if (x > a.v) { a.v = x; a.i = i; }
    }
    a = block_argmax(a, shv, shi);
    if (threadIdx.x == 0) { pv[blockIdx.x] = a.v; pi[blockIdx.x] = a.i; }
}
// stage 1 of the masked variant: only entries whose vertex lies OUTSIDE the ball |pos[i] - pos[centre]| <= radius compete
__global__ void argmax_outside_stage1_kernel(const double* __restrict__ v, int64_t n, int64_t chunk, int64_t index_offset,
                                             const double* __restrict__ pos, const int64_t* __restrict__ centre, double r2,
                                             double* __restrict__ pv, long long* __restrict__ pi) {
    __shared__ double shv[33];
    __shared__ long long shi[33];
    const int64_t i0 = (int64_t)blockIdx.x * chunk, i1 = min(n, i0 + chunk);
    const int64_t c = centre[0];
    const double cx = pos[3 * c], cy = pos[3 * c + 1], cz = pos[3 * c + 2];
    ArgMax a;
    a.v = -INFINITY;
    a.i = 0x7fffffffffffffffLL;
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        const double* p = pos + 3 * (i + index_offset);
        const double dx = p[0] - cx, dy = p[1] - cy, dz = p[2] - cz;
        const double x = v[i];
        if (dx * dx + dy * dy + dz * dz > r2 && x > a.v) { a.v = x; a.i = i; }
    }
    a = block_argmax(a, shv, shi);
    if (threadIdx.x == 0) { pv[blockIdx.x] = a.v; pi[blockIdx.x] = a.i; }
}
__global__ void argmax_stage2_kernel(const double* __restrict__ pv, const long long* __restrict__ pi, int64_t nb,
                                     int64_t index_offset, double* __restrict__ out_val, int64_t* __restrict__ out_idx) {
    __shared__ double shv[33];
    __shared__ long long shi[33];
    ArgMax a;
    a.v = -INFINITY;
    a.i = 0x7fffffffffffffffLL;
    for (int64_t b = threadIdx.x; b < nb; b += blockDim.x) {
        ArgMax t;
        t.v = pv[b];
        t.i = pi[b];
        a = argmax_combine(a, t);
    }
    a = block_argmax(a, shv, shi);
    if (threadIdx.x == 0) {
        out_val[0] = a.v;
        out_idx[0] = a.i == 0x7fffffffffffffffLL ? (int64_t)a.i : (int64_t)a.i + index_offset;   // no candidate: INT64_MAX
    }
}

// second stage of the masked argmax: (value, global index bit pattern) as ONE 16-byte record, ready for an all-gather
__global__ void argmax_pair_stage2_kernel(const double* __restrict__ pv, const long long* __restrict__ pi, int64_t nb,
                                          int64_t index_offset, double* __restrict__ out_pair) {
    __shared__ double shv[33];
    __shared__ long long shi[33];
    ArgMax a;
    a.v = -INFINITY;
    a.i = 0x7fffffffffffffffLL;
    for (int64_t b = threadIdx.x; b < nb; b += blockDim.x) {
        ArgMax t;
        t.v = pv[b];
        t.i = pi[b];
        a = argmax_combine(a, t);
    }
    a = block_argmax(a, shv, shi);
    if (threadIdx.x == 0) {
        out_pair[0] = a.v;
        const long long gi = a.i == 0x7fffffffffffffffLL ? a.i : a.i + index_offset;   // no candidate: INT64_MAX
        out_pair[1] = __longlong_as_double(gi);
    }
}
// largest value over `world` records, ties to the lowest index (np.argmax over the concatenated shards); no candidate
// anywhere (-inf): the fallback index
__global__ void select_pair_kernel(const double* __restrict__ all, int world, const int64_t* __restrict__ fallback,
                                   int64_t* __restrict__ out_idx) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double bv = -INFINITY;
    long long bi = 0x7fffffffffffffffLL;
    for (int r = 0; r < world; ++r) {
        const double v = all[2 * r];
        const long long i = __double_as_longlong(all[2 * r + 1]);
        if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
    }
    out_idx[0] = (bv == -INFINITY || bi == 0x7fffffffffffffffLL) ? fallback[0] : (int64_t)bi;
}

// g_part[fs][c] = sum_{f in slab} w[f] R[f][c]; two columns per thread, 8 frames in flight
constexpr int GEMV_WCHUNK = 2048;
__global__ void __launch_bounds__(256) deflate_gemv_kernel(const double* __restrict__ R, int64_t F, int64_t C,
                                                          const double* __restrict__ w, int64_t ldw,
                                                          int64_t frames_per_slab, double* __restrict__ g_part) {
    __shared__ double ws[GEMV_WCHUNK];
    const int64_t c = 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x);
    const int64_t f0 = (int64_t)blockIdx.y * frames_per_slab, f1 = min(F, f0 + frames_per_slab);
    double2 acc = make_double2(0.0, 0.0);
    for (int64_t fb = f0; fb < f1; fb += GEMV_WCHUNK) {
        const int64_t fe = min(f1, fb + GEMV_WCHUNK);
        __syncthreads();
        for (int64_t f = fb + threadIdx.x; f < fe; f += blockDim.x) ws[f - fb] = w[f * ldw];
        __syncthreads();
        if (c < C) {
            const double* p = R + fb * C + c;
            int64_t f = fb;
            for (; f + 8 <= fe; f += 8, p += 8 * C) {
                double2 x[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) x[u] = ld2_stream(p + u * C);
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const double wf = ws[f - fb + u];
                    acc.x = fma(wf, x[u].x, acc.x);
                    acc.y = fma(wf, x[u].y, acc.y);
                }
            }
            for (; f < fe; ++f, p += C) {
                const double2 x = ld2_stream(p);
                const double wf = ws[f - fb];
                acc.x = fma(wf, x.x, acc.x);
                acc.y = fma(wf, x.y, acc.y);
            }
        }
    }
    if (c < C) st2(g_part + (int64_t)blockIdx.y * C + c, acc);
}

// finish component: see asb_deflate_finish. g holds `slabs` partial slabs of 3*ld each when slabs > 1.
__global__ void deflate_finish_kernel(const double* __restrict__ g, int64_t slabs, const double* __restrict__ Cp,
                                      int64_t npend, const double* __restrict__ coef, const double* __restrict__ s,
                                      const double* __restrict__ scal, double* __restrict__ cout,
                                      double* __restrict__ e, int64_t N, int64_t ld, double* __restrict__ partial) {
    __shared__ double sh[33];
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const double ww = scal[1];
    double en = 0.0;
    if (n < ld) {
        double gd[3] = {0, 0, 0};
        if (n < N) {
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                double a = g[d * ld + n];
                for (int64_t k = 1; k < slabs; ++k) a += g[k * 3 * ld + d * ld + n];
                for (int64_t j = 0; j < npend; ++j) a -= coef[j] * Cp[(j * 3 + d) * ld + n];
                gd[d] = a;
            }
            const double sn = s ? s[n] : 1.0;
            const double g2 = gd[0] * gd[0] + gd[1] * gd[1] + gd[2] * gd[2];
#pragma unroll
            for (int d = 0; d < 3; ++d) cout[d * ld + n] = s ? __ddiv_rn(__dmul_rn(gd[d], sn), ww) : __ddiv_rn(gd[d], ww);
            en = e[n] - (2.0 * sn - sn * sn) * g2 / ww;
            if (en < 0.0) en = 0.0;
            e[n] = en;
        } else {
#pragma unroll
            for (int d = 0; d < 3; ++d) cout[d * ld + n] = 0.0;
        }
    }
    const double t = block_sum(en, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// meas[0] = sigma0, meas[2] = sum_b partial[b], meas[1] = sqrt(meas[2])
__global__ void deflate_meas_kernel(const double* __restrict__ partial, int64_t nb, const double* __restrict__ scal,
                                    double* __restrict__ meas) {
    __shared__ double sh[33];
    double s = 0.0;
    for (int64_t b = threadIdx.x; b < nb; b += blockDim.x) s += partial[b];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) {
        meas[0] = scal[0];
        meas[1] = sqrt(s);
        meas[2] = s;
    }
}

// R[f][.][n] -= sum_j Wb[f][j] * Cb[j][.][n]; exact energies of the new R. One vertex per thread.
template <int NB, int FU = 4>
__global__ void __launch_bounds__(128) rank_update_kernel(double* __restrict__ R, int64_t F, int64_t N, int64_t ld,
                                                         const double* __restrict__ Wb, int64_t ldw,
                                                         const double* __restrict__ Cb, int nb,
                                                         int64_t frames_per_slab, double* __restrict__ e_part) {
    constexpr int FCH = 128;
    __shared__ double wsh[FCH * NB];
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t f0 = (int64_t)blockIdx.y * frames_per_slab, f1 = min(F, f0 + frames_per_slab);
    const bool act = n < N;
    double c[NB][3];
#pragma unroll
    for (int j = 0; j < NB; ++j)
#pragma unroll
        for (int d = 0; d < 3; ++d) c[j][d] = (act && j < nb) ? Cb[((int64_t)j * 3 + d) * ld + n] : 0.0;
    double e = 0.0;
    for (int64_t fb = f0; fb < f1; fb += FCH) {
        const int64_t fe = min(f1, fb + FCH);
        __syncthreads();
        for (int64_t t = threadIdx.x; t < (fe - fb) * NB; t += blockDim.x) {
            const int64_t f = fb + t / NB;
            const int j = (int)(t % NB);
            wsh[t] = j < nb ? Wb[f * ldw + j] : 0.0;
        }
        __syncthreads();
        if (act) {
            // FU frames per trip with all 3 * FU loads issued before the arithmetic: one frame per trip left 3 loads in flight
            // per thread (r01 ncu: 31 % occupancy, 4.5 TB/s; FU = 4: 5.02 TB/s, FU = 8: 5.34 TB/s at NB = 8); the per-element operation order and the order of the energy
            // sum over f are unchanged
            double* p = R + fb * 3 * ld + n;
            int64_t f = fb;
            auto one = [&](double x, double y, double z, const double* wr, double* q) {
#pragma unroll
                for (int j = 0; j < NB; ++j) {
                    // `R -= outer(wk, ck)`: product rounded, then subtracted (no fma), j ascending
                    const double wj = wr[j];
                    x = __dsub_rn(x, __dmul_rn(wj, c[j][0]));
                    y = __dsub_rn(y, __dmul_rn(wj, c[j][1]));
                    z = __dsub_rn(z, __dmul_rn(wj, c[j][2]));
                }
                q[0] = x;
                q[ld] = y;
                q[2 * ld] = z;
                e += __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
            };
            for (; f + FU <= fe; f += FU, p += FU * 3 * ld) {
                double v[FU][3];
#pragma unroll
                for (int u = 0; u < FU; ++u)
#pragma unroll
                    for (int d = 0; d < 3; ++d) v[u][d] = p[(u * 3 + d) * ld];
#pragma unroll
                for (int u = 0; u < FU; ++u) one(v[u][0], v[u][1], v[u][2], wsh + (f - fb + u) * NB, p + u * 3 * ld);
            }
            for (; f < fe; ++f, p += 3 * ld) one(p[0], p[ld], p[2 * ld], wsh + (f - fb) * NB, p);
        }
    }
    if (n < ld) e_part[(int64_t)blockIdx.y * ld + n] = e;
}

// choose how many frame slabs to split a column-parallel pass into so that the grid fills the machine
inline int64_t pick_slabs(int64_t col_blocks, int64_t F, int64_t max_slabs) {
    const int64_t target = 4LL * sm_count();
    int64_t s = 1;
    if (col_blocks < target) s = cdiv(target, col_blocks);
    if (s > max_slabs) s = max_slabs;
    if (s > F) s = F;
    if (s < 1) s = 1;
    return s;
}

int sum_partials(const double* partial, int64_t nb, int stride, double* out, cudaStream_t st) {
    sum_partials_kernel<<<1, 256, 0, st>>>(partial, nb, stride, out);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_version(void) { return 100; }
const char* asb_last_error(void) { return g_err; }
int64_t asb_launch_count(void) { return (int64_t)g_launches.load(); }
void asb_reset_launch_count(void) { g_launches.store(0); }
void asb_add_launch_count(int64_t n) { g_launches.fetch_add((long long)n, std::memory_order_relaxed); }

int asb_device_info(int* sms, int* major, int* minor) {
    int dev = 0;
    ASB_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp p;
    ASB_CUDA(cudaGetDeviceProperties(&p, dev));
    if (sms) *sms = p.multiProcessorCount;
    if (major) *major = p.major;
    if (minor) *minor = p.minor;
    return ASB_OK;
}

// generic scratch: up to 64 frame slabs of `cols` doubles, plus block partials
int64_t asb_workspace_bytes(int64_t rows, int64_t cols) {
    (void)rows;
    return (64 * cols + 65536) * (int64_t)sizeof(double);
}

int asb_ingest_f32(const float* src, const double* row_scale, double* X, int64_t R, int64_t N, int64_t ld,
                   asb_stream_t stream) {
    ASB_REQUIRE(src && X && R > 0 && N > 0 && ld >= N && ld % 32 == 0 && R < 65536, "ingest: bad arguments");
    dim3 grid((unsigned)cdiv(ld, 256), (unsigned)R);
    ingest_kernel<float><<<grid, 256, 0, S(stream)>>>(src, row_scale, X, N, ld);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}
int asb_ingest_f64(const double* src, const double* row_scale, double* X, int64_t R, int64_t N, int64_t ld,
                   asb_stream_t stream) {
    ASB_REQUIRE(src && X && R > 0 && N > 0 && ld >= N && ld % 32 == 0 && R < 65536, "ingest: bad arguments");
    dim3 grid((unsigned)cdiv(ld, 256), (unsigned)R);
    ingest_kernel<double><<<grid, 256, 0, S(stream)>>>(src, row_scale, X, N, ld);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}
int asb_export_aos(const double* T, const double* row_scale, double* dst, int64_t R, int64_t N, int64_t ld,
                   asb_stream_t stream) {
    ASB_REQUIRE(T && dst && R > 0 && N > 0 && ld >= N && R < 65536, "export: bad arguments");
    dim3 grid((unsigned)cdiv(N, 256), (unsigned)R);
    export_aos_kernel<<<grid, 256, 0, S(stream)>>>(T, row_scale, dst, N, ld);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}
int asb_pack_bin(const double* T, double* out, int64_t K, int64_t N, int64_t ld, asb_stream_t stream) {
    ASB_REQUIRE(T && out && K > 0 && N > 0 && ld >= N && K < 65536, "pack_bin: bad arguments");
    dim3 grid((unsigned)cdiv(N, 256), (unsigned)K, 3);
    pack_bin_kernel<<<grid, 256, 0, S(stream)>>>(T, out, K, N, ld);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_col_mean(const double* X, double* mean, int64_t F, int64_t ld, asb_stream_t stream) {
    ASB_REQUIRE(X && mean && F > 0 && ld > 0, "col_mean: bad arguments");
    const int64_t C = 3 * ld;
    col_mean_kernel<<<(unsigned)cdiv(C, 128), 128, 0, S(stream)>>>(X, mean, F, C);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_center_stats(const double* X, const double* mean, int64_t F, int64_t N, int64_t ld, double* out2, void* ws,
                     asb_stream_t stream) {
    ASB_REQUIRE(X && mean && out2 && ws && F > 0 && N > 0 && ld % 32 == 0, "center_stats: bad arguments");
    const int64_t C = 3 * ld;
    int64_t gx = std::min<int64_t>(cdiv(C / 2, 256), 2LL * sm_count());
    int64_t gy = std::min<int64_t>(F, std::max<int64_t>(1, 8LL * sm_count() / gx));
    const int64_t fpb = cdiv(F, gy);
    gy = cdiv(F, fpb);
    double* partial = reinterpret_cast<double*>(ws);
    ASB_REQUIRE(gx * gy * 2 <= 65536, "center_stats: workspace");
    center_stats_kernel<<<dim3((unsigned)gx, (unsigned)gy), 256, 0, S(stream)>>>(X, mean, F, C, fpb, partial);
    ASB_LAUNCH_CHECK();
    return sum_partials(partial, gx * gy, 2, out2, S(stream));
}

int asb_center_apply(double* X, const double* mean, double scale, int64_t F, int64_t N, int64_t ld,
                     asb_stream_t stream) {
    ASB_REQUIRE(X && mean && F > 0 && F < 65536 && N > 0 && ld % 32 == 0, "center_apply: bad arguments");
    const int64_t C = 3 * ld;
    const int64_t gx = std::min<int64_t>(cdiv(C / 2, 256), 64);
    center_apply_kernel<<<dim3((unsigned)gx, (unsigned)F), 256, 0, S(stream)>>>(X, mean, scale, F, C);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_unstandardize(double* T, const double* mean, double scale, int64_t R, int64_t N, int64_t ld,
                      asb_stream_t stream) {
    ASB_REQUIRE(T && mean && R > 0 && R < 65536 && N > 0 && ld >= N, "unstandardize: bad arguments");
    unstandardize_kernel<<<dim3((unsigned)cdiv(N, 256), (unsigned)R), 256, 0, S(stream)>>>(T, mean, scale, N, ld);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_vertex_energy(const double* R, double* e, int64_t F, int64_t N, int64_t ld, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(R && e && ws && F > 0 && N > 0 && ld >= N, "vertex_energy: bad arguments");
    const int64_t gx = cdiv(ld, 128);
    const int64_t slabs = pick_slabs(gx, F, 64);
    const int64_t fps = cdiv(F, slabs);
    const int64_t ns = cdiv(F, fps);
    double* part = ns == 1 ? e : reinterpret_cast<double*>(ws);
    vertex_energy_kernel<<<dim3((unsigned)gx, (unsigned)ns), 128, 0, S(stream)>>>(R, part, F, N, ld, fps);
    ASB_LAUNCH_CHECK();
    if (ns > 1) {
        sum_slabs_kernel<<<(unsigned)cdiv(ld, 256), 256, 0, S(stream)>>>(part, e, ld, ns);
        ASB_LAUNCH_CHECK();
    }
    return ASB_OK;
}

// ---- allocation-free form of asb_vertex_energy (no counterpart in the reference; see include/asb.h) ----------------------
// The slab partials live in a file-scope device array instead of a caller workspace.  In host code `g_static_scratch` is only
// a stand-in for the symbol: its device address is looked up once per device with cudaGetSymbolAddress (that device current)
// and that pointer is what the kernels receive.
#define ASB_STATIC_SCRATCH_DOUBLES 131072   // 1 MiB per GPU that loads the library, held until exit
static __device__ __align__(256) double g_static_scratch[ASB_STATIC_SCRATCH_DOUBLES];

static int static_scratch(double** out) {
    constexpr int MAXDEV = 64;
    static double* ptrs[MAXDEV] = {};
    static std::mutex mu;
    int dev = -1;
    ASB_CUDA(cudaGetDevice(&dev));
    ASB_REQUIRE(dev >= 0 && dev < MAXDEV, "static scratch: device ordinal out of range");
    std::lock_guard<std::mutex> guard(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        ASB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        ASB_REQUIRE(p != nullptr, "static scratch: no device address for the scratch symbol");
        ptrs[dev] = reinterpret_cast<double*>(p);
    }
    *out = ptrs[dev];
    return ASB_OK;
}

int64_t asb_static_scratch_doubles(void) { return ASB_STATIC_SCRATCH_DOUBLES; }

int asb_vertex_energy_static(const double* R, double* e, int64_t F, int64_t N, int64_t ld, asb_stream_t stream) {
    ASB_REQUIRE(R && e && F > 0 && N > 0 && ld >= N, "vertex_energy_static: bad arguments");
    // the same split into frame slabs as asb_vertex_energy, so the two forms add in the same order (bit-equal results)
    const int64_t gx = cdiv(ld, 128);
    const int64_t slabs = pick_slabs(gx, F, 64);
    const int64_t fps = cdiv(F, slabs);
    const int64_t ns = cdiv(F, fps);
    ASB_REQUIRE(ns == 1 || ns * ld <= (int64_t)ASB_STATIC_SCRATCH_DOUBLES,
                "vertex_energy_static: frame-slab partials (slabs x ld doubles) exceed the 131072-double static scratch; "
                "use asb_vertex_energy with a caller workspace");
    double* part = e;
    if (ns > 1) {
        const int rc = static_scratch(&part);
        if (rc != ASB_OK) return rc;
    }
    // every partial in [0, ns * ld) is written by the first kernel before the second reads it: nothing to clear
    vertex_energy_kernel<<<dim3((unsigned)gx, (unsigned)ns), 128, 0, S(stream)>>>(R, part, F, N, ld, fps);
    ASB_LAUNCH_CHECK();
    if (ns > 1) {
        sum_slabs_kernel<<<(unsigned)cdiv(ld, 256), 256, 0, S(stream)>>>(part, e, ld, ns);
        ASB_LAUNCH_CHECK();
    }
    return ASB_OK;
}

int asb_group_sum(const double* v, int64_t n, int64_t group, double* out, asb_stream_t stream) {
    ASB_REQUIRE(v && out && n > 0 && group >= 1 && n % group == 0, "group_sum: n must be a multiple of the group size");
    group_sum_kernel<<<(unsigned)cdiv(n / group, 256), 256, 0, S(stream)>>>(v, n / group, group, out);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_argmax_first(const double* v, int64_t n, int64_t index_offset, double* out_val, int64_t* out_idx, void* ws,
                     asb_stream_t stream) {
    ASB_REQUIRE(v && out_val && out_idx && ws && n > 0, "argmax: bad arguments");
    int64_t nb = std::min<int64_t>(cdiv(n, 1024), 2048);
    const int64_t chunk = cdiv(n, nb);
    nb = cdiv(n, chunk);
    double* pv = reinterpret_cast<double*>(ws);
    long long* pi = reinterpret_cast<long long*>(pv + 2048);
    argmax_stage1_kernel<<<(unsigned)nb, 256, 0, S(stream)>>>(v, n, chunk, pv, pi);
    ASB_LAUNCH_CHECK();
    argmax_stage2_kernel<<<1, 256, 0, S(stream)>>>(pv, pi, nb, index_offset, out_val, out_idx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_argmax_outside_ball(const double* v, int64_t n, int64_t index_offset, const double* pos, const int64_t* centre,
                            double radius, double* out_pair, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(v && pos && centre && out_pair && ws && n > 0 && radius >= 0.0, "argmax_outside_ball: bad arguments");
    int64_t nb = std::min<int64_t>(cdiv(n, 1024), 2048);
    const int64_t chunk = cdiv(n, nb);
    nb = cdiv(n, chunk);
    double* pv = reinterpret_cast<double*>(ws);
    long long* pi = reinterpret_cast<long long*>(pv + 2048);
    argmax_outside_stage1_kernel<<<(unsigned)nb, 256, 0, S(stream)>>>(v, n, chunk, index_offset, pos, centre, radius * radius, pv, pi);
    ASB_LAUNCH_CHECK();
    argmax_pair_stage2_kernel<<<1, 256, 0, S(stream)>>>(pv, pi, nb, index_offset, out_pair);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_select_pair(const double* all, int64_t world, const int64_t* fallback, int64_t* out_idx, asb_stream_t stream) {
    ASB_REQUIRE(all && fallback && out_idx && world > 0 && world < (1 << 20), "select_pair: bad arguments");
    select_pair_kernel<<<1, 32, 0, S(stream)>>>(all, (int)world, fallback, out_idx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_deflate_gemv(const double* R, int64_t F, int64_t N, int64_t ld, const double* w, int64_t ldw, double* g,
                     void* ws, asb_stream_t stream) {
    ASB_REQUIRE(R && w && g && ws && F > 0 && N > 0 && ld % 32 == 0, "deflate_gemv: bad arguments");
    const int64_t C = 3 * ld;
    const int64_t gx = cdiv(C / 2, 256);
    const int64_t slabs = pick_slabs(gx, F, 64);
    const int64_t fps = cdiv(F, slabs);
    const int64_t ns = cdiv(F, fps);
    double* part = ns == 1 ? g : reinterpret_cast<double*>(ws);
    deflate_gemv_kernel<<<dim3((unsigned)gx, (unsigned)ns), 256, 0, S(stream)>>>(R, F, C, w, ldw, fps, part);
    ASB_LAUNCH_CHECK();
    if (ns > 1) {
        sum_slabs_kernel<<<(unsigned)cdiv(C, 256), 256, 0, S(stream)>>>(part, g, C, ns);
        ASB_LAUNCH_CHECK();
    }
    return ASB_OK;
}

int asb_deflate_finish(const double* g, const double* Cp, int64_t npend, const double* coef, const double* s,
                       const double* scal, double* cout, double* e, int64_t N, int64_t ld, double* meas, void* ws,
                       asb_stream_t stream) {
    ASB_REQUIRE(g && scal && cout && e && meas && ws && N > 0 && ld >= N && npend >= 0, "deflate_finish: bad arguments");
    ASB_REQUIRE(npend == 0 || (Cp && coef), "deflate_finish: pending block pointers");
    const int64_t nb = cdiv(ld, 256);
    double* partial = reinterpret_cast<double*>(ws);
    deflate_finish_kernel<<<(unsigned)nb, 256, 0, S(stream)>>>(g, 1, Cp, npend, coef, s, scal, cout, e, N, ld, partial);
    ASB_LAUNCH_CHECK();
    deflate_meas_kernel<<<1, 256, 0, S(stream)>>>(partial, nb, scal, meas);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_rank_update(double* R, int64_t F, int64_t N, int64_t ld, const double* Wb, int64_t ldw, const double* Cb,
                    int64_t nb, double* e, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(R && Wb && Cb && e && ws && F > 0 && N > 0 && ld >= N && nb >= 1 && nb <= 16, "rank_update: bad arguments");
    const int64_t gx = cdiv(ld, 128);
    const int64_t slabs = pick_slabs(gx, F, 64);
    const int64_t fps = cdiv(F, slabs);
    const int64_t ns = cdiv(F, fps);
    double* part = ns == 1 ? e : reinterpret_cast<double*>(ws);
    dim3 grid((unsigned)gx, (unsigned)ns);
    if (nb <= 4)
        rank_update_kernel<4><<<grid, 128, 0, S(stream)>>>(R, F, N, ld, Wb, ldw, Cb, (int)nb, fps, part);
    else if (nb <= 8) {
        static const int fu = [] {
            const char* e = std::getenv("ASB_RANK_FU");
            return e ? std::atoi(e) : 8;   // frames in flight per thread: 8 -> 17.97 ms, 4 -> 19.13 ms per 2 S pass (B200, 1M x 2000)
        }();
        if (fu == 2)
            rank_update_kernel<8, 2><<<grid, 128, 0, S(stream)>>>(R, F, N, ld, Wb, ldw, Cb, (int)nb, fps, part);
        else if (fu == 4)
            rank_update_kernel<8, 4><<<grid, 128, 0, S(stream)>>>(R, F, N, ld, Wb, ldw, Cb, (int)nb, fps, part);
        else
            rank_update_kernel<8, 8><<<grid, 128, 0, S(stream)>>>(R, F, N, ld, Wb, ldw, Cb, (int)nb, fps, part);
    }
    else
        rank_update_kernel<16><<<grid, 128, 0, S(stream)>>>(R, F, N, ld, Wb, ldw, Cb, (int)nb, fps, part);
    ASB_LAUNCH_CHECK();
    if (ns > 1) {
        sum_slabs_kernel<<<(unsigned)cdiv(ld, 256), 256, 0, S(stream)>>>(part, e, ld, ns);
        ASB_LAUNCH_CHECK();
    }
    return ASB_OK;
}

}  // extern "C"
