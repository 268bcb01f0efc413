// Small replicated dense work of the basis build: rank-1 SVD of a 3 x F trajectory (k4/k5), Gram-form
// Gauss-Seidel weight sweep (k10), K x K SPD inverse (k12), component-centre argmax (k11), DEIM
// gather / solve / residual (k19), CGS2 helpers (k18), the F x F symmetric eigensolve (k15, cuSOLVER).
#include <cooperative_groups.h>
#include <cusolverDn.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace asb {

// ------------------------------------------------------------------------------------------ k4/k5
__global__ void gather_traj_kernel(const double* __restrict__ R, int64_t F, int64_t N, int64_t ld,
                                   const int64_t* __restrict__ idx_dev, int64_t row0, const double* __restrict__ Wp,
                                   int64_t ldw, const double* __restrict__ Cp, int64_t npend, double* __restrict__ traj) {
    __shared__ double cp[16 * 3];
    const int64_t idx = idx_dev[0] - row0;
    if (idx < 0 || idx >= N) return;
    for (int t = threadIdx.x; t < npend * 3; t += blockDim.x) cp[t] = Cp[(int64_t)t * ld + idx];  // t = j*3+d
    __syncthreads();
    for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            double v = R[(f * 3 + d) * ld + idx];
            for (int j = 0; j < npend; ++j) v = __dsub_rn(v, __dmul_rn(Wp[f * ldw + j], cp[j * 3 + d]));
            traj[d * F + f] = v;
        }
    }
}

// Multi-GPU pick, step 1: this rank's candidate packed for ONE all-gather: cand[0] = local maximum, cand[1] = its global
// index (int64 bit pattern), cand[2 + d F + f] = the candidate vertex's trajectory (corrected for the pending block).
__global__ void gather_candidate_kernel(const double* __restrict__ R, int64_t F, int64_t N, int64_t ld,
                                        const double* __restrict__ val_dev, const int64_t* __restrict__ idx_dev, int64_t row0,
                                        const double* __restrict__ Wp, int64_t ldw, const double* __restrict__ Cp,
                                        int64_t npend, double* __restrict__ cand) {
    __shared__ double cp[16 * 3];
    const int64_t idx = idx_dev[0] - row0;
    const bool live = idx >= 0 && idx < N;
    if (threadIdx.x == 0) {
        cand[0] = live ? val_dev[0] : -INFINITY;
        cand[1] = __longlong_as_double(live ? (long long)idx_dev[0] : 0x7fffffffffffffffLL);
    }
    if (!live) {
        for (int64_t t = threadIdx.x; t < 3 * F; t += blockDim.x) cand[2 + t] = 0.0;
        return;
    }
    for (int t = threadIdx.x; t < npend * 3; t += blockDim.x) cp[t] = Cp[(int64_t)t * ld + idx];
    __syncthreads();
    for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            double v = R[(f * 3 + d) * ld + idx];
            for (int j = 0; j < npend; ++j) v = __dsub_rn(v, __dmul_rn(Wp[f * ldw + j], cp[j * 3 + d]));
            cand[2 + d * F + f] = v;
        }
    }
}

// step 2 (after the all-gather): the winner over the ranks -- largest value, ties to the lowest global index (np.argmax) --
// and its trajectory.  One block.
__global__ void select_candidate_kernel(const double* __restrict__ all, int64_t world, int64_t stride, int64_t F,
                                        double* __restrict__ out_val, int64_t* __restrict__ out_idx, double* __restrict__ traj) {
    __shared__ int win;
    if (threadIdx.x == 0) {
        ArgMax a;
        a.v = -INFINITY;
        a.i = 0x7fffffffffffffffLL;
        int w = 0;
        for (int64_t r = 0; r < world; ++r) {
            ArgMax t;
            t.v = all[r * stride];
            t.i = __double_as_longlong(all[r * stride + 1]);
            if (t.v > a.v || (t.v == a.v && t.i < a.i)) { a = t; w = (int)r; }
        }
        win = w;
        out_val[0] = a.v;
        out_idx[0] = (int64_t)a.i;
    }
    __syncthreads();
    const double* src = all + (int64_t)win * stride + 2;
    for (int64_t t = threadIdx.x; t < 3 * F; t += blockDim.x) traj[t] = src[t];
}

// one block; traj: 3 x F
__global__ void traj_rank1_kernel(const double* __restrict__ traj, int64_t F, int local_support,
                                  double* __restrict__ wout, int64_t ldw, const double* __restrict__ Wp,
                                  int64_t npend, double* __restrict__ coef, double* __restrict__ scal) {
    __shared__ double sh[33];
    __shared__ double su[4];
    double a[6] = {0, 0, 0, 0, 0, 0};  // xx xy xz yy yz zz
    for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
        const double x = traj[f], y = traj[F + f], z = traj[2 * F + f];
        a[0] = fma(x, x, a[0]);
        a[1] = fma(x, y, a[1]);
        a[2] = fma(x, z, a[2]);
        a[3] = fma(y, y, a[3]);
        a[4] = fma(y, z, a[4]);
        a[5] = fma(z, z, a[5]);
    }
    for (int i = 0; i < 6; ++i) a[i] = block_sum(a[i], sh);
    if (threadIdx.x == 0) {
        double A[9] = {a[0], a[1], a[2], a[1], a[3], a[4], a[2], a[4], a[5]}, ev[3], V[9];
        jacobi_eig3(A, ev, V);
        int m = 0;
        if (ev[1] > ev[m]) m = 1;
        if (ev[2] > ev[m]) m = 2;
        double u0 = V[0 * 3 + m], u1 = V[1 * 3 + m], u2 = V[2 * 3 + m];
        // deterministic sign: largest-magnitude entry of the left singular vector is positive
        double big = u0;
        if (fabs(u1) > fabs(big)) big = u1;
        if (fabs(u2) > fabs(big)) big = u2;
        if (big < 0) { u0 = -u0; u1 = -u1; u2 = -u2; }
        su[0] = u0; su[1] = u1; su[2] = u2;
        su[3] = sqrt(fmax(ev[m], 0.0));
    }
    __syncthreads();
    const double u0 = su[0], u1 = su[1], u2 = su[2];
    // w = u^T T  (= sigma0 * Vt[0])
    double mx = 0.0, mn = 0.0;  // max(w), max(-w)
    for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
        const double w = u0 * traj[f] + u1 * traj[F + f] + u2 * traj[2 * F + f];
        wout[f * ldw] = w;
        mx = fmax(mx, w);
        mn = fmax(mn, -w);
    }
    if (local_support) {
        mx = block_max(mx, sh);
        mn = block_max(mn, sh);
        double np = 0.0, nn = 0.0;
        for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
            const double w = wout[f * ldw];
            double p = fmax(0.0, w), q = fmax(0.0, -w);
            if (mx != 0.0) p = p / mx;
            if (mn != 0.0) q = q / mn;
            np = fma(p, p, np);
            nn = fma(q, q, nn);
        }
        np = block_sum(np, sh);
        nn = block_sum(nn, sh);
        const bool pos = sqrt(np) > sqrt(nn);
        for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
            const double w = wout[f * ldw];
            double p = pos ? fmax(0.0, w) : fmax(0.0, -w);
            const double m = pos ? mx : mn;
            if (m != 0.0) p = p / m;
            wout[f * ldw] = p;
        }
    }
    __syncthreads();
    double ww = 0.0;
    for (int64_t f = threadIdx.x; f < F; f += blockDim.x) {
        const double w = wout[f * ldw];
        ww = fma(w, w, ww);
    }
    ww = block_sum(ww, sh);
    if (threadIdx.x == 0) { scal[0] = su[3]; scal[1] = ww; }
    for (int j = 0; j < npend; ++j) {
        double c = 0.0;
        for (int64_t f = threadIdx.x; f < F; f += blockDim.x) c = fma(wout[f * ldw], Wp[f * ldw + j], c);
        c = block_sum(c, sh);
        if (threadIdx.x == 0) coef[j] = c;
    }
}

// ------------------------------------------------------------------------------------------ small GEMM
// D[i][j] = alpha * sum_k A(i,k) B(k,j) (+ diag_add on the diagonal); generic strides, 16x16 tiles.
__global__ void small_gemm_kernel(int64_t m, int64_t n, int64_t k, double alpha, const double* __restrict__ A,
                                  int64_t sai, int64_t sak, const double* __restrict__ B, int64_t sbk, int64_t sbj,
                                  double diag_add, double* __restrict__ D, int64_t ldd) {
    __shared__ double As[16][17], Bs[16][17];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int64_t i = (int64_t)blockIdx.y * 16 + ty, j = (int64_t)blockIdx.x * 16 + tx;
    double acc = 0.0;
    for (int64_t k0 = 0; k0 < k; k0 += 16) {
        const int64_t ka = k0 + tx, kb = k0 + ty;
        As[ty][tx] = (i < m && ka < k) ? A[i * sai + ka * sak] : 0.0;
        Bs[ty][tx] = (kb < k && j < n) ? B[kb * sbk + j * sbj] : 0.0;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) acc = fma(As[ty][kk], Bs[kk][tx], acc);
        __syncthreads();
    }
    if (i < m && j < n) D[i * ldd + j] = alpha * acc + (i == j ? diag_add : 0.0);
}

// ------------------------------------------------------------------------------------------ SPD inverse
// single block; H (K x K row-major) -> lower Cholesky factor in place; Hinv = H^-1. T = scratch K x K (ldi) = Linv.
__global__ void spd_inverse_kernel(double* __restrict__ H, int64_t ldh, int64_t K, double* __restrict__ Hinv,
                                   int64_t ldi, double* __restrict__ T, int* __restrict__ info) {
    __shared__ double sh[33];
    __shared__ int bad;
    if (threadIdx.x == 0) bad = 0;
    __syncthreads();
    for (int64_t j = 0; j < K; ++j) {
        // diagonal
        double s = 0.0;
        for (int64_t k = threadIdx.x; k < j; k += blockDim.x) s = fma(H[j * ldh + k], H[j * ldh + k], s);
        s = block_sum(s, sh);
        const double d = H[j * ldh + j] - s;
        if (!(d > 0.0)) {
            if (threadIdx.x == 0 && bad == 0) bad = (int)j + 1;
        }
        const double ljj = sqrt(d > 0.0 ? d : 1.0);
        __syncthreads();
        if (threadIdx.x == 0) H[j * ldh + j] = ljj;
        // column j below the diagonal (left-looking)
        for (int64_t i = j + 1 + threadIdx.x; i < K; i += blockDim.x) {
            double t0 = 0.0, t1 = 0.0;
            int64_t k = 0;
            for (; k + 2 <= j; k += 2) {
                t0 = fma(H[i * ldh + k], H[j * ldh + k], t0);
                t1 = fma(H[i * ldh + k + 1], H[j * ldh + k + 1], t1);
            }
            if (k < j) t0 = fma(H[i * ldh + k], H[j * ldh + k], t0);
            H[i * ldh + j] = (H[i * ldh + j] - (t0 + t1)) / ljj;
        }
        __syncthreads();
    }
    // T = L^-1 (lower): thread per column c, forward substitution down the rows
    for (int64_t c = threadIdx.x; c < K; c += blockDim.x) {
        for (int64_t i = 0; i < K; ++i) {
            double v;
            if (i < c) {
                v = 0.0;
            } else {
                double t0 = (i == c) ? 1.0 : 0.0;
                for (int64_t k = c; k < i; ++k) t0 = fma(-H[i * ldh + k], T[k * ldi + c], t0);
                v = t0 / H[i * ldh + i];
            }
            T[i * ldi + c] = v;
        }
    }
    __syncthreads();
    // Hinv = T^T T
    for (int64_t t = threadIdx.x; t < K * K; t += blockDim.x) {
        const int64_t a = t / K, b = t % K;
        const int64_t i0 = a > b ? a : b;
        double s = 0.0;
        for (int64_t i = i0; i < K; ++i) s = fma(T[i * ldi + a], T[i * ldi + b], s);
        Hinv[a * ldi + b] = s;
    }
    if (threadIdx.x == 0) info[0] = bad;
}

// In-place inversion of an SPD matrix by K symmetric sweeps on the packed lower triangle held in shared memory
// (K (K + 1) / 2 doubles: 160.8 KB at K = 200).  Sweep k:  a_ij -= a_ik a_jk / a_kk (i, j != k),  a_ik /= a_kk,  a_kk = -1 / a_kk;
// after all K sweeps the triangle holds -A^-1.  For an SPD matrix every pivot is a Schur-complement diagonal (> 0), so no
// pivoting is needed.  One CTA of 1024 threads: a warp owns rows i = warp, warp + 32, ...; two barriers per sweep.
// The left-looking global-memory Cholesky above takes 3.3 ms at K = 200 (one replicated launch per SPLOCS iteration: 8 % of
// the SPLOCS phase on 8 GPUs); this one ~0.06 ms.
__global__ void __launch_bounds__(1024) spd_inverse_sweep_kernel(const double* __restrict__ H, int64_t ldh, int K,
                                                                 double* __restrict__ Hinv, int64_t ldi, int* __restrict__ info) {
    extern __shared__ double tri[];                 // packed lower triangle: (i, j), j <= i, at i (i + 1) / 2 + j
    double* colk = tri + (size_t)K * (K + 1) / 2;   // column k of the current sweep
    __shared__ int bad;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) bad = 0;
    for (int i = warp; i < K; i += 32)
        for (int j = lane; j <= i; j += 32) tri[i * (i + 1) / 2 + j] = H[(int64_t)i * ldh + j];
    __syncthreads();
    for (int k = 0; k < K; ++k) {
        for (int i = threadIdx.x; i < K; i += blockDim.x) colk[i] = i >= k ? tri[i * (i + 1) / 2 + k] : tri[k * (k + 1) / 2 + i];
        __syncthreads();
        const double d = colk[k];
        if (!(d > 0.0) && threadIdx.x == 0 && bad == 0) bad = k + 1;
        const double inv = 1.0 / d;
        for (int i = warp; i < K; i += 32) {
            double* row = tri + i * (i + 1) / 2;
            if (i == k) {
                for (int j = lane; j < k; j += 32) row[j] = colk[j] * inv;
                if (lane == 0) row[k] = -inv;
            } else {
                const double aik = colk[i] * inv;
                for (int j = lane; j <= i; j += 32) row[j] = j == k ? aik : fma(-aik, colk[j], row[j]);
            }
        }
        __syncthreads();
    }
    for (int i = warp; i < K; i += 32)
        for (int j = lane; j <= i; j += 32) {
            const double v = -tri[i * (i + 1) / 2 + j];
            Hinv[(int64_t)i * ldi + j] = v;
            Hinv[(int64_t)j * ldi + i] = v;
        }
    if (threadIdx.x == 0) info[0] = bad;
}

// ------------------------------------------------------------------------------------------ k10 weights sweep
// Cooperative kernel, one grid-wide max per component.  See asb_weights_sweep.
__global__ void __launch_bounds__(256) weights_sweep_kernel(const double* __restrict__ P, int64_t ldp,
                                                           const double* __restrict__ Q, int64_t ldq,
                                                           double* __restrict__ W, int64_t ldw, int64_t F, int64_t K,
                                                           int64_t rpb, double* __restrict__ maxbuf) {
    extern __shared__ double smem[];
    __shared__ double sh[33];
    cg::grid_group grid = cg::this_grid();
    const int64_t sld = K + 1;
    double* sW = smem;                 // rpb x (K+1)
    double* sOpt = smem + rpb * sld;   // rpb
    const int64_t r0 = (int64_t)blockIdx.x * rpb;
    const int64_t nr = max((int64_t)0, min(rpb, F - r0));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int64_t t = threadIdx.x; t < nr * K; t += blockDim.x) {
        const int64_t r = t / K, j = t % K;
        sW[r * sld + j] = W[(r0 + r) * ldw + j];
    }
    __syncthreads();
    int it = 0;
    for (int64_t k = 0; k < K; ++k) {
        const double qkk = Q[k * ldq + k];
        if (qkk <= 1.e-8) {  // posComponents.py:147-150: activation zeroed, residual left as is
            for (int64_t r = threadIdx.x; r < nr; r += blockDim.x) W[(r0 + r) * ldw + k] = 0.0;
            continue;
        }
        double lmax = 0.0;
        for (int64_t r = warp; r < nr; r += nw) {
            double part = 0.0;
            for (int64_t j = lane; j < K; j += 32) part = fma(sW[r * sld + j], Q[k * ldq + j], part);
            part = warp_sum(part);
            const double opt = (P[(r0 + r) * ldp + k] - part + sW[r * sld + k] * qkk) / qkk;
            if (lane == 0) sOpt[r] = opt;
            lmax = fmax(lmax, fmax(0.0, opt));
        }
        lmax = block_max(lmax, sh);
        double* mb = maxbuf + (int64_t)(it & 1) * gridDim.x;
        if (threadIdx.x == 0) mb[blockIdx.x] = lmax;
        grid.sync();
        double gm = 0.0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) gm = fmax(gm, mb[b]);
        gm = block_max(gm, sh);
        for (int64_t r = threadIdx.x; r < nr; r += blockDim.x) {
            double x = fmax(0.0, sOpt[r]);
            if (gm != 0.0) x = x / gm;
            sW[r * sld + k] = x;
            W[(r0 + r) * ldw + k] = x;
        }
        __syncthreads();
        ++it;
    }
}

// ------------------------------------------------------------------------------------------ k11 centres
// grid (chunks, K): partial (max |c_n|^2, first idx, sum Lambda |c_n|)
__global__ void centers_stage1_kernel(const double* __restrict__ C, const double* __restrict__ Lambda, int64_t N,
                                      int64_t ld, int64_t chunk, double* __restrict__ pv, long long* __restrict__ pi,
                                      double* __restrict__ ps) {
    __shared__ double shv[33];
    __shared__ long long shi[33];
    const int64_t k = blockIdx.y;
    const int64_t n0 = (int64_t)blockIdx.x * chunk, n1 = min(N, n0 + chunk);
    const double* c = C + k * 3 * ld;
    ArgMax a;
    a.v = -INFINITY;
    a.i = 0x7fffffffffffffffLL;
    double sp = 0.0;
    for (int64_t n = n0 + threadIdx.x; n < n1; n += blockDim.x) {
        const double x = c[n], y = c[ld + n], z = c[2 * ld + n];
        const double v = __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
        if (v > a.v) { a.v = v; a.i = n; }
        if (Lambda) sp = fma(Lambda[k * ld + n], sqrt(v), sp);
    }
    a = block_argmax(a, shv, shi);
    sp = block_sum(sp, shv);
    if (threadIdx.x == 0) {
        const int64_t o = k * gridDim.x + blockIdx.x;
        pv[o] = a.v;
        pi[o] = a.i;
        ps[o] = sp;
    }
}
__global__ void centers_stage2_kernel(const double* __restrict__ pv, const long long* __restrict__ pi,
                                      const double* __restrict__ ps, int64_t chunks, int64_t index_offset,
                                      double* __restrict__ val, int64_t* __restrict__ idx, double* __restrict__ sp) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= gridDim.x * (int64_t)blockDim.x) return;
    ArgMax a;
    a.v = -INFINITY;
    a.i = 0x7fffffffffffffffLL;
    double s = 0.0;
    for (int64_t c = 0; c < chunks; ++c) {
        ArgMax t;
        t.v = pv[k * chunks + c];
        t.i = pi[k * chunks + c];
        a = argmax_combine(a, t);
        s += ps[k * chunks + c];
    }
    val[k] = a.v;
    idx[k] = (int64_t)a.i + index_offset;
    if (sp) sp[k] = s;
}

// ------------------------------------------------------------------------------------------ k19 DEIM
__global__ void deim_gather_kernel(const double* __restrict__ C, int64_t ld, int64_t N, int64_t row0,
                                   const int64_t* __restrict__ Pt, int64_t k, int64_t rhs, double* __restrict__ Vp,
                                   int64_t ldv) {
    // Vp[d][j][l] = C[l < k ? l : rhs][d][Pt[j]-row0], j < k, l <= k
    const int64_t total = 3 * k * (k + 1);
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t l = t % (k + 1), j = (t / (k + 1)) % k, d = t / ((k + 1) * k);
        const int64_t row = Pt[j] - row0;
        if (row >= 0 && row < N) Vp[(d * ldv + j) * ldv + l] = C[((l < k ? l : rhs) * 3 + d) * ld + row];
    }
}

// one block per dimension: Gaussian elimination with partial pivoting on the k x (k+1) augmented system
__global__ void __launch_bounds__(256) deim_solve_kernel(const double* __restrict__ Vp, int64_t ldv, int64_t k,
                                                        double* __restrict__ a, int64_t lda_a,
                                                        double* __restrict__ work, int* __restrict__ info) {
    __shared__ double shv[33];
    __shared__ long long shi[33];
    const int d = blockIdx.x;
    const int64_t w = k + 1;
    double* A = work + (int64_t)d * k * w;
    const double* src = Vp + (int64_t)d * ldv * ldv;
    for (int64_t t = threadIdx.x; t < k * w; t += blockDim.x) A[t] = src[(t / w) * ldv + (t % w)];
    __syncthreads();
    int bad = 0;
    for (int64_t c = 0; c < k; ++c) {
        ArgMax p;
        p.v = -1.0;
        p.i = 0x7fffffffffffffffLL;
        for (int64_t i = c + threadIdx.x; i < k; i += blockDim.x) {
            const double v = fabs(A[i * w + c]);
            if (v > p.v) { p.v = v; p.i = i; }
        }
        p = block_argmax(p, shv, shi);
        if (!(p.v > 0.0)) { bad = (int)c + 1; break; }
        const int64_t pr = p.i;
        if (pr != c) {
            for (int64_t j = threadIdx.x; j < w; j += blockDim.x) {
                const double t = A[c * w + j];
                A[c * w + j] = A[pr * w + j];
                A[pr * w + j] = t;
            }
        }
        __syncthreads();
        const double piv = A[c * w + c];
        const int64_t rows = k - c - 1, cols = w - c - 1;
        // multipliers first (column c), then the trailing update
        for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) A[(c + 1 + i) * w + c] /= piv;
        __syncthreads();
        for (int64_t t = threadIdx.x; t < rows * cols; t += blockDim.x) {
            const int64_t i = c + 1 + t / cols, j = c + 1 + t % cols;
            A[i * w + j] = fma(-A[i * w + c], A[c * w + j], A[i * w + j]);
        }
        __syncthreads();
    }
    if (bad == 0) {
        // back substitution on U x = y (y = last column); thread 0 walks, others help with dot products
        for (int64_t i = k - 1; i >= 0; --i) {
            double s = 0.0;
            for (int64_t j = i + 1 + threadIdx.x; j < k; j += blockDim.x) s = fma(A[i * w + j], a[d * lda_a + j], s);
            s = block_sum(s, shv);
            if (threadIdx.x == 0) a[d * lda_a + i] = (A[i * w + k] - s) / A[i * w + i];
            __syncthreads();
        }
    }
    if (threadIdx.x == 0 && bad) atomicMax(info, bad);
}

// e[n] = |sum_l a[d][l] C[l][d][n] - C[k][d][n]|^2 summed over d; block partial max|r|
__global__ void __launch_bounds__(256) deim_residual_kernel(const double* __restrict__ C, int64_t ld, int64_t N,
                                                           int64_t k, int64_t target, int accumulate,
                                                           const double* __restrict__ a, int64_t lda_a,
                                                           double* __restrict__ e, double* __restrict__ pmax) {
    extern __shared__ double sa[];  // 3 x k
    __shared__ double sh[33];
    for (int64_t t = threadIdx.x; t < 3 * k; t += blockDim.x) sa[t] = a[(t / k) * lda_a + (t % k)];
    __syncthreads();
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double m = 0.0;
    if (n < N) {
        double r[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            double s0 = 0.0, s1 = 0.0;
            int64_t l = 0;
            for (; l + 2 <= k; l += 2) {
                s0 = fma(sa[d * k + l], C[(l * 3 + d) * ld + n], s0);
                s1 = fma(sa[d * k + l + 1], C[((l + 1) * 3 + d) * ld + n], s1);
            }
            if (l < k) s0 = fma(sa[d * k + l], C[(l * 3 + d) * ld + n], s0);
            const double v = C[(target * 3 + d) * ld + n];
            r[d] = k == 0 ? v : (s0 + s1) - v;
            m = fmax(m, fabs(r[d]));
        }
        const double en = __dadd_rn(__dadd_rn(__dmul_rn(r[0], r[0]), __dmul_rn(r[1], r[1])), __dmul_rn(r[2], r[2]));
        e[n] = accumulate ? __dadd_rn(e[n], en) : en;
    }
    m = block_max(m, sh);
    if (threadIdx.x == 0) pmax[blockIdx.x] = accumulate ? fmax(pmax[blockIdx.x], m) : m;
}
__global__ void max_partials_kernel(const double* __restrict__ p, int64_t nb, double* __restrict__ out) {
    __shared__ double sh[33];
    double m = 0.0;
    for (int64_t b = threadIdx.x; b < nb; b += blockDim.x) m = fmax(m, p[b]);
    m = block_max(m, sh);
    if (threadIdx.x == 0) out[0] = m;
}

// ------------------------------------------------------------------------------------------ k18 CGS helpers
// h_part[block][j] = sum_{n in block range} T[j][d][n] * T[k][d][n], j < k
__global__ void __launch_bounds__(256) cgs_dots_kernel(const double* __restrict__ T, int64_t k, int64_t d, int64_t N,
                                                      int64_t ld, int64_t chunk, double* __restrict__ hpart) {
    __shared__ double sh[33];
    const int64_t n0 = (int64_t)blockIdx.x * chunk, n1 = min(N, n0 + chunk);
    const double* tk = T + (k * 3 + d) * ld;
    for (int64_t j = 0; j < k; ++j) {
        const double* tj = T + (j * 3 + d) * ld;
        double s = 0.0;
        for (int64_t n = n0 + threadIdx.x; n < n1; n += blockDim.x) s = fma(tj[n], tk[n], s);
        s = block_sum(s, sh);
        if (threadIdx.x == 0) hpart[(int64_t)blockIdx.x * k + j] = s;
    }
}
// t_k -= sum_j h[j] t_j ; nrm partial of the new t_k
__global__ void __launch_bounds__(256) cgs_project_kernel(double* __restrict__ T, int64_t k, int64_t d, int64_t N,
                                                         int64_t ld, const double* __restrict__ h,
                                                         double* __restrict__ npart) {
    __shared__ double sh[33];
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double v2 = 0.0;
    if (n < N) {
        double v = T[(k * 3 + d) * ld + n];
        for (int64_t j = 0; j < k; ++j) v = fma(-h[j], T[(j * 3 + d) * ld + n], v);
        T[(k * 3 + d) * ld + n] = v;
        v2 = v * v;
    }
    v2 = block_sum(v2, sh);
    if (threadIdx.x == 0) npart[blockIdx.x] = v2;
}
__global__ void row_scale_kernel(double* __restrict__ T, int64_t k, int64_t d, int64_t ld, const double* __restrict__ nrm2) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n < ld) T[(k * 3 + d) * ld + n] = T[(k * 3 + d) * ld + n] / sqrt(nrm2[0]);
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_gather_traj(const double* R, int64_t F, int64_t N, int64_t ld, const int64_t* idx_dev, int64_t row0,
                    const double* Wp, int64_t ldw, const double* Cp, int64_t npend, double* traj, asb_stream_t stream) {
    ASB_REQUIRE(R && idx_dev && traj && F > 0 && N > 0 && npend >= 0 && npend <= 16, "gather_traj: bad arguments");
    ASB_REQUIRE(npend == 0 || (Wp && Cp), "gather_traj: pending block pointers");
    gather_traj_kernel<<<1, 256, 0, S(stream)>>>(R, F, N, ld, idx_dev, row0, Wp, ldw, Cp, npend, traj);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_gather_candidate(const double* R, int64_t F, int64_t N, int64_t ld, const double* val_dev, const int64_t* idx_dev,
                         int64_t row0, const double* Wp, int64_t ldw, const double* Cp, int64_t npend, double* cand,
                         asb_stream_t stream) {
    ASB_REQUIRE(R && val_dev && idx_dev && cand && F > 0 && N >= 0 && npend >= 0 && npend <= 16, "gather_candidate: bad arguments");
    ASB_REQUIRE(npend == 0 || (Wp && Cp), "gather_candidate: pending block pointers");
    gather_candidate_kernel<<<1, 256, 0, S(stream)>>>(R, F, N, ld, val_dev, idx_dev, row0, Wp, ldw, Cp, npend, cand);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_select_candidate(const double* all, int64_t world, int64_t stride, int64_t F, double* out_val, int64_t* out_idx,
                         double* traj, asb_stream_t stream) {
    ASB_REQUIRE(all && out_val && out_idx && traj && world > 0 && F > 0 && stride >= 2 + 3 * F, "select_candidate: bad arguments");
    select_candidate_kernel<<<1, 256, 0, S(stream)>>>(all, world, stride, F, out_val, out_idx, traj);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_traj_rank1(const double* traj, int64_t F, int local_support, double* wout, int64_t ldw, const double* Wp,
                   int64_t npend, double* coef, double* scal, asb_stream_t stream) {
    ASB_REQUIRE(traj && wout && scal && F > 0 && npend >= 0 && npend <= 16, "traj_rank1: bad arguments");
    ASB_REQUIRE(npend == 0 || (Wp && coef), "traj_rank1: pending block pointers");
    traj_rank1_kernel<<<1, 256, 0, S(stream)>>>(traj, F, local_support, wout, ldw, Wp, npend, coef, scal);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_small_gemm(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha, const double* A, int64_t lda,
                   const double* B, int64_t ldb, double* D, int64_t ldd, asb_stream_t stream) {
    ASB_REQUIRE(A && B && D && m > 0 && n > 0 && k > 0, "small_gemm: bad arguments");
    dim3 grid((unsigned)cdiv(n, 16), (unsigned)cdiv(m, 16)), block(16, 16);
    small_gemm_kernel<<<grid, block, 0, S(stream)>>>(m, n, k, alpha, A, transA ? 1 : lda, transA ? lda : 1, B,
                                                     transB ? 1 : ldb, transB ? ldb : 1, 0.0, D, ldd);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_gram_small(const double* W, int64_t ldw, int64_t F, int64_t K, double diag_add, double* G, int64_t ldg,
                   asb_stream_t stream) {
    ASB_REQUIRE(W && G && F > 0 && K > 0, "gram_small: bad arguments");
    dim3 grid((unsigned)cdiv(K, 16), (unsigned)cdiv(K, 16)), block(16, 16);
    // G = W^T W : A(i,k) = W[k][i], B(k,j) = W[k][j]
    small_gemm_kernel<<<grid, block, 0, S(stream)>>>(K, K, F, 1.0, W, 1, ldw, W, ldw, 1, diag_add, G, ldg);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_spd_inverse(double* H, int64_t ldh, int64_t K, double* Hinv, int64_t ldi, double* scratch, int* info_dev,
                    asb_stream_t stream) {
    ASB_REQUIRE(H && Hinv && scratch && info_dev && K > 0 && K <= 2048, "spd_inverse: bad arguments");
    const size_t smem = ((size_t)K * (K + 1) / 2 + K) * sizeof(double);
    if (smem <= 226 * 1024) {                       // K <= 239: packed triangle in shared memory
        static bool attr_set = false;
        if (!attr_set) {
            ASB_CUDA(cudaFuncSetAttribute(spd_inverse_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
            attr_set = true;
        }
        spd_inverse_sweep_kernel<<<1, 1024, smem, S(stream)>>>(H, ldh, (int)K, Hinv, ldi, info_dev);
        ASB_LAUNCH_CHECK();
        return ASB_OK;
    }
    spd_inverse_kernel<<<1, 512, 0, S(stream)>>>(H, ldh, K, Hinv, ldi, scratch, info_dev);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_weights_sweep(const double* P, int64_t ldp, const double* Q, int64_t ldq, double* W, int64_t ldw, int64_t F,
                      int64_t K, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(P && Q && W && ws && F > 0 && K > 0, "weights_sweep: bad arguments");
    const int sms = sm_count();
    int64_t rpb = cdiv(F, sms);
    const int64_t smem = (rpb * (K + 1) + rpb) * (int64_t)sizeof(double);
    ASB_REQUIRE(smem <= 220 * 1024, "weights_sweep: F*K too large for the single-wave sweep");
    static bool attr_set = false;
    if (!attr_set) {
        ASB_CUDA(cudaFuncSetAttribute(weights_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        attr_set = true;
    }
    const int64_t grid = cdiv(F, rpb);
    double* maxbuf = reinterpret_cast<double*>(ws);
    void* args[] = {(void*)&P, (void*)&ldp, (void*)&Q, (void*)&ldq, (void*)&W, (void*)&ldw, (void*)&F, (void*)&K,
                    (void*)&rpb, (void*)&maxbuf};
    ASB_CUDA(cudaLaunchCooperativeKernel((void*)weights_sweep_kernel, dim3((unsigned)grid), dim3(256), args, (size_t)smem,
                                         S(stream)));
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_component_centers(const double* C, const double* Lambda, int64_t K, int64_t N, int64_t ld, int64_t index_offset,
                          double* val, int64_t* idx, double* sp, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(C && val && idx && ws && K > 0 && K <= 65535 && N > 0 && ld >= N, "component_centers: bad arguments");
    int64_t chunks = std::max<int64_t>(1, std::min<int64_t>(64, cdiv(4LL * sm_count(), K)));
    chunks = std::min<int64_t>(chunks, cdiv(N, 256));
    const int64_t chunk = cdiv(N, chunks);
    chunks = cdiv(N, chunk);
    double* pv = reinterpret_cast<double*>(ws);
    long long* pi = reinterpret_cast<long long*>(pv + K * chunks);
    double* ps = reinterpret_cast<double*>(pi + K * chunks);
    centers_stage1_kernel<<<dim3((unsigned)chunks, (unsigned)K), 256, 0, S(stream)>>>(C, Lambda, N, ld, chunk, pv, pi, ps);
    ASB_LAUNCH_CHECK();
    // one thread per component; grid*block == K exactly is not guaranteed -> use (K,1) blocks of 1 thread groups
    centers_stage2_kernel<<<(unsigned)K, 1, 0, S(stream)>>>(pv, pi, ps, chunks, index_offset, val, idx, Lambda ? sp : nullptr);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_deim_gather_rhs(const double* C, int64_t ld, int64_t N, int64_t row0, const int64_t* Pt, int64_t k, int64_t rhs,
                        double* Vp, int64_t ldv, asb_stream_t stream) {
    ASB_REQUIRE(C && Pt && Vp && k >= 1 && ldv >= k + 1 && rhs >= 0, "deim_gather: bad arguments");
    const int64_t total = 3 * k * (k + 1);
    deim_gather_kernel<<<(unsigned)std::min<int64_t>(cdiv(total, 256), 1024), 256, 0, S(stream)>>>(C, ld, N, row0, Pt, k, rhs,
                                                                                                Vp, ldv);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_deim_gather(const double* C, int64_t ld, int64_t N, int64_t row0, const int64_t* Pt, int64_t k, double* Vp,
                    int64_t ldv, asb_stream_t stream) {
    return asb_deim_gather_rhs(C, ld, N, row0, Pt, k, k, Vp, ldv, stream);
}

int asb_deim_solve(const double* Vp, int64_t ldv, int64_t k, double* a, int64_t lda_a, void* ws, int* info_dev,
                   asb_stream_t stream) {
    ASB_REQUIRE(Vp && a && ws && info_dev && k >= 1 && ldv >= k + 1 && lda_a >= k, "deim_solve: bad arguments");
    deim_solve_kernel<<<3, 256, 0, S(stream)>>>(Vp, ldv, k, a, lda_a, reinterpret_cast<double*>(ws), info_dev);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_deim_residual_block(const double* C, int64_t ld, int64_t N, int64_t k, int64_t target, int accumulate,
                            const double* a, int64_t lda_a, double* e, double* maxabs, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(C && e && maxabs && ws && N > 0 && k >= 0 && target >= 0 && (k == 0 || a), "deim_residual: bad arguments");
    ASB_REQUIRE(3 * k * 8 <= 48 * 1024, "deim_residual: k too large");
    const int64_t nb = cdiv(N, 256);
    double* pmax = reinterpret_cast<double*>(ws);
    deim_residual_kernel<<<(unsigned)nb, 256, (size_t)(3 * std::max<int64_t>(k, 1) * 8), S(stream)>>>(
        C, ld, N, k, target, accumulate, a, lda_a, e, pmax);
    ASB_LAUNCH_CHECK();
    max_partials_kernel<<<1, 256, 0, S(stream)>>>(pmax, nb, maxabs);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_deim_residual(const double* C, int64_t ld, int64_t N, int64_t k, const double* a, int64_t lda_a, double* e,
                      double* maxabs, void* ws, asb_stream_t stream) {
    return asb_deim_residual_block(C, ld, N, k, k, 0, a, lda_a, e, maxabs, ws, stream);
}

int asb_cgs_dots(const double* T, int64_t k, int64_t d, int64_t N, int64_t ld, double* h, void* ws, asb_stream_t stream) {
    ASB_REQUIRE(T && h && ws && k >= 1 && d >= 0 && d < 3 && N > 0, "cgs_dots: bad arguments");
    int64_t nb = std::min<int64_t>(cdiv(N, 1024), 2LL * sm_count());
    const int64_t chunk = cdiv(N, nb);
    nb = cdiv(N, chunk);
    double* hpart = reinterpret_cast<double*>(ws);
    cgs_dots_kernel<<<(unsigned)nb, 256, 0, S(stream)>>>(T, k, d, N, ld, chunk, hpart);
    ASB_LAUNCH_CHECK();
    return sum_partials(hpart, nb, (int)k, h, S(stream));
}

int asb_cgs_project(double* T, int64_t k, int64_t d, int64_t N, int64_t ld, const double* h, double* nrm2, void* ws,
                    asb_stream_t stream) {
    ASB_REQUIRE(T && nrm2 && ws && k >= 0 && d >= 0 && d < 3 && N > 0 && (k == 0 || h), "cgs_project: bad arguments");
    const int64_t nb = cdiv(N, 256);
    double* npart = reinterpret_cast<double*>(ws);
    cgs_project_kernel<<<(unsigned)nb, 256, 0, S(stream)>>>(T, k, d, N, ld, h, npart);
    ASB_LAUNCH_CHECK();
    return sum_partials(npart, nb, 1, nrm2, S(stream));
}

int asb_row_scale(double* T, int64_t k, int64_t d, int64_t ld, const double* nrm2, asb_stream_t stream) {
    ASB_REQUIRE(T && nrm2 && ld > 0, "row_scale: bad arguments");
    row_scale_kernel<<<(unsigned)cdiv(ld, 256), 256, 0, S(stream)>>>(T, k, d, ld, nrm2);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

// ---------------------------------------------------------------------------------------------- cuSOLVER syevd
static cusolverDnHandle_t g_solver = nullptr;

int64_t asb_syevd_workspace_bytes(int64_t n) {
    if (!g_solver && cusolverDnCreate(&g_solver) != CUSOLVER_STATUS_SUCCESS) return -1;
    int lwork = 0;
    if (cusolverDnDsyevd_bufferSize(g_solver, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, (int)n, nullptr, (int)n,
                                    nullptr, &lwork) != CUSOLVER_STATUS_SUCCESS)
        return -1;
    return (int64_t)lwork * 8 + 256;
}

int asb_syevd(double* A, int64_t lda, int64_t n, double* w, void* ws, int64_t ws_bytes, int* info_dev,
              asb_stream_t stream) {
    ASB_REQUIRE(A && w && ws && info_dev && n > 0 && lda >= n, "syevd: bad arguments");
    if (!g_solver && cusolverDnCreate(&g_solver) != CUSOLVER_STATUS_SUCCESS)
        return fail(ASB_ERR_CUDA, "cusolverDnCreate failed", __FILE__, __LINE__);
    if (cusolverDnSetStream(g_solver, S(stream)) != CUSOLVER_STATUS_SUCCESS)
        return fail(ASB_ERR_CUDA, "cusolverDnSetStream failed", __FILE__, __LINE__);
    const int lwork = (int)((ws_bytes - 256) / 8);
    cusolverStatus_t st = cusolverDnDsyevd(g_solver, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, (int)n, A, (int)lda,
                                           w, reinterpret_cast<double*>(ws), lwork, info_dev);
    if (st != CUSOLVER_STATUS_SUCCESS) return fail(ASB_ERR_NUMERIC, "cusolverDnDsyevd failed", __FILE__, __LINE__);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ASB_OK;
}

}  // extern "C"
