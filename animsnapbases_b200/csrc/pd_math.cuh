// Per-element arithmetic of the projective-dynamics local step: small symmetric eigen-problems and the
// projections built on them.  Host + device inline functions so that the same arithmetic can be compiled for the CPU and
// checked against the oracle where no GPU is present (tests/test_host_logic.py::test_pd_math_on_host); pd_local.cu wraps
// them in one-thread-per-element kernels.
//
// The reference works through LAPACK's SVD  F = U diag(s) V^T  (projective_dynamics/Constraint_projections.py:419, :547,
// :682).  Everything it needs can be written with the right singular pairs only:  u_i = F v_i / s_i,  (s_i^2, v_i) the
// eigen-pairs of F^T F (cyclic Jacobi, a handful of rotations for 3 x 3), sorted descending like LAPACK's s:
//     polar rotation        U V^T           = sum_i (1 / s_i)   (F v_i) v_i^T
//     clamped deformation   U diag(c) V^T   = sum_i (c_i / s_i) (F v_i) v_i^T,   c = clip(s)
// The result does not depend on the choice of basis inside a repeated singular value.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define ASB_HD __host__ __device__ __forceinline__
#else
#define ASB_HD inline
#endif

namespace asb {
namespace pd {

// Eigen-decomposition of a symmetric 3 x 3 matrix A (row-major, overwritten): A = V diag(w) V^T, w descending, eigenvector i
// in column i of V (row-major 3 x 3).  Cyclic Jacobi; 3 x 3 converges quadratically, 8 sweeps are far more than needed.
ASB_HD void sym_eig3(double A[9], double w[3], double V[9]) {
    for (int i = 0; i < 9; ++i) V[i] = (i % 4 == 0) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 8; ++sweep) {
        const double off = fabs(A[1]) + fabs(A[2]) + fabs(A[5]);
        const double dia = fabs(A[0]) + fabs(A[4]) + fabs(A[8]);
        if (off <= 1e-300 || off <= 1e-18 * dia) break;
        for (int pair = 0; pair < 3; ++pair) {
            const int p = pair == 2 ? 1 : 0, q = pair == 0 ? 1 : 2;
            const double apq = A[3 * p + q];
            if (apq == 0.0) continue;
            const double theta = (A[3 * q + q] - A[3 * p + p]) / (2.0 * apq);
            const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
            for (int k = 0; k < 3; ++k) {          // A <- A J
                const double akp = A[3 * k + p], akq = A[3 * k + q];
                A[3 * k + p] = c * akp - s * akq;
                A[3 * k + q] = s * akp + c * akq;
            }
            for (int k = 0; k < 3; ++k) {          // A <- J^T A
                const double apk = A[3 * p + k], aqk = A[3 * q + k];
                A[3 * p + k] = c * apk - s * aqk;
                A[3 * q + k] = s * apk + c * aqk;
            }
            for (int k = 0; k < 3; ++k) {          // V <- V J
                const double vkp = V[3 * k + p], vkq = V[3 * k + q];
                V[3 * k + p] = c * vkp - s * vkq;
                V[3 * k + q] = s * vkp + c * vkq;
            }
        }
    }
    w[0] = A[0]; w[1] = A[4]; w[2] = A[8];
    for (int a = 0; a < 2; ++a)                    // sort descending (3 elements)
        for (int b = 0; b < 2 - a; ++b)
            if (w[b] < w[b + 1]) {
                const double tw = w[b]; w[b] = w[b + 1]; w[b + 1] = tw;
                for (int k = 0; k < 3; ++k) { const double tv = V[3 * k + b]; V[3 * k + b] = V[3 * k + b + 1]; V[3 * k + b + 1] = tv; }
            }
}

ASB_HD double det3(const double F[9]) {
    return F[0] * (F[4] * F[8] - F[5] * F[7]) - F[1] * (F[3] * F[8] - F[5] * F[6]) + F[2] * (F[3] * F[7] - F[4] * F[6]);
}

// out = sum_i g_i (F v_i) v_i^T  for the right singular pairs of F (row-major 3 x 3); g is filled by `gain(s, g)`.
template <class Gain>
ASB_HD void spectral_map3(const double F[9], double out[9], Gain gain) {
    double G[9], w[3], V[9], s[3], g[3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) G[3 * i + j] = F[i] * F[j] + F[3 + i] * F[3 + j] + F[6 + i] * F[6 + j];   // F^T F
    sym_eig3(G, w, V);
    for (int i = 0; i < 3; ++i) s[i] = sqrt(fmax(w[i], 0.0));
    gain(s, g);
    for (int i = 0; i < 9; ++i) out[i] = 0.0;
    for (int i = 0; i < 3; ++i) {
        double Fv[3];
        for (int r = 0; r < 3; ++r) Fv[r] = F[3 * r] * V[i] + F[3 * r + 1] * V[3 + i] + F[3 * r + 2] * V[6 + i];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) out[3 * r + c] = fma(g[i] * Fv[r], V[3 * c + i], out[3 * r + c]);
    }
}

struct UnitGain {          // U V^T
    ASB_HD void operator()(const double s[3], double g[3]) const {
        for (int i = 0; i < 3; ++i) g[i] = 1.0 / s[i];
    }
};
struct ClampGain {         // U diag(clip(s)) V^T, last one negated for an inverted element
    double smin, smax;
    bool inverted;
    ASB_HD void operator()(const double s[3], double g[3]) const {
        for (int i = 0; i < 3; ++i) g[i] = fmin(fmax(s[i], smin), smax) / s[i];
        if (inverted) g[2] = -g[2];
    }
};

// Constraint_projections.py:680-687: R = U V^T, third COLUMN of R negated when det R < 0 (the reference's choice).
ASB_HD void polar_rotation3(const double F[9], double R[9]) {
    spectral_map3(F, R, UnitGain());
    if (det3(R) < 0.0) { R[2] = -R[2]; R[5] = -R[5]; R[8] = -R[8]; }
}

// Constraint_projections.py:547-553: Fhat = U diag(clip(s)) V^T, the smallest clipped value negated for an inverted element.
ASB_HD void clamped_deformation3(const double F[9], double smin, double smax, double Fhat[9]) {
    ClampGain gain;
    gain.smin = smin; gain.smax = smax; gain.inverted = det3(F) < 0.0;
    spectral_map3(F, Fhat, gain);
}

// 2 x 2 version (Constraint_projections.py:419-421): Fhat = U diag(clip(s)) V^T of A (row-major 2 x 2), closed-form eigenpairs.
ASB_HD void clamped_deformation2(const double A[4], double smin, double smax, double Fhat[4]) {
    const double a = A[0] * A[0] + A[2] * A[2], b = A[0] * A[1] + A[2] * A[3], d = A[1] * A[1] + A[3] * A[3];   // A^T A
    const double half = 0.5 * (a - d), rad = sqrt(half * half + b * b), mean = 0.5 * (a + d);
    double v0x, v0y;                                  // eigenvector of the larger eigenvalue mean + rad
    if (rad <= 1e-300 * (1.0 + mean)) { v0x = 1.0; v0y = 0.0; }
    else if (half >= 0.0) { v0x = half + rad; v0y = b; }
    else { v0x = b; v0y = rad - half; }
    const double nrm = sqrt(v0x * v0x + v0y * v0y);
    v0x /= nrm; v0y /= nrm;
    const double vx[2] = {v0x, -v0y}, vy[2] = {v0y, v0x};          // v_0 = (v0x, v0y), v_1 = (-v0y, v0x)
    const double s[2] = {sqrt(fmax(mean + rad, 0.0)), sqrt(fmax(mean - rad, 0.0))};
    for (int i = 0; i < 4; ++i) Fhat[i] = 0.0;
    for (int i = 0; i < 2; ++i) {
        const double g = fmin(fmax(s[i], smin), smax) / s[i];
        const double Avx = A[0] * vx[i] + A[1] * vy[i], Avy = A[2] * vx[i] + A[3] * vy[i];
        Fhat[0] = fma(g * Avx, vx[i], Fhat[0]); Fhat[1] = fma(g * Avx, vy[i], Fhat[1]);
        Fhat[2] = fma(g * Avy, vx[i], Fhat[2]); Fhat[3] = fma(g * Avy, vy[i], Fhat[3]);
    }
}

}  // namespace pd
}  // namespace asb
