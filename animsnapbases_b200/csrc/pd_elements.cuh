// One element of the projective-dynamics local step: gather its vertices from a planar position frame
// q[d * ldq + v], project, write its `rows` snapshot rows into the planar frame out[d * ldo + rows * t + j].
// Host + device (see pd_math.cuh); pd_local.cu calls pd_element from one thread per (element, frame).
#pragma once
#include <stdint.h>

#include "pd_math.cuh"

namespace asb {
namespace pd {

enum Kind { TET_ROTATION = 0, TET_STRAIN = 1, TRI_STRAIN = 2, EDGE_SPRING = 3 };

ASB_HD int kind_nodes(int kind) { return kind <= TET_STRAIN ? 4 : (kind == TRI_STRAIN ? 3 : 2); }
ASB_HD int kind_rows(int kind) { return kind <= TET_STRAIN ? 3 : (kind == TRI_STRAIN ? 2 : 1); }
ASB_HD int kind_rest(int kind) { return kind <= TET_STRAIN ? 9 : (kind == TRI_STRAIN ? 10 : 1); }   // doubles of rest data

// rest data per element:  tets: DmInv (3 x 3 row-major, Constraint_projections.py:504, :639);  triangles: P (3 x 2 row-major,
// :372-374) then DmInv (2 x 2 row-major, :381);  edges: rest length d (:280).
ASB_HD void pd_element(int kind, const double* __restrict__ q, int64_t ldq, const int32_t* __restrict__ nodes,
                       const double* __restrict__ rest, double smin, double smax, double* __restrict__ out, int64_t ldo,
                       int64_t t) {
    if (kind == TET_ROTATION || kind == TET_STRAIN) {
        double Ds[9], F[9], M[9];
        for (int r = 0; r < 3; ++r) {
            const double x4 = q[r * ldq + nodes[3]];
            for (int c = 0; c < 3; ++c) Ds[3 * r + c] = q[r * ldq + nodes[c]] - x4;        // columns x1-x4, x2-x4, x3-x4
        }
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c)
                F[3 * r + c] = Ds[3 * r] * rest[c] + Ds[3 * r + 1] * rest[3 + c] + Ds[3 * r + 2] * rest[6 + c];   // Ds DmInv
        if (kind == TET_ROTATION) {
            polar_rotation3(F, M);                                                        // rows = R^T (:687)
            for (int j = 0; j < 3; ++j)
                for (int d = 0; d < 3; ++d) out[d * ldo + 3 * t + j] = M[3 * d + j];
        } else {
            clamped_deformation3(F, smin, smax, M);                                       // rows = Fhat (:554)
            for (int j = 0; j < 3; ++j)
                for (int d = 0; d < 3; ++d) out[d * ldo + 3 * t + j] = M[3 * j + d];
        }
    } else if (kind == TRI_STRAIN) {
        const double* P = rest;            // 3 x 2
        const double* Di = rest + 6;       // 2 x 2
        double Ds[6], D2[4], A[4], Fh[4];
        for (int r = 0; r < 3; ++r) {
            const double x1 = q[r * ldq + nodes[0]];
            Ds[2 * r] = q[r * ldq + nodes[1]] - x1;
            Ds[2 * r + 1] = q[r * ldq + nodes[2]] - x1;
        }
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b) D2[2 * a + b] = P[a] * Ds[b] + P[2 + a] * Ds[2 + b] + P[4 + a] * Ds[4 + b];   // P^T Ds
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b) A[2 * a + b] = D2[2 * a] * Di[b] + D2[2 * a + 1] * Di[2 + b];
        clamped_deformation2(A, smin, smax, Fh);
        for (int j = 0; j < 2; ++j)
            for (int d = 0; d < 3; ++d) out[d * ldo + 2 * t + j] = P[2 * d] * Fh[j] + P[2 * d + 1] * Fh[2 + j];       // (P Fhat)^T
    } else {
        double sp[3], len2 = 0.0;
        for (int d = 0; d < 3; ++d) {
            sp[d] = q[d * ldq + nodes[1]] - q[d * ldq + nodes[0]];
            len2 += sp[d] * sp[d];
        }
        const double len = sqrt(len2);
        // 0.5 (pi2 - pi1) with pi1 = p1 + delta n, pi2 = p2 - delta n, delta = (len - d) / 2  (:307-311), i.e. n d / 2
        const double scale = len == 0.0 ? 0.0 : 0.5 * (len - 2.0 * (0.5 * (len - rest[0]))) / len;
        for (int d = 0; d < 3; ++d) out[d * ldo + t] = scale * sp[d];
    }
}

// Vertex bending (Constraint_projections.py:197-216): constrained vertex i = verts[i] with the star nbr[ptr[i] .. ptr[i+1]) and
// cotangent weights w (already divided by the Voronoi area, :169); rest = [rest mean curvature, rest normal (3), rest
// normal . rest mean-curvature vector] (:184-187).  One row per constrained vertex.
ASB_HD void pd_vertex_bending(const double* __restrict__ q, int64_t ldq, int32_t v, const int32_t* __restrict__ nbr,
                              const double* __restrict__ w, int deg, const double* __restrict__ rest, int prevent_flips,
                              double* __restrict__ out, int64_t ldo, int64_t i) {
    double star[3] = {0.0, 0.0, 0.0};
    const double qv[3] = {q[v], q[ldq + v], q[2 * ldq + v]};
    for (int j = 0; j < deg; ++j) {
        const int32_t n = nbr[j];
        for (int d = 0; d < 3; ++d) star[d] += (qv[d] - q[d * ldq + n]) * w[j];                  // :200-202
    }
    const double norm = sqrt(star[0] * star[0] + star[1] * star[1] + star[2] * star[2]);
    double c[3];
    if (norm < 1e-10) { for (int d = 0; d < 3; ++d) c[d] = rest[1 + d] * rest[0]; }               // :205-206
    else { const double s = rest[0] / norm; for (int d = 0; d < 3; ++d) c[d] = star[d] * s; }     // :207-208
    if (prevent_flips) {                                                                          // :210-213
        const double dot = rest[1] * c[0] + rest[2] * c[1] + rest[3] * c[2];
        if (norm > 1e-5 && dot * rest[4] < 0.0) { c[0] = -c[0]; c[1] = -c[1]; c[2] = -c[2]; }
    }
    for (int d = 0; d < 3; ++d) out[d * ldo + i] = c[d];
}

}  // namespace pd
}  // namespace asb
