// Projective-dynamics local step as a snapshot producer (projective_dynamics/Constraint_projections.py
// get_pi of TetDeformationGradientConstraint :669-687, TetStrainConstraint :534-554, TriStrainConstraint :407-426,
// EdgeSpringConstraint :291-312, stacked like Simulators.get_group_ST_p :354-356).
//
// HBM-bound gather/scatter work: one thread per (element, frame); a thread gathers 2-4 vertices (24 B each and per
// dimension plane, served by L2 for the ~6 elements that share a vertex), keeps the 3 x 3 / 2 x 2 problem in registers
// (pd_math.cuh: Jacobi eigen-pairs of F^T F instead of an SVD) and writes its rows of the planar snapshot frame.
// Algorithmic bytes per element and frame: rest data (72 / 80 / 8 B) + node ids (16 / 12 / 8 B) + rows * 24 B written + the
// element's share of the vertex planes (24 B * |V| / e).
#include "common.cuh"
#include "pd_elements.cuh"

namespace asb {

__global__ void __launch_bounds__(128) pd_project_kernel(int kind, const double* __restrict__ q, int64_t ldq, const int32_t* __restrict__ elems,
                                                        const double* __restrict__ rest, int64_t e, double smin, double smax,
                                                        double* __restrict__ out, int64_t ldo) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= e) return;
    const int64_t f = blockIdx.y;
    pd::pd_element(kind, q + f * 3 * ldq, ldq, elems + t * pd::kind_nodes(kind), rest + t * pd::kind_rest(kind), smin, smax,
                   out + f * 3 * ldo, ldo, t);
}

__global__ void __launch_bounds__(128) pd_vertex_bending_kernel(const double* __restrict__ q, int64_t ldq, const int32_t* __restrict__ verts,
                                                               const int32_t* __restrict__ ptr, const int32_t* __restrict__ nbr,
                                                               const double* __restrict__ w, const double* __restrict__ rest, int64_t n,
                                                               int prevent_flips, double* __restrict__ out, int64_t ldo) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t f = blockIdx.y;
    const int32_t p0 = ptr[i];
    pd::pd_vertex_bending(q + f * 3 * ldq, ldq, verts[i], nbr + p0, w + p0, ptr[i + 1] - p0, rest + 5 * i, prevent_flips,
                          out + f * 3 * ldo, ldo, i);
}

}  // namespace asb

extern "C" {
using namespace asb;

int asb_pd_project(int kind, const double* q, int64_t ldq, int64_t frames, const int32_t* elems, const double* rest, int64_t e,
                   double sigma_min, double sigma_max, double* out, int64_t ldo, asb_stream_t stream) {
    ASB_REQUIRE(q && elems && rest && out && e > 0 && frames > 0 && frames < 65536, "pd_project: bad arguments");
    ASB_REQUIRE(kind >= 0 && kind <= 3, "pd_project: kind must be 0 (tet rotation), 1 (tet strain), 2 (triangle strain) or 3 (edge spring)");
    ASB_REQUIRE(ldo >= e * pd::kind_rows(kind) && ldq > 0, "pd_project: output plane too short");
    pd_project_kernel<<<dim3((unsigned)cdiv(e, 128), (unsigned)frames), 128, 0, S(stream)>>>(kind, q, ldq, elems, rest, e, sigma_min,
                                                                                           sigma_max, out, ldo);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_pd_vertex_bending(const double* q, int64_t ldq, int64_t frames, const int32_t* verts, const int32_t* star_ptr,
                          const int32_t* star_nbr, const double* star_w, const double* rest, int64_t n, int prevent_flips,
                          double* out, int64_t ldo, asb_stream_t stream) {
    ASB_REQUIRE(q && verts && star_ptr && star_nbr && star_w && rest && out && n > 0 && frames > 0 && frames < 65536 && ldo >= n && ldq > 0,
                "pd_vertex_bending: bad arguments");
    pd_vertex_bending_kernel<<<dim3((unsigned)cdiv(n, 128), (unsigned)frames), 128, 0, S(stream)>>>(q, ldq, verts, star_ptr, star_nbr,
                                                                                                  star_w, rest, n, prevent_flips, out, ldo);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // extern "C"
