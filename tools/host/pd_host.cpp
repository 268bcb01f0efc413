// CPU build of the projective-dynamics local-step arithmetic (animsnapbases_b200/csrc/pd_elements.cuh, the same inline
// functions the CUDA kernels call) so that it can be checked against the oracle on a machine without a GPU.
// Test infrastructure: g++ -O2 -shared -fPIC -I animsnapbases_b200/csrc tools/host/pd_host.cpp -o <tmp>/libpd_host.so
#include "pd_elements.cuh"

extern "C" void pd_host(int kind, const double* q, int64_t ldq, int64_t frames, const int32_t* elems, const double* rest,
                        int64_t e, double smin, double smax, double* out, int64_t ldo) {
    const int nodes = asb::pd::kind_nodes(kind), nrest = asb::pd::kind_rest(kind);
    for (int64_t f = 0; f < frames; ++f)
        for (int64_t t = 0; t < e; ++t)
            asb::pd::pd_element(kind, q + f * 3 * ldq, ldq, elems + t * nodes, rest + t * nrest, smin, smax, out + f * 3 * ldo, ldo, t);
}

extern "C" void pd_host_vertex_bending(const double* q, int64_t ldq, int64_t frames, const int32_t* verts, const int32_t* ptr,
                                       const int32_t* nbr, const double* w, const double* rest, int64_t n, int prevent_flips,
                                       double* out, int64_t ldo) {
    for (int64_t f = 0; f < frames; ++f)
        for (int64_t i = 0; i < n; ++i)
            asb::pd::pd_vertex_bending(q + f * 3 * ldq, ldq, verts[i], nbr + ptr[i], w + ptr[i], ptr[i + 1] - ptr[i], rest + 5 * i,
                                       prevent_flips, out + f * 3 * ldo, ldo, i);
}
