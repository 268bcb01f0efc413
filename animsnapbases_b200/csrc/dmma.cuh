// FP64 tensor-core primitives shared by the DMMA kernels (dmma_gemm.cu, admm.cu).
// sm_100a has no fp64 kind of tcgen05.mma: every f64 mma.sync shape lowers to SASS DMMA.8x8x4, so m8n8k4 is the
// native instruction.  Fragment layout (lane = 4*lr + lc): A[row lr][k lc], B[k lc][col lr], D[row lr][cols 2lc, 2lc+1].
#pragma once
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

// m-tiles (8 rows each) of a `mtiles`-tile operand owned by row-warp `wm` of `WM`: first (mtiles % WM) warps get one more.
__device__ __forceinline__ void mtile_range(int mtiles, int WM, int wm, int& first, int& count) {
    const int q = mtiles / WM, r = mtiles % WM;
    count = q + (wm < r ? 1 : 0);
    first = wm * q + min(wm, r);
}

}  // namespace asb
