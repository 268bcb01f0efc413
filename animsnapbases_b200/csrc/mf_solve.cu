// Solve phase of the multifrontal Cholesky factorisation (animsnapbases_b200/mfsolve.py): the two triangular solves of
// a heat-method query (utils/support.py:182, :205 -- SuperLU `solve` in the reference) as one dense pass per tree level.
//
// Every front t keeps the dense operator  Mt = [ L11^-1 ; -L21 L11^-1 ]  (rows [0,n1) and [row2, row2+n2), row stride ldm):
//   forward  (leaves -> root):  f = rhs[own] (+ children's update vectors, gathered through `src`)
//                               y[own] = Mt[:n1] f1 ;  upd_t = f2 + Mt[row2:] f1
//   backward (root -> leaves):  x[own] = Mt^T [ y[own] ; x[bnd] ]
// HBM-bound: each launch streams the operators of one level once per right-hand side; all reductions are fixed-order.
#include "common.cuh"

namespace asb {

constexpr int MF_ROWS = 32;      // rows (forward) / columns (backward) of Mt per CTA; mfsolve.ROWS_PER_BLOCK
constexpr int MF_CHUNK = 1024;   // staged vector chunk (doubles)

struct FrontDesc {
    long long own_start, n1, n2, mt_off, slot_off, upd_off, ldm, row2;
};

__device__ __forceinline__ double gather_children(const int32_t* __restrict__ src0, const int32_t* __restrict__ src1,
                                                  const double* upd, long long slot) {
    double f = 0.0;
    const int s0 = src0[slot];
    if (s0 >= 0) f += upd[s0];
    const int s1 = src1[slot];
    if (s1 >= 0) f += upd[s1];
    return f;
}

// grid.x = work item (front, first row), grid.y = right-hand side.  256 threads = 8 warps x 4 rows.
__global__ void __launch_bounds__(256) mf_forward_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                        const int32_t* __restrict__ src, int64_t slots,
                                                        const int32_t* __restrict__ work, const double* __restrict__ rhs,
                                                        double* __restrict__ y, double* __restrict__ upd, int64_t ldx,
                                                        int64_t ldu) {
    __shared__ double fs[MF_CHUNK];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    const int r0 = work[2 * blockIdx.x + 1];
    const int64_t b = blockIdx.y;
    const double* rb = rhs + b * ldx + d.own_start;
    double* ub = upd + b * ldu;
    const int32_t* src0 = src + d.slot_off;
    const int32_t* src1 = src + slots + d.slot_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    const double* rowp[4];
    int jmax[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = r0 + warp * 4 + i;
        const long long srow = r < n1 ? r : d.row2 + (r - n1);
        rowp[i] = vals + d.mt_off + srow * d.ldm;
        jmax[i] = r >= n ? 0 : (r < n1 ? r + 1 : n1);   // L11^-1 is lower triangular
    }
    // columns needed by this CTA: own rows only reach column r0 + MF_ROWS - 1
    const int cmax = (r0 + MF_ROWS <= n1) ? r0 + MF_ROWS : n1;
    for (int c0 = 0; c0 < cmax; c0 += MF_CHUNK) {
        const int len = min(MF_CHUNK, cmax - c0);
        __syncthreads();
        for (int j = threadIdx.x; j < len; j += blockDim.x) fs[j] = rb[c0 + j] + gather_children(src0, src1, ub, c0 + j);
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int je = min(len, jmax[i] - c0);
            const double* p = rowp[i] + c0;
            double a = acc[i];
            for (int j = lane; j < je; j += 32) a = fma(p[j], fs[j], a);
            acc[i] = a;
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double s = warp_sum(acc[i]);
        const int r = r0 + warp * 4 + i;
        if (lane == 0 && r < n) {
            if (r < n1)
                y[b * ldx + d.own_start + r] = s;
            else
                ub[d.upd_off + (r - n1)] = gather_children(src0, src1, ub, r) + s;
        }
    }
}

// grid.x = work item (front, first column), grid.y = right-hand side.  256 threads = 32 columns x 8 row groups.
__global__ void __launch_bounds__(256) mf_backward_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                         const int32_t* __restrict__ bnd, const int32_t* __restrict__ work,
                                                         const double* __restrict__ y, double* __restrict__ x, int64_t ldx) {
    __shared__ double vs[MF_CHUNK];
    __shared__ double red[8][33];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    const int c0 = work[2 * blockIdx.x + 1];
    const int64_t b = blockIdx.y;
    const double* yb = y + b * ldx + d.own_start;
    double* xb = x + b * ldx;
    const int32_t* bl = bnd + d.upd_off;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    const int col = c0 + tx;
    const double* mp = vals + d.mt_off + col;
    double acc = 0.0;
    // own rows r < c0 hold zeros in all of this CTA's columns (lower triangular): start at c0
    for (int rb0 = c0; rb0 < n; rb0 += MF_CHUNK) {
        const int len = min(MF_CHUNK, n - rb0);
        __syncthreads();
        for (int j = threadIdx.x; j < len; j += blockDim.x) {
            const int r = rb0 + j;
            vs[j] = r < n1 ? yb[r] : xb[bl[r - n1]];
        }
        __syncthreads();
        if (col < n1) {
            for (int j = ty; j < len; j += 8) {
                const int r = rb0 + j;
                if (r < n1 && r < col) continue;
                const long long srow = r < n1 ? r : d.row2 + (r - n1);
                acc = fma(mp[srow * d.ldm], vs[j], acc);
            }
        }
    }
    red[ty][tx] = acc;
    __syncthreads();
    if (ty == 0 && col < n1) {
        double s = red[0][tx];
#pragma unroll
        for (int k = 1; k < 8; ++k) s += red[k][tx];
        xb[d.own_start + col] = s;
    }
}

__global__ void gather_rows_kernel(const double* __restrict__ in, const int32_t* __restrict__ map, double* __restrict__ out,
                                   int64_t n, int64_t ldin, int64_t ldout) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[(int64_t)blockIdx.y * ldout + i] = in[(int64_t)blockIdx.y * ldin + map[i]];
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_mf_forward(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* work,
                   int64_t nwork, const double* rhs, double* y, double* upd, int64_t nrhs, int64_t ldx, int64_t ldu,
                   asb_stream_t stream) {
    ASB_REQUIRE(desc && vals && src && work && rhs && y && upd && nwork > 0 && nrhs > 0 && nrhs < 65536, "mf_forward: bad arguments");
    mf_forward_kernel<<<dim3((unsigned)nwork, (unsigned)nrhs), 256, 0, S(stream)>>>(
        reinterpret_cast<const FrontDesc*>(desc), vals, src, slots, work, rhs, y, upd, ldx, ldu);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_mf_backward(const int64_t* desc, const double* vals, const int32_t* bnd, const int32_t* work, int64_t nwork,
                    const double* y, double* x, int64_t nrhs, int64_t ldx, asb_stream_t stream) {
    ASB_REQUIRE(desc && vals && bnd && work && y && x && nwork > 0 && nrhs > 0 && nrhs < 65536, "mf_backward: bad arguments");
    mf_backward_kernel<<<dim3((unsigned)nwork, (unsigned)nrhs), 256, 0, S(stream)>>>(
        reinterpret_cast<const FrontDesc*>(desc), vals, bnd, work, y, x, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

int asb_gather_rows(const double* in, const int32_t* map, double* out, int64_t n, int64_t nrhs, int64_t ldin, int64_t ldout,
                    asb_stream_t stream) {
    ASB_REQUIRE(in && map && out && n > 0 && nrhs > 0 && nrhs < 65536, "gather_rows: bad arguments");
    gather_rows_kernel<<<dim3((unsigned)cdiv(n, 256), (unsigned)nrhs), 256, 0, S(stream)>>>(in, map, out, n, ldin, ldout);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

}  // extern "C"
