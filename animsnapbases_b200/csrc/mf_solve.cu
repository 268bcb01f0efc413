// Solve phase of the multifrontal Cholesky factorisation (animsnapbases_b200/mfsolve.py): the two triangular solves of
// a heat-method query (utils/support.py:182, :205 -- SuperLU `solve` in the reference) as one dense pass per tree level.
//
// Every front t keeps the dense operator  Mt = [ L11^-1 ; -L21 L11^-1 ]  (rows [0,n1) and [row2, row2+n2), row stride ldm):
//   forward  (leaves -> root):  f = rhs[own] (+ children's update vectors, gathered through `src`)
//                               y[own] = Mt[:n1] f1 ;  upd_t = f2 + Mt[row2:] f1
//   backward (root -> leaves):  x[own] = Mt^T [ y[own] ; x[bnd] ]
// HBM-bound: each launch streams the operators of one level once per right-hand side; all reductions are fixed-order.
#include "dmma.cuh"
#include <cstdlib>

namespace asb {

constexpr int MF_ROWS = 32;      // rows (forward) / columns (backward) of Mt per CTA; mfsolve.ROWS_PER_BLOCK
constexpr int MF_CHUNK = 1024;   // staged vector chunk (doubles)
constexpr int MF_MRHS_MIN = 3;
constexpr int MF_LEVEL_COLS = 6;   // level table row: forward work list, its length, backward work list, its length, max n1, max n1 + n2   // from this many right-hand sides on, the register-blocked multi-RHS kernels run

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

// grid.x = work item (front, first row), grid.y = right-hand side.  RPW rows per warp: 256 threads = 8 warps x 4 rows for the
// wide levels (thousands of CTAs, bandwidth bound), 1024 threads = 32 warps x 1 row for the top of the tree, where a level has
// fewer CTAs than the machine has SMs and the pass is bound by the latency of one CTA (r01: 15 of the 30 launches of a
// single-source solve ran 32..128 CTAs for 20-30 us each).  The rows of a warp advance together, so their loads are in
// flight at the same time.
template <int RPW>
__device__ __forceinline__ void mf_forward_block(const FrontDesc& d, const int r0, const int64_t b, double* fs,
                                                 const double* __restrict__ vals, const int32_t* __restrict__ src,
                                                 int64_t slots, const double* __restrict__ rhs, double* __restrict__ y,
                                                 double* __restrict__ upd, int64_t ldx, int64_t ldu) {
    const double* rb = rhs + b * ldx + d.own_start;
    double* ub = upd + b * ldu;
    const int32_t* src0 = src + d.slot_off;
    const int32_t* src1 = src + slots + d.slot_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    double acc[RPW];
    const double* rowp[RPW];
    int jmax[RPW], jall = 0;
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
        const int r = r0 + warp * RPW + i;
        const long long srow = r < n1 ? r : d.row2 + (r - n1);
        rowp[i] = vals + d.mt_off + (r < n ? srow : 0) * d.ldm;
        jmax[i] = r >= n ? 0 : (r < n1 ? r + 1 : n1);   // L11^-1 is lower triangular
        jall = max(jall, jmax[i]);
        acc[i] = 0.0;
    }
    // columns needed by this CTA: own rows only reach column r0 + MF_ROWS - 1
    const int cmax = (r0 + MF_ROWS <= n1) ? r0 + MF_ROWS : n1;
    for (int c0 = 0; c0 < cmax; c0 += MF_CHUNK) {
        const int len = min(MF_CHUNK, cmax - c0);
        __syncthreads();
        for (int j = threadIdx.x; j < len; j += blockDim.x) fs[j] = rb[c0 + j] + gather_children(src0, src1, ub, c0 + j);
        __syncthreads();
        const int je = min(len, jall - c0);
#pragma unroll 2
        for (int j = lane; j < je; j += 32) {
            const double f = fs[j];
            double a[RPW];
#pragma unroll
            for (int i = 0; i < RPW; ++i) a[i] = (c0 + j < jmax[i]) ? rowp[i][c0 + j] : 0.0;
#pragma unroll
            for (int i = 0; i < RPW; ++i) acc[i] = fma(a[i], f, acc[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
        const double s = warp_sum(acc[i]);
        const int r = r0 + warp * RPW + i;
        if (lane == 0 && r < n) {
            if (r < n1)
                y[b * ldx + d.own_start + r] = s;
            else
                ub[d.upd_off + (r - n1)] = gather_children(src0, src1, ub, r) + s;
        }
    }
}

template <int RPW>
__global__ void __launch_bounds__(1024 / RPW) mf_forward_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                               const int32_t* __restrict__ src, int64_t slots,
                                                               const int32_t* __restrict__ work, const double* __restrict__ rhs,
                                                               double* __restrict__ y, double* __restrict__ upd, int64_t ldx,
                                                               int64_t ldu) {
    __shared__ double fs[MF_CHUNK];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    mf_forward_block<RPW>(d, work[2 * blockIdx.x + 1], blockIdx.y, fs, vals, src, slots, rhs, y, upd, ldx, ldu);
}

// Forward pass of a DELTA right-hand side: only the fronts on the path from the source's front to the root see a
// non-zero vector (everything else of y and upd stays at its memset zero), so a level is at most one front per
// right-hand side.  pathfront[b][level] = that front or -1 (mf_path_kernel); grid.x = row blocks of the largest front of
// the level, grid.y = right-hand side.
__global__ void __launch_bounds__(1024) mf_forward_path_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                             const int32_t* __restrict__ src, int64_t slots,
                                                             const int32_t* __restrict__ pathfront, int level, int nlevels,
                                                             const double* __restrict__ rhs, double* __restrict__ y,
                                                             double* __restrict__ upd, int64_t ldx, int64_t ldu) {
    __shared__ double fs[MF_CHUNK];
    const int f = pathfront[(int64_t)blockIdx.y * nlevels + level];
    if (f < 0) return;
    const FrontDesc d = desc[f];
    const int r0 = blockIdx.x * MF_ROWS;
    if (r0 >= d.n1 + d.n2) return;
    mf_forward_block<1>(d, r0, blockIdx.y, fs, vals, src, slots, rhs, y, upd, ldx, ldu);
}

// One thread per right-hand side: unit entry at the source's elimination position, and the chain of fronts above it.
__global__ void mf_path_kernel(const int64_t* __restrict__ idx, const int32_t* __restrict__ order,
                               const int64_t* __restrict__ fstart, const int32_t* __restrict__ parent,
                               const int32_t* __restrict__ depth, int64_t nfronts, int nlevels, int nrhs,
                               int32_t* __restrict__ pathfront, double* __restrict__ rp, int64_t ldx) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nrhs) return;
    const int64_t p = order[idx[b]];
    int64_t lo = 0, hi = nfronts - 1;           // last front with fstart <= p (fronts are stored in elimination order)
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) >> 1;
        if (fstart[mid] <= p) lo = mid; else hi = mid - 1;
    }
    rp[(int64_t)b * ldx + p] = 1.0;
    for (int f = (int)lo; f >= 0; f = parent[f]) pathfront[(int64_t)b * nlevels + depth[f]] = f;
}

// grid.x = work item (front, first column), grid.y = right-hand side.  32 columns x RG row groups: RG = 8 (256 threads) for
// the wide levels, RG = 32 (1024 threads) for the top of the tree (see mf_forward_block).
template <int RG>
__global__ void __launch_bounds__(32 * RG) mf_backward_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                             const int32_t* __restrict__ bnd, const int32_t* __restrict__ work,
                                                             const double* __restrict__ y, double* __restrict__ x, int64_t ldx) {
    __shared__ double vs[MF_CHUNK];
    __shared__ double red[RG][33];
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
            // rows r < col of the own block hold stored zeros (L11^-1 is lower triangular): no skip test, so that the
            // strided loads of four rows are in flight together
#pragma unroll 4
            for (int j = ty; j < len; j += RG) {
                const int r = rb0 + j;
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
        for (int k = 1; k < RG; ++k) s += red[k][tx];
        xb[d.own_start + col] = s;
    }
}

// ------------------------------------------------------------------------------------------ small fronts, one RHS
// The bottom levels of the tree hold tens of thousands of fronts with n1 <= 64 unknowns: per work item (32 rows / 32 columns
// of one front) there are a few KB of operator and a chain of dependent loads (descriptor -> index lists -> vector ->
// operator).  The CTA-per-item kernels above keep 8 items in flight per SM and pay two CTA barriers plus a shuffle reduction
// per row (r02 ncu, 1M-vertex grid, leaf level: forward 339 us for 409 MB, backward 181 us for 448 MB).  Here ONE WARP owns
// an item, 8 warps per CTA, no shared memory and no barrier: 64 items in flight per SM.
//   forward : lane = row, walks its own row of Mt (contiguous), f[j] is broadcast from lane j;  sum over j ascending
//   backward: lane = column, every row of Mt is one coalesced 256-byte read, v[r] is broadcast;  sum over r ascending
constexpr int MF_SMALL_WARPS = 8;

__global__ void __launch_bounds__(32 * MF_SMALL_WARPS) mf_forward_small_kernel(const FrontDesc* __restrict__ desc,
                                                                              const double* __restrict__ vals,
                                                                              const int32_t* __restrict__ src, int64_t slots,
                                                                              const int32_t* __restrict__ work, int64_t nwork,
                                                                              const double* __restrict__ rhs, double* __restrict__ y,
                                                                              double* __restrict__ upd, int64_t ldx, int64_t ldu) {
    const int64_t item = (int64_t)blockIdx.x * MF_SMALL_WARPS + (threadIdx.x >> 5);
    if (item >= nwork) return;
    const int lane = threadIdx.x & 31;
    const FrontDesc d = desc[work[2 * item]];
    const int r0 = work[2 * item + 1];
    const int64_t b = blockIdx.y;
    const double* rb = rhs + b * ldx + d.own_start;
    double* ub = upd + b * ldu;
    const int32_t* src0 = src + d.slot_off;
    const int32_t* src1 = src + slots + d.slot_off;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    const int r = r0 + lane;
    const long long srow = r < n1 ? r : d.row2 + (r - n1);
    const double* rowp = vals + d.mt_off + (r < n ? srow : 0) * d.ldm;
    // columns this item needs: own rows only reach column r0 + 31 (L11^-1 is lower triangular, the rest is stored zeros)
    const int cmax = (r0 + MF_ROWS <= n1) ? r0 + MF_ROWS : n1;
    double acc = 0.0;
    for (int c0 = 0; c0 < cmax; c0 += 32) {
        const int len = min(32, cmax - c0);
        double f = 0.0;
        if (lane < len) f = rb[c0 + lane] + gather_children(src0, src1, ub, c0 + lane);
        // ldm is a multiple of 8 and c0 of 32: 16-byte loads
        const double2* rp2 = reinterpret_cast<const double2*>(rowp + c0);
        const int len2 = len >> 1;
#pragma unroll 4
        for (int j2 = 0; j2 < len2; ++j2) {
            const double2 a = rp2[j2];
            acc = fma(a.x, __shfl_sync(0xffffffffu, f, 2 * j2), acc);
            acc = fma(a.y, __shfl_sync(0xffffffffu, f, 2 * j2 + 1), acc);
        }
        if (len & 1) acc = fma(rowp[c0 + len - 1], __shfl_sync(0xffffffffu, f, len - 1), acc);
    }
    if (r < n) {
        if (r < n1)
            y[b * ldx + d.own_start + r] = acc;
        else
            ub[d.upd_off + (r - n1)] = gather_children(src0, src1, ub, r) + acc;
    }
}

__global__ void __launch_bounds__(32 * MF_SMALL_WARPS) mf_backward_small_kernel(const FrontDesc* __restrict__ desc,
                                                                               const double* __restrict__ vals,
                                                                               const int32_t* __restrict__ bnd,
                                                                               const int32_t* __restrict__ work, int64_t nwork,
                                                                               const double* __restrict__ y, double* __restrict__ x,
                                                                               int64_t ldx) {
    const int64_t item = (int64_t)blockIdx.x * MF_SMALL_WARPS + (threadIdx.x >> 5);
    if (item >= nwork) return;
    const int lane = threadIdx.x & 31;
    const FrontDesc d = desc[work[2 * item]];
    const int c0 = work[2 * item + 1];
    const int64_t b = blockIdx.y;
    const double* yb = y + b * ldx + d.own_start;
    double* xb = x + b * ldx;
    const int32_t* bl = bnd + d.upd_off;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    const int col = c0 + lane;
    const bool col_ok = col < n1;
    const double* mp = vals + d.mt_off + (col_ok ? col : 0);
    double acc = 0.0;
    // own rows r < c0 hold zeros in all of this item's columns (lower triangular): start at c0
    for (int rb0 = c0; rb0 < n; rb0 += 32) {
        const int len = min(32, n - rb0);
        double v = 0.0;
        long long srow = 0;
        if (lane < len) {
            const int r = rb0 + lane;
            v = r < n1 ? yb[r] : xb[bl[r - n1]];
            srow = r < n1 ? r : d.row2 + (r - n1);
        }
#pragma unroll 4
        for (int j = 0; j < len; ++j) {
            const long long sr = __shfl_sync(0xffffffffu, srow, j);
            const double vj = __shfl_sync(0xffffffffu, v, j);
            acc = fma(mp[sr * d.ldm], vj, acc);
        }
    }
    if (col_ok) xb[d.own_start + col] = acc;
}

// ------------------------------------------------------------------------------------------ batched right-hand sides
// A SPLOCS iteration asks for the distance fields of all new component centres at once (tens of right-hand sides).
// The single-RHS kernels above re-stream the 0.94 GB operator once per right-hand side (measured: 4 514 solves = 17 TB,
// 4.9 s of a 21 s build).  These variants keep MF_NB right-hand sides in registers per CTA, so a level's operators are
// read once per group of MF_NB and the pass becomes FP64-FMA bound (2 n1 (n1 + n2) flops per front and RHS).
constexpr int MF_NB = 8;          // right-hand sides per CTA; 4 rows x 8 RHS = 32 accumulators = one per lane after the reduce
constexpr int MF_FCHUNK = 512;    // forward: staged columns per RHS (8 x 512 doubles = 32 KB)
constexpr int MF_BCHUNK = 256;    // backward: staged rows (256 x 8 doubles = 16 KB)

// v[0..31] per lane -> lane l returns sum over the warp of v[l]  (31 shuffles instead of 32 x 5; fixed order)
__device__ __forceinline__ double warp_transpose_sum32(double (&v)[32], int lane) {
#pragma unroll
    for (int h = 16; h >= 1; h >>= 1) {
        const bool upper = (lane & h) != 0;
#pragma unroll
        for (int i = 0; i < h; ++i) {
            const double send = upper ? v[i] : v[i + h];
            const double keep = upper ? v[i + h] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, h);
        }
    }
    return v[0];
}

// grid.x = work item (front, first row), grid.y = group of MF_NB right-hand sides.  256 threads = 8 warps x 4 rows.
__global__ void __launch_bounds__(256) mf_forward_mrhs_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                             const int32_t* __restrict__ src, int64_t slots,
                                                             const int32_t* __restrict__ work, const double* __restrict__ rhs,
                                                             double* __restrict__ y, double* __restrict__ upd, int64_t ldx,
                                                             int64_t ldu, int nrhs) {
    __shared__ double fs[MF_NB][MF_FCHUNK];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    const int r0 = work[2 * blockIdx.x + 1];
    const int64_t b0 = (int64_t)blockIdx.y * MF_NB;
    const int nb = min(MF_NB, nrhs - (int)b0);
    const double* rb = rhs + b0 * ldx + d.own_start;
    double* ub = upd + b0 * ldu;
    const int32_t* src0 = src + d.slot_off;
    const int32_t* src1 = src + slots + d.slot_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    double acc[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) acc[q] = 0.0;
    const double* rowp[4];
    int jmax[4], jall = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = r0 + warp * 4 + i;
        const long long srow = r < n1 ? r : d.row2 + (r - n1);
        rowp[i] = vals + d.mt_off + (r < n ? srow : 0) * d.ldm;
        jmax[i] = r >= n ? 0 : (r < n1 ? r + 1 : n1);   // L11^-1 is lower triangular
        jall = max(jall, jmax[i]);
    }
    const int cmax = (r0 + MF_ROWS <= n1) ? r0 + MF_ROWS : n1;
    for (int c0 = 0; c0 < cmax; c0 += MF_FCHUNK) {
        const int len = min(MF_FCHUNK, cmax - c0);
        __syncthreads();
        for (int j = threadIdx.x; j < len; j += blockDim.x) {
            const int s0 = src0[c0 + j], s1 = src1[c0 + j];
#pragma unroll
            for (int b = 0; b < MF_NB; ++b) {
                double f = 0.0;
                if (b < nb) {
                    f = rb[b * ldx + c0 + j];
                    double g = 0.0;                      // same order as gather_children: (0 + u0) + u1, then rhs + that
                    if (s0 >= 0) g += ub[b * ldu + s0];
                    if (s1 >= 0) g += ub[b * ldu + s1];
                    f += g;
                }
                fs[b][j] = f;
            }
        }
        __syncthreads();
        const int je = min(len, jall - c0);
        for (int j = lane; j < je; j += 32) {
            double a[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = (c0 + j < jmax[i]) ? rowp[i][c0 + j] : 0.0;
#pragma unroll
            for (int b = 0; b < MF_NB; ++b) {
                const double f = fs[b][j];
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[i * MF_NB + b] = fma(a[i], f, acc[i * MF_NB + b]);
            }
        }
    }
    const double s = warp_transpose_sum32(acc, lane);   // lane = i * MF_NB + b
    const int i = lane / MF_NB, b = lane % MF_NB;
    const int r = r0 + warp * 4 + i;
    if (r < n && b < nb) {
        if (r < n1)
            y[(b0 + b) * ldx + d.own_start + r] = s;
        else
            ub[b * ldu + d.upd_off + (r - n1)] = gather_children(src0, src1, ub + b * ldu, r) + s;
    }
}

// DMMA variant of the batched forward pass: Y (rows x 8 RHS) = Mt (rows x columns) F (columns x 8 RHS) per front on the
// FP64 tensor pipe.  The FMA kernel above spends ~85 % of its instructions outside the FMAs on the leaf fronts (n1 <= 64:
// two column steps per lane, then a 32-value transpose-reduce); here a warp owns 8 rows, reads its A fragments straight
// from global memory (lane (lr, lc) -> Mt[row lr][k + lc]: four 8-byte lanes fill one 32-byte sector) and needs no
// cross-lane reduction: D[row lr][rhs 2lc, 2lc+1] is the result.  grid.x = work item (front, first of 32 rows),
// grid.y = group of 8 right-hand sides; 128 threads = 4 warps x 8 rows.
// The staged chunk FCH is a template parameter: 512 columns (33 KB) suit the top of the tree, but on the bottom levels (n1 <= 64,
// tens of thousands of 32 x 32 items per group) that much shared memory left 6 CTAs of 4 warps per SM -- r02 ncu, 40 sources:
// leaf level 2.08 ms for 2.3 GB.  FCH + 4 == 4 mod 16 keeps the B fragments conflict-free (lane -> fs[lr][k + lc]).
template <int FCH>
__global__ void __launch_bounds__(128) mf_forward_dmma_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                             const int32_t* __restrict__ src, int64_t slots,
                                                             const int32_t* __restrict__ work, const double* __restrict__ rhs,
                                                             double* __restrict__ y, double* __restrict__ upd, int64_t ldx,
                                                             int64_t ldu, int nrhs) {
    static_assert(FCH % 16 == 0, "chunks are consumed two k-steps at a time; FCH + 4 must be 4 mod 16");
    __shared__ double fs[MF_NB][FCH + 4];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    const int r0 = work[2 * blockIdx.x + 1];
    const int64_t b0 = (int64_t)blockIdx.y * MF_NB;
    const int nb = min(MF_NB, nrhs - (int)b0);
    const double* rb = rhs + b0 * ldx + d.own_start;
    double* ub = upd + b0 * ldu;
    const int32_t* src0 = src + d.slot_off;
    const int32_t* src1 = src + slots + d.slot_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    // the lane's row of the A fragment
    const int ra = r0 + warp * 8 + lr;
    const long long sra = ra < n1 ? ra : d.row2 + (ra - n1);
    const double* arow = vals + d.mt_off + (ra < n ? sra : 0) * d.ldm;
    const int jmax_a = ra >= n ? 0 : (ra < n1 ? ra + 1 : n1);          // L11^-1 is lower triangular
    // columns this warp needs: its last live row decides (warp-uniform)
    const int rl = min(r0 + warp * 8 + 7, n - 1);
    const int jmax_w = r0 + warp * 8 >= n ? 0 : (rl < n1 ? rl + 1 : n1);
    const int cmax = (r0 + MF_ROWS <= n1) ? r0 + MF_ROWS : n1;          // columns the CTA stages
    double acc0 = 0.0, acc1 = 0.0, odd0 = 0.0, odd1 = 0.0;   // two independent DMMA chains (even / odd k-steps)
    for (int c0 = 0; c0 < cmax; c0 += FCH) {
        const int len = min(FCH, cmax - c0);
        const int len4 = (len + 7) & ~7;
        __syncthreads();
        for (int j = threadIdx.x; j < len4; j += blockDim.x) {
            int s0 = -1, s1 = -1;
            if (j < len) {
                s0 = src0[c0 + j];
                s1 = src1[c0 + j];
            }
#pragma unroll
            for (int b = 0; b < MF_NB; ++b) {
                double f = 0.0;
                if (b < nb && j < len) {
                    f = rb[b * ldx + c0 + j];
                    double g = 0.0;
                    if (s0 >= 0) g += ub[b * ldu + s0];
                    if (s1 >= 0) g += ub[b * ldu + s1];
                    f += g;
                }
                fs[b][j] = f;
            }
        }
        __syncthreads();
        const int ke = min(len4, ((jmax_w - c0) + 7) & ~7);
        const double* bp = &fs[lr][lc];
        for (int k0 = 0; k0 < ke; k0 += 8) {
            const int col = c0 + k0 + lc;
            const double a0 = col < jmax_a ? arow[col] : 0.0;
            const double a1 = col + 4 < jmax_a ? arow[col + 4] : 0.0;
            dmma(acc0, acc1, a0, bp[k0]);
            dmma(odd0, odd1, a1, bp[k0 + 4]);
        }
    }
    acc0 += odd0;
    acc1 += odd1;
    // D fragment: row lr of the warp's 8 rows, right-hand sides 2 lc and 2 lc + 1
    const int r = r0 + warp * 8 + lr;
    if (r < n) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int b = 2 * lc + h;
            if (b >= nb) continue;
            const double s = h ? acc1 : acc0;
            if (r < n1)
                y[(b0 + b) * ldx + d.own_start + r] = s;
            else
                ub[b * ldu + d.upd_off + (r - n1)] = gather_children(src0, src1, ub + b * ldu, r) + s;
        }
    }
}

// grid.x = work item (front, first column), grid.y = group of MF_NB right-hand sides.  256 threads = 32 columns x 8 row groups.
__global__ void __launch_bounds__(256) mf_backward_mrhs_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                              const int32_t* __restrict__ bnd, const int32_t* __restrict__ work,
                                                              const double* __restrict__ y, double* __restrict__ x, int64_t ldx,
                                                              int nrhs) {
    __shared__ __align__(16) double vs[MF_BCHUNK][MF_NB];
    __shared__ double red[8][32][MF_NB + 1];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    const int c0 = work[2 * blockIdx.x + 1];
    const int64_t b0 = (int64_t)blockIdx.y * MF_NB;
    const int nb = min(MF_NB, nrhs - (int)b0);
    const double* yb = y + b0 * ldx + d.own_start;
    double* xb = x + b0 * ldx;
    const int32_t* bl = bnd + d.upd_off;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    const int col = c0 + tx;
    const double* mp = vals + d.mt_off + col;
    double acc[MF_NB];
#pragma unroll
    for (int b = 0; b < MF_NB; ++b) acc[b] = 0.0;
    // own rows r < c0 hold zeros in all of this CTA's columns (lower triangular): start at c0
    for (int rb0 = c0; rb0 < n; rb0 += MF_BCHUNK) {
        const int len = min(MF_BCHUNK, n - rb0);
        __syncthreads();
        for (int e = threadIdx.x; e < len * MF_NB; e += blockDim.x) {
            const int j = e / MF_NB, b = e % MF_NB;
            const int r = rb0 + j;
            double v = 0.0;
            if (b < nb) v = r < n1 ? yb[b * ldx + r] : xb[b * ldx + bl[r - n1]];
            vs[j][b] = v;
        }
        __syncthreads();
        if (col < n1) {
            for (int j = ty; j < len; j += 8) {
                const int r = rb0 + j;
                if (r < n1 && r < col) continue;
                const long long srow = r < n1 ? r : d.row2 + (r - n1);
                const double m = mp[srow * d.ldm];
                const double2* v2 = reinterpret_cast<const double2*>(vs[j]);
#pragma unroll
                for (int b = 0; b < MF_NB / 2; ++b) {
                    const double2 v = v2[b];
                    acc[2 * b] = fma(m, v.x, acc[2 * b]);
                    acc[2 * b + 1] = fma(m, v.y, acc[2 * b + 1]);
                }
            }
        }
    }
#pragma unroll
    for (int b = 0; b < MF_NB; ++b) red[ty][tx][b] = acc[b];
    __syncthreads();
    {   // thread -> (column tx, right-hand side ty): MF_NB == 8 row groups
        const int b = ty;
        if (col < n1 && b < nb) {
            double s = red[0][tx][b];
#pragma unroll
            for (int k = 1; k < 8; ++k) s += red[k][tx][b];
            xb[b * ldx + d.own_start + col] = s;
        }
    }
}

// DMMA variant of the batched backward pass: X (columns x 8 RHS) = Mt^T (columns x rows) V (rows x 8 RHS).  A warp owns 8
// columns; lane (lr, lc) reads Mt[row k + lc][column lr] (eight consecutive doubles per row: two full sectors), the B
// fragment comes from the staged vector block vs[row][rhs] (row stride 12 doubles: conflict-free), D[column lr][rhs 2lc,
// 2lc+1] is the result.  grid.x = work item (front, first of 32 columns), grid.y = group of 8 right-hand sides; 128 threads.
constexpr int MF_VLD = 12;

template <int BCH>
__global__ void __launch_bounds__(128) mf_backward_dmma_kernel(const FrontDesc* __restrict__ desc, const double* __restrict__ vals,
                                                              const int32_t* __restrict__ bnd, const int32_t* __restrict__ work,
                                                              const double* __restrict__ y, double* __restrict__ x, int64_t ldx,
                                                              int nrhs) {
    __shared__ double vs[BCH * MF_VLD];
    const FrontDesc d = desc[work[2 * blockIdx.x]];
    const int c0 = work[2 * blockIdx.x + 1];
    const int64_t b0 = (int64_t)blockIdx.y * MF_NB;
    const int nb = min(MF_NB, nrhs - (int)b0);
    const double* yb = y + b0 * ldx + d.own_start;
    double* xb = x + b0 * ldx;
    const int32_t* bl = bnd + d.upd_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    const int n1 = (int)d.n1, n = (int)(d.n1 + d.n2);
    const int cola = c0 + warp * 8 + lr;                  // the lane's column of the A fragment (and of D)
    const bool col_ok = cola < n1;
    const double* mp = vals + d.mt_off + (col_ok ? cola : 0);
    double acc0 = 0.0, acc1 = 0.0, odd0 = 0.0, odd1 = 0.0;
    // own rows r < c0 hold zeros in all of this CTA's columns (lower triangular): start at c0 (a multiple of 32)
    for (int rb0 = c0; rb0 < n; rb0 += BCH) {
        const int len = min(BCH, n - rb0);
        const int len8 = (len + 7) & ~7;
        __syncthreads();
        for (int e = threadIdx.x; e < len8 * MF_NB; e += blockDim.x) {
            const int j = e / MF_NB, b = e % MF_NB;
            const int r = rb0 + j;
            double v = 0.0;
            if (b < nb && j < len) v = r < n1 ? yb[b * ldx + r] : xb[b * ldx + bl[r - n1]];
            vs[j * MF_VLD + b] = v;
        }
        __syncthreads();
        const double* bp = vs + lc * MF_VLD + lr;
        for (int k0 = 0; k0 < len8; k0 += 8) {
            const int ra = rb0 + k0 + lc, rb = ra + 4;
            const long long sa = ra < n1 ? ra : d.row2 + (ra - n1);
            const long long sb = rb < n1 ? rb : d.row2 + (rb - n1);
            const double a0 = (col_ok && ra < n) ? mp[sa * d.ldm] : 0.0;
            const double a1 = (col_ok && rb < n) ? mp[sb * d.ldm] : 0.0;
            dmma(acc0, acc1, a0, bp[k0 * MF_VLD]);
            dmma(odd0, odd1, a1, bp[(k0 + 4) * MF_VLD]);
        }
    }
    acc0 += odd0;
    acc1 += odd1;
    if (col_ok) {
        if (2 * lc < nb) xb[(2 * lc) * ldx + d.own_start + cola] = acc0;
        if (2 * lc + 1 < nb) xb[(2 * lc + 1) * ldx + d.own_start + cola] = acc1;
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

// ASB_MF_TALL=0 keeps the 256-thread kernels on every level (A/B switch for profiles/)
static bool mf_tall() {
    static const int v = [] {
        const char* e = std::getenv("ASB_MF_TALL");
        return e ? std::atoi(e) : 1;
    }();
    return v != 0;
}

// max_n1 / max_n: the largest n1 / n1 + n2 of the fronts in the work list (0 = unknown): they pick the staged chunk of the
// batched kernels and send the single-RHS passes of levels made of many small fronts to the warp-per-item kernels
static int mf_forward_impl(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* work,
                           int64_t nwork, const double* rhs, double* y, double* upd, int64_t nrhs, int64_t ldx, int64_t ldu,
                           int64_t max_n1, cudaStream_t st) {
    static const int fwd_dmma = [] {
        const char* e = std::getenv("ASB_MF_FWD_DMMA");
        return e ? std::atoi(e) : 1;
    }();
    static const int small = [] {
        const char* e = std::getenv("ASB_MF_SMALL");
        return e ? std::atoi(e) : 1;
    }();
    const FrontDesc* fd = reinterpret_cast<const FrontDesc*>(desc);
    const dim3 gridg((unsigned)nwork, (unsigned)cdiv(nrhs, MF_NB));
    if (nrhs >= MF_MRHS_MIN && fwd_dmma) {
        if (small && max_n1 > 0 && max_n1 <= 64)
            mf_forward_dmma_kernel<64><<<gridg, 128, 0, st>>>(fd, vals, src, slots, work, rhs, y, upd, ldx, ldu, (int)nrhs);
        else if (small && max_n1 > 0 && max_n1 <= 128)
            mf_forward_dmma_kernel<128><<<gridg, 128, 0, st>>>(fd, vals, src, slots, work, rhs, y, upd, ldx, ldu, (int)nrhs);
        else
            mf_forward_dmma_kernel<MF_FCHUNK><<<gridg, 128, 0, st>>>(fd, vals, src, slots, work, rhs, y, upd, ldx, ldu, (int)nrhs);
    } else if (nrhs >= MF_MRHS_MIN)
        mf_forward_mrhs_kernel<<<gridg, 256, 0, st>>>(fd, vals, src, slots, work, rhs, y, upd, ldx, ldu, (int)nrhs);
    else if (nwork * nrhs <= 2LL * sm_count() && mf_tall())
        mf_forward_kernel<1><<<dim3((unsigned)nwork, (unsigned)nrhs), 1024, 0, st>>>(fd, vals, src, slots, work, rhs, y, upd, ldx, ldu);
    else if (small && max_n1 > 0 && max_n1 <= 64 && nwork >= 16LL * sm_count())
        mf_forward_small_kernel<<<dim3((unsigned)cdiv(nwork, MF_SMALL_WARPS), (unsigned)nrhs), 32 * MF_SMALL_WARPS, 0, st>>>(
            fd, vals, src, slots, work, nwork, rhs, y, upd, ldx, ldu);
    else
        mf_forward_kernel<4><<<dim3((unsigned)nwork, (unsigned)nrhs), 256, 0, st>>>(fd, vals, src, slots, work, rhs, y, upd, ldx, ldu);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

static int mf_backward_impl(const int64_t* desc, const double* vals, const int32_t* bnd, const int32_t* work, int64_t nwork,
                            const double* y, double* x, int64_t nrhs, int64_t ldx, int64_t max_n, cudaStream_t st) {
    static const int bwd_dmma = [] {
        const char* e = std::getenv("ASB_MF_BWD_DMMA");
        return e ? std::atoi(e) : 1;
    }();
    static const int small = [] {
        const char* e = std::getenv("ASB_MF_SMALL");
        return e ? std::atoi(e) : 1;
    }();
    const FrontDesc* fd = reinterpret_cast<const FrontDesc*>(desc);
    const dim3 gridg((unsigned)nwork, (unsigned)cdiv(nrhs, MF_NB));
    if (nrhs >= MF_MRHS_MIN && bwd_dmma) {
        if (small && max_n > 0 && max_n <= 64)
            mf_backward_dmma_kernel<64><<<gridg, 128, 0, st>>>(fd, vals, bnd, work, y, x, ldx, (int)nrhs);
        else if (small && max_n > 0 && max_n <= 128)
            mf_backward_dmma_kernel<128><<<gridg, 128, 0, st>>>(fd, vals, bnd, work, y, x, ldx, (int)nrhs);
        else
            mf_backward_dmma_kernel<MF_BCHUNK><<<gridg, 128, 0, st>>>(fd, vals, bnd, work, y, x, ldx, (int)nrhs);
    } else if (nrhs >= MF_MRHS_MIN)
        mf_backward_mrhs_kernel<<<gridg, 256, 0, st>>>(fd, vals, bnd, work, y, x, ldx, (int)nrhs);
    else if (nwork * nrhs <= 2LL * sm_count() && mf_tall())
        mf_backward_kernel<32><<<dim3((unsigned)nwork, (unsigned)nrhs), 1024, 0, st>>>(fd, vals, bnd, work, y, x, ldx);
    else if (small && max_n > 0 && max_n <= 256 && nwork >= 16LL * sm_count())
        mf_backward_small_kernel<<<dim3((unsigned)cdiv(nwork, MF_SMALL_WARPS), (unsigned)nrhs), 32 * MF_SMALL_WARPS, 0, st>>>(
            fd, vals, bnd, work, nwork, y, x, ldx);
    else
        mf_backward_kernel<8><<<dim3((unsigned)nwork, (unsigned)nrhs), 256, 0, st>>>(fd, vals, bnd, work, y, x, ldx);
    ASB_LAUNCH_CHECK();
    return ASB_OK;
}

extern "C" {

int asb_mf_forward(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* work,
                   int64_t nwork, const double* rhs, double* y, double* upd, int64_t nrhs, int64_t ldx, int64_t ldu,
                   asb_stream_t stream) {
    ASB_REQUIRE(desc && vals && src && work && rhs && y && upd && nwork > 0 && nrhs > 0 && nrhs < 65536, "mf_forward: bad arguments");
    return mf_forward_impl(desc, vals, src, slots, work, nwork, rhs, y, upd, nrhs, ldx, ldu, 0, S(stream));
}

int asb_mf_backward(const int64_t* desc, const double* vals, const int32_t* bnd, const int32_t* work, int64_t nwork,
                    const double* y, double* x, int64_t nrhs, int64_t ldx, asb_stream_t stream) {
    ASB_REQUIRE(desc && vals && bnd && work && y && x && nwork > 0 && nrhs > 0 && nrhs < 65536, "mf_backward: bad arguments");
    return mf_backward_impl(desc, vals, bnd, work, nwork, y, x, nrhs, ldx, 0, S(stream));
}

int asb_mf_solve(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* bnd,
                 const int64_t* levels, int64_t nlevels, const double* rhs, double* y, double* x, double* upd, int64_t nrhs,
                 int64_t ldx, int64_t ldu, asb_stream_t stream) {
    ASB_REQUIRE(levels && nlevels > 0 && nlevels < 4096, "mf_solve: bad level table");
    ASB_REQUIRE(desc && vals && src && bnd && rhs && y && x && upd && nrhs > 0 && nrhs < 65536, "mf_solve: bad arguments");
    for (int64_t l = nlevels - 1; l >= 0; --l) {
        const int64_t* lv = levels + MF_LEVEL_COLS * l;
        if (lv[1] <= 0) continue;
        const int rc = mf_forward_impl(desc, vals, src, slots, reinterpret_cast<const int32_t*>(lv[0]), lv[1], rhs, y, upd, nrhs, ldx,
                                       ldu, lv[4], S(stream));
        if (rc != ASB_OK) return rc;
    }
    for (int64_t l = 0; l < nlevels; ++l) {
        const int64_t* lv = levels + MF_LEVEL_COLS * l;
        if (lv[3] <= 0) continue;
        const int rc = mf_backward_impl(desc, vals, bnd, reinterpret_cast<const int32_t*>(lv[2]), lv[3], y, x, nrhs, ldx, lv[5],
                                        S(stream));
        if (rc != ASB_OK) return rc;
    }
    return ASB_OK;
}

int asb_mf_solve_delta(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* bnd,
                       const int64_t* levels, int64_t nlevels, const int64_t* maxblk, const int64_t* fstart,
                       const int32_t* parent, const int32_t* depth, int64_t nfronts, const int32_t* order, const int64_t* idx,
                       int32_t* pathfront, double* rp, double* y, double* x, double* upd, int64_t nrhs, int64_t ldx,
                       int64_t ldu, asb_stream_t stream) {
    ASB_REQUIRE(desc && vals && src && bnd && levels && maxblk && fstart && parent && depth && order && idx && pathfront && rp &&
                    y && x && upd && nlevels > 0 && nlevels < 4096 && nrhs > 0 && nrhs < 65536 && nfronts > 0,
                "mf_solve_delta: bad arguments");
    const cudaStream_t st = S(stream);
    ASB_CUDA(cudaMemsetAsync(rp, 0, sizeof(double) * nrhs * ldx, st));
    ASB_CUDA(cudaMemsetAsync(y, 0, sizeof(double) * nrhs * ldx, st));
    ASB_CUDA(cudaMemsetAsync(upd, 0, sizeof(double) * nrhs * ldu, st));
    ASB_CUDA(cudaMemsetAsync(pathfront, 0xFF, sizeof(int32_t) * nrhs * nlevels, st));
    mf_path_kernel<<<(unsigned)cdiv(nrhs, 64), 64, 0, st>>>(idx, order, fstart, parent, depth, nfronts, (int)nlevels, (int)nrhs,
                                                           pathfront, rp, ldx);
    ASB_LAUNCH_CHECK();
    for (int64_t l = nlevels - 1; l >= 0; --l) {
        mf_forward_path_kernel<<<dim3((unsigned)maxblk[l], (unsigned)nrhs), 1024, 0, st>>>(
            reinterpret_cast<const FrontDesc*>(desc), vals, src, slots, pathfront, (int)l, (int)nlevels, rp, y, upd, ldx, ldu);
        ASB_LAUNCH_CHECK();
    }
    for (int64_t l = 0; l < nlevels; ++l) {
        const int64_t* lv = levels + MF_LEVEL_COLS * l;
        if (lv[3] <= 0) continue;
        const int rc = mf_backward_impl(desc, vals, bnd, reinterpret_cast<const int32_t*>(lv[2]), lv[3], y, x, nrhs, ldx, lv[5], st);
        if (rc != ASB_OK) return rc;
    }
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
