// Shared helpers for libasb (sm_100a).  See include/asb.h for the C-ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "asb.h"

namespace asb {

extern thread_local char g_err[512];
extern std::atomic<long long> g_launches;

inline int fail(int code, const char* what, const char* file, int line) {
    snprintf(g_err, sizeof(g_err), "%s (%s:%d)", what, file, line);
    return code;
}

#define ASB_REQUIRE(cond, msg)                                                   \
    do {                                                                         \
        if (!(cond)) return ::asb::fail(ASB_ERR_ARG, msg, __FILE__, __LINE__);   \
    } while (0)

#define ASB_CUDA(expr)                                                                        \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) return ::asb::fail(ASB_ERR_CUDA, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// every kernel launch goes through this: counts launches and surfaces launch-configuration errors
#define ASB_LAUNCH_CHECK()                                                                    \
    do {                                                                                      \
        ::asb::g_launches.fetch_add(1, std::memory_order_relaxed);                            \
        cudaError_t _e = cudaPeekAtLastError();                                               \
        if (_e != cudaSuccess) {                                                              \
            cudaGetLastError();                                                               \
            return ::asb::fail(ASB_ERR_CUDA, cudaGetErrorString(_e), __FILE__, __LINE__);     \
        }                                                                                     \
    } while (0)

inline cudaStream_t S(asb_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

int sm_count();
// out[j] = sum_b partial[b*stride + j] (fixed order, one block)
int sum_partials(const double* partial, int64_t nb, int stride, double* out, cudaStream_t st);

// ------------------------------------------------------------------------------------------ device helpers
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum (all threads get the result). blockDim.x multiple of 32, <= 1024. `sh` >= 33 doubles.
__device__ __forceinline__ double block_sum(double v, double* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        double t = lane < nw ? sh[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) sh[32] = t;
    }
    __syncthreads();
    return sh[32];
}
__device__ __forceinline__ double block_max(double v, double* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        double t = lane < nw ? sh[lane] : -INFINITY;
        t = warp_max(t);
        if (lane == 0) sh[32] = t;
    }
    __syncthreads();
    return sh[32];
}

// (value, first index) argmax combine: larger value wins, ties go to the lower index (np.argmax).
struct ArgMax {
    double v;
    long long i;
};
__device__ __forceinline__ ArgMax argmax_combine(ArgMax a, ArgMax b) {
    if (b.v > a.v || (b.v == a.v && b.i < a.i)) return b;
    return a;
}
__device__ __forceinline__ ArgMax warp_argmax(ArgMax a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ArgMax b;
        b.v = __shfl_xor_sync(0xffffffffu, a.v, o);
        b.i = __shfl_xor_sync(0xffffffffu, a.i, o);
        a = argmax_combine(a, b);
    }
    return a;
}
// `shv` >= 33 doubles, `shi` >= 33 long longs. Result valid in all threads.
__device__ __forceinline__ ArgMax block_argmax(ArgMax a, double* shv, long long* shi) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    a = warp_argmax(a);
    __syncthreads();
    if (lane == 0) { shv[w] = a.v; shi[w] = a.i; }
    __syncthreads();
    if (w == 0) {
        ArgMax t;
        t.v = lane < nw ? shv[lane] : -INFINITY;
        t.i = lane < nw ? shi[lane] : 0x7fffffffffffffffLL;
        t = warp_argmax(t);
        if (lane == 0) { shv[32] = t.v; shi[32] = t.i; }
    }
    __syncthreads();
    ArgMax r;
    r.v = shv[32];
    r.i = shi[32];
    return r;
}

__device__ __forceinline__ double2 ld2(const double* p) { return *reinterpret_cast<const double2*>(p); }
__device__ __forceinline__ void st2(double* p, double2 v) { *reinterpret_cast<double2*>(p) = v; }
// streaming (read-once) 16-byte load: do not pollute L1
__device__ __forceinline__ double2 ld2_stream(const double* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

// 3x3 symmetric eigen-decomposition by cyclic Jacobi (one thread).  a: in (row-major 3x3, symmetric),
// out: eigenvalues in ev[3], eigenvectors as COLUMNS of V (row-major 3x3).
__device__ inline void jacobi_eig3(const double* a_in, double* ev, double* V) {
    double a[3][3], v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) a[i][j] = a_in[i * 3 + j];
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
        double diag = fabs(a[0][0]) + fabs(a[1][1]) + fabs(a[2][2]);
        if (off <= 1e-300 || off <= 1e-22 * diag) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (a[p][q] == 0.0) continue;
                double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
                double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 3; ++k) {  // A <- A J
                    double akp = a[k][p], akq = a[k][q];
                    a[k][p] = c * akp - s * akq;
                    a[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 3; ++k) {  // A <- J^T A
                    double apk = a[p][k], aqk = a[q][k];
                    a[p][k] = c * apk - s * aqk;
                    a[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 3; ++k) {
                    double vkp = v[k][p], vkq = v[k][q];
                    v[k][p] = c * vkp - s * vkq;
                    v[k][q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < 3; ++i) {
        ev[i] = a[i][i];
        for (int j = 0; j < 3; ++j) V[i * 3 + j] = v[i][j];
    }
}

}  // namespace asb
