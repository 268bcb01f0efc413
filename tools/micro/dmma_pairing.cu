// Micro-benchmark: DMMA.8x8x4 issue rate as a function of how many n-tiles share one A fragment (r01 question: the ADMM
// kernel has 3 n-tiles per m-tile and sits at ~25 cycles per DMMA and sub-partition, the GEMMs with 4 n-tiles at ~19).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_pairing dmma_pairing.cu && ./dmma_pairing
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int MT, int NT, bool J_OUTER, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) k(double* out, int iters, double seed) {
    double acc[MT][NT][2], a[MT], b[NT];
#pragma unroll
    for (int i = 0; i < MT; ++i) {
        a[i] = seed + i + threadIdx.x * 1e-3;
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    }
#pragma unroll
    for (int j = 0; j < NT; ++j) b[j] = seed * 0.5 + j;
    for (int it = 0; it < iters; ++it) {
        if (J_OUTER) {
#pragma unroll
            for (int j = 0; j < NT; ++j)
#pragma unroll
                for (int i = 0; i < MT; ++i) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        } else {
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
                for (int j = 0; j < NT; ++j) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) s += acc[i][j][0] + acc[i][j][1];
    if (s == 12345.678) out[0] = s;
}

template <int MT, int NT, bool J_OUTER, int MAXT = 512>
void run(int warps, int sms, double mhz) {
    double* out;
    cudaMalloc(&out, 8);
    const int iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MT, NT, J_OUTER, MAXT><<<sms, warps * 32>>>(out, 100, 1.0);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MT, NT, J_OUTER, MAXT><<<sms, warps * 32>>>(out, iters, 1.0);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double dmma_per_subpart = (double)iters * MT * NT * warps / 4.0;
    const double cycles = ms * 1e-3 * mhz * 1e6;
    const double tflops = (double)iters * MT * NT * warps * sms * 512.0 / (ms * 1e-3) / 1e12;
    printf("MT=%d NT=%d order=%s warps/SM=%2d : %.2f cycles per DMMA and sub-partition, %.1f TFLOP/s%s\n", MT, NT,
           J_OUTER ? "j-outer(B shared)" : "i-outer(A shared)", warps, cycles / dmma_per_subpart, tflops,
           cudaGetLastError() == cudaSuccess ? "" : "  [CUDA error]");
    cudaFree(out);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    const double mhz = p.clockRate / 1000.0;
    printf("%s, %d SMs, %.0f MHz (nominal)\n", p.name, sms, mhz);
    for (int warps : {4, 8, 12, 16}) {
        run<4, 2, false>(warps, sms, mhz);
        run<4, 3, false>(warps, sms, mhz);
        run<4, 4, false>(warps, sms, mhz);
        run<4, 6, false>(warps, sms, mhz);
        run<4, 3, true>(warps, sms, mhz);
        run<3, 4, true>(warps, sms, mhz);
    }
    run<7, 3, false, 256>(8, sms, mhz);
    run<7, 4, false, 256>(8, sms, mhz);
    run<7, 4, false, 384>(12, sms, mhz);
    run<7, 3, false, 384>(12, sms, mhz);
    run<6, 3, false, 256>(8, sms, mhz);
    return 0;
}
