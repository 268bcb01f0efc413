// Micro-benchmark 2: which ingredient of the ADMM slab loop costs DMMA issue rate?  Same instruction mix as
// admm_solve_kernel<Pair, 7> (7 m-tiles x 3 n-tiles per warp, 4 warps per CTA, 2 CTAs per SM), added one at a time:
//   mode 0  fragments in registers (peak)            mode 1  + A/B fragments from shared memory every k-step
//   mode 2  + __syncthreads every 2 k-steps (slab)   mode 3  + cp.async ring (7 x 16 B per thread and slab, L2-resident source)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_feed dmma_feed.cu && ./dmma_feed
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}

constexpr int ROWS = 200, ALD = 12, DLD = 28, STAGES = 3;

template <int MT, int NT, int MODE, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) k(double* out, const double* __restrict__ Hr, int slabs) {
    extern __shared__ __align__(16) double smem[];
    double* Ds = smem;                       // [ROWS][DLD]
    double* As = smem + ROWS * DLD;          // STAGES x [ROWS][ALD]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    for (int e = threadIdx.x; e < ROWS * DLD + STAGES * ROWS * ALD; e += blockDim.x) smem[e] = 1e-3 * (e % 97);
    __syncthreads();
    double acc[MT][NT][2], a[MT], b[NT];
#pragma unroll
    for (int i = 0; i < MT; ++i) {
        a[i] = 1.0 + i;
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    }
#pragma unroll
    for (int j = 0; j < NT; ++j) b[j] = 0.5 + j;
    const int mt0 = (warp % 4) * 6;
    const double* afrag = As + (mt0 * 8 + lr) * ALD + lc;
    const double* bfrag = Ds + lc * DLD + lr;
    const int srow = threadIdx.x / 4, scc = (threadIdx.x % 4) * 2;
    const int rpp = blockDim.x / 4;
    int cs = 0, ls = STAGES - 1;
    for (int s = 0; s < slabs; ++s) {
        if (MODE >= 3) asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 2));
        if (MODE >= 2) __syncthreads();
        if (MODE >= 3) {
            const double* q = Hr + (size_t)srow * 208 + (s % 25) * 8 + scc;
            double* dst = As + ls * ROWS * ALD + srow * ALD + scc;
            for (int r = srow, t = 0; r < ROWS; r += rpp, ++t) cp_async16(dst + t * rpp * ALD, q + (size_t)t * rpp * 208);
            asm volatile("cp.async.commit_group;");
            ls = ls + 1 == STAGES ? 0 : ls + 1;
        }
        const double* as = afrag + cs * ROWS * ALD;
        const double* ds = bfrag + (s % 25) * 8 * DLD;
        cs = cs + 1 == STAGES ? 0 : cs + 1;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
            if (MODE >= 1) {
#pragma unroll
                for (int j = 0; j < NT; ++j) b[j] = ds[kk * 4 * DLD + j * 8];
            }
#pragma unroll
            for (int i = 0; i < MT; ++i) {
                if (MODE >= 1) a[i] = as[i * 8 * ALD + kk * 4];
#pragma unroll
                for (int j = 0; j < NT; ++j) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            }
        }
    }
    double sum = 0.0;
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) sum += acc[i][j][0] + acc[i][j][1];
    if (sum == 12345.678) out[0] = sum;
}

template <int MT, int NT, int MODE, int WARPS>
void run(int ctas_per_sm, int sms, double mhz, const double* Hr, double* out) {
    const int slabs = 20000;
    const size_t smem = (size_t)(ROWS * DLD + STAGES * ROWS * ALD) * 8 + (ctas_per_sm == 1 ? 60 * 1024 : 0);
    cudaFuncSetAttribute(k<MT, NT, MODE, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MT, NT, MODE, WARPS><<<sms * ctas_per_sm, WARPS * 32, smem>>>(out, Hr, 200);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MT, NT, MODE, WARPS><<<sms * ctas_per_sm, WARPS * 32, smem>>>(out, Hr, slabs);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double per_sp = (double)slabs * 2 * MT * NT * WARPS * ctas_per_sm / 4.0;
    printf("MT=%d NT=%d warps/CTA=%2d CTAs/SM=%d mode=%d : %.2f cycles per DMMA and sub-partition  (%s)\n", MT, NT, WARPS,
           ctas_per_sm, MODE, ms * 1e-3 * mhz * 1e6 / per_sp, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    const double mhz = p.clockRate / 1000.0;
    double *Hr, *out;
    cudaMalloc(&Hr, 208 * 208 * 8 * 2);
    cudaMemset(Hr, 0, 208 * 208 * 8 * 2);
    cudaMalloc(&out, 8);
    printf("%s, %d SMs\n", p.name, sms);
    run<7, 3, 0, 4>(2, sms, mhz, Hr, out);
    run<7, 3, 1, 4>(2, sms, mhz, Hr, out);
    run<7, 3, 2, 4>(2, sms, mhz, Hr, out);
    run<7, 3, 3, 4>(2, sms, mhz, Hr, out);
    run<7, 3, 1, 4>(1, sms, mhz, Hr, out);
    run<7, 3, 3, 4>(1, sms, mhz, Hr, out);
    run<7, 4, 1, 4>(2, sms, mhz, Hr, out);
    run<7, 4, 3, 4>(2, sms, mhz, Hr, out);
    run<4, 3, 1, 16>(1, sms, mhz, Hr, out);
    run<4, 3, 2, 16>(1, sms, mhz, Hr, out);
    run<4, 3, 3, 16>(1, sms, mhz, Hr, out);
    run<4, 6, 3, 8>(1, sms, mhz, Hr, out);
    return 0;
}
