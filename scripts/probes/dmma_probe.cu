// Microbenchmark: what limits DMMA.8x8x4 throughput on sm_100a?  (development probe, not product)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_probe dmma_probe.cu && ./dmma_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// MODE 0: registers only.  1: LDS.64 fragments each kk.  2: + __syncthreads every 4 kk.
template <int MODE, int MI, int NJ>
__global__ void probe(double* out, int iters) {
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = 1e-3 * i;
    __syncthreads();
    double acc[MI][NJ][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    double fa[MI], fb[NJ];
#pragma unroll
    for (int i = 0; i < MI; ++i) fa[i] = sm[lane + 32 * i];
#pragma unroll
    for (int j = 0; j < NJ; ++j) fb[j] = sm[lane + 32 * j + 512];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            if (MODE >= 1) {
                const double* s = sm + ((it * 4 + kk) & 3) * 1024 + (warp & 3) * 32;
#pragma unroll
                for (int i = 0; i < MI; ++i) fa[i] = s[i * 32 + lane];
#pragma unroll
                for (int j = 0; j < NJ; ++j) fb[j] = s[512 + j * 32 + lane];
            }
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) dmma(acc[i][j][0], acc[i][j][1], fa[i], fb[j]);
        }
        if (MODE >= 2) __syncthreads();
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) s += acc[i][j][0] + acc[i][j][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE, int MI, int NJ>
void run(int warps, int ctas_per_sm, double* out) {
    const int iters = 20000 / (MI * NJ / 4);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaFuncSetAttribute(probe<MODE, MI, NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    const int grid = 148 * ctas_per_sm;
    probe<MODE, MI, NJ><<<grid, warps * 32, 40000>>>(out, iters);
    cudaEventRecord(a);
    probe<MODE, MI, NJ><<<grid, warps * 32, 40000>>>(out, iters);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    const double flops = 2.0 * 256 * MI * NJ * 4.0 * iters * warps * grid;
    printf("mode %d  warp tile %dx%d  warps/CTA %2d  CTAs/SM %d : %7.2f TFLOP/s  (%s)\n", MODE, MI * 8,
           NJ * 8, warps, ctas_per_sm, flops / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    double* out;
    cudaMalloc(&out, 148 * 4 * 1024 * sizeof(double));
    for (int w : {4, 8, 16, 32}) run<0, 4, 4>(w, 1, out);
    for (int w : {4, 8, 16}) run<0, 4, 4>(w, 2, out);
    for (int w : {4, 8, 16}) run<0, 2, 4>(w, 1, out);
    for (int w : {4, 8}) run<0, 8, 4>(w, 1, out);
    for (int w : {4, 8, 16, 32}) run<1, 4, 4>(w, 1, out);
    for (int w : {4, 8}) run<1, 8, 4>(w, 1, out);
    for (int w : {4, 8, 16, 32}) run<2, 4, 4>(w, 1, out);
    for (int w : {4, 8}) run<2, 8, 4>(w, 1, out);
    for (int w : {8}) run<2, 4, 4>(w, 2, out);
    return 0;
}
