// Development probe: per-SM issue rates of the scalar instructions the int8 contraction's epilogue
// is made of (FP64 FMA/ADD, 64-bit integer -> FP64 conversions, IMAD.WIDE, FFMA), 8 independent
// chains per thread, 16 warps per SM.  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o alu_rate_probe alu_rate_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(512) k(int iters, double* out, long long* cyc) {
    double d[8]; float f[8]; long long l[8]; int w[8];
    for (int i = 0; i < 8; ++i) { d[i] = threadIdx.x + i; f[i] = threadIdx.x + i; l[i] = threadIdx.x * 977 + i; w[i] = threadIdx.x + i; }
    const double m = 1.0000001, a = 1e-9;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) d[i] = fma(d[i], m, a);
            if (OP == 1) d[i] = d[i] + a;
            if (OP == 2) { d[i] = (double)l[i]; l[i] += w[i]; }                       // I2F.F64.S64 + IADD
            if (OP == 3) { d[i] = (double)w[i]; w[i] += it; }                         // I2F.F64.S32 + IADD
            if (OP == 4) l[i] = (long long)w[i] * 256 + l[i];                          // IMAD.WIDE
            if (OP == 5) f[i] = fmaf(f[i], 1.0000001f, 1e-9f);
            if (OP == 6) { l[i] = __double_as_longlong(d[i]) > l[i] ? l[i] + 1 : l[i] - 1; } // 64-bit int compare/select
            if (OP == 7) d[i] = d[i] * m;
        }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int i = 0; i < 8; ++i) s += d[i] + f[i] + (double)l[i] + w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name) {
    const int grid = 148, iters = 4096;
    double* out; long long* cyc;
    cudaMalloc(&out, grid * 512 * 8); cudaMalloc(&cyc, grid * 8);
    k<OP><<<grid, 512>>>(64, out, cyc);
    k<OP><<<grid, 512>>>(iters, out, cyc);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-28s %6.1f thread-ops / clk / SM\n", name, 512.0 * 8 * iters / (double)h[0]);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("DFMA"); run<1>("DADD"); run<7>("DMUL"); run<2>("I2F.F64.S64 (+IADD)"); run<3>("I2F.F64.S32 (+IADD)");
    run<4>("IMAD.WIDE"); run<5>("FFMA"); run<6>("64-bit compare+select");
    return 0;
}
