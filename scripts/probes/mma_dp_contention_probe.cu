// Development probe: does scalar FP64 (DFMA) throughput drop while the tensor pipe is busy with
// tcgen05.mma kind::i8?  One CTA per SM: warp 0 issues M=128 N=64 K=32 int8 MMAs back to back (or
// idles), 16 more warps run 8 independent DFMA (or IMAD / FFMA) chains each.  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_dp_contention_probe mma_dp_contention_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}

template <int OP>
__global__ void __launch_bounds__(544) probe(int mma_iters, int alu_iters, double* out, long long* cyc) {
    extern __shared__ __align__(1024) uint8_t dyn[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 + 64) * 256 / 16; i += blockDim.x) reinterpret_cast<int4*>(dyn)[i] = make_int4(0x01020304, 0x05060708, 0x090a0b0c, 0x0d0e0f10);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    if (warp == 0) {
        if (tid == 0 && mma_iters > 0) {
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            const uint64_t da0 = make_desc(smem_u32(dyn), 128, 2048), db0 = make_desc(smem_u32(dyn + 128 * 256), 128, 2048);
            const long long t0 = clock64();
            for (int it = 0; it < mma_iters; ++it) {
                const uint32_t d = tmem + (uint32_t)((it % 6) * 64);
#pragma unroll
                for (int ks = 0; ks < 8; ++ks)
                    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                                 ::"r"(d), "l"(da0 + ks * 16), "l"(db0 + ks * 16), "r"(idesc), "r"((it >= 6 || ks > 0) ? 1u : 0u) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
            asm volatile("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                         " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
            cyc[blockIdx.x * 2] = clock64() - t0;
        }
    } else {
        double d[8]; float f[8]; int w[8];
        for (int i = 0; i < 8; ++i) { d[i] = tid + i; f[i] = tid + i; w[i] = tid + i; }
        const long long t0 = clock64();
        for (int it = 0; it < alu_iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (OP == 0) d[i] = fma(d[i], 1.0000001, 1e-9);
                if (OP == 1) w[i] = w[i] * 3 + it;
                if (OP == 2) f[i] = fmaf(f[i], 1.0000001f, 1e-9f);
            }
        }
        const long long t1 = clock64();
        double s = 0;
        for (int i = 0; i < 8; ++i) s += d[i] + f[i] + w[i];
        out[blockIdx.x * blockDim.x + tid] = s;
        if (tid == 32) cyc[blockIdx.x * 2 + 1] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

template <int OP>
void run(const char* name, int mma_iters, int alu_iters) {
    const int grid = 148;
    double* out; long long* cyc;
    cudaMalloc(&out, grid * 544 * 8); cudaMalloc(&cyc, grid * 16);
    cudaMemset(cyc, 0, grid * 16);
    const int smem = (128 + 64) * 256;
    cudaFuncSetAttribute(probe<OP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe<OP><<<grid, 544, smem>>>(mma_iters ? 64 : 0, alu_iters ? 64 : 0, out, cyc);
    probe<OP><<<grid, 544, smem>>>(mma_iters, alu_iters, out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[296];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-6s mma_iters %6d alu_iters %6d: %s", name, mma_iters, alu_iters, cudaGetErrorString(e));
    if (mma_iters) printf("  %6.1f clk/MMA", (double)h[0] / (mma_iters * 8.0));
    if (alu_iters) printf("  %6.1f thread-ops/clk/SM (16 warps)", 512.0 * 8 * alu_iters / (double)h[1]);
    printf("\n");
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("DFMA", 0, 8192);  run<0>("DFMA", 4096, 0);  run<0>("DFMA", 4096, 8192 * 4);
    run<1>("IMAD", 0, 8192);  run<1>("IMAD", 4096, 8192 * 8);
    run<2>("FFMA", 0, 8192);  run<2>("FFMA", 4096, 8192 * 8);
    return 0;
}
