// Development probe: issue rate of tcgen05.mma (cta_group::1, M = 128) from shared memory on every
// SM at once, for kind::i8 / kind::f8f6f4 / kind::f16, N in {64, 128, 256}, operands in the
// no-swizzle core-matrix layout or the 128-byte-swizzle layout.  Values are irrelevant (smem is
// zero); only cycles per instruction are reported.  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_rate_probe umma_rate_probe.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46) | ((uint64_t)layout << 61);
}

template <int KIND>
__device__ __forceinline__ void mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    if (KIND == 0)
        asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                     ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else if (KIND == 1)
        asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n}\n"
                     ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                     ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

// swz: 0 = no swizzle (8 x 16-byte core matrices, LBO 128, SBO 2048), 1 = 128-byte swizzle (SBO 1024)
template <int KIND>
__global__ void __launch_bounds__(128) probe(int N, int swz, int iters, int n_acc, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t dyn[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    uint8_t* sA = dyn;                 // 128 rows x 256 bytes
    uint8_t* sB = dyn + 128 * 256;     // N rows x 256 bytes
    for (int i = tid; i < (128 + N) * 256 / 16; i += 128) reinterpret_cast<int4*>(dyn)[i] = make_int4(0, 0, 0, 0);
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
    if (tid == 0) {
        uint32_t idesc;
        if (KIND == 0) idesc = (2u << 4) | (1u << 7) | (1u << 10);        // D s32, A/B s8
        else if (KIND == 1) idesc = (1u << 4);                             // D f32, A/B e4m3
        else idesc = (1u << 4) | (1u << 7) | (1u << 10);                   // D f32, A/B bf16
        idesc |= ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem + (uint32_t)((it % n_acc) * N);
            for (int ks = 0; ks < 8; ++ks) {
                uint64_t da, db;
                if (swz) {
                    const uint32_t offA = (ks >> 2) * 128 * 128 + (ks & 3) * 32, offB = (ks >> 2) * N * 128 + (ks & 3) * 32;
                    da = make_desc(smem_u32(sA) + offA, 16, 1024, 2);
                    db = make_desc(smem_u32(sB) + offB, 16, 1024, 2);
                } else {
                    da = make_desc(smem_u32(sA) + ks * 256, 128, 2048, 0);
                    db = make_desc(smem_u32(sB) + ks * 256, 128, 2048, 0);
                }
                mma<KIND>(d, da, db, idesc, (it >= n_acc || ks > 0) ? 1u : 0u);
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                     " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
        cycles[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

template <int KIND>
void run(const char* name, int N, int swz, int n_acc, int grid) {
    const int iters = 2048;
    long long* d_cyc;
    cudaMalloc(&d_cyc, grid * sizeof(long long));
    const int smem = (128 + N) * 256;
    cudaFuncSetAttribute(probe<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<KIND><<<grid, 128, smem>>>(N, swz, 64, n_acc, d_cyc);           // warm
    cudaEventRecord(e0);
    probe<KIND><<<grid, 128, smem>>>(N, swz, iters, n_acc, d_cyc);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    std::vector<long long> c(grid);
    cudaMemcpy(c.data(), d_cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    std::sort(c.begin(), c.end());
    const double per = (double)c[grid / 2] / (iters * 8.0);
    const double ops = 2.0 * 128 * N * (KIND == 2 ? 16 : 32) * 8.0 * iters * grid;
    printf("%-8s N=%3d %s acc=%d grid=%3d: %s  %7.1f clk/mma (median CTA), %8.3f ms, %7.1f Top/s\n", name, N,
           swz ? "swz128" : "noswz ", n_acc, grid, cudaGetErrorString(e), per, ms, ops / ms / 1e9);
    cudaFree(d_cyc);
}

int main() {
    for (int grid : {1, 148}) {
        for (int swz = 0; swz < 2; ++swz)
            for (int N : {64, 80, 96, 112, 128, 256}) {
                run<0>("i8", N, swz, 1, grid);
                run<1>("f8f6f4", N, swz, 1, grid);
                run<2>("f16", N, swz, 1, grid);
            }
        run<0>("i8", 64, 0, 6, grid);
        run<0>("i8", 64, 1, 6, grid);
    }
    return 0;
}
