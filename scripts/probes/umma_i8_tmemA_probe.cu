// Development probe: tcgen05.mma kind::i8 with the A operand in TENSOR MEMORY (copied there from the
// shared-memory core-matrix image by tcgen05.cp.128x256b), B from shared memory; checks the result
// against a CPU dot product and times the instruction (M = 128, N = 64, K = 32).  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_i8_tmemA_probe umma_i8_tmemA_probe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 64, K = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__host__ __device__ inline int core_offset(int row, int k) {
    return ((row >> 3) * (K / 16) + (k >> 4)) * 128 + (row & 7) * 16 + (k & 15);
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}

// mode 0: correctness (one pass, results written); mode 1: timing of `iters` x 8 MMAs with A in TMEM;
// mode 2: the same with A from shared memory
template <int mode>
__global__ void __launch_bounds__(128) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                             int32_t* __restrict__ D, int iters, long long* cyc) {
    extern __shared__ __align__(1024) int8_t dyn[];
    int8_t* sA = dyn;
    int8_t* sB = dyn + M * K;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < M * K / 16; i += 128) reinterpret_cast<int4*>(sA)[i] = reinterpret_cast<const int4*>(A)[i];
    for (int i = tid; i < N * K / 16; i += 128) reinterpret_cast<int4*>(sB)[i] = reinterpret_cast<const int4*>(B)[i];
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
    const uint32_t tmem_a = tmem + 384;                 // 64 columns: 128 rows x 256 bytes of A
    if (tid == 0) {
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint64_t da0 = make_desc(smem_u32(sA), 128, 2048), db0 = make_desc(smem_u32(sB), 128, 2048);
        if (mode != 2)
            for (int ks = 0; ks < K / 32; ++ks)
                asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tmem_a + ks * 8), "l"(da0 + ks * 16) : "memory");
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem + (uint32_t)((it % 6) * 64);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const uint32_t acc = (it >= 6 || ks > 0) ? 1u : 0u;
                if (mode == 2)
                    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                                 ::"r"(d), "l"(da0 + ks * 16), "l"(db0 + ks * 16), "r"(idesc), "r"(acc) : "memory");
                else
                    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n}\n"
                                 ::"r"(d), "r"(tmem_a + ks * 8), "l"(db0 + ks * 16), "r"(idesc), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                     " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
        if (cyc) cyc[blockIdx.x] = clock64() - t0;
    }
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (mode == 0)
        for (int c0 = 0; c0 < N; c0 += 16) {
            uint32_t r[16];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int j = 0; j < 16; ++j) D[(warp * 32 + lane) * N + c0 + j] = (int32_t)r[j];
        }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main() {
    std::vector<int8_t> a(M * K), b(N * K), ap(M * K), bp(N * K);
    srand(1);
    for (auto& v : a) v = (int8_t)(rand() % 255 - 127);
    for (auto& v : b) v = (int8_t)(rand() % 255 - 127);
    for (int r = 0; r < M; ++r) for (int k = 0; k < K; ++k) ap[core_offset(r, k)] = a[r * K + k];
    for (int r = 0; r < N; ++r) for (int k = 0; k < K; ++k) bp[core_offset(r, k)] = b[r * K + k];
    std::vector<int32_t> want(M * N);
    for (int r = 0; r < M; ++r) for (int c = 0; c < N; ++c) {
        int32_t s = 0;
        for (int k = 0; k < K; ++k) s += (int32_t)a[r * K + k] * (int32_t)b[c * K + k];
        want[r * N + c] = s;
    }
    int8_t *dA, *dB; int32_t* dD; long long* dC;
    cudaMalloc(&dA, M * K); cudaMalloc(&dB, N * K); cudaMalloc(&dD, M * N * 4); cudaMalloc(&dC, 148 * 8);
    cudaMemcpy(dA, ap.data(), M * K, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, bp.data(), N * K, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (M + N) * K);
    cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (M + N) * K);
    cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (M + N) * K);
    cudaMemset(dD, 0xff, M * N * 4);
    probe<0><<<1, 128, (M + N) * K>>>(dA, dB, dD, 1, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<int32_t> got(M * N);
    cudaMemcpy(got.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int i = 0; i < M * N; ++i) bad += got[i] != want[i];
    printf("A in TMEM via tcgen05.cp.128x256b: %s, mismatches %ld / %d (got %d %d %d, want %d %d %d)\n", cudaGetErrorString(e),
           bad, M * N, got[0], got[1], got[N], want[0], want[1], want[N]);
    for (int mode = 1; mode <= 2; ++mode) {
        const int iters = 2048;
        if (mode == 1) probe<1><<<148, 128, (M + N) * K>>>(dA, dB, dD, iters, dC);
        else probe<2><<<148, 128, (M + N) * K>>>(dA, dB, dD, iters, dC);
        e = cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, dC, sizeof(h), cudaMemcpyDeviceToHost);
        printf("%s: %s, %.1f clk per MMA (M=128 N=64 K=32, 148 CTAs)\n", mode == 1 ? "A from TMEM" : "A from smem", cudaGetErrorString(e),
               (double)h[0] / (iters * 8.0));
    }
    return 0;
}
