// Development probe: one tcgen05.mma kind::i8 tile (M=128, N=64, K=256) from shared memory in the
// canonical K-major no-swizzle layout (8x16-byte core matrices), accumulator in TMEM, read back with
// tcgen05.ld and compared with a CPU dot product.  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_i8_probe umma_i8_probe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 64, K = 256;          // K in bytes (int8 elements)
constexpr int KSTEP = 32;                         // bytes of K per MMA instruction

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// element (row, k) of a [rows x K] K-major operand in core-matrix order:
// [row / 8][k / 16][row % 8][k % 16]
__host__ __device__ inline int core_offset(int row, int k) {
    return ((row >> 3) * (K / 16) + (k >> 4)) * 128 + (row & 7) * 16 + (k & 15);
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;                       // descriptor version (sm_100)
    return d;                                     // layout_type 0 = no swizzle, base_offset 0
}

__global__ void __launch_bounds__(128) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                             int32_t* __restrict__ D, uint32_t lbo, uint32_t sbo) {
    extern __shared__ __align__(128) int8_t dyn[];
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
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    // make the generic-proxy smem writes visible to the async (tensor core) proxy
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        // instruction descriptor: D = S32 (2 << 4), A = B = signed int8 (1 << 7, 1 << 10), K-major both,
        // N >> 3 at bit 17, M >> 4 at bit 24
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        for (int ks = 0; ks < K / KSTEP; ++ks) {
            const uint64_t da = make_desc(smem_u32(sA) + ks * 256, lbo, sbo);
            const uint64_t db = make_desc(smem_u32(sB) + ks * 256, lbo, sbo);
            const uint32_t accumulate = ks > 0;
            asm volatile(
                "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                " tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    // everybody waits for the MMAs
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
        " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // warp w reads TMEM lanes 32w .. 32w+31, 64 columns, 16 at a time
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
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem));
}

int main() {
    std::vector<int8_t> a(M * K), b(N * K), ap(M * K), bp(N * K);
    srand(1);
    for (auto& v : a) v = (int8_t)(rand() % 129 - 64);
    for (auto& v : b) v = (int8_t)(rand() % 129 - 64);
    for (int r = 0; r < M; ++r) for (int k = 0; k < K; ++k) ap[core_offset(r, k)] = a[r * K + k];
    for (int r = 0; r < N; ++r) for (int k = 0; k < K; ++k) bp[core_offset(r, k)] = b[r * K + k];
    std::vector<int32_t> want(M * N);
    for (int r = 0; r < M; ++r) for (int c = 0; c < N; ++c) {
        int32_t s = 0;
        for (int k = 0; k < K; ++k) s += (int32_t)a[r * K + k] * (int32_t)b[c * K + k];
        want[r * N + c] = s;
    }
    int8_t *dA, *dB; int32_t* dD;
    cudaMalloc(&dA, M * K); cudaMalloc(&dB, N * K); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, ap.data(), M * K, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, bp.data(), N * K, cudaMemcpyHostToDevice);
    const uint32_t cfg[2][2] = {{128, 2048}, {2048, 128}};     // (LBO, SBO) candidates
    for (int t = 0; t < 2; ++t) {
        cudaMemset(dD, 0xff, M * N * 4);
        cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (M + N) * K);
        probe<<<1, 128, (M + N) * K>>>(dA, dB, dD, cfg[t][0], cfg[t][1]);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<int32_t> got(M * N);
        cudaMemcpy(got.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
        long bad = 0;
        for (int i = 0; i < M * N; ++i) bad += got[i] != want[i];
        printf("LBO %u SBO %u: %s, mismatches %ld / %d  (got[0..3] %d %d %d %d, want %d %d %d %d)\n", cfg[t][0], cfg[t][1],
               cudaGetErrorString(e), bad, M * N, got[0], got[1], got[2], got[3], want[0], want[1], want[2], want[3]);
    }
    return 0;
}
