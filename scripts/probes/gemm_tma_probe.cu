// Development probe: K5 main loop fed by 1-D bulk async copies (cp.async.bulk + mbarrier) issued
// by a dedicated producer warp, operands pre-packed in global memory as [k/4][row][4] so that a
// k-group slice of a tile is one contiguous TM*32-byte run.  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gemm_tma_probe gemm_tma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int TM, int TN, int BK, int STAGES, int WM, int WN>
__global__ void __launch_bounds__(WM * WN * 32 + 32, 1)
gemm(const double* __restrict__ Ap, const double* __restrict__ Bp, long long rowsA, long long rowsB,
     int tiles_m, int tiles_n, int kdim, double* out) {
    constexpr int CONSUMERS = WM * WN * 32;
    constexpr int MI = TM / WM / 8, NJ = TN / WN / 8;
    constexpr int OPA = BK * TM, OPB = BK * TN, STAGE = OPA + OPB;
    extern __shared__ __align__(128) double smem[];
    __shared__ __align__(8) uint64_t full[STAGES], empty[STAGES];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_tiles = tiles_m * tiles_n, kchunks = kdim / BK;
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], WM * WN); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == WM * WN) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const long long r0 = (long long)(tile % tiles_m) * TM, c0 = (long long)(tile / tiles_m) * TN;
                for (int kc = 0; kc < kchunks; ++kc) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_expect_tx(&full[stage], STAGE * 8);
                    double* st = smem + stage * STAGE;
#pragma unroll
                    for (int kg = 0; kg < BK / 4; ++kg) {
                        const long long g = (long long)kc * (BK / 4) + kg;
                        bulk_g2s(st + kg * TM * 4, Ap + (g * rowsA + r0) * 4, TM * 32, &full[stage]);
                        bulk_g2s(st + OPA + kg * TN * 4, Bp + (g * rowsB + c0) * 4, TN * 32, &full[stage]);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        return;
    }
    const int wm = warp / WN, wn = warp % WN;
    double acc[MI][NJ][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    int stage = 0; uint32_t phase = 0;
    double total = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int kc = 0; kc < kchunks; ++kc) {
            mbar_wait(&full[stage], phase);
            const double* sA = smem + stage * STAGE;
            const double* sB = sA + OPA;
#pragma unroll
            for (int kk = 0; kk < BK / 4; ++kk) {
                double fa[MI], fb[NJ];
#pragma unroll
                for (int i = 0; i < MI; ++i) fa[i] = sA[(kk * TM + wm * (TM / WM) + i * 8) * 4 + lane];
#pragma unroll
                for (int j = 0; j < NJ; ++j) fb[j] = sB[(kk * TN + wn * (TN / WN) + j * 8) * 4 + lane];
#pragma unroll
                for (int i = 0; i < MI; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) dmma(acc[i][j][0], acc[i][j][1], fa[i], fb[j]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) { total += acc[i][j][0] + acc[i][j][1]; acc[i][j][0] = acc[i][j][1] = 0.0; }
    }
    out[blockIdx.x * CONSUMERS + tid] = total;
}

__global__ void fill(double* p, size_t n, double v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

template <int TM, int TN, int BK, int STAGES, int WM, int WN>
void run(const double* A, const double* B, int rows, int cols, int kdim, double* out) {
    auto k = gemm<TM, TN, BK, STAGES, WM, WN>;
    const int smem = STAGES * (TM + TN) * BK * 8;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int tm = rows / TM, tn = cols / TN, thr = WM * WN * 32 + 32;
    k<<<148, thr, smem>>>(A, B, rows, cols, tm, tn, kdim, out);
    cudaEventRecord(a);
    k<<<148, thr, smem>>>(A, B, rows, cols, tm, tn, kdim, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    double h = 0;
    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    printf("BULK tile %3dx%3d BK %2d stages %d warps %dx%d (+1 producer, smem %3d KB): %6.2f TFLOP/s %s  (check %.3f vs %.3f)\n",
           TM, TN, BK, STAGES, WM, WN, smem / 1024, 2.0 * rows * cols * kdim / ms / 1e9,
           cudaGetErrorString(cudaGetLastError()), h, 0.0);
}

int main() {
    const int rows = 1 << 15, cols = 1 << 16, kdim = 256;
    double *A, *B, *out;
    cudaMalloc(&A, (size_t)rows * kdim * 8);
    cudaMalloc(&B, (size_t)cols * kdim * 8);
    cudaMalloc(&out, 148 * 1024 * 8);
    fill<<<1024, 256>>>(A, (size_t)rows * kdim, 1.0);
    fill<<<1024, 256>>>(B, (size_t)cols * kdim, 0.5);
    // every output = kdim * 0.5 = 128; thread 0 of CTA 0 sums 32 accumulators per tile over its tiles
    run<128, 128, 16, 4, 4, 4>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 2, 4>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 6, 2, 4>(A, B, rows, cols, kdim, out);
    run<128, 128, 32, 3, 2, 4>(A, B, rows, cols, kdim, out);
    run<128, 256, 16, 4, 2, 4>(A, B, rows, cols, kdim, out);
    run<256, 128, 16, 4, 4, 2>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 4, 2>(A, B, rows, cols, kdim, out);
    return 0;
}
