// Development probe: the K5 main loop alone (cp.async ring + LDS + DMMA), tile shape / warp
// layout / stage count as template parameters, no MIS epilogue.  Not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gemm_probe gemm_probe.cu && ./gemm_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp16(void* smem, const void* g) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(g));
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void waitg() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// NOLOAD: skip the global loads (smem holds garbage) to isolate the cost of the copy engine.
template <int TM, int TN, int BK, int STAGES, int WM, int WN, bool NOLOAD, bool NOSYNC, bool REMAP>
__global__ void __launch_bounds__(WM * WN * 32, 1)
gemm(const double* __restrict__ A, const double* __restrict__ B, int tiles_m, int tiles_n, int kdim,
     double* out) {
    constexpr int THREADS = WM * WN * 32;
    constexpr int MI = TM / WM / 8, NJ = TN / WN / 8;
    constexpr int OPA = BK * TM, OPB = BK * TN, STAGE = OPA + OPB;
    constexpr int CHUNKS = (TM + TN) * BK / 2;          // 16-byte pieces per stage
    static_assert(CHUNKS % THREADS == 0, "loader shape");
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp / WN, wn = warp % WN;
    const int n_tiles = tiles_m * tiles_n, kchunks = kdim / BK;
    int p_tile = blockIdx.x, p_kc = 0, p_stage = 0;
    auto issue = [&]() {
        if (p_tile < n_tiles) {
            if (!NOLOAD) {
                const int tr = p_tile % tiles_m, tc = p_tile / tiles_m;
                double* st = smem + p_stage * STAGE;
#pragma unroll
                for (int j = 0; j < CHUNKS / THREADS; ++j) {
                    const int q = tid + j * THREADS;
                    const bool isB = q >= TM * BK / 2;
                    const int qq = isB ? q - TM * BK / 2 : q;
                    int row, c16;
                    if (REMAP) {        // quarter-warp = 4 rows x 32 B of one k group: 128 B of distinct banks
                        static_assert(!REMAP || BK == 16, "remap assumes BK 16");
                        const int l = qq & 31, blk = qq >> 5;
                        row = blk * 4 + ((l >> 1) & 3);
                        c16 = (l & 1) + 2 * (l >> 3);
                    } else { row = qq / (BK / 2); c16 = qq % (BK / 2); }
                    const int kg = c16 >> 1, within = (c16 & 1) * 2;
                    double* dst = st + (isB ? OPA : 0) + (kg * (isB ? TN : TM) + row) * 4 + within;
                    const long long grow = (long long)(isB ? tc * TN : tr * TM) + row;
                    cp16(dst, (isB ? B : A) + grow * kdim + p_kc * BK + c16 * 2);
                }
            }
            if (++p_kc == kchunks) { p_kc = 0; p_tile += gridDim.x; }
        }
        p_stage = p_stage + 1 == STAGES ? 0 : p_stage + 1;
        commit();
    };
    for (int s = 0; s < STAGES - 1; ++s) issue();
    double acc[MI][NJ][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    int c_stage = 0;
    double total = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int kc = 0; kc < kchunks; ++kc) {
            waitg<STAGES - 2>();
            if (!NOSYNC) __syncthreads();
            issue();
            const double* sA = smem + c_stage * STAGE;
            const double* sB = sA + OPA;
            c_stage = c_stage + 1 == STAGES ? 0 : c_stage + 1;
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
        }
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) { total += acc[i][j][0] + acc[i][j][1]; acc[i][j][0] = acc[i][j][1] = 0.0; }
    }
    waitg<0>();
    out[blockIdx.x * THREADS + tid] = total;
}

template <int TM, int TN, int BK, int STAGES, int WM, int WN, bool NOLOAD = false, bool NOSYNC = false, bool REMAP = false>
void run(const double* A, const double* B, int rows, int cols, int kdim, double* out) {
    auto k = gemm<TM, TN, BK, STAGES, WM, WN, NOLOAD, NOSYNC, REMAP>;
    const int smem = STAGES * (TM + TN) * BK * 8;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int tm = rows / TM, tn = cols / TN;
    k<<<148, WM * WN * 32, smem>>>(A, B, tm, tn, kdim, out);
    cudaEventRecord(a);
    k<<<148, WM * WN * 32, smem>>>(A, B, tm, tn, kdim, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    printf("tile %3dx%3d BK %2d stages %d warps %dx%d (%3d thr, smem %3d KB) %s: %6.2f TFLOP/s %s\n", TM, TN, BK,
           STAGES, WM, WN, WM * WN * 32, smem / 1024, NOLOAD ? (NOSYNC ? "NOLOAD+NOSYNC" : "NOLOAD") : (REMAP ? "REMAP" : "      "),
           2.0 * rows * cols * kdim / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int rows = 1 << 15, cols = 1 << 16, kdim = 256;
    double *A, *B, *out;
    cudaMalloc(&A, (size_t)rows * kdim * 8);
    cudaMalloc(&B, (size_t)cols * kdim * 8);
    cudaMalloc(&out, 148 * 1024 * 8);
    cudaMemset(A, 0, (size_t)rows * kdim * 8);
    cudaMemset(B, 0, (size_t)cols * kdim * 8);
    run<128, 128, 16, 4, 4, 4>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 4, 4, true, false>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 4, 4, true, true>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 2, 4, true, false>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 2, 4, true, true>(A, B, rows, cols, kdim, out);
    run<128, 128, 32, 3, 4, 4>(A, B, rows, cols, kdim, out);
    run<128, 128, 32, 3, 2, 4>(A, B, rows, cols, kdim, out);
    run<256, 128, 16, 3, 4, 4>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 4, 4, false, false, true>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 4, 2, 4, false, false, true>(A, B, rows, cols, kdim, out);
    run<128, 128, 16, 6, 4, 4, false, false, true>(A, B, rows, cols, kdim, out);
    run<128, 64, 16, 4, 4, 2, false, false, true>(A, B, rows, cols, kdim, out);
    return 0;
}
