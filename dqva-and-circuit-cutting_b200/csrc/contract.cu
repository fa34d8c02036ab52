// K5: two-operand contraction over the cut-basis index, FP64 on the DMMA pipe, with the MIS
// epilogue (K8) fused so the 2^n-entry distribution never has to exist in HBM.
//
//   P~[r, c] = scale * sum_k A[r, k] * Bt[c, k]
//
// Replaces the per-bitstring tensor-network contraction of _recombine_using_networks
// (qsplit/qsplit_mlrecon_methods.py:412-461) and, for two subcircuits, the 4^K rank-1 updates
// of cutqc/build.c:91-169.
//
// Bound: FP64 tensor pipe.  Algorithmic FLOPs per output = 2 * kdim.  Operands are tiny next to
// the output (2 * 2^17 * 256 * 8 B = 0.5 GB for 2^34 outputs) and stay L2-resident per column
// panel, so HBM only matters when d_P_out is requested (8 or 4 bytes per output, written once).
//
// Tiling: persistent CTAs (one per SM), 128x128 output tile, 16 warps as 4x4, warp tile 32x32 =
// 4x4 mma.m8n8k4 accumulators (64 registers).  Operands stream through a 4-stage cp.async ring of
// 16-wide k slices stored as [k/4][row][4] so that every fragment load is 256 contiguous bytes
// (bank-conflict free).  The ring runs across tile boundaries, so the register-only epilogue of
// tile t overlaps the loads of tile t+1.
#include "qdc_common.cuh"

namespace {

constexpr int TM = 128, TN = 128, BK = 16, STAGES = 4, THREADS = 512;
constexpr int OPERAND_DOUBLES = (BK / 4) * TM * 4;          // 2048 doubles = 16 KB
constexpr int STAGE_DOUBLES = 2 * OPERAND_DOUBLES;           // A then B
constexpr int SMEM_BYTES = STAGES * STAGE_DOUBLES * 8;       // 128 KB
constexpr int PANEL_W = 16;                                  // tile columns per L2 panel

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

__device__ __forceinline__ void dmma_8x8x4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

struct TileCoord { long long r0, c0; };

__device__ __forceinline__ TileCoord tile_coord(long long t, long long tiles_m, long long tiles_n) {
    // column panels of PANEL_W tiles; inside a panel walk rows, then the panel's columns
    const long long per_panel = tiles_m * PANEL_W;
    const long long panel = t / per_panel;
    const long long rem = t - panel * per_panel;
    const long long pw = (tiles_n - panel * PANEL_W) < PANEL_W ? (tiles_n - panel * PANEL_W) : PANEL_W;
    const long long r = rem / pw, c = panel * PANEL_W + rem % pw;
    return {r * TM, c * TN};
}

struct ContractArgs {
    const double* A;        // rows row_lo.. (already offset), [n_rows][kdim]
    const double* Bt;       // [n_cols][kdim]
    long long row_lo, n_rows, n_cols;
    int kdim;
    double scale;
    const double* d_norm;
    MisTables tab;
    CandSink sink;
    void* P_out;
    double* partials;
};

// STORE: 0 none, 1 united order, 2 circuit-order scatter.  OUT32: write float instead of double.
template <int STORE, bool OUT32>
__global__ void __launch_bounds__(THREADS, 1) contract_f64_kernel(ContractArgs a) {
    extern __shared__ __align__(16) double smem[];
    __shared__ double scratch[3 * (THREADS / 32)];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const long long tiles_m = (a.n_rows + TM - 1) / TM, tiles_n = (a.n_cols + TN - 1) / TN;
    // tiles_n is padded to whole panels only logically: the last panel may be narrower
    const long long n_tiles = tiles_m * tiles_n;
    const int kchunks = a.kdim / BK;
    const long long my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long total = my_tiles * kchunks;
    const double out_scale = a.scale * (a.d_norm ? 1.0 / *a.d_norm : 1.0);

    // ---- producer: 4 x 16-byte cp.async per thread per stage --------------------------------
    auto issue = [&](long long g) {
        if (g < total) {
            const long long ti = g / kchunks;
            const int kc = (int)(g - ti * kchunks);
            const TileCoord tc = tile_coord(blockIdx.x + ti * gridDim.x, tiles_m, tiles_n);
            double* st = smem + (g % STAGES) * STAGE_DOUBLES;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int q = tid + j * THREADS;             // 0..2047
                const int op = q >> 10;                      // 0 = A, 1 = B
                const int qq = q & 1023;
                const int row = qq >> 3, c16 = qq & 7;       // 8 x 16 B per row slice
                const int kg = c16 >> 1, within = (c16 & 1) * 2;
                double* dst = st + op * OPERAND_DOUBLES + (kg * TM + row) * 4 + within;
                const long long grow = (op == 0 ? tc.r0 : tc.c0) + row;
                const long long lim = op == 0 ? a.n_rows : a.n_cols;
                const double* base = op == 0 ? a.A : a.Bt;
                const bool ok = grow < lim;
                const double* src = base + (ok ? grow : 0) * (long long)a.kdim + kc * BK + c16 * 2;
                cp_async16(dst, src, ok ? 16 : 0);
            }
        }
        cp_async_commit();
    };

#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) issue(s);

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    MisAccum macc = {0.0, 0.0, 0.0};

    for (long long g = 0; g < total; ++g) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        issue(g + STAGES - 1);

        const double* sA = smem + (g % STAGES) * STAGE_DOUBLES;
        const double* sB = sA + OPERAND_DOUBLES;
#pragma unroll
        for (int kk = 0; kk < BK / 4; ++kk) {
            double fa[4], fb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) fa[i] = sA[(kk * TM + wm * 32 + i * 8) * 4 + lane];
#pragma unroll
            for (int j = 0; j < 4; ++j) fb[j] = sB[(kk * TN + wn * 32 + j * 8) * 4 + lane];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma_8x8x4(acc[i][j][0], acc[i][j][1], fa[i], fb[j]);
        }

        const long long ti = g / kchunks;
        if (g - ti * kchunks == kchunks - 1) {
            // ---- epilogue of tile ti, registers only --------------------------------------
            const TileCoord tc = tile_coord(blockIdx.x + ti * gridDim.x, tiles_m, tiles_n);
            const long long cbase = tc.c0 + wn * 32 + 2 * (lane & 3);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const long long r = tc.r0 + wm * 32 + i * 8 + (lane >> 2);
                const bool rok = r < a.n_rows;
                const long long rg = a.row_lo + (rok ? r : 0);
                const uint64_t rx = __ldg(a.tab.row_x + rg), rf = __ldg(a.tab.row_forb + rg);
                const uint32_t rw = __ldg(a.tab.row_wok + rg);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const long long c = cbase + j * 8;
                    const double p0 = acc[i][j][0] * out_scale, p1 = acc[i][j][1] * out_scale;
                    acc[i][j][0] = acc[i][j][1] = 0.0;
                    if (rok && c < a.n_cols) {        // n_cols even or 1: c+1 valid unless n_cols==1
                        const bool two = c + 1 < a.n_cols;
                        const uint64_t cx0 = __ldg(a.tab.col_x + c);
                        const uint32_t cw0 = __ldg(a.tab.col_wok + c);
                        mis_accumulate(macc, p0, rx, rf, rw, cx0, cw0, a.sink);
                        uint64_t cx1 = 0;
                        if (two) {
                            cx1 = __ldg(a.tab.col_x + c + 1);
                            mis_accumulate(macc, p1, rx, rf, rw, cx1, __ldg(a.tab.col_wok + c + 1), a.sink);
                        }
                        if (STORE == 1) {
                            const long long o = r * a.n_cols + c;
                            if (OUT32) {
                                float* out = (float*)a.P_out;
                                if (two) *reinterpret_cast<float2*>(out + o) = make_float2((float)p0, (float)p1);
                                else out[o] = (float)p0;
                            } else {
                                double* out = (double*)a.P_out;
                                if (two) *reinterpret_cast<double2*>(out + o) = make_double2(p0, p1);
                                else out[o] = p0;
                            }
                        } else if (STORE == 2) {
                            if (OUT32) {
                                float* out = (float*)a.P_out;
                                out[rx | cx0] = (float)p0;
                                if (two) out[rx | cx1] = (float)p1;
                            } else {
                                double* out = (double*)a.P_out;
                                out[rx | cx0] = p0;
                                if (two) out[rx | cx1] = p1;
                            }
                        }
                    }
                }
            }
        }
    }
    cp_async_wait<0>();
    mis_block_store(macc, scratch, a.partials);
}

// ---- norm: column sums --------------------------------------------------------------------
constexpr int NORM_BLOCKS = 64;

__global__ void __launch_bounds__(256) colsum_kernel(const double* __restrict__ X, long long n_rows,
                                                     int kdim, double* __restrict__ out /*[NORM_BLOCKS][kdim]*/) {
    // thread t owns column (t % kdim) when kdim <= 256, else loops over columns
    const long long rows_per_block = (n_rows + gridDim.x - 1) / gridDim.x;
    const long long r_begin = blockIdx.x * rows_per_block;
    const long long r_end = r_begin + rows_per_block < n_rows ? r_begin + rows_per_block : n_rows;
    extern __shared__ double sh[];                                  // [256]
    for (int k0 = 0; k0 < kdim; k0 += 256) {
        const int kw = kdim - k0 < 256 ? kdim - k0 : 256;           // columns this pass
        const int lanes_per_col = 256 / kw;                          // >= 1 (kw divides 256 or kw == 256)
        const int k = threadIdx.x % kw, sub = threadIdx.x / kw;
        double s = 0.0;
        if (sub < lanes_per_col)
            for (long long r = r_begin + sub; r < r_end; r += lanes_per_col) s += X[r * kdim + k0 + k];
        sh[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x < kw) {
            double t = 0.0;
            for (int j = 0; j < lanes_per_col; ++j) t += sh[j * kw + threadIdx.x];
            out[(long long)blockIdx.x * kdim + k0 + threadIdx.x] = t;
        }
        __syncthreads();
    }
}

__global__ void norm_finalize_kernel(const double* __restrict__ pa, const double* __restrict__ pb,
                                     int kdim, double scale, double* __restrict__ norm) {
    __shared__ double sh[8];
    double s = 0.0;
    for (int k = threadIdx.x; k < kdim; k += blockDim.x) {
        double ca = 0.0, cb = 0.0;
        for (int b = 0; b < NORM_BLOCKS; ++b) { ca += pa[(long long)b * kdim + k]; cb += pb[(long long)b * kdim + k]; }
        s += ca * cb;
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        *norm = t * scale;
    }
}

}  // namespace

extern "C" int qdc_norm_workspace_doubles(int kdim) { return 2 * NORM_BLOCKS * kdim; }

extern "C" int qdc_contract_norm(const double* d_A, int64_t n_rows_total, const double* d_Bt,
                                 int64_t n_cols, int kdim, double scale, double* d_colsums,
                                 double* d_norm, void* stream) {
    QDC_REQUIRE(d_A && d_Bt && d_colsums && d_norm, "null buffer");
    QDC_REQUIRE(kdim >= 1 && (kdim <= 256 ? 256 % kdim == 0 : kdim % 256 == 0), "kdim");
    cudaStream_t s = (cudaStream_t)stream;
    double* pa = d_colsums;
    double* pb = d_colsums + (long long)NORM_BLOCKS * kdim;
    colsum_kernel<<<NORM_BLOCKS, 256, 256 * sizeof(double), s>>>(d_A, n_rows_total, kdim, pa);
    QDC_CHECK_LAUNCH();
    colsum_kernel<<<NORM_BLOCKS, 256, 256 * sizeof(double), s>>>(d_Bt, n_cols, kdim, pb);
    QDC_CHECK_LAUNCH();
    norm_finalize_kernel<<<1, 256, 0, s>>>(pa, pb, kdim, scale, d_norm);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

template <int STORE, bool OUT32>
static int launch_contract(const ContractArgs& args, int grid, cudaStream_t s) {
    auto kern = contract_f64_kernel<STORE, OUT32>;
    QDC_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    kern<<<grid, THREADS, SMEM_BYTES, s>>>(args);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_contract_f64(const double* d_A, const double* d_Bt, int64_t row_lo,
                                int64_t n_rows, int64_t n_cols, int kdim, double scale,
                                const double* d_norm, double threshold,
                                const uint64_t* d_row_x, const uint64_t* d_row_forb,
                                const uint32_t* d_row_wok, const uint64_t* d_col_x,
                                const uint32_t* d_col_wok, void* d_P_out, int store_mode,
                                int out_dtype, double* d_sums, uint64_t* d_cand_x,
                                double* d_cand_p, unsigned long long* d_cand_count,
                                int64_t cand_cap, double* d_partials, void* stream) {
    QDC_REQUIRE(d_A && d_Bt && d_sums && d_partials, "null buffer");
    QDC_REQUIRE(d_row_x && d_row_forb && d_row_wok && d_col_x && d_col_wok, "null tables");
    QDC_REQUIRE(d_cand_x && d_cand_p && d_cand_count && cand_cap >= 0, "candidate sink");
    QDC_REQUIRE(kdim >= BK && kdim % BK == 0, "kdim must be a multiple of 16");
    QDC_REQUIRE(n_rows >= 0 && n_cols >= 1 && row_lo >= 0, "shape");
    QDC_REQUIRE(n_cols == 1 || n_cols % 2 == 0, "n_cols must be 1 or even");
    QDC_REQUIRE((d_P_out == nullptr) == (store_mode == 0), "store_mode / d_P_out mismatch");
    QDC_REQUIRE(store_mode >= 0 && store_mode <= 2 && (out_dtype == 0 || out_dtype == 1), "modes");
    if (n_rows == 0) return QDC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    ContractArgs args;
    args.A = d_A + row_lo * (long long)kdim;
    args.Bt = d_Bt;
    args.row_lo = row_lo; args.n_rows = n_rows; args.n_cols = n_cols; args.kdim = kdim;
    args.scale = scale; args.d_norm = d_norm;
    args.tab = {d_row_x, d_row_forb, d_row_wok, d_col_x, d_col_wok};
    args.sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
    args.P_out = d_P_out; args.partials = d_partials;
    const long long n_tiles = ((n_rows + TM - 1) / TM) * ((n_cols + TN - 1) / TN);
    const int sms = qdc_sm_count();
    const int grid = (int)(n_tiles < sms ? n_tiles : sms);
    int rc;
    if (store_mode == 0) rc = launch_contract<0, false>(args, grid, s);
    else if (store_mode == 1) rc = out_dtype ? launch_contract<1, true>(args, grid, s)
                                              : launch_contract<1, false>(args, grid, s);
    else rc = out_dtype ? launch_contract<2, true>(args, grid, s)
                        : launch_contract<2, false>(args, grid, s);
    if (rc != QDC_OK) return rc;
    return qdc_launch_finalize_sums(d_partials, grid, d_sums, s);
}
