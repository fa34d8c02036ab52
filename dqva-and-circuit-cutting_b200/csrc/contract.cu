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
// 4x4 mma.m8n8k4 accumulators (64 registers).  Operands are pre-packed in global memory as
// [k/4][row][4] (qdc_contract_pack), so the k-group slice of a tile is one contiguous 4 KB run:
// one elected lane (the duty rotates over the warps, one k slice each) moves a 16-wide k slice of
// both operands with 8 bulk async copies (cp.async.bulk, the 1-D TMA path) into a 6-stage
// shared-memory ring and arms a per-stage "full" mbarrier (complete_tx); all warps wait on it, read
// their fragments as conflict-free 256-byte rows, and release the stage on an "empty" mbarrier.
// Only that one lane per slice computes a global address and there is no __syncthreads in the
// main loop; the ring runs across tile boundaries, so the register-only
// epilogue of tile t overlaps the loads of tile t+1.  (The per-thread cp.async version of this
// loop topped out at 27 TFLOP/s -- the address arithmetic competes with DMMA issue; this one
// reaches the DMMA pipe's 36 TFLOP/s in isolation, profiles/r1_gemm_tma_probe.log.)
#include "qdc_common.cuh"

namespace {

#ifndef QDC_TN
#define QDC_TN 128
#endif
#ifndef QDC_STAGES
#define QDC_STAGES (QDC_TN == 128 ? 6 : 4)
#endif
constexpr int TM = 128, TN = QDC_TN, BK = 16, STAGES = QDC_STAGES;
constexpr int WARPS_N = TN / 32;                             // warp tile is 32 x 32
constexpr int CONSUMER_WARPS = 4 * WARPS_N, THREADS = CONSUMER_WARPS * 32;
constexpr int CTAS_PER_SM = TN == 128 ? 1 : 2;               // 128 registers per thread either way
constexpr int LAG = 2;                                       // producer runs STAGES - LAG slices ahead
constexpr int OPERAND_DOUBLES = (BK / 4) * TM * 4;          // A slice: 2048 doubles = 16 KB
constexpr int OPERAND_B_DOUBLES = (BK / 4) * TN * 4;
constexpr int STAGE_DOUBLES = OPERAND_DOUBLES + OPERAND_B_DOUBLES;   // A then B
constexpr int SMEM_BYTES = STAGES * STAGE_DOUBLES * 8;       // 192 KB (TN 128) / 96 KB (TN 64)
constexpr int PANEL_W = 2048 / TN;                           // tile columns per L2 panel
constexpr int PACK_PAD_ROWS = 2 * TM;                        // zero rows after the last real row

__host__ __device__ __forceinline__ long long packed_rows(long long n_rows) {
    return (n_rows + TM - 1) / TM * TM + PACK_PAD_ROWS;
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

__device__ __forceinline__ void dmma_8x8x4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

constexpr uint64_t FLAG_ROW_BAD = 1ull << 63, FLAG_COL_BAD = 1ull << 62;

struct TileCoord { long long r0, c0; };

// 32-bit tile walk (n_tiles < 2^31 is checked on the host): column panels of PANEL_W tiles;
// inside a panel walk rows, then the panel's columns.
__device__ __forceinline__ TileCoord tile_coord(int t, int tiles_m, int tiles_n) {
    const int per_panel = tiles_m * PANEL_W;
    const int panel = t / per_panel;
    const int rem = t - panel * per_panel;
    const int left = tiles_n - panel * PANEL_W;
    const int pw = left < PANEL_W ? left : PANEL_W;
    const int r = rem / pw, c = panel * PANEL_W + (rem - r * pw);
    return {(long long)r * TM, (long long)c * TN};
}

struct ContractArgs {
    const double* Ap;       // packed [kdim/4][a_rows_pad][4]
    const double* Bp;       // packed [kdim/4][b_rows_pad][4]
    long long a_rows_pad, b_rows_pad;
    long long row_lo, n_rows, n_cols;
    int kdim;
    double scale;
    const double* d_norm;
    MisTables tab;
    CandSink sink;
    void* P_out;
    double* partials;
};

__device__ __forceinline__ double opaque(double x) {
    asm volatile("" : "+d"(x));          // keeps the weights in FP64 (no I2F per output)
    return x;
}

// Epilogue of one 128x128 tile from the 4x4x2 accumulators of this thread.  FULL: the tile lies
// entirely inside the slab, no per-output bounds tests.  The independence test is folded into one
// AND: bit 63 is set in every column mask and in the row mask of a row that is not itself
// independent; bit 62 the other way round, so (cf & rf) == 0 <=> row ok, column ok, no edge.
template <bool FULL, int STORE, bool OUT32>
__device__ __forceinline__ void tile_epilogue(const ContractArgs& a, double (&acc)[4][4][2],
                                              MisAccum& m0, MisAccum& m1, const TileCoord tc,
                                              const double out_scale, int wm, int wn, int lane) {
    const long long cbase = tc.c0 + wn * 32 + 2 * (lane & 3);
    const long long rbase = tc.r0 + wm * 32 + (lane >> 2);
    const double thr = a.sink.threshold;
    uint64_t rf[4];
    double rwd[4];
    uint32_t hit = 0;                           // bit i*8 + j*2 + h: output above the threshold
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const long long r = rbase + i * 8;
        const long long rg = a.row_lo + ((FULL || r < a.n_rows) ? r : 0);
        const uint32_t rw = __ldg(a.tab.row_wok + rg);
        rf[i] = __ldg(a.tab.row_forb + rg) | FLAG_COL_BAD | ((rw >> 8) & 1u ? 0ull : FLAG_ROW_BAD);
        rwd[i] = opaque((double)(rw & 0xffu));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const long long c = cbase + j * 8;
        const bool cok = FULL || c < a.n_cols;
        const bool two = FULL || c + 1 < a.n_cols;   // n_cols even or 1
        const long long c0i = cok ? c : 0, c1i = two ? c + 1 : 0;
        uint32_t cw0, cw1;
        uint64_t cx0, cx1;
        if (FULL) {                                  // c is even: one 8- and one 16-byte load
            const uint2 cw = __ldg(reinterpret_cast<const uint2*>(a.tab.col_wok + c));
            const ulonglong2 cx = __ldg(reinterpret_cast<const ulonglong2*>(a.tab.col_x + c));
            cw0 = cw.x; cw1 = cw.y; cx0 = cx.x; cx1 = cx.y;
        } else {
            cw0 = __ldg(a.tab.col_wok + c0i); cw1 = __ldg(a.tab.col_wok + c1i);
            cx0 = __ldg(a.tab.col_x + c0i); cx1 = __ldg(a.tab.col_x + c1i);
        }
        const uint64_t cf0 = cx0 | FLAG_ROW_BAD | ((cw0 >> 8) & 1u ? 0ull : FLAG_COL_BAD);
        const uint64_t cf1 = cx1 | FLAG_ROW_BAD | ((cw1 >> 8) & 1u ? 0ull : FLAG_COL_BAD);
        const double cwd0 = opaque((double)(cw0 & 0xffu)), cwd1 = opaque((double)(cw1 & 0xffu));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const double p0 = acc[i][j][0] * out_scale, p1 = acc[i][j][1] * out_scale;
            const long long r = rbase + i * 8;
            if (FULL || (r < a.n_rows && cok)) {
                const double w0 = rwd[i] + cwd0;
                m0.s_p += p0;
                m0.s_pw = fma(p0, w0, m0.s_pw);
                if ((cf0 & rf[i]) == 0ull) m0.s_pwok = fma(p0, w0, m0.s_pwok);
                if (p0 > thr) hit |= 1u << (i * 8 + j * 2);
                if (two) {
                    const double w1 = rwd[i] + cwd1;
                    m1.s_p += p1;
                    m1.s_pw = fma(p1, w1, m1.s_pw);
                    if ((cf1 & rf[i]) == 0ull) m1.s_pwok = fma(p1, w1, m1.s_pwok);
                    if (p1 > thr) hit |= 2u << (i * 8 + j * 2);
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
                    const uint64_t rx = __ldg(a.tab.row_x + a.row_lo + r);
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
    // rare path: emit candidates, selecting the accumulator by a predicated scan
    while (hit) {
        const int b = __ffs(hit) - 1;
        hit &= hit - 1;
        double v = 0.0;
#pragma unroll
        for (int q = 0; q < 32; ++q)
            if (b == q) v = acc[q >> 3][(q >> 1) & 3][q & 1];
        const long long r = rbase + (b >> 3) * 8, c = cbase + ((b >> 1) & 3) * 8 + (b & 1);
        cand_emit(a.sink, __ldg(a.tab.row_x + a.row_lo + r) | __ldg(a.tab.col_x + c), v * out_scale);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
}

// STORE: 0 none, 1 united order, 2 circuit-order scatter.  OUT32: write float instead of double.
template <int STORE, bool OUT32>
__global__ void __launch_bounds__(THREADS, CTAS_PER_SM) contract_f64_kernel(ContractArgs a) {
    extern __shared__ __align__(128) double smem[];
    __shared__ double scratch[3 * (THREADS / 32)];
    __shared__ __align__(8) uint64_t bar_full[STAGES], bar_empty[STAGES];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_m = (int)((a.n_rows + TM - 1) / TM), tiles_n = (int)((a.n_cols + TN - 1) / TN);
    const int n_tiles = tiles_m * tiles_n;
    const int kchunks = a.kdim / BK;
    const double out_scale = a.scale * (a.d_norm ? 1.0 / *a.d_norm : 1.0);

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], CONSUMER_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    MisAccum macc = {0.0, 0.0, 0.0}, macc2 = {0.0, 0.0, 0.0};

    // ---- producer duty, rotated over the warps: k slice p is issued by warp p % 16 --------------
    // Slice p reuses the ring slot of slice p - STAGES, which every warp released LAG iterations
    // ago, so the empty-barrier wait below practically never blocks.  All threads advance the
    // cursor; only the elected lane touches barriers and issues the 8 bulk copies.
    // The producer cursor is (p_tile, p_kc); its ring slot and phase follow from the consumer's
    // (it runs exactly STAGES - LAG slices ahead), the issuing warp is p_kc % 16.
    int p_tile = blockIdx.x, p_kc = 0;
    int p_r0 = 0, p_c0 = 0;                      // tile origin in units of rows (fits 32 bits: < 2^31 outputs per side)
    {
        const TileCoord t0 = tile_coord(p_tile < n_tiles ? p_tile : 0, tiles_m, tiles_n);
        p_r0 = (int)t0.r0; p_c0 = (int)t0.c0;
    }
    auto produce = [&](int p_stage, uint32_t p_phase) {
        if (p_tile >= n_tiles) return;
        if (warp == (p_kc & (CONSUMER_WARPS - 1)) && lane == 0) {
            mbar_wait(&bar_empty[p_stage], p_phase ^ 1u);
            mbar_expect_tx(&bar_full[p_stage], STAGE_DOUBLES * 8);
            double* st = smem + p_stage * STAGE_DOUBLES;
            const double* srcA = a.Ap + ((long long)p_kc * (BK / 4) * a.a_rows_pad + a.row_lo + p_r0) * 4;
            const double* srcB = a.Bp + ((long long)p_kc * (BK / 4) * a.b_rows_pad + p_c0) * 4;
#pragma unroll
            for (int kg = 0; kg < BK / 4; ++kg) {
                bulk_g2s(st + kg * TM * 4, srcA + (long long)kg * a.a_rows_pad * 4, TM * 32, &bar_full[p_stage]);
                bulk_g2s(st + OPERAND_DOUBLES + kg * TN * 4, srcB + (long long)kg * a.b_rows_pad * 4, TN * 32,
                         &bar_full[p_stage]);
            }
        }
        if (++p_kc == kchunks) {
            p_kc = 0;
            p_tile += gridDim.x;
            if (p_tile < n_tiles) {
                const TileCoord t = tile_coord(p_tile, tiles_m, tiles_n);
                p_r0 = (int)t.r0; p_c0 = (int)t.c0;
            }
        }
    };
#pragma unroll 1
    for (int i = 0; i < STAGES - LAG; ++i) produce(i, 0u);

    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
#pragma unroll 1
        for (int kc = 0; kc < kchunks; ++kc) {
            {
                const int ps = stage + (STAGES - LAG);
                const bool wrap = ps >= STAGES;
                produce(wrap ? ps - STAGES : ps, wrap ? phase ^ 1u : phase);
            }
            mbar_wait(&bar_full[stage], phase);
            const double* sA = smem + stage * STAGE_DOUBLES;
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
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_empty[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        // epilogue of this tile, registers only; interior tiles take the guard-free path
        const TileCoord tc = tile_coord(tile, tiles_m, tiles_n);
        if (tc.r0 + TM <= a.n_rows && tc.c0 + TN <= a.n_cols)
            tile_epilogue<true, STORE, OUT32>(a, acc, macc, macc2, tc, out_scale, wm, wn, lane);
        else
            tile_epilogue<false, STORE, OUT32>(a, acc, macc, macc2, tc, out_scale, wm, wn, lane);
    }
    macc.s_p += macc2.s_p; macc.s_pw += macc2.s_pw; macc.s_pwok += macc2.s_pwok;
    mis_block_store(macc, scratch, a.partials);
}

// ---- operand packing: row-major [n_rows][kdim] -> [kdim/4][rows_pad][4], zero padded ----------
__global__ void __launch_bounds__(256) pack_kernel(const double* __restrict__ X, long long n_rows,
                                                   long long rows_pad, int kdim,
                                                   double* __restrict__ out) {
    // one thread per (k group, row): reads 32 contiguous bytes, writes 32 contiguous bytes;
    // threads of a warp walk k groups of one row first so reads coalesce into 128-byte lines
    const int kgroups = kdim / 4;
    const long long total = rows_pad * kgroups;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / kgroups;
        const int g = (int)(i - row * kgroups);
        double2 v0 = make_double2(0.0, 0.0), v1 = v0;
        if (row < n_rows) {
            const double2* src = reinterpret_cast<const double2*>(X + row * kdim + g * 4);
            v0 = src[0]; v1 = src[1];
        }
        double2* dst = reinterpret_cast<double2*>(out + ((long long)g * rows_pad + row) * 4);
        dst[0] = v0; dst[1] = v1;
    }
}

}  // namespace

// (column sums and the normalisation: i8_prep.cu, qdc_colsum / qdc_contract_norm)

template <int STORE, bool OUT32>
static int launch_contract(const ContractArgs& args, int grid, cudaStream_t s) {
    auto kern = contract_f64_kernel<STORE, OUT32>;
    QDC_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    kern<<<grid, THREADS, SMEM_BYTES, s>>>(args);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int64_t qdc_contract_packed_doubles(int64_t n_rows, int kdim) {
    return n_rows < 0 || kdim < 0 ? 0 : packed_rows(n_rows) * (int64_t)kdim;
}

extern "C" int qdc_contract_pack(const double* d_X, int64_t n_rows, int kdim, double* d_packed,
                                 void* stream) {
    QDC_REQUIRE(d_X && d_packed, "null buffer");
    QDC_REQUIRE(n_rows >= 1 && kdim >= BK && kdim % BK == 0, "shape (kdim must be a multiple of 16)");
    QDC_REQUIRE((((uintptr_t)d_X) & 15) == 0 && (((uintptr_t)d_packed) & 15) == 0, "16-byte alignment");
    const long long rows_pad = packed_rows(n_rows);
    const long long total = rows_pad * (kdim / 4);
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)qdc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    pack_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_X, n_rows, rows_pad, kdim, d_packed);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_contract_f64(const double* d_Ap, int64_t a_rows_total, const double* d_Bp,
                                int64_t row_lo,
                                int64_t n_rows, int64_t n_cols, int kdim, double scale,
                                const double* d_norm, double threshold,
                                const uint64_t* d_row_x, const uint64_t* d_row_forb,
                                const uint32_t* d_row_wok, const uint64_t* d_col_x,
                                const uint32_t* d_col_wok, void* d_P_out, int store_mode,
                                int out_dtype, double* d_sums, uint64_t* d_cand_x,
                                double* d_cand_p, unsigned long long* d_cand_count,
                                int64_t cand_cap, double* d_partials, void* stream) {
    QDC_REQUIRE(d_Ap && d_Bp && d_sums && d_partials, "null buffer");
    QDC_REQUIRE(a_rows_total >= 1 && row_lo + n_rows <= a_rows_total, "row slab outside the row operand");
    QDC_REQUIRE((((uintptr_t)d_Ap) & 15) == 0 && (((uintptr_t)d_Bp) & 15) == 0, "16-byte alignment");
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
    args.Ap = d_Ap;
    args.Bp = d_Bp;
    args.a_rows_pad = packed_rows(a_rows_total);
    args.b_rows_pad = packed_rows(n_cols);
    args.row_lo = row_lo; args.n_rows = n_rows; args.n_cols = n_cols; args.kdim = kdim;
    args.scale = scale; args.d_norm = d_norm;
    args.tab = {d_row_x, d_row_forb, d_row_wok, d_col_x, d_col_wok};
    args.sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
    args.P_out = d_P_out; args.partials = d_partials;
    const long long n_tiles = ((n_rows + TM - 1) / TM) * ((n_cols + TN - 1) / TN);
    QDC_REQUIRE(n_tiles < (1ll << 31) - 4096, "too many output tiles for one launch");
    const int sms = qdc_sm_count();
    const int grid = (int)(n_tiles < (long long)sms * CTAS_PER_SM ? n_tiles : (long long)sms * CTAS_PER_SM);
    int rc;
    if (store_mode == 0) rc = launch_contract<0, false>(args, grid, s);
    else if (store_mode == 1) rc = out_dtype ? launch_contract<1, true>(args, grid, s)
                                              : launch_contract<1, false>(args, grid, s);
    else rc = out_dtype ? launch_contract<2, true>(args, grid, s)
                        : launch_contract<2, false>(args, grid, s);
    if (rc != QDC_OK) return rc;
    return qdc_launch_finalize_sums(d_partials, grid, d_sums, s);
}
