// K5 (alternative): the contraction on the 5th-generation tensor cores (tcgen05.mma kind::i8,
// accumulators in TMEM) through an error-free split of the FP64 operands into signed 7-bit
// slices ("Ozaki scheme"), with the same fused MIS epilogue as contract.cu.
//
//   A[r, k] = 2^e_r  * 2^-6 * sum_i s_i[r, k] 2^-7i ,   s_i integer in [-64, 64]   (i < S)
//   Bt[c, k] = 2^f_c * 2^-6 * sum_j t_j[c, k] 2^-7j
//   P~[r, c] = scale * 2^(e_r + f_c - 12) * sum_d 2^-7d  D_d[r, c],   D_d = sum_{i+j=d} s_i . t_j
//
// Every D_d is an exact int32 dot product (|D_d| <= (d+1) * kdim * 2^12 < 2^23 for kdim <= 256);
// X = sum_d D_d 2^(7(S-1-d)) is formed exactly in int64 and rounded to FP64 once.  What is
// dropped: the operand residuals (|.| <= 2^(-7S-1) relative to the row maximum) and the slice
// pairs with i + j >= S; for S = 6 the result differs from the FP64 product by at most
//   |err| <= 2^(e_r + f_c) * scale * kdim * (S + 2) * 2^(-7S)   (= 4.7e-10 * 2^(e_r+f_c) * scale for kdim 256)
// i.e. < 3e-11 for the normalised K = 4 fragment tensors (scale = 1/16), inside the 1e-10 bar of
// the FP64 path; S is a run-time choice (6 or 7 slices fit the 512 TMEM columns as 64-column
// accumulators; 6 keeps all A slices of a row block resident in shared memory).
//
// Structure (one persistent CTA per SM, 10 warps):
//   warp 0  producer: bulk async copies (cp.async.bulk + mbarrier complete_tx).  The S slices of
//           a 128-row block of A (S * 128 * kdim bytes) are loaded once and stay resident while
//           the CTA walks all column blocks; the slices t_j of a 64-column block stream through a
//           2-stage ring.
//   warp 1  MMA issuer (one lane): for j (ring stage) for i <= S-1-j: 128x64x32 tcgen05.mma steps
//           into accumulator d = i + j (TMEM columns 64 d .. 64 d + 63); order d is complete
//           after the first pair of group j = d and is committed to its own mbarrier, so the
//           epilogue drains orders while later pairs still run.
//   warps 2-9  epilogue: tcgen05.ld the finished order, X = 128 X + D in registers (32 outputs
//           per thread), release the accumulator; after the last order: one int64 -> FP64
//           conversion, the row / column scale, and the MIS sums / candidates of qdc_common.cuh.
// Operands are pre-sliced and stored in the K-major core-matrix order the MMA reads
// ([8-row group][16-byte k chunk][8 rows][16 bytes]), so a block is one contiguous run.
#include "qdc_common.cuh"

namespace {

constexpr int TM = 128, TN = 64;
constexpr int MAX_S = 7;
constexpr int B_STAGES = 2;
constexpr int N_EPI_WARPS = 8;
constexpr int THREADS = (2 + N_EPI_WARPS) * 32;
constexpr int TMEM_COLS = 512;

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
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor, K-major, no swizzle: core matrix = 8 rows x 16 bytes (128
// contiguous bytes); LBO = distance between core matrices along K, SBO = between 8-row groups
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}

struct I8Args {
    const int8_t* A8;        // [S][row blocks][128 x kdim core order]
    const int8_t* B8;        // [S][col blocks][64 x kdim core order]
    const double* row_scale; // 2^e_r per row (padded rows: 0)
    const double* col_scale; // 2^f_c per column
    long long n_rows, n_cols, row_lo;     // row_lo: offset of this slab in the MIS row tables / united index
    int n_rb, n_cb, kdim, S;
    double out_scale;        // scale * 2^(-12 - 7(S-1))
    const double* d_norm;
    MisTables tab;
    CandSink sink;
    double* P_out;           // optional united slab [n_rows][n_cols]
    double* partials;
};

__global__ void __launch_bounds__(THREADS, 1) contract_i8_kernel(const I8Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t a_full, a_empty, b_full[B_STAGES], b_empty[B_STAGES],
        acc_full[MAX_S], acc_empty[MAX_S];
    __shared__ uint32_t tmem_base_sh;
    __shared__ double scratch[3 * (THREADS / 32)];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = a.S, kdim = a.kdim;
    const uint32_t a_slice_bytes = TM * kdim, b_slice_bytes = TN * kdim;
    uint8_t* sA = smem;                                   // [S][a_slice_bytes]
    uint8_t* sB = smem + (size_t)S * a_slice_bytes;       // [B_STAGES][b_slice_bytes]

    if (tid == 0) {
        mbar_init(&a_full, 1); mbar_init(&a_empty, 1);
        for (int s = 0; s < B_STAGES; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
        for (int d = 0; d < MAX_S; ++d) { mbar_init(&acc_full[d], 1); mbar_init(&acc_empty[d], N_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sh)), "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_sh;

    MisAccum macc = {0.0, 0.0, 0.0};

    if (warp == 0) {
        // ================= producer =================
        if (lane == 0) {
            int b_stage = 0; uint32_t b_phase = 0, rb_phase = 0;
            for (int rb = blockIdx.x; rb < a.n_rb; rb += gridDim.x, rb_phase ^= 1u) {
                mbar_wait(&a_empty, rb_phase ^ 1u);
                mbar_expect_tx(&a_full, (uint32_t)S * a_slice_bytes);
                for (int i = 0; i < S; ++i)
                    bulk_g2s(sA + (size_t)i * a_slice_bytes,
                             a.A8 + ((size_t)i * a.n_rb + rb) * a_slice_bytes, a_slice_bytes, &a_full);
                for (int cb = 0; cb < a.n_cb; ++cb) {
                    for (int j = 0; j < S; ++j) {
                        mbar_wait(&b_empty[b_stage], b_phase ^ 1u);
                        mbar_expect_tx(&b_full[b_stage], b_slice_bytes);
                        bulk_g2s(sB + (size_t)b_stage * b_slice_bytes,
                                 a.B8 + ((size_t)j * a.n_cb + cb) * b_slice_bytes, b_slice_bytes, &b_full[b_stage]);
                        if (++b_stage == B_STAGES) { b_stage = 0; b_phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
            const uint32_t sbo = (uint32_t)(kdim / 16) * 128u;
            const int ksteps = kdim / 32;
            int b_stage = 0; uint32_t b_phase = 0, rb_phase = 0, tile_phase = 0;
            for (int rb = blockIdx.x; rb < a.n_rb; rb += gridDim.x, rb_phase ^= 1u) {
                mbar_wait(&a_full, rb_phase);
                for (int cb = 0; cb < a.n_cb; ++cb, tile_phase ^= 1u) {
                    for (int j = 0; j < S; ++j) {
                        mbar_wait(&b_full[b_stage], b_phase);
                        tc_fence_after();
                        const uint32_t b_addr = smem_u32(sB + (size_t)b_stage * b_slice_bytes);
                        for (int i = 0; i + j < S; ++i) {
                            const int d = i + j;
                            if (j == 0) {                              // first touch of order d in this tile
                                mbar_wait(&acc_empty[d], tile_phase ^ 1u);
                                tc_fence_after();
                            }
                            const uint32_t a_addr = smem_u32(sA + (size_t)i * a_slice_bytes);
                            const uint32_t tmem_d = tmem + (uint32_t)d * TN;
                            for (int ks = 0; ks < ksteps; ++ks)
                                umma_i8(tmem_d, smem_desc(a_addr + ks * 256, 128, sbo),
                                        smem_desc(b_addr + ks * 256, 128, sbo), idesc, (j > 0 || ks > 0) ? 1u : 0u);
                            if (i == 0) umma_commit(&acc_full[j]);     // order j is complete
                        }
                        umma_commit(&b_empty[b_stage]);
                        if (++b_stage == B_STAGES) { b_stage = 0; b_phase ^= 1u; }
                    }
                }
                umma_commit(&a_empty);
            }
        }
    } else {
        // ================= epilogue =================
        const int ew = warp - 2;
        const int quarter = warp & 3;                      // TMEM lanes 32 * (warp % 4) .. + 31
        const int half = ew >> 2;                          // column half of the 64-column accumulator
        const double inv_norm = a.d_norm ? 1.0 / *a.d_norm : 1.0;
        const double thr = a.sink.threshold;
        uint32_t tile_phase = 0;
        for (int rb = blockIdx.x; rb < a.n_rb; rb += gridDim.x) {
            const long long r_loc = (long long)rb * TM + quarter * 32 + lane;       // row inside the slab
            const bool row_ok = r_loc < a.n_rows;
            const long long r = a.row_lo + r_loc;                                   // row in the tables
            const uint64_t rx = row_ok ? a.tab.row_x[r] : 0ull;
            const uint64_t rforb = row_ok ? a.tab.row_forb[r] : 0ull;
            const uint32_t rwok = row_ok ? a.tab.row_wok[r] : 0u;
            const double rs = row_ok ? a.row_scale[r_loc] * a.out_scale * inv_norm : 0.0;
            for (int cb = 0; cb < a.n_cb; ++cb, tile_phase ^= 1u) {
                long long X[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) X[c] = 0;
                for (int d = 0; d < S; ++d) {
                    mbar_wait(&acc_full[d], tile_phase);
                    tc_fence_after();
                    uint32_t v[32];
                    const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(d * TN + half * 32);
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[d]);
#pragma unroll
                    for (int c = 0; c < 32; ++c) X[c] = X[c] * 128 + (long long)(int32_t)v[c];
                }
                // scale, score (mis_accumulate of qdc_common.cuh), optional store
                const long long c0 = (long long)cb * TN + half * 32;
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const long long col = c0 + c;
                    if (!row_ok || col >= a.n_cols) continue;
                    const double p = (double)X[c] * rs * a.col_scale[col];
                    mis_accumulate(macc, p, rx, rforb, rwok, a.tab.col_x[col], a.tab.col_wok[col], a.sink);
                    if (a.P_out) a.P_out[r_loc * a.n_cols + col] = p;
                }
                (void)thr;
            }
        }
    }
    // teardown
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
    mis_block_store(macc, scratch, a.partials);
}

// ---- slicing pre-pass -----------------------------------------------------------------------
// one warp per row: exponent from the row maximum, then S slices of every element, written in
// core-matrix order.  All arithmetic is exact (scaling by powers of two, rint, subtraction).
__device__ __forceinline__ size_t core_offset(int row_in_block, int k, int kdim) {
    return ((size_t)(row_in_block >> 3) * (kdim / 16) + (k >> 4)) * 128 + (row_in_block & 7) * 16 + (k & 15);
}

__global__ void __launch_bounds__(256) slice_kernel(const double* __restrict__ X, long long n_rows, int kdim_src,
                                                    int kdim, int block_rows, long long n_blocks, int S,
                                                    int8_t* __restrict__ out, double* __restrict__ scale_out) {
    const long long row = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long rows_pad = n_blocks * block_rows;
    if (row >= rows_pad) return;
    const bool live = row < n_rows;
    double mx = 0.0;
    if (live)
        for (int k = lane; k < kdim_src; k += 32) mx = fmax(mx, fabs(X[row * kdim_src + k]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    int e = 0;
    if (mx > 0.0) frexp(mx, &e);                       // mx = m 2^e, m in [0.5, 1)
    const double up = ldexp(1.0, 6 - e);               // a' 2^6, |a'| < 1
    if (lane == 0) scale_out[row] = live ? ldexp(1.0, e) : 0.0;
    const long long blk = row / block_rows;
    const int rib = (int)(row - blk * block_rows);
    const size_t slice_stride = (size_t)n_blocks * block_rows * kdim;
    int8_t* base = out + (size_t)blk * block_rows * kdim;
    for (int k = lane; k < kdim; k += 32) {
        double t = (live && k < kdim_src) ? X[row * kdim_src + k] * up : 0.0;
        const size_t off = core_offset(rib, k, kdim);
        for (int i = 0; i < S; ++i) {
            const double s = rint(t);
            base[(size_t)i * slice_stride + off] = (int8_t)(int)s;
            t = (t - s) * 128.0;
        }
    }
}

}  // namespace

extern "C" int64_t qdc_contract_i8_workspace_bytes(int64_t n_rows, int64_t n_cols, int kdim, int n_slices) {
    if (n_rows < 0 || n_cols < 0 || kdim < 1 || n_slices < 1) return 0;
    const int64_t kd = (kdim + 31) / 32 * 32;
    const int64_t rb = (n_rows + TM - 1) / TM, cb = (n_cols + TN - 1) / TN;
    const int64_t a8 = (rb * TM * kd * n_slices + 255) / 256 * 256, b8 = (cb * TN * kd * n_slices + 255) / 256 * 256;
    return a8 + b8 + (rb * TM + cb * TN) * 8 + 256;
}

extern "C" int qdc_contract_i8(const double* d_A, int64_t a_rows_total, const double* d_Bt, int64_t row_lo,
                               int64_t n_rows, int64_t n_cols, int kdim, int n_slices, double scale,
                               const double* d_norm, double threshold,
                               const uint64_t* d_row_x, const uint64_t* d_row_forb,
                               const uint32_t* d_row_wok, const uint64_t* d_col_x,
                               const uint32_t* d_col_wok, void* d_P_out, int store_mode,
                               void* d_workspace, double* d_sums, uint64_t* d_cand_x, double* d_cand_p,
                               unsigned long long* d_cand_count, int64_t cand_cap,
                               double* d_partials, void* stream) {
    QDC_REQUIRE(d_A && d_Bt && d_sums && d_partials && d_workspace, "null buffer");
    QDC_REQUIRE(a_rows_total >= 1 && row_lo >= 0 && row_lo + n_rows <= a_rows_total, "row slab outside the row operand");
    QDC_REQUIRE(d_row_x && d_row_forb && d_row_wok && d_col_x && d_col_wok, "null tables");
    QDC_REQUIRE(d_cand_x && d_cand_p && d_cand_count && cand_cap >= 0, "candidate sink");
    QDC_REQUIRE(kdim >= 1 && kdim <= 256, "the int8 path keeps a row block resident: kdim <= 256");
    QDC_REQUIRE(n_slices >= 2 && n_slices <= 6, "n_slices in 2..6");
    QDC_REQUIRE(n_rows >= 0 && n_cols >= 1, "shape");
    QDC_REQUIRE(store_mode == 0 || (store_mode == 1 && d_P_out), "store_mode 0 (none) or 1 (united slab, float64)");
    if (n_rows == 0) return QDC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int kd = (kdim + 31) / 32 * 32;
    const long long n_rb = (n_rows + TM - 1) / TM, n_cb = (n_cols + TN - 1) / TN;
    QDC_REQUIRE(n_rb < (1ll << 31) && n_cb < (1ll << 31), "too many tiles");
    uint8_t* ws = (uint8_t*)d_workspace;
    QDC_REQUIRE((((uintptr_t)ws) & 255) == 0, "workspace must be 256-byte aligned");
    const size_t a8_bytes = ((size_t)n_rb * TM * kd * n_slices + 255) / 256 * 256;
    const size_t b8_bytes = ((size_t)n_cb * TN * kd * n_slices + 255) / 256 * 256;
    int8_t* A8 = (int8_t*)ws;
    int8_t* B8 = (int8_t*)(ws + a8_bytes);
    double* rs = (double*)(ws + a8_bytes + b8_bytes);
    double* cs = rs + n_rb * TM;
    {
        const long long warps = n_rb * TM;
        slice_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(d_A + row_lo * kdim, n_rows, kdim, kd, TM, n_rb,
                                                                          n_slices, A8, rs);
        QDC_CHECK_LAUNCH();
        const long long warps_b = n_cb * TN;
        slice_kernel<<<(unsigned)((warps_b * 32 + 255) / 256), 256, 0, s>>>(d_Bt, n_cols, kdim, kd, TN, n_cb, n_slices, B8, cs);
        QDC_CHECK_LAUNCH();
    }
    I8Args a;
    a.A8 = A8; a.B8 = B8; a.row_scale = rs; a.col_scale = cs;
    a.n_rows = n_rows; a.n_cols = n_cols; a.row_lo = row_lo;
    a.n_rb = (int)n_rb; a.n_cb = (int)n_cb; a.kdim = kd; a.S = n_slices;
    a.out_scale = scale * ldexp(1.0, -12 - 7 * (n_slices - 1));
    a.d_norm = d_norm;
    a.tab = {d_row_x, d_row_forb, d_row_wok, d_col_x, d_col_wok};
    a.sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
    a.P_out = store_mode ? (double*)d_P_out : nullptr;
    a.partials = d_partials;
    const size_t smem = (size_t)n_slices * TM * kd + (size_t)B_STAGES * TN * kd;
    QDC_REQUIRE(smem <= 227 * 1024 - 2048, "shared memory");
    QDC_CHECK_CUDA(cudaFuncSetAttribute(contract_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int sms = qdc_sm_count();
    const int grid = (int)(n_rb < sms ? n_rb : sms);
    contract_i8_kernel<<<grid, THREADS, smem, s>>>(a);
    QDC_CHECK_LAUNCH();
    return qdc_launch_finalize_sums(d_partials, grid, d_sums, s);
}
