// K5 (alternative): the contraction on the 5th-generation tensor cores (tcgen05.mma kind::i8,
// accumulators in TMEM) through an error-free split of the FP64 operands into 8-bit digits
// ("Ozaki scheme"), with the same fused MIS epilogue as contract.cu.
//
// Per row r, 2^e_r is a power-of-two ceiling of the row maximum and x' = x 2^(7 - e_r) lies in
// [-127, 127].  x' (rounded to S radix-256 digits) is written with balanced digits:
//   x' ~ sum_{i=0..S-1} s_i 256^-i ,   s_i integer in [-128, 127]   (signed bytes),
// obtained as the two's-complement digits of x' + 128 (256^-1 + ... + 256^-(S-1)) minus 128 (no
// carries, no overflow); likewise t_j for the columns of Bt.  Every slice pair is one exact
// integer GEMM:
//   P~[r, c] = scale * 2^(e_r + f_c - 14) * sum_d 256^-d D_d[r, c],   D_d = sum_{i+j=d} s_i . t_j
// |D_d| <= (d+1) * kdim * 2^14 < 2^25 for kdim <= 256: exact in the int32 accumulators.
// X = sum_d D_d 256^(S-1-d) is accumulated by FP64 FMAs (exact while |X| < 2^53; at most two
// roundings of relative size 2^-53 for S = 6).  What is dropped: the operand rounding (<= 2^(-8S) relative to 2^e_r) and the
// slice pairs with i + j >= S; both have random sign (balanced digits), and
//   |err| <= 2^(e_r + f_c) * scale * kdim * (S + 1.1) * 2^(-8S)
//   S = 5: 1.4e-9, S = 6: 6.5e-12  (x 2^(e_r + f_c) * scale, kdim 256); FP64 itself: 2.8e-14.
// S is a run-time choice: 6 orders of 64 columns fit the 512 TMEM columns, and S slices of a row
// block stay resident in shared memory (S = 6, kdim = 256: 192 KB, which leaves a 2-stage ring for
// the streamed operand; S = 5: 4 stages).
//
// Structure (one persistent CTA per SM; tile = 128 rows x TN columns, TN = 64, or 96 at S = 5 where
// 5 x 96 accumulator columns fit TMEM -- an N = 96 instruction costs what N = 64 does; 2 + TN / 4 warps):
//   warp 0  producer: bulk async copies (cp.async.bulk + mbarrier complete_tx).  The S slices of
//           a 128-row block of A (S * 128 * kdim bytes) are loaded once and stay resident while
//           the CTA walks all column blocks; the slices t_j of a 64-column block stream through a
//           ring whose depth is whatever shared memory is left (2..8 stages).
//   warp 1  MMA issuer (one lane): for j (ring stage) for i <= S-1-j: 128x64x32 tcgen05.mma steps
//           into accumulator d = i + j (TMEM columns 64 d .. 64 d + 63); order d is complete
//           after the first pair of group j = d and is committed to its own mbarrier, so the
//           epilogue drains orders while later pairs still run.  The first n_ta resident slices are
//           copied to the TMEM columns the accumulators leave free (tcgen05.cp.128x256b) at the
//           start of a work item; their pairs take A from tensor memory and read only the B tile
//           from shared memory (the N = 64 instruction is bound by the 6 KB operand read).
//   warps 2..  epilogue (4 TMEM lane quarters x TN / 16 column groups of 16): tcgen05.ld two finished orders at a time, X = fma(X, 256, D_d) in FP64
//           (int32 -> FP64 through the 2^52 trick, no conversion unit), release the
//           accumulators; then, branch-free per output:
//           q = X cs_c, four running sums per row (sum q, sum q w_c, both again under the
//           independence mask) and a threshold bit; the row factor (2^e_r, normalisation, row
//           weight) is applied once per row block, candidates are emitted on a rare slow path.
// Operands are pre-sliced and stored in the K-major core-matrix order the MMA reads
// ([8-row group][16-byte k chunk][8 rows][16 bytes]), so a block is one contiguous run.
#include "qdc_common.cuh"
#include <stdlib.h>

namespace {

constexpr int TM = 128;                       // tile rows; tile columns TN and columns per epilogue thread CPT are
                                              // template parameters of the kernel: (64, 16), or (96, 32) where S * 96 <= 512
constexpr int MAX_S = 6;
constexpr int MAX_STAGES = 8;                 // ring of streamed-operand slices
constexpr int FIRST_EPI_WARP = 2;             // warp 0: producer, warp 1: MMA issuer
constexpr int MAX_WARPS = 32;
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
// polling costs shared-memory cycles the tensor core wants for its operands: waiters that are
// not on the critical path back off between polls
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity, uint32_t ns) {
    uint32_t done = 0;
    while (true) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (done) break;
        __nanosleep(ns);
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
    return pred != 0;
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

// the same with the A operand in tensor memory (128 lanes x 8 columns = 128 rows x 32 bytes)
__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n}\n"
        ::"r"(tmem_d), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// 128 rows x 32 bytes of a K-major core-matrix image (the descriptor the MMA would take) -> 8 TMEM columns
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t tmem_dst, uint64_t sdesc) {
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tmem_dst), "l"(sdesc) : "memory");
}

constexpr uint64_t FLAG_ROW_BAD = 1ull << 63, FLAG_COL_BAD = 1ull << 62;   // as in contract.cu

// CPT consecutive TMEM columns of this warp's 32 lanes (one accumulator row per thread)
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}

__device__ __forceinline__ double opaque(double x) {
    asm volatile("" : "+d"(x));
    return x;
}

// exact int32 -> FP64 without the conversion unit: 2^52 + 2^31 + y has y + 2^31 in its low word
__device__ __forceinline__ double i32_to_f64(int32_t y) {
    return __hiloint2double(0x43300000, y ^ (int)0x80000000) - 4503601774854144.0;
}

// exact int64 -> FP64 for |y| < 2^51, same trick (the constant only touches the high word)
__device__ __forceinline__ double i64_to_f64(long long y) {
    return __longlong_as_double(y + 0x4338000000000000ll) - 6755399441055744.0;
}

struct I8Args {
    const uint8_t* A8;       // [S][row blocks][128 x kdim core order]
    const uint8_t* B8;       // [S][col blocks][64 x kdim core order]
    const double* row_scale; // 2^(e_r - 7) per row (padded rows: 0)
    const double2* colq;     // per padded column: (2^(f_c - 7), 2^(f_c - 7) * weight); padded columns: 0
    const uint64_t* colx;    // per padded column: col_x | bit 63 | (column not independent ? bit 62 : 0)
    long long n_rows, n_cols, row_lo;     // row_lo: offset of this slab in the MIS row tables / united index
    int n_rb, n_cb, kdim, n_stages;
    int tmem_a;              // 1: leading resident slices are read from tensor memory (QDC_I8_TMEM_A=0 turns it off)
    int n_chunks, chunk_len;  // work item = (row block, chunk of chunk_len column blocks); item = chunk * n_rb + rb
    double out_scale;        // scale * 256^-(S-1)
    const double* d_norm;
    MisTables tab;
    CandSink sink;
    void* P_out;             // optional united slab [n_rows][n_cols], float64 or (out32) float32
    int out32;
    double* partials;
};

// S = number of slices per operand (compile time: every loop over slice pairs and orders unrolls)
template <int S, int TN, int CPT>
__global__ void __launch_bounds__((FIRST_EPI_WARP + 4 * (TN / CPT)) * 32, 1) contract_i8_kernel(const I8Args a) {
    constexpr int N_EPI_WARPS = 4 * (TN / CPT);
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t a_full, a_empty, b_full[MAX_STAGES], b_empty[MAX_STAGES],
        acc_full[MAX_S], acc_empty[MAX_S];
    __shared__ uint32_t tmem_base_sh;
    __shared__ double scratch[3 * MAX_WARPS];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kdim = a.kdim;
    const uint32_t a_slice_bytes = TM * kdim, b_slice_bytes = TN * kdim;
    uint8_t* sA = smem;                                   // [S][a_slice_bytes]
    uint8_t* sB = smem + (size_t)S * a_slice_bytes;       // [n_stages][b_slice_bytes]

    if (tid == 0) {
        mbar_init(&a_full, 1); mbar_init(&a_empty, 1);
        for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
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

    double cta_pwok = 0.0;                                // this thread's share of sum p w [independent]
    const int n_items = a.n_rb * a.n_chunks;

    if (warp == 0) {
        // ================= producer =================
        const bool leader = elect_one();
        int b_stage = 0; uint32_t b_phase = 0, rb_phase = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, rb_phase ^= 1u) {
            const int rb = item % a.n_rb, cb0 = (item / a.n_rb) * a.chunk_len;
            const int cb1 = min(a.n_cb, cb0 + a.chunk_len);
            mbar_wait_relaxed(&a_empty, rb_phase ^ 1u, 200);
            if (leader) {
                mbar_expect_tx(&a_full, (uint32_t)S * a_slice_bytes);
#pragma unroll
                for (int i = 0; i < S; ++i)
                    bulk_g2s(sA + (size_t)i * a_slice_bytes,
                             a.A8 + ((size_t)i * a.n_rb + rb) * a_slice_bytes, a_slice_bytes, &a_full);
            }
            for (int cb = cb0; cb < cb1; ++cb) {
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    mbar_wait_relaxed(&b_empty[b_stage], b_phase ^ 1u, 64);
                    if (leader) {
                        mbar_expect_tx(&b_full[b_stage], b_slice_bytes);
                        bulk_g2s(sB + (size_t)b_stage * b_slice_bytes,
                                 a.B8 + ((size_t)j * a.n_cb + cb) * b_slice_bytes, b_slice_bytes, &b_full[b_stage]);
                    }
                    if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        // The whole warp walks the loops (so that addresses and descriptors live in uniform
        // registers); one elected lane issues the MMAs and the commits.
        const bool leader = elect_one();
        // D = s32, A = B = signed bytes (bits 7, 10), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
        const int ksteps = kdim / 32;
        // descriptor of address 0; a slice at byte address x is desc0 + (x >> 4), and a step of
        // 32 k (two 16-byte chunks = 256 bytes in core-matrix order) adds 16
        const uint64_t desc0 = smem_desc(0, 128, (uint32_t)(kdim / 16) * 128u);
        const uint64_t a_desc0 = desc0 + (smem_u32(sA) >> 4), b_desc0 = desc0 + (smem_u32(sB) >> 4);
        // TMEM columns [S * TN, 512) hold the first n_ta resident slices (kdim / 4 columns each)
        const uint32_t tmem_a0 = tmem + (uint32_t)(S * TN);
        const int n_ta = a.tmem_a ? min(S, (TMEM_COLS - S * TN) / (kdim / 4)) : 0;
        int b_stage = 0; uint32_t b_phase = 0, rb_phase = 0, tile_phase = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, rb_phase ^= 1u) {
            const int cb0 = (item / a.n_rb) * a.chunk_len, cb1 = min(a.n_cb, cb0 + a.chunk_len);
            mbar_wait(&a_full, rb_phase);
            tc_fence_after();
            // The leading slices of the resident operand (the most used ones: slice i enters S - i
            // pairs) go to the TMEM columns the accumulators leave free: an MMA that takes A from
            // tensor memory reads only the 2 KB of its B tile from shared memory instead of 6 KB
            // (the N = 64 instruction is bound by that read).  tcgen05 operations of one thread
            // execute in issue order, so the copies wait for the previous item's MMAs.
            if (leader)
                for (int i = 0; i < n_ta; ++i)
                    for (int ks = 0; ks < ksteps; ++ks)
                        tmem_cp_128x256b(tmem_a0 + (uint32_t)(i * ksteps + ks) * 8u,
                                         a_desc0 + (uint32_t)i * (a_slice_bytes >> 4) + ks * 16);
            for (int cb = cb0; cb < cb1; ++cb, tile_phase ^= 1u) {
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    mbar_wait(&b_full[b_stage], b_phase);
                    tc_fence_after();
                    const uint64_t b_desc = b_desc0 + (uint32_t)b_stage * (b_slice_bytes >> 4);
#pragma unroll
                    for (int i = 0; i + j < S; ++i) {
                        const int d = i + j;
                        if (j == 0) {                              // first touch of order d in this tile
                            mbar_wait(&acc_empty[d], tile_phase ^ 1u);
                            tc_fence_after();
                        }
                        const uint64_t a_desc = a_desc0 + (uint32_t)i * (a_slice_bytes >> 4);
                        const uint32_t tmem_d = tmem + (uint32_t)d * TN;
                        if (leader) {
                            if (i < n_ta) {
                                const uint32_t ta = tmem_a0 + (uint32_t)(i * ksteps) * 8u;
                                umma_i8_ts(tmem_d, ta, b_desc, idesc, j > 0 ? 1u : 0u);
#pragma unroll
                                for (int ks = 1; ks < 8; ++ks)
                                    if (ks < ksteps) umma_i8_ts(tmem_d, ta + ks * 8, b_desc + ks * 16, idesc, 1u);
                            } else {
                                umma_i8(tmem_d, a_desc, b_desc, idesc, j > 0 ? 1u : 0u);
#pragma unroll
                                for (int ks = 1; ks < 8; ++ks)
                                    if (ks < ksteps) umma_i8(tmem_d, a_desc + ks * 16, b_desc + ks * 16, idesc, 1u);
                            }
                            if (i == 0) umma_commit(&acc_full[j]);     // order j is complete
                        }
                    }
                    if (leader) umma_commit(&b_empty[b_stage]);
                    if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                }
            }
            if (leader) umma_commit(&a_empty);
        }
    } else if (warp >= FIRST_EPI_WARP) {
        // ================= epilogue =================
        const int ew = warp - FIRST_EPI_WARP;
        const int quarter = warp & 3;                      // TMEM lanes 32 * (warp % 4) .. + 31
        const int part = ew >> 2;                          // which CPT columns of the 64-column accumulator
        const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(part * CPT);
        uint32_t tile_phase = 0;
        // (few registers here: whatever only the rare paths and the end of an item need -- row
        // factor, row bits, exact threshold -- is recomputed there from the tables)
        auto row_factor = [&](int r_loc) -> double {
            return r_loc < a.n_rows ? a.row_scale[r_loc] * a.out_scale * (a.d_norm ? 1.0 / *a.d_norm : 1.0) : 0.0;
        };
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int rb = item % a.n_rb, cb0 = (item / a.n_rb) * a.chunk_len;
            const int cb1 = min(a.n_cb, cb0 + a.chunk_len);
            const int r_loc = rb * TM + quarter * 32 + lane;                        // row inside the slab
            const bool row_ok = r_loc < a.n_rows;
            const uint32_t rwok = row_ok ? __ldg(a.tab.row_wok + a.row_lo + r_loc) : 0u;
            const uint64_t rf = (row_ok ? __ldg(a.tab.row_forb + a.row_lo + r_loc) : 0ull) | FLAG_COL_BAD |
                                ((rwok >> 8) & 1u ? 0ull : FLAG_ROW_BAD);
            // Candidate filter on the high words only: q = X cs_c with cs_c a power of two, so
            // hi(q) = hi(X) + hi(cs_c) - hi(1.0) for X > 0, and q > thr / rs needs
            // hi(X) + hi(cs_c) >= hi(thr / rs) + hi(1.0)   (all terms non-negative: unsigned, no overflow).
            // Exact test on the rare slow path.
            uint32_t thr_hi = 0xffffffffu;
            {
                const double rs = row_factor(r_loc);
                if (rs > 0.0) thr_hi = (uint32_t)__double2hiint(a.sink.threshold / rs) + 0x3ff00000u;
            }
            double s_ok = 0.0, t_ok = 0.0;              // sum X cs_c and sum X cs_c w_c under the independence mask
            for (int cb = cb0; cb < cb1; ++cb, tile_phase ^= 1u) {
                // X = sum_d D_d 256^(S-1-d), two orders per round.  D_0 256 + D_1 fits int32
                // (|D_0| <= 2^22); later pairs are formed in int64 (|.| < 2^34) and folded in by
                // an FMA; integers enter FP64 through the 2^52 trick, exactly.
                double X[CPT];
#pragma unroll
                for (int d = 0; d < S; d += 2) {
                    constexpr int dummy = 0; (void)dummy;
                    const bool two = d + 1 < S;
                    mbar_wait_relaxed(&acc_full[two ? d + 1 : d], tile_phase, 32);
                    tc_fence_after();
                    if constexpr (CPT == 16) {
                    uint32_t v0[CPT], v1[CPT];
                    tmem_ld(lane_base + (uint32_t)(d * TN), v0);
                    if (two) tmem_ld(lane_base + (uint32_t)((d + 1) * TN), v1);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&acc_empty[d]);
                        if (two) mbar_arrive(&acc_empty[d + 1]);
                    }
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        if (d == 0)
                            X[c] = i32_to_f64(two ? (int32_t)v0[c] * 256 + (int32_t)v1[c] : (int32_t)v0[c]);
                        else if (two)
                            X[c] = fma(X[c], 65536.0, i64_to_f64((long long)(int32_t)v0[c] * 256 + (long long)(int32_t)v1[c]));
                        else
                            X[c] = fma(X[c], 256.0, i32_to_f64((int32_t)v0[c]));
                    }
                    } else {                                       // 16 columns at a time (registers)
#pragma unroll
                    for (int h = 0; h < CPT / 16; ++h) {
                        uint32_t v0[16], v1[16];
                        tmem_ld(lane_base + (uint32_t)(d * TN + h * 16), v0);
                        if (two) tmem_ld(lane_base + (uint32_t)((d + 1) * TN + h * 16), v1);
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                        for (int c = 0; c < 16; ++c) {
                            double& x = X[h * 16 + c];
                            if (d == 0)
                                x = i32_to_f64(two ? (int32_t)v0[c] * 256 + (int32_t)v1[c] : (int32_t)v0[c]);
                            else if (two)
                                x = fma(x, 65536.0, i64_to_f64((long long)(int32_t)v0[c] * 256 + (long long)(int32_t)v1[c]));
                            else
                                x = fma(x, 256.0, i32_to_f64((int32_t)v0[c]));
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&acc_empty[d]);
                        if (two) mbar_arrive(&acc_empty[d + 1]);
                    }
                    }
                }
                const int c0 = cb * TN + part * CPT;
                bool any = false;
#pragma unroll
                for (int c = 0; c < CPT; ++c) {
                    const double2 cq = __ldg(a.colq + c0 + c);
                    const uint64_t cf = __ldg(a.colx + c0 + c);
                    // if ((cf & rf) == 0) { s_ok += X cs_c;  t_ok += X cs_c w_c; }
                    asm("{\n .reg .pred p;\n .reg .b64 t;\n and.b64 t, %3, %4;\n setp.eq.u64 p, t, 0;\n"
                        " @p fma.rn.f64 %0, %2, %5, %0;\n @p fma.rn.f64 %1, %2, %6, %1;\n}\n"
                        : "+d"(s_ok), "+d"(t_ok) : "d"(X[c]), "l"(cf), "l"(rf), "d"(cq.x), "d"(cq.y));
                    any |= (uint32_t)max(__double2hiint(X[c]), 0) + (uint32_t)__double2hiint(cq.x) >= thr_hi;
                    if ((c & 3) == 3) asm volatile("" ::: "memory");   // keeps the table loads of 4 columns in flight, not 16
                }
                if (any) {                                  // rare: some output may be above the candidate threshold
                    const double rs = row_factor(r_loc);
                    const double thr_q = a.sink.threshold / rs;
                    const uint64_t rx = __ldg(a.tab.row_x + a.row_lo + r_loc);
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        // (laundered: nothing of the fast loop is kept alive for this path)
                        const double q = opaque(X[c]) * ((const volatile double*)(a.colq + c0 + c))[0];
                        if (q > thr_q)
                            cand_emit(a.sink, rx | (*(const volatile uint64_t*)(a.colx + c0 + c) & ~(FLAG_ROW_BAD | FLAG_COL_BAD)), q * rs);
                    }
                }
                if (a.P_out && row_ok) {
                    const double rs = row_factor(r_loc);
#pragma unroll
                    for (int c = 0; c < CPT; ++c)
                        if (c0 + c < a.n_cols)
                        {
                            const double p = opaque(X[c]) * ((const volatile double*)(a.colq + c0 + c))[0] * rs;
                            const long long o = (long long)r_loc * a.n_cols + c0 + c;
                            if (a.out32) ((float*)a.P_out)[o] = (float)p;
                            else ((double*)a.P_out)[o] = p;
                        }
                }
            }
            // sum p and sum p w do not need the outputs at all (moments_kernel); here only the masked sum
            cta_pwok += row_factor(r_loc) * fma((double)(rwok & 0xffu), s_ok, t_ok);
        }
    }
    // teardown
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
    MisAccum macc = {0.0, cta_pwok, 0.0};
    mis_block_store(macc, scratch, a.partials);
}

// ---- slicing pre-pass -----------------------------------------------------------------------
// one warp per row: exponent from the row maximum, then S radix-256 digits of every element
// (two's complement: signed leading digit, unsigned fraction digits; rounded to nearest by adding
// half a unit of the last digit before truncating), written in core-matrix order.  Scaling by
// powers of two, floor and the subtractions are exact.
__device__ __forceinline__ size_t core_offset(int row_in_block, int k, int kdim) {
    return ((size_t)(row_in_block >> 3) * (kdim / 16) + (k >> 4)) * 128 + (row_in_block & 7) * 16 + (k & 15);
}

__global__ void __launch_bounds__(256) slice_kernel(const double* __restrict__ X, long long n_rows, int kdim_src,
                                                    int kdim, int block_rows, long long n_blocks, int S,
                                                    uint8_t* __restrict__ out, double* __restrict__ scale_out) {
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
    if (mx * ldexp(1.0, 7 - e) > 127.0) ++e;           // keeps the leading digit inside [-128, 127]
    const double up = ldexp(1.0, 7 - e);               // |x'| <= 127
    // offset = 128 * sum_{i=1..S-1} 256^-i (turns unsigned fraction digits into balanced ones)
    // + half a unit of the last digit (round to nearest)
    double offset = ldexp(1.0, -8 * (S - 1) - 1);
    for (int i = S - 1; i >= 1; --i) offset += ldexp(1.0, 7 - 8 * i);
    if (lane == 0) scale_out[row] = live ? ldexp(1.0, e - 7) : 0.0;
    const long long blk = row / block_rows;
    const int rib = (int)(row - blk * block_rows);
    const size_t slice_stride = (size_t)n_blocks * block_rows * kdim;
    uint8_t* base = out + (size_t)blk * block_rows * kdim;
    for (int k = lane; k < kdim; k += 32) {
        const size_t off = core_offset(rib, k, kdim);
        if (!(live && k < kdim_src)) {                 // padding: exact zeros in every slice
            for (int i = 0; i < S; ++i) base[(size_t)i * slice_stride + off] = 0;
            continue;
        }
        const double t = X[row * kdim_src + k] * up + offset;
        const double s0 = floor(t);                    // in [-127, 127]
        base[off] = (uint8_t)(int8_t)(int)s0;
        double f = t - s0;                             // in [0, 1)
        for (int i = 1; i < S; ++i) {
            f *= 256.0;
            const double u = floor(f);                 // in [0, 255]; the digit is u - 128
            base[(size_t)i * slice_stride + off] = (uint8_t)(int8_t)((int)u - 128);
            f -= u;
        }
    }
}

// per padded column: (2^f_c, 2^f_c * weight) and the flagged mask the epilogue ANDs with the row's
__global__ void __launch_bounds__(256) coltab_kernel(const double* __restrict__ col_scale,
                                                     const uint64_t* __restrict__ col_x,
                                                     const uint32_t* __restrict__ col_wok, long long n_cols,
                                                     long long cols_pad, double2* __restrict__ colq,
                                                     uint64_t* __restrict__ colx) {
    const long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= cols_pad) return;
    if (c < n_cols) {
        const uint32_t wok = col_wok[c];
        const double cs = col_scale[c];
        colq[c] = make_double2(cs, cs * (double)(wok & 0xffu));
        colx[c] = col_x[c] | FLAG_ROW_BAD | ((wok >> 8) & 1u ? 0ull : FLAG_COL_BAD);
    } else {
        colq[c] = make_double2(0.0, 0.0);
        colx[c] = FLAG_ROW_BAD | FLAG_COL_BAD;
    }
}

// ---- sums that need no outputs ----------------------------------------------------------------
// sum_{r,c} P~ = scale sum_k (sum_r A[r,k]) (sum_c Bt[c,k]) and, with per-row / per-column weights,
// sum P~ (w_r + w_c) = scale sum_k (sum_r w_r A[r,k]) (sum_c Bt[c,k]) + (sum_r A[r,k]) (sum_c w_c Bt[c,k]):
// two passes over the FP64 operands instead of two more FMAs per output.  One warp per row, lane
// l owns k = l, l + 32, ...; per-CTA partials are combined in a fixed order (deterministic).
constexpr int MOM_CTAS = 296, MOM_THREADS = 256;

__global__ void __launch_bounds__(MOM_THREADS) moments_kernel(const double* __restrict__ X, long long n_rows, int kdim,
                                                              const uint32_t* __restrict__ wok,
                                                              double* __restrict__ partial /* [gridDim.x][2][kdim] */) {
    __shared__ double red[2][MOM_THREADS / 32][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long gw = (long long)blockIdx.x * (MOM_THREADS / 32) + warp, nw = (long long)gridDim.x * (MOM_THREADS / 32);
    for (int k0 = 0; k0 < kdim; k0 += 32) {
        const int k = k0 + lane;
        double s = 0.0, w = 0.0;
        if (k < kdim)
            for (long long r = gw; r < n_rows; r += nw) {
                const double x = X[r * kdim + k];
                s += x;
                w = fma((double)(__ldg(wok + r) & 0xffu), x, w);
            }
        red[0][warp][lane] = s;
        red[1][warp][lane] = w;
        __syncthreads();
        if (warp < 2 && k < kdim) {
            double t = 0.0;
            for (int i = 0; i < MOM_THREADS / 32; ++i) t += red[warp][i][lane];
            partial[((size_t)blockIdx.x * 2 + warp) * kdim + k] = t;
        }
        __syncthreads();
    }
}

// one CTA: sums[0] += scale / norm * sum_k sA sB, sums[2] += scale / norm * sum_k (wA sB + sA wB)
__global__ void __launch_bounds__(256) moments_finish_kernel(const double* __restrict__ pa, const double* __restrict__ pb,
                                                             int n_part, int kdim, double scale,
                                                             const double* __restrict__ d_norm, double* __restrict__ sums) {
    __shared__ double sp[256], spw[256];
    const int k = threadIdx.x;
    double sA = 0.0, wA = 0.0, sB = 0.0, wB = 0.0;
    if (k < kdim)
        for (int i = 0; i < n_part; ++i) {
            sA += pa[((size_t)i * 2 + 0) * kdim + k];
            wA += pa[((size_t)i * 2 + 1) * kdim + k];
            sB += pb[((size_t)i * 2 + 0) * kdim + k];
            wB += pb[((size_t)i * 2 + 1) * kdim + k];
        }
    sp[k] = sA * sB;
    spw[k] = fma(wA, sB, sA * wB);
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (k < o) { sp[k] += sp[k + o]; spw[k] += spw[k + o]; }
        __syncthreads();
    }
    if (k == 0) {
        const double f = scale * (d_norm ? 1.0 / *d_norm : 1.0);
        sums[0] += f * sp[0];
        sums[2] += f * spw[0];
    }
}

}  // namespace

template <int S, int TN, int CPT>
int launch_i8(const I8Args& a, int grid, size_t smem, cudaStream_t s) {
    QDC_CHECK_CUDA(cudaFuncSetAttribute(contract_i8_kernel<S, TN, CPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    contract_i8_kernel<S, TN, CPT><<<grid, (FIRST_EPI_WARP + 4 * (TN / CPT)) * 32, smem, s>>>(a);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

// tile width: 96 columns where the S accumulators of 96 columns fit the 512 of TMEM and two stages of
// the streamed operand fit beside the resident slices (an N = 96 instruction costs what N = 64 does)
static int i8_tile_n(int n_slices, int kd) {
    const char* m = getenv("QDC_I8_TN");
    if (m && atoi(m) == 64) return 64;
    if (n_slices == 5 && (size_t)n_slices * TM * kd + 2 * (size_t)96 * kd <= 227 * 1024 - 2048) return 96;
    return 64;
}

extern "C" int64_t qdc_contract_i8_workspace_bytes(int64_t n_rows, int64_t n_cols, int kdim, int n_slices) {
    if (n_rows < 0 || n_cols < 0 || kdim < 1 || n_slices < 1) return 0;
    const int64_t kd = (kdim + 31) / 32 * 32;
    const int64_t TN = i8_tile_n(n_slices, (int)kd);
    const int64_t rb = (n_rows + TM - 1) / TM, cb = (n_cols + TN - 1) / TN;
    const int64_t a8 = (rb * TM * kd * n_slices + 255) / 256 * 256, b8 = (cb * TN * kd * n_slices + 255) / 256 * 256;
    return a8 + b8 + rb * TM * 8 + cb * TN * 32 + 2 * 296 * 2 * 256 * 8 + 256;
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
    QDC_REQUIRE(threshold >= 0.0, "candidate threshold must be non-negative");
    QDC_REQUIRE(store_mode == 0 || ((store_mode == 1 || store_mode == 2) && d_P_out),
                "store_mode 0 (none), 1 (united slab, float64) or 2 (united slab, float32)");
    if (n_rows == 0) return QDC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int kd = (kdim + 31) / 32 * 32;
    const int TN = i8_tile_n(n_slices, kd);
    const long long n_rb = (n_rows + TM - 1) / TM, n_cb = (n_cols + TN - 1) / TN;
    QDC_REQUIRE(n_rb < (1ll << 31) && n_cb < (1ll << 31), "too many tiles");
    uint8_t* ws = (uint8_t*)d_workspace;
    QDC_REQUIRE((((uintptr_t)ws) & 255) == 0, "workspace must be 256-byte aligned");
    const size_t a8_bytes = ((size_t)n_rb * TM * kd * n_slices + 255) / 256 * 256;
    const size_t b8_bytes = ((size_t)n_cb * TN * kd * n_slices + 255) / 256 * 256;
    uint8_t* A8 = ws;
    uint8_t* B8 = ws + a8_bytes;
    double* rs = (double*)(ws + a8_bytes + b8_bytes);
    double* cs = rs + n_rb * TM;
    double2* colq = (double2*)(cs + n_cb * TN);
    uint64_t* colx = (uint64_t*)(colq + n_cb * TN);
    double* mom_a = (double*)(colx + n_cb * TN);
    double* mom_b = mom_a + (size_t)MOM_CTAS * 2 * 256;
    {
        const long long warps = n_rb * TM;
        slice_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(d_A + row_lo * kdim, n_rows, kdim, kd, TM, n_rb,
                                                                          n_slices, A8, rs);
        QDC_CHECK_LAUNCH();
        const long long warps_b = n_cb * TN;
        slice_kernel<<<(unsigned)((warps_b * 32 + 255) / 256), 256, 0, s>>>(d_Bt, n_cols, kdim, kd, TN, n_cb, n_slices, B8, cs);
        QDC_CHECK_LAUNCH();
        coltab_kernel<<<(unsigned)((warps_b + 255) / 256), 256, 0, s>>>(cs, d_col_x, d_col_wok, n_cols, warps_b, colq, colx);
        QDC_CHECK_LAUNCH();
    }
    I8Args a;
    a.A8 = A8; a.B8 = B8; a.row_scale = rs; a.colq = colq; a.colx = colx;
    a.n_rows = n_rows; a.n_cols = n_cols; a.row_lo = row_lo;
    a.n_rb = (int)n_rb; a.n_cb = (int)n_cb; a.kdim = kd;
    a.out_scale = ldexp(scale, -8 * (n_slices - 1));   // the row / column factors carry 2^(e - 7)
    a.d_norm = d_norm;
    a.tab = {d_row_x, d_row_forb, d_row_wok, d_col_x, d_col_wok};
    a.sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
    a.P_out = store_mode ? d_P_out : nullptr;
    a.out32 = store_mode == 2;
    a.partials = d_partials;
    const size_t smem_max = 227 * 1024 - 2048, a_res = (size_t)n_slices * TM * kd, b_slice = (size_t)TN * kd;
    QDC_REQUIRE(a_res + 2 * b_slice <= smem_max, "shared memory");
    size_t n_stages = (smem_max - a_res) / b_slice;
    if (n_stages > MAX_STAGES) n_stages = MAX_STAGES;
    { const char* m = getenv("QDC_I8_STAGES"); if (m && atoi(m) >= 2 && (size_t)atoi(m) < n_stages) n_stages = atoi(m); }
    a.n_stages = (int)n_stages;
    const size_t smem = a_res + n_stages * b_slice;
    const int sms = qdc_sm_count();
    // Work items = row blocks x column chunks, dealt round-robin to one CTA per SM.  Splitting the
    // column walk costs one more load of the resident slices per item, so chunks stay >= 64 tiles;
    // among 1, 2, 4, ... chunks take the split that wastes the least of the last wave.
    int n_chunks = 1;
    {
        double best = 1e30;
        for (int c = 1; c <= 64 && (n_cb + c - 1) / c >= (c == 1 ? 1 : 64); c *= 2) {
            const long long items = n_rb * c, waves = (items + sms - 1) / sms;
            const double waste = (double)(waves * sms) / (double)items + 0.002 * c;
            if (waste < best - 1e-9) { best = waste; n_chunks = c; }
        }
    }
    { const char* m = getenv("QDC_I8_TMEM_A"); a.tmem_a = m ? atoi(m) : 1; }
    a.n_chunks = n_chunks;
    a.chunk_len = (int)((n_cb + n_chunks - 1) / n_chunks);
    const long long n_items = n_rb * n_chunks;
    QDC_REQUIRE(n_items < (1ll << 31), "too many work items");
    const int grid = (int)(n_items < sms ? n_items : sms);
    int rc = QDC_ERR_ARG;
    switch (n_slices) {
        case 2: rc = launch_i8<2, 64, 16>(a, grid, smem, s); break;
        case 3: rc = launch_i8<3, 64, 16>(a, grid, smem, s); break;
        case 4: rc = launch_i8<4, 64, 16>(a, grid, smem, s); break;
        case 5: rc = TN == 96 ? launch_i8<5, 96, 16>(a, grid, smem, s) : launch_i8<5, 64, 16>(a, grid, smem, s); break;
        case 6: rc = launch_i8<6, 64, 16>(a, grid, smem, s); break;
    }
    if (rc != QDC_OK) return rc;
    rc = qdc_launch_finalize_sums(d_partials, grid, d_sums, s);     // the masked sum (slots 0, 2 are zero)
    if (rc != QDC_OK) return rc;
    moments_kernel<<<MOM_CTAS, MOM_THREADS, 0, s>>>(d_A + row_lo * kdim, n_rows, kdim, d_row_wok + row_lo, mom_a);
    QDC_CHECK_LAUNCH();
    moments_kernel<<<MOM_CTAS, MOM_THREADS, 0, s>>>(d_Bt, n_cols, kdim, d_col_wok, mom_b);
    QDC_CHECK_LAUNCH();
    moments_finish_kernel<<<1, 256, 0, s>>>(mom_a, mom_b, MOM_CTAS, kdim, scale, d_norm, d_sums);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}
