// K5 (alternative): the contraction on the 5th-generation tensor cores (tcgen05.mma kind::i8,
// accumulators in TMEM) through an error-free split of the FP64 operands into 8-bit digits
// ("Ozaki scheme"), with the MIS scoring of contract.cu fused into the epilogue.
//
// Per row r, 2^e_r is a power-of-two ceiling of the row maximum and x' = x 2^(7 - e_r) lies in
// [-127, 127].  x' (rounded to S radix-256 digits) is written with balanced digits:
//   x' ~ sum_{i=0..S-1} s_i 256^-i ,   s_i integer in [-128, 127]   (signed bytes),
// likewise t_j for the columns of Bt (i8_prep.cu).  Every slice pair is one exact integer GEMM:
//   P~[r, c] = scale * 2^(e_r + f_c - 14) * sum_d 256^-d D_d[r, c],   D_d = sum_{i+j=d} s_i . t_j
// |D_d| <= (d+1) * kdim * 2^14 < 2^25 for kdim <= 256: exact in the int32 accumulators.  What is
// dropped: the operand rounding (<= 2^(-8S) relative to 2^e_r, 1/256 more when S is below the stored
// count) and the slice pairs with i + j >= S; both have random sign (balanced digits), and
//   |err| <= 2^(e_r + f_c) * scale * kdim * (S + 1.1) * 2^(-8S)
//   S = 5: 1.4e-9, S = 6: 6.5e-12  (x 2^(e_r + f_c) * scale, kdim 256); FP64 itself: 2.8e-14.
//
// Structure (one persistent CTA per SM; tile = 128 rows x TN columns; 4 + TN / 4 warps):
//   warp 0  producer: bulk async copies (cp.async.bulk + mbarrier complete_tx).  The S slices of
//           a 128-row block of A (S * 128 * kdim bytes) are loaded once and stay resident while
//           the CTA walks its column blocks; the slices t_j of a TN-column block stream through a
//           ring whose depth is whatever shared memory is left (2..8 stages).
//   warps 1-3  MMA issuers (one lane each): for j (ring stage) for i <= S-1-j: the 128 x TN x 32
//           tcgen05.mma steps of slice pair (i, j) go into accumulator d = i + j (TMEM columns
//           TN d .. TN d + TN - 1) and are issued by the warp that owns order d (owner_of: S = 5 ->
//           {4}, {3, 0}, {2, 1}, five pairs each).  Order d is complete after the first pair of group
//           j = d and is committed to its own mbarrier, so the epilogue drains orders while later
//           pairs still run.  Where TMEM columns are left, the leading resident slices are copied
//           there (tcgen05.cp, by issuer 0) and their pairs take A from tensor memory.  Why three
//           warps: one warp needed ~880 instructions per tile around the 120 MMAs and never waited
//           on anything -- the tensor pipe waited for it (profiles/r2_i8_issue_bound.md).
//   warps 4..  epilogue (4 TMEM lane quarters x TN / 16 column groups of 16), two forms:
//     EPI 0 (nothing stored): ALL-INTEGER.  Scalar FP64 all but stops while the tensor pipe runs
//           int8 MMAs (profiles/r1_mma_dp_contention_probe.log: DFMA 64 -> 0.6 per clk per SM; IMAD is
//           unaffected), so the steady state has no FP64 instruction at all.  Per output and order
//           one IMAD.WIDE: acc_d += D_d[c] * m_c with m_c = [row, column independent] *
//           (w_r + w_c) << lsh_c, formed once per column and tile (one 64-bit AND); the S sums of
//           a thread's 16 columns are combined by a 128-bit Horner step per tile, shifted to the
//           unit of the largest column exponent (rounded; at most half a unit of the last order per
//           16 columns, part of the stated bound) and added to a 128-bit total that enters FP64
//           once per work item.  Candidates for the top-k: a superset filter on the leading
//           order alone (|X - D_0 256^(S-1)| < 2^15.01 256^(S-1)), one integer compare per output;
//           flagged outputs are re-evaluated exactly from the FP64 operands (rare path).
//     EPI 1 (united slab stored): per-output FP64 fold X = fma(X, 65536, pair) and FP64 scoring, as
//           in round 1; the stored value is P~ itself.
// sum p and sum p w never need the outputs: i8_plan_kernel forms them from column sums.
#include "qdc_common.cuh"
#include "i8_layout.cuh"
#include <stdlib.h>
#include <type_traits>

int qdc_i8_launch_plan(const void* d_opA, int64_t n_rows, const void* d_opB, int64_t n_cols, int kdim, int n_stored,
                       int n_slices, double tol, int n_nodes, double scale, int norm_mode, const double* d_norm_in,
                       const double* d_extra, int n_extra, double* d_res, double* d_partials, int n_partials,
                       cudaStream_t s);
int qdc_launch_finish(const double* d_partials, int n_partials, double* d_sums, const uint64_t* d_cand_x,
                      double* d_cand_p, const unsigned long long* d_cand_count, long long cand_cap, int k,
                      uint64_t* d_top_x, double* d_top_p, cudaStream_t stream);

// Timeline tracing (diagnostic builds only: make EXTRA=-DQDC_I8_TRACE): CTA 0 records (event, clock64) for a few
// tiles -- which wait of which warp the tensor pipe's idle time hides behind (scripts/trace_i8.py reads it).
#ifdef QDC_I8_TRACE
__device__ unsigned long long g_i8_trace[8][1024];
__device__ int g_i8_trace_n[8];
#define QDC_TR(who, tile, cond, code)                                                                   \
    if (blockIdx.x == 0 && (tile) >= 200u && (tile) < 204u && (cond)) {                                  \
        const int k_ = g_i8_trace_n[who]++;                                                              \
        if (k_ < 512) { g_i8_trace[who][2 * k_] = (unsigned long long)(code) | ((unsigned long long)(tile) << 32); g_i8_trace[who][2 * k_ + 1] = clock64(); } \
    }
extern "C" int qdc_i8_debug_trace(unsigned long long* h_events, int* h_counts) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(h_events, g_i8_trace, sizeof(g_i8_trace));
    cudaMemcpyFromSymbol(h_counts, g_i8_trace_n, sizeof(g_i8_trace_n));
    int zero[8] = {};
    cudaMemcpyToSymbol(g_i8_trace_n, zero, sizeof(zero));
    return 0;
}
#else
#define QDC_TR(who, tile, cond, code)
#endif

namespace {

constexpr int TM = 128;                       // tile rows
constexpr int cpt_of(int tn) { return tn == 128 ? 32 : 16; }     // columns per epilogue thread
// accumulator slots of TN columns: one per order while S TN <= 512 TMEM columns, else 4 that rotate --
// order 0 is complete after the tile's very first slice pair and drained long before order 4 starts
constexpr int slots_of(int S, int tn) { return S * tn <= 512 ? S : 4; }
constexpr int MAX_S = 6;
constexpr int MAX_STAGES = 8;                 // ring of streamed-operand slices
constexpr int N_ISSUERS = 3;                  // MMA-issuing warps; each owns whole accumulator orders (owner_of)
constexpr int FIRST_EPI_WARP = 1 + N_ISSUERS; // warp 0: producer, warps 1..3: MMA issuers
// Which issuer warp owns order d (all pairs (i, j) with i + j = d): orders by decreasing pair count to the
// least loaded warp.  S = 5: {4}, {3, 0}, {2, 1} = 5 / 5 / 5 pairs; S = 6: {5, 0}, {4, 1}, {3, 2} = 7 / 7 / 7.
constexpr int owner_of(int S, int d) {
    int load[N_ISSUERS] = {};
    int own = 0;
    for (int dd = S - 1; dd >= 0; --dd) {
        int w = 0;
        for (int k = 1; k < N_ISSUERS; ++k)
            if (load[k] < load[w]) w = k;
        load[w] += dd + 1;
        if (dd == d) own = w;
    }
    return own;
}
constexpr int MAX_WARPS = 32;
constexpr int TMEM_COLS = 512;
constexpr int E_ZERO = -(1 << 20);            // as in i8_prep.cu
constexpr int D0_MARGIN = (1 << 15) + (1 << 9);   // |X / 256^(S-1) - D_0| < 2^15.01
constexpr int BAR_A_FULL = 0, BAR_A_EMPTY = 1, BAR_B_FULL = 2, BAR_B_EMPTY = BAR_B_FULL + 8, BAR_ACC_FULL = BAR_B_EMPTY + 8,
              BAR_ACC_EMPTY = BAR_ACC_FULL + 6, BAR_TA_FULL = BAR_ACC_EMPTY + 6, BAR_TA_FREE = BAR_TA_FULL + 1,
              BAR_CT_FULL = BAR_TA_FREE + 1, BAR_CT_EMPTY = BAR_CT_FULL + 4,
              N_BARS = BAR_CT_EMPTY + 4;                                           // (8 = MAX_STAGES, 6 = MAX_S, 4 = CT_STAGES)
// pair form: the column table of a tile (TN entries of 16 bytes + the group exponents) is staged in shared
// memory by the producer, CT_STAGES tiles ahead (see the epilogue)
constexpr int CT_STAGES = 4;
constexpr int ct_stage_bytes(int tn) { return tn * 16 + 16; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// All barriers live in ONE shared array and are addressed as base + 8 k, the base computed once per
// thread: taking the address of a __shared__ variable at every use costs an S2UR SR_CgaCtaId (a
// ~100-clock special-register read) plus address arithmetic each time -- measured at 11 % of the
// MMA-issuing warp's time when every commit did it (profiles/r2_i8_issue_bound.md).
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done;
}
// A barrier that never completes is a protocol bug: trap (the launch fails with an error) instead
// of hanging the device.  try_wait suspends the thread for a while by itself, so the limits below
// are seconds.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try(bar, parity))
        if (++spins > (1u << 26)) __trap();
}
// polling costs shared-memory cycles the tensor core wants for its operands: waiters that are
// not on the critical path back off between polls
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity, uint32_t ns) {
    uint32_t spins = 0;
    while (!mbar_try(bar, parity)) {
        __nanosleep(ns);
        if (++spins > (1u << 25)) __trap();
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// ---- CTA pair (cta_group::2) helpers ------------------------------------------------------------------
// commit that arrives on the barrier at the same shared-memory offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// the shared::cluster address of this CTA's shared-memory offset `addr` in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// Waits on barriers that also receive arrivals from the peer CTA use the plain (CTA-scope) try_wait, as
// the cluster barriers of CUTLASS's 2-SM pipelines do: what the wait guards is read by the tensor core
// (async proxy) from shared memory that was written before the arrival was sent, never by this thread.
// A cluster-scope acquire here was measured at ~800 clocks per wait, five of them on the serial chain of
// the warp that issues the last order (profiles/r2_i8_timeline.md).
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) { mbar_wait(bar, parity); }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
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
// The 8 k steps of one slice pair at kdim = 256, issued from ONE asm block whose only register
// inputs are the same for every pair of a tile (TMEM base, descriptor of slice 0, descriptor of the
// streamed stage, instruction descriptor); slice, order and k step enter as immediates.  A k step of
// 32 advances a descriptor's address field (16-byte units) by 16, a slice by TM kdim / 16; the
// field cannot carry out of its 14 bits (shared memory ends below 2^14 units), so plain 64-bit adds
// on the descriptor make the next one -- one UIADD3.64 each, no high-word copies.  Written this way
// on purpose: when the adds are visible to the compiler it hoists all S x 8 loop-invariant
// descriptors out of the tile loop into VECTOR registers and pays two R2UR per descriptor per use
// (19 issue slots per MMA on the one warp that issues them, profiles/r2_i8_issue_bound.md).
// (The adds form an in-place chain: only one descriptor pair is live at a time.  With independent
// adds ptxas computes all 16 descriptors of a block first, runs out of the 63 uniform registers,
// spills them to vector registers and pays R2UR again.)
#define QDC_SS_STEP " add.s64 da, da, 16;\n add.s64 db, db, 16;\n tcgen05.mma.cta_group::1.kind::i8 [d], da, db, %3, q;\n"
template <int A_OFF, bool ACC>
__device__ __forceinline__ void umma_pair_ss(uint32_t tmem_d, uint64_t a_desc0, uint64_t b_desc, uint32_t idesc) {
    asm volatile(
        "{\n .reg .pred p, q;\n .reg .b32 d;\n .reg .b64 da, db;\n"
        " setp.ne.b32 p, %4, 0;\n setp.eq.b32 q, %3, %3;\n"
        " mov.b32 d, %0;\n add.s64 da, %1, %5;\n mov.b64 db, %2;\n"
        " tcgen05.mma.cta_group::1.kind::i8 [d], da, db, %3, p;\n"
        QDC_SS_STEP QDC_SS_STEP QDC_SS_STEP QDC_SS_STEP QDC_SS_STEP QDC_SS_STEP QDC_SS_STEP
        "}\n"
        ::"r"(tmem_d), "l"(a_desc0), "l"(b_desc), "r"(idesc), "n"(ACC ? 1 : 0), "n"(A_OFF) : "memory");
}
// the same block for a CTA pair: M = 256 (128 rows in each CTA), the streamed tile split between the two
// CTAs' shared memories (TN / 2 columns each, same offset), issued by the leader CTA only
#define QDC_SS2_STEP " add.s64 da, da, 16;\n add.s64 db, db, 16;\n tcgen05.mma.cta_group::2.kind::i8 [d], da, db, %3, q;\n"
template <int A_OFF, bool ACC>
__device__ __forceinline__ void umma_pair_ss2(uint32_t tmem_d, uint64_t a_desc0, uint64_t b_desc, uint32_t idesc) {
    asm volatile(
        "{\n .reg .pred p, q;\n .reg .b32 d;\n .reg .b64 da, db;\n"
        " setp.ne.b32 p, %4, 0;\n setp.eq.b32 q, %3, %3;\n"
        " mov.b32 d, %0;\n add.s64 da, %1, %5;\n mov.b64 db, %2;\n"
        " tcgen05.mma.cta_group::2.kind::i8 [d], da, db, %3, p;\n"
        QDC_SS2_STEP QDC_SS2_STEP QDC_SS2_STEP QDC_SS2_STEP QDC_SS2_STEP QDC_SS2_STEP QDC_SS2_STEP
        "}\n"
        ::"r"(tmem_d), "l"(a_desc0), "l"(b_desc), "r"(idesc), "n"(ACC ? 1 : 0), "n"(A_OFF) : "memory");
}
#define QDC_TS_STEP " add.u32 ta, ta, 8;\n add.s64 db, db, 16;\n tcgen05.mma.cta_group::1.kind::i8 [d], [ta], db, %2, q;\n"
template <bool ACC>
__device__ __forceinline__ void umma_pair_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t b_desc, uint32_t idesc) {
    asm volatile(
        "{\n .reg .pred p, q;\n .reg .b32 d, ta;\n .reg .b64 db;\n"
        " setp.ne.b32 p, %4, 0;\n setp.eq.b32 q, %2, %2;\n"
        " mov.b32 d, %0;\n mov.b32 ta, %3;\n mov.b64 db, %1;\n"
        " tcgen05.mma.cta_group::1.kind::i8 [d], [ta], db, %2, p;\n"
        QDC_TS_STEP QDC_TS_STEP QDC_TS_STEP QDC_TS_STEP QDC_TS_STEP QDC_TS_STEP QDC_TS_STEP
        "}\n"
        ::"r"(tmem_d), "l"(b_desc), "r"(idesc), "r"(tmem_a), "n"(ACC ? 1 : 0) : "memory");
}

// compile-time loop: f(std::integral_constant<int, I>) for I in [I0, N)
template <int I0, int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (I0 < N) {
        f(std::integral_constant<int, I0>{});
        static_for<I0 + 1, N>(f);
    }
}

// 128 rows x 32 bytes of a K-major core-matrix image (the descriptor the MMA would take) -> 8 TMEM columns
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t tmem_dst, uint64_t sdesc) {
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tmem_dst), "l"(sdesc) : "memory");
}

constexpr uint64_t FLAG_ROW_BAD = 1ull << 63, FLAG_COL_BAD = 1ull << 62;   // as in contract.cu

// 16 consecutive TMEM columns of this warp's 32 lanes (one accumulator row per thread)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
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

// signed 128-bit integers as (lo, hi) pairs: everything inline, no helper calls, no local memory
struct I128 { unsigned long long lo; long long hi; };
__device__ __forceinline__ void i128_horner256(I128& y, long long a) {            // y = 256 y + a
    y.hi = (long long)(((unsigned long long)y.hi << 8) | (y.lo >> 56));
    const unsigned long long l8 = y.lo << 8, l = l8 + (unsigned long long)a;
    y.hi += (a >> 63) + (l < l8 ? 1 : 0);
    y.lo = l;
}
__device__ __forceinline__ void i128_add(I128& t, const I128& y) {
    const unsigned long long l = t.lo + y.lo;
    t.hi += y.hi + (l < t.lo ? 1 : 0);
    t.lo = l;
}
// y = round_half_up(y / 2^sh), 0 <= sh <= 120
__device__ __forceinline__ void i128_shr_round(I128& y, int sh) {
    if (sh == 0) return;
    I128 half;
    if (sh <= 64) { half.lo = 1ull << (sh - 1); half.hi = 0; } else { half.lo = 0; half.hi = 1ll << (sh - 65); }
    i128_add(y, half);
    if (sh >= 64) { y.lo = (unsigned long long)(y.hi >> (sh - 64)); y.hi >>= 63; }
    else { y.lo = (y.lo >> sh) | ((unsigned long long)y.hi << (64 - sh)); y.hi >>= sh; }
}

struct I8Args {
    const uint8_t* A8;       // [S stored][rows_pad][kd] core order
    const uint8_t* B8;
    size_t a_slice_stride, b_slice_stride;
    const double* row_scale; // 2^(e_r - 7) per row of the slab (padded rows: 0)
    const double2* colq;     // per padded column: (2^(f_c - 7), 2^(f_c - 7) * weight); padded columns: 0
    const uint4* coltab;     // per padded column: (x_lo, x_hi | flags, weight << lsh, 1 << lsh)
    const int* gref;         // per I8_GROUP columns: reference exponent
    const double* A_src;     // the FP64 operands themselves (exact re-evaluation of flagged candidates)
    const double* B_src;
    int kdim_src;
    long long n_rows, n_cols, row_lo;     // row_lo: offset of this slab in the MIS row tables / united index
    int n_rb, n_cb, kdim, n_stages;       // n_rb: work units along the rows (row blocks; pair form: 256-row units)
    int a_blocks_pad;        // 128-row blocks the prepared row operand holds (pair form: a unit's second half may lie beyond)
    int n_chunks, chunk_len;  // work item = (row block, chunk of chunk_len column blocks); item = chunk * n_rb + rb
    int check_slices;        // 1: run only if res[I8_RES_SLICES] == S (the count was chosen on the device)
    double out_scale;        // scale * 256^-(S-1)
    const double* res;       // result / control block: norm, slice count, largest column exponent
    int use_norm;
    MisTables tab;
    CandSink sink;
    void* P_out;             // optional united slab [n_rows][n_cols], float64 or (out32) float32
    int out32;
    double* partials;
};

// exact evaluation of the flagged outputs of one thread's columns (bit c of `flags` = column c0 + c) from the FP64 operands.
// (Everything by value: a reference to the kernel's parameter struct would force a local-memory
// copy of it, and the hot loop would read its fields through LDL.)
struct FlagCtx {
    const double* A_src; const double* B_src; const uint4* coltab; const uint64_t* row_x;
    long long n_cols; int kdim_src;
    unsigned long long* flagged;     // counter of the deferred list (result block, I8_RES_FLAGGED)
    unsigned long long* lost;        // accepted candidates without a slot (I8_RES_LOST)
};
__device__ __noinline__ void emit_flagged(FlagCtx x, CandSink sink, long long r_tab, long long r_loc, long long c0,
                                          uint32_t flags, double f) {
    // Flagged outputs are NOT evaluated here: scalar FP64 all but stops while the tensor pipe runs (a 256-term
    // dot product took ~18 us, on a warp the next tile's accumulators were waiting for: 4096 candidates cost
    // 75 ms, profiles/r2_peaked_probe.log).  They are appended as (local row, column) to the candidate arrays and
    // evaluated by i8_reeval_kernel after the contraction; only when the arrays are full does the exact test run
    // here, and what passes it then is counted as lost.
    const uint64_t rx = __ldg(x.row_x + r_tab);
    const double* ar = x.A_src + r_loc * x.kdim_src;
    while (flags) {
        const int c = __ffs(flags) - 1;
        flags &= flags - 1;
        if (c0 + c >= x.n_cols) continue;
        {
            const unsigned long long slot = atomicAdd(x.flagged, 1ull);
            if ((long long)slot < sink.cap) {
                sink.x[slot] = ((uint64_t)r_loc << 32) | (uint64_t)(c0 + c);
                continue;
            }
        }
        const double* br = x.B_src + (c0 + c) * x.kdim_src;
        double acc = 0.0;
        for (int k = 0; k < x.kdim_src; ++k) acc = fma(ar[k], br[k], acc);
        const double p = acc * f;
        if (p > sink.threshold) { atomicAdd(sink.count, 1ull); atomicAdd(x.lost, 1ull); }
    }
    (void)rx;
}

// exact evaluation of the deferred list, one warp per entry: p = scale / norm * <A[r], B[c]> in FP64 from the operands
// themselves; entries above the threshold become (united index, p), the others are marked consumed for the top-k
__global__ void __launch_bounds__(256) i8_reeval_kernel(FlagCtx x, CandSink sink, const double* __restrict__ res, int use_norm,
                                                        double scale, long long row_lo) {
    const unsigned long long n_flag = *x.flagged;
    const long long n = (long long)n_flag < sink.cap ? (long long)n_flag : sink.cap;
    const double f = use_norm ? scale / res[I8_RES_NORM] : scale;
    const int lane = threadIdx.x & 31;
    for (long long e = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); e < n; e += (long long)gridDim.x * 8) {
        const uint64_t code = sink.x[e];
        const long long r_loc = (long long)(code >> 32), c = (long long)(code & 0xffffffffull);
        const double* ar = x.A_src + r_loc * x.kdim_src;
        const double* br = x.B_src + c * x.kdim_src;
        double acc = 0.0;
        for (int k = lane; k < x.kdim_src; k += 32) acc = fma(ar[k], br[k], acc);
        acc = warp_sum(acc);
        const double p = acc * f;
        if (lane == 0) {
            if (p > sink.threshold) {
                const uint4 ct = __ldg(x.coltab + c);
                const uint64_t cx = ((uint64_t)ct.y << 32 | ct.x) & ~(FLAG_ROW_BAD | FLAG_COL_BAD);
                sink.x[e] = __ldg(x.row_x + row_lo + r_loc) | cx;
                sink.p[e] = p;
                atomicAdd(sink.count, 1ull);
            } else {
                sink.x[e] = ~0ull;
                sink.p[e] = -1.0 / 0.0;
            }
        }
    }
}

// S = number of slices per operand in use (compile time: every loop over slice pairs and orders unrolls)
// PAIR: two CTAs of a cluster form a CTA pair (cta_group::2): the unit of work is 256 rows (128 in each CTA, both
// resident), every tcgen05.mma is M = 256 issued by the leader CTA (rank 0) for both, and the streamed tile is
// split -- each CTA loads TN / 2 of its columns and the MMA reads both halves -- so a CTA's shared memory
// serves 6 KB instead of 8 KB per 64-clock instruction and takes half of the streamed bytes.  Cross-CTA
// protocol: "full" barriers of the leader count 2 (its own copy's expect_tx + a forwarded arrival from the peer,
// whose idle issuer warp 0 waits on the peer's own barrier and arrives remotely), "empty" and "accumulator
// full" barriers are signalled in both CTAs by multicast commits, the leader's "accumulator empty" barriers
// count the epilogue warps of both CTAs (the peer's arrive remotely).
template <int S, int TN, int EPI, bool PAIR = false>
__global__ void __launch_bounds__((FIRST_EPI_WARP + 4 * (TN / cpt_of(TN))) * 32, 1) contract_i8_kernel(const I8Args a) {
    constexpr int CPT = cpt_of(TN), N_EPI_WARPS = 4 * (TN / CPT), NSLOT = slots_of(S, TN);
    static_assert(NSLOT * TN <= TMEM_COLS && S - NSLOT <= 1, "accumulator slots");
    static_assert(!PAIR || (EPI == 0 && TN % 32 == 0), "the CTA-pair form is the scoring-only kernel");
    // EPI 2: the all-integer scoring of EPI 0 plus the outputs themselves as float32 (CutQC's build): the orders of an
    // output are folded in FP32 (FFMA keeps its rate beside the tensor pipe, scalar FP64 does not) and each thread
    // stores its 16 consecutive floats as four 16-byte vectors
    constexpr int TNL = PAIR ? TN / 2 : TN;            // columns of the streamed tile this CTA loads
    static_assert(EPI == 0 || (CPT == 16 && NSLOT == S), "the storing epilogue keeps 16 columns per thread and one slot per order");
    // order d of this CTA's tile number `seq` is accumulation number g = seq S + d: it lives in slot g % NSLOT
    // and is that slot's use number g / NSLOT (the parity of its two barriers)
    auto slot_of = [](uint32_t seq, int d) -> uint32_t { return NSLOT == S ? (uint32_t)d : (seq * S + d) % NSLOT; };
    auto par_of = [](uint32_t seq, int d) -> uint32_t { return NSLOT == S ? (seq & 1u) : (((seq * S + d) / NSLOT) & 1u); };
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[N_BARS];
    uint32_t bar0 = smem_u32(bars);
    asm volatile("" : "+r"(bar0));                     // opaque: never re-derived from the symbol (S2UR + arithmetic per use)
    const uint32_t a_full = bar0 + 8u * BAR_A_FULL, a_empty = bar0 + 8u * BAR_A_EMPTY;
    const uint32_t b_full0 = bar0 + 8u * BAR_B_FULL, b_empty0 = bar0 + 8u * BAR_B_EMPTY;
    const uint32_t acc_full0 = bar0 + 8u * BAR_ACC_FULL, acc_empty0 = bar0 + 8u * BAR_ACC_EMPTY;
    __shared__ uint32_t tmem_base_sh;
    __shared__ double scratch[3 * MAX_WARPS];

    if (a.check_slices && (int)a.res[I8_RES_SLICES] != S) return;     // another instantiation owns this call

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kdim = a.kdim;
    const uint32_t a_slice_bytes = TM * kdim, b_slice_bytes = TNL * kdim;
    const uint32_t crank = PAIR ? cluster_rank() : 0u;                        // 0 = leader of the pair
    const int unit0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;       // first work item of this CTA (pair)
    const int unit_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    uint32_t sA_addr = smem_u32(smem);                    // [S][a_slice_bytes], then [n_stages][b_slice_bytes]
    asm volatile("" : "+r"(sA_addr));
    const uint32_t sB_addr = sA_addr + (uint32_t)S * a_slice_bytes;
    const uint32_t sCT_addr = sB_addr + (uint32_t)a.n_stages * b_slice_bytes;   // (pair form) CT_STAGES column-table stages
    const uint32_t ct_full0 = bar0 + 8u * BAR_CT_FULL, ct_empty0 = bar0 + 8u * BAR_CT_EMPTY;
    constexpr uint32_t CTB = ct_stage_bytes(TN);

    if (tid == 0) {
        const int full_count = (PAIR && crank == 0) ? 2 : 1;                  // leader: own copy + the peer's forwarded arrival
        mbar_init(a_full, full_count); mbar_init(a_empty, N_ISSUERS);
        mbar_init(bar0 + 8u * BAR_TA_FULL, 1); mbar_init(bar0 + 8u * BAR_TA_FREE, N_ISSUERS - 1);
        for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(b_full0 + 8u * s, full_count); mbar_init(b_empty0 + 8u * s, N_ISSUERS); }
        for (int d = 0; d < MAX_S; ++d) { mbar_init(acc_full0 + 8u * d, 1); mbar_init(acc_empty0 + 8u * d, (PAIR ? 2 : 1) * N_EPI_WARPS); }
        for (int c = 0; c < CT_STAGES; ++c) { mbar_init(ct_full0 + 8u * c, 1); mbar_init(ct_empty0 + 8u * c, N_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        if constexpr (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sh)), "r"(TMEM_COLS));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sh)), "r"(TMEM_COLS));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
    }
    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();           // (pair: the peer's barriers are initialised too)
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, tmem_base_sh, 0);     // (lane-0 broadcast: known warp-uniform)

    double cta_pwok = 0.0;                                // this thread's share of sum p w [independent]
    const int n_items = a.n_rb * a.n_chunks;

    if (warp == 0) {
        // ================= producer =================
        const bool leader = elect_one();
        int b_stage = 0; uint32_t b_phase = 0, rb_phase = 0, p_seq = 0;
        (void)p_seq;
        for (int item = unit0; item < n_items; item += unit_step, rb_phase ^= 1u) {
            const int rb = PAIR ? 2 * (item % a.n_rb) + (int)crank : item % a.n_rb, cb0 = (item / a.n_rb) * a.chunk_len;
            const int cb1 = min(a.n_cb, cb0 + a.chunk_len);
            mbar_wait_relaxed(a_empty, rb_phase ^ 1u, 200);
            if (leader) {
                // (pair: the second half of the last unit may lie beyond the padded operand: nothing is loaded,
                // its MMA rows are garbage that the epilogue never looks at)
                const bool have = !PAIR || rb < a.a_blocks_pad;
                mbar_expect_tx(a_full, have ? (uint32_t)S * a_slice_bytes : 0u);
#pragma unroll
                for (int i = 0; i < S; ++i)
                    if (have)
                        bulk_g2s(sA_addr + (uint32_t)i * a_slice_bytes,
                                 a.A8 + (size_t)i * a.a_slice_stride + (size_t)rb * a_slice_bytes, a_slice_bytes, a_full);
            }
            for (int cb = cb0; cb < cb1; ++cb, ++p_seq) {
                if constexpr (PAIR) {
                    const uint32_t cs = p_seq & (CT_STAGES - 1), cpar = (p_seq / CT_STAGES) & 1u;
                    mbar_wait_relaxed(ct_empty0 + 8u * cs, cpar ^ 1u, 64);
                    if (leader) {
                        mbar_expect_tx(ct_full0 + 8u * cs, CTB);
                        bulk_g2s(sCT_addr + cs * CTB, a.coltab + (size_t)cb * TN, TN * 16, ct_full0 + 8u * cs);
                        bulk_g2s(sCT_addr + cs * CTB + TN * 16, a.gref + (size_t)cb * (TN / I8_GROUP), 16, ct_full0 + 8u * cs);
                    }
                }
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    mbar_wait_relaxed(b_empty0 + 8u * b_stage, b_phase ^ 1u, 64);
                    QDC_TR(0, p_seq, leader, 0x100 + j);          // stage free, copy of slice j issued now
                    if (leader) {
                        mbar_expect_tx(b_full0 + 8u * b_stage, b_slice_bytes);
                        bulk_g2s(sB_addr + (uint32_t)b_stage * b_slice_bytes,
                                 a.B8 + (size_t)j * a.b_slice_stride + ((size_t)cb * (PAIR ? 2 : 1) + crank) * b_slice_bytes,
                                 b_slice_bytes, b_full0 + 8u * b_stage);
                    }
                    if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                }
            }
        }
    } else if (warp < FIRST_EPI_WARP) {
        // ================= MMA issuers =================
        // N_ISSUERS warps walk the same loops; issuer `iw` issues the slice pairs of the orders it owns
        // (MMAs into different accumulators are independent, so the warps need no ordering among
        // themselves) and EVERY issuer commits every ring stage and the end of every work item, so
        // those barriers count N_ISSUERS arrivals whatever the pair split.  One issuing warp was the
        // bottleneck: ~880 instructions per tile around 120 MMAs at ~10 clocks each
        // (profiles/r2_i8_issue_bound.md).  In each warp all lanes walk the loops (addresses and
        // descriptors stay warp-uniform); one elected lane issues the MMAs and the commits.
        const int iw = warp - 1;
        const bool leader = elect_one();
        // D = s32, A = B = signed bytes (bits 7, 10), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
        const int ksteps = kdim / 32;
        // descriptor of address 0; a slice at byte address x is desc0 + (x >> 4), and a step of
        // 32 k (two 16-byte chunks = 256 bytes in core-matrix order) adds 16
        const uint64_t desc0 = smem_desc(0, 128, (uint32_t)(kdim / 16) * 128u);
        const uint64_t a_desc0 = desc0 + (sA_addr >> 4), b_desc0 = desc0 + (sB_addr >> 4);
        const uint32_t ta_full = bar0 + 8u * BAR_TA_FULL, ta_free = bar0 + 8u * BAR_TA_FREE;
        int b_stage = 0; uint32_t b_phase = 0, rb_phase = 0, seq = 0;          // seq: tiles this CTA has walked
        if constexpr (PAIR) {
            // ---- CTA pair, kdim = 256 (the launcher guarantees it) ----
            constexpr uint32_t A_SLICE16 = TM * 256 / 16, B_SLICE16 = TNL * 256 / 16;
            if (crank != 0) {
                // peer CTA: issues nothing.  Its issuer warp 0 forwards "my copy has landed" to the leader's barriers.
                if (iw == 0) {
                    const uint32_t r_a_full = mapa(a_full, 0), r_b_full0 = mapa(b_full0, 0);
                    for (int item = unit0; item < n_items; item += unit_step, rb_phase ^= 1u) {
                        const int cb0 = (item / a.n_rb) * a.chunk_len, cb1 = min(a.n_cb, cb0 + a.chunk_len);
                        mbar_wait(a_full, rb_phase);
                        if (leader) mbar_arrive_remote(r_a_full);
                        for (int cb = cb0; cb < cb1; ++cb)
                            for (int j = 0; j < S; ++j) {
                                mbar_wait(b_full0 + 8u * b_stage, b_phase);
                                if (leader) mbar_arrive_remote(r_b_full0 + 8u * b_stage);
                                if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                            }
                    }
                }
            } else {
                const uint32_t idesc2 = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)((2 * TM) >> 4) << 24);
                for (int item = unit0; item < n_items; item += unit_step, rb_phase ^= 1u) {
                    const int cb0 = (item / a.n_rb) * a.chunk_len, cb1 = min(a.n_cb, cb0 + a.chunk_len);
                    mbar_wait_cluster(a_full, rb_phase);
                    tc_fence_after();
                    for (int cb = cb0; cb < cb1; ++cb, ++seq) {
                        static_for<0, S>([&](auto jc) {
                            constexpr int j = decltype(jc)::value;
                            mbar_wait_cluster(b_full0 + 8u * b_stage, b_phase);
                            tc_fence_after();
                            QDC_TR(1 + iw, seq, leader, 0x200 + j);       // slice j is there
                            const uint64_t b_desc = b_desc0 + (uint32_t)b_stage * B_SLICE16;
                            static_for<0, S - j>([&](auto ic) {
                                constexpr int i = decltype(ic)::value, d = i + j;
                                if (iw == owner_of(S, d)) {
                                    const uint32_t slot = slot_of(seq, d);
                                    if (j == 0) {                      // first touch of order d in this tile: both epilogues are done with the slot
                                        mbar_wait_cluster(acc_empty0 + 8u * slot, par_of(seq, d) ^ 1u);
                                        tc_fence_after();
                                        QDC_TR(1 + iw, seq, leader, 0x300 + d);   // the slot of order d is free
                                    }
                                    if (leader) {
                                        umma_pair_ss2<i * (int)A_SLICE16, (j > 0)>(tmem + slot * TN, a_desc0, b_desc, idesc2);
                                        QDC_TR(1 + iw, seq, true, 0x400 + i * 16 + j);   // pair (i, j) issued
                                        if (i == 0) umma_commit_pair(acc_full0 + 8u * slot);
                                    }
                                }
                            });
                            if (leader) umma_commit_pair(b_empty0 + 8u * b_stage);
                            if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                        });
                    }
                    if (leader) umma_commit_pair(a_empty);
                }
            }
        } else if (ksteps == 8) {
            // ---- kdim = 256: everything but the ring stage is a compile-time constant ----
            // TMEM columns [S TN, 512) hold the first NTA resident slices (64 columns each); their pairs
            // take A from tensor memory and read only the streamed tile from shared memory
            constexpr int NTA = (TMEM_COLS - NSLOT * TN) / 64 < S ? (TMEM_COLS - NSLOT * TN) / 64 : S;
            constexpr uint32_t A_SLICE16 = TM * 256 / 16, B_SLICE16 = TN * 256 / 16;
            for (int item = unit0; item < n_items; item += unit_step, rb_phase ^= 1u) {
                const int cb0 = (item / a.n_rb) * a.chunk_len, cb1 = min(a.n_cb, cb0 + a.chunk_len);
                mbar_wait(a_full, rb_phase);
                tc_fence_after();
                if constexpr (NTA > 0) {
                    if (iw == 0) {
                        // issuer 0 copies the slices once the other issuers' MMAs of the previous item are
                        // done with them (its own tcgen05 operations execute in issue order anyway)
                        mbar_wait(ta_free, rb_phase ^ 1u);
                        tc_fence_after();
                        if (leader) {
                            for (int i = 0; i < NTA; ++i)
                                for (int ks = 0; ks < 8; ++ks)
                                    tmem_cp_128x256b(tmem + (uint32_t)(NSLOT * TN + (i * 8 + ks) * 8), a_desc0 + (uint32_t)i * A_SLICE16 + ks * 16);
                            umma_commit(ta_full);
                        }
                    } else {
                        mbar_wait(ta_full, rb_phase);
                        tc_fence_after();
                    }
                }
                for (int cb = cb0; cb < cb1; ++cb, ++seq) {
                    static_for<0, S>([&](auto jc) {
                        constexpr int j = decltype(jc)::value;
                        mbar_wait(b_full0 + 8u * b_stage, b_phase);
                        tc_fence_after();
                        QDC_TR(1 + iw, seq, leader, 0x200 + j);
                        const uint64_t b_desc = b_desc0 + (uint32_t)b_stage * B_SLICE16;
                        static_for<0, S - j>([&](auto ic) {
                            constexpr int i = decltype(ic)::value, d = i + j;
                            if (iw == owner_of(S, d)) {
                                const uint32_t slot = slot_of(seq, d);
                                if (j == 0) {                      // first touch of order d in this tile
                                    mbar_wait(acc_empty0 + 8u * slot, par_of(seq, d) ^ 1u);
                                    tc_fence_after();
                                    QDC_TR(1 + iw, seq, leader, 0x300 + d);
                                }
                                QDC_TR(1 + iw, seq, leader, 0x400 + i * 16 + j);
                                if (leader) {
                                    if constexpr (i < NTA) umma_pair_ts<(j > 0)>(tmem + slot * TN, tmem + (uint32_t)(NSLOT * TN + i * 64), b_desc, idesc);
                                    else umma_pair_ss<i * (int)A_SLICE16, (j > 0)>(tmem + slot * TN, a_desc0, b_desc, idesc);
                                    if (i == 0) umma_commit(acc_full0 + 8u * slot);   // order j is complete (d = j here)
                                }
                            }
                        });
                        if (leader) umma_commit(b_empty0 + 8u * b_stage);
                        if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                    });
                }
                if (leader) {
                    umma_commit(a_empty);
                    if (NTA > 0 && iw != 0) umma_commit(ta_free);
                }
            }
        } else {
            // ---- any kdim <= 256 (multiple of 32): the same walk with run-time step counts, all pairs
            // issued by issuer 0 (the others only keep the barrier counts) ----
            const uint32_t tmem_a0 = tmem + (uint32_t)(NSLOT * TN);
            const int n_ta = min(S, (TMEM_COLS - NSLOT * TN) / (kdim / 4));
            const bool mine = iw == 0;
            for (int item = unit0; item < n_items; item += unit_step, rb_phase ^= 1u) {
                const int cb0 = (item / a.n_rb) * a.chunk_len, cb1 = min(a.n_cb, cb0 + a.chunk_len);
                mbar_wait(a_full, rb_phase);
                tc_fence_after();
                if (leader && mine)
                    for (int i = 0; i < n_ta; ++i)
                        for (int ks = 0; ks < ksteps; ++ks)
                            tmem_cp_128x256b(tmem_a0 + (uint32_t)(i * ksteps + ks) * 8u,
                                             a_desc0 + (uint32_t)i * (a_slice_bytes >> 4) + ks * 16);
                for (int cb = cb0; cb < cb1; ++cb, ++seq) {
                    for (int j = 0; j < S; ++j) {
                        mbar_wait(b_full0 + 8u * b_stage, b_phase);
                        tc_fence_after();
                        const uint64_t b_desc = b_desc0 + (uint32_t)b_stage * (b_slice_bytes >> 4);
                        for (int i = 0; mine && i + j < S; ++i) {
                            const int d = i + j;
                            const uint32_t slot = slot_of(seq, d);
                            if (j == 0) {
                                mbar_wait(acc_empty0 + 8u * slot, par_of(seq, d) ^ 1u);
                                tc_fence_after();
                            }
                            const uint64_t a_desc = a_desc0 + (uint32_t)i * (a_slice_bytes >> 4);
                            const uint32_t tmem_d = tmem + slot * TN;
                            if (leader) {
                                if (i < n_ta) {
                                    const uint32_t ta = tmem_a0 + (uint32_t)(i * ksteps) * 8u;
                                    for (int ks = 0; ks < ksteps; ++ks)
                                        umma_i8_ts(tmem_d, ta + ks * 8, b_desc + ks * 16, idesc, (j > 0 || ks > 0) ? 1u : 0u);
                                } else {
                                    for (int ks = 0; ks < ksteps; ++ks)
                                        umma_i8(tmem_d, a_desc + ks * 16, b_desc + ks * 16, idesc, (j > 0 || ks > 0) ? 1u : 0u);
                                }
                                if (i == 0) umma_commit(acc_full0 + 8u * slot);
                            }
                        }
                        if (leader) umma_commit(b_empty0 + 8u * b_stage);
                        if (++b_stage == a.n_stages) { b_stage = 0; b_phase ^= 1u; }
                    }
                }
                if (leader) umma_commit(a_empty);
            }
        }
    } else if (warp >= FIRST_EPI_WARP) {
        // ================= epilogue =================
        const int ew = warp - FIRST_EPI_WARP;
        const int quarter = warp & 3;                      // TMEM lanes 32 * (warp % 4) .. + 31
        const int part = ew >> 2;                          // which CPT columns of the TN-column accumulator
        const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(part * CPT);
        uint32_t tile_phase = 0, seq = 0;
        const double inv_norm = a.use_norm ? 1.0 / a.res[I8_RES_NORM] : 1.0;
        // (pair: the leader's issuers wait for the epilogues of both CTAs; the peer's release is a remote arrival)
        const uint32_t acc_empty_rel = (PAIR && crank != 0) ? mapa(acc_empty0, 0) : acc_empty0;
        for (int item = unit0; item < n_items; item += unit_step) {
            const int rb = PAIR ? 2 * (item % a.n_rb) + (int)crank : item % a.n_rb, cb0 = (item / a.n_rb) * a.chunk_len;
            const int cb1 = min(a.n_cb, cb0 + a.chunk_len);
            const int r_loc = rb * TM + quarter * 32 + lane;                        // row inside the slab
            const bool row_ok = r_loc < a.n_rows;
            const uint32_t rwok = row_ok ? __ldg(a.tab.row_wok + a.row_lo + r_loc) : 0u;
            const uint64_t rf = (row_ok ? __ldg(a.tab.row_forb + a.row_lo + r_loc) : 0ull) | FLAG_COL_BAD |
                                ((rwok >> 8) & 1u ? 0ull : FLAG_ROW_BAD);
            const double rs = row_ok ? a.row_scale[r_loc] * a.out_scale * inv_norm : 0.0;   // row factor
            if constexpr (EPI != 1) {
                // ---------- all-integer masked sums ----------
                const int w_r = (int)(rwok & 0xffu);
                const int emax = (int)a.res[I8_RES_EMAX];
                // candidate filter: p ~ D_0 256^(S-1) 2^(f_c - 7) rs > threshold needs
                // D_0 >= T 2^-f_c - margin with T = threshold 2^7 / (256^(S-1) rs) = tm 2^tq, tm in [2^30, 2^31);
                // f_c <= gref of the column group
                int tm = 0x7fffffff, tq = 1 << 24;
                if (rs > 0.0 && a.sink.threshold * 128.0 / rs < 1e300) {               // (an infinite threshold: nothing is ever a candidate)
                    int ex;
                    const double fr = frexp(a.sink.threshold * 128.0 / rs, &ex);      // (rs carries 256^-(S-1): T = this / 256^(S-1))
                    if (fr > 0.0) { tm = (int)(fr * 2147483648.0); tq = ex - 31 - 8 * (S - 1); }
                    else { tm = 0; tq = -(1 << 24); }                                // threshold 0: every output is a candidate
                }
                // Per order d one persistent 64-bit sum run[d] += D_d[c] m_c (the IMAD.WIDE chain accumulates
                // straight into it).  Columns of one 16-group share the unit 2^(gref - 20 - 7); while the
                // thread's next tile has the same gref nothing else happens.  When it changes (or after 256
                // columns: |run| < 256 x 2^25 x 2^29 < 2^63) the orders are folded by a
                // 128-bit Horner step, shifted to the unit of the largest column exponent and added to the
                // 128-bit total, which enters FP64 once per work item.
                I128 total = {0ull, 0ll};
                long long run[S];
#pragma unroll
                for (int d = 0; d < S; ++d) run[d] = 0;
                int cur_gr = E_ZERO, n_run = 0;
                auto flush = [&]() {
                    I128 y = {(unsigned long long)run[0], run[0] >> 63};
#pragma unroll
                    for (int d = 1; d < S; ++d) i128_horner256(y, run[d]);
                    int sh = emax - cur_gr;
                    sh = sh < 0 ? 0 : (sh > 120 ? 120 : sh);
                    i128_shr_round(y, sh);
                    i128_add(total, y);
#pragma unroll
                    for (int d = 0; d < S; ++d) run[d] = 0;
                    n_run = 0;
                };
                int gr_next = __ldg(a.gref + (cb0 * TN + part * CPT) / I8_GROUP);
                for (int cb = cb0; cb < cb1; ++cb, ++seq) {
                    const int c0 = cb * TN + part * CPT;
                    QDC_TR(4 + (ew == 0 ? 0 : 1), seq, lane == 0 && (ew == 0 || ew == N_EPI_WARPS - 1), 0x700);   // tile begins (previous one finished)
                    // the next tile's column-table lines (CPT x 16 bytes) and group exponent go to L1 now: the
                    // table is walked once per row block, so every tile would otherwise start with a chain of
                    // L2 round trips before its first order can be drained
                    // (the group exponent is fetched one tile ahead: a load whose result is needed at once costs an
                    // L2 round trip per tile on the warp's serial chain -- 1.8 % of all stall samples, r2_i8_timeline.md)
                    int gr;
                    uint32_t ct_base = 0;
                    if constexpr (PAIR) {
                        // pair form: this tile's column table was staged in shared memory by the producer (the L1 that
                        // 224 KB of shared memory leave cannot hold it next to the spill stack: every tile began with
                        // L2 round trips on the serial chain of all 16 epilogue warps)
                        const uint32_t cs = seq & (CT_STAGES - 1);
                        mbar_wait(ct_full0 + 8u * cs, (seq / CT_STAGES) & 1u);
                        ct_base = sCT_addr + cs * CTB;
                        asm volatile("ld.shared.s32 %0, [%1];" : "=r"(gr) : "r"(ct_base + TN * 16 + (uint32_t)(part * CPT / I8_GROUP) * 4u));
                    } else {
                        gr = gr_next;
                        if (cb + 1 < cb1) gr_next = __ldg(a.gref + (c0 + TN) / I8_GROUP);
                    }
                    if (gr != cur_gr || n_run == 256 / CPT) { flush(); cur_gr = gr; } // warp-uniform
                    ++n_run;
                    int tcmp;                               // an output may be a candidate when D_0 >= tcmp
                    {
                        const int sh = gr - tq;             // T 2^-gref = tm >> sh
                        const int tg = (gr == E_ZERO || sh < 0) ? 0x7fffffff : (sh > 31 ? 0 : (tm >> sh));
                        tcmp = max(tg, (int)0x80000000 + D0_MARGIN) - D0_MARGIN;
                    }
                    int m[CPT];                             // [independent] (w_r + w_c) << lsh_c
                    float fo[EPI == 2 ? CPT : 1];           // EPI 2: the output itself, folded over the orders in FP32
                    float fsc[EPI == 2 ? CPT : 1];          //        and its column factor 2^(f_c - 7) = 2^(gref - 27) (1 << lsh_c)
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        uint4 ct;                                       // (x_lo, x_hi | flags, w_c << lsh_c, 1 << lsh_c)
                        if constexpr (PAIR)
                            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(ct.x), "=r"(ct.y), "=r"(ct.z), "=r"(ct.w)
                                         : "r"(ct_base + (uint32_t)(part * CPT + c) * 16u));
                        else
                            ct = __ldg(a.coltab + c0 + c);
                        const bool ok = ((ct.x & (uint32_t)rf) | (ct.y & (uint32_t)(rf >> 32))) == 0u;
                        m[c] = ok ? w_r * (int)ct.w + (int)ct.z : 0;
                        if constexpr (EPI == 2) fsc[c] = gr == E_ZERO ? 0.0f : ldexpf((float)ct.w, gr - I8_LSH_MAX - 7);
                    }
                    if constexpr (PAIR) {                               // the stage may be refilled
                        __syncwarp();
                        if (lane == 0) mbar_arrive(ct_empty0 + 8u * (seq & (CT_STAGES - 1)));
                    }
                    uint32_t flags = 0;
                    QDC_TR(4 + (ew == 0 ? 0 : 1), seq, lane == 0 && (ew == 0 || ew == N_EPI_WARPS - 1), 0x701 + (m[0] & 0));   // column factors ready
#pragma unroll
                    for (int d = 0; d < S; ++d) {
                        const uint32_t slot = slot_of(seq, d);
                        mbar_wait_relaxed(acc_full0 + 8u * slot, par_of(seq, d), 32);
                        tc_fence_after();
                        QDC_TR(4 + (ew == 0 ? 0 : 1), seq, lane == 0 && (ew == 0 || ew == N_EPI_WARPS - 1), 0x500 + d);   // order d seen complete
                        long long r = run[d];
#pragma unroll
                        for (int h = 0; h < CPT / 16; ++h) {            // 16 columns at a time (registers)
                            uint32_t v[16];
                            tmem_ld16(lane_base + slot * TN + (uint32_t)(h * 16), v);
                            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                            if (h == CPT / 16 - 1) {                    // everything of this order is in registers: release the slot
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) {
                                    if (PAIR && crank != 0) mbar_arrive_remote(acc_empty_rel + 8u * slot);
                                    else mbar_arrive(acc_empty0 + 8u * slot);
                                }
                                QDC_TR(4 + (ew == 0 ? 0 : 1), seq, lane == 0 && (ew == 0 || ew == N_EPI_WARPS - 1), 0x600 + d);   // slot released
                            }
#pragma unroll
                            for (int c = 0; c < 16; ++c) r += (long long)(int32_t)v[c] * (long long)m[h * 16 + c];
                            if constexpr (EPI == 2) {
#pragma unroll
                                for (int c = 0; c < 16; ++c)
                                    fo[h * 16 + c] = d == 0 ? (float)(int32_t)v[c] : fmaf(fo[h * 16 + c], 256.0f, (float)(int32_t)v[c]);
                            }
                            if (d == 0) {                   // candidate filter: one max tree, compares only on a hit
                                int mx = (int32_t)v[0];
#pragma unroll
                                for (int c = 1; c < 16; ++c) mx = max(mx, (int32_t)v[c]);
                                if (mx >= tcmp) {
#pragma unroll
                                    for (int c = 0; c < 16; ++c)
                                        if ((int32_t)v[c] >= tcmp) flags |= 1u << (h * 16 + c);
                                }
                            }
                        }
                        run[d] = r;
                        // the next tile's column-table lines (CPT x 16 bytes) go to L1 while the last orders are still
                        // running: issued a whole tile ahead they were evicted again before use (L1 is what the 224 KB
                        // of shared memory leave)
                        if (!PAIR && d == (S > 2 ? S - 3 : 0) && cb + 1 < cb1 && lane < CPT / 8)
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.coltab + c0 + TN + lane * 8));
                    }
                    if constexpr (EPI == 2) {
                        // P~[r, c] = (sum_d D_d 256^(S-1-d)) * 2^(f_c - 7) * rs, as float32
                        if (row_ok) {
                            const float rsf = (float)rs;
                            float* out = reinterpret_cast<float*>(a.P_out) + (long long)r_loc * a.n_cols + c0;
                            if ((a.n_cols & 3) == 0 && c0 + CPT <= a.n_cols) {
#pragma unroll
                                for (int c = 0; c < CPT; c += 4)
                                    __stcs(reinterpret_cast<float4*>(out + c),
                                           make_float4(fo[c] * rsf * fsc[c], fo[c + 1] * rsf * fsc[c + 1],
                                                       fo[c + 2] * rsf * fsc[c + 2], fo[c + 3] * rsf * fsc[c + 3]));
                            } else {
#pragma unroll
                                for (int c = 0; c < CPT; ++c)
                                    if (c0 + c < a.n_cols) out[c] = fo[c] * rsf * fsc[c];
                            }
                        }
                    }
                    if (flags && row_ok) {
                        const FlagCtx fx = {a.A_src, a.B_src, a.coltab, a.tab.row_x, a.n_cols, a.kdim_src,
                                            reinterpret_cast<unsigned long long*>(const_cast<double*>(a.res)) + I8_RES_FLAGGED,
                                            reinterpret_cast<unsigned long long*>(const_cast<double*>(a.res)) + I8_RES_LOST};
                        emit_flagged(fx, a.sink, a.row_lo + r_loc, r_loc, c0, flags, a.out_scale * inv_norm * ldexp(1.0, 8 * (S - 1)));
                    }
                }
                flush();
                // once per work item: the 128-bit total enters FP64 (unit 2^(emax - lsh_max - 7) rs)
                const double tot = (double)total.hi * 18446744073709551616.0 + (double)total.lo;
                cta_pwok += rs * ldexp(tot, emax - I8_LSH_MAX - 7);
            } else {
                // ---------- per-output FP64 (the slab is stored) ----------
                // Candidate filter on the high words only: q = X cs_c with cs_c a power of two, so
                // hi(q) = hi(X) + hi(cs_c) - hi(1.0) for X > 0, and q > thr / rs needs
                // hi(X) + hi(cs_c) >= hi(thr / rs) + hi(1.0).  Exact test on the rare slow path.
                uint32_t thr_hi = 0xffffffffu;
                if (rs > 0.0) thr_hi = (uint32_t)__double2hiint(a.sink.threshold / rs) + 0x3ff00000u;
                double s_ok = 0.0, t_ok = 0.0;              // sum X cs_c and sum X cs_c w_c under the independence mask
                for (int cb = cb0; cb < cb1; ++cb, tile_phase ^= 1u) {
                    // X = sum_d D_d 256^(S-1-d), two orders per round.  D_0 256 + D_1 fits int32
                    // (|D_0| <= 2^22); later pairs are formed in int64 (|.| < 2^34) and folded in by
                    // an FMA; integers enter FP64 through the 2^52 trick, exactly.
                    double X[CPT];
#pragma unroll
                    for (int d = 0; d < S; d += 2) {
                        const bool two = d + 1 < S;
                        mbar_wait_relaxed(acc_full0 + 8u * (two ? d + 1 : d), tile_phase, 32);
                        tc_fence_after();
                        uint32_t v0[CPT], v1[CPT];
                        tmem_ld16(lane_base + (uint32_t)(d * TN), v0);
                        if (two) tmem_ld16(lane_base + (uint32_t)((d + 1) * TN), v1);
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            mbar_arrive(acc_empty0 + 8u * d);
                            if (two) mbar_arrive(acc_empty0 + 8u * (d + 1));
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
                    }
                    const int c0 = cb * TN + part * CPT;
                    bool any = false;
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        const double2 cq = __ldg(a.colq + c0 + c);
                        const uint4 ct = __ldg(a.coltab + c0 + c);
                        const uint64_t cf = (uint64_t)ct.y << 32 | ct.x;
                        // if ((cf & rf) == 0) { s_ok += X cs_c;  t_ok += X cs_c w_c; }
                        asm("{\n .reg .pred p;\n .reg .b64 t;\n and.b64 t, %3, %4;\n setp.eq.u64 p, t, 0;\n"
                            " @p fma.rn.f64 %0, %2, %5, %0;\n @p fma.rn.f64 %1, %2, %6, %1;\n}\n"
                            : "+d"(s_ok), "+d"(t_ok) : "d"(X[c]), "l"(cf), "l"(rf), "d"(cq.x), "d"(cq.y));
                        any |= (uint32_t)max(__double2hiint(X[c]), 0) + (uint32_t)__double2hiint(cq.x) >= thr_hi;
                        if ((c & 3) == 3) asm volatile("" ::: "memory");   // keeps the table loads of 4 columns in flight, not 16
                    }
                    if (any) {                                  // rare: some output may be above the candidate threshold
                        const double thr_q = a.sink.threshold / rs;
                        const uint64_t rx = __ldg(a.tab.row_x + a.row_lo + r_loc);
#pragma unroll
                        for (int c = 0; c < CPT; ++c) {
                            // (laundered: nothing of the fast loop is kept alive for this path)
                            const double q = opaque(X[c]) * ((const volatile double*)(a.colq + c0 + c))[0];
                            if (q > thr_q && c0 + c < a.n_cols) {
                                const uint64_t cx = ((const volatile uint64_t*)(a.coltab + c0 + c))[0];
                                cand_emit(a.sink, rx | (cx & ~(FLAG_ROW_BAD | FLAG_COL_BAD)), q * rs);
                            }
                        }
                    }
                    if (a.P_out && row_ok) {
#pragma unroll
                        for (int c = 0; c < CPT; ++c)
                            if (c0 + c < a.n_cols) {
                                const double p = opaque(X[c]) * ((const volatile double*)(a.colq + c0 + c))[0] * rs;
                                const long long o = (long long)r_loc * a.n_cols + c0 + c;
                                if (a.out32) ((float*)a.P_out)[o] = (float)p;
                                else ((double*)a.P_out)[o] = p;
                            }
                    }
                }
                cta_pwok += rs * fma((double)(rwok & 0xffu), s_ok, t_ok);
            }
        }
    }
    // teardown
    tc_fence_before();
    if constexpr (PAIR) {
        cluster_sync_all();                               // neither CTA's shared / tensor memory is touched by the other any more
        if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
    } else {
        __syncthreads();
        if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
    }
    MisAccum macc = {0.0, cta_pwok, 0.0};
    mis_block_store(macc, scratch, a.partials);
}

template <int S, int TN, int EPI>
int launch_i8(const I8Args& a, int grid, size_t smem, cudaStream_t s) {
    QDC_CHECK_CUDA(cudaFuncSetAttribute(contract_i8_kernel<S, TN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    contract_i8_kernel<S, TN, EPI><<<grid, (FIRST_EPI_WARP + 4 * (TN / cpt_of(TN))) * 32, smem, s>>>(a);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

// CTA-pair launch: clusters of 2 (the pair shares a TPC), grid even
template <int S, int TN>
int launch_i8_pair(const I8Args& a, int grid, size_t smem, cudaStream_t s) {
    auto kern = contract_i8_kernel<S, TN, 0, true>;
    QDC_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((FIRST_EPI_WARP + 4 * (TN / cpt_of(TN))) * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    QDC_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
    return QDC_OK;
}

// Tile width.  The M = 128 MMA instruction runs at the tensor pipe's full rate only from N = 128 on
// (64 clocks; N = 96: 56.8, N = 64: ~47, profiles/r1_umma_rate_probe.log), so 128 columns wherever the
// accumulators fit tensor memory (S <= 4 directly, S = 5 through 4 rotating slots) and two stages of
// the streamed operand fit beside the resident slices -- for the scoring-only epilogue; the storing
// one keeps 16 columns per thread: 96 columns at S = 5, else 64.  QDC_I8_TN overrides (64 / 96 / 128).
int i8_tile_n(int n_slices, int kd, int store) {
    const size_t smem_max = 227 * 1024 - 2048, a_res = (size_t)n_slices * TM * kd;
    auto fits = [&](int tn) { return a_res + 2 * (size_t)tn * kd <= smem_max; };
    if (const char* m = getenv("QDC_I8_TN")) {
        const int tn = atoi(m);
        if (tn == 64) return 64;
        if (tn == 96 && n_slices == 5 && fits(96) && store != 2) return 96;
        if (tn == 128 && !store && n_slices <= 5 && fits(128)) return 128;
    }
    if (!store && n_slices <= 5 && fits(128)) return 128;
    if (store == 2) return 64;                           // float32 outputs (EPI 2): 16 more live registers per thread than 72 leave
    if (n_slices == 5 && fits(96)) return 96;
    return 64;
}

// one instantiation for one slice count: geometry, work split, launch
int launch_for(I8Args a, int n_slices, int store, cudaStream_t s, bool allow_pair = true) {
    const I8Args a_in = a;                     // (for the fall-back below: the geometry fields are rewritten here)
    const int kd = a.kdim;
    const int TN = i8_tile_n(n_slices, kd, store ? (a.out32 ? 2 : 1) : 0);
    // CTA pairs (cta_group::2): the scoring-only kernel at 128 columns and kdim 256, at least two 256-row units
    bool pair = allow_pair && !store && TN == 128 && kd == 256 && n_slices >= 3 && n_slices <= 5 && a.n_rows > 2 * TM;
    if (const char* m = getenv("QDC_I8_PAIR")) pair = pair && atoi(m) != 0;
    const long long n_blocks = (a.n_rows + TM - 1) / TM;
    const long long n_rb = pair ? (n_blocks + 1) / 2 : n_blocks, n_cb = (a.n_cols + TN - 1) / TN;
    QDC_REQUIRE(n_rb < (1ll << 31) && n_cb < (1ll << 31), "too many tiles");
    a.n_rb = (int)n_rb; a.n_cb = (int)n_cb;
    a.a_blocks_pad = (int)((n_blocks * TM + I8_ROW_ALIGN - 1) / I8_ROW_ALIGN * I8_ROW_ALIGN / TM);
    const size_t smem_max = 227 * 1024 - 2048, a_res = (size_t)n_slices * TM * kd, b_slice = (size_t)(pair ? TN / 2 : TN) * kd;
    QDC_REQUIRE(a_res + 2 * b_slice <= smem_max, "shared memory");
    const size_t ct_bytes = pair ? (size_t)CT_STAGES * ct_stage_bytes(TN) : 0;
    size_t n_stages = (smem_max - a_res - ct_bytes) / b_slice;
    if (n_stages > MAX_STAGES) n_stages = MAX_STAGES;
    QDC_REQUIRE(n_stages >= 2, "shared memory");
    { const char* m = getenv("QDC_I8_STAGES"); if (m && atoi(m) >= 2 && (size_t)atoi(m) < n_stages) n_stages = atoi(m); }
    a.n_stages = (int)n_stages;
    const size_t smem = a_res + n_stages * b_slice + ct_bytes;
    const int sms = pair ? qdc_sm_count() / 2 : qdc_sm_count();          // schedulable units: SMs, or SM pairs
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
    a.n_chunks = n_chunks;
    a.chunk_len = (int)((n_cb + n_chunks - 1) / n_chunks);
    a.out_scale = ldexp(a.out_scale, -8 * (n_slices - 1));   // the row / column factors carry 2^(e - 7)
    const long long n_items = n_rb * n_chunks;
    QDC_REQUIRE(n_items < (1ll << 31), "too many work items");
    const int grid = (int)(n_items < sms ? n_items : sms) * (pair ? 2 : 1);
    if (pair) {
        int rc = QDC_ERR_UNSUPPORTED;
        switch (n_slices) {
            case 3: rc = launch_i8_pair<3, 128>(a, grid, smem, s); break;
            case 4: rc = launch_i8_pair<4, 128>(a, grid, smem, s); break;
            case 5: rc = launch_i8_pair<5, 128>(a, grid, smem, s); break;
        }
        if (rc == QDC_OK) return rc;
        // a cluster launch the device configuration refuses (no pair of SMs free for a 2-CTA cluster, e.g. under a
        // partitioned GPU): the single-CTA kernel computes the same numbers
        (void)cudaGetLastError();
        return launch_for(a_in, n_slices, store, s, false);
    }
#define QDC_I8_CASE(SS, TT)                                                                         \
    if (n_slices == SS && TN == TT && !(TT == 96 && store && a.out32))                               \
        return !store ? launch_i8<SS, TT, 0>(a, grid, smem, s) : (a.out32 ? launch_i8<SS, TT, 2>(a, grid, smem, s) : launch_i8<SS, TT, 1>(a, grid, smem, s));
    QDC_I8_CASE(2, 64) QDC_I8_CASE(3, 64) QDC_I8_CASE(4, 64) QDC_I8_CASE(5, 64) QDC_I8_CASE(5, 96) QDC_I8_CASE(6, 64)
#undef QDC_I8_CASE
    if (!store && TN == 128) {
        switch (n_slices) {
            case 2: return launch_i8<2, 128, 0>(a, grid, smem, s);
            case 3: return launch_i8<3, 128, 0>(a, grid, smem, s);
            case 4: return launch_i8<4, 128, 0>(a, grid, smem, s);
            case 5: return launch_i8<5, 128, 0>(a, grid, smem, s);
        }
    }
    qdc_set_error("no int8 contraction kernel for %d slices, tile width %d", n_slices, TN);
    return QDC_ERR_UNSUPPORTED;
}

}  // namespace

extern "C" int qdc_contract_i8_prepared(const void* d_opA, int64_t n_rows, const void* d_opB, int64_t n_cols,
                                        int kdim, int n_stored, int n_slices, double tol, int n_nodes,
                                        double scale, int norm_mode, const double* d_norm_in,
                                        const double* d_extra_colsums, int n_extra, double threshold,
                                        const double* d_A_src, const double* d_Bt_src, int64_t row_lo,
                                        const uint64_t* d_row_x, const uint64_t* d_row_forb,
                                        const uint32_t* d_row_wok, void* d_P_out, int store_mode,
                                        double* d_result, uint64_t* d_cand_x, double* d_cand_p, int64_t cand_cap,
                                        int k, uint64_t* d_top_x, double* d_top_p, double* d_partials, void* stream) {
    QDC_REQUIRE(d_opA && d_opB && d_result && d_partials, "null buffer");
    QDC_REQUIRE((((uintptr_t)d_opA) & 255) == 0 && (((uintptr_t)d_opB) & 255) == 0, "prepared operands must be 256-byte aligned");
    QDC_REQUIRE(d_row_x && d_row_forb && d_row_wok, "null tables");
    QDC_REQUIRE(d_cand_x && d_cand_p && cand_cap >= 0, "candidate sink");
    QDC_REQUIRE(kdim >= 1 && kdim <= 256, "the int8 path keeps a row block resident: kdim <= 256");
    QDC_REQUIRE(n_stored >= 2 && n_stored <= 6, "stored slices in 2..6");
    QDC_REQUIRE(n_slices == 0 || (n_slices >= 2 && n_slices <= n_stored), "n_slices: 0 (chosen from the bounds) or 2..stored");
    QDC_REQUIRE(n_rows >= 1 && n_cols >= 1 && row_lo >= 0, "shape");
    QDC_REQUIRE(threshold >= 0.0, "candidate threshold must be non-negative");
    QDC_REQUIRE(norm_mode >= 0 && norm_mode <= 2 && (norm_mode != 2 || d_norm_in), "norm_mode 0 (none), 1 (from the operands), 2 (given)");
    QDC_REQUIRE(store_mode == 0 || ((store_mode == 1 || store_mode == 2) && d_P_out),
                "store_mode 0 (none), 1 (united slab, float64) or 2 (united slab, float32)");
    QDC_REQUIRE(store_mode == 1 || (d_A_src && d_Bt_src), "the FP64 operands are needed to re-evaluate flagged candidates");
    QDC_REQUIRE(k >= 0 && k <= 1024 && (k == 0 || (d_top_x && d_top_p)), "k");
    cudaStream_t s = (cudaStream_t)stream;
    const int sms = qdc_sm_count();
    int rc = qdc_i8_launch_plan(d_opA, n_rows, d_opB, n_cols, kdim, n_stored, n_slices, tol, n_nodes, scale,
                                norm_mode, d_norm_in, d_extra_colsums, n_extra, d_result, d_partials, sms, s);
    if (rc != QDC_OK) return rc;
    unsigned long long* d_cand_count = reinterpret_cast<unsigned long long*>(d_result) + I8_RES_CAND;
    {
        const I8Layout LA = i8_layout(n_rows, kdim, n_stored), LB = i8_layout(n_cols, kdim, n_stored);
        const uint8_t* pa = (const uint8_t*)d_opA;
        const uint8_t* pb = (const uint8_t*)d_opB;
        I8Args a;
        a.A8 = pa + LA.off_slices; a.B8 = pb + LB.off_slices;
        a.a_slice_stride = LA.slice_stride; a.b_slice_stride = LB.slice_stride;
        a.row_scale = (const double*)(pa + LA.off_scale);
        a.colq = (const double2*)(pb + LB.off_colq);
        a.coltab = (const uint4*)(pb + LB.off_coltab);
        a.gref = (const int*)(pb + LB.off_gref);
        a.A_src = d_A_src; a.B_src = d_Bt_src; a.kdim_src = kdim;
        a.n_rows = n_rows; a.n_cols = n_cols; a.row_lo = row_lo; a.kdim = LA.kd;
        a.out_scale = scale;
        a.res = d_result; a.use_norm = norm_mode != 0;
        a.tab = {d_row_x, d_row_forb, d_row_wok, nullptr, nullptr};
        a.sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
        a.P_out = store_mode ? d_P_out : nullptr;
        a.out32 = store_mode == 2;
        a.partials = d_partials;
        a.check_slices = n_slices == 0;
        for (int S = (n_slices ? n_slices : 3); S <= (n_slices ? n_slices : n_stored); ++S) {
            rc = launch_for(a, S, store_mode != 0, s);
            if (rc != QDC_OK) return rc;
        }
    }
    unsigned long long* d_list_count = d_cand_count;           // entries of the candidate arrays the top-k walks
    if (store_mode != 1) {
        // the integer-scoring kernels only flagged outputs: evaluate them exactly now that no MMA runs
        unsigned long long* d_flagged = reinterpret_cast<unsigned long long*>(d_result) + I8_RES_FLAGGED;
        const FlagCtx fx = {d_A_src, d_Bt_src, (const uint4*)((const uint8_t*)d_opB + i8_layout(n_cols, kdim, n_stored).off_coltab),
                            d_row_x, (long long)n_cols, kdim, d_flagged,
                            reinterpret_cast<unsigned long long*>(d_result) + I8_RES_LOST};
        const CandSink sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
        i8_reeval_kernel<<<sms * 4, 256, 0, s>>>(fx, sink, d_result, norm_mode != 0, scale, (long long)row_lo);
        QDC_CHECK_LAUNCH();
        d_list_count = d_flagged;
    }
    // the masked sum (slots 0 and 2 of the partials are zero) and the top-k, one launch
    return qdc_launch_finish(d_partials, sms, d_result + I8_RES_SUMS, d_cand_x, d_cand_p, d_list_count, cand_cap, k,
                             d_top_x, d_top_p, s);
}
