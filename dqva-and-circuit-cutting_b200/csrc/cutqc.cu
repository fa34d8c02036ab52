// CutQC-style reconstruction stages as bandwidth-bound kernels (K3c, K4c, K6, K7).
//
// Replaces the reference's four C programs, which exchange `%e` text files between processes:
//   measure.c           :123-161  signed marginal over the measured ("O") qubits
//   vertical_collapse.c :27-104   kron term = sum_i coefficient_i * measured instance_i
//   merge.c             :28-122   dynamic-definition binning (merged / active / zoomed qubits)
//   build.c             :69-196   0.5^K * sum_terms kron(v_s0, v_s1, ...) -- here: operands
//                                 gathered per subcircuit (+ a Khatri-Rao product when there are
//                                 more than two), then the tensor-core contraction of contract.cu
// All vectors stay FP64 on the device (the reference computes in float32 and rounds to 7 digits
// at every file hop).  Bound: HBM; algorithmic bytes are stated per kernel.
#include "qdc_common.cuh"

namespace {

// deposit the low bits of v into the set positions of mask (ascending)
__device__ __forceinline__ uint64_t deposit_bits(uint64_t v, uint64_t mask) {
    uint64_t out = 0;
    while (mask) {
        const uint64_t low = mask & (0 - mask);
        if (v & 1ull) out |= low;
        v >>= 1;
        mask ^= low;
    }
    return out;
}

// ---- measure: out[i][e] = sum_o (-1)^popc(x & sign_i) raw[i][x], x = dep(e, keep) | dep(o, ~keep) ---
// Algorithmic bytes: sizeof(T) * 2^d read + 8 * 2^eff written per instance.
// One team of TEAM lanes per output; lanes stride over the O states (coalesced when the lowest
// qubits are measured ones), fixed-order shuffle reduction -> deterministic.
template <typename T, int TEAM>
__global__ void __launch_bounds__(256) measure_kernel(const T* __restrict__ raw, int d, int n_eff,
                                                      long long n_inst, uint64_t keep_mask,
                                                      const uint64_t* __restrict__ sign_masks,
                                                      double* __restrict__ out) {
    const uint64_t all = (1ull << d) - 1ull, o_mask = all & ~keep_mask;
    const long long n_o = 1ll << (d - n_eff);
    const long long total = n_inst << n_eff;
    const int sub = threadIdx.x % TEAM, lane = threadIdx.x & 31;
    // the loop bound is warp-uniform (first team of the warp) so the shuffles below always see
    // all 32 lanes; teams past the end compute nothing and store nothing
    for (long long t0 = (blockIdx.x * (long long)blockDim.x + threadIdx.x - lane) / TEAM; t0 < total;
         t0 += (long long)gridDim.x * blockDim.x / TEAM) {
        const long long team = t0 + lane / TEAM;
        const bool valid = team < total;
        const long long inst = valid ? team >> n_eff : 0;
        const uint64_t e = (uint64_t)team & ((1ull << n_eff) - 1ull);
        const uint64_t base = deposit_bits(e, keep_mask);
        const uint64_t sign = sign_masks[inst];
        const T* src = raw + (inst << d);
        double acc = 0.0;
        for (long long o = sub; valid && o < n_o; o += TEAM) {
            const uint64_t x = base | deposit_bits((uint64_t)o, o_mask);
            const double v = (double)src[x];
            acc += (__popcll((long long)(x & sign)) & 1) ? -v : v;
        }
#pragma unroll
        for (int s = TEAM / 2; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
        if (sub == 0 && valid) out[team] = acc;
    }
}

// ---- vertical collapse: out[t][j] = sum_{i in term t} coeff_i * measured[inst_i][j] -------------
// CSR over terms.  Algorithmic bytes: 8 * len * (instances referenced + 1) per term.
__global__ void __launch_bounds__(256) collapse_kernel(const double* __restrict__ measured,
                                                       long long len, int n_terms,
                                                       const int32_t* __restrict__ term_ptr,
                                                       const int32_t* __restrict__ inst_idx,
                                                       const int32_t* __restrict__ coeff,
                                                       double* __restrict__ out,
                                                       int32_t* __restrict__ nonzero) {
    const int t = blockIdx.y;
    const int lo = term_ptr[t], hi = term_ptr[t + 1];
    bool any = false;
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < len;
         j += (long long)gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int k = lo; k < hi; ++k) acc += (double)coeff[k] * measured[(long long)inst_idx[k] * len + j];
        out[(long long)t * len + j] = acc;
        any |= fabs(acc) > 1e-16;                         // vertical_collapse.c:65-75
    }
    if (__syncthreads_or(any) && threadIdx.x == 0 && nonzero) atomicOr(nonzero + t, 1);
}

// ---- merge: out[v][a] = sum_m in[v][dep(a, active) | dep(m, merged) | zoom_val] -----------------
// Algorithmic bytes: 8 * 2^(active + merged) read + 8 * 2^active written per vector.
template <int TEAM>
__global__ void __launch_bounds__(256) merge_kernel(const double* __restrict__ in, int n_eff,
                                                    int n_active, int n_merged, long long n_vec,
                                                    uint64_t active_mask, uint64_t merged_mask,
                                                    uint64_t zoom_val, double* __restrict__ out) {
    const long long total = n_vec << n_active;
    const long long n_m = 1ll << n_merged;
    const int sub = threadIdx.x % TEAM, lane = threadIdx.x & 31;
    for (long long t0 = (blockIdx.x * (long long)blockDim.x + threadIdx.x - lane) / TEAM; t0 < total;
         t0 += (long long)gridDim.x * blockDim.x / TEAM) {
        const long long team = t0 + lane / TEAM;
        const bool valid = team < total;
        const long long v = valid ? team >> n_active : 0;
        const uint64_t a = (uint64_t)team & ((1ull << n_active) - 1ull);
        const uint64_t base = deposit_bits(a, active_mask) | zoom_val;
        const double* src = in + (v << n_eff);
        double acc = 0.0;
        for (long long m = sub; valid && m < n_m; m += TEAM)
            acc += src[base | deposit_bits((uint64_t)m, merged_mask)];
#pragma unroll
        for (int s = TEAM / 2; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
        if (sub == 0 && valid) out[team] = acc;
    }
}

// ---- tiled signed marginal (measure and merge at HBM speed) ---------------------------------------
//   out[v][a] = sum_m (-1)^popc(x & sign_v) in[v][x],   x = dep(a, keep) | dep(m, summed) | fixed
// The per-element bit deposits of the kernels above cost ~100 instructions per element (5-15 % of
// the HBM rate).  Here the index is split at bit TILE_BITS: one CTA owns (vector, high kept bits)
// and walks the high summed bits; each step loads one contiguous tile of 2^TILE_BITS elements
// (coalesced 16-byte loads, sign of the low bits applied on the way) into shared memory, and the
// low bits are resolved through two small tables, so an element costs one shared-memory read, an
// OR and an add.  Accumulation order is fixed (deterministic).  Algorithmic bytes: the input once,
// the output once.
constexpr int TILE_BITS = 12, MARG_THREADS = 256, MARG_ACC_MAX = (1 << TILE_BITS) / MARG_THREADS;

struct MargArgs {
    const void* in;
    double* out;
    const uint64_t* sign_masks;      // per vector, or null
    long long n_vec;
    int d, tile_bits;                // total index bits; low bits inside a tile (<= TILE_BITS)
    uint64_t keep_mask, sum_mask, fixed_val;
};

// Round 2 (ncu: the kernel was issue-bound at 62 / 100 thread-instructions per element, profiles/r2_marginal_ncu_full.csv):
// ACC = accumulator slots a thread really needs (1, 4 or 16: the 16 predicated slots of which 1-4 were live cost a
// third of the instructions), the high kept bits of a work item are deposited through two 256-entry tables instead
// of a bit loop per item, the high summed bits advance by the carry trick, and the sign of a loaded element is a
// per-thread constant times a per-iteration one.
template <typename T, bool LOOP, int ACC>
__global__ void __launch_bounds__(MARG_THREADS, LOOP ? 4 : 3) marginal_kernel(const MargArgs a) {
    __shared__ double tile[1 << TILE_BITS];
    // the two low-bit tables share one array: 2^n_klo + 2^n_slo <= 2^12 + 1 entries (n_klo + n_slo <= 12)
    __shared__ uint16_t lut_pool[(1 << TILE_BITS) + 256];
    __shared__ uint64_t dep_lo[256], dep_hi[256];
    constexpr int VEC = 16 / (int)sizeof(T), VEC_BITS = VEC == 4 ? 2 : 1;
    const int L = a.tile_bits, tid = threadIdx.x;
    const uint64_t low = (1ull << L) - 1ull;
    // trailing summed bits that fall inside one 16-byte load are added up in registers on the way
    // in (tz of them): the tile in shared memory then starts at the first bit that is not summed
    int tz = 0;
    while (tz < VEC_BITS && tz < L && ((a.sum_mask >> tz) & 1ull)) ++tz;
    const int Lt = L - tz;                                         // bits of the compacted tile
    const uint64_t keep_lo = (a.keep_mask & low) >> tz, sum_lo = (a.sum_mask & low) >> tz, fixed_lo = (a.fixed_val & low) >> tz;
    const uint64_t keep_hi = a.keep_mask >> L, sum_hi = a.sum_mask >> L, fixed_hi = a.fixed_val >> L;
    const int n_klo = __popcll(keep_lo), n_slo = __popcll(sum_lo), n_khi = __popcll(keep_hi), n_shi = __popcll(sum_hi);
    const int n_keep = n_klo + n_khi;
    uint16_t* lut_a = lut_pool;
    uint16_t* lut_m = lut_pool + (1 << n_klo);
    for (int i = tid; i < (1 << n_klo); i += MARG_THREADS) lut_a[i] = (uint16_t)(deposit_bits((uint64_t)i, keep_lo) | fixed_lo);
    for (int i = tid; i < (1 << n_slo); i += MARG_THREADS) lut_m[i] = (uint16_t)deposit_bits((uint64_t)i, sum_lo);
    for (int i = tid; i < 256; i += MARG_THREADS) {
        dep_lo[i] = deposit_bits((uint64_t)i, keep_hi);
        dep_hi[i] = deposit_bits((uint64_t)i << 8, keep_hi);
    }
    // Work split of the low bits: with 2^n_klo >= 256 outputs per tile a thread owns outputs
    // tid, tid + 256, ... and walks all 2^n_slo summed states; with fewer outputs P = 256 / 2^n_klo
    // adjacent lanes share one output and split the summed states (their partial sums meet in a
    // shuffle reduction after the last tile, P <= 32).
    const int n_out = 1 << n_klo;
    int p_bits = 0;
    while ((n_out << p_bits) < MARG_THREADS && p_bits < 5 && p_bits < n_slo) ++p_bits;
    const int P = 1 << p_bits;
    const int my_part = tid & (P - 1), my_out0 = tid >> p_bits;     // outputs my_out0 + j * (256 / P)
    const int out_stride = MARG_THREADS >> p_bits;
    const int m_per = (1 << n_slo) >> p_bits, m0 = my_part * m_per;
    uint32_t la[ACC];
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ACC; ++j) la[j] = lut_a[(my_out0 + j * out_stride) & (n_out - 1)];
    // persistent CTAs: work item = (vector, high kept bits); the tables above are built once
    const long long n_work = a.n_vec << n_khi;
    long long w = blockIdx.x;
    // the next tile's loads are issued right after this tile has been written to shared memory, so they are in
    // flight during the table phase (U 16-byte loads per thread = one whole tile)
    constexpr int U = (1 << TILE_BITS) / (VEC * MARG_THREADS);
    int4 raw[U];
    bool pref = false;
    const bool vec_path = (1 << L) >= VEC * MARG_THREADS;
    auto load_tile = [&](const T* p) {
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (tid * VEC + u * VEC * MARG_THREADS < (1 << L))
                raw[u] = __ldcs(reinterpret_cast<const int4*>(p + tid * VEC + u * VEC * MARG_THREADS));
    };
    do {
    const long long v = w >> n_khi;
    const uint64_t a_hi = (uint64_t)w & ((1ull << n_khi) - 1ull);
    const uint64_t x_keep_hi = (n_khi <= 16 ? (dep_lo[a_hi & 255] | dep_hi[(a_hi >> 8) & 255]) : deposit_bits(a_hi, keep_hi)) | fixed_hi;
    const uint64_t sign = a.sign_masks ? a.sign_masks[v] : 0ull;
    const uint32_t sign_lo = (uint32_t)(sign & low);
    const T* src = reinterpret_cast<const T*>(a.in) + (v << a.d);
    double acc[ACC];
#pragma unroll
    for (int j = 0; j < ACC; ++j) acc[j] = 0.0;
    // sign of element i + k of a tile = parity((i + k) & sign_lo): i = tid VEC + it VEC 256, the two parts share no bits
    double sgn_k[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) sgn_k[k] = (__popc((uint32_t)(tid * VEC + k) & sign_lo) & 1) ? -1.0 : 1.0;
    uint64_t x_sum_hi = 0;                                          // deposit of m_hi into sum_hi, advanced by the carry trick
    for (long long m_hi = 0; m_hi < (1ll << n_shi); ++m_hi, x_sum_hi = ((x_sum_hi | ~sum_hi) + 1ull) & sum_hi) {
        const uint64_t x_hi = x_keep_hi | x_sum_hi;
        const double hi_sign = (__popcll((long long)((x_hi << L) & sign)) & 1) ? -1.0 : 1.0;
        const T* t_src = src + (x_hi << L);
        if ((1 << L) >= VEC * MARG_THREADS) {
            if (!pref) load_tile(t_src);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = tid * VEC + u * VEC * MARG_THREADS;
                if (i >= (1 << L)) break;
                T e[VEC];
                *reinterpret_cast<int4*>(e) = raw[u];
                double f[VEC];
                const double it_sign = (__popc((uint32_t)(i - tid * VEC) & sign_lo) & 1) ? -hi_sign : hi_sign;
#pragma unroll
                for (int k = 0; k < VEC; ++k) f[k] = (double)e[k] * (sgn_k[k] * it_sign);
                if (tz == 2 && VEC == 4) tile[i >> 2] = (f[0] + f[1]) + (f[2 % VEC] + f[3 % VEC]);
                else if (tz == 1) {
#pragma unroll
                    for (int k = 0; k < VEC; k += 2) tile[(i + k) >> 1] = f[k] + f[k + 1];
                } else {
#pragma unroll
                    for (int k = 0; k < VEC; ++k) tile[i + k] = f[k];
                }
            }
        } else {                                                   // small vectors: tz = 0 path only
            for (int i = tid; i < (1 << L); i += MARG_THREADS) {
                const double f = (__popc((uint32_t)i & sign_lo) & 1) ? -hi_sign * (double)t_src[i] : hi_sign * (double)t_src[i];
                if (tz == 0) tile[i] = f;
                else atomicAdd(&tile[i >> tz], f);                 // (never taken: see launch_marginal)
            }
        }
        __syncthreads();
        pref = false;
        if (vec_path) {                                            // (uniform over the CTA)
            if (m_hi + 1 < (1ll << n_shi)) {
                const uint64_t nx = ((x_sum_hi | ~sum_hi) + 1ull) & sum_hi;
                load_tile(src + ((x_keep_hi | nx) << L));
                pref = true;
            } else if (LOOP && w + gridDim.x < n_work) {
                const long long wn = w + gridDim.x;
                const uint64_t an = (uint64_t)wn & ((1ull << n_khi) - 1ull);
                const uint64_t xk = (n_khi <= 16 ? (dep_lo[an & 255] | dep_hi[(an >> 8) & 255]) : deposit_bits(an, keep_hi)) | fixed_hi;
                load_tile(reinterpret_cast<const T*>(a.in) + ((wn >> n_khi) << a.d) + (xk << L));   // (its m_hi = 0: no high summed bits set)
                pref = true;
            }
        }
        for (int m = m0; m < m0 + m_per; ++m) {
            const uint32_t lm = lut_m[m];
#pragma unroll
            for (int j = 0; j < ACC; ++j)
                if (my_out0 + j * out_stride < n_out) acc[j] += tile[la[j] | lm];
        }
        __syncthreads();
    }
    double* dst = a.out + (v << n_keep) + ((long long)a_hi << n_klo);
#pragma unroll
    for (int j = 0; j < ACC; ++j) {
        double r = acc[j];
        for (int o = 1; o < P; o <<= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (my_part == 0 && my_out0 + j * out_stride < n_out) dst[my_out0 + j * out_stride] = r;
    }
    w += gridDim.x;
    } while (LOOP && w < n_work);
    (void)Lt;
}

// The same sum when NO summed (and no fixed-to-one) bit lies between the trailing bits a 16-byte load covers and
// bit TILE_BITS: every element of a tile then belongs to its own output (summed bits are either added inside the
// loaded vector or are high bits, i.e. other tiles), so a thread keeps its U loads' outputs in registers across
// the high summed bits and writes them back itself -- no shared-memory tile, no table phase, no barrier: a
// streaming pass.  Arithmetic and summation order are those of marginal_kernel (bit-identical results).
// TZ (trailing summed bits inside a load) is a template parameter: a thread keeps VEC >> TZ sums per load, and with
// only those live more CTAs fit an SM -- the kernel has nothing but its loads in flight to hide HBM latency with.
template <typename T, int TZ>
__global__ void __launch_bounds__(MARG_THREADS, TZ == 0 ? 3 : (TZ == 1 && sizeof(T) == 4 ? 4 : 6)) marginal_direct_kernel(const MargArgs a) {
    constexpr int tz = TZ;
    constexpr int VEC = 16 / (int)sizeof(T), U = (1 << TILE_BITS) / (VEC * MARG_THREADS);
    constexpr int L = TILE_BITS;
    const int tid = threadIdx.x;
    const uint64_t low = (1ull << L) - 1ull;
    const uint64_t keep_hi = a.keep_mask >> L, sum_hi = a.sum_mask >> L, fixed_hi = a.fixed_val >> L;
    const int n_khi = __popcll(keep_hi), n_shi = __popcll(sum_hi), n_klo = L - tz, n_keep = n_klo + n_khi;
    const long long n_work = a.n_vec << n_khi;
    for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
        const long long v = w >> n_khi;
        const uint64_t a_hi = (uint64_t)w & ((1ull << n_khi) - 1ull);
        const uint64_t x_keep_hi = deposit_bits(a_hi, keep_hi) | fixed_hi;
        const uint64_t sign = a.sign_masks ? a.sign_masks[v] : 0ull;
        const uint32_t sign_lo = (uint32_t)(sign & low);
        const T* src = reinterpret_cast<const T*>(a.in) + (v << a.d);
        double acc[U][VEC >> TZ];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < (VEC >> TZ); ++k) acc[u][k] = 0.0;
        double sgn_k[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) sgn_k[k] = (__popc((uint32_t)(tid * VEC + k) & sign_lo) & 1) ? -1.0 : 1.0;
        uint64_t x_sum_hi = 0;
        for (long long m_hi = 0; m_hi < (1ll << n_shi); ++m_hi, x_sum_hi = ((x_sum_hi | ~sum_hi) + 1ull) & sum_hi) {
            const uint64_t x_hi = x_keep_hi | x_sum_hi;
            const double hi_sign = (__popcll((long long)((x_hi << L) & sign)) & 1) ? -1.0 : 1.0;
            const T* t_src = src + (x_hi << L);
            int4 raw[U];
#pragma unroll
            for (int u = 0; u < U; ++u) raw[u] = __ldcs(reinterpret_cast<const int4*>(t_src + tid * VEC + u * VEC * MARG_THREADS));
#pragma unroll
            for (int u = 0; u < U; ++u) {
                T e[VEC];
                *reinterpret_cast<int4*>(e) = raw[u];
                double f[VEC];
                const double it_sign = (__popc((uint32_t)(u * VEC * MARG_THREADS) & sign_lo) & 1) ? -hi_sign : hi_sign;
#pragma unroll
                for (int k = 0; k < VEC; ++k) f[k] = (double)e[k] * (sgn_k[k] * it_sign);
                if constexpr (TZ == 2) acc[u][0] += (f[0] + f[1]) + (f[2 % VEC] + f[3 % VEC]);
                else if constexpr (TZ == 1) {
#pragma unroll
                    for (int k = 0; k < VEC; k += 2) acc[u][k >> 1] += f[k] + f[k + 1];
                } else {
#pragma unroll
                    for (int k = 0; k < VEC; ++k) acc[u][k] += f[k];
                }
            }
        }
        double* dst = a.out + (v << n_keep) + ((long long)a_hi << n_klo);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            double* o = dst + ((tid * VEC + u * VEC * MARG_THREADS) >> tz);
            constexpr int n_out = VEC >> TZ;                           // 4, 2 or 1 consecutive outputs
            if constexpr (n_out == 1) __stcs(o, acc[u][0]);
            else {
#pragma unroll
                for (int k = 0; k < n_out; k += 2)
                    __stcs(reinterpret_cast<double2*>(o + k), make_double2(acc[u][k], acc[u][k + 1]));
            }
        }
    }
}

// true when the tiled kernel applies (enough index bits for a tile, a launchable grid)
template <typename T>
bool launch_marginal(const void* in, double* out, const uint64_t* sign_masks, long long n_vec, int d,
                     uint64_t keep_mask, uint64_t sum_mask, uint64_t fixed_val, cudaStream_t s) {
    const int L = d < TILE_BITS ? d : TILE_BITS;
    if (L < 10) return false;                              // a tile is at least one 16-byte load per thread
    const int n_khi = __builtin_popcountll(keep_mask >> L);
    const long long work = n_vec << n_khi;
    if (work <= 0 || n_khi > 40) return false;
    // one tile per work item (no high summed bits): a few resident CTAs per SM walk the items, so
    // the tables and the launch cost are paid once per many tiles; several tiles per item: one CTA
    // each, the hardware scheduler balances them
    const int n_shi = __builtin_popcountll(sum_mask >> L);
    MargArgs a = {in, out, sign_masks, n_vec, d, L, keep_mask, sum_mask, fixed_val};
    const long long resident = 4ll * qdc_sm_count();
    // accumulator slots a thread needs (the kernel's own work split, restated): outputs per tile over the threads
    // that do not share an output
    int acc;
    {
        const int vec_bits = sizeof(T) == 4 ? 2 : 1;
        const uint64_t low = (1ull << L) - 1ull;
        int tz = 0;
        while (tz < vec_bits && tz < L && ((sum_mask >> tz) & 1ull)) ++tz;
        const int n_klo = __builtin_popcountll((keep_mask & low) >> tz), n_slo = __builtin_popcountll((sum_mask & low) >> tz);
        int p_bits = 0;
        while (((1 << n_klo) << p_bits) < MARG_THREADS && p_bits < 5 && p_bits < n_slo) ++p_bits;
        const int need = ((1 << n_klo) << p_bits) / MARG_THREADS;
        acc = need <= 1 ? 1 : (need <= 4 ? 4 : MARG_ACC_MAX);
    }
    {
        // streaming form: all low bits kept except the trailing ones a load adds up (QDC_MARG_DIRECT=0: tiled form)
        const int vec_bits = sizeof(T) == 4 ? 2 : 1;
        int tz = 0;
        while (tz < vec_bits && ((sum_mask >> tz) & 1ull)) ++tz;
        const uint64_t low = (1ull << TILE_BITS) - 1ull;
        const char* knob = getenv("QDC_MARG_DIRECT");
        const bool direct_ok = !(knob && atoi(knob) == 0);
        if (direct_ok && L == TILE_BITS && ((keep_mask & low) >> tz) == (low >> tz) && (fixed_val & low) == 0 &&
            (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
            const unsigned grid = (unsigned)(work < (1ll << 30) ? work : (1ll << 30));
            if (tz == 0) marginal_direct_kernel<T, 0><<<grid, MARG_THREADS, 0, s>>>(a);
            else if (tz == 1) marginal_direct_kernel<T, 1><<<grid, MARG_THREADS, 0, s>>>(a);
            else if constexpr (sizeof(T) == 4) marginal_direct_kernel<T, 2><<<grid, MARG_THREADS, 0, s>>>(a);
            return true;
        }
    }
    const bool loop = (n_shi == 0 && work > resident) || work >= (1ll << 31);
    const unsigned grid = loop ? (unsigned)resident : (unsigned)work;
#define QDC_MARG(LOOPV, ACCV) marginal_kernel<T, LOOPV, ACCV><<<grid, MARG_THREADS, 0, s>>>(a)
    if (loop) { if (acc == 1) QDC_MARG(true, 1); else if (acc == 4) QDC_MARG(true, 4); else QDC_MARG(true, MARG_ACC_MAX); }
    else { if (acc == 1) QDC_MARG(false, 1); else if (acc == 4) QDC_MARG(false, 4); else QDC_MARG(false, MARG_ACC_MAX); }
#undef QDC_MARG
    return true;
}

// ---- operand gather: A[i][t] = vecs[idx[t]][i]  (idx < 0 or t >= n_terms: 0) ---------------------
// 32x32 shared-memory transpose so both the vector reads and the operand writes are coalesced.
__global__ void __launch_bounds__(256) gather_kernel(const double* __restrict__ vecs, long long len,
                                                     const int32_t* __restrict__ idx, int n_terms,
                                                     int kdim, double* __restrict__ A) {
    __shared__ double tile[32][33];
    const long long i0 = (long long)blockIdx.x * 32;
    const int t0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
    for (int r = ty; r < 32; r += 8) {
        const int t = t0 + r;
        const int src = t < n_terms ? idx[t] : -1;
        const long long i = i0 + tx;
        tile[r][tx] = (src >= 0 && i < len) ? vecs[(long long)src * len + i] : 0.0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const long long i = i0 + r;
        const int t = t0 + tx;
        if (i < len && t < kdim) A[i * kdim + t] = tile[tx][r];
    }
}

// ---- Khatri-Rao: W[(i, j)][t] = A0[i][t] * A1[j][t] ---------------------------------------------
__global__ void __launch_bounds__(256) khatri_rao_kernel(const double* __restrict__ A0, long long m0,
                                                         const double* __restrict__ A1, long long m1,
                                                         int kdim, double* __restrict__ W) {
    const long long total = m0 * m1 * (kdim / 2);
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const int t2 = (int)(e % (kdim / 2));
        const long long ij = e / (kdim / 2);
        const long long j = ij % m1, i = ij / m1;
        const double2 a = reinterpret_cast<const double2*>(A0 + i * kdim)[t2];
        const double2 b = reinterpret_cast<const double2*>(A1 + j * kdim)[t2];
        reinterpret_cast<double2*>(W + ij * kdim)[t2] = make_double2(a.x * b.x, a.y * b.y);
    }
}

int grid_for(long long threads_needed, int block) {
    long long blocks = (threads_needed + block - 1) / block;
    const long long cap = (long long)qdc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

extern "C" int qdc_cutqc_measure(const void* d_raw, int dtype, int n_qubits, int64_t n_inst,
                                 uint64_t keep_mask, const uint64_t* d_sign_masks, double* d_out,
                                 void* stream) {
    QDC_REQUIRE(d_raw && d_sign_masks && d_out, "null buffer");
    QDC_REQUIRE(n_qubits >= 1 && n_qubits <= 40 && n_inst >= 1, "shape");
    QDC_REQUIRE(dtype == 0 || dtype == 1, "dtype (0 f64, 1 f32)");
    QDC_REQUIRE((keep_mask >> n_qubits) == 0, "keep_mask");
    const int n_eff = __builtin_popcountll(keep_mask);
    const int n_o = n_qubits - n_eff;
    cudaStream_t s = (cudaStream_t)stream;
    {
        const uint64_t all = (1ull << n_qubits) - 1ull;
        const bool done = dtype == 0
            ? launch_marginal<double>(d_raw, d_out, d_sign_masks, n_inst, n_qubits, keep_mask, all & ~keep_mask, 0ull, s)
            : launch_marginal<float>(d_raw, d_out, d_sign_masks, n_inst, n_qubits, keep_mask, all & ~keep_mask, 0ull, s);
        if (done) { QDC_CHECK_LAUNCH(); return QDC_OK; }
    }
    const long long teams = n_inst << n_eff;
#define QDC_LAUNCH_MEASURE(T, TEAM)                                                              \
    measure_kernel<T, TEAM><<<grid_for(teams * TEAM, 256), 256, 0, s>>>(                         \
        (const T*)d_raw, n_qubits, n_eff, n_inst, keep_mask, d_sign_masks, d_out)
    if (dtype == 0) {
        if (n_o >= 5) QDC_LAUNCH_MEASURE(double, 32); else if (n_o >= 2) QDC_LAUNCH_MEASURE(double, 4);
        else QDC_LAUNCH_MEASURE(double, 1);
    } else {
        if (n_o >= 5) QDC_LAUNCH_MEASURE(float, 32); else if (n_o >= 2) QDC_LAUNCH_MEASURE(float, 4);
        else QDC_LAUNCH_MEASURE(float, 1);
    }
#undef QDC_LAUNCH_MEASURE
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_cutqc_collapse(const double* d_measured, int64_t len, int n_terms,
                                  const int32_t* d_term_ptr, const int32_t* d_inst_idx,
                                  const int32_t* d_coeff, double* d_out, int32_t* d_nonzero,
                                  void* stream) {
    QDC_REQUIRE(d_measured && d_term_ptr && d_inst_idx && d_coeff && d_out, "null buffer");
    QDC_REQUIRE(len >= 1 && n_terms >= 1 && n_terms <= 65535, "shape (at most 65535 terms per call)");
    cudaStream_t s = (cudaStream_t)stream;
    if (d_nonzero) QDC_CHECK_CUDA(cudaMemsetAsync(d_nonzero, 0, sizeof(int32_t) * n_terms, s));
    long long bx = (len + 255) / 256;
    if (bx > 1024) bx = 1024;
    collapse_kernel<<<dim3((unsigned)bx, (unsigned)n_terms), 256, 0, s>>>(d_measured, len, n_terms, d_term_ptr,
                                                                        d_inst_idx, d_coeff, d_out, d_nonzero);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_cutqc_merge(const double* d_in, int64_t n_vec, int n_effective,
                               uint64_t active_mask, uint64_t zoom_mask, uint64_t zoom_val,
                               double* d_out, void* stream) {
    QDC_REQUIRE(d_in && d_out, "null buffer");
    QDC_REQUIRE(n_effective >= 0 && n_effective <= 40 && n_vec >= 1, "shape");
    const uint64_t all = n_effective == 0 ? 0ull : ((1ull << n_effective) - 1ull);
    QDC_REQUIRE((active_mask & ~all) == 0 && (zoom_mask & ~all) == 0 && (active_mask & zoom_mask) == 0 &&
                (zoom_val & ~zoom_mask) == 0, "masks");
    const uint64_t merged_mask = all & ~active_mask & ~zoom_mask;
    const int n_active = __builtin_popcountll(active_mask), n_merged = __builtin_popcountll(merged_mask);
    cudaStream_t s = (cudaStream_t)stream;
    if (launch_marginal<double>(d_in, d_out, nullptr, n_vec, n_effective, active_mask, merged_mask, zoom_val, s)) {
        QDC_CHECK_LAUNCH();
        return QDC_OK;
    }
    const long long teams = n_vec << n_active;
    if (n_merged >= 5)
        merge_kernel<32><<<grid_for(teams * 32, 256), 256, 0, s>>>(d_in, n_effective, n_active, n_merged, n_vec,
                                                                   active_mask, merged_mask, zoom_val, d_out);
    else if (n_merged >= 2)
        merge_kernel<4><<<grid_for(teams * 4, 256), 256, 0, s>>>(d_in, n_effective, n_active, n_merged, n_vec,
                                                                 active_mask, merged_mask, zoom_val, d_out);
    else
        merge_kernel<1><<<grid_for(teams, 256), 256, 0, s>>>(d_in, n_effective, n_active, n_merged, n_vec,
                                                             active_mask, merged_mask, zoom_val, d_out);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_cutqc_gather(const double* d_vecs, int64_t len, const int32_t* d_idx, int n_terms,
                                int kdim, double* d_A, void* stream) {
    QDC_REQUIRE(d_vecs && d_idx && d_A, "null buffer");
    QDC_REQUIRE(len >= 1 && n_terms >= 1 && kdim >= n_terms, "shape");
    const long long gx = (len + 31) / 32;
    QDC_REQUIRE(gx < (1ll << 31) && (kdim + 31) / 32 <= 65535, "grid");
    gather_kernel<<<dim3((unsigned)gx, (unsigned)((kdim + 31) / 32)), 256, 0, (cudaStream_t)stream>>>(
        d_vecs, len, d_idx, n_terms, kdim, d_A);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_khatri_rao(const double* d_A0, int64_t m0, const double* d_A1, int64_t m1, int kdim,
                              double* d_W, void* stream) {
    QDC_REQUIRE(d_A0 && d_A1 && d_W, "null buffer");
    QDC_REQUIRE(m0 >= 1 && m1 >= 1 && kdim >= 2 && kdim % 2 == 0, "shape (kdim even)");
    QDC_REQUIRE(((((uintptr_t)d_A0) | ((uintptr_t)d_A1) | ((uintptr_t)d_W)) & 15) == 0, "16-byte alignment");
    khatri_rao_kernel<<<grid_for(m0 * m1 * (kdim / 2), 256), 256, 0, (cudaStream_t)stream>>>(d_A0, m0, d_A1, m1,
                                                                                            kdim, d_W);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}
