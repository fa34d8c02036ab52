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
