// Penalised MIS objective of QAOA+ over a materialised distribution or over amplitudes:
//   sum_x p(x),  sum_x p(x) * #edges(x),  sum_x p(x) * |x|
// with #edges(x) = number of graph edges whose two ends are both 1 in x, so that the reference's
// expectation_value(counts, G, Lambda) (ansatz/qaoa_plus.py:40-53) = sum p|x| - Lambda sum p #edges.
// HBM-bound: sizeof(T) per output read once.
//
// #edges is decomposed so that no thread loops over edges per element: x = hi | lo with `lo` the
// low L index bits; edges(x) = E_lo[lo] (global table, 2^L u16) + E_hi(hi) (once per CTA chunk)
// + cross(hi, lo), and cross = t0[lo & 63] + t1[lo >> 6] from two 64-entry shared tables built
// per chunk from c_v = popc(neighbours(v) & hi) for the L low wires v.
#include "qdc_common.cuh"

namespace {

constexpr int LO_BITS = 12;
constexpr int THREADS = 256;

struct PenaltyArgs {
    uint64_t adj[64];
    uint64_t node_mask;
    uint64_t flip;
    int n_bits, lo_bits;
    long long n_chunks;
};

__global__ void penalty_table_kernel(PenaltyArgs a, uint16_t* __restrict__ e_lo) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << a.lo_bits)) return;
    const uint64_t x = (uint64_t)i & a.node_mask;
    int twice = 0;
    for (uint64_t rem = x; rem; rem &= rem - 1) twice += __popcll(a.adj[__ffsll((long long)rem) - 1] & x);
    e_lo[i] = (uint16_t)(twice >> 1) | (uint16_t)(__popcll(x) << 10);      // edges < 1024, weight in the top 6 bits
}

template <typename T> struct Load;
template <> struct Load<double> { static __device__ __forceinline__ double p(const double* d, long long i) { return __ldcs(d + i); } };
template <> struct Load<float> { static __device__ __forceinline__ double p(const float* d, long long i) { return (double)__ldcs(d + i); } };
template <> struct Load<double2> {
    static __device__ __forceinline__ double p(const double2* d, long long i) {
        const double2 v = __ldcs(d + i);
        return v.x * v.x + v.y * v.y;
    }
};

template <typename T>
__global__ void __launch_bounds__(THREADS, 4)
penalty_reduce_kernel(PenaltyArgs a, const T* __restrict__ data, const uint16_t* __restrict__ e_lo,
                      double* __restrict__ partials) {
    __shared__ int t0[64], t1[64];
    __shared__ int hi_const[2];
    __shared__ double scratch[3 * THREADS / 32];
    const int L = a.lo_bits, n_lo = 1 << L;
    const uint64_t lo_flip = a.flip & (uint64_t)(n_lo - 1);
    MisAccum acc = {0.0, 0.0, 0.0};
    for (long long chunk = blockIdx.x; chunk < a.n_chunks; chunk += gridDim.x) {
        const uint64_t hi = ((((uint64_t)chunk << L) ^ a.flip) & ~(uint64_t)(n_lo - 1)) & a.node_mask;
        __syncthreads();
        if (threadIdx.x < 128) {
            const int half = threadIdx.x >> 6, e = threadIdx.x & 63;
            int s = 0;
            for (int b = 0; b < 6; ++b) {
                const int v = half * 6 + b;
                if (((e >> b) & 1) && v < L && ((a.node_mask >> v) & 1)) s += __popcll(a.adj[v] & hi);
            }
            (half ? t1 : t0)[e] = s;
        } else if (threadIdx.x == 128) {
            int twice = 0;
            for (uint64_t rem = hi; rem; rem &= rem - 1) twice += __popcll(a.adj[__ffsll((long long)rem) - 1] & hi);
            hi_const[0] = twice >> 1;
            hi_const[1] = __popcll(hi);
        }
        __syncthreads();
        const int e_hi = hi_const[0], w_hi = hi_const[1];
        const T* base = data + (chunk << L);
        for (int i = threadIdx.x; i < n_lo; i += THREADS) {
            const double p = Load<T>::p(base, i);
            const int lo = i ^ (int)lo_flip;
            const int tab = e_lo[lo];
            const int edges = (tab & 1023) + e_hi + t0[lo & 63] + t1[lo >> 6];
            const int w = (tab >> 10) + w_hi;
            acc.s_p += p;
            acc.s_pwok += p * (double)edges;
            acc.s_pw += p * (double)w;
        }
    }
    mis_block_store(acc, scratch, partials);
}

}  // namespace

extern "C" int qdc_penalty_workspace_bytes(void) { return (1 << LO_BITS) * (int)sizeof(uint16_t); }

extern "C" int qdc_penalty_reduce(const void* d_P, int dtype, int n_bits, uint64_t index_flip,
                                  const uint64_t* h_adj, int n_wires, uint64_t node_mask,
                                  void* d_workspace, double* d_sums, double* d_partials, void* stream) {
    QDC_REQUIRE(d_P && h_adj && d_workspace && d_sums && d_partials, "null pointer");
    QDC_REQUIRE(dtype >= 0 && dtype <= 2, "dtype: 0 f64, 1 f32, 2 complex128 amplitudes");
    QDC_REQUIRE(n_bits >= 1 && n_bits <= 40, "n_bits");
    QDC_REQUIRE(n_wires >= 0 && n_wires <= 64, "n_wires");
    QDC_REQUIRE(n_bits == 64 || (index_flip >> n_bits) == 0, "index_flip");
    PenaltyArgs a;
    for (int i = 0; i < 64; ++i) a.adj[i] = i < n_wires ? (h_adj[i] & node_mask) : 0ull;
    for (int i = 0; i < n_wires; ++i) QDC_REQUIRE(__builtin_popcountll(a.adj[i]) < 1024, "degree");
    a.node_mask = node_mask;
    a.flip = index_flip;
    a.n_bits = n_bits;
    a.lo_bits = n_bits < LO_BITS ? n_bits : LO_BITS;
    a.n_chunks = 1ll << (n_bits - a.lo_bits);
    cudaStream_t s = (cudaStream_t)stream;
    uint16_t* e_lo = (uint16_t*)d_workspace;
    penalty_table_kernel<<<((1 << a.lo_bits) + 255) / 256, 256, 0, s>>>(a, e_lo);
    QDC_CHECK_LAUNCH();
    long long want = (long long)qdc_sm_count() * 8;
    int grid = (int)(a.n_chunks < want ? a.n_chunks : want);
    if (grid > QDC_MAX_PARTIALS) grid = QDC_MAX_PARTIALS;
    if (dtype == 0) penalty_reduce_kernel<double><<<grid, THREADS, 0, s>>>(a, (const double*)d_P, e_lo, d_partials);
    else if (dtype == 1) penalty_reduce_kernel<float><<<grid, THREADS, 0, s>>>(a, (const float*)d_P, e_lo, d_partials);
    else penalty_reduce_kernel<double2><<<grid, THREADS, 0, s>>>(a, (const double2*)d_P, e_lo, d_partials);
    QDC_CHECK_LAUNCH();
    return qdc_launch_finalize_sums(d_partials, grid, d_sums, s);
}
