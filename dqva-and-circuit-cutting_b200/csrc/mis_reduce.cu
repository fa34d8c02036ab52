// K8: MIS objective reduction + candidate emission + top-k (HBM-bound).
//
// Replaces, per output bitstring, strip_ancillas (utils/helper_funcs.py:3-16), hamming_weight
// (:18-19), is_indset (utils/graph_funcs.py:107-118) as summed by avg_cost (mis.py:125-126) and
// filtered/sorted by top_probs (mis.py:232-233).
//
// Algorithmic bytes: sizeof(T) per output read once; nothing written but 3 doubles per CTA.
#include <stdlib.h>

#include "qdc_common.cuh"

// ------------------------------------------------------------------------------------------
// tables
// ------------------------------------------------------------------------------------------
struct BitTable {
    int nbits;
    int bitpos[64];
    uint64_t adj[64];
    uint64_t node_mask;
};

__global__ void mis_table_kernel(BitTable t, int bit_lo, long long count, uint64_t flip, uint64_t* __restrict__ xs,
                                 uint64_t* __restrict__ forb, uint32_t* __restrict__ wok) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
         i += (long long)gridDim.x * blockDim.x) {
        uint64_t x = 0;
        for (int b = 0; b < t.nbits; ++b)
            if (((i ^ (long long)flip) >> b) & 1) x |= 1ull << t.bitpos[bit_lo + b];
        const uint64_t nodes = x & t.node_mask;
        uint64_t f = 0;
        uint64_t rem = nodes;
        while (rem) {
            const int a = __ffsll((long long)rem) - 1;
            rem &= rem - 1;
            f |= t.adj[a];
        }
        const uint32_t ok = (nodes & f) == 0ull ? 1u : 0u;
        xs[i] = x;
        if (forb) forb[i] = f & t.node_mask;
        wok[i] = (uint32_t)__popcll((long long)nodes) | (ok << 8);
    }
}

extern "C" int qdc_mis_build_tables(int row_bits, int col_bits, const int* h_bitpos,
                                    const uint64_t* h_adj, int n_wires, uint64_t node_mask,
                                    uint64_t index_flip, uint64_t* d_row_x, uint64_t* d_row_forb, uint32_t* d_row_wok,
                                    uint64_t* d_col_x, uint32_t* d_col_wok, void* stream) {
    QDC_REQUIRE(row_bits >= 0 && col_bits >= 0 && row_bits + col_bits <= 62, "bit counts");
    QDC_REQUIRE(n_wires >= 0 && n_wires <= 62, "n_wires");
    QDC_REQUIRE((node_mask >> 62) == 0, "node_mask uses reserved bits");
    QDC_REQUIRE(h_bitpos && h_adj, "null tables");
    QDC_REQUIRE((index_flip >> (row_bits + col_bits)) == 0, "index_flip");
    BitTable t;
    for (int i = 0; i < 64; ++i) { t.bitpos[i] = 0; t.adj[i] = 0; }
    uint64_t seen = 0;
    for (int b = 0; b < row_bits + col_bits; ++b) {
        QDC_REQUIRE(h_bitpos[b] >= 0 && h_bitpos[b] < 62, "bitpos range (wires 62, 63 are reserved flag bits)");
        QDC_REQUIRE(!((seen >> h_bitpos[b]) & 1), "bitpos must be injective");
        seen |= 1ull << h_bitpos[b];
        t.bitpos[b] = h_bitpos[b];
    }
    for (int i = 0; i < n_wires; ++i) t.adj[i] = h_adj[i];
    t.node_mask = node_mask;
    cudaStream_t s = (cudaStream_t)stream;
    const int threads = 256;
    {
        const long long n = 1ll << col_bits;
        t.nbits = col_bits;
        int blocks = (int)((n + threads - 1) / threads);
        if (blocks > 148 * 8) blocks = 148 * 8;
        mis_table_kernel<<<blocks, threads, 0, s>>>(t, 0, n, index_flip & (uint64_t)(n - 1), d_col_x, nullptr, d_col_wok);
        QDC_CHECK_LAUNCH();
    }
    {
        const long long n = 1ll << row_bits;
        t.nbits = row_bits;
        int blocks = (int)((n + threads - 1) / threads);
        if (blocks > 148 * 8) blocks = 148 * 8;
        mis_table_kernel<<<blocks, threads, 0, s>>>(t, col_bits, n, (index_flip >> col_bits) & (uint64_t)(n - 1), d_row_x, d_row_forb, d_row_wok);
        QDC_CHECK_LAUNCH();
    }
    return QDC_OK;
}

// ------------------------------------------------------------------------------------------
// reduction over a materialised slab
// ------------------------------------------------------------------------------------------
template <typename T> struct Vec;
template <> struct Vec<double> { using type = double2; static constexpr int N = 2; };
template <> struct Vec<float>  { using type = float4;  static constexpr int N = 4; };
struct Amp { double re, im; };            // complex128 amplitude: p = |amp|^2 (uncut cost, mis.py:639-665)
template <> struct Vec<Amp>    { using type = double2; static constexpr int N = 1; };

template <typename T>
__device__ __forceinline__ typename Vec<T>::type ld_stream(const T* p);
template <> __device__ __forceinline__ double2 ld_stream<double>(const double* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
template <> __device__ __forceinline__ float4 ld_stream<float>(const float* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

template <> __device__ __forceinline__ double2 ld_stream<Amp>(const Amp* p) {
    return ld_stream<double>(reinterpret_cast<const double*>(p));
}

template <typename T> __device__ __forceinline__ void unpack(const typename Vec<T>::type& v, double* o);
template <> __device__ __forceinline__ void unpack<double>(const double2& v, double* o) { o[0] = v.x; o[1] = v.y; }
template <> __device__ __forceinline__ void unpack<Amp>(const double2& v, double* o) { o[0] = v.x * v.x + v.y * v.y; }
template <> __device__ __forceinline__ void unpack<float>(const float4& v, double* o) {
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
}

// A thread owns VN adjacent columns for its whole life: their table entries (col_x, col_wok) sit in
// registers, so the steady state reads nothing but the distribution itself (16 bytes per thread
// per row, UNROLL rows in flight) plus three warp-uniform row-table words per row.  A CTA covers
// a strip of up to 256 * VN columns; when the table is narrower, the spare threads take
// neighbouring rows.  n_cols is a power of two >= VN.
template <typename T, int UNROLL>
__global__ void __launch_bounds__(256, 4) mis_reduce_kernel(
    const T* __restrict__ P, long long row_lo, long long n_rows, int col_shift,
    const double* __restrict__ d_norm, MisTables tab, CandSink sink, double* __restrict__ partials) {
    constexpr int VN = Vec<T>::N;
    using V = typename Vec<T>::type;
    __shared__ double scratch[3 * 8];
    const double inv_norm = d_norm ? 1.0 / *d_norm : 1.0;
    const long long n_cols = 1ll << col_shift;
    const long long strip_cols = n_cols < 256ll * VN ? n_cols : 256ll * VN;
    const int rows_per_pass = (int)(256ll * VN / strip_cols);
    const long long e0 = (long long)threadIdx.x * VN;
    const long long c = (long long)blockIdx.x * strip_cols + e0 % strip_cols;
    const int rsub = (int)(e0 / strip_cols);
    // per-column constants in registers: bits, weight as a double (no I2F per output), and the
    // independence test folded into one AND (flag bits 62 / 63 carry "column / row is not itself
    // independent", as in the contraction epilogue)
    constexpr uint64_t FLAG_ROW_BAD = 1ull << 63, FLAG_COL_BAD = 1ull << 62;
    uint64_t cx[VN], cf[VN];
    double cwd[VN];
#pragma unroll
    for (int j = 0; j < VN; ++j) {
        cx[j] = __ldg(tab.col_x + c + j);
        const uint32_t cw = __ldg(tab.col_wok + c + j);
        cf[j] = cx[j] | FLAG_ROW_BAD | ((cw >> 8) & 1u ? 0ull : FLAG_COL_BAD);
        cwd[j] = (double)(cw & 0xffu);
    }
    const double thr = sink.threshold;
    MisAccum acc = {0.0, 0.0, 0.0};
    // Rows are taken in batches of RB per CTA pass; the first RB * rows_per_pass threads fold each
    // row's table words (flags, weight as a double) into shared memory once, so the streaming loop
    // pays two LDS per row instead of three global loads and an int->double conversion per thread.
    constexpr int RB = 32;
    __shared__ uint64_t sh_rfl[RB * 4];
    __shared__ double sh_rwd[RB * 4];
    const long long r_step = (long long)gridDim.y * rows_per_pass;       // rows between passes
    const int rpp = rows_per_pass > 4 ? 4 : rows_per_pass;               // (host guarantees <= 4 here)
    for (long long rb0 = (long long)blockIdx.y * rows_per_pass; rb0 < n_rows; rb0 += r_step * RB) {
        __syncthreads();
        if (threadIdx.x < RB * rpp) {
            const int k = threadIdx.x / rpp, sub = threadIdx.x % rpp;
            const long long r = rb0 + (long long)k * r_step + sub;
            uint64_t fl = 0; double wd = 0.0;
            if (r < n_rows) {
                const uint32_t rw = __ldg(tab.row_wok + row_lo + r);
                fl = __ldg(tab.row_forb + row_lo + r) | FLAG_COL_BAD | ((rw >> 8) & 1u ? 0ull : FLAG_ROW_BAD);
                wd = (double)(rw & 0xffu);
            }
            sh_rfl[threadIdx.x] = fl; sh_rwd[threadIdx.x] = wd;
        }
        __syncthreads();
        for (int k0 = 0; k0 < RB; k0 += UNROLL) {
            V vals[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const long long r = rb0 + (long long)(k0 + u) * r_step + rsub;
                if (r < n_rows) vals[u] = ld_stream<T>(P + ((r << col_shift) + c));
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const long long r = rb0 + (long long)(k0 + u) * r_step + rsub;
                if (r >= n_rows) break;
                const uint64_t rfl = sh_rfl[(k0 + u) * rpp + rsub];
                const double rwd = sh_rwd[(k0 + u) * rpp + rsub];
                double p[VN];
                unpack<T>(vals[u], p);
#pragma unroll
                for (int j = 0; j < VN; ++j) {
                    const double pj = p[j] * inv_norm, w = rwd + cwd[j];
                    acc.s_p += pj;
                    acc.s_pw = fma(pj, w, acc.s_pw);
                    if ((cf[j] & rfl) == 0ull) acc.s_pwok = fma(pj, w, acc.s_pwok);
                    if (pj > thr) cand_emit(sink, __ldg(tab.row_x + row_lo + r) | cx[j], pj);
                }
            }
        }
    }
    // one partial triple per CTA, linear CTA index
    const int cta = blockIdx.y * gridDim.x + blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double a0 = warp_sum(acc.s_p), a1 = warp_sum(acc.s_pwok), a2 = warp_sum(acc.s_pw);
    if (lane == 0) { scratch[warp] = a0; scratch[8 + warp] = a1; scratch[16 + warp] = a2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double x = 0, y = 0, z = 0;
        for (int w = 0; w < 8; ++w) { x += scratch[w]; y += scratch[8 + w]; z += scratch[16 + w]; }
        partials[3 * cta] = x; partials[3 * cta + 1] = y; partials[3 * cta + 2] = z;
    }
}

// scalar fallback for tiny tables (n_cols < vector width)
template <typename T> __device__ __forceinline__ double prob_of(const T& v) { return (double)v; }
template <> __device__ __forceinline__ double prob_of<Amp>(const Amp& v) { return v.re * v.re + v.im * v.im; }

template <typename T>
__global__ void __launch_bounds__(256) mis_reduce_small_kernel(
    const T* __restrict__ P, long long row_lo, long long n_rows, int col_shift,
    const double* __restrict__ d_norm, MisTables tab, CandSink sink, double* __restrict__ partials) {
    __shared__ double scratch[3 * 8];
    const double inv_norm = d_norm ? 1.0 / *d_norm : 1.0;
    const long long n = n_rows << col_shift;
    MisAccum acc = {0.0, 0.0, 0.0};
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x) {
        const long long r = row_lo + (e >> col_shift), c = e & ((1ll << col_shift) - 1);
        mis_accumulate(acc, prob_of<T>(P[e]) * inv_norm, tab.row_x[r], tab.row_forb[r], tab.row_wok[r],
                       tab.col_x[c], tab.col_wok[c], sink);
    }
    mis_block_store(acc, scratch, partials);
}

__global__ void finalize_sums_kernel(const double* __restrict__ partials, int n, double* sums) {
    __shared__ double scratch[3 * 8];
    double a = 0, b = 0, c = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        a += partials[3 * i]; b += partials[3 * i + 1]; c += partials[3 * i + 2];
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    if (lane == 0) { scratch[warp] = a; scratch[8 + warp] = b; scratch[16 + warp] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double x = 0, y = 0, z = 0;
        for (int w = 0; w < 8; ++w) { x += scratch[w]; y += scratch[8 + w]; z += scratch[16 + w]; }
        sums[0] += x; sums[1] += y; sums[2] += z;
    }
}

int qdc_launch_finalize_sums(const double* d_partials, int n_partials, double* d_sums,
                             cudaStream_t stream) {
    finalize_sums_kernel<<<1, 256, 0, stream>>>(d_partials, n_partials, d_sums);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_reduce_partials(void) { return QDC_MAX_PARTIALS; }

extern "C" int qdc_mis_reduce(const void* d_P, int dtype, int64_t row_lo, int64_t n_rows,
                              int64_t n_cols, const double* d_norm, double threshold,
                              const uint64_t* d_row_x, const uint64_t* d_row_forb,
                              const uint32_t* d_row_wok, const uint64_t* d_col_x,
                              const uint32_t* d_col_wok, double* d_sums, uint64_t* d_cand_x,
                              double* d_cand_p, unsigned long long* d_cand_count,
                              int64_t cand_cap, double* d_partials, void* stream) {
    QDC_REQUIRE(d_P && d_sums && d_partials, "null buffer");
    QDC_REQUIRE(dtype >= 0 && dtype <= 2, "dtype");
    QDC_REQUIRE(n_rows >= 0 && n_cols > 0 && (n_cols & (n_cols - 1)) == 0, "n_cols power of two");
    QDC_REQUIRE(d_cand_x && d_cand_p && d_cand_count && cand_cap >= 0, "candidate sink");
    if (n_rows == 0) return QDC_OK;
    int col_shift = 0;
    while ((1ll << col_shift) < n_cols) ++col_shift;
    MisTables tab = {d_row_x, d_row_forb, d_row_wok, d_col_x, d_col_wok};
    CandSink sink = {d_cand_x, d_cand_p, d_cand_count, (long long)cand_cap, threshold};
    cudaStream_t s = (cudaStream_t)stream;
    const long long total = n_rows * n_cols;
    const int vn = dtype == 0 ? 2 : dtype == 1 ? 4 : 1;
    int blocks;
    QDC_REQUIRE(dtype != 2 || (((uintptr_t)d_P) & 15) == 0, "complex128 input must be 16-byte aligned");
    if (n_cols >= vn && n_cols * 4 >= 256ll * vn && (((uintptr_t)d_P) & 15) == 0) {
        constexpr int UNROLL = 4;
        const long long strip_cols = n_cols < 256ll * vn ? n_cols : 256ll * vn;
        const long long gx = n_cols / strip_cols;
        const int rows_per_pass = (int)(256ll * vn / strip_cols);
        QDC_REQUIRE(gx <= QDC_MAX_PARTIALS, "table too wide for one launch");
        // one resident wave (4 CTAs of 256 threads per SM): no tail, every CTA streams rows
        long long gy = ((long long)qdc_sm_count() * 4) / gx * (dtype == 1 ? 3 : 2);   // measured best, profiles/r1_k8_tuning.log
        if (const char* env = getenv("QDC_K8_WAVES")) gy = ((long long)qdc_sm_count() * 4) / gx * atoi(env);
        const long long gy_need = (n_rows + (long long)rows_per_pass * UNROLL - 1) / ((long long)rows_per_pass * UNROLL);
        if (gy > gy_need) gy = gy_need;
        if (gy * gx > QDC_MAX_PARTIALS) gy = QDC_MAX_PARTIALS / gx;
        if (gy < 1) gy = 1;
        if (gy > 65535) gy = 65535;
        blocks = (int)(gx * gy);
        const dim3 grid((unsigned)gx, (unsigned)gy);
        if (dtype == 0)
            mis_reduce_kernel<double, UNROLL><<<grid, 256, 0, s>>>(
                (const double*)d_P, row_lo, n_rows, col_shift, d_norm, tab, sink, d_partials);
        else if (dtype == 2)
            mis_reduce_kernel<Amp, UNROLL><<<grid, 256, 0, s>>>(
                (const Amp*)d_P, row_lo, n_rows, col_shift, d_norm, tab, sink, d_partials);
        else
            mis_reduce_kernel<float, UNROLL><<<grid, 256, 0, s>>>(
                (const float*)d_P, row_lo, n_rows, col_shift, d_norm, tab, sink, d_partials);
    } else {
        blocks = (int)((total + 255) / 256);
        if (blocks > 1024) blocks = 1024;
        if (dtype == 0)
            mis_reduce_small_kernel<double><<<blocks, 256, 0, s>>>(
                (const double*)d_P, row_lo, n_rows, col_shift, d_norm, tab, sink, d_partials);
        else if (dtype == 2)
            mis_reduce_small_kernel<Amp><<<blocks, 256, 0, s>>>(
                (const Amp*)d_P, row_lo, n_rows, col_shift, d_norm, tab, sink, d_partials);
        else
            mis_reduce_small_kernel<float><<<blocks, 256, 0, s>>>(
                (const float*)d_P, row_lo, n_rows, col_shift, d_norm, tab, sink, d_partials);
    }
    QDC_CHECK_LAUNCH();
    return qdc_launch_finalize_sums(d_partials, blocks, d_sums, s);
}

// ------------------------------------------------------------------------------------------
// top-k: k rounds of a single-CTA arg-max with (p desc, x asc) order
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ bool better(double p, uint64_t x, double q, uint64_t y) {
    return p > q || (p == q && x < y);
}

// optionally first: sums[j] += sum_i partials[3 i + j] in a fixed order (what finalize_sums_kernel
// does), so that scoring kernels need one finishing launch, not two
__global__ void __launch_bounds__(1024) topk_kernel(const uint64_t* __restrict__ cx, double* cp,
                                                    const unsigned long long* __restrict__ count,
                                                    long long cap, int k, uint64_t* top_x,
                                                    double* top_p, const double* __restrict__ partials,
                                                    int n_partials, double* sums) {
    __shared__ double sp[32];
    __shared__ uint64_t sx[32];
    __shared__ long long si[32];
    if (partials) {
        __shared__ double fs[3 * 32];
        double a = 0, b = 0, c = 0;
        for (int i = threadIdx.x; i < n_partials; i += blockDim.x) {
            a += partials[3 * i]; b += partials[3 * i + 1]; c += partials[3 * i + 2];
        }
        a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
        if ((threadIdx.x & 31) == 0) { fs[threadIdx.x >> 5] = a; fs[32 + (threadIdx.x >> 5)] = b; fs[64 + (threadIdx.x >> 5)] = c; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double x = 0, y = 0, z = 0;
            for (int w = 0; w < 32; ++w) { x += fs[w]; y += fs[32 + w]; z += fs[64 + w]; }
            sums[0] += x; sums[1] += y; sums[2] += z;
        }
    }
    const long long n = (long long)*count < cap ? (long long)*count : cap;
    const double NEG = -1.0 / 0.0;
    for (int round = 0; round < k; ++round) {
        double bp = NEG; uint64_t bx = ~0ull; long long bi = -1;
        for (long long i = threadIdx.x; i < n; i += blockDim.x) {
            const double p = cp[i]; const uint64_t x = cx[i];
            if (p != NEG && (bi < 0 || better(p, x, bp, bx))) { bp = p; bx = x; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double q = __shfl_xor_sync(0xffffffffu, bp, o);
            const uint64_t y = __shfl_xor_sync(0xffffffffu, bx, o);
            const long long j = __shfl_xor_sync(0xffffffffu, bi, o);
            if (j >= 0 && (bi < 0 || better(q, y, bp, bx))) { bp = q; bx = y; bi = j; }
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) { sp[warp] = bp; sx[warp] = bx; si[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            bp = sp[lane]; bx = sx[lane]; bi = si[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double q = __shfl_xor_sync(0xffffffffu, bp, o);
                const uint64_t y = __shfl_xor_sync(0xffffffffu, bx, o);
                const long long j = __shfl_xor_sync(0xffffffffu, bi, o);
                if (j >= 0 && (bi < 0 || better(q, y, bp, bx))) { bp = q; bx = y; bi = j; }
            }
            if (lane == 0) {
                if (bi >= 0) { top_x[round] = bx; top_p[round] = bp; cp[bi] = NEG; }
                else { top_x[round] = ~0ull; top_p[round] = NEG; }
            }
        }
        __syncthreads();
    }
}

extern "C" int qdc_topk_select(const uint64_t* d_cand_x, double* d_cand_p,
                               const unsigned long long* d_cand_count, int64_t cand_cap, int k,
                               uint64_t* d_top_x, double* d_top_p, void* stream) {
    QDC_REQUIRE(k >= 1 && k <= 1024, "k");
    QDC_REQUIRE(cand_cap >= 0 && d_top_x && d_top_p && d_cand_count, "buffers");
    QDC_REQUIRE(d_cand_x && d_cand_p, "candidates");
    topk_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(d_cand_x, d_cand_p, d_cand_count, cand_cap, k,
                                                      d_top_x, d_top_p, nullptr, 0, nullptr);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

// finalize the partial sums and select the top-k in one launch (k = 0: sums only)
int qdc_launch_finish(const double* d_partials, int n_partials, double* d_sums, const uint64_t* d_cand_x,
                      double* d_cand_p, const unsigned long long* d_cand_count, long long cand_cap, int k,
                      uint64_t* d_top_x, double* d_top_p, cudaStream_t stream) {
    topk_kernel<<<1, 1024, 0, stream>>>(d_cand_x, d_cand_p, d_cand_count, cand_cap, k, d_top_x, d_top_p,
                                        d_partials, n_partials, d_sums);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}
