// Operand preparation for the int8-sliced contraction (contract_i8.cu) and the small kernels around
// it: everything the 4^K contraction needs from an FP64 fragment tensor in ONE pass over it.
//
// Replaces, on the int8 path, the separate passes of round 1 (column sums for combined_norm,
// qsplit_mlrecon_methods.py:471-472; slicing; the moments that give sum p and sum p w; the row
// maxima the a-posteriori error bounds need) and the host synchronisation that picked the slice
// count: `i8_prep_kernel` reads the operand once, `i8_plan_kernel` (one CTA) turns the per-CTA
// partials into the normalisation, the two unmasked sums, the error bounds and the slice count --
// all on the device, nothing is read back.
//
// Slicing (see contract_i8.cu for the arithmetic): per row r a power-of-two exponent e_r with
// |x| 2^(7 - e_r) <= 127; v = rint(x 2^(7 - e_r) 256^(S-1)) is a signed integer of at most 7 + 8 (S-1)
// bits whose balanced radix-256 digits (two's complement of v + 128 (256^(S-2) + ... + 1), low S-1
// bytes minus 128) are the S slices.  Truncating to the first S' < S slices leaves a remainder of at
// most 128.5 units of the last kept digit (rounding directly to S' digits: 128), which the stated
// bound (S' + 1.1) kdim 2^(-8 S') covers, so one 6-slice buffer serves every slice count.
//
// Column role (the operand streamed through the tiles): an epilogue thread sums 16 or 32 adjacent
// columns in integers, so those columns need one common unit.  Every 32-row group of the operand
// gets the exponent of its largest row as reference; a row keeps its own exponent down to 2^-20 of
// that (the epilogue shifts its terms left by lsh = 20 - (ref - e_r)), smaller rows are sliced with
// exponent ref - 20 (they lose relative precision, their absolute error stays that of a row at
// 2^-20 of the group's largest; the statistics below use the exponent actually applied).
#include "qdc_common.cuh"
#include "i8_layout.cuh"
#include <mutex>

namespace {

constexpr int PREP_THREADS = 256;
constexpr int PREP_TILE = 32;                           // rows per CTA pass = one column group
constexpr uint64_t FLAG_ROW_BAD = 1ull << 63, FLAG_COL_BAD = 1ull << 62;
constexpr int E_ZERO = -(1 << 20);                      // exponent of an all-zero row / group

struct PrepArgs {
    const double* X;
    long long n_rows, rows_pad;
    int kdim, kd, S, vec, grouped;
    const uint32_t* wok;        // weight | ok << 8 per row (nullptr: weight 0, ok)
    const uint64_t* xbits;      // circuit-order bits per row (column role) or nullptr
    uint8_t* slices;
    size_t slice_stride;
    double* scale;
    double2* colq;
    uint4* coltab;
    int* gref;
    double* part;
};

__device__ __forceinline__ double2 ld_stream2(const double* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ double pow2(int e) {         // 2^e for e in the normal range
    return __hiloint2double((e + 1023) << 20, 0);
}

__global__ void __launch_bounds__(PREP_THREADS, 2) i8_prep_kernel(const PrepArgs a) {
    extern __shared__ __align__(16) uint8_t tile[];     // [S][32 * kd] digits (16-byte slots swizzled); later the moment reduction
    __shared__ int rowE[PREP_TILE];
    __shared__ double red_stat[PREP_THREADS / 32][3];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kd = a.kd, S = a.S, cpg = kd >> 4;        // 16-byte k chunks per 8-row group
    const int tile_bytes = PREP_TILE * kd;
    long long off = 0;                                  // 128 (256^(S-2) + ... + 1)
    for (int i = 1; i < S; ++i) off = off * 256 + 128;
    double s[8], w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i] = 0.0; w[i] = 0.0; }
    double st_max = 0.0, st_sum = 0.0;
    int st_emax = E_ZERO;

    const long long n_tiles = a.rows_pad / PREP_TILE;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const long long r0 = t * PREP_TILE + 4 * warp;
        double x[4][8];
        int e[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const long long r = r0 + q;
            const double* row = a.X + r * a.kdim;
            uint32_t hi = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = 2 * lane + 64 * j;
                double2 v = make_double2(0.0, 0.0);
                if (r < a.n_rows && k < a.kdim) {
                    if (a.vec) v = ld_stream2(row + k);
                    else { v.x = row[k]; if (k + 1 < a.kdim) v.y = row[k + 1]; }
                }
                x[q][2 * j] = v.x; x[q][2 * j + 1] = v.y;
                hi = max(hi, max((uint32_t)__double2hiint(fabs(v.x)), (uint32_t)__double2hiint(fabs(v.y))));
            }
            hi = __reduce_max_sync(0xffffffffu, hi);    // high word of the row maximum: exponent + 20 mantissa bits
            // mx = m 2^ee with m in [0.5, 1); one more when m 128 could exceed 127 (decided on the
            // 20 leading mantissa bits, conservatively)
            const int be = (int)(hi >> 20);
            int ee = be - 1022 + ((hi & 0xFFFFFu) >= 0xFC000u ? 1 : 0);
            if (be == 0 || ee < -900) ee = E_ZERO;      // zero (or vanishing) row: all digits 0
            e[q] = ee;
            if (lane == 0) rowE[4 * warp + q] = ee;
        }
        __syncthreads();
        int gmax = E_ZERO;
        if (a.grouped) {                                    // one group = the 32 rows of this pass
            static_assert(PREP_TILE == I8_GROUP, "a preparation pass covers exactly one column group");
            gmax = (int)__reduce_max_sync(0xffffffffu, (unsigned)(rowE[lane] - E_ZERO)) + E_ZERO;
            if (a.gref && warp == 0 && lane == 0) a.gref[t] = gmax;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const long long r = r0 + q;
            const bool live = r < a.n_rows;
            int f = e[q], lsh = 0;
            if (a.grouped && gmax != E_ZERO) {
                const int down = (f == E_ZERO) ? I8_LSH_MAX : min(gmax - f, I8_LSH_MAX);
                f = gmax - down;
                lsh = I8_LSH_MAX - down;
            }
            const bool zero = (e[q] == E_ZERO);          // digits are all 0 either way
            const double up = zero ? 0.0 : pow2(7 - f + 8 * (S - 1));
            const uint32_t wk = (live && a.wok) ? __ldg(a.wok + r) : (live ? 0x100u : 0u);
            const double wt = (double)(wk & 0xffu);
            const int rit = 4 * warp + q, g = rit >> 3, r7 = rit & 7;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = 2 * lane + 64 * j;
                if (k >= kd) continue;
                const double x0 = x[q][2 * j], x1 = x[q][2 * j + 1];
                s[2 * j] += x0; s[2 * j + 1] += x1;
                w[2 * j] = fma(wt, x0, w[2 * j]); w[2 * j + 1] = fma(wt, x1, w[2 * j + 1]);
                const long long u0 = __double2ll_rn(x0 * up) + off, u1 = __double2ll_rn(x1 * up) + off;
                const int cc = g * cpg + (k >> 4);
                uint8_t* dst = tile + ((cc * 8 + (r7 ^ (cc & 7))) << 4) + (k & 15);
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    if (i < S) {
                        const int sh = 8 * (S - 1 - i);
                        uint32_t b0 = (uint32_t)(u0 >> sh) & 0xffu, b1 = (uint32_t)(u1 >> sh) & 0xffu;
                        if (i) { b0 ^= 0x80u; b1 ^= 0x80u; }
                        *reinterpret_cast<uint16_t*>(dst + (size_t)i * tile_bytes) = (uint16_t)(b0 | (b1 << 8));
                    }
                }
            }
            if (lane == 0) {
                const double sc = (live && !zero) ? pow2(f - 7) : 0.0;
                if (a.scale) a.scale[r] = sc;
                if (a.colq) a.colq[r] = make_double2(sc, sc * wt);
                if (a.coltab) {
                    uint64_t cx = FLAG_ROW_BAD | FLAG_COL_BAD;
                    if (live) cx = (a.xbits ? __ldg(a.xbits + r) : 0ull) | FLAG_ROW_BAD | ((wk >> 8) & 1u ? 0ull : FLAG_COL_BAD);
                    a.coltab[r] = make_uint4((uint32_t)cx, (uint32_t)(cx >> 32), (wk & 0xffu) << lsh, 1u << lsh);
                }
                if (live && !zero) {
                    const double p2 = pow2(f);
                    st_max = fmax(st_max, p2);
                    if ((wk >> 8) & 1u) st_sum += p2;
                    st_emax = max(st_emax, f);
                }
            }
        }
        __syncthreads();
        // copy out: slice i of this tile is one contiguous run of 32 * kd bytes
        const int units = tile_bytes >> 4;
        for (int i = 0; i < S; ++i) {
            uint4* dst = reinterpret_cast<uint4*>(a.slices + (size_t)i * a.slice_stride + (size_t)t * tile_bytes);
            const uint8_t* src = tile + (size_t)i * tile_bytes;
            for (int u = tid; u < units; u += PREP_THREADS) {
                const int cc = u >> 3;
                dst[u] = *reinterpret_cast<const uint4*>(src + (((cc << 3) + ((u & 7) ^ (cc & 7))) << 4));
            }
        }
        __syncthreads();
    }
    // per-CTA partials, fixed order: [2][kd] column sums, then the statistics
    double* red = reinterpret_cast<double*>(tile);      // [8 warps][2][kd]
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k = 2 * lane + 64 * j;
        if (k < kd) {
            red[(warp * 2 + 0) * kd + k] = s[2 * j]; red[(warp * 2 + 0) * kd + k + 1] = s[2 * j + 1];
            red[(warp * 2 + 1) * kd + k] = w[2 * j]; red[(warp * 2 + 1) * kd + k + 1] = w[2 * j + 1];
        }
    }
    st_max = __shfl_sync(0xffffffffu, st_max, 0); st_sum = __shfl_sync(0xffffffffu, st_sum, 0);
    st_emax = __shfl_sync(0xffffffffu, st_emax, 0);
    if (lane == 0) { red_stat[warp][0] = st_max; red_stat[warp][1] = st_sum; red_stat[warp][2] = (double)st_emax; }
    __syncthreads();
    double* out = a.part + (size_t)blockIdx.x * (2 * kd + I8_STAT);
    for (int i = tid; i < 2 * kd; i += PREP_THREADS) {
        double acc = 0.0;
#pragma unroll
        for (int wv = 0; wv < PREP_THREADS / 32; ++wv) acc += red[wv * 2 * kd + i];
        out[i] = acc;
    }
    if (tid == 0) {
        double m = 0.0, sm = 0.0, em = (double)E_ZERO;
        for (int wv = 0; wv < PREP_THREADS / 32; ++wv) {
            m = fmax(m, red_stat[wv][0]); sm += red_stat[wv][1]; em = fmax(em, red_stat[wv][2]);
        }
        out[2 * kd + 0] = m; out[2 * kd + 1] = sm; out[2 * kd + 2] = em; out[2 * kd + 3] = 0.0;
    }
}

// ---- column sums alone (rows of the row operand outside this rank's slab; the FP64 path's norm) ----
// partial[cta][k] = sum over this CTA's rows of X[r][k]; every thread streams 16-byte loads, 8 in flight
constexpr int COLSUM_THREADS = 256;

__global__ void __launch_bounds__(COLSUM_THREADS) colsum_kernel(const double* __restrict__ X, long long r_begin, long long r_end,
                                                                long long r2_begin, long long r2_end, int kdim, int vec,
                                                                double* __restrict__ partial) {
    extern __shared__ double sh[];                      // [COLSUM_THREADS * 2]
    const int tid = threadIdx.x;
    const int cw = vec ? 2 : 1;                         // columns per thread
    for (int k0 = 0; k0 < kdim; k0 += COLSUM_THREADS * cw) {
        const int kw = min(kdim - k0, COLSUM_THREADS * cw);      // columns this pass
        const int tpr = (kw + cw - 1) / cw;             // threads per row
        const int rpp = COLSUM_THREADS / tpr;           // rows per pass of the CTA
        const int c = (tid % tpr) * cw, sub = tid / tpr;
        double s0 = 0.0, s1 = 0.0;
        if (sub < rpp) {
            for (int range = 0; range < 2; ++range) {
                const long long rb = range ? r2_begin : r_begin, re = range ? r2_end : r_end;
                const long long step = (long long)gridDim.x * rpp;
                long long r = rb + (long long)blockIdx.x * rpp + sub;
                for (; r + 7 * step < re; r += 8 * step) {
                    double2 v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const double* p = X + (r + u * step) * kdim + k0 + c;
                        if (vec) v[u] = ld_stream2(p); else { v[u].x = *p; v[u].y = 0.0; }
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) { s0 += v[u].x; s1 += v[u].y; }
                }
                for (; r < re; r += step) {
                    const double* p = X + r * kdim + k0 + c;
                    if (vec) { const double2 v = ld_stream2(p); s0 += v.x; s1 += v.y; } else s0 += *p;
                }
            }
        }
        sh[2 * tid] = s0; sh[2 * tid + 1] = s1;
        __syncthreads();
        if (tid < tpr) {
            double t0 = 0.0, t1 = 0.0;
            for (int j = 0; j < rpp; ++j) { t0 += sh[2 * (j * tpr + tid)]; t1 += sh[2 * (j * tpr + tid) + 1]; }
            partial[(size_t)blockIdx.x * kdim + k0 + c] = t0;
            if (vec) partial[(size_t)blockIdx.x * kdim + k0 + c + 1] = t1;
        }
        __syncthreads();
    }
}

// norm = scale * sum_k (sum_parts pa[.][k]) (sum_parts pb[.][k]); one CTA
__global__ void __launch_bounds__(1024) norm_finalize_kernel(const double* __restrict__ pa, const double* __restrict__ pb,
                                                             int n_part, int kdim, double scale, double* __restrict__ norm) {
    __shared__ double sh[32];
    double acc = 0.0;
    for (int k = threadIdx.x; k < kdim; k += blockDim.x) {
        double ca = 0.0, cb = 0.0;
#pragma unroll 8
        for (int b = 0; b < n_part; ++b) { ca += pa[(size_t)b * kdim + k]; cb += pb[(size_t)b * kdim + k]; }
        acc += ca * cb;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int wv = 0; wv < (int)(blockDim.x >> 5); ++wv) t += sh[wv];
        *norm = t * scale;
    }
}

// ---- plan: partials -> normalisation, unmasked sums, error bounds, slice count (one CTA) ----------
struct PlanArgs {
    const double* partA;        // [I8_PREP_GRID][2 kd + I8_STAT]  (this rank's slab of the row operand)
    const double* partB;        // the column operand
    const double* extra;        // [n_extra][kdim_extra] column sums of row-operand rows outside the slab, or nullptr
    int n_extra, kdim_extra;    // kdim_extra: the contraction length itself (kd is its padding to 32)
    int kd;
    double scale;
    int norm_mode;              // 0: none, 1: from the operands, 2: *norm_in
    const double* norm_in;
    int n_slices, n_stored;     // n_slices 0: smallest count in 3..n_stored whose bounds are below tol
    double tol;
    int n_nodes;
    long long n_cols;
    double* res;                // I8_RES_DOUBLES
    double* partials;           // masked-sum partials of the contraction kernel: zeroed here
    int n_partials;
};

__global__ void __launch_bounds__(1024) i8_plan_kernel(const PlanArgs a) {
    __shared__ double col[5][256];                      // sA, wA, sB, wB, sExtra per k
    __shared__ double red[3][32];
    __shared__ double stat[4];
    const int tid = threadIdx.x, kd = a.kd, stride = 2 * kd + I8_STAT;
    for (int i = tid; i < 3 * a.n_partials; i += blockDim.x) a.partials[i] = 0.0;
    // thread (k, which) sums one column of one array over the partial rows, 8 loads in flight
    for (int i = tid; i < 5 * kd; i += blockDim.x) {
        const int which = i / kd, k = i - which * kd;
        double acc = 0.0;
        if (which < 4) {
            const double* p = (which < 2 ? a.partA : a.partB) + (which & 1) * kd + k;
#pragma unroll 8
            for (int b = 0; b < I8_PREP_GRID; ++b) acc += p[(size_t)b * stride];
        } else if (a.extra && k < a.kdim_extra) {
#pragma unroll 8
            for (int b = 0; b < a.n_extra; ++b) acc += a.extra[(size_t)b * a.kdim_extra + k];
        }
        col[which][k] = acc;
    }
    if (tid < 64) {                                     // statistics: warp 0 -> A, warp 1 -> B
        const double* p = (tid < 32 ? a.partA : a.partB) + 2 * kd;
        double m = 0.0, sm = 0.0, em = -1e300;
        for (int b = tid & 31; b < I8_PREP_GRID; b += 32) {
            m = fmax(m, p[(size_t)b * stride]); sm += p[(size_t)b * stride + 1]; em = fmax(em, p[(size_t)b * stride + 2]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            sm += __shfl_xor_sync(0xffffffffu, sm, o);
            em = fmax(em, __shfl_xor_sync(0xffffffffu, em, o));
        }
        if ((tid & 31) == 0) { red[0][tid >> 5] = m; red[1][tid >> 5] = sm; red[2][tid >> 5] = em; }
    }
    __syncthreads();
    if (tid < 64) {
        if (tid == 0) { stat[0] = red[0][0] * red[0][1]; stat[1] = red[1][0] * red[1][1]; stat[2] = red[2][1]; stat[3] = red[0][1] * red[1][0]; }
    }
    __syncthreads();
    // three dot products over k: slab total, all-rows total, weighted slab total
    double d_slab = 0.0, d_all = 0.0, d_w = 0.0;
    for (int k = tid; k < kd; k += blockDim.x) {
        const double sA = col[0][k], wA = col[1][k], sB = col[2][k], wB = col[3][k], sX = col[4][k];
        d_slab += sA * sB;
        d_all += (sA + sX) * sB;
        d_w += fma(wA, sB, sA * wB);
    }
    d_slab = warp_sum(d_slab); d_all = warp_sum(d_all); d_w = warp_sum(d_w);
    __syncthreads();
    if ((tid & 31) == 0) { red[0][tid >> 5] = d_slab; red[1][tid >> 5] = d_all; red[2][tid >> 5] = d_w; }
    __syncthreads();
    if (tid == 0) {
        double t_slab = 0.0, t_all = 0.0, t_w = 0.0;
        for (int wv = 0; wv < (int)(blockDim.x >> 5); ++wv) { t_slab += red[0][wv]; t_all += red[1][wv]; t_w += red[2][wv]; }
        double norm = 1.0;
        if (a.norm_mode == 1) norm = a.scale * t_all;
        else if (a.norm_mode == 2) norm = *a.norm_in;
        const double f = a.scale / norm;
        a.res[I8_RES_SUMS + 0] = f * t_slab;
        a.res[I8_RES_SUMS + 1] = 0.0;
        a.res[I8_RES_SUMS + 2] = f * t_w;
        a.res[I8_RES_NORM] = norm;
        // a-posteriori bounds (contract_i8.cu): per output 2^(e_r + f_c) |f| kdim (S + 1.1) 2^(-8S) at the
        // largest row and column; on the masked sum every output counts at most n_nodes times and
        // only independent rows and columns enter; the integer fold of the epilogue adds at most
        // half a unit of the last order (2^-6 of 2^(e_r + f_max) |f| 2^(-8S)) per 16 columns
        const double af = fabs(f), fold = ldexp(0.5 * (double)((a.n_cols + I8_FOLD_COLS - 1) / I8_FOLD_COLS), -6 - I8_LSH_MAX);
        int chosen = a.n_slices;
        double b_out = 0.0, b_cost = 0.0;
        for (int S = (a.n_slices ? a.n_slices : 3); S <= (a.n_slices ? a.n_slices : a.n_stored); ++S) {
            const double g = ldexp((double)a.kdim_extra * ((double)S + 1.1), -8 * S), u = ldexp(1.0, -8 * S);
            b_out = stat[0] * af * g;
            b_cost = (double)a.n_nodes * af * (stat[1] * g + stat[3] * fold * u);
            chosen = S;
            if (b_out <= a.tol && b_cost <= a.tol) break;
        }
        a.res[I8_RES_SLICES] = (double)chosen;
        a.res[I8_RES_BOUND_OUT] = b_out;
        a.res[I8_RES_BOUND_COST] = b_cost;
        reinterpret_cast<unsigned long long*>(a.res)[I8_RES_CAND] = 0ull;
        a.res[I8_RES_EMAX] = stat[2];
        for (int i = I8_RES_EMAX + 1; i < I8_RES_DOUBLES; ++i) a.res[i] = 0.0;
    }
}

}  // namespace

// ---- C-ABI -----------------------------------------------------------------------------------
extern "C" int64_t qdc_i8_operand_bytes(int64_t n_rows, int kdim, int n_slices) {
    if (n_rows < 0 || kdim < 1 || kdim > 256 || n_slices < 2 || n_slices > 6) return 0;
    return (int64_t)i8_layout(n_rows, kdim, n_slices).total;
}

extern "C" int qdc_i8_prepare(const double* d_X, int64_t n_rows, int kdim, int n_slices, int column_role,
                              const uint32_t* d_wok, const uint64_t* d_x, void* d_operand, void* stream) {
    QDC_REQUIRE(d_X && d_operand, "null buffer");
    QDC_REQUIRE(n_rows >= 1 && kdim >= 1 && kdim <= 256, "shape (the int8 path keeps a row block resident: kdim <= 256)");
    QDC_REQUIRE(n_slices >= 2 && n_slices <= 6, "n_slices in 2..6");
    QDC_REQUIRE((((uintptr_t)d_operand) & 255) == 0, "operand buffer must be 256-byte aligned");
    const I8Layout L = i8_layout(n_rows, kdim, n_slices);
    uint8_t* base = (uint8_t*)d_operand;
    PrepArgs a;
    a.X = d_X; a.n_rows = n_rows; a.rows_pad = L.rows_pad; a.kdim = kdim; a.kd = L.kd; a.S = n_slices;
    a.vec = (kdim % 2 == 0) && ((((uintptr_t)d_X) & 15) == 0);
    a.grouped = column_role ? 1 : 0;
    a.wok = d_wok; a.xbits = d_x;
    a.slices = base + L.off_slices; a.slice_stride = L.slice_stride;
    a.scale = (double*)(base + L.off_scale);
    a.colq = (double2*)(base + L.off_colq);
    a.coltab = (uint4*)(base + L.off_coltab);
    a.gref = (int*)(base + L.off_gref);
    a.part = (double*)(base + L.off_part);
    const size_t smem = (size_t)L.kd * (PREP_TILE * n_slices > 128 ? PREP_TILE * n_slices : 128);
    QDC_CHECK_CUDA(cudaFuncSetAttribute(i8_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    i8_prep_kernel<<<I8_PREP_GRID, PREP_THREADS, smem, (cudaStream_t)stream>>>(a);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_colsum_partials(void) { return I8_PREP_GRID; }

extern "C" int qdc_colsum(const double* d_X, int64_t r0_begin, int64_t r0_end, int64_t r1_begin, int64_t r1_end,
                          int kdim, double* d_partials, void* stream) {
    QDC_REQUIRE(d_X && d_partials, "null buffer");
    QDC_REQUIRE(kdim >= 1 && r0_begin >= 0 && r0_end >= r0_begin && r1_begin >= 0 && r1_end >= r1_begin, "shape");
    const int vec = (kdim % 2 == 0) && ((((uintptr_t)d_X) & 15) == 0);
    colsum_kernel<<<I8_PREP_GRID, COLSUM_THREADS, 2 * COLSUM_THREADS * sizeof(double), (cudaStream_t)stream>>>(
        d_X, r0_begin, r0_end, r1_begin, r1_end, kdim, vec, d_partials);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_norm_workspace_doubles(int kdim) { return 2 * I8_PREP_GRID * (kdim > 0 ? kdim : 1); }

extern "C" int qdc_contract_norm(const double* d_A, int64_t n_rows_total, const double* d_Bt,
                                 int64_t n_cols, int kdim, double scale, double* d_colsums,
                                 double* d_norm, void* stream) {
    QDC_REQUIRE(d_A && d_Bt && d_colsums && d_norm, "null buffer");
    QDC_REQUIRE(kdim >= 1 && n_rows_total >= 0 && n_cols >= 0, "shape");
    double* pa = d_colsums;
    double* pb = d_colsums + (size_t)I8_PREP_GRID * kdim;
    int rc = qdc_colsum(d_A, 0, n_rows_total, 0, 0, kdim, pa, stream);
    if (rc != QDC_OK) return rc;
    rc = qdc_colsum(d_Bt, 0, n_cols, 0, 0, kdim, pb, stream);
    if (rc != QDC_OK) return rc;
    norm_finalize_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(pa, pb, I8_PREP_GRID, kdim, scale, d_norm);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

// ---- the normalisation without any allocation -----------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid it):
// the column-sum partials live in a file-scope device array instead of a caller workspace.
// 2 operands x 296 partial rows x 256 columns x 8 B = 1 212 416 bytes, held by every GPU that loads
// the library until the process exits.  All calls on a device share the array and it keeps the
// last call's contents: every entry a call reads is written by that call first (each colsum CTA
// stores its whole row of partials, zeros included), so nothing has to be cleared; two calls that
// overlap on the GPU (different streams) race on it.
constexpr int NORM_STATIC_MAX_KDIM = 256;
static __device__ __align__(256) double g_norm_scratch[2 * I8_PREP_GRID * NORM_STATIC_MAX_KDIM];

extern "C" int qdc_contract_norm_static_max_kdim(void) { return NORM_STATIC_MAX_KDIM; }

extern "C" int qdc_contract_norm_static(const double* d_A, int64_t n_rows_total, const double* d_Bt,
                                        int64_t n_cols, int kdim, double scale, double* d_norm, void* stream) {
    QDC_REQUIRE(d_A && d_Bt && d_norm, "null buffer");
    QDC_REQUIRE(kdim >= 1 && n_rows_total >= 0 && n_cols >= 0, "shape");
    if (kdim > NORM_STATIC_MAX_KDIM) {
        qdc_set_error("qdc_contract_norm_static: kdim %d exceeds the static scratch (at most %d columns); "
                      "use qdc_contract_norm with a caller-owned workspace", kdim, NORM_STATIC_MAX_KDIM);
        return QDC_ERR_ARG;
    }
    // In host code `g_norm_scratch` is only a stand-in for the symbol, not a device address: the
    // address is looked up once per device, with that device current, and cached.
    constexpr int MAX_DEVICES = 64;
    static double* scratch_of[MAX_DEVICES] = {};
    static std::mutex lock;
    int dev = -1;
    QDC_CHECK_CUDA(cudaGetDevice(&dev));
    QDC_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "device ordinal");
    double* scratch;
    {
        std::lock_guard<std::mutex> guard(lock);
        if (!scratch_of[dev]) {
            void* p = nullptr;
            QDC_CHECK_CUDA(cudaGetSymbolAddress(&p, g_norm_scratch));
            scratch_of[dev] = (double*)p;
        }
        scratch = scratch_of[dev];
    }
    double* pa = scratch;
    double* pb = scratch + (size_t)I8_PREP_GRID * kdim;
    int rc = qdc_colsum(d_A, 0, n_rows_total, 0, 0, kdim, pa, stream);
    if (rc != QDC_OK) return rc;
    rc = qdc_colsum(d_Bt, 0, n_cols, 0, 0, kdim, pb, stream);
    if (rc != QDC_OK) return rc;
    norm_finalize_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(pa, pb, I8_PREP_GRID, kdim, scale, d_norm);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

// internal: used by qdc_contract_i8_prepared (contract_i8.cu)
int qdc_i8_launch_plan(const void* d_opA, int64_t n_rows, const void* d_opB, int64_t n_cols, int kdim, int n_stored,
                       int n_slices, double tol, int n_nodes, double scale, int norm_mode, const double* d_norm_in,
                       const double* d_extra, int n_extra, double* d_res, double* d_partials, int n_partials,
                       cudaStream_t s) {
    const I8Layout LA = i8_layout(n_rows, kdim, n_stored), LB = i8_layout(n_cols, kdim, n_stored);
    PlanArgs p;
    p.partA = (const double*)((const uint8_t*)d_opA + LA.off_part);
    p.partB = (const double*)((const uint8_t*)d_opB + LB.off_part);
    p.extra = d_extra; p.n_extra = n_extra; p.kdim_extra = kdim;
    p.kd = LA.kd; p.scale = scale; p.norm_mode = norm_mode; p.norm_in = d_norm_in;
    p.n_slices = n_slices; p.n_stored = n_stored; p.tol = tol; p.n_nodes = n_nodes; p.n_cols = n_cols;
    p.res = d_res; p.partials = d_partials; p.n_partials = n_partials;
    i8_plan_kernel<<<1, 1024, 0, s>>>(p);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}
