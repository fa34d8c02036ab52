// K4b: maximum_likelihood_model on the device -- the nearest positive-semidefinite fragment model
// (qsplit/qsplit_mlrecon_methods.py:320-356, arXiv:1106.5458), replacing the batched library eigh of
// round 1 with kernels of this repository.
//
// Reference algorithm, per fragment: eigen-decompose every Choi block Lambda_b (D x D Hermitian, D =
// 2^(cut ends) <= 16); sort ALL eigenvalues of the fragment; walk up from the smallest one: zero
// it and spread its (negative) value evenly over the ones that are left, until the smallest one
// left is non-negative (:333-343); rebuild every block from its eigenvectors and the adjusted
// eigenvalues through to_projector (:45-46, 354-356), which conjugates: the result is
// conj(V diag V^+).
//
// The sequential walk is soft-thresholding: lambda' = max(lambda - t0, 0), where t0 >= 0 is the
// root of h(t) = sum_i max(lambda_i - t, 0) - sum_i lambda_i (piecewise linear, convex,
// decreasing; every step of the walk keeps the total, and the eigenvalues still alive have all
// received the same shift).  Newton from t = 0 climbs to the root monotonically and lands on it
// exactly once the set {lambda > t} stops changing -- a handful of global reductions, no sort.
//
// Pipeline (all launches unconditional, nothing is read back by the host):
//   1  ml_psd_kernel      per block (one warp): Lambda_b from the Pauli tensor row; Cholesky of
//                         Lambda_b + eps I with eps = 1e-14 trace: fails => the fragment needs the
//                         full treatment (one global flag).  Exact (infinite-shot) data is PSD,
//                         and then t0 = 0 and the model is just conj(Lambda): kernels 2, 3 return
//                         at once and 4 writes the conjugate.
//   2  ml_jacobi_kernel   per block (one warp): parallel-order cyclic Jacobi on the Hermitian block
//                         in shared memory (D/2 disjoint rotations per step, D-1 steps per sweep,
//                         stops at off-diagonal norm <= 1e-15 of the diagonal's); eigenvalues and
//                         eigenvectors go to the workspace.
//   3  ml_threshold_kernel x MAX_NEWTON: one Newton step each, grid-wide reduction in a fixed
//                         order (last CTA done folds the per-CTA partials): deterministic.
//   4  ml_rebuild_kernel  per block: conj(sum_k max(lambda_k - t0, 0) v_k v_k^+) back to Pauli
//                         coordinates.
// The Pauli <-> matrix change of basis uses two sparse tables made by the host from the same basis
// the direct model uses (every entry of Lambda touches D tensor components and vice versa), so the
// axis / transposition conventions live in ONE place (qsplit_mlrecon_methods._pauli_choi_basis).
#include "qdc_common.cuh"

namespace {

constexpr int ML_WARPS = 4;                  // blocks (warps) per CTA
constexpr int MAX_SWEEPS = 16;
constexpr int MAX_NEWTON = 12;
constexpr int THR_THREADS = 256, THR_MAX_CTAS = 1024;

struct MlState {                             // head of the workspace (256 bytes reserved)
    int need_full;                           // some block is not PSD (or the caller forces the full path)
    int done;                                // Newton converged
    unsigned int counter;                    // CTAs finished in the current threshold launch
    int newton_steps;
    double t, total, abs_total;
    int conj_not_diag;                       // set when Lambda -> conj(Lambda) is NOT a sign flip of tensor components
};

struct MlTables {
    const int* fwd_k;        // [(i D + j) D + t] tensor component
    const double2* fwd_c;    //                   coefficient (2^-q folded in): Lambda[i][j] = sum_t T[k_t] c_t
    const int* inv_ij;       // [kappa D + t]     (i << 8) | j
    const double2* inv_c;    //                   T'[kappa] = Re sum_t c_t L'[j_t][i_t]
};

__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ double2 cmulc(double2 a, double2 b) { return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }   // a conj(b)

// round-robin tournament: D/2 disjoint pairs for step in [0, D-1); every pair once per sweep
template <int D>
__device__ __forceinline__ void pair_of(int step, int k, int& p, int& q) {
    constexpr int M = D > 1 ? D - 1 : 1;                       // (D = 1 has no pairs; keeps the modulus legal)
    const int a = (k == 0) ? D - 1 : (step + k - 1) % M;
    const int b = (step + (D - 1 - k) - 1) % M;                // partner = slot D-1-k of the tournament table
    p = a < b ? a : b;
    q = a < b ? b : a;
}

// Lambda (D x D) of this warp's block from its tensor row
template <int D>
__device__ __forceinline__ void build_lambda(const double* __restrict__ Trow, const MlTables& tab, double* sT, double2* sA, int lane) {
    for (int e = lane; e < D * D; e += 32) sT[e] = Trow[e];
    __syncwarp();
    for (int e = lane; e < D * D; e += 32) {
        double2 acc = make_double2(0.0, 0.0);
#pragma unroll 4
        for (int t = 0; t < D; ++t) {
            const double x = sT[__ldg(tab.fwd_k + e * D + t)];
            const double2 c = __ldg(tab.fwd_c + e * D + t);
            acc.x = fma(x, c.x, acc.x);
            acc.y = fma(x, c.y, acc.y);
        }
        sA[e] = acc;
    }
    __syncwarp();
}

// ---- the change-of-basis tables in shared memory -------------------------------------------------
// The host tables are generic (index, complex coefficient) lists, 160 KB at D = 16: far more than the
// L1 that the per-warp blocks leave, so every lookup of every block went to L2 (412 GB of L2 -> SM
// traffic for the 2^19 blocks of a C4 fragment: 45 ms for what is arithmetically a 3 ms job).  For a
// Pauli basis every coefficient is s * {1, -1, i, -i, 0} with one magnitude s, so an entry packs
// into 4 bytes (index, re sign, im sign): 16 KB per table, built once per CTA (components within 1e-12
// of +-s count as +-s: a relative change of the result below 1e-12).  Tables that do not have that
// form keep the global-memory path.
struct PackedTab { const uchar4* e; double s; bool ok; };

template <int D>
__device__ PackedTab pack_table(const int* __restrict__ idx, const double2* __restrict__ coef, bool inverse,
                                uchar4* dst, unsigned long long* s_max, int* s_bad) {
    constexpr int N = D * D * D;
    __syncthreads();                                   // (a previous call's result has been read by everyone)
    if (threadIdx.x == 0) { *s_max = 0ull; *s_bad = 0; }
    __syncthreads();
    double m = 0.0;
    for (int e = threadIdx.x; e < N; e += blockDim.x) {
        const double2 c = __ldg(coef + e);
        m = fmax(m, fmax(fabs(c.x), fabs(c.y)));
    }
    atomicMax(s_max, (unsigned long long)__double_as_longlong(m));     // non-negative doubles order like integers
    __syncthreads();
    const double sc = __longlong_as_double((long long)*s_max);
    bool bad = false;
    for (int e = threadIdx.x; e < N; e += blockDim.x) {
        const double2 c = __ldg(coef + e);
        const int k = __ldg(idx + e);
        // (the host basis is built from normalised factors: components are +-s up to a few ulp)
        const double tol = 1e-12 * sc;
        const int re = fabs(fabs(c.x) - sc) <= tol ? (c.x > 0.0 ? 1 : -1) : 0, im = fabs(fabs(c.y) - sc) <= tol ? (c.y > 0.0 ? 1 : -1) : 0;
        if ((re == 0 && fabs(c.x) > tol) || (im == 0 && fabs(c.y) > tol)) bad = true;
        int code = k;
        if (inverse) { code = ((k >> 8) << 4) | (k & 0xff); if ((k >> 8) >= 16 || (k & 0xff) >= 16) bad = true; }
        else if (k < 0 || k > 255) bad = true;
        // stored [t][entry]: the 32 lanes of a warp work on 32 consecutive entries at the same t, so the table
        // reads are conflict-free ([entry][t] put them 16 words apart: a 16-way bank conflict on every read)
        dst[(e % D) * (D * D) + e / D] = make_uchar4((unsigned char)code, (unsigned char)(signed char)re, (unsigned char)(signed char)im, 0);
    }
    if (bad) atomicOr(s_bad, 1);
    __syncthreads();
    PackedTab t = {dst, sc, *s_bad == 0};
    return t;
}

__device__ __forceinline__ double signed_pick(int sgn, double x) { return sgn > 0 ? x : (sgn < 0 ? -x : 0.0); }

// build_lambda from the packed table
template <int D>
__device__ __forceinline__ void build_lambda_packed(const double* __restrict__ Trow, const PackedTab& tab, double* sT, double2* sA, int lane) {
    for (int e = lane; e < D * D; e += 32) sT[e] = Trow[e];
    __syncwarp();
    for (int e = lane; e < D * D; e += 32) {
        double ax = 0.0, ay = 0.0;
#pragma unroll 4
        for (int t = 0; t < D; ++t) {
            const uchar4 q = tab.e[t * D * D + e];
            const double x = sT[q.x];
            ax += signed_pick((signed char)q.y, x);
            ay += signed_pick((signed char)q.z, x);
        }
        sA[e] = make_double2(ax * tab.s, ay * tab.s);
    }
    __syncwarp();
}

template <int D>
__global__ void __launch_bounds__(ML_WARPS * 32) ml_psd_kernel(const double* __restrict__ T, long long n_blocks, long long ld,
                                                               MlTables tab, MlState* st) {
    __shared__ double sT[ML_WARPS][D * D];
    __shared__ double2 sA[ML_WARPS][D * D];
    __shared__ uchar4 sFwd[D * D * D];
    __shared__ unsigned long long s_max;
    __shared__ int s_bad;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PackedTab fwd = pack_table<D>(tab.fwd_k, tab.fwd_c, false, sFwd, &s_max, &s_bad);
    bool bad = false;
    for (long long b = (long long)blockIdx.x * ML_WARPS + warp; b < n_blocks; b += (long long)gridDim.x * ML_WARPS) {
        double2* A = sA[warp];
        if (fwd.ok) build_lambda_packed<D>(T + b * ld, fwd, sT[warp], A, lane);
        else build_lambda<D>(T + b * ld, tab, sT[warp], A, lane);
        double tr = 0.0;
        for (int i = 0; i < D; ++i) tr += A[i * D + i].x;
        const double eps = 1e-14 * fabs(tr) + 1e-300;
        if (tr < -eps) { bad = true; continue; }
        __syncwarp();
        if (lane < D) A[lane * D + lane].x += eps;
        __syncwarp();
        // right-looking Cholesky of the shifted block: fails on the first non-positive pivot
        for (int k = 0; k < D; ++k) {
            const double pivot = A[k * D + k].x;
            if (!(pivot > 0.0)) { bad = true; break; }
            const double inv = 1.0 / pivot;
            for (int e = lane; e < D * D; e += 32) {
                const int i = e / D, j = e % D;
                if (i > k && j > k) {
                    // A[i][k] conj(A[j][k]) read from ROW k (the trailing block stays Hermitian: both triangles are
                    // updated): A[k][j] conj(A[k][i]).  Column reads A[j * D + k] put the 16 values of j 256 bytes
                    // apart -- a 16-way bank conflict that made this kernel 16.7 ms for 2^19 blocks.
                    const double2 u = cmulc(A[k * D + j], A[k * D + i]);
                    A[e].x -= u.x * inv;
                    A[e].y -= u.y * inv;
                }
            }
            __syncwarp();
        }
        __syncwarp();
    }
    if (bad && lane == 0) atomicOr(&st->need_full, 1);
}

template <int D>
__global__ void __launch_bounds__(ML_WARPS * 32) ml_jacobi_kernel(const double* __restrict__ T, long long n_blocks, long long ld,
                                                                  MlTables tab, const MlState* st,
                                                                  double* __restrict__ evals, double2* __restrict__ vecs) {
    __shared__ double sT[ML_WARPS][D * D];
    __shared__ double2 sA[ML_WARPS][D * D];
    __shared__ double2 sV[ML_WARPS][D * D];
    __shared__ double sRot[ML_WARPS][(D / 2 > 0 ? D / 2 : 1) * 3];          // c, s phi (re, im)
    if (!st->need_full) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (long long b = (long long)blockIdx.x * ML_WARPS + warp; b < n_blocks; b += (long long)gridDim.x * ML_WARPS) {
        double2* A = sA[warp];
        double2* V = sV[warp];
        double* rot = sRot[warp];
        build_lambda<D>(T + b * ld, tab, sT[warp], A, lane);
        for (int e = lane; e < D * D; e += 32) V[e] = make_double2(e / D == e % D ? 1.0 : 0.0, 0.0);
        __syncwarp();
        for (int sweep = 0; sweep < MAX_SWEEPS && D > 1; ++sweep) {
            double off = 0.0, dg = 0.0;
            for (int e = lane; e < D * D; e += 32) {
                const double m2 = A[e].x * A[e].x + A[e].y * A[e].y;
                if (e / D == e % D) dg += m2; else off += m2;
            }
            off = warp_sum(off); dg = warp_sum(dg);
            if (off <= 1e-30 * dg || off == 0.0) break;
            for (int step = 0; step < D - 1; ++step) {
                if (lane < D / 2) {
                    int p, q;
                    pair_of<D>(step, lane, p, q);
                    const double a = A[p * D + p].x, d = A[q * D + q].x;
                    const double2 bq = A[p * D + q];
                    const double ab = sqrt(bq.x * bq.x + bq.y * bq.y);
                    double c = 1.0, sx = 0.0, sy = 0.0;
                    if (ab > 0.0) {
                        const double tau = (d - a) / (2.0 * ab);
                        const double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
                        c = 1.0 / sqrt(1.0 + t * t);
                        const double s = t * c;
                        sx = s * bq.x / ab; sy = s * bq.y / ab;            // s phi, phi = b / |b|
                    }
                    rot[lane * 3] = c; rot[lane * 3 + 1] = sx; rot[lane * 3 + 2] = sy;
                }
                __syncwarp();
                // rows:  A <- J^+ A   ((J^+ A)[p] = c A[p] - s phi A[q],  (J^+ A)[q] = s conj(phi) A[p] + c A[q])
                for (int it = lane; it < (D / 2) * D; it += 32) {
                    const int k = it / D, j = it % D;
                    int p, q;
                    pair_of<D>(step, k, p, q);
                    const double c = rot[k * 3];
                    const double2 sp = make_double2(rot[k * 3 + 1], rot[k * 3 + 2]);
                    const double2 x = A[p * D + j], y = A[q * D + j];
                    const double2 u = cmul(sp, y), w = cmulc(x, sp);       // s phi y ;  x conj(s phi) = s conj(phi) x
                    A[p * D + j] = make_double2(c * x.x - u.x, c * x.y - u.y);
                    A[q * D + j] = make_double2(w.x + c * y.x, w.y + c * y.y);
                }
                __syncwarp();
                // columns:  A <- A J,  V <- V J   ((A J)[.][p] = c A[.][p] - s conj(phi) A[.][q],  (A J)[.][q] = s phi A[.][p] + c A[.][q])
                for (int it = lane; it < (D / 2) * D * 2; it += 32) {
                    const int which = it / ((D / 2) * D), r = it % ((D / 2) * D);
                    const int k = r / D, i = r % D;
                    int p, q;
                    pair_of<D>(step, k, p, q);
                    double2* M = which ? V : A;
                    const double c = rot[k * 3];
                    const double2 sp = make_double2(rot[k * 3 + 1], rot[k * 3 + 2]);
                    const double2 x = M[i * D + p], y = M[i * D + q];
                    const double2 u = cmulc(y, sp), w = cmul(x, sp);       // y conj(s phi) ;  x s phi
                    M[i * D + p] = make_double2(c * x.x - u.x, c * x.y - u.y);
                    M[i * D + q] = make_double2(w.x + c * y.x, w.y + c * y.y);
                }
                __syncwarp();
                if (lane < D / 2) {                                       // what the rotation annihilates, exactly
                    int p, q;
                    pair_of<D>(step, lane, p, q);
                    A[p * D + q] = make_double2(0.0, 0.0);
                    A[q * D + p] = make_double2(0.0, 0.0);
                    A[p * D + p].y = 0.0; A[q * D + q].y = 0.0;
                }
                __syncwarp();
            }
        }
        for (int k = lane; k < D; k += 32) evals[b * D + k] = A[k * D + k].x;
        for (int e = lane; e < D * D; e += 32) vecs[b * D * D + e] = V[e];
        __syncwarp();
    }
}

// one Newton step on h(t) = sum max(lambda - t, 0) - sum lambda; the last CTA to finish folds the
// per-CTA partials in CTA order (deterministic) and updates the state
__global__ void __launch_bounds__(THR_THREADS) ml_threshold_kernel(const double* __restrict__ evals, long long n, MlState* st,
                                                                   double* __restrict__ partials /* [grid][4] */) {
    __shared__ double red[4][THR_THREADS / 32];
    __shared__ bool last;
    if (!st->need_full || st->done) return;
    const double t = st->t;
    double s_pos = 0.0, cnt = 0.0, s_all = 0.0, s_abs = 0.0;
    for (long long i = (long long)blockIdx.x * THR_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * THR_THREADS) {
        const double v = evals[i];
        if (v > t) { s_pos += v - t; cnt += 1.0; }
        s_all += v; s_abs += fabs(v);
    }
    s_pos = warp_sum(s_pos); cnt = warp_sum(cnt); s_all = warp_sum(s_all); s_abs = warp_sum(s_abs);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { red[0][warp] = s_pos; red[1][warp] = cnt; red[2][warp] = s_all; red[3][warp] = s_abs; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0, c = 0, d = 0;
        for (int w = 0; w < THR_THREADS / 32; ++w) { a += red[0][w]; b += red[1][w]; c += red[2][w]; d += red[3][w]; }
        partials[blockIdx.x * 4 + 0] = a; partials[blockIdx.x * 4 + 1] = b;
        partials[blockIdx.x * 4 + 2] = c; partials[blockIdx.x * 4 + 3] = d;
        __threadfence();
        last = atomicAdd(&st->counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        double a = 0, b = 0, c = 0, d = 0;
        for (unsigned g = 0; g < gridDim.x; ++g) {
            a += partials[g * 4 + 0]; b += partials[g * 4 + 1]; c += partials[g * 4 + 2]; d += partials[g * 4 + 3];
        }
        st->total = c; st->abs_total = d;
        const double h = a - c;
        if (!(h > 1e-16 * d) || b <= 0.0) st->done = 1;
        else { st->t = t + h / b; st->newton_steps += 1; }
        st->counter = 0u;
        __threadfence();
    }
}

template <int D>
__global__ void __launch_bounds__(ML_WARPS * 32) ml_rebuild_kernel(const double* __restrict__ T, long long n_blocks, long long ld,
                                                                   MlTables tab, const MlState* st,
                                                                   const double* __restrict__ evals, const double2* __restrict__ vecs,
                                                                   const double* __restrict__ sgn, double* __restrict__ T_out, long long ld_out) {
    // dynamic shared memory (rebuild_smem_bytes<D>()): [inverse table][L per warp][V per warp][T per warp][w per warp];
    // on the exact-data path V is not needed and its space holds the packed forward table
    extern __shared__ __align__(16) unsigned char ml_dyn[];
    constexpr int TABB = (D * D * D * 4 + 15) / 16 * 16;
    uchar4* sInv = reinterpret_cast<uchar4*>(ml_dyn);
    double2* sA_all = reinterpret_cast<double2*>(ml_dyn + TABB);
    double2* sV_all = sA_all + ML_WARPS * D * D;
    double* sT_all = reinterpret_cast<double*>(sV_all + ML_WARPS * D * D);
    double* sW_all = sT_all + ML_WARPS * D * D;
    __shared__ unsigned long long s_max;
    __shared__ int s_bad;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool full = st->need_full != 0;
    const double t0 = st->t;
    if (!full && !st->conj_not_diag) {
        // exact data, Pauli basis: the model is the conjugate, i.e. a sign per tensor component (ml_conj_map_kernel)
        const long long total = n_blocks * (D * D);
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
            const long long b = i / (D * D);
            const int e = (int)(i - b * (D * D));
            T_out[b * ld_out + e] = sgn[e] * T[b * ld + e];
        }
        return;
    }
    const PackedTab inv = pack_table<D>(tab.inv_ij, tab.inv_c, true, sInv, &s_max, &s_bad);
    PackedTab fwd = {nullptr, 0.0, false};
    if (!full) fwd = pack_table<D>(tab.fwd_k, tab.fwd_c, false, reinterpret_cast<uchar4*>(sV_all), &s_max, &s_bad);   // (uniform branch)
    double* sTw = sT_all + warp * D * D;
    double* sWw = sW_all + warp * D;
    for (long long b = (long long)blockIdx.x * ML_WARPS + warp; b < n_blocks; b += (long long)gridDim.x * ML_WARPS) {
        double2* L = sA_all + warp * D * D;
        if (!full) {
            if (fwd.ok) build_lambda_packed<D>(T + b * ld, fwd, sTw, L, lane);
            else build_lambda<D>(T + b * ld, tab, sTw, L, lane);
            for (int e = lane; e < D * D; e += 32) L[e].y = -L[e].y;           // to_projector conjugates (:45-46)
        } else {
            double2* V = sV_all + warp * D * D;
            for (int e = lane; e < D * D; e += 32) V[e] = vecs[b * D * D + e];
            for (int k = lane; k < D; k += 32) sWw[k] = fmax(evals[b * D + k] - t0, 0.0);
            __syncwarp();
            for (int e = lane; e < D * D; e += 32) {
                const int i = e / D, j = e % D;
                double2 acc = make_double2(0.0, 0.0);
                for (int k = 0; k < D; ++k) {
                    const double2 u = cmulc(V[i * D + k], V[j * D + k]);
                    acc.x = fma(sWw[k], u.x, acc.x);
                    acc.y = fma(sWw[k], u.y, acc.y);
                }
                L[e] = make_double2(acc.x, -acc.y);                         // conj(V diag V^+)
            }
        }
        __syncwarp();
        for (int kap = lane; kap < D * D; kap += 32) {
            double acc = 0.0;
            if (inv.ok) {
#pragma unroll 4
                for (int t = 0; t < D; ++t) {
                    const uchar4 q = inv.e[t * D * D + kap];                    // (i << 4 | j, re sign, im sign)
                    const double2 l = L[(q.x & 15) * D + (q.x >> 4)];           // L'[j][i]
                    acc += signed_pick((signed char)q.y, l.x) - signed_pick((signed char)q.z, l.y);
                }
                acc *= inv.s;
            } else {
#pragma unroll 4
                for (int t = 0; t < D; ++t) {
                    const int ij = __ldg(tab.inv_ij + kap * D + t);
                    const double2 c = __ldg(tab.inv_c + kap * D + t);
                    const double2 l = L[(ij & 0xff) * D + (ij >> 8)];           // L'[j][i]
                    acc += c.x * l.x - c.y * l.y;
                }
            }
            T_out[b * ld_out + kap] = acc;
        }
        __syncwarp();
    }
}

__global__ void ml_init_kernel(MlState* st, int force_full) {
    st->need_full = force_full > 0 ? 1 : 0; st->done = 0; st->counter = 0u; st->newton_steps = 0;
    st->t = 0.0; st->total = 0.0; st->abs_total = 0.0; st->conj_not_diag = 0;
}

// What the exact-data path does to a tensor row is the fixed linear map  T -> inv(conj(fwd(T))).  For a
// basis of Hermitian Pauli products conj(P) = +-P (minus for an odd number of Y), so the map is a sign
// per component and the "rebuild" of 2^19 blocks is one streaming pass.  This kernel applies the map to
// the unit vector of component kappa = blockIdx.x through the caller's tables, stores the diagonal
// entry in sgn[kappa] and raises conj_not_diag if anything else is non-zero or the entry is not +-1
// (then the table path runs).  D^2 CTAs of D^2 threads, microseconds.
template <int D>
__global__ void __launch_bounds__(D * D) ml_conj_map_kernel(MlTables tab, MlState* st, double* __restrict__ sgn) {
    __shared__ double2 L[D * D];
    const int kappa = blockIdx.x, e = threadIdx.x;
    double2 acc = make_double2(0.0, 0.0);
    for (int t = 0; t < D; ++t)
        if (tab.fwd_k[e * D + t] == kappa) { const double2 c = tab.fwd_c[e * D + t]; acc.x += c.x; acc.y += c.y; }
    L[e] = make_double2(acc.x, -acc.y);                                   // conj(Lambda(e_kappa))
    __syncthreads();
    double out = 0.0;
    for (int t = 0; t < D; ++t) {
        const int ij = tab.inv_ij[e * D + t];
        const double2 c = tab.inv_c[e * D + t];
        const double2 l = L[(ij & 0xff) * D + (ij >> 8)];
        out += c.x * l.x - c.y * l.y;
    }
    if (e == kappa) {
        sgn[kappa] = out;
        if (fabs(fabs(out) - 1.0) > 1e-12) atomicOr(&st->conj_not_diag, 1);
    } else if (fabs(out) > 1e-12) atomicOr(&st->conj_not_diag, 1);
}

template <int D>
int ml_run(const double* T, long long n, long long ld, MlTables tab, double* T_out, long long ld_out, uint8_t* ws,
           int force_full, cudaStream_t s) {
    MlState* st = (MlState*)ws;
    double* partials = (double*)(ws + 256);
    double* evals = partials + 4 * THR_MAX_CTAS;
    double2* vecs = (double2*)(evals + ((n * D + 1) / 2) * 2);
    const long long want = (n + ML_WARPS - 1) / ML_WARPS;
    const int grid = (int)(want < 148ll * 16 ? want : 148ll * 16);
    ml_init_kernel<<<1, 1, 0, s>>>(st, force_full);
    QDC_CHECK_LAUNCH();
    if (force_full <= 0) {
        // (sgn lives in the Newton partials: they are only touched when the full path runs, the signs only when it does not)
        ml_conj_map_kernel<D><<<D * D, D * D, 0, s>>>(tab, st, partials);
        QDC_CHECK_LAUNCH();
        if (force_full == 0) {                             // (< 0: the caller vouches for positive semidefinite blocks)
            ml_psd_kernel<D><<<grid, ML_WARPS * 32, 0, s>>>(T, n, ld, tab, st);
            QDC_CHECK_LAUNCH();
        }
    }
    ml_jacobi_kernel<D><<<grid, ML_WARPS * 32, 0, s>>>(T, n, ld, tab, st, evals, vecs);
    QDC_CHECK_LAUNCH();
    const long long nv = n * D, tw = (nv + THR_THREADS * 8 - 1) / (THR_THREADS * 8);
    const int tgrid = (int)(tw < 1 ? 1 : (tw < THR_MAX_CTAS ? tw : THR_MAX_CTAS));
    for (int it = 0; it < MAX_NEWTON; ++it) {
        ml_threshold_kernel<<<tgrid, THR_THREADS, 0, s>>>(evals, nv, st, partials);
        QDC_CHECK_LAUNCH();
    }
    constexpr size_t rebuild_smem = (size_t)((D * D * D * 4 + 15) / 16 * 16) + (size_t)ML_WARPS * D * D * (16 + 16 + 8) + (size_t)ML_WARPS * D * 8;
    QDC_CHECK_CUDA(cudaFuncSetAttribute(ml_rebuild_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rebuild_smem));
    ml_rebuild_kernel<D><<<grid, ML_WARPS * 32, rebuild_smem, s>>>(T, n, ld, tab, st, evals, vecs, partials, T_out, ld_out);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

}  // namespace

extern "C" int64_t qdc_ml_workspace_bytes(int64_t n_blocks, int dim) {
    if (n_blocks < 0 || dim < 1 || dim > 16) return 0;
    const int64_t ev = ((n_blocks * dim + 1) / 2) * 2;
    return 256 + 4 * THR_MAX_CTAS * 8 + ev * 8 + n_blocks * dim * dim * 16;
}

extern "C" int qdc_ml_project(const double* d_T, int64_t n_blocks, int dim, int64_t ld,
                              const int32_t* d_fwd_k, const double* d_fwd_c, const int32_t* d_inv_ij,
                              const double* d_inv_c, double* d_T_out, int64_t ld_out, void* d_workspace,
                              int force_full, void* stream) {
    QDC_REQUIRE(d_T && d_T_out && d_workspace && d_fwd_k && d_fwd_c && d_inv_ij && d_inv_c, "null buffer");
    QDC_REQUIRE(dim == 1 || dim == 2 || dim == 4 || dim == 8 || dim == 16, "block dimension 2^q, q = 0..4 (at most 4 cut ends per fragment)");
    QDC_REQUIRE(n_blocks >= 0 && ld >= (int64_t)dim * dim && ld_out >= (int64_t)dim * dim, "shape");
    QDC_REQUIRE((((uintptr_t)d_workspace) & 15) == 0 && (((uintptr_t)d_fwd_c) & 15) == 0 && (((uintptr_t)d_inv_c) & 15) == 0,
                "16-byte alignment");
    if (n_blocks == 0) return QDC_OK;
    const MlTables tab = {d_fwd_k, (const double2*)d_fwd_c, d_inv_ij, (const double2*)d_inv_c};
    cudaStream_t s = (cudaStream_t)stream;
    uint8_t* ws = (uint8_t*)d_workspace;
    switch (dim) {
        case 1: return ml_run<1>(d_T, n_blocks, ld, tab, d_T_out, ld_out, ws, force_full, s);
        case 2: return ml_run<2>(d_T, n_blocks, ld, tab, d_T_out, ld_out, ws, force_full, s);
        case 4: return ml_run<4>(d_T, n_blocks, ld, tab, d_T_out, ld_out, ws, force_full, s);
        case 8: return ml_run<8>(d_T, n_blocks, ld, tab, d_T_out, ld_out, ws, force_full, s);
        default: return ml_run<16>(d_T, n_blocks, ld, tab, d_T_out, ld_out, ws, force_full, s);
    }
}
