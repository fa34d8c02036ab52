/* ORACLE (test infrastructure, not product code): plain-C restatement of the reference's
 * recombination + MIS scoring, used (i) as a second, independent checker for the CUDA path at
 * sizes numpy handles slowly and (ii) as the timed CPU baseline ("port") in bench.py.
 *
 * What it restates, per output bitstring:
 *   value   = Re sum_kappa prod_f T_f[b_f; kappa]   qsplit/qsplit_mlrecon_methods.py:412-461
 *             (real Pauli-coordinate form, two operands; SURVEY.md A.3 "GEMM form")
 *   key     = united bits permuted to circuit order  qsplit_mlrecon_methods.py:423-424, 532-536
 *   weight  = number of '1' among node wires          utils/helper_funcs.py:3-19
 *   indset  = no two chosen nodes adjacent            utils/graph_funcs.py:107-118
 *   cost    = sum prob * weight over independent sets mis.py:125-126
 *   top     = p > threshold, best by p                mis.py:232-233
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may load this.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline int is_indset_bits(uint64_t x, const uint64_t *adj) {
    /* for idx in set bits ascending: reject if an already chosen node is a neighbour
       (graph_funcs.py:109-117)  <=>  x & adj[idx] != 0 for some set idx */
    uint64_t rem = x;
    while (rem) {
        int idx = __builtin_ctzll(rem);
        rem &= rem - 1;
        if (x & adj[idx]) return 0;
    }
    return 1;
}

static inline uint64_t permute_bits(uint64_t u, const int *bitpos, int nbits) {
    uint64_t x = 0;
    for (int b = 0; b < nbits; ++b)
        if ((u >> b) & 1) x |= 1ull << bitpos[b];
    return x;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* sums[0..2] = sum p, sum p*w*indset, sum p*w;  best_x/best_p = arg-max entry above threshold
 * (ties: smaller x).  P_out (optional, circuit order, length 2^nbits) receives every value. */
void oracle_contract_score(const double *A, const double *Bt, int64_t row_lo, int64_t n_rows,
                           int64_t n_cols, int kdim, double scale, double inv_norm,
                           const int *bitpos, int nbits, int col_bits, const uint64_t *adj,
                           uint64_t node_mask, double threshold, double *sums,
                           uint64_t *best_x, double *best_p, double *P_out) {
    double s0 = 0, s1 = 0, s2 = 0;
    double gbp = -INFINITY;
    uint64_t gbx = ~0ull;
#pragma omp parallel
    {
        double bp = -INFINITY;
        uint64_t bx = ~0ull;
#pragma omp for reduction(+ : s0, s1, s2) schedule(static)
        for (int64_t r = row_lo; r < row_lo + n_rows; ++r) {
            const double *a = A + r * (int64_t)kdim;
            for (int64_t c = 0; c < n_cols; ++c) {
                const double *b = Bt + c * (int64_t)kdim;
                double acc = 0;
                for (int k = 0; k < kdim; ++k) acc += a[k] * b[k];
                const double p = acc * scale * inv_norm;
                const uint64_t u = ((uint64_t)r << col_bits) | (uint64_t)c;
                const uint64_t x = permute_bits(u, bitpos, nbits);
                const uint64_t nodes = x & node_mask;
                const int w = __builtin_popcountll(nodes);
                s0 += p;
                s2 += p * w;
                if (is_indset_bits(nodes, adj)) s1 += p * w;
                if (p > threshold && (p > bp || (p == bp && x < bx))) { bp = p; bx = x; }
                if (P_out) P_out[x] = p;
            }
        }
#pragma omp critical
        if (bp > gbp || (bp == gbp && bx < gbx)) { gbp = bp; gbx = bx; }
    }
    sums[0] = s0; sums[1] = s1; sums[2] = s2;
    *best_x = gbx; *best_p = gbp;
}

/* score an already formed distribution slab P[lo .. lo+count) in circuit order */
void oracle_score(const double *P, int64_t lo, int64_t count, const uint64_t *adj,
                  uint64_t node_mask, double *sums) {
    double s0 = 0, s1 = 0, s2 = 0;
#pragma omp parallel for reduction(+ : s0, s1, s2) schedule(static)
    for (int64_t i = 0; i < count; ++i) {
        const uint64_t nodes = (uint64_t)(lo + i) & node_mask;
        const double p = P[i];
        const int w = __builtin_popcountll(nodes);
        s0 += p; s2 += p * w;
        if (is_indset_bits(nodes, adj)) s1 += p * w;
    }
    sums[0] = s0; sums[1] = s1; sums[2] = s2;
}
