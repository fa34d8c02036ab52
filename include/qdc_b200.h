/*
 * qdc_b200.h -- C-ABI of libqdc_b200.so: the B200 (sm_100a) implementation of the Quantum
 * Divide-and-Conquer hot path of teaguetomesh/dqva-and-circuit-cutting.
 *
 * The reference has no FFI of its own: its seams are Python call signatures and,
 * for cutqc, four C executables.  Each entry point below names the reference code it replaces
 * (file:line relative to the reference tree).  Conventions:
 *
 *   - every pointer named d_* is DEVICE memory owned by the caller (PyTorch tensors in the
 *     Python host layer); h_* is host memory; the library allocates nothing persistent,
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is
 *     enqueued asynchronously on it unless the function documents a host result,
 *   - return value 0 = success, negative = error; qdc_last_error() gives a thread-local message;
 *     nothing throws, nothing calls exit(),
 *   - bit conventions are qiskit's: integer index x has bit i <-> wire i.
 */
#ifndef QDC_B200_H
#define QDC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QDC_OK              0
#define QDC_ERR_ARG        -1
#define QDC_ERR_CUDA       -2
#define QDC_ERR_OVERFLOW   -3   /* candidate buffer too small */
#define QDC_ERR_UNSUPPORTED -4

/* ---- library ---------------------------------------------------------------------------- */
int         qdc_version(void);
const char* qdc_last_error(void);
/* sm_count / compute capability of the current device */
int         qdc_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- K8: MIS scoring tables ---------------------------------------------------------------
 * Replaces the per-bitstring Python of hamming_weight (utils/helper_funcs.py:18-19) and
 * is_indset (utils/graph_funcs.py:107-118) as applied by avg_cost (mis.py:125-126).
 *
 * The recombined distribution is indexed by a "united" integer u = row * 2^col_bits + col
 * (row = fragments F-1..1, col = fragment 0; qsplit_mlrecon_methods.py:423).  bitpos[b] is the
 * full-circuit wire carried by united bit b (the inverse of united_axis_permutation,
 * qsplit_mlrecon_methods.py:532-536).  adj[i] = neighbour mask of node i, node_mask = wires that
 * are graph nodes (ancilla wires excluded: strip_ancillas, utils/helper_funcs.py:3-16).
 *
 * Outputs (device): d_row_x[r], d_col_x[c] = circuit-order bits contributed by that row/col;
 * d_row_forb[r] = OR of adj over the nodes set in row r; d_row_wok[r], d_col_wok[c] =
 * weight | (self-independent << 8).  Then for u=(r,c):
 *   x = row_x|col_x, weight = w_r + w_c, indset = ok_r & ok_c & !(col_x & row_forb).
 */
int qdc_mis_build_tables(int row_bits, int col_bits, const int* h_bitpos,
                         const uint64_t* h_adj, int n_wires, uint64_t node_mask,
                         uint64_t* d_row_x, uint64_t* d_row_forb, uint32_t* d_row_wok,
                         uint64_t* d_col_x, uint32_t* d_col_wok, void* stream);

/* ---- K8: fused MIS reduction over a materialised distribution -----------------------------
 * Replaces strip_ancillas + avg_cost + the p > threshold filter of top_probs
 * (mis.py:114, 125-126, 232).  d_P is rows x cols (row-major, united order), dtype 0 = f64,
 * 1 = f32; rows [row_lo, row_lo+n_rows) of the table are used (slab owned by this GPU).
 * Every entry is first divided by *d_norm (device scalar, NULL = 1; the combined_norm of
 * qsplit_mlrecon_methods.py:471-472) so no host synchronisation is needed between the two.
 * d_sums[0..2] += (sum p, sum p*weight*indset, sum p*weight); entries with p > threshold are
 * appended to (d_cand_x, d_cand_p) through the counter d_cand_count (capacity cand_cap).
 * d_partials: workspace of at least 3 * qdc_reduce_partials() doubles.
 */
int qdc_reduce_partials(void);
int qdc_mis_reduce(const void* d_P, int dtype, int64_t row_lo, int64_t n_rows, int64_t n_cols,
                   const double* d_norm, double threshold,
                   const uint64_t* d_row_x, const uint64_t* d_row_forb, const uint32_t* d_row_wok,
                   const uint64_t* d_col_x, const uint32_t* d_col_wok,
                   double* d_sums, uint64_t* d_cand_x, double* d_cand_p,
                   unsigned long long* d_cand_count, int64_t cand_cap,
                   double* d_partials, void* stream);

/* Top-k of the first min(*d_cand_count, cand_cap) candidates by (p descending, x ascending) --
 * sorted(...)[:cutoff] of mis.py:232-233 with a deterministic tie-break.  k <= 1024; unused
 * output slots get x = ~0, p = -inf.  Destroys d_cand_p.  The count is read on the device, so
 * no host synchronisation separates scoring from selection. */
int qdc_topk_select(const uint64_t* d_cand_x, double* d_cand_p,
                    const unsigned long long* d_cand_count, int64_t cand_cap, int k,
                    uint64_t* d_top_x, double* d_top_p, void* stream);

/* ---- K5: two-operand contraction over the cut-basis index ---------------------------------
 * Replaces the itertools.product loop of _recombine_using_networks
 * (qsplit_mlrecon_methods.py:412-461) and cutqc build.c:91-169 for two operands:
 *   P~[r, c] = scale * sum_k A[r, k] * Bt[c, k],  r < n_rows (rows row_lo.. of A), c < n_cols.
 * d_A: [*, kdim] row-major (row operand, fragments F-1..1), d_Bt: [n_cols, kdim] row-major
 * (column operand, fragment 0), kdim = 4^K, kdim % 16 == 0 required (pad with zeros).
 * FP64 via DMMA (mma.sync m8n8k4).  The MIS epilogue of qdc_mis_reduce is applied to every
 * tile in registers; P is only written when d_P_out != NULL:
 *   store_mode 1: d_P_out[(r - row_lo) * n_cols + c]         (united order, slab-local)
 *   store_mode 2: d_P_out[row_x[r] | col_x[c]]               (full-circuit order, scatter)
 * out_dtype 0 = f64, 1 = f32.
 */
int qdc_contract_f64(const double* d_A, const double* d_Bt, int64_t row_lo, int64_t n_rows,
                     int64_t n_cols, int kdim, double scale, const double* d_norm, double threshold,
                     const uint64_t* d_row_x, const uint64_t* d_row_forb,
                     const uint32_t* d_row_wok, const uint64_t* d_col_x,
                     const uint32_t* d_col_wok,
                     void* d_P_out, int store_mode, int out_dtype,
                     double* d_sums, uint64_t* d_cand_x, double* d_cand_p,
                     unsigned long long* d_cand_count, int64_t cand_cap,
                     double* d_partials, void* stream);

/* sum over all (r, c) of P~ without forming it: scale * sum_k (sum_r A[r,k]) (sum_c Bt[c,k]).
 * This is combined_norm of qsplit_mlrecon_methods.py:471.  d_colsums: workspace of
 * qdc_norm_workspace_doubles(kdim) doubles (fixed-order partial sums: the result is
 * run-to-run deterministic); d_norm: 1 double. */
int qdc_norm_workspace_doubles(int kdim);
int qdc_contract_norm(const double* d_A, int64_t n_rows_total, const double* d_Bt,
                      int64_t n_cols, int kdim, double scale, double* d_colsums,
                      double* d_norm, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* QDC_B200_H */
