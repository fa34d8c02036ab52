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
 * index_flip: entry u of the scored array holds united index u ^ index_flip (the flip frame of
 * the statevector engine, qdc_sv_info); 0 for the contraction.
 */
int qdc_mis_build_tables(int row_bits, int col_bits, const int* h_bitpos,
                         const uint64_t* h_adj, int n_wires, uint64_t node_mask,
                         uint64_t index_flip, uint64_t* d_row_x, uint64_t* d_row_forb, uint32_t* d_row_wok,
                         uint64_t* d_col_x, uint32_t* d_col_wok, void* stream);

/* ---- K8: fused MIS reduction over a materialised distribution -----------------------------
 * Replaces strip_ancillas + avg_cost + the p > threshold filter of top_probs
 * (mis.py:114, 125-126, 232).  d_P is rows x cols (row-major, united order), dtype 0 = f64,
 * 1 = f32, 2 = complex128 amplitudes scored as |amp|^2 (the uncut cost f(params), mis.py:639-665); rows [row_lo, row_lo+n_rows) of the table are used (slab owned by this GPU).
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

/* Penalised objective of QAOA+ (ansatz/qaoa_plus.py:40-53: expectation_value = sum_x p(x) (|x| -
 * Lambda * #edges of the graph with both ends 1 in x)) over 2^n_bits entries, entry i holding
 * basis state i ^ index_flip, bit b of the state = wire b.  dtype as qdc_mis_reduce (2: complex128
 * amplitudes, p = |amp|^2).  h_adj[w] = neighbour mask of wire w, node_mask = wires that count.
 * d_sums (3 doubles, accumulated into: zero them first) = [sum p, sum p * #edges, sum p * |x|];
 * d_workspace: qdc_penalty_workspace_bytes() bytes; d_partials as qdc_mis_reduce.  One pass,
 * sizeof(entry) bytes read per output, deterministic summation order. */
int qdc_penalty_workspace_bytes(void);
int qdc_penalty_reduce(const void* d_P, int dtype, int n_bits, uint64_t index_flip,
                       const uint64_t* h_adj, int n_wires, uint64_t node_mask,
                       void* d_workspace, double* d_sums, double* d_partials, void* stream);

/* Top-k of the first min(*d_cand_count, cand_cap) candidates by (p descending, x ascending) --
 * sorted(...)[:cutoff] of mis.py:232-233 with a deterministic tie-break.  k <= 1024; unused
 * output slots get x = ~0, p = -inf.  Destroys d_cand_p.  The count is read on the device, so
 * no host synchronisation separates scoring from selection. */
int qdc_topk_select(const uint64_t* d_cand_x, double* d_cand_p,
                    const unsigned long long* d_cand_count, int64_t cand_cap, int k,
                    uint64_t* d_top_x, double* d_top_p, void* stream);

/* ---- K1-K3: batched statevector engine -----------------------------------------------------
 * Replaces the simulator the reference reaches through qiskit.execute(circuits, backend, shots)
 * (qsplit/qsplit_mlrecon_methods.py:20-42; mis.py:649) for the gate set its ansatz modules emit
 * (ansatz/dqv_cut_ansatz.py:127-171, ansatz/qaoa_plus.py:8-37) and for the tomography
 * preparation / basis-change gates of partial_tomography (qsplit_mlrecon_methods.py:116-162).
 *
 * Every gate of that set is "a 2x2 on one target, applied where the control qubits have given
 * values": x/h/s/sdg/rx/ry/rz/u1/u3 (no controls), cx/mcx/crx/cz (controls = 1) and the
 * open-controlled partial mixer (controls = 0).  The multi-qubit `Id` is a no-op.
 * angle = scale * params[set][param_slot] + offset (param_slot < 0: angle = offset).
 */
#define QDC_G_ID   0
#define QDC_G_X    1
#define QDC_G_Y    2
#define QDC_G_Z    3
#define QDC_G_H    4
#define QDC_G_S    5
#define QDC_G_SDG  6
#define QDC_G_T    7
#define QDC_G_TDG  8
#define QDC_G_RX   9
#define QDC_G_RY   10
#define QDC_G_RZ   11
#define QDC_G_U1   12
#define QDC_G_U2   13   /* explicit constant 2x2 in m[] (u3 with bound angles, SIC preparations) */

typedef struct qdc_gate {
    int32_t  kind;        /* QDC_G_* */
    int32_t  target;      /* qubit the 2x2 acts on */
    uint64_t ctrl_mask;   /* control qubits (bit q <-> qubit q) */
    uint64_t ctrl_val;    /* required control values, subset of ctrl_mask */
    int32_t  param_slot;  /* column of the parameter table, or -1 */
    int32_t  reserved;
    double   scale, offset;
    double   m[8];        /* QDC_G_U2: row-major (re, im) of [[m00, m01], [m10, m11]] */
} qdc_gate;

typedef struct qdc_sv_program qdc_sv_program;   /* opaque: compiled gate schedule */

/* Host-side compilation (no device work except one small upload):
 *   1. X gates are absorbed into a per-qubit flip frame (control values toggle, targets are
 *      conjugated); what is left of the frame at the end is returned as `flip_mask`:
 *      amplitude of basis state x is stored at index x ^ flip_mask.
 *   2. with QDC_SV_FUSE_ANCILLA, mcx(C -> a) . G(a -> q) . mcx(C -> a) on an ancilla `a` known
 *      to be |0> becomes G on q controlled by C (the reference's partial mixer,
 *      ansatz/dqv_cut_ansatz.py:127-151); `zero_mask` = qubits that start in |0>.
 *   3. gates are list-scheduled (respecting commutation) into groups whose non-diagonal targets
 *      fit a tile of 2^tile_bits amplitudes held in shared memory; the tile always contains the
 *      `coalesce_bits` lowest qubits so global accesses are contiguous runs.
 * tile_bits <= 0 selects the default (min(n_qubits, 12)); tile_bits is clamped to n_qubits. */
#define QDC_SV_FUSE_ANCILLA 1
int  qdc_sv_compile(int n_qubits, const qdc_gate* gates, int n_gates, uint64_t zero_mask,
                    int tile_bits, int coalesce_bits, int flags, qdc_sv_program** out);
void qdc_sv_free(qdc_sv_program* prog);
/* any pointer may be NULL */
int  qdc_sv_info(const qdc_sv_program* prog, int* n_qubits, int* n_groups, int* n_ops,
                 int* n_matrices, int* tile_bits, uint64_t* flip_mask);
/* Schedule dump for host-side tests: group_first_op[n_groups + 1], group_local_mask[n_groups],
 * and per op: op_kind (0 general, 1 diagonal), op_target, op_ctrl_mask, op_ctrl_val (full-index
 * masks), op_src_gate (index into the input list of the gate that supplies the matrix) and
 * op_conj_x (1: the matrix is X.M.X). */
int  qdc_sv_dump(const qdc_sv_program* prog, int32_t* group_first_op, uint64_t* group_local_mask,
                 int32_t* op_kind, int32_t* op_target, uint64_t* op_ctrl_mask,
                 uint64_t* op_ctrl_val, int32_t* op_src_gate, int32_t* op_conj_x);

/* Run the program on a batch of states.  Batch element b = init * n_sets + set, init in
 * [0, n_init), uses parameter row `set` and the initial product state number init_lo + init:
 * qubit h_prep_qubits[j] starts in the 2-vector h_code_table[digit_j] (re0, im0, re1, im1), digit_j
 * = the j-th most significant base-n_codes digit of (init_lo + init) -- itertools.product order of
 * the SIC / cutqc preparation labels (qsplit_mlrecon_methods.py:148; cutqc/evaluator.py:34-67);
 * all other qubits start in |0>.  d_state: [n_init * n_sets][2^n_qubits] complex128 (written).
 * d_matrices: workspace of n_sets * n_matrices * 8 doubles. */
int  qdc_sv_run(const qdc_sv_program* prog, void* d_state, int64_t init_lo, int64_t n_init,
                int64_t n_sets, const double* d_params, int n_params,
                const int32_t* h_prep_qubits, int n_prep, const double* h_code_table, int n_codes,
                double* d_matrices, void* stream);

/* The uncut cost closure f(params) of solve_mis_dqva / solve_mis_qaoa (mis.py:639-665, 485-511, 315-341:
 * simulate the statevector, then sum p(x) |x| over the outcomes) in one pass: qdc_sv_run for circuits
 * without preparation variants, with the MIS reduction FUSED INTO THE LAST GATE GROUP -- every tile is
 * scored from shared memory before (or, with keep_state = 0, instead of) being written back, so the final
 * state is neither re-read (16 B / amplitude) nor, if unwanted, stored (16 B / amplitude).
 *   d_sums[set][3] = sum p, sum p w [x independent], sum p w;  w = popcount(x & node_mask), independence
 *   with respect to d_adj (bit i of d_adj[j] <=> edge (i, j); n_qubits device words; ancilla bits outside
 *   node_mask are ignored, as strip_ancillas does).  Written, not accumulated; summation order fixed.
 * d_state: [n_sets][2^n] complex128 working buffer (between groups the state lives there; its final
 * contents are the state only when keep_state != 0).  d_workspace: qdc_sv_score_workspace_bytes(prog,
 * n_sets) bytes, 8-byte aligned.  Launches the group kernels + a 4096-entry table kernel + one
 * finishing kernel; allocates nothing. */
int64_t qdc_sv_score_workspace_bytes(const qdc_sv_program* prog, int64_t n_sets);
int  qdc_sv_run_score(const qdc_sv_program* prog, void* d_state, int64_t n_sets, const double* d_params,
                      int n_params, double* d_matrices, const uint64_t* d_adj, uint64_t node_mask,
                      int keep_state, void* d_workspace, double* d_sums, void* stream);

/* p[b][x] = |amp_b(x)|^2, undoing the flip frame: d_probs[b * 2^n + x] (float64). */
int  qdc_sv_probs(const void* d_state, int n_qubits, int64_t n_batch, uint64_t flip_mask,
                  double* d_probs, void* stream);

/* ---- finite-shot tomography: sample the variants of one measurement-basis choice ------------------
 * Replaces run_circuits on qasm_simulator (qsplit_mlrecon_methods.py:20-42, jobs of <= 8192 shots)
 * followed by the regrouping of organize_tomography_data (:169-213): `shots` outcomes are drawn from
 * |amp|^2 of every state of the batch and counted straight into the organised table
 *   d_counts[(set * 2^(n - m) + b) * ld + d_columns[init * 2^m + v]] += 1.0
 * where, for the drawn basis state x (flip frame undone), b = the bits of the qubits outside
 * meas_mask (ascending qubit order) and v = the bits of the qubits in meas_mask.  d_columns maps
 * (preparation index, measured value) to the column of the (prep label, measured state) pair in the
 * caller's label list; d_counts is accumulated into (zero it before the first basis choice).
 * d_state: [n_init * n_sets][2^n_qubits] complex128, state = init * n_sets + set (qdc_sv_run order).
 * Draw j of state (init, set) uses the counter-based generator Philox4x32-10 keyed by `seed` and
 * `stream_id` (pass the index of the basis choice): the table is a function of the arguments, not of
 * the launch order.  d_workspace: qdc_sv_sample_workspace_doubles(n_qubits, n_init * n_sets) doubles.
 * At most 65535 states per call.  Launches 3 kernels (mass, scan, draw); allocates nothing. */
int64_t qdc_sv_sample_workspace_doubles(int n_qubits, int64_t n_states);
int  qdc_sv_sample_counts(const void* d_state, int n_qubits, int64_t n_init, int64_t n_sets,
                          uint64_t flip_mask, int64_t shots, uint64_t seed, uint64_t stream_id,
                          uint64_t meas_mask, const int32_t* d_columns, double* d_counts, int64_t ld,
                          double* d_workspace, void* stream);

/* ---- K3 + K4: fragment tensor in the Pauli basis straight from the amplitudes ---------------
 * Replaces, in the exact (infinite-shot) limit, the 3^m measurement variants + organize +
 * least-squares inversion + Choi reshaping of qsplit_mlrecon_methods.py:116-317, 428-440 (and
 * cutqc's measure.c / vertical_collapse.c basis change, cutqc/post_process.py:20-30):
 *   T[set][b][kappa] = sum_s prod_j frame[a_j][s_j] * <psi_{s,set}| Pi_b (x) P_(a') |psi_{s,set}>
 * b = value of the non-measured qubits (ascending qubit order), a = Pauli index (I, X, Y, Z) per
 * cut axis, prep axes first (order of h_prep_qubits used in qdc_sv_run) then meas axes
 * (h_meas_qubits); kappa = sum_axis a_axis * h_axis_stride[axis].  h_frame[a * n_codes + s] are the
 * expansion coefficients of Pauli a in the preparation projectors.  d_state holds
 * n_codes^n_prep * n_sets states laid out as by qdc_sv_run (init_lo = 0).
 * n_codes = 2 selects the BASIS-INPUT form: d_state holds the 2^n_prep evolutions of the
 * computational-basis inputs (qdc_sv_run with the code table {|0>, |1>}); every preparation is a
 * linear combination of them, so T[set][b][a] = Tr(P_a M_b), (M_b)_{x'x} = conj(psi_x'(b))
 * psi_x(b), gives the same numbers from 2^p instead of 4^p simulations (h_frame unused;
 * n_prep + n_meas <= 4).
 * d_T: [n_sets][2^(n_qubits - n_meas)][ld] float64, ld >= 4^(n_prep + n_meas). */
int  qdc_sv_pauli_tensor(const void* d_state, int n_qubits, int64_t n_sets, uint64_t flip_mask,
                         int n_prep, int n_codes, const double* h_frame,
                         const int32_t* h_meas_qubits, int n_meas, const int64_t* h_axis_stride,
                         double* d_T, int64_t ld, void* stream);

/* ---- K4b: nearest positive-semidefinite fragment model ----------------------------------------
 * Replaces maximum_likelihood_model (qsplit/qsplit_mlrecon_methods.py:320-356): per Choi block
 * (dim x dim Hermitian, dim = 2^(cut ends of the fragment) in {1, 2, 4, 8, 16}) scipy.linalg.eigh
 * (:330-331), the global sort-and-clip of ALL the fragment's eigenvalues (:333-343) and the rebuild
 * through to_projector (:45-46, 354-356: the complex conjugate of V diag V^+).
 *   d_T    [n_blocks][ld]     Pauli-basis coordinates of the blocks (qdc_sv_pauli_tensor layout)
 *   d_T_out[n_blocks][ld_out] the projected model, same coordinates
 * The change of basis is given as two sparse tables (each block entry touches dim components and
 * vice versa), made by the host from the basis of the direct model:
 *   Lambda[i][j] = sum_{t < dim} d_T[b][d_fwd_k[(i dim + j) dim + t]] * fwd_c[(i dim + j) dim + t]
 *   T'[kappa]    = Re sum_{t < dim} inv_c[kappa dim + t] * L'[j_t][i_t],  d_inv_ij = (i_t << 8) | j_t
 * (coefficients as interleaved (re, im) doubles, 16-byte aligned).
 * One warp per block: Cholesky test of Lambda + 1e-14 trace I; if every block passes (exact data)
 * the result is conj(Lambda).  Otherwise parallel-order Jacobi per block, the clipping threshold
 * t0 (root of sum max(lambda - t, 0) = sum lambda, equal to the reference's sequential walk) by
 * Newton steps over deterministic grid-wide reductions, and the rebuild.  force_full > 0 skips
 * the test and takes the full path; force_full < 0 skips the test and takes the exact-data path
 * (for blocks the caller knows to be positive semidefinite by construction: models of exactly
 * simulated fragments, where the projection changes nothing beyond rounding).  d_workspace: qdc_ml_workspace_bytes(n_blocks, dim) bytes, 16-byte aligned; its first
 * 40 bytes report {int need_full, done, -, newton_steps; double t0, sum lambda, sum |lambda|}.
 * Nothing is read back by the host. */
int64_t qdc_ml_workspace_bytes(int64_t n_blocks, int dim);
int qdc_ml_project(const double* d_T, int64_t n_blocks, int dim, int64_t ld,
                   const int32_t* d_fwd_k, const double* d_fwd_c, const int32_t* d_inv_ij,
                   const double* d_inv_c, double* d_T_out, int64_t ld_out, void* d_workspace,
                   int force_full, void* stream);

/* ---- K5: two-operand contraction over the cut-basis index ---------------------------------
 * Replaces the itertools.product loop of _recombine_using_networks
 * (qsplit_mlrecon_methods.py:412-461) and cutqc build.c:91-169 for two operands:
 *   P~[r, c] = scale * sum_k A[r, k] * Bt[c, k],  r in [row_lo, row_lo + n_rows), c < n_cols.
 * A: [a_rows_total, kdim] (row operand, fragments F-1..1), Bt: [n_cols, kdim] (column operand,
 * fragment 0), kdim = 4^K, kdim % 16 == 0 required (pad with zeros).  Both operands are passed
 * PACKED: qdc_contract_pack turns a row-major [n_rows][kdim] array into the k-group-major layout
 * [kdim/4][rows_pad][4] (zero padded; qdc_contract_packed_doubles(n_rows, kdim) doubles) whose tile
 * slices are contiguous, so the kernel is fed by bulk async copies (TMA 1-D path) alone.
 * FP64 via DMMA (mma.sync m8n8k4).  The MIS epilogue of qdc_mis_reduce is applied to every
 * tile in registers; P is only written when d_P_out != NULL:
 *   store_mode 1: d_P_out[(r - row_lo) * n_cols + c]         (united order, slab-local)
 *   store_mode 2: d_P_out[row_x[r] | col_x[c]]               (full-circuit order, scatter)
 * out_dtype 0 = f64, 1 = f32.
 */
int64_t qdc_contract_packed_doubles(int64_t n_rows, int kdim);
int qdc_contract_pack(const double* d_X, int64_t n_rows, int kdim, double* d_packed, void* stream);
int qdc_contract_f64(const double* d_Ap, int64_t a_rows_total, const double* d_Bp, int64_t row_lo,
                     int64_t n_rows, int64_t n_cols, int kdim, double scale, const double* d_norm,
                     double threshold,
                     const uint64_t* d_row_x, const uint64_t* d_row_forb,
                     const uint32_t* d_row_wok, const uint64_t* d_col_x,
                     const uint32_t* d_col_wok,
                     void* d_P_out, int store_mode, int out_dtype,
                     double* d_sums, uint64_t* d_cand_x, double* d_cand_p,
                     unsigned long long* d_cand_count, int64_t cand_cap,
                     double* d_partials, void* stream);

/* The same contraction + fused scoring on the 5th-generation tensor cores (tcgen05.mma kind::i8,
 * TMEM accumulators): the FP64 operands are split into radix-256 digits per element (balanced
 * digits in [-128, 127]) with one power-of-two scale per row; all digit-pair products of order
 * i + j < n_slices are accumulated exactly in int32 and combined exactly in integers.  Error bound
 * (stated, not measured):
 * |P~ - P~_exact| <= 2^(e_r + f_c) * scale * kdim * (n_slices + 1.1) * 2^(-8 n_slices), 2^e_r / 2^f_c
 * = power-of-two ceilings of the row maxima; kdim = 256: 1.4e-9 (5 slices), 6.5e-12 (6 slices)
 * times 2^(e_r + f_c) * scale -- FP64 accumulation itself: 2.8e-14.  kdim <= 256.
 *
 * Step 1, once per operand and evaluation (qdc_i8_prepare): ONE pass over the row-major FP64
 * operand d_X[n_rows][kdim] writes, into a caller-owned 256-byte aligned buffer of
 * qdc_i8_operand_bytes(n_rows, kdim, n_slices) bytes, the n_slices digit planes in the tensor
 * cores' core-matrix order, the row scales, the scoring tables of the column role, the column sums
 * (sum_r x, sum_r weight_r x: they give combined_norm, qsplit_mlrecon_methods.py:471-472, sum p and
 * sum p w without touching the outputs) and the statistics of the error bounds.
 * d_wok: weight | ok << 8 per row (qdc_mis_build_tables; NULL: weight 0, ok).  column_role != 0 for
 * the operand of fragment 0 (streamed through the tiles): d_x = its col_x table; every 32 adjacent
 * columns then share an integer unit (a column more than 2^20 below the largest of its group is
 * sliced at that floor; the bounds use the exponent actually applied).
 *
 * Step 2 (qdc_contract_i8_prepared): plan + contraction + finish, no host synchronisation.
 *   n_stored: digit planes the operands hold; n_slices: planes used, 2..n_stored, or 0 = the
 *   smallest count in 3..n_stored whose a-posteriori bounds for THESE operands -- per output, and
 *   on the masked sum (n_nodes * (sum_{r ok} 2^e_r) (sum_{c ok} 2^f_c) * ... ) -- are both <= tol,
 *   chosen on the device (the kernels of the other counts return at once).
 *   norm_mode 0: none; 1: the total of the distribution, from the column sums of both operands
 *   plus d_extra_colsums (n_extra x kdim partial column sums of row-operand rows that are not in
 *   this slab: qdc_colsum; NULL if the slab is the whole operand); 2: *d_norm_in.
 *   d_A_src / d_Bt_src: the FP64 rows the operands were prepared from (store_mode 0: outputs the
 *   integer candidate filter flags are re-evaluated exactly from them).
 *   row tables: indexed row_lo + local row.  store_mode 0, 1 = united slab float64, 2 = float32.
 *   d_result: 16 doubles [sum p, sum p w [indset], sum p w, norm, slices used, bound per output,
 *   bound on the masked sum, candidate count (uint64), largest column exponent, flagged outputs
 *   (uint64), lost candidates (uint64), ...] -- written, not accumulated.  d_cand_*: candidate sink as
 *   qdc_mis_reduce; k, d_top_*: as qdc_topk_select (k = 0: no selection).  With store_mode 0 the
 *   contraction only FLAGS outputs (an integer test on the leading digit order: a superset of the
 *   candidates) into the candidate arrays and a follow-up kernel evaluates them exactly from the FP64
 *   operands, so cand_cap should leave room for the superset (the Python side passes up to 2^22
 *   entries); entry [10] counts candidates that found no room (0 = the result is complete).
 *   d_partials: 3 * qdc_reduce_partials() doubles. */
int64_t qdc_i8_operand_bytes(int64_t n_rows, int kdim, int n_slices);
int qdc_i8_prepare(const double* d_X, int64_t n_rows, int kdim, int n_slices, int column_role,
                   const uint32_t* d_wok, const uint64_t* d_x, void* d_operand, void* stream);
int qdc_contract_i8_prepared(const void* d_opA, int64_t n_rows, const void* d_opB, int64_t n_cols,
                             int kdim, int n_stored, int n_slices, double tol, int n_nodes,
                             double scale, int norm_mode, const double* d_norm_in,
                             const double* d_extra_colsums, int n_extra, double threshold,
                             const double* d_A_src, const double* d_Bt_src, int64_t row_lo,
                             const uint64_t* d_row_x, const uint64_t* d_row_forb,
                             const uint32_t* d_row_wok, void* d_P_out, int store_mode,
                             double* d_result, uint64_t* d_cand_x, double* d_cand_p, int64_t cand_cap,
                             int k, uint64_t* d_top_x, double* d_top_p, double* d_partials, void* stream);

/* Column sums of rows [r0_begin, r0_end) and [r1_begin, r1_end) of a row-major FP64 matrix:
 * d_partials[qdc_colsum_partials()][kdim], one row per CTA, to be added in row order (fixed order:
 * run-to-run deterministic).  Full grid, 16-byte loads. */
int qdc_colsum_partials(void);
int qdc_colsum(const double* d_X, int64_t r0_begin, int64_t r0_end, int64_t r1_begin, int64_t r1_end,
               int kdim, double* d_partials, void* stream);

/* sum over all (r, c) of P~ without forming it: scale * sum_k (sum_r A[r,k]) (sum_c Bt[c,k]).
 * This is combined_norm of qsplit_mlrecon_methods.py:471.  d_colsums: workspace of
 * qdc_norm_workspace_doubles(kdim) doubles (fixed-order partial sums: the result is
 * run-to-run deterministic); d_norm: 1 double. */
int qdc_norm_workspace_doubles(int kdim);
int qdc_contract_norm(const double* d_A, int64_t n_rows_total, const double* d_Bt,
                      int64_t n_cols, int kdim, double scale, double* d_colsums,
                      double* d_norm, void* stream);

/* qdc_contract_norm without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  Same inputs, same result (bit for bit) as
 * qdc_contract_norm, but the column-sum partials live in a file-scope device array of the library
 * (1 212 416 bytes per GPU that loads it, held until the process exits) instead of a caller
 * workspace; the call itself performs no cudaMalloc / cudaFree and calls no allocating routine.
 * Limit: kdim <= qdc_contract_norm_static_max_kdim() = 256; larger kdim returns QDC_ERR_ARG with a
 * message naming the limit.  All calls on one GPU share the scratch: calls that overlap on the GPU
 * (different streams) race -- keep them on one stream or order them with events.  The scratch keeps
 * the previous call's contents; every entry a call reads it has written first, nothing needs
 * clearing.  The first call on each device looks up the array's address (cudaGetSymbolAddress):
 * make it outside stream capture. */
int qdc_contract_norm_static_max_kdim(void);
int qdc_contract_norm_static(const double* d_A, int64_t n_rows_total, const double* d_Bt,
                             int64_t n_cols, int kdim, double scale, double* d_norm, void* stream);

/* ---- CutQC-style stages (K3c, K4c, K6, K7) --------------------------------------------------
 * In-memory equivalents of the reference's four C programs (which talk through %e text files;
 * the CLI/file-compatible shims live in the Python package `qdc_b200.cutqc`).  Bit conventions:
 * index bit q <-> subcircuit qubit q (measure.c decToBinary/combine_effective_O_state are
 * big-endian arrays of the same integer).  All vectors are float64 on the device.
 *
 * measure (cutqc/measure.c:123-161): d_raw [n_inst][2^n_qubits] (dtype 0 f64, 1 f32);
 *   out[i][e] = sum over the non-kept qubits of (-1)^popcount(x & sign_mask_i) * raw[i][x], where
 *   keep_mask = qubits measured in "comp" (they form the effective index e, ascending) and
 *   sign_mask_i = qubits of instance i whose measurement label is neither "I" nor "comp". */
int qdc_cutqc_measure(const void* d_raw, int dtype, int n_qubits, int64_t n_inst,
                      uint64_t keep_mask, const uint64_t* d_sign_masks, double* d_out,
                      void* stream);
/* vertical collapse (cutqc/vertical_collapse.c:27-104): CSR over kron terms,
 *   out[t][j] = sum_{k in [term_ptr[t], term_ptr[t+1])} coeff[k] * measured[inst_idx[k]][j];
 *   d_nonzero[t] = 1 iff some |out[t][j]| > 1e-16 (the early-termination test, :65-75). */
int qdc_cutqc_collapse(const double* d_measured, int64_t len, int n_terms,
                       const int32_t* d_term_ptr, const int32_t* d_inst_idx,
                       const int32_t* d_coeff, double* d_out, int32_t* d_nonzero, void* stream);
/* merge (cutqc/merge.c:28-122): d_in [n_vec][2^n_effective]; qubits in active_mask keep their
 *   relative order in the output index, qubits in zoom_mask must equal zoom_val, all others are
 *   summed out: d_out [n_vec][2^popcount(active_mask)]. */
int qdc_cutqc_merge(const double* d_in, int64_t n_vec, int n_effective, uint64_t active_mask,
                    uint64_t zoom_mask, uint64_t zoom_val, double* d_out, void* stream);
/* build, step 1 (cutqc/build.c:97-160): operand of one subcircuit, A[i][t] = vecs[idx[t]][i] for
 *   summation term t (idx[t] < 0: the kron file is missing, the term is zero, build.c:35-67),
 *   zero for t in [n_terms, kdim).  d_vecs [*][len], d_A [len][kdim] row-major. */
int qdc_cutqc_gather(const double* d_vecs, int64_t len, const int32_t* d_idx, int n_terms,
                     int kdim, double* d_A, void* stream);
/* build, step 2 for more than two subcircuits: W[(i, j)][t] = A0[i][t] * A1[j][t] (the growing
 *   cblas_sger product of build.c:142, kept factored per term).  Step 3 is qdc_contract_f64 with
 *   store_mode 1 and scale 0.5^num_cuts (build.c:169). */
int qdc_khatri_rao(const double* d_A0, int64_t m0, const double* d_A1, int64_t m1, int kdim,
                   double* d_W, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* QDC_B200_H */
