// K1-K3: batched statevector engine (complex128), batch = initial-state variant x parameter set.
//
// Replaces the simulator behind qiskit.execute(circuits, backend, shots) on the reference's hot
// path (qsplit/qsplit_mlrecon_methods.py:20-42, 116-162; mis.py:649) for the gate set of its
// ansatz modules (ansatz/dqv_cut_ansatz.py:127-171; ansatz/qaoa_plus.py:8-37): every gate is a
// 2x2 on one target applied where the control qubits hold given values.
//
// Design (B200): the gate list is compiled on the host into GROUPS.  One group = the gates whose
// non-diagonal targets all lie in a set L of `t` qubits; a CTA owns one tile = the 2^t
// amplitudes that differ only in the qubits of L, keeps it in shared memory, applies the whole
// group there (controls outside L are constant per tile -> the gate is applied or skipped for the
// whole tile; diagonal gates need no locality at all), and touches global memory once per group
// (16 B read + 16 B write per amplitude; the first group builds the initial product state in
// shared memory and reads nothing).  L always contains the lowest `c` qubits, so global accesses
// are runs of 2^c * 16 B.  Fragments of <= t qubits run their whole circuit in ONE launch with
// one CTA per batch element.  Bound: shared-memory bandwidth inside a group (32 B per amplitude
// per non-diagonal gate), L2/HBM between groups (algorithmic bytes per group = 32 B * 2^n * batch).
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "qdc_common.cuh"

namespace {

constexpr int MAX_RUNS = 24;
constexpr int MAX_FIXED = 16;
constexpr int MAX_PREP = 10;
constexpr int MAX_CODES = 6;
constexpr int MAX_DIAG_RUN = 32;      // <= smallest CTA size

struct __align__(16) SvOp {   // 64 bytes
    uint32_t kind;            // 0 general 2x2, 1 diagonal
    int32_t tbit;             // tile-local bit of the target, -1: diagonal on a non-local qubit
    int32_t gtarget;          // full-index qubit of the target
    int32_t mat;              // matrix slot
    uint32_t lmask, lval;     // controls inside the tile (tile-local bit positions)
    uint64_t gmask, gval;     // controls outside the tile (full-index positions)
    uint8_t n_fixed;          // entries of fixed[]: (target bit for kind 0) + local control bits, ascending
    uint8_t diag_run;         // kind 1 without controls: consecutive such ops starting here
    uint8_t pad[2];
    uint8_t fixed[MAX_FIXED];
    uint32_t fmask;           // OR of 1 << fixed[k]
};
static_assert(sizeof(SvOp) == 64, "SvOp layout");

struct MatDesc {              // how to build matrix slot i for a parameter set
    int32_t kind, slot, conj_x, pad;
    double scale, offset;
    double m[8];
};

struct Group {
    int first_op, n_ops;
    uint64_t local_mask;
    int n_runs;
    uint8_t run_pos[MAX_RUNS], run_len[MAX_RUNS];
};

struct GroupArgs {
    double2* state;
    const SvOp* ops;
    const double* mats;       // [n_sets][n_mats][8]
    long long n_tiles;        // tiles per state
    long long init_lo;
    uint64_t prep_mask, hole_mask;
    int n, t, n_sets, n_mats, op_lo, op_hi, first, n_runs, n_prep, n_codes;
    int bulk;                 // stage the tile with bulk copies (runs of >= 64 bytes)
    // fused MIS scoring of the final state (last group of qdc_sv_run_score): sum p, sum p w [independent], sum p w
    // per CTA into score_partials[3 blockIdx.x + .], w = popcount of the node bits, straight from the tile
    const uint64_t* score_tab;    // per tile-local index: logical node bits | w << 48 | [independent among themselves] << 63
    const uint64_t* score_adj;    // adjacency masks of the graph nodes (bit i of adj[j] <=> edge (i, j))
    double* score_partials;
    uint64_t score_flip, score_nodes;
    int score, store;         // store = 0: the state is not written back (only its score is wanted)
    uint8_t run_pos[MAX_RUNS], run_len[MAX_RUNS];
    uint8_t prep_q[MAX_PREP];
    double code[MAX_CODES][4];
};

__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 cfma(double2 a, double2 b, double2 c) {   // a*b + c
    return make_double2(fma(a.x, b.x, fma(-a.y, b.y, c.x)), fma(a.x, b.y, fma(a.y, b.x, c.y)));
}

// tile-local index -> full index with zeros in the non-local positions
__device__ __forceinline__ uint64_t expand_runs(uint64_t l, int n_runs, const uint8_t* pos,
                                                const uint8_t* len) {
    for (int r = 0; r < n_runs; ++r) {
        const int p = pos[r], w = len[r];
        l = ((l >> p) << (p + w)) | (l & ((1ull << p) - 1ull));
    }
    return l;
}

// deposit the bits of i into the zero positions of fm (ascending); registers only
__device__ __forceinline__ uint32_t deposit_bits(uint32_t i, uint32_t fm) {
    while (fm) {
        const uint32_t below = (fm & (0u - fm)) - 1u;          // bits under the lowest fixed position
        i = ((i & ~below) << 1) | (i & below);
        fm &= fm - 1u;
    }
    return i;
}

__device__ __forceinline__ uint32_t insert_fixed(uint32_t i, int n_fixed, const uint8_t* fixed) {
    for (int f = 0; f < n_fixed; ++f) {
        const int p = fixed[f];
        i = ((i >> p) << (p + 1)) | (i & ((1u << p) - 1u));
    }
    return i;
}

// ---- bulk-copy staging of a tile (cp.async.bulk, one instruction per contiguous run) ------------------
// A tile is 2^t amplitudes whose full indices differ only in the group's local qubits; its lowest
// 2^p0 local indices (p0 = position of the lowest non-local qubit >= the coalesce bits) are contiguous
// in HBM, so the tile is 2^(t - p0) runs of 16 << p0 bytes: each is ONE bulk copy HBM -> shared
// memory (completion counted in bytes on an mbarrier) and one bulk copy back, instead of a load,
// a 64-bit index walk and a store per amplitude.
__device__ __forceinline__ uint32_t sv_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sv_mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sv_smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void sv_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sv_smem_u32(bar)), "r"(bytes) : "memory");
}
// bounded wait: a copy that never lands traps instead of hanging the device
__device__ __forceinline__ void sv_mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(sv_smem_u32(bar)), "r"(parity) : "memory");
        if (spin > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void sv_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(sv_smem_u32(dst)), "l"(src), "r"(bytes), "r"(sv_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void sv_bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(dst), "r"(sv_smem_u32(src)), "r"(bytes) : "memory");
}

constexpr int OPS_CHUNK = 32;          // ops staged in shared memory at a time
constexpr int TAB_LO_BITS = 6;          // diagonal-run phase table over the low tile bits

// Index walks use the carry trick: if x = expand(l) has zeros at the positions of mask M, then
// expand(l + s) = ((x | M) + expand(s)) & ~M -- three integer ops per step instead of a loop.
__global__ void __launch_bounds__(512) sv_group_kernel(const GroupArgs a) {
    extern __shared__ double2 tile[];
    __shared__ SvOp sh_ops[OPS_CHUNK];
    __shared__ double sh_mats[OPS_CHUNK][8];
    __shared__ double2 sh_tab[(1 << TAB_LO_BITS) + (1 << (13 - TAB_LO_BITS))];
    __shared__ __align__(8) uint64_t sh_bar;

    const long long blk = blockIdx.x;
    const long long batch = blk / a.n_tiles;
    const long long tile_id = blk - batch * a.n_tiles;
    uint64_t base = 0;
    {
        uint64_t rem = (uint64_t)tile_id;
        for (int r = 0; r < a.n_runs; ++r) {
            const int w = a.run_len[r];
            base |= (rem & ((1ull << w) - 1ull)) << a.run_pos[r];
            rem >>= w;
        }
    }
    double2* st = a.state + (batch << a.n);
    const uint32_t T = 1u << a.t;
    const uint32_t nthr = blockDim.x, tid = threadIdx.x;
    const uint64_t hole = a.hole_mask;
    const uint64_t x_first = expand_runs(tid, a.n_runs, a.run_pos, a.run_len);
    const uint64_t x_step = expand_runs(nthr, a.n_runs, a.run_pos, a.run_len);
    // contiguous runs of the tile in HBM (see the bulk-copy note above)
    const int p0 = a.n_runs ? (a.run_pos[0] < a.t ? a.run_pos[0] : a.t) : a.t;
    const uint32_t n_chunks = T >> p0, chunk_bytes = 16u << p0;
    const bool bulk = a.bulk != 0;

    if (a.first) {
        // initial product state: prep qubit j in code[digit_j], every other qubit in |0>.  All but
        // 2^n_prep amplitudes of a state are zero: clear the tile, then place the few that fall in it.
        for (uint32_t l = tid; l < T; l += nthr) tile[l] = make_double2(0.0, 0.0);
        __syncthreads();
        if ((base & ~a.prep_mask) == 0ull) {                       // uniform over the CTA
            long long v = a.init_lo + batch / a.n_sets;
            int digit[MAX_PREP];
#pragma unroll
            for (int j = MAX_PREP - 1; j >= 0; --j)
                if (j < a.n_prep) { digit[j] = (int)(v % a.n_codes); v /= a.n_codes; }
            for (uint32_t k = tid; k < (1u << a.n_prep); k += nthr) {
                uint64_t full = 0;
                double2 amp = make_double2(1.0, 0.0);
#pragma unroll 1
                for (int j = 0; j < a.n_prep; ++j) {
                    const int bit = (int)((k >> j) & 1u);
                    full |= (uint64_t)bit << a.prep_q[j];
                    amp = cmul(amp, make_double2(a.code[digit[j]][2 * bit], a.code[digit[j]][2 * bit + 1]));
                }
                if ((full & hole) != base) continue;               // another tile's amplitude
                uint64_t l = full & ~hole;
                for (int r = a.n_runs - 1; r >= 0; --r) {          // full index -> tile-local index
                    const int p = a.run_pos[r], w = a.run_len[r];
                    l = ((l >> (p + w)) << p) | (l & ((1ull << p) - 1ull));
                }
                tile[l] = amp;
            }
        }
    } else if (bulk) {
        if (tid == 0) sv_mbar_init(&sh_bar, 1);
        __syncthreads();
        if (tid == 0) sv_mbar_expect_tx(&sh_bar, T * 16u);
        // (every thread issues its own copy: a few lanes issuing them all in a loop was measured 1.5x slower for the
        // whole kernel -- the issue of a bulk copy is slow, its parallelism across threads is what hides it)
        for (uint32_t j = tid; j < n_chunks; j += nthr)
            sv_bulk_g2s(tile + ((size_t)j << p0), st + (expand_runs((uint64_t)j << p0, a.n_runs, a.run_pos, a.run_len) | base),
                        chunk_bytes, &sh_bar);
        // only warp 0 polls the barrier; the others sleep at the __syncthreads that follows (256 spinning threads were
        // 13 % of the issued instructions)
        if (tid < 32) sv_mbar_wait(&sh_bar, 0);
        __syncthreads();
    } else {
        // all loads of a batch are issued before the first one is consumed (8 in flight per thread)
        constexpr int U = 8;
        uint64_t x = x_first;
        for (uint32_t l = tid; l < T; l += nthr * U) {
            double2 v[U];
            uint64_t xs = x;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (l + u * nthr < T) v[u] = __ldcs(st + (xs | base));
                xs = ((xs | hole) + x_step) & ~hole;
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (l + u * nthr < T) tile[l + u * nthr] = v[u];
            x = xs;
        }
    }

    const double* mats = a.mats + (size_t)(batch % a.n_sets) * a.n_mats * 8;
    for (int cb = a.op_lo; cb < a.op_hi; cb += OPS_CHUNK) {
        const int n_chunk = a.op_hi - cb < OPS_CHUNK ? a.op_hi - cb : OPS_CHUNK;
        __syncthreads();                                   // tile ready / previous chunk consumed
        {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(a.ops + cb);
            uint32_t* dst = reinterpret_cast<uint32_t*>(sh_ops);
            for (uint32_t w = tid; w < (uint32_t)n_chunk * 16u; w += nthr) dst[w] = src[w];
            const double* msrc = mats + (size_t)cb * 8;    // matrix slot == op index
            double* mdst = &sh_mats[0][0];
            for (uint32_t w = tid; w < (uint32_t)n_chunk * 8u; w += nthr) mdst[w] = msrc[w];
        }
        __syncthreads();
        for (int i = 0; i < n_chunk;) {
            const SvOp& op = sh_ops[i];
            if (!(op.kind == 1 && op.diag_run >= 1) && (base & op.gmask) != op.gval) { ++i; continue; }   // uniform over the CTA (a run checks per op)
            const double* m = sh_mats[i];
            if (op.kind == 0) {
                const double2 m00 = make_double2(m[0], m[1]), m01 = make_double2(m[2], m[3]);
                const double2 m10 = make_double2(m[4], m[5]), m11 = make_double2(m[6], m[7]);
                const uint32_t count = 1u << (a.t - op.n_fixed), tb = 1u << op.tbit;
                const uint32_t fm = op.fmask, lval = op.lval;
                uint32_t e = deposit_bits(tid, fm);
                if (count <= nthr) {                               // controlled gates: at most one pair per thread
                    if (tid < count) {
                        const uint32_t l0 = e | lval, l1 = l0 | tb;
                        const double2 a0 = tile[l0], a1 = tile[l1];
                        tile[l0] = cfma(m00, a0, cmul(m01, a1));
                        tile[l1] = cfma(m10, a0, cmul(m11, a1));
                    }
                } else {
                    const uint32_t e_step = deposit_bits(nthr, fm);
                    for (uint32_t k = tid; k < count; k += nthr, e = ((e | fm) + e_step) & ~fm) {
                        const uint32_t l0 = e | lval, l1 = l0 | tb;
                        const double2 a0 = tile[l0], a1 = tile[l1];
                        tile[l0] = cfma(m00, a0, cmul(m01, a1));
                        tile[l1] = cfma(m10, a0, cmul(m11, a1));
                    }
                }
                ++i;
            } else if (op.diag_run >= 1) {
                // joint pass over a run of diagonal gates that depend on at most ONE tile-local bit each (uncontrolled,
                // or one control with control and target not both in the tile): the phase of tile index l
                // factors into a table over its low bits times a table over its high bits (the
                // latter also carries the constant factor of the gates on non-local qubits)
                const int run = op.diag_run;
                const int h = a.t < TAB_LO_BITS ? a.t : TAB_LO_BITS;
                const uint32_t n_lo = 1u << h, n_hi = 1u << (a.t - h);
                for (uint32_t w = tid; w < n_lo + n_hi; w += nthr) {
                    const bool hi = w >= n_lo;
                    const uint32_t idx = hi ? w - n_lo : w;
                    double2 f = make_double2(1.0, 0.0);
                    for (int k = 0; k < run; ++k) {
                        const SvOp& o = sh_ops[i + k];
                        const double* mk = sh_mats[i + k];
                        if ((base & o.gmask) != o.gval) continue;          // a control outside the tile that is not satisfied
                        int bit = -1;
                        if (o.lmask) {
                            // one control inside the tile and the target outside (the compiler admits no other
                            // controlled gate to a run): a 1-qubit diagonal diag(1, d) or diag(d, 1) on the control bit
                            const int cb = __ffs((int)o.lmask) - 1;
                            int cbit = -1;
                            if (!hi && cb < h) cbit = (idx >> cb) & 1;
                            else if (hi && cb >= h) cbit = (idx >> (cb - h)) & 1;
                            if (cbit >= 0 && cbit == (int)((o.lval >> cb) & 1u)) bit = (int)((base >> o.gtarget) & 1ull);
                        } else if (o.tbit < 0) { if (hi) bit = (int)((base >> o.gtarget) & 1ull); }
                        else if (!hi && o.tbit < h) bit = (idx >> o.tbit) & 1;
                        else if (hi && o.tbit >= h) bit = (idx >> (o.tbit - h)) & 1;
                        if (bit >= 0) f = cmul(f, bit ? make_double2(mk[6], mk[7]) : make_double2(mk[0], mk[1]));
                    }
                    sh_tab[w] = f;
                }
                __syncthreads();
                for (uint32_t l = tid; l < T; l += nthr)
                    tile[l] = cmul(tile[l], cmul(sh_tab[l & (n_lo - 1u)], sh_tab[n_lo + (l >> h)]));
                i += run;
            } else {
                // controlled diagonal gate: fixed[] = local control bits
                const double2 d0 = make_double2(m[0], m[1]), d1 = make_double2(m[6], m[7]);
                const uint32_t count = 1u << (a.t - op.n_fixed);
                const bool local = op.tbit >= 0;
                const double2 dg = ((base >> op.gtarget) & 1ull) ? d1 : d0;
                const uint32_t fm = op.fmask, lval = op.lval;
                uint32_t e = deposit_bits(tid, fm);
                const uint32_t e_step = count > nthr ? deposit_bits(nthr, fm) : 0u;
                for (uint32_t k = tid; k < count; k += nthr, e = ((e | fm) + e_step) & ~fm) {
                    const uint32_t l = e | lval;
                    const double2 d = local ? (((l >> op.tbit) & 1u) ? d1 : d0) : dg;
                    tile[l] = cmul(tile[l], d);
                }
                ++i;
            }
            __syncthreads();
        }
    }
    if (a.score) {
        // ---- MIS objective of this tile, from shared memory: no second pass over the state ----
        __shared__ uint64_t sh_forb;
        __shared__ int sh_bw;
        __shared__ double sh_red[3][16];
        __syncthreads();
        if (tid < 32) {
            // the tile's fixed (non-local) bits: their weight, whether they are independent among themselves,
            // and the union of their neighbourhoods (a local node conflicts with them iff it lies in it)
            const uint64_t xb = (base ^ (a.score_flip & hole)) & a.score_nodes;
            uint64_t f = 0;
            bool bad = false;
            for (int i = (int)tid; i < a.n; i += 32)
                if ((xb >> i) & 1ull) { const uint64_t ai = a.score_adj[i]; f |= ai; bad |= (ai & xb) != 0ull; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) f |= __shfl_xor_sync(0xffffffffu, f, o);
            bad = __any_sync(0xffffffffu, bad);
            if (tid == 0) { sh_forb = f; sh_bw = __popcll(xb) | (bad ? 0 : 256); }
        }
        __syncthreads();
        const uint64_t forb_b = sh_forb;
        const int wb = sh_bw & 255;
        const bool okb = (sh_bw >> 8) != 0;
        double s_p = 0.0, s_pwok = 0.0, s_pw = 0.0;
        for (uint32_t l = tid; l < T; l += nthr) {
            const uint64_t e = __ldg(a.score_tab + l);
            const double2 v = tile[l];
            const double p = v.x * v.x + v.y * v.y;
            const double pw = p * (double)((int)((e >> 48) & 0x7full) + wb);
            s_p += p; s_pw += pw;
            if (okb && (e >> 63) && ((e & 0xffffffffffffull) & forb_b) == 0ull) s_pwok += pw;
        }
        s_p = warp_sum(s_p); s_pwok = warp_sum(s_pwok); s_pw = warp_sum(s_pw);
        if ((tid & 31) == 0) { sh_red[0][tid >> 5] = s_p; sh_red[1][tid >> 5] = s_pwok; sh_red[2][tid >> 5] = s_pw; }
        __syncthreads();
        if (tid < 3) {
            double t = 0.0;
            for (uint32_t w = 0; w < (nthr >> 5); ++w) t += sh_red[tid][w];
            a.score_partials[3 * (size_t)blk + tid] = t;
        }
        if (!a.store) return;
    }
    if (bulk) {
        // the gate passes wrote the tile through the generic proxy: order them before the bulk reads
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        for (uint32_t j = tid; j < n_chunks; j += nthr)
            sv_bulk_s2g(st + (expand_runs((uint64_t)j << p0, a.n_runs, a.run_pos, a.run_len) | base),
                        tile + ((size_t)j << p0), chunk_bytes);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory is released at exit
        return;
    }
    __syncthreads();
    {
        uint64_t x = x_first;
        for (uint32_t l = tid; l < T; l += nthr, x = ((x | hole) + x_step) & ~hole) st[x | base] = tile[l];
    }
}

// score table of a group: entry l = logical node bits of tile-local index l | popcount << 48 | [no edge inside] << 63
__global__ void sv_score_table_kernel(uint64_t* __restrict__ tab, int t, int n_runs, GroupArgs a) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= (1u << t)) return;
    const uint64_t local_mask = expand_runs((1ull << t) - 1ull, n_runs, a.run_pos, a.run_len);
    const uint64_t x = (expand_runs(l, n_runs, a.run_pos, a.run_len) ^ (a.score_flip & local_mask)) & a.score_nodes;
    bool ok = true;
    for (uint64_t m = x; m; m &= m - 1) {
        const int i = __ffsll((long long)m) - 1;
        if (a.score_adj[i] & x) { ok = false; break; }
    }
    tab[l] = x | ((uint64_t)__popcll(x) << 48) | ((uint64_t)(ok ? 1 : 0) << 63);
}

// sums[set][j] = sum over the tiles of a state of partials[3 tile + j], fixed order (deterministic)
__global__ void __launch_bounds__(1024) sv_score_finish_kernel(const double* __restrict__ partials, long long n_tiles,
                                                               double* __restrict__ sums) {
    __shared__ double red[3][32];
    const double* p = partials + 3 * (size_t)blockIdx.x * n_tiles;
    double a = 0.0, b = 0.0, c = 0.0;
    for (long long i = threadIdx.x; i < n_tiles; i += blockDim.x) { a += p[3 * i]; b += p[3 * i + 1]; c += p[3 * i + 2]; }
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = a; red[1][threadIdx.x >> 5] = b; red[2][threadIdx.x >> 5] = c; }
    __syncthreads();
    if (threadIdx.x < 3) {
        double t = 0.0;
        for (int w = 0; w < 32; ++w) t += red[threadIdx.x][w];
        sums[3 * blockIdx.x + threadIdx.x] = t;
    }
}

// one thread per (parameter set, matrix slot)
__global__ void sv_matrix_kernel(const MatDesc* __restrict__ desc, int n_mats, long long n_sets,
                                 const double* __restrict__ params, int n_params,
                                 double* __restrict__ mats) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n_sets * n_mats) return;
    const long long set = i / n_mats;
    const MatDesc d = desc[i - set * n_mats];
    const double ang = d.scale * (d.slot >= 0 ? params[set * n_params + d.slot] : 0.0) + d.offset;
    double m[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const double r = 0.70710678118654752440;
    double s, c;
    switch (d.kind) {
        case QDC_G_ID: m[0] = 1; m[6] = 1; break;
        case QDC_G_X: m[2] = 1; m[4] = 1; break;
        case QDC_G_Y: m[3] = -1; m[5] = 1; break;
        case QDC_G_Z: m[0] = 1; m[6] = -1; break;
        case QDC_G_H: m[0] = r; m[2] = r; m[4] = r; m[6] = -r; break;
        case QDC_G_S: m[0] = 1; m[7] = 1; break;
        case QDC_G_SDG: m[0] = 1; m[7] = -1; break;
        case QDC_G_T: m[0] = 1; m[6] = r; m[7] = r; break;
        case QDC_G_TDG: m[0] = 1; m[6] = r; m[7] = -r; break;
        case QDC_G_RX: sincos(0.5 * ang, &s, &c); m[0] = c; m[3] = -s; m[5] = -s; m[6] = c; break;
        case QDC_G_RY: sincos(0.5 * ang, &s, &c); m[0] = c; m[2] = -s; m[4] = s; m[6] = c; break;
        case QDC_G_RZ: sincos(0.5 * ang, &s, &c); m[0] = c; m[1] = -s; m[6] = c; m[7] = s; break;
        case QDC_G_U1: sincos(ang, &s, &c); m[0] = 1; m[6] = c; m[7] = s; break;
        default:
            for (int k = 0; k < 8; ++k) m[k] = d.m[k];
    }
    if (d.conj_x) {                       // X.M.X: swap rows and columns
        double t;
        t = m[0]; m[0] = m[6]; m[6] = t;  t = m[1]; m[1] = m[7]; m[7] = t;
        t = m[2]; m[2] = m[4]; m[4] = t;  t = m[3]; m[3] = m[5]; m[5] = t;
    }
    double* out = mats + i * 8;
    for (int k = 0; k < 8; ++k) out[k] = m[k];
}

__global__ void __launch_bounds__(256) sv_probs_kernel(const double2* __restrict__ st, int n,
                                                       long long total, uint64_t flip,
                                                       double* __restrict__ probs) {
    const uint64_t mask = (1ull << n) - 1ull;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const uint64_t x = (uint64_t)i & mask;
        const double2 v = st[((uint64_t)i & ~mask) | (x ^ flip)];
        probs[i] = v.x * v.x + v.y * v.y;
    }
}

// ---- fragment tensor in the Pauli basis -----------------------------------------------------
struct TensorArgs {
    const double2* state;
    double* T;
    long long n_sets, n_b, ld;
    uint64_t flip;
    int n, n_prep, n_meas, bb;      // bb = final-bit values per CTA
    int n_basis;                    // basis-input mode: leading axes that select one of 2^n_basis states
    int n_runs;                     // runs of measured-qubit positions (holes of the b deposit)
    uint8_t run_pos[MAX_RUNS], run_len[MAX_RUNS];
    uint8_t meas_q[8];
    long long stride[12];           // kappa stride per axis (prep axes, then meas axes)
    double frame[16];               // frame[a * 4 + s]
};

__global__ void __launch_bounds__(256) sv_pauli_tensor_kernel(const TensorArgs a) {
    extern __shared__ double2 sh_raw[];
    const int p = a.n_prep, m = a.n_meas;
    const int n_s = 1 << (2 * p), n_mu = 1 << m, n_pm = 1 << (2 * m);
    const int n_kappa = n_s * n_pm;
    double2* amp = sh_raw;                                        // [bb][n_s][n_mu]
    double2* C = sh_raw + (size_t)a.bb * n_s * n_mu;             // [bb][n_s][n_pm] complex work array (m > 0)
    double* R = reinterpret_cast<double*>(C);                    // [bb][n_s][n_pm], aliases C
    const long long set = blockIdx.y;
    const long long b0 = (long long)blockIdx.x * a.bb;
    const int tid = threadIdx.x, nthr = blockDim.x;

    // 1. gather amplitudes (bl fastest: neighbouring b are neighbouring addresses)
    for (int idx = tid; idx < a.bb * n_s * n_mu; idx += nthr) {
        const int bl = idx % a.bb, rest = idx / a.bb;
        const int mu = rest % n_mu, s = rest / n_mu;
        const long long b = b0 + bl;
        double2 v = make_double2(0.0, 0.0);
        if (b < a.n_b) {
            uint64_t x = expand_runs((uint64_t)b, a.n_runs, a.run_pos, a.run_len);
            for (int j = 0; j < m; ++j) x |= (uint64_t)((mu >> j) & 1) << a.meas_q[j];
            v = a.state[(((long long)s * a.n_sets + set) << a.n) | (long long)(x ^ a.flip)];
        }
        amp[((size_t)bl * n_s + s) * n_mu + mu] = v;
    }
    __syncthreads();

    if (m == 0) {
        // no measured qubit in this fragment: the only "Pauli" is the identity, <I> = |amp|^2
        for (int idx = tid; idx < a.bb * n_s; idx += nthr) {
            const double2 v = amp[idx];
            R[idx] = v.x * v.x + v.y * v.y;
        }
        __syncthreads();
    } else if (m <= 2) {
        // few measured qubits: direct sums  R = Re <psi_b| P |psi_b>  (at most 4 terms each)
        for (int idx = tid; idx < a.bb * n_s * n_pm; idx += nthr) {
            const int pm = idx % n_pm, rest = idx / n_pm;            // rest = bl * n_s + s
            uint32_t xmask = 0, zmask = 0;
            int n_y = 0;
            for (int j = 0; j < m; ++j) {                             // axis j: digit j, most significant first
                const int aj = (pm >> (2 * (m - 1 - j))) & 3;         // 0 I, 1 X, 2 Y, 3 Z
                if (aj == 1 || aj == 2) xmask |= 1u << j;
                if (aj == 2 || aj == 3) zmask |= 1u << j;
                n_y += aj == 2;
            }
            const double2* v = amp + (size_t)rest * n_mu;
            double re = 0.0, im = 0.0;
            for (int mu = 0; mu < n_mu; ++mu) {
                const double2 ket = v[mu], bra = v[mu ^ xmask];      // conj(bra) * ket
                double tr = bra.x * ket.x + bra.y * ket.y, ti = bra.x * ket.y - bra.y * ket.x;
                if (__popc(mu & zmask) & 1) { tr = -tr; ti = -ti; }
                re += tr; im += ti;
            }
            const int ny = n_y & 3;                                   // multiply by i^n_y, keep Re
            R[idx] = ny == 0 ? re : ny == 1 ? -im : ny == 2 ? -re : im;
        }
        __syncthreads();
    } else {
        // 2. Pauli coordinates of rho_b = |psi_b><psi_b| on the measured qubits: outer product in the
        //    matrix-unit basis (digit j = 2 * row bit + column bit of measured qubit j), then one
        //    4-point butterfly per qubit (e00, e01, e10, e11) -> (I, X, Y, Z):
        //    I = r00 + r11, X = r01 + r10, Y = i (r01 - r10), Z = r00 - r11   (Tr(rho P))
        for (int idx = tid; idx < a.bb * n_s * n_pm; idx += nthr) {
            const int dgt = idx % n_pm, rest = idx / n_pm;            // rest = bl * n_s + s
            uint32_t row = 0, col = 0;
            for (int j = 0; j < m; ++j) {
                const int aj = (dgt >> (2 * (m - 1 - j))) & 3;
                row |= (uint32_t)(aj >> 1) << j;
                col |= (uint32_t)(aj & 1) << j;
            }
            const double2 r = amp[(size_t)rest * n_mu + row], c = amp[(size_t)rest * n_mu + col];
            C[idx] = make_double2(r.x * c.x + r.y * c.y, r.y * c.x - r.x * c.y);     // r * conj(c)
        }
        __syncthreads();
        for (int j = 0; j < m; ++j) {
            const int inner = 1 << (2 * (m - 1 - j));
            const int n_vec = a.bb * n_s * n_pm / 4;
            for (int idx = tid; idx < n_vec; idx += nthr) {
                const int lo = idx % inner, hi = idx / inner;
                double2* q = C + (size_t)hi * inner * 4 + lo;
                const double2 r00 = q[0], r01 = q[inner], r10 = q[2 * inner], r11 = q[3 * inner];
                q[0] = make_double2(r00.x + r11.x, r00.y + r11.y);
                q[inner] = make_double2(r01.x + r10.x, r01.y + r10.y);
                q[2 * inner] = make_double2(-(r01.y - r10.y), r01.x - r10.x);
                q[3 * inner] = make_double2(r00.x - r11.x, r00.y - r11.y);
            }
            __syncthreads();
        }
        // keep the real parts, compacted in place through registers (R aliases C; a chunk only
        // overwrites complex entries that this or an earlier chunk has already read)
        for (int base_idx = 0; base_idx < a.bb * n_kappa; base_idx += 8 * nthr) {
            double keep[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = base_idx + tid + u * nthr;
                keep[u] = idx < a.bb * n_kappa ? C[idx].x : 0.0;
            }
            __syncthreads();
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = base_idx + tid + u * nthr;
                if (idx < a.bb * n_kappa) R[idx] = keep[u];
            }
            __syncthreads();
        }
    }

    // 3. preparation frame inverse along each prep axis: out[a] = sum_s frame[a][s] in[s]
    for (int j = 0; j < p; ++j) {
        const int inner = n_pm << (2 * (p - 1 - j));              // stride of digit j inside kappa
        const int n_vec = a.bb * n_kappa / 4;
        for (int idx = tid; idx < n_vec; idx += nthr) {
            const int lo = idx % inner, hi = idx / inner;
            double* q = R + (size_t)hi * inner * 4 + lo;
            const double v0 = q[0], v1 = q[inner], v2 = q[2 * inner], v3 = q[3 * inner];
#pragma unroll
            for (int aa = 0; aa < 4; ++aa)
                q[aa * inner] = a.frame[aa * 4] * v0 + a.frame[aa * 4 + 1] * v1 +
                                a.frame[aa * 4 + 2] * v2 + a.frame[aa * 4 + 3] * v3;
        }
        __syncthreads();
    }

    // 4. write T[set][b][kappa_out]
    const int n_axes = p + m;
    for (int idx = tid; idx < a.bb * n_kappa; idx += nthr) {
        const int kap = idx % n_kappa, bl = idx / n_kappa;
        const long long b = b0 + bl;
        if (b >= a.n_b) continue;
        long long ko = 0;
        for (int ax = 0; ax < n_axes; ++ax)
            ko += (long long)((kap >> (2 * (n_axes - 1 - ax))) & 3) * a.stride[ax];
        a.T[(set * a.n_b + b) * a.ld + ko] = R[idx];
    }
}

// ---- fragment tensor, specialised layouts ------------------------------------------------------
// (a) preparation-only fragment (m == 0, p <= 4): T[b][kappa] = sum_s F[kappa][s] |amp_s(b)|^2.
//     A CTA owns 32 consecutive b: a warp reads 512 contiguous bytes of one state, probabilities
//     go to R[s][bl] with a row pitch of 33 doubles, so the frame butterflies (lanes = bl) and
//     the kappa-contiguous write (lanes = kappa) are both bank-conflict free.
__device__ __forceinline__ void pauli4(double2& r00, double2& r01, double2& r10, double2& r11) {
    const double2 i_ = make_double2(r00.x + r11.x, r00.y + r11.y);
    const double2 x_ = make_double2(r01.x + r10.x, r01.y + r10.y);
    const double2 y_ = make_double2(-(r01.y - r10.y), r01.x - r10.x);
    const double2 z_ = make_double2(r00.x - r11.x, r00.y - r11.y);
    r00 = i_; r01 = x_; r10 = y_; r11 = z_;
}

template <int N>   // N = number of axes (1 or 2) held in registers: v[4^N], digit 0 most significant
__device__ __forceinline__ void pauli_axes(double2* v) {
    if (N == 1) {
        pauli4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) pauli4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
#pragma unroll
        for (int e = 0; e < 4; ++e) pauli4(v[e], v[4 + e], v[8 + e], v[12 + e]);
    }
}

constexpr int PT_BB = 32, PT_PITCH = 33, PT_UNROLL = 8;

__device__ __forceinline__ void frame4(const double* f, double& v0, double& v1, double& v2, double& v3) {
    const double a0 = f[0] * v0 + f[1] * v1 + f[2] * v2 + f[3] * v3;
    const double a1 = f[4] * v0 + f[5] * v1 + f[6] * v2 + f[7] * v3;
    const double a2 = f[8] * v0 + f[9] * v1 + f[10] * v2 + f[11] * v3;
    const double a3 = f[12] * v0 + f[13] * v1 + f[14] * v2 + f[15] * v3;
    v0 = a0; v1 = a1; v2 = a2; v3 = a3;
}

template <int M>      // M = measured (exit) qubits of the fragment, 0..2
__global__ void __launch_bounds__(256, 3) sv_prep_tensor_kernel(const TensorArgs a) {
    extern __shared__ double R[];                                  // [n_kappa][PT_PITCH]
    __shared__ double fr[16];
    constexpr int N_PM = 1 << (2 * M), N_MU = 1 << M;
    const int p = a.n_prep, n_s = 1 << (2 * p), n_kappa = n_s * N_PM;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long set = blockIdx.y, b0 = (long long)blockIdx.x * PT_BB;
    if (tid < 16) fr[tid] = a.frame[tid];
    const long long b = b0 + lane;
    const bool live = b < a.n_b;
    if (M == 0) {
        // 1. |amp|^2, PT_UNROLL independent loads in flight per thread
        const long long off = (long long)((uint64_t)b ^ a.flip);
        for (int s0 = warp; s0 < n_s; s0 += 8 * PT_UNROLL) {
            double2 v[PT_UNROLL];
#pragma unroll
            for (int u = 0; u < PT_UNROLL; ++u) {
                const int s = s0 + 8 * u;
                v[u] = (live && s < n_s) ? __ldcs(a.state + ((((long long)s * a.n_sets + set) << a.n) | off))
                                         : make_double2(0.0, 0.0);
            }
#pragma unroll
            for (int u = 0; u < PT_UNROLL; ++u) {
                const int s = s0 + 8 * u;
                if (s < n_s) R[s * PT_PITCH + lane] = v[u].x * v[u].x + v[u].y * v[u].y;
            }
        }
    } else {
        // 1. per (state, b): the 2^M amplitudes -> 4^M Pauli coordinates of the measured qubits
        const uint64_t xb = expand_runs((uint64_t)(live ? b : 0), a.n_runs, a.run_pos, a.run_len);
        constexpr int U = M == 1 ? 4 : 2;                          // states per thread per round
        for (int s0 = warp; s0 < n_s; s0 += 8 * U) {
            double2 am[U][N_MU];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int s = s0 + 8 * u;
#pragma unroll
                for (int mu = 0; mu < N_MU; ++mu) {
                    uint64_t x = xb;
#pragma unroll
                    for (int j = 0; j < M; ++j) x |= (uint64_t)((mu >> j) & 1) << a.meas_q[j];
                    am[u][mu] = (live && s < n_s)
                        ? __ldcs(a.state + ((((long long)s * a.n_sets + set) << a.n) | (long long)(x ^ a.flip)))
                        : make_double2(0.0, 0.0);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int s = s0 + 8 * u;
                if (s >= n_s) continue;
                double2 v[N_PM];
#pragma unroll
                for (int e = 0; e < N_PM; ++e) {
                    int row = 0, col = 0;
#pragma unroll
                    for (int j = 0; j < M; ++j) {
                        const int d = (e >> (2 * (M - 1 - j))) & 3;
                        row |= (d >> 1) << j; col |= (d & 1) << j;
                    }
                    const double2 r = am[u][row], c = am[u][col];
                    v[e] = make_double2(r.x * c.x + r.y * c.y, r.y * c.x - r.x * c.y);
                }
                pauli_axes<M == 0 ? 1 : M>(v);
#pragma unroll
                for (int e = 0; e < N_PM; ++e) R[(s * N_PM + e) * PT_PITCH + lane] = v[e].x;
            }
        }
    }
    __syncthreads();
    // 2. frame inverse along every axis; two axes per pass while at least two remain
    int j = 0;
    for (; j + 1 < p; j += 2) {
        const int inner = N_PM << (2 * (p - 2 - j));               // stride of digit j + 1
        const int n_item = (n_kappa >> 4) * PT_BB;
        for (int idx = tid; idx < n_item; idx += 256) {
            const int bl = idx & 31, vec = idx >> 5;
            const int lo = vec % inner, hi = vec / inner;
            double* q = R + (size_t)(hi * 16 * inner + lo) * PT_PITCH + bl;
            const int st = inner * PT_PITCH;
            double v[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) v[e] = q[e * st];
#pragma unroll
            for (int e = 0; e < 4; ++e) frame4(fr, v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);   // digit j + 1
#pragma unroll
            for (int e = 0; e < 4; ++e) frame4(fr, v[e], v[4 + e], v[8 + e], v[12 + e]);                   // digit j
#pragma unroll
            for (int e = 0; e < 16; ++e) q[e * st] = v[e];
        }
        __syncthreads();
    }
    if (j < p) {                                                   // odd p: the last prep axis (stride 4^M)
        const int n_item = (n_kappa >> 2) * PT_BB;
        for (int idx = tid; idx < n_item; idx += 256) {
            const int bl = idx & 31, vec = idx >> 5;
            const int lo = vec % N_PM, hi = vec / N_PM;
            double* q = R + (size_t)(hi * 4 * N_PM + lo) * PT_PITCH + bl;
            constexpr int st = N_PM * PT_PITCH;
            double v0 = q[0], v1 = q[st], v2 = q[2 * st], v3 = q[3 * st];
            frame4(fr, v0, v1, v2, v3);
            q[0] = v0; q[st] = v1; q[2 * st] = v2; q[3 * st] = v3;
        }
        __syncthreads();
    }
    // 3. T[set][b][kappa_out], lanes along kappa
    for (int idx = tid; idx < PT_BB * n_kappa; idx += 256) {
        const int kap = idx % n_kappa, bl = idx / n_kappa;
        if (b0 + bl >= a.n_b) break;
        long long ko = 0;
        for (int ax = 0; ax < p + M; ++ax) ko += (long long)((kap >> (2 * (p + M - 1 - ax))) & 3) * a.stride[ax];
        __stcs(a.T + (set * a.n_b + b0 + bl) * a.ld + ko, R[kap * PT_PITCH + bl]);
    }
}

// (b) measurement-only fragment (p == 0, m = M_HI + M_LO <= 4): per b the 2^m amplitudes give
//     4^m real Pauli coordinates.  Pass 1: a thread forms the matrix-unit products for the low
//     M_LO measured axes in registers and applies their (e00,e01,e10,e11)->(I,X,Y,Z) butterflies;
//     pass 2 (after one exchange through shared memory, pitch 4^M_LO + 1 complex) does the same for
//     the high axes and stores the real parts straight to T, 4^M_LO consecutive doubles per lane group.
template <int M_HI, int M_LO>
__global__ void __launch_bounds__(256) sv_meas_tensor_kernel(const TensorArgs a) {
    constexpr int M = M_HI + M_LO, N_MU = 1 << M, N_LO = 1 << (2 * M_LO), N_HI = 1 << (2 * M_HI);
    constexpr int PITCH = N_LO + 1;
    extern __shared__ double2 sh_raw[];
    double2* amp = sh_raw;                                         // [bb][N_MU]
    double2* C = sh_raw + (size_t)a.bb * N_MU;                     // [bb][N_HI][PITCH]
    const int tid = threadIdx.x, nthr = blockDim.x;
    const long long set = blockIdx.y, b0 = (long long)blockIdx.x * a.bb;
    for (int idx = tid; idx < a.bb * N_MU; idx += nthr) {
        const int bl = idx % a.bb, mu = idx / a.bb;
        const long long b = b0 + bl;
        double2 v = make_double2(0.0, 0.0);
        if (b < a.n_b) {
            // axis j <-> bit j of mu: the first n_basis axes pick the basis-input state (first
            // prep qubit = most significant digit of the state number), the rest a measured qubit
            uint64_t x = expand_runs((uint64_t)b, a.n_runs, a.run_pos, a.run_len);
            long long init = 0;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const uint64_t bit = (uint64_t)((mu >> j) & 1);
                if (j < a.n_basis) init |= (long long)bit << (a.n_basis - 1 - j);
                else x |= bit << a.meas_q[j - a.n_basis];
            }
            v = __ldcs(a.state + (((init * a.n_sets + set) << a.n) | (long long)(x ^ a.flip)));
        }
        amp[bl * N_MU + mu] = v;
    }
    __syncthreads();
    // axis j (most significant digit first) <-> measured qubit j <-> bit j of mu.
    // high axes = 0 .. M_HI-1, low axes = M_HI .. M-1.
    if (M_HI == 0) {
        for (int idx = tid; idx < a.bb; idx += nthr) {
            const double2* am = amp + idx * N_MU;
            double2 v[N_LO];
#pragma unroll
            for (int e = 0; e < N_LO; ++e) {
                int row = 0, col = 0;
#pragma unroll
                for (int j = 0; j < M_LO; ++j) {
                    const int d = (e >> (2 * (M_LO - 1 - j))) & 3;
                    row |= (d >> 1) << j; col |= (d & 1) << j;
                }
                const double2 r = am[row], c = am[col];
                v[e] = make_double2(r.x * c.x + r.y * c.y, r.y * c.x - r.x * c.y);
            }
            pauli_axes<M_LO>(v);
            const long long b = b0 + idx;
            if (b < a.n_b) {
#pragma unroll
                for (int e = 0; e < N_LO; ++e) {
                    long long ko = 0;
                    bool neg = false;                 // Tr(P^T N) on input axes: Y^T = -Y
#pragma unroll
                    for (int j = 0; j < M_LO; ++j) {
                        const int d = (e >> (2 * (M_LO - 1 - j))) & 3;
                        ko += (long long)d * a.stride[j];
                        neg ^= (j < a.n_basis) && d == 2;
                    }
                    a.T[(set * a.n_b + b) * a.ld + ko] = neg ? -v[e].x : v[e].x;
                }
            }
        }
        return;
    }
    // pass 1: item = (bl, hi digits); products over the low axes, low butterflies
    for (int idx = tid; idx < a.bb * N_HI; idx += nthr) {
        const int h = idx % N_HI, bl = idx / N_HI;
        int row_h = 0, col_h = 0;
#pragma unroll
        for (int j = 0; j < M_HI; ++j) {
            const int d = (h >> (2 * (M_HI - 1 - j))) & 3;
            row_h |= (d >> 1) << j; col_h |= (d & 1) << j;
        }
        const double2* am = amp + bl * N_MU;
        double2 v[N_LO];
#pragma unroll
        for (int e = 0; e < N_LO; ++e) {
            int row = row_h, col = col_h;
#pragma unroll
            for (int j = 0; j < M_LO; ++j) {
                const int d = (e >> (2 * (M_LO - 1 - j))) & 3;
                row |= (d >> 1) << (M_HI + j); col |= (d & 1) << (M_HI + j);
            }
            const double2 r = am[row], c = am[col];
            v[e] = make_double2(r.x * c.x + r.y * c.y, r.y * c.x - r.x * c.y);
        }
        pauli_axes<M_LO>(v);
        // exchange layout C[bl][lo digits][hi digits] with pitch N_HI + 1 would serve pass 2's
        // reads; we use [bl][e][h] so that this write has lanes (h fastest) on consecutive words
        double2* dst = C + (size_t)bl * N_LO * (N_HI + 1) + h;
#pragma unroll
        for (int e = 0; e < N_LO; ++e) dst[e * (N_HI + 1)] = v[e];
    }
    __syncthreads();
    // pass 2: item = (bl, lo digits); high butterflies, real parts to T
    for (int idx = tid; idx < a.bb * N_LO; idx += nthr) {
        const int e = idx % N_LO, bl = idx / N_LO;
        const double2* src = C + (size_t)bl * N_LO * (N_HI + 1) + e * (N_HI + 1);
        double2 v[N_HI];
#pragma unroll
        for (int h = 0; h < N_HI; ++h) v[h] = src[h];
        pauli_axes<M_HI>(v);
        const long long b = b0 + bl;
        if (b >= a.n_b) continue;
        long long ko_lo = 0;
        bool neg_lo = false;
#pragma unroll
        for (int j = 0; j < M_LO; ++j) {
            const int d = (e >> (2 * (M_LO - 1 - j))) & 3;
            ko_lo += (long long)d * a.stride[M_HI + j];
            neg_lo ^= (M_HI + j < a.n_basis) && d == 2;
        }
        double* out = a.T + (set * a.n_b + b) * a.ld + ko_lo;
#pragma unroll
        for (int h = 0; h < N_HI; ++h) {
            long long ko = 0;
            bool neg = neg_lo;
#pragma unroll
            for (int j = 0; j < M_HI; ++j) {
                const int d = (h >> (2 * (M_HI - 1 - j))) & 3;
                ko += (long long)d * a.stride[j];
                neg ^= (j < a.n_basis) && d == 2;
            }
            __stcs(out + ko, neg ? -v[h].x : v[h].x);
        }
    }
    (void)PITCH;
}

// ---- host: compiler ---------------------------------------------------------------------------
struct LGate {                 // lowered gate (after the X frame)
    int src, kind, target, conj;
    uint64_t cmask, cval;
    bool diag, tgt_zero;
};

bool kind_is_diag(const qdc_gate& g) {
    switch (g.kind) {
        case QDC_G_ID: case QDC_G_Z: case QDC_G_S: case QDC_G_SDG: case QDC_G_T: case QDC_G_TDG:
        case QDC_G_RZ: case QDC_G_U1: return true;
        case QDC_G_U2: return g.m[2] == 0.0 && g.m[3] == 0.0 && g.m[4] == 0.0 && g.m[5] == 0.0;
        default: return false;
    }
}

}  // namespace

struct qdc_sv_program {
    int n, t, c;
    uint64_t flip;
    std::vector<SvOp> ops;
    std::vector<MatDesc> mats;
    std::vector<Group> groups;
    std::vector<int> op_src, op_conj;        // for qdc_sv_dump
    std::vector<uint64_t> op_cmask, op_cval;
    SvOp* d_ops = nullptr;
    MatDesc* d_mats = nullptr;
};

extern "C" void qdc_sv_free(qdc_sv_program* prog) {
    if (!prog) return;
    if (prog->d_ops) cudaFree(prog->d_ops);
    if (prog->d_mats) cudaFree(prog->d_mats);
    delete prog;
}

extern "C" int qdc_sv_compile(int n_qubits, const qdc_gate* gates, int n_gates, uint64_t zero_mask,
                              int tile_bits, int coalesce_bits, int flags, qdc_sv_program** out) {
    QDC_REQUIRE(out != nullptr, "out");
    *out = nullptr;
    QDC_REQUIRE(n_qubits >= 1 && n_qubits <= 40, "n_qubits in [1, 40]");
    QDC_REQUIRE(n_gates >= 0 && (gates != nullptr || n_gates == 0), "gates");
    const int n = n_qubits;
    int t = tile_bits <= 0 ? 12 : tile_bits;
    if (t > n) t = n;
    QDC_REQUIRE(t >= 1 && t <= 13, "tile_bits in [1, 13] (2^13 * 16 B = 128 KB of shared memory)");
    int c = coalesce_bits < 0 ? 5 : coalesce_bits;          // 512-byte runs for the bulk copies
    if (c > t) c = t;
    if (t < n && c >= t) c = t - 1;      // leave room for at least one non-local target
    const uint64_t all = n == 64 ? ~0ull : ((1ull << n) - 1ull);

    // ---- pass 1: X frame ------------------------------------------------------------------
    std::vector<LGate> lg;
    uint64_t flip = 0;
    for (int gi = 0; gi < n_gates; ++gi) {
        const qdc_gate& g = gates[gi];
        QDC_REQUIRE(g.kind >= QDC_G_ID && g.kind <= QDC_G_U2, "gate kind");
        QDC_REQUIRE(g.target >= 0 && g.target < n, "gate target");
        QDC_REQUIRE((g.ctrl_mask & ~all) == 0 && (g.ctrl_val & ~g.ctrl_mask) == 0, "gate controls");
        QDC_REQUIRE(!((g.ctrl_mask >> g.target) & 1ull), "target among controls");
        if (g.kind == QDC_G_ID) continue;
        if (g.kind == QDC_G_X && g.ctrl_mask == 0) { flip ^= 1ull << g.target; continue; }
        LGate l;
        l.src = gi; l.kind = g.kind; l.target = g.target;
        l.cmask = g.ctrl_mask; l.cval = g.ctrl_val ^ (flip & g.ctrl_mask);
        l.conj = (int)((flip >> g.target) & 1ull);
        l.diag = kind_is_diag(g);
        l.tgt_zero = false;
        lg.push_back(l);
    }

    // ---- pass 2: ancilla compute / act / uncompute -> one controlled gate -----------------
    std::vector<LGate> fg;
    {
        uint64_t zero = zero_mask & all;          // stored-frame |0> qubits
        for (size_t i = 0; i < lg.size();) {
            const LGate& a = lg[i];
            bool fused = false;
            if ((flags & QDC_SV_FUSE_ANCILLA) && a.kind == QDC_G_X && a.cmask != 0 && !a.conj &&
                ((zero >> a.target) & 1ull) && i + 2 < lg.size()) {
                const LGate& b = lg[i + 1];
                const LGate& cgate = lg[i + 2];
                const uint64_t abit = 1ull << a.target;
                if (b.cmask == abit && b.cval == abit && !((a.cmask >> b.target) & 1ull) &&
                    b.target != a.target && cgate.kind == QDC_G_X && cgate.target == a.target &&
                    cgate.cmask == a.cmask && cgate.cval == a.cval && !cgate.conj) {
                    LGate f = b;
                    f.cmask = a.cmask; f.cval = a.cval;
                    fg.push_back(f);
                    if (!f.diag) zero &= ~(1ull << f.target);
                    i += 3;
                    fused = true;
                }
            }
            // cx(a -> b) . D(b) . cx(a -> b) with D diagonal (the ZZ rotations of QAOA+'s phase separator,
            // ansatz/qaoa_plus.py:20-24) is diagonal as a whole: D on b where the cx is inactive, X D X where it is
            // active.  Two controlled diagonal gates instead of two non-diagonal ones + one: diagonal gates need no
            // tile locality, so the phase separator stops dictating the grouping.
            if (!fused && a.kind == QDC_G_X && __builtin_popcountll(a.cmask) == 1 && i + 2 < lg.size()) {
                const LGate& d = lg[i + 1];
                const LGate& c2 = lg[i + 2];
                if (d.diag && d.cmask == 0 && d.target == a.target && c2.kind == QDC_G_X && c2.target == a.target &&
                    c2.cmask == a.cmask && c2.cval == a.cval) {
                    LGate g0 = d, g1 = d;
                    g0.cmask = a.cmask; g0.cval = a.cval ^ a.cmask;        // control at the value that leaves the cx inactive
                    g1.cmask = a.cmask; g1.cval = a.cval; g1.conj ^= 1;    // active: X D X
                    fg.push_back(g0);
                    fg.push_back(g1);
                    i += 3;
                    fused = true;
                }
            }
            if (!fused) {
                fg.push_back(a);
                if (!a.diag) zero &= ~(1ull << a.target);
                ++i;
            }
        }
    }

    // ---- pass 3: list scheduling into tile-local groups --------------------------------------
    qdc_sv_program* prog = new (std::nothrow) qdc_sv_program();
    QDC_REQUIRE(prog != nullptr, "out of host memory");
    prog->n = n; prog->t = t; prog->c = c; prog->flip = flip;
    const uint64_t low = c >= 64 ? ~0ull : ((1ull << c) - 1ull);
    std::vector<char> done(fg.size(), 0);
    size_t n_done = 0;
    std::vector<std::vector<int>> group_gates;
    std::vector<uint64_t> group_L;
    while (n_done < fg.size()) {
        uint64_t L = low & all, skipped_uses = 0, skipped_nd_targets = 0;
        std::vector<int> mine;
        for (size_t i = 0; i < fg.size(); ++i) {
            if (done[i]) continue;
            const LGate& g = fg[i];
            const uint64_t tb = 1ull << g.target, uses = tb | g.cmask, nd = g.diag ? 0ull : tb;
            bool take = !(nd & skipped_uses) && !(uses & skipped_nd_targets);
            if (take && !g.diag && !(L & tb)) {
                if (__builtin_popcountll(L) < t) L |= tb; else take = false;
            }
            if (take) { mine.push_back((int)i); done[i] = 1; ++n_done; }
            else { skipped_uses |= uses; skipped_nd_targets |= nd; }
        }
        group_gates.push_back(mine);
        group_L.push_back(L);
    }
    if (group_gates.empty()) { group_gates.push_back({}); group_L.push_back(low & all); }

    for (size_t gidx = 0; gidx < group_gates.size(); ++gidx) {
        uint64_t L = group_L[gidx];
        for (int q = 0; q < n && __builtin_popcountll(L) < t; ++q) L |= 1ull << q;   // pad to t qubits
        Group gr;
        gr.first_op = (int)prog->ops.size();
        gr.local_mask = L;
        gr.n_runs = 0;
        for (int q = 0; q < n;) {
            if ((L >> q) & 1ull) { ++q; continue; }
            int e = q;
            while (e < n && !((L >> e) & 1ull)) ++e;
            if (gr.n_runs >= MAX_RUNS) { qdc_sv_free(prog); QDC_REQUIRE(false, "too many tile runs"); }
            gr.run_pos[gr.n_runs] = (uint8_t)q; gr.run_len[gr.n_runs] = (uint8_t)(e - q);
            ++gr.n_runs;
            q = e;
        }
        auto local_bit = [&](int q) { return __builtin_popcountll(L & ((1ull << q) - 1ull)); };
        // Diagonal gates commute with each other: inside every block of consecutive diagonal gates those that can
        // share a table pass (they depend on at most one tile-local bit: uncontrolled, or one control with control
        // and target not both local) are moved behind the others, so that they form long runs (QAOA+'s phase
        // separator: 180 diagonal gates in one group become ~15 separate passes + 6 table passes instead of 135).
        auto tabled = [&](const LGate& g) {
            if (!g.diag) return false;
            const int nc = __builtin_popcountll(g.cmask);
            if (nc == 0) return true;
            if (nc > 1) return false;
            return !(((L >> g.target) & 1ull) && (L & g.cmask));
        };
        {
            std::vector<int>& gg = group_gates[gidx];
            for (size_t b = 0; b < gg.size();) {
                if (!fg[gg[b]].diag) { ++b; continue; }
                size_t e = b;
                while (e < gg.size() && fg[gg[e]].diag) ++e;
                std::stable_partition(gg.begin() + b, gg.begin() + e, [&](int gi) { return !tabled(fg[gi]); });
                b = e;
            }
        }
        for (int gi : group_gates[gidx]) {
            const LGate& g = fg[gi];
            SvOp op;
            memset(&op, 0, sizeof(op));
            op.kind = g.diag ? 1u : 0u;
            op.gtarget = g.target;
            op.tbit = ((L >> g.target) & 1ull) ? local_bit(g.target) : -1;
            op.mat = (int)prog->mats.size();
            std::vector<int> fixed;
            if (!g.diag) fixed.push_back(op.tbit);
            for (int q = 0; q < n; ++q) {
                if (!((g.cmask >> q) & 1ull)) continue;
                if ((L >> q) & 1ull) {
                    const int lb = local_bit(q);
                    op.lmask |= 1u << lb;
                    if ((g.cval >> q) & 1ull) op.lval |= 1u << lb;
                    fixed.push_back(lb);
                } else {
                    op.gmask |= 1ull << q;
                    op.gval |= g.cval & (1ull << q);
                }
            }
            std::sort(fixed.begin(), fixed.end());
            if ((int)fixed.size() > MAX_FIXED) { qdc_sv_free(prog); QDC_REQUIRE(false, "too many local controls"); }
            op.n_fixed = (uint8_t)fixed.size();
            for (size_t k = 0; k < fixed.size(); ++k) {
                op.fixed[k] = (uint8_t)fixed[k];
                op.fmask |= 1u << fixed[k];
            }
            prog->ops.push_back(op);
            MatDesc md;
            memset(&md, 0, sizeof(md));
            const qdc_gate& src = gates[g.src];
            md.kind = src.kind; md.slot = src.param_slot; md.conj_x = g.kind == QDC_G_X ? 0 : g.conj;
            md.scale = src.scale; md.offset = src.offset;
            for (int k = 0; k < 8; ++k) md.m[k] = src.m[k];
            prog->mats.push_back(md);
            prog->op_src.push_back(g.src);
            prog->op_conj.push_back(md.conj_x);
            prog->op_cmask.push_back(g.cmask);
            prog->op_cval.push_back(g.cval);
        }
        gr.n_ops = (int)prog->ops.size() - gr.first_op;
        // runs of uncontrolled diagonal ops (they commute with each other: one joint pass)
        for (int i = gr.first_op; i < gr.first_op + gr.n_ops;) {
            SvOp& o = prog->ops[i];
            auto in_run = [&](const SvOp& x) {
                if (x.kind != 1) return false;
                const int nc = __builtin_popcount(x.lmask) + __builtin_popcountll(x.gmask);
                return nc == 0 || (nc == 1 && !(x.lmask != 0 && x.tbit >= 0));
            };
            if (in_run(o)) {
                int e = i;
                // a run must not straddle a staging chunk of the kernel (OPS_CHUNK ops from first_op)
                const int chunk_end = gr.first_op + ((i - gr.first_op) / OPS_CHUNK + 1) * OPS_CHUNK;
                while (e < gr.first_op + gr.n_ops && e < chunk_end && e - i < MAX_DIAG_RUN && in_run(prog->ops[e])) ++e;
                o.diag_run = (uint8_t)(e - i);
                i = e;
            } else {
                ++i;
            }
        }
        prog->groups.push_back(gr);
    }

    if (!prog->ops.empty()) {
        cudaError_t e1 = cudaMalloc(&prog->d_ops, prog->ops.size() * sizeof(SvOp));
        cudaError_t e2 = cudaMalloc(&prog->d_mats, prog->mats.size() * sizeof(MatDesc));
        if (e1 == cudaSuccess && e2 == cudaSuccess) {
            e1 = cudaMemcpy(prog->d_ops, prog->ops.data(), prog->ops.size() * sizeof(SvOp), cudaMemcpyHostToDevice);
            e2 = cudaMemcpy(prog->d_mats, prog->mats.data(), prog->mats.size() * sizeof(MatDesc), cudaMemcpyHostToDevice);
        }
        if (e1 != cudaSuccess || e2 != cudaSuccess) {
            // no device (e.g. host-only schedule tests): keep the host schedule, run() will refuse
            cudaGetLastError();
            if (prog->d_ops) { cudaFree(prog->d_ops); prog->d_ops = nullptr; }
            if (prog->d_mats) { cudaFree(prog->d_mats); prog->d_mats = nullptr; }
            cudaGetLastError();
        }
    }
    *out = prog;
    return QDC_OK;
}

extern "C" int qdc_sv_info(const qdc_sv_program* prog, int* n_qubits, int* n_groups, int* n_ops,
                           int* n_matrices, int* tile_bits, uint64_t* flip_mask) {
    QDC_REQUIRE(prog != nullptr, "program");
    if (n_qubits) *n_qubits = prog->n;
    if (n_groups) *n_groups = (int)prog->groups.size();
    if (n_ops) *n_ops = (int)prog->ops.size();
    if (n_matrices) *n_matrices = (int)prog->mats.size();
    if (tile_bits) *tile_bits = prog->t;
    if (flip_mask) *flip_mask = prog->flip;
    return QDC_OK;
}

extern "C" int qdc_sv_dump(const qdc_sv_program* prog, int32_t* group_first_op,
                           uint64_t* group_local_mask, int32_t* op_kind, int32_t* op_target,
                           uint64_t* op_ctrl_mask, uint64_t* op_ctrl_val, int32_t* op_src_gate,
                           int32_t* op_conj_x) {
    QDC_REQUIRE(prog != nullptr, "program");
    const size_t ng = prog->groups.size();
    for (size_t g = 0; g < ng; ++g) {
        if (group_first_op) group_first_op[g] = prog->groups[g].first_op;
        if (group_local_mask) group_local_mask[g] = prog->groups[g].local_mask;
    }
    if (group_first_op) group_first_op[ng] = (int32_t)prog->ops.size();
    for (size_t i = 0; i < prog->ops.size(); ++i) {
        if (op_kind) op_kind[i] = (int32_t)prog->ops[i].kind;
        if (op_target) op_target[i] = prog->ops[i].gtarget;
        if (op_ctrl_mask) op_ctrl_mask[i] = prog->op_cmask[i];
        if (op_ctrl_val) op_ctrl_val[i] = prog->op_cval[i];
        if (op_src_gate) op_src_gate[i] = prog->op_src[i];
        if (op_conj_x) op_conj_x[i] = prog->op_conj[i];
    }
    return QDC_OK;
}

struct ScoreReq {                 // fused scoring of the final state (qdc_sv_run_score); tab == nullptr: none
    uint64_t* tab; const uint64_t* adj; double* partials; double* sums; uint64_t nodes; int keep_state;
};

static int sv_run_impl(const qdc_sv_program* prog, void* d_state, int64_t init_lo, int64_t n_init,
                       int64_t n_sets, const double* d_params, int n_params,
                       const int32_t* h_prep_qubits, int n_prep, const double* h_code_table,
                       int n_codes, double* d_matrices, const ScoreReq& score, void* stream) {
    QDC_REQUIRE(prog != nullptr && d_state != nullptr, "program / state");
    QDC_REQUIRE(n_init >= 1 && n_sets >= 1 && init_lo >= 0, "batch");
    QDC_REQUIRE(n_prep >= 0 && n_prep <= MAX_PREP, "n_prep");
    QDC_REQUIRE(n_prep == 0 || (h_prep_qubits && h_code_table && n_codes >= 1 && n_codes <= MAX_CODES),
                "preparation table");
    QDC_REQUIRE(prog->ops.empty() || (prog->d_ops && prog->d_mats), "program has no device copy (compiled without a GPU)");
    QDC_REQUIRE(prog->mats.empty() || d_matrices != nullptr, "matrix workspace");
    cudaStream_t s = (cudaStream_t)stream;
    const int n_mats = (int)prog->mats.size();
    if (n_mats > 0) {
        for (const MatDesc& md : prog->mats)
            QDC_REQUIRE(md.slot < n_params && (md.slot < 0 || d_params != nullptr), "parameter slot out of range");
        const long long total = n_sets * n_mats;
        sv_matrix_kernel<<<(unsigned)((total + 127) / 128), 128, 0, s>>>(prog->d_mats, n_mats, n_sets,
                                                                         d_params, n_params, d_matrices);
        QDC_CHECK_LAUNCH();
    }
    GroupArgs a;
    memset(&a, 0, sizeof(a));
    a.state = (double2*)d_state;
    a.ops = prog->d_ops;
    a.mats = d_matrices;
    a.n = prog->n; a.t = prog->t;
    a.n_tiles = 1ll << (prog->n - prog->t);
    a.init_lo = init_lo;
    a.n_sets = (int)n_sets; a.n_mats = n_mats;
    a.n_prep = n_prep; a.n_codes = n_prep ? n_codes : 1;
    for (int j = 0; j < n_prep; ++j) {
        QDC_REQUIRE(h_prep_qubits[j] >= 0 && h_prep_qubits[j] < prog->n, "prep qubit");
        QDC_REQUIRE(!((a.prep_mask >> h_prep_qubits[j]) & 1ull), "duplicate prep qubit");
        a.prep_q[j] = (uint8_t)h_prep_qubits[j];
        a.prep_mask |= 1ull << h_prep_qubits[j];
    }
    if (n_prep)
        for (int k = 0; k < n_codes; ++k)
            for (int e = 0; e < 4; ++e) a.code[k][e] = h_code_table[4 * k + e];
    const long long n_batch = n_init * n_sets;
    const long long grid = n_batch * a.n_tiles;
    QDC_REQUIRE(grid < (1ll << 31), "batch * tiles exceeds one grid: split the batch");
    const size_t smem = (size_t)16 << prog->t;
    QDC_CHECK_CUDA(cudaFuncSetAttribute(sv_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // measured on B200 (profiles/r1_time_statevec_threads.log): several small CTAs per SM overlap
    // their load / gate / store phases better than one or two large ones
    const int max_threads = prog->t >= 13 ? 512 : prog->t == 12 ? 256 : 128;
    int threads = 1 << (prog->t > 1 ? prog->t - 1 : 0);
    if (threads > max_threads) threads = max_threads;
    if (threads < 32) threads = 32;
    if (const char* env = getenv("QDC_SV_THREADS")) {       // tuning knob (power of two, 32..512)
        const int v = atoi(env);
        if (v >= 32 && v <= 512 && (v & (v - 1)) == 0) threads = v;
    }
    QDC_CHECK_CUDA(cudaFuncSetAttribute(sv_group_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                        cudaSharedmemCarveoutMaxShared));
    bool use_bulk = true;
    if (const char* env = getenv("QDC_SV_BULK")) use_bulk = atoi(env) != 0;   // tuning knob: 0 = per-element staging
    for (size_t g = 0; g < prog->groups.size(); ++g) {
        const Group& gr = prog->groups[g];
        a.op_lo = gr.first_op; a.op_hi = gr.first_op + gr.n_ops;
        a.first = g == 0;
        a.n_runs = gr.n_runs;
        a.hole_mask = ~gr.local_mask;                      // bits above n never receive a carry-out that matters
        for (int r = 0; r < gr.n_runs; ++r) { a.run_pos[r] = gr.run_pos[r]; a.run_len[r] = gr.run_len[r]; }
        {
            const int p0 = gr.n_runs ? (gr.run_pos[0] < prog->t ? gr.run_pos[0] : prog->t) : prog->t;
            a.bulk = use_bulk && p0 >= 2;                  // >= 64-byte runs; shorter ones go element by element
        }
        a.score = 0; a.store = 1;
        if (score.tab && g + 1 == prog->groups.size()) {
            // the last group also scores its tiles (and may skip writing the state back)
            a.score = 1; a.store = score.keep_state;
            a.score_tab = score.tab; a.score_adj = score.adj; a.score_partials = score.partials;
            a.score_flip = prog->flip; a.score_nodes = score.nodes;
            sv_score_table_kernel<<<((1u << prog->t) + 255) / 256, 256, 0, s>>>(score.tab, prog->t, gr.n_runs, a);
            QDC_CHECK_LAUNCH();
        }
        sv_group_kernel<<<(unsigned)grid, threads, smem, s>>>(a);
        QDC_CHECK_LAUNCH();
    }
    if (score.tab) {
        sv_score_finish_kernel<<<(unsigned)n_batch, 1024, 0, s>>>(score.partials, a.n_tiles, score.sums);
        QDC_CHECK_LAUNCH();
    }
    return QDC_OK;
}

extern "C" int qdc_sv_run(const qdc_sv_program* prog, void* d_state, int64_t init_lo, int64_t n_init,
                          int64_t n_sets, const double* d_params, int n_params,
                          const int32_t* h_prep_qubits, int n_prep, const double* h_code_table,
                          int n_codes, double* d_matrices, void* stream) {
    const ScoreReq none = {nullptr, nullptr, nullptr, nullptr, 0ull, 1};
    return sv_run_impl(prog, d_state, init_lo, n_init, n_sets, d_params, n_params, h_prep_qubits, n_prep,
                       h_code_table, n_codes, d_matrices, none, stream);
}

extern "C" int64_t qdc_sv_score_workspace_bytes(const qdc_sv_program* prog, int64_t n_sets) {
    if (!prog || n_sets < 1) return -1;
    const long long n_tiles = 1ll << (prog->n - prog->t);
    return (int64_t)(8ll << prog->t) + 3 * 8 * n_sets * n_tiles;
}

extern "C" int qdc_sv_run_score(const qdc_sv_program* prog, void* d_state, int64_t n_sets, const double* d_params,
                                int n_params, double* d_matrices, const uint64_t* d_adj, uint64_t node_mask,
                                int keep_state, void* d_workspace, double* d_sums, void* stream) {
    QDC_REQUIRE(prog != nullptr && d_adj && d_workspace && d_sums, "null buffer");
    QDC_REQUIRE((((uintptr_t)d_workspace) & 7) == 0, "workspace alignment");
    QDC_REQUIRE(prog->n <= 48 && (prog->n >= 64 || (node_mask >> prog->n) == 0), "node_mask");
    QDC_REQUIRE(!prog->groups.empty(), "a program without gate groups has no final group to score in");
    ScoreReq score;
    score.tab = (uint64_t*)d_workspace;
    score.partials = (double*)((uint8_t*)d_workspace + (8ll << prog->t));
    score.adj = d_adj; score.sums = d_sums; score.nodes = node_mask; score.keep_state = keep_state ? 1 : 0;
    return sv_run_impl(prog, d_state, 0, 1, n_sets, d_params, n_params, nullptr, 0, nullptr, 1, d_matrices, score, stream);
}

extern "C" int qdc_sv_probs(const void* d_state, int n_qubits, int64_t n_batch, uint64_t flip_mask,
                            double* d_probs, void* stream) {
    QDC_REQUIRE(d_state && d_probs, "null buffer");
    QDC_REQUIRE(n_qubits >= 1 && n_qubits <= 40 && n_batch >= 1, "shape");
    QDC_REQUIRE((flip_mask >> n_qubits) == 0, "flip_mask");
    const long long total = n_batch << n_qubits;
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)qdc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    sv_probs_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const double2*)d_state, n_qubits,
                                                                       total, flip_mask, d_probs);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}

extern "C" int qdc_sv_pauli_tensor(const void* d_state, int n_qubits, int64_t n_sets,
                                   uint64_t flip_mask, int n_prep, int n_codes, const double* h_frame,
                                   const int32_t* h_meas_qubits, int n_meas,
                                   const int64_t* h_axis_stride, double* d_T, int64_t ld, void* stream) {
    QDC_REQUIRE(d_state && d_T, "null buffer");
    QDC_REQUIRE(n_qubits >= 1 && n_qubits <= 40 && n_sets >= 1, "shape");
    QDC_REQUIRE(n_prep >= 0 && n_meas >= 0 && n_meas <= 6 && n_prep + n_meas <= 6, "at most 6 cut axes per fragment");
    QDC_REQUIRE(n_meas <= n_qubits, "n_meas");
    QDC_REQUIRE(n_prep == 0 || n_codes == 2 || (n_codes == 4 && h_frame != nullptr),
                "n_codes 4 (preparation states + frame) or 2 (basis inputs)");
    QDC_REQUIRE(n_prep == 0 || n_codes != 2 || n_prep + n_meas <= 4, "basis-input mode handles at most 4 cut axes");
    QDC_REQUIRE(n_meas == 0 || h_meas_qubits != nullptr, "meas qubits");
    QDC_REQUIRE((flip_mask >> n_qubits) == 0, "flip_mask");
    TensorArgs a;
    memset(&a, 0, sizeof(a));
    a.state = (const double2*)d_state; a.T = d_T;
    a.n_sets = n_sets; a.n = n_qubits; a.n_prep = n_prep; a.n_meas = n_meas; a.flip = flip_mask;
    a.n_b = 1ll << (n_qubits - n_meas);
    const int n_axes = n_prep + n_meas;
    const long long n_kappa = 1ll << (2 * n_axes);
    QDC_REQUIRE(ld >= n_kappa, "ld");
    uint64_t mm = 0;
    for (int j = 0; j < n_meas; ++j) {
        QDC_REQUIRE(h_meas_qubits[j] >= 0 && h_meas_qubits[j] < n_qubits && !((mm >> h_meas_qubits[j]) & 1ull), "meas qubit");
        mm |= 1ull << h_meas_qubits[j];
        a.meas_q[j] = (uint8_t)h_meas_qubits[j];
    }
    for (int q = 0; q < n_qubits;) {              // runs of measured positions = holes of the b deposit
        if (!((mm >> q) & 1ull)) { ++q; continue; }
        int e = q;
        while (e < n_qubits && ((mm >> e) & 1ull)) ++e;
        a.run_pos[a.n_runs] = (uint8_t)q; a.run_len[a.n_runs] = (uint8_t)(e - q);
        ++a.n_runs;
        q = e;
    }
    long long nat = n_kappa;
    for (int ax = 0; ax < n_axes; ++ax) {
        nat /= 4;
        a.stride[ax] = h_axis_stride ? h_axis_stride[ax] : nat;
        QDC_REQUIRE(a.stride[ax] >= 0 && 3 * a.stride[ax] < ld, "axis stride");
    }
    if (n_prep && n_codes == 4) for (int k = 0; k < 16; ++k) a.frame[k] = h_frame[k];
    const bool basis = n_prep > 0 && n_codes == 2;
    a.n_basis = basis ? n_prep : 0;
    a.ld = ld;
    if (!basis && n_prep >= 1 && n_meas <= 2 && n_prep + n_meas <= 4 && !getenv("QDC_SV_TENSOR_GENERIC")) {
        const size_t smem = (size_t)n_kappa * PT_PITCH * sizeof(double);
        const long long gx = (a.n_b + PT_BB - 1) / PT_BB;
        QDC_REQUIRE(gx < (1ll << 31) && n_sets <= 65535, "grid");
        const dim3 grid((unsigned)gx, (unsigned)n_sets);
        cudaStream_t st = (cudaStream_t)stream;
#define QDC_PREP_CASE(MM)                                                                                   \
        {                                                                                                   \
            QDC_CHECK_CUDA(cudaFuncSetAttribute(sv_prep_tensor_kernel<MM>,                                  \
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
            sv_prep_tensor_kernel<MM><<<grid, 256, smem, st>>>(a);                                          \
        }
        if (n_meas == 0) QDC_PREP_CASE(0)
        else if (n_meas == 1) QDC_PREP_CASE(1)
        else QDC_PREP_CASE(2)
#undef QDC_PREP_CASE
        QDC_CHECK_LAUNCH();
        return QDC_OK;
    }
    if (basis || (n_prep == 0 && n_meas >= 1 && n_meas <= 4 && !getenv("QDC_SV_TENSOR_GENERIC"))) {
        const int n_ax = basis ? n_prep + n_meas : n_meas;        // amplitude-index axes
        const int m_lo = n_ax < 2 ? n_ax : 2, m_hi = n_ax - m_lo;
        const int n_lo = 1 << (2 * m_lo), n_hi = 1 << (2 * m_hi);
        // b values per CTA.  With 4 axes the exchange buffer is 4.3 KB per b: 8 of them and 128 threads put six
        // CTAs on an SM instead of three, in different phases (gather / products / exchange / stores)
        int bb = n_ax <= 2 ? 256 : (n_ax == 3 ? 64 : 8);
        if (const char* env = getenv("QDC_SV_TENSOR_BB")) { const int v = atoi(env); if (v >= 1 && v <= 256) bb = v; }
        if (bb > a.n_b) bb = (int)a.n_b;
        a.bb = bb;
        int mt_threads = bb * (m_hi ? (n_hi > n_lo ? n_hi : n_lo) : 1);
        mt_threads = mt_threads < 64 ? 64 : (mt_threads > 256 ? 256 : (mt_threads + 31) / 32 * 32);
        const size_t smem = (size_t)bb * ((size_t)(1 << n_ax) + (m_hi ? (size_t)n_lo * (n_hi + 1) : 0)) * 16;
        const long long gx = (a.n_b + bb - 1) / bb;
        QDC_REQUIRE(gx < (1ll << 31) && n_sets <= 65535, "grid");
        const dim3 grid((unsigned)gx, (unsigned)n_sets);
        cudaStream_t st = (cudaStream_t)stream;
        const int n_meas_axes = n_ax;
#define QDC_MEAS_CASE(HI, LO)                                                                               \
        {                                                                                                   \
            QDC_CHECK_CUDA(cudaFuncSetAttribute(sv_meas_tensor_kernel<HI, LO>,                              \
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
            sv_meas_tensor_kernel<HI, LO><<<grid, mt_threads, smem, st>>>(a);                               \
        }
        if (n_meas_axes == 1) QDC_MEAS_CASE(0, 1)
        else if (n_meas_axes == 2) QDC_MEAS_CASE(0, 2)
        else if (n_meas_axes == 3) QDC_MEAS_CASE(1, 2)
        else QDC_MEAS_CASE(2, 2)
#undef QDC_MEAS_CASE
        QDC_CHECK_LAUNCH();
        return QDC_OK;
    }
    int bb = (int)(2048 / n_kappa);
    if (bb < 1) bb = 1;
    if (bb > a.n_b) bb = (int)a.n_b;
    a.bb = bb;
    const size_t n_s = (size_t)1 << (2 * n_prep), n_mu = (size_t)1 << n_meas;
    const size_t smem = (size_t)bb * n_s * n_mu * 16 + (size_t)bb * n_kappa * (n_meas >= 3 ? 16 : 8);
    QDC_REQUIRE(smem <= 200 * 1024, "shared memory");
    QDC_CHECK_CUDA(cudaFuncSetAttribute(sv_pauli_tensor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long gx = (a.n_b + bb - 1) / bb;
    QDC_REQUIRE(gx < (1ll << 31) && n_sets <= 65535, "grid");
    a.ld = ld;
    dim3 grid((unsigned)gx, (unsigned)n_sets);
    sv_pauli_tensor_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(a);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}
