// Finite-shot tomography on the device: draw `shots` measurement outcomes from each of a batch of
// statevectors and file them straight into the organised counts table.
//
// What it replaces: the reference runs every fragment variant on qasm_simulator in jobs of at most
// 8192 shots (qsplit_mlrecon_methods.py:20-42), then regroups the per-variant count dicts by the
// bits of the non-cut qubits into {final bits: {(prep label, measured state): count}}
// (organize_tomography_data, :169-213).  Here the variants of one measurement-basis choice are a
// batch of statevectors (all 4^p preparations x all parameter sets) and one call samples all of
// them: no probability vector is stored, no per-variant histogram exists, the table is the output.
//
// Method (inverse-CDF by a three-level walk; every level is a coalesced read):
//   mass pass   : |amp|^2 summed per tile of 256 amplitudes and per super-tile of 256 tiles
//                 (one CTA per super-tile, one streaming read of the states: 16 B / amplitude);
//   super scan  : inclusive running sums of the super-tile masses per state (<= 2^(n-16) values);
//   draw        : one warp per shot.  u = Philox4x32-10(seed; shot, init, set, stream) in [0, 1),
//                 target = u * (mass of the state); binary search in the super scan, then a
//                 32-wide scan walk over the 256 tile masses of that super-tile, then the same walk
//                 over the 256 amplitudes of that tile; outcome -> (final bits b, column) -> one
//                 atomicAdd of 1.0 (sums of small integers in FP64 are exact, so the table does not
//                 depend on the order of the atomics: same seed, same table).
// Rounding: the walk re-adds masses in another order than the mass pass did; a target that falls off
// the end of a level by that rounding lands on the last non-zero entry of the level.
#include "qdc_common.cuh"

namespace {

constexpr int TILE = 256;            // amplitudes per tile
constexpr int SUPER = 256;           // tiles per super-tile

struct SampleShape {
    int n;                           // qubits
    long long n_amp;                 // 2^n
    int tile;                        // min(TILE, n_amp)
    long long n_tiles;               // n_amp / tile
    int tiles_per_super;             // min(SUPER, n_tiles)
    long long n_super;               // n_tiles / tiles_per_super
};

__host__ __device__ inline SampleShape shape_of(int n) {
    SampleShape s;
    s.n = n;
    s.n_amp = 1ll << n;
    s.tile = s.n_amp < TILE ? (int)s.n_amp : TILE;
    s.n_tiles = s.n_amp / s.tile;
    s.tiles_per_super = s.n_tiles < SUPER ? (int)s.n_tiles : SUPER;
    s.n_super = s.n_tiles / s.tiles_per_super;
    return s;
}

// ---- mass pass --------------------------------------------------------------------------------
// grid = (n_super, n_states); 256 threads = 8 warps, warp w sums tiles w, w + 8, ...
__global__ void __launch_bounds__(256) sample_mass_kernel(const double2* __restrict__ states, SampleShape sh,
                                                          double* __restrict__ tile_mass,
                                                          double* __restrict__ super_mass) {
    __shared__ double warp_part[8];
    const long long state = blockIdx.y, sup = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const double2* base = states + state * sh.n_amp + sup * sh.tiles_per_super * (long long)sh.tile;
    double* tm = tile_mass + state * sh.n_tiles + sup * sh.tiles_per_super;
    double mine = 0.0;
    for (int t = warp; t < sh.tiles_per_super; t += 8) {
        const double2* p = base + (long long)t * sh.tile;
        double acc = 0.0;
        for (int i = lane; i < sh.tile; i += 32) {
            const double2 v = __ldcs(p + i);
            acc += v.x * v.x + v.y * v.y;
        }
        acc = warp_sum(acc);
        if (lane == 0) tm[t] = acc;
        mine += acc;                                   // identical in every lane after warp_sum
    }
    if (lane == 0) warp_part[warp] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += warp_part[w];
        super_mass[state * sh.n_super + sup] = s;
    }
}

// ---- super scan: one warp per state, inclusive running sum in place ------------------------------
__global__ void __launch_bounds__(128) sample_scan_kernel(double* __restrict__ super_mass, long long n_super,
                                                          long long n_states) {
    const long long state = blockIdx.x * 4ll + (threadIdx.x >> 5);
    if (state >= n_states) return;
    const int lane = threadIdx.x & 31;
    double* p = super_mass + state * n_super;
    double carry = 0.0;
    for (long long i = 0; i < n_super; i += 32) {
        double v = (i + lane < n_super) ? p[i + lane] : 0.0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double up = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += up;
        }
        v += carry;
        if (i + lane < n_super) p[i + lane] = v;
        carry = __shfl_sync(0xffffffffu, v, 31);
    }
}

// ---- Philox4x32-10 (Salmon et al., SC'11), counter-based: no state to carry between shots -------------
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

__device__ __forceinline__ double philox_uniform(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
    uint32_t c[4] = {c0, c1, c2, c3};
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        philox_round(c, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    const uint64_t bits = ((uint64_t)c[0] << 32) | c[1];
    return (double)(bits >> 11) * (1.0 / 9007199254740992.0);             // 53 bits, [0, 1)
}

// Walk `count` masses (count a multiple of 32 or < 32) for the first index whose running sum
// exceeds `target`; returns the index and leaves in `target` the remainder inside that entry.
// All lanes take part and return the same values.  `load(i)` gives the mass of entry i.
template <class Load>
__device__ __forceinline__ int warp_walk(int count, double& target, Load load) {
    const int lane = threadIdx.x & 31;
    int last_nz = -1;
    double last_nz_before = 0.0;
    double before = 0.0;                               // mass of the entries already passed
    for (int i = 0; i < count; i += 32) {
        const double m = (i + lane < count) ? load(i + lane) : 0.0;
        double s = m;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double up = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += up;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, m > 0.0 && before + s > target);
        if (hit) {
            const int l = __ffs(hit) - 1;
            const double incl = __shfl_sync(0xffffffffu, s, l), mass = __shfl_sync(0xffffffffu, m, l);
            target -= before + (incl - mass);
            if (target < 0.0) target = 0.0;
            return i + l;
        }
        const unsigned nz = __ballot_sync(0xffffffffu, m > 0.0);
        if (nz) {
            const int l = 31 - __clz(nz);
            const double incl = __shfl_sync(0xffffffffu, s, l), mass = __shfl_sync(0xffffffffu, m, l);
            last_nz = i + l;
            last_nz_before = before + (incl - mass);
        }
        before += __shfl_sync(0xffffffffu, s, 31);
    }
    // fell off the end by rounding: the last entry that has any mass, remainder = all of it
    if (last_nz >= 0) {
        target -= last_nz_before;
        return last_nz;
    }
    return -1;                                         // level without mass
}

struct DrawArgs {
    const double2* states;
    const double* tile_mass;
    const double* super_cum;
    const int32_t* columns;      // [n_init][2^m]
    double* counts;              // [n_sets][2^(n - m)][ld]
    SampleShape sh;
    long long n_init, n_sets, shots, ld;
    uint64_t flip, meas_mask, seed, stream_id;
    int n_meas;
};

__device__ __forceinline__ uint64_t extract_bits(uint64_t x, uint64_t mask) {
    uint64_t out = 0;
    int j = 0;
    while (mask) {
        const int q = __ffsll((long long)mask) - 1;
        out |= ((x >> q) & 1ull) << j;
        ++j;
        mask &= mask - 1;
    }
    return out;
}

// grid.x covers shots in groups of 8 (one warp each), grid.y = states
__global__ void __launch_bounds__(256) sample_draw_kernel(const DrawArgs a) {
    const long long state = blockIdx.y;
    const long long shot = blockIdx.x * 8ll + (threadIdx.x >> 5);
    if (shot >= a.shots) return;
    const int lane = threadIdx.x & 31;
    const long long init = state / a.n_sets, set = state % a.n_sets;
    const SampleShape& sh = a.sh;
    const double* cum = a.super_cum + state * sh.n_super;
    const double total = cum[sh.n_super - 1];
    if (!(total > 0.0)) return;                        // a state without norm draws nothing
    const double u = philox_uniform(a.seed ^ (a.stream_id * 0x9E3779B97F4A7C15ull), (uint32_t)shot,
                                    (uint32_t)((uint64_t)shot >> 32) ^ ((uint32_t)init << 8), (uint32_t)set,
                                    (uint32_t)a.stream_id);
    double target = u * total;
    // level 1: first super-tile whose running sum exceeds the target
    long long lo = 0, hi = sh.n_super - 1;
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (cum[mid] > target) hi = mid; else lo = mid + 1;
    }
    const long long sup = lo;
    target -= sup ? cum[sup - 1] : 0.0;
    if (target < 0.0) target = 0.0;
    // level 2: tile inside the super-tile
    const double* tm = a.tile_mass + state * sh.n_tiles + sup * sh.tiles_per_super;
    int t = warp_walk(sh.tiles_per_super, target, [&](int i) { return tm[i]; });
    if (t < 0) {                                       // rounding put the target in an empty super-tile:
        for (long long s2 = sh.n_super - 1; s2 >= 0 && t < 0; --s2) {   // take the last one with mass
            tm = a.tile_mass + state * sh.n_tiles + s2 * sh.tiles_per_super;
            target = 1e300;
            t = warp_walk(sh.tiles_per_super, target, [&](int i) { return tm[i]; });
            if (t >= 0) lo = s2;
        }
        if (t < 0) return;
        target = 1e300;
    }
    // level 3: amplitude inside the tile
    const long long tile_first = (lo * sh.tiles_per_super + t) * (long long)sh.tile;
    const double2* amp = a.states + state * sh.n_amp + tile_first;
    const int k = warp_walk(sh.tile, target, [&](int i) { const double2 v = amp[i]; return v.x * v.x + v.y * v.y; });
    if (k < 0) return;
    if (lane == 0) {
        const uint64_t x = (uint64_t)(tile_first + k) ^ a.flip;          // logical basis state
        const uint64_t keep = (~a.meas_mask) & (((uint64_t)1 << sh.n) - 1ull);
        const uint64_t b = extract_bits(x, keep), mid = extract_bits(x, a.meas_mask);
        const int col = a.columns[init * (1ll << a.n_meas) + (long long)mid];
        atomicAdd(a.counts + ((unsigned long long)set * (1ull << (sh.n - a.n_meas)) + b) * a.ld + col, 1.0);
    }
}

}  // namespace

extern "C" int64_t qdc_sv_sample_workspace_doubles(int n_qubits, int64_t n_states) {
    if (n_qubits < 1 || n_qubits > 40 || n_states < 1) return -1;
    const SampleShape sh = shape_of(n_qubits);
    return n_states * (sh.n_tiles + sh.n_super);
}

extern "C" int qdc_sv_sample_counts(const void* d_state, int n_qubits, int64_t n_init, int64_t n_sets,
                                    uint64_t flip_mask, int64_t shots, uint64_t seed, uint64_t stream_id,
                                    uint64_t meas_mask, const int32_t* d_columns, double* d_counts, int64_t ld,
                                    double* d_workspace, void* stream) {
    QDC_REQUIRE(d_state && d_columns && d_counts && d_workspace, "null buffer");
    QDC_REQUIRE(n_qubits >= 1 && n_qubits <= 40 && n_init >= 1 && n_sets >= 1, "shape");
    QDC_REQUIRE(shots >= 0 && shots < (1ll << 40), "shots");
    QDC_REQUIRE((flip_mask >> n_qubits) == 0 && (meas_mask >> n_qubits) == 0, "masks");
    QDC_REQUIRE(ld >= 1, "ld");
    const long long n_states = n_init * n_sets;
    QDC_REQUIRE(n_states <= 65535, "at most 65535 states per call (grid.y)");
    if (shots == 0) return QDC_OK;
    const SampleShape sh = shape_of(n_qubits);
    QDC_REQUIRE(sh.n_super <= 0x7fffffffll, "state too large");
    cudaStream_t st = (cudaStream_t)stream;
    double* tile_mass = d_workspace;
    double* super_mass = d_workspace + n_states * sh.n_tiles;
    sample_mass_kernel<<<dim3((unsigned)sh.n_super, (unsigned)n_states), 256, 0, st>>>(
        (const double2*)d_state, sh, tile_mass, super_mass);
    QDC_CHECK_LAUNCH();
    sample_scan_kernel<<<(unsigned)((n_states + 3) / 4), 128, 0, st>>>(super_mass, sh.n_super, n_states);
    QDC_CHECK_LAUNCH();
    DrawArgs a;
    a.states = (const double2*)d_state; a.tile_mass = tile_mass; a.super_cum = super_mass;
    a.columns = d_columns; a.counts = d_counts; a.sh = sh;
    a.n_init = n_init; a.n_sets = n_sets; a.shots = shots; a.ld = ld;
    a.flip = flip_mask; a.meas_mask = meas_mask; a.seed = seed; a.stream_id = stream_id;
    a.n_meas = __builtin_popcountll(meas_mask);
    const long long gx = (shots + 7) / 8;
    QDC_REQUIRE(gx <= 0x7fffffffll, "shots");
    sample_draw_kernel<<<dim3((unsigned)gx, (unsigned)n_states), 256, 0, st>>>(a);
    QDC_CHECK_LAUNCH();
    return QDC_OK;
}
