// Layout of a PREPARED operand of the int8-sliced contraction (qdc_i8_prepare -> qdc_contract_i8_prepared).
//
// One pass over a row-major FP64 operand X[n_rows][kdim] produces, in ONE caller-owned device buffer:
//   slices   int8  [n_slices][rows_pad][kd]   radix-256 balanced digits, K-major core-matrix order
//                                             ([8-row group][16-byte k chunk][8 rows][16 bytes]); the
//                                             order does not depend on the tile width, so one buffer
//                                             serves 64-, 96- and 128-row tiles (rows_pad % 384 == 0)
//   scale    f64   [rows_pad]                 2^(e_r - 7): x = scale * sum_i digit_i 256^-i
//   colq     f64x2 [rows_pad]                 (scale, scale * weight)               (column role)
//   coltab   u32x4 [rows_pad]                 (x_lo, x_hi | flags, weight << lsh, 1 << lsh) (column role)
//   gref     i32   [rows_pad / 32]            exponent of the largest row of each 32-row group
//   part     f64   [PREP_GRID][2 kd + 4]      per-CTA column sums (sum x, sum weight x) and the
//                                             statistics the error bounds need
// kd = kdim rounded up to 32 (zero padded).  Nothing here is visible to callers but the byte count.
#pragma once
#include <stdint.h>
#include <stddef.h>

constexpr int I8_PREP_GRID = 296;        // CTAs of the preparation pass = partial rows of `part`
constexpr int I8_ROW_ALIGN = 384;        // lcm of the tile heights / widths in use (128, 96, 64)
constexpr int I8_GROUP = 32;             // adjacent columns that share one integer unit (gref): what one epilogue thread sums
constexpr int I8_FOLD_COLS = 16;         // fewest columns an epilogue thread folds (and rounds) at a time: enters the error bound
constexpr int I8_LSH_MAX = 20;           // a column keeps its own exponent down to 2^-20 of its group's largest
constexpr int I8_STAT = 4;               // per-CTA statistics: max 2^e, sum_{ok} 2^e, max e (as a double), spare

struct I8Layout {
    int kd, n_slices;
    long long rows_pad;
    size_t slice_stride;                 // bytes between slices
    size_t off_slices, off_scale, off_colq, off_coltab, off_gref, off_part, total;
};

inline size_t i8_align256(size_t x) { return (x + 255) / 256 * 256; }

inline I8Layout i8_layout(long long n_rows, int kdim, int n_slices) {
    I8Layout L;
    L.kd = (kdim + 31) / 32 * 32;
    L.n_slices = n_slices;
    L.rows_pad = (n_rows + I8_ROW_ALIGN - 1) / I8_ROW_ALIGN * I8_ROW_ALIGN;
    if (L.rows_pad == 0) L.rows_pad = I8_ROW_ALIGN;
    L.slice_stride = (size_t)L.rows_pad * L.kd;
    size_t o = 0;
    L.off_slices = o;  o = i8_align256(o + L.slice_stride * n_slices);
    L.off_scale = o;   o = i8_align256(o + (size_t)L.rows_pad * 8);
    L.off_colq = o;    o = i8_align256(o + (size_t)L.rows_pad * 16);
    L.off_coltab = o;  o = i8_align256(o + (size_t)L.rows_pad * 16);
    L.off_gref = o;    o = i8_align256(o + (size_t)(L.rows_pad / I8_GROUP) * 4);
    L.off_part = o;    o = i8_align256(o + (size_t)I8_PREP_GRID * (2 * L.kd + I8_STAT) * 8);
    L.total = o;
    return L;
}

// result / control block of one contraction call (16 doubles, caller-owned, device)
constexpr int I8_RES_SUMS = 0;           // [0..2] sum p, sum p w [indset], sum p w
constexpr int I8_RES_NORM = 3;           // the normalisation the kernels divide by
constexpr int I8_RES_SLICES = 4;         // slice count in use (chosen on the device when n_slices = 0)
constexpr int I8_RES_BOUND_OUT = 5;      // a-posteriori bound per output for that count
constexpr int I8_RES_BOUND_COST = 6;     // ... and on the masked sum
constexpr int I8_RES_CAND = 7;           // candidate counter (unsigned 64-bit integer in the same slot size)
constexpr int I8_RES_EMAX = 8;           // largest column exponent (integer valued)
constexpr int I8_RES_FLAGGED = 9;        // outputs the integer filter flagged (uint64): entries of the candidate arrays before re-evaluation
constexpr int I8_RES_LOST = 10;          // candidates above the threshold that found no room in the arrays (uint64)
constexpr int I8_RES_DOUBLES = 16;
