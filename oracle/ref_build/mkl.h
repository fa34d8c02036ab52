/* Minimal stand-in for Intel MKL's <mkl.h>, written for this repo so that the reference's
 * cutqc C programs (cutqc/build.c, merge.c, vertical_collapse.c) compile with gcc where MKL is
 * not installed.  Only the symbols those files use are provided (build.c:80-81,142,144,154,169;
 * vertical_collapse.c:62), as plain C loops -- this is NOT MKL and is slower than MKL. */
#ifndef QDC_MKL_SHIM_H
#define QDC_MKL_SHIM_H
typedef long long MKL_INT;
typedef enum { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_LAYOUT;

static inline void cblas_sger(CBLAS_LAYOUT layout, MKL_INT m, MKL_INT n, float alpha,
                              const float *x, MKL_INT incx, const float *y, MKL_INT incy,
                              float *a, MKL_INT lda) {
    (void)layout; /* row-major only: a[i*lda + j] += alpha * x[i] * y[j] */
    for (MKL_INT i = 0; i < m; ++i) {
        const float xi = alpha * x[i * incx];
        float *row = a + i * lda;
        for (MKL_INT j = 0; j < n; ++j) row[j] += xi * y[j * incy];
    }
}
static inline void cblas_scopy(MKL_INT n, const float *x, MKL_INT incx, float *y, MKL_INT incy) {
    for (MKL_INT i = 0; i < n; ++i) y[i * incy] = x[i * incx];
}
static inline void cblas_sscal(MKL_INT n, float a, float *x, MKL_INT incx) {
    for (MKL_INT i = 0; i < n; ++i) x[i * incx] *= a;
}
static inline void cblas_saxpy(MKL_INT n, float a, const float *x, MKL_INT incx, float *y,
                               MKL_INT incy) {
    for (MKL_INT i = 0; i < n; ++i) y[i * incy] += a * x[i * incx];
}
static inline void vsAdd(MKL_INT n, const float *a, const float *b, float *r) {
    for (MKL_INT i = 0; i < n; ++i) r[i] = a[i] + b[i];
}
#endif
