// Shared device/host helpers for libqdc_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/qdc_b200.h"

#define QDC_MAX_PARTIALS 4096          // upper bound on reducing-grid size (CTAs)
#define QDC_NUM_SMS_B200 148

void qdc_set_error(const char* fmt, ...);

#define QDC_CHECK_CUDA(expr)                                                        \
    do {                                                                            \
        cudaError_t _e = (expr);                                                    \
        if (_e != cudaSuccess) {                                                    \
            qdc_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,              \
                          cudaGetErrorString(_e));                                  \
            return QDC_ERR_CUDA;                                                    \
        }                                                                           \
    } while (0)

#define QDC_CHECK_LAUNCH() QDC_CHECK_CUDA(cudaGetLastError())

#define QDC_REQUIRE(cond, msg)                                                      \
    do {                                                                            \
        if (!(cond)) {                                                              \
            qdc_set_error("%s:%d argument check failed: %s (%s)", __FILE__,         \
                          __LINE__, #cond, msg);                                    \
            return QDC_ERR_ARG;                                                     \
        }                                                                           \
    } while (0)

int qdc_sm_count();

// ---- MIS epilogue shared by the fused contraction (K5) and the standalone reduction (K8) ----
// Per row r: row_x (circuit-order bits), row_forb (union of neighbour masks), row_wok
// (weight | ok << 8).  Per column likewise (no forb needed: the cross test is one-sided).
struct MisTables {
    const uint64_t* __restrict__ row_x;
    const uint64_t* __restrict__ row_forb;
    const uint32_t* __restrict__ row_wok;
    const uint64_t* __restrict__ col_x;
    const uint32_t* __restrict__ col_wok;
};

struct MisAccum {
    double s_p, s_pwok, s_pw;
};

struct CandSink {
    uint64_t* x;
    double* p;
    unsigned long long* count;
    long long cap;
    double threshold;
};

__device__ __forceinline__ void cand_emit(const CandSink& sink, uint64_t x, double p) {
    unsigned long long slot = atomicAdd(sink.count, 1ull);
    if ((long long)slot < sink.cap) {
        sink.x[slot] = x;
        sink.p[slot] = p;
    }
}

__device__ __forceinline__ void mis_accumulate(MisAccum& acc, double p, uint64_t rx, uint64_t rforb,
                                               uint32_t rwok, uint64_t cx, uint32_t cwok,
                                               const CandSink& sink) {
    const uint32_t w = (rwok & 0xffu) + (cwok & 0xffu);
    const bool ok = ((rwok & cwok) >> 8) & 1u ? ((cx & rforb) == 0ull) : false;
    const double pw = p * (double)w;
    acc.s_p += p;
    acc.s_pw += pw;
    if (ok) acc.s_pwok += pw;
    if (p > sink.threshold) {
        unsigned long long slot = atomicAdd(sink.count, 1ull);
        if ((long long)slot < sink.cap) {
            sink.x[slot] = rx | cx;
            sink.p[slot] = p;
        }
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-level reduction of the three accumulators into partials[3*blockIdx.x + {0,1,2}].
// `scratch` needs 3 * (blockDim.x / 32) doubles of shared memory.
__device__ __forceinline__ void mis_block_store(const MisAccum& acc, double* scratch,
                                                double* __restrict__ partials) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    double a = warp_sum(acc.s_p), b = warp_sum(acc.s_pwok), c = warp_sum(acc.s_pw);
    if (lane == 0) {
        scratch[warp] = a;
        scratch[nwarp + warp] = b;
        scratch[2 * nwarp + warp] = c;
    }
    __syncthreads();
    if (warp == 0) {
        double x = 0, y = 0, z = 0;
        for (int i = lane; i < nwarp; i += 32) {
            x += scratch[i];
            y += scratch[nwarp + i];
            z += scratch[2 * nwarp + i];
        }
        x = warp_sum(x); y = warp_sum(y); z = warp_sum(z);
        if (lane == 0) {
            partials[3 * blockIdx.x + 0] = x;
            partials[3 * blockIdx.x + 1] = y;
            partials[3 * blockIdx.x + 2] = z;
        }
    }
}

// launched with one CTA of 256 threads: sums[j] += sum_i partials[3*i + j], fixed order.
int qdc_launch_finalize_sums(const double* d_partials, int n_partials, double* d_sums,
                             cudaStream_t stream);
