// Shared helpers for the turbo_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/turbo_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "turbo_b200 is written for sm_100a (B200) only"
#endif

namespace tb200 {

constexpr int TILE = 128;  // padding granularity and blocked-factorisation block size

extern thread_local char g_err[512];
extern int64_t g_launches;

inline int fail(const char* what, const char* file, int line) {
    snprintf(g_err, sizeof(g_err), "%s (%s:%d)", what, file, line);
    return -1;
}

#define TB_CHECK(cond, msg)                                        \
    do {                                                           \
        if (!(cond)) return ::tb200::fail(msg, __FILE__, __LINE__); \
    } while (0)

#define TB_CUDA(expr)                                                             \
    do {                                                                          \
        cudaError_t e__ = (expr);                                                 \
        if (e__ != cudaSuccess) return ::tb200::fail(cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

// every launch goes through this so bench.py can report gpu_launches
#define TB_LAUNCH(kernel, grid, block, smem, stream, ...)                         \
    do {                                                                          \
        kernel<<<grid, block, smem, stream>>>(__VA_ARGS__);                       \
        ++::tb200::g_launches;                                                    \
        cudaError_t e__ = cudaGetLastError();                                     \
        if (e__ != cudaSuccess) return ::tb200::fail(cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

inline int pad_n(int n) { return (n + TILE - 1) / TILE * TILE; }

// ---------------------------------------------------------------------------------------------
// hyper-parameters: theta (log space, sklearn order) -> physical values
// ---------------------------------------------------------------------------------------------
struct Hyper {
    double sf2;    // ConstantKernel value (1 if absent)
    double noise;  // WhiteKernel level (0 if absent)
};

__host__ __device__ inline int spec_p(const tb200_kernel_spec& s) {
    return (s.has_const ? 1 : 0) + (s.ard ? s.d : 1) + (s.has_white ? 1 : 0);
}

__device__ inline Hyper unpack_hyper(const tb200_kernel_spec& s, const double* theta) {
    Hyper h;
    h.sf2 = s.has_const ? exp(theta[0]) : 1.0;
    const int nl = s.ard ? s.d : 1;
    h.noise = s.has_white ? exp(theta[(s.has_const ? 1 : 0) + nl]) : 0.0;
    return h;
}

__device__ inline double length_scale(const tb200_kernel_spec& s, const double* theta, int k) {
    return exp(theta[(s.has_const ? 1 : 0) + (s.ard ? k : 0)]);
}

// stationary shape g as a function of the squared scaled distance r2 (sklearn:kernels.py:1561-1570,
// :1722-1729).  Matern families take sqrt first, exactly as cdist/pdist 'euclidean' does.
__device__ __forceinline__ double shape_fn(int kind, double r2) {
    if (kind == TB200_RBF) return exp(-0.5 * r2);
    const double r = sqrt(r2);
    if (kind == TB200_MATERN12) return exp(-r);
    if (kind == TB200_MATERN32) {
        const double t = r * 1.7320508075688772;  // math.sqrt(3)
        return (1.0 + t) * exp(-t);
    }
    const double t = r * 2.23606797749979;  // math.sqrt(5)
    return (1.0 + t + t * t * (1.0 / 3.0)) * exp(-t);  // sklearn divides by 3.0: <= 1 ulp apart
}

// h(r2) with dK/dlog(l_k) = sf2 * h * D_k, D_k = ((x_k - x'_k)/l_k)^2 (sklearn:kernels.py:1581-1584,
// :1757-1771)
__device__ __forceinline__ double shape_grad_fn(int kind, double r2) {
    if (kind == TB200_RBF) return exp(-0.5 * r2);
    if (kind == TB200_MATERN12) {
        const double r = sqrt(r2);
        return r != 0.0 ? exp(-r) / r : 0.0;
    }
    if (kind == TB200_MATERN32) return 3.0 * exp(-sqrt(3.0 * r2));
    const double t = sqrt(5.0 * r2);
    return 5.0 / 3.0 * (t + 1.0) * exp(-t);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace tb200
