// Shared helpers for the turbo_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <mutex>

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
// once-per-DEVICE configuration.  cudaFuncSetAttribute (dynamic shared-memory opt-in, carve-out) acts on the current
// device's context: a process that touches a second GPU must repeat it there.  One DeviceOnce per attribute group;
// mutex-guarded because the caller may be any host thread (the reference's Qt GUI runs the optimiser on a QThread,
// turbo/gui/qt.py:139-147).
// ---------------------------------------------------------------------------------------------
constexpr int MAX_DEVICES = 64;
struct DeviceOnce {
    std::mutex m;
    bool done[MAX_DEVICES] = {};
};
template <class F>
inline int device_once(DeviceOnce& o, F&& configure) {
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CHECK(dev >= 0 && dev < MAX_DEVICES, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(o.m);
    if (!o.done[dev]) {
        if (configure()) return -1;
        o.done[dev] = true;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// The ONE debug hook of the library: the environment variable TB200_DEBUG, a comma-separated list of key=value pairs,
// parsed once on first use (never on the call path afterwards; tb200_debug_reload() re-reads it for tests).
//   fit_engine=dmma      keep the batched fit on the FP64 DMMA GEMM (the cross-check the tests compare the INT8 fit with)
//   fit_gmax=7|8|9       slice-pair cut-off of every INT8 fit stage (default 8/7/7: potrf, trtri, K^-1)
//   fit_gmax_trtri=..    the inverse factor alone;   fit_gmax_kinv=..  the K^-1 product alone
// Everything else that used to be switchable (stream priorities, K* staging, look-ahead, schedule thresholds) was an
// A/B question of round 1 and is settled; the measurements are under profiles/r1_*.
// ---------------------------------------------------------------------------------------------
struct DebugConfig {
    bool fit_dmma = false;
    int fit_gmax[3] = {8, 7, 7};
};
const DebugConfig& debug_config();
void debug_config_reload();

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

// Branch-free sqrt and exp(-t) for the pair loops.  The library versions carry slow-path branches and calls (denormals,
// overflow) that end a basic block per evaluation, so the 16 independent evaluations of an unrolled pair loop ran one
// after the other on their dependent chains; these two are straight-line code the compiler interleaves.
// They also stay off the XU pipe (conversions, MUFU; 16 lanes/clk/SM): ncu showed the K* kernel at 92 % XU busy with
// four XU operations per pair (double->float, MUFU.RSQ, float->double, double->u64), 36 % FP64.  sqrt_pos uses ONE
// (MUFU.RSQ64H through rsqrt.approx.ftz.f64) and the fixed-point conversions below use none.
// sqrt_pos: a >= 0.  Reciprocal square root to ~2^-22 from the high word, two coupled Newton steps (-> 2^-44 -> 2^-88)
// and a residual correction: <= 1 ulp.  a < 1e-290 (incl. 0 and subnormals, which the approximation flushes) returns 0;
// a > 1e300 returns 1e150 (every caller's result is 0 there); NaN propagates.
__device__ __forceinline__ double sqrt_pos(double a) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    double g = a * y, h = 0.5 * y;
    double r = fma(-g, h, 0.5);
    g = fma(g, r, g);
    h = fma(h, r, h);
    r = fma(-g, h, 0.5);
    g = fma(g, r, g);
    h = fma(h, r, h);
    g = fma(fma(-g, g, a), h, g);
    g = a < 1e-290 ? 0.0 : g;
    return a > 1e300 ? 1e150 : g;
}

// rn(x * 2^(32 + S)) as a 64-bit integer for |x * 2^S| < 2^31, without a conversion instruction: the high part
// rn(x 2^S) and the rounded remainder (x 2^S - high) 2^32 are read out of the mantissa of (value + 1.5 * 2^52).
// The remainder is exact (one FMA), so the result is rounded once.  S = 24: unsigned digits of g in [0, 1] (2^56 g);
// S = 23: balanced digits of |a| < 1 (2^55 a).
template <int S>
__device__ __forceinline__ long long fixed_point_64(double x) {
    const double M = 6755399441055744.0;       // 1.5 * 2^52
    const double scale = (double)(1 << S);
    const double t = fma(x, scale, M);
    const int hi = __double2loint(t);          // low word of the mantissa = rn(x 2^S) (two's complement)
    const double rem = fma(x, scale, -(t - M));  // in [-0.5, 0.5], exact
    const double u = fma(rem, 4294967296.0, M);
    const long long lo = (__double_as_longlong(u) & 0x000FFFFFFFFFFFFFll) - (1ll << 51);  // in [-2^31, 2^31]
    return ((long long)hi << 32) + lo;
}

// exp_neg(t) = exp(-t), t >= 0: k = round(-t / ln 2) by the 1.5 * 2^52 shift, two-term Cody-Waite reduction, degree-13
// Taylor polynomial on |r| <= ln(2)/2 (remainder 4e-18), scaling through the exponent field.  Results below 2^-1021
// are flushed to 0 (absolute error 5e-308).  About 1 ulp; NaN propagates.
__device__ __forceinline__ double exp_neg(double t) {
    const double x = -(t > 750.0 ? 750.0 : t);
    const double kd = fma(x, 1.4426950408889634, 6755399441055744.0);
    const int k = __double2loint(kd);
    const double kf = kd - 6755399441055744.0;
    double r = fma(kf, -6.93147180369123816490e-01, x);
    r = fma(kf, -1.90821492927058770002e-10, r);
    double p = 1.6059043836821613e-10;              // 1/13!
    p = fma(p, r, 2.08767569878681e-09);            // 1/12!
    p = fma(p, r, 2.505210838544172e-08);           // 1/11!
    p = fma(p, r, 2.755731922398589e-07);           // 1/10!
    p = fma(p, r, 2.7557319223985893e-06);          // 1/9!
    p = fma(p, r, 2.48015873015873e-05);            // 1/8!
    p = fma(p, r, 1.984126984126984e-04);           // 1/7!
    p = fma(p, r, 1.388888888888889e-03);           // 1/6!
    p = fma(p, r, 8.333333333333333e-03);           // 1/5!
    p = fma(p, r, 4.1666666666666664e-02);          // 1/4!
    p = fma(p, r, 1.6666666666666666e-01);          // 1/3!
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    const double scaled = __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
    const double v = k < -1020 ? 0.0 : scaled;
    return t != t ? t : v;
}

// stationary shape g as a function of the squared scaled distance r2 (sklearn:kernels.py:1561-1570,
// :1722-1729).  Matern families take sqrt first, exactly as cdist/pdist 'euclidean' does.
__device__ __forceinline__ double shape_fn(int kind, double r2) {
    if (kind == TB200_RBF) return exp_neg(0.5 * r2);
    const double r = sqrt_pos(r2);
    if (kind == TB200_MATERN12) return exp_neg(r);
    if (kind == TB200_MATERN32) {
        const double t = r * 1.7320508075688772;  // math.sqrt(3)
        return (1.0 + t) * exp_neg(t);
    }
    const double t = r * 2.23606797749979;  // math.sqrt(5)
    return (1.0 + t + t * t * (1.0 / 3.0)) * exp_neg(t);  // sklearn divides by 3.0: <= 1 ulp apart
}

// h(r2) with dK/dlog(l_k) = sf2 * h * D_k, D_k = ((x_k - x'_k)/l_k)^2 (sklearn:kernels.py:1581-1584,
// :1757-1771)
__device__ __forceinline__ double shape_grad_fn(int kind, double r2) {
    if (kind == TB200_RBF) return exp_neg(0.5 * r2);
    if (kind == TB200_MATERN12) {
        const double r = sqrt_pos(r2);
        return r2 != 0.0 ? exp_neg(r) / r : 0.0;
    }
    if (kind == TB200_MATERN32) return 3.0 * exp_neg(sqrt_pos(3.0 * r2));
    const double t = sqrt_pos(5.0 * r2);
    return 5.0 / 3.0 * (t + 1.0) * exp_neg(t);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace tb200
