// GP fit path: Gram matrix, blocked FP64 Cholesky, triangular inverse, alpha, log-marginal
// likelihood and its gradient -- batched over R hyper-parameter restarts.
//
// Follows sklearn:_gpr.py:584-651 (log_marginal_likelihood) and :349-367 (final factor) behind
// turbo/modules/surrogates.py:294-326.  All dense O(n^3) work is DMMA GEMM (gemm_f64.cuh):
//   potrf : right-looking, 128-wide panels; the 128x128 diagonal block is factored AND inverted in
//           shared memory so the panel solve is a GEMM with Dinv^T;
//   trtri : right-looking TRSM of the identity (same dependency shape as potrf), GEMM only;
//   K^-1  : Linv^T Linv, one triangular GEMM;
//   grad  : 0.5 * sum_ij (alpha alpha^T - K^-1)_ij dK_ij/dtheta with dK recomputed on the fly
//           (sklearn materialises an (n,n,p) tensor: kernels.py:1753-1771).
#include <stdlib.h>
#include <string.h>

#include <string>

#include "gemm_f64.cuh"
#include "ozaki_gemm.cuh"

namespace tb200 {

thread_local char g_err[512] = "";
int64_t g_launches = 0;

// TB200_DEBUG (see common.cuh): parsed once; the call paths only read the struct
static DebugConfig g_debug;
static std::once_flag g_debug_once;
static void parse_debug_config() {
    DebugConfig c;
    const char* env = getenv("TB200_DEBUG");
    std::string text = env ? env : "";
    size_t pos = 0;
    while (pos < text.size()) {
        size_t end = text.find(',', pos);
        if (end == std::string::npos) end = text.size();
        const std::string item = text.substr(pos, end - pos);
        pos = end + 1;
        const size_t eq = item.find('=');
        if (eq == std::string::npos) continue;
        const std::string key = item.substr(0, eq), val = item.substr(eq + 1);
        const int iv = atoi(val.c_str());
        const bool g_ok = iv >= 7 && iv <= 9;
        if (key == "fit_engine") c.fit_dmma = !val.empty() && val[0] == 'd';
        else if (key == "fit_gmax" && g_ok) c.fit_gmax[0] = c.fit_gmax[1] = c.fit_gmax[2] = iv;
        else if (key == "fit_gmax_trtri" && g_ok) c.fit_gmax[1] = iv;
        else if (key == "fit_gmax_kinv" && g_ok) c.fit_gmax[2] = iv;
    }
    g_debug = c;
}
const DebugConfig& debug_config() {
    std::call_once(g_debug_once, parse_debug_config);
    return g_debug;
}
void debug_config_reload() {
    debug_config();
    parse_debug_config();
}

// ---------------------------------------------------------------------------------------------
// XsT[r][k][j] = X[j][k] / length_scale_k(theta_r)   (j < n; 0 in the padding)
// ---------------------------------------------------------------------------------------------
__global__ void scale_points_kernel(const double* __restrict__ X, int n, int NP, tb200_kernel_spec spec,
                                    const double* __restrict__ theta, int p, double* __restrict__ XsT) {
    const int r = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= NP) return;
    const double* th = theta + (long)r * p;
    double* out = XsT + (long)r * spec.d * NP;
    for (int k = 0; k < spec.d; ++k) {
        const double ls = length_scale(spec, th, k);
        out[(long)k * NP + j] = (j < n) ? X[(long)j * spec.d + k] / ls : 0.0;
    }
}

// ---------------------------------------------------------------------------------------------
// K1: lower tiles (64x64) of K = sf2 * g(r) [+ noise on the diagonal] + jitter*I; identity padding.
// sklearn:kernels.py:1558-1565 / :1713-1743 (pdist on X/l, diagonal forced to 1), :1315/:1438,
// sklearn:_gpr.py:589.
// ---------------------------------------------------------------------------------------------
constexpr int GT = 64;

__device__ __forceinline__ void lower_tile(int l, int& ti, int& tj) {
    ti = (int)((sqrt(8.0 * l + 1.0) - 1.0) * 0.5);
    while ((long)(ti + 1) * (ti + 2) / 2 <= l) ++ti;
    while ((long)ti * (ti + 1) / 2 > l) --ti;
    tj = l - ti * (ti + 1) / 2;
}

__global__ void __launch_bounds__(256) gram_train_kernel(const double* __restrict__ XsT, int n, int NP,
                                                         tb200_kernel_spec spec, const double* __restrict__ theta,
                                                         int p, double jitter, double* __restrict__ K) {
    extern __shared__ double sh[];  // xi[d][64], xj[d][64]
    const int d = spec.d;
    double* xi = sh;
    double* xj = sh + d * GT;
    const int r = blockIdx.y;
    int ti, tj;
    lower_tile(blockIdx.x, ti, tj);
    const double* xs = XsT + (long)r * d * NP;
    for (int e = threadIdx.x; e < d * GT; e += blockDim.x) {
        const int k = e / GT, c = e - k * GT;
        xi[e] = xs[(long)k * NP + ti * GT + c];
        xj[e] = xs[(long)k * NP + tj * GT + c];
    }
    __syncthreads();
    const Hyper h = unpack_hyper(spec, theta + (long)r * p);
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    double r2[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) r2[a][b] = 0.0;
    // thread (ty, tx) owns rows ty*4 + a and columns tx + 16*b: the 16 tx of a half-warp read 16 CONSECUTIVE doubles of
    // xj per load (conflict-free; the tx*4 + b mapping of round 1 was a 4-way bank conflict on every xj load:
    // 21 M conflicts per lml_grad launch, profiles/r2_fit_small_kernels_ncu.md) and write 128 contiguous bytes of K
    for (int k = 0; k < d; ++k) {
        double va[4], vb[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) va[a] = xi[k * GT + ty * 4 + a];
#pragma unroll
        for (int b = 0; b < 4; ++b) vb[b] = xj[k * GT + tx + 16 * b];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const double df = va[a] - vb[b];
                r2[a][b] += df * df;
            }
    }
    double* Kr = K + (long)r * NP * NP;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int i = ti * GT + ty * 4 + a;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int j = tj * GT + tx + 16 * b;
            double v;
            if (i >= n || j >= n) {
                v = (i == j) ? 1.0 : 0.0;
            } else if (i == j) {
                v = h.sf2 * 1.0 + h.noise;
                v += jitter;
            } else {
                v = h.sf2 * shape_fn(spec.kind, r2[a][b]);
            }
            Kr[(long)i * NP + j] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// 128x128 diagonal block: Cholesky in shared memory, then in-place inverse of the factor.
//   A block (kb,kb) <- L_kk (upper part zeroed);  Dinv[r][kb] <- L_kk^-1 (lower, upper zeroed)
//   info[r] = first failing pivot (1-based global index) if not already set; the block is then
//   completed with pivot 1 so the rest of the batch keeps running.
// ---------------------------------------------------------------------------------------------
constexpr int DB = 128, DLD = 129;
constexpr int DP = 8;   // columns per panel (factor) / rows per block (inverse)
constexpr int DT = 4;   // threads per row (factor) / per column (inverse)
constexpr int DIAG_THREADS = DB * DT;
// S[128][129] (factor, row stride 129: a thread walking its own row and the broadcast of a pivot row are both
// conflict-free) + X packed lower triangle [k(k+1)/2 + c] (inverse) + reciprocal diagonal + the 8x8 panel block
constexpr size_t DIAG_SMEM = (size_t)(DB * DLD + DB * (DB + 1) / 2 + DB + DP * (DP + 1) + DP) * sizeof(double);

// Cholesky factor and inverse factor of one 128x128 diagonal block per CTA, everything in shared memory.  The block is
// a chain of dependent columns, so the kernel is latency-bound and what counts is the fixed cost per step: a first
// version with one column per step (two barriers, a reduction and a square root each) took 199-225 us at any thread
// count (648 K warp instructions at 1.7 IPC with 1024 threads, 313 K at 0.75 with 256: ncu, round 1).  Here a step
// is a PANEL of 8 columns:
//   factor : four threads own row t.  Left-looking: they form A[t, j0..j0+7] - sum_{k<j0} L[t,k] L[j0+q,k] for the 8
//            columns of the panel at once (one load of L[t,k] feeds 8 FMAs against broadcast pivot rows), the 8 rows
//            of the diagonal 8x8 block publish theirs, ONE thread factorises that 8x8 block in registers, and every
//            row below solves its 8 entries against it in registers.  Three barriers per panel, 48 in all.
//   inverse: four threads own column c of X = L^-1; forward substitution in blocks of 8 rows:
//            y = delta - L[i0..i0+7, c0..i0) x, then the 8x8 triangular solve in registers with the reciprocal
//            diagonal kept from the factorisation.  A warp (8 columns) needs no barrier.
// LAPACK dpotf2 scales by the reciprocal pivot as well.  Measured: 85.7 us per launch (164 K warp instructions per CTA,
// IPC 1.0); what is left is the serial 8x8 factorisation of each panel and the short late panels.
// A/B of round 2 (tools/gpu_round2_o.sh): the same kernel with S column-major and the eight pivot entries fetched as four
// 16-byte loads was SLOWER (LML+gradient at 8 restarts 11.9 -> 12.4 ms with row stride 132, 13.3 ms with 136): a 16-byte
// shared-memory load is served a quarter-warp at a time, so four of them cost more wavefronts than eight 8-byte
// broadcasts.  The row-major layout below stays.
__device__ __forceinline__ double diag_group_sum(double v) {
#pragma unroll
    for (int o = DT / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(DIAG_THREADS, 1) diag_potrf_inv_kernel(double* __restrict__ Abase, long ld,
                                                                          long strideA, int kb,
                                                                          double* __restrict__ Dinv, long strideD,
                                                                          int* __restrict__ info) {
    extern __shared__ double S[];
    double* Xp = S + DB * DLD;
    double* rdiag = Xp + DB * (DB + 1) / 2;  // 1 / L[i,i]
    double* Bd = rdiag + DB;                 // [8][9] diagonal block of the current panel
    double* rd8 = Bd + DP * (DP + 1);        // reciprocal pivots of the current panel
    const int tid = threadIdx.x;
    const int r = blockIdx.x;
    double* A = Abase + (long)r * strideA + (long)kb * DB * ld + (long)kb * DB;
    for (int e = tid; e < DB * DB; e += DIAG_THREADS) {
        const int i = e >> 7, j = e & 127;
        S[i * DLD + j] = (j <= i) ? A[(long)i * ld + j] : 0.0;
    }
    __syncthreads();
    const int t = tid / DT, h = tid % DT;
    const double* mine = S + t * DLD;
    for (int j0 = 0; j0 < DB; j0 += DP) {
        double v[DP];
#pragma unroll
        for (int q = 0; q < DP; ++q) v[q] = 0.0;
        if (t >= j0) {
            const double* pr = S + j0 * DLD;
            for (int k = h; k < j0; k += DT) {
                const double m = mine[k];
#pragma unroll
                for (int q = 0; q < DP; ++q) v[q] = fma(m, pr[q * DLD + k], v[q]);
            }
        }
#pragma unroll
        for (int q = 0; q < DP; ++q) v[q] = mine[j0 + q] - diag_group_sum(v[q]);
        if (t >= j0 && t < j0 + DP && h == 0) {
#pragma unroll
            for (int q = 0; q < DP; ++q) Bd[(t - j0) * (DP + 1) + q] = v[q];
        }
        __syncthreads();
        if (tid == 0) {
            double b[DP][DP];
#pragma unroll
            for (int i = 0; i < DP; ++i)
#pragma unroll
                for (int q = 0; q <= i; ++q) b[i][q] = Bd[i * (DP + 1) + q];
#pragma unroll
            for (int q = 0; q < DP; ++q) {
                double d = b[q][q];
#pragma unroll
                for (int pp = 0; pp < q; ++pp) d = fma(-b[q][pp], b[q][pp], d);
                if (!(d > 0.0)) {  // also catches NaN, as LAPACK dpotrf does
                    atomicCAS(&info[r], 0, kb * DB + j0 + q + 1);
                    d = 1.0;
                }
                // reciprocal square root + one Newton correction each (the IEEE sqrt and division sequences are ~250
                // dependent cycles apiece, and this thread is the critical path of the panel): <= 1 ulp
                double ri = rsqrt(d);
                double rd = d * ri;
                rd = fma(fma(-rd, rd, d), 0.5 * ri, rd);
                ri = fma(fma(-rd, ri, 1.0), ri, ri);
                b[q][q] = rd;
                rd8[q] = ri;
                rdiag[j0 + q] = ri;
#pragma unroll
                for (int i = q + 1; i < DP; ++i) {
                    double sacc = b[i][q];
#pragma unroll
                    for (int pp = 0; pp < q; ++pp) sacc = fma(-b[i][pp], b[q][pp], sacc);
                    b[i][q] = sacc * ri;
                }
            }
#pragma unroll
            for (int i = 0; i < DP; ++i)
#pragma unroll
                for (int q = 0; q <= i; ++q) Bd[i * (DP + 1) + q] = b[i][q];
        }
        __syncthreads();
        if (t >= j0 && h == 0) {
            double* dst = S + t * DLD + j0;
            if (t < j0 + DP) {  // a row of the diagonal block: the factor of the 8x8 block itself
                for (int q = 0; q <= t - j0; ++q) dst[q] = Bd[(t - j0) * (DP + 1) + q];
            } else {
                double x[DP];
#pragma unroll
                for (int q = 0; q < DP; ++q) {
                    double sacc = v[q];
#pragma unroll
                    for (int pp = 0; pp < q; ++pp) sacc = fma(-x[pp], Bd[q * (DP + 1) + pp], sacc);
                    x[q] = sacc * rd8[q];
                }
#pragma unroll
                for (int q = 0; q < DP; ++q) dst[q] = x[q];
            }
        }
        __syncthreads();  // panel final: the rows below are complete through column j0 + 7
    }
    for (int e = tid; e < DB * DB; e += DIAG_THREADS) {
        const int i = e >> 7, j = e & 127;
        A[(long)i * ld + j] = S[i * DLD + j];  // upper part already 0
    }
    {
        // the lanes of a warp (8 columns c0..c0+7) walk the same rows; lane (c, h) takes k = c0 + h, c0 + h + 4, ...
        // of the warp's common range [c0, i0) and adds nothing where its column starts below row k (c > k)
        const int c = t, c0 = (tid & ~31) / DT;
        for (int i0 = c0; i0 < DB; i0 += DP) {
            double y[DP];
#pragma unroll
            for (int q = 0; q < DP; ++q) y[q] = 0.0;
            const double* lb = S + i0 * DLD;
            for (int k = c0 + h; k < i0; k += DT) {
                const double x = (k >= c) ? Xp[k * (k + 1) / 2 + c] : 0.0;
#pragma unroll
                for (int q = 0; q < DP; ++q) y[q] = fma(lb[q * DLD + k], x, y[q]);
            }
            double xs[DP];
#pragma unroll
            for (int q = 0; q < DP; ++q) {
                const int i = i0 + q;
                double sacc = ((i == c) ? 1.0 : 0.0) - diag_group_sum(y[q]);
#pragma unroll
                for (int pp = 0; pp < q; ++pp) sacc = fma(-lb[q * DLD + i0 + pp], xs[pp], sacc);
                xs[q] = (i >= c) ? sacc * rdiag[i] : 0.0;
            }
            if (h == 0) {
#pragma unroll
                for (int q = 0; q < DP; ++q)
                    if (i0 + q >= c) Xp[(i0 + q) * (i0 + q + 1) / 2 + c] = xs[q];
            }
            __syncwarp();
        }
    }
    __syncthreads();
    double* D = Dinv + (long)r * strideD + (long)kb * DB * DB;
    for (int e = tid; e < DB * DB; e += DIAG_THREADS) {
        const int i = e >> 7, j = e & 127;
        D[e] = (j <= i) ? Xp[i * (i + 1) / 2 + j] : 0.0;
    }
}

// identity (NP x NP) per batch member
__global__ void set_identity_kernel(double* __restrict__ M, int NP) {
    const long r = blockIdx.y;
    const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long)NP * NP) return;
    const int i = (int)(e / NP), j = (int)(e - (long)i * NP);
    M[r * NP * NP + e] = (i == j) ? 1.0 : 0.0;
}

// diag(L) and sum(log diag L) per batch member
__global__ void logdet_kernel(const double* __restrict__ L, int NP, double* __restrict__ Ldiag,
                              double* __restrict__ logdet) {
    __shared__ double red[32];
    const int r = blockIdx.x;
    const double* Lr = L + (long)r * NP * NP;
    double s = 0.0;
    for (int i = threadIdx.x; i < NP; i += blockDim.x) {
        const double v = Lr[(long)i * NP + i];
        if (Ldiag) Ldiag[(long)r * NP + i] = v;
        s += log(v);
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        logdet[r] = t;
    }
}

// z = Linv * y  (row i: sum_{k<=i}); one warp per row
__global__ void __launch_bounds__(256) trmv_lower_kernel(const double* __restrict__ Linv, int NP,
                                                         const double* __restrict__ y, long strideY,
                                                         double* __restrict__ z) {
    const int r = blockIdx.y;
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= NP) return;
    const double* row = Linv + (long)r * NP * NP + (long)i * NP;
    const double* yr = y + (long)r * strideY;
    double s = 0.0;
    for (int k = lane; k <= i; k += 32) s += row[k] * yr[k];
    s = warp_sum(s);
    if (lane == 0) z[(long)r * NP + i] = s;
}

// alpha = Linv^T * z (column i: sum_{k>=i}); 64 columns per CTA, 4 k-groups
__global__ void __launch_bounds__(256) trmv_lower_t_kernel(const double* __restrict__ Linv, int NP,
                                                           const double* __restrict__ z,
                                                           double* __restrict__ alpha) {
    __shared__ double red[4][64];
    const int r = blockIdx.y;
    const int c = threadIdx.x & 63, kg = threadIdx.x >> 6;
    const int i = blockIdx.x * 64 + c;
    const double* M = Linv + (long)r * NP * NP;
    const double* zr = z + (long)r * NP;
    double s = 0.0;
    for (int k = blockIdx.x * 64 + kg; k < NP; k += 4) s += M[(long)k * NP + i] * zr[k];
    red[kg][c] = s;
    __syncthreads();
    if (kg == 0) alpha[(long)r * NP + i] = (red[0][c] + red[1][c]) + (red[2][c] + red[3][c]);
}

// lml[r] = -0.5 y.alpha - logdet - n/2 log(2 pi)   (sklearn:_gpr.py:613-617); -inf if info[r]
__global__ void lml_kernel(const double* __restrict__ y, long strideY, const double* __restrict__ alpha, int NP,
                           int n, const double* __restrict__ logdet, const int* __restrict__ info,
                           double* __restrict__ lml) {
    __shared__ double red[32];
    const int r = blockIdx.x;
    double s = 0.0;
    for (int i = threadIdx.x; i < NP; i += blockDim.x) s += y[(long)r * strideY + i] * alpha[(long)r * NP + i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        double v = -0.5 * t;
        v -= logdet[r];
        v -= n / 2.0 * log(2 * 3.141592653589793);
        lml[r] = info[r] ? -INFINITY : v;
    }
}

// ---------------------------------------------------------------------------------------------
// gradient trace: partial[r][tile][q] over lower 64x64 tiles of W = alpha alpha^T - K^-1
// ---------------------------------------------------------------------------------------------
// g(r2) and h(r2) of common.cuh's shape_fn / shape_grad_fn from ONE sqrt and ONE exp
template <int KIND>
__device__ __forceinline__ void shape_pair(double r2, double& g, double& h) {
    if (KIND == TB200_RBF) {
        g = h = exp_neg(0.5 * r2);
    } else if (KIND == TB200_MATERN12) {
        const double r = sqrt_pos(r2);
        g = exp_neg(r);
        h = r2 != 0.0 ? g / r : 0.0;
    } else if (KIND == TB200_MATERN32) {
        const double t = sqrt_pos(3.0 * r2);  // shape_grad_fn's form; shape_fn's r*sqrt(3) differs by <= 1 ulp
        const double e = exp_neg(t);
        g = (1.0 + t) * e;
        h = 3.0 * e;
    } else {
        const double t = sqrt_pos(5.0 * r2);
        const double e = exp_neg(t);
        g = (1.0 + t + t * t * (1.0 / 3.0)) * e;
        h = 5.0 / 3.0 * (t + 1.0) * e;
    }
}

// Each thread owns a 4x4 block of pairs.  Pass 1: r2 and the pair weights f = 2 w sf2 h(r2) stay in registers;
// pass 2 walks the dimensions ONCE MORE as the outer loop, so the per-dimension sums never live in a dynamically
// indexed (= local-memory) array: one scalar per dimension is reduced over the warp and parked in shared memory.
template <int KIND>
__global__ void __launch_bounds__(256) lml_grad_kernel(const double* __restrict__ XsT, int n, int NP,
                                                       tb200_kernel_spec spec, const double* __restrict__ theta,
                                                       int p, const double* __restrict__ Kinv,
                                                       const double* __restrict__ alpha,
                                                       double* __restrict__ partial, int ntiles) {
    extern __shared__ double sh[];  // xi[d][64], xj[d][64], red[8][p]
    const int d = spec.d;
    double* xi = sh;
    double* xj = sh + d * GT;
    double* red = sh + 2 * d * GT;
    const int r = blockIdx.y;
    int ti, tj;
    lower_tile(blockIdx.x, ti, tj);
    const double* xs = XsT + (long)r * d * NP;
    for (int e = threadIdx.x; e < d * GT; e += blockDim.x) {
        const int k = e / GT, c = e - k * GT;
        xi[e] = xs[(long)k * NP + ti * GT + c];
        xj[e] = xs[(long)k * NP + tj * GT + c];
    }
    __syncthreads();
    const Hyper hy = unpack_hyper(spec, theta + (long)r * p);
    const double* Kr = Kinv + (long)r * NP * NP;
    const double* al = alpha + (long)r * NP;
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nl = spec.ard ? d : 1;
    const int off_ls = spec.has_const ? 1 : 0;
    const int i0 = ti * GT + ty * 4, j0 = tj * GT + tx;  // columns j0 + 16 b (see gram_train_kernel)

    // w = alpha_i alpha_j - Kinv_ij for the valid lower pairs (loads first: they overlap the distance loop)
    double w[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int i = i0 + a;
        const double ai = i < n ? al[i] : 0.0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int j = j0 + 16 * b;
            w[a][b] = (i < n && j <= i) ? ai * al[j] - Kr[(long)i * NP + j] : 0.0;
        }
    }
    double r2[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) r2[a][b] = 0.0;
    for (int k = 0; k < d; ++k) {
        double xa[4], xb[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { xa[a] = xi[k * GT + ty * 4 + a]; xb[a] = xj[k * GT + tx + 16 * a]; }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const double df = xa[a] - xb[b];
                r2[a][b] += df * df;
            }
    }
    double g_const = 0.0, g_noise = 0.0, g_iso = 0.0;
    double f[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int i = i0 + a, j = j0 + 16 * b;
            f[a][b] = 0.0;
            if (!(i < n && j <= i)) continue;
            if (i == j) {
                g_const += w[a][b] * hy.sf2;  // dK/dlog(const) on the diagonal = sf2 * 1
                g_noise += w[a][b] * hy.noise;
                continue;
            }
            double g, h;
            shape_pair<KIND>(r2[a][b], g, h);
            const double w2 = 2.0 * w[a][b];  // (i,j) and (j,i)
            g_const += w2 * hy.sf2 * g;
            f[a][b] = w2 * hy.sf2 * h;
            g_iso += f[a][b] * r2[a][b];
        }
    // warp sums in a fixed order, one slot per (warp, hyper-parameter)
    {
        const double vc = warp_sum(g_const), vn = warp_sum(g_noise);
        if (lane == 0) {
            if (spec.has_const) red[warp * p] = vc;
            if (spec.has_white) red[warp * p + off_ls + nl] = vn;
        }
    }
    if (spec.ard) {
        for (int k = 0; k < d; ++k) {
            double xa[4], xb[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) { xa[a] = xi[k * GT + ty * 4 + a]; xb[a] = xj[k * GT + tx + 16 * a]; }
            double acc = 0.0;
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const double df = xa[a] - xb[b];
                    acc += f[a][b] * (df * df);
                }
            acc = warp_sum(acc);
            if (lane == 0) red[warp * p + off_ls + k] = acc;
        }
    } else {
        const double v = warp_sum(g_iso);
        if (lane == 0) red[warp * p + off_ls] = v;
    }
    __syncthreads();
    if (threadIdx.x < p) {
        double t = 0.0;
        for (int wq = 0; wq < 8; ++wq) t += red[wq * p + threadIdx.x];
        partial[((long)r * ntiles + blockIdx.x) * p + threadIdx.x] = t;
    }
}

__global__ void lml_grad_reduce_kernel(const double* __restrict__ partial, int ntiles, int p,
                                       const int* __restrict__ info, double* __restrict__ grad) {
    __shared__ double red[32];
    const int r = blockIdx.x, q = blockIdx.y;
    double s = 0.0;
    for (int t = threadIdx.x; t < ntiles; t += blockDim.x) s += partial[((long)r * ntiles + t) * p + q];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        grad[(long)r * p + q] = info[r] ? 0.0 : 0.5 * t;
    }
}

__global__ void pad_vector_kernel(const double* __restrict__ y, int n, int NP, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < NP) out[i] = (i < n) ? y[i] : 0.0;
}

// ---------------------------------------------------------------------------------------------
// INT8 engine of the batched fit (ozaki_gemm.cuh): scales and byte slicing of L and L^-1
// ---------------------------------------------------------------------------------------------
// balanced s8 digits of a, |a| <= 0.98:  a ~= sum_{s=0..6} b_s 2^(-7-8s), b_s in [-128, 127].
// X = rn(a 2^55) is a 56-bit signed integer.  Its balanced base-256 digits obey digit = (int8)(low byte),
// X <- (X + 128) >> 8; adding 128 at every byte position at once propagates exactly those carries, so byte j of
// (X + 0x80..80) ^ 0x80..80 is the two's-complement digit of weight 2^(8j): digit s of the GEMM is byte 6 - s.
// Four FP64 operations per element and no conversion (common.cuh fixed_point_64); the per-digit FP64 round / subtract
// chain this replaces made the slicing kernels FP64-pipe bound (7 x 5 FP64-class operations per element).
__device__ __forceinline__ unsigned long long fit_slice_signed(double a) {
    const long long x = fixed_point_64<23>(a);  // rn(a 2^55), rounded once, no conversion instruction
    return ((unsigned long long)x + 0x0080808080808080ull) ^ 0x0080808080808080ull;
}

__device__ __forceinline__ double pow2_cover(double mx) {
    if (!(mx > 0.0) || isinf(mx)) return 1.0;  // empty / failed member: any scale
    int e = 0;
    const double f = frexp(mx, &e);
    if (f > 0.98) ++e;
    return ldexp(1.0, e);
}

// rows of L: |L[r,k]| <= sqrt(K[r,r]) (sum_k L[r,k]^2 = K[r,r]), known from the Gram diagonal before factoring
__global__ void rowscale_from_diag_kernel(const double* __restrict__ K, int NP, double* __restrict__ scale) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x, rs = blockIdx.y;
    if (r >= NP) return;
    const double dg = K[(size_t)rs * NP * NP + (size_t)r * NP + r];
    scale[(size_t)rs * NP + r] = pow2_cover(sqrt(fabs(dg)));
}

// One warp = one (8 operand rows) x (64 of K) piece = 4 core matrices per slice, written as 128 contiguous bytes
// per core matrix.  TRANSPOSED = false: operand row = matrix row, K = matrix column (rows of L / of X = L^-1);
// TRANSPOSED = true: operand row = matrix column, K = matrix row (columns of L^-1, columns of an L panel).
// Pieces are enumerated over operand row groups [rg0, rg1) x k-blocks [kb0, kb1); with `lower` set, pieces whose
// k-block lies entirely before the 128-row block of the operand rows are skipped (never read by the tri GEMM).
// Operand-row scale: `scale` (power-of-two doubles) or 1.  tileRow0: operand row of tile 0 of `out`.
template <bool TRANSPOSED>
__global__ void __launch_bounds__(256) slice_operand_kernel(const double* __restrict__ M, int NP, long mStride,
                                                            const double* __restrict__ scale,
                                                            int rg0, int rg1, int kb0, int kb1, int lower, int tileRow0,
                                                            uint8_t* __restrict__ out, long sliceStride,
                                                            long restartStride, int kblks) {
    const int lane = threadIdx.x & 31;
    const long piece = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int nkb = kb1 - kb0;
    if (piece >= (long)(rg1 - rg0) * nkb) return;
    const int rg = rg0 + (int)(piece / nkb), kb = kb0 + (int)(piece % nkb);
    if (lower && kb < 2 * (rg / 16)) return;
    const int rs = blockIdx.y;
    const int orow = rg * 8 + (lane & 7);       // operand row
    const int k0 = kb * 64 + (lane >> 3) * 16;  // first of this lane's 16 K indices
    const double* src = M + (size_t)rs * mStride;
    const double inv = scale ? 1.0 / scale[(size_t)rs * NP + orow] : 1.0;  // exact: power of two
    uint32_t lo[16], hi[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const double v = TRANSPOSED ? src[(size_t)(k0 + q) * NP + orow] : src[(size_t)orow * NP + k0 + q];
        const unsigned long long y = fit_slice_signed(v * inv);
        lo[q] = (uint32_t)y;
        hi[q] = (uint32_t)(y >> 32);
    }
    const size_t off = (size_t)rs * restartStride + ((size_t)((orow - tileRow0) / 64) * kblks + kb) * i8::B_TILE +
                       (size_t)((orow % 64) >> 3) * i8::SBO + (lane >> 3) * i8::LBO + (lane & 7) * 16;
#pragma unroll
    for (int sl = 0; sl < i8::S; ++sl) {
        // digit sl = byte (6 - sl) of the value: bytes 2,1,0 of hi, then 3,2,1,0 of lo
        const int byte = (sl < 3) ? 2 - sl : 6 - sl;
        uint32_t w[4];
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const uint32_t* srcw = (sl < 3) ? hi : lo;
            const uint32_t p01 = __byte_perm(srcw[4 * v], srcw[4 * v + 1], byte | ((4 + byte) << 4));
            const uint32_t p23 = __byte_perm(srcw[4 * v + 2], srcw[4 * v + 3], byte | ((4 + byte) << 4));
            w[v] = __byte_perm(p01, p23, 0x5410);
        }
        *reinterpret_cast<uint4*>(out + (size_t)sl * sliceStride + off) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---------------------------------------------------------------------------------------------
// Recursive inverse of the factor (trtri_batched_i8): helpers.  Maxima are kept as upper bounds of the HIGH WORD of a
// non-negative double (order-preserving as an unsigned integer: atomicMax / redux work on 32 bits); the GEMM epilogues
// update them (ozaki_gemm.cuh), these kernels seed them and turn them into power-of-two scales.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned abs_hi_word(double v) {
    const unsigned hw = (unsigned)__double2hiint(v) & 0x7fffffffu;
    return hw >= 0x7ff00000u ? 0u : hw;  // NaN / Inf never enters a scale
}

// Linv diagonal tiles <- Dinv blocks (everything else of Linv was zeroed), and the row / column maxima of every block
// (the level-0 operands of the recursion).  256 threads: thread = column j (two threads per column), rows alternate.
__global__ void __launch_bounds__(256) place_diag_blocks_max_kernel(const double* __restrict__ Dinv, long strideD,
                                                                    double* __restrict__ Linv, int NP,
                                                                    unsigned* __restrict__ rowmax32,
                                                                    unsigned* __restrict__ colmax32) {
    const int kb = blockIdx.x, r = blockIdx.y, lane = threadIdx.x & 31;
    const double* D = Dinv + (long)r * strideD + (long)kb * TILE * TILE;
    double* out = Linv + (long)r * NP * NP + (long)kb * TILE * NP + (long)kb * TILE;
    unsigned cm = 0;
    for (int e = threadIdx.x; e < TILE * TILE; e += 256) {
        const int i = e >> 7, j = e & 127;
        const double v = D[e];
        out[(long)i * NP + j] = v;
        const unsigned hw = abs_hi_word(v);
        cm = hw > cm ? hw : cm;
        const unsigned rm = __reduce_max_sync(0xffffffffu, hw);  // the warp holds 32 consecutive columns of row i
        if (lane == 0 && rm > 0) atomicMax(rowmax32 + (size_t)r * NP + kb * TILE + i, rm + 1);
    }
    if (cm > 0) atomicMax(colmax32 + (size_t)r * NP + kb * TILE + (threadIdx.x & 127), cm + 1);
}

// power-of-two covers of high-word maxima: out[e] = 2^k with 0.98 * 2^k >= max (1 where nothing was recorded)
__global__ void cover32_kernel(const unsigned* __restrict__ hi32, long total, double* __restrict__ out) {
    const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < total) out[e] = pow2_cover(__hiloint2double((int)hi32[e], 0));
}

// Digit planes of the operands of one recursion level (pairs of diagonal blocks A, C of h64 64-tiles, see
// GemmI8Params::recur), written into the ONE plane buffer of the fit at tile (operand row / 64, k-block):
//   MODE 0: rows of inv(C)      operand row = matrix row i in C,    k in C, k-block <= 128-row tile of i  scale = row cover
//   MODE 1: columns of inv(A)   operand row = matrix column j in A, k in A, k-block >= tile of j   scale = column cover
//   MODE 2: columns of T        operand row = matrix column j in A, k in C                         scale = cover of T's columns
// The three regions and the planes of the rows of L (strictly below the diagonal blocks, written by the factorisation
// and read by the first product of the level) never overlap within a level: rows of L and of inv(C) lie below or on
// the diagonal, column operands above or on it, and A and C are different diagonal blocks.
// One warp = 8 operand rows x 64 of K, as in slice_operand_kernel.
template <int MODE>
__global__ void __launch_bounds__(256) slice_level_kernel(const double* __restrict__ M, int NP, long mStride,
                                                          const double* __restrict__ scale, int h64, int npairs,
                                                          uint8_t* __restrict__ out, long sliceStride,
                                                          long restartStride, int kblks) {
    const int lane = threadIdx.x & 31;
    const long piece = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const long per = (long)(8 * h64) * h64;  // row groups x k-blocks of one pair
    if (piece >= per * npairs) return;
    const int pr = (int)(piece / per), q = (int)(piece - (long)pr * per);
    const int lrg = q / h64, lkb = q - lrg * h64;
    const int a0 = 2 * pr * h64, c0 = a0 + h64;
    const int rg = (MODE == 0 ? c0 : a0) * 8 + lrg;  // operand row group (8 rows)
    const int kb = (MODE == 1 ? a0 : c0) + lkb;
    if (rg * 8 >= NP || kb >= kblks) return;          // C cut short by the end of the matrix
    // inv(C) is lower triangular; the GEMM reads its A operand in 128-row tiles up to the k-block of the tile's LAST
    // row, so the zero 64x64 piece right of the diagonal in the upper half of a tile is sliced too
    if (MODE == 0 && kb > 2 * (rg / 16) + 1) return;
    if (MODE == 1 && kb < rg / 8) return;
    const int rs = blockIdx.y;
    const int orow = rg * 8 + (lane & 7);
    const int k0 = kb * 64 + (lane >> 3) * 16;
    const double* src = M + (size_t)rs * mStride;
    const double inv = 1.0 / scale[(size_t)rs * NP + orow];  // exact: power of two
    uint32_t lo[16], hi[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) {
        const double v = MODE == 0 ? src[(size_t)orow * NP + k0 + u] : src[(size_t)(k0 + u) * NP + orow];
        const unsigned long long y = fit_slice_signed(v * inv);
        lo[u] = (uint32_t)y;
        hi[u] = (uint32_t)(y >> 32);
    }
    const size_t off = (size_t)rs * restartStride + ((size_t)(orow / 64) * kblks + kb) * i8::B_TILE +
                       (size_t)((orow % 64) >> 3) * i8::SBO + (lane >> 3) * i8::LBO + (lane & 7) * 16;
#pragma unroll
    for (int sl = 0; sl < i8::S; ++sl) {
        const int byte = (sl < 3) ? 2 - sl : 6 - sl;
        uint32_t w[4];
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const uint32_t* srcw = (sl < 3) ? hi : lo;
            const uint32_t p01 = __byte_perm(srcw[4 * v], srcw[4 * v + 1], byte | ((4 + byte) << 4));
            const uint32_t p23 = __byte_perm(srcw[4 * v + 2], srcw[4 * v + 3], byte | ((4 + byte) << 4));
            w[v] = __byte_perm(p01, p23, 0x5410);
        }
        *reinterpret_cast<uint4*>(out + (size_t)sl * sliceStride + off) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---------------------------------------------------------------------------------------------
// incremental refit: append one training point to a factor at fixed hyper-parameters
//   L' = [[L, 0], [l^T, lam]],  l = L^-1 k,  lam^2 = k(x,x) + jitter - l.l
//   L'^-1 = [[L^-1, 0], [-(L^-T l)^T / lam, 1 / lam]]
// The padded row n_old of Linv (identity) becomes the new row; O(n^2) instead of the O(n^3) refactorisation.
// ---------------------------------------------------------------------------------------------
__global__ void kcol_kernel(const double* __restrict__ x_new, const double* __restrict__ XsT, int n, int NP,
                            tb200_kernel_spec spec, const double* __restrict__ theta, double* __restrict__ kv) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= NP) return;
    double v = 0.0;
    if (j < n) {
        double r2 = 0.0;
        for (int k = 0; k < spec.d; ++k) {
            const double df = x_new[k] / length_scale(spec, theta, k) - XsT[(long)k * NP + j];
            r2 += df * df;
        }
        v = unpack_hyper(spec, theta).sf2 * shape_fn(spec.kind, r2);
    }
    kv[j] = v;
}

__global__ void __launch_bounds__(256) append_row_kernel(const double* __restrict__ x_new, const double* __restrict__ l,
                                                         const double* __restrict__ u, int n_old, int NP,
                                                         tb200_kernel_spec spec, const double* __restrict__ theta,
                                                         double jitter, double* __restrict__ XsT,
                                                         double* __restrict__ Linv, double* __restrict__ Ldiag,
                                                         int* __restrict__ info) {
    __shared__ double red[8];
    __shared__ double lam_s;
    double s = 0.0;
    for (int j = threadIdx.x; j < n_old; j += blockDim.x) s = fma(l[j], l[j], s);
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        const Hyper h = unpack_hyper(spec, theta);
        const double lam2 = (h.sf2 + h.noise + jitter) - t;  // diagonal of K as gram_train_kernel writes it
        const bool ok = lam2 > 0.0;                             // false for NaN too
        lam_s = ok ? sqrt(lam2) : 0.0;
        *info = ok ? 0 : n_old + 1;                             // first failing pivot, 1-based (LAPACK convention)
    }
    __syncthreads();
    const double lam = lam_s;
    if (lam == 0.0) return;
    double* row = Linv + (long)n_old * NP;
    for (int j = threadIdx.x; j < n_old; j += blockDim.x) row[j] = -u[j] / lam;
    if (threadIdx.x == 0) {
        row[n_old] = 1.0 / lam;
        Ldiag[n_old] = lam;
    }
    for (int k = threadIdx.x; k < spec.d; k += blockDim.x)
        XsT[(long)k * NP + n_old] = x_new[k] / length_scale(spec, theta, k);
}

__global__ void logdet_from_diag_kernel(const double* __restrict__ Ldiag, int n, double* __restrict__ logdet) {
    __shared__ double red[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += log(Ldiag[i]);
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        *logdet = t;
    }
}

// ---------------------------------------------------------------------------------------------
// host drivers
// ---------------------------------------------------------------------------------------------
struct FitBuffers {
    double* K;       // [R][NP][NP]   Gram -> L -> (K^-1)
    double* Linv;    // [R][NP][NP]
    double* Dinv;    // [R][NB][128][128]
    double* XsT;     // [R][d][NP]
    double* ypad;    // [NP]
    double* z;       // [R][NP]
    double* alpha;   // [R][NP]
    double* logdet;  // [R]
    double* partial; // [R][ntiles64][p]
    uint8_t* sliced; // [R][7][NP/64][NP/64][4096]  byte slices of L rows, then of L^-1 columns (INT8 engine)
    double* scale;   // [R][NP] power-of-two row / column scales of the sliced operand
    unsigned long long* tmax;  // [R][NP] 8 bytes each: running column and row maxima of the inverse (2 x 32-bit high words)
    unsigned long long* smax;  // [R][NP] column maxima of T of the current recursion level (32-bit high words)
    double* sigma;   // [R][NP] power-of-two column scales of T
    double* colcov;  // [R][NP] power-of-two column scales of the inverse factor so far (recursive inverse)
    double* rowcov;  // [R][NP] ... row scales
    mutable i8::SchedPool sched;  // scheduler counters of the persistent INT8 GEMM launches (zeroed per fit call)
};

// slice products kept by the fit GEMMs, per stage: 9 (34 products, truncation 2^-64: below FP64 rounding), 8 (28
// products, 2^-56: the size of FP64 rounding itself) or 7 (21 products, 2^-48).  Defaults: Cholesky column update 8 (its
// error is amplified by the rest of the factorisation: 7 costs three digits of the LML on ill-conditioned members,
// 3.4e-9), inverse factor 7 (recursive inverse, round 2: worst LML difference 5e-11, gradient 3.9e-10 of max|grad| on
// the hard members), K^-1 = Linv^T Linv 7 (feeds only the gradient: 1.6e-10 of max|grad|).  Measured against the FP64
// DMMA path in profiles/r1_fit_precision.md and profiles/r2_fit_precision.md (parity bars: 1e-6).
// TB200_DEBUG=fit_gmax=.. sets all three stages, fit_gmax_trtri / fit_gmax_kinv one stage (common.cuh).
// FIT_TRTRI_MODEL: the inverse factor of a model that will PREDICT (tb200_factor, or any LML-only evaluation): never below
// 28 products.  With 21 the predictive mean of an ill-conditioned model (n = 12400, d = 3) moved by 2.8e-7 against the
// FP64 DMMA path (tests/test_gpu_i8.py); the 21-product inverse is for the optimiser's LML+gradient evaluations only.
enum FitStage { FIT_POTRF = 0, FIT_TRTRI = 1, FIT_KINV = 2, FIT_TRTRI_MODEL = 3 };
static int fit_gmax(int stage) {
    const DebugConfig& c = debug_config();
    if (stage == FIT_TRTRI_MODEL) return c.fit_gmax[FIT_TRTRI] < 8 ? 8 : c.fit_gmax[FIT_TRTRI];
    return c.fit_gmax[stage];
}
static int fit_sm_count() {  // CTAs of a persistent grid: one per SM of the current device
    static int sms[MAX_DEVICES] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return 0;
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        sms[dev] = v;  // racing threads write the same value
    }
    return sms[dev];
}
static int launch_fit_gemm_i8(const FitBuffers& fb, const i8::GemmI8Params& g, int ntiles, int R, cudaStream_t st,
                              int stage) {
    const int sms = fit_sm_count();
    switch (fit_gmax(stage)) {
        case 7: return i8::launch_gemm_i8<7>(g, ntiles, R, fb.sched, sms, st);
        case 8: return i8::launch_gemm_i8<8>(g, ntiles, R, fb.sched, sms, st);
        default: return i8::launch_gemm_i8<9>(g, ntiles, R, fb.sched, sms, st);
    }
}

static size_t align256(size_t b) { return (b + 255) / 256 * 256; }

static bool fit_dmma_forced() { return debug_config().fit_dmma; }

static bool use_left_looking(int NB, int R) {
    // R * (n/128) at which the batched schedule takes over: measured at n = 4096, the batched INT8 schedule wins from
    // R = 4 (profiles/r1_fit_schedule_threshold.md).  One to three matrices: a block column must offer about one CTA
    // per SM by itself (n >= 12288 for a single matrix) and the products must be on the INT8 engine; otherwise the
    // right-looking schedule has more tiles per step
    return R >= 4 ? (long)R * NB >= 128 : (!fit_dmma_forced() && (long)R * NB >= 96);
}

// The batched (left-looking) schedule runs its long-K products on the INT8 tensor cores (k-ranges above 8192 in
// several exact INT32 windows, ozaki_gemm.cuh).  TB200_DEBUG=fit_engine=dmma keeps everything on the FP64 DMMA GEMM.
static bool fit_uses_i8(int NP, int R) {
    return !fit_dmma_forced() && NP <= TB200_I8_MAX_N && use_left_looking(NP / TILE, R);
}

size_t fit_workspace_layout(int n, int d, int R, int p, bool own_linv, bool own_xst, bool own_alpha, FitBuffers* fb,
                            char* base) {
    const size_t NP = pad_n(n), NB = NP / TILE;
    const size_t nt64 = (NP / GT) * (NP / GT + 1) / 2;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char* ptr = base ? base + off : nullptr;
        off += align256(bytes);
        return reinterpret_cast<double*>(ptr);
    };
    double* K = take(sizeof(double) * R * NP * NP);
    double* Linv = own_linv ? take(sizeof(double) * R * NP * NP) : nullptr;
    double* Dinv = take(sizeof(double) * R * NB * TILE * TILE);
    double* XsT = own_xst ? take(sizeof(double) * R * d * NP) : nullptr;
    double* ypad = take(sizeof(double) * NP);
    double* z = take(sizeof(double) * R * NP);
    double* alpha = own_alpha ? take(sizeof(double) * R * NP) : nullptr;
    double* logdet = take(sizeof(double) * R);
    double* partial = take(sizeof(double) * R * nt64 * (p > 0 ? p : 1));
    uint8_t* sliced = nullptr;
    double* scale = nullptr;
    const bool i8fit = fit_uses_i8((int)NP, R);
    if (i8fit) {
        sliced = reinterpret_cast<uint8_t*>(take((size_t)i8::S * R * NP * NP));
        scale = take(sizeof(double) * R * NP);
    }
    unsigned long long* tmax = nullptr; unsigned long long* smax = nullptr; double* sigma = nullptr;
    unsigned* sched = nullptr;
    double* colcov = nullptr; double* rowcov = nullptr;
    if (i8fit) {
        tmax = reinterpret_cast<unsigned long long*>(take(sizeof(double) * R * NP));
        smax = reinterpret_cast<unsigned long long*>(take(sizeof(double) * R * NP));
        sigma = take(sizeof(double) * R * NP);
        sched = reinterpret_cast<unsigned*>(take(sizeof(unsigned) * 2 * i8::SCHED_POOL));
        colcov = take(sizeof(double) * R * NP);
        rowcov = take(sizeof(double) * R * NP);
    }
    if (fb) {
        fb->sched.base = sched; fb->sched.next = 0;
        fb->colcov = colcov; fb->rowcov = rowcov;
        fb->sliced = sliced; fb->scale = scale; fb->tmax = tmax; fb->smax = smax; fb->sigma = sigma;
        fb->K = K; fb->Linv = Linv; fb->Dinv = Dinv; fb->XsT = XsT; fb->ypad = ypad; fb->z = z;
        fb->alpha = alpha; fb->logdet = logdet; fb->partial = partial;
    }
    return off;
}

// Two schedules of the same blocked algorithms, chosen by how many independent matrices there are:
//  * several matrices (batched restarts, R >= 4 and R * n/128 >= 128): LEFT-looking.  Every tile accumulates its whole update in
//    registers over a long K range and is written once, so the GEMMs run at long-K efficiency and
//    each tile of C is read/written O(1) times instead of once per 128-wide step;
//  * few matrices (the final factor, n large): RIGHT-looking, which exposes (NB-k)^2/2 tiles per
//    step to fill 148 SMs from a single matrix.

static int launch_diag(const FitBuffers& fb, int NP, int R, int k, int* info, cudaStream_t st) {
    const int NB = NP / TILE;
    const long sK = (long)NP * NP, sD = (long)NB * TILE * TILE;
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(diag_potrf_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)DIAG_SMEM));
            return 0;
        }))
        return -1;
    TB_LAUNCH(diag_potrf_inv_kernel, dim3(R), dim3(DIAG_THREADS), DIAG_SMEM, st, fb.K, (long)NP, sK, k, fb.Dinv, sD, info);
    return 0;
}

// panel: A[i,k] <- A[i,k] * Dinv_k^T  (i > k), in place
static int launch_panel(const FitBuffers& fb, int NP, int R, int k, cudaStream_t st) {
    const int NB = NP / TILE, rem = NB - k - 1;
    const long sK = (long)NP * NP, sD = (long)NB * TILE * TILE;
    double* panel = fb.K + (long)(k + 1) * TILE * NP + (long)k * TILE;
    GemmParams p = gemm_params(panel, NP, fb.Dinv + (long)k * TILE * TILE, TILE, panel, NP, rem, 1, TILE, 1.0, 0.0);
    p.sA1 = sK; p.sB1 = sD; p.sC1 = sK;
    p.b_tri = 1;  // Dinv_k is lower triangular: C[i,j] = sum_{k<=j} A[i,k] Dinv[j,k]
    return launch_gemm<false, true>(p, R, 1, st);
}

// high-priority side stream of the batched fit (look-ahead of the diagonal block), one per device
struct FitStreams {
    bool ready = false;
    cudaStream_t side, look;
    cudaEvent_t fork, join, fork_look, look_done;
};
static FitStreams g_fit_streams[MAX_DEVICES];
static std::mutex g_fit_streams_mutex;
static int fit_streams_for_device(FitStreams** out) {
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CHECK(dev >= 0 && dev < MAX_DEVICES, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_fit_streams_mutex);
    FitStreams& f = g_fit_streams[dev];
    if (!f.ready) {
        int lo = 0, hi = 0;
        TB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        TB_CUDA(cudaStreamCreateWithPriority(&f.side, cudaStreamNonBlocking, hi));
        TB_CUDA(cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming));
        TB_CUDA(cudaEventCreateWithFlags(&f.join, cudaEventDisableTiming));
        TB_CUDA(cudaStreamCreateWithPriority(&f.look, cudaStreamNonBlocking, hi));
        TB_CUDA(cudaEventCreateWithFlags(&f.fork_look, cudaEventDisableTiming));
        TB_CUDA(cudaEventCreateWithFlags(&f.look_done, cudaEventDisableTiming));
        f.ready = true;
    }
    *out = &f;
    return 0;
}

// Update of diagonal tiles by ONE panel: A[j,j] -= L[j,k] L[j,k]^T for `count` block rows j = j0.. (one 128^3 DMMA
// product per tile and restart).  Look-ahead schedule of the batched INT8 fit: the long-K left-looking product of
// diagonal tile k+1 over the panels 0..k-1 runs on a third stream while block column k is still being worked on, so
// after panel k only this single-panel update and the latency-bound 128x128 factorisation remain on the critical path.
static int launch_diag_update(const FitBuffers& fb, int NP, int R, int k, int j0, int count, cudaStream_t st) {
    const long sK = (long)NP * NP;
    const double* rows = fb.K + (long)j0 * TILE * NP + (long)k * TILE;
    double* c = fb.K + (long)j0 * TILE * NP + (long)j0 * TILE;
    GemmParams p = gemm_params(rows, NP, rows, NP, c, NP, 1, 1, TILE, -1.0, 1.0);
    p.sA1 = sK; p.sB1 = sK; p.sC1 = sK;
    p.sA2 = (long)TILE * NP; p.sB2 = (long)TILE * NP; p.sC2 = (long)TILE * NP + TILE;
    return launch_gemm<false, true>(p, R, count, st);
}

// blocked Cholesky of R matrices in fb.K (lower), Dinv blocks kept for trtri
static int potrf_batched(const FitBuffers& fb, int NP, int R, int* info, cudaStream_t st) {
    const int NB = NP / TILE;
    const long sK = (long)NP * NP;
    const bool left = use_left_looking(NB, R);
    const bool i8fit = fit_uses_i8(NP, R) && fb.sliced != nullptr;
    if (i8fit)
        TB_LAUNCH(rowscale_from_diag_kernel, dim3((NP + 255) / 256, R), dim3(256), 0, st, fb.K, NP, fb.scale);
    FitStreams* rl = nullptr;  // side stream of the right-looking diagonal-tile schedule (batched INT8 fit only)
    if (i8fit && NB > 1 && fit_streams_for_device(&rl)) rl = nullptr;
    for (int k = 0; k < NB; ++k) {
        bool diag_done = false;
        if (rl && k > 0) {
            // The diagonal tile (k,k) was completed and factorised on the side stream right after panel k-1; only the
            // rows below take the long-K INT8 product here.
            if (NB - k > 1) {
                i8::GemmI8Params g{};
                g.S = fb.sliced; g.restartStride = (long)i8::S * sK; g.sliceStride = sK;
                g.scale = fb.scale; g.scaleStride = NP;
                g.C = fb.K; g.ldc = NP; g.cStride = sK;
                g.kblks = NP / i8::KB; g.tri = 0;
                g.rowTile0 = k + 1; g.colTile0 = 2 * k; g.nColTiles = 2; g.kbLo = 0; g.kbHi = 2 * k;
                g.alpha = -1.0; g.beta = 1.0;
                if (launch_fit_gemm_i8(fb, g, (NB - k - 1) * 2, R, st, FIT_POTRF)) return -1;
            }
            TB_CUDA(cudaStreamWaitEvent(st, rl->join, 0));
            diag_done = true;
        } else if (i8fit && k > 0) {
            // same block-column update on the INT8 tensor cores: rows of L are already sliced up to column k*128.
            // Look-ahead: the diagonal tile's update and its 128x128 factorisation (a latency-bound 64-CTA kernel) run
            // on a high-priority side stream while the main stream updates the rows below.
            i8::GemmI8Params g{};
            g.S = fb.sliced; g.restartStride = (long)i8::S * sK; g.sliceStride = sK;
            g.scale = fb.scale; g.scaleStride = NP;
            g.C = fb.K; g.ldc = NP; g.cStride = sK;
            g.kblks = NP / i8::KB; g.tri = 0;
            g.rowTile0 = k; g.colTile0 = 2 * k; g.nColTiles = 2; g.kbLo = 0; g.kbHi = 2 * k;
            g.alpha = -1.0; g.beta = 1.0;
            FitStreams* fs = nullptr;
            if (NB - k > 1 && !fit_streams_for_device(&fs)) {
                TB_CUDA(cudaEventRecord(fs->fork, st));
                TB_CUDA(cudaStreamWaitEvent(fs->side, fs->fork, 0));
                if (launch_fit_gemm_i8(fb, g, 2, R, fs->side, FIT_POTRF)) return -1;           // diagonal tile (it = k)
                if (launch_diag(fb, NP, R, k, info, fs->side)) return -1;
                TB_CUDA(cudaEventRecord(fs->join, fs->side));
                g.rowTile0 = k + 1;
                if (launch_fit_gemm_i8(fb, g, (NB - k - 1) * 2, R, st, FIT_POTRF)) return -1;  // rows below
                TB_CUDA(cudaStreamWaitEvent(st, fs->join, 0));
                diag_done = true;
            } else {
                if (launch_fit_gemm_i8(fb, g, (NB - k) * 2, R, st, FIT_POTRF)) return -1;
            }
        } else if (left && k > 0) {
            // block column k: A[i,k] -= L[i,0:k] L[k,0:k]^T  (i >= k), K = k*128 accumulated in registers
            const double* rows = fb.K + (long)k * TILE * NP;
            double* col = fb.K + (long)k * TILE * NP + (long)k * TILE;
            GemmParams p = gemm_params(rows, NP, rows, NP, col, NP, NB - k, 1, k * TILE, -1.0, 1.0);
            p.sA1 = sK; p.sB1 = sK; p.sC1 = sK;
            if (launch_gemm<false, true>(p, R, 1, st)) return -1;
        }
        if (!diag_done && launch_diag(fb, NP, R, k, info, st)) return -1;
        const int rem = NB - k - 1;
        if (rem == 0) break;
        if (launch_panel(fb, NP, R, k, st)) return -1;
        if (rl) {
            // side stream: add panel k to diagonal tile k+1 (panels 0..k-1 came from the look-ahead product) and
            // factorise it
            TB_CUDA(cudaEventRecord(rl->fork, st));
            TB_CUDA(cudaStreamWaitEvent(rl->side, rl->fork, 0));
            if (k >= 1) TB_CUDA(cudaStreamWaitEvent(rl->side, rl->look_done, 0));
            if (launch_diag_update(fb, NP, R, k, k + 1, 1, rl->side)) return -1;
            if (launch_diag(fb, NP, R, k + 1, info, rl->side)) return -1;
            TB_CUDA(cudaEventRecord(rl->join, rl->side));
        }
        if (i8fit) {
            // slice the finished panel L[(k+1)*128:, k*128:(k+1)*128] (k-blocks 2k, 2k+1 of those rows)
            const int rg0 = (k + 1) * (TILE / 8), rg1 = NP / 8;
            const long pieces = (long)(rg1 - rg0) * 2;
            TB_LAUNCH(slice_operand_kernel<false>, dim3((unsigned)((pieces + 7) / 8), R), dim3(256), 0, st, fb.K, NP, sK,
                      fb.scale, rg0, rg1, 2 * k, 2 * k + 2, 0, 0, fb.sliced, sK,
                      (long)i8::S * sK, NP / i8::KB);
        }
        if (rl && k + 2 < NB) {
            // look-ahead stream: diagonal tile k+2 -= L[k+2, 0:(k+1)*128] L[k+2, 0:(k+1)*128]^T (panels 0..k, all sliced)
            TB_CUDA(cudaEventRecord(rl->fork_look, st));
            TB_CUDA(cudaStreamWaitEvent(rl->look, rl->fork_look, 0));
            i8::GemmI8Params g{};
            g.S = fb.sliced; g.restartStride = (long)i8::S * sK; g.sliceStride = sK;
            g.scale = fb.scale; g.scaleStride = NP;
            g.C = fb.K; g.ldc = NP; g.cStride = sK;
            g.kblks = NP / i8::KB; g.tri = 0;
            g.rowTile0 = k + 2; g.colTile0 = 2 * (k + 2); g.nColTiles = 2; g.kbLo = 0; g.kbHi = 2 * (k + 1);
            g.alpha = -1.0; g.beta = 1.0;
            if (launch_fit_gemm_i8(fb, g, 2, R, rl->look, FIT_POTRF)) return -1;
            TB_CUDA(cudaEventRecord(rl->look_done, rl->look));
        }
        if (!left) {
            // trailing: A[i,j] -= A[i,k] A[j,k]^T  (k < j <= i)
            const double* panel = fb.K + (long)(k + 1) * TILE * NP + (long)k * TILE;
            double* trail = fb.K + (long)(k + 1) * TILE * NP + (long)(k + 1) * TILE;
            GemmParams p = gemm_params(panel, NP, panel, NP, trail, NP, rem, rem, TILE, -1.0, 1.0);
            p.order = ORDER_LOWER;
            p.sA1 = sK; p.sB1 = sK; p.sC1 = sK;
            if (launch_gemm<false, true>(p, R, 1, st)) return -1;
        }
    }
    return 0;
}

// Linv diagonal tiles <- Dinv blocks (everything else of Linv was zeroed)
__global__ void __launch_bounds__(256) place_diag_blocks_kernel(const double* __restrict__ Dinv, long strideD,
                                                                double* __restrict__ Linv, int NP) {
    const int kb = blockIdx.x, r = blockIdx.y;
    const double* D = Dinv + (long)r * strideD + (long)kb * TILE * TILE;
    double* out = Linv + (long)r * NP * NP + (long)kb * TILE * NP + (long)kb * TILE;
    for (int e = threadIdx.x; e < TILE * TILE; e += blockDim.x) {
        const int i = e >> 7, j = e & 127;
        out[(long)i * NP + j] = D[e];
    }
}

// Linv = L^-1 on the INT8 tensor cores by RECURSIVE DOUBLING (round 2; round 1 walked the block columns from the right:
// 31 dependent steps of 8 launches at n = 4096, most of them latency-bound, 6.0 of the 14.7 ms of an LML+gradient
// evaluation at 8 restarts).  The inverses of the 128x128 diagonal blocks are there (Dinv).  Level h = 128, 256, ...:
// the matrix is cut into pairs of neighbouring diagonal blocks A, C of size h (C possibly cut short by the matrix
// end) with B = L[C, A] between them, and
//        inv [[A, 0], [B, C]] = [[inv A, 0], [-inv(C) B inv(A), inv C]]
// so every pair needs two products,  T = B inv(A)  and  X = -inv(C) T,  and ALL pairs of a level and all restarts run in
// the same two launches: log2(n / 128) levels of ~8 launches instead of n / 128 steps.  The flop count is the one of
// the column recurrence (n^3 / 3): the triangular shape of inv(A) and inv(C) is used through the k-ranges.
// Operands (all in the ONE plane buffer, slice_level_kernel): rows of L -- sliced by the factorisation with their a-priori
// scale sqrt(K_rr), each block B is read by exactly one level; columns of inv(A), rows of inv(C), columns of T -- sliced
// per level with the power-of-two cover of their current maxima, which the GEMM epilogues keep up to date (running row
// and column maxima of the inverse; T's column maxima per level), so no pass over the matrix measures anything.
// All scalings are powers of two: nothing is rounded outside the kept slice pairs, as in ozaki_gemm.cuh.
// On exit fb.scale holds the column covers of Linv -- the operand scaling of the K^-1 = Linv^T Linv step that follows.
static int trtri_batched_i8(const FitBuffers& fb, double* Linv, int NP, int R, int stage, cudaStream_t st) {
    const int NB = NP / TILE, kblks = NP / i8::KB;
    const long sK = (long)NP * NP, sD = (long)NB * TILE * TILE;
    const long total = (long)R * NP;
    // high-word maxima: columns / rows of the inverse so far, columns of T (the two 8-byte arrays of round 1 hold them)
    unsigned* colmax32 = reinterpret_cast<unsigned*>(fb.tmax);
    unsigned* rowmax32 = colmax32 + total;
    unsigned* tmax32 = reinterpret_cast<unsigned*>(fb.smax);
    double* tcov = fb.sigma;
    TB_CUDA(cudaMemsetAsync(Linv, 0, sizeof(double) * (size_t)R * NP * NP, st));
    TB_CUDA(cudaMemsetAsync(fb.tmax, 0, sizeof(double) * (size_t)total, st));
    TB_LAUNCH(place_diag_blocks_max_kernel, dim3(NB, R), dim3(256), 0, st, fb.Dinv, sD, Linv, NP, rowmax32, colmax32);
    const unsigned cgrid = (unsigned)((total + 255) / 256);
    for (int h64 = 2; h64 < kblks; h64 *= 2) {
        const int npairs = (kblks - h64 + 2 * h64 - 1) / (2 * h64);  // pairs whose C is not empty: (2p+1) h64 < kblks
        const long pieces = (long)(8 * h64) * h64 * npairs;
        const dim3 sgrid((unsigned)((pieces + 7) / 8), R);
        const int ntiles = npairs * (h64 / 2) * h64;
        TB_LAUNCH(cover32_kernel, dim3(cgrid), dim3(256), 0, st, colmax32, total, fb.colcov);
        TB_LAUNCH(cover32_kernel, dim3(cgrid), dim3(256), 0, st, rowmax32, total, fb.rowcov);
        TB_LAUNCH(slice_level_kernel<1>, sgrid, dim3(256), 0, st, Linv, NP, sK, fb.colcov, h64, npairs, fb.sliced, sK,
                  (long)i8::S * sK, kblks);
        TB_LAUNCH(slice_level_kernel<0>, sgrid, dim3(256), 0, st, Linv, NP, sK, fb.rowcov, h64, npairs, fb.sliced, sK,
                  (long)i8::S * sK, kblks);
        TB_CUDA(cudaMemsetAsync(tmax32, 0, sizeof(unsigned) * (size_t)total, st));
        i8::GemmI8Params g{};
        g.S = fb.sliced; g.restartStride = (long)i8::S * sK; g.sliceStride = sK;
        g.scaleStride = NP;
        g.C = Linv; g.ldc = NP; g.cStride = sK;
        g.kblks = kblks; g.tri = 0; g.h64 = h64;
        g.beta = 0.0;
        // T = L[C, A] inv(A), written where X will be
        g.recur = 1; g.scale = fb.scale; g.scaleB = fb.colcov; g.alpha = 1.0;
        g.rowmax32 = nullptr; g.colmax32 = tmax32;
        if (launch_fit_gemm_i8(fb, g, ntiles, R, st, stage)) return -1;
        TB_LAUNCH(cover32_kernel, dim3(cgrid), dim3(256), 0, st, tmax32, total, tcov);
        TB_LAUNCH(slice_level_kernel<2>, sgrid, dim3(256), 0, st, Linv, NP, sK, tcov, h64, npairs, fb.sliced, sK,
                  (long)i8::S * sK, kblks);
        // X = -inv(C) T
        g.recur = 2; g.scale = fb.rowcov; g.scaleB = tcov; g.alpha = -1.0;
        g.rowmax32 = rowmax32; g.colmax32 = colmax32;
        if (launch_fit_gemm_i8(fb, g, ntiles, R, st, stage)) return -1;
    }
    TB_LAUNCH(cover32_kernel, dim3(cgrid), dim3(256), 0, st, colmax32, total, fb.scale);
    return 0;
}

// Linv = L^-1 for R factors, from the inverted diagonal blocks; GEMM only
static int trtri_batched(const FitBuffers& fb, double* Linv, int NP, int R, bool for_model, cudaStream_t st) {
    const int NB = NP / TILE;
    const long sK = (long)NP * NP, sD = (long)NB * TILE * TILE;
    if (fit_uses_i8(NP, R) && fb.sliced != nullptr)
        return trtri_batched_i8(fb, Linv, NP, R, for_model ? FIT_TRTRI_MODEL : FIT_TRTRI, st);
    if (use_left_looking(NB, R)) {
        // row block i:  X[i,c] = -Dinv_i * sum_{k=c}^{i-1} L[i,k] X[k,c]  (c < i),  X[i,i] = Dinv_i
        TB_CUDA(cudaMemsetAsync(Linv, 0, sizeof(double) * (size_t)R * NP * NP, st));
        TB_LAUNCH(place_diag_blocks_kernel, dim3(NB, R), dim3(256), 0, st, fb.Dinv, sD, Linv, NP);
        for (int i = 1; i < NB; ++i) {
            const double* Lrow = fb.K + (long)i * TILE * NP;
            double* Xrow = Linv + (long)i * TILE * NP;
            GemmParams p = gemm_params(Lrow, NP, Linv, NP, Xrow, NP, 1, i, i * TILE, -1.0, 0.0);
            p.klo_mode = KLO_BY_TJ;  // X[k,c] = 0 for k < c
            p.sA1 = sK; p.sB1 = sK; p.sC1 = sK;
            if (launch_gemm<false, false>(p, R, 1, st)) return -1;
            GemmParams q = gemm_params(fb.Dinv + (long)i * TILE * TILE, TILE, Xrow, NP, Xrow, NP, 1, i, TILE, 1.0, 0.0);
            q.sA1 = sD; q.sB1 = sK; q.sC1 = sK;
            if (launch_gemm<false, false>(q, R, 1, st)) return -1;
        }
        return 0;
    }
    // right-looking TRSM of the identity
    {
        const long total = (long)NP * NP;
        TB_LAUNCH(set_identity_kernel, dim3((unsigned)((total + 255) / 256), R), dim3(256), 0, st, Linv, NP);
    }
    for (int j = 0; j < NB; ++j) {
        // X[j, 0:j+1] = Dinv_j * Acc[j, 0:j+1]   (in place; the diagonal block of Acc is I)
        {
            double* rowj = Linv + (long)j * TILE * NP;
            GemmParams p = gemm_params(fb.Dinv + (long)j * TILE * TILE, TILE, rowj, NP, rowj, NP, 1, j + 1, TILE, 1.0, 0.0);
            p.sA1 = sD; p.sB1 = sK; p.sC1 = sK;
            if (launch_gemm<false, false>(p, R, 1, st)) return -1;
        }
        const int rem = NB - j - 1;
        if (rem == 0) break;
        // Acc[i, 0:j+1] -= L[i,j] * X[j, 0:j+1]   (i > j)
        {
            const double* Lcol = fb.K + (long)(j + 1) * TILE * NP + (long)j * TILE;
            const double* rowj = Linv + (long)j * TILE * NP;
            double* acc = Linv + (long)(j + 1) * TILE * NP;
            GemmParams p = gemm_params(Lcol, NP, rowj, NP, acc, NP, rem, j + 1, TILE, -1.0, 1.0);
            p.sA1 = sK; p.sB1 = sK; p.sC1 = sK;
            if (launch_gemm<false, false>(p, R, 1, st)) return -1;
        }
    }
    return 0;
}

// scale -> gram -> potrf -> logdet -> trtri -> z, alpha -> lml [-> K^-1 -> grad]
static int fit_pipeline(const double* X, const double* y_n, int n, const tb200_kernel_spec& spec, const double* theta,
                        int R, double jitter, bool with_grad, const FitBuffers& fb, double* XsT, double* Linv,
                        double* alpha, double* Ldiag, double* lml, double* grad, int* info, cudaStream_t st) {
    const int NP = pad_n(n), d = spec.d, p = spec_p(spec);
    const int nt64 = (NP / GT) * (NP / GT + 1) / 2;
    TB_CUDA(cudaMemsetAsync(info, 0, sizeof(int) * R, st));
    if (fb.sched.base) TB_CUDA(cudaMemsetAsync(fb.sched.base, 0, sizeof(unsigned) * 2 * i8::SCHED_POOL, st));
    TB_LAUNCH(pad_vector_kernel, dim3((NP + 255) / 256), dim3(256), 0, st, y_n, n, NP, fb.ypad);
    TB_LAUNCH(scale_points_kernel, dim3((NP + 255) / 256, R), dim3(256), 0, st, X, n, NP, spec, theta, p, XsT);
    const size_t gsm = sizeof(double) * 2 * d * GT;
    {
        // d up to TB200_MAX_D = 64 needs 64 KB (gram) / 70 KB (gradient) of dynamic shared memory: opt in once
        static DeviceOnce once;
        if (device_once(once, [] {
                const int mx = (int)(sizeof(double) * (2 * TB200_MAX_D * GT + 8 * (TB200_MAX_D + 2)));
                TB_CUDA(cudaFuncSetAttribute(gram_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
                TB_CUDA(cudaFuncSetAttribute(lml_grad_kernel<TB200_RBF>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
                TB_CUDA(cudaFuncSetAttribute(lml_grad_kernel<TB200_MATERN12>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
                TB_CUDA(cudaFuncSetAttribute(lml_grad_kernel<TB200_MATERN32>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
                TB_CUDA(cudaFuncSetAttribute(lml_grad_kernel<TB200_MATERN52>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
                return 0;
            }))
            return -1;
    }
    TB_LAUNCH(gram_train_kernel, dim3(nt64, R), dim3(256), gsm, st, XsT, n, NP, spec, theta, p, jitter, fb.K);
    if (potrf_batched(fb, NP, R, info, st)) return -1;
    TB_LAUNCH(logdet_kernel, dim3(R), dim3(256), 0, st, fb.K, NP, Ldiag, fb.logdet);
    if (trtri_batched(fb, Linv, NP, R, !with_grad, st)) return -1;
    TB_LAUNCH(trmv_lower_kernel, dim3(NP / 8, R), dim3(256), 0, st, Linv, NP, fb.ypad, 0L, fb.z);
    TB_LAUNCH(trmv_lower_t_kernel, dim3(NP / 64, R), dim3(256), 0, st, Linv, NP, fb.z, alpha);
    TB_LAUNCH(lml_kernel, dim3(R), dim3(256), 0, st, fb.ypad, 0L, alpha, NP, n, fb.logdet, info, lml);
    if (with_grad) {
        // K^-1 (lower tiles) = Linv^T Linv, written over L
        if (fit_uses_i8(NP, R) && fb.sliced != nullptr) {
            // INT8 tensor cores: operand rows = columns of Linv (K index = row), per-column power-of-two scales
            const long sK = (long)NP * NP;
            const int NB = NP / TILE;
            // column scales of Linv: left in fb.scale by the INT8 trtri recurrence
            const long pieces = (long)(NP / 8) * (NP / i8::KB);
            TB_LAUNCH(slice_operand_kernel<true>, dim3((unsigned)((pieces + 7) / 8), R), dim3(256), 0, st, Linv, NP, sK,
                      fb.scale, 0, NP / 8, 0, NP / i8::KB, 1, 0, fb.sliced, sK,
                      (long)i8::S * sK, NP / i8::KB);
            i8::GemmI8Params g{};
            g.S = fb.sliced; g.restartStride = (long)i8::S * sK; g.sliceStride = sK;
            g.scale = fb.scale; g.scaleStride = NP;
            g.C = fb.K; g.ldc = NP; g.cStride = sK;
            g.kblks = NP / i8::KB; g.tri = 1;
            g.alpha = 1.0; g.beta = 0.0;
            if (launch_fit_gemm_i8(fb, g, NB * (NB + 1), R, st, FIT_KINV)) return -1;
        } else {
            GemmParams q = gemm_params(Linv, NP, Linv, NP, fb.K, NP, NP / TILE, NP / TILE, NP, 1.0, 0.0);
            q.order = ORDER_LOWER;
            q.klo_mode = KLO_BY_TI;
            q.sA1 = q.sB1 = q.sC1 = (long)NP * NP;
            if (launch_gemm<true, false>(q, R, 1, st)) return -1;
        }
        const size_t sm = sizeof(double) * (2 * d * GT + 8 * p);
#define TB_LML_GRAD(KIND)                                                                                     \
    TB_LAUNCH(lml_grad_kernel<KIND>, dim3(nt64, R), dim3(256), sm, st, XsT, n, NP, spec, theta, p, fb.K, alpha, \
              fb.partial, nt64)
        switch (spec.kind) {
            case TB200_RBF: TB_LML_GRAD(TB200_RBF); break;
            case TB200_MATERN12: TB_LML_GRAD(TB200_MATERN12); break;
            case TB200_MATERN32: TB_LML_GRAD(TB200_MATERN32); break;
            default: TB_LML_GRAD(TB200_MATERN52); break;
        }
#undef TB_LML_GRAD
        TB_LAUNCH(lml_grad_reduce_kernel, dim3(R, p), dim3(256), 0, st, fb.partial, nt64, p, info, grad);
    }
    return 0;
}

// tb200_shutdown: the fit path's side / look-ahead streams and events of every device
int fit_streams_destroy() {
    std::lock_guard<std::mutex> lock(g_fit_streams_mutex);
    int cur = 0;
    cudaGetDevice(&cur);
    for (int dev = 0; dev < MAX_DEVICES; ++dev) {
        FitStreams& f = g_fit_streams[dev];
        if (!f.ready) continue;
        cudaSetDevice(dev);
        cudaStreamSynchronize(f.side);
        cudaStreamSynchronize(f.look);
        cudaStreamDestroy(f.side);
        cudaStreamDestroy(f.look);
        cudaEventDestroy(f.fork); cudaEventDestroy(f.join); cudaEventDestroy(f.fork_look); cudaEventDestroy(f.look_done);
        f.ready = false;
    }
    cudaSetDevice(cur);
    return 0;
}

}  // namespace tb200

using namespace tb200;

extern "C" {

int tb200_version(void) { return 200; }

int tb200_debug_reload(void) {
    debug_config_reload();
    return 0;
}

const char* tb200_last_error(void) { return g_err; }
int tb200_padded_n(int n) { return pad_n(n); }
int64_t tb200_launch_count(void) { return g_launches; }

size_t tb200_fit_workspace_bytes(int n, int d, int R, int with_grad) {
    (void)with_grad;
    return fit_workspace_layout(n, d, R, d + 2, true, true, true, nullptr, nullptr);
}

static int check_spec(const tb200_kernel_spec* s) {
    TB_CHECK(s != nullptr, "kernel spec is NULL");
    TB_CHECK(s->kind >= TB200_RBF && s->kind <= TB200_MATERN52, "unknown kernel kind");
    TB_CHECK(s->d >= 1 && s->d <= TB200_MAX_D, "d out of range (1..64)");
    return 0;
}

int tb200_lml_batched(const double* X, const double* y_n, int n, const tb200_kernel_spec* spec, const double* theta,
                      int R, double jitter, int with_grad, double* lml, double* grad, int* info, void* work,
                      size_t work_bytes, void* stream) {
    if (check_spec(spec)) return -1;
    TB_CHECK(n >= 1 && R >= 1, "n and R must be >= 1");
    TB_CHECK(X && y_n && theta && lml && info && work, "NULL pointer argument");
    TB_CHECK(!with_grad || grad, "grad is NULL but with_grad is set");
    FitBuffers fb;
    const size_t need = fit_workspace_layout(n, spec->d, R, spec_p(*spec), true, true, true, &fb, (char*)work);
    TB_CHECK(work_bytes >= need, "workspace too small (see tb200_fit_workspace_bytes)");
    return fit_pipeline(X, y_n, n, *spec, theta, R, jitter, with_grad != 0, fb, fb.XsT, fb.Linv, fb.alpha, nullptr,
                        lml, grad, info, (cudaStream_t)stream);
}

int tb200_factor(const double* X, const double* y_n, int n, const tb200_kernel_spec* spec, const double* theta,
                 double jitter, double* XsT, double* Linv, double* alpha, double* Ldiag, double* lml, int* info,
                 void* work, size_t work_bytes, void* stream) {
    if (check_spec(spec)) return -1;
    TB_CHECK(n >= 1, "n must be >= 1");
    TB_CHECK(X && y_n && theta && XsT && Linv && alpha && lml && info && work, "NULL pointer argument");
    FitBuffers fb;
    const size_t need = fit_workspace_layout(n, spec->d, 1, spec_p(*spec), false, false, false, &fb, (char*)work);
    TB_CHECK(work_bytes >= need, "workspace too small (see tb200_fit_workspace_bytes)");
    return fit_pipeline(X, y_n, n, *spec, theta, 1, jitter, false, fb, XsT, Linv, alpha, Ldiag, lml, nullptr, info,
                        (cudaStream_t)stream);
}

size_t tb200_append_workspace_bytes(int n) { return align256(sizeof(double) * pad_n(n)) * 4 + 512; }

int tb200_factor_append(const double* x_new, double* XsT, double* Linv, double* Ldiag, int n_old,
                        const tb200_kernel_spec* spec, const double* theta, double jitter, int* info, void* work,
                        size_t work_bytes, void* stream) {
    if (check_spec(spec)) return -1;
    TB_CHECK(x_new && XsT && Linv && Ldiag && theta && info && work, "NULL pointer argument");
    TB_CHECK(n_old >= 1, "n_old must be >= 1");
    const int NP = pad_n(n_old);
    TB_CHECK(n_old < NP, "no padded row left: rebuild the factor (tb200_factor) for the larger n");
    TB_CHECK(work_bytes >= tb200_append_workspace_bytes(n_old), "workspace too small (tb200_append_workspace_bytes)");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t vec = align256(sizeof(double) * NP);
    double* kv = (double*)work;
    double* l = (double*)((char*)work + vec);
    double* u = (double*)((char*)work + 2 * vec);
    TB_LAUNCH(kcol_kernel, dim3((NP + 255) / 256), dim3(256), 0, st, x_new, XsT, n_old, NP, *spec, theta, kv);
    TB_LAUNCH(trmv_lower_kernel, dim3(NP / 8, 1), dim3(256), 0, st, Linv, NP, kv, 0L, l);
    TB_LAUNCH(trmv_lower_t_kernel, dim3(NP / 64, 1), dim3(256), 0, st, Linv, NP, l, u);
    TB_LAUNCH(append_row_kernel, dim3(1), dim3(256), 0, st, x_new, l, u, n_old, NP, *spec, theta, jitter, XsT, Linv, Ldiag,
              info);
    return 0;
}

int tb200_alpha_lml(const double* y_n, int n, const double* Linv, const double* Ldiag, double* alpha, double* lml,
                    void* work, size_t work_bytes, void* stream) {
    TB_CHECK(y_n && Linv && Ldiag && alpha && lml && work, "NULL pointer argument");
    TB_CHECK(n >= 1, "n must be >= 1");
    TB_CHECK(work_bytes >= tb200_append_workspace_bytes(n), "workspace too small (tb200_append_workspace_bytes)");
    cudaStream_t st = (cudaStream_t)stream;
    const int NP = pad_n(n);
    const size_t vec = align256(sizeof(double) * NP);
    double* ypad = (double*)work;
    double* z = (double*)((char*)work + vec);
    double* logdet = (double*)((char*)work + 3 * vec);
    int* info0 = (int*)((char*)work + 3 * vec + 256);
    TB_CUDA(cudaMemsetAsync(info0, 0, sizeof(int), st));
    TB_LAUNCH(pad_vector_kernel, dim3((NP + 255) / 256), dim3(256), 0, st, y_n, n, NP, ypad);
    TB_LAUNCH(trmv_lower_kernel, dim3(NP / 8, 1), dim3(256), 0, st, Linv, NP, ypad, 0L, z);
    TB_LAUNCH(trmv_lower_t_kernel, dim3(NP / 64, 1), dim3(256), 0, st, Linv, NP, z, alpha);
    TB_LAUNCH(logdet_from_diag_kernel, dim3(1), dim3(256), 0, st, Ldiag, n, logdet);
    TB_LAUNCH(lml_kernel, dim3(1), dim3(256), 0, st, ypad, 0L, alpha, NP, n, logdet, info0, lml);
    return 0;
}

// tb200_alpha_lml for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the same kernels, with the scratch vectors in a file-scope device array instead of a caller workspace.
// No counterpart in the reference.  The array is held by every GPU that
// loads the library until exit (512.5 KB).  All calls on a device share it: calls that overlap on the GPU race.
// Everything in it is rewritten by each call before it is read (ypad, z, logdet by their kernels, the info word by the
// memset below), so the previous call's contents never matter.
constexpr int ALPHA_LML_STATIC_MAX_N = 16384;
constexpr size_t ALPHA_LML_STATIC_BYTES = 4 * (sizeof(double) * ALPHA_LML_STATIC_MAX_N) + 512;
static __device__ __align__(256) unsigned char g_alpha_lml_scratch[ALPHA_LML_STATIC_BYTES];
// In host code the array's name is a host-side stand-in, not a device address: the address is resolved once per
// device, with that device current.
static void* g_alpha_lml_scratch_ptr[tb200::MAX_DEVICES] = {};
static std::mutex g_alpha_lml_scratch_mutex;

int tb200_alpha_lml_static(const double* y_n, int n, const double* Linv, const double* Ldiag, double* alpha, double* lml,
                           void* stream) {
    TB_CHECK(n >= 1, "n must be >= 1");
    TB_CHECK(n <= ALPHA_LML_STATIC_MAX_N,
             "tb200_alpha_lml_static: n above 16384, the limit of the built-in scratch (use tb200_alpha_lml with a workspace)");
    TB_CHECK(tb200_append_workspace_bytes(n) <= ALPHA_LML_STATIC_BYTES, "tb200_alpha_lml_static: scratch too small");
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CHECK(dev >= 0 && dev < tb200::MAX_DEVICES, "device ordinal out of range");
    void* ptr = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_alpha_lml_scratch_mutex);
        if (!g_alpha_lml_scratch_ptr[dev]) {
            void* q = nullptr;
            TB_CUDA(cudaGetSymbolAddress(&q, g_alpha_lml_scratch));
            TB_CHECK(q != nullptr, "tb200_alpha_lml_static: no device address for the scratch array");
            g_alpha_lml_scratch_ptr[dev] = q;
        }
        ptr = g_alpha_lml_scratch_ptr[dev];
    }
    return tb200_alpha_lml(y_n, n, Linv, Ldiag, alpha, lml, ptr, ALPHA_LML_STATIC_BYTES, stream);
}

int tb200_dgemm(int ta, int tb, int M, int N, int K, double alpha, const double* A, int lda, const double* B, int ldb,
                double beta, double* C, int ldc, void* stream) {
    TB_CHECK(M > 0 && N > 0 && K > 0 && M % G_TM == 0 && N % G_TN == 0 && K % G_TK == 0,
             "tb200_dgemm: M,N multiples of 128 and K multiple of 16 required");
    GemmParams p = gemm_params(A, lda, B, ldb, C, ldc, M / G_TM, N / G_TN, K, alpha, beta);
    cudaStream_t st = (cudaStream_t)stream;
    if (!ta && !tb) return launch_gemm<false, false>(p, 1, 1, st);
    if (!ta && tb) return launch_gemm<false, true>(p, 1, 1, st);
    if (ta && !tb) return launch_gemm<true, false>(p, 1, 1, st);
    return launch_gemm<true, true>(p, 1, 1, st);
}

}  // extern "C"
