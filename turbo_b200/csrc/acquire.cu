// Predict + acquisition + argmax over candidate batches (the headline path).
//
// Replaces acq(random_x) + argsort of turbo/modules/auxiliary_optimisers.py:60-66:
//   EI/PI/UCB.FunctionInstance.__call__ (acquisition_functions.py:336-358, :225-247, :147-158)
//   -> ModelInstance.predict (surrogates.py:332-338) -> sklearn predict (sklearn:_gpr.py:446-500).
//
// Per batch of `batch` candidates (a multiple of 128 sized so its scratch stays in L2):
//   1. [optional] generate the candidates on device from (seed, global index)
//   2. kstar_mean_kernel : Ks[c][j] = sf2 * g(|xc/l - xj/l|) into an L2-resident scratch, and the
//                          predictive mean  mu_c = sum_j Ks[c][j] alpha_j   (fixed-order reduction)
//   3. DMMA GEMM         : V = Linv * Ks^T (lower-triangular A) with a column sum-of-squares
//                          epilogue -> partial[row tile][c]   (V itself is never stored)
//   4. acquire_epilogue  : var = k** - sum partial, clamp, de-normalise, EI/PI/UCB, running argmax
// The candidate matrix, K*, V, mu, sigma and the acquisition values are not written to HBM unless
// the caller asks for them (mu_out / sigma_out / acq_out).
#include <stdlib.h>

#include <vector>

#include "gemm_f64.cuh"
#include "ozaki_i8.cuh"

namespace tb200 {

// ---------------------------------------------------------------------------------------------
// counter-based candidate generation (Philox4x32-10), identical for any sharding of the index range
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
    for (int round = 0; round < 10; ++round) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// uniform double in [0,1) with 53 random bits for (global index, dimension, stream)
__host__ __device__ inline double uniform01(uint64_t seed, uint64_t idx, uint32_t dim, uint32_t stream) {
    uint32_t o[4];
    philox4x32_10((uint32_t)idx, (uint32_t)(idx >> 32), dim, stream, (uint32_t)seed, (uint32_t)(seed >> 32), o);
    const uint64_t bits = ((uint64_t)o[0] << 21) ^ (uint64_t)(o[1] >> 11);  // 53 bits
    return (double)(bits & ((1ull << 53) - 1)) * (1.0 / 9007199254740992.0);
}

// keyed bijection of [0, m): 4-round Feistel on the enclosing power-of-two domain + cycle walking
__host__ __device__ inline uint64_t permute_index(uint64_t seed, uint32_t dim, uint64_t i, uint64_t m) {
    int bits = 2;
    while ((1ull << bits) < m) ++bits;
    if (bits & 1) ++bits;
    const int half = bits / 2;
    const uint64_t mask = (1ull << half) - 1;
    uint64_t x = i;
    do {
        uint64_t l = x >> half, r = x & mask;
        for (uint32_t round = 0; round < 4; ++round) {
            uint32_t o[4];
            philox4x32_10((uint32_t)r, (uint32_t)(r >> 32), dim, 0x5eed0000u + round, (uint32_t)seed,
                          (uint32_t)(seed >> 32), o);
            const uint64_t f = (((uint64_t)o[0] << 32) | o[1]) & mask;
            const uint64_t nl = r, nr = l ^ f;
            l = nl; r = nr;
        }
        x = (l << half) | r;
    } while (x >= m);
    return x;
}

struct CandGen {
    int mode;
    uint64_t seed;
    int64_t m_total;
    double lo[TB200_MAX_D];
    double hi[TB200_MAX_D];
};

inline CandGen make_gen(const tb200_cand_spec& c, int d) {
    CandGen g;
    g.mode = c.mode; g.seed = c.seed; g.m_total = c.m_total;
    for (int k = 0; k < TB200_MAX_D; ++k) { g.lo[k] = k < d ? c.lo[k] : 0.0; g.hi[k] = k < d ? c.hi[k] : 0.0; }
    return g;
}

__host__ __device__ inline double gen_coord(const CandGen& g, double lo, double hi, uint64_t idx, uint32_t dim) {
    const double u = uniform01(g.seed, idx, dim, 0u);
    if (g.mode == TB200_CAND_UNIFORM) {
#ifdef __CUDA_ARCH__
        return __dadd_rn(lo, __dmul_rn(hi - lo, u));  // no FMA contraction: bit-identical to the NumPy mirror
#else
        return lo + (hi - lo) * u;
#endif
    }
    // LHS (naive_selectors.py:75-78): lo + range * (stratum + u) / m with a per-dimension permutation
    const uint64_t s = permute_index(g.seed, dim, idx, (uint64_t)g.m_total);
    return lo + (hi - lo) * ((double)s + u) / (double)g.m_total;
}

__global__ void generate_candidates_kernel(const __grid_constant__ CandGen g, int d, int64_t index_base, int64_t m,
                                           double* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= m * d) return;
    const int64_t c = e / d;
    const int k = (int)(e - c * d);
    out[e] = gen_coord(g, g.lo[k], g.hi[k], (uint64_t)(index_base + c), (uint32_t)k);
}

// ---------------------------------------------------------------------------------------------
// K* rows and predictive mean.  CPB candidates per CTA; each thread walks training points j.
// ---------------------------------------------------------------------------------------------
constexpr int CPB = 8;

__global__ void __launch_bounds__(256) kstar_mean_kernel(const double* __restrict__ Xc, int64_t m_valid,
                                                         const double* __restrict__ XsT, int n, int NP,
                                                         tb200_kernel_spec spec, const double* __restrict__ theta,
                                                         const double* __restrict__ alpha, double* __restrict__ Ks,
                                                         double* __restrict__ mean, int jchunk, long mean_stride) {
    extern __shared__ double sh[];  // xc[CPB][d] scaled candidates, red[8][CPB]
    const int d = spec.d;
    double* xc = sh;
    double* red = sh + CPB * d;
    const int64_t c0 = (int64_t)blockIdx.x * CPB;
    for (int e = threadIdx.x; e < CPB * d; e += blockDim.x) {
        const int c = e / d, k = e - c * d;
        const int64_t ci = c0 + c;
        // sklearn divides the points by the length scale before cdist (kernels.py:1566, :1720)
        xc[e] = (ci < m_valid) ? Xc[ci * d + k] / length_scale(spec, theta, k) : 0.0;
    }
    __syncthreads();
    const Hyper h = unpack_hyper(spec, theta);
    double msum[CPB];
#pragma unroll
    for (int c = 0; c < CPB; ++c) msum[c] = 0.0;
    // blockIdx.y owns training points [jlo, jhi): enough CTAs to fill the SMs at batch ~ 2048
    const int jlo = blockIdx.y * jchunk, jhi = min(NP, jlo + jchunk);
    for (int j = jlo + threadIdx.x; j < jhi; j += blockDim.x) {
        double r2[CPB];
#pragma unroll
        for (int c = 0; c < CPB; ++c) r2[c] = 0.0;
        for (int k = 0; k < d; ++k) {
            const double xt = XsT[(long)k * NP + j];
#pragma unroll
            for (int c = 0; c < CPB; ++c) {
                const double df = xc[c * d + k] - xt;
                r2[c] += df * df;
            }
        }
        const double a = alpha[j];
        const bool live = j < n;
#pragma unroll
        for (int c = 0; c < CPB; ++c) {
            const double kv = live ? h.sf2 * shape_fn(spec.kind, r2[c]) : 0.0;
            Ks[(c0 + c) * NP + j] = kv;
            msum[c] += kv * a;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < CPB; ++c) {
        const double v = warp_sum(msum[c]);
        if (lane == 0) red[warp * CPB + c] = v;
    }
    __syncthreads();
    if (threadIdx.x < CPB) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w * CPB + threadIdx.x];
        mean[blockIdx.y * mean_stride + c0 + threadIdx.x] = t;  // partial mean; the epilogue adds the splits in order
    }
}

// ---------------------------------------------------------------------------------------------
// INT8 engine (ozaki_i8.cuh): byte slicing of the two operands
// ---------------------------------------------------------------------------------------------
// digits of b in [0,1]:  b ~= sum_{s=1..7} u_s 2^(-8s), u_s in [0,255]; exact arithmetic (scaling by 256,
// floor and the subtraction are error-free in FP64); the last digit is rounded
__device__ __forceinline__ void slice_unsigned(double t, uint8_t (&u)[i8::S]) {
#pragma unroll
    for (int s = 0; s < i8::S; ++s) {
        t *= 256.0;
        int q = (s == i8::S - 1) ? __double2int_rd(t + 0.5) : __double2int_rd(t);
        q = min(q, 255);
        u[s] = (uint8_t)q;
        t -= (double)q;
    }
}
// balanced digits of a, |a| <= 0.98:  a = sum_{s=1..7} b_s 2^(1-8s), b_s in [-128,127].  Every step is exact in
// FP64 (power-of-two scaling, integer subtraction).  Invariant: the remainder stays in [-128/255, 127/255], so the
// next digit 256*rem fits in s8; the last digit is rounded.
__device__ __forceinline__ void slice_signed(double a, uint8_t (&u)[i8::S]) {
    double t = a * 128.0;
#pragma unroll
    for (int s = 0; s < i8::S; ++s) {
        if (s > 0) t *= 256.0;
        int q = __double2int_rn(t);
        q = max(q, -128);
        q = min(q, 127);
        double rem = t - (double)q;
        if (s < i8::S - 1 && rem > 127.0 / 255.0 && q < 127) { q += 1; rem -= 1.0; }
        u[s] = (uint8_t)(int8_t)q;
        t = rem;
    }
}

// K* rows of CPB candidates as byte slices in the tiled layout ozaki_i8 reads, plus the predictive mean.
// One CTA = 8 candidates (one core-matrix row group) x jchunk training points; 16 consecutive j of the 8
// candidates form one contiguous 128-byte core matrix per slice, staged in shared memory and written
// with 16-byte stores.
__global__ void __launch_bounds__(256) kstar_mean_i8_kernel(const double* __restrict__ Xc, int64_t m_valid,
                                                            const double* __restrict__ XsT, int n, int NP,
                                                            tb200_kernel_spec spec, const double* __restrict__ theta,
                                                            const double* __restrict__ alpha,
                                                            uint8_t* __restrict__ KsS, long slice_stride,
                                                            double* __restrict__ mean, int jchunk, long mean_stride) {
    extern __shared__ double sh[];  // xc[CPB][d], red[8][CPB], then bytes[S][16 chunks][CPB][16]
    const int d = spec.d;
    double* xc = sh;
    double* red = sh + CPB * d;
    uint8_t* stage = reinterpret_cast<uint8_t*>(red + 8 * CPB);
    const int64_t c0 = (int64_t)blockIdx.x * CPB;
    for (int e = threadIdx.x; e < CPB * d; e += blockDim.x) {
        const int c = e / d, k = e - c * d;
        const int64_t ci = c0 + c;
        xc[e] = (ci < m_valid) ? Xc[ci * d + k] / length_scale(spec, theta, k) : 0.0;
    }
    __syncthreads();
    const Hyper h = unpack_hyper(spec, theta);
    const int kblks = NP / i8::KB;
    const int ct = (int)(c0 / i8::TM), rg = (int)((c0 % i8::TM) >> 3);
    double msum[CPB];
#pragma unroll
    for (int c = 0; c < CPB; ++c) msum[c] = 0.0;
    const int jlo = blockIdx.y * jchunk, jhi = min(NP, jlo + jchunk);
    for (int jb = jlo; jb < jhi; jb += 256) {
        const int j = jb + threadIdx.x;
        if (j < jhi) {
            double r2[CPB];
#pragma unroll
            for (int c = 0; c < CPB; ++c) r2[c] = 0.0;
            for (int k = 0; k < d; ++k) {
                const double xt = XsT[(long)k * NP + j];
#pragma unroll
                for (int c = 0; c < CPB; ++c) {
                    const double df = xc[c * d + k] - xt;
                    r2[c] += df * df;
                }
            }
            const double a = alpha[j];
            const bool live = j < n;
#pragma unroll
            for (int c = 0; c < CPB; ++c) {
                const double g = live ? shape_fn(spec.kind, r2[c]) : 0.0;
                msum[c] += (h.sf2 * g) * a;
                uint8_t u[i8::S];
                slice_unsigned(g, u);
#pragma unroll
                for (int s = 0; s < i8::S; ++s)
                    stage[((s * 16 + (threadIdx.x >> 4)) * CPB + c) * 16 + (threadIdx.x & 15)] = u[s];
            }
        }
        __syncthreads();
        // 7 slices x (up to 16) core matrices of 128 B -> global, 16 B per thread
        const int nchunks = min(16, (jhi - jb) >> 4);
        for (int e = threadIdx.x; e < i8::S * nchunks * CPB; e += 256) {
            const int row = e % CPB, jc = (e / CPB) % nchunks, sl = e / (CPB * nchunks);
            const int jj = jb + jc * 16;
            const size_t off = (size_t)sl * slice_stride + ((size_t)ct * kblks + jj / i8::KB) * i8::A_TILE +
                               (size_t)rg * i8::SBO + ((jj % i8::KB) >> 4) * i8::LBO + row * 16;
            *reinterpret_cast<uint4*>(KsS + off) =
                *reinterpret_cast<const uint4*>(stage + ((sl * 16 + jc) * CPB + row) * 16);
        }
        __syncthreads();
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < CPB; ++c) {
        const double v = warp_sum(msum[c]);
        if (lane == 0) red[warp * CPB + c] = v;
    }
    __syncthreads();
    if (threadIdx.x < CPB) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w * CPB + threadIdx.x];
        mean[blockIdx.y * mean_stride + c0 + threadIdx.x] = t;
    }
}

// rowscale[r] = sf2 * 2^(e_r) with 0.98 * 2^(e_r) >= max_k |Linv[r,k]|  (one warp per row)
__global__ void __launch_bounds__(256) linv_rowscale_kernel(const double* __restrict__ Linv, int NP,
                                                            tb200_kernel_spec spec, const double* __restrict__ theta,
                                                            double* __restrict__ rowscale) {
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= NP) return;
    double mx = 0.0;
    for (int k = lane; k <= r; k += 32) mx = fmax(mx, fabs(Linv[(long)r * NP + k]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) {
        int e = 0;
        const double f = frexp(mx, &e);  // mx = f * 2^e, f in [0.5, 1)
        if (f > 0.98) ++e;               // |Linv / 2^e| <= 0.98: the leading balanced digit stays within s8
        rowscale[NP + r] = (double)e;  // second half of the buffer keeps the exponent for the slicing kernel
        rowscale[r] = unpack_hyper(spec, theta).sf2 * ldexp(1.0, e);
    }
}

// Linv (lower) -> byte slices in the tiled layout: thread = (row, 16-byte K chunk) of the lower tiles
__global__ void __launch_bounds__(256) linv_slice_kernel(const double* __restrict__ Linv, int NP,
                                                         const double* __restrict__ rowscale, uint8_t* __restrict__ out,
                                                         long slice_stride) {
    const int chunks_per_row = NP / 16;
    const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long)NP * chunks_per_row) return;
    const int r = (int)(e / chunks_per_row), kc = (int)(e % chunks_per_row);
    const int rt = r / i8::TN, kb = (kc * 16) / i8::KB;
    if (kb > rt) return;  // tiles above the diagonal are never read
    const double sc = ldexp(1.0, -(int)rowscale[NP + r]);
    __align__(16) uint8_t bytes[i8::S][16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        uint8_t u[i8::S];
        slice_signed(Linv[(long)r * NP + kc * 16 + q] * sc, u);
#pragma unroll
        for (int s = 0; s < i8::S; ++s) bytes[s][q] = u[s];
    }
    const int kblks = NP / i8::KB;
    const size_t off = ((size_t)rt * kblks + kb) * i8::B_TILE + (size_t)((r % i8::TN) >> 3) * i8::SBO +
                       (((kc * 16) % i8::KB) >> 4) * i8::LBO + (r & 7) * 16;
#pragma unroll
    for (int s = 0; s < i8::S; ++s)
        *reinterpret_cast<uint4*>(out + (size_t)s * slice_stride + off) = *reinterpret_cast<const uint4*>(bytes[s]);
}

// ---------------------------------------------------------------------------------------------
// variance, acquisition value and running argmax for one batch (single CTA)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double norm_cdf(double z) { return 0.5 * erfc(-z * 0.7071067811865476); }
__device__ __forceinline__ double norm_pdf(double z) { return exp(-z * z / 2.0) / 2.5066282746310002; }

__device__ __forceinline__ bool better(double v, int64_t i, double bv, int64_t bi) {
    // NaN never wins; ties keep the lowest index; (−inf, −1) is the empty state
    if (isnan(v)) return false;
    if (bi < 0) return true;
    return v > bv || (v == bv && i < bi);
}

__global__ void __launch_bounds__(1024) acquire_epilogue_kernel(
    const double* __restrict__ mean, int nsplit, const double* __restrict__ partial, int ntiles, long ldp, int64_t batch,
    int64_t m_valid, int64_t local_base, int64_t index_base, tb200_kernel_spec spec,
    const double* __restrict__ theta, tb200_acq_spec acq,
    double* __restrict__ mu_out, double* __restrict__ sigma_out, double* __restrict__ acq_out,
    double* __restrict__ best_val, int64_t* __restrict__ best_idx) {
    __shared__ double sv[32];
    __shared__ int64_t si[32];
    double bv = -INFINITY;
    int64_t bi = -1;
    const Hyper hy = unpack_hyper(spec, theta);
    const double prior_var = hy.sf2 + hy.noise;  // kernel_.diag(X): const*1 + white (kernels.py:1315, :1438)
    for (int64_t c = threadIdx.x; c < batch && c < m_valid; c += blockDim.x) {
        double vv = 0.0;
        for (int t = 0; t < ntiles; ++t) vv += partial[(long)t * ldp + c];
        double var = prior_var - vv;                // kernel_.diag(X) - einsum(V.T, V)
        if (var < 0.0) var = 0.0;                   // sklearn:_gpr.py:485-491
        var = var * (acq.y_std * acq.y_std);        // :494
        const double sigma = sqrt(var);
        double ms = 0.0;
        for (int t = 0; t < nsplit; ++t) ms += mean[(long)t * ldp + c];
        const double mu = acq.y_std * ms + acq.y_mean;  // :450
        double a;
        if (acq.acq == TB200_UCB) {
            a = isinf(acq.param) ? sigma : acq.sign * mu + acq.param * sigma;
        } else if (sigma != 0.0) {
            const double diff = acq.sign * (mu - acq.incumbent) - acq.param;
            const double Z = diff / sigma;
            a = (acq.acq == TB200_PI) ? norm_cdf(Z) : diff * norm_cdf(Z) + sigma * norm_pdf(Z);
        } else {
            a = 0.0;
        }
        const int64_t lc = local_base + c;
        if (mu_out) mu_out[lc] = mu;
        if (sigma_out) sigma_out[lc] = sigma;
        if (acq_out) acq_out[lc] = a;
        const int64_t gi = index_base + lc;
        if (better(a, gi, bv, bi)) { bv = a; bi = gi; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi >= 0 && better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sv[warp] = bv; si[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double gv = *best_val;
        int64_t gi = *best_idx;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
            if (si[w] >= 0 && better(sv[w], si[w], gv, gi)) { gv = sv[w]; gi = si[w]; }
        *best_val = gv;
        *best_idx = gi;
    }
}

__global__ void best_reset_kernel(double* best_val, int64_t* best_idx) {
    *best_val = -INFINITY;
    *best_idx = -1;
}

// ---------------------------------------------------------------------------------------------
// FP64 peak microbenchmarks: register-resident DMMA / DFMA chains on every SM
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dmma_peak_kernel(double* out, int iters) {
    double c[16][2];
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i][0] = c[i][1] = 0.0;
    const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) dmma884(c[i][0], c[i][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
    if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, int iters) {
    double c[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = i;
    const double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i] = fma(c[i], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += c[i];
    if (s == 123.456) out[0] = s;
}

// optional CUDA-event timing of the dominant kernel (the variance GEMM), for bench.py's roofline
static bool g_time_gemm = false;
static std::vector<cudaEvent_t> g_events;
static size_t g_events_used = 0;

static cudaEvent_t next_event() {
    if (g_events_used == g_events.size()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        g_events.push_back(e);
    }
    return g_events[g_events_used++];
}

struct AcqBuffers {
    double* cand;     // [batch][d]   generated candidates
    double* Ks;       // [batch][NP]
    double* mean;     // [batch]
    double* partial;  // [NB][batch]
    double* best_val; // [1] running best of this slot
    int64_t* best_idx;
};

static size_t a256(size_t b) { return (b + 255) / 256 * 256; }

// Two slots: consecutive candidate batches alternate between two internal streams so that the tail
// of one batch's triangular GEMM, its epilogue and the next batch's K* kernel overlap.
constexpr int ACQ_SLOTS = 2;
constexpr int KSPLIT_MAX = 8;  // K* kernel: training points split over up to 8 CTAs per candidate group

static size_t acquire_layout(int n, int d, int64_t batch, AcqBuffers* ab, char* base, bool use_i8 = false) {
    const size_t NP = pad_n(n), NB = NP / TILE;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char* ptr = base ? base + off : nullptr;
        off += a256(bytes);
        return reinterpret_cast<double*>(ptr);
    };
    for (int s = 0; s < ACQ_SLOTS; ++s) {
        double* cand = take(sizeof(double) * batch * d);
        // DMMA engine: K* in FP64 (batch x NP); INT8 engine: 7 byte slices of it (tiled), partials per 64 rows
        double* Ks = take(use_i8 ? (size_t)i8::S * batch * NP : sizeof(double) * batch * NP);
        double* mean = take(sizeof(double) * KSPLIT_MAX * batch);
        double* partial = take(sizeof(double) * (use_i8 ? NP / i8::TN : NB) * batch);
        double* bv = take(sizeof(double));
        int64_t* bi = reinterpret_cast<int64_t*>(take(sizeof(int64_t)));
        if (ab) { ab[s].cand = cand; ab[s].Ks = Ks; ab[s].mean = mean; ab[s].partial = partial;
                  ab[s].best_val = bv; ab[s].best_idx = bi; }
    }
    return off;
}

// merge the per-slot bests into the caller's running best (read-modify-write, single thread)
__global__ void best_merge_kernel(const double* v0, const int64_t* i0, const double* v1, const int64_t* i1,
                                  double* best_val, int64_t* best_idx) {
    double gv = *best_val;
    int64_t gi = *best_idx;
    if (*i0 >= 0 && better(*v0, *i0, gv, gi)) { gv = *v0; gi = *i0; }
    if (v1 && *i1 >= 0 && better(*v1, *i1, gv, gi)) { gv = *v1; gi = *i1; }
    *best_val = gv;
    *best_idx = gi;
}

// internal streams / events, one set per device, created on first use
struct AcqStreams {
    bool ready = false;
    cudaStream_t s[ACQ_SLOTS];
    cudaEvent_t fork, join[ACQ_SLOTS];
};
static AcqStreams g_streams[64];
static int g_acq_streams = -1;  // -1: read TB200_ACQ_STREAMS on first use

static int acq_streams_for_device(AcqStreams** out) {
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CHECK(dev >= 0 && dev < 64, "device ordinal out of range");
    AcqStreams& a = g_streams[dev];
    if (!a.ready) {
        for (int i = 0; i < ACQ_SLOTS; ++i) {
            TB_CUDA(cudaStreamCreateWithFlags(&a.s[i], cudaStreamNonBlocking));
            TB_CUDA(cudaEventCreateWithFlags(&a.join[i], cudaEventDisableTiming));
        }
        TB_CUDA(cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming));
        a.ready = true;
    }
    *out = &a;
    return 0;
}

}  // namespace tb200

using namespace tb200;

extern "C" {

int64_t tb200_default_batch(int n) {
    // 64 MB of K* per slot: large enough that one launch is >= 3 waves of GEMM tiles at n = 4096
    // (measured: 2048 per slot beats 1024 by 3 % even though two slots + Linv exceed the 126 MB L2;
    // the GEMM re-reads K* tiles from L2/HBM at < 5 % of HBM bandwidth), never below 256 candidates
    // (tile balance) nor above 4096
    const int64_t NP = pad_n(n);
    int64_t b = (64ll << 20) / (NP * 8);
    b = b / 128 * 128;
    if (b < 256) b = 256;
    if (b > 4096) b = 4096;
    return b;
}

size_t tb200_acquire_workspace_bytes(int n, int d, int64_t batch) {
    if (batch <= 0) batch = tb200_default_batch(n);
    return acquire_layout(n, d, batch, nullptr, nullptr);
}

int tb200_best_reset(double* best_val, int64_t* best_idx, void* stream) {
    TB_CHECK(best_val && best_idx, "NULL pointer argument");
    TB_LAUNCH(best_reset_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, best_val, best_idx);
    return 0;
}

int tb200_generate_candidates(const tb200_cand_spec* cand, int d, int64_t index_base, int64_t m, double* out,
                              void* stream) {
    TB_CHECK(cand && out, "NULL pointer argument");
    TB_CHECK(cand->mode == TB200_CAND_UNIFORM || cand->mode == TB200_CAND_LHS, "generator mode required");
    TB_CHECK(d >= 1 && d <= TB200_MAX_D, "d out of range");
    TB_CHECK(cand->mode != TB200_CAND_LHS || cand->m_total >= index_base + m, "LHS: m_total smaller than the range");
    if (m <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const CandGen g = make_gen(*cand, d);
    const int64_t total = m * d;
    TB_LAUNCH(generate_candidates_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, g, d, index_base, m,
              out);
    return 0;
}

struct I8Factor {
    const uint8_t* linvS;    // NULL: DMMA engine
    const double* rowscale;
    int gmax;
};

static int acquire_impl(const double* Xc, int64_t m, int64_t index_base, const tb200_cand_spec* cand, const double* XsT,
                        const double* Linv, const I8Factor& f8, const double* alpha, int n,
                        const tb200_kernel_spec* spec, const double* theta, const tb200_acq_spec* acq, double* mu_out,
                        double* sigma_out, double* acq_out, double* best_val, int64_t* best_idx, int64_t batch,
                        void* work, size_t work_bytes, void* stream) {
    const bool use_i8 = f8.linvS != nullptr;
    TB_CHECK(spec && acq && XsT && alpha && theta && best_val && best_idx && work, "NULL pointer argument");
    TB_CHECK(use_i8 ? f8.rowscale != nullptr : Linv != nullptr, "NULL factor pointer");
    TB_CHECK(spec->kind >= TB200_RBF && spec->kind <= TB200_MATERN52, "unknown kernel kind");
    TB_CHECK(spec->d >= 1 && spec->d <= TB200_MAX_D, "d out of range (1..64)");
    TB_CHECK(acq->acq >= TB200_EI && acq->acq <= TB200_UCB, "unknown acquisition function");
    TB_CHECK(!use_i8 || (f8.gmax == 8 || f8.gmax == 9), "i8 engine: gmax must be 8 or 9");
    TB_CHECK(!use_i8 || pad_n(n) <= TB200_I8_MAX_N, "i8 engine: n > 4096 would overflow the INT32 accumulators");
    const int mode = cand ? cand->mode : TB200_CAND_GIVEN;
    if (m <= 0) return 0;
    TB_CHECK(mode != TB200_CAND_GIVEN || Xc, "Xc is NULL and no generator given");
    TB_CHECK(mode != TB200_CAND_LHS || cand->m_total >= index_base + m, "LHS: m_total smaller than the range");
    if (m <= 0) return 0;
    if (batch <= 0) batch = tb200_default_batch(n);
    TB_CHECK(batch % 128 == 0, "batch must be a multiple of 128");
    cudaStream_t st = (cudaStream_t)stream;
    const int NP = pad_n(n), NB = NP / TILE, d = spec->d;
    AcqBuffers ab[ACQ_SLOTS];
    const size_t need = acquire_layout(n, d, batch, ab, (char*)work, use_i8);
    TB_CHECK(work_bytes >= need, "workspace too small (see tb200_acquire_workspace_bytes)");
    const int ntiles = use_i8 ? NP / i8::TN : NB;  // row tiles of the variance partials
    const long slice_stride = (long)batch * NP;   // bytes between K* byte slices (i8 engine)

    if (g_acq_streams < 0) {
        const char* e = getenv("TB200_ACQ_STREAMS");
        g_acq_streams = (e && atoi(e) == 1) ? 1 : ACQ_SLOTS;
    }
    // one batch, serial mode, or GEMM timing requested: everything on the caller's stream
    const int64_t nbatches = (m + batch - 1) / batch;
    const int slots = (nbatches < 2 || g_time_gemm) ? 1 : g_acq_streams;
    AcqStreams* as = nullptr;
    cudaStream_t ss[ACQ_SLOTS] = {st, st};
    if (slots > 1) {
        if (acq_streams_for_device(&as)) return -1;
        TB_CUDA(cudaEventRecord(as->fork, st));
        for (int i = 0; i < slots; ++i) {
            ss[i] = as->s[i];
            TB_CUDA(cudaStreamWaitEvent(ss[i], as->fork, 0));
        }
    }
    for (int i = 0; i < slots; ++i)
        TB_LAUNCH(best_reset_kernel, dim3(1), dim3(1), 0, ss[i], ab[i].best_val, ab[i].best_idx);

    const size_t ksm = sizeof(double) * (CPB * d + 8 * CPB);
    // split the training points so one K* launch is ~2k CTAs (3 resident per SM): NB/ksplit tiles each
    int ksplit = NB < KSPLIT_MAX ? NB : KSPLIT_MAX;
    const int jchunk = (NB + ksplit - 1) / ksplit * TILE;
    ksplit = (NP + jchunk - 1) / jchunk;
    int64_t bno = 0;
    for (int64_t off = 0; off < m; off += batch, ++bno) {
        const int sl = (int)(bno % slots);
        const AcqBuffers& w = ab[sl];
        cudaStream_t cs = ss[sl];
        const int64_t mb = (m - off < batch) ? (m - off) : batch;          // valid candidates
        const int64_t mbp = (mb + 127) / 128 * 128;                          // padded to GEMM tiles
        const double* xc = Xc ? Xc + off * d : nullptr;
        if (mode != TB200_CAND_GIVEN) {
            const CandGen g = make_gen(*cand, d);
            const int64_t total = mb * d;
            TB_LAUNCH(generate_candidates_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, cs, g, d,
                      index_base + off, mb, w.cand);
            xc = w.cand;
        }
        if (use_i8) {
            const size_t ksm8 = ksm + (size_t)i8::S * 16 * CPB * 16;
            TB_LAUNCH(kstar_mean_i8_kernel, dim3((unsigned)(mbp / CPB), ksplit), dim3(256), ksm8, cs, xc, mb, XsT, n, NP,
                      *spec, theta, alpha, reinterpret_cast<uint8_t*>(w.Ks), slice_stride, w.mean, jchunk, (long)batch);
            i8::Params ip;
            ip.candS = reinterpret_cast<const uint8_t*>(w.Ks);
            ip.linvS = f8.linvS;
            ip.rowscale = f8.rowscale;
            ip.partial = w.partial;
            ip.ldp = batch;
            ip.candSliceStride = slice_stride;
            ip.linvSliceStride = (long)NP * NP;
            ip.kblks = NP / i8::KB;
            ip.rowTiles = NP / i8::TN;
            ip.candTiles = (int)(mbp / i8::TM);
            if (g_time_gemm) TB_CUDA(cudaEventRecord(next_event(), cs));
            if (f8.gmax == 8 ? i8::launch_colsumsq_i8<8>(ip, cs) : i8::launch_colsumsq_i8<9>(ip, cs)) return -1;
            if (g_time_gemm) TB_CUDA(cudaEventRecord(next_event(), cs));
        } else {
            TB_LAUNCH(kstar_mean_kernel, dim3((unsigned)(mbp / CPB), ksplit), dim3(256), ksm, cs, xc, mb, XsT, n, NP,
                      *spec, theta, alpha, w.Ks, w.mean, jchunk, (long)batch);
            GemmParams gp = gemm_params(Linv, NP, w.Ks, NP, w.partial, batch, NB, (int)(mbp / 128), NP, 1.0, 0.0);
            gp.order = ORDER_RECT_HEAVY_LAST_ROW_FIRST;
            gp.khi_mode = KHI_BY_TI;
            gp.a_tri_diag = 1;  // Linv is lower triangular
            gp.epi = EPI_COLSUMSQ;
            if (g_time_gemm) TB_CUDA(cudaEventRecord(next_event(), cs));
            if (launch_gemm<false, true>(gp, 1, 1, cs)) return -1;
            if (g_time_gemm) TB_CUDA(cudaEventRecord(next_event(), cs));
        }
        TB_LAUNCH(acquire_epilogue_kernel, dim3(1), dim3(1024), 0, cs, w.mean, ksplit, w.partial, ntiles, (long)batch, mbp, mb,
                  off, index_base, *spec, theta, *acq, mu_out, sigma_out, acq_out, w.best_val, w.best_idx);
    }
    if (slots > 1) {
        for (int i = 0; i < slots; ++i) {
            TB_CUDA(cudaEventRecord(as->join[i], ss[i]));
            TB_CUDA(cudaStreamWaitEvent(st, as->join[i], 0));
        }
    }
    TB_LAUNCH(best_merge_kernel, dim3(1), dim3(1), 0, st, ab[0].best_val, ab[0].best_idx,
              slots > 1 ? ab[1].best_val : (const double*)nullptr, slots > 1 ? ab[1].best_idx : (const int64_t*)nullptr,
              best_val, best_idx);
    return 0;
}

int tb200_acquire(const double* Xc, int64_t m, int64_t index_base, const tb200_cand_spec* cand, const double* XsT,
                  const double* Linv, const double* alpha, int n, const tb200_kernel_spec* spec, const double* theta,
                  const tb200_acq_spec* acq, double* mu_out, double* sigma_out, double* acq_out, double* best_val,
                  int64_t* best_idx, int64_t batch, void* work, size_t work_bytes, void* stream) {
    const I8Factor none{nullptr, nullptr, 0};
    return acquire_impl(Xc, m, index_base, cand, XsT, Linv, none, alpha, n, spec, theta, acq, mu_out, sigma_out, acq_out,
                        best_val, best_idx, batch, work, work_bytes, stream);
}

int tb200_acquire_i8(const double* Xc, int64_t m, int64_t index_base, const tb200_cand_spec* cand, const double* XsT,
                     const uint8_t* linvS, const double* rowscale, int gmax, const double* alpha, int n,
                     const tb200_kernel_spec* spec, const double* theta, const tb200_acq_spec* acq, double* mu_out,
                     double* sigma_out, double* acq_out, double* best_val, int64_t* best_idx, int64_t batch, void* work,
                     size_t work_bytes, void* stream) {
    TB_CHECK(linvS && rowscale, "NULL pointer argument");
    const I8Factor f8{linvS, rowscale, gmax};
    return acquire_impl(Xc, m, index_base, cand, XsT, nullptr, f8, alpha, n, spec, theta, acq, mu_out, sigma_out,
                        acq_out, best_val, best_idx, batch, work, work_bytes, stream);
}

size_t tb200_acquire_i8_workspace_bytes(int n, int d, int64_t batch) {
    if (batch <= 0) batch = tb200_default_batch(n);
    return acquire_layout(n, d, batch, nullptr, nullptr, true);
}

size_t tb200_i8_factor_bytes(int n) { return (size_t)i8::S * pad_n(n) * pad_n(n); }

int tb200_i8_prepare(const double* Linv, int n, const tb200_kernel_spec* spec, const double* theta, uint8_t* linvS,
                     double* rowscale, void* stream) {
    TB_CHECK(Linv && spec && theta && linvS && rowscale, "NULL pointer argument");
    const int NP = pad_n(n);
    TB_CHECK(NP <= TB200_I8_MAX_N, "i8 engine: n > 4096 would overflow the INT32 accumulators");
    cudaStream_t st = (cudaStream_t)stream;
    TB_LAUNCH(linv_rowscale_kernel, dim3(NP / 8), dim3(256), 0, st, Linv, NP, *spec, theta, rowscale);
    const long total = (long)NP * (NP / 16);
    TB_LAUNCH(linv_slice_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, Linv, NP, rowscale, linvS,
              (long)NP * NP);
    return 0;
}

int tb200_timing_enable(int on) {
    g_time_gemm = on != 0;
    g_events_used = 0;
    return 0;
}

int tb200_timing_read(double* gemm_ms, int64_t* gemm_launches) {
    TB_CHECK(gemm_ms && gemm_launches, "NULL pointer argument");
    double total = 0.0;
    for (size_t i = 0; i + 1 < g_events_used; i += 2) {
        TB_CUDA(cudaEventSynchronize(g_events[i + 1]));
        float t = 0.f;
        TB_CUDA(cudaEventElapsedTime(&t, g_events[i], g_events[i + 1]));
        total += t;
    }
    *gemm_ms = total;
    *gemm_launches = (int64_t)(g_events_used / 2);
    g_events_used = 0;
    return 0;
}

int tb200_fp64_peak(double ms, double* dmma_tflops, double* dfma_tflops) {
    TB_CHECK(dmma_tflops && dfma_tflops, "NULL pointer argument");
    int dev = 0, sms = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    double* out = nullptr;
    TB_CUDA(cudaMalloc((void**)&out, 64));
    cudaEvent_t e0, e1;
    TB_CUDA(cudaEventCreate(&e0));
    TB_CUDA(cudaEventCreate(&e1));
    const int ctas = sms * 2;
    for (int which = 0; which < 2; ++which) {
        int iters = 2000;
        double best = 0.0;
        for (int rep = 0; rep < 6; ++rep) {
            TB_CUDA(cudaEventRecord(e0, 0));
            if (which == 0) { TB_LAUNCH(dmma_peak_kernel, dim3(ctas), dim3(256), 0, 0, out, iters); }
            else { TB_LAUNCH(dfma_peak_kernel, dim3(ctas), dim3(256), 0, 0, out, iters); }
            TB_CUDA(cudaEventRecord(e1, 0));
            TB_CUDA(cudaEventSynchronize(e1));
            float t = 0.f;
            TB_CUDA(cudaEventElapsedTime(&t, e0, e1));
            // flops: DMMA.8x8x4 = 256 FMA per warp instr; DFMA = 32 FMA per warp instr
            const double fma = (which == 0 ? 256.0 : 32.0) * 16.0 * iters * 8.0 * ctas;
            const double tf = 2.0 * fma / (t * 1e-3) / 1e12;
            if (rep > 0 && tf > best) best = tf;
            if (t < ms) iters = (int)(iters * (ms / (t > 1e-3 ? t : 1e-3)) * 1.1) + 1;
        }
        (which == 0 ? *dmma_tflops : *dfma_tflops) = best;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return 0;
}

}  // extern "C"
