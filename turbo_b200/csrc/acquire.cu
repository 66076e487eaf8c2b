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
// digits of b in [0,1]:  b ~= sum_{s=1..7} u_s 2^(-8s), u_s in [0,255]: b is rounded to 56-bit fixed point (one
// exact power-of-two scaling + one conversion) and the digits are its bytes; b = 1 maps to 1 - 2^-56
__device__ __forceinline__ void slice_unsigned(double t, uint8_t (&u)[i8::S]) {
    unsigned long long x = __double2ull_rn(t * 72057594037927936.0);  // 2^56
    x = min(x, 72057594037927935ull);
#pragma unroll
    for (int s = 0; s < i8::S; ++s) u[s] = (uint8_t)(x >> (48 - 8 * s));
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

// K* rows of 8 candidates as byte slices in the tiled layout ozaki_i8 reads, plus the predictive mean.
// Thread = (candidate c = tid & 7, group of 16 consecutive training points): its 16 values of one slice are exactly
// one 16-byte row of a core matrix, and the 8 candidates of the CTA (one core-matrix row group) make the 128-byte core
// matrix contiguous, so every slice is written with one 16-byte store per thread straight from registers -- no
// shared-memory staging, no barriers in the loop, the candidate's coordinates are the only shared-memory reads.
// (v1 walked one training point x 8 candidates per thread and staged single bytes through shared memory: 223
// instructions per pair, 64 of them FP64 -- profiles/r1_kstar_ncu.md.)  KIND is a template parameter so the
// 16-way unrolled body carries one kernel family.
constexpr int KS_THREADS = 128;  // 16 point groups x 8 candidates; small CTAs share an SM with a variance CTA
constexpr int KS_JG = 16;        // training points per thread and pass
constexpr int KS_CTAS_PER_SM = 4;  // persistent grid: 3 x 128 threads x 80 registers fit beside one variance CTA, a 4th
                                   // where none is resident (measured in round 1: n=1024 +10 %, n=4096 unchanged)

template <int KIND>
__global__ void __launch_bounds__(KS_THREADS) kstar_mean_i8_kernel(
    const double* __restrict__ Xc, int64_t m_valid, const double* __restrict__ XsT, int n, int NP,
    tb200_kernel_spec spec, const double* __restrict__ theta, const double* __restrict__ alpha,
    uint8_t* __restrict__ KsS, long slice_stride, int nplanes, double* __restrict__ mean, int jchunk, long mean_stride,
    int ngroups, int ksplit) {
    extern __shared__ double sh[];  // xc[d][CPB] scaled candidates, red[4][CPB]
    const int d = spec.d;
    double* xc = sh;
    double* red = sh + CPB * d;
    const double sf2 = unpack_hyper(spec, theta).sf2;
    const int c = threadIdx.x & 7, jg = threadIdx.x >> 3;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int kblks = NP / i8::KB;
    // persistent: a few CTAs per SM walk the (candidate group, point split) items.  The block scheduler places a
    // later grid only after every CTA of the grids ahead of it; a small grid is placed at once next to the resident
    // variance CTAs instead of holding back (or being held back by) the 1024-CTA variance grid.
    for (int item = blockIdx.x; item < ngroups * ksplit; item += gridDim.x) {
        const int grp = item / ksplit, split = item - grp * ksplit;
        const int64_t c0 = (int64_t)grp * CPB;
        __syncthreads();  // previous item's xc / red are no longer read
        for (int e = threadIdx.x; e < CPB * d; e += KS_THREADS) {
            const int k = e / CPB, cc = e - k * CPB;
            const int64_t ci = c0 + cc;
            // sklearn divides the points by the length scale before cdist (kernels.py:1566, :1720)
            xc[e] = (ci < m_valid) ? Xc[ci * d + k] / length_scale(spec, theta, k) : 0.0;
        }
        __syncthreads();
        const int ct = (int)(c0 / i8::TM), rg = (int)((c0 % i8::TM) >> 3);
        uint8_t* const tile_base = KsS + (size_t)ct * kblks * i8::A_TILE + (size_t)rg * i8::SBO + c * 16;
        double msum = 0.0;
        const int jlo = split * jchunk, jhi = min(NP, jlo + jchunk);
        for (int j0 = jlo + jg * KS_JG; j0 < jhi; j0 += (KS_THREADS / 8) * KS_JG) {
            double r2[KS_JG];
#pragma unroll
            for (int q = 0; q < KS_JG; ++q) r2[q] = 0.0;
#pragma unroll 2
            for (int k = 0; k < d; ++k) {  // two dimensions in flight: the next 8 loads issue under this one's FMAs
                const double xk = xc[k * CPB + c];
                const double2* xt = reinterpret_cast<const double2*>(XsT + (long)k * NP + j0);
#pragma unroll
                for (int q = 0; q < KS_JG / 2; ++q) {
                    const double2 v = __ldg(xt + q);
                    const double d0 = xk - v.x, d1 = xk - v.y;
                    r2[2 * q] = fma(d0, d0, r2[2 * q]);
                    r2[2 * q + 1] = fma(d1, d1, r2[2 * q + 1]);
                }
            }
            uint32_t lo[KS_JG], hi[KS_JG];
            const double2* al = reinterpret_cast<const double2*>(alpha + j0);
#pragma unroll
            for (int q = 0; q < KS_JG; q += 2) {
                const double2 a = __ldg(al + (q >> 1));
                // evaluated for the padding rows too and masked afterwards: straight-line code over the 16 pairs
                const double s0 = shape_fn(KIND, r2[q]), s1 = shape_fn(KIND, r2[q + 1]);
                const double g0 = (j0 + q < n) ? s0 : 0.0;
                const double g1 = (j0 + q + 1 < n) ? s1 : 0.0;
                msum = fma(sf2 * g0, a.x, msum);
                msum = fma(sf2 * g1, a.y, msum);
                // 56-bit fixed point of g in [0,1] (see slice_unsigned): its bytes are the 7 digits
                unsigned long long x0 = (unsigned long long)fixed_point_64<24>(g0);
                unsigned long long x1 = (unsigned long long)fixed_point_64<24>(g1);
                x0 = min(x0, 72057594037927935ull);
                x1 = min(x1, 72057594037927935ull);
                lo[q] = (uint32_t)x0; hi[q] = (uint32_t)(x0 >> 32);
                lo[q + 1] = (uint32_t)x1; hi[q + 1] = (uint32_t)(x1 >> 32);
            }
            uint8_t* dst = tile_base + (size_t)(j0 / i8::KB) * i8::A_TILE + ((j0 % i8::KB) >> 4) * i8::LBO;
#pragma unroll
            for (int sl = 0; sl < i8::S; ++sl) {
                // digit sl = byte (6 - sl) of the fixed-point value: bytes 2,1,0 of hi, then 3,2,1,0 of lo
                const int byte = (sl < 3) ? 2 - sl : 6 - sl;
                uint32_t w[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint32_t* src = (sl < 3) ? hi : lo;
                    const uint32_t p01 = __byte_perm(src[4 * v], src[4 * v + 1], byte | ((4 + byte) << 4));
                    const uint32_t p23 = __byte_perm(src[4 * v + 2], src[4 * v + 3], byte | ((4 + byte) << 4));
                    w[v] = __byte_perm(p01, p23, 0x5410);
                }
                // planes the variance kernel's cut-off never reads (the 7th at 21 products) are not written
                if (sl < nplanes) *reinterpret_cast<uint4*>(dst + (size_t)sl * slice_stride) = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
        // lanes with the same candidate (lane & 7) hold partial means: fold lane bits 3 and 4, then the 4 warps in order
        msum += __shfl_xor_sync(0xffffffffu, msum, 8);
        msum += __shfl_xor_sync(0xffffffffu, msum, 16);
        if (lane < CPB) red[warp * CPB + lane] = msum;
        __syncthreads();
        if (threadIdx.x < CPB) {
            double t = 0.0;
            for (int w = 0; w < KS_THREADS / 32; ++w) t += red[w * CPB + threadIdx.x];
            mean[split * mean_stride + c0 + threadIdx.x] = t;  // partial mean; the epilogue adds the splits in order
        }
    }
}

// The same K* rows and partial means with the training points STAGED in shared memory (n <= 4096).  ncu of the kernel
// above: 44 % of the stall samples are the loads of the scaled training points -- next to the variance kernel's 170 KB
// of shared memory L1 is 28 KB, so every pass fetched its 16 x d values per thread from L2 and waited for them
// (profiles/r1_kstar_ncu_full.md).  Here a CTA owns a contiguous range of items in SPLIT-MAJOR order (item = chunk of
// 256 training points x group of 8 candidates): the chunk (d x 256 doubles + its alpha) is staged once and the CTA
// walks candidate groups under it.  Points are stored 16 to a row of 18 doubles so that the four point groups a warp
// reads with one 16-byte load fall into different banks.  One partial mean per chunk (at most KS_STAGED_SPLITS).
// NC = candidates per item (8 or 16): the 128 threads are NC candidates x 128/NC point groups of 16, so a chunk is
// 256 points (NC = 8) or 128 points (NC = 16; half the shared memory per CTA, which keeps two CTAs per SM up to
// d = 21; at d = 10 it would allow a third CTA, which measured slower in the pipeline: 379.5 vs 375.5 ms per step).
constexpr int KS_STAGED_SPLITS = 32;   // partial means per candidate (mean buffer rows): n <= 4096 at 128-point chunks
__host__ __device__ constexpr int ks_chunk(int NC) { return KS_THREADS / NC * KS_JG; }
__host__ __device__ constexpr int ks_row(int NC) { return KS_THREADS / NC * (KS_JG + 2); }  // 16 points to a padded row of 18 doubles

static size_t kstar_staged_smem(int d, int NC) {
    return sizeof(double) * ((size_t)d * ks_row(NC) + ks_row(NC) + (size_t)NC * d + (KS_THREADS / 32) * NC);
}

template <int KIND, int NC>
__global__ void __launch_bounds__(KS_THREADS) kstar_mean_i8_staged_kernel(
    const double* __restrict__ Xc, int64_t m_valid, const double* __restrict__ XsT, int n, int NP,
    tb200_kernel_spec spec, const double* __restrict__ theta, const double* __restrict__ alpha,
    uint8_t* __restrict__ KsS, long slice_stride, int nplanes, double* __restrict__ mean, long mean_stride, int ngroups,
    int nsplit) {
    constexpr int CHUNK = ks_chunk(NC), ROW = ks_row(NC);
    extern __shared__ double sh[];
    const int d = spec.d;
    double* xs = sh;                    // [d][ROW] scaled training points of the current chunk
    double* al = xs + (size_t)d * ROW;  // [ROW] alpha of the chunk, same padding
    double* xc = al + ROW;              // [d][NC] scaled candidates of the current group
    double* red = xc + NC * d;
    const double sf2 = unpack_hyper(spec, theta).sf2;
    const int c = threadIdx.x % NC, jg = threadIdx.x / NC;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int kblks = NP / i8::KB;
    const int total = ngroups * nsplit;
    const int per = (total + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo_item = (int)blockIdx.x * per, hi_item = min(total, lo_item + per);
    int staged = -1;
    for (int item = lo_item; item < hi_item; ++item) {
        const int split = item / ngroups, grp = item - split * ngroups;
        const int64_t c0 = (int64_t)grp * NC;
        const int jlo = split * CHUNK;
        __syncthreads();  // previous item's xs / xc / red are no longer read
        if (split != staged) {
            for (int e = threadIdx.x; e < (d + 1) * CHUNK; e += KS_THREADS) {
                const int k = e / CHUNK, j = e - k * CHUNK;
                const int pos = j + (j / KS_JG) * 2;
                const bool in = jlo + j < NP;
                if (k < d) xs[(size_t)k * ROW + pos] = in ? __ldg(XsT + (long)k * NP + jlo + j) : 0.0;
                else al[pos] = in ? __ldg(alpha + jlo + j) : 0.0;
            }
            staged = split;
        }
        for (int e = threadIdx.x; e < NC * d; e += KS_THREADS) {
            const int k = e / NC, cc = e - k * NC;
            const int64_t ci = c0 + cc;
            xc[e] = (ci < m_valid) ? Xc[ci * d + k] / length_scale(spec, theta, k) : 0.0;
        }
        __syncthreads();
        // candidate c of the item: row group (c0 % 128) / 8 + c / 8 of candidate tile c0 / 128, row c % 8 in the group
        const int ct = (int)(c0 / i8::TM), rg = (int)((c0 % i8::TM) >> 3) + (c >> 3);
        uint8_t* const tile_base = KsS + (size_t)ct * kblks * i8::A_TILE + (size_t)rg * i8::SBO + (c & 7) * 16;
        double msum = 0.0;
        const int j0 = jlo + jg * KS_JG;
        if (j0 < NP) {
            const double* xrow = xs + jg * (KS_JG + 2);
            double r2[KS_JG];
#pragma unroll
            for (int q = 0; q < KS_JG; ++q) r2[q] = 0.0;
#pragma unroll 2
            for (int k = 0; k < d; ++k) {
                const double xk = xc[k * NC + c];
                const double2* xt = reinterpret_cast<const double2*>(xrow + (size_t)k * ROW);
#pragma unroll
                for (int q = 0; q < KS_JG / 2; ++q) {
                    const double2 v = xt[q];
                    const double d0 = xk - v.x, d1 = xk - v.y;
                    r2[2 * q] = fma(d0, d0, r2[2 * q]);
                    r2[2 * q + 1] = fma(d1, d1, r2[2 * q + 1]);
                }
            }
            uint32_t lo[KS_JG], hi[KS_JG];
            const double2* ap = reinterpret_cast<const double2*>(al + jg * (KS_JG + 2));
#pragma unroll
            for (int q = 0; q < KS_JG; q += 2) {
                const double2 a = ap[q >> 1];
                const double s0 = shape_fn(KIND, r2[q]), s1 = shape_fn(KIND, r2[q + 1]);
                const double g0 = (j0 + q < n) ? s0 : 0.0;
                const double g1 = (j0 + q + 1 < n) ? s1 : 0.0;
                msum = fma(sf2 * g0, a.x, msum);
                msum = fma(sf2 * g1, a.y, msum);
                unsigned long long x0 = (unsigned long long)fixed_point_64<24>(g0);
                unsigned long long x1 = (unsigned long long)fixed_point_64<24>(g1);
                x0 = min(x0, 72057594037927935ull);
                x1 = min(x1, 72057594037927935ull);
                lo[q] = (uint32_t)x0; hi[q] = (uint32_t)(x0 >> 32);
                lo[q + 1] = (uint32_t)x1; hi[q + 1] = (uint32_t)(x1 >> 32);
            }
            uint8_t* dst = tile_base + (size_t)(j0 / i8::KB) * i8::A_TILE + ((j0 % i8::KB) >> 4) * i8::LBO;
#pragma unroll
            for (int sl = 0; sl < i8::S; ++sl) {
                const int byte = (sl < 3) ? 2 - sl : 6 - sl;
                uint32_t w[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint32_t* src = (sl < 3) ? hi : lo;
                    const uint32_t p01 = __byte_perm(src[4 * v], src[4 * v + 1], byte | ((4 + byte) << 4));
                    const uint32_t p23 = __byte_perm(src[4 * v + 2], src[4 * v + 3], byte | ((4 + byte) << 4));
                    w[v] = __byte_perm(p01, p23, 0x5410);
                }
                // planes the variance kernel's cut-off never reads (the 7th at 21 products) are not written
                if (sl < nplanes) *reinterpret_cast<uint4*>(dst + (size_t)sl * slice_stride) = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
        // lanes with the same candidate (lane % NC) hold partial means: fold the point-group bits of the lane index
#pragma unroll
        for (int o = NC; o < 32; o <<= 1) msum += __shfl_xor_sync(0xffffffffu, msum, o);
        if (lane < NC) red[warp * NC + lane] = msum;
        __syncthreads();
        if (threadIdx.x < NC) {
            double t = 0.0;
            for (int w = 0; w < KS_THREADS / 32; ++w) t += red[w * NC + threadIdx.x];
            mean[split * mean_stride + c0 + threadIdx.x] = t;  // partial mean; the epilogue adds the splits in order
        }
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
// "too close to an existing sample" over a whole candidate batch.  The reference tests ONE selected point
// against the trial set after the argmax, turbo/utils.py:111-130 `close_to_any`:
//     np.any(np.sum((xs - x)**2, axis=1) <= tol)
// and throws the Bayesian trial away when it is too close (turbo/optimiser.py:349-352, turbo/modules/fallback.py:49-54).
// Here every candidate gets its smallest squared distance to the n points so the argmax can skip such candidates in
// the same pass.  The sum over the d attributes follows NumPy's contiguous reduction bit for bit: fewer than 8 terms
// are added left to right; otherwise 8 running sums over strides of 8, combined pairwise, then the tail left to right
// (no FMA contraction: NumPy squares, then adds).  One thread per candidate; the points are staged through shared
// memory in chunks and read as broadcasts.  FP64-pipe bound: 3 n d operations per candidate.
// ---------------------------------------------------------------------------------------------
constexpr int MSD_THREADS = 128, MSD_CHUNK = 64;
static int g_sm_count = 148;  // refreshed by configure_carveout()

__global__ void __launch_bounds__(MSD_THREADS) min_sqdist_kernel(const double* __restrict__ Xc, int64_t m,
                                                                 const double* __restrict__ X, int n, int d,
                                                                 double* __restrict__ min_out,
                                                                 int64_t* __restrict__ arg_out, int jchunk) {
    extern __shared__ double msd_sh[];
    double* xs = msd_sh;                     // [d][MSD_THREADS]: column t belongs to thread t
    double* xt = msd_sh + (size_t)d * MSD_THREADS;  // [MSD_CHUNK][d]
    const int64_t c0 = (int64_t)blockIdx.x * MSD_THREADS;
    const int64_t rows = (m - c0 < MSD_THREADS) ? (m - c0) : MSD_THREADS;
    for (int64_t e = threadIdx.x; e < (int64_t)MSD_THREADS * d; e += MSD_THREADS) {
        const int64_t r = e / d;
        const int k = (int)(e - r * d);
        xs[(size_t)k * MSD_THREADS + r] = (r < rows) ? Xc[c0 * d + e] : 0.0;
    }
    double best = INFINITY;
    int64_t arg = -1;
    const double* mine = xs + threadIdx.x;
    const int d8 = d - (d & 7);
    // blockIdx.y owns the points [jlo, jhi): with more than one such range the results meet in an atomic min on the
    // bit pattern (non-negative doubles order like unsigned integers; min_out starts at +inf, arg_out is not kept)
    const int jlo = (int)blockIdx.y * jchunk;
    const int jhi = (jlo + jchunk < n) ? (jlo + jchunk) : n;
    for (int j0 = jlo; j0 < jhi; j0 += MSD_CHUNK) {
        const int cnt = (jhi - j0 < MSD_CHUNK) ? (jhi - j0) : MSD_CHUNK;
        __syncthreads();
        for (int e = threadIdx.x; e < cnt * d; e += MSD_THREADS) xt[e] = X[(size_t)j0 * d + e];
        __syncthreads();
        for (int j = 0; j < cnt; ++j) {
            const double* t = xt + (size_t)j * d;
            double sum;
            if (d < 8) {
                sum = 0.0;
                for (int k = 0; k < d; ++k) {
                    const double df = __dsub_rn(t[k], mine[(size_t)k * MSD_THREADS]);
                    sum = __dadd_rn(sum, __dmul_rn(df, df));
                }
            } else {
                double r[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const double df = __dsub_rn(t[u], mine[(size_t)u * MSD_THREADS]);
                    r[u] = __dmul_rn(df, df);
                }
                for (int k = 8; k < d8; k += 8) {
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const double df = __dsub_rn(t[k + u], mine[(size_t)(k + u) * MSD_THREADS]);
                        r[u] = __dadd_rn(r[u], __dmul_rn(df, df));
                    }
                }
                sum = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                                __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
                for (int k = d8; k < d; ++k) {
                    const double df = __dsub_rn(t[k], mine[(size_t)k * MSD_THREADS]);
                    sum = __dadd_rn(sum, __dmul_rn(df, df));
                }
            }
            if (sum < best) { best = sum; arg = j0 + j; }  // NaN never wins; ties keep the first point
        }
    }
    if (threadIdx.x < rows) {
        if (gridDim.y > 1) {
            if (best < INFINITY)
                atomicMin(reinterpret_cast<unsigned long long*>(min_out + c0 + threadIdx.x),
                          (unsigned long long)__double_as_longlong(best));
        } else {
            min_out[c0 + threadIdx.x] = best;
            if (arg_out) arg_out[c0 + threadIdx.x] = arg;
        }
    }
}

__global__ void fill_inf_kernel(double* __restrict__ out, int64_t count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = INFINITY;
}

static size_t msd_smem(int d) { return sizeof(double) * (size_t)d * (MSD_THREADS + MSD_CHUNK); }

static int launch_min_sqdist(const double* Xc, int64_t m, const double* X, int n, int d, double* min_out,
                             int64_t* arg_out, cudaStream_t st) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(min_sqdist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)msd_smem(TB200_MAX_D)));
            return 0;
        }))
        return -1;
    if (m <= 0) return 0;
    // few candidates (one pipeline batch): split the points over blockIdx.y so the launch fills the machine
    const int64_t ctas = (m + MSD_THREADS - 1) / MSD_THREADS;
    int split = 1;
    if (!arg_out && n > MSD_CHUNK && ctas < 8 * g_sm_count) {
        const int64_t want = (8 * (int64_t)g_sm_count + ctas - 1) / ctas;
        const int chunks = (n + MSD_CHUNK - 1) / MSD_CHUNK;
        split = (int)(want < chunks ? want : chunks);
    }
    int jchunk = n > 0 ? ((n + split - 1) / split + MSD_CHUNK - 1) / MSD_CHUNK * MSD_CHUNK : MSD_CHUNK;
    split = n > 0 ? (n + jchunk - 1) / jchunk : 1;
    if (split > 1)
        TB_LAUNCH(fill_inf_kernel, dim3((unsigned)((m + 255) / 256)), dim3(256), 0, st, min_out, m);
    TB_LAUNCH(min_sqdist_kernel, dim3((unsigned)ctas, (unsigned)split), dim3(MSD_THREADS), msd_smem(d), st, Xc, m, X, n,
              d, min_out, arg_out, jchunk);
    return 0;
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

// One CTA of 128 threads per 128 candidates: small enough to share an SM with a resident variance-kernel CTA
// (a single 1024-thread CTA could not, and stalled the two-stream schedule until the other slot's GEMM had
// drained).  CTA b keeps its own running best in best_val[b] / best_idx[b]: launches of one slot are ordered by
// their stream, so the read-modify-write needs no atomics and the result is deterministic.
constexpr int EPI_THREADS = 128;

__global__ void __launch_bounds__(EPI_THREADS) acquire_epilogue_kernel(
    const double* __restrict__ mean, int nsplit, const double* __restrict__ partial, int ntiles, long ldp, int64_t batch,
    int64_t m_valid, int64_t local_base, int64_t index_base, tb200_kernel_spec spec,
    const double* __restrict__ theta, tb200_acq_spec acq,
    double* __restrict__ mu_out, double* __restrict__ sigma_out, double* __restrict__ acq_out,
    double* __restrict__ best_val, int64_t* __restrict__ best_idx, const double* __restrict__ min_sqdist,
    double close_tol) {
    __shared__ double sv[EPI_THREADS / 32];
    __shared__ int64_t si[EPI_THREADS / 32];
    double bv = -INFINITY;
    int64_t bi = -1;
    const Hyper hy = unpack_hyper(spec, theta);
    const double prior_var = hy.sf2 + hy.noise;  // kernel_.diag(X): const*1 + white (kernels.py:1315, :1438)
    const int64_t c = (int64_t)blockIdx.x * EPI_THREADS + threadIdx.x;
    if (c < batch && c < m_valid) {
        // 4 independent chains over the row tiles, combined in a fixed order
        double v4[4] = {0.0, 0.0, 0.0, 0.0};
        int t = 0;
        for (; t + 16 <= ntiles; t += 16) {  // 16 loads in flight
            double ld[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) ld[u] = partial[(long)(t + u) * ldp + c];
#pragma unroll
            for (int u = 0; u < 16; ++u) v4[u & 3] += ld[u];
        }
        for (; t + 4 <= ntiles; t += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) v4[u] += partial[(long)(t + u) * ldp + c];
        }
        for (; t < ntiles; ++t) v4[0] += partial[(long)t * ldp + c];
        const double vv = (v4[0] + v4[1]) + (v4[2] + v4[3]);
        double var = prior_var - vv;                // kernel_.diag(X) - einsum(V.T, V)
        if (var < 0.0) var = 0.0;                   // sklearn:_gpr.py:485-491
        var = var * (acq.y_std * acq.y_std);        // :494
        const double sigma = sqrt(var);
        double ms = 0.0;
        for (int u = 0; u < nsplit; ++u) ms += mean[(long)u * ldp + c];
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
        // a candidate within close_tol of an existing sample keeps its value in acq_out but cannot be selected
        const bool too_close = min_sqdist != nullptr && min_sqdist[c] <= close_tol;
        if (!too_close && better(a, gi, bv, bi)) { bv = a; bi = gi; }
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
        double gv = best_val[blockIdx.x];
        int64_t gi = best_idx[blockIdx.x];
        for (int w = 0; w < EPI_THREADS / 32; ++w)
            if (si[w] >= 0 && better(sv[w], si[w], gv, gi)) { gv = sv[w]; gi = si[w]; }
        best_val[blockIdx.x] = gv;
        best_idx[blockIdx.x] = gi;
    }
}

// ---------------------------------------------------------------------------------------------
// acquisition value AND gradient w.r.t. the point (analytic replacement for the numerical
// gradients of the L-BFGS-B stage of auxiliary_optimisers.py:69-112).  One CTA per point p:
//   dk_j/dx = -sf2 h(r2_j) (x/l - x_j/l) / l          (h = shape_grad_fn, dK/dlog l_k = sf2 h D_k)
//   dmu/dx  = y_std sum_j alpha_j dk_j/dx,   dvar_n/dx = -2 sum_j w_j dk_j/dx,  w = Linv^T (Linv k*)
// written as xs*SA - sum_j A_j xs_j with A_j = alpha_j sf2 h_j (B_j = w_j sf2 h_j) so that no thread keeps
// per-dimension accumulators.  V = Linv K*^T and W = Linv^T V come from the DMMA GEMM (NP x 128, ld 128).
// ---------------------------------------------------------------------------------------------
constexpr int GRAD_PMAX = 128;  // points per call (one GEMM column tile)

__global__ void __launch_bounds__(256) acq_grad_kernel(
    const double* __restrict__ Xp, int P, const double* __restrict__ XsT, int n, int NP, tb200_kernel_spec spec,
    const double* __restrict__ theta, const double* __restrict__ alpha, const double* __restrict__ mean, int nsplit,
    long mean_stride, const double* __restrict__ V, const double* __restrict__ W, double* __restrict__ Ag,
    double* __restrict__ Bg, tb200_acq_spec acq, double* __restrict__ val_out, double* __restrict__ grad_out,
    double* __restrict__ mu_out, double* __restrict__ sigma_out) {
    extern __shared__ double sh[];  // xs[d], red[4][8], res[8], dsum[2][d]
    const int d = spec.d, p = blockIdx.x;
    double* xs = sh;
    double* red = sh + d;
    double* res = red + 32;
    double* dsum = res + 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < d; k += blockDim.x) xs[k] = Xp[(long)p * d + k] / length_scale(spec, theta, k);
    __syncthreads();
    const Hyper hy = unpack_hyper(spec, theta);
    double vv = 0.0, sa = 0.0, sb = 0.0;
    for (int j = threadIdx.x; j < NP; j += blockDim.x) {
        const double v = V[(long)j * GRAD_PMAX + p];
        vv = fma(v, v, vv);
        double r2 = 0.0;
        for (int k = 0; k < d; ++k) {
            const double df = xs[k] - XsT[(long)k * NP + j];
            r2 += df * df;
        }
        const double h = (j < n) ? hy.sf2 * shape_grad_fn(spec.kind, r2) : 0.0;
        const double a = alpha[j] * h, b = W[(long)j * GRAD_PMAX + p] * h;
        Ag[(long)p * NP + j] = a;
        Bg[(long)p * NP + j] = b;
        sa += a;
        sb += b;
    }
    vv = warp_sum(vv); sa = warp_sum(sa); sb = warp_sum(sb);
    if (lane == 0) { red[warp] = vv; red[8 + warp] = sa; red[16 + warp] = sb; }
    __syncthreads();  // also orders this CTA's Ag/Bg writes before the reads below
    if (threadIdx.x == 0) {
        double t0 = 0.0, t1 = 0.0, t2 = 0.0;
        for (int w = 0; w < 8; ++w) { t0 += red[w]; t1 += red[8 + w]; t2 += red[16 + w]; }
        double ms = 0.0;
        for (int t = 0; t < nsplit; ++t) ms += mean[(long)t * mean_stride + p];
        res[0] = t0; res[1] = t1; res[2] = t2; res[3] = ms;
    }
    __syncthreads();
    // per-dimension dot products sum_j A_j xs_j[k], sum_j B_j xs_j[k]: one warp per dimension (strided)
    for (int k = warp; k < d; k += 8) {
        double da = 0.0, db = 0.0;
        for (int j = lane; j < NP; j += 32) {
            const double xt = XsT[(long)k * NP + j];
            da = fma(Ag[(long)p * NP + j], xt, da);
            db = fma(Bg[(long)p * NP + j], xt, db);
        }
        da = warp_sum(da); db = warp_sum(db);
        if (lane == 0) { dsum[k] = da; dsum[d + k] = db; }
    }
    __syncthreads();
    const double prior_var = hy.sf2 + hy.noise;
    double var = prior_var - res[0];
    const bool clamped = var <= 0.0;
    if (var < 0.0) var = 0.0;
    var *= acq.y_std * acq.y_std;
    const double sigma = sqrt(var);
    const double mu = acq.y_std * res[3] + acq.y_mean;
    double a, cm, cs;  // value, d a / d mu, d a / d sigma
    if (acq.acq == TB200_UCB) {
        if (isinf(acq.param)) { a = sigma; cm = 0.0; cs = 1.0; }
        else { a = acq.sign * mu + acq.param * sigma; cm = acq.sign; cs = acq.param; }
    } else if (sigma != 0.0) {
        const double diff = acq.sign * (mu - acq.incumbent) - acq.param;
        const double Z = diff / sigma;
        if (acq.acq == TB200_PI) {
            a = norm_cdf(Z);
            cm = norm_pdf(Z) * acq.sign / sigma;
            cs = -norm_pdf(Z) * diff / (sigma * sigma);
        } else {
            a = diff * norm_cdf(Z) + sigma * norm_pdf(Z);
            cm = norm_cdf(Z) * acq.sign;  // dEI/dmu = s Phi(Z), dEI/dsigma = phi(Z)
            cs = norm_pdf(Z);
        }
    } else {
        a = 0.0; cm = 0.0; cs = 0.0;
    }
    if (threadIdx.x == 0) {
        val_out[p] = a;
        if (mu_out) mu_out[p] = mu;
        if (sigma_out) sigma_out[p] = sigma;
    }
    for (int k = threadIdx.x; k < d; k += blockDim.x) {
        const double il = 1.0 / length_scale(spec, theta, k);
        const double dm = -(xs[k] * res[1] - dsum[k]) * il;              // d(mean_n)/dx_k
        const double dvn = 2.0 * (xs[k] * res[2] - dsum[d + k]) * il;     // d(var_n)/dx_k
        const double dmu = acq.y_std * dm;
        const double dsig = (clamped || sigma == 0.0) ? 0.0 : acq.y_std * acq.y_std * dvn / (2.0 * sigma);
        grad_out[(long)p * d + k] = cm * dmu + cs * dsig;
    }
}

// probe of the per-model slice cut-off (tb200_i8_probe_points): P training points evenly spaced through X, then the
// midpoints between each of them and its successor
__global__ void probe_points_kernel(const double* __restrict__ X, int n, int d, int P, double* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= 2 * P * d) return;
    const int q = e / d, k = e - q * d;
    const int pnt = q < P ? q : q - P;
    const long i0 = P > 1 ? ((long)pnt * (n - 1) + (P - 1) / 2) / (P - 1) : 0;
    const long i1 = (i0 + 1) % n;
    const double a = X[i0 * d + k];
    out[e] = q < P ? a : 0.5 * (a + X[i1 * d + k]);
}

// out[0] = max_i |a_i^2 - b_i^2|  (one CTA; NaN propagates as +inf so a broken probe never selects the cheap engine)
__global__ void __launch_bounds__(256) max_sqdiff_kernel(const double* __restrict__ a, const double* __restrict__ b,
                                                         int64_t m, double* __restrict__ out) {
    __shared__ double red[8];
    double mx = 0.0;
    for (int64_t i = threadIdx.x; i < m; i += blockDim.x) {
        const double dv = fabs(a[i] * a[i] - b[i] * b[i]);
        mx = (dv != dv) ? INFINITY : fmax(mx, dv);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t = fmax(t, red[w]);
        out[0] = t;
    }
}

__global__ void best_reset_kernel(double* best_val, int64_t* best_idx, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        best_val[i] = -INFINITY;
        best_idx[i] = -1;
    }
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

// diagnostic trace (tb200_trace_enable): an event before and after every kernel of tb200_acquire, on the
// stream it is launched on, WITHOUT changing the two-stream schedule -- shows what actually overlaps
static bool g_trace = false;
struct TraceRec { cudaEvent_t e; int tag; };
static std::vector<TraceRec> g_trace_recs;
static void trace_mark(int tag, cudaStream_t st) {
    if (!g_trace || g_trace_recs.size() >= 4096) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    g_trace_recs.push_back({e, tag});
}

int fit_streams_destroy();  // fit.cu

struct AcqBuffers {
    double* cand;     // [batch][d]   generated candidates
    double* Ks;       // [batch][NP]
    double* mean;     // [batch]
    double* partial;  // [NB][batch]
    double* best_val; // [1] running best of this slot
    int64_t* best_idx;
    double* mind;     // [batch]      smallest squared distance to an excluded point (tb200_exclusion)
    unsigned* sched;  // [2]          tile / finished-CTA counters of the persistent variance kernel (shared by the slots)
};

static size_t a256(size_t b) { return (b + 255) / 256 * 256; }

// Two buffer slots and two internal streams split BY ROLE: a high-priority stream runs the small kernels (candidate
// generation, K* + mean, epilogue) and a low-priority stream runs the variance kernels back to back.  The block
// scheduler dispatches a later grid only once every CTA of the earlier grids has been placed, so with one stream
// per slot the small kernels of batch b+1 queued behind the 1024-CTA variance grid of batch b and nothing overlapped
// (tools/trace_acquire.py); with the priority split their CTAs are placed first and share the SMs with the resident
// variance CTAs (tensor pipe vs FP64 pipe).  Order in the high-priority stream: K*_0 K*_1 [G_0] E_0 K*_2 [G_1] E_1 ...
// ([G_b] = wait for the variance kernel of batch b), which also protects the slot buffers: K*_{b+2} overwrites
// Ks/mean of slot b%2 only after G_b and E_b, and G_{b+2} (which waits for K*_{b+2}) overwrites `partial` after E_b.
constexpr int ACQ_SLOTS = 3;
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
        double* mean = take(sizeof(double) * (KSPLIT_MAX > KS_STAGED_SPLITS ? KSPLIT_MAX : KS_STAGED_SPLITS) * batch);
        double* partial = take(sizeof(double) * (use_i8 ? NP / i8::TN : NB) * batch);
        double* mind = take(sizeof(double) * batch);
        if (ab) { ab[s].cand = cand; ab[s].Ks = Ks; ab[s].mean = mean; ab[s].partial = partial; ab[s].mind = mind; }
    }
    // running best per epilogue CTA, all slots contiguous (one reset, one merge)
    const size_t nb = (size_t)(batch / EPI_THREADS);
    double* bv = take(sizeof(double) * ACQ_SLOTS * nb);
    int64_t* bi = reinterpret_cast<int64_t*>(take(sizeof(int64_t) * ACQ_SLOTS * nb));
    unsigned* sched = reinterpret_cast<unsigned*>(take(256));
    if (ab)
        for (int s = 0; s < ACQ_SLOTS; ++s) {
            ab[s].best_val = bv + s * nb;
            ab[s].best_idx = bi + s * nb;
            ab[s].sched = sched;
        }
    return off;
}

// merge the per-CTA bests of the slots into the caller's running best (read-modify-write, one warp; the
// (value, lowest index) order makes the result independent of how the entries are visited)
__global__ void best_merge_kernel(const double* v, const int64_t* idx, int count, double* best_val,
                                  int64_t* best_idx) {
    double bv = -INFINITY;
    int64_t bi = -1;
    for (int e = threadIdx.x; e < count; e += 32)
        if (idx[e] >= 0 && better(v[e], idx[e], bv, bi)) { bv = v[e]; bi = idx[e]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi >= 0 && better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (threadIdx.x == 0 && bi >= 0 && better(bv, bi, *best_val, *best_idx)) {
        *best_val = bv;
        *best_idx = bi;
    }
}

// internal streams / events, one set per device, created on first use
struct AcqStreams {
    bool ready = false;
    cudaStream_t small, var;                       // high priority: K*/epilogue; low priority: variance kernels
    cudaEvent_t fork, join_small, join_var;
    cudaEvent_t kdone[ACQ_SLOTS], vdone[ACQ_SLOTS];
};
static AcqStreams g_streams[MAX_DEVICES];
static std::mutex g_streams_mutex;

static int acq_streams_for_device(AcqStreams** out) {
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CHECK(dev >= 0 && dev < MAX_DEVICES, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_streams_mutex);
    AcqStreams& a = g_streams[dev];
    if (!a.ready) {
        int lo = 0, hi = 0;  // numerically lower = higher priority
        TB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        // the small kernels get the high priority: their few CTAs are placed at once and the variance grid keeps
        // filling the rest of each SM (the flipped assignment measured slower in round 1)
        TB_CUDA(cudaStreamCreateWithPriority(&a.small, cudaStreamNonBlocking, hi));
        TB_CUDA(cudaStreamCreateWithPriority(&a.var, cudaStreamNonBlocking, lo));
        for (int i = 0; i < ACQ_SLOTS; ++i) {
            TB_CUDA(cudaEventCreateWithFlags(&a.kdone[i], cudaEventDisableTiming));
            TB_CUDA(cudaEventCreateWithFlags(&a.vdone[i], cudaEventDisableTiming));
        }
        TB_CUDA(cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming));
        TB_CUDA(cudaEventCreateWithFlags(&a.join_small, cudaEventDisableTiming));
        TB_CUDA(cudaEventCreateWithFlags(&a.join_var, cudaEventDisableTiming));
        a.ready = true;
    }
    *out = &a;
    return 0;
}

// All kernels that share the SMs with the variance kernel ask for the same (maximum) shared-memory carve-out: an SM
// whose resident CTAs were launched with a small carve-out cannot take a 174 KB variance CTA until it has drained and
// been reconfigured, which serialised the "concurrent" K* and variance kernels (tools/trace_acquire.py).
static DeviceOnce g_carveout_once;
static int configure_carveout_device() {
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CUDA(cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev));
    const int mx = (int)cudaSharedmemCarveoutMaxShared;
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_kernel<TB200_RBF>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_kernel<TB200_MATERN12>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_kernel<TB200_MATERN32>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_kernel<TB200_MATERN52>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
#define TB_STAGED_ATTR(KIND)                                                                                            \
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_staged_kernel<KIND, 8>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));  \
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_staged_kernel<KIND, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                                 (int)kstar_staged_smem(TB200_MAX_D, 8)));                                               \
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_staged_kernel<KIND, 16>, cudaFuncAttributePreferredSharedMemoryCarveout, mx)); \
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_i8_staged_kernel<KIND, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                 (int)kstar_staged_smem(TB200_MAX_D, 16)))
    TB_STAGED_ATTR(TB200_RBF);
    TB_STAGED_ATTR(TB200_MATERN12);
    TB_STAGED_ATTR(TB200_MATERN32);
    TB_STAGED_ATTR(TB200_MATERN52);
#undef TB_STAGED_ATTR
    TB_CUDA(cudaFuncSetAttribute(kstar_mean_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    TB_CUDA(cudaFuncSetAttribute(acquire_epilogue_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    TB_CUDA(cudaFuncSetAttribute(generate_candidates_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    TB_CUDA(cudaFuncSetAttribute(best_reset_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    return 0;
}
static int configure_carveout() { return device_once(g_carveout_once, configure_carveout_device); }

template <int V>
static int run_mma_list(int kblocks, int planes, const uint8_t* src, long src_bytes, int sms, long long* d_cyc) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(i8::i8_mma_list_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(i8::STAGE_BYTES + 1024)));
            TB_CUDA(cudaFuncSetAttribute(i8::i8_mma_pipe_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)i8::SMEM_BYTES));
            return 0;
        }))
        return -1;
    if (planes < 0) {
        TB_LAUNCH(i8::i8_mma_list_kernel<V>, dim3(sms), dim3(64), i8::STAGE_BYTES + 1024, 0, kblocks, d_cyc);
    } else {
        TB_LAUNCH(i8::i8_mma_pipe_kernel<V>, dim3(sms), dim3(64), i8::SMEM_BYTES, 0, kblocks, planes, src, src_bytes,
                  d_cyc);
    }
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
    TB_LAUNCH(best_reset_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, best_val, best_idx, 1);
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
                        void* work, size_t work_bytes, void* stream, const tb200_exclusion* excl = nullptr) {
    const bool use_i8 = f8.linvS != nullptr;
    if (excl && (excl->n <= 0 || excl->tol < 0.0)) excl = nullptr;  // nothing to exclude / disabled (fallback.py:51-52)
    TB_CHECK(!excl || excl->X, "exclusion: X is NULL");
    TB_CHECK(spec && acq && XsT && alpha && theta && best_val && best_idx && work, "NULL pointer argument");
    TB_CHECK(use_i8 ? f8.rowscale != nullptr : Linv != nullptr, "NULL factor pointer");
    TB_CHECK(spec->kind >= TB200_RBF && spec->kind <= TB200_MATERN52, "unknown kernel kind");
    TB_CHECK(spec->d >= 1 && spec->d <= TB200_MAX_D, "d out of range (1..64)");
    TB_CHECK(acq->acq >= TB200_EI && acq->acq <= TB200_UCB, "unknown acquisition function");
    TB_CHECK(!use_i8 || (f8.gmax >= 7 && f8.gmax <= 9), "i8 engine: gmax must be 7, 8 or 9");
    TB_CHECK(!use_i8 || pad_n(n) <= TB200_I8_MAX_N, "i8 engine: n above TB200_I8_MAX_N");
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

    // one batch, or GEMM timing requested: everything on the caller's stream
    const int64_t nbatches = (m + batch - 1) / batch;
    const int slots = (nbatches < 2 || g_time_gemm) ? 1 : ACQ_SLOTS;
    if (configure_carveout()) return -1;
    AcqStreams* as = nullptr;
    cudaStream_t s_small = st, s_var = st;
    if (slots > 1) {
        if (acq_streams_for_device(&as)) return -1;
        s_small = as->small;
        s_var = as->var;
        TB_CUDA(cudaEventRecord(as->fork, st));
        TB_CUDA(cudaStreamWaitEvent(s_small, as->fork, 0));
        TB_CUDA(cudaStreamWaitEvent(s_var, as->fork, 0));
    }
    TB_LAUNCH(best_reset_kernel, dim3(1), dim3(64), 0, s_small, ab[0].best_val, ab[0].best_idx,
              (int)(ACQ_SLOTS * (batch / EPI_THREADS)));
    // the persistent variance kernel leaves its counters zeroed for the next launch; the workspace is the caller's
    if (use_i8) TB_CUDA(cudaMemsetAsync(ab[0].sched, 0, 2 * sizeof(unsigned), s_var));

    const size_t ksm = sizeof(double) * (CPB * d + 8 * CPB);
    // split the training points so one K* launch is ~2k CTAs (3 resident per SM): NB/ksplit tiles each
    int ksplit = NB < KSPLIT_MAX ? NB : KSPLIT_MAX;
    int jchunk = (NB + ksplit - 1) / ksplit * TILE;
    // the INT8 K* kernel covers 256 points per pass of its 128 threads: a shorter chunk would idle half of them
    if (use_i8 && jchunk < 2 * TILE) jchunk = 2 * TILE;
    ksplit = (NP + jchunk - 1) / jchunk;
    // staged K* kernel (training-point chunk in shared memory): 1024 <= n <= 4096 and at least two CTAs of it fit into
    // the ~57 KB the variance kernel leaves on an SM.  Measured
    // (profiles/r1_kstar_ncu_full.md): n=4096, d=10 (variance-bound step) 388.8 -> 376.2 ms per 2^21 candidates.  (In
    // round 1, with 256-point chunks, four direct-load CTAs per SM beat three staged ones at n=1024, d=6: 5.49e7 vs
    // 5.00e7 /s; with 128-point chunks the staged kernel wins there too, see below.)
    int staged_ctas = 0, staged_nc = 8;
    if (use_i8 && NP >= 1024) {
        // Three CTAs of the 128-point chunk where they fit (d <= 12) -- with the integer-pipe epilogue of the persistent
        // variance kernel the FP64 pipe belongs to K* alone and a third CTA pays: 308.0 vs 310.8 ms per step at n=4096,
        // d=10 (a fourth: 308.2; round 1, when the variance epilogue competed for the FP64 pipe, a third CTA lost,
        // 379.5 vs 375.5 ms).  Otherwise two CTAs: the 256-point chunk, or the 128-point chunk where only that fits twice.
        auto fit_of = [&](int nc, int cap) {
            if ((NP + ks_chunk(nc) - 1) / ks_chunk(nc) > KS_STAGED_SPLITS) return 0;
            const int f = (int)((57 * 1024) / (kstar_staged_smem(d, nc) + 1024));
            return f < cap ? f : cap;
        };
        // n < 2048 (the step is K*-bound there): up to four -- cfg2 (n=1024, d=6) 6.60e7 candidates/s against 6.55e7 with
        // three and 6.39e7 with four CTAs of the direct-load kernel (tools/gpu_round2_z.sh)
        const int cap = NP < 2048 ? 4 : 3;
        if (fit_of(16, cap) >= 3) {
            staged_nc = 16;
            staged_ctas = fit_of(16, cap);
        } else {
            const int f8 = fit_of(8, 2), f16 = fit_of(16, 2);
            staged_nc = (f8 < 2 && f16 >= 2) ? 16 : 8;
            const int f = staged_nc == 16 ? f16 : f8;
            if (f >= 2) staged_ctas = f;
        }
    }
    const int staged_splits = (NP + ks_chunk(staged_nc) - 1) / ks_chunk(staged_nc);
    const int mean_splits = staged_ctas ? staged_splits : ksplit;

    struct Batch { int sl; int64_t off, mb, mbp; };
    auto batch_of = [&](int64_t bno) {
        Batch q;
        q.sl = (int)(bno % slots);
        q.off = bno * batch;
        q.mb = (m - q.off < batch) ? (m - q.off) : batch;  // valid candidates
        q.mbp = (q.mb + 127) / 128 * 128;                  // padded to GEMM tiles
        return q;
    };
    // candidates (optional) + K* rows + mean of one batch, on the small-kernel stream
    auto front = [&](int64_t bno) -> int {
        const Batch q = batch_of(bno);
        const AcqBuffers& w = ab[q.sl];
        const double* xc = Xc ? Xc + q.off * d : nullptr;
        if (mode != TB200_CAND_GIVEN) {
            const CandGen g = make_gen(*cand, d);
            const int64_t total = q.mb * d;
            TB_LAUNCH(generate_candidates_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, s_small, g, d,
                      index_base + q.off, q.mb, w.cand);
            xc = w.cand;
        }
        if (excl) {
            double* mo = excl->min_out ? excl->min_out + q.off : w.mind;
            if (launch_min_sqdist(xc, q.mb, excl->X, excl->n, d, mo, nullptr, s_small)) return -1;
        }
        trace_mark(q.sl * 16 + 0, s_small);
        if (use_i8) {
            const size_t ksm8 = sizeof(double) * (CPB * d + (KS_THREADS / 32) * CPB);
            const int ngroups = (int)(q.mbp / CPB);
            const int kctas = g_sm_count * (staged_ctas ? staged_ctas : KS_CTAS_PER_SM);
            const int st_groups = (int)(q.mbp / staged_nc);
            const int kitems = staged_ctas ? st_groups * staged_splits : ngroups * ksplit;
            const dim3 kgrid((unsigned)(kitems < kctas ? kitems : kctas));
            const size_t ksm_st = kstar_staged_smem(d, staged_nc);
            const int kplanes = i8::planes_used(f8.gmax);
#define TB_KSTAR_ST(KIND, NC)                                                                                          \
    TB_LAUNCH((kstar_mean_i8_staged_kernel<KIND, NC>), kgrid, dim3(KS_THREADS), ksm_st, s_small, xc, q.mb, XsT, n, NP, \
              *spec, theta, alpha, reinterpret_cast<uint8_t*>(w.Ks), slice_stride, kplanes, w.mean, (long)batch,       \
              st_groups, staged_splits)
#define TB_KSTAR_I8(KIND)                                                                                              \
    if (staged_ctas && staged_nc == 16) {                                                                              \
        TB_KSTAR_ST(KIND, 16);                                                                                         \
    } else if (staged_ctas) {                                                                                          \
        TB_KSTAR_ST(KIND, 8);                                                                                          \
    } else {                                                                                                           \
        TB_LAUNCH(kstar_mean_i8_kernel<KIND>, kgrid, dim3(KS_THREADS), ksm8, s_small, xc, q.mb, XsT, n, NP, *spec,     \
                  theta, alpha, reinterpret_cast<uint8_t*>(w.Ks), slice_stride, kplanes, w.mean, jchunk, (long)batch,  \
                  ngroups, ksplit);                                                                                    \
    }
            switch (spec->kind) {
                case TB200_RBF: TB_KSTAR_I8(TB200_RBF); break;
                case TB200_MATERN12: TB_KSTAR_I8(TB200_MATERN12); break;
                case TB200_MATERN32: TB_KSTAR_I8(TB200_MATERN32); break;
                default: TB_KSTAR_I8(TB200_MATERN52); break;
            }
#undef TB_KSTAR_I8
#undef TB_KSTAR_ST
        } else {
            TB_LAUNCH(kstar_mean_kernel, dim3((unsigned)(q.mbp / CPB), ksplit), dim3(256), ksm, s_small, xc, q.mb, XsT, n,
                      NP, *spec, theta, alpha, w.Ks, w.mean, jchunk, (long)batch);
        }
        trace_mark(q.sl * 16 + 4, s_small);
        return 0;
    };
    // the variance contraction of one batch, on the variance stream
    auto variance = [&](int64_t bno) -> int {
        const Batch q = batch_of(bno);
        const AcqBuffers& w = ab[q.sl];
        if (g_time_gemm) TB_CUDA(cudaEventRecord(next_event(), s_var));
        trace_mark(q.sl * 16 + 1, s_var);
        if (use_i8) {
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
            ip.candTiles = (int)(q.mbp / i8::TM);
            ip.sched = w.sched;
            const int rc = f8.gmax == 7   ? i8::launch_colsumsq_i8<7>(ip, g_sm_count, s_var)
                           : f8.gmax == 8 ? i8::launch_colsumsq_i8<8>(ip, g_sm_count, s_var)
                                          : i8::launch_colsumsq_i8<9>(ip, g_sm_count, s_var);
            if (rc) return -1;
        } else {
            GemmParams gp = gemm_params(Linv, NP, w.Ks, NP, w.partial, batch, NB, (int)(q.mbp / 128), NP, 1.0, 0.0);
            gp.order = ORDER_RECT_HEAVY_LAST_ROW_FIRST;
            gp.khi_mode = KHI_BY_TI;
            gp.a_tri_diag = 1;  // Linv is lower triangular
            gp.epi = EPI_COLSUMSQ;
            if (launch_gemm<false, true>(gp, 1, 1, s_var)) return -1;
        }
        trace_mark(q.sl * 16 + 2, s_var);
        if (g_time_gemm) TB_CUDA(cudaEventRecord(next_event(), s_var));
        return 0;
    };
    auto epilogue = [&](int64_t bno) -> int {
        const Batch q = batch_of(bno);
        const AcqBuffers& w = ab[q.sl];
        TB_LAUNCH(acquire_epilogue_kernel, dim3((unsigned)(q.mbp / EPI_THREADS)), dim3(EPI_THREADS), 0, s_small, w.mean,
                  mean_splits, w.partial, ntiles, (long)batch, q.mbp, q.mb, q.off, index_base, *spec, theta, *acq, mu_out,
                  sigma_out, acq_out, w.best_val, w.best_idx,
                  excl ? (excl->min_out ? excl->min_out + q.off : w.mind) : nullptr, excl ? excl->tol : 0.0);
        trace_mark(q.sl * 16 + 3, s_small);
        return 0;
    };

    if (slots == 1) {
        for (int64_t bno = 0; bno < nbatches; ++bno)
            if (front(bno) || variance(bno) || epilogue(bno)) return -1;
    } else {
        auto pipeline = [&]() -> int {
            for (int64_t bno = 0; bno < slots && bno < nbatches; ++bno) {
                if (front(bno)) return -1;
                TB_CUDA(cudaEventRecord(as->kdone[bno % slots], s_small));
            }
            for (int64_t bno = 0; bno < nbatches; ++bno) {
                const int sl = (int)(bno % slots);
                TB_CUDA(cudaStreamWaitEvent(s_var, as->kdone[sl], 0));
                if (variance(bno)) return -1;
                TB_CUDA(cudaEventRecord(as->vdone[sl], s_var));
                TB_CUDA(cudaStreamWaitEvent(s_small, as->vdone[sl], 0));
                if (epilogue(bno)) return -1;
                if (bno + slots < nbatches) {
                    if (front(bno + slots)) return -1;
                    TB_CUDA(cudaEventRecord(as->kdone[sl], s_small));
                }
            }
            TB_CUDA(cudaEventRecord(as->join_small, s_small));
            TB_CUDA(cudaEventRecord(as->join_var, s_var));
            TB_CUDA(cudaStreamWaitEvent(st, as->join_small, 0));
            TB_CUDA(cudaStreamWaitEvent(st, as->join_var, 0));
            return 0;
        };
        if (pipeline()) {
            // a failed launch or CUDA call part-way: the two internal streams were forked from the caller's stream and
            // may still be reading the workspace the caller is about to free.  Drain them before reporting the error
            // (the message of the first failure is already in g_err).
            cudaStreamSynchronize(s_small);
            cudaStreamSynchronize(s_var);
            return -1;
        }
    }
    TB_LAUNCH(best_merge_kernel, dim3(1), dim3(32), 0, st, ab[0].best_val, ab[0].best_idx,
              (int)(ACQ_SLOTS * (batch / EPI_THREADS)), best_val, best_idx);
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

// the same two calls with an exclusion set: candidates whose squared distance to any of excl->X's rows
// is <= excl->tol (turbo/utils.py:111-130) are evaluated but never selected
int tb200_acquire_excl(const tb200_exclusion* excl, const double* Xc, int64_t m, int64_t index_base,
                       const tb200_cand_spec* cand, const double* XsT, const double* Linv, const double* alpha, int n,
                       const tb200_kernel_spec* spec, const double* theta, const tb200_acq_spec* acq, double* mu_out,
                       double* sigma_out, double* acq_out, double* best_val, int64_t* best_idx, int64_t batch,
                       void* work, size_t work_bytes, void* stream) {
    const I8Factor none{nullptr, nullptr, 0};
    return acquire_impl(Xc, m, index_base, cand, XsT, Linv, none, alpha, n, spec, theta, acq, mu_out, sigma_out, acq_out,
                        best_val, best_idx, batch, work, work_bytes, stream, excl);
}

int tb200_acquire_i8_excl(const tb200_exclusion* excl, const double* Xc, int64_t m, int64_t index_base,
                          const tb200_cand_spec* cand, const double* XsT, const uint8_t* linvS, const double* rowscale,
                          int gmax, const double* alpha, int n, const tb200_kernel_spec* spec, const double* theta,
                          const tb200_acq_spec* acq, double* mu_out, double* sigma_out, double* acq_out,
                          double* best_val, int64_t* best_idx, int64_t batch, void* work, size_t work_bytes,
                          void* stream) {
    TB_CHECK(linvS && rowscale, "NULL pointer argument");
    const I8Factor f8{linvS, rowscale, gmax};
    return acquire_impl(Xc, m, index_base, cand, XsT, nullptr, f8, alpha, n, spec, theta, acq, mu_out, sigma_out,
                        acq_out, best_val, best_idx, batch, work, work_bytes, stream, excl);
}

int tb200_min_sqdist(const double* Xq, int64_t m, const double* X, int n, int d, double* min_out, int64_t* arg_out,
                     void* stream) {
    TB_CHECK((m <= 0 || (Xq && min_out)) && (n <= 0 || X), "NULL pointer argument");
    TB_CHECK(d >= 1 && d <= TB200_MAX_D, "d out of range (1..64)");
    TB_CHECK(n >= 0, "n negative");
    return launch_min_sqdist(Xq, m, X, n, d, min_out, arg_out, (cudaStream_t)stream);
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
    TB_CHECK(NP <= TB200_I8_MAX_N, "i8 engine: n above TB200_I8_MAX_N");
    cudaStream_t st = (cudaStream_t)stream;
    TB_LAUNCH(linv_rowscale_kernel, dim3(NP / 8), dim3(256), 0, st, Linv, NP, *spec, theta, rowscale);
    const long total = (long)NP * (NP / 16);
    TB_LAUNCH(linv_slice_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, Linv, NP, rowscale, linvS,
              (long)NP * NP);
    return 0;
}

// the per-epilogue-CTA running bests of the LAST tb200_acquire(_i8) call on `work` (entry e covers the candidates
// whose position within their batch is in [128 e', 128 e'+128) of one buffer slot): distinct good candidates, used as
// starting points of the gradient polish (the reference takes the best `start_from_best` random samples,
// auxiliary_optimisers.py:73-75).  Unused entries have index -1.
int tb200_acquire_local_best_count(int64_t batch, int n) {
    if (batch <= 0) batch = tb200_default_batch(n);
    return (int)(ACQ_SLOTS * (batch / EPI_THREADS));
}

int tb200_acquire_local_bests(void* work, int n, int d, int64_t batch, int use_i8, double* vals_out, int64_t* idx_out,
                              void* stream) {
    TB_CHECK(work && vals_out && idx_out, "NULL pointer argument");
    if (batch <= 0) batch = tb200_default_batch(n);
    AcqBuffers ab[ACQ_SLOTS];
    acquire_layout(n, d, batch, ab, (char*)work, use_i8 != 0);
    const size_t cnt = (size_t)ACQ_SLOTS * (batch / EPI_THREADS);
    TB_CUDA(cudaMemcpyAsync(vals_out, ab[0].best_val, sizeof(double) * cnt, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    TB_CUDA(cudaMemcpyAsync(idx_out, ab[0].best_idx, sizeof(int64_t) * cnt, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}

size_t tb200_acquire_grad_workspace_bytes(int n, int d) {
    (void)d;
    const size_t NP = pad_n(n);
    return a256(sizeof(double) * GRAD_PMAX * NP) * 5 + a256(sizeof(double) * KSPLIT_MAX * GRAD_PMAX);
}

int tb200_acquire_grad(const double* Xp, int P, const double* XsT, const double* Linv, const double* alpha, int n,
                       const tb200_kernel_spec* spec, const double* theta, const tb200_acq_spec* acq, double* val_out,
                       double* grad_out, double* mu_out, double* sigma_out, void* work, size_t work_bytes, void* stream) {
    TB_CHECK(Xp && XsT && Linv && alpha && spec && theta && acq && val_out && grad_out && work, "NULL pointer argument");
    TB_CHECK(spec->kind >= TB200_RBF && spec->kind <= TB200_MATERN52, "unknown kernel kind");
    TB_CHECK(spec->d >= 1 && spec->d <= TB200_MAX_D, "d out of range (1..64)");
    TB_CHECK(acq->acq >= TB200_EI && acq->acq <= TB200_UCB, "unknown acquisition function");
    TB_CHECK(P >= 0 && P <= GRAD_PMAX, "at most 128 points per call");
    if (P == 0) return 0;
    TB_CHECK(work_bytes >= tb200_acquire_grad_workspace_bytes(n, spec->d), "workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int NP = pad_n(n), NB = NP / TILE, d = spec->d;
    char* base = (char*)work;
    const size_t mat = a256(sizeof(double) * GRAD_PMAX * NP);
    double* Ks = (double*)base;
    double* V = (double*)(base + mat);
    double* W = (double*)(base + 2 * mat);
    double* Ag = (double*)(base + 3 * mat);
    double* Bg = (double*)(base + 4 * mat);
    double* mean = (double*)(base + 5 * mat);
    // K* rows + mean partials of the (padded) 128 points
    const size_t ksm = sizeof(double) * (CPB * d + 8 * CPB);
    int ksplit = NB < KSPLIT_MAX ? NB : KSPLIT_MAX;
    const int jchunk = (NB + ksplit - 1) / ksplit * TILE;
    ksplit = (NP + jchunk - 1) / jchunk;
    TB_LAUNCH(kstar_mean_kernel, dim3(GRAD_PMAX / CPB, ksplit), dim3(256), ksm, st, Xp, (int64_t)P, XsT, n, NP, *spec, theta,
              alpha, Ks, mean, jchunk, (long)GRAD_PMAX);
    // V = Linv Ks^T (NP x 128), W = Linv^T V
    GemmParams g1 = gemm_params(Linv, NP, Ks, NP, V, GRAD_PMAX, NB, 1, NP, 1.0, 0.0);
    g1.khi_mode = KHI_BY_TI;
    g1.a_tri_diag = 1;
    if (launch_gemm<false, true>(g1, 1, 1, st)) return -1;
    GemmParams g2 = gemm_params(Linv, NP, V, GRAD_PMAX, W, GRAD_PMAX, NB, 1, NP, 1.0, 0.0);
    g2.klo_mode = KLO_BY_TI;
    if (launch_gemm<true, false>(g2, 1, 1, st)) return -1;
    const size_t gsm = sizeof(double) * (d + 32 + 8 + 2 * d);
    TB_LAUNCH(acq_grad_kernel, dim3(P), dim3(256), gsm, st, Xp, P, XsT, n, NP, *spec, theta, alpha, mean, ksplit,
              (long)GRAD_PMAX, V, W, Ag, Bg, *acq, val_out, grad_out, mu_out, sigma_out);
    return 0;
}

int tb200_i8_probe_points(const double* X, int n, int d, int P, double* out, void* stream) {
    TB_CHECK(X && out, "NULL pointer argument");
    TB_CHECK(n >= 1 && d >= 1 && d <= TB200_MAX_D && P >= 1 && P <= n, "tb200_i8_probe_points: 1 <= P <= n required");
    const int total = 2 * P * d;
    TB_LAUNCH(probe_points_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, (cudaStream_t)stream, X, n, d, P,
              out);
    return 0;
}

int tb200_max_sqdiff(const double* a, const double* b, int64_t m, double* out, void* stream) {
    TB_CHECK(a && b && out && m >= 0, "NULL pointer argument");
    TB_LAUNCH(max_sqdiff_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, a, b, m, out);
    return 0;
}

int tb200_i8_peak(double ms, int random_digits, double* int8_tops) {
    return tb200_i8_mma_rate(ms, random_digits, i8::PEAK_N, int8_tops);
}

int tb200_i8_mma_rate(double ms, int random_digits, int n_cols, double* int8_tops) {
    TB_CHECK(int8_tops, "NULL pointer argument");
    TB_CHECK(n_cols >= 16 && n_cols <= i8::PEAK_N && n_cols % 16 == 0, "n_cols: a multiple of 16 in 16..256");
    int dev = 0, sms = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(i8::i8_peak_kernel<i8::PEAK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)i8::PEAK_SMEM));
            return 0;
        }))
        return -1;
    cudaEvent_t e0, e1;
    TB_CUDA(cudaEventCreate(&e0));
    TB_CUDA(cudaEventCreate(&e1));
    int batches = 64;
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        TB_CUDA(cudaEventRecord(e0, 0));
        TB_LAUNCH(i8::i8_peak_kernel<i8::PEAK_N>, dim3(sms), dim3(64), i8::PEAK_SMEM, 0, batches, random_digits, n_cols);
        TB_CUDA(cudaEventRecord(e1, 0));
        TB_CUDA(cudaEventSynchronize(e1));
        float t = 0.f;
        TB_CUDA(cudaEventElapsedTime(&t, e0, e1));
        // one MMA = M128 x N x K32 multiply-accumulates, 2 operations each
        const double ops = 2.0 * i8::TM * n_cols * 32.0 * i8::PEAK_BATCH * (double)batches * sms;
        const double tops = ops / (t * 1e-3) / 1e12;
        if (rep > 0 && tops > best) best = tops;
        if (t < ms) batches = (int)(batches * (ms / (t > 1e-3 ? t : 1e-3)) * 1.1) + 2;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *int8_tops = best;
    return 0;
}

int tb200_i8_mma_list(int variant, int kblocks, int fill_planes, double* cycles_per_kblock, int* mmas_per_step,
                      int* columns_per_step) {
    TB_CHECK(cycles_per_kblock && mmas_per_step && columns_per_step, "NULL pointer argument");
    TB_CHECK(variant >= 0 && variant < i8::MMA_LIST_VARIANTS && kblocks >= 8, "tb200_i8_mma_list: variant 0..13, >= 8 k-blocks");
    TB_CHECK(fill_planes >= -1 && fill_planes <= i8::S, "tb200_i8_mma_list: fill_planes -1 (resident operands) .. 7");
    int dev = 0, sms = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    long long* d_cyc = nullptr;
    uint8_t* d_src = nullptr;
    const long src_bytes = 64l << 20;  // L2-resident scratch the fills read from (the values do not matter)
    TB_CUDA(cudaMalloc(&d_cyc, sizeof(long long) * sms));
    if (fill_planes > 0) {
        TB_CUDA(cudaMalloc(&d_src, src_bytes));
        TB_CUDA(cudaMemset(d_src, 0x5a, src_bytes));
    }
    double best = 0.0;
    int rc = 0;
    const int skipped = fill_planes < 0 ? 2 : 4;  // k-blocks before the clock is started
    for (int rep = 0; rep < 3 && rc == 0; ++rep) {
        int n = 0, cols = 0;
#define TB_RUN(V)                                                                \
    case V:                                                                      \
        rc = run_mma_list<V>(kblocks, fill_planes, d_src, src_bytes, sms, d_cyc); \
        n = i8::MmaList<V>::n;                                                   \
        for (int i = 0; i < n; ++i) cols += i8::MmaList<V>::ops[i].n;            \
        break
        switch (variant) {
            TB_RUN(0); TB_RUN(1); TB_RUN(2); TB_RUN(3); TB_RUN(4); TB_RUN(5); TB_RUN(6);
            TB_RUN(7); TB_RUN(8); TB_RUN(9); TB_RUN(10); TB_RUN(11); TB_RUN(12); TB_RUN(13);
        }
#undef TB_RUN
        if (rc) break;
        *mmas_per_step = n;
        *columns_per_step = cols;
        std::vector<long long> h(sms);
        if (cudaMemcpy(h.data(), d_cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost) != cudaSuccess) {
            rc = fail(cudaGetErrorString(cudaGetLastError()), __FILE__, __LINE__);
            break;
        }
        double sum = 0.0;
        for (int i = 0; i < sms; ++i) sum += (double)h[i];
        const double per = sum / sms / (double)(kblocks - skipped);
        if (rep == 0 || per < best) best = per;
    }
    cudaFree(d_cyc);
    if (d_src) cudaFree(d_src);
    *cycles_per_kblock = best;
    return rc;
}

int tb200_shutdown(void) {
    {
        std::lock_guard<std::mutex> lock(g_streams_mutex);
        int cur = 0;
        cudaGetDevice(&cur);
        for (int dev = 0; dev < MAX_DEVICES; ++dev) {
            AcqStreams& a = g_streams[dev];
            if (!a.ready) continue;
            cudaSetDevice(dev);
            cudaStreamSynchronize(a.small);
            cudaStreamSynchronize(a.var);
            cudaStreamDestroy(a.small);
            cudaStreamDestroy(a.var);
            for (int i = 0; i < ACQ_SLOTS; ++i) { cudaEventDestroy(a.kdone[i]); cudaEventDestroy(a.vdone[i]); }
            cudaEventDestroy(a.fork); cudaEventDestroy(a.join_small); cudaEventDestroy(a.join_var);
            a.ready = false;
        }
        cudaSetDevice(cur);
    }
    for (auto& e : g_events) cudaEventDestroy(e);
    g_events.clear();
    g_events_used = 0;
    for (auto& r : g_trace_recs) cudaEventDestroy(r.e);
    g_trace_recs.clear();
    return fit_streams_destroy();
}

int tb200_trace_enable(int on) {
    g_trace = on != 0;
    for (auto& r : g_trace_recs) cudaEventDestroy(r.e);
    g_trace_recs.clear();
    return 0;
}

int tb200_trace_read(double* t_ms, int* tags, int cap) {
    TB_CHECK(t_ms && tags, "NULL pointer argument");
    int k = 0;
    for (size_t i = 0; i < g_trace_recs.size() && k < cap; ++i, ++k) {
        TB_CUDA(cudaEventSynchronize(g_trace_recs[i].e));
        float t = 0.f;
        TB_CUDA(cudaEventElapsedTime(&t, g_trace_recs[0].e, g_trace_recs[i].e));
        t_ms[k] = t;
        tags[k] = g_trace_recs[i].tag;
    }
    return k;
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
