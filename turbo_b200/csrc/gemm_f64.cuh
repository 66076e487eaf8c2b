// FP64 tensor-core (DMMA.8x8x4) tile GEMM used by every dense contraction of the path:
// Cholesky panel/trailing updates, triangular inverse, K^-1 = Linv^T Linv, and the predictive
// variance V = Linv * Kstar^T with a fused column sum-of-squares epilogue.
//
// sm_100a has no FP64 kind on tcgen05/TMEM; FP64 tensor work is warp-level mma.sync m8n8k4
// (every larger f64 shape lowers to DMMA.8x8x4 in SASS) and DMMA shares the FP64 pipe with DFMA
// (measured: half DMMA / half DFMA warps = 34 TFLOP/s, no concurrency), so the only lever is to keep
// that one pipe issuing back to back.  128x128x16 CTA tile, 16 warps of 32x32 (4 per scheduler),
// 4-stage cp.async (LDGSTS) pipeline where tile kt+1 is guaranteed landed one iteration early so
// the first fragments of the next k-tile are fetched BEFORE the per-tile barrier: the DMMA pipe
// never waits on an LDS after a barrier (profiles/r1_gemm_sweep.md: 29.8 -> 33.1 TFLOP/s).
#pragma once
#include "common.cuh"

namespace tb200 {

constexpr int G_TM = 128, G_TN = 128, G_TK = 16, G_STAGES = 4, G_THREADS = 512;
constexpr int G_WM = 4, G_WN = 4;  // warp grid; warp tile 32 x 32
constexpr int G_MI = G_TM / G_WM / 8, G_NJ = G_TN / G_WN / 8;
constexpr int G_LDN = G_TK + 4;    // row stride (doubles) of a [row][k] tile: conflict-free fragment loads
constexpr int G_LDT = G_TM + 4;    // row stride of a [k][row] tile
constexpr int G_TILE_ELEMS = (G_TM * G_LDN > G_TK * G_LDT) ? G_TM * G_LDN : G_TK * G_LDT;
constexpr size_t G_SMEM_BYTES = size_t(G_STAGES) * 2 * G_TILE_ELEMS * sizeof(double);

enum : int { EPI_STORE = 0, EPI_COLSUMSQ = 1 };
enum : int { ORDER_RECT = 0, ORDER_RECT_HEAVY_LAST_ROW_FIRST = 1, ORDER_LOWER = 2 };
enum : int { KLO_ZERO = 0, KLO_BY_TJ = 1, KLO_BY_TI = 2 };
enum : int { KHI_FULL = 0, KHI_BY_TI = 1 };

struct GemmParams {
    const double* A;
    const double* B;
    double* C;          // EPI_STORE: C tile; EPI_COLSUMSQ: partial[(ti) * ldc + col]
    long lda, ldb, ldc;
    long sA1, sB1, sC1;  // batch strides for blockIdx.y
    long sA2, sB2, sC2;  // batch strides for blockIdx.z
    int tilesM, tilesN;  // tile counts (M = tilesM*128, N = tilesN*128)
    int K;               // multiple of 16
    double alpha, beta;
    int order, klo_mode, khi_mode, epi;
    int a_tri_diag;      // with KHI_BY_TI: A's diagonal 128-block is lower triangular -> warps skip its zero k-range
    int b_tri;           // K == 128, one column tile: B is a lower-triangular 128-block.  1: stored [n][k] (TB), zero for
                         // k > n -> a warp's columns need k < (wn+1)*32;  2: stored [k][n], zero for k < n -> k >= wn*32
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// copy one operand tile (128 rows x 16 k) into shared memory.
//  TRANS = false: memory is [row][k] (k contiguous)  -> smem [row][G_LDN]
//  TRANS = true : memory is [k][row] (row contiguous) -> smem [k][G_LDT]
template <bool TRANS>
__device__ __forceinline__ void load_tile(double* s, const double* g, long ld, int row0, int k0, int tid) {
    if (!TRANS) {
#pragma unroll
        for (int i = 0; i < G_TM * G_TK / 2 / G_THREADS; ++i) {
            const int c = tid + i * G_THREADS;  // 1024 chunks of 16 B
            const int r = c >> 3, kc = c & 7;
            cp_async16(s + r * G_LDN + kc * 2, g + (long)(row0 + r) * ld + k0 + kc * 2);
        }
    } else {
#pragma unroll
        for (int i = 0; i < G_TM * G_TK / 2 / G_THREADS; ++i) {
            const int c = tid + i * G_THREADS;
            const int kr = c >> 6, cc = c & 63;
            cp_async16(s + kr * G_LDT + cc * 2, g + (long)(k0 + kr) * ld + row0 + cc * 2);
        }
    }
}

template <bool TRANS>
__device__ __forceinline__ double frag(const double* s, int row, int k) {
    return TRANS ? s[k * G_LDT + row] : s[row * G_LDN + k];
}

// C[ti,tj] (+)= alpha * sum_k opA[ti rows, k] * opB[k, tj cols]
//   TA=false: A stored [m][k];  TA=true: A stored [k][m]
//   TB=false: B stored [k][n];  TB=true: B stored [n][k]
template <bool TA, bool TB>
__global__ void __launch_bounds__(G_THREADS, 1) gemm_f64_kernel(const GemmParams p) {
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // 4 x 4 warps, warp tile 32 x 32; the four warps of one scheduler (warp % 4) cover all four wm, so the
    // triangular skip below removes the same amount of work from every scheduler
    // (b_tri skips by wn instead, so there the four warps of a scheduler cover all four wn)
    const bool by_wn = p.b_tri != 0;
    const int wm = by_wn ? warp % G_WM : warp / G_WN, wn = by_wn ? warp / G_WM : warp % G_WN;
    const int g = lane >> 2, kk = lane & 3;

    int ti, tj;
    if (p.order == ORDER_LOWER) {
        const int l = blockIdx.x;
        ti = (int)((sqrt(8.0 * l + 1.0) - 1.0) * 0.5);
        while ((long)(ti + 1) * (ti + 2) / 2 <= l) ++ti;
        while ((long)ti * (ti + 1) / 2 > l) --ti;
        tj = l - ti * (ti + 1) / 2;
    } else {
        ti = blockIdx.x / p.tilesN;
        tj = blockIdx.x - ti * p.tilesN;
        if (p.order == ORDER_RECT_HEAVY_LAST_ROW_FIRST) ti = p.tilesM - 1 - ti;
    }
    int klo = 0, khi = p.K;
    if (p.klo_mode == KLO_BY_TJ) klo = tj * G_TN;
    if (p.klo_mode == KLO_BY_TI) klo = ti * G_TM;
    if (p.khi_mode == KHI_BY_TI) khi = min(p.K, (ti + 1) * G_TM);
    const int nk = (khi - klo) / G_TK;

    const double* A = p.A + blockIdx.y * p.sA1 + blockIdx.z * p.sA2;
    const double* B = p.B + blockIdx.y * p.sB1 + blockIdx.z * p.sB2;
    double* C = p.C + blockIdx.y * p.sC1 + blockIdx.z * p.sC2;
    const int row0 = ti * G_TM, col0 = tj * G_TN;

    double acc[G_MI][G_NJ][2];
#pragma unroll
    for (int i = 0; i < G_MI; ++i)
#pragma unroll
        for (int j = 0; j < G_NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto sA = [&](int st) { return smem + (size_t)st * 2 * G_TILE_ELEMS; };
    auto sB = [&](int st) { return smem + (size_t)st * 2 * G_TILE_ELEMS + G_TILE_ELEMS; };
    auto issue = [&](int kt) {
        if (kt < nk) {
            const int st = kt % G_STAGES, k0 = klo + kt * G_TK;
            load_tile<TA>(sA(st), A, p.lda, row0, k0, tid);
            // B: TB=true is [n][k] (k contiguous) -> the non-transposed tile loader
            load_tile<!TB>(sB(st), B, p.ldb, col0, k0, tid);
        }
        cp_async_commit();
    };
    const int arow = wm * (G_TM / G_WM) + g, bcol = wn * (G_TN / G_WN) + g;
    auto load_frags = [&](int st, int ks, double (&af)[G_MI], double (&bf)[G_NJ]) {
        const double* a_s = sA(st);
        const double* b_s = sB(st);
#pragma unroll
        for (int i = 0; i < G_MI; ++i) af[i] = frag<TA>(a_s, arow + i * 8, ks * 4 + kk);
#pragma unroll
        for (int j = 0; j < G_NJ; ++j) bf[j] = frag<!TB>(b_s, bcol + j * 8, ks * 4 + kk);
    };

#pragma unroll
    for (int s = 0; s < G_STAGES - 1; ++s) issue(s);

    double af[G_MI], bf[G_NJ];
    cp_async_wait<G_STAGES - 2>();
    __syncthreads();
    load_frags(0, 0, af, bf);
    auto sync_and_issue = [&](int kt) {
        // after this barrier tile kt+1 has landed for every thread and stage (kt-1)%STAGES is free
        cp_async_wait<G_STAGES - 3>();
        __syncthreads();
        issue(kt + G_STAGES - 1);
    };
    auto compute = [&](int kt) {
        const int st = kt % G_STAGES, stn = (kt + 1) % G_STAGES;
#pragma unroll
        for (int ks = 0; ks < G_TK / 4; ++ks) {
            double an[G_MI], bn[G_NJ];
            if (ks + 1 < G_TK / 4) load_frags(st, ks + 1, an, bn);
            else if (kt + 1 < nk) load_frags(stn, 0, an, bn);
#pragma unroll
            for (int i = 0; i < G_MI; ++i)
#pragma unroll
                for (int j = 0; j < G_NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
#pragma unroll
            for (int i = 0; i < G_MI; ++i) af[i] = an[i];
#pragma unroll
            for (int j = 0; j < G_NJ; ++j) bf[j] = bn[j];
        }
    };
    // main k-range, then (a_tri_diag) the lower-triangular diagonal block where rows wm*32..+31 are zero
    // for k >= (wm+1)*32: those warps skip the DMMAs but keep taking part in barriers and copies
    const int nk_main = p.a_tri_diag ? nk - G_TM / G_TK : nk;
    if (p.b_tri) {
        // warp columns wn*32..+31 of a lower-triangular B block: 2 k-tiles of 16 per 32 columns
        const int kt_lo = p.b_tri == 2 ? wn * (G_TN / G_WN / G_TK) : 0;
        const int kt_hi = p.b_tri == 1 ? (wn + 1) * (G_TN / G_WN / G_TK) : nk;
        for (int kt = 0; kt < nk; ++kt) {
            sync_and_issue(kt);
            if (kt == kt_lo && kt > 0) load_frags(kt % G_STAGES, 0, af, bf);  // the pipeline of fragments starts here
            if (kt >= kt_lo && kt < kt_hi) compute(kt);
        }
    } else
    for (int kt = 0; kt < nk_main; ++kt) {
        sync_and_issue(kt);
        compute(kt);
    }
    if (p.a_tri_diag) {
        const int kt_end = nk_main + (wm + 1) * (G_TM / G_WM / G_TK);
        for (int kt = nk_main; kt < nk; ++kt) {
            sync_and_issue(kt);
            if (kt < kt_end) compute(kt);
        }
    }
    cp_async_wait<0>();
    __syncthreads();  // all operand reads done: in-place C (aliasing A or B tiles of this CTA) is safe

    if (p.epi == EPI_STORE) {
#pragma unroll
        for (int i = 0; i < G_MI; ++i) {
            const long r = row0 + arow + i * 8;
#pragma unroll
            for (int j = 0; j < G_NJ; ++j) {
                double2* dst = reinterpret_cast<double2*>(C + r * p.ldc + col0 + wn * (G_TN / G_WN) + j * 8 + kk * 2);
                double2 v;
                v.x = p.alpha * acc[i][j][0];
                v.y = p.alpha * acc[i][j][1];
                if (p.beta != 0.0) {
                    const double2 o = *dst;
                    v.x += p.beta * o.x;
                    v.y += p.beta * o.y;
                }
                *dst = v;
            }
        }
    } else {  // EPI_COLSUMSQ: partial[ti][col] = sum over the tile's 128 rows of acc^2 (fixed order)
        double* red = smem;  // [G_WM warps along M][128 cols]
#pragma unroll
        for (int j = 0; j < G_NJ; ++j) {
            double s0 = 0.0, s1 = 0.0;
#pragma unroll
            for (int i = 0; i < G_MI; ++i) {
                s0 += acc[i][j][0] * acc[i][j][0];
                s1 += acc[i][j][1] * acc[i][j][1];
            }
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                s0 += __shfl_xor_sync(0xffffffffu, s0, o);
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            }
            if (g == 0) {
                red[wm * G_TN + wn * (G_TN / G_WN) + j * 8 + kk * 2] = s0;
                red[wm * G_TN + wn * (G_TN / G_WN) + j * 8 + kk * 2 + 1] = s1;
            }
        }
        __syncthreads();
        if (tid < G_TN) {
            const double s = (red[tid] + red[G_TN + tid]) + (red[2 * G_TN + tid] + red[3 * G_TN + tid]);
            C[(long)ti * p.ldc + col0 + tid] = s;
        }
    }
}

template <bool TA, bool TB>
inline int launch_gemm(const GemmParams& p, int batch1, int batch2, cudaStream_t stream) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(gemm_f64_kernel<TA, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)G_SMEM_BYTES));
            TB_CUDA(cudaFuncSetAttribute(gemm_f64_kernel<TA, TB>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared));
            return 0;
        }))
        return -1;
    const long tiles = (p.order == ORDER_LOWER) ? (long)p.tilesM * (p.tilesM + 1) / 2 : (long)p.tilesM * p.tilesN;
    if (tiles <= 0 || batch1 <= 0 || batch2 <= 0) return 0;
    dim3 grid((unsigned)tiles, (unsigned)batch1, (unsigned)batch2);
    TB_LAUNCH((gemm_f64_kernel<TA, TB>), grid, dim3(G_THREADS), G_SMEM_BYTES, stream, p);
    return 0;
}

inline GemmParams gemm_params(const double* A, long lda, const double* B, long ldb, double* C, long ldc, int tilesM,
                              int tilesN, int K, double alpha, double beta) {
    GemmParams p{};
    p.A = A; p.B = B; p.C = C;
    p.lda = lda; p.ldb = ldb; p.ldc = ldc;
    p.tilesM = tilesM; p.tilesN = tilesN; p.K = K;
    p.alpha = alpha; p.beta = beta;
    p.order = ORDER_RECT; p.klo_mode = KLO_ZERO; p.khi_mode = KHI_FULL; p.epi = EPI_STORE; p.a_tri_diag = 0; p.b_tri = 0;
    return p;
}

}  // namespace tb200
