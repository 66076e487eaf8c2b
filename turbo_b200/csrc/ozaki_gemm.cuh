// FP64 matrix products of the batched fit path on the INT8 tensor cores: C = beta*C + alpha * A * B^T where A and B
// are row blocks of ONE byte-sliced operand (the Cholesky factor L for the left-looking column update of potrf, the
// columns of L^-1 for K^-1 = L^-T L^-1; sklearn:_gpr.py:586-651 computes both with LAPACK potrf/potrs).
// Same exact scheme as ozaki_i8.cuh (balanced s8 digits on both sides here, INT32 accumulation in TMEM, Linv-style
// per-row power-of-two scales, slices stacked along N), with an FP64 tile store instead of the column sum of squares.
//
// Operand layout: [restart][slice 0..6][64-row tile][k-block of 64][4 KB canonical K-major core-matrix tile]; a 128-row
// A tile is two consecutive 64-row tiles (two bulk copies per slice), so one buffer serves both operands.
#pragma once
#include "ozaki_i8.cuh"

namespace tb200 {
namespace i8 {

struct GemmI8Params {
    const uint8_t* S;      // sliced A operand (and B operand unless SB is set)
    long restartStride;    // bytes between restarts
    long sliceStride;      // bytes between slices of one restart
    const double* scale;   // [restart][rows]: 2^(e_r), |x[r,k]| <= 0.98 * 2^(e_r); NULL = 1 (A side)
    long scaleStride;
    const uint8_t* SB;     // optional separate B operand (a panel buffer): tile index = jt - colTile0
    long restartStrideB, sliceStrideB;
    const double* scaleB;  // scales of the B rows, indexed by absolute output column; NULL = `scale`
    double* C;             // FP64 output, row-major
    long ldc, cStride;
    int kblks;             // k-blocks per 64-row tile (= K extent of the buffer / 64)
    int tri;               // 1: lower-triangular tile set (it, jt <= 2*it+1), k-blocks [2*it, kblks)  (X^T X of lower X^T)
                           // 0: rectangle rowTile0.. x colTile0.., k-blocks [kbLo, kbHi)
    int rowTile0, colTile0, nColTiles;  // 128-row tiles / 64-col tiles (rectangle mode)
    int kbLo, kbHi;
    int kbHiByRow;         // rectangle mode: k-blocks [kbLo, 2*(it+1)) (A is lower triangular: X[i,k] = 0 for k > i)
    double alpha, beta;
    // One launch accumulates at most PASS_KB k-blocks (8192 terms) exactly in INT32.  Longer k-ranges are cut into
    // windows [kbWinLo, kbWinHi) launched in increasing order: a CTA whose range misses the window does nothing, and
    // a window that does not contain the start of the CTA's range accumulates onto C (beta = 1) instead of applying
    // the caller's beta.  kbWinHi = 0: no window.
    int kbWinLo, kbWinHi;
};

template <int GMAX>
__global__ void __launch_bounds__(THREADS, 1) gemm_i8_kernel(const GemmI8Params p) {
    constexpr int NG = GMAX - 1;
    static_assert(NG * TN <= TMEM_COLS, "accumulators exceed TMEM");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], done_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int it, jt, kb_lo, kb_hi;
    if (p.tri) {
        // heavy tiles first: the k-range [2*it, kblks) shrinks with it
        it = 0;
        int rem = (int)blockIdx.x;
        while (rem >= 2 * it + 2) { rem -= 2 * it + 2; ++it; }
        jt = rem;
        kb_lo = 2 * it;
        kb_hi = p.kblks;
    } else {
        it = p.rowTile0 + (int)blockIdx.x / p.nColTiles;
        jt = p.colTile0 + (int)blockIdx.x % p.nColTiles;
        kb_lo = p.kbLo;
        kb_hi = p.kbHiByRow ? 2 * (it + 1) : p.kbHi;
    }
    double beta = p.beta;
    if (p.kbWinHi > 0) {
        if (kb_lo < p.kbWinLo) { kb_lo = p.kbWinLo; beta = 1.0; }
        if (kb_hi > p.kbWinHi) kb_hi = p.kbWinHi;
    }
    const int nkb = kb_hi - kb_lo;
    if (nkb <= 0) return;  // uniform over the CTA, before any barrier or TMEM allocation
    const int rs = (int)blockIdx.y;  // restart

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&done_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                     "r"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // ===== producer: per slice two 4 KB copies for the 128 A rows and one for the 64 B rows =====
        const uint8_t* base = p.S + (size_t)rs * p.restartStride;
        const uint8_t* a0_src = base + ((size_t)(2 * it) * p.kblks) * B_TILE;
        const uint8_t* a1_src = base + ((size_t)(2 * it + 1) * p.kblks) * B_TILE;
        const uint8_t* b_src = p.SB ? p.SB + (size_t)rs * p.restartStrideB + ((size_t)(jt - p.colTile0) * p.kblks) * B_TILE
                                    : base + ((size_t)jt * p.kblks) * B_TILE;
        const long b_slice = p.SB ? p.sliceStrideB : p.sliceStride;
        for (int i = 0; i < nkb; ++i) {
            const int s = i % STAGES;
            const size_t kb = (size_t)(kb_lo + i);
            if (i >= STAGES) mbar_wait(&empty_bar[s], ((i / STAGES) - 1) & 1);
            if (elect_one()) {
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], STAGE_BYTES);
#pragma unroll
                for (int sl = 0; sl < S; ++sl) {
                    const size_t so = (size_t)sl * p.sliceStride + kb * B_TILE;
                    bulk_g2s(st + sl * A_TILE, a0_src + so, B_TILE, &full_bar[s]);
                    bulk_g2s(st + sl * A_TILE + B_TILE, a1_src + so, B_TILE, &full_bar[s]);
                    bulk_g2s(st + S * A_TILE + sl * B_TILE, b_src + (size_t)sl * b_slice + kb * B_TILE, B_TILE,
                             &full_bar[s]);
                }
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        for (int i = 0; i < nkb; ++i) {
            const int s = i % STAGES;
            mbar_wait(&full_bar[s], (i / STAGES) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a0 = smem_u32(smem + (size_t)s * STAGE_BYTES);
            const uint64_t ad0 = make_desc(a0);
            const uint64_t bd0 = make_desc(a0 + S * A_TILE);
            if (elect_one()) {
                if (i == 0) {
                    issue_step<GMAX, 1, true, 0, 1>(tmem_base, ad0, bd0);
                } else {
                    issue_step<GMAX, 1, false, 0, 1>(tmem_base, ad0, bd0);
                }
                issue_step<GMAX, 1, false, 0, 1>(tmem_base, ad0 + (uint64_t)((2 * LBO) >> 4),
                                                 bd0 + (uint64_t)((2 * LBO) >> 4));
                umma_commit(&empty_bar[s]);
                if (i == nkb - 1) umma_commit(&done_bar);
            }
            __syncwarp();
        }
    } else {
        // ===== epilogue: thread = output row (TMEM lane), 64 output columns =====
        const int q = warp & 3;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        const int row = it * TM + q * 32 + lane;
        const double srow = p.alpha * (p.scale ? p.scale[(size_t)rs * p.scaleStride + row] : 1.0);
        double* crow = p.C + (size_t)rs * p.cStride + (size_t)row * p.ldc + (size_t)jt * TN;
        const double* scol = (p.scaleB ? p.scaleB : p.scale) + (size_t)rs * p.scaleStride + jt * TN;
        mbar_wait(&done_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
        for (int cc = 0; cc < TN / 8; ++cc) {
            uint32_t r[NG][8];
#pragma unroll
            for (int g = 0; g < NG; ++g) tmem_ld8(lane_base + g * TN + cc * 8, r[g]);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            double out[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                // both operands carry digits b_s 2^(1-8s): group g = i+j has weight 2^(2-8g), smallest first
                double v = 0.0;
#pragma unroll
                for (int g = NG - 1; g >= 0; --g) v = fma((double)(int)r[g][c], ldexp(1.0, 2 - 8 * (g + 2)), v);
                out[c] = v * (srow * __ldg(scol + cc * 8 + c));
            }
            double2* dst = reinterpret_cast<double2*>(crow + cc * 8);
            if (beta != 0.0) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const double2 old = dst[c];
                    out[2 * c] = fma(beta, old.x, out[2 * c]);
                    out[2 * c + 1] = fma(beta, old.y, out[2 * c + 1]);
                }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) dst[c] = make_double2(out[2 * c], out[2 * c + 1]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

template <int GMAX>
inline int launch_gemm_i8(const GemmI8Params& p, int ntiles, int R, cudaStream_t stream) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(gemm_i8_kernel<GMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)SMEM_BYTES));
            return 0;
        }))
        return -1;
    if (ntiles <= 0 || R <= 0) return 0;
    // the longest k-range any CTA of this launch can have
    const int span = p.tri ? p.kblks : (p.kbHiByRow ? p.kblks : p.kbHi) - p.kbLo;
    if (span <= PASS_KB) {
        TB_LAUNCH((gemm_i8_kernel<GMAX>), dim3((unsigned)ntiles, (unsigned)R), dim3(THREADS), SMEM_BYTES, stream, p);
        return 0;
    }
    const int lo0 = p.tri ? 0 : p.kbLo, hi0 = p.tri ? p.kblks : (p.kbHiByRow ? p.kblks : p.kbHi);
    for (int w = lo0; w < hi0; w += PASS_KB) {
        GemmI8Params q = p;
        q.kbWinLo = w;
        q.kbWinHi = (w + PASS_KB < hi0) ? w + PASS_KB : hi0;
        TB_LAUNCH((gemm_i8_kernel<GMAX>), dim3((unsigned)ntiles, (unsigned)R), dim3(THREADS), SMEM_BYTES, stream, q);
    }
    return 0;
}

}  // namespace i8
}  // namespace tb200
