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
    // Recursive inverse of the triangular factor (fit.cu trtri_batched_i8), level h = 64 * h64: the matrix is cut into
    // pairs of diagonal blocks (A = tiles [2p h64, (2p+1) h64), C = the next h64 tiles, possibly cut short by the matrix
    // end); tile set = rows of C x columns of A of every pair.  recur = 1: T = L[C,A] inv(A), k over A from the column
    // tile on (inv(A) is lower triangular); recur = 2: X = -inv(C) T, k over C up to the row tile.  0: off.
    int recur, h64;
    // optional running maxima of |C| after this launch, as upper bounds of the HIGH WORD of the double (exponent + 20
    // mantissa bits, rounded up: all a power-of-two cover needs), one per output row / column; atomicMax, so the caller
    // zeroes them.  Written only by the launch (k-window) that completes a tile.
    unsigned* rowmax32;
    unsigned* colmax32;
    int ntiles, R;         // tiles per restart and restarts (filled by launch_gemm_i8)
    unsigned* sched;       // [2] tile counter + finished-CTA counter of this launch (zero at launch, left zero)
};

// tile t of a launch (restart-major; inside a restart the tiles with the longest k-range come first, so the dynamic
// scheduler ends a launch on its cheapest tiles) -> restart, row tile, column tile, k-range and beta
struct GemmTile {
    int rs, it, jt, kb_lo, nkb;
    bool final;  // this launch's k-window reaches the end of the tile's k-range: C holds the finished value afterwards
    double beta;
};
__device__ __forceinline__ GemmTile gemm_tile(const GemmI8Params& p, int t) {
    GemmTile g;
    g.rs = t / p.ntiles;
    int bx = t - g.rs * p.ntiles, kb_hi;
    if (p.recur) {
        const int per = (p.h64 / 2) * p.h64;  // tiles of one pair: h64/2 row tiles of 128 x h64 column tiles of 64
        const int pr = bx / per, q = bx - pr * per;
        const int a0 = 2 * pr * p.h64, c0 = a0 + p.h64;  // first 64-tile of A and of C
        g.it = c0 / 2 + q / p.h64;
        g.jt = a0 + q % p.h64;
        if (2 * g.it >= p.kblks) {  // C is cut short by the end of the matrix
            g.kb_lo = kb_hi = 0;
        } else if (p.recur == 1) {
            g.kb_lo = g.jt;
            kb_hi = c0;
        } else {
            g.kb_lo = c0;
            kb_hi = 2 * (g.it + 1);
        }
    } else if (p.tri) {
        // heavy tiles first: the k-range [2*it, kblks) shrinks with it
        g.it = 0;
        while (bx >= 2 * g.it + 2) { bx -= 2 * g.it + 2; ++g.it; }
        g.jt = bx;
        g.kb_lo = 2 * g.it;
        kb_hi = p.kblks;
    } else {
        int r = bx / p.nColTiles;
        // a k-range that grows with the row (trtri): last rows first
        if (p.kbHiByRow) r = p.ntiles / p.nColTiles - 1 - r;
        g.it = p.rowTile0 + r;
        g.jt = p.colTile0 + bx % p.nColTiles;
        g.kb_lo = p.kbLo;
        kb_hi = p.kbHiByRow ? 2 * (g.it + 1) : p.kbHi;
    }
    g.beta = p.beta;
    g.final = true;
    if (p.kbWinHi > 0) {
        if (g.kb_lo < p.kbWinLo) { g.kb_lo = p.kbWinLo; g.beta = 1.0; }
        if (kb_hi > p.kbWinHi) { kb_hi = p.kbWinHi; g.final = false; }
    }
    g.nkb = kb_hi - g.kb_lo;
    return g;
}

// PERSISTENT (round 2), same pipeline as colsumsq_i8_kernel: at most one CTA per SM takes tiles from a global counter;
// the producer warp runs on into the next tile's loads while the MMAs of the current one are still queued, TMEM and the
// barriers are set up once per CTA, and the epilogue hands TMEM back as soon as its last tcgen05.ld has landed.  Round
// 1 launched one CTA per tile and paid ~15 k cycles of set-up, first-load latency and epilogue per tile -- as much as
// the MMAs of a 10-k-block tile, which is what the block columns at both ends of potrf / trtri consist of.
// Tiles whose k-range misses the launch's window are skipped by the producer and never enter the pipeline.
template <int GMAX>
__global__ void __launch_bounds__(THREADS, 1) gemm_i8_kernel(const GemmI8Params p) {
    constexpr int NG = GMAX - 1;
    static_assert(NG * TN <= TMEM_COLS, "accumulators exceed TMEM");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], done_bar, tfree_bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ volatile int tile_ring[RING];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = p.ntiles * p.R;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&done_bar, 1);
        mbar_init(&tfree_bar, TM);
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
        uint32_t it = 0, ti = 0;  // k-blocks issued, tiles published
        for (;;) {
            int t = 0;
            if (lane == 0) t = (int)atomicAdd(p.sched, 1u);
            t = __shfl_sync(0xffffffffu, t, 0);
            const bool last = t >= total;
            GemmTile g{};
            if (!last) {
                g = gemm_tile(p, t);
                if (g.nkb <= 0) continue;  // outside this launch's k-window
            }
            const int nkb = last ? 1 : g.nkb;
            const uint8_t* base = p.S + (size_t)g.rs * p.restartStride;
            const uint8_t* a0_src = base + ((size_t)(2 * g.it) * p.kblks) * B_TILE;
            const uint8_t* a1_src = base + ((size_t)(2 * g.it + 1) * p.kblks) * B_TILE;
            const uint8_t* b_src =
                p.SB ? p.SB + (size_t)g.rs * p.restartStrideB + ((size_t)(g.jt - p.colTile0) * p.kblks) * B_TILE
                     : base + ((size_t)g.jt * p.kblks) * B_TILE;
            const long b_slice = p.SB ? p.sliceStrideB : p.sliceStride;
            for (int i = 0; i < nkb; ++i, ++it) {
                const int s = it % STAGES;
                const size_t kb = (size_t)(g.kb_lo + i);
                if (it >= STAGES) mbar_wait(&empty_bar[s], ((it / STAGES) - 1) & 1);
                if (elect_one()) {
                    if (i == 0) tile_ring[ti % RING] = last ? -1 : t;
                    if (last) {
                        mbar_arrive(&full_bar[s]);
                    } else {
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
                }
                __syncwarp();
            }
            ++ti;
            if (last) break;
        }
        if (lane == 0) {
            // every CTA has taken its end marker once the last one arrives here: leave the pair zeroed for its next user
            if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
                atomicExch(p.sched, 0u);
                atomicExch(p.sched + 1, 0u);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        uint32_t it = 0;
        for (uint32_t ti = 0;; ++ti) {
            mbar_wait(&full_bar[it % STAGES], (it / STAGES) & 1);  // first stage of the tile: its id is visible now
            const int t = tile_ring[ti % RING];
            if (t < 0) {
                if (ti > 0) mbar_wait(&tfree_bar, (ti - 1) & 1);
                if (elect_one()) mbar_arrive(&done_bar);
                __syncwarp();
                break;
            }
            const int nkb = gemm_tile(p, t).nkb;
            for (int i = 0; i < nkb; ++i, ++it) {
                const int s = it % STAGES;
                if (i == 0 && ti > 0) mbar_wait(&tfree_bar, (ti - 1) & 1);  // previous tile's accumulators drained
                if (i > 0) mbar_wait(&full_bar[s], (it / STAGES) & 1);
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
        }
    } else {
        // ===== epilogue: thread = output row (TMEM lane), 64 output columns =====
        const int q = warp & 3;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        for (uint32_t ti = 0;; ++ti) {
            mbar_wait(&done_bar, ti & 1);
            const int t = tile_ring[ti % RING];
            if (t < 0) break;
            const GemmTile g = gemm_tile(p, t);
            const int row = g.it * TM + q * 32 + lane;
            const double srow = p.alpha * (p.scale ? p.scale[(size_t)g.rs * p.scaleStride + row] : 1.0);
            double* crow = p.C + (size_t)g.rs * p.cStride + (size_t)row * p.ldc + (size_t)g.jt * TN;
            const double* scol = (p.scaleB ? p.scaleB : p.scale) + (size_t)g.rs * p.scaleStride + g.jt * TN;
            const double beta = g.beta;
            const bool want_max = g.final && (p.rowmax32 != nullptr || p.colmax32 != nullptr);
            unsigned rmax = 0;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // Phase 1: drain.  All 64 columns are recombined (integer pipe + a few FP64 additions) into registers and
            // TMEM goes back to the MMA warp BEFORE the tile is scaled, merged with C and stored: with short k-ranges
            // (the triangular products of the inverse: 16 k-blocks on average) the FP64 store phase was a fifth of a
            // tile's time during which the tensor pipe waited (ncu: 63 % busy, profiles/r2_gemm_i8_large_ncu.md).
            double v[TN];
#pragma unroll
            for (int cc = 0; cc < TN / 8; ++cc) {
                uint32_t r[NG][8];
#pragma unroll
                for (int gg = 0; gg < NG; ++gg) tmem_ld8(lane_base + gg * TN + cc * 8, r[gg]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int c = 0; c < 8; ++c) v[cc * 8 + c] = recombine<NG>(r, c);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&tfree_bar);  // the accumulators are in registers: the next tile's MMAs may start
            // Phase 2: scale, merge, store, maxima
#pragma unroll
            for (int cc = 0; cc < TN / 8; ++cc) {
                double out[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    // both operands carry digits b_s 2^(1-8s): group g = i+j has weight 2^(2-8g), smallest first;
                    // recombine() weighs with 2^(1-8g), the missing factor 2 is exact
                    out[c] = v[cc * 8 + c] * (2.0 * srow * __ldg(scol + cc * 8 + c));
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
                if (want_max) {
                    unsigned cm = 0;
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        unsigned hw = (unsigned)__double2hiint(out[c]) & 0x7fffffffu;
                        if (hw >= 0x7ff00000u) hw = 0;  // NaN / Inf of a failed member never enters a scale
                        rmax = hw > rmax ? hw : rmax;
                        const unsigned m = __reduce_max_sync(0xffffffffu, hw);
                        if (lane == c) cm = m;
                    }
                    if (p.colmax32 && lane < 8 && cm > 0)
                        atomicMax(p.colmax32 + (size_t)g.rs * p.scaleStride + g.jt * TN + cc * 8 + lane, cm + 1);
                }
            }
            if (want_max && p.rowmax32 && rmax > 0) atomicMax(p.rowmax32 + (size_t)g.rs * p.scaleStride + row, rmax + 1);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// Scheduler counter pairs of the persistent launches: a pool in the caller's workspace, zeroed once per fit call and
// handed out round-robin.  A pair is zero again when its launch has finished; launches of one fit call that may be in
// flight together (main, side and look-ahead stream) are at most a few iterations apart, far less than the pool size.
constexpr int SCHED_POOL = 128;
struct SchedPool {
    unsigned* base = nullptr;  // [SCHED_POOL][2]
    int next = 0;
    unsigned* take() {
        unsigned* q = base + 2 * next;
        next = (next + 1) % SCHED_POOL;
        return q;
    }
};

template <int GMAX>
inline int launch_gemm_i8(const GemmI8Params& p0, int ntiles, int R, SchedPool& pool, int sm_count, cudaStream_t stream) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(gemm_i8_kernel<GMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)SMEM_BYTES));
            return 0;
        }))
        return -1;
    if (ntiles <= 0 || R <= 0) return 0;
    TB_CHECK(pool.base != nullptr && sm_count > 0, "fit GEMM: scheduler counters missing");
    GemmI8Params p = p0;
    p.ntiles = ntiles;
    p.R = R;
    const long total = (long)ntiles * R;
    const unsigned grid = (unsigned)(total < sm_count ? total : sm_count);
    // the longest k-range any CTA of this launch can have
    const int span = p.recur ? p.h64 : p.tri ? p.kblks : (p.kbHiByRow ? p.kblks : p.kbHi) - p.kbLo;
    if (span <= PASS_KB) {
        p.sched = pool.take();
        TB_LAUNCH((gemm_i8_kernel<GMAX>), dim3(grid), dim3(THREADS), SMEM_BYTES, stream, p);
        return 0;
    }
    const int lo0 = (p.tri || p.recur) ? 0 : p.kbLo;
    const int hi0 = (p.tri || p.recur) ? p.kblks : (p.kbHiByRow ? p.kblks : p.kbHi);
    for (int w = lo0; w < hi0; w += PASS_KB) {
        GemmI8Params q = p;
        q.kbWinLo = w;
        q.kbWinHi = (w + PASS_KB < hi0) ? w + PASS_KB : hi0;
        q.sched = pool.take();
        TB_LAUNCH((gemm_i8_kernel<GMAX>), dim3(grid), dim3(THREADS), SMEM_BYTES, stream, q);
    }
    return 0;
}

}  // namespace i8
}  // namespace tb200
