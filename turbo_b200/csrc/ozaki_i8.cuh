// Predictive-variance contraction on the INT8 tensor cores (tcgen05.mma kind::i8, TMEM accumulators):
// an error-free-transformation ("Ozaki scheme") evaluation of  V = Linv * Kstar^T  followed by the
// column sums of squares, as an alternative engine to the DMMA GEMM of gemm_f64.cuh.
//
// Why: FP64 has no tcgen05 kind, and DMMA shares the 64 FMA/clk/SM FP64 pipe with DFMA (37 TFLOP/s
// measured).  The 5th-gen tensor cores do 8192 INT8 MAC/clk/SM with exact INT32 accumulation, so a
// product of two FP64 matrices can be assembled EXACTLY from byte slices:
//
//   Kstar[c,k] / sf2          = sum_{i=1..7} a_i[c,k] * 2^(-8i)      a_i in [0,255]     (u8)
//   Linv[r,k]  / 2^(e_r)      = sum_{j=1..7} b_j[r,k] * 2^(1-8j)     b_j in [-128,127]  (s8, balanced digits)
//   V[c,r] = sf2 * 2^(e_r) * sum_g 2^(1-8g) * G_g[c,r],   G_g = sum_{i+j=g} sum_k a_i[c,k] b_j[r,k]
//
// Every G_g is an integer matrix product accumulated exactly in INT32 TMEM (7 pairs * 8192 * 255*128
// < 2^31, hence n <= 8192 per accumulation pass).  Pairs with i+j > GMAX are dropped: GMAX = 8 keeps 28
// products and truncates at ~2^-54 of (row max * sf2) per term; GMAX = 9 keeps 34 and is
// indistinguishable from an FP64 GEMM (tests/test_gpu_i8.py).
//
// Kernel shape: one CTA per (64 Linv rows) x (128 candidates) tile.  A-operand (M = 128) = one candidate
// slice, B-operand = the Linv-row slices j = 1..GMAX-i STACKED along N (all digits are s8, so one
// instruction descriptor serves every slice; the 64-row slice tiles are adjacent in shared memory, so the
// stack is one canonical K-major operand of N = 64*(GMAX-i) rows).  One tcgen05.mma therefore produces the
// products of a_i with several b_j at once and lands them in the accumulators of groups g = i+1 .. GMAX,
// which are adjacent TMEM column blocks: 10 MMAs of N = 64..224 per K = 32 step instead of 28 of N = 64.
// That matters because an M128 x N64 MMA re-reads 6 KB of operands from shared memory for 32 cycles of
// tensor work and the shared-memory pipe, not the tensor pipe, was the limiter (profiles/r1_i8_ncu_full.md:
// L1/TEX 78 %, tensor 73 %); stacking cuts operand reads per K step from 168 KB to 96 KB.
// Both operands are K-major in the canonical no-swizzle core-matrix layout, pre-tiled in global memory so
// that one k-block (64 bytes of K) of one slice is ONE contiguous bulk-async copy (cp.async.bulk, the TMA
// engine, no tensor map needed).  Warp 0 = copy producer, warp 1 = MMA issuer (one elected lane) + TMEM
// owner, warps 2-5 = epilogue: each thread owns one candidate (TMEM lane) and reduces its 64 columns in
// registers, so no cross-thread reduction.
#pragma once
#include "common.cuh"

namespace tb200 {
namespace i8 {

constexpr int S = 7;                 // byte slices per operand
constexpr int KB = 64;               // bytes of K per pipeline stage
constexpr int TM = 128;              // candidates per CTA tile (UMMA M)
constexpr int TN = 64;               // Linv rows per CTA tile (UMMA N)
constexpr int A_TILE = TM * KB;      // 8192 B
constexpr int B_TILE = TN * KB;      // 4096 B
constexpr int STAGE_BYTES = S * (A_TILE + B_TILE);  // 86016
constexpr int STAGES = 2;
constexpr int THREADS = 192;
constexpr int TMEM_COLS = 512;
constexpr int LBO = 128;             // bytes between the two 16-byte K chunks of one core-matrix row group
constexpr int SBO = (KB / 16) * 128; // bytes between groups of 8 rows
constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024;

// byte offset of element (row r, k in [0,64)) inside a canonical K-major no-swizzle tile
__host__ __device__ inline int tile_offset(int r, int k) {
    return (r >> 3) * SBO + (k >> 4) * LBO + (r & 7) * 16 + (k & 15);
}

struct Params {
    const uint8_t* candS;    // [S][candTiles][kblks][A_TILE]
    const uint8_t* linvS;    // [S][rowTiles][kblks][B_TILE]
    const double* rowscale;  // [rowTiles*64]  sf2 * 2^(e_r)
    double* partial;         // [rowTiles][ldp]
    long ldp;
    long candSliceStride, linvSliceStride;  // bytes
    int kblks;               // K / 64 (tiles per row of tiles)
    int rowTiles, candTiles;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// try_wait suspends the thread in hardware between polls.  The bound is a compile-time option and it is in WALL-CLOCK
// time (%globaltimer): -DTB200_MBAR_TRAP_SECONDS=<s> traps after <s> seconds in one wait, so a protocol bug ends the
// launch instead of hanging the device; without the macro the wait is unbounded.  build.sh passes 30 s: no wait of a
// correct launch is within four orders of magnitude of that (the longest kernel runs for milliseconds), so neither ncu
// replay nor time-slicing can trip it -- unlike round 1's bound of 2^24 POLLS, which could.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
#ifdef TB200_MBAR_TRAP_SECONDS
    if (mbar_try_wait(bar, parity)) return;
    uint64_t t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (!mbar_try_wait(bar, parity)) {
        uint64_t t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > (uint64_t)(TB200_MBAR_TRAP_SECONDS) * 1000000000ull) __trap();
    }
#else
    while (!mbar_try_wait(bar, parity)) {
    }
#endif
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
    // K-major, SWIZZLE_NONE: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48)
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(SBO >> 4) << 32) |
           (1ull << 46);
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
        :
        : "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}

// instruction descriptor: D = S32 (c_format 2 @4), A u8 (0 @7), B s8 (1 @10), K-major both,
// N>>3 @17, M>>4 @24
__host__ __device__ constexpr uint32_t idesc(int n_cols, int a_signed = 0) {
    return (2u << 4) | ((uint32_t)a_signed << 7) | (1u << 10) | ((uint32_t)(n_cols >> 3) << 17) |
           ((uint32_t)(TM >> 4) << 24);
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// MMAs of one candidate-slice tile (descriptor `ad`) against rows [N_LO, N_HI) of the stacked Linv slices (descriptor
// of row 0 = `bd0`; 8 rows further = +SBO bytes = +SBO/16 in the descriptor's address field) -> TMEM columns
// d_tmem + [N_LO, N_HI).  All resolved at compile time so the issuing lane executes a handful of uniform-register adds
// per MMA.  Chunking: BALANCED chunks of <= 256 columns (multiples of 16).  Measured cost of one M128 x N x K32 kind::i8
// MMA (tb200_i8_mma_rate, profiles/r2_i8_mma_rate.md): ~N/2 cycles from N = 128 up (72 / 84 / 98 / 114 / 129 cycles at
// N = 128 / 160 / 192 / 224 / 256) over a floor of ~50 cycles per instruction (58 at N = 64), so 384 columns are
// cheaper as 192 + 192 (196 cycles) than as 256 + 128 (201), 320 as 160 + 160 (168) than as 256 + 64 (187): an A/B of
// the two orders on the variance kernel gave 0.2541 vs 0.2555 ms (profiles/r2_ab_chunking.md).
template <int N_LO, int N_HI, int ACC, int ASGN = 0>
__device__ __forceinline__ void issue_range(uint32_t d_tmem, uint64_t ad, uint64_t bd0) {
    if constexpr (N_HI > N_LO) {
        constexpr int width = N_HI - N_LO;
        constexpr int nchunks = (width + 255) / 256;
        constexpr int chunk = ((width / nchunks) + 15) & ~15;
        constexpr int nc = chunk < width ? chunk : width;
        umma_i8(d_tmem + N_LO, ad, bd0 + (uint64_t)((N_LO >> 3) * (SBO >> 4)), idesc(nc, ASGN), (uint32_t)ACC);
        issue_range<N_LO + nc, N_HI, ACC, ASGN>(d_tmem, ad, bd0);
    }
}

// all slice products of one K = 32 step.  FIRST: no accumulator holds a value yet -- groups [0, COVERED) have been
// written by earlier slices of this step, the rest are overwritten (accumulate = 0).
template <int GMAX, int I, bool FIRST, int COVERED, int ASGN = 0>
__device__ __forceinline__ void issue_step(uint32_t tmem_base, uint64_t ad0, uint64_t bd0) {
    if constexpr (I <= S) {
        constexpr int J = (GMAX - I < S) ? GMAX - I : S;  // Linv slices j = 1..J pair with a_I
        if constexpr (J > 0) {
            // a_I * b_j lands in group I+j-2: columns (I-1)*TN + (j-1)*TN of the accumulator block
            const uint32_t d = tmem_base + (I - 1) * TN;
            const uint64_t ad = ad0 + (uint64_t)((I - 1) * (A_TILE >> 4));
            if constexpr (!FIRST) {
                issue_range<0, J * TN, 1, ASGN>(d, ad, bd0);
                issue_step<GMAX, I + 1, false, 0, ASGN>(tmem_base, ad0, bd0);
            } else {
                constexpr int split0 = (COVERED - (I - 1)) * TN;  // columns below `split` are initialised
                constexpr int split = split0 < 0 ? 0 : (split0 > J * TN ? J * TN : split0);
                issue_range<0, split, 1, ASGN>(d, ad, bd0);
                issue_range<split, J * TN, 0, ASGN>(d, ad, bd0);
                constexpr int cov = (I - 1 + J > COVERED) ? I - 1 + J : COVERED;
                issue_step<GMAX, I + 1, true, cov, ASGN>(tmem_base, ad0, bd0);
            }
        } else {
            issue_step<GMAX, I + 1, FIRST, COVERED, ASGN>(tmem_base, ad0, bd0);
        }
    }
}

// MULTI = false: the whole k-range of a tile (<= PASS_KB k-blocks, n <= 8192) is one exact INT32 accumulation pass.
// MULTI = true : longer k-ranges are cut into passes of PASS_KB k-blocks; after each pass the epilogue threads drain the
// accumulators into 64 FP64 registers per thread (the rows of V must be complete before they are squared) and hand
// TMEM back to the MMA warp.  One drain costs ~3k cycles against ~230k cycles of MMAs per pass.
constexpr int PASS_KB = 128;  // 8192 of K: 7 pairs * 8192 * 255 * 128 < 2^31

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// digit planes a kernel with cut-off GMAX reads: plane s pairs with planes 1..GMAX-s, so plane 7 is unused at GMAX = 7
__host__ __device__ constexpr int planes_used(int gmax) { return gmax - 1 < S ? gmax - 1 : S; }

template <int GMAX, bool MULTI>
__global__ void __launch_bounds__(THREADS, 1) colsumsq_i8_kernel(const Params p) {
    constexpr int NG = GMAX - 1;  // groups g = 2..GMAX, one 64-column INT32 accumulator each
    constexpr int SU = planes_used(GMAX);
    static_assert(NG * TN <= TMEM_COLS, "accumulators exceed TMEM");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], done_bar, tfree_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // heavy row tiles first (k-range grows with the row tile)
    const int rt = p.rowTiles - 1 - (int)(blockIdx.x / p.candTiles);
    const int ct = blockIdx.x % p.candTiles;
    const int nkb = rt + 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&done_bar, 1);
        mbar_init(&tfree_bar, TM);  // every epilogue thread arrives once it has drained its TMEM lane
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
        // ===== producer: 14 bulk copies per k-block, issued by one elected lane of a converged warp =====
        const uint8_t* a_src = p.candS + ((size_t)ct * p.kblks) * A_TILE;
        const uint8_t* b_src = p.linvS + ((size_t)rt * p.kblks) * B_TILE;
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % STAGES;
            if (kb >= STAGES) mbar_wait(&empty_bar[s], ((kb / STAGES) - 1) & 1);
            if (elect_one()) {
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], SU * (A_TILE + B_TILE));
#pragma unroll
                for (int sl = 0; sl < SU; ++sl) {
                    bulk_g2s(st + sl * A_TILE, a_src + (size_t)sl * p.candSliceStride + (size_t)kb * A_TILE, A_TILE,
                             &full_bar[s]);
                    bulk_g2s(st + S * A_TILE + sl * B_TILE, b_src + (size_t)sl * p.linvSliceStride + (size_t)kb * B_TILE,
                             B_TILE, &full_bar[s]);
                }
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the warp stays converged, one elected lane issues =====
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % STAGES;
            const bool pass_start = MULTI ? (kb % PASS_KB == 0) : (kb == 0);
            if (MULTI && pass_start && kb > 0) {
                mbar_wait(&tfree_bar, ((kb / PASS_KB) - 1) & 1);  // accumulators of the previous pass drained
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            mbar_wait(&full_bar[s], (kb / STAGES) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a0 = smem_u32(smem + (size_t)s * STAGE_BYTES);
            const uint64_t ad0 = make_desc(a0);
            const uint64_t bd0 = make_desc(a0 + S * A_TILE);
            if (elect_one()) {
                // K = 32 per MMA: the second half of the 64-byte k-block is 2 LBO further (descriptor units of 16 B)
                if (pass_start) {
                    issue_step<GMAX, 1, true, 0>(tmem_base, ad0, bd0);
                } else {
                    issue_step<GMAX, 1, false, 0>(tmem_base, ad0, bd0);
                }
                issue_step<GMAX, 1, false, 0>(tmem_base, ad0 + (uint64_t)((2 * LBO) >> 4), bd0 + (uint64_t)((2 * LBO) >> 4));
                umma_commit(&empty_bar[s]);  // frees the stage once these MMAs have read it
                if (kb == nkb - 1 || (MULTI && (kb + 1) % PASS_KB == 0)) umma_commit(&done_bar);
            }
            __syncwarp();
        }
    } else {
        // ===== epilogue: thread = candidate (TMEM lane), 64 columns = Linv rows of this tile =====
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        double acc = 0.0;
        if constexpr (!MULTI) {
            mbar_wait(&done_bar, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
            for (int cc = 0; cc < TN / 8; ++cc) {
                uint32_t r[NG][8];
#pragma unroll
                for (int g = 0; g < NG; ++g) tmem_ld8(lane_base + g * TN + cc * 8, r[g]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    // sum_g 2^(1-8(g+2)) G_g, smallest weight first
                    double v = 0.0;
#pragma unroll
                    for (int g = NG - 1; g >= 0; --g) v = fma((double)(int)r[g][c], ldexp(1.0, 1 - 8 * (g + 2)), v);
                    v *= __ldg(p.rowscale + rt * TN + cc * 8 + c);
                    acc = fma(v, v, acc);
                }
            }
        } else {
            double vrow[TN];  // V[candidate, 64 rows] / rowscale, summed over the passes (registers: fully unrolled)
            const int npass = (nkb + PASS_KB - 1) / PASS_KB;
            for (int pass = 0; pass < npass; ++pass) {
                mbar_wait(&done_bar, pass & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int cc = 0; cc < TN / 8; ++cc) {
                    uint32_t r[NG][8];
#pragma unroll
                    for (int g = 0; g < NG; ++g) tmem_ld8(lane_base + g * TN + cc * 8, r[g]);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        double v = 0.0;
#pragma unroll
                        for (int g = NG - 1; g >= 0; --g) v = fma((double)(int)r[g][c], ldexp(1.0, 1 - 8 * (g + 2)), v);
                        vrow[cc * 8 + c] = (pass == 0) ? v : vrow[cc * 8 + c] + v;
                    }
                }
                if (pass + 1 < npass) {
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(&tfree_bar);
                }
            }
#pragma unroll
            for (int c = 0; c < TN; ++c) {
                const double v = vrow[c] * __ldg(p.rowscale + rt * TN + c);
                acc = fma(v, v, acc);
            }
        }
        p.partial[(long)rt * p.ldp + (long)ct * TM + q * 32 + lane] = acc;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

template <int GMAX, bool MULTI>
inline int launch_colsumsq_i8_impl(const Params& p, cudaStream_t stream) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(colsumsq_i8_kernel<GMAX, MULTI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)SMEM_BYTES));
            // same (maximum) carve-out as the small kernels that share the SMs with it: CTAs launched under different
            // shared-memory configurations cannot be resident on one SM at the same time
            TB_CUDA(cudaFuncSetAttribute(colsumsq_i8_kernel<GMAX, MULTI>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared));
            return 0;
        }))
        return -1;
    TB_LAUNCH((colsumsq_i8_kernel<GMAX, MULTI>), dim3((unsigned)(p.rowTiles * p.candTiles)), dim3(THREADS), SMEM_BYTES,
              stream, p);
    return 0;
}

template <int GMAX>
inline int launch_colsumsq_i8(const Params& p, cudaStream_t stream) {
    // the heaviest tile spans rowTiles k-blocks
    return p.rowTiles <= PASS_KB ? launch_colsumsq_i8_impl<GMAX, false>(p, stream)
                                 : launch_colsumsq_i8_impl<GMAX, true>(p, stream);
}

// ---------------------------------------------------------------------------------------------
// INT8 tensor peak of this device: the roofline denominator of the two kernels above, MEASURED instead of inferred from
// the bf16 figure.  One CTA per SM; one elected lane issues tcgen05.mma kind::i8 M128 x N256 x K32 instructions on
// operands that stay resident in shared memory, alternating between the two 256-column halves of TMEM.  `random`
// fills the operands with pseudo-random digits: the instruction rate does not depend on the values, but the POWER does,
// and on a 1 kW part the clock the power limit allows is what sets the sustained rate (zero digits ran at 1.90 GHz and
// 4.60 POP/s in round 2; the variance kernel's step runs at 1.65-1.70 GHz).  Two mbarriers alternate so a batch is always queued behind the
// one in flight and the pipe never drains.  No global traffic, no epilogue: nothing but the MMA rate is measured.
// ---------------------------------------------------------------------------------------------
constexpr int PEAK_N = 256, PEAK_BATCH = 16;
constexpr size_t PEAK_SMEM = (size_t)(TM + PEAK_N) * KB + 1024;  // A: 128 x 64 B, B: 256 x 64 B (canonical K-major)

template <int N_COLS>
__global__ void __launch_bounds__(64, 1) i8_peak_kernel(int batches, int random, int n_cols) {
    static_assert(N_COLS == PEAK_N, "one instantiation (a template only so that both translation units may include it)");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar[2];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    for (int e = threadIdx.x; e < (int)((TM + PEAK_N) * KB / 16); e += blockDim.x) {
        uint32_t w[4] = {0u, 0u, 0u, 0u};
        if (random) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {  // a 32-bit mix of (CTA, word index): every byte is a uniform digit
                uint32_t x = (uint32_t)(e * 4 + q) * 2654435761u + blockIdx.x * 40503u + 12345u;
                x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
                w[q] = x;
            }
        }
        reinterpret_cast<uint4*>(smem)[e] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> tensor-core (async-proxy) reads
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
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
        const uint64_t ad = make_desc(smem_u32(smem));
        const uint64_t bd = make_desc(smem_u32(smem + TM * KB));
        for (int b = 0; b < batches; ++b) {
            if (b >= 2) mbar_wait(&bar[b & 1], ((b >> 1) - 1) & 1);  // the batch issued two rounds ago has retired
            if (elect_one()) {
#pragma unroll
                for (int i = 0; i < PEAK_BATCH; ++i)
                    umma_i8(tmem_base + (uint32_t)((i & 1) * PEAK_N), ad + (uint64_t)((i & 1) * ((2 * LBO) >> 4)),
                            bd + (uint64_t)((i & 1) * ((2 * LBO) >> 4)), idesc(n_cols), (b > 0 || i > 1) ? 1u : 0u);
                umma_commit(&bar[b & 1]);
            }
            __syncwarp();
        }
        for (int b = (batches > 2 ? batches - 2 : 0); b < batches; ++b) mbar_wait(&bar[b & 1], (b >> 1) & 1);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

}  // namespace i8
}  // namespace tb200
