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

// Layout of the VARIANCE kernel's operands (K* digit planes, Linv digit planes) and its pipeline granularity: one stage
// = one tile of V_KB bytes of K per plane.  64: 2 stages of 84 KB (round 1).  32: 4 stages of 42 KB -- the same shared
// memory, but a stage is refilled while THREE others are consumed, so the L2 round trip of a refill is covered by
// 3 x 896 cycles of MMAs instead of 1 x 1792 (the kernel measured 2336 cycles per 64 bytes of K against 1792 of MMAs).
// The batched-fit GEMM (ozaki_gemm.cuh) keeps the 64-byte tiles above.
#ifndef TB200_VAR_KB
#define TB200_VAR_KB 64
#endif
namespace v {
constexpr int KB = TB200_VAR_KB;
static_assert(KB == 32 || KB == 64, "variance tiles hold 32 or 64 bytes of K");
constexpr int A_TILE = TM * KB;
constexpr int B_TILE = TN * KB;
constexpr int SBO = (KB / 16) * 128;
constexpr int STAGE_BYTES = S * (A_TILE + B_TILE);
constexpr int STAGES = (2 * 86016) / STAGE_BYTES;
constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024;
constexpr int PASS_KB = 8192 / KB;  // k-blocks of one exact INT32 accumulation pass
}  // namespace v

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
    int kblks;               // K / v::KB (tiles per row of tiles)
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
// multicast variants (thread-block clusters): the copy lands at the same CTA-relative offset in every CTA of `mask` and
// signals the mbarrier at the same offset in each of them; the commit arrives on that barrier in every CTA of `mask`
__device__ __forceinline__ void bulk_g2s_multicast(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar,
                                                   uint16_t mask) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
        : "memory");
}
__device__ __forceinline__ void umma_commit_multicast(uint64_t* bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(mask)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, int sbo = SBO) {
    // K-major, SWIZZLE_NONE: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48)
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
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
// d_tmem + [N_LO, N_HI).  Balanced chunks of <= 256 columns (multiples of 16), all resolved at compile time so the
// issuing lane executes a handful of uniform-register adds per MMA.
template <int N_LO, int N_HI, int ACC, int ASGN = 0, int KBT = KB>
__device__ __forceinline__ void issue_range(uint32_t d_tmem, uint64_t ad, uint64_t bd0) {
    if constexpr (N_HI > N_LO) {
        constexpr int width = N_HI - N_LO;
        constexpr int nchunks = (width + 255) / 256;
        constexpr int chunk = ((width / nchunks) + 15) & ~15;
        constexpr int nc = chunk < width ? chunk : width;
        constexpr int sbo = (KBT / 16) * 128;
        umma_i8(d_tmem + N_LO, ad, bd0 + (uint64_t)((N_LO >> 3) * (sbo >> 4)), idesc(nc, ASGN), (uint32_t)ACC);
        issue_range<N_LO + nc, N_HI, ACC, ASGN, KBT>(d_tmem, ad, bd0);
    }
}

// all slice products of one K = 32 step.  FIRST: no accumulator holds a value yet -- groups [0, COVERED) have been
// written by earlier slices of this step, the rest are overwritten (accumulate = 0).
template <int GMAX, int I, bool FIRST, int COVERED, int ASGN = 0, int KBT = KB>
__device__ __forceinline__ void issue_step(uint32_t tmem_base, uint64_t ad0, uint64_t bd0) {
    if constexpr (I <= S) {
        constexpr int J = (GMAX - I < S) ? GMAX - I : S;  // Linv slices j = 1..J pair with a_I
        if constexpr (J > 0) {
            // a_I * b_j lands in group I+j-2: columns (I-1)*TN + (j-1)*TN of the accumulator block
            const uint32_t d = tmem_base + (I - 1) * TN;
            const uint64_t ad = ad0 + (uint64_t)((I - 1) * ((TM * KBT) >> 4));
            if constexpr (!FIRST) {
                issue_range<0, J * TN, 1, ASGN, KBT>(d, ad, bd0);
                issue_step<GMAX, I + 1, false, 0, ASGN, KBT>(tmem_base, ad0, bd0);
            } else {
                constexpr int split0 = (COVERED - (I - 1)) * TN;  // columns below `split` are initialised
                constexpr int split = split0 < 0 ? 0 : (split0 > J * TN ? J * TN : split0);
                issue_range<0, split, 1, ASGN, KBT>(d, ad, bd0);
                issue_range<split, J * TN, 0, ASGN, KBT>(d, ad, bd0);
                constexpr int cov = (I - 1 + J > COVERED) ? I - 1 + J : COVERED;
                issue_step<GMAX, I + 1, true, cov, ASGN, KBT>(tmem_base, ad0, bd0);
            }
        } else {
            issue_step<GMAX, I + 1, FIRST, COVERED, ASGN, KBT>(tmem_base, ad0, bd0);
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

// PAIR = true: launched in clusters of two CTAs that work on the SAME candidate tile and on two ADJACENT Linv row tiles
// (2q, 2q+1).  Both need the same K* digit planes: each CTA fetches every other plane and multicasts it into both shared
// memories, so a pair reads the candidate operand -- two thirds of a stage -- from L2 once instead of twice (84 KB -> 56
// KB per CTA and 64 bytes of K).  The two k-ranges differ by one 64-byte block; the shorter one is extended by the first
// tile above the diagonal of its row tile, which tb200_i8_prepare writes as zero digits (Linv is lower triangular), so
// both CTAs run the same loop and every barrier has a fixed number of arrivals: a stage is refilled once BOTH CTAs'
// MMAs have released it (their commits arrive on both CTAs' empty barriers).
template <int GMAX, bool MULTI, bool PAIR = false>
__global__ void __launch_bounds__(THREADS, 1) colsumsq_i8_kernel(const Params p) {
    constexpr int NG = GMAX - 1;  // groups g = 2..GMAX, one 64-column INT32 accumulator each
    constexpr int SU = planes_used(GMAX);
    static_assert(NG * TN <= TMEM_COLS, "accumulators exceed TMEM");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[v::STAGES], empty_bar[v::STAGES], done_bar, tfree_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // heavy row tiles first (k-range grows with the row tile)
    int rt, ct, nkb;
    uint32_t crank = 0;
    if constexpr (PAIR) {
        crank = cluster_ctarank();  // = blockIdx.x & 1 (cluster dimensions 2 x 1 x 1)
        const int idx = (int)(blockIdx.x >> 1);
        const int rp = p.rowTiles / 2 - 1 - idx / p.candTiles;
        rt = 2 * rp + (int)crank;
        ct = idx % p.candTiles;
        nkb = (2 * rp + 2) * (TN / v::KB);  // the k-range of the pair's longer row tile
    } else {
        rt = p.rowTiles - 1 - (int)(blockIdx.x / p.candTiles);
        ct = blockIdx.x % p.candTiles;
        nkb = (rt + 1) * (TN / v::KB);  // k-blocks covering columns [0, 64 (rt + 1))
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < v::STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], PAIR ? 2 : 1);
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
    if constexpr (PAIR) cluster_sync_all();  // the peer's barriers exist before anything is multicast to them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // ===== producer: 14 bulk copies per k-block, issued by one elected lane of a converged warp =====
        const uint8_t* a_src = p.candS + ((size_t)ct * p.kblks) * v::A_TILE;
        const uint8_t* b_src = p.linvS + ((size_t)rt * p.kblks) * v::B_TILE;
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % v::STAGES;
            if (kb >= v::STAGES) mbar_wait(&empty_bar[s], ((kb / v::STAGES) - 1) & 1);
            if (elect_one()) {
                uint8_t* st = smem + (size_t)s * v::STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], SU * (v::A_TILE + v::B_TILE));
#pragma unroll
                for (int sl = 0; sl < SU; ++sl) {
                    if constexpr (PAIR) {
                        if ((uint32_t)(sl & 1) == crank)  // my half of the planes, for both CTAs
                            bulk_g2s_multicast(st + sl * v::A_TILE,
                                               a_src + (size_t)sl * p.candSliceStride + (size_t)kb * v::A_TILE, v::A_TILE,
                                               &full_bar[s], (uint16_t)3);
                    } else {
                        bulk_g2s(st + sl * v::A_TILE, a_src + (size_t)sl * p.candSliceStride + (size_t)kb * v::A_TILE,
                                 v::A_TILE, &full_bar[s]);
                    }
                    bulk_g2s(st + S * v::A_TILE + sl * v::B_TILE,
                             b_src + (size_t)sl * p.linvSliceStride + (size_t)kb * v::B_TILE, v::B_TILE, &full_bar[s]);
                }
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the warp stays converged, one elected lane issues =====
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % v::STAGES;
            const bool pass_start = MULTI ? (kb % v::PASS_KB == 0) : (kb == 0);
            if (MULTI && pass_start && kb > 0) {
                mbar_wait(&tfree_bar, ((kb / v::PASS_KB) - 1) & 1);  // accumulators of the previous pass drained
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            mbar_wait(&full_bar[s], (kb / v::STAGES) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a0 = smem_u32(smem + (size_t)s * v::STAGE_BYTES);
            const uint64_t ad0 = make_desc(a0, v::SBO);
            const uint64_t bd0 = make_desc(a0 + S * v::A_TILE, v::SBO);
            if (elect_one()) {
                if (pass_start) {
                    issue_step<GMAX, 1, true, 0, 0, v::KB>(tmem_base, ad0, bd0);
                } else {
                    issue_step<GMAX, 1, false, 0, 0, v::KB>(tmem_base, ad0, bd0);
                }
                // K = 32 per MMA: the second half of a 64-byte k-block is 2 LBO further (descriptor units of 16 B)
                if constexpr (v::KB == 64)
                    issue_step<GMAX, 1, false, 0, 0, v::KB>(tmem_base, ad0 + (uint64_t)((2 * LBO) >> 4),
                                                            bd0 + (uint64_t)((2 * LBO) >> 4));
                // frees the stage once these MMAs have read it (PAIR: in both CTAs, whose producers write into it)
                if constexpr (PAIR) umma_commit_multicast(&empty_bar[s], (uint16_t)3);
                else umma_commit(&empty_bar[s]);
                if (kb == nkb - 1 || (MULTI && (kb + 1) % v::PASS_KB == 0)) umma_commit(&done_bar);
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
            const int npass = (nkb + v::PASS_KB - 1) / v::PASS_KB;
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
    if constexpr (PAIR) cluster_sync_all();  // no CTA leaves while its peer may still signal its barriers
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

template <int GMAX, bool MULTI, bool PAIR>
inline int launch_colsumsq_i8_impl(const Params& p, cudaStream_t stream) {
    static DeviceOnce once;
    if (device_once(once, [] {
            TB_CUDA(cudaFuncSetAttribute(colsumsq_i8_kernel<GMAX, MULTI, PAIR>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v::SMEM_BYTES));
            // same (maximum) carve-out as the small kernels that share the SMs with it: CTAs launched under different
            // shared-memory configurations cannot be resident on one SM at the same time
            TB_CUDA(cudaFuncSetAttribute(colsumsq_i8_kernel<GMAX, MULTI, PAIR>,
                                         cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared));
            return 0;
        }))
        return -1;
    if constexpr (PAIR) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(p.rowTiles * p.candTiles));  // rowTiles is even (n is padded to 128)
        cfg.blockDim = dim3(THREADS);
        cfg.dynamicSmemBytes = v::SMEM_BYTES;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        TB_CUDA(cudaLaunchKernelEx(&cfg, colsumsq_i8_kernel<GMAX, MULTI, PAIR>, p));
        ++g_launches;
    } else {
        TB_LAUNCH((colsumsq_i8_kernel<GMAX, MULTI, PAIR>), dim3((unsigned)(p.rowTiles * p.candTiles)), dim3(THREADS),
                  v::SMEM_BYTES, stream, p);
    }
    return 0;
}

template <int GMAX>
inline int launch_colsumsq_i8(const Params& p, cudaStream_t stream, bool pair) {
    // the heaviest tile spans 64 * rowTiles bytes of K; one exact INT32 pass covers 8192
    const bool multi = p.rowTiles * TN > 8192;
    if (pair) return multi ? launch_colsumsq_i8_impl<GMAX, true, true>(p, stream)
                           : launch_colsumsq_i8_impl<GMAX, false, true>(p, stream);
    return multi ? launch_colsumsq_i8_impl<GMAX, true, false>(p, stream)
                 : launch_colsumsq_i8_impl<GMAX, false, false>(p, stream);
}

// ---------------------------------------------------------------------------------------------
// INT8 tensor peak of this device: the roofline denominator of the two kernels above, MEASURED instead of inferred from
// the bf16 figure.  One CTA per SM; one elected lane issues tcgen05.mma kind::i8 M128 x N256 x K32 instructions on
// operands that stay resident in shared memory (zero digits: the tensor pipe's rate does not depend on the values),
// alternating between the two 256-column halves of TMEM.  Two mbarriers alternate so a batch is always queued behind the
// one in flight and the pipe never drains.  No global traffic, no epilogue: nothing but the MMA rate is measured.
// ---------------------------------------------------------------------------------------------
constexpr int PEAK_N = 256, PEAK_BATCH = 16;
constexpr size_t PEAK_SMEM = (size_t)(TM + PEAK_N) * KB + 1024;  // A: 128 x 64 B, B: 256 x 64 B (canonical K-major)

template <int N_COLS>
__global__ void __launch_bounds__(64, 1) i8_peak_kernel(int batches) {
    static_assert(N_COLS == PEAK_N, "one instantiation (a template only so that both translation units may include it)");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar[2];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    for (int e = threadIdx.x; e < (int)((TM + PEAK_N) * KB / 16); e += blockDim.x)
        reinterpret_cast<uint4*>(smem)[e] = make_uint4(0u, 0u, 0u, 0u);
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
                            bd + (uint64_t)((i & 1) * ((2 * LBO) >> 4)), idesc(PEAK_N), (b > 0 || i > 1) ? 1u : 0u);
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
