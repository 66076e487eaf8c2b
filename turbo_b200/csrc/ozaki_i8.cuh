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
    unsigned* sched;         // [2] tile counter + finished-CTA counter of the persistent grid: zero at launch
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
// the same MMA with a collector hint for the A operand: COLL = 1 keeps A in the collector buffer after this instruction
// (collector::a::fill), 2 takes it from there instead of reading shared memory again (collector::a::lastuse)
template <int COLL>
__device__ __forceinline__ void umma_i8_coll(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
    if constexpr (COLL == 1) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8.collector::a::fill [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
            :
            : "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
            : "memory");
    } else if constexpr (COLL == 2) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8.collector::a::lastuse [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
            :
            : "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
            : "memory");
    } else {
        umma_i8(d_tmem, adesc, bdesc, idesc, accumulate);
    }
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
template <int N_LO, int N_HI, int ACC, int ASGN = 0, int FIRST_LO = N_LO>
__device__ __forceinline__ void issue_range(uint32_t d_tmem, uint64_t ad, uint64_t bd0) {
    if constexpr (N_HI > N_LO) {
        constexpr int width = N_HI - N_LO;
        constexpr int nchunks = (width + 255) / 256;
        constexpr int chunk = ((width / nchunks) + 15) & ~15;
        constexpr int nc = chunk < width ? chunk : width;
        // a range cut into several chunks reads the same A tile for each: the first keeps it in the collector buffer,
        // the last takes it from there (4 KB of shared-memory reads less per extra chunk)
        constexpr int coll = (nc < width) ? 1 : (N_LO > FIRST_LO ? 2 : 0);
        umma_i8_coll<coll>(d_tmem + N_LO, ad, bd0 + (uint64_t)((N_LO >> 3) * (SBO >> 4)), idesc(nc, ASGN), (uint32_t)ACC);
        issue_range<N_LO + nc, N_HI, ACC, ASGN, FIRST_LO>(d_tmem, ad, bd0);
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
constexpr int RING = 8;       // tile ids in flight between the producer and the epilogue (at most 4 are live)

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// digit planes a kernel with cut-off GMAX reads: plane s pairs with planes 1..GMAX-s, so plane 7 is unused at GMAX = 7
__host__ __device__ constexpr int planes_used(int gmax) { return gmax - 1 < S ? gmax - 1 : S; }

// sum_g 2^(1-8(g+2)) G_g for one column of the NG accumulator groups -- on the INTEGER pipe as far as it goes.  The
// round-1 epilogue converted every accumulator (I2F.F64, a quarter-rate instruction: 48 per 8 columns) and chained NG
// DFMAs; next to the K* kernel, which lives on the FP64 pipe, that made the drain of the accumulators -- the only time a
// persistent CTA's tensor pipe waits -- the contended part.  Here three neighbouring groups are merged exactly into one
// 64-bit integer, T = G_g 2^16 + G_{g+1} 2^8 + G_{g+2} (|T| < 2^48), and T 2^E is turned into a double without the
// conversion unit: adding T to the bit pattern of 1.5 2^(52+E) gives the double (1.5 2^52 + T) 2^E, and subtracting
// 1.5 2^(52+E) is exact.  Per column: ~10 integer instructions and NG/3 + 1 FP64 additions instead of NG conversions and
// NG DFMAs.
template <int E>
__device__ __forceinline__ double s64_times_pow2(long long t) {
    constexpr long long C = ((long long)(1075 + E) << 52) | (1ll << 51);
    return __longlong_as_double(t + C) - __longlong_as_double(C);
}
// groups [G0, NG) of column c, three at a time from the least significant end; E_LAST = weight exponent of group NG-1
template <int NG, int HI>
__device__ __forceinline__ double recombine_from(const uint32_t (&r)[NG][8], int c, double v) {
    if constexpr (HI < 0) {
        return v;
    } else {
        constexpr int LO = HI - 2 < 0 ? 0 : HI - 2;  // merge groups LO..HI (1 to 3 of them)
        long long t = (long long)(int)r[HI][c];
        if constexpr (HI - 1 >= LO) t += (long long)(int)r[HI - 1][c] << 8;
        if constexpr (HI - 2 >= LO) t += (long long)(int)r[HI - 2][c] << 16;
        const double x = s64_times_pow2<1 - 8 * (HI + 2)>(t);
        return recombine_from<NG, LO - 1>(r, c, HI == NG - 1 ? x : v + x);
    }
}
template <int NG>
__device__ __forceinline__ double recombine(const uint32_t (&r)[NG][8], int c) {
    return recombine_from<NG, NG - 1>(r, c, 0.0);
}

// PERSISTENT kernel (round 2): one CTA per SM walks (64 Linv rows) x (128 candidates) tiles handed out heavy-first by a
// global counter.  Round 1 launched one CTA per tile; ncu showed 25 % of the SM cycles outside the MMAs -- CTA launch,
// barrier/TMEM set-up, the first loads' latency and the epilogue, ~15 k cycles per tile, none of it overlapped because a
// 173 KB CTA has no neighbour (profiles/r2_variance_ncu_full.md).  Now the producer warp runs straight on into the next
// tile's loads, TMEM and barriers are set up once, and only the drain of the accumulators (one TMEM buffer: 6 x 64
// columns do not fit twice) separates two tiles' MMAs.
//   warp 0 : fetches the next tile id (atomicAdd), publishes it in `tile_ring` with the first stage of the tile
//   warp 1 : MMA issuer; starts a tile (or a pass) once the epilogue has drained the previous one (tfree_bar)
//   2..5   : epilogue; releases TMEM right after its last tcgen05.ld, before the arithmetic of the last columns
// p.sched = {next tile, CTAs finished}: zero at launch; the last CTA to finish zeroes it again for the next launch
// (launches that share one sched pair must be stream-ordered -- they are: one variance stream per call).
template <int GMAX, bool MULTI>
__global__ void __launch_bounds__(THREADS, 1) colsumsq_i8_kernel(const Params p) {
    constexpr int NG = GMAX - 1;  // groups g = 2..GMAX, one 64-column INT32 accumulator each
    constexpr int SU = planes_used(GMAX);
    static_assert(NG * TN <= TMEM_COLS, "accumulators exceed TMEM");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], done_bar, tfree_bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ volatile int tile_ring[RING];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = p.rowTiles * p.candTiles;

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
        // ===== producer: 2*SU bulk copies per k-block, issued by one elected lane of a converged warp =====
        uint32_t it = 0;  // k-blocks issued so far (stage = it % STAGES)
        for (uint32_t ti = 0;; ++ti) {
            int t = 0;
            if (lane == 0) t = (int)atomicAdd(p.sched, 1u);
            t = __shfl_sync(0xffffffffu, t, 0);
            const bool last = t >= ntiles;  // the end marker travels through the pipeline like a tile of one k-block
            // heavy row tiles first (the k-range grows with the row tile)
            const int rt = last ? 0 : p.rowTiles - 1 - t / p.candTiles;
            const int ct = last ? 0 : t % p.candTiles;
            const int nkb = rt + 1;
            const uint8_t* a_src = p.candS + ((size_t)ct * p.kblks) * A_TILE;
            const uint8_t* b_src = p.linvS + ((size_t)rt * p.kblks) * B_TILE;
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % STAGES;
                if (it >= STAGES) mbar_wait(&empty_bar[s], ((it / STAGES) - 1) & 1);
                if (elect_one()) {
                    // the tile id is published by the arrive on the first stage's barrier (release -> the waiter's acquire)
                    if (kb == 0) tile_ring[ti % RING] = last ? -1 : t;
                    if (last) {
                        mbar_arrive(&full_bar[s]);
                    } else {
                        uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                        mbar_expect_tx(&full_bar[s], SU * (A_TILE + B_TILE));
#pragma unroll
                        for (int sl = 0; sl < SU; ++sl) {
                            bulk_g2s(st + sl * A_TILE, a_src + (size_t)sl * p.candSliceStride + (size_t)kb * A_TILE,
                                     A_TILE, &full_bar[s]);
                            bulk_g2s(st + S * A_TILE + sl * B_TILE,
                                     b_src + (size_t)sl * p.linvSliceStride + (size_t)kb * B_TILE, B_TILE, &full_bar[s]);
                        }
                    }
                }
                __syncwarp();
            }
            if (last) break;
        }
        if (lane == 0) {
            // every CTA has taken its end marker once the last one arrives here: hand the next launch a zeroed pair
            if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
                atomicExch(p.sched, 0u);
                atomicExch(p.sched + 1, 0u);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the warp stays converged, one elected lane issues =====
        uint32_t it = 0, passes = 0;  // k-blocks consumed, accumulation passes committed
        for (uint32_t ti = 0;; ++ti) {
            mbar_wait(&full_bar[it % STAGES], (it / STAGES) & 1);  // first stage of the tile: its id is visible now
            const int t = tile_ring[ti % RING];
            if (t < 0) {
                // the epilogue must have consumed the last real completion of done_bar before the end marker's plain
                // arrive may complete another phase
                if (passes > 0) mbar_wait(&tfree_bar, (passes - 1) & 1);
                if (elect_one()) mbar_arrive(&done_bar);
                __syncwarp();
                break;
            }
            const int nkb = p.rowTiles - t / p.candTiles;
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % STAGES;
                const bool pass_start = MULTI ? (kb % PASS_KB == 0) : (kb == 0);
                const bool pass_end = kb == nkb - 1 || (MULTI && (kb + 1) % PASS_KB == 0);
                if (pass_start && passes > 0) mbar_wait(&tfree_bar, (passes - 1) & 1);  // previous accumulators drained
                if (kb > 0) mbar_wait(&full_bar[s], (it / STAGES) & 1);
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
                    issue_step<GMAX, 1, false, 0>(tmem_base, ad0 + (uint64_t)((2 * LBO) >> 4),
                                                  bd0 + (uint64_t)((2 * LBO) >> 4));
                    umma_commit(&empty_bar[s]);  // frees the stage once these MMAs have read it
                    if (pass_end) umma_commit(&done_bar);
                }
                __syncwarp();
                if (pass_end) ++passes;
            }
        }
    } else {
        // ===== epilogue: thread = candidate (TMEM lane), 64 columns = Linv rows of this tile =====
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        uint32_t drains = 0;  // completions of done_bar consumed
        for (uint32_t ti = 0;; ++ti) {
            mbar_wait(&done_bar, drains & 1);
            const int t = tile_ring[ti % RING];
            if (t < 0) break;
            const int rt = p.rowTiles - 1 - t / p.candTiles, ct = t % p.candTiles;
            const double* rs = p.rowscale + rt * TN;
            double acc = 0.0;
            if constexpr (!MULTI) {
                ++drains;
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
                for (int cc = 0; cc < TN / 8; ++cc) {
                    uint32_t r[NG][8];
#pragma unroll
                    for (int g = 0; g < NG; ++g) tmem_ld8(lane_base + g * TN + cc * 8, r[g]);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (cc == TN / 8 - 1) {  // the accumulators are in registers: the next tile's MMAs may start
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        mbar_arrive(&tfree_bar);
                    }
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const double v = recombine<NG>(r, c) * __ldg(rs + cc * 8 + c);
                        acc = fma(v, v, acc);
                    }
                }
            } else {
                double vrow[TN];  // V[candidate, 64 rows] / rowscale, summed over the passes (registers: fully unrolled)
                const int nkb = rt + 1;
                const int npass = (nkb + PASS_KB - 1) / PASS_KB;
                for (int pass = 0; pass < npass; ++pass) {
                    if (pass > 0) mbar_wait(&done_bar, drains & 1);
                    ++drains;
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                    for (int cc = 0; cc < TN / 8; ++cc) {
                        uint32_t r[NG][8];
#pragma unroll
                        for (int g = 0; g < NG; ++g) tmem_ld8(lane_base + g * TN + cc * 8, r[g]);
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const double v = recombine<NG>(r, c);
                            vrow[cc * 8 + c] = (pass == 0) ? v : vrow[cc * 8 + c] + v;
                        }
                    }
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(&tfree_bar);
                }
#pragma unroll
                for (int c = 0; c < TN; ++c) {
                    const double v = vrow[c] * __ldg(rs + c);
                    acc = fma(v, v, acc);
                }
            }
            p.partial[(long)rt * p.ldp + (long)ct * TM + q * 32 + lane] = acc;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

template <int GMAX, bool MULTI>
inline int launch_colsumsq_i8_impl(const Params& p, int sm_count, cudaStream_t stream) {
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
    const int tiles = p.rowTiles * p.candTiles;
    TB_CHECK(p.sched != nullptr && sm_count > 0, "variance kernel: scheduler counters missing");
    TB_LAUNCH((colsumsq_i8_kernel<GMAX, MULTI>), dim3((unsigned)(tiles < sm_count ? tiles : sm_count)), dim3(THREADS),
              SMEM_BYTES, stream, p);
    return 0;
}

// sm_count: CTAs of the persistent grid (one per SM; 173 KB of shared memory keeps a second one off the SM)
template <int GMAX>
inline int launch_colsumsq_i8(const Params& p, int sm_count, cudaStream_t stream) {
    // the heaviest tile spans rowTiles k-blocks
    return p.rowTiles <= PASS_KB ? launch_colsumsq_i8_impl<GMAX, false>(p, sm_count, stream)
                                 : launch_colsumsq_i8_impl<GMAX, true>(p, sm_count, stream);
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

// ---------------------------------------------------------------------------------------------
// Diagnostic (tb200_i8_mma_list): cycles one k-block costs when an arbitrary LIST of MMAs -- the variance kernel's own
// (A plane, accumulator columns, N, first B row) list or a re-ordered / re-chunked variant of it -- is issued
// back to back on operands that stay resident in shared memory.  No loads, no epilogue: what is measured is what the
// tensor pipe makes of that instruction mix, dependencies between overlapping accumulator ranges included.  A k-block
// is two K = 32 steps like in the kernels above.
// ---------------------------------------------------------------------------------------------
struct MmaOp { int plane, col0, n, brow0, coll = 0; };
constexpr int MMA_LIST_MAX = 64;
constexpr int MMA_LIST_VARIANTS = 14;

// compile-time lists: the issuing lane executes nothing but uniform-register moves between two MMAs, as in the kernels
template <int V> struct MmaList;
#define TB_MMA_LIST(V, ...)                                                    \
    template <> struct MmaList<V> {                                            \
        static constexpr MmaOp ops[] = {__VA_ARGS__};                          \
        static constexpr int n = (int)(sizeof(ops) / sizeof(MmaOp));           \
    }
// 0: the variance kernel's list at 21 products (plane i-1 x stacked planes 1..7-i, balanced chunks)
TB_MMA_LIST(0, {0, 0, 192, 0}, {0, 192, 192, 192}, {1, 64, 160, 0}, {1, 224, 160, 160}, {2, 128, 256, 0}, {3, 192, 192, 0}, {4, 256, 128, 0}, {5, 320, 64, 0});
// 1: the same instructions, ONE A plane;  2: one A plane and B rows from 0
TB_MMA_LIST(1, {0, 0, 192, 0}, {0, 192, 192, 192}, {0, 64, 160, 0}, {0, 224, 160, 160}, {0, 128, 256, 0}, {0, 192, 192, 0}, {0, 256, 128, 0}, {0, 320, 64, 0});
TB_MMA_LIST(2, {0, 0, 192, 0}, {0, 192, 192, 0}, {0, 64, 160, 0}, {0, 224, 160, 0}, {0, 128, 256, 0}, {0, 192, 192, 0}, {0, 256, 128, 0}, {0, 320, 64, 0});
// 3: the kernel's operands and N sequence, accumulators alternating between the aligned halves of TMEM
TB_MMA_LIST(3, {0, 0, 192, 0}, {0, 256, 192, 192}, {1, 0, 160, 0}, {1, 256, 160, 160}, {2, 0, 256, 0}, {3, 256, 192, 0}, {4, 0, 128, 0}, {5, 256, 64, 0});
// 4: uniform: seven MMAs of N = 192 (the same 1344 columns), planes in turn, alternating halves
TB_MMA_LIST(4, {0, 0, 192, 0}, {1, 256, 192, 0}, {2, 0, 192, 0}, {3, 256, 192, 0}, {4, 0, 192, 0}, {5, 256, 192, 0}, {0, 0, 192, 192});
// 5: the kernel's list with every accumulator range starting at column 0
TB_MMA_LIST(5, {0, 0, 192, 0}, {0, 0, 192, 192}, {1, 0, 160, 0}, {1, 0, 160, 160}, {2, 0, 256, 0}, {3, 0, 192, 0}, {4, 0, 128, 0}, {5, 0, 64, 0});
// 6: the kernel's list, short instructions first
TB_MMA_LIST(6, {5, 320, 64, 0}, {4, 256, 128, 0}, {3, 192, 192, 0}, {2, 128, 256, 0}, {1, 64, 160, 0}, {1, 224, 160, 160}, {0, 0, 192, 0}, {0, 192, 192, 192});
// 7: chunks of 256 + remainder
TB_MMA_LIST(7, {0, 0, 256, 0}, {0, 256, 128, 256}, {1, 64, 256, 0}, {1, 320, 64, 256}, {2, 128, 256, 0}, {3, 192, 192, 0}, {4, 256, 128, 0}, {5, 320, 64, 0});
// 8: the kernel's list at 28 products
TB_MMA_LIST(8, {0, 0, 224, 0}, {0, 224, 224, 224}, {1, 64, 192, 0}, {1, 256, 192, 192}, {2, 128, 160, 0}, {2, 288, 160, 160}, {3, 192, 256, 0}, {4, 256, 192, 0}, {5, 320, 128, 0}, {6, 384, 64, 0});
// 9: 21 products, short and long instructions interleaved
TB_MMA_LIST(9, {0, 0, 192, 0}, {5, 320, 64, 0}, {0, 192, 192, 192}, {4, 256, 128, 0}, {1, 64, 160, 0}, {1, 224, 160, 160}, {2, 128, 256, 0}, {3, 192, 192, 0});
// 10: six MMAs of N = 224 (= 1344 columns);  11: five of 256 + one of 64
TB_MMA_LIST(10, {0, 0, 224, 0}, {1, 256, 224, 0}, {2, 0, 224, 0}, {3, 256, 224, 0}, {4, 0, 224, 0}, {5, 256, 224, 0});
TB_MMA_LIST(11, {0, 0, 256, 0}, {1, 256, 256, 0}, {2, 0, 256, 0}, {3, 256, 256, 0}, {4, 0, 256, 0}, {5, 256, 64, 0});
// 12: the kernel's list with the A operand of a two-chunk range kept in the collector for the second chunk
TB_MMA_LIST(12, {0, 0, 192, 0, 1}, {0, 192, 192, 192, 2}, {1, 64, 160, 0, 1}, {1, 224, 160, 160, 2}, {2, 128, 256, 0}, {3, 192, 192, 0}, {4, 256, 128, 0}, {5, 320, 64, 0});
// 13: the same, short and long instructions interleaved (list 9 with the hints)
TB_MMA_LIST(13, {0, 0, 192, 0, 1}, {0, 192, 192, 192, 2}, {5, 320, 64, 0}, {1, 64, 160, 0, 1}, {1, 224, 160, 160, 2}, {4, 256, 128, 0}, {2, 128, 256, 0}, {3, 192, 192, 0});
#undef TB_MMA_LIST

template <int V, int I = 0>
__device__ __forceinline__ void issue_list(uint32_t tmem_base, uint64_t ad0, uint64_t bd0, uint32_t acc) {
    if constexpr (I < MmaList<V>::n) {
        constexpr MmaOp o = MmaList<V>::ops[I];
        umma_i8_coll<o.coll>(tmem_base + (uint32_t)o.col0, ad0 + (uint64_t)(o.plane * (A_TILE >> 4)),
                             bd0 + (uint64_t)((o.brow0 >> 3) * (SBO >> 4)), idesc(o.n), acc);
        issue_list<V, I + 1>(tmem_base, ad0, bd0, acc);
    }
}

template <int V>
__global__ void __launch_bounds__(64, 1) i8_mma_list_kernel(int kblocks, long long* __restrict__ cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar[2];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    for (int e = threadIdx.x; e < (int)(STAGE_BYTES / 16); e += blockDim.x) {
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t x = (uint32_t)(e * 4 + q) * 2654435761u + blockIdx.x * 40503u + 12345u;
            x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
            w[q] = x;
        }
        reinterpret_cast<uint4*>(smem)[e] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
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
        const uint64_t ad0 = make_desc(smem_u32(smem));
        const uint64_t bd0 = make_desc(smem_u32(smem + S * A_TILE));
        const uint64_t ko = (uint64_t)((2 * LBO) >> 4);
        long long t0 = 0;
        for (int b = 0; b < kblocks; ++b) {
            if (b >= 2) mbar_wait(&bar[b & 1], ((b >> 1) - 1) & 1);
            if (b == 2) t0 = clock64();  // the first two k-blocks fill the queue
            if (elect_one()) {
                issue_list<V>(tmem_base, ad0, bd0, b > 0 ? 1u : 0u);
                issue_list<V>(tmem_base, ad0 + ko, bd0 + ko, 1u);
                umma_commit(&bar[b & 1]);
            }
            __syncwarp();
        }
        for (int b = (kblocks > 2 ? kblocks - 2 : 0); b < kblocks; ++b) mbar_wait(&bar[b & 1], (b >> 1) & 1);
        const long long t1 = clock64();
        if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// The same lists inside the kernels' two-stage pipeline: warp 0 refills the stage the MMAs are not reading with
// `planes` digit planes of both operands (cp.async.bulk from an L2-resident scratch buffer; 0 = no copies, the barrier
// is completed by a plain arrive), warp 1 issues list V on the other stage.  No epilogue, no tiles: the difference to
// i8_mma_list_kernel is exactly the fill traffic into shared memory and the stage hand-over.
template <int V>
__global__ void __launch_bounds__(64, 1) i8_mma_pipe_kernel(int kblocks, int planes, const uint8_t* __restrict__ src,
                                                           long src_bytes, long long* __restrict__ cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    for (int e = threadIdx.x; e < (int)(STAGES * STAGE_BYTES / 16); e += blockDim.x) {
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t x = (uint32_t)(e * 4 + q) * 2654435761u + blockIdx.x * 40503u + 12345u;
            x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
            w[q] = x;
        }
        reinterpret_cast<uint4*>(smem)[e] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
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
        const long span = src_bytes - (long)STAGE_BYTES;
        long off = ((long)blockIdx.x * 7919 * 4096) % span;
        for (int it = 0; it < kblocks; ++it) {
            const int s = it % STAGES;
            if (it >= STAGES) mbar_wait(&empty_bar[s], ((it / STAGES) - 1) & 1);
            if (elect_one()) {
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                if (planes == 0) {
                    mbar_arrive(&full_bar[s]);
                } else {
                    mbar_expect_tx(&full_bar[s], planes * (A_TILE + B_TILE));
                    for (int sl = 0; sl < planes; ++sl) {
                        bulk_g2s(st + sl * A_TILE, src + off + (size_t)sl * A_TILE, A_TILE, &full_bar[s]);
                        bulk_g2s(st + S * A_TILE + sl * B_TILE, src + off + S * A_TILE + (size_t)sl * B_TILE, B_TILE,
                                 &full_bar[s]);
                    }
                }
            }
            __syncwarp();
            off = (off + 196608) % span;  // span is a multiple of 4096: every copy stays 16-byte aligned
        }
    } else {
        const uint64_t ko = (uint64_t)((2 * LBO) >> 4);
        long long t0 = 0;
        for (int it = 0; it < kblocks; ++it) {
            const int s = it % STAGES;
            mbar_wait(&full_bar[s], (it / STAGES) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (it == 4) t0 = clock64();
            const uint32_t a0 = smem_u32(smem + (size_t)s * STAGE_BYTES);
            const uint64_t ad0 = make_desc(a0);
            const uint64_t bd0 = make_desc(a0 + S * A_TILE);
            if (elect_one()) {
                issue_list<V>(tmem_base, ad0, bd0, it > 0 ? 1u : 0u);
                issue_list<V>(tmem_base, ad0 + ko, bd0 + ko, 1u);
                umma_commit(&empty_bar[s]);
            }
            __syncwarp();
        }
        // the last commits: every MMA has retired once both stages are "empty" again
        for (int it = kblocks - STAGES; it < kblocks; ++it) mbar_wait(&empty_bar[it % STAGES], (it / STAGES) & 1);
        const long long t1 = clock64();
        if (threadIdx.x == 32) cycles[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

}  // namespace i8
}  // namespace tb200
