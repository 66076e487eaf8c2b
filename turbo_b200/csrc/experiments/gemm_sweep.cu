// Experiment harness (not part of the library): sweeps tile shapes / warp layouts / pipeline depths
// of the FP64 DMMA GEMM on the predict-variance shape (V = Linv * Ks^T, lower-triangular A,
// column sum-of-squares epilogue) and measures whether DMMA and DFMA issue concurrently.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o build/gemm_sweep gemm_sweep.cu
//   build/gemm_sweep [n=4096] [batch=2048]
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e = (x);                                                           \
        if (e != cudaSuccess) {                                                        \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                   \
        }                                                                              \
    } while (0)

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

// ------------------------------------------------------------------------------------------
// generic NT kernel: A [M][K] (k contiguous), B [N][K] (k contiguous); A lower-triangular in
// 128-blocks: row tile ti only needs k < (ti+1)*TMrows...; epilogue: column sums of squares
// ------------------------------------------------------------------------------------------
template <int TM, int TN, int TK, int STAGES, int WM, int WN, int MINB, int PREF>
__global__ void __launch_bounds__(WM* WN * 32, MINB)
    gemm_nt_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ P, int lda, int ldb,
                   int ldp, int tilesM, int tilesN, int K) {
    constexpr int THREADS = WM * WN * 32;
    constexpr int LD = TK + 4;
    constexpr int A_ELEMS = TM * LD, B_ELEMS = TN * LD;
    constexpr int MI = TM / WM / 8, NJ = TN / WN / 8;
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp % WM, wn = warp / WM;
    const int g = lane >> 2, kk = lane & 3;

    int ti = blockIdx.x / tilesN;
    const int tj = blockIdx.x - ti * tilesN;
    ti = tilesM - 1 - ti;  // heavy rows first
    const int row0 = ti * TM, col0 = tj * TN;
    int khi = (row0 + TM + 127) / 128 * 128;
    if (khi > K) khi = K;
    const int nk = khi / TK;

    double acc[MI][NJ][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto sA = [&](int st) { return smem + (size_t)st * (A_ELEMS + B_ELEMS); };
    auto sB = [&](int st) { return smem + (size_t)st * (A_ELEMS + B_ELEMS) + A_ELEMS; };
    auto issue = [&](int kt) {
        if (kt < nk) {
            const int st = kt % STAGES, k0 = kt * TK;
            constexpr int CPR = TK / 2;  // 16-byte chunks per row
            double* a_s = sA(st);
            double* b_s = sB(st);
#pragma unroll
            for (int c = tid; c < TM * CPR; c += THREADS) {
                const int r = c / CPR, kc = c % CPR;
                cp_async16(a_s + r * LD + kc * 2, A + (long)(row0 + r) * lda + k0 + kc * 2);
            }
#pragma unroll
            for (int c = tid; c < TN * CPR; c += THREADS) {
                const int r = c / CPR, kc = c % CPR;
                cp_async16(b_s + r * LD + kc * 2, B + (long)(col0 + r) * ldb + k0 + kc * 2);
            }
        }
        cp_async_commit();
    };

#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) issue(s);

    if (PREF == 0) {
        for (int kt = 0; kt < nk; ++kt) {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            issue(kt + STAGES - 1);
            const double* a_s = sA(kt % STAGES) + (wm * (TM / WM) + g) * LD + kk;
            const double* b_s = sB(kt % STAGES) + (wn * (TN / WN) + g) * LD + kk;
#pragma unroll
            for (int ks = 0; ks < TK / 4; ++ks) {
                double af[MI], bf[NJ];
#pragma unroll
                for (int i = 0; i < MI; ++i) af[i] = a_s[i * 8 * LD + ks * 4];
#pragma unroll
                for (int j = 0; j < NJ; ++j) bf[j] = b_s[j * 8 * LD + ks * 4];
#pragma unroll
                for (int i = 0; i < MI; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
            }
        }
    } else {
        // PREF: tile kt+1 is guaranteed landed one iteration early (needs STAGES >= 4), so the first
        // fragments of the next tile are fetched BEFORE the barrier and the DMMA pipe never waits on LDS
        double af[MI], bf[NJ];
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            const double* a_s = sA(0) + (wm * (TM / WM) + g) * LD + kk;
            const double* b_s = sB(0) + (wn * (TN / WN) + g) * LD + kk;
#pragma unroll
            for (int i = 0; i < MI; ++i) af[i] = a_s[i * 8 * LD];
#pragma unroll
            for (int j = 0; j < NJ; ++j) bf[j] = b_s[j * 8 * LD];
        }
        for (int kt = 0; kt < nk; ++kt) {
            // tile kt+1 landed for everyone after this barrier; stage (kt-1)%STAGES is free
            cp_async_wait<STAGES - 3>();
            __syncthreads();
            issue(kt + STAGES - 1);
            const double* a_s = sA(kt % STAGES) + (wm * (TM / WM) + g) * LD + kk;
            const double* b_s = sB(kt % STAGES) + (wn * (TN / WN) + g) * LD + kk;
            const double* a_n = sA((kt + 1) % STAGES) + (wm * (TM / WM) + g) * LD + kk;
            const double* b_n = sB((kt + 1) % STAGES) + (wn * (TN / WN) + g) * LD + kk;
#pragma unroll
            for (int ks = 0; ks < TK / 4; ++ks) {
                double an[MI], bn[NJ];
                if (ks + 1 < TK / 4) {
#pragma unroll
                    for (int i = 0; i < MI; ++i) an[i] = a_s[i * 8 * LD + (ks + 1) * 4];
#pragma unroll
                    for (int j = 0; j < NJ; ++j) bn[j] = b_s[j * 8 * LD + (ks + 1) * 4];
                } else if (kt + 1 < nk) {
#pragma unroll
                    for (int i = 0; i < MI; ++i) an[i] = a_n[i * 8 * LD];
#pragma unroll
                    for (int j = 0; j < NJ; ++j) bn[j] = b_n[j * 8 * LD];
                }
#pragma unroll
                for (int i = 0; i < MI; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
#pragma unroll
                for (int i = 0; i < MI; ++i) af[i] = an[i];
#pragma unroll
                for (int j = 0; j < NJ; ++j) bf[j] = bn[j];
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();

    // column sums of squares over the tile's TM rows -> P[ti][col]
    double* red = smem;  // [WM][TN]
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            s0 += acc[i][j][0] * acc[i][j][0];
            s1 += acc[i][j][1] * acc[i][j][1];
        }
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, o);
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        }
        if (g == 0) {
            red[wm * TN + wn * (TN / WN) + j * 8 + kk * 2] = s0;
            red[wm * TN + wn * (TN / WN) + j * 8 + kk * 2 + 1] = s1;
        }
    }
    __syncthreads();
    for (int c = tid; c < TN; c += THREADS) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < WM; ++w) s += red[w * TN + c];
        P[(long)ti * ldp + col0 + c] = s;
    }
}

// ------------------------------------------------------------------------------------------
// mixed-pipe microbenchmark: warps with (warp & 1) == 0 run DMMA chains, the others DFMA chains
// mode 0: all DMMA, 1: all DFMA, 2: half/half
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) mixed_peak_kernel(double* out, int iters, int mode) {
    const int warp = threadIdx.x >> 5;
    const bool use_dmma = mode == 0 || (mode == 2 && (warp & 1) == 0);
    double s = 0.0;
    if (use_dmma) {
        double c[16][2];
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i][0] = c[i][1] = 0.0;
        const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) dmma884(c[i][0], c[i][1], a, b);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
    } else {
        double c[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i] = i;
        const double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
        for (int it = 0; it < iters * 8; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) c[i] = fma(c[i], a, b);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) s += c[i];
    }
    if (s == 123.456) out[0] = s;
}

static float time_it(void (*fn)(void*), void* ctx, int reps) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    fn(ctx);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0));
        fn(ctx);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float t;
        CK(cudaEventElapsedTime(&t, e0, e1));
        if (t < best) best = t;
    }
    CK(cudaGetLastError());
    return best;
}

struct Ctx {
    const double *A, *B;
    double* P;
    int n, batch;
};

template <int TM, int TN, int TK, int STAGES, int WM, int WN, int MINB, int PREF>
static void launch_variant(void* c) {
    Ctx* x = (Ctx*)c;
    constexpr size_t smem = (size_t)STAGES * (TM + TN) * (TK + 4) * sizeof(double);
    static bool cfg = false;
    if (!cfg) {
        CK(cudaFuncSetAttribute(gemm_nt_kernel<TM, TN, TK, STAGES, WM, WN, MINB, PREF>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cfg = true;
    }
    const int tilesM = x->n / TM, tilesN = x->batch / TN;
    gemm_nt_kernel<TM, TN, TK, STAGES, WM, WN, MINB, PREF>
        <<<tilesM * tilesN, WM * WN * 32, smem>>>(x->A, x->B, x->P, x->n, x->n, x->batch, tilesM, tilesN, x->n);
}

template <int TM, int TN, int TK, int STAGES, int WM, int WN, int MINB, int PREF>
static void run_variant(Ctx& x, const std::vector<double>& ref, std::vector<double>& host) {
    const float ms = time_it(launch_variant<TM, TN, TK, STAGES, WM, WN, MINB, PREF>, &x, 8);
    const int tilesM = x.n / TM;
    CK(cudaMemcpy(host.data(), x.P, sizeof(double) * tilesM * x.batch, cudaMemcpyDeviceToHost));
    // total column sums vs reference (sum over row tiles)
    double maxrel = 0.0;
    for (int c = 0; c < x.batch; ++c) {
        double s = 0.0;
        for (int t = 0; t < tilesM; ++t) s += host[(size_t)t * x.batch + c];
        if (!ref.empty()) {
            const double r = fabs(s - ref[c]) / fabs(ref[c]);
            if (r > maxrel) maxrel = r;
        }
    }
    const double flop = (double)x.n * x.n * x.batch;
    printf("TM=%3d TN=%3d TK=%2d ST=%d warps=%dx%d minb=%d pref=%d : %8.3f ms  %6.2f TFLOP/s  maxrel=%.2e\n", TM, TN,
           TK, STAGES, WM, WN, MINB, PREF, ms, flop / (ms * 1e-3) / 1e12, maxrel);
    fflush(stdout);
}

int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 4096;
    const int batch = argc > 2 ? atoi(argv[2]) : 2048;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    printf("SMs %d, n %d, batch %d\n", sms, n, batch);

    // ---- mixed pipe test ----
    {
        double* out;
        CK(cudaMalloc(&out, 64));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        for (int threads = 256; threads <= 512; threads += 256)
            for (int mode = 0; mode < 3; ++mode) {
                const int iters = 20000;
                float best = 1e30f;
                for (int rep = 0; rep < 4; ++rep) {
                    CK(cudaEventRecord(e0));
                    mixed_peak_kernel<<<sms, threads>>>(out, iters, mode);
                    CK(cudaEventRecord(e1));
                    CK(cudaEventSynchronize(e1));
                    float t;
                    CK(cudaEventElapsedTime(&t, e0, e1));
                    if (rep > 0 && t < best) best = t;
                }
                const int warps = threads / 32;
                const double wd = mode == 0 ? warps : mode == 1 ? 0 : warps / 2;
                const double wf = mode == 1 ? warps : mode == 0 ? 0 : warps / 2;
                const double fma = (wd * 256.0 * 16.0 * iters + wf * 32.0 * 16.0 * iters * 8.0) * sms;
                printf("mixed peak threads=%d mode=%d (0 dmma,1 dfma,2 half/half): %.3f ms  %.2f TFLOP/s\n", threads,
                       mode, best, 2.0 * fma / (best * 1e-3) / 1e12);
            }
        CK(cudaFree(out));
    }

    // ---- data: A lower-triangular random, B random ----
    std::vector<double> hA((size_t)n * n, 0.0), hB((size_t)batch * n);
    srand(1);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) hA[(size_t)i * n + j] = (rand() / (double)RAND_MAX - 0.5) / sqrt((double)n);
    for (size_t e = 0; e < hB.size(); ++e) hB[e] = rand() / (double)RAND_MAX - 0.5;
    double *A, *B, *P;
    CK(cudaMalloc(&A, sizeof(double) * n * n));
    CK(cudaMalloc(&B, sizeof(double) * batch * n));
    CK(cudaMalloc(&P, sizeof(double) * (n / 32) * batch));
    CK(cudaMemcpy(A, hA.data(), sizeof(double) * n * n, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(B, hB.data(), sizeof(double) * batch * n, cudaMemcpyHostToDevice));
    Ctx x{A, B, P, n, batch};
    std::vector<double> host((size_t)(n / 32) * batch), ref;

    // reference column sums from the baseline variant (first 64 columns cross-checked on the host)
    run_variant<128, 128, 16, 3, 4, 2, 1, 0>(x, ref, host);
    ref.resize(batch);
    for (int c = 0; c < batch; ++c) {
        double s = 0.0;
        for (int t = 0; t < n / 128; ++t) s += host[(size_t)t * batch + c];
        ref[c] = s;
    }
    {
        double maxrel = 0.0;
        for (int c = 0; c < 8; ++c) {
            double s = 0.0;
            for (int i = 0; i < n; ++i) {
                double v = 0.0;
                for (int k = 0; k <= i; ++k) v += hA[(size_t)i * n + k] * hB[(size_t)c * n + k];
                s += v * v;
            }
            maxrel = fmax(maxrel, fabs(s - ref[c]) / fabs(s));
        }
        printf("baseline vs host (8 columns): maxrel %.2e\n", maxrel);
    }
    //            TM   TN  TK ST WM WN MINB PREF
    run_variant<128, 128, 16, 3, 4, 2, 1, 0>(x, ref, host);
    run_variant<128, 128, 16, 4, 4, 2, 1, 0>(x, ref, host);
    run_variant<128, 128, 16, 4, 4, 2, 1, 1>(x, ref, host);
    run_variant<128, 128, 16, 5, 4, 2, 1, 1>(x, ref, host);
    run_variant<128, 128, 32, 3, 4, 2, 1, 0>(x, ref, host);
    run_variant<128, 128, 8, 4, 4, 2, 1, 0>(x, ref, host);
    run_variant<128, 128, 8, 6, 4, 2, 1, 1>(x, ref, host);
    run_variant<128, 128, 16, 3, 4, 4, 1, 0>(x, ref, host);   // 16 warps, warp tile 32x32
    run_variant<128, 128, 16, 4, 4, 4, 1, 1>(x, ref, host);
    run_variant<128, 128, 16, 3, 2, 4, 1, 0>(x, ref, host);   // 8 warps, warp tile 64x32
    run_variant<64, 128, 16, 3, 2, 4, 2, 0>(x, ref, host);    // 2 CTAs/SM, warp tile 32x32
    run_variant<64, 128, 16, 4, 2, 4, 2, 1>(x, ref, host);
    run_variant<128, 64, 16, 3, 4, 2, 2, 0>(x, ref, host);    // 2 CTAs/SM, warp tile 32x32
    run_variant<128, 64, 16, 4, 4, 2, 2, 1>(x, ref, host);
    run_variant<64, 64, 16, 4, 2, 2, 3, 0>(x, ref, host);     // 4 warps, 3 CTAs/SM
    run_variant<128, 256, 16, 3, 4, 4, 1, 0>(x, ref, host);   // 16 warps, warp tile 32x64, big CTA tile
    run_variant<256, 128, 16, 3, 4, 4, 1, 0>(x, ref, host);   // 16 warps, warp tile 64x32
    return 0;
}
