// Experiment / bring-up harness for ozaki_i8.cuh (not part of the library): random byte slices in the
// canonical tiled layout, tcgen05 kernel vs a plain integer reference kernel, then timing at the
// predict-variance shape.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o build/i8_test i8_test.cu
//   build/i8_test [n=512] [batch=256] [check_cand_tiles=all]
#include <stdlib.h>

#include <vector>

#include "../ozaki_i8.cuh"

namespace tb200 {
thread_local char g_err[512] = "";
int64_t g_launches = 0;
}  // namespace tb200
using namespace tb200;

#define CK(x)                                                                              \
    do {                                                                                   \
        cudaError_t e = (x);                                                               \
        if (e != cudaSuccess) {                                                            \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                       \
        }                                                                                  \
    } while (0)

// reference: one thread per (candidate, Linv row) of the checked candidate tiles
template <int GMAX>
__global__ void reference_kernel(const i8::Params p, int ct0, int nct, double* out /*[nct*128][rowTiles*64]*/) {
    const int rows = p.rowTiles * i8::TN;
    const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long)nct * i8::TM * rows) return;
    const int r = (int)(e % rows);
    const int cl = (int)(e / rows);
    const int ct = ct0 + cl / i8::TM, c = cl % i8::TM;
    const int rt = r / i8::TN, rr = r % i8::TN;
    long long G[GMAX + 1];
    for (int g = 0; g <= GMAX; ++g) G[g] = 0;
    for (int kb = 0; kb <= rt; ++kb) {
        for (int k = 0; k < i8::KB; ++k) {
            int a[i8::S], b[i8::S];
            for (int s = 0; s < i8::S; ++s) {
                a[s] = p.candS[(size_t)s * p.candSliceStride + ((size_t)ct * p.kblks + kb) * i8::A_TILE + i8::tile_offset(c, k)];
                const uint8_t bv =
                    p.linvS[(size_t)s * p.linvSliceStride + ((size_t)rt * p.kblks + kb) * i8::B_TILE + i8::tile_offset(rr, k)];
                b[s] = (int)(int8_t)bv;  // balanced digits: every Linv slice is s8
            }
            for (int i = 1; i <= i8::S; ++i)
                for (int j = 1; j <= i8::S; ++j)
                    if (i + j <= GMAX) G[i + j] += (long long)a[i - 1] * b[j - 1];
        }
    }
    double v = 0.0;
    for (int g = GMAX; g >= 2; --g) v = fma((double)G[g], ldexp(1.0, 1 - 8 * g), v);
    v *= p.rowscale[r];
    out[e] = v;
}

template <int GMAX>
static int run(int n, int batch, int check_tiles) {
    const int kblks = n / 64, rowTiles = n / 64, candTiles = batch / 128;
    const size_t candSlice = (size_t)candTiles * kblks * i8::A_TILE, linvSlice = (size_t)rowTiles * kblks * i8::B_TILE;
    std::vector<uint8_t> hc(candSlice * i8::S), hl(linvSlice * i8::S);
    srand(7);
    for (auto& v : hc) v = (uint8_t)(rand() & 255);
    for (auto& v : hl) v = (uint8_t)(rand() & 255);
    std::vector<double> hs(n);
    for (int r = 0; r < n; ++r) hs[r] = ldexp(1.0, (rand() % 9) - 4);
    uint8_t *dc, *dl;
    double *ds, *dp, *dref;
    CK(cudaMalloc(&dc, hc.size()));
    CK(cudaMalloc(&dl, hl.size()));
    CK(cudaMalloc(&ds, sizeof(double) * n));
    CK(cudaMalloc(&dp, sizeof(double) * rowTiles * batch));
    CK(cudaMemcpy(dc, hc.data(), hc.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dl, hl.data(), hl.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ds, hs.data(), sizeof(double) * n, cudaMemcpyHostToDevice));
    CK(cudaMemset(dp, 0xff, sizeof(double) * rowTiles * batch));
    i8::Params p{dc, dl, ds, dp, batch, (long)candSlice, (long)linvSlice, kblks, rowTiles, candTiles};

    if (i8::launch_colsumsq_i8<GMAX>(p, 0)) {
        printf("launch failed: %s\n", g_err);
        return 1;
    }
    CK(cudaDeviceSynchronize());
    std::vector<double> hp((size_t)rowTiles * batch);
    CK(cudaMemcpy(hp.data(), dp, sizeof(double) * hp.size(), cudaMemcpyDeviceToHost));

    if (check_tiles > candTiles) check_tiles = candTiles;
    const long nref = (long)check_tiles * i8::TM * n;
    CK(cudaMalloc(&dref, sizeof(double) * nref));
    reference_kernel<GMAX><<<(unsigned)((nref + 255) / 256), 256>>>(p, 0, check_tiles, dref);
    CK(cudaDeviceSynchronize());
    std::vector<double> hr(nref);
    CK(cudaMemcpy(hr.data(), dref, sizeof(double) * nref, cudaMemcpyDeviceToHost));
    double maxrel = 0.0;
    int bad = 0;
    for (int cl = 0; cl < check_tiles * i8::TM; ++cl)
        for (int rt = 0; rt < rowTiles; ++rt) {
            double s = 0.0;
            for (int rr = 0; rr < i8::TN; ++rr) {
                const double v = hr[(size_t)cl * n + rt * i8::TN + rr];
                s = fma(v, v, s);
            }
            const double got = hp[(size_t)rt * batch + cl];
            const double rel = fabs(got - s) / (fabs(s) + 1e-300);
            if (!(rel < 1e-12)) {
                if (bad < 8) printf("  mismatch cand %d rowtile %d: got %.17g want %.17g\n", cl, rt, got, s);
                ++bad;
            }
            if (rel > maxrel) maxrel = rel;
        }
    printf("GMAX=%d n=%d batch=%d: checked %d cand tiles x %d row tiles, mismatches %d, maxrel %.3e\n", GMAX, n, batch,
           check_tiles, rowTiles, bad, maxrel);

    // timing
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
        CK(cudaEventRecord(e0));
        i8::launch_colsumsq_i8<GMAX>(p, 0);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float t;
        CK(cudaEventElapsedTime(&t, e0, e1));
        if (t < best) best = t;
    }
    int pairs = 0;
    for (int i = 1; i <= i8::S; ++i)
        for (int j = 1; j <= i8::S; ++j) pairs += (i + j <= GMAX);
    const double flop64 = (double)n * n * batch;                    // FP64-equivalent algorithmic flop
    const double macs = 0.5 * (double)n * (n + 64) * batch * pairs;  // INT8 MACs issued
    printf("  %.3f ms: %.2f TFLOP/s FP64-equivalent, %.1f TMAC/s INT8 (%d pairs), %.3e cand/s\n", best,
           flop64 / (best * 1e-3) / 1e12, macs / (best * 1e-3) / 1e12, pairs, batch / (best * 1e-3));
    cudaFree(dc); cudaFree(dl); cudaFree(ds); cudaFree(dp); cudaFree(dref);
    return bad != 0;
}

int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 512;
    const int batch = argc > 2 ? atoi(argv[2]) : 256;
    const int check = argc > 3 ? atoi(argv[3]) : 1 << 30;
    int rc = run<8>(n, batch, check);
    rc |= run<9>(n, batch, check);
    return rc;
}
