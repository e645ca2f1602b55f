// Standalone check of the tcgen05 tap-GEMM (vgemm_umma) against the CUDA-core tile kernel (vgemm_simt)
// on random data: shapes of the vocoder (taps with negative row offsets, N = 96, ragged M), plus timing.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "../../qwen3-tts.cpp_b200/csrc/voc_gemm.cuh"
using namespace q3t;
namespace q3t { int set_err(q3t_ctx *, int c, const std::string &) { return c; } int cuda_check(q3t_ctx *, cudaError_t e, const char *) { return e != cudaSuccess; } }

static float frand() { return (float) rand() / RAND_MAX * 2.f - 1.f; }

static bool run_case(q3t_ctx * ctx, int M, int N, int Kc, int ntaps, const int * offs, int a_rows, bool bias, bool time_it) {
    const int Ktot = ntaps * Kc;
    std::vector<__half> hA((size_t) a_rows * Kc), hW((size_t) N * Ktot);
    for (auto & x : hA) x = __float2half(frand() * 0.5f);
    for (auto & x : hW) x = __float2half(frand() * 0.1f);
    std::vector<float> hb(N);
    for (auto & x : hb) x = frand();
    __half * dA, * dW; float * db, * o1, * o2;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dW, hW.size() * 2); cudaMalloc(&db, N * 4);
    cudaMalloc(&o1, (size_t) M * N * 4); cudaMalloc(&o2, (size_t) M * N * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dW, hW.data(), hW.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(db, hb.data(), N * 4, cudaMemcpyHostToDevice);
    cudaMemset(o1, 0xff, (size_t) M * N * 4); cudaMemset(o2, 0xff, (size_t) M * N * 4);
    VGemm p{};
    p.A = dA; p.lda = Kc; p.a_rows = a_rows; p.M = M; p.N = N; p.ntaps = ntaps; p.Kc = Kc;
    for (int i = 0; i < ntaps; ++i) p.tap_off[i] = offs[i];
    p.W = dW; p.bias = bias ? db : nullptr; p.pmod = N; p.out32 = o1; p.ld32 = N;
    dim3 grid((M + VG_BM - 1) / VG_BM, (N + VG_BN - 1) / VG_BN);
    vgemm_simt<<<grid, 256>>>(p);
    p.out32 = o2;
    setenv("Q3T_VOC_SIMT", "0", 1);
    voc_gemm(ctx, p);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return false; }
    std::vector<float> r1((size_t) M * N), r2((size_t) M * N);
    cudaMemcpy(r1.data(), o1, r1.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(r2.data(), o2, r2.size() * 4, cudaMemcpyDeviceToHost);
    double maxd = 0, maxv = 0; size_t bad = 0;
    for (size_t i = 0; i < r1.size(); ++i) {
        const double d = fabs((double) r1[i] - (double) r2[i]);
        if (!(d <= 1e-3 * (1.0 + fabs(r1[i])))) ++bad;
        if (d > maxd) maxd = d;
        if (fabs(r1[i]) > maxv) maxv = fabs(r1[i]);
    }
    printf("M=%7d N=%4d Kc=%4d taps=%d: max|diff| %.3g (max|ref| %.3g) mismatches %zu %s\n", M, N, Kc, ntaps, maxd, maxv, bad, bad ? "FAIL" : "ok");
    if (time_it && !bad) {
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        for (int k = 0; k < 2; ++k) {
            cudaEventRecord(a);
            for (int i = 0; i < 10; ++i) { if (k == 0) vgemm_simt<<<grid, 256>>>(p); else voc_gemm(ctx, p); }
            cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b);
            printf("    %-5s %8.3f ms  %7.1f TFLOP/s\n", k == 0 ? "simt" : "umma", ms / 10, 2.0 * M * N * Ktot / (ms / 10 * 1e-3) / 1e12);
        }
    }
    cudaFree(dA); cudaFree(dW); cudaFree(db); cudaFree(o1); cudaFree(o2);
    return bad == 0;
}

// fused epilogue: in-place residual (res aliases out32), bias, snake -> F16; the tensor-core kernel uses the SFU
// sine (snake_fast), the CUDA-core kernel libm-accurate sinf: out16 must agree to F16 rounding
static bool run_epi_case(q3t_ctx * ctx, int M, int N, int Kc) {
    std::vector<__half> hA((size_t) M * Kc), hW((size_t) N * Kc);
    for (auto & x : hA) x = __float2half(frand() * 0.5f);
    for (auto & x : hW) x = __float2half(frand() * 0.1f);
    std::vector<float> hb(N), hea(N), heb(N), hres((size_t) M * N);
    for (int i = 0; i < N; ++i) { hb[i] = frand(); hea[i] = expf(frand() * 0.3f) * 3.0f; heb[i] = expf(frand() * 0.3f); }
    for (auto & x : hres) x = frand() * 4.0f;
    __half * dA, * dW, * h1, * h2; float * db, * dea, * deb, * o1, * o2;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dW, hW.size() * 2); cudaMalloc(&db, N * 4); cudaMalloc(&dea, N * 4); cudaMalloc(&deb, N * 4);
    cudaMalloc(&o1, hres.size() * 4); cudaMalloc(&o2, hres.size() * 4); cudaMalloc(&h1, hres.size() * 2); cudaMalloc(&h2, hres.size() * 2);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(dW, hW.data(), hW.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(db, hb.data(), N * 4, cudaMemcpyHostToDevice); cudaMemcpy(dea, hea.data(), N * 4, cudaMemcpyHostToDevice); cudaMemcpy(deb, heb.data(), N * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(o1, hres.data(), hres.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(o2, hres.data(), hres.size() * 4, cudaMemcpyHostToDevice);
    VGemm p{};
    p.A = dA; p.lda = Kc; p.a_rows = M; p.M = M; p.N = N; p.ntaps = 1; p.Kc = Kc; p.tap_off[0] = 0;
    p.W = dW; p.bias = db; p.pmod = N; p.act = VACT_SNAKE; p.ea = dea; p.eb = deb;
    p.res = o1; p.ldres = N; p.out32 = o1; p.ld32 = N; p.out16 = h1; p.ld16 = N;
    dim3 grid((M + VG_BM - 1) / VG_BM, (N + VG_BN - 1) / VG_BN);
    vgemm_simt<<<grid, 256>>>(p);
    p.res = o2; p.out32 = o2; p.out16 = h2;
    voc_gemm(ctx, p);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return false; }
    std::vector<float> r1(hres.size()), r2(hres.size()); std::vector<__half> g1(hres.size()), g2(hres.size());
    cudaMemcpy(r1.data(), o1, r1.size() * 4, cudaMemcpyDeviceToHost); cudaMemcpy(r2.data(), o2, r2.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(g1.data(), h1, g1.size() * 2, cudaMemcpyDeviceToHost); cudaMemcpy(g2.data(), h2, g2.size() * 2, cudaMemcpyDeviceToHost);
    size_t bad32 = 0, bad16 = 0; double m16 = 0;
    for (size_t i = 0; i < r1.size(); ++i) {
        if (!(fabs((double) r1[i] - r2[i]) <= 1e-3 * (1.0 + fabs(r1[i])))) ++bad32;
        const double a = __half2float(g1[i]), b = __half2float(g2[i]), d = fabs(a - b);
        if (!(d <= 2.5e-3 * (1.0 + fabs(a)))) ++bad16;
        if (d > m16) m16 = d;
    }
    printf("epilogue M=%7d N=%4d Kc=%4d: out32 mismatches %zu, out16 mismatches %zu (max|diff| %.3g) %s\n", M, N, Kc, bad32, bad16, m16, (bad32 || bad16) ? "FAIL" : "ok");
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    for (int i = 0; i < 10; ++i) voc_gemm(ctx, p);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    printf("    umma epilogue-bound launch %8.3f ms  (%.0f GB/s of activation traffic)\n", ms / 10, (double) M * N * (2.0 * Kc / N + 4 + 4 + 2) / (ms / 10 * 1e-3) / 1e9);
    cudaFree(dA); cudaFree(dW); cudaFree(db); cudaFree(dea); cudaFree(deb); cudaFree(o1); cudaFree(o2); cudaFree(h1); cudaFree(h2);
    return !(bad32 || bad16);
}

int main() {
    q3t_ctx ctx; ctx.stream = nullptr;
    bool ok = true;
    const int o1[1] = {0}, o7[7] = {-6, -5, -4, -3, -2, -1, 0}, o2[2] = {0, 1}, o7d[7] = {-18, -15, -12, -9, -6, -3, 0};
    ok &= run_case(&ctx, 128, 128, 64, 1, o1, 128, false, false);
    ok &= run_case(&ctx, 300, 96, 96, 1, o1, 300, true, false);
    ok &= run_case(&ctx, 1000, 1536, 512, 1, o1, 1000, true, false);
    ok &= run_case(&ctx, 777, 768, 768, 7, o7, 777, true, false);
    ok &= run_case(&ctx, 2559, 768, 1536, 2, o2, 2560, false, false);
    ok &= run_case(&ctx, 5000, 192, 192, 7, o7d, 5000, true, false);
    const int o7e[7] = {-54, -45, -36, -27, -18, -9, 0}, ofar[2] = {-80, 0}, o3[3] = {-2, -1, 0};
    ok &= run_case(&ctx, 3000, 384, 384, 7, o7e, 3000, true, false);     // dilation 9: 182-row shared A stage
    ok &= run_case(&ctx, 1000, 128, 64, 2, ofar, 1000, true, false);      // taps too far apart: per-step A tiles
    ok &= run_case(&ctx, 130, 512, 1024, 3, o3, 130, false, false);
    ok &= run_epi_case(&ctx, 1000, 96, 96);
    ok &= run_epi_case(&ctx, 598485, 96, 96);
    ok &= run_epi_case(&ctx, 199500, 192, 192);
    if (ok) {
        run_case(&ctx, 20480, 768, 768, 7, o7, 20480, true, true);
        run_case(&ctx, 102400, 384, 384, 7, o7, 102400, true, true);
        run_case(&ctx, 409600, 192, 192, 7, o7, 409600, true, true);
        run_case(&ctx, 614400, 96, 96, 7, o7, 614400, true, true);
    }
    printf(ok ? "ALL OK\n" : "FAILED\n");
    return ok ? 0 : 1;
}
