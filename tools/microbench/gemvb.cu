// Microbenchmark: the inner body of gemv_b (R rows x NB sequences of FHFMA chains fed from shared memory)
// in isolation, to separate issue rate / operand-selector effects from everything else in lm_batch_kernel.
#include <cstdio>
#include <cuda_fp16.h>
__device__ __forceinline__ float flo(unsigned w, unsigned x, float acc) {
    asm("{.reg .b16 wl, wh, xl, xh; mov.b32 {wl, wh}, %1; mov.b32 {xl, xh}, %2; fma.rn.f32.f16 %0, wl, xl, %0;}" : "+f"(acc) : "r"(w), "r"(x));
    return acc;
}
__device__ __forceinline__ float fhi(unsigned w, unsigned x, float acc) {
    asm("{.reg .b16 wl, wh, xl, xh; mov.b32 {wl, wh}, %1; mov.b32 {xl, xh}, %2; fma.rn.f32.f16 %0, wh, xh, %0;}" : "+f"(acc) : "r"(w), "r"(x));
    return acc;
}
__device__ __forceinline__ unsigned g4(const uint4 & v, int e) { return e == 0 ? v.x : e == 1 ? v.y : e == 2 ? v.z : v.w; }
template <int R, int NB, int MODE>
__global__ void __launch_bounds__(384, 1) k(float * out, long long * cyc, int iters) {
    extern __shared__ uint4 sm[];
    uint4 * xh = sm;                 // [NB][128]
    uint4 * wr = sm + NB * 128;      // [R*12][128]
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    for (int i = t; i < NB * 128 + R * 12 * 128; i += blockDim.x) sm[i] = make_uint4(0x3c003800u + i, 0x34003c00u, 0x3c003c00u, 0x38003400u);
    __syncthreads();
    float acc[R][NB];
    for (int i = 0; i < R; ++i) for (int q = 0; q < NB; ++q) acc[i][q] = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int u = 0; u < 4; ++u) {
            uint4 xv[NB];
#pragma unroll
            for (int q = 0; q < NB; ++q) xv[q] = xh[q * 128 + u * 32 + lane];
#pragma unroll
            for (int i = 0; i < R; ++i) {
                const uint4 wv = wr[((warp * R + i) * 4 + u) * 32 + lane];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const unsigned w2 = g4(wv, e);
#pragma unroll
                    for (int q = 0; q < NB; ++q) {
                        if (MODE == 0) { acc[i][q] = flo(w2, g4(xv[q], e), acc[i][q]); acc[i][q] = fhi(w2, g4(xv[q], e), acc[i][q]); }
                        if (MODE == 1) { acc[i][q] = flo(w2, g4(xv[q], e), acc[i][q]); acc[i][q] = flo(w2 >> 16, g4(xv[q], e) >> 16, acc[i][q]); }
                        if (MODE == 2) { const float2 wf = __half22float2(*reinterpret_cast<const __half2 *>(&w2));
                                         const unsigned xx = g4(xv[q], e); const float2 xf = __half22float2(*reinterpret_cast<const __half2 *>(&xx));
                                         acc[i][q] = fmaf(wf.x, xf.x, acc[i][q]); acc[i][q] = fmaf(wf.y, xf.y, acc[i][q]); }
                    }
                }
            }
        }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < R; ++i) for (int q = 0; q < NB; ++q) s += acc[i][q];
    out[blockIdx.x * blockDim.x + t] = s;
    if (t == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int R, int NB, int MODE> void run(const char * name, int threads) {
    float * out; long long * cyc; cudaMalloc(&out, 4 * 1024 * 148); cudaMalloc(&cyc, 8 * 148);
    const int smem = (NB * 128 + R * 12 * 128) * 16;
    cudaFuncSetAttribute(k<R, NB, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int iters = 16;
    k<R, NB, MODE><<<1, threads, smem>>>(out, cyc, iters);
    k<R, NB, MODE><<<1, threads, smem>>>(out, cyc, iters);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double fma_per_warp = (double) iters * 4 * R * 4 * NB * 2;
    printf("%-22s R=%d NB=%d threads=%3d: %8lld cycles, %6.1f cycles per tile (4 u-steps), %.2f cycles per FMA per warp, %.1f lane-FMA/clk/SM (%s)\n",
           name, R, NB, threads, c, (double) c / iters, c / fma_per_warp, fma_per_warp * threads / c, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<4, 8, 0>("FHFMA lo/hi", 32); run<4, 8, 0>("FHFMA lo/hi", 128); run<4, 8, 0>("FHFMA lo/hi", 352);
    run<3, 8, 0>("FHFMA lo/hi", 352); run<1, 8, 0>("FHFMA lo/hi", 352);
    run<4, 8, 1>("FHFMA lo + shifted", 352);
    run<4, 8, 2>("cvt + FFMA", 352);
    run<4, 4, 0>("FHFMA lo/hi", 352); run<4, 2, 0>("FHFMA lo/hi", 352);
    return 0;
}
