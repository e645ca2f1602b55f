// Microbenchmark: latency / throughput of FHFMA (fma.rn.f32.f16) vs FFMA on sm_100a.
#include <cstdio>
#include <cuda_fp16.h>
__device__ __forceinline__ float fh(unsigned w, unsigned x, float acc) {
    asm volatile("{.reg .b16 wl, wh, xl, xh; mov.b32 {wl, wh}, %1; mov.b32 {xl, xh}, %2; fma.rn.f32.f16 %0, wl, xl, %0;}" : "+f"(acc) : "r"(w), "r"(x));
    return acc;
}
template <int ILP, int MODE>
__global__ void k(unsigned w, unsigned x, float * out, long long * cyc) {
    float acc[ILP];
    for (int i = 0; i < ILP; ++i) acc[i] = threadIdx.x * 1e-9f + i;
    const float wf = __uint_as_float(w | 0x3f000000u), xf = __uint_as_float(x | 0x3f000000u);
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 64; ++it) {
#pragma unroll
        for (int r = 0; r < 16; ++r)
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                if (MODE == 0) acc[i] = fh(w, x, acc[i]);
                else asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(wf), "f"(xf));
            }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < ILP; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int ILP, int MODE> void run(const char * name, int threads) {
    float * out; long long * cyc; cudaMalloc(&out, 4 * 1024); cudaMalloc(&cyc, 8);
    k<ILP, MODE><<<1, threads>>>(0x3c003c00u, 0x3c003c00u, out, cyc);
    k<ILP, MODE><<<1, threads>>>(0x3c003c00u, 0x3c003c00u, out, cyc);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double n = 64.0 * 16 * ILP;
    printf("%-6s ILP=%d threads=%4d: %8lld cycles, %.2f cycles per instr per warp, %.1f lane-ops/clk/SM\n", name, ILP, threads, c, c / n, n * threads / c);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<1, 0>("FHFMA", 32);  run<1, 1>("FFMA", 32);
    run<2, 0>("FHFMA", 32);  run<4, 0>("FHFMA", 32); run<8, 0>("FHFMA", 32);
    run<4, 1>("FFMA", 32);   run<8, 1>("FFMA", 32);
    run<4, 0>("FHFMA", 128); run<4, 0>("FHFMA", 512); run<8, 0>("FHFMA", 512);
    run<4, 1>("FFMA", 128);  run<4, 1>("FFMA", 512);  run<8, 1>("FFMA", 512);
    return 0;
}
