// Microbenchmark 2 (round 2): what limits the LL packet all-to-all among 148 persistent CTAs?
// In lm_kernel one poll round of the x exchange takes ~1150 cycles and the data of the slowest producer is only
// visible ~1000 cycles after a CTA's own store; polling earlier makes the whole frame slower.  Hypothesis: the 148 CTAs
// poll the SAME cache lines, the L2 slice serialises the requests per line, and the producers' stores queue behind
// the pollers' loads.  Variants: R replicas of the packet buffer (CTA b polls replica b % R, producers store to all R),
// first-poll delay D, number of polling threads.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
struct __align__(8) Pkt { float v; unsigned e; };
__device__ __forceinline__ void pk_store(Pkt * p, float v, unsigned ep) {
    const unsigned long long u = (unsigned long long) __float_as_uint(v) | ((unsigned long long) ep << 32);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(u) : "memory");
}
template <int R, int NT>
__global__ void __launch_bounds__(384, 1) k(Pkt * pk, int N, int iters, int delay, int work, long long * cyc, int * rounds_out, float * sink) {
    __shared__ float xs[4096];
    const int t = threadIdx.x, nb = gridDim.x, b = blockIdx.x;
    const int r0 = (int) ((long long) b * N / nb), r1 = (int) ((long long) (b + 1) * N / nb);
    const Pkt * mine = pk + (size_t) (b % R) * 4096;
    long long total = 0; int rounds = 0;
    float acc = (float) b;
    for (int it = 1; it <= iters; ++it) {
        const unsigned ep = (unsigned) it;
        for (int i = 0; i < work + (b * 37 % 64); ++i) acc = __fmaf_rn(acc, 1.0001f, 0.5f);     // dependent work with a little per-CTA skew
        __syncthreads();
        const long long t0 = clock64();
        if (t < r1 - r0) {
#pragma unroll
            for (int r = 0; r < R; ++r) pk_store(pk + (size_t) r * 4096 + r0 + t, acc + t, ep);
        }
        if (delay > 0) { const long long c0 = clock64(); while (clock64() - c0 < delay) { } }
        if (t < NT) {
            for (int i = t; i < N / 2; i += NT) {
                unsigned long long u0, u1;
                while (true) {
                    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(u0), "=l"(u1) : "l"(mine + 2 * i) : "memory");
                    ++rounds;
                    if ((int) ((unsigned) (u0 >> 32) - ep) >= 0 && (int) ((unsigned) (u1 >> 32) - ep) >= 0) break;
                }
                xs[2 * i] = __uint_as_float((unsigned) u0); xs[2 * i + 1] = __uint_as_float((unsigned) u1);
            }
        }
        __syncthreads();
        total += clock64() - t0;
        acc += xs[(t * 7) % N];
    }
    if (t == 0) { cyc[b] = total; }
    atomicAdd(rounds_out, rounds);
    sink[b * blockDim.x + t] = acc;
}
__global__ void warm(float * s, long long cycles) { const long long t0 = clock64(); float a = threadIdx.x; while (clock64() - t0 < cycles) a = a * 1.0001f + 0.5f; s[blockIdx.x * blockDim.x + threadIdx.x] = a; }
template <int R, int NT> void run(int N, int delay, int work) {
    const int nb = 148, iters = 300;
    Pkt * pk; float * sink; long long * cyc; int * rounds;
    cudaMalloc(&pk, (size_t) R * 4096 * 8); cudaMemset(pk, 0, (size_t) R * 4096 * 8);
    cudaMalloc(&cyc, 8 * nb); cudaMalloc(&sink, 4 * nb * 384); cudaMalloc(&rounds, 4); cudaMemset(rounds, 0, 4);
    int n = N, it = iters, dl = delay, wk = work;
    void * args[] = { &pk, &n, &it, &dl, &wk, &cyc, &rounds, &sink };
    cudaLaunchCooperativeKernel((const void *) k<R, NT>, dim3(nb), dim3(384), args, 0, 0);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; int hr = 0; cudaMemcpy(h, cyc, 8 * nb, cudaMemcpyDeviceToHost); cudaMemcpy(&hr, rounds, 4, cudaMemcpyDeviceToHost);
    long long mn = h[0], mx = h[0], sm = 0; for (int i = 0; i < nb; ++i) { if (h[i] < mn) mn = h[i]; if (h[i] > mx) mx = h[i]; sm += h[i]; }
    const double polls = (double) (N / 2) * nb * iters;
    printf("R=%d NT=%3d N=%4d delay=%4d: cycles per exchange (store -> all polled) min %5lld mean %5lld max %5lld, %.2f rounds per pair  (%s)\n",
           R, NT, N, delay, mn / iters, sm / nb / iters, mx / iters, hr / polls, cudaGetErrorString(e));
    cudaFree(pk); cudaFree(cyc); cudaFree(sink); cudaFree(rounds);
}
int main() {
    float * s; cudaMalloc(&s, 4 * 148 * 256);
    warm<<<148, 256>>>(s, 3000000000LL); cudaDeviceSynchronize();          // ~1.5 s: clocks up before anything is timed
    for (int rep = 0; rep < 2; ++rep)
    for (int N : {1024, 3072}) {
        for (int delay : {0, 400, 800, 1200}) {
            run<1, 352>(N, delay, 1500); run<2, 352>(N, delay, 1500); run<4, 352>(N, delay, 1500);
        }
        run<1, 128>(N, 400, 1500); run<4, 128>(N, 400, 1500);
    }
    return 0;
}
