// Microbenchmark: latency of an all-to-all exchange of N floats among 148 persistent CTAs through L2.
// Variants: 0 = LL packets {value, epoch} polled by all threads (16-byte loads)
//           1 = per-producer-CTA flag (st.release) polled by one warp, then plain data loads
//           2 = atomic counter barrier (red.release + ld.acquire spin by thread 0) + plain data loads
//           3 = LL packets, but only ONE warp per CTA polls (loop), others wait at a CTA barrier
//           4 = 32-byte sector packets {epoch, v0, v1, v2 | v3, v4, v5, epoch}: one 256-bit store / load per 6 values,
//               each 16-byte half carries its own epoch (valid under 16-byte access atomicity, like variant 0)
//           5 = 32-byte sector packets {v0 .. v6, epoch}: 7 values per 256-bit access (needs 32-byte atomicity)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
struct __align__(8) Pkt { float v; unsigned e; };
__device__ __forceinline__ void pk_store(Pkt * p, float v, unsigned ep) {
    const unsigned long long u = (unsigned long long) __float_as_uint(v) | ((unsigned long long) ep << 32);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(u) : "memory");
}
__device__ __forceinline__ float2 pk_poll2(const Pkt * p, unsigned ep) {
    unsigned long long u0, u1;
    do { asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(u0), "=l"(u1) : "l"(p) : "memory"); }
    while ((int) ((unsigned) (u0 >> 32) - ep) < 0 || (int) ((unsigned) (u1 >> 32) - ep) < 0);   // monotone: a fast CTA may already be one round ahead
    return make_float2(__uint_as_float((unsigned) u0), __uint_as_float((unsigned) u1));
}
__device__ __forceinline__ unsigned ld_acq(const unsigned * p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_rlx(const unsigned * p) { unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_rel(unsigned * p, unsigned v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }

struct __align__(32) Pk32 { unsigned w[8]; };
__device__ __forceinline__ void st32(Pk32 * p, const unsigned (&w)[8]) {
    asm volatile("st.relaxed.gpu.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]),
                 "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
}
__device__ __forceinline__ void ld32(const Pk32 * p, unsigned (&w)[8]) {
    asm volatile("ld.relaxed.gpu.global.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]),
                 "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p) : "memory");
}

template <int VAR, int STRIDE>
__global__ void __launch_bounds__(512, 1) k(Pkt * pk, float * data, unsigned * flags, unsigned * counter, int N, int iters, long long * cyc, float * sink, int work) {
    __shared__ float xs[4096];
    const int t = threadIdx.x, nb = gridDim.x, b = blockIdx.x;
    const int r0 = (int) ((long long) b * N / nb), r1 = (int) ((long long) (b + 1) * N / nb);
    long long total = 0;
    float acc = 0.f;
    for (int it = 1; it <= iters; ++it) {
        const unsigned ep = (unsigned) it;
        // "compute": a little dependent work so CTAs arrive roughly together
        for (int i = 0; i < work; ++i) acc = __fmaf_rn(acc, 1.0001f, 0.5f);
        __syncthreads();
        const long long t0 = clock64();
        if (VAR == 0 || VAR == 3) {
            // STRIDE > 16: the 16-byte pairs are spread out so that consecutive pairs fall into different L2 slices
            constexpr int S8 = STRIDE / 8;      // stride of a pair in Pkt units
            if (t < r1 - r0) pk_store(pk + (size_t) ((r0 + t) >> 1) * S8 + ((r0 + t) & 1), acc + t, ep);
            if (VAR == 0) {
                for (int i = t; i < N / 2; i += blockDim.x) { const float2 f = pk_poll2(pk + (size_t) i * S8, ep); xs[2 * i] = f.x; xs[2 * i + 1] = f.y; }
            } else {
                if (t < 32) for (int i = t; i < N / 2; i += 32) { const float2 f = pk_poll2(pk + (size_t) i * S8, ep); xs[2 * i] = f.x; xs[2 * i + 1] = f.y; }
            }
        } else if (VAR == 4 || VAR == 5) {
            // CTA b owns packets [b * PPC, (b + 1) * PPC): PPC packets of VPP values cover its <= ceil(N / nb) rows
            constexpr int VPP = VAR == 4 ? 6 : 7;
            Pk32 * pk32 = reinterpret_cast<Pk32 *>(pk);
            const int rows_max = (N + nb - 1) / nb, PPC = (rows_max + VPP - 1) / VPP;
            if (t < PPC) {
                unsigned w[8];
                if (VAR == 4) { w[0] = ep; w[7] = ep; for (int j = 0; j < 6; ++j) w[1 + j] = __float_as_uint(acc + t * 6 + j); }
                else          { w[7] = ep; for (int j = 0; j < 7; ++j) w[j] = __float_as_uint(acc + t * 7 + j); }
                st32(pk32 + (size_t) (b * PPC + t) * (STRIDE / 32 > 0 ? STRIDE / 32 : 1), w);
            }
            for (int i = t; i < nb * PPC; i += blockDim.x) {
                unsigned w[8];
                if (VAR == 4) { do { ld32(pk32 + (size_t) i * (STRIDE / 32 > 0 ? STRIDE / 32 : 1), w); } while ((int) (w[0] - ep) < 0 || (int) (w[7] - ep) < 0); }
                else          { do { ld32(pk32 + (size_t) i * (STRIDE / 32 > 0 ? STRIDE / 32 : 1), w); } while ((int) (w[7] - ep) < 0); }
                const int cta = i / PPC, j0 = (i % PPC) * VPP;
                const int base = (int) ((long long) cta * N / nb);
                for (int j = 0; j < VPP; ++j) if (base + j0 + j < N) xs[base + j0 + j] = __uint_as_float(w[VAR == 4 ? 1 + j : j]);
            }
        } else if (VAR == 1) {
            if (t < 32) {
                for (int i = t; i < r1 - r0; i += 32) data[r0 + i] = acc + i;
                __syncwarp();
                if (t == 0) st_rel(flags + b, ep);          // release: orders this thread's prior stores only -> fence for the warp
            }
            if (t < 32) {
                for (int i = t; i < nb; i += 32) while ((int) (ld_acq(flags + i) - ep) < 0) { }
            }
            __syncthreads();
            for (int i = t; i < N / 4; i += blockDim.x) { const float4 f = __ldcg(reinterpret_cast<const float4 *>(data) + i); xs[4 * i] = f.x; xs[4 * i + 1] = f.y; xs[4 * i + 2] = f.z; xs[4 * i + 3] = f.w; }
        } else {
            if (t < r1 - r0) data[r0 + t] = acc + t;
            __syncthreads();
            if (t == 0) {
                __threadfence();
                asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(counter) : "memory");
                while (ld_acq(counter) < (unsigned) (it * nb)) { }
            }
            __syncthreads();
            for (int i = t; i < N / 4; i += blockDim.x) { const float4 f = __ldcg(reinterpret_cast<const float4 *>(data) + i); xs[4 * i] = f.x; xs[4 * i + 1] = f.y; xs[4 * i + 2] = f.z; xs[4 * i + 3] = f.w; }
        }
        __syncthreads();
        const long long t1 = clock64();
        total += t1 - t0;
        acc += xs[(t * 7) % N];
    }
    if (t == 0) cyc[b] = total;
    sink[b * blockDim.x + t] = acc;
}
template <int VAR, int STRIDE = 16> void run(int N, int work) {
    const int nb = 148, iters = 200;
    Pkt * pk; float * data, * sink; unsigned * flags, * counter; long long * cyc;
    const size_t pkb = (size_t) 4096 * 1024 + 65536; cudaMalloc(&pk, pkb); cudaMemset(pk, 0, pkb); cudaMalloc(&data, 4 * 4096); cudaMalloc(&flags, 4 * 512); cudaMemset(flags, 0, 2048);
    cudaMalloc(&counter, 4); cudaMemset(counter, 0, 4); cudaMalloc(&cyc, 8 * nb); cudaMalloc(&sink, 4 * nb * 512);
    int n = N, it = iters, wk = work;
    void * args[] = { &pk, &data, &flags, &counter, &n, &it, &cyc, &sink, &wk };
    cudaLaunchCooperativeKernel((const void *) k<VAR, STRIDE>, dim3(nb), dim3(512), args, 0, 0);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, 8 * nb, cudaMemcpyDeviceToHost);
    long long mn = h[0], mx = h[0]; for (int i = 1; i < nb; ++i) { if (h[i] < mn) mn = h[i]; if (h[i] > mx) mx = h[i]; }
    printf("variant %d stride %4d N=%4d work=%5d: cycles per exchange min %lld max %lld  (%s)\n", VAR, STRIDE, N, work, mn / iters, mx / iters, cudaGetErrorString(e));
    cudaFree(pk); cudaFree(data); cudaFree(flags); cudaFree(counter); cudaFree(cyc); cudaFree(sink);
}
__global__ void __launch_bounds__(384, 1) kclu(int * out) { if (threadIdx.x == 0) out[blockIdx.x] = 1; }
static void cluster_report() {
    // how many CTAs of 384 threads / 200 KB shared memory can be resident at once for each cluster size
    cudaFuncSetAttribute(kclu, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(kclu, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cs * 64); cfg.blockDim = dim3(384); cfg.dynamicSmemBytes = 200 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kclu, &cfg);
        printf("cluster size %2d: max active clusters %d -> %d CTAs (%s)\n", cs, n, n * cs, cudaGetErrorString(e));
    }
}
int main() {
    cluster_report();
    for (int N : {1024, 2048, 3072}) for (int work : {0, 2000}) {
        run<0>(N, work); run<0, 256>(N, work); run<0, 512>(N, work); run<0, 1024>(N, work);
        run<4, 32>(N, work); run<4, 256>(N, work); run<4, 1024>(N, work);
        run<5, 32>(N, work); run<5, 256>(N, work); run<5, 1024>(N, work);
        if (N == 1024) { run<1>(N, work); run<2>(N, work); }
    }
    return 0;
}
