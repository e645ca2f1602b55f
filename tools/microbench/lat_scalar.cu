// Dependent-operation latencies of the scalar sections of lm_kernel (one warp, clock64 around 64-deep chains), sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bin/lat_scalar lat_scalar.cu && bin/lat_scalar
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CHAIN(NAME, INIT, STEP)                                                     \
    __global__ void NAME(long long * out, double seedd, float seedf) {             \
        INIT;                                                                       \
        PIN(a);                                                                     \
        long long t0; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t0) :: "memory");  \
        PIN(a);                                                                     \
        _Pragma("unroll") for (int i = 0; i < 64; ++i) { STEP; PIN(a); }            \
        long long t1; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1) :: "memory");  \
        if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = (long long) (SINK); }    \
    }

#define SINK (a * 1e3)
#define PIN(x) asm volatile("" : "+d"(x) :: "memory")
CHAIN(k_dadd, double a = seedd; double b = seedd * 1e-3, a = __dadd_rn(a, b))
CHAIN(k_dfma, double a = seedd; double b = 1.0000001, a = __fma_rn(a, b, 1e-9))
CHAIN(k_drcp, double a = seedd + 2.0, a = __drcp_rn(a) + 1.5)
CHAIN(k_ddiv, double a = seedd + 2.0, a = 1.0 / a + 1.5)
CHAIN(k_f2d_dadd, double a = seedd; float f = seedf, { a = __dadd_rn(a, (double) f); f = (float) a; })
CHAIN(k_shfl_d, double a = seedd, a = __dadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1)))
#undef SINK
#undef PIN
#define SINK (a * 1e3f)
#define PIN(x) asm volatile("" : "+f"(x) :: "memory")
CHAIN(k_fadd, float a = seedf; float b = seedf * 1e-3f, a = __fadd_rn(a, b))
CHAIN(k_ffma, float a = seedf; float b = 1.0000001f, a = __fmaf_rn(a, b, 1e-9f))
CHAIN(k_fdiv, float a = seedf + 2.0f, a = __fdiv_rn(1.0f, a) + 1.5f)
CHAIN(k_fsqrt, float a = seedf + 2.0f, a = __fsqrt_rn(a) + 1.5f)
CHAIN(k_shfl_f, float a = seedf, a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1)))
CHAIN(k_f2h, float a = seedf, a = __half2float(__float2half_rn(a)) + 1e-3f)
CHAIN(k_expf, float a = seedf * 1e-3f, a = expf(a) - 1.0f)

int main() {
    long long * d; cudaMalloc(&d, 16);
    struct { const char * n; void (*k)(long long *, double, float); } ks[] = {
        {"DADD", k_dadd}, {"DFMA", k_dfma}, {"__drcp_rn + DADD", k_drcp}, {"1.0 / x (double) + DADD", k_ddiv},
        {"F2F.F64.F32 + DADD + F2F.F32.F64", k_f2d_dadd}, {"SHFL (2 x 32 bit) + DADD", k_shfl_d},
        {"FADD", k_fadd}, {"FFMA", k_ffma}, {"__fdiv_rn(1, x) + FADD", k_fdiv}, {"__fsqrt_rn + FADD", k_fsqrt},
        {"SHFL + FADD", k_shfl_f}, {"F2F.F16.F32 + back + FADD", k_f2h}, {"expf + FADD", k_expf}};
    for (auto & e : ks) {
        for (int warps = 1; warps <= 8; warps *= 8) {
            e.k<<<1, 32 * warps>>>(d, 1.25, 1.25f); e.k<<<1, 32 * warps>>>(d, 1.25, 1.25f);
            long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            printf("%-36s %d warp(s): %6.1f cycles per dependent step\n", e.n, warps, h[0] / 64.0);
        }
    }
    return 0;
}
