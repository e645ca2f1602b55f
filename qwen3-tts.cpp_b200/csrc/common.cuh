// Device-side helpers shared by the sm_100a kernels of the engine.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace q3t {

// ---- fixed 0.6B shapes the decode kernels are specialised on (src/tts_transformer.h:58-99)
constexpr int H      = 1024;   // hidden_size
constexpr int NH     = 16;     // attention heads
constexpr int NKV    = 8;      // key/value heads
constexpr int HD     = 128;    // head_dim
constexpr int FFN    = 3072;   // intermediate_size
constexpr int QKV_N  = (NH + 2 * NKV) * HD;   // 4096 fused rows: q | k | v
constexpr int NCB    = 16;     // codebooks

// ---- cache-policy loads ------------------------------------------------------------------
// weights: read-only, streamed once per pass -> no L1 allocation
__device__ __forceinline__ uint4 ld_weight(const void * p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// activations written by other CTAs inside the same launch: must not be served by a stale L1 line
__device__ __forceinline__ float ld_act(const float * p) { return __ldcg(p); }
__device__ __forceinline__ float4 ld_act4(const float * p) { return __ldcg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ int ld_act_i(const int * p) { return __ldcg(p); }

// TMA-unit driven L2 prefetch of a contiguous byte range (fire and forget); size multiple of 16
__device__ __forceinline__ void l2_prefetch_bulk(const void * p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}

// ---- acquire / release on global counters ---------------------------------------------------
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long * p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add_u64(unsigned long long * p, unsigned long long v) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// ---- warp reductions ----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// round-to-nearest-even to F16 and back: the oracle's `from_float` before every F16 mat-mul
// (ggml/src/ggml-cpu/ggml-cpu.c:1291-1331)
__device__ __forceinline__ float r16(float v) { return __half2float(__float2half_rn(v)); }

// 8 halves (one 16-byte vector) dotted with 8 halves, F32 accumulate, products exact in F32
__device__ __forceinline__ float dot8_hh(const uint4 & w, const uint4 & x, float acc) {
    const __half2 * wh = reinterpret_cast<const __half2 *>(&w);
    const __half2 * xh = reinterpret_cast<const __half2 *>(&x);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 a = __half22float2(wh[i]);
        float2 b = __half22float2(xh[i]);
        acc = fmaf(a.x, b.x, acc);
        acc = fmaf(a.y, b.y, acc);
    }
    return acc;
}
// 8 halves (weights, exact in F32) dotted with 8 floats
__device__ __forceinline__ float dot8_hf(const uint4 & w, const float4 & x0, const float4 & x1, float acc) {
    const __half2 * wh = reinterpret_cast<const __half2 *>(&w);
    float2 a0 = __half22float2(wh[0]), a1 = __half22float2(wh[1]);
    float2 a2 = __half22float2(wh[2]), a3 = __half22float2(wh[3]);
    acc = fmaf(a0.x, x0.x, acc); acc = fmaf(a0.y, x0.y, acc);
    acc = fmaf(a1.x, x0.z, acc); acc = fmaf(a1.y, x0.w, acc);
    acc = fmaf(a2.x, x1.x, acc); acc = fmaf(a2.y, x1.y, acc);
    acc = fmaf(a3.x, x1.z, acc); acc = fmaf(a3.y, x1.w, acc);
    return acc;
}

// ------------------------------------------------------------------------------- exact math
// ggml_v_expf, AVX2+FMA variant, one lane (ggml/src/ggml-cpu/vec.h:1230-1267)
__device__ __forceinline__ float v_expf(float x) {
    const float r = 0x1.8p23f;
    const float z = __fmaf_rn(x, 0x1.715476p+0f, r);
    const float n = __fsub_rn(z, r);
    const float b = __fmaf_rn(-n, 0x1.7f7d1cp-20f, __fmaf_rn(-n, 0x1.62e4p-1f, x));
    const unsigned e = __float_as_uint(z) << 23;
    const float k = __uint_as_float(e + 0x3f800000u);
    const float u = __fmul_rn(b, b);
    const float j = __fmaf_rn(__fmaf_rn(__fmaf_rn(0x1.0e4020p-7f, b, 0x1.573e2ep-5f), u,
                                        __fmaf_rn(0x1.555e66p-3f, b, 0x1.fffdb6p-2f)),
                              u, __fmul_rn(0x1.ffffecp-1f, b));
    if (!(fabsf(n) > 126.0f)) return __fmaf_rn(j, k, k);
    const unsigned g = (n <= 0.0f) ? 0x82000000u : 0u;
    const float s1 = __uint_as_float(g + 0x7f000000u);
    const float s2 = __uint_as_float(e - g);
    if (fabsf(n) > 192.0f) return __fmul_rn(s1, s1);
    return __fmul_rn(__fmaf_rn(s2, j, s2), s1);
}
// scalar tail of the SIMD loops calls libm expf (vec.cpp:611-615).  This is glibc 2.39's
// expf (sysdeps/ieee754/flt-32/e_expf.c, the FMA ifunc variant x86-64 hosts select): exp(x) =
// 2^(k/32) * 2^(r/32) with a 32-entry table and a cubic in double precision, then one rounding to
// float.  The table is 2^(i/32) correctly rounded minus (i << 47); checked exhaustively on the
// host against libm over all 2.2e9 floats in [-87, 88]: 2 mismatches (1e-9 per call), where
// round-to-float of a double exp() disagrees with libm 7.7e-5 of the time.
static __device__ __constant__ unsigned long long k_exp2f_tab[32] = {
    0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,
    0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,
    0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,
    0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,
    0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,
    0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,
    0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,
    0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL,
};
__device__ __forceinline__ float tail_expf(float x) {
    if (!(fabsf(x) < 88.0f)) {                              // e_expf.c special cases
        if (x != x) return x + x;
        if (x > 0x1.62e42ep6f) return __int_as_float(0x7f800000);
        if (x < -0x1.9fe368p6f) return 0.0f;
    }
    const double xd = (double) x;
    const double z  = __dmul_rn(0x1.71547652b82fep+5, xd);         // x * 32/ln2
    double kd = __dadd_rn(z, 0x1.8p+52);
    const unsigned long long ki = (unsigned long long) __double_as_longlong(kd);
    kd = __dsub_rn(kd, 0x1.8p+52);
    const double r = __dsub_rn(z, kd);
    const unsigned long long t = k_exp2f_tab[ki & 31] + (ki << 47);
    const double s  = __longlong_as_double((long long) t);
    const double zz = __fma_rn(0x1.c6af84b912394p-5 / 32768.0, r, 0x1.ebfce50fac4f3p-3 / 1024.0);
    const double r2 = __dmul_rn(r, r);
    double y = __fma_rn(0x1.62e42ff0c52d6p-1 / 32.0, r, 1.0);
    y = __fma_rn(zz, r2, y);
    return (float) __dmul_rn(y, s);
}

// the oracle's horizontal reduction of the 32 accumulator lanes, lane = k%32
// ((s0+s2)+(s1+s3) per AVX lane -> lane l + lane l+4 -> hadd, hadd)
__device__ __forceinline__ float reduce32_lanes(float v) {
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 16));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 8));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 4));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 2));
    return v;
}


} // namespace q3t
