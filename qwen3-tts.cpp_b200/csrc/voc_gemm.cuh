// "Tap GEMM" used by every dense contraction of the vocoder (conv1d, ConvTranspose1d, linear):
//
//   acc[m][n] = sum_{tap < ntaps} sum_{c < Kc}  A[m + tap_off[tap]][c] * W[n][tap*Kc + c]
//
// A is a channels-last F16 activation [a_rows][lda] (rows outside [0, a_rows) read as zero = the
// causal left padding of the reference, src/audio_tokenizer_decoder.cpp:511,562,706,759), W is
// F16 [N][ntaps*Kc], accumulation is F32 (products of two F16 values are exact in F32, like
// ggml_vec_dot_f16).  Fused epilogue, in the reference's operation order:
//
//   v = acc (+ bias[n % pmod]) (* gamma[n % pmod]) (+ res[m][n])     -> out32 (optional)
//   v = act(v): none | snake(ea, eb)[n % pmod] | gelu-through-F16    -> out16 = F16(v) (optional)
//   swiglu: columns (2i, 2i+1) = (gate_i, up_i) -> out16[m][i] = F16(silu(gate) * up)
#pragma once

#include "common.cuh"
#include "engine.h"

namespace q3t {

enum { VACT_NONE = 0, VACT_SNAKE = 1, VACT_GELU = 2 };

struct VGemm {
    const __half * A; int lda; int a_rows;
    int M, N;
    int ntaps; int Kc; int tap_off[8];
    const __half * W;
    const float * bias; const float * gamma; int pmod;
    const float * res; int ldres;
    float * out32; int ld32;
    __half * out16; int ld16;
    int act; const float * ea; const float * eb;
    int swiglu;
};

// snake(x) = x + exp(-beta) * sin^2(exp(alpha) * x)   (src/audio_tokenizer_decoder.cpp:375-402)
__device__ __forceinline__ float snake_f(float x, float ea, float eb) {
    const float s = sinf(x * ea);
    return x + (s * s) * eb;
}
// ggml_vec_gelu_f32 with GGML_GELU_FP16 (ggml/src/ggml-cpu/vec.h:996-1009): table lookup on F16(x),
// table entry = F16(gelu_f32(F32(F16(x))))
__device__ __forceinline__ float gelu_f16_table(float x) {
    if (x <= -10.0f) return 0.0f;
    if (x >= 10.0f) return x;
    const float xr = r16(x);
    const float g = 0.5f * xr * (1.0f + tanhf(0.79788456080286535587989211986876f * xr * (1.0f + 0.044715f * xr * xr)));
    return r16(g);
}
__device__ __forceinline__ float silu_f(float g) { return g / (1.0f + expf(-g)); }

__device__ __forceinline__ void vgemm_epilogue_pair(const VGemm & p, size_t m, int n, float v0, float v1, bool ok1) {
    // two adjacent columns n, n+1 of row m (n even)
    if (p.swiglu) {
        p.out16[m * p.ld16 + (n >> 1)] = __float2half_rn(silu_f(v0) * v1);
        return;
    }
    float v[2] = {v0, v1};
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        if (e == 1 && !ok1) break;
        const int nn = n + e, pi = nn % p.pmod;
        float x = v[e];
        if (p.bias)  x += p.bias[pi];
        if (p.gamma) x *= p.gamma[pi];
        if (p.res)   x += p.res[m * p.ldres + nn];
        if (p.out32) p.out32[m * p.ld32 + nn] = x;
        if (p.out16) {
            if (p.act == VACT_SNAKE) x = snake_f(x, p.ea[pi], p.eb[pi]);
            else if (p.act == VACT_GELU) x = gelu_f16_table(x);
            p.out16[m * p.ld16 + nn] = __float2half_rn(x);
        }
    }
}

// ---- v1: CUDA-core tiled kernel (64 x 64 tile, BK 32, 256 threads, 4 x 4 per thread)
constexpr int VG_BM = 64, VG_BN = 64, VG_BK = 32;

__global__ void __launch_bounds__(256)
vgemm_simt(const VGemm p) {
    __shared__ float As[VG_BK][VG_BM + 4];
    __shared__ float Ws[VG_BK][VG_BN + 4];
    const int tid = threadIdx.x;
    const size_t m0 = (size_t) blockIdx.x * VG_BM;
    const int n0 = blockIdx.y * VG_BN;
    const int tx = tid & 15, ty = tid >> 4;          // tx -> n (4 cols: tx*4..), ty -> m (4 rows)
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const int Ktot = p.ntaps * p.Kc;
    // loader mapping: 64 rows x 32 cols of halves = 64 x 4 uint4; thread -> (row = tid/4, seg = tid%4)
    const int lr = tid >> 2, ls = tid & 3;
    for (int tap = 0; tap < p.ntaps; ++tap) {
        const long arow = (long) m0 + lr + p.tap_off[tap];
        const bool a_ok = arow >= 0 && arow < p.a_rows && (m0 + lr) < (size_t) p.M;
        const __half * ap = p.A + (size_t) (a_ok ? arow : 0) * p.lda;
        const bool w_ok = n0 + lr < p.N;
        const __half * wp = p.W + (size_t) (w_ok ? n0 + lr : 0) * Ktot + (size_t) tap * p.Kc;
        for (int k0 = 0; k0 < p.Kc; k0 += VG_BK) {
            uint4 av = make_uint4(0, 0, 0, 0), wv = make_uint4(0, 0, 0, 0);
            if (a_ok) av = *reinterpret_cast<const uint4 *>(ap + k0 + ls * 8);
            if (w_ok) wv = *reinterpret_cast<const uint4 *>(wp + k0 + ls * 8);
            const __half * ah = reinterpret_cast<const __half *>(&av);
            const __half * wh = reinterpret_cast<const __half *>(&wv);
            __syncthreads();
#pragma unroll
            for (int e = 0; e < 8; ++e) { As[ls * 8 + e][lr] = __half2float(ah[e]); Ws[ls * 8 + e][lr] = __half2float(wh[e]); }
            __syncthreads();
#pragma unroll 8
            for (int k = 0; k < VG_BK; ++k) {
                const float4 a4 = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
                const float4 w4 = *reinterpret_cast<const float4 *>(&Ws[k][tx * 4]);
                const float a[4] = {a4.x, a4.y, a4.z, a4.w}, w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const size_t m = m0 + ty * 4 + i;
        if (m >= (size_t) p.M) continue;
#pragma unroll
        for (int j = 0; j < 4; j += 2) {
            const int n = n0 + tx * 4 + j;
            if (n >= p.N) continue;
            vgemm_epilogue_pair(p, m, n, acc[i][j], acc[i][j + 1], n + 1 < p.N);
        }
    }
}

inline void voc_gemm(q3t_ctx * ctx, const VGemm & p) {
    dim3 grid((unsigned) ((p.M + VG_BM - 1) / VG_BM), (unsigned) ((p.N + VG_BN - 1) / VG_BN));
    vgemm_simt<<<grid, 256, 0, ctx->stream>>>(p);
}

} // namespace q3t
