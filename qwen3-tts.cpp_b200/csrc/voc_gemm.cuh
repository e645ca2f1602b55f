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
//   act relu / relu+tanh (speaker encoder) are applied BEFORE out32: there the F32 stream is the activated one
//   swiglu: columns (2i, 2i+1) = (gate_i, up_i) -> out16[m][i] = F16(silu(gate) * up)
#pragma once

#include <algorithm>
#include <cuda.h>           // CUtensorMap (types only: the encoder is fetched with cudaGetDriverEntryPoint, no -lcuda)
#include "common.cuh"
#include "engine.h"

namespace q3t {

enum { VACT_NONE = 0, VACT_SNAKE = 1, VACT_GELU = 2, VACT_RELU = 3, VACT_RELU_TANH = 4 };   // 3, 4: speaker encoder (spk_encoder.inc)

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
// tensor-core epilogue variant: two-term Cody-Waite reduction to [-pi, pi] and the SFU sine (abs. error
// < 5e-7 there); the result is rounded to F16 (2^-11 relative) right after, and the waveform bar is an SNR.
__device__ __forceinline__ float snake_fast(float x, float ea, float eb) {
    const float y = x * ea;
    const float k = rintf(y * 0.15915494309189535f);
    float r = fmaf(k, -6.28318548202514648f, y);
    r = fmaf(k, 1.7484555e-7f, r);
    const float sn = __sinf(r);
    return x + (sn * sn) * eb;
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
        const float sv = silu_f(v0) * v1;
        p.out16[m * p.ld16 + (n >> 1)] = __float2half_rn(sv);
        if (p.out32) p.out32[m * p.ld32 + (n >> 1)] = sv;          // un-rounded copy (Q8_0 files: the next mat-mul quantises it)
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
        if (p.act >= VACT_RELU) { x = fmaxf(x, 0.0f); if (p.act == VACT_RELU_TANH) x = tanhf(x); }
        if (p.out32) p.out32[m * p.ld32 + nn] = x;
        if (p.out16) {
            if (p.act == VACT_SNAKE) x = snake_f(x, p.ea[pi], p.eb[pi]);
            else if (p.act == VACT_GELU) x = gelu_f16_table(x);
            p.out16[m * p.ld16 + nn] = __float2half_rn(x);
        }
    }
}

// the same epilogue with the per-column parameters of the pair (n, n+1) fetched once (a thread of the
// tensor-core kernel keeps its column pair for all rows of the tile)
struct VEpiCols { float bias[2], gamma[2], ea[2], eb[2]; };
__device__ __forceinline__ void vgemm_epi_cols(const VGemm & p, int n, bool ok1, VEpiCols & c) {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        const int pi = (n + ((e == 1 && ok1) ? 1 : 0)) % p.pmod;
        c.bias[e]  = p.bias ? p.bias[pi] : 0.0f;
        c.gamma[e] = p.gamma ? p.gamma[pi] : 1.0f;
        c.ea[e] = (p.act == VACT_SNAKE && p.out16) ? p.ea[pi] : 0.0f;
        c.eb[e] = (p.act == VACT_SNAKE && p.out16) ? p.eb[pi] : 0.0f;
    }
}
// residual of the pair (issued early so that several rows' loads are in flight)
__device__ __forceinline__ float2 vgemm_load_res(const VGemm & p, size_t m, int n, bool ok1) {
    float2 r = make_float2(0.0f, 0.0f);
    if (p.res) {
        if (ok1 && ((p.ldres & 1) == 0)) r = *reinterpret_cast<const float2 *>(p.res + m * p.ldres + n);
        else { r.x = p.res[m * p.ldres + n]; if (ok1) r.y = p.res[m * p.ldres + n + 1]; }
    }
    return r;
}
// EPI selects a compile-time specialisation of the fused epilogue (the generic one tests ~10 run-time flags
// per row, and the hot launches are instruction-issue bound):
//   VEPI_GENERIC     every combination, flags read from VGemm
//   VEPI_BIAS_SNAKE  out16 = F16(snake(acc + bias))                              (k7 conv of a residual unit)
//   VEPI_RES_SNAKE   x = acc + bias + res; [out32 = x;] out16 = F16(snake(x))    (1x1 conv of a residual unit)
//   VEPI_O32_SNAKE   x = acc + bias; out32 = x; out16 = F16(snake(x))            (ConvTranspose)
// the specialised kinds require even N / leading dimensions (checked by voc_gemm).
enum { VEPI_GENERIC = 0, VEPI_BIAS_SNAKE = 1, VEPI_RES_SNAKE = 2, VEPI_O32_SNAKE = 3 };

template <int EPI>
__device__ __forceinline__ void vgemm_epilogue_cols(const VGemm & p, size_t m, int n, float v0, float v1, bool ok1, const VEpiCols & c, float2 rr) {
    constexpr bool G = EPI == VEPI_GENERIC;
    if (G && p.swiglu) {
        const float sv = silu_f(v0) * v1;
        p.out16[m * p.ld16 + (n >> 1)] = __float2half_rn(sv);
        if (p.out32) p.out32[m * p.ld32 + (n >> 1)] = sv;          // un-rounded copy (Q8_0 files: the next mat-mul quantises it)
        return;
    }
    const bool has_bias = G ? p.bias != nullptr : true;
    const bool has_gamma = G ? p.gamma != nullptr : false;
    const bool has_res = G ? p.res != nullptr : EPI == VEPI_RES_SNAKE;
    const bool has_o32 = G ? p.out32 != nullptr : (EPI == VEPI_BIAS_SNAKE ? false : EPI == VEPI_O32_SNAKE ? true : p.out32 != nullptr);
    const bool has_o16 = G ? p.out16 != nullptr : true;
    const int  act = G ? p.act : VACT_SNAKE;
    float x[2] = {v0, v1};
    const float r[2] = {rr.x, rr.y};
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        if (has_bias)  x[e] += c.bias[e];
        if (has_gamma) x[e] *= c.gamma[e];
        if (has_res)   x[e] += r[e];
        if (G && act >= VACT_RELU) { x[e] = fmaxf(x[e], 0.0f); if (act == VACT_RELU_TANH) x[e] = tanhf(x[e]); }   // ReLU outputs are the F32 stream too
    }
    if (has_o32) {
        if (!G || (ok1 && ((p.ld32 & 1) == 0))) *reinterpret_cast<float2 *>(p.out32 + m * p.ld32 + n) = make_float2(x[0], x[1]);
        else { p.out32[m * p.ld32 + n] = x[0]; if (ok1) p.out32[m * p.ld32 + n + 1] = x[1]; }
    }
    if (has_o16) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            if (act == VACT_SNAKE) x[e] = snake_fast(x[e], c.ea[e], c.eb[e]);
            else if (act == VACT_GELU) x[e] = gelu_f16_table(x[e]);
        }
        if (!G || (ok1 && ((p.ld16 & 1) == 0))) *reinterpret_cast<__half2 *>(p.out16 + m * p.ld16 + n) = __floats2half2_rn(x[0], x[1]);
        else { p.out16[m * p.ld16 + n] = __float2half_rn(x[0]); if (ok1) p.out16[m * p.ld16 + n + 1] = __float2half_rn(x[1]); }
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

// ---- v2: 5th-generation tensor cores.  128 x BN output tile per CTA, accumulator in TMEM (128 lanes x BN
// columns of F32), operands staged in shared memory by cp.async in the no-swizzle K-major UMMA layout
//   offset(row, k16B) = (row / 8) * SBO + k16B * LBO + (row % 8) * 16      (core matrix = 8 rows x 16 bytes)
// and consumed by tcgen05.mma.kind::f16 (F16 x F16 -> F32) issued by one thread.  The A tile of tap t is the
// activation matrix read at row offset tap_off[t]; rows outside [0, a_rows) are zero-filled by cp.async
// (= the causal left padding of the reference), so no im2col buffer exists.
// Warp-specialised 4-stage ring with no CTA barrier in the K loop: warps 0-7 are loaders (cp.async, three steps
// in flight; when a step of the warp has landed: cp.async.wait_group, fence.proxy.async, one arrive per warp on
// the stage's `full` mbarrier), lane 0 of warp 8 waits on `full`, issues the step's MMAs and hands the stage back
// through tcgen05.commit on its `empty` mbarrier, which the loaders wait on before they overwrite it.
// Epilogue: tcgen05.ld (each warp owns its 32 TMEM lanes = 32 output rows), fused bias / layer-scale /
// residual / snake / gelu / swiglu as in vgemm_epilogue_pair.
constexpr int VU_BM = 128, VU_BK = 32, VU_STAGES = 4, VU_THREADS = 256;      // VU_THREADS: loader / epilogue threads
constexpr int VU_LAUNCH = VU_THREADS + 32;                                   // + the MMA warp
constexpr int VU_AHEAD = VU_STAGES - 1;                                      // steps a loader keeps in flight
constexpr int VU_LBO = 128, VU_SBO = (VU_BK / 8) * 128;            // bytes: 4 k-chunks of 16 B per 8-row group
constexpr int VU_A_STAGE = VU_BM * VU_BK * 2;                       // 8 KB: A tile of one tap
// A operand ring.  Its layout keeps the rows of a k-chunk contiguous: offset(row, k16B) = k16B * LBO_A + row * 16
// (SBO = 128, LBO_A = rows * 16), so the tile "shifted down by d rows" is the same bytes with the descriptor start
// address advanced by 16 d.  A k-block of a multi-tap conv is therefore loaded ONCE for rows
// [m0 + min(tap_off), m0 + 127 + max(tap_off)] and every tap's MMA reads it at its own row shift (7x less A
// traffic for the k7 convs, whose L2 -> SM operand stream is what bounds the 128 x 128 tile).  Taps too far apart
// (span > VU_SPAN_MAX) and 1-tap GEMMs load a 128-row tile per step as before.
// MT = 2: the CTA owns two M tiles (256 rows, two accumulators in TMEM): every B stage feeds two MMAs, which
// halves the W traffic per flop; used when the grid still fills the GPU.
constexpr int VU_SPAN_MAX = 64;
__host__ __device__ constexpr int vu_ra_shared(int mt) { return VU_BM * mt + VU_SPAN_MAX; }            // 192 / 320 rows
__host__ __device__ constexpr int vu_a_ring(int mt) {                                                  // 36 KB / 64 KB
    return 3 * vu_ra_shared(mt) * VU_BK * 2 > VU_STAGES * VU_BM * mt * VU_BK * 2 ? 3 * vu_ra_shared(mt) * VU_BK * 2
                                                                                  : VU_STAGES * VU_BM * mt * VU_BK * 2;
}
__host__ __device__ constexpr int vu_b_stage(int bn) { return bn * VU_BK * 2; }

__device__ __forceinline__ uint32_t vu_smem(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t vu_desc(uint32_t saddr) {
    // SmemDescriptor (cute/arch/mma_sm100_desc.hpp): start >> 4 [0,14), LBO >> 4 [16,30), SBO >> 4 [32,46),
    // version 1 [46,48), layout type SWIZZLE_NONE [61,64)
    return (uint64_t) ((saddr & 0x3FFFFu) >> 4) | ((uint64_t) (VU_LBO >> 4) << 16) | ((uint64_t) (VU_SBO >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ uint64_t vu_desc2(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t) ((saddr & 0x3FFFFu) >> 4) | ((uint64_t) (lbo >> 4) << 16) | ((uint64_t) (sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void vu_cp16(uint32_t dst, const void * src, bool ok) {
    const int n = ok ? 16 : 0;                                      // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(n) : "memory");
}

__device__ __forceinline__ void vu_mbar_wait(unsigned long long * bar, uint32_t parity) {
    asm volatile("{ .reg .pred q; W_%=: mbarrier.try_wait.parity.shared::cta.b64 q, [%0], %1; @q bra D_%=; bra W_%=; D_%=: }"
                 :: "r"(vu_smem(bar)), "r"(parity) : "memory");
}

// TMA tensor-map load of a [box rows][4 k-chunks][8 halves] tile (cp.async.bulk.tensor.3d, SASS UTMALDG): coordinates
// {0, row, k / 8}; rows outside [0, rows) -- the causal left pad of a conv, the tail of the last tile -- arrive as zeros
__device__ __forceinline__ void vu_tma_3d(uint32_t dst, const CUtensorMap * tm, int c1, int c2, unsigned long long * bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"(dst), "l"(tm), "r"(0), "r"(c1), "r"(c2), "r"(vu_smem(bar)) : "memory");
}

template <int EPI>
__global__ void __launch_bounds__(VU_LAUNCH)
vgemm_umma(const VGemm p, const int BN, const int MT, const int use_tma, const int a_box_rows, const int a_box_kc,
           const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW) {
    extern __shared__ __align__(1024) unsigned char vu_sm[];
    __shared__ unsigned long long mbar_full[VU_STAGES], mbar_empty[VU_STAGES];
    __shared__ unsigned long long mbar_done;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const long m0 = (long) blockIdx.x * VU_BM * MT;
    const int n0 = blockIdx.y * BN;
    const int b_stage = vu_b_stage(BN);
    unsigned char * smA = vu_sm;                                    // A ring (see VU_A_RING)
    unsigned char * smB = vu_sm + vu_a_ring(MT);                    // [stage][BN x 32 halves]
    const int tmem_cols = BN * MT <= 32 ? 32 : BN * MT <= 64 ? 64 : BN * MT <= 128 ? 128 : 256;   // tile j: columns j * BN ..

    if (tid == 0) {
        for (int i = 0; i < VU_STAGES; ++i) {
            // cp.async path: one arrival per loader warp; TMA path: the producer thread's arrive.expect_tx
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(vu_smem(&mbar_full[i])), "r"(use_tma ? 1 : VU_THREADS / 32));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(vu_smem(&mbar_empty[i])));
        }
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(vu_smem(&mbar_done)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(vu_smem(&tmem_base_s)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;

    // step = kb * ntaps + tap (k-block major): the taps of a k-block share one A stage when their offsets are close
    const int n_steps = p.ntaps * (p.Kc / VU_BK);
    const int Ktot = p.ntaps * p.Kc;
    int min_off = p.tap_off[0], max_off = p.tap_off[0];
    for (int i = 1; i < p.ntaps; ++i) { min_off = min(min_off, p.tap_off[i]); max_off = max(max_off, p.tap_off[i]); }
    const bool share = p.ntaps > 1 && max_off - min_off <= VU_SPAN_MAX;
    const int ra = share ? vu_ra_shared(MT) : VU_BM * MT;            // rows of an A stage
    const uint32_t lbo_a = (uint32_t) ra * 16, a_stage_b = (uint32_t) ra * VU_BK * 2;
    const int n_a_stages = share ? 3 : VU_STAGES;
    // W stage layout.  cp.async path: 8-row groups of 4 k-chunks (LBO 128, SBO 512).  TMA path: what a {8, BN, 4} box
    // lands as, [k-chunk][row][8 halves] (LBO = BN * 16, SBO 128) -- the A ring's layout.
    const uint32_t lbo_b = use_tma ? (uint32_t) BN * 16 : (uint32_t) VU_LBO, sbo_b = use_tma ? 128u : (uint32_t) VU_SBO;
    const int ra_used = VU_BM * MT + (share ? max_off - min_off : 0);
    // instruction descriptor (kind::f16): D = F32 [4,6) = 1, A = B = F16, both K-major, N >> 3 [17,23), M >> 4 [24,29)
    const uint32_t idesc = (1u << 4) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (VU_BM >> 4) << 24);

    // per-thread loader state, fixed for the whole K loop: the thread owns 16-byte chunk (row r, k-chunk kc) of
    // A rows r + 64 i (i < 5) and of up to two W rows; only the row origin and the k origin change per step
    const int l_r = tid >> 2, l_kc = tid & 3;
    const uint32_t l_soff = (uint32_t) ((l_r >> 3) * VU_SBO + l_kc * VU_LBO + (l_r & 7) * 16);   // B: rows r and r + 64: + 8 * SBO
    const uint32_t l_aoff = (uint32_t) l_kc * lbo_a + (uint32_t) l_r * 16;                        // A: rows r + 64 i: + 1024 i
    const __half * l_a = p.A + l_kc * 8;
    const bool l_w_ok0 = l_r < BN && n0 + l_r < p.N, l_w_ok1 = l_r + 64 < BN && n0 + l_r + 64 < p.N;
    const __half * l_w0 = p.W + (size_t) (l_w_ok0 ? n0 + l_r : 0) * Ktot + l_kc * 8;
    const __half * l_w1 = p.W + (size_t) (l_w_ok1 ? n0 + l_r + 64 : 0) * Ktot + l_kc * 8;
    const bool l_w_has0 = l_r < BN, l_w_has1 = l_r + 64 < BN;

    auto load_step = [&](int step) {
        const int stage = step % VU_STAGES;
        const int kb = step / p.ntaps, tap = step - kb * p.ntaps, k0 = kb * VU_BK;
        if (!share || tap == 0) {
            const int ast = (share ? kb : step) % n_a_stages;
            const uint32_t a_dst = vu_smem(smA) + (uint32_t) ast * a_stage_b + l_aoff;
            const long ar = m0 + l_r + (share ? min_off : p.tap_off[tap]);
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                if (l_r + 64 * i < ra_used) {
                    const long a = ar + 64 * i;
                    const bool ok = a >= 0 && a < p.a_rows;
                    vu_cp16(a_dst + 1024 * i, l_a + (size_t) (ok ? a : 0) * p.lda + k0, ok);
                }
            }
        }
        const uint32_t b_dst = vu_smem(smB + stage * b_stage) + l_soff;
        const size_t wcol = (size_t) tap * p.Kc + k0;
        if (l_w_has0) vu_cp16(b_dst, l_w0 + wcol, l_w_ok0);
        if (l_w_has1) vu_cp16(b_dst + 8 * VU_SBO, l_w1 + wcol, l_w_ok1);
    };

    if (use_tma && tid == 0) {
        // ---- TMA producer: one thread walks the steps; a step's operands are 1..5 tensor-map copies of A (64 rows each, only when
        //      the A stage changes) and one of W, all completing on the stage's `full` barrier
        {
            // an A stage is [4 k-chunks][ra rows][16 B]; the A box is {8 halves, a_box_rows rows, a_box_kc k-chunks}: the whole
            // stage in one copy when ra <= 256 (box limit), else half stages one k-chunk at a time
            const int row_parts = ra / a_box_rows, kc_parts = (VU_BK / 8) / a_box_kc;
            const uint32_t a_bytes = (uint32_t) ra * VU_BK * 2;
            for (int it = 0; it < n_steps; ++it) {
                const int stage = it % VU_STAGES;
                if (it >= VU_STAGES) vu_mbar_wait(&mbar_empty[stage], (uint32_t) ((it / VU_STAGES) - 1) & 1u);
                const int kb = it / p.ntaps, tap = it - kb * p.ntaps, k0 = kb * VU_BK;
                const bool load_a = !share || tap == 0;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                             :: "r"(vu_smem(&mbar_full[stage])), "r"((uint32_t) b_stage + (load_a ? a_bytes : 0u)) : "memory");
                if (load_a) {
                    const int ast = (share ? kb : it) % n_a_stages;
                    const long ar = m0 + (share ? min_off : p.tap_off[tap]);
                    for (int i = 0; i < row_parts; ++i)
                        for (int kc = 0; kc < kc_parts; ++kc)
                            vu_tma_3d(vu_smem(smA) + (uint32_t) ast * a_stage_b + (uint32_t) (kc * a_box_kc) * lbo_a + (uint32_t) i * a_box_rows * 16,
                                      &tmA, (int) (ar + (long) i * a_box_rows), k0 / 8 + kc * a_box_kc, &mbar_full[stage]);
                }
                vu_tma_3d(vu_smem(smB + stage * b_stage), &tmW, n0, (tap * p.Kc + k0) / 8, &mbar_full[stage]);
            }
        }
    } else if (!use_tma && warp < VU_THREADS / 32) {
        // ---- loaders: iteration `it` issues the loads of step `it` and publishes step it - (VU_AHEAD - 1)
        for (int it = 0; it < n_steps + VU_AHEAD - 1; ++it) {
            if (it < n_steps) {
                if (it >= VU_STAGES) vu_mbar_wait(&mbar_empty[it % VU_STAGES], (uint32_t) ((it / VU_STAGES) - 1) & 1u);
                load_step(it);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");                          // possibly empty: keeps the group count = it + 1
            const int ready = it - (VU_AHEAD - 1);
            if (ready >= 0) {
                asm volatile("cp.async.wait_group %0;" :: "n"(VU_AHEAD - 1) : "memory");  // this thread's part of step `ready` landed
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");              // generic-proxy writes -> async proxy (UMMA)
                __syncwarp();
                if ((tid & 31) == 0)
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(vu_smem(&mbar_full[ready % VU_STAGES])) : "memory");
            }
        }
    } else if (tid == VU_THREADS) {
        // ---- MMA issuer
        for (int step = 0; step < n_steps; ++step) {
            const int stage = step % VU_STAGES;
            vu_mbar_wait(&mbar_full[stage], (uint32_t) (step / VU_STAGES) & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int kb = step / p.ntaps, tap = step - kb * p.ntaps;
            const uint32_t a_base = vu_smem(smA) + (uint32_t) ((share ? kb : step) % n_a_stages) * a_stage_b
                                  + (share ? (uint32_t) (p.tap_off[tap] - min_off) * 16u : 0u);
            const uint32_t b_base = vu_smem(smB + stage * b_stage);
#pragma unroll
            for (int kk = 0; kk < VU_BK / 16; ++kk) {
                const uint64_t ad = vu_desc2(a_base + kk * 2 * lbo_a, lbo_a, 128u), bd = vu_desc2(b_base + kk * 2 * lbo_b, lbo_b, sbo_b);
                const uint32_t accum = (step > 0 || kk > 0) ? 1u : 0u;
                asm volatile("{ .reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, q; }"
                             :: "r"(tmem_d), "l"(ad), "l"(bd), "r"(idesc), "r"(accum) : "memory");
                if (MT == 2) {                                   // second M tile: A rows + 128, accumulator columns + BN
                    const uint64_t ad1 = vu_desc2(a_base + kk * 2 * lbo_a + VU_BM * 16, lbo_a, 128u);
                    asm volatile("{ .reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, q; }"
                                 :: "r"(tmem_d + (uint32_t) BN), "l"(ad1), "l"(bd), "r"(idesc), "r"(accum) : "memory");
                }
            }
            // frees this stage when the MMAs above (and all earlier ones) have read their operands
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(vu_smem(&mbar_empty[stage])) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(vu_smem(&mbar_done)) : "memory");
    }
    // ---- epilogue: TMEM -> registers (thread = output row) -> shared memory (padded rows) -> coalesced fused
    //      epilogue (consecutive threads = consecutive column pairs; per-column parameters live in registers)
    vu_mbar_wait(&mbar_done, 0u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    float * stg = reinterpret_cast<float *>(vu_sm);                   // the operand ring is idle now
    const int sld = BN + 4;                                           // row stride (floats): 16-byte aligned, conflict-free float4 rows
    for (int mt = 0; mt < MT; ++mt) {
    const long m0t = m0 + (long) mt * VU_BM;
    if (m0t >= p.M) break;                                            // uniform over the CTA
    if (mt > 0) __syncthreads();                                      // the staging buffer is being read by the previous tile's stores
    if (warp < VU_THREADS / 32) {
        // a warp may touch TMEM lanes 32 * (warp % 4) .. + 31 only: warps w and w + 4 share a lane quadrant and
        // take alternate 16-column chunks
        const int quad = warp & 3, lane = tid & 31;
        const int row = quad * 32 + lane;
        for (int c0 = (warp >> 2) * 16; c0 < BN; c0 += 32) {
            uint32_t v[16];
            const uint32_t taddr = tmem_d + ((uint32_t) (quad * 32) << 16) + (uint32_t) (mt * BN + c0);
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                           "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                         : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            float4 * dst = reinterpret_cast<float4 *>(stg + (size_t) row * sld + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    {
        const int npairs = BN >> 1, rpp = VU_THREADS / npairs;        // rows handled per sweep
        if (tid < rpp * npairs) {
            const int pr = tid % npairs, r0 = tid / npairs;
            const int n = n0 + 2 * pr;
            if (n < p.N) {
                const bool ok1 = n + 1 < p.N;
                VEpiCols ec;
                vgemm_epi_cols(p, n, ok1, ec);
                // a thread reads and writes only its own elements (the residual may alias out32), so the
                // residual loads of the next rows can be issued ahead of the stores of the current ones
                constexpr int U = 4;
                for (int rb = r0; rb < VU_BM; rb += rpp * U) {
                    float2 rr[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const size_t m = (size_t) m0t + rb + u * rpp;
                        rr[u] = ((EPI == VEPI_GENERIC || EPI == VEPI_RES_SNAKE) && rb + u * rpp < VU_BM && m < (size_t) p.M)
                              ? vgemm_load_res(p, m, n, ok1) : make_float2(0.0f, 0.0f);
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int r = rb + u * rpp;
                        const size_t m = (size_t) m0t + r;
                        if (r < VU_BM && m < (size_t) p.M) {
                            const float2 v = *reinterpret_cast<const float2 *>(stg + (size_t) r * sld + 2 * pr);
                            vgemm_epilogue_cols<EPI>(p, m, n, v.x, v.y, ok1, ec, rr[u]);
                        }
                    }
                }
            }
        }
    }
    }   // M tiles
    __syncthreads();                                                  // every warp is done with TMEM
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_d), "r"(tmem_cols) : "memory");
}

inline bool voc_use_umma() {
    static int v = -1;
    if (v < 0) { const char * e = getenv("Q3T_VOC_SIMT"); v = (e && atoi(e)) ? 0 : 1; }
    return v != 0;
}

typedef CUresult (*vu_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline vu_encode_fn vu_encoder() {
    static vu_encode_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (vu_encode_fn) p;
    }
    return fn;
}
// [rows][ld] F16 matrix (k contiguous) as the rank-3 tensor {8 halves, rows, K / 8}: a box {8, box_rows, box_kc} lands in shared
// memory as [k-chunk][row][16 B] = the no-swizzle K-major core-matrix layout the UMMA descriptors describe (LBO = box_rows * 16, SBO 128)
inline bool vu_make_map(CUtensorMap * tm, const __half * base, long rows, long K, long ld, int box_rows, int box_kc) {
    vu_encode_fn enc = vu_encoder();
    if (!enc || ((uintptr_t) base & 15) || (ld % 8) || (K % 8) || box_rows > 256 || rows <= 0) return false;
    const cuuint64_t dims[3] = {8, (cuuint64_t) rows, (cuuint64_t) (K / 8)};
    const cuuint64_t strides[2] = {(cuuint64_t) ld * 2, 16};
    const cuuint32_t box[3] = {8, (cuuint32_t) box_rows, (cuuint32_t) box_kc};
    const cuuint32_t es[3] = {1, 1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<__half *>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

inline void voc_gemm(q3t_ctx * ctx, const VGemm & p) {
    if (voc_use_umma() && p.Kc % VU_BK == 0 && p.N >= 16 && p.N % 2 == 0 && p.lda % 8 == 0) {
        // widest tile <= 128 columns that divides N (no half-empty last tile), else 128 / N rounded up to 16
        int BN = p.N >= 128 ? 128 : ((p.N + 15) / 16) * 16;
        for (int c = 128; c >= 64; c -= 16) if (p.N % c == 0) { BN = c; break; }
        // two M tiles per CTA when the grid still gives every SM its two resident CTAs
        const long tiles256 = (long) ((p.M + 2 * VU_BM - 1) / (2 * VU_BM)) * ((p.N + BN - 1) / BN);
        static int mt_env = -1, mt_k = 1024, tma_env = -1;
        if (mt_env < 0) {
            const char * e = getenv("Q3T_VOC_MT"); mt_env = e ? atoi(e) : 0;
            if ((e = getenv("Q3T_VOC_MT_K"))) mt_k = atoi(e);
            e = getenv("Q3T_VOC_TMA"); tma_env = e ? atoi(e) : 1;      // 0: operands staged with cp.async (the round-1 path, kept for A/B runs)
        }
        const int MT = mt_env ? mt_env : (tiles256 >= 2 * 148 && p.ntaps * p.Kc >= mt_k ? 2 : 1);   // short K: the epilogue dominates, keep 3 CTAs per SM
        const int smem = std::max(vu_a_ring(MT) + VU_STAGES * vu_b_stage(BN), VU_BM * (BN + 4) * 4);
        const int smem_max = std::max(vu_a_ring(2) + VU_STAGES * vu_b_stage(128), VU_BM * (128 + 4) * 4);
        static bool attr_dev[64] = {};                     // function attributes are per device
        bool & attr = attr_dev[ctx->device & 63];
        if (!attr) {
            cudaFuncSetAttribute(vgemm_umma<VEPI_GENERIC>,    cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
            cudaFuncSetAttribute(vgemm_umma<VEPI_BIAS_SNAKE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
            cudaFuncSetAttribute(vgemm_umma<VEPI_RES_SNAKE>,  cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
            cudaFuncSetAttribute(vgemm_umma<VEPI_O32_SNAKE>,  cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
            attr = true;
        }
        // TMA tensor maps of both operands (the kernel's stage geometry decides the A box)
        int min_off = p.tap_off[0], max_off = p.tap_off[0];
        for (int i = 1; i < p.ntaps; ++i) { min_off = std::min(min_off, p.tap_off[i]); max_off = std::max(max_off, p.tap_off[i]); }
        const bool share = p.ntaps > 1 && max_off - min_off <= VU_SPAN_MAX;
        const int ra = share ? vu_ra_shared(MT) : VU_BM * MT;
        const int a_box_rows = ra <= 256 ? ra : ra / 2, a_box_kc = ra <= 256 ? VU_BK / 8 : 1;
        CUtensorMap tmA, tmW;
        memset(&tmA, 0, sizeof(tmA)); memset(&tmW, 0, sizeof(tmW));
        const int use_tma = tma_env && vu_make_map(&tmA, p.A, p.a_rows, p.Kc, p.lda, a_box_rows, a_box_kc) &&
                            vu_make_map(&tmW, p.W, p.N, (long) p.ntaps * p.Kc, (long) p.ntaps * p.Kc, BN, VU_BK / 8) ? 1 : 0;
        dim3 grid((unsigned) ((p.M + VU_BM * MT - 1) / (VU_BM * MT)), (unsigned) ((p.N + BN - 1) / BN));
        const bool snake16 = p.bias && !p.gamma && !p.swiglu && p.out16 && p.act == VACT_SNAKE && (p.ld16 & 1) == 0;
        if (snake16 && !p.res && !p.out32)
            vgemm_umma<VEPI_BIAS_SNAKE><<<grid, VU_LAUNCH, smem, ctx->stream>>>(p, BN, MT, use_tma, a_box_rows, a_box_kc, tmA, tmW);
        else if (snake16 && p.res && (p.ldres & 1) == 0 && (!p.out32 || (p.ld32 & 1) == 0))
            vgemm_umma<VEPI_RES_SNAKE><<<grid, VU_LAUNCH, smem, ctx->stream>>>(p, BN, MT, use_tma, a_box_rows, a_box_kc, tmA, tmW);
        else if (snake16 && !p.res && p.out32 && (p.ld32 & 1) == 0)
            vgemm_umma<VEPI_O32_SNAKE><<<grid, VU_LAUNCH, smem, ctx->stream>>>(p, BN, MT, use_tma, a_box_rows, a_box_kc, tmA, tmW);
        else
            vgemm_umma<VEPI_GENERIC><<<grid, VU_LAUNCH, smem, ctx->stream>>>(p, BN, MT, use_tma, a_box_rows, a_box_kc, tmA, tmW);
        return;
    }
    dim3 grid((unsigned) ((p.M + VG_BM - 1) / VG_BM), (unsigned) ((p.N + VG_BN - 1) / VG_BN));
    vgemm_simt<<<grid, 256, 0, ctx->stream>>>(p);
}

} // namespace q3t
