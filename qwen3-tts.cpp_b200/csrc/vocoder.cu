// Vocoder (audio_tokenizer_decoder): 16-codebook RVQ codes -> 24 kHz waveform on sm_100a.
//
// Replaces AudioTokenizerDecoder::{load_model tensor binding, normalize_codebooks, build_graph,
// decode} (src/audio_tokenizer_decoder.cpp:40-73, 118-316, 622-879) and the ggml ops behind them.
//
// Data layout: every activation is channels-last [time][channel] so that
//   * a Conv1d (k taps, dilation d, causal left pad) is a GEMM whose A operand is the same
//     activation matrix read at k row offsets  (no im2col buffer, the reference materialises one:
//     ggml/src/ggml.c:4409-4457);
//   * a ConvTranspose1d with K = 2*stride and the reference's (K - stride) trim on both sides
//     (src/audio_tokenizer_decoder.cpp:596-609) has exactly two taps per output sample:
//     y[q*s + r][o] = sum_c x[q][c] w[c][o][r+s] + x[q+1][c] w[c][o][r], i.e. one GEMM with
//     K = 2*Cin, N = s*Cout whose row-major output IS the time-major upsampled signal.
// Numerics follow the reference's rounding points: every weight op takes
// F16 weights and F16-rounded activations with F32 accumulation (im2col / conv_transpose / mul_mat
// all convert the activation to F16), everything element-wise is F32, the transformer's attention
// is F32 throughout, GELU goes through F16 like ggml's lookup table.  Elementwise work is fused
// into the GEMM epilogues: bias, layer scale, residual, snake -> the next conv's F16 operand.
#include "common.cuh"
#include "engine.h"
#include "voc_gemm.cuh"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

using namespace q3t;

namespace q3t {

// ------------------------------------------------------------------------------- model
constexpr int VQ_DIM = 256, VQ_OUT = 512, LATENT = 1024, TH = 512, TNH = 16, THD = 64, TFFN = 1024, TL = 8;
constexpr int DEC_DIM = 1536;
constexpr float T_EPS = 1e-5f, T_THETA = 10000.0f;

struct VTfmLayer {
    const float * attn_norm, * ffn_norm, * attn_scale, * ffn_scale;
    const __half * wqkv;      // [3072][512]  q | k | v
    const __half * wo;        // [512][1024]
    const __half * wgu;       // [2048][512]  row 2i = gate_i, 2i+1 = up_i
    const __half * wdown;     // [512][1024]
};
struct VUpsample {
    const __half * convt;     // [2*1024][1024]   n = k*1024 + o
    const float * convt_b;
    const __half * dw;        // [1024][7] F16 (as in the file)
    const float * dw_b, * ln_w, * ln_b;
    const __half * pw1;       // [4096][1024]
    const float * pw1_b;
    const __half * pw2;       // [1024][4096]
    const float * pw2_b, * gamma;
};
struct VSnake { const float * ea, * eb; };      // exp(alpha), exp(-beta)  (src/audio_tokenizer_decoder.cpp:383-397)
struct VResUnit {
    VSnake act1, act2;
    const __half * w1;        // [C][7*C]  k-major taps
    const float * b1;
    const __half * w2;        // [C][C]
    const float * b2;
    int dilation;
};
struct VDecBlock {
    int cin, cout, stride;
    VSnake snake;
    const __half * convt;     // [s*Cout][2*Cin]
    const float * convt_b;
    VResUnit res[3];
};

} // namespace q3t

struct q3t_vocoder {
    // RVQ
    const __half * codebooks[16];      // [2048][256]
    int codebook_size = 2048;
    const __half * vq_proj;            // [512][16*256]
    const __half * pre_conv;           // [1024][3*512]
    const float * pre_conv_b;
    const __half * in_proj;            // [512][1024]
    const float * in_proj_b;
    VTfmLayer tfm[TL];
    const float * tfm_norm;
    const __half * out_proj;           // [1024][512]
    const float * out_proj_b;
    VUpsample up[2];
    const __half * dec0;               // [1536][7*1024]
    const float * dec0_b;
    VDecBlock blk[4];
    VSnake dec5;
    const __half * dec6;               // [7*96]  k-major
    const float * dec6_b;
    // Q8_0 matrices in the file (scripts/convert_tokenizer_to_gguf.py:286-297): the reference then quantises the ACTIVATIONS of
    // those mat-muls to Q8_0 too; the F16 operand is passed through the same quantise / de-quantise before the GEMM
    bool q8_in_proj = false, q8_tfm[TL] = {}, q8_out_proj = false, q8_pw[2] = {};
    float * rope = nullptr;            // [n_rope][64]: cos[32] | sin[32]
    int n_rope = 0;
    std::vector<void *> owned;         // repacked weights
    // scratch arena
    unsigned char * arena = nullptr;
    size_t arena_bytes = 0;
    int * codes_dev = nullptr; int codes_cap = 0;
};

namespace q3t {

// ------------------------------------------------------------------------------- small kernels
__global__ void k_repack_conv(const __half * __restrict__ w, __half * __restrict__ out, int Cout, int Cin, int K) {
    // ggml [K, Cin, Cout] (w[(o*Cin + c)*K + k])  ->  out[o][k*Cin + c]
    const size_t n = (size_t) Cout * Cin * K;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const int c = (int) (i % Cin); const int k = (int) ((i / Cin) % K); const int o = (int) (i / ((size_t) Cin * K));
        out[i] = w[((size_t) o * Cin + c) * K + k];
    }
}
__global__ void k_repack_convt2tap(const __half * __restrict__ w, __half * __restrict__ out, int Cin, int Cout, int s) {
    // ggml [K=2s, Cout, Cin] (w[(c*Cout + o)*K + k]) -> out[r*Cout + o][tap*Cin + c], tap0 <- k=r+s, tap1 <- k=r
    const size_t n = (size_t) s * Cout * 2 * Cin;
    const int K = 2 * s;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const int c = (int) (i % Cin); const int tap = (int) ((i / Cin) % 2);
        const int row = (int) (i / ((size_t) 2 * Cin)); const int o = row % Cout, r = row / Cout;
        const int k = tap == 0 ? r + s : r;
        out[i] = w[((size_t) c * Cout + o) * K + k];
    }
}
__global__ void k_repack_convt_k2s2(const __half * __restrict__ w, __half * __restrict__ out, int Cin, int Cout) {
    // ggml [K=2, Cout, Cin] -> out[k*Cout + o][c]
    const size_t n = (size_t) 2 * Cout * Cin;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const int c = (int) (i % Cin); const int row = (int) (i / Cin); const int o = row % Cout, k = row / Cout;
        out[i] = w[((size_t) c * Cout + o) * 2 + k];
    }
}
__global__ void k_copy_rows_h(const __half * __restrict__ src, __half * __restrict__ dst, int rows, int K, int dst_stride_rows, int dst_row0, int dst_ld, int dst_col0) {
    // dst[(dst_row0 + r*dst_stride_rows)][dst_col0 + k] = src[r][k]
    const size_t n = (size_t) rows * K;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const int k = (int) (i % K); const int r = (int) (i / K);
        dst[((size_t) dst_row0 + (size_t) r * dst_stride_rows) * dst_ld + dst_col0 + k] = src[i];
    }
}
__global__ void k_exp(const float * __restrict__ x, float * __restrict__ y, int n, float sign) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = expf(sign * x[i]);
}
// Activation side of ggml's mul_mat with Q8_0 weights (AVX2 quantize_row_q8_0, arch/x86/quants.c:290-385): per block of 32,
// d = F16(amax / 127), q = round-half-even(x * 127 / amax); the operand becomes q * d (one warp per block, in place)
// src32 (optional): the un-rounded activations, which is what the reference quantises (rounding to F16 first would flip
// a few percent of the quants by one step)
__global__ void k_fake_quant_q8(__half * __restrict__ a, const float * __restrict__ src32, long n_blocks) {
    const int lane = threadIdx.x & 31;
    const long w0 = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long) gridDim.x * blockDim.x) >> 5;
    for (long b = w0; b < n_blocks; b += nw) {
        const float v = src32 ? src32[b * 32 + lane] : __half2float(a[b * 32 + lane]);
        float amax = fabsf(v);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float d  = __half2float(__float2half_rn(__fdiv_rn(amax, 127.0f)));
        const float id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
        const float q = rintf(__fmul_rn(v, id));
        a[b * 32 + lane] = __float2half_rn(__fmul_rn(q, d));
    }
}

// normalize_codebooks (src/audio_tokenizer_decoder.cpp:40-64): row / max(usage, 1e-5), re-rounded to F16
__global__ void k_normalize_codebook(__half * cb, const float * __restrict__ usage, int rows, int dim) {
    const size_t n = (size_t) rows * dim;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        float u = usage[i / dim];
        if (u < 1e-5f) u = 1e-5f;
        cb[i] = __float2half_rn(__half2float(cb[i]) * (1.0f / u));
    }
}

// RVQ gather: A[t][cb*256 + j] = codebook_cb[code[t][cb]][j]   (get_rows, :648-661)
struct GatherArgs { const __half * cb[16]; };
__global__ void k_rvq_gather(const GatherArgs g, const int * __restrict__ codes, int cb_size, __half * __restrict__ A) {
    const int t = blockIdx.x, cb = blockIdx.y;
    int code = codes[t * 16 + cb];
    code = min(max(code, 0), cb_size - 1);
    const uint4 * src = reinterpret_cast<const uint4 *>(g.cb[cb] + (size_t) code * VQ_DIM);
    uint4 * dst = reinterpret_cast<uint4 *>(A + ((size_t) t * 16 + cb) * VQ_DIM);
    if (threadIdx.x < VQ_DIM / 8) dst[threadIdx.x] = src[threadIdx.x];
}

// RMSNorm over 512 channels * gamma -> F16 (the next mat-mul's operand); one warp per row
__global__ void k_rmsnorm512(const float * __restrict__ x, const float * __restrict__ gamma, __half * __restrict__ out, int T, float eps,
                             float * __restrict__ o32 = nullptr) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= T) return;
    const float * xr = x + (size_t) row * TH;
    float v[16]; double sq = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) { v[i] = xr[lane + 32 * i]; sq += (double) (v[i] * v[i]); }
    sq = warp_sum_d(sq);
    const float scale = 1.0f / sqrtf((float) (sq / TH) + eps);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float y = (v[i] * scale) * gamma[lane + 32 * i];
        out[(size_t) row * TH + lane + 32 * i] = __float2half_rn(y);
        if (o32) o32[(size_t) row * TH + lane + 32 * i] = y;
    }
}

// NEOX RoPE on q and k (heads of 64, pairs (i, i+32)), in place on the F32 qkv buffer [T][3072]
__global__ void k_rope_qk(float * __restrict__ qkv, const float * __restrict__ rope, int T) {
    const int t = blockIdx.x;
    const float * rr = rope + (size_t) t * THD;
    for (int i = threadIdx.x; i < 2 * TNH * 32; i += blockDim.x) {       // q heads then k heads
        const int hh = i >> 5, d = i & 31;
        float * p = qkv + (size_t) t * 3 * LATENT + hh * THD;
        const float c = rr[d], s = rr[32 + d];
        const float x0 = p[d], x1 = p[d + 32];
        p[d] = __fmaf_rn(x0, c, -__fmul_rn(x1, s));         // contraction of the reference build (g++ -O3 -mfma)
        p[d + 32] = __fmaf_rn(x0, s, __fmul_rn(x1, c));
    }
}

// causal attention, F32 (the vocoder's K/Q/V never pass through F16: :450-470); out F16 (operand of the
// o-projection).
constexpr int VA_QW = 4, VA_QB = 4 * VA_QW;          // queries per warp / per block (4 warps)
__global__ void __launch_bounds__(128)
k_attention(const float * __restrict__ qkv, __half * __restrict__ out, int T, float * __restrict__ o32 = nullptr) {
    // Causal soft-max attention of 16 consecutive queries of one head.  K/V tiles of 32 positions are staged in
    // shared memory once per block (K transposed: lane = key position in the score loop), a warp carries four
    // queries through the tile (running max / sum, F32), lane = output dims (lane, lane + 32) in the P.V loop.
    // Tiles start at absolute multiples of 32 and masked keys contribute exact zeros, so a query's result does
    // not depend on how many rows the call has (streaming decode == whole decode, bit for bit).
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, h = blockIdx.y;
    const int tq0 = blockIdx.x * VA_QB + warp * VA_QW;
    __shared__ float Ks[THD * 33];
    __shared__ __align__(16) float Vs[32 * THD];
    __shared__ float qs[4][VA_QW][THD];
#pragma unroll
    for (int qi = 0; qi < VA_QW; ++qi) {
        const int t = tq0 + qi;
        qs[warp][qi][lane]      = t < T ? qkv[(size_t) t * 3072 + h * THD + lane] * 0.125f : 0.f;        // 1/sqrt(64), exact
        qs[warp][qi][lane + 32] = t < T ? qkv[(size_t) t * 3072 + h * THD + lane + 32] * 0.125f : 0.f;
    }
    const float * kb = qkv + LATENT + h * THD;
    const float * vb = qkv + 2 * LATENT + h * THD;
    float m[VA_QW], l[VA_QW], o0[VA_QW], o1[VA_QW];
#pragma unroll
    for (int qi = 0; qi < VA_QW; ++qi) { m[qi] = -INFINITY; l[qi] = 0.f; o0[qi] = 0.f; o1[qi] = 0.f; }
    const int t_last = min(T - 1, (int) blockIdx.x * VA_QB + VA_QB - 1);
    const int ntiles = t_last / 32 + 1;
    for (int tile = 0; tile < ntiles; ++tile) {
        const int j0 = tile * 32;
        __syncthreads();                                         // the previous tile has been consumed
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = threadIdx.x + 128 * i, row = idx >> 4, c4 = idx & 15;
            float4 k4 = make_float4(0.f, 0.f, 0.f, 0.f), v4 = k4;
            if (j0 + row < T) {
                k4 = reinterpret_cast<const float4 *>(kb + (size_t) (j0 + row) * 3072)[c4];
                v4 = reinterpret_cast<const float4 *>(vb + (size_t) (j0 + row) * 3072)[c4];
            }
            Ks[(4 * c4) * 33 + row] = k4.x; Ks[(4 * c4 + 1) * 33 + row] = k4.y; Ks[(4 * c4 + 2) * 33 + row] = k4.z; Ks[(4 * c4 + 3) * 33 + row] = k4.w;
            reinterpret_cast<float4 *>(Vs + row * THD)[c4] = v4;
        }
        __syncthreads();
        if (j0 > tq0 + VA_QW - 1 || tq0 >= T) continue;          // warp-uniform: nothing visible to this warp's queries
        float sc[VA_QW];
#pragma unroll
        for (int qi = 0; qi < VA_QW; ++qi) sc[qi] = 0.f;
#pragma unroll 16
        for (int d = 0; d < THD; ++d) {
            const float kd = Ks[d * 33 + lane];
#pragma unroll
            for (int qi = 0; qi < VA_QW; ++qi) sc[qi] = fmaf(kd, qs[warp][qi][d], sc[qi]);
        }
        float pq[VA_QW];
#pragma unroll
        for (int qi = 0; qi < VA_QW; ++qi) {
            const int t = tq0 + qi;
            const bool valid = j0 + lane <= t && t < T;
            const float sv = valid ? sc[qi] : -INFINITY;
            const float mn = fmaxf(m[qi], warp_max(sv));
            float p = 0.f, corr = 1.f;
            if (mn != -INFINITY) { p = valid ? expf(sv - mn) : 0.f; corr = expf(m[qi] - mn); }
            l[qi] = l[qi] * corr + warp_sum(p);
            o0[qi] *= corr; o1[qi] *= corr;
            m[qi] = mn; pq[qi] = p;
        }
        const int nj = min(32, tq0 + VA_QW - j0);                // keys beyond the warp's last query have p = 0
        for (int j = 0; j < nj; ++j) {
            const float v0 = Vs[j * THD + lane], v1 = Vs[j * THD + lane + 32];
#pragma unroll
            for (int qi = 0; qi < VA_QW; ++qi) {
                const float pj = __shfl_sync(0xffffffffu, pq[qi], j);
                o0[qi] = fmaf(pj, v0, o0[qi]); o1[qi] = fmaf(pj, v1, o1[qi]);
            }
        }
    }
#pragma unroll
    for (int qi = 0; qi < VA_QW; ++qi) {
        const int t = tq0 + qi;
        if (t < T) {
            const float inv = 1.0f / l[qi];
            out[(size_t) t * LATENT + h * THD + lane]      = __float2half_rn(o0[qi] * inv);
            out[(size_t) t * LATENT + h * THD + lane + 32] = __float2half_rn(o1[qi] * inv);
            if (o32) { o32[(size_t) t * LATENT + h * THD + lane] = o0[qi] * inv; o32[(size_t) t * LATENT + h * THD + lane + 32] = o1[qi] * inv; }
        }
    }
}

// depthwise causal conv k7 (+bias) on F16-rounded input, LayerNorm over the 1024 channels
// (eps 1e-6, ggml_norm: mean, then variance of the centred values), * w + b -> F16
// (apply_upsample_block, :509-529).  One block per time step, 256 threads x 4 channels.
__global__ void __launch_bounds__(256)
k_dwconv_ln(const float * __restrict__ x /* [L][1024] F32 */, const __half * __restrict__ w /* [1024][7] */, const float * __restrict__ b,
            const float * __restrict__ ln_w, const float * __restrict__ ln_b, __half * __restrict__ out, int L, float * __restrict__ o32 = nullptr) {
    const int t = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    __shared__ float red[8];
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = tid + 256 * i;
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const int tt = t - 6 + k;
            if (tt >= 0) acc = fmaf(__half2float(w[c * 7 + k]), r16(x[(size_t) tt * LATENT + c]), acc);
        }
        v[i] = acc + b[c];
    }
    float s = v[0] + v[1] + v[2] + v[3];
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) tot += red[i];
    const float mean = tot / LATENT;
    __syncthreads();
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[i] -= mean; q += v[i] * v[i]; }
    q = warp_sum(q);
    if (lane == 0) red[warp] = q;
    __syncthreads();
    tot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) tot += red[i];
    const float scale = 1.0f / sqrtf(tot / LATENT + 1e-6f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = tid + 256 * i;
        const float y = (v[i] * scale) * ln_w[c] + ln_b[c];
        out[(size_t) t * LATENT + c] = __float2half_rn(y);
        if (o32) o32[(size_t) t * LATENT + c] = y;
    }
}

// final causal conv k7 96 -> 1 (+bias) and tanh (:784-792): one thread per output sample;
// a = F16 snake output [L][96], w = [7*96] k-major.
__global__ void __launch_bounds__(256)
k_final_conv_tanh(const __half * __restrict__ a, const __half * __restrict__ w, const float * __restrict__ b, float * __restrict__ out, long L) {
    __shared__ float ws[7 * 96];
    for (int i = threadIdx.x; i < 7 * 96; i += blockDim.x) ws[i] = __half2float(w[i]);
    __syncthreads();
    const long t = blockIdx.x * (long) blockDim.x + threadIdx.x;
    if (t >= L) return;
    float acc = 0.f;
    for (int k = 0; k < 7; ++k) {
        const long tt = t - 6 + k;
        if (tt < 0) continue;
        const uint4 * row = reinterpret_cast<const uint4 *>(a + (size_t) tt * 96);
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const uint4 u = row[j];
            const __half2 * h2 = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(h2[e]);
                acc = fmaf(f.x, ws[k * 96 + j * 8 + 2 * e], acc);
                acc = fmaf(f.y, ws[k * 96 + j * 8 + 2 * e + 1], acc);
            }
        }
    }
    out[t] = tanhf(acc + b[0]);
}

// ------------------------------------------------------------------------------- load / repack
static const q3t_host_tensor * vfind(q3t_ctx * ctx, const std::string & name) {
    auto it = ctx->raw.find(name);
    return it == ctx->raw.end() ? nullptr : &it->second;
}

struct VLoad {
    q3t_ctx * ctx; q3t_vocoder * v; int rc = 0; int64_t bytes = 0;
    const q3t_host_tensor * get(const std::string & name, int32_t type, int64_t n_elem) {
        if (rc) return nullptr;
        const q3t_host_tensor * t = vfind(ctx, name);
        if (!t) { rc = set_err(ctx, Q3T_ERR_FORMAT, "missing tensor: " + name); return nullptr; }
        if (t->type != type) { rc = set_err(ctx, Q3T_ERR_FORMAT, "unexpected dtype for tensor: " + name + (type == Q3T_TYPE_F16 ? " (want F16)" : " (want F32)")); return nullptr; }
        if (t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3] != n_elem) { rc = set_err(ctx, Q3T_ERR_FORMAT, "unexpected shape for tensor: " + name); return nullptr; }
        bytes += (int64_t) t->nbytes;
        return t;
    }
    const __half * h(const std::string & name, int64_t n) { auto t = get(name, Q3T_TYPE_F16, n); return t ? (const __half *) t->dev : nullptr; }
    const float * f(const std::string & name, int64_t n) { auto t = get(name, Q3T_TYPE_F32, n); return t ? (const float *) t->dev : nullptr; }
    template <typename T> T * alloc(size_t n) {
        if (rc) return nullptr;
        void * p = nullptr;
        if (cudaMalloc(&p, std::max<size_t>(n * sizeof(T), 16)) != cudaSuccess) { rc = set_err(ctx, Q3T_ERR_CUDA, "cudaMalloc failed (vocoder weights)"); return nullptr; }
        v->owned.push_back(p);
        return (T *) p;
    }
    const __half * conv(const std::string & name, int Cout, int Cin, int K) {
        const __half * w = h(name, (int64_t) Cout * Cin * K);
        __half * o = alloc<__half>((size_t) Cout * Cin * K);
        if (rc) return nullptr;
        k_repack_conv<<<1024, 256, 0, ctx->stream>>>(w, o, Cout, Cin, K);
        return o;
    }
    VSnake snake(const std::string & prefix, const char * a, const char * b, int C) {
        VSnake s{nullptr, nullptr};
        const float * al = f(prefix + a, C), * be = f(prefix + b, C);
        float * ea = alloc<float>(C), * eb = alloc<float>(C);
        if (rc) return s;
        k_exp<<<(C + 255) / 256, 256, 0, ctx->stream>>>(al, ea, C, 1.0f);
        k_exp<<<(C + 255) / 256, 256, 0, ctx->stream>>>(be, eb, C, -1.0f);
        s.ea = ea; s.eb = eb;
        return s;
    }
};

void vocoder_free(q3t_ctx * ctx) {
    q3t_vocoder * v = ctx->voc;
    if (!v) return;
    cudaStreamSynchronize(ctx->stream);
    for (void * p : v->owned) cudaFree(p);
    if (v->arena) cudaFree(v->arena);
    if (v->codes_dev) cudaFree(v->codes_dev);
    if (v->rope) cudaFree(v->rope);
    if (ctx->d_samples) { cudaFree(ctx->d_samples); ctx->d_samples = nullptr; ctx->samples_cap = 0; }
    delete v;
    ctx->voc = nullptr;
    ctx->voc_bytes = 0;
}

int vocoder_finalize(q3t_ctx * ctx) {
    vocoder_free(ctx);
    q3t_vocoder * v = new q3t_vocoder();
    ctx->voc = v;
    VLoad L{ctx, v};
    const std::string P = "tok_dec.";
    cudaStream_t st = ctx->stream;

    // ---- RVQ: codebooks (optionally usage-normalised) and the K-concatenated output projection
    for (int cb = 0; cb < 16; ++cb) {
        const std::string base = cb == 0 ? P + "vq_first.0" : P + "vq_rest." + std::to_string(cb - 1);
        const q3t_host_tensor * probe = vfind(ctx, base + ".codebook");
        const q3t_host_tensor * t = L.get(base + ".codebook", Q3T_TYPE_F16, probe ? probe->ne[0] * probe->ne[1] * probe->ne[2] * probe->ne[3] : 0);
        if (L.rc) break;
        if (t->ne[0] != VQ_DIM) { L.rc = set_err(ctx, Q3T_ERR_FORMAT, "codebook dim must be 256: " + base); break; }
        if (cb == 0) v->codebook_size = (int) t->ne[1];
        else if (t->ne[1] != v->codebook_size) { L.rc = set_err(ctx, Q3T_ERR_FORMAT, "codebook sizes differ"); break; }
        v->codebooks[cb] = (const __half *) t->dev;
        if (const q3t_host_tensor * u = vfind(ctx, base + ".usage")) {
            if (u->type == Q3T_TYPE_F32 && u->ne[0] == t->ne[1])
                k_normalize_codebook<<<512, 256, 0, st>>>((__half *) t->dev, (const float *) u->dev, (int) t->ne[1], VQ_DIM);
        }
    }
    {
        const __half * p1 = L.h(P + "vq_first.output_proj.weight", (int64_t) VQ_OUT * VQ_DIM);
        const __half * p2 = L.h(P + "vq_rest.output_proj.weight", (int64_t) VQ_OUT * VQ_DIM);
        __half * w = L.alloc<__half>((size_t) VQ_OUT * 16 * VQ_DIM);
        if (!L.rc) {
            for (int cb = 0; cb < 16; ++cb)
                k_copy_rows_h<<<256, 256, 0, st>>>(cb == 0 ? p1 : p2, w, VQ_OUT, VQ_DIM, 1, 0, 16 * VQ_DIM, cb * VQ_DIM);
            v->vq_proj = w;
        }
    }
    v->pre_conv = L.conv(P + "pre_conv.weight", LATENT, VQ_OUT, 3);
    v->pre_conv_b = L.f(P + "pre_conv.bias", LATENT);
    v->in_proj = L.h(P + "pre_tfm.input_proj.weight", (int64_t) TH * LATENT);
    v->in_proj_b = L.f(P + "pre_tfm.input_proj.bias", TH);
    auto was_q8 = [&](const std::string & name) { const q3t_host_tensor * t = vfind(ctx, name); return t && t->src_type == Q3T_TYPE_Q8_0; };
    v->q8_in_proj = was_q8(P + "pre_tfm.input_proj.weight");
    v->q8_out_proj = was_q8(P + "pre_tfm.output_proj.weight");
    for (int l = 0; l < TL && !L.rc; ++l) {
        const std::string b = P + "pre_tfm.blk." + std::to_string(l) + ".";
        VTfmLayer & y = v->tfm[l];
        y.attn_norm = L.f(b + "attn_norm.weight", TH); y.ffn_norm = L.f(b + "ffn_norm.weight", TH);
        y.attn_scale = L.f(b + "attn_scale", TH); y.ffn_scale = L.f(b + "ffn_scale", TH);
        const __half * q = L.h(b + "attn_q.weight", (int64_t) LATENT * TH), * k = L.h(b + "attn_k.weight", (int64_t) LATENT * TH),
                     * vv = L.h(b + "attn_v.weight", (int64_t) LATENT * TH);
        const __half * g = L.h(b + "ffn_gate.weight", (int64_t) TFFN * TH), * u = L.h(b + "ffn_up.weight", (int64_t) TFFN * TH);
        y.wo = L.h(b + "attn_output.weight", (int64_t) TH * LATENT);
        y.wdown = L.h(b + "ffn_down.weight", (int64_t) TH * TFFN);
        __half * wqkv = L.alloc<__half>((size_t) 3 * LATENT * TH), * wgu = L.alloc<__half>((size_t) 2 * TFFN * TH);
        if (L.rc) break;
        k_copy_rows_h<<<256, 256, 0, st>>>(q, wqkv, LATENT, TH, 1, 0, TH, 0);
        k_copy_rows_h<<<256, 256, 0, st>>>(k, wqkv, LATENT, TH, 1, LATENT, TH, 0);
        k_copy_rows_h<<<256, 256, 0, st>>>(vv, wqkv, LATENT, TH, 1, 2 * LATENT, TH, 0);
        k_copy_rows_h<<<256, 256, 0, st>>>(g, wgu, TFFN, TH, 2, 0, TH, 0);
        k_copy_rows_h<<<256, 256, 0, st>>>(u, wgu, TFFN, TH, 2, 1, TH, 0);
        y.wqkv = wqkv; y.wgu = wgu;
        v->q8_tfm[l] = was_q8(b + "attn_q.weight");
    }
    v->tfm_norm = L.f(P + "pre_tfm.norm.weight", TH);
    v->out_proj = L.h(P + "pre_tfm.output_proj.weight", (int64_t) LATENT * TH);
    v->out_proj_b = L.f(P + "pre_tfm.output_proj.bias", LATENT);
    for (int i = 0; i < 2 && !L.rc; ++i) {
        const std::string b = P + "upsample." + std::to_string(i) + ".";
        VUpsample & u = v->up[i];
        const __half * ct = L.h(b + "conv.weight", (int64_t) LATENT * LATENT * 2);
        __half * ctp = L.alloc<__half>((size_t) 2 * LATENT * LATENT);
        if (L.rc) break;
        k_repack_convt_k2s2<<<1024, 256, 0, st>>>(ct, ctp, LATENT, LATENT);
        u.convt = ctp;
        u.convt_b = L.f(b + "conv.bias", LATENT);
        u.dw = L.h(b + "dwconv.weight", (int64_t) LATENT * 7);
        u.dw_b = L.f(b + "dwconv.bias", LATENT);
        u.ln_w = L.f(b + "norm.weight", LATENT); u.ln_b = L.f(b + "norm.bias", LATENT);
        u.pw1 = L.h(b + "pwconv1.weight", (int64_t) 4 * LATENT * LATENT); u.pw1_b = L.f(b + "pwconv1.bias", 4 * LATENT);
        u.pw2 = L.h(b + "pwconv2.weight", (int64_t) 4 * LATENT * LATENT); u.pw2_b = L.f(b + "pwconv2.bias", LATENT);
        u.gamma = L.f(b + "gamma", LATENT);
        v->q8_pw[i] = was_q8(b + "pwconv1.weight");
    }
    v->dec0 = L.conv(P + "dec.0.conv.weight", DEC_DIM, LATENT, 7);
    v->dec0_b = L.f(P + "dec.0.conv.bias", DEC_DIM);
    const int rates[4] = {8, 5, 4, 3};
    const int dil[3] = {1, 3, 9};
    int cin = DEC_DIM;
    for (int i = 0; i < 4 && !L.rc; ++i) {
        const std::string b = P + "dec." + std::to_string(i + 1) + ".";
        VDecBlock & B = v->blk[i];
        B.cin = cin; B.cout = cin / 2; B.stride = rates[i];
        B.snake = L.snake(b, "snake.alpha", "snake.beta", cin);
        const __half * ct = L.h(b + "conv_t.weight", (int64_t) cin * B.cout * 2 * B.stride);
        __half * ctp = L.alloc<__half>((size_t) B.stride * B.cout * 2 * cin);
        if (L.rc) break;
        k_repack_convt2tap<<<1024, 256, 0, st>>>(ct, ctp, cin, B.cout, B.stride);
        B.convt = ctp;
        B.convt_b = L.f(b + "conv_t.bias", B.cout);
        for (int r = 0; r < 3 && !L.rc; ++r) {
            const std::string rb = b + "res." + std::to_string(r + 2) + ".";
            VResUnit & R = B.res[r];
            R.dilation = dil[r];
            R.act1 = L.snake(rb, "act1.alpha", "act1.beta", B.cout);
            R.act2 = L.snake(rb, "act2.alpha", "act2.beta", B.cout);
            R.w1 = L.conv(rb + "conv1.weight", B.cout, B.cout, 7); R.b1 = L.f(rb + "conv1.bias", B.cout);
            R.w2 = L.h(rb + "conv2.weight", (int64_t) B.cout * B.cout); R.b2 = L.f(rb + "conv2.bias", B.cout);
        }
        cin = B.cout;
    }
    v->dec5 = L.snake(P + "dec.5.", "snake.alpha", "snake.beta", 96);
    v->dec6 = L.conv(P + "dec.6.conv.weight", 1, 96, 7);
    v->dec6_b = L.f(P + "dec.6.conv.bias", 1);
    if (L.rc) { vocoder_free(ctx); return L.rc; }
    int rc = cuda_check(ctx, cudaGetLastError(), "vocoder repack");
    if (!rc) rc = cuda_check(ctx, cudaStreamSynchronize(st), "vocoder repack");
    if (rc) { vocoder_free(ctx); return rc; }
    ctx->voc_bytes = L.bytes;
    ctx->voc_ready = true;
    return Q3T_OK;
}

// RoPE table with the reference's recurrence (ggml/src/ggml-cpu/ops.cpp:5548-5562), theta 1e4, 64 dims
static int ensure_rope(q3t_ctx * ctx, int T) {
    q3t_vocoder * v = ctx->voc;
    if (v->n_rope >= T) return Q3T_OK;
    const int n = std::max(T, 512);
    if (v->rope) { cudaStreamSynchronize(ctx->stream); cudaFree(v->rope); v->rope = nullptr; v->n_rope = 0; }
    std::vector<float> tab((size_t) n * THD);
    const float theta_scale = powf(T_THETA, -2.0f / (float) THD);
    for (int p = 0; p < n; ++p) {
        float theta = (float) p;
        for (int i = 0; i < THD / 2; ++i) { tab[(size_t) p * THD + i] = cosf(theta); tab[(size_t) p * THD + 32 + i] = sinf(theta); theta *= theta_scale; }
    }
    int rc = cuda_check(ctx, cudaMalloc(&v->rope, tab.size() * 4), "cudaMalloc rope");
    if (rc) return rc;
    rc = cuda_check(ctx, cudaMemcpyAsync(v->rope, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, ctx->stream), "rope upload");
    cudaStreamSynchronize(ctx->stream);
    v->n_rope = n;
    return rc;
}

// ------------------------------------------------------------------------------- forward
struct Arena {
    unsigned char * base; size_t off = 0, cap;
    template <typename T> T * take(size_t n) {
        off = (off + 255) & ~(size_t) 255;
        T * p = reinterpret_cast<T *>(base + off);
        off += n * sizeof(T);
        return p;
    }
};

static size_t arena_need(int T) {
    // generous closed form: the decoder blocks dominate (see vocoder_run for the carve-up)
    size_t b = 0;
    const size_t t = (size_t) T;
    b += t * (4096 * 2 + 512 * 2 + 1024 * 2 + 512 * 4 + 512 * 2 + 3072 * 4 + 1024 * 2 + 1024 * 2 + 1024 * 2);
    b += 4 * t * (1024 * 4 + 1024 * 2 + 4096 * 2 + 1024 * 2) * 2;
    b += 4 * t * 1536 * 2;
    b += 4 * t * 1024 * 4;                                     // f32s (Q8_0 files)
    size_t L = 4 * t; int C = 1536; const int rates[4] = {8, 5, 4, 3};
    for (int i = 0; i < 4; ++i) { L = L * rates[i]; C /= 2; b += L * C * (4 + 4 + 2 + 2); }
    return b + (1 << 20);
}

// t_from > 0 (streaming, q3t_vocoder_decode_from): the RVQ / pre-conv / pre-transformer stages run on all T
// frames (full causal attention has no finite window), the convolutional stack only on latent frames
// [t_from, T); every conv is causal, so local sample j of the slice is global sample j + 1920 * t_from and is
// exact once j is past the stack's left context (< 10 frames, VOC_STREAM_CONTEXT).
int64_t vocoder_run(q3t_ctx * ctx, const int32_t * codes_dev, int T, float ** samples_dev, int t_from) {
    q3t_vocoder * v = ctx->voc;
    if (!v || !ctx->voc_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (T <= 0) return set_err(ctx, Q3T_ERR_ARG, "n_frames must be > 0");
    cudaStream_t st = ctx->stream;
    int rc = ensure_rope(ctx, T);
    if (rc) return rc;
    const size_t need_b = arena_need(T);
    if (v->arena_bytes < need_b) {
        cudaStreamSynchronize(st);
        if (v->arena) cudaFree(v->arena);
        v->arena = nullptr; v->arena_bytes = 0;
        if ((rc = cuda_check(ctx, cudaMalloc(&v->arena, need_b), "cudaMalloc vocoder scratch"))) return rc;
        v->arena_bytes = need_b;
    }
    if (t_from < 0 || t_from >= T) return set_err(ctx, Q3T_ERR_ARG, "internal: bad stream origin");
    const int64_t n_out = 1920LL * (T - t_from) - 555;
    if (ctx->samples_cap < n_out) {
        cudaStreamSynchronize(st);
        if (ctx->d_samples) cudaFree(ctx->d_samples);
        ctx->d_samples = nullptr; ctx->samples_cap = 0;
        if ((rc = cuda_check(ctx, cudaMalloc(&ctx->d_samples, (size_t) (n_out + 1024) * 4), "cudaMalloc samples"))) return rc;
        ctx->samples_cap = n_out + 1024;
    }
    Arena A{v->arena, 0, v->arena_bytes};
    int64_t & nl = ctx->timing.kernel_launches;
    // developer aid: Q3T_VOC_DEBUG_DIR=<dir> writes stage outputs as raw files (tools/dev/dev_voc_trace.py)
    const char * dbg_dir = getenv("Q3T_VOC_DEBUG_DIR");
    auto dump = [&](const char * name, const void * dev, size_t bytes) {
        if (!dbg_dir) return;
        std::vector<unsigned char> h(bytes);
        cudaStreamSynchronize(st);
        cudaMemcpy(h.data(), dev, bytes, cudaMemcpyDeviceToHost);
        const std::string path = std::string(dbg_dir) + "/" + name + ".bin";
        if (FILE * f = fopen(path.c_str(), "wb")) { fwrite(h.data(), 1, bytes, f); fclose(f); }
    };

    const bool any_q8 = v->q8_in_proj || v->q8_out_proj || v->q8_pw[0] || v->q8_pw[1] || v->q8_tfm[0];
    float * f32s = any_q8 ? A.take<float>((size_t) 4 * T * LATENT) : nullptr;      // un-rounded copy of the operand being quantised
    auto fq = [&](bool on, __half * a, const float * src32, size_t n_elem) {       // Q8_0 files: the reference's activation quantisation
        if (!on) return;
        const long nb = (long) (n_elem / 32);
        k_fake_quant_q8<<<(unsigned) std::min<long>((nb + 7) / 8, 148 * 8), 256, 0, st>>>(a, src32, nb); ++nl;
    };
    // ---- RVQ lookup + projections as one K = 16*256 GEMM -> F16 latent [T][512]
    __half * a_rvq = A.take<__half>((size_t) T * 16 * VQ_DIM);
    __half * lat16 = A.take<__half>((size_t) T * VQ_OUT);
    {
        GatherArgs g; for (int i = 0; i < 16; ++i) g.cb[i] = v->codebooks[i];
        k_rvq_gather<<<dim3(T, 16), 32, 0, st>>>(g, codes_dev, v->codebook_size, a_rvq); ++nl;
        VGemm p{}; p.A = a_rvq; p.lda = 16 * VQ_DIM; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = 16 * VQ_DIM;
        p.W = v->vq_proj; p.N = VQ_OUT; p.out16 = lat16; p.ld16 = VQ_OUT; p.pmod = VQ_OUT;
        voc_gemm(ctx, p); ++nl;
        dump("lat16", lat16, (size_t) T * VQ_OUT * 2);
    }
    // ---- pre_conv: causal k3 512 -> 1024 (+b) -> F16
    __half * pc16 = A.take<__half>((size_t) T * LATENT);
    {
        VGemm p{}; p.A = lat16; p.lda = VQ_OUT; p.a_rows = T; p.M = T; p.ntaps = 3; p.Kc = VQ_OUT;
        p.tap_off[0] = -2; p.tap_off[1] = -1; p.tap_off[2] = 0;
        p.W = v->pre_conv; p.N = LATENT; p.bias = v->pre_conv_b; p.pmod = LATENT; p.out16 = pc16; p.ld16 = LATENT;
        if (v->q8_in_proj) { p.out32 = f32s; p.ld32 = LATENT; }
        voc_gemm(ctx, p); ++nl;
        dump("pc16", pc16, (size_t) T * LATENT * 2);
    }
    // ---- transformer input projection -> F32 residual stream [T][512]
    float * x32 = A.take<float>((size_t) T * TH);
    fq(v->q8_in_proj, pc16, f32s, (size_t) T * LATENT);
    {
        VGemm p{}; p.A = pc16; p.lda = LATENT; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = LATENT;
        p.W = v->in_proj; p.N = TH; p.bias = v->in_proj_b; p.pmod = TH; p.out32 = x32; p.ld32 = TH;
        voc_gemm(ctx, p); ++nl;
        dump("tfm_in", x32, (size_t) T * TH * 4);
    }
    __half * n16 = A.take<__half>((size_t) T * TH);
    float * qkv = A.take<float>((size_t) T * 3 * LATENT);
    __half * att16 = A.take<__half>((size_t) T * LATENT);
    __half * h16 = A.take<__half>((size_t) T * TFFN);
    for (int l = 0; l < TL; ++l) {
        const VTfmLayer & y = v->tfm[l];
        k_rmsnorm512<<<(T + 7) / 8, 256, 0, st>>>(x32, y.attn_norm, n16, T, T_EPS, v->q8_tfm[l] ? f32s : nullptr); ++nl;
        fq(v->q8_tfm[l], n16, f32s, (size_t) T * TH);
        { VGemm p{}; p.A = n16; p.lda = TH; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = TH; p.W = y.wqkv; p.N = 3 * LATENT; p.pmod = 3 * LATENT;
          p.out32 = qkv; p.ld32 = 3 * LATENT; voc_gemm(ctx, p); ++nl; }
        k_rope_qk<<<T, 256, 0, st>>>(qkv, v->rope, T); ++nl;
        k_attention<<<dim3((T + VA_QB - 1) / VA_QB, TNH), 128, 0, st>>>(qkv, att16, T, v->q8_tfm[l] ? f32s : nullptr); ++nl;
        fq(v->q8_tfm[l], att16, f32s, (size_t) T * LATENT);
        { VGemm p{}; p.A = att16; p.lda = LATENT; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = LATENT; p.W = y.wo; p.N = TH; p.pmod = TH;
          p.gamma = y.attn_scale; p.res = x32; p.ldres = TH; p.out32 = x32; p.ld32 = TH; voc_gemm(ctx, p); ++nl; }
        k_rmsnorm512<<<(T + 7) / 8, 256, 0, st>>>(x32, y.ffn_norm, n16, T, T_EPS, v->q8_tfm[l] ? f32s : nullptr); ++nl;
        fq(v->q8_tfm[l], n16, f32s, (size_t) T * TH);
        { VGemm p{}; p.A = n16; p.lda = TH; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = TH; p.W = y.wgu; p.N = 2 * TFFN; p.pmod = 2 * TFFN;
          p.swiglu = 1; p.out16 = h16; p.ld16 = TFFN; if (v->q8_tfm[l]) { p.out32 = f32s; p.ld32 = TFFN; } voc_gemm(ctx, p); ++nl; }
        fq(v->q8_tfm[l], h16, f32s, (size_t) T * TFFN);
        { VGemm p{}; p.A = h16; p.lda = TFFN; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = TFFN; p.W = y.wdown; p.N = TH; p.pmod = TH;
          p.gamma = y.ffn_scale; p.res = x32; p.ldres = TH; p.out32 = x32; p.ld32 = TH; voc_gemm(ctx, p); ++nl; }
    }
    dump("tfm_x", x32, (size_t) T * TH * 4);
    k_rmsnorm512<<<(T + 7) / 8, 256, 0, st>>>(x32, v->tfm_norm, n16, T, T_EPS, v->q8_out_proj ? f32s : nullptr); ++nl;
    fq(v->q8_out_proj, n16, f32s, (size_t) T * TH);
    __half * cur16 = A.take<__half>((size_t) T * LATENT);
    { VGemm p{}; p.A = n16; p.lda = TH; p.a_rows = T; p.M = T; p.ntaps = 1; p.Kc = TH; p.W = v->out_proj; p.N = LATENT; p.pmod = LATENT;
      p.bias = v->out_proj_b; p.out16 = cur16; p.ld16 = LATENT; voc_gemm(ctx, p); ++nl; }
    dump("tfm_out16", cur16, (size_t) T * LATENT * 2);

    // ---- two ConvNeXt-style x2 upsample blocks
    cur16 += (size_t) t_from * LATENT;
    int L = T - t_from;
    for (int i = 0; i < 2; ++i) {
        const VUpsample & u = v->up[i];
        const int L2 = 2 * L;
        float * res32 = A.take<float>((size_t) L2 * LATENT);
        __half * d16 = A.take<__half>((size_t) L2 * LATENT);
        __half * g16 = A.take<__half>((size_t) L2 * 4 * LATENT);
        __half * y16 = A.take<__half>((size_t) L2 * LATENT);
        { VGemm p{}; p.A = cur16; p.lda = LATENT; p.a_rows = L; p.M = L; p.ntaps = 1; p.Kc = LATENT; p.W = u.convt; p.N = 2 * LATENT; p.pmod = LATENT;
          p.bias = u.convt_b; p.out32 = res32; p.ld32 = 2 * LATENT; voc_gemm(ctx, p); ++nl; }
        k_dwconv_ln<<<L2, 256, 0, st>>>(res32, u.dw, u.dw_b, u.ln_w, u.ln_b, d16, L2, v->q8_pw[i] ? f32s : nullptr); ++nl;
        fq(v->q8_pw[i], d16, f32s, (size_t) L2 * LATENT);
        { VGemm p{}; p.A = d16; p.lda = LATENT; p.a_rows = L2; p.M = L2; p.ntaps = 1; p.Kc = LATENT; p.W = u.pw1; p.N = 4 * LATENT; p.pmod = 4 * LATENT;
          p.bias = u.pw1_b; p.act = VACT_GELU; p.out16 = g16; p.ld16 = 4 * LATENT; voc_gemm(ctx, p); ++nl; }
        fq(v->q8_pw[i], g16, nullptr, (size_t) L2 * 4 * LATENT);      // GELU goes through the F16 table in the reference: g16 IS its operand
        { VGemm p{}; p.A = g16; p.lda = 4 * LATENT; p.a_rows = L2; p.M = L2; p.ntaps = 1; p.Kc = 4 * LATENT; p.W = u.pw2; p.N = LATENT; p.pmod = LATENT;
          p.bias = u.pw2_b; p.gamma = u.gamma; p.res = res32; p.ldres = LATENT; p.out16 = y16; p.ld16 = LATENT; voc_gemm(ctx, p); ++nl; }
        cur16 = y16; L = L2;
        dump(i == 0 ? "up0_16" : "up1_16", y16, (size_t) L2 * LATENT * 2);
        if (i == 0) { dump("up0_res32", res32, (size_t) L2 * LATENT * 4); dump("up0_d16", d16, (size_t) L2 * LATENT * 2); dump("up0_g16", g16, (size_t) L2 * 4 * LATENT * 2); }
    }
    // ---- dec.0: causal k7 1024 -> 1536 (+b), then dec.1's snake -> F16
    __half * a16 = A.take<__half>((size_t) L * DEC_DIM);
    { VGemm p{}; p.A = cur16; p.lda = LATENT; p.a_rows = L; p.M = L; p.ntaps = 7; p.Kc = LATENT;
      for (int k = 0; k < 7; ++k) p.tap_off[k] = k - 6;
      p.W = v->dec0; p.N = DEC_DIM; p.pmod = DEC_DIM; p.bias = v->dec0_b; p.act = VACT_SNAKE; p.ea = v->blk[0].snake.ea; p.eb = v->blk[0].snake.eb;
      p.out16 = a16; p.ld16 = DEC_DIM; voc_gemm(ctx, p); ++nl; }
    dump("dec0_snake16", a16, (size_t) L * DEC_DIM * 2);
    // ---- 4 decoder blocks: ConvT (x s) + 3 residual units
    for (int i = 0; i < 4; ++i) {
        const VDecBlock & B = v->blk[i];
        const int Lo = (L - 1) * B.stride, C = B.cout;
        float * xa = A.take<float>((size_t) Lo * C), * xb = A.take<float>((size_t) Lo * C);
        __half * s1 = A.take<__half>((size_t) Lo * C), * s2 = A.take<__half>((size_t) Lo * C);
        { VGemm p{}; p.A = a16; p.lda = B.cin; p.a_rows = L; p.M = L - 1; p.ntaps = 2; p.Kc = B.cin; p.tap_off[0] = 0; p.tap_off[1] = 1;
          p.W = B.convt; p.N = B.stride * C; p.pmod = C; p.bias = B.convt_b; p.out32 = xa; p.ld32 = B.stride * C;
          p.act = VACT_SNAKE; p.ea = B.res[0].act1.ea; p.eb = B.res[0].act1.eb; p.out16 = s1; p.ld16 = B.stride * C; voc_gemm(ctx, p); ++nl; }
        { char nm[32]; snprintf(nm, sizeof nm, "blk%d_convt32", i + 1); dump(nm, xa, (size_t) Lo * C * 4); }
        { char nm[32]; snprintf(nm, sizeof nm, "blk%d_convt32", i + 1); dump(nm, xa, (size_t) Lo * C * 4); }
        for (int r = 0; r < 3; ++r) {
            const VResUnit & R = B.res[r];
            { VGemm p{}; p.A = s1; p.lda = C; p.a_rows = Lo; p.M = Lo; p.ntaps = 7; p.Kc = C;
              for (int k = 0; k < 7; ++k) p.tap_off[k] = (k - 6) * R.dilation;
              p.W = R.w1; p.N = C; p.pmod = C; p.bias = R.b1; p.act = VACT_SNAKE; p.ea = R.act2.ea; p.eb = R.act2.eb; p.out16 = s2; p.ld16 = C;
              voc_gemm(ctx, p); ++nl; }
            const VSnake & nx = r < 2 ? B.res[r + 1].act1 : (i < 3 ? v->blk[i + 1].snake : v->dec5);
            { VGemm p{}; p.A = s2; p.lda = C; p.a_rows = Lo; p.M = Lo; p.ntaps = 1; p.Kc = C;
              p.W = R.w2; p.N = C; p.pmod = C; p.bias = R.b2; p.res = xa; p.ldres = C; p.out32 = (r < 2) ? xb : nullptr; p.ld32 = C;
              p.act = VACT_SNAKE; p.ea = nx.ea; p.eb = nx.eb; p.out16 = s1; p.ld16 = C; voc_gemm(ctx, p); ++nl; }
            if (r < 2) { char nm[32]; snprintf(nm, sizeof nm, "blk%d_res%d_32", i + 1, r); dump(nm, xb, (size_t) Lo * C * 4); }
            if (r < 2) { char nm[32]; snprintf(nm, sizeof nm, "blk%d_res%d_32", i + 1, r); dump(nm, xb, (size_t) Lo * C * 4); }
            std::swap(xa, xb);
        }
        a16 = s1; L = Lo;
    }
    // ---- dec.6: causal k7 96 -> 1 (+b), tanh
    k_final_conv_tanh<<<(unsigned) ((L + 255) / 256), 256, 0, st>>>(a16, v->dec6, v->dec6_b, ctx->d_samples, (long) L); ++nl;
    rc = cuda_check(ctx, cudaGetLastError(), "vocoder launch");
    if (rc) return rc;
    if (L != n_out) return set_err(ctx, Q3T_ERR_STATE, "internal: vocoder length mismatch");
    *samples_dev = ctx->d_samples;
    ctx->last_n_samples = n_out;
    return n_out;
}

} // namespace q3t

// ------------------------------------------------------------------------------- C ABI
extern "C" int64_t q3t_vocoder_n_samples(int32_t n_frames) { return n_frames > 0 ? 1920LL * n_frames - 555 : 0; }

extern "C" int64_t q3t_vocoder_decode(q3t_ctx * ctx, const int32_t * codes, int32_t n_frames, float * samples_out, int64_t cap) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->voc_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!codes || n_frames <= 0) return set_err(ctx, Q3T_ERR_ARG, "codes is null or n_frames <= 0");
    cudaSetDevice(ctx->device);
    q3t_vocoder * v = ctx->voc;
    if (v->codes_cap < n_frames * 16) {
        cudaStreamSynchronize(ctx->stream);
        if (v->codes_dev) cudaFree(v->codes_dev);
        v->codes_dev = nullptr; v->codes_cap = 0;
        int rc = cuda_check(ctx, cudaMalloc(&v->codes_dev, (size_t) (n_frames + 64) * 16 * 4), "cudaMalloc codes");
        if (rc) return rc;
        v->codes_cap = (n_frames + 64) * 16;
    }
    int rc = cuda_check(ctx, cudaMemcpyAsync(v->codes_dev, codes, (size_t) n_frames * 16 * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D codes");
    if (rc) return rc;
    cudaEventRecord(ctx->ev[2], ctx->stream);
    float * smp = nullptr;
    const int64_t n = vocoder_run(ctx, v->codes_dev, n_frames, &smp, 0);
    if (n < 0) return n;
    cudaEventRecord(ctx->ev[3], ctx->stream);
    if (samples_out) {
        rc = cuda_check(ctx, cudaMemcpyAsync(samples_out, smp, (size_t) std::min<int64_t>(n, cap) * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H samples");
        if (rc) return rc;
    }
    rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "vocoder sync");
    if (rc) return rc;
    float ms = 0.f; cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]); ctx->timing.vocoder_ms = ms;
    return n;
}

// Streaming decode: `codes` holds all n_frames frames received so far, of which the first n_done were already
// decoded by earlier calls; returns the samples that became final: global samples [n(n_done), n(n_frames)) with
// n(T) = 1920 T - 555 (n(0) = 0).  The waveform is identical to decode() of the whole utterance (causal stack).
constexpr int VOC_STREAM_CONTEXT = 16;      // latent frames of left context re-run per chunk (the stack needs < 10)
extern "C" int64_t q3t_vocoder_decode_from(q3t_ctx * ctx, const int32_t * codes, int32_t n_frames, int32_t n_done,
                                           float * samples_out, int64_t cap) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->voc_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!codes || n_frames <= 0) return set_err(ctx, Q3T_ERR_ARG, "codes is null or n_frames <= 0");
    if (n_done < 0 || n_done >= n_frames) return set_err(ctx, Q3T_ERR_ARG, "n_done must be in [0, n_frames)");
    cudaSetDevice(ctx->device);
    q3t_vocoder * v = ctx->voc;
    if (v->codes_cap < n_frames * 16) {
        cudaStreamSynchronize(ctx->stream);
        if (v->codes_dev) cudaFree(v->codes_dev);
        v->codes_dev = nullptr; v->codes_cap = 0;
        int rc = cuda_check(ctx, cudaMalloc(&v->codes_dev, (size_t) (n_frames + 64) * 16 * 4), "cudaMalloc codes");
        if (rc) return rc;
        v->codes_cap = (n_frames + 64) * 16;
    }
    int rc = cuda_check(ctx, cudaMemcpyAsync(v->codes_dev, codes, (size_t) n_frames * 16 * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D codes");
    if (rc) return rc;
    const int t_from = std::max(0, n_done - VOC_STREAM_CONTEXT);
    cudaEventRecord(ctx->ev[2], ctx->stream);
    float * smp = nullptr;
    const int64_t n_loc = vocoder_run(ctx, v->codes_dev, n_frames, &smp, t_from);
    if (n_loc < 0) return n_loc;
    cudaEventRecord(ctx->ev[3], ctx->stream);
    const int64_t g0 = n_done > 0 ? 1920LL * n_done - 555 : 0;          // first global sample not yet emitted
    const int64_t skip = g0 - 1920LL * t_from;
    const int64_t n = n_loc - skip;
    if (samples_out) {
        rc = cuda_check(ctx, cudaMemcpyAsync(samples_out, smp + skip, (size_t) std::min<int64_t>(n, cap) * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H samples");
        if (rc) return rc;
    }
    rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "vocoder sync");
    if (rc) return rc;
    float ms = 0.f; cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]); ctx->timing.vocoder_ms = ms;
    return n;
}

extern "C" int64_t q3t_fetch_samples(q3t_ctx * ctx, float * samples_out, int64_t cap) {
    if (!ctx || !samples_out) return Q3T_ERR_ARG;
    const int64_t n = std::min<int64_t>(ctx->last_n_samples, cap);
    if (n <= 0) return 0;
    int rc = cuda_check(ctx, cudaMemcpyAsync(samples_out, ctx->d_samples, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H samples");
    if (rc) return rc;
    rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "sync");
    return rc ? rc : n;
}

#include "spk_encoder.inc"

namespace q3t { void spk_release(q3t_ctx * ctx) { spk_free(ctx); } }
