// On-device prompt assembly: text-embedding gather, text-projection MLP and the prefill /
// trailing row layout.  Replaces project_text_tokens, lookup_embedding_rows and the host
// assembly in build_prefill_graph (src/tts_transformer.cpp:793-982, 984-1122), which the
// reference runs as four tiny ggml graphs plus host memcpys per utterance.
#include "common.cuh"
#include "engine.h"

#include <algorithm>

namespace q3t {

// rows of the F16 text embedding table -> contiguous F16 matrix [T][K]
__global__ void gather_rows_h(const __half * __restrict__ table, const int * __restrict__ ids, int K, __half * __restrict__ out) {
    const int r = blockIdx.x;
    const uint4 * src = reinterpret_cast<const uint4 *>(table + (size_t) ids[r] * K);
    uint4 * dst = reinterpret_cast<uint4 *>(out + (size_t) r * K);
    for (int i = threadIdx.x; i < K / 8; i += blockDim.x) dst[i] = src[i];
}

// out[t][n] = act( W[n][:] . X[t][:] + bias[n] );  W, X F16 (X already F16-rounded = the oracle's
// from_float before an F16 mat-mul).  One warp per output row; lane = accumulator lane k%32 and
// the k/32 chain is walked with FMAs, then folded in the oracle's order (bit-exact with
// ggml_vec_dot_f16, see common.cuh).  8 tokens per sweep so a weight row is streamed once per 8
// tokens.  SILU=1 applies ggml_v_silu and stores the result rounded to F16 (it is the next
// mat-mul's input); SILU=0 stores F32.
template <int SILU>
__global__ void __launch_bounds__(256)
proj_rows(const __half * __restrict__ W, const __half * __restrict__ X, const float * __restrict__ bias,
          int N, int K, int T, float * __restrict__ out_f, __half * __restrict__ out_h) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * 8 + warp;
    if (n >= N) return;
    const __half * wrow = W + (size_t) n * K;
    for (int t0 = 0; t0 < T; t0 += 8) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int c = 0; c < K / 32; ++c) {
            const float wv = __half2float(wrow[c * 32 + lane]);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (t0 + j < T) acc[j] = __fmaf_rn(wv, __half2float(X[(size_t) (t0 + j) * K + c * 32 + lane]), acc[j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v = reduce32_lanes(acc[j]);
            if (lane == 0 && t0 + j < T) {
                v = __fadd_rn(v, bias[n]);
                if (SILU) {
                    v = __fdiv_rn(v, __fadd_rn(1.0f, v_expf(__fsub_rn(0.0f, v))));
                    out_h[(size_t) (t0 + j) * N + n] = __float2half_rn(v);
                } else out_f[(size_t) (t0 + j) * N + n] = v;
            }
        }
    }
}

// ---- Q8_0 text projection (fc1/fc2 quantised by scripts/convert_tts_to_gguf.py:244-270)
// activation rows -> Q8_0 (quantize_row_q8_0, AVX2 flavour: d = amax/127 as F16, q = rne(x * 127/amax));
// one warp per block of 32
template <typename T>
__global__ void quant_rows_q8(const T * __restrict__ X, int K, signed char * __restrict__ xq, float * __restrict__ xd) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const size_t row = blockIdx.x;
    for (int b = warp; b < K / 32; b += nw) {
        const float v = (float) X[row * K + b * 32 + lane];
        float amax = fabsf(v);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
        xq[row * K + b * 32 + lane] = (signed char) __float2int_rn(__fmul_rn(v, id));
        if (lane == 0) xd[row * (K / 32) + b] = __half2float(__float2half_rn(d));
    }
}

// out[t][n] = act( q8dot(W[n], X[t]) + bias[n] ) in the order of ggml_vec_dot_q8_0_q8_0 (AVX2,
// arch/x86/quants.c:1029-1046): accumulator lane j takes bytes 4j..4j+3 of every block.  One warp
// per output row: lane = (token slot g = lane/8, accumulator lane j = lane%8), 4 tokens per sweep.
// W is the raw GGUF tensor (34-byte blocks, 2-byte aligned).
template <int SILU>
__global__ void __launch_bounds__(256)
proj_rows_q8(const unsigned char * __restrict__ W, const signed char * __restrict__ xq, const float * __restrict__ xd,
             const float * __restrict__ bias, int N, int K, int T, float * __restrict__ out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 3, j = lane & 7;
    const int n = blockIdx.x * 8 + warp;
    if (n >= N) return;
    const int nb = K / 32;
    const unsigned char * wrow = W + (size_t) n * nb * 34;
    for (int t0 = 0; t0 < T; t0 += 4) {
        const int t = t0 + g < T ? t0 + g : T - 1;
        float acc = 0.0f;
        for (int b = 0; b < nb; ++b) {
            const unsigned short * wb = reinterpret_cast<const unsigned short *>(wrow + (size_t) b * 34);
            const float dw = __half2float(__ushort_as_half(wb[0]));
            const unsigned wq = (unsigned) wb[1 + 2 * j] | ((unsigned) wb[2 + 2 * j] << 16);
            const unsigned xv = *reinterpret_cast<const unsigned *>(xq + (size_t) t * K + b * 32 + 4 * j);
            const float d = __fmul_rn(dw, xd[(size_t) t * nb + b]);
            acc = __fmaf_rn(d, (float) __dp4a((int) wq, (int) xv, 0), acc);
        }
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 4));
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 2));
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 1));
        if (j == 0 && t0 + g < T) {
            float v = __fadd_rn(acc, bias[n]);
            if (SILU) v = __fdiv_rn(v, __fadd_rn(1.0f, v_expf(__fsub_rn(0.0f, v))));
            out[(size_t) (t0 + g) * N + n] = v;
        }
    }
}

struct AsmArgs {
    const float * proj;       // [7 + tc][H]: bos, eos, pad, role0..2, first_text, trailing...
    const __half * codec_embd;
    const float * spk;        // may be null
    int codec_ids[6];         // prefill codec tokens then pad, bos
    int n_codec_prefill;      // 3 or 4
    int has_spk;
    int prefill_len, trailing_len;
    float * prefill, * trailing, * pad;
};

// one block per output row: prefill rows, then trailing rows, then the pad row
__global__ void assemble_prompt(const AsmArgs a) {
    const int row = blockIdx.x;
    const int L = a.n_codec_prefill + a.has_spk + 2;            // codec_input_len
    for (int h = threadIdx.x; h < H; h += blockDim.x) {
        auto codec_in = [&](int t) -> float {                   // codec_input_embedding[t][h]
            if (t < a.n_codec_prefill) return __half2float(a.codec_embd[(size_t) a.codec_ids[t] * H + h]);
            if (a.has_spk && t == a.n_codec_prefill) return a.spk[h];
            const int k = t - a.n_codec_prefill - a.has_spk;    // 0: codec_pad, 1: codec_bos
            return __half2float(a.codec_embd[(size_t) a.codec_ids[a.n_codec_prefill + k] * H + h]);
        };
        if (row < a.prefill_len) {
            float v;
            if (row < 3) v = a.proj[(size_t) (3 + row) * H + h];                       // role tokens
            else if (row < a.prefill_len - 1) {
                const int t = row - 3;                                                 // overlay rows
                const float ov = (t == L - 2) ? a.proj[0 * H + h] : a.proj[2 * H + h]; // tts_bos : tts_pad
                v = __fadd_rn(ov, codec_in(t));
            } else v = __fadd_rn(a.proj[(size_t) 6 * H + h], codec_in(L - 1));         // first text + codec_bos
            a.prefill[(size_t) row * H + h] = v;
        } else if (row < a.prefill_len + a.trailing_len) {
            const int t = row - a.prefill_len;
            a.trailing[(size_t) t * H + h] = (t == a.trailing_len - 1) ? a.proj[1 * H + h]      // tts_eos
                                                                       : a.proj[(size_t) (7 + t) * H + h];
        } else {
            a.pad[h] = a.proj[2 * H + h];
        }
    }
}

__global__ void fill_ids(int * dst, const int * __restrict__ tokens, int n_tokens, int bos, int eos, int pad, int tc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) dst[0] = bos;
    if (i == 1) dst[1] = eos;
    if (i == 2) dst[2] = pad;
    if (i >= 3 && i < 7) dst[i] = tokens[i - 3];
    if (i >= 7 && i < 7 + tc) dst[i] = tokens[4 + (i - 7)];
}

static int project_ids(q3t_ctx * ctx, const int * ids_dev, int T, float * out_dev) {
    const q3t_tts_config & c = ctx->cfg;
    const int K = c.text_embd_dim;
    auto get = [&](const char * n) -> const q3t_host_tensor * {
        auto it = ctx->raw.find(n);
        return it == ctx->raw.end() ? nullptr : &it->second;
    };
    const q3t_host_tensor * emb = get("talker.text_embd.weight");
    const q3t_host_tensor * w1 = get("talker.text_proj.fc1.weight"), * b1 = get("talker.text_proj.fc1.bias");
    const q3t_host_tensor * w2 = get("talker.text_proj.fc2.weight"), * b2 = get("talker.text_proj.fc2.bias");
    if (!emb || !w1 || !b1 || !w2 || !b2) return set_err(ctx, Q3T_ERR_FORMAT, "missing text embedding / projection tensors");
    const bool q8 = w1->type == Q3T_TYPE_Q8_0;
    if (emb->type != Q3T_TYPE_F16 || (w1->type != Q3T_TYPE_F16 && !q8) || w2->type != w1->type)
        return set_err(ctx, Q3T_ERR_FORMAT, "text projection: text_embd must be F16, fc1/fc2 both F16 or both Q8_0");
    if (K % 8 != 0 || emb->ne[0] != K || w1->ne[0] != K || w1->ne[1] != K || w2->ne[0] != K || w2->ne[1] != H)
        return set_err(ctx, Q3T_ERR_FORMAT, "unexpected text projection shapes");
    __half * xh = nullptr, * hh = nullptr;
    int rc = cuda_check(ctx, cudaMalloc(&xh, (size_t) T * K * 2), "cudaMalloc");
    if (rc) return rc;
    rc = cuda_check(ctx, cudaMalloc(&hh, (size_t) T * K * 2), "cudaMalloc");
    if (rc) { cudaFree(xh); return rc; }
    gather_rows_h<<<T, 128, 0, ctx->stream>>>((const __half *) emb->dev, ids_dev, K, xh);
    signed char * xq = nullptr; float * xd = nullptr, * hf = nullptr;
    if (!q8) {
        proj_rows<1><<<(K + 7) / 8, 256, 0, ctx->stream>>>((const __half *) w1->dev, xh, (const float *) b1->dev, K, K, T, nullptr, hh);
        proj_rows<0><<<(H + 7) / 8, 256, 0, ctx->stream>>>((const __half *) w2->dev, hh, (const float *) b2->dev, H, K, T, out_dev, nullptr);
        ctx->timing.kernel_launches += 3;
    } else {
        rc = cuda_check(ctx, cudaMalloc(&xq, (size_t) T * K), "cudaMalloc");
        if (!rc) rc = cuda_check(ctx, cudaMalloc(&xd, (size_t) T * (K / 32) * 4), "cudaMalloc");
        if (!rc) rc = cuda_check(ctx, cudaMalloc(&hf, (size_t) T * K * 4), "cudaMalloc");
        if (!rc) {
            quant_rows_q8<__half><<<T, 256, 0, ctx->stream>>>(xh, K, xq, xd);
            proj_rows_q8<1><<<(K + 7) / 8, 256, 0, ctx->stream>>>((const unsigned char *) w1->dev, xq, xd, (const float *) b1->dev, K, K, T, hf);
            quant_rows_q8<float><<<T, 256, 0, ctx->stream>>>(hf, K, xq, xd);
            proj_rows_q8<0><<<(H + 7) / 8, 256, 0, ctx->stream>>>((const unsigned char *) w2->dev, xq, xd, (const float *) b2->dev, H, K, T, out_dev);
            ctx->timing.kernel_launches += 5;
        }
    }
    if (!rc) rc = cuda_check(ctx, cudaGetLastError(), "text projection launch");
    cudaStreamSynchronize(ctx->stream);
    cudaFree(xh); cudaFree(hh); cudaFree(xq); cudaFree(xd); cudaFree(hf);
    return rc;
}

int prompt_project(q3t_ctx * ctx, const int32_t * tokens_dev, int n_tokens, float * out_dev) {
    return project_ids(ctx, tokens_dev, n_tokens, out_dev);
}

int prompt_build(q3t_ctx * ctx, const int32_t * tokens_dev, int n_tokens, const float * spk_dev, int language_id,
                 float * prefill_dev, int * prefill_len, float * trailing_dev, int * trailing_len, float * pad_dev) {
    const q3t_tts_config & c = ctx->cfg;
    const int tc = std::max(0, n_tokens - 9);                   // trailing_token_count (:1105)
    const int T = 7 + tc;
    int * ids = nullptr; float * proj = nullptr;
    int rc = cuda_check(ctx, cudaMalloc(&ids, (size_t) T * 4), "cudaMalloc");
    if (rc) return rc;
    rc = cuda_check(ctx, cudaMalloc(&proj, (size_t) T * H * 4), "cudaMalloc");
    if (rc) { cudaFree(ids); return rc; }
    fill_ids<<<(T + 127) / 128, 128, 0, ctx->stream>>>(ids, tokens_dev, n_tokens, c.tts_bos_token_id, c.tts_eos_token_id,
                                                       c.tts_pad_token_id, tc);
    rc = project_ids(ctx, ids, T, proj);
    if (!rc) {
        AsmArgs a{};
        a.proj = proj; a.codec_embd = ctx->hm.codec_embd; a.spk = spk_dev; a.has_spk = spk_dev ? 1 : 0;
        if (language_id < 0) {
            a.n_codec_prefill = 3;
            a.codec_ids[0] = c.codec_nothink_id; a.codec_ids[1] = c.codec_think_bos_id; a.codec_ids[2] = c.codec_think_eos_id;
        } else {
            a.n_codec_prefill = 4;
            a.codec_ids[0] = c.codec_think_id; a.codec_ids[1] = c.codec_think_bos_id;
            a.codec_ids[2] = language_id;      a.codec_ids[3] = c.codec_think_eos_id;
        }
        a.codec_ids[a.n_codec_prefill] = c.codec_pad_id; a.codec_ids[a.n_codec_prefill + 1] = c.codec_bos_id;
        const int L = a.n_codec_prefill + a.has_spk + 2;
        a.prefill_len = 3 + (L - 1) + 1;
        a.trailing_len = tc + 1;
        a.prefill = prefill_dev; a.trailing = trailing_dev; a.pad = pad_dev;
        assemble_prompt<<<a.prefill_len + a.trailing_len + 1, 256, 0, ctx->stream>>>(a);
        ctx->timing.kernel_launches += 2;
        rc = cuda_check(ctx, cudaGetLastError(), "assemble_prompt launch");
        *prefill_len = a.prefill_len; *trailing_len = a.trailing_len;
    }
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ids); cudaFree(proj);
    return rc;
}


// ------------------------------------------------------------------------------------------------
// q3t_project_text_static_dev: the text projection above for callers that cannot allocate (CUDA graph
// capture, tight serving loops).  No counterpart in the reference.
// The intermediates live in ONE file-scope device array instead of cudaMalloc'd buffers:
// every GPU that loads the library holds STATIC_SCRATCH_BYTES (1.14 MiB) until exit, all calls on a
// device share it (it keeps the last call's contents), so calls that overlap on the GPU race.
constexpr int    STATIC_MAX_TOKENS = 64;                 // rows per call
constexpr int    STATIC_MAX_K      = 2048;               // text_embd_dim of the 0.6B model
constexpr size_t SCR_XH   = 0;                                                      // half  [64][2048] gathered rows
constexpr size_t SCR_HH   = SCR_XH + (size_t) STATIC_MAX_TOKENS * STATIC_MAX_K * 2; // half  [64][2048] silu(fc1)
constexpr size_t SCR_HF   = SCR_HH + (size_t) STATIC_MAX_TOKENS * STATIC_MAX_K * 2; // float [64][2048] silu(fc1), Q8_0 path
constexpr size_t SCR_XQ   = SCR_HF + (size_t) STATIC_MAX_TOKENS * STATIC_MAX_K * 4; // int8  [64][2048] quantised rows
constexpr size_t SCR_XD   = SCR_XQ + (size_t) STATIC_MAX_TOKENS * STATIC_MAX_K;     // float [64][64]   block scales
constexpr size_t SCR_FLAG = SCR_XD + (size_t) STATIC_MAX_TOKENS * (STATIC_MAX_K / 32) * 4;  // int: bad-token flag
constexpr size_t STATIC_SCRATCH_BYTES = SCR_FLAG + 256;
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

// In host code `g_static_scratch` is only a handle: the device address is asked for once per device
// (cudaGetSymbolAddress with that device current) and kept here.
constexpr int STATIC_MAX_DEVICES = 64;
static unsigned char * g_static_scratch_ptr[STATIC_MAX_DEVICES] = {};

static int static_scratch_for(q3t_ctx * ctx, unsigned char ** out) {
    if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) return set_err(ctx, Q3T_ERR_ARG, "device index out of range for the static scratch");
    int rc = cuda_check(ctx, cudaSetDevice(ctx->device), "cudaSetDevice");
    if (rc) return rc;
    if (!g_static_scratch_ptr[ctx->device]) {
        void * p = nullptr;
        rc = cuda_check(ctx, cudaGetSymbolAddress(&p, g_static_scratch), "cudaGetSymbolAddress(g_static_scratch)");
        if (rc) return rc;
        if (!p) return set_err(ctx, Q3T_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        g_static_scratch_ptr[ctx->device] = (unsigned char *) p;
    }
    *out = g_static_scratch_ptr[ctx->device];
    return Q3T_OK;
}

// gather with a range test (the ids are device memory, the host cannot validate them): a bad id gives a
// zero row and raises the flag
__global__ void gather_rows_h_checked(const __half * __restrict__ table, const int * __restrict__ ids, int K, int vocab,
                                      __half * __restrict__ out, int * __restrict__ flag) {
    const int r = blockIdx.x;
    const int id = ids[r];
    uint4 * dst = reinterpret_cast<uint4 *>(out + (size_t) r * K);
    if (id < 0 || id >= vocab) {
        if (threadIdx.x == 0) atomicOr(flag, 1);
        for (int i = threadIdx.x; i < K / 8; i += blockDim.x) dst[i] = make_uint4(0u, 0u, 0u, 0u);
        return;
    }
    const uint4 * src = reinterpret_cast<const uint4 *>(table + (size_t) id * K);
    for (int i = threadIdx.x; i < K / 8; i += blockDim.x) dst[i] = src[i];
}

} // namespace q3t

extern "C" int q3t_project_text_static_max_tokens(void) { return q3t::STATIC_MAX_TOKENS; }

extern "C" int q3t_project_text_static_dev(q3t_ctx * ctx, const int32_t * text_tokens_dev, int32_t n_tokens, float * out_dev) {
    using namespace q3t;
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (n_tokens <= 0) return Q3T_OK;
    if (!text_tokens_dev || !out_dev) return set_err(ctx, Q3T_ERR_ARG, "text_tokens is null");
    if (n_tokens > STATIC_MAX_TOKENS)
        return set_err(ctx, Q3T_ERR_ARG, "q3t_project_text_static_dev: n_tokens exceeds the static scratch (max 64 tokens per call)");
    const q3t_tts_config & c = ctx->cfg;
    const int K = c.text_embd_dim, T = n_tokens;
    if (K > STATIC_MAX_K || K % 32 != 0)
        return set_err(ctx, Q3T_ERR_ARG, "q3t_project_text_static_dev: text_embd_dim exceeds the static scratch (max 2048)");
    auto get = [&](const char * n) -> const q3t_host_tensor * {
        auto it = ctx->raw.find(n);
        return it == ctx->raw.end() ? nullptr : &it->second;
    };
    const q3t_host_tensor * emb = get("talker.text_embd.weight");
    const q3t_host_tensor * w1 = get("talker.text_proj.fc1.weight"), * b1 = get("talker.text_proj.fc1.bias");
    const q3t_host_tensor * w2 = get("talker.text_proj.fc2.weight"), * b2 = get("talker.text_proj.fc2.bias");
    if (!emb || !w1 || !b1 || !w2 || !b2) return set_err(ctx, Q3T_ERR_FORMAT, "missing text embedding / projection tensors");
    const bool q8 = w1->type == Q3T_TYPE_Q8_0;
    if (emb->type != Q3T_TYPE_F16 || (w1->type != Q3T_TYPE_F16 && !q8) || w2->type != w1->type)
        return set_err(ctx, Q3T_ERR_FORMAT, "text projection: text_embd must be F16, fc1/fc2 both F16 or both Q8_0");
    if (emb->ne[0] != K || w1->ne[0] != K || w1->ne[1] != K || w2->ne[0] != K || w2->ne[1] != H)
        return set_err(ctx, Q3T_ERR_FORMAT, "unexpected text projection shapes");

    unsigned char * scr = nullptr;
    int rc = static_scratch_for(ctx, &scr);          // device address of g_static_scratch on ctx->device
    if (rc) return rc;
    __half * xh = (__half *) (scr + SCR_XH), * hh = (__half *) (scr + SCR_HH);
    float * hf = (float *) (scr + SCR_HF), * xd = (float *) (scr + SCR_XD);
    signed char * xq = (signed char *) (scr + SCR_XQ);
    int * flag = (int *) (scr + SCR_FLAG);
    // the array keeps the previous call's contents; every buffer is fully overwritten for the T rows in
    // use before it is read, only the flag has to start at zero
    rc = cuda_check(ctx, cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream), "cudaMemsetAsync");
    if (rc) return rc;
    gather_rows_h_checked<<<T, 128, 0, ctx->stream>>>((const __half *) emb->dev, text_tokens_dev, K, (int) emb->ne[1], xh, flag);
    if (!q8) {
        proj_rows<1><<<(K + 7) / 8, 256, 0, ctx->stream>>>((const __half *) w1->dev, xh, (const float *) b1->dev, K, K, T, nullptr, hh);
        proj_rows<0><<<(H + 7) / 8, 256, 0, ctx->stream>>>((const __half *) w2->dev, hh, (const float *) b2->dev, H, K, T, out_dev, nullptr);
        ctx->timing.kernel_launches += 3;
    } else {
        quant_rows_q8<__half><<<T, 256, 0, ctx->stream>>>(xh, K, xq, xd);
        proj_rows_q8<1><<<(K + 7) / 8, 256, 0, ctx->stream>>>((const unsigned char *) w1->dev, xq, xd, (const float *) b1->dev, K, K, T, hf);
        quant_rows_q8<float><<<T, 256, 0, ctx->stream>>>(hf, K, xq, xd);
        proj_rows_q8<0><<<(H + 7) / 8, 256, 0, ctx->stream>>>((const unsigned char *) w2->dev, xq, xd, (const float *) b2->dev, H, K, T, out_dev);
        ctx->timing.kernel_launches += 5;
    }
    rc = cuda_check(ctx, cudaGetLastError(), "text projection launch");
    int bad = 0;
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(&bad, flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    if (!rc) rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "cudaStreamSynchronize");
    if (!rc && bad) rc = set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
    return rc;
}
