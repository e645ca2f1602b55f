// Engine context, weight upload / private HBM repack, KV cache and the q3t_* C-ABI entry points
// for the talker + code predictor.  Replaces (reference, file:line):
//   init_preferred_backend / ggml_backend_sched_new      src/gguf_loader.cpp:27-58, src/tts_transformer.cpp:123-150
//   create_tensors / load_tensor_data                    src/tts_transformer.cpp:346-683
//   init_kv_cache / init_code_pred_kv_cache              src/tts_transformer.cpp:685-791
//   forward_prefill / forward_step / predict_codes_autoregressive / generate   :1831-2108, :2310-2851
#include "common.cuh"
#include "engine.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <chrono>

using namespace q3t;

static std::string g_create_err;

namespace q3t {

int set_err(q3t_ctx * ctx, int code, const std::string & msg) {
    if (ctx) ctx->err = msg; else g_create_err = msg;
    return code;
}

int cuda_check(q3t_ctx * ctx, cudaError_t e, const char * what) {
    if (e == cudaSuccess) return Q3T_OK;
    return set_err(ctx, Q3T_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

} // namespace q3t

#define CK(call) do { int _rc = cuda_check(ctx, (call), #call); if (_rc != Q3T_OK) return _rc; } while (0)

constexpr size_t PKT_TOTAL = QKV_N + NH * HD + H + FFN + H + 3072 + 2 * 512;   // packets of all exchange buffers

template <typename T>
static int dev_alloc(q3t_ctx * ctx, T ** out, size_t count, bool zero = false, bool owned = true) {
    void * p = nullptr;
    int rc = cuda_check(ctx, cudaMalloc(&p, std::max<size_t>(count * sizeof(T), 16)), "cudaMalloc");
    if (rc != Q3T_OK) return rc;
    if (zero) cudaMemsetAsync(p, 0, std::max<size_t>(count * sizeof(T), 16), ctx->stream);
    if (owned) ctx->owned.push_back(p);
    *out = reinterpret_cast<T *>(p);
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ context
extern "C" int q3t_ctx_create(int device, q3t_ctx ** out) {
    if (!out) return Q3T_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0)
        return set_err(nullptr, Q3T_ERR_CUDA, std::string("no CUDA device available (this engine has no CPU fallback): ") +
                                               (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0 || device >= n) return set_err(nullptr, Q3T_ERR_ARG, "device index out of range");
    q3t_ctx * ctx = new q3t_ctx();
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return set_err(nullptr, Q3T_ERR_CUDA, "cudaSetDevice failed"); }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    ctx->sm_count = prop.multiProcessorCount;
    ctx->cc_major = prop.major; ctx->cc_minor = prop.minor;
    if (prop.major != 10) {
        delete ctx;
        return set_err(nullptr, Q3T_ERR_CUDA, "this engine is compiled for sm_100a (B200) only; device reports cc " +
                                               std::to_string(prop.major) + "." + std::to_string(prop.minor));
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx; return set_err(nullptr, Q3T_ERR_CUDA, "cudaStreamCreate failed");
    }
    for (auto & ev : ctx->ev) cudaEventCreate(&ev);
    *out = ctx;
    return Q3T_OK;
}

static void free_raw(q3t_ctx * ctx, const char * prefix) {
    for (auto it = ctx->raw.begin(); it != ctx->raw.end();) {
        if (it->first.compare(0, strlen(prefix), prefix) == 0) { cudaFree(it->second.dev); it = ctx->raw.erase(it); }
        else ++it;
    }
}

static void batch_cache_free(q3t_ctx * ctx);

extern "C" int q3t_unload(q3t_ctx * ctx, int32_t which) {
    if (!ctx) return Q3T_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (which & 1) {
        batch_cache_free(ctx);
        for (void * p : ctx->owned) cudaFree(p);
        ctx->owned.clear();
        if (ctx->kv_alloc) { cudaFree(ctx->kv_alloc); ctx->kv_alloc = nullptr; }
        ctx->n_ctx = 0;
        free_raw(ctx, "talker."); free_raw(ctx, "code_pred.");
        ctx->tts_ready = false;
        ctx->hm = DevModel{};
        ctx->d_passes = nullptr; ctx->passes_cap = 0;
        ctx->d_prefill = nullptr; ctx->d_hidden_all = nullptr; ctx->d_trailing = nullptr; ctx->trailing_cap = 0;
        ctx->d_pad = nullptr; ctx->d_codes = nullptr; ctx->codes_cap = 0; ctx->d_tokens = nullptr; ctx->tokens_cap = 0;
        ctx->d_spk = nullptr; ctx->d_tmp = nullptr; ctx->tmp_cap = 0;
    }
    if (which & 2) {
        vocoder_free(ctx);
        free_raw(ctx, "tok_dec.");
        ctx->voc_ready = false;
    }
    if (which & 4) {
        spk_release(ctx);
        free_raw(ctx, "spk_enc.");
    }
    return Q3T_OK;
}

extern "C" void q3t_ctx_destroy(q3t_ctx * ctx) {
    if (!ctx) return;
    q3t_unload(ctx, 7);
    for (auto & kv : ctx->raw) cudaFree(kv.second.dev);
    ctx->raw.clear();
    for (auto & ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    if (ctx->d_prof) cudaFree(ctx->d_prof);
    if (ctx->d_prof_n) cudaFree(ctx->d_prof_n);
    if (ctx->h_done) cudaFreeHost(ctx->h_done);
    if (ctx->h_states) cudaFreeHost(ctx->h_states);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char * q3t_last_error(q3t_ctx * ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" int q3t_device_info(q3t_ctx * ctx, int32_t * sm_count, int32_t * cc_major, int32_t * cc_minor) {
    if (!ctx) return Q3T_ERR_ARG;
    if (sm_count) *sm_count = ctx->sm_count;
    if (cc_major) *cc_major = ctx->cc_major;
    if (cc_minor) *cc_minor = ctx->cc_minor;
    return Q3T_OK;
}

extern "C" int q3t_set_tts_config(q3t_ctx * ctx, const q3t_tts_config * cfg) {
    if (!ctx || !cfg) return Q3T_ERR_ARG;
    // the decode kernels are specialised on the 0.6B per-layer shapes (SURVEY.md §8 constants)
    if (cfg->hidden_size != H || cfg->n_attention_heads != NH || cfg->n_key_value_heads != NKV ||
        cfg->head_dim != HD || cfg->intermediate_size != FFN || cfg->n_codebooks != NCB)
        return set_err(ctx, Q3T_ERR_FORMAT, "unsupported model shape: kernels are built for hidden 1024, 16/8 heads x 128, ffn 3072, 16 codebooks");
    if (cfg->codec_vocab_size > 3072 || cfg->code_pred_vocab_size > 3072 || cfg->codec_vocab_size < 1024)
        return set_err(ctx, Q3T_ERR_FORMAT, "unsupported vocabulary size (max 3072)");
    if (cfg->n_layers < 1 || cfg->code_pred_layers < 1) return set_err(ctx, Q3T_ERR_FORMAT, "layer counts must be >= 1");
    ctx->cfg = *cfg;
    ctx->have_cfg = true;
    return Q3T_OK;
}

static size_t type_row_bytes(int32_t type, int64_t ne0) {
    switch (type) {
        case Q3T_TYPE_F32:  return (size_t) ne0 * 4;
        case Q3T_TYPE_F16:  return (size_t) ne0 * 2;
        case Q3T_TYPE_Q8_0: return (ne0 % 32) ? 0 : (size_t) (ne0 / 32) * 34;
        case Q3T_TYPE_Q4_K: return (ne0 % 256) ? 0 : (size_t) (ne0 / 256) * 144;
        default: return 0;
    }
}

// ---- GGUF variants that the kernels do not consume natively are converted to F16 on the host at
// load: F32 matrices (`convert_*_to_gguf.py --type f32`), Q4_K talker / code-predictor matrices
// (scripts/convert_tts_to_gguf.py:306-317) and Q8_0 vocoder matrices (scripts/convert_tokenizer_to_gguf.py:286-297).
// The reference multiplies these types with other activation roundings (F32 x F32, Q4_K x Q8_K, Q8_0 x Q8_0), so such
// files are supported within the logit / SNR tolerance, not bit for bit; F16 files and Q8_0 talker files keep the
// bit-exact native paths.
static inline float half_bits_to_float(uint16_t h) { __half v; memcpy(&v, &h, 2); return __half2float(v); }

// block_q4_K (ggml/src/ggml-common.h: 256 weights = {half d, half dmin, u8 scales[12] (8 x 6-bit scale + 8 x 6-bit min),
// u8 qs[128]}); value = d * scale * q - dmin * min per sub-block of 32
static void dequant_q4_k_row(const uint8_t * src, int64_t n, float * dst) {
    for (int64_t b = 0; b < n / 256; ++b, src += 144, dst += 256) {
        uint16_t dh, mh;
        memcpy(&dh, src, 2); memcpy(&mh, src + 2, 2);
        const float d = half_bits_to_float(dh), dmin = half_bits_to_float(mh);
        const uint8_t * sc = src + 4, * q = src + 16;
        for (int j = 0; j < 8; ++j) {
            int scale, mn;
            if (j < 4) { scale = sc[j] & 63; mn = sc[j + 4] & 63; }
            else       { scale = (sc[j + 4] & 0xF) | ((sc[j - 4] >> 6) << 4); mn = (sc[j + 4] >> 4) | ((sc[j] >> 6) << 4); }
            const float dl = d * (float) scale, ml = dmin * (float) mn;
            const uint8_t * qq = q + 32 * (j >> 1);
            for (int l = 0; l < 32; ++l) dst[32 * j + l] = dl * (float) ((j & 1) ? (qq[l] >> 4) : (qq[l] & 0xF)) - ml;
        }
    }
}
static void dequant_q8_0_row(const uint8_t * src, int64_t n, float * dst) {
    for (int64_t b = 0; b < n / 32; ++b, src += 34, dst += 32) {
        uint16_t dh; memcpy(&dh, src, 2);
        const float d = half_bits_to_float(dh);
        for (int l = 0; l < 32; ++l) dst[l] = (float) (int8_t) src[2 + l] * d;
    }
}

// does this tensor have to become F16 before the engine can use it?
static bool wants_f16_conversion(const char * name, int32_t type, const int64_t ne[4]) {
    const bool matrix = ne[1] * ne[2] * ne[3] > 1;            // 1-D tensors (norms, biases, snake parameters) stay F32
    if (!matrix) return false;
    if (type == Q3T_TYPE_F32 || type == Q3T_TYPE_Q4_K) return true;
    if (type == Q3T_TYPE_Q8_0 && !strncmp(name, "tok_dec.", 8)) return true;   // the tensor-core vocoder consumes F16 operands
    return false;
}

extern "C" int q3t_upload_tensor(q3t_ctx * ctx, const char * name, int32_t type, const int64_t ne[4],
                                 const void * data, size_t nbytes) {
    if (!ctx || !name || !ne || !data) return Q3T_ERR_ARG;
    cudaSetDevice(ctx->device);
    const bool mine = !strncmp(name, "talker.", 7) || !strncmp(name, "code_pred.", 10) || !strncmp(name, "tok_dec.", 8) ||
                      !strncmp(name, "spk_enc.", 8);
    if (!mine) return 1;       // tok_enc.* ...: not on this path (src/tts_transformer.cpp:375-377)
    for (int i = 0; i < 4; ++i) if (ne[i] <= 0 || ne[i] > (int64_t) 1 << 40) return set_err(ctx, Q3T_ERR_FORMAT, std::string("bad shape for ") + name);
    const size_t rb = type_row_bytes(type, ne[0]);
    if (rb == 0) return set_err(ctx, Q3T_ERR_FORMAT, std::string("unsupported tensor type for ") + name);
    const size_t rows = (size_t) (ne[1] * ne[2] * ne[3]);
    const size_t expect = rb * rows;
    if (expect != nbytes) return set_err(ctx, Q3T_ERR_FORMAT, std::string("size mismatch for ") + name);
    auto it = ctx->raw.find(name);
    if (it != ctx->raw.end()) { cudaFree(it->second.dev); ctx->raw.erase(it); }
    q3t_host_tensor t{};
    t.type = type; t.src_type = type; t.nbytes = nbytes;
    for (int i = 0; i < 4; ++i) t.ne[i] = ne[i];
    std::vector<__half> conv;
    if (wants_f16_conversion(name, type, ne)) {
        conv.resize((size_t) ne[0] * rows);
        std::vector<float> rowf((size_t) ne[0]);
        for (size_t r = 0; r < rows; ++r) {
            const uint8_t * src = static_cast<const uint8_t *>(data) + r * rb;
            const float * f = rowf.data();
            if (type == Q3T_TYPE_F32) f = reinterpret_cast<const float *>(src);
            else if (type == Q3T_TYPE_Q4_K) dequant_q4_k_row(src, ne[0], rowf.data());
            else dequant_q8_0_row(src, ne[0], rowf.data());
            __half * o = conv.data() + r * (size_t) ne[0];
            for (int64_t k = 0; k < ne[0]; ++k) o[k] = __float2half_rn(f[k]);
        }
        data = conv.data();
        nbytes = conv.size() * sizeof(__half);
        t.type = Q3T_TYPE_F16; t.nbytes = nbytes;
        ctx->converted_tensors++;
    }
    CK(cudaMalloc(&t.dev, std::max<size_t>(nbytes, 16)));
    cudaError_t e = cudaMemcpyAsync(t.dev, data, nbytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && !conv.empty()) e = cudaStreamSynchronize(ctx->stream);       // the staging vector dies with this call
    if (e != cudaSuccess) { cudaFree(t.dev); return cuda_check(ctx, e, "tensor upload"); }
    ctx->raw[name] = t;
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ finalize (repack)
static const q3t_host_tensor * find_raw(q3t_ctx * ctx, const std::string & name) {
    auto it = ctx->raw.find(name);
    return it == ctx->raw.end() ? nullptr : &it->second;
}

static int need(q3t_ctx * ctx, const std::string & name, int32_t type, int64_t ne0, int64_t ne1, const q3t_host_tensor ** out) {
    const q3t_host_tensor * t = find_raw(ctx, name);
    if (!t) return set_err(ctx, Q3T_ERR_FORMAT, "missing tensor: " + name);
    if (type >= 0 && t->type != type) return set_err(ctx, Q3T_ERR_FORMAT, "unexpected dtype for tensor: " + name);
    if (t->ne[0] != ne0 || t->ne[1] != ne1) return set_err(ctx, Q3T_ERR_FORMAT, "unexpected shape for tensor: " + name);
    *out = t;
    return Q3T_OK;
}

// copies `rows` rows of `row_bytes` into dst with a destination pitch (row interleave for gate/up)
static int copy_rows(q3t_ctx * ctx, void * dst, size_t dst_pitch, const void * src, size_t row_bytes, size_t rows) {
    return cuda_check(ctx, cudaMemcpy2DAsync(dst, dst_pitch, src, row_bytes, row_bytes, rows, cudaMemcpyDeviceToDevice, ctx->stream),
                      "repack copy");
}

// "chain-major" row layout of every mat-vec weight (private device layout): element k = 32j + a of
// a row (a = ggml accumulator lane, j = step of its FMA chain) moves to (j/8)*256 + a*8 + j%8, so
// that a warp reads 512 contiguous bytes and lane a receives 8 consecutive steps of ITS chain.
__global__ void k_chain_pack(const __half * __restrict__ src, __half * __restrict__ dst, size_t rows, int K) {
    const size_t n = rows * (size_t) K;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / K; const int k = (int) (i % K);
        const int j = k >> 5, a = k & 31;
        dst[r * K + (size_t) ((j >> 3) * 256 + a * 8 + (j & 7))] = src[i];
    }
}
static int chain_pack_inplace(q3t_ctx * ctx, __half * w, size_t rows, int K) {
    __half * tmp = nullptr;
    CK(cudaMalloc(&tmp, rows * (size_t) K * 2));
    k_chain_pack<<<1024, 256, 0, ctx->stream>>>(w, tmp, rows, K);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(w, tmp, rows * (size_t) K * 2, cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaFree(tmp));
    return Q3T_OK;
}

// Q8_0 rows (34-byte blocks {half d; int8 qs[32]}, ggml/src/ggml-common.h:219-224) -> ring records of
// 1024 weights: [1024 int8][32 F16 block scales] = 1088 B, K/1024 records per row, rows at
// dst_row0 + r * dst_row_step (fused q|k|v planes, gate/up interleave).
//   mode 0 (int8 dot, lm_kernel.cu gemv_q8):      value e of block b -> (b/4)*128 + (e/4)*16 + (b%4)*4 + e%4
//   mode 1 (F32 chain, lm_kernel.cu gemv_q8_f32): value e of block b -> (b/8)*256 + e*8 + b%8, stored as q + 128
//           (biased: the kernel turns the byte into a float with one byte-permute instead of an int->float conversion)
__global__ void k_q8_pack(const unsigned char * __restrict__ raw, unsigned char * __restrict__ dst, size_t n_rows, int K,
                          size_t dst_row0, int dst_row_step, int mode) {
    const int nb = K / 32;
    const size_t n = n_rows * (size_t) nb;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / nb; const int b = (int) (i % nb);
        const unsigned char * src = raw + i * 34;
        unsigned char * rec = dst + ((dst_row0 + r * (size_t) dst_row_step) * (size_t) (K / 1024) + (size_t) (b >> 5)) * 1088;
        const int bl = b & 31;
        rec[1024 + bl * 2] = src[0]; rec[1024 + bl * 2 + 1] = src[1];
        for (int e = 0; e < 32; ++e) {
            const int off = mode == 0 ? (bl >> 2) * 128 + (e >> 2) * 16 + (bl & 3) * 4 + (e & 3)
                                      : (bl >> 3) * 256 + e * 8 + (bl & 7);
            rec[off] = mode == 0 ? src[2 + e] : (unsigned char) (src[2 + e] ^ 0x80u);
        }
    }
}
static int q8_pack(q3t_ctx * ctx, const q3t_host_tensor * t, unsigned char * plane, size_t n_rows, int K,
                   size_t dst_row0, int dst_row_step, int mode) {
    k_q8_pack<<<1024, 256, 0, ctx->stream>>>((const unsigned char *) t->dev, plane, n_rows, K, dst_row0, dst_row_step, mode);
    return cuda_check(ctx, cudaGetLastError(), "q8 repack");
}

static int build_stack_q8(q3t_ctx * ctx, const char * prefix, int n_layers, StackW & W, int64_t & bytes) {
    const int32_t wt = Q3T_TYPE_Q8_0;
    const size_t REC = 1088;
    unsigned char * wqkv, * wo, * wgu, * wdown;
    float * an, * fn, * qn, * kn;
    int rc;
    if ((rc = dev_alloc(ctx, &wqkv,  (size_t) n_layers * QKV_N * REC))) return rc;
    if ((rc = dev_alloc(ctx, &wo,    (size_t) n_layers * H * 2 * REC))) return rc;
    if ((rc = dev_alloc(ctx, &wgu,   (size_t) n_layers * 2 * FFN * REC))) return rc;
    if ((rc = dev_alloc(ctx, &wdown, (size_t) n_layers * H * 3 * REC))) return rc;
    if ((rc = dev_alloc(ctx, &an, (size_t) n_layers * H))) return rc;
    if ((rc = dev_alloc(ctx, &fn, (size_t) n_layers * H))) return rc;
    if ((rc = dev_alloc(ctx, &qn, (size_t) n_layers * HD))) return rc;
    if ((rc = dev_alloc(ctx, &kn, (size_t) n_layers * HD))) return rc;
    for (int l = 0; l < n_layers; ++l) {
        const std::string b = std::string(prefix) + ".blk." + std::to_string(l) + ".";
        const q3t_host_tensor * q, * k, * v, * o, * g, * u, * d, * t;
        if ((rc = need(ctx, b + "attn_q.weight", wt, H, NH * HD, &q))) return rc;
        if ((rc = need(ctx, b + "attn_k.weight", wt, H, NKV * HD, &k))) return rc;
        if ((rc = need(ctx, b + "attn_v.weight", wt, H, NKV * HD, &v))) return rc;
        if ((rc = need(ctx, b + "attn_output.weight", wt, NH * HD, H, &o))) return rc;
        if ((rc = need(ctx, b + "ffn_gate.weight", wt, H, FFN, &g))) return rc;
        if ((rc = need(ctx, b + "ffn_up.weight", wt, H, FFN, &u))) return rc;
        if ((rc = need(ctx, b + "ffn_down.weight", wt, FFN, H, &d))) return rc;
        const size_t a0 = (size_t) l * QKV_N;
        if ((rc = q8_pack(ctx, q, wqkv, NH * HD,  H, a0, 1, 0))) return rc;
        if ((rc = q8_pack(ctx, k, wqkv, NKV * HD, H, a0 + NH * HD, 1, 0))) return rc;
        if ((rc = q8_pack(ctx, v, wqkv, NKV * HD, H, a0 + (NH + NKV) * HD, 1, 0))) return rc;
        if ((rc = q8_pack(ctx, o, wo, H, NH * HD, (size_t) l * H, 1, 0))) return rc;
        if ((rc = q8_pack(ctx, g, wgu, FFN, H, (size_t) l * 2 * FFN, 2, 0))) return rc;          // even rows: gate
        if ((rc = q8_pack(ctx, u, wgu, FFN, H, (size_t) l * 2 * FFN + 1, 2, 0))) return rc;      // odd rows: up
        if ((rc = q8_pack(ctx, d, wdown, H, FFN, (size_t) l * H, 1, 1))) return rc;
        if ((rc = need(ctx, b + "attn_norm.weight", Q3T_TYPE_F32, H, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, an + (size_t) l * H, H * 4, t->dev, H * 4, 1))) return rc;
        if ((rc = need(ctx, b + "ffn_norm.weight", Q3T_TYPE_F32, H, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, fn + (size_t) l * H, H * 4, t->dev, H * 4, 1))) return rc;
        if ((rc = need(ctx, b + "attn_q_norm.weight", Q3T_TYPE_F32, HD, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, qn + (size_t) l * HD, HD * 4, t->dev, HD * 4, 1))) return rc;
        if ((rc = need(ctx, b + "attn_k_norm.weight", Q3T_TYPE_F32, HD, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, kn + (size_t) l * HD, HD * 4, t->dev, HD * 4, 1))) return rc;
    }
    W.n_layers = n_layers; W.wtype = wt;
    W.wqkv = wqkv; W.wo = wo; W.wgu = wgu; W.wdown = wdown;
    W.sqkv = W.so = W.sgu = W.sdown = nullptr;      // block scales travel inside the records
    W.attn_norm = an; W.ffn_norm = fn; W.q_norm = qn; W.k_norm = kn;
    bytes = (int64_t) n_layers * ((int64_t) QKV_N + 2 * H + 2LL * FFN + 3 * H) * (int64_t) REC
          + (int64_t) n_layers * (2 * H + 2 * HD) * 4;
    return Q3T_OK;
}

static int build_stack(q3t_ctx * ctx, const char * prefix, int n_layers, StackW & W, int64_t & bytes) {
    // dtype of the big matrices follows the file (F16 or Q8_0); all seven must agree
    const q3t_host_tensor * probe = find_raw(ctx, std::string(prefix) + ".blk.0.attn_q.weight");
    if (!probe) return set_err(ctx, Q3T_ERR_FORMAT, std::string("missing tensor: ") + prefix + ".blk.0.attn_q.weight");
    const int32_t wt = probe->type;
    if (wt == Q3T_TYPE_Q8_0) return build_stack_q8(ctx, prefix, n_layers, W, bytes);
    if (wt != Q3T_TYPE_F16) return set_err(ctx, Q3T_ERR_FORMAT, "layer weights must be F16 or Q8_0");
    W.n_layers = n_layers; W.wtype = wt;
    __half * wqkv, * wo, * wgu, * wdown;
    float * an, * fn, * qn, * kn;
    int rc;
    if ((rc = dev_alloc(ctx, &wqkv,  (size_t) n_layers * QKV_N * H))) return rc;
    if ((rc = dev_alloc(ctx, &wo,    (size_t) n_layers * H * NH * HD))) return rc;
    if ((rc = dev_alloc(ctx, &wgu,   (size_t) n_layers * 2 * FFN * H))) return rc;
    if ((rc = dev_alloc(ctx, &wdown, (size_t) n_layers * H * FFN))) return rc;
    if ((rc = dev_alloc(ctx, &an, (size_t) n_layers * H))) return rc;
    if ((rc = dev_alloc(ctx, &fn, (size_t) n_layers * H))) return rc;
    if ((rc = dev_alloc(ctx, &qn, (size_t) n_layers * HD))) return rc;
    if ((rc = dev_alloc(ctx, &kn, (size_t) n_layers * HD))) return rc;
    for (int l = 0; l < n_layers; ++l) {
        const std::string b = std::string(prefix) + ".blk." + std::to_string(l) + ".";
        const q3t_host_tensor * q, * k, * v, * o, * g, * u, * d, * t;
        if ((rc = need(ctx, b + "attn_q.weight", wt, H, NH * HD, &q))) return rc;
        if ((rc = need(ctx, b + "attn_k.weight", wt, H, NKV * HD, &k))) return rc;
        if ((rc = need(ctx, b + "attn_v.weight", wt, H, NKV * HD, &v))) return rc;
        if ((rc = need(ctx, b + "attn_output.weight", wt, NH * HD, H, &o))) return rc;
        if ((rc = need(ctx, b + "ffn_gate.weight", wt, H, FFN, &g))) return rc;
        if ((rc = need(ctx, b + "ffn_up.weight", wt, H, FFN, &u))) return rc;
        if ((rc = need(ctx, b + "ffn_down.weight", wt, FFN, H, &d))) return rc;
        __half * dq = wqkv + (size_t) l * QKV_N * H;
        if ((rc = copy_rows(ctx, dq, H * 2, q->dev, H * 2, NH * HD))) return rc;
        if ((rc = copy_rows(ctx, dq + (size_t) NH * HD * H, H * 2, k->dev, H * 2, NKV * HD))) return rc;
        if ((rc = copy_rows(ctx, dq + (size_t) (NH + NKV) * HD * H, H * 2, v->dev, H * 2, NKV * HD))) return rc;
        if ((rc = copy_rows(ctx, wo + (size_t) l * H * NH * HD, NH * HD * 2, o->dev, NH * HD * 2, H))) return rc;
        __half * dg = wgu + (size_t) l * 2 * FFN * H;
        if ((rc = copy_rows(ctx, dg, 2 * H * 2, g->dev, H * 2, FFN))) return rc;          // even rows: gate
        if ((rc = copy_rows(ctx, dg + H, 2 * H * 2, u->dev, H * 2, FFN))) return rc;      // odd rows: up
        if ((rc = copy_rows(ctx, wdown + (size_t) l * H * FFN, FFN * 2, d->dev, FFN * 2, H))) return rc;
        if ((rc = need(ctx, b + "attn_norm.weight", Q3T_TYPE_F32, H, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, an + (size_t) l * H, H * 4, t->dev, H * 4, 1))) return rc;
        if ((rc = need(ctx, b + "ffn_norm.weight", Q3T_TYPE_F32, H, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, fn + (size_t) l * H, H * 4, t->dev, H * 4, 1))) return rc;
        if ((rc = need(ctx, b + "attn_q_norm.weight", Q3T_TYPE_F32, HD, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, qn + (size_t) l * HD, HD * 4, t->dev, HD * 4, 1))) return rc;
        if ((rc = need(ctx, b + "attn_k_norm.weight", Q3T_TYPE_F32, HD, 1, &t))) return rc;
        if ((rc = copy_rows(ctx, kn + (size_t) l * HD, HD * 4, t->dev, HD * 4, 1))) return rc;
    }
    for (int l = 0; l < n_layers; ++l) {
        if ((rc = chain_pack_inplace(ctx, wqkv + (size_t) l * QKV_N * H, QKV_N, H))) return rc;
        if ((rc = chain_pack_inplace(ctx, wo + (size_t) l * H * NH * HD, H, NH * HD))) return rc;
        if ((rc = chain_pack_inplace(ctx, wgu + (size_t) l * 2 * FFN * H, 2 * FFN, H))) return rc;
        if ((rc = chain_pack_inplace(ctx, wdown + (size_t) l * H * FFN, H, FFN))) return rc;
    }
    W.wqkv = wqkv; W.wo = wo; W.wgu = wgu; W.wdown = wdown;
    W.sqkv = W.so = W.sgu = W.sdown = nullptr;
    W.attn_norm = an; W.ffn_norm = fn; W.q_norm = qn; W.k_norm = kn;
    bytes = (int64_t) n_layers * ((int64_t) QKV_N * H + (int64_t) H * NH * HD + 2LL * FFN * H + (int64_t) H * FFN) * 2
          + (int64_t) n_layers * (2 * H + 2 * HD) * 4;
    return Q3T_OK;
}

static void drop_stack_raw(q3t_ctx * ctx, const char * prefix) {
    const std::string p = std::string(prefix) + ".blk.";
    cudaStreamSynchronize(ctx->stream);
    free_raw(ctx, p.c_str());
}

static int finalize_tts(q3t_ctx * ctx) {
    if (!ctx->have_cfg) return set_err(ctx, Q3T_ERR_STATE, "q3t_set_tts_config must precede q3t_finalize_weights");
    const q3t_tts_config & c = ctx->cfg;
    DevModel & m = ctx->hm;
    m = DevModel{};
    int rc;
    if ((rc = build_stack(ctx, "talker", c.n_layers, m.talker, ctx->talker_bytes))) return rc;
    drop_stack_raw(ctx, "talker");
    if ((rc = build_stack(ctx, "code_pred", c.code_pred_layers, m.cp, ctx->cp_bytes))) return rc;
    drop_stack_raw(ctx, "code_pred");
    if (m.talker.wtype != m.cp.wtype)
        return set_err(ctx, Q3T_ERR_FORMAT, "talker and code predictor layer weights must share one dtype (F16 or Q8_0)");
    const q3t_host_tensor * t;
    if ((rc = need(ctx, "talker.output_norm.weight", Q3T_TYPE_F32, H, 1, &t))) return rc;
    m.talker.out_norm = (const float *) t->dev;
    if ((rc = need(ctx, "code_pred.output_norm.weight", Q3T_TYPE_F32, H, 1, &t))) return rc;
    m.cp.out_norm = (const float *) t->dev;
    if ((rc = need(ctx, "talker.codec_head.weight", Q3T_TYPE_F16, H, c.codec_vocab_size, &t))) return rc;
    if ((rc = chain_pack_inplace(ctx, (__half *) t->dev, c.codec_vocab_size, H))) return rc;   // only lm_kernel reads the head
    m.codec_head.w = (const __half *) t->dev; m.codec_head.vocab = c.codec_vocab_size;
    if ((rc = need(ctx, "talker.codec_embd.weight", Q3T_TYPE_F16, H, c.codec_vocab_size, &t))) return rc;
    m.codec_embd = (const __half *) t->dev;
    for (int i = 0; i < NCB - 1; ++i) {
        if ((rc = need(ctx, "code_pred.lm_head." + std::to_string(i) + ".weight", Q3T_TYPE_F16, H, c.code_pred_vocab_size, &t))) return rc;
        if ((rc = chain_pack_inplace(ctx, (__half *) t->dev, c.code_pred_vocab_size, H))) return rc;
        m.cp_head[i].w = (const __half *) t->dev; m.cp_head[i].vocab = c.code_pred_vocab_size;
        if ((rc = need(ctx, "code_pred.codec_embd." + std::to_string(i) + ".weight", Q3T_TYPE_F16, H, c.code_pred_vocab_size, &t))) return rc;
        m.cp_embd[i] = (const __half *) t->dev;
    }
    ctx->talker_bytes += (int64_t) c.codec_vocab_size * H * 2 + H * 4;
    ctx->cp_bytes     += (int64_t) c.code_pred_vocab_size * H * 2 + H * 4 + H * 2;   // one head + norm + one embedding row per pass
    m.eps = c.rms_norm_eps;

    // code-predictor KV: 16 positions (src/tts_transformer.cpp:2336-2341) + F32 shadow of positions 0,1
    const int L2 = c.code_pred_layers;
    __half * ck, * cv; float * kf, * vf;
    if ((rc = dev_alloc(ctx, &ck, (size_t) L2 * NKV * 16 * HD, true))) return rc;
    if ((rc = dev_alloc(ctx, &cv, (size_t) L2 * NKV * 16 * HD, true))) return rc;
    if ((rc = dev_alloc(ctx, &kf, (size_t) L2 * 2 * NKV * HD, true))) return rc;
    if ((rc = dev_alloc(ctx, &vf, (size_t) L2 * 2 * NKV * HD, true))) return rc;
    m.cp.kcache = ck; m.cp.vcache = cv; m.cp.kf32 = kf; m.cp.vf32 = vf; m.cp.n_ctx = 16;

    // activations / scratch
    // all packet buffers live in one allocation so one memset re-arms them (reset_state)
    Pkt * pk = nullptr;
    if ((rc = dev_alloc(ctx, &pk, PKT_TOTAL, true))) return rc;
    m.x_qkv = pk; m.x_att = m.x_qkv + QKV_N; m.x_x1 = m.x_att + NH * HD; m.x_m = m.x_x1 + H; m.x_x = m.x_m + FFN;
    m.x_logit = m.x_x + H; m.x_cand = m.x_logit + 3072;
    if ((rc = dev_alloc(ctx, &m.x_sc, (size_t) NH * LM_ATT_MAX_KV, true))) return rc;
    {   // rows longer than this are scored by 8 CTAs per head (0 = never); Q3T_WIDE_MIN overrides for experiments
        const char * e = getenv("Q3T_WIDE_MIN");
        m.wide_min = e ? atoi(e) : 48;
    }
    if ((rc = dev_alloc(ctx, &m.hidden, H, true))) return rc;
    if ((rc = dev_alloc(ctx, &m.logits, 3072, true))) return rc;
    if ((rc = dev_alloc(ctx, &m.seen, 4096, true))) return rc;
    if ((rc = dev_alloc(ctx, &m.st, 1, true))) return rc;
    ctx->passes_cap = 64;
    if ((rc = dev_alloc(ctx, &ctx->d_passes, ctx->passes_cap))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->d_prefill, 16 * H, true))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->d_hidden_all, 64 * H, true))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->d_pad, H, true))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->d_spk, H, true))) return rc;
    ctx->tmp_cap = 8 * 3072;
    if ((rc = dev_alloc(ctx, &ctx->d_tmp, ctx->tmp_cap, true))) return rc;
    m.pad_embed = ctx->d_pad;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->tts_ready = true;
    ctx->grid = 0;
    return Q3T_OK;
}

extern "C" int q3t_finalize_weights(q3t_ctx * ctx, int32_t which) {
    if (!ctx) return Q3T_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (which == 1) return finalize_tts(ctx);
    if (which == 2) return vocoder_finalize(ctx);
    if (which == 3) return spk_finalize(ctx);
    return set_err(ctx, Q3T_ERR_ARG, "q3t_finalize_weights: which must be 1 (tts), 2 (vocoder) or 3 (speaker encoder)");
}

extern "C" int64_t q3t_weight_bytes(q3t_ctx * ctx, int32_t which) {
    if (!ctx) return 0;
    return which == 1 ? ctx->talker_bytes : which == 2 ? ctx->cp_bytes : ctx->voc_bytes;
}

// ------------------------------------------------------------------------------ KV cache + RoPE table
// K/V are F16, private layout [layer][kv_head][n_ctx][128] (the reference keeps [128][8][n_ctx],
// src/tts_transformer.cpp:714-724; a head-major layout makes a chunk of one head contiguous).
// The RoPE table reproduces ggml_rope_cache_init (ggml/src/ggml-cpu/ops.cpp:5548-5562) with the
// host libm so cos/sin are bit-identical to the oracle's: theta_i = pos * s^i by repeated multiply.
extern "C" int q3t_kv_init(q3t_ctx * ctx, int32_t n_ctx) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (n_ctx <= 0) return set_err(ctx, Q3T_ERR_ARG, "n_ctx must be > 0");
    if (n_ctx > LM_ATT_MAX_KV) return set_err(ctx, Q3T_ERR_ARG, "n_ctx exceeds the attention kernel's limit (4608)");
    cudaSetDevice(ctx->device);
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->kv_alloc) { cudaFree(ctx->kv_alloc); ctx->kv_alloc = nullptr; ctx->n_ctx = 0; }
    const int L = ctx->cfg.n_layers;
    const size_t per = (size_t) L * NKV * n_ctx * HD;                    // halves per K (or V)
    const int n_rope = std::max(n_ctx, 16);
    const size_t rope_bytes = (size_t) n_rope * HD * sizeof(float);
    // K | V | V again, dealt in 16-dim slices for the wide attention path (lm_kernel.cu attention_wide) | RoPE table
    const size_t total = per * 3 * sizeof(__half) + rope_bytes;
    CK(cudaMalloc(&ctx->kv_alloc, total));
    CK(cudaMemsetAsync(ctx->kv_alloc, 0, per * 3 * sizeof(__half), ctx->stream));
    DevModel & m = ctx->hm;
    m.talker.kcache = (__half *) ctx->kv_alloc;
    m.talker.vcache = m.talker.kcache + per;
    m.talker.vslice = m.talker.vcache + per;
    m.talker.kf32 = nullptr; m.talker.vf32 = nullptr;
    m.talker.n_ctx = n_ctx;
    float * rope_dev = (float *) (m.talker.vslice + per);
    std::vector<float> tab((size_t) n_rope * HD);
    const float theta_scale = powf(ctx->cfg.rope_theta, -2.0f / (float) HD);
    for (int p = 0; p < n_rope; ++p) {
        float theta = (float) p;
        for (int i = 0; i < HD / 2; ++i) {
            tab[(size_t) p * HD + i]          = cosf(theta);
            tab[(size_t) p * HD + HD / 2 + i] = sinf(theta);
            theta *= theta_scale;
        }
    }
    CK(cudaMemcpyAsync(rope_dev, tab.data(), rope_bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    m.rope = rope_dev; m.n_rope_pos = n_rope;
    ctx->n_ctx = n_ctx;
    return Q3T_OK;
}

extern "C" int q3t_kv_n_ctx(q3t_ctx * ctx) { return ctx ? ctx->n_ctx : 0; }

// ------------------------------------------------------------------------------ program helpers
static int reset_state(q3t_ctx * ctx, const GenState & init, bool clear_seen) {
    DevModel & m = ctx->hm;
    GenState st = init;
    st.ep_base = 0; st.pending_ep = 0;
    st.error = 0;
    CK(cudaMemsetAsync(m.x_qkv, 0, PKT_TOTAL * sizeof(Pkt), ctx->stream));     // epoch 0 = never written
    if (m.talker.vslice && m.x_sc) CK(cudaMemsetAsync(m.x_sc, 0, (size_t) NH * LM_ATT_MAX_KV * sizeof(Pkt), ctx->stream));
    if (clear_seen) CK(cudaMemsetAsync(m.seen, 0, 4096, ctx->stream));
    CK(cudaMemcpyAsync(m.st, &st, sizeof(GenState), cudaMemcpyHostToDevice, ctx->stream));
    return Q3T_OK;
}

static GenState default_state(q3t_ctx * ctx) {
    GenState st{};
    st.frame = 0; st.n_past = 0; st.max_len = 1 << 30; st.done = 0; st.trailing_len = 0;
    st.eos_id = ctx->cfg.codec_eos_id;
    st.suppress_start = ctx->cfg.codec_vocab_size - 1024;      // src/tts_transformer.cpp:2658
    st.codec_vocab = ctx->cfg.codec_vocab_size;
    st.rep_penalty = 1.0f; st.temperature = 0.0f; st.top_k = 0;
    st.seed = 0; st.draw = 0; st.pending = -1;
    return st;
}

static int upload_passes(q3t_ctx * ctx, const std::vector<PassDesc> & v, int offset = 0) {
    if ((int) v.size() + offset > ctx->passes_cap) return set_err(ctx, Q3T_ERR_ARG, "program too long");
    return cuda_check(ctx, cudaMemcpyAsync(ctx->d_passes + offset, v.data(), v.size() * sizeof(PassDesc),
                                           cudaMemcpyHostToDevice, ctx->stream), "program upload");
}

static PassDesc make_pass(int stack, int xmode, int arg, int pos, int f32attn, int head, int n_layers, int process) {
    PassDesc p{};
    p.stack = stack; p.xmode = xmode; p.arg = arg; p.pos = pos; p.f32attn = f32attn; p.head = head;
    p.n_layers = n_layers; p.process = process; p.width = 0; p.nonorm = 0;
    p.xsrc = nullptr; p.hidden_dst = nullptr; p.logits_dst = nullptr;
    return p;
}

// the 16 code-predictor passes of one frame (src/tts_transformer.cpp:2398-2569): prefill of
// [hidden, codec_embd[cb0]] at positions 0,1 (run as two causal single-token passes, F32
// attention numerics), then 14 steps at positions 2..15 with per-step input table and head.
static void append_codepred(q3t_ctx * ctx, std::vector<PassDesc> & v, int first_xmode, const float * hidden_src) {
    const int L = ctx->cfg.code_pred_layers;
    PassDesc p0 = make_pass(1, first_xmode, 0, 0, 1, -1, L, 0);
    p0.xsrc = hidden_src;
    v.push_back(p0);
    v.push_back(make_pass(1, XM_CB0_EMBD, 0, 1, 1, 1, L, 0));
    for (int s = 1; s < NCB - 1; ++s) v.push_back(make_pass(1, XM_CP_EMBD, s, s + 1, 0, s + 1, L, 0));
}

static int check_device_error(q3t_ctx * ctx) {
    GenState st;
    CK(cudaMemcpyAsync(&st, ctx->hm.st, sizeof(GenState), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (st.error) return set_err(ctx, Q3T_ERR_DEVICE, "device watchdog: a grid barrier timed out inside lm_kernel");
    return Q3T_OK;
}

static int ensure_ctx(q3t_ctx * ctx, int need_rows, int min_default) {
    // forward_prefill/forward_step lazily create a cache (src/tts_transformer.cpp:1847-1857, 2005-2015)
    if (ctx->n_ctx == 0) {
        int rc = q3t_kv_init(ctx, std::max(256, min_default));
        if (rc) return rc;
    }
    if (need_rows > ctx->n_ctx) return set_err(ctx, Q3T_ERR_CTX_FULL, "Context length exceeded");
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ talker prefill / step
static int run_talker_rows(q3t_ctx * ctx, const float * embd_dev, int n_tokens, int n_past, bool all_heads, int process,
                           float * hidden_all_dev /* [n_tokens][H] or null */) {
    const int L = ctx->cfg.n_layers;
    int done = 0;
    while (done < n_tokens) {
        const int n = std::min(n_tokens - done, ctx->passes_cap);
        std::vector<PassDesc> v;
        for (int i = 0; i < n; ++i) {
            const int r = done + i;
            const bool last = r == n_tokens - 1;
            PassDesc p = make_pass(0, XM_BUF, 0, n_past + r, 0, (all_heads || last) ? 0 : -1, L, last ? process : 0);
            p.xsrc = embd_dev + (size_t) r * H;
            // the reference scores the whole prefill block per row; a single token's row is just its n_kv entries
            // (width 0: the form the frame program uses, which may take the wide attention path on long rows)
            p.width = n_tokens == 1 ? 0 : n_past + n_tokens;
            if (hidden_all_dev && (all_heads || last)) p.hidden_dst = hidden_all_dev + (size_t) r * H;
            v.push_back(p);
        }
        int rc = upload_passes(ctx, v);
        if (rc) return rc;
        if ((rc = lm_launch(ctx, ctx->d_passes, n, 0))) return rc;
        done += n;
    }
    return Q3T_OK;
}

extern "C" int q3t_talker_prefill(q3t_ctx * ctx, const float * embd, int32_t n_tokens, int32_t n_past,
                                  float * hidden_out, float * logits_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!embd) return set_err(ctx, Q3T_ERR_ARG, "prefill_embd is null");
    if (n_tokens <= 0) return set_err(ctx, Q3T_ERR_ARG, "n_tokens must be > 0");
    if (n_past < 0) return set_err(ctx, Q3T_ERR_ARG, "n_past must be >= 0");
    cudaSetDevice(ctx->device);
    int rc = ensure_ctx(ctx, n_past + n_tokens, n_past + n_tokens + 16);
    if (rc) return rc;
    float * d_in = nullptr, * d_hid = nullptr;
    CK(cudaMalloc(&d_in, (size_t) n_tokens * H * 4));
    if ((rc = cuda_check(ctx, cudaMalloc(&d_hid, (size_t) n_tokens * H * 4), "cudaMalloc"))) { cudaFree(d_in); return rc; }
    rc = cuda_check(ctx, cudaMemcpyAsync(d_in, embd, (size_t) n_tokens * H * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D");
    if (!rc) rc = reset_state(ctx, default_state(ctx), false);
    if (!rc) rc = run_talker_rows(ctx, d_in, n_tokens, n_past, true, 0, d_hid);
    if (!rc && hidden_out) rc = cuda_check(ctx, cudaMemcpyAsync(hidden_out, d_hid, (size_t) n_tokens * H * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    if (!rc && logits_out) rc = cuda_check(ctx, cudaMemcpyAsync(logits_out, ctx->hm.logits, (size_t) ctx->cfg.codec_vocab_size * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    if (!rc) rc = check_device_error(ctx);
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_in); cudaFree(d_hid);
    return rc;
}

extern "C" int q3t_talker_step(q3t_ctx * ctx, const float * embd, int32_t n_past, float * logits_out, float * hidden_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!embd) return set_err(ctx, Q3T_ERR_ARG, "step_embd is null");
    if (n_past < 0) return set_err(ctx, Q3T_ERR_ARG, "n_past must be >= 0");
    cudaSetDevice(ctx->device);
    int rc = ensure_ctx(ctx, n_past + 1, n_past + 1 + 16);
    if (rc) return rc;
    CK(cudaMemcpyAsync(ctx->d_tmp, embd, H * 4, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = reset_state(ctx, default_state(ctx), false))) return rc;
    if ((rc = run_talker_rows(ctx, ctx->d_tmp, 1, n_past, true, 0, nullptr))) return rc;
    if (logits_out) CK(cudaMemcpyAsync(logits_out, ctx->hm.logits, (size_t) ctx->cfg.codec_vocab_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (hidden_out) CK(cudaMemcpyAsync(hidden_out, ctx->hm.hidden, H * 4, cudaMemcpyDeviceToHost, ctx->stream));
    return check_device_error(ctx);
}

// ------------------------------------------------------------------------------ code predictor, one frame
extern "C" int q3t_codepred_frame(q3t_ctx * ctx, const float * hidden, int32_t cb0, float temperature, int32_t top_k,
                                  uint64_t seed, int32_t * codes15_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!hidden || !codes15_out) return set_err(ctx, Q3T_ERR_ARG, "null argument");
    if (cb0 < 0 || cb0 >= ctx->cfg.codec_vocab_size) return set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
    cudaSetDevice(ctx->device);
    int rc;
    if (ctx->n_ctx == 0 && (rc = q3t_kv_init(ctx, 256))) return rc;     // the RoPE table lives with the KV allocation
    CK(cudaMemcpyAsync(ctx->d_tmp, hidden, H * 4, cudaMemcpyHostToDevice, ctx->stream));
    GenState st = default_state(ctx);
    st.temperature = temperature; st.top_k = top_k; st.seed = seed;
    st.cur[0] = cb0;
    if ((rc = reset_state(ctx, st, false))) return rc;
    std::vector<PassDesc> v;
    append_codepred(ctx, v, XM_BUF, ctx->d_tmp);
    v.push_back(make_pass(1, XM_SELECT_ONLY, 0, 0, 0, -1, 0, 0));
    if ((rc = upload_passes(ctx, v))) return rc;
    if ((rc = lm_launch(ctx, ctx->d_passes, (int) v.size(), 0))) return rc;
    GenState out;
    CK(cudaMemcpyAsync(&out, ctx->hm.st, sizeof(GenState), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (out.error) return set_err(ctx, Q3T_ERR_DEVICE, "device watchdog: a grid barrier timed out inside lm_kernel");
    for (int i = 0; i < NCB - 1; ++i) codes15_out[i] = out.cur[i + 1];
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ sampler in isolation
extern "C" int q3t_sample_logits(q3t_ctx * ctx, const float * logits, int32_t vocab, int32_t process,
                                 const int32_t * seen_ids, int32_t n_seen, float repetition_penalty,
                                 float temperature, int32_t top_k, uint64_t seed, int32_t n_draws, int32_t * tokens_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!logits || !tokens_out || vocab <= 0 || vocab > 3072 || n_draws <= 0) return set_err(ctx, Q3T_ERR_ARG, "q3t_sample_logits: bad argument");
    if (!(temperature > 0.0f)) return set_err(ctx, Q3T_ERR_ARG, "q3t_sample_logits: temperature must be > 0 (greedy selection is a decode-kernel path)");
    if (n_seen > 0 && !seen_ids) return set_err(ctx, Q3T_ERR_ARG, "q3t_sample_logits: seen_ids is null");
    cudaSetDevice(ctx->device);
    std::vector<Pkt> pk((size_t) vocab);
    for (int i = 0; i < vocab; ++i) { pk[i].v = logits[i]; pk[i].e = 1u; }
    std::vector<unsigned char> seen(4096, 0);
    for (int i = 0; i < n_seen; ++i) if (seen_ids[i] >= 0 && seen_ids[i] < 4096) seen[seen_ids[i]] = 1;
    Pkt * d_pk = nullptr; unsigned char * d_seen = nullptr; int * d_out = nullptr;
    int rc = cuda_check(ctx, cudaMalloc(&d_pk, pk.size() * sizeof(Pkt)), "cudaMalloc");
    if (!rc) rc = cuda_check(ctx, cudaMalloc(&d_seen, seen.size()), "cudaMalloc");
    if (!rc) rc = cuda_check(ctx, cudaMalloc(&d_out, (size_t) n_draws * 4), "cudaMalloc");
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(d_pk, pk.data(), pk.size() * sizeof(Pkt), cudaMemcpyHostToDevice, ctx->stream), "H2D");
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(d_seen, seen.data(), seen.size(), cudaMemcpyHostToDevice, ctx->stream), "H2D");
    if (!rc) rc = lm_sample_logits(ctx, d_pk, d_seen, vocab, process, repetition_penalty, temperature, top_k, seed, 0ull, n_draws, d_out);
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(tokens_out, d_out, (size_t) n_draws * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    if (!rc) rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "sample_logits_kernel");
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_pk); cudaFree(d_seen); cudaFree(d_out);
    return rc;
}

// ------------------------------------------------------------------------------ legacy predict_codes
// src/tts_transformer.cpp:2130-2180 + build_code_pred_graph :1413-1532: ONE token through the five
// code-predictor layers without RoPE or KV cache, then all 15 heads on the un-normed layer output.
// With a single key the soft-max is exactly 1.0 and the mixed value is the token's own un-rounded V
// row, which is what the F32-attention pass numerics give at position 0 (the q/k-norm and the
// rotation cannot change a one-element soft-max).  Input: hidden + sum_cb f32(code_pred_embd[cb][prev[cb]])
// added in codebook order (:1445-1451).  Not on the timed path: the stack is simply re-run per head.
extern "C" int q3t_codepred_parallel(q3t_ctx * ctx, const float * hidden, const int32_t * prev_codes, float * logits_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!hidden || !logits_out) return set_err(ctx, Q3T_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    int rc;
    if (ctx->n_ctx == 0 && (rc = q3t_kv_init(ctx, 256))) return rc;     // the RoPE table lives with the KV allocation
    const int V = ctx->cfg.code_pred_vocab_size;
    std::vector<float> x(hidden, hidden + H);
    if (prev_codes) {
        std::vector<__half> row(H);
        for (int cb = 0; cb < NCB - 1; ++cb) {
            if (prev_codes[cb] < 0 || prev_codes[cb] >= V) return set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
            CK(cudaMemcpyAsync(row.data(), ctx->hm.cp_embd[cb] + (size_t) prev_codes[cb] * H, H * 2, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            for (int k = 0; k < H; ++k) x[k] = x[k] + __half2float(row[k]);
        }
    }
    float * d_log = nullptr;
    CK(cudaMalloc(&d_log, (size_t) (NCB - 1) * V * 4));
    rc = cuda_check(ctx, cudaMemcpyAsync(ctx->d_tmp, x.data(), H * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D");
    if (!rc) rc = reset_state(ctx, default_state(ctx), false);
    if (!rc) {
        std::vector<PassDesc> v;
        for (int cb = 0; cb < NCB - 1; ++cb) {
            PassDesc p = make_pass(1, XM_BUF, 0, 0, 1, cb + 1, ctx->cfg.code_pred_layers, 0);
            p.xsrc = ctx->d_tmp; p.nonorm = 1; p.logits_dst = d_log + (size_t) cb * V;
            v.push_back(p);
        }
        rc = upload_passes(ctx, v);
        if (!rc) rc = lm_launch(ctx, ctx->d_passes, (int) v.size(), 0);
    }
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(logits_out, d_log, (size_t) (NCB - 1) * V * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    if (!rc) rc = check_device_error(ctx);
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_log);
    return rc;
}

// ------------------------------------------------------------------------------ prompt staging
static int ensure_buf_i(q3t_ctx * ctx, int ** buf, int * cap, int need_n) {
    if (*cap >= need_n) return Q3T_OK;
    int * p = nullptr;
    int rc = dev_alloc(ctx, &p, (size_t) need_n, true);
    if (rc) return rc;
    *buf = p; *cap = need_n;          // the old buffer stays in `owned` until unload (tiny)
    return Q3T_OK;
}
static int ensure_buf_f(q3t_ctx * ctx, float ** buf, int * cap, int need_n) {
    if (*cap >= need_n) return Q3T_OK;
    float * p = nullptr;
    int rc = dev_alloc(ctx, &p, (size_t) need_n, true);
    if (rc) return rc;
    *buf = p; *cap = need_n;
    return Q3T_OK;
}

extern "C" int q3t_stage_prompt(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens,
                                const float * speaker_embd, int32_t language_id) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!text_tokens) return set_err(ctx, Q3T_ERR_ARG, "text_tokens is null");
    if (n_tokens < 4) return set_err(ctx, Q3T_ERR_ARG, "Need at least 4 text tokens for generation");
    for (int i = 0; i < n_tokens; ++i)
        if (text_tokens[i] < 0 || text_tokens[i] >= ctx->cfg.text_vocab_size) return set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
    if (language_id >= ctx->cfg.codec_vocab_size) return set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
    cudaSetDevice(ctx->device);
    int rc;
    if ((rc = ensure_buf_i(ctx, &ctx->d_tokens, &ctx->tokens_cap, n_tokens + 8))) return rc;
    int trailing_rows = std::max(0, n_tokens - 9) + 1;
    int tcap = ctx->trailing_cap / H;
    if (tcap < trailing_rows) {
        int capf = ctx->trailing_cap;
        if ((rc = ensure_buf_f(ctx, &ctx->d_trailing, &capf, (trailing_rows + 16) * H))) return rc;
        ctx->trailing_cap = capf;
    }
    CK(cudaMemcpyAsync(ctx->d_tokens, text_tokens, (size_t) n_tokens * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (speaker_embd) CK(cudaMemcpyAsync(ctx->d_spk, speaker_embd, H * 4, cudaMemcpyHostToDevice, ctx->stream));
    int pl = 0, tl = 0;
    if ((rc = prompt_build(ctx, ctx->d_tokens, n_tokens, speaker_embd ? ctx->d_spk : nullptr, language_id,
                           ctx->d_prefill, &pl, ctx->d_trailing, &tl, ctx->d_pad))) return rc;
    ctx->staged_prefill_len = pl; ctx->staged_trailing_len = tl;
    ctx->hm.trailing = ctx->d_trailing; ctx->hm.pad_embed = ctx->d_pad;
    return Q3T_OK;
}

extern "C" int q3t_build_prompt(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                                int32_t language_id, float * prefill_out, int32_t * prefill_len,
                                float * trailing_out, int32_t * trailing_len, float * pad_out) {
    int rc = q3t_stage_prompt(ctx, text_tokens, n_tokens, speaker_embd, language_id);
    if (rc) return rc;
    if (prefill_len) *prefill_len = ctx->staged_prefill_len;
    if (trailing_len) *trailing_len = ctx->staged_trailing_len;
    if (prefill_out) CK(cudaMemcpyAsync(prefill_out, ctx->d_prefill, (size_t) ctx->staged_prefill_len * H * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (trailing_out) CK(cudaMemcpyAsync(trailing_out, ctx->d_trailing, (size_t) ctx->staged_trailing_len * H * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (pad_out) CK(cudaMemcpyAsync(pad_out, ctx->d_pad, H * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return Q3T_OK;
}

extern "C" int q3t_project_text(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens, float * out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (n_tokens <= 0) return Q3T_OK;
    if (!text_tokens || !out) return set_err(ctx, Q3T_ERR_ARG, "text_tokens is null");
    for (int i = 0; i < n_tokens; ++i)
        if (text_tokens[i] < 0 || text_tokens[i] >= ctx->cfg.text_vocab_size) return set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
    cudaSetDevice(ctx->device);
    int rc;
    if ((rc = ensure_buf_i(ctx, &ctx->d_tokens, &ctx->tokens_cap, n_tokens + 8))) return rc;
    float * d_out = nullptr;
    CK(cudaMalloc(&d_out, (size_t) n_tokens * H * 4));
    rc = cuda_check(ctx, cudaMemcpyAsync(ctx->d_tokens, text_tokens, (size_t) n_tokens * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D");
    if (!rc) rc = prompt_project(ctx, ctx->d_tokens, n_tokens, d_out);
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(out, d_out, (size_t) n_tokens * H * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_out);
    return rc;
}

extern "C" int q3t_codec_embed(q3t_ctx * ctx, const int32_t * ids, int32_t n, float * out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (n <= 0) return Q3T_OK;
    if (!ids || !out) return set_err(ctx, Q3T_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    std::vector<__half> row(H);
    for (int i = 0; i < n; ++i) {
        if (ids[i] < 0 || ids[i] >= ctx->cfg.codec_vocab_size) return set_err(ctx, Q3T_ERR_ARG, "Embedding token ID out of range");
        CK(cudaMemcpyAsync(row.data(), ctx->hm.codec_embd + (size_t) ids[i] * H, H * 2, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < H; ++k) out[(size_t) i * H + k] = __half2float(row[k]);
    }
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ generate
static int run_generate(q3t_ctx * ctx, int max_len, float rep_penalty, float temperature, int top_k, uint64_t seed,
                        int * n_frames_out) {
    const q3t_tts_config & c = ctx->cfg;
    const int M = ctx->staged_prefill_len;
    if (M <= 0) return set_err(ctx, Q3T_ERR_STATE, "no prompt staged");
    int rc;
    // KV sizing rule of the reference (src/tts_transformer.cpp:2630-2636)
    const int required_ctx = M + max_len + 8;
    if (ctx->n_ctx < required_ctx || ctx->n_ctx > std::max(required_ctx * 2, 512))
        if ((rc = q3t_kv_init(ctx, required_ctx))) return rc;
    if ((rc = ensure_buf_i(ctx, &ctx->d_codes, &ctx->codes_cap, max_len * NCB))) return rc;
    ctx->hm.codes_out = ctx->d_codes;

    GenState st = default_state(ctx);
    st.max_len = max_len; st.n_past = M; st.trailing_len = ctx->staged_trailing_len;
    st.rep_penalty = rep_penalty; st.temperature = temperature; st.top_k = top_k; st.seed = seed;
    if ((rc = reset_state(ctx, st, true))) return rc;

    cudaEventRecord(ctx->ev[0], ctx->stream);
    // prefill: M causal single-token talker passes in one launch; only the last needs the head
    if ((rc = run_talker_rows(ctx, ctx->d_prefill, M, 0, false, 1, nullptr))) return rc;
    cudaEventRecord(ctx->ev[1], ctx->stream);

    // frame program: 16 code-predictor passes + the talker step (src/tts_transformer.cpp:2664-2804)
    std::vector<PassDesc> v;
    append_codepred(ctx, v, XM_HIDDEN, nullptr);
    v.push_back(make_pass(0, XM_TALKER_NEXT, 0, -1, 0, 0, c.n_layers, 1));
    if ((rc = upload_passes(ctx, v))) return rc;

    // enqueue frames in batches; poll `done` one batch behind so the GPU never idles on the host
    const int BATCH = 32;
    if (!ctx->h_done) CK(cudaMallocHost(&ctx->h_done, 2 * sizeof(int)));   // pinned once: page-locking per call stalls for 100s of ms now and then
    int * h_done = ctx->h_done;
    h_done[0] = h_done[1] = 0;
    cudaEvent_t evs[2];
    cudaEventCreateWithFlags(&evs[0], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&evs[1], cudaEventDisableTiming);
    int launched = 0, batch = 0;
    bool stop = false;
    rc = Q3T_OK;
    while (launched < max_len && !stop && !rc) {
        const int n = std::min(BATCH, max_len - launched);
        for (int i = 0; i < n && !rc; ++i) rc = lm_launch(ctx, ctx->d_passes, (int) v.size(), 1);
        launched += n;
        const int slot = batch & 1;
        if (batch >= 1) {                          // wait for the previous batch's flag
            cudaEventSynchronize(evs[slot ^ 1]);
            if (h_done[slot ^ 1]) stop = true;
        }
        cudaMemcpyAsync(&h_done[slot], &ctx->hm.st->done, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
        cudaEventRecord(evs[slot], ctx->stream);
        ++batch;
    }
    cudaEventRecord(ctx->ev[2], ctx->stream);
    cudaEventDestroy(evs[0]); cudaEventDestroy(evs[1]);
    GenState out;
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(&out, ctx->hm.st, sizeof(GenState), cudaMemcpyDeviceToHost, ctx->stream), "D2H state");
    if (!rc) rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "generate sync");
    if (rc) return rc;
    if (out.error) return set_err(ctx, Q3T_ERR_DEVICE, "device watchdog: a grid barrier timed out inside lm_kernel");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]); ctx->timing.prefill_ms = ms;
    cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); ctx->timing.frames_ms = ms;
    ctx->timing.n_frames = out.frame;
    ctx->last_n_frames = out.frame;
    *n_frames_out = out.frame;
    return Q3T_OK;
}

extern "C" int q3t_generate(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                            int32_t max_len, int32_t language_id, float repetition_penalty, float temperature,
                            int32_t top_k, uint64_t seed, int32_t * codes_out, int32_t * n_frames_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!text_tokens) return set_err(ctx, Q3T_ERR_ARG, "text_tokens is null");
    if (n_tokens < 4) return set_err(ctx, Q3T_ERR_ARG, "Need at least 4 text tokens for generation");
    if (n_frames_out) *n_frames_out = 0;
    if (max_len <= 0) return Q3T_OK;                        // src/tts_transformer.cpp:2604-2607
    int rc = q3t_stage_prompt(ctx, text_tokens, n_tokens, speaker_embd, language_id);
    if (rc) return rc;
    int nf = 0;
    if ((rc = run_generate(ctx, max_len, repetition_penalty, temperature, top_k, seed, &nf))) return rc;
    if (codes_out && nf > 0) {
        CK(cudaMemcpyAsync(codes_out, ctx->d_codes, (size_t) nf * NCB * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    if (n_frames_out) *n_frames_out = nf;
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ streaming generate
// The generate loop of src/tts_transformer.cpp:2577-2851 cut into chunks of frames (the reference
// pipeline only hands codes to the vocoder when the whole utterance is done, src/qwen3_tts.cpp:345-396).  begin = prompt
// staging + prefill + frame program; next = up to max_frames more frame launches, returns the codes of the new frames.
// The codes are those of q3t_generate with the same arguments (same launches, same state), chunk size has no effect.
extern "C" int q3t_stream_begin(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                                int32_t max_len, int32_t language_id, float repetition_penalty, float temperature,
                                int32_t top_k, uint64_t seed) {
    if (!ctx) return Q3T_ERR_ARG;
    ctx->stream_max_len = 0; ctx->stream_launched = 0; ctx->stream_fetched = 0; ctx->stream_done = 1;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!text_tokens) return set_err(ctx, Q3T_ERR_ARG, "text_tokens is null");
    if (n_tokens < 4) return set_err(ctx, Q3T_ERR_ARG, "Need at least 4 text tokens for generation");
    if (max_len <= 0) return Q3T_OK;
    int rc = q3t_stage_prompt(ctx, text_tokens, n_tokens, speaker_embd, language_id);
    if (rc) return rc;
    const q3t_tts_config & c = ctx->cfg;
    const int M = ctx->staged_prefill_len;
    const int required_ctx = M + max_len + 8;
    if (ctx->n_ctx < required_ctx || ctx->n_ctx > std::max(required_ctx * 2, 512))
        if ((rc = q3t_kv_init(ctx, required_ctx))) return rc;
    if ((rc = ensure_buf_i(ctx, &ctx->d_codes, &ctx->codes_cap, max_len * NCB))) return rc;
    ctx->hm.codes_out = ctx->d_codes;
    GenState st = default_state(ctx);
    st.max_len = max_len; st.n_past = M; st.trailing_len = ctx->staged_trailing_len;
    st.rep_penalty = repetition_penalty; st.temperature = temperature; st.top_k = top_k; st.seed = seed;
    if ((rc = reset_state(ctx, st, true))) return rc;
    if ((rc = run_talker_rows(ctx, ctx->d_prefill, M, 0, false, 1, nullptr))) return rc;
    std::vector<PassDesc> v;
    append_codepred(ctx, v, XM_HIDDEN, nullptr);
    v.push_back(make_pass(0, XM_TALKER_NEXT, 0, -1, 0, 0, c.n_layers, 1));
    if ((rc = upload_passes(ctx, v))) return rc;
    ctx->stream_n_passes = (int) v.size();
    ctx->stream_max_len = max_len; ctx->stream_done = 0;
    return Q3T_OK;
}

extern "C" int q3t_stream_next(q3t_ctx * ctx, int32_t max_frames, int32_t * codes_out, int32_t * n_new_out, int32_t * done_out) {
    if (!ctx || !n_new_out || !done_out) return Q3T_ERR_ARG;
    *n_new_out = 0; *done_out = 1;
    if (ctx->stream_done || ctx->stream_max_len <= 0) return Q3T_OK;
    if (max_frames <= 0) return set_err(ctx, Q3T_ERR_ARG, "max_frames must be > 0");
    cudaSetDevice(ctx->device);
    const int n = std::min(max_frames, ctx->stream_max_len - ctx->stream_launched);
    int rc = Q3T_OK;
    for (int i = 0; i < n && !rc; ++i) rc = lm_launch(ctx, ctx->d_passes, ctx->stream_n_passes, 1);
    if (rc) return rc;
    ctx->stream_launched += n;
    GenState out;
    CK(cudaMemcpyAsync(&out, ctx->hm.st, sizeof(GenState), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (out.error) return set_err(ctx, Q3T_ERR_DEVICE, "device watchdog: a grid barrier timed out inside lm_kernel");
    const int n_new = out.frame - ctx->stream_fetched;
    if (n_new > 0 && codes_out) {
        CK(cudaMemcpyAsync(codes_out, ctx->d_codes + (size_t) ctx->stream_fetched * NCB, (size_t) n_new * NCB * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    ctx->stream_fetched = out.frame;
    ctx->last_n_frames = out.frame;
    ctx->timing.n_frames = out.frame;
    ctx->stream_done = (out.done || ctx->stream_launched >= ctx->stream_max_len) ? 1 : 0;
    *n_new_out = n_new > 0 ? n_new : 0;
    *done_out = ctx->stream_done;
    return Q3T_OK;
}

// ------------------------------------------------------------------------------ speaker encoder
extern "C" int q3t_speaker_encode(q3t_ctx * ctx, const float * samples, int32_t n_samples, float * embedding_out, int32_t * n_frames_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->spk_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (!samples || !embedding_out || n_samples <= 0) return set_err(ctx, Q3T_ERR_ARG, "samples is null or n_samples <= 0");
    cudaSetDevice(ctx->device);
    float * d_in = nullptr, * d_out = nullptr;
    int rc = cuda_check(ctx, cudaMalloc(&d_in, (size_t) n_samples * 4), "cudaMalloc");
    if (!rc) rc = cuda_check(ctx, cudaMalloc(&d_out, 1024 * 4), "cudaMalloc");
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(d_in, samples, (size_t) n_samples * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D");
    int nf = 0;
    if (!rc) rc = spk_run(ctx, d_in, n_samples, d_out, &nf);
    if (!rc) rc = cuda_check(ctx, cudaMemcpyAsync(embedding_out, d_out, 1024 * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H");
    if (!rc) rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "speaker encoder");
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_in); cudaFree(d_out);
    if (n_frames_out) *n_frames_out = nf;
    return rc;
}

// ------------------------------------------------------------------------------ batched generate
// n_seq independent utterances, greedy, in waves of Q3T_MAX_BATCH sequences (two interleaved lock-step groups of <= 8)
// through lm_batch_kernel (lm_batch.inc).  Each sequence gets its own packet block, KV caches, state,
// repetition set, trailing rows and code buffer; the prompt and the prefill of a sequence run through
// the batch-1 path into those buffers (once per utterance, ~4 ms), then the frames are batched.
struct BatchSeqBuf {
    Pkt * pk = nullptr; float * hidden = nullptr, * logits = nullptr; unsigned char * seen = nullptr;
    GenState * st = nullptr; int * codes = nullptr; float * trailing = nullptr; float * prefill = nullptr;
    __half * kv = nullptr;          // talker K then V
    __half * ck = nullptr, * cv = nullptr; float * kf = nullptr, * vf = nullptr;
    void release() {
        cudaFree(pk); cudaFree(hidden); cudaFree(logits); cudaFree(seen); cudaFree(st); cudaFree(codes);
        cudaFree(trailing); cudaFree(prefill); cudaFree(kv); cudaFree(ck); cudaFree(cv); cudaFree(kf); cudaFree(vf);
        *this = BatchSeqBuf{};
    }
};

// The buffers of a wave are kept in the context between calls (allocating and freeing ~1.3 GB of KV per call cost ~80 ms
// of the 64-utterance job); they are re-made when a call needs a longer context or more code rows.
struct BatchCache {
    BatchSeqBuf buf[Q3T_MAX_BATCH];
    float * rope = nullptr;
    int n_ctx = 0, max_len = 0;
};
static void batch_cache_free(q3t_ctx * ctx) {
    BatchCache * bc = static_cast<BatchCache *>(ctx->batch_cache);
    if (!bc) return;
    for (auto & b : bc->buf) b.release();
    cudaFree(bc->rope);
    delete bc;
    ctx->batch_cache = nullptr;
}

static int rope_table(q3t_ctx * ctx, int n_rope, float ** out_dev) {
    std::vector<float> tab((size_t) n_rope * HD);
    const float theta_scale = powf(ctx->cfg.rope_theta, -2.0f / (float) HD);
    for (int p = 0; p < n_rope; ++p) {
        float theta = (float) p;
        for (int i = 0; i < HD / 2; ++i) {
            tab[(size_t) p * HD + i]          = cosf(theta);
            tab[(size_t) p * HD + HD / 2 + i] = sinf(theta);
            theta *= theta_scale;
        }
    }
    CK(cudaMalloc(out_dev, tab.size() * sizeof(float)));
    CK(cudaMemcpyAsync(*out_dev, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return Q3T_OK;
}

extern "C" int q3t_generate_batch(q3t_ctx * ctx, int32_t n_seq, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                  const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                  float repetition_penalty, int32_t * const * codes_out, int32_t * n_frames_out) {
    return q3t_generate_batch_sampled(ctx, n_seq, text_tokens, n_tokens, speaker_embd, max_len, language_id, repetition_penalty,
                                      0.0f, 0, 0, codes_out, n_frames_out);
}

extern "C" int q3t_generate_batch_sampled(q3t_ctx * ctx, int32_t n_seq, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                          const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                          float repetition_penalty, float temperature, int32_t top_k, uint64_t seed,
                                          int32_t * const * codes_out, int32_t * n_frames_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    if (n_seq <= 0 || !text_tokens || !n_tokens || !codes_out || !n_frames_out) return set_err(ctx, Q3T_ERR_ARG, "null argument");
    for (int i = 0; i < n_seq; ++i) n_frames_out[i] = 0;
    if (max_len <= 0) return Q3T_OK;
    cudaSetDevice(ctx->device);
    const q3t_tts_config & c = ctx->cfg;
    const int L = c.n_layers, L2 = c.code_pred_layers;
    const DevModel saved = ctx->hm;
    int rc = Q3T_OK;
    BatchCache * bc = static_cast<BatchCache *>(ctx->batch_cache);
    if (bc && (bc->n_ctx < 16 + max_len + 8 || bc->max_len < max_len)) { batch_cache_free(ctx); bc = nullptr; }
    if (!bc) {
        bc = new BatchCache();
        bc->n_ctx = 16 + max_len + 8;                                // prefill is <= 10 rows (src/tts_transformer.cpp:2630)
        bc->max_len = max_len;
        ctx->batch_cache = bc;
        if ((rc = rope_table(ctx, std::max(bc->n_ctx, 16), &bc->rope))) { batch_cache_free(ctx); return rc; }
    }
    const int n_ctx = bc->n_ctx;                                     // the cached buffers' row stride
    float * rope = bc->rope;
    BatchSeqBuf * buf = bc->buf;
    DevModel views[Q3T_MAX_BATCH];
    int prefill_len[Q3T_MAX_BATCH] = {};
    // wall-clock split of the call (Q3T_BATCH_TRACE=1 prints it): setup, prompt + prefill, frames, copy-out
    const bool trace = getenv("Q3T_BATCH_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t0) { return std::chrono::duration<double, std::milli>(now() - t0).count(); };
    double w_setup = 0, w_prefill = 0, w_frames = 0, w_out = 0;
    const auto t_call = now();
    // pinned once per context: page-locking per call stalls for 100s of ms now and then
    if (!ctx->h_states && cudaMallocHost(&ctx->h_states, Q3T_MAX_BATCH * sizeof(GenState)) != cudaSuccess)
        return set_err(ctx, Q3T_ERR_CUDA, "cudaMallocHost failed");
    GenState * hst = static_cast<GenState *>(ctx->h_states);
    const size_t per = (size_t) L * NKV * n_ctx * HD;
    float ms_frames = 0.f, ms_prefill = 0.f;
    auto fail = [&](int code) {
        const auto t0 = now();
        ctx->hm = saved;
        if (code != Q3T_OK) batch_cache_free(ctx);                   // after an error nothing is assumed about the buffers
        if (trace) fprintf(stderr, "[q3t_generate_batch] %d utterances: setup %.1f ms, prompt+prefill %.1f ms, frames %.1f ms, copy-out %.1f ms, release %.1f ms, total %.1f ms\n",
                           n_seq, w_setup, w_prefill, w_frames, w_out, ms_since(t0), ms_since(t_call));
        return code;
    };

    // sequences per launch: up to 8 run as one lock-step group, up to 16 as two interleaved groups (Q3T_BATCH_WAVE: tuning knob)
    int wave = Q3T_MAX_BATCH;
    if (const char * e = getenv("Q3T_BATCH_WAVE")) wave = std::max(1, std::min(Q3T_MAX_BATCH, atoi(e)));
    for (int w0 = 0; w0 < n_seq && !rc; w0 += wave) {
        const int nb = std::min(wave, n_seq - w0);
        cudaEventRecord(ctx->ev[0], ctx->stream);
        auto t_w = now();
        for (int q = 0; q < nb && !rc; ++q) {
            BatchSeqBuf & b = buf[q];
            if (!b.pk) {
                const auto t_a = now();
                bool ok = cudaMalloc(&b.pk, PKT_TOTAL * sizeof(Pkt)) == cudaSuccess && cudaMalloc(&b.hidden, H * 4) == cudaSuccess &&
                          cudaMalloc(&b.logits, 3072 * 4) == cudaSuccess && cudaMalloc(&b.seen, 4096) == cudaSuccess &&
                          cudaMalloc(&b.st, sizeof(GenState)) == cudaSuccess && cudaMalloc(&b.codes, (size_t) bc->max_len * NCB * 4) == cudaSuccess &&
                          cudaMalloc(&b.trailing, (size_t) 4096 * H * 4) == cudaSuccess && cudaMalloc(&b.prefill, (size_t) 16 * H * 4) == cudaSuccess &&
                          cudaMalloc(&b.kv, per * 2 * sizeof(__half)) == cudaSuccess &&
                          cudaMalloc(&b.ck, (size_t) L2 * NKV * 16 * HD * 2) == cudaSuccess && cudaMalloc(&b.cv, (size_t) L2 * NKV * 16 * HD * 2) == cudaSuccess &&
                          cudaMalloc(&b.kf, (size_t) L2 * 2 * NKV * HD * 4) == cudaSuccess && cudaMalloc(&b.vf, (size_t) L2 * 2 * NKV * HD * 4) == cudaSuccess;
                if (!ok) return fail(set_err(ctx, Q3T_ERR_CUDA, "cudaMalloc failed (batch buffers)"));
                w_setup += ms_since(t_a);
            }
            const int i = w0 + q;
            if ((rc = q3t_stage_prompt(ctx, text_tokens[i], n_tokens[i], speaker_embd ? speaker_embd[i] : nullptr, language_id))) break;
            const int M = ctx->staged_prefill_len, TL = ctx->staged_trailing_len;
            if (M > 16 || TL > 4096) { rc = set_err(ctx, Q3T_ERR_ARG, "prompt too long for the batched path"); break; }
            cudaMemcpyAsync(b.trailing, ctx->d_trailing, (size_t) TL * H * 4, cudaMemcpyDeviceToDevice, ctx->stream);
            cudaMemcpyAsync(b.prefill, ctx->d_prefill, (size_t) M * H * 4, cudaMemcpyDeviceToDevice, ctx->stream);
            // (the talker K/V rows need no clearing: a row is written at its own step before any later step reads it)
            cudaMemsetAsync(b.ck, 0, (size_t) L2 * NKV * 16 * HD * 2, ctx->stream);
            cudaMemsetAsync(b.cv, 0, (size_t) L2 * NKV * 16 * HD * 2, ctx->stream);
            // view of the model for this sequence
            DevModel & m = ctx->hm;
            m = saved;
            m.x_qkv = b.pk; m.x_att = m.x_qkv + QKV_N; m.x_x1 = m.x_att + NH * HD; m.x_m = m.x_x1 + H; m.x_x = m.x_m + FFN;
            m.x_logit = m.x_x + H; m.x_cand = m.x_logit + 3072;
            m.hidden = b.hidden; m.logits = b.logits; m.seen = b.seen; m.st = b.st; m.codes_out = b.codes;
            m.trailing = b.trailing; m.pad_embed = ctx->d_pad; m.prefill_x = b.prefill;
            m.talker.vslice = nullptr;                              // the batched kernel keeps one CTA per (sequence, head)
            m.talker.kcache = b.kv; m.talker.vcache = b.kv + per; m.talker.kf32 = nullptr; m.talker.vf32 = nullptr; m.talker.n_ctx = n_ctx;
            m.cp.kcache = b.ck; m.cp.vcache = b.cv; m.cp.kf32 = b.kf; m.cp.vf32 = b.vf; m.cp.n_ctx = 16;
            m.rope = rope; m.n_rope_pos = std::max(n_ctx, 16);
            GenState st = default_state(ctx);
            st.max_len = max_len; st.n_past = M; st.trailing_len = TL;
            // utterance i draws from its own stream: seed + i (= what q3t_generate produces for it with that seed)
            st.rep_penalty = repetition_penalty; st.temperature = temperature > 0.0f ? temperature : 0.0f; st.top_k = top_k;
            st.seed = seed + (uint64_t) i;
            if ((rc = reset_state(ctx, st, true))) break;
            views[q] = m;
            prefill_len[q] = M;
        }
        // prefill: when every sequence of the wave has the same number of prompt rows (9 without, 10 with a speaker vector) the
        // rows run in lock-step through the batched kernel (M passes for the whole wave instead of M per sequence on the batch-1
        // kernel: ~14 ms instead of ~67 ms for 16 sequences); mixed waves fall back to one sequence at a time
        bool same_m = !rc;
        for (int q = 1; q < nb && same_m; ++q) same_m = prefill_len[q] == prefill_len[0];
        if (!rc && same_m) {
            const int M = prefill_len[0];
            std::vector<PassDesc> pv;
            for (int r = 0; r < M; ++r) {
                const bool last = r == M - 1;
                PassDesc p = make_pass(0, XM_BUF, r, r, 0, last ? 0 : -1, L, last ? 1 : 0);
                p.width = M;                                        // the reference scores the whole prefill block per row (run_talker_rows)
                pv.push_back(p);
            }
            if (!(rc = upload_passes(ctx, pv))) rc = lm_launch_batch(ctx, views, nb, ctx->d_passes, M);
        } else if (!rc) {
            const int saved_nctx = ctx->n_ctx;
            ctx->n_ctx = n_ctx;                                     // run_talker_rows checks the context bound
            for (int q = 0; q < nb && !rc; ++q) {
                ctx->hm = views[q];
                rc = run_talker_rows(ctx, buf[q].prefill, prefill_len[q], 0, false, 1, nullptr);
            }
            ctx->n_ctx = saved_nctx;
        }
        ctx->hm = saved;
        if (rc) break;
        cudaEventRecord(ctx->ev[1], ctx->stream);

        std::vector<PassDesc> v;
        append_codepred(ctx, v, XM_HIDDEN, nullptr);
        v.push_back(make_pass(0, XM_TALKER_NEXT, 0, -1, 0, 0, L, 1));
        if ((rc = upload_passes(ctx, v))) break;
        // frames: every launch advances all live sequences by one frame; poll `done` every 32 launches
        if (trace) cudaStreamSynchronize(ctx->stream);
        w_prefill += ms_since(t_w); t_w = now();
        int launched = 0;
        bool all_done = false;
        while (launched < max_len && !all_done && !rc) {
            const int n = std::min(32, max_len - launched);
            for (int k = 0; k < n && !rc; ++k) rc = lm_launch_batch(ctx, views, nb, ctx->d_passes, (int) v.size());
            launched += n;
            if (rc) break;
            for (int q = 0; q < nb; ++q) cudaMemcpyAsync(&hst[q], buf[q].st, sizeof(GenState), cudaMemcpyDeviceToHost, ctx->stream);
            if ((rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "batch sync"))) break;
            all_done = true;
            for (int q = 0; q < nb; ++q) {
                if (hst[q].error) rc = set_err(ctx, Q3T_ERR_DEVICE, "device watchdog: an exchange timed out inside lm_batch_kernel");
                all_done = all_done && hst[q].done;
            }
        }
        if (rc) break;
        cudaEventRecord(ctx->ev[2], ctx->stream);
        w_frames += ms_since(t_w); t_w = now();
        for (int q = 0; q < nb; ++q) {
            const int nf = hst[q].frame;
            n_frames_out[w0 + q] = nf;
            if (nf > 0 && codes_out[w0 + q])
                cudaMemcpyAsync(codes_out[w0 + q], buf[q].codes, (size_t) nf * NCB * 4, cudaMemcpyDeviceToHost, ctx->stream);
        }
        if ((rc = cuda_check(ctx, cudaStreamSynchronize(ctx->stream), "batch sync"))) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]); ms_prefill += ms;
        cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); ms_frames += ms;
        w_out += ms_since(t_w);
    }
    ctx->timing.prefill_ms = ms_prefill; ctx->timing.frames_ms = ms_frames;
    return fail(rc);
}

extern "C" int q3t_run_staged(q3t_ctx * ctx, int32_t max_len, float repetition_penalty, float temperature, int32_t top_k,
                              uint64_t seed, int32_t with_vocoder, int32_t * n_frames_out, int64_t * n_samples_out) {
    if (!ctx) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    cudaSetDevice(ctx->device);
    int nf = 0;
    int rc = run_generate(ctx, max_len, repetition_penalty, temperature, top_k, seed, &nf);
    if (rc) return rc;
    if (n_frames_out) *n_frames_out = nf;
    if (n_samples_out) *n_samples_out = 0;
    if (with_vocoder && nf > 0) {
        if (!ctx->voc_ready) return set_err(ctx, Q3T_ERR_STATE, "vocoder not loaded");
        cudaEventRecord(ctx->ev[2], ctx->stream);
        float * smp = nullptr;
        const int64_t ns = vocoder_run(ctx, ctx->d_codes, nf, &smp, 0);
        if (ns < 0) return (int) ns;
        cudaEventRecord(ctx->ev[3], ctx->stream);
        CK(cudaStreamSynchronize(ctx->stream));
        float ms = 0.f; cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]); ctx->timing.vocoder_ms = ms;
        ctx->last_n_samples = ns;
        if (n_samples_out) *n_samples_out = ns;
    }
    return Q3T_OK;
}

extern "C" int q3t_fetch_codes(q3t_ctx * ctx, int32_t * codes_out, int32_t n_frames) {
    if (!ctx || !codes_out) return Q3T_ERR_ARG;
    if (n_frames < 0) return set_err(ctx, Q3T_ERR_ARG, "n_frames must be >= 0");
    if (n_frames == 0) return Q3T_OK;
    if (n_frames > ctx->last_n_frames) return set_err(ctx, Q3T_ERR_ARG, "more frames requested than generated");
    CK(cudaMemcpyAsync(codes_out, ctx->d_codes, (size_t) n_frames * NCB * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return Q3T_OK;
}

// developer aid: arm the in-kernel phase profiler for CTA `cta` (cap records), or disarm with cap <= 0
extern "C" int q3t_profile_arm(q3t_ctx * ctx, int32_t cta, int32_t cap) {
    if (!ctx || !ctx->tts_ready) return Q3T_ERR_STATE;
    cudaSetDevice(ctx->device);
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->d_prof) { cudaFree(ctx->d_prof); cudaFree(ctx->d_prof_n); ctx->d_prof = nullptr; ctx->d_prof_n = nullptr; ctx->prof_cap = 0; }
    ctx->hm.prof = nullptr; ctx->hm.prof_n = nullptr; ctx->hm.prof_cap = 0; ctx->hm.prof_cta = 0;
    if (cap <= 0) return Q3T_OK;
    CK(cudaMalloc(&ctx->d_prof, (size_t) cap * 8 * 512));
    CK(cudaMalloc(&ctx->d_prof_n, 4 * 512));
    CK(cudaMemset(ctx->d_prof_n, 0, 4 * 512));
    ctx->prof_cap = cap;
    ctx->hm.prof = ctx->d_prof; ctx->hm.prof_n = ctx->d_prof_n; ctx->hm.prof_cap = cap; ctx->hm.prof_cta = 0; ctx->hm.prof_frame = cta;
    return Q3T_OK;
}
// copies the records (clock64 << 8 | tag) of the selected CTA for the LAST launch out; returns the count
extern "C" int q3t_profile_read(q3t_ctx * ctx, uint64_t * out, int32_t cap) {
    if (!ctx || !ctx->d_prof) return Q3T_ERR_STATE;
    if (!out || cap < 0) return Q3T_ERR_ARG;
    CK(cudaStreamSynchronize(ctx->stream));
    const int cta = ctx->hm.prof_cta;
    unsigned n = 0;
    CK(cudaMemcpy(&n, ctx->d_prof_n + cta, 4, cudaMemcpyDeviceToHost));
    if ((int) n > ctx->prof_cap) n = ctx->prof_cap;
    if ((int) n > cap) n = cap;
    if (n) CK(cudaMemcpy(out, ctx->d_prof + (size_t) cta * ctx->prof_cap, (size_t) n * 8, cudaMemcpyDeviceToHost));
    return (int) n;
}
extern "C" int q3t_profile_select(q3t_ctx * ctx, int32_t cta) {
    if (!ctx || !ctx->d_prof || cta < 0 || cta >= 512) return Q3T_ERR_STATE;
    ctx->hm.prof_cta = cta;
    return Q3T_OK;
}

extern "C" int q3t_get_timing(q3t_ctx * ctx, q3t_timing * out) {
    if (!ctx || !out) return Q3T_ERR_ARG;
    *out = ctx->timing;
    return Q3T_OK;
}

// mean device time of the talker decode step at a fixed position (roofline figure)
extern "C" int q3t_bench_talker_step(q3t_ctx * ctx, int32_t n_past, int32_t iters, float * us_per_step) {
    if (!ctx || !us_per_step || iters <= 0) return Q3T_ERR_ARG;
    if (!ctx->tts_ready) return set_err(ctx, Q3T_ERR_STATE, "Model not loaded");
    cudaSetDevice(ctx->device);
    int rc = ensure_ctx(ctx, n_past + 1, n_past + 1 + 16);
    if (rc) return rc;
    if ((rc = reset_state(ctx, default_state(ctx), false))) return rc;
    std::vector<PassDesc> v;
    PassDesc p = make_pass(0, XM_BUF, 0, n_past, 0, 0, ctx->cfg.n_layers, 0);
    p.xsrc = ctx->d_tmp;
    const int per_launch = std::min(iters, 16);
    for (int i = 0; i < per_launch; ++i) v.push_back(p);
    if ((rc = upload_passes(ctx, v))) return rc;
    if ((rc = lm_launch(ctx, ctx->d_passes, per_launch, 0))) return rc;          // warm-up
    const int launches = (iters + per_launch - 1) / per_launch;
    cudaEventRecord(ctx->ev[0], ctx->stream);
    for (int i = 0; i < launches; ++i) if ((rc = lm_launch(ctx, ctx->d_passes, per_launch, 0))) return rc;
    cudaEventRecord(ctx->ev[1], ctx->stream);
    CK(cudaStreamSynchronize(ctx->stream));
    float ms = 0.f; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    *us_per_step = ms * 1000.0f / (float) (launches * per_launch);
    return check_device_error(ctx);
}
