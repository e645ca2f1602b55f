// TEST INFRASTRUCTURE — not part of the product.
//
// C-ABI shim over the UNMODIFIED reference classes (qwen3_tts::TTSTransformer,
// qwen3_tts::AudioTokenizerDecoder, qwen3_tts::AudioTokenizerEncoder) so that the
// Python tests / bench `cpu_baseline` leg can drive the reference's own ggml-CPU
// path through ctypes.  It is compiled by oracle/Makefile.ref together with the
// reference sources *where they lie* under /root/reference (nothing is copied)
// into oracle/_ref/libq3t_ref.so.
//
// Reference interfaces driven here:
//   src/tts_transformer.h:190-280        (TTSTransformer public API)
//   src/audio_tokenizer_decoder.h:159-179 (AudioTokenizerDecoder public API)
//   src/audio_tokenizer_encoder.h         (speaker x-vector, used once for a fixture)
//   src/gguf_loader.cpp:27-58             (process-wide backend singleton -> thread count)
//
// `private` is opened up only so the harness can call build_prefill_graph()
// (src/tts_transformer.cpp:984-1122), which has no public entry point; access
// specifiers do not change the class layout.

#define private public
#include "tts_transformer.h"
#include "audio_tokenizer_decoder.h"
#undef private
#include "audio_tokenizer_encoder.h"
#include "gguf_loader.h"
#include "ggml-cpu.h"

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

using namespace qwen3_tts;

static std::string g_err;

extern "C" {

const char * ref_last_error() { return g_err.c_str(); }

// The CLI's -j never reaches ggml (src/main.cpp:123); the backend is a process-wide
// singleton (src/gguf_loader.cpp:27-58), so take a reference and set threads on it.
int ref_set_threads(int n) {
    ggml_backend_t be = init_preferred_backend("harness", &g_err);
    if (!be) return -1;
    ggml_backend_cpu_set_n_threads(be, n);
    // keep the extra reference for the lifetime of the process so the setting sticks
    return 0;
}

// ---------------- TTSTransformer ----------------
void * ref_tts_load(const char * path) {
    auto * t = new TTSTransformer();
    if (!t->load_model(path)) { g_err = t->get_error(); delete t; return nullptr; }
    return t;
}
void ref_tts_free(void * h) { delete (TTSTransformer *) h; }

int ref_tts_config(void * h, int32_t * out, int n) {
    const auto & c = ((TTSTransformer *) h)->get_config();
    int32_t v[] = { c.text_vocab_size, c.text_embd_dim, c.hidden_size, c.n_layers,
                    c.n_attention_heads, c.n_key_value_heads, c.intermediate_size, c.head_dim,
                    c.codec_vocab_size, c.n_codebooks, c.code_pred_layers, c.code_pred_vocab_size,
                    c.codec_pad_id, c.codec_bos_id, c.codec_eos_id };
    int m = (int) (sizeof(v) / sizeof(v[0]));
    for (int i = 0; i < n && i < m; ++i) out[i] = v[i];
    return m;
}

int ref_tts_init_kv(void * h, int32_t n_ctx) {
    auto * t = (TTSTransformer *) h;
    if (!t->init_kv_cache(n_ctx)) { g_err = t->get_error(); return -1; }
    return 0;
}

// prompt assembly: returns prefill_len, fills caller buffers (sized generously by caller)
int ref_tts_build_prompt(void * h, const int32_t * tokens, int32_t n_tokens, const float * spk,
                         int32_t language_id, float * prefill_out, int32_t * prefill_len,
                         float * trailing_out, int32_t * trailing_len, float * pad_out) {
    auto * t = (TTSTransformer *) h;
    std::vector<float> pre, trail, pad;
    if (!t->build_prefill_graph(tokens, n_tokens, spk, language_id, pre, trail, pad)) {
        g_err = t->get_error(); return -1;
    }
    const int H = t->get_config().hidden_size;
    *prefill_len = (int32_t) (pre.size() / H);
    *trailing_len = (int32_t) (trail.size() / H);
    memcpy(prefill_out, pre.data(), pre.size() * sizeof(float));
    memcpy(trailing_out, trail.data(), trail.size() * sizeof(float));
    memcpy(pad_out, pad.data(), pad.size() * sizeof(float));
    return 0;
}

int ref_tts_prefill(void * h, const float * embd, int32_t n_tokens, int32_t n_past,
                    float * hidden_out /*[n_tokens*H]*/, float * logits_out /*[V]*/) {
    auto * t = (TTSTransformer *) h;
    std::vector<float> hid, logits;
    if (!t->forward_prefill(embd, n_tokens, n_past, hid, &logits)) { g_err = t->get_error(); return -1; }
    if (hidden_out) memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    return 0;
}

int ref_tts_step(void * h, const float * embd, int32_t n_past, float * logits_out, float * hidden_out) {
    auto * t = (TTSTransformer *) h;
    std::vector<float> logits, hid;
    if (!t->forward_step(embd, n_past, logits, &hid)) { g_err = t->get_error(); return -1; }
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    if (hidden_out) memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    return 0;
}

int ref_tts_predict_codes_ar(void * h, const float * hidden, int32_t cb0, int32_t * out15,
                             float temperature, int32_t top_k) {
    auto * t = (TTSTransformer *) h;
    std::vector<int32_t> out;
    if (!t->predict_codes_autoregressive(hidden, cb0, out, temperature, top_k)) { g_err = t->get_error(); return -1; }
    memcpy(out15, out.data(), out.size() * sizeof(int32_t));
    return (int) out.size();
}

// legacy parallel-heads formulation (src/tts_transformer.cpp:2130-2180)
int ref_tts_predict_codes(void * h, const float * hidden, const int32_t * prev_codes, float * out) {
    auto * t = (TTSTransformer *) h;
    std::vector<float> o;
    if (!t->predict_codes(hidden, prev_codes, o)) { g_err = t->get_error(); return -1; }
    memcpy(out, o.data(), o.size() * sizeof(float));
    return (int) o.size();
}

int ref_tts_generate(void * h, const int32_t * tokens, int32_t n_tokens, const float * spk,
                     int32_t max_len, int32_t * codes_out, int32_t language_id,
                     float rep_penalty, float temperature, int32_t top_k) {
    auto * t = (TTSTransformer *) h;
    std::vector<int32_t> out;
    if (!t->generate(tokens, n_tokens, spk, max_len, out, language_id, rep_penalty, temperature, top_k)) {
        g_err = t->get_error(); return -1;
    }
    memcpy(codes_out, out.data(), out.size() * sizeof(int32_t));
    return (int) (out.size() / t->get_config().n_codebooks);
}

// legacy wrappers (src/tts_transformer.cpp:1963-1991, 2110-2120)
int ref_tts_forward_text(void * h, const int32_t * tokens, int32_t n_tokens, const float * spk, int32_t n_past, float * hidden_out) {
    auto * t = (TTSTransformer *) h;
    std::vector<float> hid;
    if (!t->forward_text(tokens, n_tokens, spk, n_past, hid)) { g_err = t->get_error(); return -1; }
    memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    return (int) hid.size();
}
int ref_tts_forward_codec(void * h, int32_t codec_token, int32_t n_past, float * logits_out) {
    auto * t = (TTSTransformer *) h;
    std::vector<float> logits;
    if (!t->forward_codec(codec_token, n_past, logits)) { g_err = t->get_error(); return -1; }
    memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    return (int) logits.size();
}

// ---------------- AudioTokenizerDecoder ----------------
void * ref_dec_load(const char * path) {
    auto * d = new AudioTokenizerDecoder();
    if (!d->load_model(path)) { g_err = d->get_error(); delete d; return nullptr; }
    return d;
}
void ref_dec_free(void * h) { delete (AudioTokenizerDecoder *) h; }

// returns n_samples; samples_out may be null to query the size (runs the decode either way)
long ref_dec_decode(void * h, const int32_t * codes, int32_t n_frames, float * samples_out, long cap) {
    auto * d = (AudioTokenizerDecoder *) h;
    std::vector<float> s;
    if (!d->decode(codes, n_frames, s)) { g_err = d->get_error(); return -1; }
    long n = (long) s.size();
    if (samples_out) memcpy(samples_out, s.data(), (size_t) (n < cap ? n : cap) * sizeof(float));
    return n;
}

// ---------------- speaker encoder (fixture generation only) ----------------
int ref_enc_embed(const char * model_path, const float * samples, int32_t n_samples, float * out, int cap) {
    AudioTokenizerEncoder enc;
    if (!enc.load_model(model_path)) { g_err = enc.get_error(); return -1; }
    std::vector<float> e;
    if (!enc.encode(samples, n_samples, e)) { g_err = enc.get_error(); return -1; }
    int n = (int) e.size();
    memcpy(out, e.data(), (size_t) (n < cap ? n : cap) * sizeof(float));
    return n;
}

} // extern "C"

// ---------------- debug: per-node trace of one talker step (bisecting numerics) ----------------
#include "ggml-backend.h"
struct trace_rec { std::string name; std::string op; int64_t ne[4]; int type; std::vector<float> data; };
static std::vector<trace_rec> g_trace;

static bool trace_cb(struct ggml_tensor * t, bool ask, void *) {
    if (ask) return true;
    trace_rec r;
    r.name = t->name; r.op = ggml_op_name(t->op); r.type = (int) t->type;
    for (int i = 0; i < 4; ++i) r.ne[i] = t->ne[i];
    const int64_t n = ggml_nelements(t);
    if (ggml_is_contiguous(t) && n <= (1 << 22)) {
        r.data.resize((size_t) n);
        if (t->type == GGML_TYPE_F32) memcpy(r.data.data(), t->data, (size_t) n * 4);
        else if (t->type == GGML_TYPE_F16) { const ggml_fp16_t * h = (const ggml_fp16_t *) t->data; for (int64_t i = 0; i < n; ++i) r.data[(size_t) i] = ggml_fp16_to_fp32(h[i]); }
        else r.data.clear();
    }
    g_trace.push_back(std::move(r));
    return true;
}

extern "C" {
int ref_tts_trace_step(void * h, const float * embd, int32_t n_past) {
    auto * t = (TTSTransformer *) h;
    g_trace.clear();
    ggml_backend_sched_set_eval_callback(t->state_.sched, trace_cb, nullptr);
    std::vector<float> logits, hid;
    bool ok = t->forward_step(embd, n_past, logits, &hid);
    ggml_backend_sched_set_eval_callback(t->state_.sched, nullptr, nullptr);
    if (!ok) { g_err = t->get_error(); return -1; }
    return (int) g_trace.size();
}
// per-node trace of one vocoder decode (named nodes: vq_output, pre_conv_output, pre_tfm_input, ...)
int ref_dec_trace(void * h, const int32_t * codes, int32_t n_frames) {
    auto * d = (AudioTokenizerDecoder *) h;
    g_trace.clear();
    ggml_backend_sched_set_eval_callback(d->state_.sched, trace_cb, nullptr);
    std::vector<float> s;
    bool ok = d->decode(codes, n_frames, s);
    ggml_backend_sched_set_eval_callback(d->state_.sched, nullptr, nullptr);
    if (!ok) { g_err = d->get_error(); return -1; }
    return (int) g_trace.size();
}
int ref_trace_info(int i, char * name, char * op, int64_t * ne, int * type, long * n) {
    if (i < 0 || i >= (int) g_trace.size()) return -1;
    strncpy(name, g_trace[i].name.c_str(), 63); name[63] = 0;
    strncpy(op, g_trace[i].op.c_str(), 31); op[31] = 0;
    for (int k = 0; k < 4; ++k) ne[k] = g_trace[i].ne[k];
    *type = g_trace[i].type; *n = (long) g_trace[i].data.size();
    return 0;
}
int ref_trace_data(int i, float * out) {
    if (i < 0 || i >= (int) g_trace.size()) return -1;
    memcpy(out, g_trace[i].data.data(), g_trace[i].data.size() * 4);
    return 0;
}
}
