// C view of the host classes (second half of include/q3t.h): what a ctypes / cgo / JNI binding
// of the reference's TTSTransformer / AudioTokenizerDecoder / GGUFLoader would call.
#include "../../include/q3t.h"
#include "audio_tokenizer_decoder.h"
#include "audio_tokenizer_encoder.h"
#include "b200_gguf_loader.h"
#include "b200_pipeline.h"
#include "tts_transformer.h"

#include <cstring>
#include <thread>
#include <vector>

using namespace qwen3_tts;
using qwen3_tts::b200::GGUFLoader;

struct q3t_gguf { GGUFLoader l; std::string err; };
struct q3t_tts  { TTSTransformer t; };
struct q3t_dec  { AudioTokenizerDecoder d; };
struct q3t_enc  { AudioTokenizerEncoder e; };

extern "C" {

// ---------------------------------------------------------------- GGUFLoader
Q3T_API q3t_gguf * q3t_gguf_open(const char * path) {
    static thread_local std::string last;
    q3t_gguf * g = new q3t_gguf();
    if (!path || !g->l.open(path)) { g->err = g->l.get_error(); }
    return g;                        // caller checks q3t_gguf_get_error(); empty string = ok
}
Q3T_API void q3t_gguf_close(q3t_gguf * g) { delete g; }
Q3T_API const char * q3t_gguf_get_error(q3t_gguf * g) { return g ? g->err.c_str() : "null loader"; }
Q3T_API int64_t q3t_gguf_get_n_tensors(q3t_gguf * g) { return g ? g->l.get_n_tensors() : 0; }
Q3T_API const char * q3t_gguf_get_tensor_name(q3t_gguf * g, int64_t idx) { return g ? g->l.get_tensor_name(idx) : nullptr; }
Q3T_API int32_t q3t_gguf_get_tensor_type(q3t_gguf * g, int64_t idx) { return g ? g->l.get_tensor_type(idx) : 0; }
Q3T_API size_t q3t_gguf_get_tensor_offset(q3t_gguf * g, int64_t idx) { return g ? g->l.get_tensor_offset(idx) : 0; }
Q3T_API size_t q3t_gguf_get_tensor_size(q3t_gguf * g, int64_t idx) { return g ? g->l.get_tensor_size(idx) : 0; }
Q3T_API int32_t q3t_gguf_get_tensor_ne(q3t_gguf * g, int64_t idx, int64_t ne_out[4]) { return (g && g->l.get_tensor_ne(idx, ne_out)) ? 0 : -1; }
Q3T_API int32_t q3t_gguf_get_u32(q3t_gguf * g, const char * key, int32_t d) { return g ? g->l.get_u32(key, d) : d; }
Q3T_API float q3t_gguf_get_f32(q3t_gguf * g, const char * key, float d) { return g ? g->l.get_f32(key, d) : d; }
Q3T_API size_t q3t_gguf_get_data_offset(q3t_gguf * g) { return g ? g->l.get_data_offset() : 0; }

// ---------------------------------------------------------------- TTSTransformer
Q3T_API q3t_tts * q3t_tts_new(void) { return new q3t_tts(); }
Q3T_API void q3t_tts_delete(q3t_tts * t) { delete t; }
Q3T_API int  q3t_tts_load_model(q3t_tts * t, const char * p) { return t && p && t->t.load_model(p); }
Q3T_API void q3t_tts_unload_model(q3t_tts * t) { if (t) t->t.unload_model(); }
Q3T_API int  q3t_tts_init_kv_cache(q3t_tts * t, int32_t n_ctx) { return t && t->t.init_kv_cache(n_ctx); }
Q3T_API void q3t_tts_clear_kv_cache(q3t_tts * t) { if (t) t->t.clear_kv_cache(); }

Q3T_API int q3t_tts_forward_prefill(q3t_tts * t, const float * e, int32_t n_tokens, int32_t n_past,
                                    float * hidden_out, float * logits_out) {
    if (!t) return 0;
    std::vector<float> hid, logits;
    if (!t->t.forward_prefill(e, n_tokens, n_past, hid, logits_out ? &logits : nullptr)) return 0;
    if (hidden_out) memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    return 1;
}
Q3T_API int q3t_tts_forward_step(q3t_tts * t, const float * e, int32_t n_past, float * logits_out, float * hidden_out) {
    if (!t) return 0;
    std::vector<float> logits, hid;
    if (!t->t.forward_step(e, n_past, logits, &hid)) return 0;
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    if (hidden_out) memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    return 1;
}
Q3T_API int q3t_tts_forward_text(q3t_tts * t, const int32_t * tokens, int32_t n_tokens, const float * spk,
                                 int32_t n_past, float * hidden_out) {
    if (!t) return 0;
    std::vector<float> hid;
    if (!t->t.forward_text(tokens, n_tokens, spk, n_past, hid)) return 0;
    if (hidden_out) memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    return 1;
}
Q3T_API int q3t_tts_forward_codec(q3t_tts * t, int32_t tok, int32_t n_past, float * logits_out) {
    if (!t) return 0;
    std::vector<float> logits;
    if (!t->t.forward_codec(tok, n_past, logits)) return 0;
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    return 1;
}
Q3T_API int q3t_tts_get_hidden_states(q3t_tts * t, float * hidden_out) {
    if (!t) return 0;
    std::vector<float> h;
    if (!t->t.get_hidden_states(h)) return 0;
    if (hidden_out) memcpy(hidden_out, h.data(), h.size() * sizeof(float));
    return 1;
}
Q3T_API int q3t_tts_predict_codes(q3t_tts * t, const float * hidden, const int32_t * prev, float * logits_out) {
    if (!t) return 0;
    std::vector<float> o;
    if (!t->t.predict_codes(hidden, prev, o)) return 0;
    if (logits_out) memcpy(logits_out, o.data(), o.size() * sizeof(float));
    return 1;
}
Q3T_API int q3t_tts_predict_codes_autoregressive(q3t_tts * t, const float * hidden, int32_t cb0, int32_t * codes15_out,
                                                 float temperature, int32_t top_k) {
    if (!t) return 0;
    std::vector<int32_t> o;
    if (!t->t.predict_codes_autoregressive(hidden, cb0, o, temperature, top_k)) return 0;
    if (codes15_out) memcpy(codes15_out, o.data(), o.size() * sizeof(int32_t));
    return 1;
}
Q3T_API int q3t_tts_generate(q3t_tts * t, const int32_t * tokens, int32_t n_tokens, const float * spk, int32_t max_len,
                             int32_t * codes_out, int32_t * n_frames_out, int32_t language_id, float rep, float temp, int32_t top_k) {
    if (!t) return 0;
    std::vector<int32_t> o;
    if (!t->t.generate(tokens, n_tokens, spk, max_len, o, language_id, rep, temp, top_k)) return 0;
    if (codes_out && !o.empty()) memcpy(codes_out, o.data(), o.size() * sizeof(int32_t));
    if (n_frames_out) *n_frames_out = (int32_t) (o.size() / (size_t) t->t.get_config().n_codebooks);
    return 1;
}
Q3T_API int q3t_tts_generate_batch(q3t_tts * t, int32_t n, const int32_t * const * tokens, const int32_t * n_tokens,
                                   const float * const * spk, int32_t max_len, int32_t * const * codes_out, int32_t * n_frames_out,
                                   int32_t language_id, float rep) {
    return q3t_tts_generate_batch_sampled(t, n, tokens, n_tokens, spk, max_len, codes_out, n_frames_out, language_id, rep, 0.0f, 0);
}
Q3T_API int q3t_tts_generate_batch_sampled(q3t_tts * t, int32_t n, const int32_t * const * tokens, const int32_t * n_tokens,
                                           const float * const * spk, int32_t max_len, int32_t * const * codes_out, int32_t * n_frames_out,
                                           int32_t language_id, float rep, float temperature, int32_t top_k) {
    if (!t || n <= 0) return 0;
    std::vector<const int32_t *> tk(tokens, tokens + n);
    std::vector<int32_t> nt(n_tokens, n_tokens + n);
    std::vector<const float *> sp;
    if (spk) sp.assign(spk, spk + n);
    std::vector<std::vector<int32_t>> o;
    if (!t->t.generate_batch(tk, nt, sp, max_len, o, language_id, rep, temperature, top_k)) return 0;
    const size_t ncb = (size_t) t->t.get_config().n_codebooks;
    for (int32_t i = 0; i < n; ++i) {
        if (codes_out && codes_out[i] && !o[i].empty()) memcpy(codes_out[i], o[i].data(), o[i].size() * sizeof(int32_t));
        if (n_frames_out) n_frames_out[i] = (int32_t) (o[i].size() / ncb);
    }
    return 1;
}
Q3T_API void q3t_tts_set_seed(q3t_tts * t, uint64_t seed) { if (t) t->t.set_seed(seed); }
Q3T_API int q3t_tts_get_config(q3t_tts * t, q3t_tts_config * o) {
    if (!t || !o) return 0;
    const auto & c = t->t.get_config();
    o->text_vocab_size = c.text_vocab_size; o->text_embd_dim = c.text_embd_dim; o->hidden_size = c.hidden_size;
    o->n_layers = c.n_layers; o->n_attention_heads = c.n_attention_heads; o->n_key_value_heads = c.n_key_value_heads;
    o->intermediate_size = c.intermediate_size; o->head_dim = c.head_dim; o->rms_norm_eps = c.rms_norm_eps;
    o->rope_theta = c.rope_theta; o->codec_vocab_size = c.codec_vocab_size; o->n_codebooks = c.n_codebooks;
    o->code_pred_layers = c.code_pred_layers; o->code_pred_vocab_size = c.code_pred_vocab_size;
    o->codec_pad_id = c.codec_pad_id; o->codec_bos_id = c.codec_bos_id; o->codec_eos_id = c.codec_eos_id;
    o->tts_bos_token_id = c.tts_bos_token_id; o->tts_eos_token_id = c.tts_eos_token_id; o->tts_pad_token_id = c.tts_pad_token_id;
    o->codec_think_id = c.codec_think_id; o->codec_nothink_id = c.codec_nothink_id;
    o->codec_think_bos_id = c.codec_think_bos_id; o->codec_think_eos_id = c.codec_think_eos_id;
    o->english_language_id = c.english_language_id;
    return 1;
}
Q3T_API const char * q3t_tts_get_error(q3t_tts * t) { return t ? t->t.get_error().c_str() : "null object"; }
Q3T_API q3t_ctx * q3t_tts_engine(q3t_tts * t) { return t ? t->t.engine() : nullptr; }

// ---------------------------------------------------------------- AudioTokenizerDecoder
Q3T_API q3t_dec * q3t_dec_new(void) { return new q3t_dec(); }
Q3T_API void q3t_dec_delete(q3t_dec * d) { delete d; }
Q3T_API int  q3t_dec_load_model(q3t_dec * d, const char * p) { return d && p && d->d.load_model(p); }
Q3T_API void q3t_dec_unload_model(q3t_dec * d) { if (d) d->d.unload_model(); }
Q3T_API int64_t q3t_dec_decode(q3t_dec * d, const int32_t * codes, int32_t n_frames, float * samples_out, int64_t cap) {
    if (!d) return -1;
    std::vector<float> s;
    if (!d->d.decode(codes, n_frames, s)) return -1;
    const int64_t n = (int64_t) s.size();
    if (samples_out) memcpy(samples_out, s.data(), (size_t) (n < cap ? n : cap) * sizeof(float));
    return n;
}
Q3T_API void q3t_dec_stream_reset(q3t_dec * d) { if (d) d->d.stream_reset(); }
Q3T_API int64_t q3t_dec_decode_append(q3t_dec * d, const int32_t * codes, int32_t n_new_frames, float * samples_out, int64_t cap) {
    if (!d) return -1;
    std::vector<float> o;
    if (!d->d.decode_append(codes, n_new_frames, o)) return -1;
    if (samples_out) memcpy(samples_out, o.data(), std::min<size_t>(o.size(), (size_t) std::max<int64_t>(cap, 0)) * sizeof(float));
    return (int64_t) o.size();
}
Q3T_API const char * q3t_dec_get_error(q3t_dec * d) { return d ? d->d.get_error().c_str() : "null object"; }

// ---------------------------------------------------------------- AudioTokenizerEncoder
Q3T_API q3t_enc * q3t_enc_new(void) { return new q3t_enc(); }
Q3T_API void q3t_enc_delete(q3t_enc * e) { delete e; }
Q3T_API int q3t_enc_load_model(q3t_enc * e, const char * model_path) { return e && model_path && e->e.load_model(model_path) ? 1 : 0; }
Q3T_API int q3t_enc_encode(q3t_enc * e, const float * samples, int32_t n_samples, float * embedding_out) {
    if (!e || !embedding_out) return 0;
    std::vector<float> v;
    if (!e->e.encode(samples, n_samples, v)) return 0;
    memcpy(embedding_out, v.data(), v.size() * sizeof(float));
    return 1;
}
Q3T_API const char * q3t_enc_get_error(q3t_enc * e) { return e ? e->e.get_error().c_str() : "null object"; }

// ---------------------------------------------------------------- streaming / batched front door (b200_pipeline.h)
Q3T_API int q3t_tts_generate_begin(q3t_tts * t, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd, int32_t max_len,
                                   int32_t language_id, float repetition_penalty, float temperature, int32_t top_k) {
    return t && t->t.generate_begin(text_tokens, n_tokens, speaker_embd, max_len, language_id, repetition_penalty, temperature, top_k) ? 1 : 0;
}
Q3T_API int q3t_tts_generate_next(q3t_tts * t, int32_t max_frames, int32_t * codes_out, int32_t * n_new_out, int32_t * done_out) {
    if (!t || !n_new_out || !done_out) return 0;
    std::vector<int32_t> c; bool done = true;
    if (!t->t.generate_next(max_frames, c, done)) return 0;
    if (codes_out && !c.empty()) memcpy(codes_out, c.data(), c.size() * sizeof(int32_t));
    *n_new_out = (int32_t) (c.size() / (size_t) t->t.get_config().n_codebooks);
    *done_out = done ? 1 : 0;
    return 1;
}
static std::string g_pipe_err;
Q3T_API const char * q3t_pipe_get_error(void) { return g_pipe_err.c_str(); }
Q3T_API int64_t q3t_pipe_synthesize_stream(q3t_tts * t, q3t_dec * d, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                                           int32_t max_len, int32_t language_id, float repetition_penalty, float temperature, int32_t top_k,
                                           int32_t first_chunk_frames, int32_t chunk_frames, float * samples_out, int64_t cap,
                                           q3t_stream_stats * stats_out) {
    if (!t || !d) { g_pipe_err = "null object"; return -1; }
    b200::synth_params p;
    p.max_audio_tokens = max_len; p.language_id = language_id; p.repetition_penalty = repetition_penalty; p.temperature = temperature; p.top_k = top_k;
    int64_t n = 0;
    b200::stream_stats st;
    auto sink = [&](const float * s, size_t k) {
        if (samples_out && n < cap) memcpy(samples_out + n, s, (size_t) std::min<int64_t>((int64_t) k, cap - n) * sizeof(float));
        n += (int64_t) k;
    };
    if (!b200::synthesize_stream(t->t, d->d, text_tokens, n_tokens, speaker_embd, p, first_chunk_frames, chunk_frames, sink, &st, g_pipe_err)) return -1;
    if (stats_out) {
        stats_out->t_first_audio_ms = st.t_first_audio_ms; stats_out->t_total_ms = st.t_total_ms;
        stats_out->n_frames = st.n_frames; stats_out->n_chunks = st.n_chunks; stats_out->n_samples = st.n_samples;
    }
    return n;
}
Q3T_API int q3t_pipe_synthesize_batch(q3t_tts * t, q3t_dec * d, int32_t n, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                      const float * const * speaker_embd, int32_t max_len, int32_t language_id, float repetition_penalty,
                                      float * const * samples_out, const int64_t * caps, int64_t * n_samples_out, q3t_batch_stats * stats_out) {
    return q3t_pipe_synthesize_batch_multi(t, &d, 1, n, text_tokens, n_tokens, speaker_embd, max_len, language_id, repetition_penalty,
                                           samples_out, caps, n_samples_out, stats_out);
}
Q3T_API int q3t_pipe_decode_batch_multi(q3t_dec * const * decs, int32_t n_dec, int32_t n, const int32_t * const * codes, const int32_t * n_frames,
                                        float * const * samples_out, const int64_t * caps, int64_t * n_samples_out) {
    if (!decs || n_dec <= 0 || n <= 0 || !codes || !n_frames || !n_samples_out) { g_pipe_err = "bad argument"; return 0; }
    std::vector<AudioTokenizerDecoder *> dv;
    for (int i = 0; i < n_dec; ++i) { if (!decs[i]) { g_pipe_err = "bad argument"; return 0; } dv.push_back(&decs[i]->d); }
    if (!samples_out || !caps) { g_pipe_err = "bad argument"; return 0; }
    // straight into the caller's buffers: decoder k takes utterances k, k + n_dec, ... on its own host thread
    std::vector<std::string> errs((size_t) n_dec);
    auto work = [&](int k) {
        for (int i = k; i < n; i += n_dec) {
            n_samples_out[i] = 0;
            if (n_frames[i] <= 0) continue;
            if (!codes[i] || !samples_out[i]) { errs[(size_t) k] = "bad argument"; return; }
            if (!dv[(size_t) k]->decode_to(codes[i], n_frames[i], samples_out[i], caps[i], n_samples_out[i])) { errs[(size_t) k] = dv[(size_t) k]->get_error(); return; }
        }
    };
    if (n_dec == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int k = 0; k < n_dec; ++k) th.emplace_back(work, k);
        for (auto & x : th) x.join();
    }
    for (const auto & e : errs) if (!e.empty()) { g_pipe_err = e; return 0; }
    return 1;
}
Q3T_API int q3t_pipe_synthesize_batch_multi(q3t_tts * t, q3t_dec * const * decs, int32_t n_dec, int32_t n, const int32_t * const * text_tokens,
                                            const int32_t * n_tokens, const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                            float repetition_penalty, float * const * samples_out, const int64_t * caps, int64_t * n_samples_out,
                                            q3t_batch_stats * stats_out) {
    if (!t || !decs || n_dec <= 0 || n <= 0 || !text_tokens || !n_tokens || !n_samples_out) { g_pipe_err = "bad argument"; return 0; }
    std::vector<AudioTokenizerDecoder *> dv;
    for (int i = 0; i < n_dec; ++i) { if (!decs[i]) { g_pipe_err = "bad argument"; return 0; } dv.push_back(&decs[i]->d); }
    b200::synth_params p;
    p.max_audio_tokens = max_len; p.language_id = language_id; p.repetition_penalty = repetition_penalty; p.temperature = 0.0f; p.top_k = 0;
    std::vector<const int32_t *> tt(text_tokens, text_tokens + n);
    std::vector<int32_t> nt(n_tokens, n_tokens + n);
    std::vector<const float *> sp;
    if (speaker_embd) sp.assign(speaker_embd, speaker_embd + n);
    std::vector<std::vector<float>> audio;
    b200::batch_stats st;
    if (!b200::synthesize_batch(t->t, dv, tt, nt, sp, p, audio, &st, g_pipe_err)) return 0;
    for (int i = 0; i < n; ++i) {
        n_samples_out[i] = (int64_t) audio[i].size();
        if (samples_out && samples_out[i] && caps)
            memcpy(samples_out[i], audio[i].data(), (size_t) std::min<int64_t>((int64_t) audio[i].size(), caps[i]) * sizeof(float));
    }
    if (stats_out) {
        stats_out->t_generate_ms = st.t_generate_ms; stats_out->t_decode_ms = st.t_decode_ms; stats_out->t_total_ms = st.t_total_ms;
        stats_out->n_frames = st.n_frames; stats_out->n_samples = st.n_samples;
    }
    return 1;
}
Q3T_API q3t_ctx * q3t_dec_engine(q3t_dec * d) { return d ? d->d.engine() : nullptr; }

} // extern "C"
