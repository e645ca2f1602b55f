#include "tts_transformer.h"
#include "b200_gguf_loader.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <initializer_list>

namespace qwen3_tts {

using b200::GGUFLoader;
using b200::init_preferred_backend;
using b200::release_preferred_backend;
using b200::load_tensor_data_from_file;

TTSTransformer::TTSTransformer() = default;

TTSTransformer::~TTSTransformer() { unload_model(); }

bool TTSTransformer::fail(int rc) {
    if (rc == Q3T_OK) return false;
    error_msg_ = engine_ ? q3t_last_error(engine_) : q3t_last_error(nullptr);
    return true;
}

void TTSTransformer::unload_model() {
    if (engine_) {
        q3t_unload(engine_, 1);
        release_preferred_backend(engine_, "TTSTransformer");
        engine_ = nullptr;
    }
    loaded_ = false;
    last_hidden_.clear();
}

// GGUF KV -> config with the reference's fallback key lists, first match wins
// (src/tts_transformer.cpp:212-344)
bool TTSTransformer::parse_config(const GGUFLoader & loader) {
    auto has = [&](const char * key) { return loader.file().find(key) != nullptr; };
    auto get_u32_any = [&](std::initializer_list<const char *> keys, int32_t default_val) -> int32_t {
        for (const char * key : keys) if (has(key)) return loader.get_u32(key, default_val);
        return default_val;
    };
    auto get_f32_any = [&](std::initializer_list<const char *> keys, float default_val) -> float {
        for (const char * key : keys) if (has(key)) return loader.get_f32(key, default_val);
        return default_val;
    };
    auto & cfg = config_;
    cfg = tts_transformer_config{};
    cfg.text_vocab_size   = get_u32_any({"qwen3-tts.text.vocab_size", "qwen3-tts.text_vocab_size"}, 151936);
    cfg.text_embd_dim     = get_u32_any({"qwen3-tts.text.embedding_dim", "qwen3-tts.text_hidden_size"}, 2048);
    cfg.hidden_size       = get_u32_any({"qwen3-tts.talker.embedding_length", "qwen3-tts.embedding_length"}, 1024);
    cfg.n_layers          = get_u32_any({"qwen3-tts.talker.block_count", "qwen3-tts.block_count"}, 28);
    cfg.n_attention_heads = get_u32_any({"qwen3-tts.talker.attention.head_count", "qwen3-tts.attention.head_count"}, 16);
    cfg.n_key_value_heads = get_u32_any({"qwen3-tts.talker.attention.head_count_kv", "qwen3-tts.attention.head_count_kv"}, 8);
    cfg.intermediate_size = get_u32_any({"qwen3-tts.talker.feed_forward_length", "qwen3-tts.feed_forward_length"}, 3072);
    cfg.head_dim          = get_u32_any({"qwen3-tts.talker.attention.key_length", "qwen3-tts.attention.key_length"}, 128);
    cfg.rms_norm_eps      = get_f32_any({"qwen3-tts.talker.attention.layer_norm_rms_epsilon", "qwen3-tts.attention.layer_norm_rms_epsilon"}, 1e-6f);
    cfg.rope_theta        = get_f32_any({"qwen3-tts.talker.rope.freq_base", "qwen3-tts.rope.freq_base"}, 1000000.0f);
    cfg.codec_vocab_size  = get_u32_any({"qwen3-tts.talker.codec_vocab_size", "qwen3-tts.vocab_size"}, 3072);
    cfg.n_codebooks       = get_u32_any({"qwen3-tts.talker.num_codebooks", "qwen3-tts.num_code_groups"}, 16);
    cfg.code_pred_layers  = get_u32_any({"qwen3-tts.code_pred.layer_count", "qwen3-tts.code_predictor.layer_count"}, 5);
    cfg.code_pred_vocab_size = get_u32_any({"qwen3-tts.code_pred.vocab_size", "qwen3-tts.code_predictor.vocab_size"}, 2048);
    cfg.codec_pad_id      = get_u32_any({"qwen3-tts.codec.pad_id"}, 2148);
    cfg.codec_bos_id      = get_u32_any({"qwen3-tts.codec.bos_id"}, 2149);
    cfg.codec_eos_id      = get_u32_any({"qwen3-tts.codec.eos_id", "qwen3-tts.codec.eos_token_id"}, 2150);
    cfg.tts_bos_token_id  = get_u32_any({"qwen3-tts.tts_bos_token_id", "qwen3-tts.tts.bos_token_id", "qwen3-tts.tts.bos_id"}, 151672);
    cfg.tts_eos_token_id  = get_u32_any({"qwen3-tts.tts_eos_token_id", "qwen3-tts.tts.eos_token_id", "qwen3-tts.tts.eos_id"}, 151673);
    cfg.tts_pad_token_id  = get_u32_any({"qwen3-tts.tts_pad_token_id", "qwen3-tts.tts.pad_token_id", "qwen3-tts.tts.pad_id"}, 151671);
    cfg.codec_think_id    = get_u32_any({"qwen3-tts.codec.think_id", "qwen3-tts.codec_think_id"}, 2154);
    cfg.codec_nothink_id  = get_u32_any({"qwen3-tts.codec.nothink_id", "qwen3-tts.codec_nothink_id"}, 2155);
    cfg.codec_think_bos_id = get_u32_any({"qwen3-tts.codec.think_bos_id", "qwen3-tts.codec_think_bos_id"}, 2156);
    cfg.codec_think_eos_id = get_u32_any({"qwen3-tts.codec.think_eos_id", "qwen3-tts.codec_think_eos_id"}, 2157);
    cfg.english_language_id = get_u32_any({"qwen3-tts.language.english_id", "qwen3-tts.codec.language.english_id", "qwen3-tts.language_id"}, 2050);
    return true;
}

static q3t_tts_config to_engine_config(const tts_transformer_config & c) {
    q3t_tts_config e{};
    e.text_vocab_size = c.text_vocab_size; e.text_embd_dim = c.text_embd_dim;
    e.hidden_size = c.hidden_size; e.n_layers = c.n_layers;
    e.n_attention_heads = c.n_attention_heads; e.n_key_value_heads = c.n_key_value_heads;
    e.intermediate_size = c.intermediate_size; e.head_dim = c.head_dim;
    e.rms_norm_eps = c.rms_norm_eps; e.rope_theta = c.rope_theta;
    e.codec_vocab_size = c.codec_vocab_size; e.n_codebooks = c.n_codebooks;
    e.code_pred_layers = c.code_pred_layers; e.code_pred_vocab_size = c.code_pred_vocab_size;
    e.codec_pad_id = c.codec_pad_id; e.codec_bos_id = c.codec_bos_id; e.codec_eos_id = c.codec_eos_id;
    e.tts_bos_token_id = c.tts_bos_token_id; e.tts_eos_token_id = c.tts_eos_token_id; e.tts_pad_token_id = c.tts_pad_token_id;
    e.codec_think_id = c.codec_think_id; e.codec_nothink_id = c.codec_nothink_id;
    e.codec_think_bos_id = c.codec_think_bos_id; e.codec_think_eos_id = c.codec_think_eos_id;
    e.english_language_id = c.english_language_id;
    return e;
}

bool TTSTransformer::load_model(const std::string & model_path) {
    unload_model();                                            // src/tts_transformer.cpp:52-53

    GGUFLoader loader;
    if (!loader.open(model_path)) { error_msg_ = loader.get_error(); return false; }
    if (!parse_config(loader)) return false;

    engine_ = init_preferred_backend("TTSTransformer", &error_msg_);
    if (!engine_) return false;
    int32_t sms = 0, maj = 0, min = 0;
    q3t_device_info(engine_, &sms, &maj, &min);
    fprintf(stderr, "  TTSTransformer backend: CUDA sm_%d%d (%d SMs)\n", maj, min, sms);

    const q3t_tts_config ecfg = to_engine_config(config_);
    if (fail(q3t_set_tts_config(engine_, &ecfg))) {
        std::string keep = error_msg_; unload_model(); error_msg_ = keep; return false;
    }
    // spk_enc.* / tok_* are skipped exactly like the reference loader (src/tts_transformer.cpp:375-377)
    const char * prefixes[] = { "talker.", "code_pred." };
    if (!load_tensor_data_from_file(loader, engine_, prefixes, 2, error_msg_)) {
        std::string keep = error_msg_; unload_model(); error_msg_ = keep; return false;
    }
    if (fail(q3t_finalize_weights(engine_, 1))) {
        std::string keep = error_msg_; unload_model(); error_msg_ = keep; return false;
    }
    loaded_ = true;
    return true;
}

bool TTSTransformer::init_kv_cache(int32_t n_ctx) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    return !fail(q3t_kv_init(engine_, n_ctx));
}

void TTSTransformer::clear_kv_cache() {}          // n_used = 0 in the reference (:735-737); positions are explicit here

bool TTSTransformer::init_code_pred_kv_cache(int32_t n_ctx) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    (void) n_ctx;                                  // the engine owns a fixed 16-position cache (:2336-2341)
    return true;
}

void TTSTransformer::clear_code_pred_kv_cache() {}

bool TTSTransformer::forward_prefill(const float * prefill_embd, int32_t n_tokens, int32_t n_past,
                                     std::vector<float> & output, std::vector<float> * logits_out) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!prefill_embd) { error_msg_ = "prefill_embd is null"; return false; }
    if (n_tokens <= 0) { error_msg_ = "n_tokens must be > 0"; return false; }
    output.resize((size_t) n_tokens * config_.hidden_size);
    std::vector<float> logits((size_t) config_.codec_vocab_size);
    if (fail(q3t_talker_prefill(engine_, prefill_embd, n_tokens, n_past, output.data(), logits.data()))) return false;
    last_hidden_.assign(output.end() - config_.hidden_size, output.end());
    if (logits_out) *logits_out = std::move(logits);
    return true;
}

bool TTSTransformer::forward_text(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                                  int32_t n_past, std::vector<float> & output) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!text_tokens) { error_msg_ = "text_tokens is null"; return false; }
    if (n_tokens <= 0) { error_msg_ = "n_tokens must be > 0"; return false; }
    std::vector<float> projected((size_t) n_tokens * config_.hidden_size);
    if (fail(q3t_project_text(engine_, text_tokens, n_tokens, projected.data()))) return false;
    if (speaker_embd) {                                        // src/tts_transformer.cpp:1980-1988
        const int32_t hs = config_.hidden_size;
        for (int32_t t = 0; t < n_tokens; ++t)
            for (int32_t h = 0; h < hs; ++h) projected[(size_t) t * hs + h] += speaker_embd[h];
    }
    return forward_prefill(projected.data(), n_tokens, n_past, output, nullptr);
}

bool TTSTransformer::forward_step(const float * step_embd, int32_t n_past, std::vector<float> & output,
                                  std::vector<float> * hidden_out) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!step_embd) { error_msg_ = "step_embd is null"; return false; }
    output.resize((size_t) config_.codec_vocab_size);
    last_hidden_.resize((size_t) config_.hidden_size);
    if (fail(q3t_talker_step(engine_, step_embd, n_past, output.data(), last_hidden_.data()))) return false;
    if (hidden_out) *hidden_out = last_hidden_;
    return true;
}

bool TTSTransformer::forward_codec(int32_t codec_token, int32_t n_past, std::vector<float> & output) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    std::vector<float> row((size_t) config_.hidden_size);
    if (fail(q3t_codec_embed(engine_, &codec_token, 1, row.data()))) return false;
    return forward_step(row.data(), n_past, output, nullptr);
}

bool TTSTransformer::get_hidden_states(std::vector<float> & hidden) const {
    if (last_hidden_.empty()) return false;
    hidden = last_hidden_;
    return true;
}

bool TTSTransformer::predict_codes(const float * hidden, const int32_t * prev_codes, std::vector<float> & output) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    output.resize((size_t) (config_.n_codebooks - 1) * config_.code_pred_vocab_size);
    return !fail(q3t_codepred_parallel(engine_, hidden, prev_codes, output.data()));
}

bool TTSTransformer::predict_codes_autoregressive(const float * hidden, int32_t codebook_0_token,
                                                  std::vector<int32_t> & output, float temperature, int32_t top_k) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    output.resize(15);
    return !fail(q3t_codepred_frame(engine_, hidden, codebook_0_token, temperature, top_k,
                                    seed_ + 0x9E3779B97F4A7C15ULL * (++draws_), output.data()));
}

bool TTSTransformer::generate(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                              int32_t max_len, std::vector<int32_t> & output, int32_t language_id,
                              float repetition_penalty, float temperature, int32_t top_k) {
    const auto t0 = std::chrono::high_resolution_clock::now();
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!text_tokens) { error_msg_ = "text_tokens is null"; return false; }
    if (n_tokens < 4) { error_msg_ = "Need at least 4 text tokens for generation"; return false; }
    if (max_len <= 0) { output.clear(); return true; }          // src/tts_transformer.cpp:2604-2607
    output.resize((size_t) max_len * config_.n_codebooks);
    int32_t n_frames = 0;
    if (fail(q3t_generate(engine_, text_tokens, n_tokens, speaker_embd, max_len, language_id, repetition_penalty,
                          temperature, top_k, seed_ + 0x9E3779B97F4A7C15ULL * (++draws_), output.data(), &n_frames))) {
        output.clear();
        return false;
    }
    output.resize((size_t) n_frames * config_.n_codebooks);
    q3t_timing tm{};
    q3t_get_timing(engine_, &tm);
    timing_.t_prefill_forward_ms = tm.prefill_ms;
    timing_.t_talker_forward_ms = tm.frames_ms;
    timing_.n_frames = tm.n_frames;
    timing_.kernel_launches = tm.kernel_launches;
    timing_.t_generate_total_ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count();
    return true;
}

bool TTSTransformer::generate_begin(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd, int32_t max_len,
                                    int32_t language_id, float repetition_penalty, float temperature, int32_t top_k) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!text_tokens) { error_msg_ = "text_tokens is null"; return false; }
    if (n_tokens < 4) { error_msg_ = "Need at least 4 text tokens for generation"; return false; }
    return !fail(q3t_stream_begin(engine_, text_tokens, n_tokens, speaker_embd, max_len, language_id, repetition_penalty,
                                  temperature, top_k, seed_ + 0x9E3779B97F4A7C15ULL * (++draws_)));
}

bool TTSTransformer::generate_next(int32_t max_frames, std::vector<int32_t> & new_codes, bool & done) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    new_codes.resize((size_t) std::max(max_frames, 0) * config_.n_codebooks);
    int32_t n_new = 0, d = 1;
    if (fail(q3t_stream_next(engine_, max_frames, new_codes.data(), &n_new, &d))) { new_codes.clear(); return false; }
    new_codes.resize((size_t) n_new * config_.n_codebooks);
    done = d != 0;
    return true;
}

bool TTSTransformer::generate_batch(const std::vector<const int32_t *> & text_tokens, const std::vector<int32_t> & n_tokens,
                                    const std::vector<const float *> & speaker_embd, int32_t max_len,
                                    std::vector<std::vector<int32_t>> & outputs, int32_t language_id, float repetition_penalty,
                                    float temperature, int32_t top_k) {
    const auto t0 = std::chrono::high_resolution_clock::now();
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    const size_t n = text_tokens.size();
    if (n == 0 || n_tokens.size() != n || (!speaker_embd.empty() && speaker_embd.size() != n)) { error_msg_ = "batch arrays disagree"; return false; }
    for (size_t i = 0; i < n; ++i) {
        if (!text_tokens[i]) { error_msg_ = "text_tokens is null"; return false; }
        if (n_tokens[i] < 4) { error_msg_ = "Need at least 4 text tokens for generation"; return false; }
    }
    outputs.assign(n, {});
    if (max_len <= 0) return true;
    std::vector<int32_t *> outp(n);
    std::vector<int32_t> nf(n, 0);
    for (size_t i = 0; i < n; ++i) { outputs[i].resize((size_t) max_len * config_.n_codebooks); outp[i] = outputs[i].data(); }
    if (fail(q3t_generate_batch_sampled(engine_, (int32_t) n, text_tokens.data(), n_tokens.data(),
                                        speaker_embd.empty() ? nullptr : speaker_embd.data(), max_len, language_id,
                                        repetition_penalty, temperature, top_k, seed_, outp.data(), nf.data()))) {
        outputs.assign(n, {});
        return false;
    }
    for (size_t i = 0; i < n; ++i) outputs[i].resize((size_t) nf[i] * config_.n_codebooks);
    q3t_timing tm{};
    q3t_get_timing(engine_, &tm);
    timing_.t_prefill_forward_ms = tm.prefill_ms;
    timing_.t_talker_forward_ms = tm.frames_ms;
    timing_.kernel_launches = tm.kernel_launches;
    timing_.t_generate_total_ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count();
    return true;
}

bool TTSTransformer::forward(const int32_t * tokens, int32_t n_tokens, int32_t n_past, std::vector<float> & output) {
    return forward_text(tokens, n_tokens, nullptr, n_past, output);
}

bool TTSTransformer::forward_with_audio(const int32_t * tokens, int32_t n_tokens, const float * audio_embd,
                                        int32_t n_audio, int32_t audio_start_pos, int32_t n_past,
                                        std::vector<float> & output) {
    (void) audio_embd; (void) n_audio; (void) audio_start_pos;  // ignored by the reference too (:2858-2866)
    return forward_text(tokens, n_tokens, nullptr, n_past, output);
}

} // namespace qwen3_tts
