// qwen3_tts::TTSTransformer with the reference's public interface (src/tts_transformer.h:190-280);
// every forward runs on the B200 engine behind include/q3t.h — there is no ggml and no CPU path.
#pragma once

#include "../../include/q3t.h"

#include <cstdint>
#include <string>
#include <vector>

namespace qwen3_tts {

// kept unconditionally (the reference's -DQWEN3_TTS_TIMING changes its class layout,
// src/tts_transformer.h:342-344; here the ABI is the same either way)
struct tts_timing {
    double t_prefill_build_ms = 0;
    double t_prefill_forward_ms = 0;
    double t_talker_forward_ms = 0;      // frames: talker + code predictor are one device program
    double t_code_pred_ms = 0;
    int32_t n_frames = 0;
    double t_generate_total_ms = 0;
    int64_t kernel_launches = 0;
};

// src/tts_transformer.h:58-99, field for field
struct tts_transformer_config {
    int32_t text_vocab_size = 151936;
    int32_t text_embd_dim = 2048;
    int32_t hidden_size = 1024;
    int32_t n_layers = 28;
    int32_t n_attention_heads = 16;
    int32_t n_key_value_heads = 8;
    int32_t intermediate_size = 3072;
    int32_t head_dim = 128;
    float rms_norm_eps = 1e-6f;
    float rope_theta = 1000000.0f;
    int32_t mrope_section[3] = {24, 20, 20};
    int32_t codec_vocab_size = 3072;
    int32_t n_codebooks = 16;
    int32_t code_pred_layers = 5;
    int32_t code_pred_vocab_size = 2048;
    int32_t codec_pad_id = 2148;
    int32_t codec_bos_id = 2149;
    int32_t codec_eos_id = 2150;
    int32_t tts_bos_token_id = 151672;
    int32_t tts_eos_token_id = 151673;
    int32_t tts_pad_token_id = 151671;
    int32_t codec_think_id = 2154;
    int32_t codec_nothink_id = 2155;
    int32_t codec_think_bos_id = 2156;
    int32_t codec_think_eos_id = 2157;
    int32_t english_language_id = 2050;
};

namespace b200 { class GGUFLoader; }

class Q3T_API TTSTransformer {
public:
    TTSTransformer();
    ~TTSTransformer();

    bool load_model(const std::string & model_path);
    void unload_model();

    bool init_kv_cache(int32_t n_ctx);
    void clear_kv_cache();
    bool init_code_pred_kv_cache(int32_t n_ctx);
    void clear_code_pred_kv_cache();

    bool forward_text(const int32_t * text_tokens, int32_t n_tokens,
                      const float * speaker_embd, int32_t n_past,
                      std::vector<float> & output);

    bool forward_prefill(const float * prefill_embd, int32_t n_tokens,
                         int32_t n_past, std::vector<float> & output,
                         std::vector<float> * logits_out = nullptr);

    bool forward_codec(int32_t codec_token, int32_t n_past, std::vector<float> & output);

    bool forward_step(const float * step_embd, int32_t n_past,
                      std::vector<float> & output,
                      std::vector<float> * hidden_out = nullptr);

    bool get_hidden_states(std::vector<float> & hidden) const;

    bool predict_codes(const float * hidden, const int32_t * prev_codes, std::vector<float> & output);

    bool predict_codes_autoregressive(const float * hidden, int32_t codebook_0_token,
                                      std::vector<int32_t> & output,
                                      float temperature = 0.9f, int32_t top_k = 50);

    bool generate(const int32_t * text_tokens, int32_t n_tokens,
                  const float * speaker_embd, int32_t max_len,
                  std::vector<int32_t> & output,
                  int32_t language_id = 2050,
                  float repetition_penalty = 1.05f,
                  float temperature = 0.9f,
                  int32_t top_k = 50);

    const tts_transformer_config & get_config() const { return config_; }
    const std::string & get_error() const { return error_msg_; }

    bool forward(const int32_t * tokens, int32_t n_tokens, int32_t n_past, std::vector<float> & output);
    bool forward_with_audio(const int32_t * tokens, int32_t n_tokens,
                            const float * audio_embd, int32_t n_audio,
                            int32_t audio_start_pos, int32_t n_past,
                            std::vector<float> & output);

    // ---- extensions (not in the reference) --------------------------------------------------
    // n independent utterances batched on the device (up to 16 sequences per kernel launch); outputs[i] receives what
    // generate(..., temperature = 0) returns for utterance i.  temperature > 0: sampled like generate(), utterance i
    // drawing from the stream seeded set_seed()'s seed + i
    bool generate_batch(const std::vector<const int32_t *> & text_tokens, const std::vector<int32_t> & n_tokens,
                        const std::vector<const float *> & speaker_embd, int32_t max_len,
                        std::vector<std::vector<int32_t>> & outputs,
                        int32_t language_id = 2050, float repetition_penalty = 1.05f, float temperature = 0.0f, int32_t top_k = 0);
    // streaming generate: begin = prompt + prefill, next = up to max_frames more frames appended to `new_codes`;
    // the concatenation equals generate() with the same arguments
    bool generate_begin(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd, int32_t max_len,
                        int32_t language_id = 2050, float repetition_penalty = 1.05f, float temperature = 0.9f, int32_t top_k = 50);
    bool generate_next(int32_t max_frames, std::vector<int32_t> & new_codes, bool & done);
    void set_seed(uint64_t seed) { seed_ = seed; draws_ = 0; }   // the reference's rng_ cannot be seeded; re-seeding restarts the stream
    q3t_ctx * engine() const { return engine_; }
    const tts_timing & last_timing() const { return timing_; }

private:
    bool parse_config(const b200::GGUFLoader & loader);
    bool fail(int rc);

    tts_transformer_config config_;
    q3t_ctx * engine_ = nullptr;
    bool loaded_ = false;
    std::string error_msg_;
    std::vector<float> last_hidden_;
    uint64_t seed_ = 0x5eed5eedULL;
    uint64_t draws_ = 0;
    tts_timing timing_;
};

} // namespace qwen3_tts
