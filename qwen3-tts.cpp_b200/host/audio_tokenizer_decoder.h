// qwen3_tts::AudioTokenizerDecoder with the reference's public interface
// (src/audio_tokenizer_decoder.h:15-29,159-179) on the B200 engine.
#pragma once

#include "../../include/q3t.h"

#include <cstdint>
#include <string>
#include <vector>

namespace qwen3_tts {

struct audio_decoder_config {                 // src/audio_tokenizer_decoder.h:15-29
    int32_t sample_rate = 24000;
    int32_t n_codebooks = 16;
    int32_t codebook_size = 2048;
    int32_t codebook_dim = 256;
    int32_t latent_dim = 1024;
    int32_t hidden_dim = 512;
    int32_t n_pre_tfm_layers = 8;
    int32_t n_heads = 16;
    int32_t ffn_dim = 1024;
    int32_t decoder_dim = 1536;
    int32_t upsample_rates[4] = {8, 5, 4, 3};
    float rms_norm_eps = 1e-5f;
    float rope_theta = 10000.0f;
};

class Q3T_API AudioTokenizerDecoder {
public:
    AudioTokenizerDecoder();
    ~AudioTokenizerDecoder();

    bool load_model(const std::string & model_path);
    void unload_model();

    // codes: [n_frames, n_codebooks] row-major -> samples in [-1,1] at 24 kHz (1920*n_frames-555 of them)
    bool decode(const int32_t * codes, int32_t n_frames, std::vector<float> & samples);

    const audio_decoder_config & get_config() const { return config_; }
    const std::string & get_error() const { return error_msg_; }

    // ---- extensions (not in the reference) --------------------------------------------------
    // streaming: append n_new_frames frames of codes to the running utterance and receive the samples that
    // became final; the concatenation over the calls equals decode() of all frames (the decoder is causal)
    void stream_reset() { stream_codes_.clear(); stream_done_ = 0; }
    bool decode_append(const int32_t * codes, int32_t n_new_frames, std::vector<float> & new_samples);
    q3t_ctx * engine() const { return engine_; }
    // decode() into a caller's buffer of `cap` floats (no intermediate vector); n_samples receives the count written
    bool decode_to(const int32_t * codes, int32_t n_frames, float * out, int64_t cap, int64_t & n_samples);

private:
    audio_decoder_config config_;
    q3t_ctx * engine_ = nullptr;
    bool loaded_ = false;
    std::string error_msg_;
    std::vector<int32_t> stream_codes_;
    int32_t stream_done_ = 0;
};

} // namespace qwen3_tts
