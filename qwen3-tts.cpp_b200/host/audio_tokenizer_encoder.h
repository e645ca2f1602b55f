// qwen3_tts::AudioTokenizerEncoder (speaker encoder, ECAPA-TDNN x-vector) with the reference's public interface
// (src/audio_tokenizer_encoder.h:13-25,104-150) on the B200 engine: log-mel front end and the whole network run on the GPU
// (csrc/spk_encoder.inc).  SURVEY.md §8f N3.
#pragma once

#include "../../include/q3t.h"

#include <cstdint>
#include <string>
#include <vector>

namespace qwen3_tts {

struct speaker_encoder_config {               // src/audio_tokenizer_encoder.h:13-25, field for field
    int32_t sample_rate = 24000;
    int32_t n_mels = 128;
    int32_t n_fft = 1024;
    int32_t hop_length = 256;
    int32_t win_length = 1024;
    int32_t embedding_dim = 1024;
    int32_t hidden_dim = 512;
    int32_t n_res2net_blocks = 3;
    int32_t res2net_scale = 8;
    float f_min = 0.0f;
    float f_max = 12000.0f;
};

class Q3T_API AudioTokenizerEncoder {
public:
    AudioTokenizerEncoder();
    ~AudioTokenizerEncoder();

    // the main TTS GGUF (spk_enc.* tensors), not the tokenizer file
    bool load_model(const std::string & model_path);

    // samples: 24 kHz, [-1, 1] -> embedding [1024]
    bool encode(const float * samples, int32_t n_samples, std::vector<float> & embedding);

    // legacy interfaces of the reference header: not used for speaker encoding, same error strings
    bool encode(const float * samples, int32_t n_samples, std::vector<int32_t> & codes, int32_t & n_frames) {
        (void) samples; (void) n_samples; (void) codes; (void) n_frames;
        error_msg_ = "Use encode(samples, n_samples, embedding) for speaker encoding";
        return false;
    }
    bool get_embeddings(const int32_t * codes, int32_t n_frames, std::vector<float> & embeddings) {
        (void) codes; (void) n_frames; (void) embeddings;
        error_msg_ = "Use encode() for speaker embedding extraction";
        return false;
    }

    const speaker_encoder_config & get_config() const { return config_; }
    const std::string & get_error() const { return error_msg_; }
    q3t_ctx * engine() const { return engine_; }

private:
    void unload();
    speaker_encoder_config config_;
    q3t_ctx * engine_ = nullptr;
    bool loaded_ = false;
    std::string error_msg_;
};

} // namespace qwen3_tts
