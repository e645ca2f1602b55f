#include "audio_tokenizer_encoder.h"
#include "b200_gguf_loader.h"

#include <cstdio>
#include <cstring>

namespace qwen3_tts {

using b200::GGUFLoader;
using b200::init_preferred_backend;
using b200::release_preferred_backend;
using b200::load_tensor_data_from_file;

AudioTokenizerEncoder::AudioTokenizerEncoder() = default;
AudioTokenizerEncoder::~AudioTokenizerEncoder() { unload(); }

void AudioTokenizerEncoder::unload() {
    if (engine_) {
        q3t_unload(engine_, 4);
        release_preferred_backend(engine_, "AudioTokenizerEncoder");
        engine_ = nullptr;
    }
    loaded_ = false;
}

bool AudioTokenizerEncoder::load_model(const std::string & model_path) {
    unload();
    GGUFLoader loader;
    if (!loader.open(model_path)) { error_msg_ = loader.get_error(); return false; }
    config_ = speaker_encoder_config{};
    config_.sample_rate   = loader.get_u32("qwen3-tts.speaker_encoder.sample_rate", 24000);        // src/audio_tokenizer_encoder.cpp:154-155
    config_.embedding_dim = loader.get_u32("qwen3-tts.speaker_encoder.embedding_length", 1024);
    int n_spk = 0;
    for (int64_t i = 0; i < loader.get_n_tensors(); ++i) {
        const char * name = loader.get_tensor_name(i);
        if (name && strncmp(name, "spk_enc.", 8) == 0) n_spk++;
    }
    if (n_spk == 0) { error_msg_ = "No speaker encoder tensors found in model"; return false; }
    if (config_.embedding_dim != 1024) { error_msg_ = "unsupported speaker embedding size (kernels are built for 1024)"; return false; }

    // a context of its own kind: the encoder must not take the transformer's or the decoder's slot
    engine_ = init_preferred_backend("AudioTokenizerEncoder", &error_msg_);
    if (!engine_) return false;
    int32_t sms = 0, maj = 0, min = 0;
    q3t_device_info(engine_, &sms, &maj, &min);
    fprintf(stderr, "  AudioTokenizerEncoder backend: CUDA sm_%d%d (%d SMs)\n", maj, min, sms);
    const char * prefixes[] = { "spk_enc." };
    if (!load_tensor_data_from_file(loader, engine_, prefixes, 1, error_msg_)) {
        std::string keep = error_msg_; unload(); error_msg_ = keep; return false;
    }
    if (q3t_finalize_weights(engine_, 3) != Q3T_OK) {
        std::string keep = q3t_last_error(engine_); unload(); error_msg_ = keep; return false;
    }
    loaded_ = true;
    return true;
}

bool AudioTokenizerEncoder::encode(const float * samples, int32_t n_samples, std::vector<float> & embedding) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    embedding.resize((size_t) config_.embedding_dim);
    if (q3t_speaker_encode(engine_, samples, n_samples, embedding.data(), nullptr) != Q3T_OK) {
        error_msg_ = q3t_last_error(engine_); embedding.clear(); return false;
    }
    return true;
}

} // namespace qwen3_tts
