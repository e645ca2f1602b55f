#include "audio_tokenizer_decoder.h"
#include "b200_gguf_loader.h"

#include <algorithm>
#include <cstdio>
#include <cstring>

namespace qwen3_tts {

using b200::GGUFLoader;
using b200::init_preferred_backend;
using b200::release_preferred_backend;
using b200::load_tensor_data_from_file;

AudioTokenizerDecoder::AudioTokenizerDecoder() = default;
AudioTokenizerDecoder::~AudioTokenizerDecoder() { unload_model(); }

void AudioTokenizerDecoder::unload_model() {
    stream_reset();
    if (engine_) {
        q3t_unload(engine_, 2);
        release_preferred_backend(engine_, "AudioTokenizerDecoder");
        engine_ = nullptr;
    }
    loaded_ = false;
}

bool AudioTokenizerDecoder::load_model(const std::string & model_path) {
    unload_model();
    GGUFLoader loader;
    if (!loader.open(model_path)) { error_msg_ = loader.get_error(); return false; }
    // same (mismatching) keys as the reference, so the defaults win there and here
    // (src/audio_tokenizer_decoder.cpp:84-86 vs scripts/convert_tokenizer_to_gguf.py:411-413)
    config_ = audio_decoder_config{};
    config_.sample_rate   = loader.get_u32("qwen3-tts.tokenizer.sample_rate", 24000);
    config_.n_codebooks   = loader.get_u32("qwen3-tts.tokenizer.num_codebooks", 16);
    config_.codebook_size = loader.get_u32("qwen3-tts.tokenizer.codebook_size", 2048);

    int dec_tensor_count = 0;
    for (int64_t i = 0; i < loader.get_n_tensors(); ++i) {
        const char * name = loader.get_tensor_name(i);
        if (name && strncmp(name, "tok_dec.", 8) == 0) dec_tensor_count++;
    }
    if (dec_tensor_count == 0) { error_msg_ = "No decoder tensors found in model"; return false; }

    engine_ = init_preferred_backend("AudioTokenizerDecoder", &error_msg_);
    if (!engine_) return false;
    int32_t sms = 0, maj = 0, min = 0;
    q3t_device_info(engine_, &sms, &maj, &min);
    fprintf(stderr, "  AudioTokenizerDecoder backend: CUDA sm_%d%d (%d SMs)\n", maj, min, sms);

    const char * prefixes[] = { "tok_dec." };
    if (!load_tensor_data_from_file(loader, engine_, prefixes, 1, error_msg_)) {
        std::string keep = error_msg_; unload_model(); error_msg_ = keep; return false;
    }
    if (q3t_finalize_weights(engine_, 2) != Q3T_OK) {
        std::string keep = q3t_last_error(engine_); unload_model(); error_msg_ = keep; return false;
    }
    loaded_ = true;
    return true;
}

bool AudioTokenizerDecoder::decode(const int32_t * codes, int32_t n_frames, std::vector<float> & samples) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!codes || n_frames <= 0) { error_msg_ = "codes is null or n_frames <= 0"; return false; }
    const int64_t n = q3t_vocoder_n_samples(n_frames);
    samples.resize((size_t) std::max<int64_t>(n, 0));
    const int64_t got = q3t_vocoder_decode(engine_, codes, n_frames, samples.data(), (int64_t) samples.size());
    if (got < 0) { error_msg_ = q3t_last_error(engine_); samples.clear(); return false; }
    samples.resize((size_t) got);
    return true;
}

bool AudioTokenizerDecoder::decode_to(const int32_t * codes, int32_t n_frames, float * out, int64_t cap, int64_t & n_samples) {
    n_samples = 0;
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!codes || n_frames <= 0 || !out) { error_msg_ = "codes or out is null or n_frames <= 0"; return false; }
    if (cap < q3t_vocoder_n_samples(n_frames)) { error_msg_ = "output buffer too small"; return false; }
    const int64_t got = q3t_vocoder_decode(engine_, codes, n_frames, out, cap);
    if (got < 0) { error_msg_ = q3t_last_error(engine_); return false; }
    n_samples = got;
    return true;
}

bool AudioTokenizerDecoder::decode_append(const int32_t * codes, int32_t n_new_frames, std::vector<float> & new_samples) {
    if (!loaded_) { error_msg_ = "Model not loaded"; return false; }
    if (!codes || n_new_frames <= 0) { error_msg_ = "codes is null or n_frames <= 0"; return false; }
    stream_codes_.insert(stream_codes_.end(), codes, codes + (size_t) n_new_frames * config_.n_codebooks);
    const int32_t total = stream_done_ + n_new_frames;
    const int64_t n = q3t_vocoder_n_samples(total) - q3t_vocoder_n_samples(stream_done_);
    new_samples.resize((size_t) std::max<int64_t>(n, 0));
    const int64_t got = q3t_vocoder_decode_from(engine_, stream_codes_.data(), total, stream_done_, new_samples.data(), (int64_t) new_samples.size());
    if (got < 0) {
        error_msg_ = q3t_last_error(engine_); new_samples.clear();
        stream_codes_.resize((size_t) stream_done_ * config_.n_codebooks);      // the chunk was not consumed
        return false;
    }
    new_samples.resize((size_t) got);
    stream_done_ = total;
    return true;
}

} // namespace qwen3_tts
