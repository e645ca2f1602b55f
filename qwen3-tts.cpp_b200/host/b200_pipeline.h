// Batched / streaming front door over the two hot-path classes.  The reference pipeline
// (Qwen3TTS::synthesize_internal, src/qwen3_tts.cpp:278-442) is one utterance at a time and decodes only when the
// whole utterance has been generated; these helpers keep its order of operations (generate -> decode) but
//   * synthesize_stream: hands chunks of frames to the vocoder as they are generated and reports the time to the first
//     audible samples (every layer of the vocoder is causal, so the streamed waveform equals the one-shot decode);
//   * synthesize_batch: pushes N utterances through the batched engine (generate_batch: 16 sequences per launch) and
//     decodes each one, with the per-phase wall times.
// Extensions: no counterpart in the reference; namespace qwen3_tts::b200.
#pragma once

#include "audio_tokenizer_decoder.h"
#include "tts_transformer.h"

#include <functional>

namespace qwen3_tts {
namespace b200 {

struct synth_params {                      // the hot-path subset of tts_params (src/qwen3_tts.h:16-44)
    int32_t max_audio_tokens = 4096;
    float   temperature = 0.9f;
    int32_t top_k = 50;
    float   repetition_penalty = 1.05f;
    int32_t language_id = 2050;
};

struct stream_stats {
    double t_first_audio_ms = 0;           // call -> first chunk of samples delivered
    double t_total_ms = 0;
    int32_t n_frames = 0, n_chunks = 0;
    int64_t n_samples = 0;
};

struct batch_stats {
    double t_generate_ms = 0, t_decode_ms = 0, t_total_ms = 0;
    int64_t n_frames = 0, n_samples = 0;
};

using sample_sink = std::function<void(const float * samples, size_t n)>;

// text tokens -> waveform, streamed: `first_chunk_frames` frames are generated before the first decode (latency knob),
// then chunks of `chunk_frames`.  The samples handed to `sink` concatenate to decode(generate(...)).
Q3T_API bool synthesize_stream(TTSTransformer & tts, AudioTokenizerDecoder & dec, const int32_t * text_tokens, int32_t n_tokens,
                               const float * speaker_embd, const synth_params & p, int32_t first_chunk_frames, int32_t chunk_frames,
                               const sample_sink & sink, stream_stats * stats, std::string & error);

// N utterances (p.temperature <= 0: greedy; else sampled, utterance i from the stream seeded set_seed()'s seed + i): audio[i] receives utterance i's waveform
Q3T_API bool synthesize_batch(TTSTransformer & tts, AudioTokenizerDecoder & dec, const std::vector<const int32_t *> & text_tokens,
                              const std::vector<int32_t> & n_tokens, const std::vector<const float *> & speaker_embd,
                              const synth_params & p, std::vector<std::vector<float>> & audio, batch_stats * stats, std::string & error);

// the same with several decoder instances: utterance i is decoded by decoder i mod N, each decoder on its own host thread and
// CUDA stream, so the small early-stage launches of one decode overlap the large late-stage GEMMs of another (64 x 320 frames:
// 0.56 -> 0.43 s with 4 instances)
Q3T_API bool synthesize_batch(TTSTransformer & tts, const std::vector<AudioTokenizerDecoder *> & decs, const std::vector<const int32_t *> & text_tokens,
                              const std::vector<int32_t> & n_tokens, const std::vector<const float *> & speaker_embd,
                              const synth_params & p, std::vector<std::vector<float>> & audio, batch_stats * stats, std::string & error);

// the decode half alone: codes[i] (frame-major, n_frames[i] x n_codebooks) -> audio[i], utterance i on decoder i mod N (one host
// thread per decoder)
Q3T_API bool decode_batch(const std::vector<AudioTokenizerDecoder *> & decs, const std::vector<const int32_t *> & codes,
                          const std::vector<int32_t> & n_frames, std::vector<std::vector<float>> & audio, std::string & error);

} // namespace b200
} // namespace qwen3_tts
