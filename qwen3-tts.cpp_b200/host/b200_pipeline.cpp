#include "b200_pipeline.h"

#include <thread>

#include <algorithm>
#include <chrono>

namespace qwen3_tts {
namespace b200 {

namespace {
using clk = std::chrono::steady_clock;
double ms_since(clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); }
}

bool synthesize_stream(TTSTransformer & tts, AudioTokenizerDecoder & dec, const int32_t * text_tokens, int32_t n_tokens,
                       const float * speaker_embd, const synth_params & p, int32_t first_chunk_frames, int32_t chunk_frames,
                       const sample_sink & sink, stream_stats * stats, std::string & error) {
    const auto t0 = clk::now();
    stream_stats st;
    if (first_chunk_frames <= 0) first_chunk_frames = 8;
    if (chunk_frames <= 0) chunk_frames = 32;
    if (!tts.generate_begin(text_tokens, n_tokens, speaker_embd, p.max_audio_tokens, p.language_id, p.repetition_penalty,
                            p.temperature, p.top_k)) { error = tts.get_error(); return false; }
    dec.stream_reset();
    std::vector<int32_t> codes;
    std::vector<float> samples;
    bool done = p.max_audio_tokens <= 0;
    while (!done) {
        if (!tts.generate_next(st.n_chunks == 0 ? first_chunk_frames : chunk_frames, codes, done)) { error = tts.get_error(); return false; }
        const int32_t nf = (int32_t) (codes.size() / (size_t) tts.get_config().n_codebooks);
        if (nf == 0) continue;
        if (!dec.decode_append(codes.data(), nf, samples)) { error = dec.get_error(); return false; }
        if (st.n_chunks == 0) st.t_first_audio_ms = ms_since(t0);
        if (sink) sink(samples.data(), samples.size());
        st.n_frames += nf; st.n_chunks += 1; st.n_samples += (int64_t) samples.size();
    }
    dec.stream_reset();
    st.t_total_ms = ms_since(t0);
    if (stats) *stats = st;
    return true;
}

bool synthesize_batch(TTSTransformer & tts, AudioTokenizerDecoder & dec, const std::vector<const int32_t *> & text_tokens,
                      const std::vector<int32_t> & n_tokens, const std::vector<const float *> & speaker_embd,
                      const synth_params & p, std::vector<std::vector<float>> & audio, batch_stats * stats, std::string & error) {
    std::vector<AudioTokenizerDecoder *> decs{&dec};
    return synthesize_batch(tts, decs, text_tokens, n_tokens, speaker_embd, p, audio, stats, error);
}

bool synthesize_batch(TTSTransformer & tts, const std::vector<AudioTokenizerDecoder *> & decs, const std::vector<const int32_t *> & text_tokens,
                      const std::vector<int32_t> & n_tokens, const std::vector<const float *> & speaker_embd,
                      const synth_params & p, std::vector<std::vector<float>> & audio, batch_stats * stats, std::string & error) {
    if (decs.empty()) { error = "no decoder"; return false; }
    const auto t0 = clk::now();
    batch_stats st;
    std::vector<std::vector<int32_t>> codes;
    if (!tts.generate_batch(text_tokens, n_tokens, speaker_embd, p.max_audio_tokens, codes, p.language_id, p.repetition_penalty,
                            p.temperature, p.top_k)) {
        error = tts.get_error(); return false;
    }
    st.t_generate_ms = ms_since(t0);
    const auto t1 = clk::now();
    audio.assign(codes.size(), {});
    const int ncb = tts.get_config().n_codebooks;
    std::vector<const int32_t *> cp(codes.size());
    std::vector<int32_t> nfv(codes.size());
    for (size_t i = 0; i < codes.size(); ++i) { cp[i] = codes[i].data(); nfv[i] = (int32_t) (codes[i].size() / (size_t) ncb); }
    if (!decode_batch(decs, cp, nfv, audio, error)) return false;
    for (size_t i = 0; i < codes.size(); ++i) {
        st.n_frames += (int64_t) (codes[i].size() / (size_t) ncb);
        st.n_samples += (int64_t) audio[i].size();
    }
    st.t_decode_ms = ms_since(t1);
    st.t_total_ms = ms_since(t0);
    if (stats) *stats = st;
    return true;
}

bool decode_batch(const std::vector<AudioTokenizerDecoder *> & decs, const std::vector<const int32_t *> & codes,
                  const std::vector<int32_t> & n_frames, std::vector<std::vector<float>> & audio, std::string & error) {
    if (decs.empty() || codes.size() != n_frames.size()) { error = "bad argument"; return false; }
    audio.assign(codes.size(), {});
    // decoder k takes utterances k, k + N, ...; one host thread per decoder (a decoder object is not thread-safe, its own calls stay serial)
    const size_t nd = decs.size();
    std::vector<std::string> errs(nd);
    auto work = [&](size_t k) {
        for (size_t i = k; i < codes.size(); i += nd) {
            if (n_frames[i] <= 0) continue;
            if (!decs[k]->decode(codes[i], n_frames[i], audio[i])) { errs[k] = decs[k]->get_error(); return; }
        }
    };
    if (nd == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (size_t k = 0; k < nd; ++k) th.emplace_back(work, k);
        for (auto & x : th) x.join();
    }
    for (const auto & e : errs) if (!e.empty()) { error = e; return false; }
    return true;
}

} // namespace b200
} // namespace qwen3_tts
