/*
 * q3t.h — C ABI of the B200-native codes-to-audio engine (libq3t_b200.so).
 *
 * This is the drop-in boundary for the hot path of clawd20130/qwen3-tts.cpp:
 * plain pointers and sizes, no C++/torch types.  Two layers are exported from
 * the same shared library:
 *
 *   (1) q3t_*   — the device engine (CUDA translation units).  This is what the
 *       reference's ggml graph build/alloc/compute sequence is replaced by.
 *       Each entry point cites the reference interface it stands in for
 *       (paths relative to the reference repository root).
 *
 *   (2) q3t_tts_* / q3t_dec_* / q3t_gguf_* — a C view of the C++ host classes
 *       qwen3_tts::TTSTransformer / AudioTokenizerDecoder / GGUFLoader that
 *       this repo re-implements with the reference's method names, argument
 *       meaning and error behaviour (src/tts_transformer.h:190-280,
 *       src/audio_tokenizer_decoder.h:159-179, src/gguf_loader.h:16-81).  It
 *       exists so non-C++ callers (ctypes, cgo, JNI ...) can bind the same API.
 *
 * All functions return 0 on success and a negative q3t_status otherwise unless
 * stated; q3t_last_error()/q3t_*_get_error() give the message.  Pointers are
 * HOST pointers unless the parameter name ends in _dev.  Outputs are caller
 * allocated.  One context / object is bound to one CUDA device and must be
 * used from one thread at a time (same rule as the reference, AGENTS.md:44-47).
 * There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef Q3T_H
#define Q3T_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define Q3T_API __declspec(dllexport)
#else
#define Q3T_API __attribute__((visibility("default")))
#endif

typedef enum q3t_status {
    Q3T_OK            =  0,
    Q3T_ERR_ARG       = -1,  /* bad argument                          */
    Q3T_ERR_CUDA      = -2,  /* CUDA runtime error / no device        */
    Q3T_ERR_STATE     = -3,  /* call order (weights not finalized...) */
    Q3T_ERR_CTX_FULL  = -4,  /* "Context length exceeded"             */
    Q3T_ERR_IO        = -5,
    Q3T_ERR_FORMAT    = -6,  /* malformed GGUF / unsupported dtype    */
    Q3T_ERR_DEVICE    = -7   /* device-side watchdog tripped          */
} q3t_status;

/* ggml type ids as stored in GGUF (ggml/include/ggml.h:390-398) */
enum { Q3T_TYPE_F32 = 0, Q3T_TYPE_F16 = 1, Q3T_TYPE_Q8_0 = 8, Q3T_TYPE_Q4_K = 12 };

typedef struct q3t_ctx q3t_ctx;

/* Mirror of tts_transformer_config (src/tts_transformer.h:58-99), POD. */
typedef struct q3t_tts_config {
    int32_t text_vocab_size, text_embd_dim;
    int32_t hidden_size, n_layers, n_attention_heads, n_key_value_heads;
    int32_t intermediate_size, head_dim;
    float   rms_norm_eps, rope_theta;
    int32_t codec_vocab_size, n_codebooks;
    int32_t code_pred_layers, code_pred_vocab_size;
    int32_t codec_pad_id, codec_bos_id, codec_eos_id;
    int32_t tts_bos_token_id, tts_eos_token_id, tts_pad_token_id;
    int32_t codec_think_id, codec_nothink_id, codec_think_bos_id, codec_think_eos_id;
    int32_t english_language_id;
} q3t_tts_config;

/* Per-phase device timings of the last generate (CUDA events), the counterpart of
 * struct tts_timing (src/tts_transformer.h:19-53). */
typedef struct q3t_timing {
    double prefill_ms, frames_ms, vocoder_ms;
    int32_t n_frames;
    int64_t kernel_launches;
} q3t_timing;

/* ------------------------------------------------------------------ engine */

/* replaces init_preferred_backend()/ggml_backend_sched_new (src/gguf_loader.cpp:27-58,
 * src/tts_transformer.cpp:123-150): binds a CUDA device, creates the stream. */
Q3T_API int  q3t_ctx_create(int device, q3t_ctx ** out);
Q3T_API void q3t_ctx_destroy(q3t_ctx * ctx);
Q3T_API const char * q3t_last_error(q3t_ctx * ctx);   /* ctx may be NULL for create errors */
Q3T_API int  q3t_device_info(q3t_ctx * ctx, int32_t * sm_count, int32_t * cc_major, int32_t * cc_minor);

/* replaces parse_config (src/tts_transformer.cpp:212-344): must precede uploads. */
Q3T_API int  q3t_set_tts_config(q3t_ctx * ctx, const q3t_tts_config * cfg);

/* replaces ggml_backend_tensor_set per tensor in load_tensor_data
 * (src/tts_transformer.cpp:648-677) and load_tensor_data_from_file
 * (src/gguf_loader.cpp:154-227).  `name` is the GGUF tensor name, `type` a Q3T_TYPE_*,
 * `ne` the ggml shape (ne[0] fastest).  Unknown names are ignored (returns 1).  F16 everywhere and Q8_0 talker /
 * code-predictor matrices are consumed natively (bit-exact paths); F32 and Q4_K matrices and Q8_0 vocoder matrices
 * (the other outputs of the reference's converters) are converted to F16 on the host here. */
Q3T_API int  q3t_upload_tensor(q3t_ctx * ctx, const char * name, int32_t type,
                               const int64_t ne[4], const void * data, size_t nbytes);
/* private device repack after the last upload (fused QKV / gate-up planes, Q8_0 split into
 * quant + scale planes, RoPE tables).  `which`: 1 = talker+code predictor, 2 = vocoder, 3 = speaker encoder
 * (q3t_unload takes the bit mask 1 | 2 | 4). */
Q3T_API int  q3t_finalize_weights(q3t_ctx * ctx, int32_t which);
Q3T_API int  q3t_unload(q3t_ctx * ctx, int32_t which);

/* init_kv_cache / clear_kv_cache (src/tts_transformer.cpp:685-737) */
Q3T_API int  q3t_kv_init(q3t_ctx * ctx, int32_t n_ctx);
Q3T_API int  q3t_kv_n_ctx(q3t_ctx * ctx);

/* forward_prefill (src/tts_transformer.cpp:1831-1961): embd f32[n_tokens*H] ->
 * hidden_out f32[n_tokens*H] (may be NULL), logits_out f32[codec_vocab] of the last row. */
Q3T_API int  q3t_talker_prefill(q3t_ctx * ctx, const float * embd, int32_t n_tokens, int32_t n_past,
                                float * hidden_out, float * logits_out);
/* forward_step (src/tts_transformer.cpp:1993-2108) */
Q3T_API int  q3t_talker_step(q3t_ctx * ctx, const float * embd, int32_t n_past,
                             float * logits_out, float * hidden_out);
/* predict_codes_autoregressive (src/tts_transformer.cpp:2310-2575): hidden f32[H], cb0 ->
 * codes15_out int32[15].  temperature<=0 = greedy.  `seed` drives the device sampler
 * (the reference's rng_ is unseedable, src/tts_transformer.h:336). */
Q3T_API int  q3t_codepred_frame(q3t_ctx * ctx, const float * hidden, int32_t cb0,
                                float temperature, int32_t top_k, uint64_t seed, int32_t * codes15_out);
/* The token selection of sample_or_argmax / the cb0 post-processing in isolation (src/tts_transformer.cpp:2349-2387,
 * :2664-2722): n_draws tokens drawn from ONE logit vector f32[vocab] (vocab <= 3072) by the very device routine the
 * decode kernels run after a head: temperature, top-k threshold (k-th largest, ties kept, `<` test), soft-max with a
 * double-precision sum, inverse-CDF draw number d of stream `seed`.  process != 0 first applies the codebook-0 rules:
 * ids >= codec_vocab-1024 except EOS set to -inf, HF repetition penalty over the ids in seen_ids.  Exists so the
 * distribution can be tested against the reference's; temperature must be > 0. */
Q3T_API int  q3t_sample_logits(q3t_ctx * ctx, const float * logits, int32_t vocab, int32_t process,
                               const int32_t * seen_ids, int32_t n_seen, float repetition_penalty,
                               float temperature, int32_t top_k, uint64_t seed, int32_t n_draws, int32_t * tokens_out);
/* predict_codes, the legacy parallel-heads formulation (src/tts_transformer.cpp:2130-2180,
 * graph :1413-1532): logits_out f32[15*code_pred_vocab]. prev_codes may be NULL. */
Q3T_API int  q3t_codepred_parallel(q3t_ctx * ctx, const float * hidden, const int32_t * prev_codes,
                                   float * logits_out);
/* build_prefill_graph (src/tts_transformer.cpp:984-1122): prompt assembly on device.
 * Outputs may be NULL; prefill_len/trailing_len are always written. */
Q3T_API int  q3t_build_prompt(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens,
                              const float * speaker_embd, int32_t language_id,
                              float * prefill_out, int32_t * prefill_len,
                              float * trailing_out, int32_t * trailing_len, float * pad_out);
/* project_text_tokens (src/tts_transformer.cpp:917-982): out f32[n_tokens*H] */
Q3T_API int  q3t_project_text(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens, float * out);
/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * computation and bit-identical result as q3t_project_text, but it allocates NOTHING: tokens and output
 * are DEVICE pointers (int32[n_tokens], f32[n_tokens*H]) supplied by the caller, and the intermediates
 * live in a file-scope `static __device__` array of the library (1.14 MiB, held by every GPU that loads
 * the library until exit; its device address is resolved once per device with cudaGetSymbolAddress).
 * Limit: n_tokens <= q3t_project_text_static_max_tokens() (64) and text_embd_dim <= 2048; larger sizes
 * fail with Q3T_ERR_ARG.  Token ids are range-checked on the device (Q3T_ERR_ARG, the bad row is zero).
 * All calls on one GPU share the array, which keeps the previous call's contents (everything that is read
 * is rewritten first, the error flag is cleared per call): calls that overlap on the GPU RACE - serialise
 * them.  The call is synchronous (returns after the work on the engine stream has finished). */
Q3T_API int  q3t_project_text_static_dev(q3t_ctx * ctx, const int32_t * text_tokens_dev, int32_t n_tokens, float * out_dev);
Q3T_API int  q3t_project_text_static_max_tokens(void);
/* lookup_embedding_rows on talker.codec_embd (src/tts_transformer.cpp:793-915) */
Q3T_API int  q3t_codec_embed(q3t_ctx * ctx, const int32_t * ids, int32_t n, float * out);

/* TTSTransformer::generate (src/tts_transformer.cpp:2577-2851), whole loop on device.
 * codes_out int32[max_len*n_codebooks] frame-major; *n_frames_out = frames produced
 * (EOS stops before emitting a frame; max_len<=0 -> 0 frames, success). */
Q3T_API int  q3t_generate(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens,
                          const float * speaker_embd, int32_t max_len, int32_t language_id,
                          float repetition_penalty, float temperature, int32_t top_k, uint64_t seed,
                          int32_t * codes_out, int32_t * n_frames_out);

/* Streaming form of q3t_generate (extension; the reference pipeline hands codes to the vocoder only
 * when the utterance is complete, src/qwen3_tts.cpp:345-396).  begin: prompt + prefill; next: up to max_frames more frames,
 * codes_out int32[max_frames*n_codebooks] receives the new frames (*n_new_out of them), *done_out = 1 after EOS / max_len.
 * The concatenation over the calls equals what q3t_generate returns for the same arguments, for any chunking. */
Q3T_API int  q3t_stream_begin(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                              int32_t max_len, int32_t language_id, float repetition_penalty, float temperature,
                              int32_t top_k, uint64_t seed);
Q3T_API int  q3t_stream_next(q3t_ctx * ctx, int32_t max_frames, int32_t * codes_out, int32_t * n_new_out, int32_t * done_out);

/* Batched generate (extension; SURVEY.md §8b "batched variants take n_seq + per-seq arrays"): n_seq
 * independent utterances, greedy, same semantics per utterance as q3t_generate with temperature 0
 * (src/tts_transformer.cpp:2577-2851).  Up to 16 sequences per kernel launch (two interleaved lock-step groups of 8);
 * codes_out[i] int32[max_len*n_codebooks], n_frames_out[i] = frames of utterance i.
 * speaker_embd may be null (no speaker for any utterance) or hold per-utterance pointers (null entries allowed).
 * Waves whose sequences all have the same number of prompt rows are also prefilled in lock-step.  The per-sequence device
 * buffers of a call (KV caches, packets, code rows: ~80 MB per sequence at max_len 320, 16 sequences) stay allocated in the
 * context for the next call and are released by q3t_unload(ctx, 1) / q3t_ctx_destroy. */
Q3T_API int  q3t_generate_batch(q3t_ctx * ctx, int32_t n_seq, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                float repetition_penalty, int32_t * const * codes_out, int32_t * n_frames_out);

/* The same with the sampled selection of q3t_generate (temperature > 0: top-k threshold with ties, double-precision
 * normalisation, src/tts_transformer.cpp:2349-2387, :2664-2722): utterance i draws from the stream seeded `seed + i`,
 * i.e. its codes equal what q3t_generate returns for it with that seed.  temperature <= 0 = q3t_generate_batch. */
Q3T_API int  q3t_generate_batch_sampled(q3t_ctx * ctx, int32_t n_seq, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                        const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                        float repetition_penalty, float temperature, int32_t top_k, uint64_t seed,
                                        int32_t * const * codes_out, int32_t * n_frames_out);

/* AudioTokenizerDecoder::decode (src/audio_tokenizer_decoder.cpp:804-879):
 * codes int32[n_frames*16] frame-major -> samples f32; returns sample count
 * (1920*n_frames-555) or a negative status.  cap = capacity of samples_out in floats. */
Q3T_API int64_t q3t_vocoder_decode(q3t_ctx * ctx, const int32_t * codes, int32_t n_frames,
                                   float * samples_out, int64_t cap);
Q3T_API int64_t q3t_vocoder_n_samples(int32_t n_frames);
/* Streaming form of the same decode (the reference pipeline decodes once per utterance,
 * src/qwen3_tts.cpp:384-396).  `codes` holds ALL n_frames frames received so far, n_done of which were
 * decoded by earlier calls; samples_out receives the samples that became final, i.e. global samples
 * [n(n_done), n(n_frames)) with n(T) = 1920*T-555, n(0) = 0.  Every layer of the decoder is causal
 * (src/audio_tokenizer_decoder.cpp:455,511,562,706,759), so the concatenation over the calls is the
 * waveform decode() returns for the whole utterance.  Returns the sample count or a negative status. */
Q3T_API int64_t q3t_vocoder_decode_from(q3t_ctx * ctx, const int32_t * codes, int32_t n_frames, int32_t n_done,
                                        float * samples_out, int64_t cap);

/* AudioTokenizerEncoder::encode (src/audio_tokenizer_encoder.cpp:703-760, mel front end :288-373, ECAPA-TDNN graph :445-701):
 * samples f32[n_samples] at 24 kHz in [-1, 1] -> embedding_out f32[1024] (the x-vector of prefill row 7).  Needs the
 * spk_enc.* tensors uploaded and q3t_finalize_weights(ctx, 3).  *n_frames_out (optional) = mel frames. */
Q3T_API int  q3t_speaker_encode(q3t_ctx * ctx, const float * samples, int32_t n_samples, float * embedding_out, int32_t * n_frames_out);

/* Device-resident variants used by bench.py `value` (inputs already in HBM):
 * run generate+decode for a prompt that was staged with q3t_stage_prompt, leave the
 * result on the device, report frames and samples. */
Q3T_API int  q3t_stage_prompt(q3t_ctx * ctx, const int32_t * text_tokens, int32_t n_tokens,
                              const float * speaker_embd, int32_t language_id);
Q3T_API int  q3t_run_staged(q3t_ctx * ctx, int32_t max_len, float repetition_penalty,
                            float temperature, int32_t top_k, uint64_t seed, int32_t with_vocoder,
                            int32_t * n_frames_out, int64_t * n_samples_out);
Q3T_API int  q3t_fetch_codes(q3t_ctx * ctx, int32_t * codes_out, int32_t n_frames);
Q3T_API int64_t q3t_fetch_samples(q3t_ctx * ctx, float * samples_out, int64_t cap);

/* Repeats the talker decode step `iters` times at position n_past (KV must hold >= n_past
 * rows) and returns the mean device time per step in microseconds (CUDA events on the
 * engine stream).  Used for the roofline figure; does not advance any state. */
Q3T_API int  q3t_bench_talker_step(q3t_ctx * ctx, int32_t n_past, int32_t iters, float * us_per_step);
Q3T_API int  q3t_get_timing(q3t_ctx * ctx, q3t_timing * out);
/* Developer aid (no reference counterpart): in-kernel phase profiler of the persistent decode
 * kernel.  arm: thread 0 of every CTA keeps up to `cap` (clock64 << 8 | tag) records of the launch
 * whose frame counter equals `cta` (< 0: the last launch), cap <= 0 disarms; read: copies the records of the selected CTA out, returns the count. */
Q3T_API int  q3t_profile_arm(q3t_ctx * ctx, int32_t cta, int32_t cap);
Q3T_API int  q3t_profile_read(q3t_ctx * ctx, uint64_t * out, int32_t cap);
Q3T_API int  q3t_profile_select(q3t_ctx * ctx, int32_t cta);   /* which CTA q3t_profile_read returns */
Q3T_API int64_t q3t_weight_bytes(q3t_ctx * ctx, int32_t which); /* 1 talker stack, 2 code-pred stack (one pass), 3 vocoder */

/* --------------------------------------------------- host classes, C view */

typedef struct q3t_gguf q3t_gguf;              /* qwen3_tts::GGUFLoader (src/gguf_loader.h:16-62) */
Q3T_API q3t_gguf * q3t_gguf_open(const char * path);
Q3T_API void       q3t_gguf_close(q3t_gguf * g);
Q3T_API const char * q3t_gguf_get_error(q3t_gguf * g);
Q3T_API int64_t    q3t_gguf_get_n_tensors(q3t_gguf * g);
Q3T_API const char * q3t_gguf_get_tensor_name(q3t_gguf * g, int64_t idx);
Q3T_API int32_t    q3t_gguf_get_tensor_type(q3t_gguf * g, int64_t idx);
Q3T_API size_t     q3t_gguf_get_tensor_offset(q3t_gguf * g, int64_t idx);
Q3T_API size_t     q3t_gguf_get_tensor_size(q3t_gguf * g, int64_t idx);
Q3T_API int32_t    q3t_gguf_get_tensor_ne(q3t_gguf * g, int64_t idx, int64_t ne_out[4]);
Q3T_API int32_t    q3t_gguf_get_u32(q3t_gguf * g, const char * key, int32_t default_val);
Q3T_API float      q3t_gguf_get_f32(q3t_gguf * g, const char * key, float default_val);
Q3T_API size_t     q3t_gguf_get_data_offset(q3t_gguf * g);

typedef struct q3t_tts q3t_tts;                /* qwen3_tts::TTSTransformer */
Q3T_API q3t_tts * q3t_tts_new(void);
Q3T_API void      q3t_tts_delete(q3t_tts * t);
Q3T_API int  q3t_tts_load_model(q3t_tts * t, const char * model_path);      /* 1 = true, 0 = false */
Q3T_API void q3t_tts_unload_model(q3t_tts * t);
Q3T_API int  q3t_tts_init_kv_cache(q3t_tts * t, int32_t n_ctx);
Q3T_API void q3t_tts_clear_kv_cache(q3t_tts * t);
Q3T_API int  q3t_tts_forward_prefill(q3t_tts * t, const float * prefill_embd, int32_t n_tokens, int32_t n_past,
                                     float * hidden_out, float * logits_out);
Q3T_API int  q3t_tts_forward_step(q3t_tts * t, const float * step_embd, int32_t n_past,
                                  float * logits_out, float * hidden_out);
Q3T_API int  q3t_tts_forward_text(q3t_tts * t, const int32_t * text_tokens, int32_t n_tokens,
                                  const float * speaker_embd, int32_t n_past, float * hidden_out);
Q3T_API int  q3t_tts_forward_codec(q3t_tts * t, int32_t codec_token, int32_t n_past, float * logits_out);
Q3T_API int  q3t_tts_get_hidden_states(q3t_tts * t, float * hidden_out);
Q3T_API int  q3t_tts_predict_codes(q3t_tts * t, const float * hidden, const int32_t * prev_codes, float * logits_out);
Q3T_API int  q3t_tts_predict_codes_autoregressive(q3t_tts * t, const float * hidden, int32_t codebook_0_token,
                                                  int32_t * codes15_out, float temperature, int32_t top_k);
Q3T_API int  q3t_tts_generate(q3t_tts * t, const int32_t * text_tokens, int32_t n_tokens,
                              const float * speaker_embd, int32_t max_len,
                              int32_t * codes_out, int32_t * n_frames_out,
                              int32_t language_id, float repetition_penalty, float temperature, int32_t top_k);
Q3T_API int  q3t_tts_generate_batch(q3t_tts * t, int32_t n, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                    const float * const * speaker_embd, int32_t max_len,
                                    int32_t * const * codes_out, int32_t * n_frames_out,
                                    int32_t language_id, float repetition_penalty);   /* extension: batched greedy generate */
Q3T_API int  q3t_tts_generate_batch_sampled(q3t_tts * t, int32_t n, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                            const float * const * speaker_embd, int32_t max_len,
                                            int32_t * const * codes_out, int32_t * n_frames_out, int32_t language_id,
                                            float repetition_penalty, float temperature, int32_t top_k);   /* extension: the same, sampled (utterance i: set_seed's seed + i) */
Q3T_API void q3t_tts_set_seed(q3t_tts * t, uint64_t seed);   /* extension: the reference has no seed API */
Q3T_API int  q3t_tts_get_config(q3t_tts * t, q3t_tts_config * out);
Q3T_API const char * q3t_tts_get_error(q3t_tts * t);
Q3T_API q3t_ctx * q3t_tts_engine(q3t_tts * t);                /* the engine context behind the object */

typedef struct q3t_enc q3t_enc;                /* qwen3_tts::AudioTokenizerEncoder (src/audio_tokenizer_encoder.h:104-150) */
Q3T_API q3t_enc * q3t_enc_new(void);
Q3T_API void      q3t_enc_delete(q3t_enc * e);
Q3T_API int  q3t_enc_load_model(q3t_enc * e, const char * model_path);                 /* the TTS GGUF (spk_enc.* tensors) */
Q3T_API int  q3t_enc_encode(q3t_enc * e, const float * samples, int32_t n_samples, float * embedding_out /* f32[1024] */);
Q3T_API const char * q3t_enc_get_error(q3t_enc * e);

typedef struct q3t_dec q3t_dec;                /* qwen3_tts::AudioTokenizerDecoder */
Q3T_API q3t_dec * q3t_dec_new(void);
Q3T_API void      q3t_dec_delete(q3t_dec * d);
Q3T_API int  q3t_dec_load_model(q3t_dec * d, const char * model_path);
Q3T_API void q3t_dec_unload_model(q3t_dec * d);
Q3T_API int64_t q3t_dec_decode(q3t_dec * d, const int32_t * codes, int32_t n_frames, float * samples_out, int64_t cap);
/* extension: streaming decode (decode_append / stream_reset of the host class) */
Q3T_API void q3t_dec_stream_reset(q3t_dec * d);
Q3T_API int64_t q3t_dec_decode_append(q3t_dec * d, const int32_t * codes, int32_t n_new_frames, float * samples_out, int64_t cap);
Q3T_API const char * q3t_dec_get_error(q3t_dec * d);
Q3T_API q3t_ctx * q3t_dec_engine(q3t_dec * d);

/* ------------------------------------------- streaming / batched front door (extensions)
 * The reference pipeline synthesises one utterance at a time and decodes after the last frame
 * (Qwen3TTS::synthesize_internal, src/qwen3_tts.cpp:278-442).  C view of host/b200_pipeline.h. */
typedef struct q3t_stream_stats {
    double  t_first_audio_ms;     /* call -> first chunk of samples available */
    double  t_total_ms;
    int32_t n_frames, n_chunks;
    int64_t n_samples;
} q3t_stream_stats;
typedef struct q3t_batch_stats {
    double  t_generate_ms, t_decode_ms, t_total_ms;
    int64_t n_frames, n_samples;
} q3t_batch_stats;
/* generate() in chunks (TTSTransformer::generate_begin / generate_next); 1 = ok, 0 = failed (q3t_tts_get_error) */
Q3T_API int  q3t_tts_generate_begin(q3t_tts * t, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                                    int32_t max_len, int32_t language_id, float repetition_penalty, float temperature, int32_t top_k);
Q3T_API int  q3t_tts_generate_next(q3t_tts * t, int32_t max_frames, int32_t * codes_out, int32_t * n_new_out, int32_t * done_out);
/* text tokens -> waveform with the vocoder fed chunk by chunk (first_chunk_frames, then chunk_frames frames at a time);
 * the samples equal decode(generate(...)); returns the sample count or -1 (q3t_pipe_get_error). */
Q3T_API int64_t q3t_pipe_synthesize_stream(q3t_tts * t, q3t_dec * d, const int32_t * text_tokens, int32_t n_tokens,
                                           const float * speaker_embd, int32_t max_len, int32_t language_id,
                                           float repetition_penalty, float temperature, int32_t top_k,
                                           int32_t first_chunk_frames, int32_t chunk_frames,
                                           float * samples_out, int64_t cap, q3t_stream_stats * stats_out);
/* n utterances through the batched engine (greedy) + one decode each; samples_out[i] has room for caps[i] floats;
 * 1 = ok, 0 = failed (q3t_pipe_get_error). */
Q3T_API int  q3t_pipe_synthesize_batch(q3t_tts * t, q3t_dec * d, int32_t n, const int32_t * const * text_tokens, const int32_t * n_tokens,
                                       const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                       float repetition_penalty, float * const * samples_out, const int64_t * caps,
                                       int64_t * n_samples_out, q3t_batch_stats * stats_out);
/* the same with n_dec decoder instances (utterance i -> decoder i mod n_dec, one host thread and CUDA stream each) */
Q3T_API int  q3t_pipe_synthesize_batch_multi(q3t_tts * t, q3t_dec * const * decs, int32_t n_dec, int32_t n, const int32_t * const * text_tokens,
                                             const int32_t * n_tokens, const float * const * speaker_embd, int32_t max_len, int32_t language_id,
                                             float repetition_penalty, float * const * samples_out, const int64_t * caps,
                                             int64_t * n_samples_out, q3t_batch_stats * stats_out);
/* the decode half alone: codes[i] int32[n_frames[i]*16] -> samples_out[i] (room for caps[i] floats), utterance i on decoder i mod n_dec */
Q3T_API int  q3t_pipe_decode_batch_multi(q3t_dec * const * decs, int32_t n_dec, int32_t n, const int32_t * const * codes, const int32_t * n_frames,
                                         float * const * samples_out, const int64_t * caps, int64_t * n_samples_out);
Q3T_API const char * q3t_pipe_get_error(void);

#ifdef __cplusplus
}
#endif
#endif /* Q3T_H */
