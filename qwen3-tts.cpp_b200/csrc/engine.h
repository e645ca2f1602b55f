// Internal layout of the engine context shared by the CUDA translation units.
// Nothing here is part of the public ABI (include/q3t.h).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/q3t.h"

namespace q3t {

// ------------------------------------------------------------------ device-visible structs

// One decoder-layer stack (talker: 28 layers, code predictor: 5) in the private HBM layout:
// all layers of one matrix kind are contiguous so a CTA's row slice of layer l is
// base + l*layer_stride + row0*K.
struct StackW {
    int n_layers;
    int wtype;                 // Q3T_TYPE_F16 or Q3T_TYPE_Q8_0 for the 7 big matrices
    // F16 planes (or int8 quant planes when wtype == Q8_0); row-major [rows][K]
    const void * wqkv;         // [L][4096][1024]   rows: q(2048) | k(1024) | v(1024)
    const void * wo;           // [L][1024][2048]
    const void * wgu;          // [L][6144][1024]   row 2i = gate_i, row 2i+1 = up_i
    const void * wdown;        // [L][1024][3072]
    // Q8_0 scale planes (half), [rows][K/32]
    const __half * sqkv, * so, * sgu, * sdown;
    const float * attn_norm;   // [L][1024]
    const float * ffn_norm;    // [L][1024]
    const float * q_norm;      // [L][128]
    const float * k_norm;      // [L][128]
    const float * out_norm;    // [1024]
    __half * kcache;           // [L][8][n_ctx][128]
    __half * vcache;
    __half * vslice;           // [L][8][8][n_ctx][16]  the V history again, dealt in 16-dim slices (wide attention; talker only)
    float  * kf32;             // [L][2][8][128]  F32 shadow of positions 0,1 (code-pred prefill numerics)
    float  * vf32;
    int n_ctx;
};

struct HeadW {                 // F16 [vocab][1024]
    const __half * w;
    int vocab;
};

enum XMode : int {
    XM_BUF = 0,        // x = xsrc[0..1024) (float, device)
    XM_HIDDEN,         // x = last talker hidden (normed); also selects cb0 from the talker head
    XM_CB0_EMBD,       // x = f32(talker.codec_embd[cb0])
    XM_CP_EMBD,        // x = f32(code_pred.codec_embd[arg-1][code[arg-1]]); code[arg-1] selected from previous head
    XM_TALKER_NEXT,    // x = sum of the 16 code embeddings + trailing/pad row; selects code[14] first
    XM_SELECT_ONLY     // no stack: only select the pending token (used to close a code-predictor run)
};

struct PassDesc {
    int stack;         // 0 talker, 1 code predictor
    int xmode;
    int arg;           // code-predictor step for XM_CP_EMBD
    int pos;           // position; <0 -> GenState::n_past
    int f32attn;       // code-predictor prefill numerics (F32 K/V/Q/P)
    int head;          // -1 none, 0 talker codec_head, 1..15 -> code_pred.lm_head[head-1]
    int n_layers;      // layers to run (<= stack layers)
    int process;       // 1: apply suppress / repetition penalty to the talker logits before select
    int width;         // soft-max row width of the reference graph (n_past + n_tokens in a prefill); 0 -> pos + 1
    int nonorm;        // 1: the head reads the raw stack output (no out_norm): legacy predict_codes graph
    const float * xsrc;
    float * hidden_dst;   // optional extra copy of the normed hidden (device)
    float * logits_dst;   // optional extra copy of the raw logits (device)
};

// Generation state living in HBM, updated by CTA 0, read by all CTAs at kernel entry.
struct GenState {
    int frame;             // frames emitted so far
    int n_past;            // talker KV rows in use
    int max_len;
    int done;              // EOS seen or max_len reached
    int trailing_len;
    int eos_id, suppress_start, codec_vocab;
    float rep_penalty, temperature;
    int top_k;
    int error;             // device watchdog
    unsigned long long seed;
    unsigned long long draw;       // sampler draw counter
    unsigned ep_base;              // packet epoch of the last exchange of the previous launch
    unsigned pending_ep;           // epoch of the head packets waiting to be selected
    int cur[16];           // codes of the frame being built
    int pending;           // 1: a head result is waiting to be selected
};

// {value, epoch} activation packet: one aligned 8-byte word (see lm_kernel.cu "LL packets")
struct __align__(8) Pkt { float v; unsigned e; };

struct DevModel {
    StackW talker, cp;
    HeadW  codec_head;
    HeadW  cp_head[15];
    const __half * codec_embd;       // [3072][1024]
    const __half * cp_embd[15];      // [2048][1024]
    const float * rope;              // [n_rope_pos][128]: cos[64] | sin[64]
    int n_rope_pos;
    float eps;
    // activations / scratch (device)
    Pkt * x_qkv;                     // [4096] fused q|k|v of the token            (phase A -> B)
    Pkt * x_att;                     // [2048] attention output                    (B -> C)
    Pkt * x_x1;                      // [1024] x + attn_output                     (C -> D)
    Pkt * x_m;                       // [3072] silu(gate) * up                     (D -> E)
    Pkt * x_x;                       // [1024] layer output                        (E -> next A / head)
    Pkt * x_logit;                   // [3072] raw head logits                     (head -> select)
    Pkt * x_cand;                    // [2*grid] per-CTA arg-max candidate {value, index}
    Pkt * x_sc;                      // [16][LM_ATT_MAX_KV] attention scores of a head (wide attention: members -> members)
    int wide_min;                    // rows longer than this take the wide attention path (0 = never)
    float * hidden;                  // [1024]
    float * logits;                  // [3072]
    unsigned char * seen;            // [3072] repetition-penalty set
    const float * trailing;          // [trailing_len][1024]
    const float * prefill_x;         // [<=16][1024] prompt rows of this sequence (batched prefill: XM_BUF reads row PassDesc::arg)
    const float * pad_embed;         // [1024]
    int * codes_out;                 // [max_len][16]
    GenState * st;
    // developer profiling (q3t_profile_arm): thread 0 of CTA prof_cta appends (clock64 << 8 | tag)
    int poll_d0, poll_d1;            // tuning: cycles before the first poll of a vector / between polls
    int l2_pin;                      // code-predictor weight slices (4 per layer) whose copies carry an L2 evict_last policy (0 = no cache hints)
    int prod_pace, prod_piece;       // producer pacing: cycles between the pieces of a slice's bulk copy / ring slots per piece (0 = one copy)
    unsigned long long * prof; int prof_cap; int prof_cta; int prof_frame;
    unsigned int * prof_n;
};

constexpr int LM_THREADS = 384;       // 11 compute warps + 1 producer warp; 168 registers per thread
constexpr int Q3T_MAX_BATCH = 16;   // sequences per batched frame launch: two interleaved lock-step groups of <= 8 (lm_batch.inc)
constexpr int LM_MAXC    = 18;       // max KV chunks per kv-head in the attention phase
constexpr int LM_MIN_GRID = 140;
constexpr int LM_ATT_MAX_KV = 4608;  // KV positions the attention phase supports (P kept in shared memory)

} // namespace q3t

// ------------------------------------------------------------------ host-side context

struct q3t_host_tensor {
    int32_t type;
    int32_t src_type;    // dtype in the GGUF file (differs from `type` when the matrix was converted to F16 at load)
    int64_t ne[4];
    void * dev;          // raw upload in GGUF layout
    size_t nbytes;
};

struct q3t_vocoder;      // defined in vocoder.cu
struct q3t_spk_encoder;  // defined in spk_encoder.inc (vocoder.cu)

struct q3t_ctx {
    int device = 0;
    int sm_count = 0;
    int cc_major = 0, cc_minor = 0;
    cudaStream_t stream = nullptr;
    std::string err;

    bool have_cfg = false;
    q3t_tts_config cfg{};

    std::map<std::string, q3t_host_tensor> raw;     // uploaded tensors by GGUF name
    bool tts_ready = false;
    int converted_tensors = 0;                      // tensors converted to F16 at load (F32 / Q4_K / Q8_0-vocoder files)
    bool voc_ready = false;

    // talker / code predictor
    q3t::DevModel hm{};               // host mirror of the device model struct
    q3t::DevModel * dm = nullptr;     // device copy
    std::vector<void *> owned;        // device allocations freed on unload(1)
    int n_ctx = 0;
    void * kv_alloc = nullptr;        // talker KV (k then v)
    int grid = 0;
    q3t::PassDesc * d_passes = nullptr;   // device program buffer
    int passes_cap = 0;
    float * d_prefill = nullptr;      // staged prompt rows [<=16][1024]
    float * d_hidden_all = nullptr;   // [<=16+..][1024]
    int staged_prefill_len = 0, staged_trailing_len = 0;
    float * d_trailing = nullptr; int trailing_cap = 0;
    float * d_pad = nullptr;
    int * d_codes = nullptr; int codes_cap = 0;
    int * d_tokens = nullptr; int tokens_cap = 0;
    float * d_spk = nullptr;
    float * d_tmp = nullptr; size_t tmp_cap = 0;   // misc staging
    int64_t talker_bytes = 0, cp_bytes = 0, voc_bytes = 0;

    // vocoder
    q3t_vocoder * voc = nullptr;
    q3t_spk_encoder * spk = nullptr; bool spk_ready = false;   // speaker encoder (ECAPA x-vector)
    float * d_samples = nullptr; int64_t samples_cap = 0; int64_t last_n_samples = 0;
    int last_n_frames = 0;
    // streaming generate (q3t_stream_begin / q3t_stream_next)
    int stream_max_len = 0, stream_launched = 0, stream_fetched = 0, stream_done = 1, stream_n_passes = 0;

    q3t_timing timing{};
    int * h_done = nullptr;           // pinned flag pair polled by the generate loop
    void * h_states = nullptr;        // pinned GenState[Q3T_MAX_BATCH] read back by the batched generate loop
    void * batch_cache = nullptr;     // per-sequence buffers of q3t_generate_batch, kept between calls (engine.cu BatchCache; freed by q3t_unload(1))
    unsigned long long * d_prof = nullptr; unsigned int * d_prof_n = nullptr; int prof_cap = 0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};

// internal entry points implemented across the .cu files
namespace q3t {
int  set_err(q3t_ctx * ctx, int code, const std::string & msg);
int  cuda_check(q3t_ctx * ctx, cudaError_t e, const char * what);
int  lm_launch(q3t_ctx * ctx, const PassDesc * passes_dev, int n_passes, int frame_mode);
int  lm_max_grid(q3t_ctx * ctx);
int  lm_sample_logits(q3t_ctx * ctx, const Pkt * logits_pk_dev, const unsigned char * seen_dev, int vocab, int process,
                      float rep_penalty, float temperature, int top_k, unsigned long long seed, unsigned long long draw0,
                      int n_draws, int * out_dev);
int  lm_launch_batch(q3t_ctx * ctx, const DevModel * seqs, int nb, const PassDesc * passes_dev, int n_passes);
int  prompt_build(q3t_ctx * ctx, const int32_t * tokens_dev, int n_tokens, const float * spk_dev,
                  int language_id, float * prefill_dev, int * prefill_len, float * trailing_dev,
                  int * trailing_len, float * pad_dev);
int  prompt_project(q3t_ctx * ctx, const int32_t * tokens_dev, int n_tokens, float * out_dev);
int  vocoder_finalize(q3t_ctx * ctx);
int  spk_finalize(q3t_ctx * ctx);
void spk_release(q3t_ctx * ctx);
int  spk_run(q3t_ctx * ctx, const float * samples_dev, int n, float * out_dev, int * n_frames_out);
void vocoder_free(q3t_ctx * ctx);
int64_t vocoder_run(q3t_ctx * ctx, const int32_t * codes_dev, int n_frames, float ** samples_dev, int t_from);
}
