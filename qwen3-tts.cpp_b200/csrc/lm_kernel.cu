// Persistent decode kernel for the talker and the code predictor (sm_100a).
//
// What it replaces in the reference: build_step_graph / build_code_pred_prefill_graph /
// build_code_pred_step_graph and the host loop that runs 16 dependent ggml graphs per frame
// (src/tts_transformer.cpp:1268-1411, 1534-1829, 2310-2575, 2664-2804).
//
// Design (B200-first):
//   * ONE cooperative launch of `grid` (= #SMs) CTAs x 512 threads runs a whole *program* of
//     passes (a pass = one token through one layer stack + optional head): a frame is
//     16 code-predictor passes + 1 talker pass, a prefill is M talker passes.
//   * Batch-1 decode is HBM/latency bound, so the weight stream is decoupled from the
//     dependency chain in two stages: one thread per CTA issues TMA-unit bulk L2 prefetches
//     (cp.async.bulk.prefetch.L2) one layer ahead, and a producer warp issues TMA bulk copies
//     (cp.async.bulk global->shared, mbarrier complete_tx) of the CTA's weight-row slice of the
//     phase after next into one of two shared-memory buffers *before* the grid barrier, so
//     barrier latency is covered by copies in flight and the compute reads only shared memory.
//   * Every phase boundary is an all-to-all dependency (GEMV output -> next GEMV input); it is
//     a release/acquire counter barrier in global memory, not a kernel boundary.
//   * BIT-EXACT with the oracle: F16 rounding points amplify any F32 re-association noise into
//     ~1e-3 logit differences, enough to flip greedy tokens of a random-init model within a few
//     hundred steps.  So every dot product is accumulated in the order of ggml-cpu's AVX2+FMA
//     build (32 accumulator lanes k%32, FMA chain over k/32, fixed reduction tree;
//     ggml/src/ggml-cpu/vec.cpp:265-390, simd-mappings.h:549-567), exp() is the same
//     polynomial (vec.h:1230-1267), soft-max sums follow the 8-lane pattern (vec.cpp:558-568).
//     oracle/q3t_oracle.c is the executable statement of these rules.
#include "common.cuh"
#include "engine.h"

#include <math_constants.h>

namespace q3t {

// ------------------------------------------------------------------------------- shared memory
constexpr int PAD_B      = 0;                          // rows are dense: chain-major rows are read 512 contiguous bytes per warp
constexpr int PITCH_1K   = 1024 * 2 + PAD_B;           // 2112
constexpr int PITCH_2K   = 2048 * 2 + PAD_B;           // 4160
constexpr int PITCH_3K   = 3072 * 2 + PAD_B;           // 6208
constexpr int MAXR_GU    = 44;                         // gate/up rows per CTA   (grid >= 140)
constexpr int MAXR_H     = 8;                          // rows of a 1024-row matrix per CTA
constexpr int BUF0_BYTES = MAXR_GU * PITCH_1K;         // qkv / gate-up / head slices     92,928 B
constexpr int BUF1_BYTES = MAXR_H * PITCH_3K;          // attn_output / ffn_down slices   49,664 B
constexpr int ATT_MAX_KV = LM_ATT_MAX_KV;              // positions the attention phase can hold (P in F16)
constexpr int VT_HALVES  = 8192;                       // V^T tile: nd dims x TB positions

struct __align__(128) Smem {
    unsigned char buf0[BUF0_BYTES];
    unsigned char buf1[BUF1_BYTES];
    float  xres[H];                // residual stream of the current token (x, then x1)
    float  xf[FFN];                // GEMV input as F32 (already F16-rounded where the oracle rounds)
    float  outv[96];               // finalised row values of the phase
    // attention
    float  qf[2][HD];              // normed+roped q of the two heads of this kv group (F32)
    float  qr[2][HD];              // same, F16-rounded, as float
    float  knew_f[HD], vnew_f[HD];
    __half knew_h[HD], vnew_h[HD];
    __half ph[2][ATT_MAX_KV];      // soft-max probabilities rounded to F16 (oracle D.4)
    __half vt[VT_HALVES];          // V^T tile [dim][position]
    double dred[2][16];
    float  fred[32];
    int    ired[32];
    int    tok;
    unsigned long long mbar[2];    // TMA completion barriers of buf0 / buf1
};

// ------------------------------------------------------------------------------- phase profiler
__device__ __forceinline__ void prof_mark(const DevModel & m, int tag) {
    if (m.prof && threadIdx.x == 0 && (int) blockIdx.x == m.prof_cta) {
        const unsigned i = atomicAdd(m.prof_n, 1u);
        if ((int) i < m.prof_cap) m.prof[i] = ((unsigned long long) clock64() << 8) | (unsigned long long) (tag & 0xff);
    }
}
#define PROF(tag) prof_mark(m, tag)

// ------------------------------------------------------------------------------- grid barriers
struct BarState {
    unsigned long long target;       // grid barrier epoch target
    unsigned long long gtarget;      // group barrier target of this CTA's kv group
    bool dead;
};

__device__ __forceinline__ bool spin_until(const unsigned long long * ctr, unsigned long long target, GenState * st) {
    long long t0 = clock64();
    unsigned spins = 0;
    while (ld_acquire_u64(ctr) < target) {
        if (((++spins) & 0x3ff) == 0) {
            if (clock64() - t0 > 4000000000LL) {      // ~2 s: another CTA died; bail out loudly
                st->error = 1;
                return false;
            }
        }
    }
    return true;
}

__device__ __forceinline__ void grid_barrier(const DevModel & m, BarState & bs) {
    __syncthreads();
    if (threadIdx.x == 0) {
        bs.target += gridDim.x;
        __threadfence();
        red_release_add_u64(m.bar, 1ULL);
        if (!bs.dead && !spin_until(m.bar, bs.target, m.st)) bs.dead = true;
    }
    __syncthreads();
}

// barrier among the `n` CTAs that share kv group g (attention soft-max statistics exchange)
__device__ __forceinline__ void group_barrier(const DevModel & m, BarState & bs, int g, int n, bool participate) {
    bs.gtarget += (unsigned long long) n;
    if (!participate) return;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        red_release_add_u64(m.gbar + g, 1ULL);
        if (!bs.dead && !spin_until(m.gbar + g, bs.gtarget, m.st)) bs.dead = true;
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------- TMA weight stream
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long * bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long * bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long * bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "W_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra D_%=;\n\t"
        "bra W_%=;\n\t"
        "D_%=:\n\t}"
        :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared::cta, completion signalled on an mbarrier (bytes multiple of 16)
__device__ __forceinline__ void tma_load_1d(void * dst_smem, const void * src_gmem, uint32_t bytes, unsigned long long * bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct WStream {                 // per-thread view of the two weight buffers
    uint32_t parity[2];          // phase parity the next wait on buffer b must observe
    bool     pending[2];         // a fetch into buffer b has been issued and not yet waited for
};

// Warp 0 stages `n_rows` rows of `row_bytes` (contiguous in HBM) into buffer `b` with the padded
// pitch.  Must be called by all threads AFTER a __syncthreads that ends all reads of that buffer.
__device__ __forceinline__ void issue_fetch(Smem & s, WStream & ws, int b, const void * src, int n_rows, int row_bytes) {
    ws.pending[b] = true;
    if (threadIdx.x < 32) {
        unsigned char * dst = b == 0 ? s.buf0 : s.buf1;
        const int pitch = row_bytes + PAD_B;
        if (threadIdx.x == 0) mbar_expect_tx(&s.mbar[b], (uint32_t) (n_rows * row_bytes));
        __syncwarp();
        for (int r = threadIdx.x; r < n_rows; r += 32)
            tma_load_1d(dst + (size_t) r * pitch, (const unsigned char *) src + (size_t) r * row_bytes, (uint32_t) row_bytes, &s.mbar[b]);
    }
}
__device__ __forceinline__ void wait_fetch(Smem & s, WStream & ws, int b) {
    mbar_wait(&s.mbar[b], ws.parity[b]);
    ws.parity[b] ^= 1u;
    ws.pending[b] = false;
}

__device__ __forceinline__ void cta_rows(int n_rows, int & r0, int & r1) {
    r0 = (int) (((long long) blockIdx.x * n_rows) / gridDim.x);
    r1 = (int) (((long long) (blockIdx.x + 1) * n_rows) / gridDim.x);
}

// ------------------------------------------------------------------------------- ordered GEMV
// Rows of chain-major F16 weights (engine.cu k_chain_pack) in shared memory times the activation
// vector s.xf (F32 values; already F16-rounded for the F16 mat-vecs, raw F32 for ffn_down).
// Lane a of a warp IS accumulator lane a of ggml_vec_dot_f16 / ggml_vec_dot_f32 (k % 32 == a): it
// keeps x[32j + a] for all j in registers and walks its FMA chain over j for R rows at a time
// (R independent chains per lane hide the FMA latency); one 16-byte shared-memory load brings 8
// consecutive chain steps, a warp reads 512 contiguous bytes per load (conflict free).  The 32
// lane sums are folded in the oracle's order (reduce32_lanes).  Result of row r -> s.outv[r].
template <int K, int R>
__device__ __forceinline__ void gemv_lane(const unsigned char * __restrict__ wbuf, int n_rows, Smem & s) {
    constexpr int NJ = K / 32, NU = K / 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= n_rows) return;
    float xr[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) xr[j] = s.xf[32 * j + lane];
    for (int r0 = warp; r0 < n_rows; r0 += 16 * R) {
        float acc[R];
        const unsigned char * wrow[R];
#pragma unroll
        for (int i = 0; i < R; ++i) {
            acc[i] = 0.0f;
            const int r = r0 + 16 * i;
            wrow[i] = wbuf + (size_t) (r < n_rows ? r : r0) * (K * 2) + lane * 16;
        }
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            uint4 wv[R];
#pragma unroll
            for (int i = 0; i < R; ++i) wv[i] = *reinterpret_cast<const uint4 *>(wrow[i] + u * 512);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
#pragma unroll
                for (int i = 0; i < R; ++i) {
                    const unsigned w2 = e == 0 ? wv[i].x : e == 1 ? wv[i].y : e == 2 ? wv[i].z : wv[i].w;
                    const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w2));
                    acc[i] = __fmaf_rn(f.x, xr[8 * u + 2 * e], acc[i]);
                    acc[i] = __fmaf_rn(f.y, xr[8 * u + 2 * e + 1], acc[i]);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const float res = reduce32_lanes(acc[i]);
            const int r = r0 + 16 * i;
            if (lane == 0 && r < n_rows) s.outv[r] = res;
        }
    }
}

// ------------------------------------------------------------------------------- RMSNorm
// y = (x * rsqrt(mean(x^2)+eps)) * gamma; sum of squares in double (oracle D.5,
// ggml/src/ggml-cpu/ops.cpp:3666-3691).  Writes float(r16(y)) to s.xf (the F16 mat-vec input)
// and optionally the unrounded y (the `hidden_states` output).
__device__ __forceinline__ void rmsnorm_to_xf(Smem & s, const float * __restrict__ gamma, float eps, float * y_out) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const float a = s.xres[t], b = s.xres[t + 512];
    double sq = (double) __fmul_rn(a, a) + (double) __fmul_rn(b, b);
    sq = warp_sum_d(sq);
    if (lane == 0) s.dred[0][warp] = sq;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) tot += s.dred[0][i];
    const float mean  = (float) (tot / (double) H);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    const float ya = __fmul_rn(__fmul_rn(a, scale), __ldg(gamma + t));
    const float yb = __fmul_rn(__fmul_rn(b, scale), __ldg(gamma + t + 512));
    s.xf[t] = r16(ya); s.xf[t + 512] = r16(yb);
    if (y_out) { y_out[t] = ya; y_out[t + 512] = yb; }
    __syncthreads();
}

// per-head RMSNorm(128)*gamma then NEOX RoPE (FMA-contracted like the reference build), one warp;
// lane l owns dims {l, l+32, l+64, l+96}; rope row = cos[64] | sin[64] (oracle D.7)
__device__ __forceinline__ void head_norm_rope(const float * __restrict__ raw, const float * __restrict__ gamma,
                                               const float * __restrict__ rope_row, float eps, bool do_norm,
                                               float (&out)[4]) {
    const int lane = threadIdx.x & 31;
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = ld_act(raw + lane + 32 * i);
    if (do_norm) {
        double sq = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) sq += (double) __fmul_rn(v[i], v[i]);
        sq = warp_sum_d(sq);
        const float mean  = (float) (sq / (double) HD);
        const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = __fmul_rn(__fmul_rn(v[i], scale), __ldg(gamma + lane + 32 * i));
        const float c0 = __ldg(rope_row + lane),      s0 = __ldg(rope_row + 64 + lane);
        const float c1 = __ldg(rope_row + lane + 32), s1 = __ldg(rope_row + 64 + lane + 32);
        // the reference build (g++ -O3 -mfma) contracts the rotation: fmsub(x0,c,x1*s), fmadd(x0,s,x1*c)
        out[0] = __fmaf_rn(v[0], c0, -__fmul_rn(v[2], s0));
        out[2] = __fmaf_rn(v[0], s0,  __fmul_rn(v[2], c0));
        out[1] = __fmaf_rn(v[1], c1, -__fmul_rn(v[3], s1));
        out[3] = __fmaf_rn(v[1], s1,  __fmul_rn(v[3], c1));
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = v[i];
    }
}

// ------------------------------------------------------------------------------- attention (phase B)
// CTA role: kv group g = blockIdx % 8 (q heads 2g, 2g+1 share kv head g — GQA by broadcast,
// ggml-cpu.c:1160-1199), split index c = blockIdx / 8 of C.  Stage 1 splits the positions for the
// scores, stage 2 splits the head dimension for P.V, so every dot product keeps the oracle's
// accumulation order (D.4): s_t = dot_f16(K_t, r16(q)) * scale ; P = softmax ; o_d =
// dot_f16(V^T_d, r16(P)) over positions (32 accumulator lanes t%32 + double-precision leftovers).
__device__ __forceinline__ int attn_splits(int n_kv, bool f32attn) {
    if (f32attn || n_kv <= 48) return 1;
    if (n_kv <= 192) return 4;
    if (n_kv <= 1024) return 8;
    return 16;
}

__device__ __forceinline__ float ld_half_cg(const __half * p) {
    const unsigned short u = __ldcg(reinterpret_cast<const unsigned short *>(p));
    return __half2float(__ushort_as_half(u));
}

__device__ void attention_phase(const DevModel & m, const StackW & W, int l, int pos, bool f32attn, int width,
                                Smem & s, BarState & bs) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int g = blockIdx.x % NKV, c = blockIdx.x / NKV;
    const int n_kv = pos + 1;
    // row width of the reference graph: a prefill scores all n_past + n_tokens cache rows and masks
    // the future ones with -inf (src/tts_transformer.cpp:1205-1227); masked entries add exact zeros,
    // but the width decides which positions take the 8-wide SIMD exp vs the libm tail and which
    // take the 32-lane part of the P.V dot vs its double-precision leftovers.
    const int wd = width > n_kv ? width : n_kv;
    const int C = attn_splits(n_kv, f32attn);
    const bool active = c < C;
    const float kq_scale = 1.0f / sqrtf((float) HD);
    const size_t kv_head_base = ((size_t) (l * NKV + g) * W.n_ctx) * HD;
    float * ao = m.part_o;                                   // [16 heads][128] attention output

    if (!active) {
        if (C > 1) group_barrier(m, bs, g, C, false);
        return;
    }
    // ---- prologue: q (2 heads), k, v of the new token; the CTA with c == 0 appends K/V to the cache
    if (warp < 4) {
        const float * rope_row = m.rope + (size_t) pos * HD;
        float o[4];
        if (warp < 2) {
            head_norm_rope(m.qkv + (2 * g + warp) * HD, W.q_norm + l * HD, rope_row, m.eps, true, o);
#pragma unroll
            for (int i = 0; i < 4; ++i) { s.qf[warp][lane + 32 * i] = o[i]; s.qr[warp][lane + 32 * i] = r16(o[i]); }
        } else if (warp == 2) {
            head_norm_rope(m.qkv + NH * HD + g * HD, W.k_norm + l * HD, rope_row, m.eps, true, o);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int d = lane + 32 * i;
                const __half hv = __float2half_rn(o[i]);
                s.knew_f[d] = o[i]; s.knew_h[d] = hv;
                if (c == 0) {
                    W.kcache[kv_head_base + (size_t) pos * HD + d] = hv;
                    if (W.kf32 && pos < 2) W.kf32[((size_t) (l * 2 + pos) * NKV + g) * HD + d] = o[i];
                }
            }
        } else {
            head_norm_rope(m.qkv + (NH + NKV) * HD + g * HD, nullptr, rope_row, m.eps, false, o);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int d = lane + 32 * i;
                const __half hv = __float2half_rn(o[i]);
                s.vnew_f[d] = o[i]; s.vnew_h[d] = hv;
                if (c == 0) {
                    W.vcache[kv_head_base + (size_t) pos * HD + d] = hv;
                    if (W.vf32 && pos < 2) W.vf32[((size_t) (l * 2 + pos) * NKV + g) * HD + d] = o[i];
                }
            }
        }
    }
    __syncthreads();

    if (f32attn) {
        // code-predictor prefill numerics: the reference graph multiplies the fresh F32 Kcur/Vcur/Q
        // and does not round P (src/tts_transformer.cpp:1624-1635): ggml_vec_dot_f32 order.
        if (warp < 2) {
            float sc_t[2];
            float mx = -CUDART_INF_F;
            const int n = pos + 1 < 2 ? pos + 1 : 2;
            for (int tt = 0; tt < n; ++tt) {
                const float * krow = (tt == pos) ? s.knew_f : (W.kf32 + ((size_t) (l * 2 + tt) * NKV + g) * HD);
                float acc = 0.0f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float kv = (tt == pos) ? krow[lane + 32 * i] : ld_act(krow + lane + 32 * i);
                    acc = __fmaf_rn(kv, s.qf[warp][lane + 32 * i], acc);
                }
                sc_t[tt] = __fmul_rn(reduce32_lanes(acc), kq_scale);
                mx = fmaxf(mx, sc_t[tt]);
            }
            double sum = 0.0;
            for (int tt = 0; tt < n; ++tt) { sc_t[tt] = tail_expf(__fsub_rn(sc_t[tt], mx)); sum += (double) sc_t[tt]; }
            const float inv = (float) (1.0 / sum);
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            for (int tt = 0; tt < n; ++tt) {
                const float * vrow = (tt == pos) ? s.vnew_f : (W.vf32 + ((size_t) (l * 2 + tt) * NKV + g) * HD);
                const float p = __fmul_rn(sc_t[tt], inv);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float vv = (tt == pos) ? vrow[lane + 32 * i] : ld_act(vrow + lane + 32 * i);
                    o[i] = __fmaf_rn(vv, p, o[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) ao[(2 * g + warp) * HD + lane + 32 * i] = o[i];
        }
        return;
    }

    float * sb0 = m.sbuf + (size_t) (2 * g) * ATT_MAX_KV;
    float * sb1 = sb0 + ATT_MAX_KV;
    // ---- stage 1: scores of this CTA's position chunk
    {
        const int Lc = (n_kv + C - 1) / C;
        const int t0 = min(c * Lc, n_kv), t1 = min(n_kv, t0 + Lc);
        float q0[4], q1[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { q0[i] = s.qr[0][lane + 32 * i]; q1[i] = s.qr[1][lane + 32 * i]; }
        float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F;
        for (int tt = t0 + warp; tt < t1; tt += 16) {
            float kv[4];
            if (tt == pos) {
#pragma unroll
                for (int i = 0; i < 4; ++i) kv[i] = __half2float(s.knew_h[lane + 32 * i]);
            } else {
                const __half * krow = W.kcache + kv_head_base + (size_t) tt * HD;
#pragma unroll
                for (int i = 0; i < 4; ++i) kv[i] = ld_half_cg(krow + lane + 32 * i);
            }
            float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
            for (int i = 0; i < 4; ++i) { a0 = __fmaf_rn(kv[i], q0[i], a0); a1 = __fmaf_rn(kv[i], q1[i], a1); }
            const float s0 = __fmul_rn(reduce32_lanes(a0), kq_scale);
            const float s1 = __fmul_rn(reduce32_lanes(a1), kq_scale);
            mx0 = fmaxf(mx0, s0); mx1 = fmaxf(mx1, s1);
            if (lane == 0) { sb0[tt] = s0; sb1[tt] = s1; }
        }
        if (lane == 0) { s.fred[warp] = mx0; s.fred[16 + warp] = mx1; }
        __threadfence_block();
        __syncthreads();
        if (t == 0) {
            float a = s.fred[0], b = s.fred[16];
            for (int i = 1; i < 16; ++i) { a = fmaxf(a, s.fred[i]); b = fmaxf(b, s.fred[16 + i]); }
            float * pm = m.part_ms + (size_t) (g * 16 + c) * 2;
            pm[0] = a; pm[1] = b;
        }
    }
    if (C > 1) group_barrier(m, bs, g, C, true);
    else { __threadfence_block(); __syncthreads(); }

    // ---- soft-max over all n_kv positions (every CTA of the group, redundantly)
    float M0 = -CUDART_INF_F, M1 = -CUDART_INF_F;
    for (int i = 0; i < C; ++i) {
        const float * pm = m.part_ms + (size_t) (g * 16 + i) * 2;
        M0 = fmaxf(M0, ld_act(pm)); M1 = fmaxf(M1, ld_act(pm + 1));
    }
    const int n8 = wd >> 3;                                  // chunks that take the SIMD exp
    const int n8v = min(n8, (n_kv + 7) >> 3);                // ... of which these hold unmasked positions
    double ds0 = 0.0, ds1 = 0.0;
    for (int ch = t; ch < n8v; ch += LM_THREADS) {          // ggml_vec_soft_max_f32, 8 lanes at a time (vec.cpp:558-568)
        float e0[8], e1[8];
        const float4 * p0 = reinterpret_cast<const float4 *>(sb0 + ch * 8);
        const float4 * p1 = reinterpret_cast<const float4 *>(sb1 + ch * 8);
        const float4 x0 = __ldcg(p0), x1 = __ldcg(p0 + 1), y0 = __ldcg(p1), y1 = __ldcg(p1 + 1);
        const float xs[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        const float ys[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const bool live = ch * 8 + j < n_kv;             // masked: exp(-inf) = +0
            e0[j] = live ? v_expf(__fsub_rn(xs[j], M0)) : 0.0f;
            e1[j] = live ? v_expf(__fsub_rn(ys[j], M1)) : 0.0f;
        }
        ds0 += (double) __fadd_rn(__fadd_rn(__fadd_rn(e0[4], e0[0]), __fadd_rn(e0[6], e0[2])),
                                  __fadd_rn(__fadd_rn(e0[5], e0[1]), __fadd_rn(e0[7], e0[3])));
        ds1 += (double) __fadd_rn(__fadd_rn(__fadd_rn(e1[4], e1[0]), __fadd_rn(e1[6], e1[2])),
                                  __fadd_rn(__fadd_rn(e1[5], e1[1]), __fadd_rn(e1[7], e1[3])));
    }
    if (n8 * 8 + t < n_kv && t < 8) {                       // scalar tail: libm expf
        const int tt = n8 * 8 + t;
        ds0 += (double) tail_expf(__fsub_rn(ld_act(sb0 + tt), M0));
        ds1 += (double) tail_expf(__fsub_rn(ld_act(sb1 + tt), M1));
    }
    ds0 = warp_sum_d(ds0); ds1 = warp_sum_d(ds1);
    if (lane == 0) { s.dred[0][warp] = ds0; s.dred[1][warp] = ds1; }
    __syncthreads();
    double S0 = 0.0, S1 = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) { S0 += s.dred[0][i]; S1 += s.dred[1][i]; }
    const float inv0 = (float) (1.0 / S0), inv1 = (float) (1.0 / S1);
    for (int tt = t; tt < n_kv; tt += LM_THREADS) {
        const float x = __fsub_rn(ld_act(sb0 + tt), M0), y = __fsub_rn(ld_act(sb1 + tt), M1);
        const bool simd = tt < n8 * 8;
        s.ph[0][tt] = __float2half_rn(__fmul_rn(simd ? v_expf(x) : tail_expf(x), inv0));
        s.ph[1][tt] = __float2half_rn(__fmul_rn(simd ? v_expf(y) : tail_expf(y), inv1));
    }
    __syncthreads();

    // ---- stage 2: o[d] for this CTA's nd = 128/C dims, both heads
    const int nd = HD / C, d0 = c * nd;
    const int TB = VT_HALVES / nd;                          // positions per V^T tile (multiple of 32)
    const int np = min(wd & ~31, n_kv);                     // positions in the 32-lane part of the dot
    const int ndg = nd >> 3;                                // 16-byte groups per position
    float acc0[8], acc1[8];                                 // up to 8 dims per warp (nd = 128)
#pragma unroll
    for (int j = 0; j < 8; ++j) { acc0[j] = 0.0f; acc1[j] = 0.0f; }
    const __half * vbase = W.vcache + kv_head_base;
    for (int tb0 = 0; tb0 < np; tb0 += TB) {
        const int tbn = min(TB, np - tb0);
        for (int idx = t; idx < tbn * ndg; idx += LM_THREADS) {
            const int tt = idx / ndg, dg = idx % ndg;
            const int tg = tb0 + tt;
            uint4 v;
            if (tg == pos) v = *reinterpret_cast<const uint4 *>(s.vnew_h + d0 + dg * 8);
            else v = __ldcg(reinterpret_cast<const uint4 *>(vbase + (size_t) tg * HD + d0 + dg * 8));
            const __half * vh = reinterpret_cast<const __half *>(&v);
#pragma unroll
            for (int e = 0; e < 8; ++e) s.vt[(dg * 8 + e) * TB + tt] = vh[e];
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int dl = warp + 16 * j;
            if (dl < nd) {
                for (int tt = lane; tt < tbn; tt += 32) {
                    const float vv = __half2float(s.vt[dl * TB + tt]);
                    acc0[j] = __fmaf_rn(vv, __half2float(s.ph[0][tb0 + tt]), acc0[j]);
                    acc1[j] = __fmaf_rn(vv, __half2float(s.ph[1][tb0 + tt]), acc1[j]);
                }
            }
        }
        __syncthreads();
    }
    // leftovers t in [np, n_kv): staged into the tile, then one thread per (head, dim) adds them
    // sequentially in double like ggml_vec_dot_f16's scalar tail (vec.cpp:381-384)
    const int nleft = n_kv - np;
    for (int idx = t; idx < nleft * ndg; idx += LM_THREADS) {
        const int tt = idx / ndg, dg = idx % ndg;
        const int tg = np + tt;
        uint4 v;
        if (tg == pos) v = *reinterpret_cast<const uint4 *>(s.vnew_h + d0 + dg * 8);
        else v = __ldcg(reinterpret_cast<const uint4 *>(vbase + (size_t) tg * HD + d0 + dg * 8));
        const __half * vh = reinterpret_cast<const __half *>(&v);
#pragma unroll
        for (int e = 0; e < 8; ++e) s.vt[(dg * 8 + e) * TB + tt] = vh[e];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int dl = warp + 16 * j;
        const float r0 = reduce32_lanes(acc0[j]), r1 = reduce32_lanes(acc1[j]);
        if (dl < nd && lane == 0) { s.xf[dl] = r0; s.xf[HD + dl] = r1; }      // xf is free during phase B
    }
    __syncthreads();
    if (t < 2 * nd) {
        const int h = t / nd, dl = t % nd;
        double sumf = (double) s.xf[h * HD + dl];
        for (int i = 0; i < nleft; ++i)
            sumf += (double) __fmul_rn(__half2float(s.vt[dl * TB + i]), __half2float(s.ph[h][np + i]));
        ao[(2 * g + h) * HD + d0 + dl] = (float) sumf;
    }
}

// --------------------------------------------------------------------------- token selection
// Reference: host argmax (src/tts_transformer.cpp:2182-2192: first maximum wins) and the
// sample_or_argmax lambda / cb0 post-processing (:2349-2387, :2664-2722).  Every CTA evaluates
// the selection redundantly on identical global data, so no extra barrier is needed.
__device__ __forceinline__ float proc_logit(const DevModel & m, float raw, int i, bool process,
                                            int suppress_start, int eos_id, float rep_penalty) {
    if (!process) return raw;
    if (i >= suppress_start && i != eos_id) return -CUDART_INF_F;
    if (rep_penalty != 1.0f && __ldcg(m.seen + i)) return raw > 0.0f ? raw / rep_penalty : raw * rep_penalty;
    return raw;
}

__device__ int select_greedy(const DevModel & m, Smem & s) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    float v = -CUDART_INF_F; int idx = 0x7fffffff;
    if (t < (int) gridDim.x) { v = ld_act(m.cand_val + t); idx = ld_act_i(m.cand_idx + t); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int   oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    if (lane == 0) { s.fred[warp] = v; s.ired[warp] = idx; }
    __syncthreads();
    v = s.fred[0]; idx = s.ired[0];
#pragma unroll
    for (int i = 1; i < 16; ++i) {
        const float ov = s.fred[i]; const int oi = s.ired[i];
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    __syncthreads();
    return idx;
}

__device__ __forceinline__ unsigned int f2key(float f) {          // order-preserving float -> uint
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ double u01(unsigned long long seed, unsigned long long n) {   // splitmix64
    unsigned long long z = seed + 0x9E3779B97F4A7C15ULL * (n + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return (double) (z >> 11) * (1.0 / 9007199254740992.0);
}

// temperature -> top-k threshold (ties kept, `<` test) -> softmax (double sum) -> inverse CDF draw
__device__ int select_sample(const DevModel & m, Smem & s, int vocab, bool process, int suppress_start,
                             int eos_id, float rep_penalty, float temperature, int top_k,
                             unsigned long long seed, unsigned long long draw) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    constexpr int PER = 6;                       // 512 * 6 = 3072 >= vocab
    float l[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int idx = t * PER + i;
        l[i] = -CUDART_INF_F;
        if (idx < vocab) l[i] = proc_logit(m, ld_act(m.logits + idx), idx, process, suppress_start, eos_id, rep_penalty) / temperature;
    }
    int * hist = reinterpret_cast<int *>(s.vt);              // 256 bins + 2 control words
    if (top_k > 0 && top_k < vocab) {
        unsigned int prefix = 0, mask = 0;
        int k_rem = top_k;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            if (t < 256) hist[t] = 0;
            __syncthreads();
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                if (t * PER + i < vocab) {
                    const unsigned int key = f2key(l[i]);
                    if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1);
                }
            }
            __syncthreads();
            if (t == 0) {
                int acc = 0, d = 255;
                for (; d > 0; --d) { if (acc + hist[d] >= k_rem) break; acc += hist[d]; }
                hist[256] = d; hist[257] = k_rem - acc;
            }
            __syncthreads();
            prefix |= ((unsigned int) hist[256]) << shift;
            mask   |= 0xffu << shift;
            k_rem   = hist[257];
            __syncthreads();
        }
        const float thr = key2f(prefix);
#pragma unroll
        for (int i = 0; i < PER; ++i) if (l[i] < thr) l[i] = -CUDART_INF_F;
    }
    float mx = -CUDART_INF_F;
#pragma unroll
    for (int i = 0; i < PER; ++i) mx = fmaxf(mx, l[i]);
    mx = warp_max(mx);
    if (lane == 0) s.fred[warp] = mx;
    __syncthreads();
    mx = s.fred[0];
#pragma unroll
    for (int i = 1; i < 16; ++i) mx = fmaxf(mx, s.fred[i]);
    float e[PER];
    double loc = 0.0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { e[i] = (t * PER + i < vocab) ? expf(l[i] - mx) : 0.0f; loc += (double) e[i]; }
    // inclusive scan of the per-thread sums (thread order == index order)
    double inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    __syncthreads();
    if (lane == 31) s.dred[0][warp] = inc;
    __syncthreads();
    double woff = 0.0, total = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) { if (i < warp) woff += s.dred[0][i]; total += s.dred[0][i]; }
    inc += woff;
    const double target = u01(seed, draw) * total;
    if (t == 0) s.tok = -1;
    __syncthreads();
    const double exc = inc - loc;
    if (target >= exc && target < inc) {
        double a = exc; int pick = -1;
#pragma unroll
        for (int i = 0; i < PER; ++i) { a += (double) e[i]; if (pick < 0 && e[i] > 0.0f && target < a) pick = t * PER + i; }
        if (pick < 0) { for (int i = PER - 1; i >= 0; --i) if (e[i] > 0.0f) { pick = t * PER + i; break; } }
        if (pick >= 0) s.tok = pick;
    }
    __syncthreads();
    if (s.tok < 0) {            // rounding at the very end of the CDF: fall back to the arg-max element
        if (t == 0) s.tok = 0;
        __syncthreads();
#pragma unroll
        for (int i = 0; i < PER; ++i) if (t * PER + i < vocab && l[i] == mx) s.tok = t * PER + i;
        __syncthreads();
    }
    const int tok = s.tok;
    __syncthreads();
    return tok;
}


// ------------------------------------------------------------------------------- weight slices
struct Slices {            // this CTA's row ranges (fixed for the whole launch)
    int a0, a1;            // qkv rows            (of 4096, K=1024)
    int o0, o1;            // attn_output rows    (of 1024, K=2048)
    int p0, p1;            // gate/up pairs       (of 3072, K=1024, 2 rows per pair)
    int d0, d1;            // ffn_down rows       (of 1024, K=3072)
};

__device__ __forceinline__ const __half * sl_a(const StackW & W, int l, const Slices & z) {
    return reinterpret_cast<const __half *>(W.wqkv) + ((size_t) l * QKV_N + z.a0) * H;
}
__device__ __forceinline__ const __half * sl_o(const StackW & W, int l, const Slices & z) {
    return reinterpret_cast<const __half *>(W.wo) + ((size_t) l * H + z.o0) * (NH * HD);
}
__device__ __forceinline__ const __half * sl_gu(const StackW & W, int l, const Slices & z) {
    return reinterpret_cast<const __half *>(W.wgu) + ((size_t) l * 2 * FFN + 2 * z.p0) * H;
}
__device__ __forceinline__ const __half * sl_d(const StackW & W, int l, const Slices & z) {
    return reinterpret_cast<const __half *>(W.wdown) + ((size_t) l * H + z.d0) * FFN;
}

__device__ __forceinline__ void l2_prefetch_range(const void * p, size_t bytes) {
    const char * c = reinterpret_cast<const char *>(p);
    while (bytes > 0) {
        const uint32_t n = bytes > 32768 ? 32768u : (uint32_t) bytes;
        l2_prefetch_bulk(c, n);
        c += n; bytes -= n;
    }
}
// one thread: pull this CTA's slices of layer l of stack W towards L2 (TMA-unit bulk prefetch)
__device__ __forceinline__ void l2_prefetch_layer(const StackW & W, int l, const Slices & z) {
    l2_prefetch_range(sl_a(W, l, z),  (size_t) (z.a1 - z.a0) * H * 2);
    l2_prefetch_range(sl_o(W, l, z),  (size_t) (z.o1 - z.o0) * NH * HD * 2);
    l2_prefetch_range(sl_gu(W, l, z), (size_t) (z.p1 - z.p0) * 2 * H * 2);
    l2_prefetch_range(sl_d(W, l, z),  (size_t) (z.d1 - z.d0) * FFN * 2);
}

// ------------------------------------------------------------------------------- one pass
// One token through `n_layers` layers of stack W (+ optional head).  s.xres holds x on entry;
// the TMA fetches of A(0) -> buf0 and C(0) -> buf1 have been issued by the caller.  On exit the
// fetches of the next pass' A(0)/C(0) (stack NW, if any) are in flight.
__device__ void run_stack(const DevModel & m, const StackW & W, const PassDesc & p, int pos,
                          const HeadW * head, bool process, const GenState & gs,
                          const Slices & z, const StackW * NW,
                          Smem & s, BarState & bs, WStream & ws) {
    const int t = threadIdx.x;
    const int nA = z.a1 - z.a0, nO = z.o1 - z.o0, nP = z.p1 - z.p0, nD = z.d1 - z.d0;
    int h0 = 0, h1 = 0;
    if (head) cta_rows(head->vocab, h0, h1);
    const int nH = h1 - h0;

    for (int l = 0; l < p.n_layers; ++l) {
        const bool last = l + 1 == p.n_layers;
        if (t == 0) {
            if (!last) l2_prefetch_layer(W, l + 1, z);
            else if (head) l2_prefetch_range(head->w + (size_t) h0 * H, (size_t) nH * H * 2);
        }
        // ---- phase A: RMSNorm + fused QKV mat-vec
        PROF(1);
        rmsnorm_to_xf(s, W.attn_norm + (size_t) l * H, m.eps, nullptr);
        PROF(2);
        wait_fetch(s, ws, 0);
        PROF(3);
        gemv_lane<1024, 2>(s.buf0, nA, s);
        __syncthreads();
        if (t < nA) m.qkv[z.a0 + t] = s.outv[t];
        issue_fetch(s, ws, 0, sl_gu(W, l, z), 2 * nP, H * 2);
        PROF(4);
        grid_barrier(m, bs);
        PROF(5);

        // ---- phase B: per-head norm + RoPE + KV append + attention
        attention_phase(m, W, l, pos, p.f32attn != 0, p.width, s, bs);
        PROF(6);
        grid_barrier(m, bs);
        PROF(7);

        // ---- phase C: attn_output mat-vec (input rounded to F16) + residual
#pragma unroll
        for (int i = 0; i < 4; ++i) s.xf[t + LM_THREADS * i] = r16(ld_act(m.part_o + t + LM_THREADS * i));
        __syncthreads();
        PROF(8);
        wait_fetch(s, ws, 1);
        PROF(9);
        gemv_lane<2048, 1>(s.buf1, nO, s);
        __syncthreads();
        if (t < nO) m.xb[z.o0 + t] = __fadd_rn(s.outv[t], s.xres[z.o0 + t]);
        issue_fetch(s, ws, 1, sl_d(W, l, z), nD, FFN * 2);
        PROF(10);
        grid_barrier(m, bs);
        PROF(11);

        // ---- phase D: RMSNorm + gate/up mat-vec + SwiGLU
        s.xres[t] = ld_act(m.xb + t); s.xres[t + 512] = ld_act(m.xb + t + 512);
        __syncthreads();
        rmsnorm_to_xf(s, W.ffn_norm + (size_t) l * H, m.eps, nullptr);
        PROF(12);
        wait_fetch(s, ws, 0);
        PROF(13);
        gemv_lane<1024, 3>(s.buf0, 2 * nP, s);
        __syncthreads();
        if (t < nP) {
            const float gt = s.outv[2 * t], up = s.outv[2 * t + 1];
            const float sl = __fdiv_rn(gt, __fadd_rn(1.0f, v_expf(__fsub_rn(0.0f, gt))));   // ggml_v_silu
            m.mbuf[z.p0 + t] = __fmul_rn(sl, up);
        }
        if (!last)      issue_fetch(s, ws, 0, sl_a(W, l + 1, z), nA, H * 2);
        else if (head)  issue_fetch(s, ws, 0, head->w + (size_t) h0 * H, nH, H * 2);
        else if (NW)    issue_fetch(s, ws, 0, sl_a(*NW, 0, z), nA, H * 2);
        PROF(14);
        grid_barrier(m, bs);
        PROF(15);

        // ---- phase E: ffn_down on F32 activations (weights exact in F32) + residual
#pragma unroll
        for (int i = 0; i < FFN / LM_THREADS; ++i) s.xf[t + LM_THREADS * i] = ld_act(m.mbuf + t + LM_THREADS * i);
        __syncthreads();
        PROF(16);
        wait_fetch(s, ws, 1);
        PROF(17);
        gemv_lane<3072, 1>(s.buf1, nD, s);
        __syncthreads();
        if (t < nD) m.xa[z.d0 + t] = __fadd_rn(s.outv[t], s.xres[z.d0 + t]);
        if (!last)      issue_fetch(s, ws, 1, sl_o(W, l + 1, z), nO, NH * HD * 2);
        else if (NW)    issue_fetch(s, ws, 1, sl_o(*NW, 0, z), nO, NH * HD * 2);
        PROF(18);
        grid_barrier(m, bs);
        PROF(19);
        s.xres[t] = ld_act(m.xa + t); s.xres[t + 512] = ld_act(m.xa + t + 512);
        __syncthreads();
    }
    PROF(20);

    if (!head) return;
    // ---- head: final RMSNorm (= hidden_states) + vocabulary mat-vec + per-CTA arg-max candidate
    rmsnorm_to_xf(s, W.out_norm, m.eps, blockIdx.x == 0 ? m.hidden : nullptr);
    if (p.hidden_dst && blockIdx.x == 0) { p.hidden_dst[t] = m.hidden[t]; p.hidden_dst[t + 512] = m.hidden[t + 512]; }
    wait_fetch(s, ws, 0);
    gemv_lane<1024, 2>(s.buf0, nH, s);
    __syncthreads();
    float pv = -CUDART_INF_F;
    if (t < nH) {
        const float raw = s.outv[t];
        m.logits[h0 + t] = raw;
        if (p.logits_dst) p.logits_dst[h0 + t] = raw;
        pv = proc_logit(m, raw, h0 + t, process, gs.suppress_start, gs.eos_id, gs.rep_penalty);
    }
    __syncthreads();
    if (t < nH) s.outv[t] = pv;
    __syncthreads();
    if (t == 0) {
        float bv = -CUDART_INF_F; int bi = 0x7fffffff;
        for (int i = 0; i < nH; ++i) if (s.outv[i] > bv) { bv = s.outv[i]; bi = h0 + i; }
        if (nH > 0 && bi == 0x7fffffff) bi = h0;       // all -inf: keep first-index semantics
        m.cand_val[blockIdx.x] = bv; m.cand_idx[blockIdx.x] = bi;
    }
    if (NW) issue_fetch(s, ws, 0, sl_a(*NW, 0, z), nA, H * 2);
    PROF(21);
    grid_barrier(m, bs);
    PROF(22);
}

// --------------------------------------------------------------------------- the kernel
// frame_mode = 1: `passes` is the 17-pass frame program; the kernel owns the generate-loop
// bookkeeping of src/tts_transformer.cpp:2664-2804 (EOS test, code emission, repetition set,
// n_past) in GenState so the identical launch can be replayed once per frame.
__global__ void __launch_bounds__(LM_THREADS, 1)
lm_kernel(const __grid_constant__ DevModel m, const PassDesc * __restrict__ passes, int n_passes, int frame_mode) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem & s = *reinterpret_cast<Smem *>(smem_raw);
    __shared__ GenState gs;
    __shared__ PassDesc pd;
    const int t = threadIdx.x;

    if (t == 0) gs = *m.st;
    __syncthreads();
    if (frame_mode && gs.done) return;

    BarState bs;
    bs.target  = gs.bar_base;
    bs.gtarget = gs.gbar_base[blockIdx.x % NKV];
    bs.dead    = false;

    Slices z;
    cta_rows(QKV_N, z.a0, z.a1);
    cta_rows(H,     z.o0, z.o1);
    cta_rows(FFN,   z.p0, z.p1);
    cta_rows(H,     z.d0, z.d1);

    WStream ws;
    ws.parity[0] = ws.parity[1] = 0u;
    ws.pending[0] = ws.pending[1] = false;
    if (t == 0) {
        mbar_init(&s.mbar[0], 1);
        mbar_init(&s.mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    int frame = gs.frame, n_past = gs.n_past, done = gs.done, pending = gs.pending;
    int pending_vocab = gs.codec_vocab, pending_process = 1;
    unsigned long long draw = gs.draw;

    for (int ip = 0; ip < n_passes; ++ip) {
        PROF(30);
        if (t == 0) pd = passes[ip];
        __syncthreads();
        const PassDesc p = pd;
        const StackW & W = p.stack == 0 ? m.talker : m.cp;

        // start pulling layer 0 of this stack towards L2 / shared memory before the (serial) selection
        if (p.xmode != XM_SELECT_ONLY) {
            if (t == 0) l2_prefetch_layer(W, 0, z);
            if (!ws.pending[0]) issue_fetch(s, ws, 0, sl_a(W, 0, z), z.a1 - z.a0, H * 2);
            if (!ws.pending[1]) issue_fetch(s, ws, 1, sl_o(W, 0, z), z.o1 - z.o0, NH * HD * 2);
        }

        // ---- select the token produced by the previous head, if one is waiting
        const bool needs_select = p.xmode == XM_HIDDEN || p.xmode == XM_CP_EMBD ||
                                  p.xmode == XM_TALKER_NEXT || p.xmode == XM_SELECT_ONLY;
        if (needs_select && pending >= 0) {
            int tok;
            if (gs.temperature <= 0.0f) tok = select_greedy(m, s);
            else tok = select_sample(m, s, pending_vocab, pending == 0 && pending_process, gs.suppress_start, gs.eos_id,
                                     gs.rep_penalty, gs.temperature, gs.top_k, gs.seed, draw++);
            if (t == 0) gs.cur[pending] = tok;
            __syncthreads();
            pending = -1;
        }
        if (p.xmode == XM_SELECT_ONLY) continue;

        if (p.xmode == XM_HIDDEN && frame_mode) {
            // src/tts_transformer.cpp:2724-2729: EOS stops before a frame is emitted
            if (gs.cur[0] == gs.eos_id) { done = 1; break; }
            if (blockIdx.x == 0 && t == 0) m.seen[gs.cur[0]] = 1;
        }
        if (p.xmode == XM_TALKER_NEXT && frame_mode) {
            if (blockIdx.x == 0 && t < NCB) m.codes_out[(size_t) frame * NCB + t] = gs.cur[t];
            ++frame;
            if (frame >= gs.max_len) { done = 1; break; }     // :2755-2757 no talker step after the last frame
        }

        // ---- form x in shared memory
        {
            float xa = 0.f, xb = 0.f;
            switch (p.xmode) {
                case XM_BUF:    xa = ld_act(p.xsrc + t); xb = ld_act(p.xsrc + t + 512); break;
                case XM_HIDDEN: xa = ld_act(m.hidden + t); xb = ld_act(m.hidden + t + 512); break;
                case XM_CB0_EMBD: {
                    const __half * r = m.codec_embd + (size_t) gs.cur[0] * H;
                    xa = __half2float(r[t]); xb = __half2float(r[t + 512]);
                } break;
                case XM_CP_EMBD: {
                    const __half * r = m.cp_embd[p.arg - 1] + (size_t) gs.cur[p.arg] * H;
                    xa = __half2float(r[t]); xb = __half2float(r[t + 512]);
                } break;
                case XM_TALKER_NEXT: {
                    // :2759-2790  codec_embd[cb0] + sum_i code_pred_embd[i-1][c_i] + trailing/pad, F32 in this order
                    const __half * r = m.codec_embd + (size_t) gs.cur[0] * H;
                    xa = __half2float(r[t]); xb = __half2float(r[t + 512]);
#pragma unroll
                    for (int cb = 1; cb < NCB; ++cb) {
                        const __half * q = m.cp_embd[cb - 1] + (size_t) gs.cur[cb] * H;
                        xa = __fadd_rn(xa, __half2float(q[t])); xb = __fadd_rn(xb, __half2float(q[t + 512]));
                    }
                    const int fr = frame - 1;      // the frame just emitted
                    const float * tr = (fr < gs.trailing_len) ? (m.trailing + (size_t) fr * H) : m.pad_embed;
                    xa = __fadd_rn(xa, ld_act(tr + t)); xb = __fadd_rn(xb, ld_act(tr + t + 512));
                } break;
                default: break;
            }
            s.xres[t] = xa; s.xres[t + 512] = xb;
        }
        __syncthreads();

        PROF(31);
        const int pos = p.pos >= 0 ? p.pos : n_past;
        const HeadW * head = p.head < 0 ? nullptr : (p.head == 0 ? &m.codec_head : &m.cp_head[p.head - 1]);
        // the next pass' first weight slices are fetched across the pass boundary
        const StackW * NW = nullptr;
        if (ip + 1 < n_passes) {
            const int nx = passes[ip + 1].xmode, ns = passes[ip + 1].stack;
            if (nx != XM_SELECT_ONLY) NW = ns == 0 ? &m.talker : &m.cp;
        }
        run_stack(m, W, p, pos, head, p.process != 0, gs, z, NW, s, bs, ws);
        if (head) { pending = p.head; pending_vocab = head->vocab; pending_process = p.process; }
        if (p.stack == 0 && p.pos < 0) ++n_past;
    }

    // drain TMA copies still in flight (early exit on EOS / max_len) before the CTA retires
    if (ws.pending[0]) wait_fetch(s, ws, 0);
    if (ws.pending[1]) wait_fetch(s, ws, 1);

    // ---- write the state back (CTA 0); barrier epochs are identical in every CTA
    if (blockIdx.x == 0 && t == 0) {
        GenState * st = m.st;
        st->frame = frame; st->n_past = n_past; st->done = done; st->pending = pending;
        st->draw = draw;
        for (int i = 0; i < NCB; ++i) st->cur[i] = gs.cur[i];
        st->bar_base = bs.target;
    }
    if (blockIdx.x < NKV && t == 0) m.st->gbar_base[blockIdx.x] = bs.gtarget;
}

// --------------------------------------------------------------------------- host launch
int lm_max_grid(q3t_ctx * ctx) {
    static bool attr_set = false;
    const size_t smem = sizeof(Smem);
    if (!attr_set) {
        if (cudaFuncSetAttribute(lm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem) != cudaSuccess) return -1;
        attr_set = true;
    }
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lm_kernel, LM_THREADS, smem) != cudaSuccess) return -1;
    if (per_sm < 1) return 0;
    return ctx->sm_count;          // one persistent CTA per SM
}

int lm_launch(q3t_ctx * ctx, const PassDesc * passes_dev, int n_passes, int frame_mode) {
    if (ctx->grid <= 0) {
        const int g = lm_max_grid(ctx);
        if (g < LM_MIN_GRID) return set_err(ctx, Q3T_ERR_CUDA, "device cannot co-schedule the persistent decode kernel (needs >= 140 SMs)");
        ctx->grid = g > 512 ? 512 : g;
    }
    DevModel m = ctx->hm;
    void * args[] = { (void *) &m, (void *) &passes_dev, (void *) &n_passes, (void *) &frame_mode };
    cudaError_t e = cudaLaunchCooperativeKernel((const void *) lm_kernel, dim3(ctx->grid), dim3(LM_THREADS),
                                                args, sizeof(Smem), ctx->stream);
    ctx->timing.kernel_launches++;
    return cuda_check(ctx, e, "lm_kernel launch");
}

} // namespace q3t
