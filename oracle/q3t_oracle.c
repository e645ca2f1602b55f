/*
 * TEST INFRASTRUCTURE — CPU restatement (plain C, scalar) of the reference's arithmetic for the
 * codes-to-audio hot path.  Not part of the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg may load this.  Parity is PINNED: tests/test_oracle_cpu.py checks
 * it against the unmodified reference (oracle/_ref) and against the golden vectors under
 * tests/golden/ that oracle/gen_golden.py produced from the reference itself.
 *
 * Every function cites the reference lines it follows (paths relative to the reference root).
 * Rounding points reproduced: D.1 act->F16 before F16 mat-mul, D.2 Q8_0
 * activation quantisation, D.3 ffn_down on F32 activations, D.4 K/V/Q/P F16 in attention, D.5
 * RMSNorm sum in double, D.7 RoPE recurrence, D.8 GELU via F16 table, D.9 im2col F16.
 * The SIMD summation order of ggml_vec_dot_* (AVX2+FMA build) and the compiler's FMA contraction
 * of the RoPE rotation are reproduced too, so logits/hidden states are BIT-EXACT, not just close.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

enum { ORC_F32 = 0, ORC_F16 = 1, ORC_Q8_0 = 8 };

/* ---- F16 <-> F32, round-to-nearest-even, bit-exact with F16C (ggml-impl.h GGML_COMPUTE_FP32_TO_FP16) */
static inline float h2f(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ff, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else { int e = -1; do { e++; man <<= 1; } while ((man & 0x400) == 0); man &= 0x3ff; bits = sign | ((uint32_t) (127 - 15 - e) << 23) | (man << 13); }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f; memcpy(&f, &bits, 4); return f;
}
static inline uint16_t f2h(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t) ((x >> 16) & 0x8000);
    x &= 0x7fffffffu;
    if (x >= 0x7f800000u) return (uint16_t) (sign | 0x7c00 | (x > 0x7f800000u ? 0x200 : 0));
    if (x >= 0x477ff000u) return (uint16_t) (sign | 0x7c00);           /* overflow -> inf */
    if (x < 0x38800000u) {                                             /* subnormal half */
        if (x < 0x33000000u) return sign;
        const int shift = 113 - (int) (x >> 23);
        uint32_t man = (x & 0x7fffffu) | 0x800000u;
        const uint32_t lsb = 1u << (shift + 13), half = lsb >> 1;
        uint32_t r = man >> (shift + 13);
        const uint32_t rem = man & (lsb - 1);
        if (rem > half || (rem == half && (r & 1))) r++;
        return (uint16_t) (sign | r);
    }
    uint32_t r = ((x - 0x38000000u) >> 13);
    const uint32_t rem = x & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (r & 1))) r++;
    return (uint16_t) (sign | r);
}
static inline float r16(float f) { return h2f(f2h(f)); }

ORC_API float orc_r16(float f) { return r16(f); }
ORC_API void orc_f16_to_f32(const uint16_t * h, float * f, long n) { for (long i = 0; i < n; ++i) f[i] = h2f(h[i]); }

/* ---- Q8_0: block = { half d; int8 qs[32]; }  (ggml/src/ggml-common.h:219-224) */
typedef struct { uint16_t d; int8_t qs[32]; } __attribute__((packed)) blk_q8_0;

/* activation quantisation as the AVX2 path does it (ggml/src/ggml-cpu/arch/x86/quants.c:290-385):
 * d = amax/127 stored as F16, values scaled by id = 127/amax and rounded half-to-even */
static void quant_row_q8_0(const float * x, blk_q8_0 * y, int k) {
    for (int b = 0; b < k / 32; ++b) {
        float amax = 0.f;
        for (int j = 0; j < 32; ++j) { const float a = fabsf(x[b * 32 + j]); if (a > amax) amax = a; }
        const float d = amax / 127.f;
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        y[b].d = f2h(d);
        for (int j = 0; j < 32; ++j) y[b].qs[j] = (int8_t) nearbyintf(x[b * 32 + j] * id);
    }
}

/* ---- dot products in the exact order of the AVX2+FMA build of ggml-cpu -----------------------
 * ggml_vec_dot_f16 / ggml_vec_dot_f32 (vec.cpp:265-390, :11-137) keep 4 accumulator registers of
 * 8 lanes: element k goes to accumulator (k%32) with a fused multiply-add, in order of k; the
 * reduction (simd-mappings.h:549-567) is (s0+s2)+(s1+s3) per lane, lane l + lane l+4, then
 * (t0+t1)+(t2+t3).  Leftovers (n%32) are added sequentially: in double for F16, float FMA for F32.
 * Reproducing this order makes the restatement (and the CUDA kernels) BIT-EXACT with the
 * reference instead of merely close: F16 rounding points amplify any re-association noise. */
static inline float reduce32(const float * a) {
    float l[8], t[4];
    for (int i = 0; i < 8; ++i) l[i] = (a[i] + a[16 + i]) + (a[8 + i] + a[24 + i]);
    for (int i = 0; i < 4; ++i) t[i] = l[i] + l[i + 4];
    return (t[0] + t[1]) + (t[2] + t[3]);
}
static float dot_f16_hh(const uint16_t * x, const uint16_t * y, int n) {      /* both F16 */
    float a[32]; for (int i = 0; i < 32; ++i) a[i] = 0.f;
    const int np = n & ~31;
    for (int i = 0; i < np; i += 32) for (int j = 0; j < 32; ++j) a[j] = fmaf(h2f(x[i + j]), h2f(y[i + j]), a[j]);
    double sumf = (double) reduce32(a);
    for (int i = np; i < n; ++i) sumf += (double) (h2f(x[i]) * h2f(y[i]));
    return (float) sumf;
}
static float dot_f32_ff(const float * x, const float * y, int n) {
    float a[32]; for (int i = 0; i < 32; ++i) a[i] = 0.f;
    const int np = n & ~31;
    for (int i = 0; i < np; i += 32) for (int j = 0; j < 32; ++j) a[j] = fmaf(x[i + j], y[i + j], a[j]);
    float sumf = reduce32(a);
    for (int i = np; i < n; ++i) sumf = fmaf(x[i], y[i], sumf);         /* g++ -O3 -mfma contracts `sumf += x*y` */
    return sumf;
}
ORC_API float orc_dot_f16(const uint16_t * x, const uint16_t * y, int n) { return dot_f16_hh(x, y, n); }
ORC_API float orc_dot_f32(const float * x, const float * y, int n) { return dot_f32_ff(x, y, n); }

/* y[m][n] = W[n][:] . x[m][:]  with the oracle's activation rounding for the weight type
 * (ggml/src/ggml-cpu/ggml-cpu.c:1139-1331; vec.cpp:265-390; arch/x86/quants.c:1012-1078) */
ORC_API void orc_linear(const void * W, int wtype, int N, int K, const float * x, int M, float * y) {
    if (wtype == ORC_F16) {
        uint16_t * xh = (uint16_t *) malloc((size_t) K * sizeof(uint16_t));
        for (int m = 0; m < M; ++m) {
            for (int k = 0; k < K; ++k) xh[k] = f2h(x[(size_t) m * K + k]);
            for (int n = 0; n < N; ++n) y[(size_t) m * N + n] = dot_f16_hh((const uint16_t *) W + (size_t) n * K, xh, K);
        }
        free(xh);
    } else if (wtype == ORC_Q8_0) {
        const int nb = K / 32;
        blk_q8_0 * xq = (blk_q8_0 *) malloc((size_t) nb * sizeof(blk_q8_0));
        for (int m = 0; m < M; ++m) {
            quant_row_q8_0(x + (size_t) m * K, xq, K);
            for (int n = 0; n < N; ++n) {
                /* ggml_vec_dot_q8_0_q8_0, AVX2 (arch/x86/quants.c:1029-1046): lane j of the 8-wide
                 * accumulator takes the int sum of bytes 4j..4j+3 of every block (maddubs + madd),
                 * acc[j] = fma(dW*dA, (float) sum4, acc[j]); hsum_float_8 (:43-49) folds
                 * ((a4+a0)+(a6+a2)) + ((a5+a1)+(a7+a3)) */
                const blk_q8_0 * w = (const blk_q8_0 *) W + (size_t) n * nb;
                float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                for (int b = 0; b < nb; ++b) {
                    const float d = h2f(w[b].d) * h2f(xq[b].d);
                    for (int j = 0; j < 8; ++j) {
                        int32_t s4 = 0;
                        for (int e = 0; e < 4; ++e) s4 += (int32_t) w[b].qs[4 * j + e] * (int32_t) xq[b].qs[4 * j + e];
                        a[j] = fmaf(d, (float) s4, a[j]);
                    }
                }
                y[(size_t) m * N + n] = ((a[4] + a[0]) + (a[6] + a[2])) + ((a[5] + a[1]) + (a[7] + a[3]));
            }
        }
        free(xq);
    } else {            /* F32 weights: plain F32 dot */
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) y[(size_t) m * N + n] = dot_f32_ff((const float *) W + (size_t) n * K, x + (size_t) m * K, K);
    }
}

/* ffn_down: ggml_cast(W -> F32) then F32 x F32, activations NOT rounded
 * (src/tts_transformer.cpp:1389-1390; ggml-quants.c:401-415 dequantize_row_q8_0) */
ORC_API void orc_linear_wf32(const void * W, int wtype, int N, int K, const float * x, int M, float * y) {
    float * wf = (float *) malloc((size_t) K * sizeof(float));
    for (int n = 0; n < N; ++n) {
        if (wtype == ORC_F16) {
            const uint16_t * w = (const uint16_t *) W + (size_t) n * K;
            for (int k = 0; k < K; ++k) wf[k] = h2f(w[k]);
        } else if (wtype == ORC_Q8_0) {
            const blk_q8_0 * w = (const blk_q8_0 *) W + (size_t) n * (K / 32);
            for (int k = 0; k < K; ++k) wf[k] = (float) w[k / 32].qs[k % 32] * h2f(w[k / 32].d);
        } else memcpy(wf, (const float *) W + (size_t) n * K, (size_t) K * sizeof(float));
        for (int m = 0; m < M; ++m) y[(size_t) m * N + n] = dot_f32_ff(wf, x + (size_t) m * K, K);
    }
    free(wf);
}

/* ggml_rms_norm + ggml_mul (ggml/src/ggml-cpu/ops.cpp:3666-3691) */
ORC_API void orc_rmsnorm(const float * x, const float * gamma, int n, float eps, float * y) {
    double sum = 0.0;
    for (int i = 0; i < n; ++i) sum += (double) (x[i] * x[i]);
    const float mean = (float) (sum / n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int i = 0; i < n; ++i) { const float t = x[i] * scale; y[i] = gamma ? t * gamma[i] : t; }
}

/* ggml_rope_ext NEOX, freq_scale 1, ext_factor 0, attn_factor 1 (ops.cpp:5548-5562, 5636-5652) */
ORC_API void orc_rope_neox(float * v, int n_dims, int pos, float freq_base) {
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    float theta = (float) pos;
    for (int i = 0; i < n_dims / 2; ++i) {
        const float c = cosf(theta), s = sinf(theta);
        const float x0 = v[i], x1 = v[i + n_dims / 2];
        /* g++ -O3 -mfma contracts the rotation: fmsub(x0, c, x1*s) and fmadd(x0, s, x1*c)
         * (verified node by node against the reference with the eval-callback tracer) */
        v[i] = fmaf(x0, c, -(x1 * s));
        v[i + n_dims / 2] = fmaf(x0, s, x1 * c);
        theta *= theta_scale;
    }
}

/* ggml_v_expf, AVX2+FMA variant (ggml/src/ggml-cpu/vec.h:1230-1267), one lane */
static inline float as_f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t as_u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float v_expf(float x) {
    const float r = 0x1.8p23f;
    const float z = fmaf(x, 0x1.715476p+0f, r);
    const float n = z - r;
    const float b = fmaf(-n, 0x1.7f7d1cp-20f, fmaf(-n, 0x1.62e4p-1f, x));
    const uint32_t e = as_u(z) << 23;
    const float k = as_f(e + as_u(1.0f));
    const int c = fabsf(n) > 126.0f;
    const float u = b * b;
    const float j = fmaf(fmaf(fmaf(0x1.0e4020p-7f, b, 0x1.573e2ep-5f), u, fmaf(0x1.555e66p-3f, b, 0x1.fffdb6p-2f)), u, 0x1.ffffecp-1f * b);
    if (!c) return fmaf(j, k, k);
    const uint32_t g = (n <= 0.0f) ? 0x82000000u : 0u;
    const float s1 = as_f(g + 0x7f000000u);
    const float s2 = as_f(e - g);
    if (fabsf(n) > 192.0f) return s1 * s1;
    return fmaf(s2, j, s2) * s1;
}
ORC_API float orc_v_expf(float x) { return v_expf(x); }
static void silu_vec(float * y, const float * x, int n);
ORC_API void orc_silu(float * y, const float * x, int n) { silu_vec(y, x, n); }
/* ggml_vec_silu_f32 (vec.cpp:397-432): 8-wide ggml_v_silu, scalar tail x/(1+expf(-x)) */
static void silu_vec(float * y, const float * x, int n) {
    int i = 0;
    for (; i + 7 < n; i += 8) for (int j = 0; j < 8; ++j) { const float v = x[i + j]; y[i + j] = v / (1.0f + v_expf(0.0f - v)); }
    for (; i < n; ++i) y[i] = x[i] / (1.0f + expf(-x[i]));
}
/* ggml_vec_soft_max_f32 (vec.cpp:548-617): y = exp(x-max); returns the sum (double) */
static double soft_max_vec(float * y, const float * x, int n, float max) {
    int i = 0; double sum = 0.0;
    for (; i + 7 < n; i += 8) {
        float v[8];
        for (int j = 0; j < 8; ++j) { v[j] = v_expf(x[i + j] - max); y[i + j] = v[j]; }
        const float a0 = v[4] + v[0], a1 = v[5] + v[1], a2 = v[6] + v[2], a3 = v[7] + v[3];
        sum += (double) ((a0 + a2) + (a1 + a3));
    }
    for (; i < n; ++i) { const float val = expf(x[i] - max); sum += (double) val; y[i] = val; }
    return sum;
}

/* -------------------------------------------------------------------------------------------
 * One decoder layer stack (talker or code predictor), M tokens at positions pos0..pos0+M-1.
 * Follows build_step_graph / build_prefill_forward_graph / build_code_pred_step_graph
 * (src/tts_transformer.cpp:1302-1393, 1157-1248, 1722-1813).  With f32_attn != 0 it follows
 * build_code_pred_prefill_graph (:1577-1657) where attention multiplies the fresh F32
 * Kcur/Vcur/Q and P is not rounded (pos0 must be 0 then).
 * KV cache layout is the reference's [n_ctx][n_kv_head][head_dim] F16 per layer (:714-724).
 */
typedef struct {
    int wtype;
    const void * wq, * wk, * wv, * wo, * wgate, * wup, * wdown;
    const float * attn_norm, * ffn_norm, * q_norm, * k_norm;
} orc_layer;

typedef struct {
    int n_layers, hidden, n_head, n_kv_head, head_dim, ffn;
    float eps, rope_theta;
    const orc_layer * layers;
    const float * out_norm;
    uint16_t * kcache, * vcache;     /* [L][n_ctx][n_kv_head][head_dim] */
    int n_ctx;
} orc_stack;

ORC_API void orc_stack_forward(const orc_stack * s, const float * x_in, int M, int pos0, int f32_attn,
                               float * hidden_out /* [M][hidden] after the final RMSNorm */) {
    const int Hd = s->hidden, NHh = s->n_head, NK = s->n_kv_head, D = s->head_dim, F = s->ffn;
    const int QN = NHh * D, KN = NK * D;
    const float kq_scale = 1.0f / sqrtf((float) D);
    float * x  = (float *) malloc((size_t) M * Hd * sizeof(float));
    float * h  = (float *) malloc((size_t) M * Hd * sizeof(float));
    float * q  = (float *) malloc((size_t) M * QN * sizeof(float));
    float * k  = (float *) malloc((size_t) M * KN * sizeof(float));
    float * v  = (float *) malloc((size_t) M * KN * sizeof(float));
    float * ao = (float *) malloc((size_t) M * QN * sizeof(float));
    float * t1 = (float *) malloc((size_t) M * Hd * sizeof(float));
    float * g  = (float *) malloc((size_t) M * F * sizeof(float));
    float * u  = (float *) malloc((size_t) M * F * sizeof(float));
    const int n_kv_max = pos0 + M;
    float * sc = (float *) malloc((size_t) n_kv_max * sizeof(float));
    float * pr = (float *) malloc((size_t) n_kv_max * sizeof(float));
    float * vt = (float *) malloc((size_t) n_kv_max * sizeof(float));
    uint16_t * ph = (uint16_t *) malloc((size_t) n_kv_max * sizeof(uint16_t));
    uint16_t * vth = (uint16_t *) malloc((size_t) n_kv_max * sizeof(uint16_t));
    uint16_t qh16[512];
    memcpy(x, x_in, (size_t) M * Hd * sizeof(float));

    for (int l = 0; l < s->n_layers; ++l) {
        const orc_layer * L = &s->layers[l];
        for (int m = 0; m < M; ++m) orc_rmsnorm(x + (size_t) m * Hd, L->attn_norm, Hd, s->eps, h + (size_t) m * Hd);
        orc_linear(L->wq, L->wtype, QN, Hd, h, M, q);
        orc_linear(L->wk, L->wtype, KN, Hd, h, M, k);
        orc_linear(L->wv, L->wtype, KN, Hd, h, M, v);
        uint16_t * kc = s->kcache + (size_t) l * s->n_ctx * KN;
        uint16_t * vc = s->vcache + (size_t) l * s->n_ctx * KN;
        for (int m = 0; m < M; ++m) {
            const int pos = pos0 + m;
            for (int hh = 0; hh < NHh; ++hh) {
                float * qh = q + (size_t) m * QN + hh * D;
                if (L->q_norm) orc_rmsnorm(qh, L->q_norm, D, s->eps, qh);
                orc_rope_neox(qh, D, pos, s->rope_theta);
            }
            for (int hh = 0; hh < NK; ++hh) {
                float * kh = k + (size_t) m * KN + hh * D;
                if (L->k_norm) orc_rmsnorm(kh, L->k_norm, D, s->eps, kh);
                orc_rope_neox(kh, D, pos, s->rope_theta);
            }
            for (int i = 0; i < KN; ++i) {          /* ggml_cpy F32 -> F16 cache view (:1337-1348) */
                kc[(size_t) pos * KN + i] = f2h(k[(size_t) m * KN + i]);
                vc[(size_t) pos * KN + i] = f2h(v[(size_t) m * KN + i]);
            }
        }
        for (int m = 0; m < M; ++m) {
            const int pos = pos0 + m, n_kv = pos + 1;        /* diag_mask_inf(n_past): causal */
            for (int hh = 0; hh < NHh; ++hh) {
                const int kvh = hh / (NHh / NK);
                const float * qh = q + (size_t) m * QN + hh * D;
                float * o = ao + (size_t) m * QN + hh * D;
                if (f32_attn) {
                    /* the graph computes all M x M scores, masks, and soft-maxes rows of length M */
                    for (int t = 0; t < M; ++t)
                        sc[t] = (t <= m) ? dot_f32_ff(k + (size_t) t * KN + kvh * D, qh, D) * kq_scale : -INFINITY;
                    float mx = -INFINITY;
                    for (int t = 0; t < M; ++t) if (sc[t] > mx) mx = sc[t];
                    const double sum = soft_max_vec(pr, sc, M, mx);
                    const float inv = (float) (1.0 / sum);
                    for (int t = 0; t < M; ++t) pr[t] *= inv;
                    for (int d = 0; d < D; ++d) {
                        for (int t = 0; t < M; ++t) vt[t] = v[(size_t) t * KN + kvh * D + d];
                        o[d] = dot_f32_ff(vt, pr, M);
                    }
                } else {
                    /* the graph scores all W = pos0+M cache rows, masks t > pos with -inf
                     * (diag_mask_inf) and soft-maxes / multiplies rows of length W
                     * (src/tts_transformer.cpp:1205-1227): W decides which positions go through
                     * the 8-wide SIMD exp vs the libm tail, and the 32-wide part of the P.V dot */
                    const int W = pos0 + M;
                    for (int d = 0; d < D; ++d) qh16[d] = f2h(qh[d]);
                    float mx = -INFINITY;
                    for (int t = 0; t < W; ++t) {
                        sc[t] = (t < n_kv) ? dot_f16_hh(kc + (size_t) t * KN + kvh * D, qh16, D) * kq_scale : -INFINITY;
                        if (sc[t] > mx) mx = sc[t];
                    }
                    const double sum = soft_max_vec(pr, sc, W, mx);      /* ops.cpp:5243-5259 */
                    const float inv = (float) (1.0 / sum);
                    for (int t = 0; t < W; ++t) ph[t] = f2h(pr[t] * inv);
                    for (int d = 0; d < D; ++d) {
                        for (int t = 0; t < W; ++t) vth[t] = vc[(size_t) t * KN + kvh * D + d];
                        o[d] = dot_f16_hh(vth, ph, W);
                    }
                }
            }
        }
        orc_linear(L->wo, L->wtype, Hd, QN, ao, M, t1);
        for (size_t i = 0; i < (size_t) M * Hd; ++i) x[i] = t1[i] + x[i];                 /* inpFF */
        for (int m = 0; m < M; ++m) orc_rmsnorm(x + (size_t) m * Hd, L->ffn_norm, Hd, s->eps, h + (size_t) m * Hd);
        orc_linear(L->wgate, L->wtype, F, Hd, h, M, g);
        orc_linear(L->wup, L->wtype, F, Hd, h, M, u);
        for (int m = 0; m < M; ++m) silu_vec(g + (size_t) m * F, g + (size_t) m * F, F);
        for (size_t i = 0; i < (size_t) M * F; ++i) g[i] = g[i] * u[i];
        orc_linear_wf32(L->wdown, L->wtype, Hd, F, g, M, t1);
        for (size_t i = 0; i < (size_t) M * Hd; ++i) x[i] = t1[i] + x[i];
    }
    if (hidden_out) {
        if (s->out_norm) for (int m = 0; m < M; ++m) orc_rmsnorm(x + (size_t) m * Hd, s->out_norm, Hd, s->eps, hidden_out + (size_t) m * Hd);
        else memcpy(hidden_out, x, (size_t) M * Hd * sizeof(float));
    }
    free(x); free(h); free(q); free(k); free(v); free(ao); free(t1); free(g); free(u); free(sc); free(pr); free(vt); free(ph); free(vth);
}

/* argmax with first-maximum-wins (src/tts_transformer.cpp:2182-2192) */
ORC_API int orc_argmax(const float * d, int n) {
    int mi = 0; float mv = d[0];
    for (int i = 1; i < n; ++i) if (d[i] > mv) { mv = d[i]; mi = i; }
    return mi;
}

/* cb0 logit post-processing of generate() (:2664-2683): suppress then HF repetition penalty */
ORC_API void orc_process_cb0_logits(float * logits, int vocab, int eos_id, const uint8_t * seen, float rep_penalty) {
    for (int i = vocab - 1024; i < vocab; ++i) if (i != eos_id) logits[i] = -INFINITY;
    if (rep_penalty != 1.0f)
        for (int i = 0; i < vocab; ++i)
            if (seen[i]) logits[i] = logits[i] > 0.0f ? logits[i] / rep_penalty : logits[i] * rep_penalty;
}
