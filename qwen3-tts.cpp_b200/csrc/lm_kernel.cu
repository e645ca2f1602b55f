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
//     dependency chain: one thread per CTA issues TMA-unit bulk L2 prefetches
//     (cp.async.bulk.prefetch.L2) one layer ahead, and every thread pulls its 16-byte slices
//     of the NEXT phase's weight rows into registers *before* waiting on the grid barrier, so
//     the barrier latency is covered by loads in flight.
//   * Every phase boundary is an all-to-all dependency (GEMV output -> next GEMV input); it is
//     implemented as a release/acquire counter barrier in global memory (one red.release +
//     one polling thread per CTA), not a kernel boundary.
//   * All rounding points of the oracle are reproduced: activations
//     are rounded to F16 before F16 mat-vecs, ffn_down runs on F32 activations, K/V are stored
//     F16, P is rounded to F16 after a *global* softmax normalisation, RMSNorm sums in double,
//     RoPE uses a host table built with the oracle's recurrence.
#include "common.cuh"
#include "engine.h"

#include <cooperative_groups.h>
#include <math_constants.h>

namespace q3t {

constexpr int MAXV   = 12;     // 16-byte weight vectors a thread may hold for one phase
constexpr int MAXU   = MAXV * 4;
constexpr int SC_MAX = 1024;   // max KV positions per attention chunk

struct __align__(16) Smem {
    float  xres[H];            // residual stream of the current token (x, then x1)
    float  xin_f[FFN];         // F32 GEMV input (ffn_down) / scratch
    __half xin_h[FFN];         // F16-rounded GEMV input
    float  red[MAXU * 4];      // per-unit warp partials
    float  outv[MAXU];         // finalised row values of the phase
    // attention
    __half qh[2][HD];          // normed+roped q of the two heads of this kv group (F16-rounded)
    float  qf[2][HD];          // same, unrounded (F32-attention mode)
    __half knew_h[HD], vnew_h[HD];
    float  knew_f[HD], vnew_f[HD];
    float  sc[2][SC_MAX];      // scores of this CTA's chunk
    float  red_o[16][2][HD];   // per-warp partial outputs
    float  wred[2][16][2];     // per-warp (max,sum) for both heads
    float  gms[2][2];          // global (max, inv_sum)
    double dred[16];
    float  fred[32];
    int    ired[32];
    int    tok;
};

// --------------------------------------------------------------------------- barriers
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

// barrier among the `n` CTAs that share kv group g (attention softmax statistics exchange)
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

// --------------------------------------------------------------------------- row slices
__device__ __forceinline__ void cta_rows(int n_rows, int & r0, int & r1) {
    r0 = (int) (((long long) blockIdx.x * n_rows) / gridDim.x);
    r1 = (int) (((long long) (blockIdx.x + 1) * n_rows) / gridDim.x);
}

// A *unit* is 1024 consecutive weights of one row (128 threads x 8 halves). A row of
// K = NSEG*1024 is NSEG consecutive units; a CTA's row slice is a contiguous unit range.
__device__ __forceinline__ void prefetch_w(uint4 (&w)[MAXV], const __half * slice, int n_units) {
    const int q = threadIdx.x >> 7, k = threadIdx.x & 127;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int u = i * 4 + q;
        if (u < n_units) w[i] = ld_weight(slice + (size_t) u * 1024 + k * 8);
    }
}

// F16 weights x F16-rounded activations (oracle D.1)
template <int NSEG>
__device__ __forceinline__ void gemv_units_h(const uint4 (&w)[MAXV], int n_units, Smem & s) {
    const int q = threadIdx.x >> 7, k = threadIdx.x & 127, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int u = i * 4 + q;
        if (u < n_units) {
            const int seg = u % NSEG;
            const uint4 xv = *reinterpret_cast<const uint4 *>(s.xin_h + seg * 1024 + k * 8);
            float p = dot8_hh(w[i], xv, 0.0f);
            p = warp_sum(p);
            if (lane == 0) s.red[u * 4 + (warp & 3)] = p;
        }
    }
}

// F16 weights (exact in F32) x F32 activations (oracle D.3: ffn_down)
template <int NSEG>
__device__ __forceinline__ void gemv_units_f(const uint4 (&w)[MAXV], int n_units, Smem & s) {
    const int q = threadIdx.x >> 7, k = threadIdx.x & 127, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int u = i * 4 + q;
        if (u < n_units) {
            const int seg = u % NSEG;
            const float4 * xp = reinterpret_cast<const float4 *>(s.xin_f + seg * 1024 + k * 8);
            float p = dot8_hf(w[i], xp[0], xp[1], 0.0f);
            p = warp_sum(p);
            if (lane == 0) s.red[u * 4 + (warp & 3)] = p;
        }
    }
}

// after __syncthreads: thread r < n_rows folds the NSEG*4 partials of row r in a fixed order
template <int NSEG>
__device__ __forceinline__ float fold_row(const Smem & s, int r) {
    float v = 0.0f;
#pragma unroll
    for (int j = 0; j < NSEG * 4; ++j) v += s.red[r * NSEG * 4 + j];
    return v;
}

// --------------------------------------------------------------------------- RMSNorm
// y = (x * rsqrt(mean(x^2)+eps)) * gamma, sum of squares in double (oracle D.5,
// ggml/src/ggml-cpu/ops.cpp:3666-3691); result rounded to F16 into xin_h, optional F32 copy.
__device__ __forceinline__ void rmsnorm_to_h(Smem & s, const float * __restrict__ gamma, float eps,
                                             float * hidden_f32_out /* smem/global or null */) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const float a = s.xres[t], b = s.xres[t + 512];
    double sq = (double) (a * a) + (double) (b * b);
    sq = warp_sum_d(sq);
    if (lane == 0) s.dred[warp] = sq;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) tot += s.dred[i];
    const float mean  = (float) (tot / (double) H);
    const float scale = 1.0f / sqrtf(mean + eps);
    const float ya = (a * scale) * __ldg(gamma + t);
    const float yb = (b * scale) * __ldg(gamma + t + 512);
    s.xin_h[t]       = __float2half_rn(ya);
    s.xin_h[t + 512] = __float2half_rn(yb);
    if (hidden_f32_out) { hidden_f32_out[t] = ya; hidden_f32_out[t + 512] = yb; }
    __syncthreads();
}

// per-head RMSNorm(128) * gamma then NEOX RoPE, one warp; lane l owns dims {l, l+32, l+64, l+96}
// (pairs (l,l+64) and (l+32,l+96)); rope row = cos[64] | sin[64] (oracle D.7).
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
        for (int i = 0; i < 4; ++i) sq += (double) (v[i] * v[i]);
        sq = warp_sum_d(sq);
        const float mean  = (float) (sq / (double) HD);
        const float scale = 1.0f / sqrtf(mean + eps);
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = (v[i] * scale) * __ldg(gamma + lane + 32 * i);
        const float c0 = __ldg(rope_row + lane),      s0 = __ldg(rope_row + 64 + lane);
        const float c1 = __ldg(rope_row + lane + 32), s1 = __ldg(rope_row + 64 + lane + 32);
        out[0] = v[0] * c0 - v[2] * s0;
        out[2] = v[0] * s0 + v[2] * c0;
        out[1] = v[1] * c1 - v[3] * s1;
        out[3] = v[1] * s1 + v[3] * c1;
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = v[i];
    }
}

// --------------------------------------------------------------------------- attention (phase B)
// CTA role: kv group g = blockIdx % 8 (q heads 2g, 2g+1 share kv head g — GQA by broadcast,
// ggml-cpu.c:1160-1199), KV chunk c = blockIdx / 8.  Softmax statistics are exchanged between
// the C chunk-CTAs of a group so that P can be rounded to F16 *after* the global normalisation,
// exactly where the oracle rounds it.
__device__ __forceinline__ int attn_chunks(int n_kv, bool f32attn) {
    if (f32attn) return 1;
    int C = (n_kv + 31) >> 5;
    const int maxc = min((int) gridDim.x / NKV, LM_MAXC);
    if (C > maxc) C = maxc;
    if (C < 1) C = 1;
    return C;
}

__device__ void attention_phase(const DevModel & m, const StackW & W, int l, int pos, bool f32attn,
                                Smem & s, BarState & bs) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int g = blockIdx.x % NKV, c = blockIdx.x / NKV;
    const int n_kv = pos + 1;
    const int C  = attn_chunks(n_kv, f32attn);
    const int Lc = (n_kv + C - 1) / C;
    const bool active = c < C;
    const int t0 = min(c * Lc, n_kv), t1 = min(n_kv, t0 + Lc);
    const bool owns_new = active && (pos >= t0) && (pos < t1);
    const float kq_scale = 1.0f / sqrtf((float) HD);
    const size_t kv_head_base = ((size_t) (l * NKV + g) * W.n_ctx) * HD;

    if (active) {
        if (warp < 4) {
            const float * rope_row = m.rope + (size_t) pos * HD;
            float o[4];
            if (warp < 2) {
                head_norm_rope(m.qkv + (2 * g + warp) * HD, W.q_norm + l * HD, rope_row, m.eps, true, o);
#pragma unroll
                for (int i = 0; i < 4; ++i) { s.qf[warp][lane + 32 * i] = o[i]; s.qh[warp][lane + 32 * i] = __float2half_rn(o[i]); }
            } else if (warp == 2) {
                head_norm_rope(m.qkv + NH * HD + g * HD, W.k_norm + l * HD, rope_row, m.eps, true, o);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int d = lane + 32 * i;
                    const __half hv = __float2half_rn(o[i]);
                    s.knew_f[d] = o[i]; s.knew_h[d] = hv;
                    if (owns_new) {
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
                    if (owns_new) {
                        W.vcache[kv_head_base + (size_t) pos * HD + d] = hv;
                        if (W.vf32 && pos < 2) W.vf32[((size_t) (l * 2 + pos) * NKV + g) * HD + d] = o[i];
                    }
                }
            }
        }
        __syncthreads();
    }

    if (f32attn) {
        // code-predictor prefill numerics: K, V, Q and P all stay F32 because the reference graph
        // multiplies the fresh F32 Kcur/Vcur, not the F16 cache (src/tts_transformer.cpp:1624-1635).
        if (active && warp < 2) {
            float sc_t[2];
            float mx = -CUDART_INF_F;
            for (int tt = 0; tt <= pos && tt < 2; ++tt) {
                const float * krow = (tt == pos) ? s.knew_f : (W.kf32 + ((size_t) (l * 2 + tt) * NKV + g) * HD);
                float d = 0.0f;
#pragma unroll
                for (int i = 0; i < 4; ++i) d = fmaf(krow[lane + 32 * i], s.qf[warp][lane + 32 * i], d);
                d = warp_sum(d) * kq_scale;
                sc_t[tt] = d;
                mx = fmaxf(mx, d);
            }
            double sum = 0.0;
            for (int tt = 0; tt <= pos && tt < 2; ++tt) { sc_t[tt] = expf(sc_t[tt] - mx); sum += (double) sc_t[tt]; }
            const float inv = (float) (1.0 / sum);
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            for (int tt = 0; tt <= pos && tt < 2; ++tt) {
                const float * vrow = (tt == pos) ? s.vnew_f : (W.vf32 + ((size_t) (l * 2 + tt) * NKV + g) * HD);
                const float p = sc_t[tt] * inv;
#pragma unroll
                for (int i = 0; i < 4; ++i) o[i] = fmaf(vrow[lane + 32 * i], p, o[i]);
            }
            float * dst = m.part_o + ((size_t) (g * LM_MAXC + 0) * 2 + warp) * HD;
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[lane + 32 * i] = o[i];
        }
        return;
    }

    const int hlf = lane >> 4, sub = lane & 15;
    const int n_loc = t1 - t0;
    float Mh[2] = {-CUDART_INF_F, -CUDART_INF_F};
    float inv_h[2] = {0.f, 0.f};

    if (active) {
        // ---- stage 1: scores s_t = (K_t . r16(q)) * scale
        float qa[8], qb[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { qa[j] = __half2float(s.qh[0][sub * 8 + j]); qb[j] = __half2float(s.qh[1][sub * 8 + j]); }
        for (int base = t0 + warp * 2; base < t1; base += 32) {
            const int tt = base + hlf;
            const bool valid = tt < t1;
            uint4 kv = make_uint4(0, 0, 0, 0);
            if (valid) {
                const __half * krow = (tt == pos) ? s.knew_h : (W.kcache + kv_head_base + (size_t) tt * HD);
                kv = *reinterpret_cast<const uint4 *>(krow + sub * 8);
            }
            const __half2 * kh = reinterpret_cast<const __half2 *>(&kv);
            float da = 0.f, db = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 kf = __half22float2(kh[j]);
                da = fmaf(kf.x, qa[2 * j], da); da = fmaf(kf.y, qa[2 * j + 1], da);
                db = fmaf(kf.x, qb[2 * j], db); db = fmaf(kf.y, qb[2 * j + 1], db);
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) { da += __shfl_xor_sync(0xffffffffu, da, o); db += __shfl_xor_sync(0xffffffffu, db, o); }
            if (valid && sub == 0) { s.sc[0][tt - t0] = da * kq_scale; s.sc[1][tt - t0] = db * kq_scale; }
        }
        __syncthreads();
        // ---- chunk-local max and sum(exp)
        float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F;
        for (int i = t; i < n_loc; i += LM_THREADS) { mx0 = fmaxf(mx0, s.sc[0][i]); mx1 = fmaxf(mx1, s.sc[1][i]); }
        mx0 = warp_max(mx0); mx1 = warp_max(mx1);
        if (lane == 0) { s.wred[0][warp][0] = mx0; s.wred[1][warp][0] = mx1; }
        __syncthreads();
        mx0 = -CUDART_INF_F; mx1 = -CUDART_INF_F;
#pragma unroll
        for (int i = 0; i < 16; ++i) { mx0 = fmaxf(mx0, s.wred[0][i][0]); mx1 = fmaxf(mx1, s.wred[1][i][0]); }
        double sm0 = 0.0, sm1 = 0.0;
        for (int i = t; i < n_loc; i += LM_THREADS) { sm0 += (double) expf(s.sc[0][i] - mx0); sm1 += (double) expf(s.sc[1][i] - mx1); }
        sm0 = warp_sum_d(sm0); sm1 = warp_sum_d(sm1);
        __syncthreads();
        if (lane == 0) { s.dred[warp] = sm0; }
        __syncthreads();
        double S0 = 0.0;
#pragma unroll
        for (int i = 0; i < 16; ++i) S0 += s.dred[i];
        __syncthreads();
        if (lane == 0) { s.dred[warp] = sm1; }
        __syncthreads();
        double S1 = 0.0;
#pragma unroll
        for (int i = 0; i < 16; ++i) S1 += s.dred[i];
        if (C == 1) {
            Mh[0] = mx0; Mh[1] = mx1;
            inv_h[0] = (float) (1.0 / S0); inv_h[1] = (float) (1.0 / S1);
        } else if (t == 0) {
            float * pm = m.part_ms + (size_t) (g * LM_MAXC + c) * 4;
            pm[0] = mx0; pm[1] = (float) S0; pm[2] = mx1; pm[3] = (float) S1;
        }
    }
    if (C > 1) {
        group_barrier(m, bs, g, C, active);
        if (active) {
            float M0 = -CUDART_INF_F, M1 = -CUDART_INF_F;
            for (int i = 0; i < C; ++i) {
                const float * pm = m.part_ms + (size_t) (g * LM_MAXC + i) * 4;
                M0 = fmaxf(M0, ld_act(pm + 0)); M1 = fmaxf(M1, ld_act(pm + 2));
            }
            double S0 = 0.0, S1 = 0.0;
            for (int i = 0; i < C; ++i) {
                const float * pm = m.part_ms + (size_t) (g * LM_MAXC + i) * 4;
                const float s0 = ld_act(pm + 1), s1 = ld_act(pm + 3);
                if (s0 > 0.f) S0 += (double) s0 * (double) expf(ld_act(pm + 0) - M0);
                if (s1 > 0.f) S1 += (double) s1 * (double) expf(ld_act(pm + 2) - M1);
            }
            Mh[0] = M0; Mh[1] = M1;
            inv_h[0] = (float) (1.0 / S0); inv_h[1] = (float) (1.0 / S1);
        }
    }
    if (!active) return;

    // ---- stage 2: o = sum_t r16(p_t) * V_t with p_t = expf(s_t - M) * (1/S)
    float aa[8], ab[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { aa[j] = 0.f; ab[j] = 0.f; }
    for (int base = t0 + warp * 2; base < t1; base += 32) {
        const int tt = base + hlf;
        if (tt < t1) {
            const __half * vrow = (tt == pos) ? s.vnew_h : (W.vcache + kv_head_base + (size_t) tt * HD);
            const uint4 vv = *reinterpret_cast<const uint4 *>(vrow + sub * 8);
            const __half2 * vh = reinterpret_cast<const __half2 *>(&vv);
            const float pa = r16(expf(s.sc[0][tt - t0] - Mh[0]) * inv_h[0]);
            const float pb = r16(expf(s.sc[1][tt - t0] - Mh[1]) * inv_h[1]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 vf = __half22float2(vh[j]);
                aa[2 * j] = fmaf(pa, vf.x, aa[2 * j]); aa[2 * j + 1] = fmaf(pa, vf.y, aa[2 * j + 1]);
                ab[2 * j] = fmaf(pb, vf.x, ab[2 * j]); ab[2 * j + 1] = fmaf(pb, vf.y, ab[2 * j + 1]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        aa[j] += __shfl_xor_sync(0xffffffffu, aa[j], 16);
        ab[j] += __shfl_xor_sync(0xffffffffu, ab[j], 16);
    }
    if (hlf == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { s.red_o[warp][0][sub * 8 + j] = aa[j]; s.red_o[warp][1][sub * 8 + j] = ab[j]; }
    }
    __syncthreads();
    if (t < 2 * HD) {
        const int h = t >> 7, d = t & 127;
        float o = 0.f;
#pragma unroll
        for (int w = 0; w < 16; ++w) o += s.red_o[w][h][d];
        m.part_o[((size_t) (g * LM_MAXC + c) * 2 + h) * HD + d] = o;
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
    int * hist = reinterpret_cast<int *>(s.sc[0]);           // 256 bins + 2 control words
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
    if (lane == 31) s.dred[warp] = inc;
    __syncthreads();
    double woff = 0.0, total = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) { if (i < warp) woff += s.dred[i]; total += s.dred[i]; }
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

// --------------------------------------------------------------------------- weight slices
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

// --------------------------------------------------------------------------- one pass
// One token through `n_layers` layers of stack W (+ optional head).  s.xres holds x on entry;
// `w` holds the phase-A weights of layer 0 on entry (prefetched by the caller).
__device__ void run_stack(const DevModel & m, const StackW & W, const PassDesc & p, int pos,
                          const HeadW * head, bool process, const GenState & gs,
                          const Slices & z, const __half * next_a0, int next_a0_units,
                          Smem & s, BarState & bs, uint4 (&w)[MAXV]) {
    const int t = threadIdx.x;
    const int nA = z.a1 - z.a0, nO = z.o1 - z.o0, nP = z.p1 - z.p0, nD = z.d1 - z.d0;
    int h0 = 0, h1 = 0;
    if (head) cta_rows(head->vocab, h0, h1);

    for (int l = 0; l < p.n_layers; ++l) {
        if (t == 0) {
            if (l + 1 < p.n_layers) l2_prefetch_layer(W, l + 1, z);
            else if (head) l2_prefetch_range(head->w + (size_t) h0 * H, (size_t) (h1 - h0) * H * 2);
        }
        // ---- phase A: RMSNorm + fused QKV mat-vec
        rmsnorm_to_h(s, W.attn_norm + (size_t) l * H, m.eps, nullptr);
        gemv_units_h<1>(w, nA, s);
        __syncthreads();
        if (t < nA) m.qkv[z.a0 + t] = fold_row<1>(s, t);
        grid_barrier(m, bs);

        // ---- phase B: per-head norm + RoPE + KV append + attention
        prefetch_w(w, sl_o(W, l, z), 2 * nO);
        attention_phase(m, W, l, pos, p.f32attn != 0, s, bs);
        grid_barrier(m, bs);

        // ---- phase C: attn_output mat-vec + residual
        {
            const int C = attn_chunks(pos + 1, p.f32attn != 0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int idx = t + LM_THREADS * i;
                const int j = idx >> 7, d = idx & 127;
                const float * src = m.part_o + ((size_t) ((j >> 1) * LM_MAXC) * 2 + (j & 1)) * HD + d;
                float o = 0.f;
                for (int c = 0; c < C; ++c) o += ld_act(src + (size_t) c * 2 * HD);
                s.xin_h[idx] = __float2half_rn(o);
            }
        }
        __syncthreads();
        gemv_units_h<2>(w, 2 * nO, s);
        __syncthreads();
        if (t < nO) m.xb[z.o0 + t] = fold_row<2>(s, t) + s.xres[z.o0 + t];
        prefetch_w(w, sl_gu(W, l, z), 2 * nP);
        grid_barrier(m, bs);

        // ---- phase D: RMSNorm + gate/up mat-vec + SwiGLU
        s.xres[t] = ld_act(m.xb + t); s.xres[t + 512] = ld_act(m.xb + t + 512);
        __syncthreads();
        rmsnorm_to_h(s, W.ffn_norm + (size_t) l * H, m.eps, nullptr);
        gemv_units_h<1>(w, 2 * nP, s);
        __syncthreads();
        if (t < nP) {
            const float gt = fold_row<1>(s, 2 * t), up = fold_row<1>(s, 2 * t + 1);
            m.mbuf[z.p0 + t] = (gt / (1.0f + expf(-gt))) * up;
        }
        prefetch_w(w, sl_d(W, l, z), 3 * nD);
        grid_barrier(m, bs);

        // ---- phase E: ffn_down on F32 activations + residual
#pragma unroll
        for (int i = 0; i < FFN / LM_THREADS; ++i) s.xin_f[t + LM_THREADS * i] = ld_act(m.mbuf + t + LM_THREADS * i);
        __syncthreads();
        gemv_units_f<3>(w, 3 * nD, s);
        __syncthreads();
        if (t < nD) m.xa[z.d0 + t] = fold_row<3>(s, t) + s.xres[z.d0 + t];
        if (l + 1 < p.n_layers) prefetch_w(w, sl_a(W, l + 1, z), nA);
        else if (head)          prefetch_w(w, head->w + (size_t) h0 * H, h1 - h0);
        else if (next_a0)       prefetch_w(w, next_a0, next_a0_units);
        grid_barrier(m, bs);
        s.xres[t] = ld_act(m.xa + t); s.xres[t + 512] = ld_act(m.xa + t + 512);
        __syncthreads();
    }

    if (!head) return;
    // ---- head: final RMSNorm (= hidden_states) + vocabulary mat-vec + per-CTA arg-max candidate
    rmsnorm_to_h(s, W.out_norm, m.eps, blockIdx.x == 0 ? m.hidden : nullptr);
    if (p.hidden_dst && blockIdx.x == 0) {
        // xres * scale * gamma was just written to m.hidden by this CTA's threads
        p.hidden_dst[t] = m.hidden[t]; p.hidden_dst[t + 512] = m.hidden[t + 512];
    }
    const int nH = h1 - h0;
    gemv_units_h<1>(w, nH, s);
    __syncthreads();
    if (t < nH) {
        const float raw = fold_row<1>(s, t);
        m.logits[h0 + t] = raw;
        if (p.logits_dst) p.logits_dst[h0 + t] = raw;
        s.outv[t] = proc_logit(m, raw, h0 + t, process, gs.suppress_start, gs.eos_id, gs.rep_penalty);
    }
    __syncthreads();
    if (t == 0) {
        float bv = -CUDART_INF_F; int bi = 0x7fffffff;
        for (int i = 0; i < nH; ++i) if (s.outv[i] > bv) { bv = s.outv[i]; bi = h0 + i; }
        if (nH > 0 && bi == 0x7fffffff) bi = h0;       // all -inf: keep first-index semantics
        m.cand_val[blockIdx.x] = bv; m.cand_idx[blockIdx.x] = bi;
    }
    if (next_a0) prefetch_w(w, next_a0, next_a0_units);
    grid_barrier(m, bs);
}

// --------------------------------------------------------------------------- the kernel
// frame_mode = 1: `passes` is the 17-pass frame program; the kernel owns the generate-loop
// bookkeeping of src/tts_transformer.cpp:2664-2804 (EOS test, code emission, repetition set,
// n_past) in GenState so the identical launch can be replayed once per frame.
__global__ void __launch_bounds__(LM_THREADS, 1)
lm_kernel(const __grid_constant__ DevModel m, const PassDesc * __restrict__ passes, int n_passes, int frame_mode) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
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

    uint4 w[MAXV];
    int frame = gs.frame, n_past = gs.n_past, done = gs.done, pending = gs.pending;
    int pending_vocab = gs.codec_vocab, pending_process = 1;
    unsigned long long draw = gs.draw;
    bool have_w = false;
    int  have_w_stack = -1;

    for (int ip = 0; ip < n_passes; ++ip) {
        if (t == 0) pd = passes[ip];
        __syncthreads();
        const PassDesc p = pd;
        const StackW & W = p.stack == 0 ? m.talker : m.cp;

        // start pulling layer 0 of this stack towards L2 / registers before the (serial) selection
        if (p.xmode != XM_SELECT_ONLY) {
            if (t == 0) l2_prefetch_layer(W, 0, z);
            if (!(have_w && have_w_stack == p.stack)) prefetch_w(w, sl_a(W, 0, z), z.a1 - z.a0);
        }
        have_w = false;

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
                        xa += __half2float(q[t]); xb += __half2float(q[t + 512]);
                    }
                    const int fr = frame - 1;      // the frame just emitted
                    const float * tr = (fr < gs.trailing_len) ? (m.trailing + (size_t) fr * H) : m.pad_embed;
                    xa += ld_act(tr + t); xb += ld_act(tr + t + 512);
                } break;
                default: break;
            }
            s.xres[t] = xa; s.xres[t + 512] = xb;
        }
        __syncthreads();

        const int pos = p.pos >= 0 ? p.pos : n_past;
        const HeadW * head = p.head < 0 ? nullptr : (p.head == 0 ? &m.codec_head : &m.cp_head[p.head - 1]);
        // next pass' first weight slice (register prefetch across the pass boundary)
        const __half * next_a0 = nullptr; int next_units = 0; int next_stack = -1;
        if (ip + 1 < n_passes) {
            const PassDesc & np = passes[ip + 1];
            if (np.xmode != XM_SELECT_ONLY) {
                next_stack = np.stack;
                const StackW & NW = np.stack == 0 ? m.talker : m.cp;
                next_a0 = sl_a(NW, 0, z); next_units = z.a1 - z.a0;
            }
        }
        run_stack(m, W, p, pos, head, p.process != 0, gs, z, next_a0, next_units, s, bs, w);
        have_w = next_a0 != nullptr; have_w_stack = next_stack;
        if (head) { pending = p.head; pending_vocab = head->vocab; pending_process = p.process; }
        if (p.stack == 0 && p.pos < 0) ++n_past;
    }

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
        if (g < LM_MIN_GRID) return set_err(ctx, Q3T_ERR_CUDA, "device cannot co-schedule the persistent decode kernel (needs >= 128 SMs)");
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
