// Persistent decode kernel for the talker and the code predictor (sm_100a).
//
// What it replaces in the reference: build_step_graph / build_code_pred_prefill_graph /
// build_code_pred_step_graph and the host loop that runs 16 dependent ggml graphs per frame
// (src/tts_transformer.cpp:1268-1411, 1534-1829, 2310-2575, 2664-2804).
//
// Design (B200-first):
//   * ONE cooperative launch of `grid` (= #SMs) CTAs x 384 threads runs a whole *program* of
//     passes (a pass = one token through one layer stack + optional head): a frame is
//     16 code-predictor passes + 1 talker pass, a prefill is M talker passes.
//   * Batch-1 decode is a chain of ~600 dependent mat-vecs per frame, so the kernel is built around
//     the LATENCY of that chain while HBM streams underneath it:
//       - warp 11 of every CTA is a producer: it walks the (deterministic) schedule of weight
//         slices of this CTA and keeps a 160 KB shared-memory ring full with TMA bulk copies
//         (cp.async.bulk global->shared, mbarrier complete_tx), gated only by ring space — the
//         weight stream never waits for the dependency chain; a slice is issued as 8 KB pieces a
//         few dozen cycles apart, with L2 cache policies (the code predictor's first slices
//         evict_last, everything read once per frame evict_first): bursts of the weight stream
//         and HBM queueing were what made the packet exchanges slow;
//       - warps 0..10 compute.  Every phase boundary is an all-to-all exchange of a small
//         activation vector; instead of a grid barrier (fence + atomic + poll + reload = 4 L2 round
//         trips) activations travel as 8-byte {value, epoch} packets (one store, one polling
//         load: the value is valid iff its epoch matches) — the NCCL "LL" idea used inside one GPU;
//       - small parameters (norm weights) are fetched one layer ahead into registers.
//   * Weights are stored chain-major (engine.cu k_chain_pack) so that lane a of a warp is ggml's
//     accumulator lane a (k % 32): it keeps x[32j + a] in registers as packed halves and walks its
//     FMA chain with FHFMA (fma.rn.f32.f16: F16 x F16 + F32 in one instruction, no conversions).
//   * BIT-EXACT with the oracle: every dot product is accumulated in the order of ggml-cpu's
//     AVX2+FMA build (32 accumulator lanes k%32, FMA chain over k/32, fixed reduction tree;
//     ggml/src/ggml-cpu/vec.cpp:265-390, simd-mappings.h:549-567), exp() is the same polynomial
//     (vec.h:1230-1267), soft-max sums follow the 8-lane pattern (vec.cpp:558-568).
//     oracle/q3t_oracle.c is the executable statement of these rules.
#include "common.cuh"
#include "engine.h"

#include <math_constants.h>
#include <stdlib.h>

namespace q3t {

// ------------------------------------------------------------------------------- shared memory
constexpr int NCW        = LM_THREADS / 32 - 1;        // compute warps (the last warp is the TMA producer)
constexpr int NCT        = NCW * 32;                   // compute threads
constexpr int SLOT_B     = 2048;                       // ring slot = one K=1024 row of F16 weights
constexpr int NSLOT      = 80;                         // 160 KB ring
constexpr int KV_CH      = 128;                        // K/V history rows per ring slice (16 slots)
constexpr int SLOT_Q8    = 1088;                       // Q8_0 record of 1024 weights: 1024 int8 + 32 F16 block scales
constexpr int NMB        = 8;                          // mbarriers cycling over the slices in flight
constexpr int ATT_MAX_KV = LM_ATT_MAX_KV;              // positions the attention phase can hold
constexpr int WIDE_G     = 8;                          // wide attention: CTAs per head,
constexpr int VS_D       = HD / WIDE_G;                // dims per member (16 halves = 32 bytes per position),
constexpr int VS_CH      = 512;                        // positions per V slice (16 KB = 8 ring slots)
constexpr int ATT_FAST_KV = 16;                        // rows up to this width take the register-fed short path (no K/V ring slices)

struct __align__(128) Smem {
    static constexpr int NSLOTS = NSLOT;
    unsigned char ring[NSLOT * SLOT_B];
    float  xres[H];                // residual stream of the current token (x, then x1)
    float  xf[FFN];                // F32 GEMV input (ffn_down) / scratch
    unsigned xh[H];                // F16 GEMV input, lane-packed: xh[m*32 + a] = {x[64m + a], x[64m + 32 + a]}
    float  outv[64];               // finalised row values of the phase
    signed char xq[2048];          // Q8_0 mat-vec input: int8 values, lane-major per group of 4 blocks (see gemv_q8)
    float  xd[64];                 // ... and its block scales (F16-rounded, as float)
    // attention (one CTA per head)
    float  sc[ATT_MAX_KV];         // scores, then exp values
    __half ph[ATT_MAX_KV];         // soft-max probabilities rounded to F16 (oracle D.4)
    float  qf[HD];                 // normed+roped q (F32)
    float  qr[HD];                 // same, F16-rounded, as float
    float  knew_f[HD], vnew_f[HD];
    __half knew_h[HD], vnew_h[HD]; // new token's K and V rounded to F16 (natural dim order)
    float  o32[HD];
    __half vt[32 * HD];            // leftover V rows (positions np..n_kv-1)
    double dred[16];
    float  fred[32];
    int    ired[32];
    int    tok;
    int    prof_cnt;
    unsigned long long mbar[NMB];  // TMA completion barriers of the slices in flight
    volatile unsigned long long consumed;   // low 32: ring slots released, high 32: slices released
    volatile unsigned tokens;               // exchanges completed by the compute warps
};

// ------------------------------------------------------------------------------- phase profiler
// thread 0 of EVERY CTA appends (clock64 << 8 | tag) to its own region of m.prof with plain stores
// (no round trip: the marks do not perturb the chain); the count lives in shared memory and is
// written to m.prof_n[cta] when the kernel ends, so the buffer holds the last launch.
__device__ __forceinline__ void prof_mark(const DevModel & m, int * cnt, int tag) {
    if (m.prof && threadIdx.x == 0 && *cnt >= 0) {
        const int i = *cnt;
        if (i < m.prof_cap) {
            m.prof[(size_t) blockIdx.x * m.prof_cap + i] = ((unsigned long long) clock64() << 8) | (unsigned long long) (tag & 0xff);
            *cnt = i + 1;
        }
    }
}
#define PROF(tag) prof_mark(m, &s.prof_cnt, tag)

// ------------------------------------------------------------------------------- CTA-level sync
// barrier among the 352 compute threads only (the producer warp runs free)
__device__ __forceinline__ void cbar() { asm volatile("bar.sync 1, %0;" :: "n"(NCT) : "memory"); }

// ------------------------------------------------------------------------------- LL packets
// An activation travels as {float value, u32 epoch} in one aligned 8-byte word: an 8-byte store is
// single-copy atomic, so a reader that sees the expected epoch also sees the value.  Every exchange
// point of the program has its own epoch (a counter that advances identically in every CTA).
struct Watch { long long t0; bool dead; int * err; };   // err: the launch's device error flag (GenState::error)

__device__ __forceinline__ void pk_store(Pkt * p, float v, unsigned ep) {
    // gpu-scope relaxed (SASS .STRONG.GPU): `volatile` would be system scope, which is far slower
    const unsigned long long u = (unsigned long long) __float_as_uint(v) | ((unsigned long long) ep << 32);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(u) : "memory");
}
__device__ __forceinline__ void spin_cycles(int n) {
    if (n <= 0) return;
    const long long t0 = clock64();
    while (clock64() - t0 < n) { }
}
__device__ __forceinline__ bool pk_give_up(const DevModel & m, Watch & w, unsigned & spins) {
    if (((++spins) & 0x3ff) != 0) return false;
    if (*reinterpret_cast<volatile int *>(w.err)) { w.dead = true; return true; }
    if (clock64() - w.t0 > 6000000000LL) { *w.err = 1; w.dead = true; return true; }   // ~3 s: a CTA died
    return false;
}
__device__ __forceinline__ float pk_poll(const DevModel & m, const Pkt * p, unsigned ep, Watch & w) {
    unsigned long long u;
    unsigned spins = 0;
    while (true) {
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(u) : "l"(p) : "memory");
        if ((unsigned) (u >> 32) == ep || w.dead) break;
        if (pk_give_up(m, w, spins)) break;
    }
    return __uint_as_float((unsigned) u);
}
// two adjacent packets (16-byte aligned) in one load
__device__ __forceinline__ float2 pk_poll2(const DevModel & m, const Pkt * p, unsigned ep, Watch & w) {
    unsigned long long u0, u1;
    unsigned spins = 0;
    while (true) {
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(u0), "=l"(u1) : "l"(p) : "memory");
        if (((unsigned) (u0 >> 32) == ep && (unsigned) (u1 >> 32) == ep) || w.dead) break;
        if (pk_give_up(m, w, spins)) break;
    }
    return make_float2(__uint_as_float((unsigned) u0), __uint_as_float((unsigned) u1));
}

// N packets per thread with every load in flight at once: a thread that polls its packets one after
// the other pays one L2 round trip per packet even when all of them have already arrived.
// OPT: optimistic first round -- plain loads and ONE accumulated epoch test (~3 instructions per packet instead of ~10
// for the predicated re-poll loop).  It pays where the packets have normally landed before the poll starts (the batched
// kernel, whose groups hide each other's exchanges: 2 % per frame); in the batch-1 kernel, which arrives at every poll
// early, the wasted round made the frame 6 % slower, so it stays off there.
template <int N, bool OPT = false>
__device__ __forceinline__ void pk_poll_n(const DevModel & m, const Pkt * p, int stride, int n_valid, unsigned ep,
                                          float (&v)[N], Watch & w) {
    unsigned long long u[N];
    unsigned pending = n_valid >= N ? (N >= 32 ? 0xffffffffu : (1u << (N & 31)) - 1u) : (1u << n_valid) - 1u;
    unsigned spins = 0;
    if (OPT && n_valid >= N) {
#pragma unroll
        for (int j = 0; j < N; ++j)
            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(u[j]) : "l"(p + (size_t) j * stride) : "memory");
        unsigned bad = 0u;
#pragma unroll
        for (int j = 0; j < N; ++j) bad |= (unsigned) (u[j] >> 32) ^ ep;
        if (bad == 0u) pending = 0u;
        else {
#pragma unroll
            for (int j = 0; j < N; ++j) if ((unsigned) (u[j] >> 32) == ep) pending &= ~(1u << j);
        }
    } else {
#pragma unroll
        for (int j = 0; j < N; ++j) u[j] = 0ull;
    }
    while (pending) {
#pragma unroll
        for (int j = 0; j < N; ++j)
            asm volatile("{ .reg .pred q; setp.ne.u32 q, %2, 0; @q ld.relaxed.gpu.global.u64 %0, [%1]; }"
                         : "+l"(u[j]) : "l"(p + (size_t) j * stride), "r"(pending & (1u << j)) : "memory");
#pragma unroll
        for (int j = 0; j < N; ++j) if ((unsigned) (u[j] >> 32) == ep) pending &= ~(1u << j);
        if (w.dead) break;
        if (pending) { if (pk_give_up(m, w, spins)) break; spin_cycles(m.poll_d1); }
    }
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = __uint_as_float((unsigned) u[j]);
}
// same for N adjacent pairs (16-byte loads); pair j of the thread is at p + 2 * j * stride_pairs
template <int N>
__device__ __forceinline__ void pk_poll2_n(const DevModel & m, const Pkt * p, int stride_pairs, int n_valid, unsigned ep,
                                           float2 (&v)[N], Watch & w) {
    unsigned long long u0[N], u1[N];
    unsigned pending = n_valid >= N ? (1u << N) - 1u : (1u << n_valid) - 1u;
    unsigned spins = 0;
#pragma unroll
    for (int j = 0; j < N; ++j) { u0[j] = 0ull; u1[j] = 0ull; }
    while (pending) {
#pragma unroll
        for (int j = 0; j < N; ++j)
            asm volatile("{ .reg .pred q; setp.ne.u32 q, %3, 0; @q ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2]; }"
                         : "+l"(u0[j]), "+l"(u1[j]) : "l"(p + (size_t) 2 * j * stride_pairs), "r"(pending & (1u << j)) : "memory");
#pragma unroll
        for (int j = 0; j < N; ++j)
            if ((unsigned) (u0[j] >> 32) == ep && (unsigned) (u1[j] >> 32) == ep) pending &= ~(1u << j);
        if (w.dead) break;
        if (pending) { if (pk_give_up(m, w, spins)) break; spin_cycles(m.poll_d1); }
    }
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = make_float2(__uint_as_float((unsigned) u0[j]), __uint_as_float((unsigned) u1[j]));
}

// ------------------------------------------------------------------------------- TMA weight ring
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
// pol: L2 cache policy of the copy (createpolicy), 0 = none
__device__ __forceinline__ void tma_load_1d(void * dst_smem, const void * src_gmem, uint32_t bytes, unsigned long long * bar,
                                            unsigned long long pol = 0ull) {
    if (pol)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                     :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
    else
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void cta_rows(int n_rows, int & r0, int & r1) {
    r0 = (int) (((long long) blockIdx.x * n_rows) / gridDim.x);
    r1 = (int) (((long long) (blockIdx.x + 1) * n_rows) / gridDim.x);
}

struct Slices {            // this CTA's row ranges (fixed for the whole launch)
    int a0, a1;            // qkv rows            (of 4096, K=1024)
    int o0, o1;            // attn_output rows    (of 1024, K=2048)
    int p0, p1;            // gate/up pairs       (of 3072, K=1024, 2 rows per pair)
    int d0, d1;            // ffn_down rows       (of 1024, K=3072)
};

// consumer-side cursor over the slice stream (identical bookkeeping runs in the producer)
struct Ring {
    unsigned slot_head;    // ring slot (unwrapped) where the next slice starts
    unsigned slice_idx;    // index of the next slice
};

__device__ __forceinline__ bool mbar_try(unsigned long long * bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// wait for the next slice; returns its first ring slot (wrapped)
template <class SM>
__device__ __forceinline__ int ring_acquire(const DevModel & m, SM & s, const Ring & rg, Watch & w) {
    unsigned long long * bar = &s.mbar[rg.slice_idx % NMB];
    const uint32_t parity = (rg.slice_idx / NMB) & 1u;
    unsigned spins = 0;
    // one lane per warp waits (every thread hitting the same mbarrier serialises for ~1 cycle each);
    // __syncwarp orders the other lanes' reads of the slice after the observed completion
    if ((threadIdx.x & 31) == 0) {
        while (!mbar_try(bar, parity)) {
            if (w.dead) break;
            if (pk_give_up(m, w, spins)) break;
        }
    }
    __syncwarp();
    return (int) (rg.slot_head % SM::NSLOTS);
}
// all compute threads are past their last read of the slice (call after a cbar())
template <class SM>
__device__ __forceinline__ void ring_release(SM & s, Ring & rg, int n_slots) {
    rg.slot_head += (unsigned) n_slots;
    rg.slice_idx += 1u;
    if (threadIdx.x == 0) {
        __threadfence_block();
        s.consumed = ((unsigned long long) rg.slice_idx << 32) | (unsigned long long) rg.slot_head;
    }
}

// producer: one slice = `n_slots` ring slots filled from `src` (contiguous in HBM); <= 2 bulk copies
template <class SM>
__device__ __forceinline__ void produce_slice(SM & s, unsigned & p_slot, unsigned & p_idx, const void * src, int n_slots,
                                              int slot_bytes = SLOT_B, int tail_bytes = SLOT_B, int pace = 0, int piece_slots = 4,
                                              unsigned long long pol = 0ull) {
    if (n_slots <= 0) return;
    // wait for ring space and for the mbarrier of slice (p_idx - NMB) to be drained by the consumers
    while (true) {
        const unsigned long long c = s.consumed;
        const unsigned c_slot = (unsigned) c, c_idx = (unsigned) (c >> 32);
        if (p_slot + (unsigned) n_slots - c_slot <= (unsigned) SM::NSLOTS && p_idx - c_idx < (unsigned) NMB) break;
    }
    unsigned long long * bar = &s.mbar[p_idx % NMB];
    const int w0 = (int) (p_slot % SM::NSLOTS);
    const int n0 = min(n_slots, SM::NSLOTS - w0);
    // order the consumers' generic-proxy reads of the recycled slots before the async-proxy writes
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (slot_bytes == SLOT_B) {
        // contiguous source; the last slot may be partly filled (tail_bytes, multiple of 16)
        const uint32_t total = (uint32_t) (n_slots - 1) * SLOT_B + (uint32_t) tail_bytes;
        const uint32_t first = min(total, (uint32_t) n0 * SLOT_B);
        mbar_expect_tx(bar, total);
        if (pace > 0) {
            // paced: pieces of `piece_slots` ring slots `pace` cycles apart.  A slice is requested the moment its ring space is
            // released, i.e. exactly when the phase's packet exchange starts; as ONE bulk copy (up to 84 KB) from each of 148
            // SMs that is an 8-12 MB burst into HBM / L2 during the latency-critical store -> poll round trip.  8 KB pieces 150
            // cycles apart: frame 1.614 -> 1.486 ms, talker step 437 -> 406 us on one box (sweep in profiles/r02b_pacing_sweep.txt)
            const uint32_t PIECE = (uint32_t) piece_slots * SLOT_B;
            for (uint32_t off = 0; off < total; off += PIECE) {
                const uint32_t n = min(PIECE, total - off);
                const uint32_t in_first = off < first ? min(n, first - off) : 0u;
                if (in_first) tma_load_1d(s.ring + (size_t) w0 * SLOT_B + off, (const unsigned char *) src + off, in_first, bar, pol);
                if (n > in_first) tma_load_1d(s.ring + (off + in_first - first), (const unsigned char *) src + off + in_first, n - in_first, bar, pol);
                if (off + PIECE < total) spin_cycles(pace);
            }
        } else {
            tma_load_1d(s.ring + (size_t) w0 * SLOT_B, src, first, bar, pol);
            if (total > first) tma_load_1d(s.ring, (const unsigned char *) src + first, total - first, bar, pol);
        }
    } else {
        mbar_expect_tx(bar, (uint32_t) n_slots * (uint32_t) slot_bytes);
        // compact records in HBM (pitch slot_bytes) land at the start of 2 KB ring slots: one copy per slot
        for (int i = 0; i < n_slots; ++i) {
            int slot = w0 + i; if (slot >= SM::NSLOTS) slot -= SM::NSLOTS;
            tma_load_1d(s.ring + (size_t) slot * SLOT_B, (const unsigned char *) src + (size_t) i * slot_bytes, (uint32_t) slot_bytes, bar, pol);
        }
    }
    p_slot += (unsigned) n_slots;
    p_idx  += 1u;
}

// first byte of this CTA's rows of layer l (F16: 2048 B per 1024 weights, Q8_0: 1088 B records)
__device__ __forceinline__ const unsigned char * sl_rows(const void * plane, int wtype, size_t row, int spr) {
    return reinterpret_cast<const unsigned char *>(plane) + row * (size_t) spr * (wtype == Q3T_TYPE_Q8_0 ? SLOT_Q8 : SLOT_B);
}
__device__ __forceinline__ const void * sl_a(const StackW & W, int l, const Slices & z) {
    return sl_rows(W.wqkv, W.wtype, (size_t) l * QKV_N + z.a0, 1);
}
__device__ __forceinline__ const void * sl_o(const StackW & W, int l, const Slices & z) {
    return sl_rows(W.wo, W.wtype, (size_t) l * H + z.o0, 2);
}
__device__ __forceinline__ const void * sl_gu(const StackW & W, int l, const Slices & z) {
    return sl_rows(W.wgu, W.wtype, (size_t) l * 2 * FFN + 2 * z.p0, 1);
}
__device__ __forceinline__ const void * sl_d(const StackW & W, int l, const Slices & z) {
    return sl_rows(W.wdown, W.wtype, (size_t) l * H + z.d0, 3);
}

// ------------------------------------------------------------------------------- ordered GEMV
// fma.rn.f32.f16 (SASS FHFMA, sm_100+): d = f32(a) * f32(b) + c with a, b F16 and one rounding —
// exactly ggml's F16C-convert + vfmadd lane step, in one instruction.
__device__ __forceinline__ float fhfma_lo(unsigned w, unsigned x, float acc) {
    asm("{.reg .b16 wl, wh, xl, xh; mov.b32 {wl, wh}, %1; mov.b32 {xl, xh}, %2; fma.rn.f32.f16 %0, wl, xl, %0;}" : "+f"(acc) : "r"(w), "r"(x));
    return acc;
}
__device__ __forceinline__ float fhfma_hi(unsigned w, unsigned x, float acc) {
    asm("{.reg .b16 wl, wh, xl, xh; mov.b32 {wl, wh}, %1; mov.b32 {xl, xh}, %2; fma.rn.f32.f16 %0, wh, xh, %0;}" : "+f"(acc) : "r"(w), "r"(x));
    return acc;
}
// 16-byte shared-memory load that keeps its place in program order
__device__ __forceinline__ uint4 lds_v4(const unsigned char * p) {
    uint4 v;
    asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ uint2 lds_v2(const unsigned char * p) {
    uint2 v;
    asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ unsigned u4_get(const uint4 & v, int e) { return e == 0 ? v.x : e == 1 ? v.y : e == 2 ? v.z : v.w; }

// Rows of chain-major F16 weights in the shared-memory ring (slice starting at slot s0, K/1024 slots
// per row) times the activation vector.  Lane a of a warp IS accumulator lane a of
// ggml_vec_dot_f16 / ggml_vec_dot_f32 (k % 32 == a): it holds x[32j + a] for all j in registers
// (packed halves from s.xh for the F16 mat-vecs, floats from s.xf for ffn_down) and walks its FMA
// chain over j for R rows at a time; one 16-byte load brings 8 consecutive chain steps and a warp
// reads 512 contiguous bytes per load.  The 32 lane sums are folded in the oracle's order
// (reduce32_lanes).  Result of row r -> s.outv[r].
template <int K, int R, bool XF32>
__device__ __forceinline__ void gemv_ring(Smem & s, int s0, int n_rows) {
    constexpr int SPR = K / 1024, NU = K / 256, NJ = K / 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp * R >= n_rows) return;
    // XF32: the activation is read from shared memory step by step (conflict-free, addresses known up front);
    // holding all K/32 = 96 values in registers left ptxas no room to issue the weight loads ahead of their use
    const float * xr = s.xf + lane;
    unsigned xp[XF32 ? 1 : NJ / 2];
    if (!XF32) {
#pragma unroll
        for (int m2 = 0; m2 < NJ / 2; ++m2) xp[m2] = s.xh[m2 * 32 + lane];
    }
    for (int rb = warp * R; rb < n_rows; rb += NCW * R) {
        float acc[R];
        int   sb[R];
#pragma unroll
        for (int i = 0; i < R; ++i) {
            acc[i] = 0.0f;
            const int r = rb + i < n_rows ? rb + i : rb;
            sb[i] = s0 + r * SPR;
        }
        // ffn_down (XF32, one row = one FMA chain per warp): ptxas issues each 16-byte weight load right before its
        // use and the chain eats the full shared-memory latency 12 times per row; volatile loads keep their program
        // order, two k-chunks ahead of the arithmetic
        uint4 wq[XF32 ? 2 : 1][R];
        if (XF32) {
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int i = 0; i < R; ++i) {
                    int slot = sb[i] + (a >> 2);
                    if (slot >= NSLOT) slot -= NSLOT;
                    wq[a][i] = lds_v4(s.ring + (size_t) slot * SLOT_B + (a & 3) * 512 + lane * 16);
                }
        }
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            uint4 wv[R];
#pragma unroll
            for (int i = 0; i < R; ++i) {
                if (XF32) {
                    wv[i] = wq[u & 1][i];
                    if (u + 2 < NU) {
                        int slot = sb[i] + ((u + 2) >> 2);
                        if (slot >= NSLOT) slot -= NSLOT;
                        wq[u & 1][i] = lds_v4(s.ring + (size_t) slot * SLOT_B + ((u + 2) & 3) * 512 + lane * 16);
                    }
                } else {
                    int slot = sb[i] + (u >> 2);
                    if (slot >= NSLOT) slot -= NSLOT;
                    wv[i] = *reinterpret_cast<const uint4 *>(s.ring + (size_t) slot * SLOT_B + (u & 3) * 512 + lane * 16);
                }
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
#pragma unroll
                for (int i = 0; i < R; ++i) {
                    const unsigned w2 = u4_get(wv[i], e);
                    if (XF32) {
                        const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w2));
                        acc[i] = __fmaf_rn(f.x, xr[32 * (8 * u + 2 * e)], acc[i]);
                        acc[i] = __fmaf_rn(f.y, xr[32 * (8 * u + 2 * e + 1)], acc[i]);
                    } else {
                        acc[i] = fhfma_lo(w2, xp[4 * u + e], acc[i]);
                        acc[i] = fhfma_hi(w2, xp[4 * u + e], acc[i]);
                    }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const float res = reduce32_lanes(acc[i]);
            if (lane == 0 && rb + i < n_rows) s.outv[rb + i] = res;
        }
    }
}

// ------------------------------------------------------------------------------- Q8_0 mat-vecs
// Activation quantisation of ggml's AVX2 quantize_row_q8_0 (arch/x86/quants.c:290-385): per block
// of 32, d = amax / 127 (stored as F16), q = round-half-even(x * (127 / amax)).  One warp per block.
// s.xf[0..K) -> s.xq (int8) + s.xd (scale as float).  Layout of s.xq per 1024 values: group c of 4
// blocks is 128 bytes, lane l (0..7) owns bytes [c*128 + l*16, +16) = its 4 values (4l..4l+3) of each
// of the 4 blocks -- the operand order of one dp4a per block in gemv_q8.  Ends with a cbar().
// N blocks of one warp at once (b0, b0 + NCW, ...; lane `lane` holds value `lane` of each): the N max-reductions
// are independent, so their shuffle latencies overlap, and the two IEEE divisions of block i (d = amax / 127,
// id = 127 / amax) are evaluated by lane i alone and broadcast -- one pair of division sequences per warp instead
// of N (this sits on the layer's critical path twice)
template <int N>
__device__ __forceinline__ void q8_blocks(Smem & s, int b0, int n_blocks, int lane, const float (&v)[N]) {
    float amax[N];
#pragma unroll
    for (int i = 0; i < N; ++i) amax[i] = fabsf(v[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int i = 0; i < N; ++i) amax[i] = fmaxf(amax[i], __shfl_xor_sync(0xffffffffu, amax[i], o));
    float am = amax[0];
#pragma unroll
    for (int i = 1; i < N; ++i) am = lane == i ? amax[i] : am;
    const float d_l  = __half2float(__float2half_rn(__fdiv_rn(am, 127.0f)));
    const float id_l = am != 0.0f ? __fdiv_rn(127.0f, am) : 0.0f;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const int b = b0 + i * NCW;
        const float id = __shfl_sync(0xffffffffu, id_l, i);
        const int q = __float2int_rn(__fmul_rn(v[i], id));
        const int bl = b & 31;
        if (b < n_blocks) {
            s.xq[(b >> 5) * 1024 + (bl >> 2) * 128 + (lane >> 2) * 16 + (bl & 3) * 4 + (lane & 3)] = (signed char) q;
            if (lane == i) s.xd[b] = d_l;
        }
    }
}
template <int K>
__device__ __forceinline__ void quant_q8(Smem & s) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int NB = K / 32, PER = (NB + NCW - 1) / NCW, G = 3;
#pragma unroll
    for (int i0 = 0; i0 < PER; i0 += G) {
        float v[G];
#pragma unroll
        for (int i = 0; i < G; ++i) {
            const int b = warp + (i0 + i) * NCW;
            v[i] = b < NB ? s.xf[b * 32 + lane] : 0.0f;
        }
        q8_blocks<G>(s, warp + i0 * NCW, NB, lane, v);
    }
    cbar();
}

// ggml_vec_dot_q8_0_q8_0, AVX2 (arch/x86/quants.c:1029-1046): accumulator lane j (0..7) takes the
// int sum of bytes 4j..4j+3 of every block: acc_j = fma(dW*dX, (float) sum4, acc_j) over the blocks
// in order, then hsum_float_8 (:43-49).  8 lanes per row, 4 rows per warp step; the ring record of
// 1024 weights is [8 groups x 128 B lane-major int8][32 F16 block scales] (engine.cu k_q8_pack).
template <int K>
__device__ __forceinline__ void gemv_q8(Smem & s, int s0, int n_rows) {
    constexpr int SPR = K / 1024;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 3, l = lane & 7;
    for (int rb = warp * 4; rb < n_rows; rb += NCW * 4) {
        const int r = rb + g < n_rows ? rb + g : rb;
        float acc = 0.0f;
#pragma unroll
        for (int ch = 0; ch < SPR; ++ch) {
            int slot = s0 + r * SPR + ch;
            slot %= NSLOT;
            const unsigned char * rec = s.ring + (size_t) slot * SLOT_B;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint4  wq = *reinterpret_cast<const uint4 *>(rec + c * 128 + l * 16);
                const uint4  xq = *reinterpret_cast<const uint4 *>(s.xq + ch * 1024 + c * 128 + l * 16);
                const uint2  ws = *reinterpret_cast<const uint2 *>(rec + 1024 + c * 8);
                const float4 xs = *reinterpret_cast<const float4 *>(s.xd + ch * 32 + c * 4);
                const float2 w01 = __half22float2(*reinterpret_cast<const __half2 *>(&ws.x));
                const float2 w23 = __half22float2(*reinterpret_cast<const __half2 *>(&ws.y));
                acc = __fmaf_rn(__fmul_rn(w01.x, xs.x), (float) __dp4a((int) wq.x, (int) xq.x, 0), acc);
                acc = __fmaf_rn(__fmul_rn(w01.y, xs.y), (float) __dp4a((int) wq.y, (int) xq.y, 0), acc);
                acc = __fmaf_rn(__fmul_rn(w23.x, xs.z), (float) __dp4a((int) wq.z, (int) xq.z, 0), acc);
                acc = __fmaf_rn(__fmul_rn(w23.y, xs.w), (float) __dp4a((int) wq.w, (int) xq.w, 0), acc);
            }
        }
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 4));
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 2));
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 1));
        if (l == 0 && rb + g < n_rows) s.outv[rb + g] = acc;
    }
}

// ffn_down with Q8_0 weights: the reference casts W to F32 (dequantize_row_q8_0: (float) q * d) and
// runs ggml_vec_dot_f32 on un-rounded activations (src/tts_transformer.cpp:1389-1390).  Lane a is
// accumulator lane a; the record keeps the int8 values chain-major ((j/8)*256 + a*8 + j%8 for block
// j of the 1024-chunk) so one 8-byte load brings 8 chain steps, and the 32 block scales behind them.
template <int K>
__device__ __forceinline__ void gemv_q8_f32(Smem & s, int s0, int n_rows) {
    constexpr int SPR = K / 1024, NU = K / 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= n_rows) return;
    // like the F16 ffn_down (gemv_ring<.., XF32>): the activation comes from shared memory step by step and the
    // weight loads are issued two 256-weight chunks ahead in program order -- with all 96 activations in registers
    // ptxas had no room to hoist them and every chunk paid the shared-memory latency inside the FMA chain
    const float * xr = s.xf + lane;
    for (int r = warp; r < n_rows; r += NCW) {
        float acc = 0.0f;
        uint2 qa[2]; uint4 sa[2];
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            int slot = s0 + r * SPR + (a >> 2);
            slot %= NSLOT;
            const unsigned char * rec = s.ring + (size_t) slot * SLOT_B;
            qa[a] = lds_v2(rec + (a & 3) * 256 + lane * 8);
            sa[a] = lds_v4(rec + 1024 + (a & 3) * 16);
        }
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            const uint2 q8 = qa[u & 1];
            const uint4 sc = sa[u & 1];
            if (u + 2 < NU) {
                int slot = s0 + r * SPR + ((u + 2) >> 2);
                slot %= NSLOT;
                const unsigned char * rec = s.ring + (size_t) slot * SLOT_B;
                qa[u & 1] = lds_v2(rec + ((u + 2) & 3) * 256 + lane * 8);
                sa[u & 1] = lds_v4(rec + 1024 + ((u + 2) & 3) * 16);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                // the record stores q + 128 (k_q8_pack mode 1): one byte-permute drops it into the mantissa of 2^23,
                // and (2^23 + 128 + q) - (2^23 + 128) is exactly (float) q -- no I2F (a quarter-rate instruction)
                const unsigned word = e < 4 ? q8.x : q8.y;
                const float qf = __fsub_rn(__uint_as_float(__byte_perm(word, 0x4B000000u, 0x7650u | (unsigned) (e & 3))), 8388736.0f);
                const unsigned sw = u4_get(sc, e >> 1);
                const float2 d2 = __half22float2(*reinterpret_cast<const __half2 *>(&sw));
                acc = __fmaf_rn(__fmul_rn(qf, (e & 1) ? d2.y : d2.x), xr[32 * (8 * u + e)], acc);
            }
        }
        const float res = reduce32_lanes(acc);
        if (lane == 0) s.outv[r] = res;
    }
}

// write x[k] rounded to F16 into the lane-packed GEMV input (see Smem::xh)
__device__ __forceinline__ void xh_put(Smem & s, int k, float v) {
    reinterpret_cast<__half *>(s.xh)[(((k >> 6) * 32 + (k & 31)) << 1) + ((k >> 5) & 1)] = __float2half_rn(v);
}

// ------------------------------------------------------------------------------- RMSNorm
// y = (x * rsqrt(mean(x^2)+eps)) * gamma; sum of squares in double (oracle D.5,
// ggml/src/ggml-cpu/ops.cpp:3666-3691).  x = s.xres; thread t owns elements t, t+352, t+704.
// Writes F16(y) into s.xh (the F16 mat-vec input) and optionally the unrounded y (hidden states).
// gamma values come in registers (fetched a layer ahead).  Ends with a cbar().
struct Gam { float g[3]; };
__device__ __forceinline__ Gam load_gamma(const float * __restrict__ gamma) {
    Gam r;
    const int t = threadIdx.x;
    r.g[0] = __ldg(gamma + t); r.g[1] = __ldg(gamma + t + NCT);
    r.g[2] = (t + 2 * NCT < H) ? __ldg(gamma + t + 2 * NCT) : 0.0f;
    return r;
}
template <bool Q8 = false>
__device__ __forceinline__ void rmsnorm_to_xh(Smem & s, const Gam & gm, float eps, float * y_out, float * y_out2) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const bool has3 = t + 2 * NCT < H;
    const float a = s.xres[t], b = s.xres[t + NCT], c = has3 ? s.xres[t + 2 * NCT] : 0.0f;
    double sq = (double) __fmul_rn(a, a) + (double) __fmul_rn(b, b) + (double) __fmul_rn(c, c);
    sq = warp_sum_d(sq);
    if (lane == 0) s.dred[warp] = sq;
    cbar();
    double ts[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int i = 0; i < NCW; ++i) ts[i & 3] += s.dred[i];
    const double tot = (ts[0] + ts[1]) + (ts[2] + ts[3]);
    const float mean  = (float) (tot * (1.0 / (double) H));       // H is a power of two: same value as tot / H
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    const float ya = __fmul_rn(__fmul_rn(a, scale), gm.g[0]);
    const float yb = __fmul_rn(__fmul_rn(b, scale), gm.g[1]);
    const float yc = has3 ? __fmul_rn(__fmul_rn(c, scale), gm.g[2]) : 0.0f;
    if (Q8) {
        // NCT is a multiple of 32: the lanes of warp w hold Q8_0 blocks w, w + NCW, w + 2 NCW of y in registers,
        // so the activation quantisation needs no trip through shared memory and no extra barrier
        static_assert(NCT % 32 == 0 && 3 * NCT >= H, "block ownership of the fused quantisation");
        const float v[3] = {ya, yb, yc};
        q8_blocks<3>(s, warp, H / 32, lane, v);
    } else {
        xh_put(s, t, ya); xh_put(s, t + NCT, yb);
        if (has3) xh_put(s, t + 2 * NCT, yc);
    }
    if (y_out)  { y_out[t] = ya;  y_out[t + NCT] = yb; if (has3) y_out[t + 2 * NCT] = yc; }
    if (y_out2) { y_out2[t] = ya; y_out2[t + NCT] = yb; if (has3) y_out2[t + 2 * NCT] = yc; }
    cbar();
}

// per-head RMSNorm(128)*gamma then NEOX RoPE (FMA-contracted like the reference build), one warp;
// lane l owns dims {l, l+32, l+64, l+96}; rope values c0,s0 (dim l) and c1,s1 (dim l+32) (oracle D.7)
__device__ __forceinline__ void head_norm_rope(float (&v)[4], const float (&gn)[4], const float (&rp)[4], float eps) {
    double sq = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) sq += (double) __fmul_rn(v[i], v[i]);
    sq = warp_sum_d(sq);
    const float mean  = (float) (sq * (1.0 / (double) HD));
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = __fmul_rn(__fmul_rn(v[i], scale), gn[i]);
    const float c0 = rp[0], s0 = rp[1], c1 = rp[2], s1 = rp[3];
    // the reference build (g++ -O3 -mfma) contracts the rotation: fmsub(x0,c,x1*s), fmadd(x0,s,x1*c)
    const float o0 = __fmaf_rn(v[0], c0, -__fmul_rn(v[2], s0));
    const float o2 = __fmaf_rn(v[0], s0,  __fmul_rn(v[2], c0));
    const float o1 = __fmaf_rn(v[1], c1, -__fmul_rn(v[3], s1));
    const float o3 = __fmaf_rn(v[1], s1,  __fmul_rn(v[3], c1));
    v[0] = o0; v[1] = o1; v[2] = o2; v[3] = o3;
}

// ------------------------------------------------------------------------------- attention (phase B)
// CTA h (< 16) owns query head h of every layer (kv head h/2 — GQA by broadcast, ggml-cpu.c:1160-1199;
// the even head of a pair appends K/V to the cache).  Every dot product keeps the oracle's order
// (D.4): s_t = dot_f16(K_t, r16(q)) * scale ; P = softmax ; o_d = dot_f16(V^T_d, r16(P)) over
// positions (32 accumulator lanes t%32 + double-precision leftovers).
// K cache rows are chain-major (element d = 32i + a at a*4 + i: one 8-byte load per lane), V rows natural.
struct AttnRegs {          // per-lane parameters fetched ahead of use (warps 0 and 1)
    float gn[4];           // q_norm (warp 0) / k_norm (warp 1) of the layer
    float rp[4];           // RoPE c0, s0, c1, s1 of the position
};

__device__ __forceinline__ uint2 ld_cg_u2(const void * p) { return __ldcg(reinterpret_cast<const uint2 *>(p)); }

// ---- reduce32_lanes() of N accumulators at once ("transposed" butterfly): at every level a lane keeps one
// half of its values and sends the other half to its partner, so the pair sums a[l] + a[l ^ mask] are formed
// once instead of twice and N values cost N - 1 + (5 - log2 N) shuffles instead of 5 N.  The pairing per
// level (xor 16, 8, 4, 1, 2) and therefore every rounding is the one of reduce32_lanes / GGML_F32x8_REDUCE.
// Returns the finished sum of accumulator `idx` (set on return); every accumulator ends in 32 / N lanes.
template <int N>
__device__ __forceinline__ float reduce32_transposed(float (&v)[N], int lane, int & idx) {
    constexpr int masks[5] = {16, 8, 4, 1, 2};
    idx = 0;
    int c = N;
#pragma unroll
    for (int lv = 0; lv < 5; ++lv) {
        const int mask = masks[lv];
        if (c > 1) {
            const int h = c >> 1;
            const bool up = (lane & mask) != 0;
#pragma unroll
            for (int k = 0; k < N / 2; ++k) {
                if (k < h) {
                    const float send = up ? v[k] : v[k + h];
                    const float keep = up ? v[k + h] : v[k];
                    v[k] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, mask));
                }
            }
            if (up) idx += h;
            c = h;
        } else {
            v[0] = __fadd_rn(v[0], __shfl_xor_sync(0xffffffffu, v[0], mask));
        }
    }
    return v[0];
}

// soft-max of a row of at most 32 entries by ONE warp (ggml_vec_soft_max_f32: 8 lanes at a time, libm tail;
// vec.cpp:548-617): lane i holds score i; returns exp(score - M) / sum for the lane (0 for lanes >= n_kv).
// n8 = chunks of the reference row width that take the SIMD exp, n8v = those holding unmasked positions.
__device__ __forceinline__ float softmax_row32(float score, float M, int lane, int n_kv, int n8, int n8v) {
    const bool live = lane < n_kv;
    const bool simd = lane < n8 * 8;
    const float x = live ? __fsub_rn(score, M) : 0.0f;
    // straight-line code (this warp is the critical path of the layer): the SIMD exp is evaluated by every
    // lane, the libm tail only when the row has one (a warp-uniform test); all sums below are selects,
    // adding +0.0 to a non-negative double is exact
    float e = live && simd ? v_expf(x) : 0.0f;
    if (n8 * 8 < n_kv) {
        const float et = tail_expf(x);
        e = live && !simd ? et : e;
    }
    // chunk sums in the oracle's order: ((e4+e0)+(e6+e2)) + ((e5+e1)+(e7+e3)) per 8 lanes
    float c = simd || !live ? e : 0.0f;              // masked lanes of a SIMD chunk add +0
    c = __fadd_rn(c, __shfl_xor_sync(0xffffffffu, c, 4));
    c = __fadd_rn(c, __shfl_xor_sync(0xffffffffu, c, 2));
    c = __fadd_rn(c, __shfl_xor_sync(0xffffffffu, c, 1));
    float cs[4], ev[7];
#pragma unroll
    for (int ch = 0; ch < 4; ++ch) cs[ch] = __shfl_sync(0xffffffffu, c, ch * 8);
#pragma unroll
    for (int i = 0; i < 7; ++i) ev[i] = __shfl_sync(0xffffffffu, e, (n8 * 8 + i) & 31);
    double S = 0.0;
#pragma unroll
    for (int ch = 0; ch < 4; ++ch) S += ch < n8v ? (double) cs[ch] : 0.0;
#pragma unroll
    for (int i = 0; i < 7; ++i) S += n8 * 8 + i < n_kv ? (double) ev[i] : 0.0;      // scalar tail, in order
    const float inv = (float) (1.0 / S);
    return __fmul_rn(e, inv);
}

// soft-max of a row longer than 32 entries by the whole CTA: s.sc[0..n_kv) scores -> s.ph[0..n_kv) probabilities
// rounded to F16 (ggml_vec_soft_max_f32: 8 lanes at a time, libm tail; vec.cpp:548-617).  One cbar() inside; the
// caller adds the one that publishes s.ph.
template <class SM>
__device__ __forceinline__ void softmax_long(SM & s, float M, int n_kv, int n8, int n8v) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    double ds = 0.0;
    for (int ch = t; ch < n8v; ch += NCT) {
        float e[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int tt = ch * 8 + j;
            e[j] = tt < n_kv ? v_expf(__fsub_rn(s.sc[tt], M)) : 0.0f;     // masked: exp(-inf) = +0
        }
        ds += (double) __fadd_rn(__fadd_rn(__fadd_rn(e[4], e[0]), __fadd_rn(e[6], e[2])),
                                 __fadd_rn(__fadd_rn(e[5], e[1]), __fadd_rn(e[7], e[3])));
#pragma unroll
        for (int j = 0; j < 8; ++j) if (ch * 8 + j < n_kv) s.sc[ch * 8 + j] = e[j];
    }
    if (t < 8 && n8 * 8 + t < n_kv) {                        // scalar tail: libm expf
        const int tt = n8 * 8 + t;
        const float e = tail_expf(__fsub_rn(s.sc[tt], M));
        ds += (double) e;
        s.sc[tt] = e;
    }
    ds = warp_sum_d(ds);
    if (lane == 0) s.dred[warp] = ds;
    cbar();
    double S = 0.0;
#pragma unroll
    for (int i = 0; i < NCW; ++i) S += s.dred[i];
    const float inv = (float) (1.0 / S);
    for (int tt = t; tt < n_kv; tt += NCT) s.ph[tt] = __float2half_rn(__fmul_rn(s.sc[tt], inv));
}

// address of history row r (0..KV_CH) of a K/V slice that starts at ring slot s0
template <class SM>
__device__ __forceinline__ const unsigned char * kv_row(const SM & s, int s0, int r) {
    int slot = s0 + (r >> 3);
    if (slot >= SM::NSLOTS) slot -= SM::NSLOTS;
    return s.ring + (size_t) slot * SLOT_B + (r & 7) * (HD * 2);
}

// The K/V rows a layer appends (plain stores by warps 1 and 2 of the even head's CTA) are read by OTHER CTAs -- the odd
// head of the pair, the members of the wide path -- from the next pass on, through L2 (ld.cg) or the TMA unit (async
// proxy): they are published right after the attention output has left, i.e. before this CTA's next packets, which every
// reader polls before it touches the history.  (Publishing once per pass instead was measured: no difference.)
__device__ __forceinline__ void publish_kv(bool writer_cta) {
    const int warp = threadIdx.x >> 5;
    if (writer_cta && (warp == 1 || warp == 2)) { asm volatile("fence.proxy.async.global;" ::: "memory"); __threadfence(); }
}

template <class SM>
__device__ void attention_head(const DevModel & m, const StackW & W, int l, int h, int pos, bool f32attn, int width,
                               const AttnRegs & ar, unsigned ep_in, unsigned ep_out, SM & s, Ring & rg, Watch & wd_) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int g = h >> 1;
    const bool writer = (h & 1) == 0;
    const int n_kv = pos + 1;
    // row width of the reference graph: a prefill scores all n_past + n_tokens cache rows and masks
    // the future ones with -inf (src/tts_transformer.cpp:1205-1227); masked entries add exact zeros,
    // but the width decides which positions take the 8-wide SIMD exp vs the libm tail and which
    // take the 32-lane part of the P.V dot vs its double-precision leftovers.
    const int wd = width > n_kv ? width : n_kv;
    const float kq_scale = 1.0f / sqrtf((float) HD);
    const size_t kv_head_base = ((size_t) (l * NKV + g) * W.n_ctx) * HD;
    Pkt * att = m.x_att + h * HD;
    const int np = min(wd & ~31, n_kv);                      // positions in the 32-lane part of the P.V dot
    // rows of at most ATT_FAST_KV entries (every code-predictor step, the first talker frames): the K/V history is a
    // few KB, so it is fetched straight from L2 into registers while q/k/v are still in flight instead of being
    // streamed through the ring: no slice waits, no ring hand-over, two CTA barriers (same arithmetic, same order)
    const bool fast = !f32attn && wd <= ATT_FAST_KV;
    const int nck = (f32attn || fast) ? 0 : (pos + KV_CH - 1) / KV_CH; // history slices (K, then V) the producer streams

    PROF(50);
    // the first K slice was requested long ago: take its mbarrier wait now, while q/k/v are still in flight
    const int s0_first = nck > 0 ? ring_acquire(m, s, rg, wd_) : 0;
    // fast path: history rows warp, warp + 11 of K (this lane's chain: 4 halves) and, in warps 4..7, column
    // d = t - 128 of V for all history rows (packed two positions per register); f32attn: row 0 of the F32 shadow
    uint2 kpre[2] = {make_uint2(0u, 0u), make_uint2(0u, 0u)};
    unsigned vpre[ATT_FAST_KV / 2];
    float kf0[4] = {0.f, 0.f, 0.f, 0.f}, vf0[4] = {0.f, 0.f, 0.f, 0.f};
    if (fast) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int r = warp + j * NCW;
            if (r < pos) kpre[j] = __ldcg(reinterpret_cast<const uint2 *>(W.kcache + kv_head_base + (size_t) r * HD) + lane);
        }
        if (warp >= 4 && warp < 8) {
            const unsigned short * vcol = reinterpret_cast<const unsigned short *>(W.vcache + kv_head_base) + (t - 128);
#pragma unroll
            for (int j = 0; j < ATT_FAST_KV / 2; ++j) {
                const unsigned lo = 2 * j     < pos ? (unsigned) __ldcg(vcol + (size_t) (2 * j) * HD)     : 0u;
                const unsigned hi = 2 * j + 1 < pos ? (unsigned) __ldcg(vcol + (size_t) (2 * j + 1) * HD) : 0u;
                vpre[j] = lo | (hi << 16);
            }
        }
    } else if (f32attn && pos == 1 && warp == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            kf0[i] = ld_act(W.kf32 + ((size_t) (l * 2) * NKV + g) * HD + lane + 32 * i);
            vf0[i] = ld_act(W.vf32 + ((size_t) (l * 2) * NKV + g) * HD + lane + 32 * i);
        }
    }
    // ---- prologue: q, k, v of the new token (warps 0, 1, 2)
    if (warp < 3) {
        const Pkt * src = m.x_qkv + (warp == 0 ? h * HD : warp == 1 ? NH * HD + g * HD : (NH + NKV) * HD + g * HD);
        float v[4];
        pk_poll_n<4>(m, src + lane, 32, 4, ep_in, v, wd_);
        if (warp == 0) {
            head_norm_rope(v, ar.gn, ar.rp, m.eps);
#pragma unroll
            for (int i = 0; i < 4; ++i) { s.qf[lane + 32 * i] = v[i]; s.qr[lane + 32 * i] = r16(v[i]); }
        } else if (warp == 1) {
            head_norm_rope(v, ar.gn, ar.rp, m.eps);
            __half hv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { hv[i] = __float2half_rn(v[i]); s.knew_f[lane + 32 * i] = v[i]; s.knew_h[lane + 32 * i] = hv[i]; }
            if (writer) {
                uint2 pk;
                pk.x = (unsigned) __half_as_ushort(hv[0]) | ((unsigned) __half_as_ushort(hv[1]) << 16);
                pk.y = (unsigned) __half_as_ushort(hv[2]) | ((unsigned) __half_as_ushort(hv[3]) << 16);
                *reinterpret_cast<uint2 *>(W.kcache + kv_head_base + (size_t) pos * HD + lane * 4) = pk;
                if (W.kf32 && pos < 2) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) W.kf32[((size_t) (l * 2 + pos) * NKV + g) * HD + lane + 32 * i] = v[i];
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int d = lane + 32 * i;
                const __half hv = __float2half_rn(v[i]);
                s.vnew_f[d] = v[i]; s.vnew_h[d] = hv;
                if (writer) {
                    W.vcache[kv_head_base + (size_t) pos * HD + d] = hv;
                    if (W.vslice) W.vslice[(((size_t) (l * NKV + g) * WIDE_G + (d >> 4)) * W.n_ctx + pos) * VS_D + (d & 15)] = hv;
                    if (W.vf32 && pos < 2) W.vf32[((size_t) (l * 2 + pos) * NKV + g) * HD + d] = v[i];
                }
            }
        }
    }
    PROF(51);
    cbar();
    PROF(52);

    if (f32attn) {
        // code-predictor prefill numerics: the reference graph multiplies the fresh F32 Kcur/Vcur/Q
        // and does not round P (src/tts_transformer.cpp:1624-1635): ggml_vec_dot_f32 order.
        if (warp == 0) {
            float sc_t[2];
            float mx = -CUDART_INF_F;
            const int n = pos + 1 < 2 ? pos + 1 : 2;
            for (int tt = 0; tt < n; ++tt) {
                float acc = 0.0f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float kv = (tt == pos) ? s.knew_f[lane + 32 * i] : kf0[i];       // tt < pos only for pos == 1: row 0
                    acc = __fmaf_rn(kv, s.qf[lane + 32 * i], acc);
                }
                sc_t[tt] = __fmul_rn(reduce32_lanes(acc), kq_scale);
                mx = fmaxf(mx, sc_t[tt]);
            }
            double sum = 0.0;
            for (int tt = 0; tt < n; ++tt) { sc_t[tt] = tail_expf(__fsub_rn(sc_t[tt], mx)); sum += (double) sc_t[tt]; }
            const float inv = (float) (1.0 / sum);
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            for (int tt = 0; tt < n; ++tt) {
                const float p = __fmul_rn(sc_t[tt], inv);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float vv = (tt == pos) ? s.vnew_f[lane + 32 * i] : vf0[i];
                    o[i] = __fmaf_rn(vv, p, o[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) pk_store(att + lane + 32 * i, o[i], ep_out);
        }
        return;                                              // (the appended K/V rows are published once per pass: publish_kv)
    }

    if (fast) {
        // ---- scores from registers: warp w has rows w and w + 11, the warp the next row would go to takes the new token
        float q[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = s.qr[lane + 32 * i];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int r = warp + j * NCW;
            if (r < pos) {
                const float2 f01 = __half22float2(*reinterpret_cast<const __half2 *>(&kpre[j].x));
                const float2 f23 = __half22float2(*reinterpret_cast<const __half2 *>(&kpre[j].y));
                float a = __fmaf_rn(f01.x, q[0], 0.0f);
                a = __fmaf_rn(f01.y, q[1], a); a = __fmaf_rn(f23.x, q[2], a); a = __fmaf_rn(f23.y, q[3], a);
                const float sv = __fmul_rn(reduce32_lanes(a), kq_scale);
                if (lane == 0) s.sc[r] = sv;
            }
        }
        if (warp == pos % NCW) {
            float a = __fmaf_rn(__half2float(s.knew_h[lane]), q[0], 0.0f);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 32]), q[1], a);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 64]), q[2], a);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 96]), q[3], a);
            const float sv = __fmul_rn(reduce32_lanes(a), kq_scale);
            if (lane == 0) s.sc[pos] = sv;
        }
        PROF(72);
        cbar();
        PROF(53);
        // ---- soft-max + P.V in warps 4..7 (each evaluates the 16-entry soft-max itself: no barrier in between);
        //      wd < 32, so every position is a double-precision leftover of ggml_vec_dot_f16 (vec.cpp:381-384)
        if (warp >= 4 && warp < 8) {
            const float sv = lane < n_kv ? s.sc[lane] : -CUDART_INF_F;
            float M = sv;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
            const float pv = softmax_row32(sv, M, lane, n_kv, wd >> 3, min(wd >> 3, (n_kv + 7) >> 3));
            const unsigned ph = (unsigned) __half_as_ushort(__float2half_rn(pv));
            PROF(54);
            const int d = t - 128;
            const float vnew = __half2float(s.vnew_h[d]);
            double sumf = 0.0;
#pragma unroll
            for (int tt = 0; tt < ATT_FAST_KV; ++tt) {
                const float p  = __half2float(__ushort_as_half((unsigned short) __shfl_sync(0xffffffffu, ph, tt)));
                const unsigned vw = vpre[tt >> 1];
                const float vh = __half2float(__ushort_as_half((unsigned short) ((tt & 1) ? (vw >> 16) : (vw & 0xffffu))));
                const float pr = __fmul_rn(tt == pos ? vnew : vh, p);
                sumf += tt < n_kv ? (double) pr : 0.0;
            }
            PROF(58);
            pk_store(att + d, (float) sumf, ep_out);
        }
        PROF(56);
        return;
    }

    // ---- scores: the K history arrives in the ring (chain-major rows: element d = 32i + a at a*4 + i);
    //      warp w takes rows w, w+11, ... of a slice; lane a walks chain a (4 steps) of dot_f16(K_t, q)
    {
        float q[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = s.qr[lane + 32 * i];
        float mx = -CUDART_INF_F;
        // the new token's row comes from shared memory; it goes to the warp that the next history row would go to
        // and is done first, so the whole score phase needs one CTA barrier after its last slice
        const int n_last = nck > 0 ? pos - (nck - 1) * KV_CH : 0;
        if (warp == n_last % NCW) {
            float a = __fmaf_rn(__half2float(s.knew_h[lane]), q[0], 0.0f);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 32]), q[1], a);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 64]), q[2], a);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 96]), q[3], a);
            const float sv = __fmul_rn(reduce32_lanes(a), kq_scale);
            mx = fmaxf(mx, sv);
            if (lane == 0) s.sc[pos] = sv;
        }
        for (int c = 0; c < nck; ++c) {
            const int s0 = c == 0 ? s0_first : ring_acquire(m, s, rg, wd_);
            PROF(59);
            const int rows = min(KV_CH, pos - c * KV_CH);
            // four rows per warp step: the 32-lane folds of the four dot products share their shuffles
            // (reduce32_transposed: 6 shuffles instead of 20, same pairing and rounding per value)
            for (int r0 = warp * 4; r0 < rows; r0 += NCW * 4) {
                float a[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = min(r0 + i, rows - 1);
                    const uint2 kk = *reinterpret_cast<const uint2 *>(kv_row(s, s0, r) + lane * 8);
                    const float2 f01 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.x));
                    const float2 f23 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.y));
                    a[i] = __fmaf_rn(f01.x, q[0], 0.0f);
                    a[i] = __fmaf_rn(f01.y, q[1], a[i]); a[i] = __fmaf_rn(f23.x, q[2], a[i]); a[i] = __fmaf_rn(f23.y, q[3], a[i]);
                }
                int idx;
                const float sv = __fmul_rn(reduce32_transposed<4>(a, lane, idx), kq_scale);
                if (r0 + idx < rows) {
                    mx = fmaxf(mx, sv);
                    s.sc[c * KV_CH + r0 + idx] = sv;         // the 8 lanes that hold row idx store the same value
                }
            }
            if (c + 1 == nck) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                if (lane == 0) s.fred[warp] = mx;
            }
            PROF(72);
            cbar();
            ring_release(s, rg, (rows + 7) >> 3);
            PROF(73);
        }
        if (nck == 0) { if (lane == 0) s.fred[warp] = mx; cbar(); }
    }
    PROF(53);
    float M = s.fred[0];
#pragma unroll
    for (int i = 1; i < NCW; ++i) M = fmaxf(M, s.fred[i]);

    // ---- soft-max over the row (ggml_vec_soft_max_f32: 8 lanes at a time, libm tail; vec.cpp:548-617)
    const int n8  = wd >> 3;                                 // chunks that take the SIMD exp
    const int n8v = min(n8, (n_kv + 7) >> 3);                // ... of which these hold unmasked positions
    if (wd <= 32) {
        // short rows (the code predictor, early talker steps): one warp, no block-level reductions
        if (warp == 0) {
            const float pv = softmax_row32(lane < n_kv ? s.sc[lane] : 0.0f, M, lane, n_kv, n8, n8v);
            if (lane < n_kv) s.ph[lane] = __float2half_rn(pv);
        }
    } else {
        softmax_long(s, M, n_kv, n8, n8v);
    }
    cbar();
    PROF(54);

    // ---- P.V (dot_f16(V^T_d, P) over positions): accumulator lane a = position % 32.  Warp w < 8 owns
    //      accumulator lanes 4w..4w+3, its lane owns dims 4*lane..4*lane+3, so a row read is 256
    //      contiguous bytes of the ring.  The 32 partial sums of a dim are folded in the oracle's order
    //      through shared memory (the score buffer is free again); positions >= np are the
    //      double-precision leftovers of ggml_vec_dot_f16 (vec.cpp:381-384), one thread per dim.
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[i][k] = 0.0f;
    int s0_last = 0, rows_last = 0;
    for (int c = 0; c < nck; ++c) {
        const int s0 = ring_acquire(m, s, rg, wd_);
        const int rows = min(KV_CH, pos - c * KV_CH);
        if (warp < 8 && c * KV_CH < np) {                    // short rows (np == 0) have no 32-lane part at all
#pragma unroll
            for (int gq = 0; gq < KV_CH / 32; ++gq) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = gq * 32 + 4 * warp + i;
                    const int tt = c * KV_CH + r;
                    if (r < rows && tt < np) {
                        const uint2 vv = *reinterpret_cast<const uint2 *>(kv_row(s, s0, r) + lane * 8);
                        const unsigned pu = (unsigned) __half_as_ushort(s.ph[tt]);
                        const unsigned p2 = pu | (pu << 16);
                        acc[i][0] = fhfma_lo(vv.x, p2, acc[i][0]); acc[i][1] = fhfma_hi(vv.x, p2, acc[i][1]);
                        acc[i][2] = fhfma_lo(vv.y, p2, acc[i][2]); acc[i][3] = fhfma_hi(vv.y, p2, acc[i][3]);
                    }
                }
            }
        }
        if (c + 1 < nck) { cbar(); ring_release(s, rg, (rows + 7) >> 3); }
        else { s0_last = s0; rows_last = rows; }             // the last slice also holds the leftover rows
    }
    PROF(55);
    if (np > 0) {
        if (warp < 8) {
            if (pos < np && ((pos & 31) >> 2) == warp) {     // the new token is the last step of its chain
                const int i = pos & 3;
                const uint2 vv = *reinterpret_cast<const uint2 *>(s.vnew_h + lane * 4);
                const unsigned pu = (unsigned) __half_as_ushort(s.ph[pos]);
                const unsigned p2 = pu | (pu << 16);
#pragma unroll
                for (int ii = 0; ii < 4; ++ii)
                    if (ii == i) {
                        acc[ii][0] = fhfma_lo(vv.x, p2, acc[ii][0]); acc[ii][1] = fhfma_hi(vv.x, p2, acc[ii][1]);
                        acc[ii][2] = fhfma_lo(vv.y, p2, acc[ii][2]); acc[ii][3] = fhfma_hi(vv.y, p2, acc[ii][3]);
                    }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
                *reinterpret_cast<float4 *>(s.sc + (4 * warp + i) * HD + 4 * lane) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
        cbar();
    }
    if (t < HD) {
        double sumf = 0.0;
        if (np > 0) {
            float lq[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                lq[i] = __fadd_rn(__fadd_rn(s.sc[i * HD + t], s.sc[(16 + i) * HD + t]), __fadd_rn(s.sc[(8 + i) * HD + t], s.sc[(24 + i) * HD + t]));
            const float t0 = __fadd_rn(lq[0], lq[4]), t1 = __fadd_rn(lq[1], lq[5]), t2 = __fadd_rn(lq[2], lq[6]), t3 = __fadd_rn(lq[3], lq[7]);
            sumf = (double) __fadd_rn(__fadd_rn(t0, t1), __fadd_rn(t2, t3));
        }
        const int c_last = nck - 1;
        PROF(57);
        // four positions per round, branch-free: the loads and products are independent, only the double additions
        // keep the reference's order.  Positions past the row read a clamped index and add +0.0 (sumf is never
        // -0.0, so that is an exact no-op); the new token's row is selected by pointer, not by a branch.
        for (int tt = np; tt < n_kv; tt += 4) {
            double pd[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int u = tt + j, uc = min(u, n_kv - 1);
                const __half * hist = reinterpret_cast<const __half *>(kv_row(s, s0_last, uc - c_last * KV_CH)) + t;
                const __half * src = uc == pos ? s.vnew_h + t : hist;
                const float pr = __fmul_rn(__half2float(*src), __half2float(s.ph[uc]));
                pd[j] = u < n_kv ? (double) pr : 0.0;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) sumf += pd[j];
        }
        PROF(58);
        pk_store(att + t, (float) sumf, ep_out);
    }
    PROF(56);
    if (nck > 0) { cbar(); ring_release(s, rg, (rows_last + 7) >> 3); }
}

// --------------------------------------------------------------------------- wide attention (long rows)
// One CTA per head streams the whole K and V history of the head through ONE SM (2 x n_kv x 256 bytes per layer:
// 1 MB at 2 000 positions) while 132 CTAs wait for its output.  Rows longer than DevModel::wide_min are dealt to
// WIDE_G = 8 CTAs per head instead (CTA b: head b % 16, member b / 16; member 0 is the classic attention CTA):
//   * scores are independent per position: member j scores its share of the K rows and publishes them as packets
//     (x_sc, one more all-to-all inside the group), every member then holds the full row and evaluates the same
//     soft-max (same code, same order: bit-identical in all members and with the one-CTA path);
//   * P.V keeps the oracle's chain order per output DIM (32 accumulator lanes t % 32 over all positions, fixed fold,
//     double-precision leftovers), so it is dealt by dims: member j owns dims 16j..16j+15 and reads the 32-byte
//     pieces of all V rows -- contiguous in the second V layout StackW::vslice that the cache writer keeps.
// Per member and layer that is n_kv x (256/8 + 32) bytes through the ring instead of n_kv x 512.
__device__ __forceinline__ bool wide_pass(const DevModel & m, const StackW & W, int f32attn, int width, int pos) {
    return W.vslice != nullptr && m.wide_min > 0 && !f32attn && width == 0 && pos + 1 > m.wide_min &&
           (int) gridDim.x >= NH * WIDE_G;
}
// K rows of member j: the history rows 0..pos-1 dealt in 8-row ring slots
__device__ __forceinline__ void wide_k_range(int pos, int j, int & r0, int & r1) {
    const int S = (pos + 7) >> 3;
    r0 = min(pos, ((S * j) / WIDE_G) * 8);
    r1 = min(pos, ((S * (j + 1)) / WIDE_G) * 8);
}
// address of byte `off` of a slice that starts at ring slot s0
template <class SM>
__device__ __forceinline__ const unsigned char * slice_byte(const SM & s, int s0, int off) {
    int slot = s0 + (off >> 11);
    if (slot >= SM::NSLOTS) slot -= SM::NSLOTS;
    return s.ring + (size_t) slot * SLOT_B + (off & (SLOT_B - 1));
}

template <class SM>
__device__ __forceinline__ void attention_wide(const DevModel & m, const StackW & W, int l, int pos,
                                            const AttnRegs & ar, unsigned ep_in, unsigned ep_out, SM & s, Ring & rg, Watch & wd_) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int h = (int) blockIdx.x % NH, j = (int) blockIdx.x / NH;
    const int g = h >> 1;
    const bool leader = j == 0, writer = leader && (h & 1) == 0;
    const int n_kv = pos + 1;
    const float kq_scale = 1.0f / sqrtf((float) HD);
    const size_t kv_head_base = ((size_t) (l * NKV + g) * W.n_ctx) * HD;
    Pkt * att = m.x_att + h * HD;
    Pkt * scp = m.x_sc + (size_t) h * ATT_MAX_KV;
    const int np = n_kv & ~31;                               // positions in the 32-lane part of the P.V dot
    int kr0, kr1;
    wide_k_range(pos, j, kr0, kr1);
    const int nck = (kr1 - kr0 + KV_CH - 1) / KV_CH;         // K slices of this member
    const int nvs = (pos + VS_CH - 1) / VS_CH;               // V slices (all positions, this member's dims)

    PROF(50);
    const int s0_first = nck > 0 ? ring_acquire(m, s, rg, wd_) : 0;
    // ---- prologue: q (every member), k (leader), v (every member: the new token is a step of the P.V chains)
    if (warp < 3 && (warp != 1 || leader)) {
        const Pkt * src = m.x_qkv + (warp == 0 ? h * HD : warp == 1 ? NH * HD + g * HD : (NH + NKV) * HD + g * HD);
        float v[4];
        pk_poll_n<4>(m, src + lane, 32, 4, ep_in, v, wd_);
        if (warp == 0) {
            head_norm_rope(v, ar.gn, ar.rp, m.eps);
#pragma unroll
            for (int i = 0; i < 4; ++i) s.qr[lane + 32 * i] = r16(v[i]);
        } else if (warp == 1) {
            head_norm_rope(v, ar.gn, ar.rp, m.eps);
            __half hv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { hv[i] = __float2half_rn(v[i]); s.knew_h[lane + 32 * i] = hv[i]; }
            if (writer) {
                uint2 pk;
                pk.x = (unsigned) __half_as_ushort(hv[0]) | ((unsigned) __half_as_ushort(hv[1]) << 16);
                pk.y = (unsigned) __half_as_ushort(hv[2]) | ((unsigned) __half_as_ushort(hv[3]) << 16);
                *reinterpret_cast<uint2 *>(W.kcache + kv_head_base + (size_t) pos * HD + lane * 4) = pk;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int d = lane + 32 * i;
                const __half hv = __float2half_rn(v[i]);
                s.vnew_h[d] = hv;
                if (writer) {
                    W.vcache[kv_head_base + (size_t) pos * HD + d] = hv;
                    W.vslice[(((size_t) (l * NKV + g) * WIDE_G + (d >> 4)) * W.n_ctx + pos) * VS_D + (d & 15)] = hv;
                }
            }
        }
    }
    PROF(51);
    cbar();
    PROF(52);

    // ---- scores of this member's K rows -> packets (K rows are chain-major: one 8-byte load per lane and row)
    {
        float q[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = s.qr[lane + 32 * i];
        if (leader && warp == NCW - 1) {                     // the new token's own score (+ the pad of an odd row)
            float a = __fmaf_rn(__half2float(s.knew_h[lane]), q[0], 0.0f);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 32]), q[1], a);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 64]), q[2], a);
            a = __fmaf_rn(__half2float(s.knew_h[lane + 96]), q[3], a);
            const float sv = __fmul_rn(reduce32_lanes(a), kq_scale);
            if (lane == 0) pk_store(scp + pos, sv, ep_in);
            if (lane == 1 && (n_kv & 1)) pk_store(scp + n_kv, 0.0f, ep_in);
        }
        for (int c = 0; c < nck; ++c) {
            const int s0 = c == 0 ? s0_first : ring_acquire(m, s, rg, wd_);
            PROF(59);
            const int rows = min(KV_CH, kr1 - kr0 - c * KV_CH);
            for (int r0 = warp * 4; r0 < rows; r0 += NCW * 4) {
                float a[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = min(r0 + i, rows - 1);
                    const uint2 kk = *reinterpret_cast<const uint2 *>(kv_row(s, s0, r) + lane * 8);
                    const float2 f01 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.x));
                    const float2 f23 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.y));
                    a[i] = __fmaf_rn(f01.x, q[0], 0.0f);
                    a[i] = __fmaf_rn(f01.y, q[1], a[i]); a[i] = __fmaf_rn(f23.x, q[2], a[i]); a[i] = __fmaf_rn(f23.y, q[3], a[i]);
                }
                int idx;
                const float sv = __fmul_rn(reduce32_transposed<4>(a, lane, idx), kq_scale);
                if ((lane & 7) == 0 && r0 + idx < rows) pk_store(scp + kr0 + c * KV_CH + r0 + idx, sv, ep_in);
            }
            PROF(72);
            cbar();
            ring_release(s, rg, (rows + 7) >> 3);
            PROF(73);
        }
    }
    // ---- the whole row from all members
    float mx = -CUDART_INF_F;
    {
        const int n_pairs = (n_kv + 1) >> 1;
        for (int base = 0; base < n_pairs; base += 4 * NCT) {
            float2 f[4];
            const int rem = n_pairs - base - t;
            const int mine = rem > 0 ? min(4, (rem - 1) / NCT + 1) : 0;
            pk_poll2_n<4>(m, scp + 2 * (base + t), NCT, mine, ep_in, f, wd_);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (i < mine) {
                    const int tt = 2 * (base + t + i * NCT);
                    s.sc[tt] = f[i].x; mx = fmaxf(mx, f[i].x);
                    if (tt + 1 < n_kv) { s.sc[tt + 1] = f[i].y; mx = fmaxf(mx, f[i].y); }
                }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) s.fred[warp] = mx;
    }
    cbar();
    PROF(53);
    float M = s.fred[0];
#pragma unroll
    for (int i = 1; i < NCW; ++i) M = fmaxf(M, s.fred[i]);
    softmax_long(s, M, n_kv, n_kv >> 3, n_kv >> 3);
    cbar();
    PROF(54);

    // ---- P.V for dims 16j..16j+15: warp c (0, 1) owns dims 16j + 8c .. + 7, its lane a is accumulator lane a
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.0f;
    int s0_last = 0, rows_last = 0;
    const int n_chain = min(np, pos);                        // history positions of the 32-lane part
    for (int cs = 0; cs < nvs; ++cs) {
        const int s0 = ring_acquire(m, s, rg, wd_);
        const int rows = min(VS_CH, pos - cs * VS_CH);
        if (warp < 2) {
#pragma unroll
            for (int gq = 0; gq < VS_CH / 32; ++gq) {
                const int r = gq * 32 + lane;
                const int tt = cs * VS_CH + r;
                if (tt < n_chain) {
                    const uint4 vv = *reinterpret_cast<const uint4 *>(slice_byte(s, s0, r * (VS_D * 2) + warp * 16));
                    const unsigned pu = (unsigned) __half_as_ushort(s.ph[tt]);
                    const unsigned p2 = pu | (pu << 16);
                    acc[0] = fhfma_lo(vv.x, p2, acc[0]); acc[1] = fhfma_hi(vv.x, p2, acc[1]);
                    acc[2] = fhfma_lo(vv.y, p2, acc[2]); acc[3] = fhfma_hi(vv.y, p2, acc[3]);
                    acc[4] = fhfma_lo(vv.z, p2, acc[4]); acc[5] = fhfma_hi(vv.z, p2, acc[5]);
                    acc[6] = fhfma_lo(vv.w, p2, acc[6]); acc[7] = fhfma_hi(vv.w, p2, acc[7]);
                }
            }
        }
        if (cs + 1 < nvs) { cbar(); ring_release(s, rg, (rows * (VS_D * 2) + SLOT_B - 1) / SLOT_B); }
        else { s0_last = s0; rows_last = rows; }             // the last slice also holds the leftover rows
    }
    PROF(55);
    if (np > 0) {
        if (warp < 2) {
            if (pos < np && lane == (pos & 31)) {            // the new token is the last step of its chain
                const uint4 vv = *reinterpret_cast<const uint4 *>(s.vnew_h + j * VS_D + warp * 8);
                const unsigned pu = (unsigned) __half_as_ushort(s.ph[pos]);
                const unsigned p2 = pu | (pu << 16);
                acc[0] = fhfma_lo(vv.x, p2, acc[0]); acc[1] = fhfma_hi(vv.x, p2, acc[1]);
                acc[2] = fhfma_lo(vv.y, p2, acc[2]); acc[3] = fhfma_hi(vv.y, p2, acc[3]);
                acc[4] = fhfma_lo(vv.z, p2, acc[4]); acc[5] = fhfma_hi(vv.z, p2, acc[5]);
                acc[6] = fhfma_lo(vv.w, p2, acc[6]); acc[7] = fhfma_hi(vv.w, p2, acc[7]);
            }
            // chain a of dim d -> s.sc[a * 16 + d] (the score buffer is free again)
            *reinterpret_cast<float4 *>(s.sc + lane * VS_D + warp * 8)     = make_float4(acc[0], acc[1], acc[2], acc[3]);
            *reinterpret_cast<float4 *>(s.sc + lane * VS_D + warp * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
        }
        cbar();
    }
    if (t < VS_D) {
        double sumf = 0.0;
        if (np > 0) {
            float lq[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                lq[i] = __fadd_rn(__fadd_rn(s.sc[i * VS_D + t], s.sc[(16 + i) * VS_D + t]), __fadd_rn(s.sc[(8 + i) * VS_D + t], s.sc[(24 + i) * VS_D + t]));
            const float t0 = __fadd_rn(lq[0], lq[4]), t1 = __fadd_rn(lq[1], lq[5]), t2 = __fadd_rn(lq[2], lq[6]), t3 = __fadd_rn(lq[3], lq[7]);
            sumf = (double) __fadd_rn(__fadd_rn(t0, t1), __fadd_rn(t2, t3));
        }
        PROF(57);
        // leftovers np..n_kv-1 in double, in order (vec.cpp:381-384); the history ones sit in the last V slice
        const int row_base = (nvs - 1) * VS_CH;
        for (int tt = np; tt < n_kv; tt += 4) {
            double pd[4];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const int u = tt + jj, uc = min(u, n_kv - 1);
                const int rr = min(max(uc - row_base, 0), rows_last - 1);
                const __half * hist = reinterpret_cast<const __half *>(slice_byte(s, s0_last, rr * (VS_D * 2))) + t;
                const __half * src = uc == pos ? s.vnew_h + j * VS_D + t : hist;
                const float pr = __fmul_rn(__half2float(*src), __half2float(s.ph[uc]));
                pd[jj] = u < n_kv ? (double) pr : 0.0;
            }
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) sumf += pd[jj];
        }
        PROF(58);
        pk_store(att + j * VS_D + t, (float) sumf, ep_out);
    }
    PROF(56);
    if (nvs > 0) { cbar(); ring_release(s, rg, (rows_last * (VS_D * 2) + SLOT_B - 1) / SLOT_B); }
}

// --------------------------------------------------------------------------- token selection
// Reference: host argmax (src/tts_transformer.cpp:2182-2192: first maximum wins) and the
// sample_or_argmax lambda / cb0 post-processing (:2349-2387, :2664-2722).  Every CTA evaluates
// the selection redundantly on the same packets, so no extra exchange is needed.
__device__ __forceinline__ float proc_logit(const DevModel & m, float raw, int i, bool process,
                                            int suppress_start, int eos_id, float rep_penalty) {
    if (!process) return raw;
    if (i >= suppress_start && i != eos_id) return -CUDART_INF_F;
    if (rep_penalty != 1.0f && __ldcg(m.seen + i)) return raw > 0.0f ? raw / rep_penalty : raw * rep_penalty;
    return raw;
}

__device__ __noinline__ int select_greedy(const DevModel & m, Smem & s, unsigned ep, Watch & w) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    float v = -CUDART_INF_F; int idx = 0x7fffffff;
    if (t < (int) gridDim.x) {
        const float2 c = pk_poll2(m, m.x_cand + 2 * t, ep, w);
        v = c.x; idx = __float_as_int(c.y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int   oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    if (lane == 0) { s.fred[warp] = v; s.ired[warp] = idx; }
    cbar();
    v = s.fred[0]; idx = s.ired[0];
#pragma unroll
    for (int i = 1; i < NCW; ++i) {
        const float ov = s.fred[i]; const int oi = s.ired[i];
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    cbar();
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
// scratch of the sampler in either kernel's shared memory: 256 histogram bins + 2 control words
__device__ __forceinline__ int * sample_hist(Smem & s) { return reinterpret_cast<int *>(s.vt); }
template <class SM>
__device__ __noinline__ int select_sample(const DevModel & m, SM & s, unsigned ep, Watch & w, int vocab, bool process, int suppress_start,
                             int eos_id, float rep_penalty, float temperature, int top_k,
                             unsigned long long seed, unsigned long long draw) {
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    constexpr int PER = (3072 + NCT - 1) / NCT;  // NCT * PER >= vocab
    float l[PER];
    {
        const int have = min(max(vocab - t * PER, 0), PER);
        pk_poll_n<PER>(m, m.x_logit + t * PER, 1, have, ep, l, w);
    }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int idx = t * PER + i;
        l[i] = idx < vocab ? proc_logit(m, l[i], idx, process, suppress_start, eos_id, rep_penalty) / temperature : -CUDART_INF_F;
    }
    int * hist = sample_hist(s);
    if (top_k > 0 && top_k < vocab) {
        unsigned int prefix = 0, mask = 0;
        int k_rem = top_k;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            if (t < 256) hist[t] = 0;
            cbar();
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                if (t * PER + i < vocab) {
                    const unsigned int key = f2key(l[i]);
                    if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1);
                }
            }
            cbar();
            if (warp == 0) {
                // highest bin d whose count, added to the bins above it, reaches k_rem (a serial scan of the 256
                // bins by one thread cost ~7 k cycles per pass): lane i owns bins 8i..8i+7, suffix sums by shuffles
                int c[8], lsum = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) { c[j] = hist[lane * 8 + j]; lsum += c[j]; }
                int inc = lsum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_down_sync(0xffffffffu, inc, o); if (lane + o < 32) inc += v; }
                const int above = inc - lsum;                 // entries in the bins above this lane's
                if (above < k_rem && inc >= k_rem) {          // exactly one lane (the group holds >= k_rem entries)
                    int acc = above, d = lane * 8;
                    bool found = false;
#pragma unroll
                    for (int j = 7; j >= 0; --j) {
                        if (!found) { if (acc + c[j] >= k_rem) { d = lane * 8 + j; found = true; } else acc += c[j]; }
                    }
                    hist[256] = d; hist[257] = k_rem - acc;
                }
            }
            cbar();
            prefix |= ((unsigned int) hist[256]) << shift;
            mask   |= 0xffu << shift;
            k_rem   = hist[257];
            cbar();
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
    cbar();
    mx = s.fred[0];
#pragma unroll
    for (int i = 1; i < NCW; ++i) mx = fmaxf(mx, s.fred[i]);
    float e[PER];
    double loc = 0.0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { e[i] = (t * PER + i < vocab) ? expf(l[i] - mx) : 0.0f; loc += (double) e[i]; }
    // inclusive scan of the per-thread sums (thread order == index order)
    double inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    cbar();
    if (lane == 31) s.dred[warp] = inc;
    cbar();
    double woff = 0.0, total = 0.0;
#pragma unroll
    for (int i = 0; i < NCW; ++i) { if (i < warp) woff += s.dred[i]; total += s.dred[i]; }
    inc += woff;
    const double target = u01(seed, draw) * total;
    if (t == 0) s.tok = -1;
    cbar();
    const double exc = inc - loc;
    if (target >= exc && target < inc) {
        double a = exc; int pick = -1;
#pragma unroll
        for (int i = 0; i < PER; ++i) { a += (double) e[i]; if (pick < 0 && e[i] > 0.0f && target < a) pick = t * PER + i; }
        if (pick < 0) { for (int i = PER - 1; i >= 0; --i) if (e[i] > 0.0f) { pick = t * PER + i; break; } }
        if (pick >= 0) s.tok = pick;
    }
    cbar();
    if (s.tok < 0) {            // rounding at the very end of the CDF: fall back to the arg-max element
        if (t == 0) s.tok = 0;
        cbar();
#pragma unroll
        for (int i = 0; i < PER; ++i) if (t * PER + i < vocab && l[i] == mx) s.tok = t * PER + i;
        cbar();
    }
    const int tok = s.tok;
    cbar();
    return tok;
}

// the compute warps have just finished an exchange: the SM's return path is idle, a good moment for the
// producer to start its next bulk copy (keeps the weight stream from delaying the latency-critical polls)
template <class SM>
__device__ __forceinline__ void token(SM & s) { if (threadIdx.x == 0) s.tokens = s.tokens + 1u; }

// ------------------------------------------------------------------------------- one pass (compute warps)
struct PassRegs {          // parameters fetched ahead of use; carried across layers and passes
    Gam ga, gf;            // attn_norm / ffn_norm of the layer about to run
    float gn[4];           // q_norm (warp 0) / k_norm (warp 1) of that layer (attention CTAs)
};

__device__ __forceinline__ void prefetch_layer_params(const StackW & W, int l, PassRegs & r) {
    r.ga = load_gamma(W.attn_norm + (size_t) l * H);
    r.gf = load_gamma(W.ffn_norm + (size_t) l * H);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if ((int) blockIdx.x < NH * WIDE_G && warp < 2) {         // attention CTAs (+ the members of the wide path)
        const float * g = (warp == 0 ? W.q_norm : W.k_norm) + (size_t) l * HD;
#pragma unroll
        for (int i = 0; i < 4; ++i) r.gn[i] = __ldg(g + lane + 32 * i);
    }
}

// poll N packets (N even) from src into dst (float) — all compute threads, every load of a thread in flight at once
template <int N>
__device__ __forceinline__ void poll_vec(const DevModel & m, const Pkt * src, unsigned ep, float * dst, Watch & w) {
    constexpr int PER = (N / 2 + NCT - 1) / NCT;
    const int t = threadIdx.x;
    const int mine = t < N / 2 ? (N / 2 - 1 - t) / NCT + 1 : 0;
    float2 f[PER];
    spin_cycles(m.poll_d0);
    pk_poll2_n<PER>(m, src + 2 * t, NCT, mine, ep, f, w);
#pragma unroll
    for (int j = 0; j < PER; ++j)
        if (j < mine) { dst[2 * (t + j * NCT)] = f[j].x; dst[2 * (t + j * NCT) + 1] = f[j].y; }
}

// One token through `n_layers` layers of stack W (+ optional head).  s.xres holds x on entry (and a
// cbar() has made it visible); pr holds layer 0's parameters.  On exit pr holds layer 0's parameters
// of the next pass' stack NW (if any).
template <bool Q8>
__device__ void run_stack(const DevModel & m, const StackW & W, const PassDesc & p, int pos,
                          const HeadW * head, bool process, const GenState & gs,
                          const Slices & z, const StackW * NW,
                          Smem & s, Ring & rg, unsigned & ep, PassRegs & pr, Watch & w) {
    const int t = threadIdx.x;
    const int nA = z.a1 - z.a0, nO = z.o1 - z.o0, nP = z.p1 - z.p0, nD = z.d1 - z.d0;
    int h0 = 0, h1 = 0;
    if (head) cta_rows(head->vocab, h0, h1);
    const int nH = h1 - h0;
    const bool attn_cta = (int) blockIdx.x < NH;

    const bool wide = wide_pass(m, W, p.f32attn, p.width, pos);
    const bool wide_cta = wide && (int) blockIdx.x < NH * WIDE_G;
    AttnRegs ar;
    if ((attn_cta || wide_cta) && t < 64) {                                // RoPE row of this position (warps 0, 1)
        const float * rope_row = m.rope + (size_t) pos * HD;
        const int lane = t & 31;
        ar.rp[0] = __ldg(rope_row + lane);      ar.rp[1] = __ldg(rope_row + 64 + lane);
        ar.rp[2] = __ldg(rope_row + lane + 32); ar.rp[3] = __ldg(rope_row + 64 + lane + 32);
    }
    Gam gh;                                                  // out_norm for the head
    gh.g[0] = gh.g[1] = gh.g[2] = 0.0f;

    for (int l = 0; l < p.n_layers; ++l) {
        const bool last = l + 1 == p.n_layers;
        const unsigned e_qkv = ep + 1, e_att = ep + 2, e_x1 = ep + 3, e_m = ep + 4, e_x = ep + 5;
        // this layer's parameters are in registers; start fetching the next ones
        const Gam ga = pr.ga, gf = pr.gf;
#pragma unroll
        for (int i = 0; i < 4; ++i) ar.gn[i] = pr.gn[i];
        if (!last) prefetch_layer_params(W, l + 1, pr);
        else {
            if (head) gh = load_gamma(W.out_norm);
            if (NW) prefetch_layer_params(*NW, 0, pr);
        }

        // ---- phase A: RMSNorm + fused QKV mat-vec
        PROF(1);
        // the slice wait comes BEFORE the packet poll in every phase: the slice was requested thousands of cycles
        // ago, and what the mbarrier wait costs (~350 cycles even when the phase has completed) then overlaps the
        // flight time of the packets instead of sitting between the last packet and the first FMA
        int s0 = ring_acquire(m, s, rg, w);
        if (l > 0) { poll_vec<H>(m, m.x_x, ep, s.xres, w); cbar(); }
        token(s);
        PROF(2);
        rmsnorm_to_xh<Q8>(s, ga, m.eps, nullptr, nullptr);
        PROF(3);
        PROF(4);
        if (Q8) gemv_q8<1024>(s, s0, nA); else gemv_ring<1024, 3, false>(s, s0, nA);
        PROF(40);
        cbar();
        PROF(41);
        if (t < nA) pk_store(m.x_qkv + z.a0 + t, s.outv[t], e_qkv);
        PROF(42);
        ring_release(s, rg, nA);
        PROF(5);

        // ---- phase B: per-head norm + RoPE + KV append + attention (CTA h owns head h)
        if (wide_cta) attention_wide(m, W, l, pos, ar, e_qkv, e_att, s, rg, w);
        else if (attn_cta && !wide) attention_head(m, W, l, (int) blockIdx.x, pos, p.f32attn != 0, p.width, ar, e_qkv, e_att, s, rg, w);
        publish_kv(attn_cta && ((int) blockIdx.x & 1) == 0);
        PROF(6);

        // ---- phase C: attn_output mat-vec (input rounded to F16) + residual
        {
            constexpr int NP = NH * HD / 2, PER = (NP + NCT - 1) / NCT;
            const int mine = t < NP ? (NP - 1 - t) / NCT + 1 : 0;
            float2 f[PER];
            spin_cycles(m.poll_d0);
            s0 = ring_acquire(m, s, rg, w);
            pk_poll2_n<PER>(m, m.x_att + 2 * t, NCT, mine, e_att, f, w);
#pragma unroll
            for (int j = 0; j < PER; ++j)
                if (j < mine) {
                    if (Q8) { s.xf[2 * (t + j * NCT)] = f[j].x; s.xf[2 * (t + j * NCT) + 1] = f[j].y; }
                    else    { xh_put(s, 2 * (t + j * NCT), f[j].x); xh_put(s, 2 * (t + j * NCT) + 1, f[j].y); }
                }
        }
        cbar();
        if (Q8) quant_q8<NH * HD>(s);
        token(s);
        PROF(7);
        PROF(8);
        if (Q8) gemv_q8<2048>(s, s0, nO); else gemv_ring<2048, 1, false>(s, s0, nO);
        cbar();
        if (t < nO) pk_store(m.x_x1 + z.o0 + t, __fadd_rn(s.outv[t], s.xres[z.o0 + t]), e_x1);
        ring_release(s, rg, nO * 2);
        PROF(9);

        // ---- phase D: RMSNorm + gate/up mat-vec + SwiGLU
        s0 = ring_acquire(m, s, rg, w);
        poll_vec<H>(m, m.x_x1, e_x1, s.xres, w);
        cbar();
        token(s);
        PROF(10);
        rmsnorm_to_xh<Q8>(s, gf, m.eps, nullptr, nullptr);
        PROF(11);
        PROF(12);
        if (Q8) gemv_q8<1024>(s, s0, 2 * nP); else gemv_ring<1024, 4, false>(s, s0, 2 * nP);
        cbar();
        if (t < nP) {
            const float gt = s.outv[2 * t], up = s.outv[2 * t + 1];
            const float sl = __fdiv_rn(gt, __fadd_rn(1.0f, v_expf(__fsub_rn(0.0f, gt))));   // ggml_v_silu
            pk_store(m.x_m + z.p0 + t, __fmul_rn(sl, up), e_m);
        }
        ring_release(s, rg, 2 * nP);
        PROF(13);

        // ---- phase E: ffn_down on F32 activations (weights exact in F32) + residual
        s0 = ring_acquire(m, s, rg, w);
        poll_vec<FFN>(m, m.x_m, e_m, s.xf, w);
        cbar();
        token(s);
        PROF(14);
        PROF(15);
        if (Q8) gemv_q8_f32<3072>(s, s0, nD); else gemv_ring<3072, 1, true>(s, s0, nD);
        cbar();
        if (t < nD) pk_store(m.x_x + z.d0 + t, __fadd_rn(s.outv[t], s.xres[z.d0 + t]), e_x);
        ring_release(s, rg, nD * 3);
        PROF(16);
        ep += 5;
    }
    if (!head) return;
    // ---- head: final RMSNorm (= hidden_states) + vocabulary mat-vec + per-CTA arg-max candidate
    const unsigned e_head = ep + 1;
    poll_vec<H>(m, m.x_x, ep, s.xres, w);
    cbar();
    token(s);
    if (p.nonorm) {
        // legacy predict_codes graph (src/tts_transformer.cpp:1517-1524): the heads multiply the un-normed
        // output of the last layer; the mat-vec rounds it to F16 like any F16-weight ggml_mul_mat
        for (int k = t; k < H; k += NCT) {
            xh_put(s, k, s.xres[k]);
            if (blockIdx.x == 0 && p.hidden_dst) p.hidden_dst[k] = s.xres[k];
        }
        cbar();
    } else {
        rmsnorm_to_xh(s, gh, m.eps, blockIdx.x == 0 ? m.hidden : nullptr, blockIdx.x == 0 ? p.hidden_dst : nullptr);
    }
    PROF(20);
    const int s0 = ring_acquire(m, s, rg, w);
    gemv_ring<1024, 2, false>(s, s0, nH);
    cbar();
    // the CTA's rows (nH <= 22 on >= 140 CTAs) all sit in warp 0: its arg-max candidate (first maximum wins, as the host
    // argmax of src/tts_transformer.cpp:2182-2192) is folded with shuffles instead of a serial scan through shared memory
    float pv = -CUDART_INF_F; int pi = 0x7fffffff;
    if (t < nH) {
        const float raw = s.outv[t];
        m.logits[h0 + t] = raw;
        if (p.logits_dst) p.logits_dst[h0 + t] = raw;
        pk_store(m.x_logit + h0 + t, raw, e_head);
        pv = proc_logit(m, raw, h0 + t, process, gs.suppress_start, gs.eos_id, gs.rep_penalty);
        pi = h0 + t;
    }
    ring_release(s, rg, nH);
    if (t < 32) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, pv, o);
            const int   oi = __shfl_xor_sync(0xffffffffu, pi, o);
            if (ov > pv || (ov == pv && oi < pi)) { pv = ov; pi = oi; }
        }
        if (t == 0) {
            pk_store(m.x_cand + 2 * blockIdx.x, pv, e_head);
            pk_store(m.x_cand + 2 * blockIdx.x + 1, __int_as_float(pi), e_head);
        }
    }
    ep += 1;
    PROF(21);
}

constexpr int PD_CACHE = 20;               // passes of a program kept in shared memory (longer programs read the rest from HBM)
__device__ __forceinline__ const PassDesc & pass_desc(const PassDesc * __restrict__ passes, const PassDesc * pds, int i) {
    return i < PD_CACHE ? pds[i] : passes[i];
}

// ------------------------------------------------------------------------------- producer warp
// Walks the same pass / layer / slice order as run_stack and keeps the ring full.  s_go: -1 until
// the compute warps know how many passes of the program will run (EOS / max_len), then that count.
__device__ void producer_loop(const DevModel & m, const PassDesc * __restrict__ passes_g, const PassDesc * pds, int n_passes,
                              const Slices & z, Smem & s, volatile int * s_go, int n_past0) {
    struct { const PassDesc * g; const PassDesc * c; __device__ const PassDesc & operator[](int i) const { return pass_desc(g, c, i); } } passes{passes_g, pds};
    unsigned p_slot = 0, p_idx = 0;
    const int nA = z.a1 - z.a0, nO = z.o1 - z.o0, nP = z.p1 - z.p0, nD = z.d1 - z.d0;
    bool aborted = false;
    const long long t_start = clock64();
    unsigned pspins = 0;
    int n_past = n_past0;
    const bool attn_cta = (int) blockIdx.x < NH;
    // L2 residency (DevModel::l2_pin > 0): the code predictor's weights are re-read by each of its 16 passes per frame, but a pass
    // (161 MB F16) does not fit the 126 MB L2 next to the talker stream, so nothing survived between passes.  The copies of its
    // first l2_pin weight slices ask L2 to keep the lines (evict_last); everything that is read once per frame (talker layers, the
    // rest of the code predictor, the heads) is marked evict_first so that it does not push them out.
    unsigned long long pol_first = 0ull, pol_last = 0ull;
    if (m.l2_pin > 0) {
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_first));
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_last));
    }
    for (int ip = 0; ip < n_passes && !aborted; ++ip) {
        int go = *s_go;
        while (go < 0 && ip > 0) {                                    // pass 0 is produced optimistically
            __nanosleep(64); go = *s_go;
            if (((++pspins) & 0xfff) == 0 && clock64() - t_start > 8000000000LL) { m.st->error = 1; aborted = true; break; }
        }
        if (aborted) break;
        if (go >= 0 && ip >= go) break;
        const int xmode = passes[ip].xmode;
        if (xmode == XM_SELECT_ONLY) continue;
        const StackW & W = passes[ip].stack == 0 ? m.talker : m.cp;
        const int n_layers = passes[ip].n_layers, hd = passes[ip].head;
        const int pos = passes[ip].pos >= 0 ? passes[ip].pos : n_past;
        if (passes[ip].stack == 0 && passes[ip].pos < 0) ++n_past;
        // attention CTAs: the K then V history of the head (rows 0..pos-1) follows the QKV slice, KV_CH rows per slice
        const bool fast = max(passes[ip].width, pos + 1) <= ATT_FAST_KV;        // same test as attention_head
        const bool wide = wide_pass(m, W, passes[ip].f32attn, passes[ip].width, pos);
        const int n_kvs = (attn_cta && !passes[ip].f32attn && !fast && !wide) ? (pos + KV_CH - 1) / KV_CH : 0;
        // wide attention: this member's share of the K rows, then its 16-dim slice of every V row
        int wk0 = 0, wk1 = 0, n_wk = 0, n_wv = 0;
        const int wh = (int) blockIdx.x % NH, wj = (int) blockIdx.x / NH;
        if (wide && (int) blockIdx.x < NH * WIDE_G) {
            wide_k_range(pos, wj, wk0, wk1);
            n_wk = (wk1 - wk0 + KV_CH - 1) / KV_CH;
            n_wv = (pos + VS_CH - 1) / VS_CH;
        }
        const int n_mid = 2 * n_kvs + n_wk + n_wv;
        for (int l = 0; l < n_layers && !aborted; ++l) {
            for (int k = 0; k < 4 + n_mid; ++k) {
                const void * src; int n_slots, slot_bytes = W.wtype == Q3T_TYPE_Q8_0 ? SLOT_Q8 : SLOT_B, tail = SLOT_B;
                // l2_pin counts weight slices of the code predictor in schedule order (4 per layer: QKV, O, gate/up, down)
                const int wk = k == 0 ? 0 : k - n_mid;            // 0..3 for the four weight slices (<= 0 for K/V history slices)
                unsigned long long pol = (passes[ip].stack == 1 && l * 4 + wk < m.l2_pin) ? pol_last : pol_first;
                if (k == 0) { src = sl_a(W, l, z); n_slots = nA; }
                else if (k <= n_wk) {
                    const int c = k - 1;
                    const int rows = min(KV_CH, wk1 - wk0 - c * KV_CH);
                    src = W.kcache + ((size_t) (l * NKV + (wh >> 1)) * W.n_ctx + (size_t) (wk0 + c * KV_CH)) * HD;
                    n_slots = (rows + 7) >> 3; slot_bytes = SLOT_B; tail = (rows - (n_slots - 1) * 8) * HD * 2; pol = 0ull;
                } else if (k <= n_wk + n_wv) {
                    const int cs = k - 1 - n_wk;
                    const int bytes = min(VS_CH, pos - cs * VS_CH) * (VS_D * 2);
                    src = W.vslice + (((size_t) (l * NKV + (wh >> 1)) * WIDE_G + wj) * W.n_ctx + (size_t) cs * VS_CH) * VS_D;
                    n_slots = (bytes + SLOT_B - 1) / SLOT_B; slot_bytes = SLOT_B; tail = bytes - (n_slots - 1) * SLOT_B; pol = 0ull;
                } else if (k <= 2 * n_kvs) {
                    const int c = (k - 1) % n_kvs;
                    const int rows = min(KV_CH, pos - c * KV_CH);
                    const __half * base = (k - 1 < n_kvs ? W.kcache : W.vcache)
                                        + ((size_t) (l * NKV + ((int) blockIdx.x >> 1)) * W.n_ctx + (size_t) c * KV_CH) * HD;
                    src = base; n_slots = (rows + 7) >> 3; slot_bytes = SLOT_B; tail = (rows - (n_slots - 1) * 8) * HD * 2; pol = 0ull;
                } else {
                    const int kk = k - n_mid;
                    src = kk == 1 ? sl_o(W, l, z) : kk == 2 ? sl_gu(W, l, z) : sl_d(W, l, z);
                    n_slots = kk == 1 ? nO * 2 : kk == 2 ? nP * 2 : nD * 3;
                }
                // ring-space wait with an abort check (EOS decided while pass 0 is being produced)
                while (true) {
                    const unsigned long long c = s.consumed;
                    if (p_slot + (unsigned) n_slots - (unsigned) c <= (unsigned) NSLOT && p_idx - (unsigned) (c >> 32) < (unsigned) NMB &&
                        true) break;
                    if (*s_go == 0) { aborted = true; break; }
                    if (((++pspins) & 0xfff) == 0 && clock64() - t_start > 8000000000LL) { m.st->error = 1; aborted = true; break; }   // watchdog
                }
                if (aborted) break;
                produce_slice(s, p_slot, p_idx, src, n_slots, slot_bytes, tail, m.prod_pace, m.prod_piece, pol);
            }
        }
        if (hd >= 0 && !aborted) {
            const HeadW & hw = hd == 0 ? m.codec_head : m.cp_head[hd - 1];
            int h0, h1; cta_rows(hw.vocab, h0, h1);
            while (true) {
                const unsigned long long c = s.consumed;
                if (p_slot + (unsigned) (h1 - h0) - (unsigned) c <= (unsigned) NSLOT && p_idx - (unsigned) (c >> 32) < (unsigned) NMB) break;
                if (*s_go == 0) { aborted = true; break; }
                if (((++pspins) & 0xfff) == 0 && clock64() - t_start > 8000000000LL) { m.st->error = 1; aborted = true; break; }
            }
            if (!aborted) produce_slice(s, p_slot, p_idx, hw.w + (size_t) h0 * H, h1 - h0, SLOT_B, SLOT_B, m.prod_pace, m.prod_piece, pol_first);
        }
    }
    if (*s_go == 0 || aborted) {
        // the compute warps stopped consuming: wait for the copies already issued before the CTA retires
        const unsigned c_idx = (unsigned) (s.consumed >> 32);
        for (unsigned i = c_idx; i < p_idx; ++i) mbar_wait(&s.mbar[i % NMB], (i / NMB) & 1u);
    }
}

// --------------------------------------------------------------------------- the kernel
// frame_mode = 1: `passes` is the 17-pass frame program; the kernel owns the generate-loop
// bookkeeping of src/tts_transformer.cpp:2664-2804 (EOS test, code emission, repetition set,
// n_past) in GenState so the identical launch can be replayed once per frame.
template <bool Q8>
__global__ void __launch_bounds__(LM_THREADS, 1)
lm_kernel(const __grid_constant__ DevModel m, const PassDesc * __restrict__ passes, int n_passes, int frame_mode) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem & s = *reinterpret_cast<Smem *>(smem_raw);
    __shared__ GenState gs;
    __shared__ PassDesc pds[PD_CACHE];     // the program (a frame is 17 passes): a PassDesc read from global memory at every pass
                                           // boundary cost two dependent L2 round trips on the critical path
    __shared__ int s_go_raw;
    volatile int * s_go = &s_go_raw;
    const int t = threadIdx.x;

    {
        const int n_ints = (n_passes < PD_CACHE ? n_passes : PD_CACHE) * (int) (sizeof(PassDesc) / sizeof(int));
        if (t < n_ints) reinterpret_cast<int *>(pds)[t] = reinterpret_cast<const int *>(passes)[t];
    }
    if (t == 0) {
        gs = *m.st;
        s_go_raw = frame_mode ? -1 : n_passes;
        s.consumed = 0ULL;
        s.tokens = 0u;
        s.prof_cnt = (m.prof_frame < 0 || m.prof_frame == gs.frame) ? 0 : -1;    // record one chosen frame only
        for (int i = 0; i < NMB; ++i) mbar_init(&s.mbar[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if ((frame_mode && gs.done) || gs.error) return;      // a tripped watchdog makes every later launch a no-op

    Slices z;
    cta_rows(QKV_N, z.a0, z.a1);
    cta_rows(H,     z.o0, z.o1);
    cta_rows(FFN,   z.p0, z.p1);
    cta_rows(H,     z.d0, z.d1);

    if (t >= NCT) {                                   // ---- producer warp: one thread drives the TMA unit
        if (t == NCT) producer_loop(m, passes, pds, n_passes, z, s, s_go, gs.n_past);
        return;
    }

    // ---- compute warps
    Watch w; w.t0 = clock64(); w.dead = false; w.err = &m.st->error;
    Ring rg; rg.slot_head = 0; rg.slice_idx = 0;
    unsigned ep = gs.ep_base;
    unsigned pending_ep = gs.pending_ep;
    int frame = gs.frame, n_past = gs.n_past, done = gs.done, pending = gs.pending;
    int pending_vocab = gs.codec_vocab, pending_process = 1;
    unsigned long long draw = gs.draw;
    PassRegs pr;
    bool pr_valid = false;

    for (int ip = 0; ip < n_passes; ++ip) {
        PROF(30);
        const PassDesc & p = pass_desc(passes, pds, ip);
        const StackW & W = p.stack == 0 ? m.talker : m.cp;
        if (!pr_valid && p.xmode != XM_SELECT_ONLY) { prefetch_layer_params(W, 0, pr); pr_valid = true; }

        // ---- select the token produced by the previous head, if one is waiting
        const bool needs_select = p.xmode == XM_HIDDEN || p.xmode == XM_CP_EMBD ||
                                  p.xmode == XM_TALKER_NEXT || p.xmode == XM_SELECT_ONLY;
        if (needs_select && pending >= 0) {
            int tok;
            if (gs.temperature <= 0.0f) tok = select_greedy(m, s, pending_ep, w);
            else tok = select_sample(m, s, pending_ep, w, pending_vocab, pending == 0 && pending_process, gs.suppress_start, gs.eos_id,
                                     gs.rep_penalty, gs.temperature, gs.top_k, gs.seed, draw++);
            if (t == 0) gs.cur[pending] = tok;
            cbar();
            pending = -1;
        }
        if (p.xmode == XM_SELECT_ONLY) continue;

        if (p.xmode == XM_HIDDEN && frame_mode) {
            // src/tts_transformer.cpp:2724-2729: EOS stops before a frame is emitted
            if (gs.cur[0] == gs.eos_id) { done = 1; if (t == 0) *s_go = 0; break; }
            if (blockIdx.x == 0 && t == 0) { m.seen[gs.cur[0]] = 1; __threadfence(); }
        }
        if (ip == 0 && frame_mode && t == 0) {
            // :2755-2757 no talker step after the last frame: the producer must not stream it either
            const bool trunc = pass_desc(passes, pds, n_passes - 1).xmode == XM_TALKER_NEXT && frame + 1 >= gs.max_len;
            *s_go = trunc ? n_passes - 1 : n_passes;
        }
        if (p.xmode == XM_TALKER_NEXT && frame_mode) {
            if (blockIdx.x == 0 && t < NCB) m.codes_out[(size_t) frame * NCB + t] = gs.cur[t];
            ++frame;
            if (frame >= gs.max_len) { done = 1; break; }
        }

        // ---- form x in shared memory (elements t, t+352, t+704)
        for (int k = t; k < H; k += NCT) {
            float xv = 0.f;
            switch (p.xmode) {
                case XM_BUF:    xv = ld_act(p.xsrc + k); break;
                case XM_HIDDEN: xv = ld_act(m.hidden + k); break;
                case XM_CB0_EMBD: xv = __half2float(m.codec_embd[(size_t) gs.cur[0] * H + k]); break;
                case XM_CP_EMBD:  xv = __half2float(m.cp_embd[p.arg - 1][(size_t) gs.cur[p.arg] * H + k]); break;
                case XM_TALKER_NEXT: {
                    // :2759-2790  codec_embd[cb0] + sum_i code_pred_embd[i-1][c_i] + trailing/pad, F32 in this order
                    float hv[NCB];
                    hv[0] = __half2float(m.codec_embd[(size_t) gs.cur[0] * H + k]);
#pragma unroll
                    for (int cb = 1; cb < NCB; ++cb) hv[cb] = __half2float(m.cp_embd[cb - 1][(size_t) gs.cur[cb] * H + k]);
                    const int fr = frame - 1;      // the frame just emitted
                    const float * tr = (fr < gs.trailing_len) ? (m.trailing + (size_t) fr * H) : m.pad_embed;
                    const float tv = ld_act(tr + k);
                    xv = hv[0];
#pragma unroll
                    for (int cb = 1; cb < NCB; ++cb) xv = __fadd_rn(xv, hv[cb]);
                    xv = __fadd_rn(xv, tv);
                } break;
                default: break;
            }
            s.xres[k] = xv;
        }
        cbar();
        PROF(31);

        const int pos = p.pos >= 0 ? p.pos : n_past;
        const HeadW * head = p.head < 0 ? nullptr : (p.head == 0 ? &m.codec_head : &m.cp_head[p.head - 1]);
        // the next pass' first-layer parameters are fetched across the pass boundary
        const StackW * NW = nullptr;
        if (ip + 1 < n_passes) {
            const int nx = pass_desc(passes, pds, ip + 1).xmode, ns = pass_desc(passes, pds, ip + 1).stack;
            if (nx != XM_SELECT_ONLY) NW = ns == 0 ? &m.talker : &m.cp;
        }
        if (!NW) pr_valid = false;
        run_stack<Q8>(m, W, p, pos, head, p.process != 0, gs, z, NW, s, rg, ep, pr, w);
        if (head) { pending = p.head; pending_vocab = head->vocab; pending_process = p.process; pending_ep = ep; }
        if (p.stack == 0 && p.pos < 0) ++n_past;
    }

    if (m.prof && t == 0 && s.prof_cnt >= 0) m.prof_n[blockIdx.x] = (unsigned) s.prof_cnt;
    // ---- write the state back (CTA 0); epochs are identical in every CTA
    if (blockIdx.x == 0 && t == 0) {
        GenState * st = m.st;
        st->frame = frame; st->n_past = n_past; st->done = done; st->pending = pending;
        st->draw = draw;
        for (int i = 0; i < NCB; ++i) st->cur[i] = gs.cur[i];
        st->ep_base = ep; st->pending_ep = pending_ep;
    }
}

#include "lm_batch.inc"

// --------------------------------------------------------------------------- sampler in isolation
// Test hook ("distribution test of the sampler in isolation"): draws n tokens from ONE logit
// vector through select_sample, i.e. exactly the code the decode kernels run after a head
// (src/tts_transformer.cpp:2349-2387 for the code predictor, :2664-2722 with suppress / repetition penalty
// for codebook 0).  The logits arrive as epoch-1 packets in a private buffer; CTA b takes draws b, b + grid, ...
__global__ void __launch_bounds__(LM_THREADS, 1)
sample_logits_kernel(const __grid_constant__ DevModel m, int vocab, int process, int suppress_start, int eos_id, float rep_penalty,
                     float temperature, int top_k, unsigned long long seed, unsigned long long draw0, int n_draws, int * __restrict__ out) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem & s = *reinterpret_cast<Smem *>(smem_raw);
    if (threadIdx.x >= NCT) return;
    Watch w; w.t0 = clock64(); w.dead = false; w.err = &m.st->error;
    for (int d = blockIdx.x; d < n_draws; d += gridDim.x) {
        const int tok = select_sample(m, s, 1u, w, vocab, process != 0, suppress_start, eos_id, rep_penalty, temperature, top_k,
                                      seed, draw0 + (unsigned long long) d);
        if (threadIdx.x == 0) out[d] = tok;
    }
}

int lm_sample_logits(q3t_ctx * ctx, const Pkt * logits_pk_dev, const unsigned char * seen_dev, int vocab, int process,
                     float rep_penalty, float temperature, int top_k, unsigned long long seed, unsigned long long draw0,
                     int n_draws, int * out_dev) {
    if (lm_max_grid(ctx) <= 0) return set_err(ctx, Q3T_ERR_CUDA, "cannot configure the decode kernels on this device");
    cudaError_t e = cudaFuncSetAttribute(sample_logits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(Smem));
    if (e != cudaSuccess) return cuda_check(ctx, e, "cudaFuncSetAttribute");
    DevModel m = ctx->hm;
    m.x_logit = const_cast<Pkt *>(logits_pk_dev);
    m.seen = const_cast<unsigned char *>(seen_dev);
    m.poll_d0 = 0; m.poll_d1 = 0;
    const int grid = n_draws < 128 ? n_draws : 128;
    sample_logits_kernel<<<grid, LM_THREADS, sizeof(Smem), ctx->stream>>>(m, vocab, process, ctx->cfg.codec_vocab_size - 1024, ctx->cfg.codec_eos_id,
                                                                         rep_penalty, temperature, top_k, seed, draw0, n_draws, out_dev);
    ctx->timing.kernel_launches++;
    return cuda_check(ctx, cudaGetLastError(), "sample_logits_kernel launch");
}


// --------------------------------------------------------------------------- host launch
int lm_max_grid(q3t_ctx * ctx) {
    // function attributes are per device: one flag per device, not per process (a process may hold contexts on several GPUs)
    static bool attr_set_dev[64] = {};
    bool & attr_set = attr_set_dev[ctx->device & 63];
    const size_t smem = sizeof(Smem);
    if (!attr_set) {
        if (cudaFuncSetAttribute(lm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem) != cudaSuccess) return -1;
        if (cudaFuncSetAttribute(lm_kernel<true>,  cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem) != cudaSuccess) return -1;
        attr_set = true;
    }
    int per_sm = 0, per_sm_q = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lm_kernel<false>, LM_THREADS, smem) != cudaSuccess) return -1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_q, lm_kernel<true>, LM_THREADS, smem) != cudaSuccess) return -1;
    if (per_sm < 1 || per_sm_q < 1) return 0;
    return ctx->sm_count;          // one persistent CTA per SM
}

// producer pacing (see produce_slice / producer_loop); environment overrides are tuning knobs
static void lm_pacing(DevModel & m, bool batched = false) {
    static int pp = -2, pc = -1;
    if (pp == -2) {
        const char * e1 = getenv("Q3T_PROD_PACE"), * e2 = getenv("Q3T_PROD_PIECE");
        pp = e1 ? atoi(e1) : -1; pc = e2 ? atoi(e2) : 4;
        if (pc < 1) pc = 1;
    }
    // the batch-1 kernel wants its refills in 8 KB pieces 80 cycles apart; the batched kernel hides its exchanges behind the
    // other group's arithmetic and is (slightly) better off with whole-slice copies (5.84 -> 5.78 ms per 16-sequence frame)
    m.prod_pace = pp >= 0 ? pp : (batched ? 0 : 80); m.prod_piece = pc;
    static int pin = -1;
    if (pin == -1) { const char * e = getenv("Q3T_L2_PIN"); pin = e ? atoi(e) : -2; }
    // measured plateau (profiles/r02c_l2_sweep.txt): 6-9 slices (47-71 MB) of the F16 code predictor, 16 of the Q8_0 one.
    // (Prefetching every slice into L2 one layer ahead of its ring copy was measured too: 1.361 -> 1.439 ms per frame.)
    m.l2_pin = pin >= 0 ? pin : (m.cp.wtype == Q3T_TYPE_Q8_0 ? 16 : 9);
}

int lm_launch(q3t_ctx * ctx, const PassDesc * passes_dev, int n_passes, int frame_mode) {
    if (ctx->grid <= 0) {
        const int g = lm_max_grid(ctx);
        if (g < LM_MIN_GRID) return set_err(ctx, Q3T_ERR_CUDA, "device cannot co-schedule the persistent decode kernel (needs >= 140 SMs)");
        ctx->grid = g > 480 ? 480 : g;
    }
    DevModel m = ctx->hm;
    static int d0 = -1, d1 = -1;
    if (d0 < 0) { const char * e0 = getenv("Q3T_POLL_D0"), * e1 = getenv("Q3T_POLL_D1"); d0 = e0 ? atoi(e0) : 150; d1 = e1 ? atoi(e1) : 0; }
    m.poll_d0 = d0; m.poll_d1 = d1;
    lm_pacing(m);
    void * args[] = { (void *) &m, (void *) &passes_dev, (void *) &n_passes, (void *) &frame_mode };
    // both stacks of a model share the weight type (the converter quantises them together)
    const bool q8 = m.talker.wtype == Q3T_TYPE_Q8_0;
    cudaError_t e = cudaLaunchCooperativeKernel(q8 ? (const void *) lm_kernel<true> : (const void *) lm_kernel<false>,
                                                dim3(ctx->grid), dim3(LM_THREADS), args, sizeof(Smem), ctx->stream);
    ctx->timing.kernel_launches++;
    return cuda_check(ctx, e, "lm_kernel launch");
}

// batched frame launch: nb sequences (views in `seqs`), instantiation = smallest NB >= group size, F16 or Q8_0 layer weights
template <int NB, bool Q8>
static cudaError_t launch_batch_t(q3t_ctx * ctx, const BatchModel & bm, const PassDesc * passes_dev, int n_passes) {
    static bool attr_set_dev[64] = {};                         // per device, see lm_max_grid
    bool & attr_set = attr_set_dev[ctx->device & 63];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(lm_batch_kernel<NB, Q8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(BSmem<NB>));
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    void * args[] = { (void *) &bm, (void *) &passes_dev, (void *) &n_passes };
    return cudaLaunchCooperativeKernel((const void *) lm_batch_kernel<NB, Q8>, dim3(ctx->grid), dim3(LM_THREADS), args,
                                       sizeof(BSmem<NB>), ctx->stream);
}

int lm_launch_batch(q3t_ctx * ctx, const DevModel * seqs, int nb, const PassDesc * passes_dev, int n_passes) {
    if (nb < 1 || nb > Q3T_MAX_BATCH) return set_err(ctx, Q3T_ERR_ARG, "batch size out of range");
    if (ctx->grid <= 0) {
        const int g = lm_max_grid(ctx);
        if (g < LM_MIN_GRID) return set_err(ctx, Q3T_ERR_CUDA, "device cannot co-schedule the persistent decode kernel (needs >= 140 SMs)");
        ctx->grid = g > 480 ? 480 : g;
    }
    // up to 8 sequences: one lock-step group; more: two groups interleaved phase by phase (lm_batch.inc)
    BatchModel bm;
    bm.nb = nb;
    bm.ng = nb > 8 ? 2 : 1;
    bm.gbase[0] = 0; bm.gbase[1] = bm.ng == 2 ? (nb + 1) / 2 : nb; bm.gbase[2] = nb;
    if (NH * bm.gbase[1] > ctx->grid) return set_err(ctx, Q3T_ERR_ARG, "batch needs one CTA per (sequence, head) attention unit of a group");
    for (int q = 0; q < Q3T_MAX_BATCH; ++q) bm.seq[q] = seqs[q < nb ? q : 0];
    lm_pacing(bm.seq[0], true);
    // both stacks of a model share the weight type (the converter quantises them together); Q8_0 models run the 8-wide instantiation
    const bool q8 = seqs[0].talker.wtype == Q3T_TYPE_Q8_0;
    cudaError_t e = q8      ? launch_batch_t<8, true>(ctx, bm, passes_dev, n_passes)
                  : nb <= 2 ? launch_batch_t<2, false>(ctx, bm, passes_dev, n_passes)
                  : nb <= 4 ? launch_batch_t<4, false>(ctx, bm, passes_dev, n_passes)
                            : launch_batch_t<8, false>(ctx, bm, passes_dev, n_passes);
    ctx->timing.kernel_launches++;
    return cuda_check(ctx, e, "lm_batch_kernel launch");
}

} // namespace q3t
