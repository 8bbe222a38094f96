// Kernel (1): fused credibility-biased FlashAttention-style prefill for sm_100a (v5.22: persistent, two query tiles
// per CTA, ping-pong softmax warpgroups, 128-key tiles).
//
//   S = Q K^T      tcgen05.mma (SS) M128 x N128 x K16, 128x128 fp32 accumulator in TMEM, operands TMA-staged in
//                  128B-swizzled shared memory. N = 128 is what keeps this product off the shared-memory read port:
//                  every instruction re-reads its 4 KB slice of Q whatever N is, so at N = 64 (v4) the pair of operands
//                  needed 192 B/clk of a 128 B/clk port and Q K^T ran at 2/3 of the tensor rate
//                  (profiles/r2_attn_ablation_and_trace.md); at N = 128 it is 8 KB per 64 cycles = 128 B/clk
//   online softmax one thread per query row straight out of TMEM (tcgen05.ld 32x32b); the per-key
//                  log-credibility bias (already x log2e, with the key-length mask folded in as
//                  -inf) is one packed FFMA2: t = S*scale*log2e + bias[k]
//   O += P V       tcgen05.mma (TS): P is written back to TMEM as bf16/fp16 over the S columns and
//                  consumed as the A operand; V is the MN-major B operand from shared memory
//
// This replaces the reference's dense [b,H,q,k] mask rewrite (utils/re_weighting.py:36-61) plus HF
// eager attention (modeling_llama.py:199-221): the bias is O(T) instead of O(H*T^2) and S/P never
// touch HBM. Biasing head h by logcred[k] before the softmax equals multiplying its probabilities
// by cred[k] and renormalising (README.md:21).
//
// CTA = 256 query rows (two 128-row tiles, "stages" 0 and 1) of one (sequence, head); 12 warps:
//   warps 0-3   softmax of stage 0 (TMEM lanes 32w..32w+31), correction and epilogue of that stage
//   warps 4-7   same for stage 1
//   warp  8     producer: TMA loads of Q0, Q1, K[j], V[j]
//   warp  9     MMA issuer (one elected lane); owns the TMEM allocation
//   warp  10    builds the per-tile bias vectors (log-credibility x log2e, -inf past the key length), 2 tiles ahead
//   warp  11    idle (warps 8-11 exist so that setmaxnreg can hand their registers to the softmax warps)
// The two stages share every K/V tile (half the shared-memory traffic per FLOP of a one-tile CTA) and
// alternate on the tensor pipe: while warpgroup s runs the softmax of tile j, the MMA warp runs
// P V (j) and Q K^T (j+1) of the other stage. TMEM: S0 | S1 | O0 | O1 (4 x 128 columns); P (64 packed columns)
// aliases the first half of S, so per stage the chain is softmax(j) -> P V(j) -> Q K^T(j+1) -> softmax(j+1).
// There is no running max in the hot loop: the exponentials of a half-tile are taken against a standing reference, its row
// sum is the overflow detector, the reference moves lazily and one tile late (see the softmax section below).
//
// PERSISTENT: one CTA per SM walks a dynamic queue of work items (256 rows of one sequence and head); the next item is
// taken three tiles before the end of the current one. On heads that carry the bias, 128-key tiles whose keys all have
// credibility 0 are never visited (live_tiles_kernel + LiveTiles). The output leaves through shared memory and TMA stores.
// The kernel is instruction-cache sensitive: code that runs once per item (the diagonal tile's body, the epilogue, the
// rare exact path) is kept small or out of line -- profiles/r2_attn_ablation_and_trace.md has the event logs behind each
// of these choices.
#include <atomic>
#include <cstdlib>
#include <mutex>
#include <type_traits>
#include <utility>

#include "attn.cuh"
#include "sm100_ptx.cuh"

namespace cram {

namespace {

constexpr int kBlockM = 128;  // rows per query tile (= TMEM lanes)
constexpr int kStages = 2;    // query tiles per CTA
constexpr int kBlockN = 128;  // keys per K/V tile
constexpr int kKV = 2;        // K ring depth = V ring depth (32 KB per slot at d = 128)
constexpr int kBiasRing = 4;  // bias vectors in flight (written with K[j], read until P V[j] is done)
constexpr int kThreads = 384;
constexpr int kSched = 4;           // scheduler ring depth (work items published ahead)
constexpr int kItemConsumers = 10;  // MMA warp, bias warp, eight softmax warps (lane 0 of each frees a slot)
constexpr int kTmemCols = 512;
constexpr float kLog2e = 1.4426950408889634f;
// Debug timeline (scripts/trace_attn.py builds a separate library with -DCRAM_ATTN_TRACE=<block>; the product
// build compiles none of this): clock64 stamps of one CTA's softmax warps, MMA issuer and producer, per tile.
#ifdef CRAM_ATTN_TRACE
__device__ long long g_trace[12][64][4];
__device__ long long g_trace_fine[16][16];  // warp 0, tiles 16..31: stamps inside softmax_tile
__device__ int g_trace_tile;
#define CRAM_TRACE(slot, j, e)                                                              \
  do {                                                                                      \
    if (blockIdx.x == CRAM_ATTN_TRACE && (j) < 64) g_trace[slot][j][e] = clock64();         \
  } while (0)
#define CRAM_TRACE_FINE(e)                                                                                   \
  do {                                                                                                       \
    if (blockIdx.x == CRAM_ATTN_TRACE && threadIdx.x == 0 && g_trace_tile >= 16 && g_trace_tile < 32)          \
      g_trace_fine[g_trace_tile - 16][e] = clock64();                                                         \
  } while (0)
// Event log of the persistent kernel (one CTA): role `slot` appends (clock, tag) pairs; scripts/trace_attn.py prints the
// timeline. tag = item ordinal of the CTA << 16 | tile << 4 | event kind.
__device__ long long g_evlog[8][2048][2];
#define CRAM_EV_DECL int ev_n = 0
#define CRAM_EV(slot, item, tile, kind)                                                        \
  do {                                                                                         \
    if (blockIdx.x == CRAM_ATTN_TRACE && ev_n < 2048) {                                        \
      g_evlog[slot][ev_n][0] = clock64();                                                      \
      g_evlog[slot][ev_n][1] = ((long long)(item) << 16) | ((long long)(tile) << 4) | (kind); \
      ++ev_n;                                                                                  \
    }                                                                                          \
  } while (0)
#else
#define CRAM_EV_DECL \
  do {               \
  } while (0)
#define CRAM_EV(slot, item, tile, kind) \
  do {                                  \
  } while (0)
#define CRAM_TRACE(slot, j, e) \
  do {                         \
  } while (0)
#define CRAM_TRACE_FINE(e) \
  do {                     \
  } while (0)
#endif

// Per-launch scratch of the persistent kernel, out of a ring of module globals (one entry per launch in flight): the
// dynamic work-queue counter, and the live-tile masks of the dead-tile pre-pass (bit j of live[seq]: the 128-key tile j
// of that sequence holds at least one key whose credibility is not zero).
constexpr int kMaxMaskSeqs = 1024;
constexpr int kMaxMaskTiles = 64;
struct LaunchScratch {
  unsigned counter, pad;
  unsigned long long live[kMaxMaskSeqs];
};
__device__ LaunchScratch g_scratch[64];

struct Barriers {
  uint64_t q_full[2];
  uint64_t k_full[kKV], k_empty[kKV];
  uint64_t v_full[kKV], v_empty[kKV];
  uint64_t bias_full[kBiasRing];
  uint64_t s_full[2];                   // [query stage] S(j) is complete (one phase per tile)
  uint64_t p_full[2][2];                // [query stage][half] P of that 64-key half is written (one phase per tile)
  uint64_t o_half[2];                   // [query stage] P V of the first half has completed (exact path of the second)
  uint64_t q_empty[2];                  // [query stage] the last Q K^T that reads this Q tile has completed
  uint64_t item_full[kSched], item_empty[kSched];  // scheduler ring: work item published / read by every consumer
  uint64_t o_final[2];                  // [query stage] single phase: the last P V has completed
};

template <int DH>
struct Smem {
  static constexpr int kQTileBytes = kBlockM * DH * 2;
  static constexpr int kQSlabBytes = kBlockM * 128;  // 128 rows x 64 elements (one 128B swizzle slab)
  static constexpr int kTileBytes = kBlockN * DH * 2;  // one K / V tile
  static constexpr int kSlabBytes = kBlockN * 128;     // 128 rows x 64 elements
  static constexpr int kQ = 0;
  static constexpr int kK = kQ + kStages * kQTileBytes;
  static constexpr int kV = kK + kKV * kTileBytes;
  static constexpr int kBias = kV + kKV * kTileBytes;
  static constexpr int kSched = kBias + kBiasRing * kBlockN * 4;  // Item [kSched]: decoded work items, published by the producer
  static constexpr int kBar = kSched + cram::kSched * 64;
  static constexpr int kTmemPtr = kBar + (int)sizeof(Barriers);
  // epilogue staging: per softmax warp one 32-row x 64-column box of the output (4 KB = 2 x kOutBoxBytes, 128-byte rows,
  // 128B swizzle) that a TMA store carries to global memory
  static constexpr int kOutBoxBytes = 32 * 32 * 2;
  static constexpr int kOut = (kTmemPtr + 16 + 1023) / 1024 * 1024;
  static constexpr int kEnd = kOut + 8 * 2 * kOutBoxBytes;
  // Dynamic smem request: this is ALL of the 227 KB a CTA can have at d = 128 (no slack for a manual round-up: the kernel
  // has no static shared memory, so the dynamic window starts 1024-aligned, and traps if that ever stops being true);
  // always above half an SM, so exactly one CTA (one 512-column TMEM allocation) is resident per SM.
  static constexpr int kRequest = kEnd > 120 * 1024 ? kEnd : 120 * 1024;
  static_assert(kEnd <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Timing ablations (scripts/dev/variants.py builds separate libraries with these; results are WRONG on purpose, the
// product build defines none): a share of the exponentials is replaced by one FMA-pipe instruction to measure how far
// the kernel is from the point where the MUFU pipe stops mattering.
__device__ __forceinline__ float ex2_fake(float x) { return fmaf(x, 0.001f, 0.5f); }
#if defined(CRAM_ABL_NOEXP)
#define CRAM_EX2_0 ex2_fake
#define CRAM_EX2_1 ex2_fake
#define CRAM_EX2_2 ex2_fake
#define CRAM_EX2_3 ex2_fake
#elif defined(CRAM_ABL_QUARTEREXP)
#define CRAM_EX2_0 ex2
#define CRAM_EX2_1 ex2_fake
#define CRAM_EX2_2 ex2_fake
#define CRAM_EX2_3 ex2_fake
#elif defined(CRAM_ABL_HALFEXP)
#define CRAM_EX2_0 ex2
#define CRAM_EX2_1 ex2_fake
#define CRAM_EX2_2 ex2
#define CRAM_EX2_3 ex2_fake
#elif defined(CRAM_ABL_3QEXP)
#define CRAM_EX2_0 ex2
#define CRAM_EX2_1 ex2
#define CRAM_EX2_2 ex2
#define CRAM_EX2_3 ex2_fake
#else
#define CRAM_EX2_0 ex2
#define CRAM_EX2_1 ex2
#define CRAM_EX2_2 ex2
#define CRAM_EX2_3 ex2
#endif
// packed fp32x2 arithmetic (FFMA2 / FADD2): two lanes per issue slot
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t r, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
template <int N>
__device__ __forceinline__ void setmaxnreg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void setmaxnreg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}

template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// ---- softmax of one K/V tile (128 keys) for one query row (= one thread) ----------------------------------------
// t = S*scale2 + bias[k] (one packed FFMA2 per two scores; this is where the credibility enters), P = 2^(t - m_ref)
// packed to T and written back over the first half of the S columns, in two 64-key halves with their own hand-off to
// the MMA warp (P V of the first half runs under the softmax of the second).
//
// m_ref is the reference the exponentials are taken against. O / l does not depend on it; it only has to keep the numbers
// in range, and bf16 / fp32 have 2^+-126 of it. So there is no running max in the hot loop at all: a half-tile is ONE
// pass (scores -> exponentials -> row sum -> pack), and its row sum is the overflow detector:
//   sum < 2^14        nothing to do (every p < 2^14: exact in bf16, inside the fp16 range);
//   sum >= 2^14       "soft": the NEXT tile starts from m_ref + log2(sum) (>= the true max), after the main loop has
//                     rescaled O and l -- lazy and delayed, off the hot path; fp16 P takes the exact path instead;
//   sum = inf / NaN   some t - m_ref exceeded 127: exact two-pass path for this half (half max, rescale, recompute from
//                     the scores, which are still intact in TMEM because P of a half is stored only after this check).
// The first tile of a row block takes its reference from the max of its first 32 keys.
template <typename T>
struct ExactAbove {
  static constexpr float value = 1.0e38f;  // bf16 P: only a non-finite sum needs the exact path
};
template <>
struct ExactAbove<__half> {
  static constexpr float value = 16384.0f;  // fp16 P saturates at 2^16
};
constexpr float kSoftAbove = 16384.0f;  // 2^14

// kDiag: the tile straddles the causal diagonal, columns > climit are masked.
template <bool kDiag>
__device__ __forceinline__ void score_chunk(const uint32_t (&raw)[32], const uint32_t bias, const int c, const uint64_t sc2,
                                            const int climit, float (&t)[32]) {
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    uint64_t bx, by;
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(bx), "=l"(by) : "r"(bias + (c * 32 + i) * 4));
    const uint64_t t01 = ffma2(pk2(__uint_as_float(raw[i + 0]), __uint_as_float(raw[i + 1])), sc2, bx);
    const uint64_t t23 = ffma2(pk2(__uint_as_float(raw[i + 2]), __uint_as_float(raw[i + 3])), sc2, by);
    upk2(t01, t[i + 0], t[i + 1]);
    upk2(t23, t[i + 2], t[i + 3]);
#ifndef CRAM_ABL_NODIAG  // timing ablation: no causal predicate on diagonal tiles (wrong results on purpose)
    if (kDiag) {
      const int col = c * 32 + i;
      t[i + 0] = (col + 0 <= climit) ? t[i + 0] : -INFINITY;
      t[i + 1] = (col + 1 <= climit) ? t[i + 1] : -INFINITY;
      t[i + 2] = (col + 2 <= climit) ? t[i + 2] : -INFINITY;
      t[i + 3] = (col + 3 <= climit) ? t[i + 3] : -INFINITY;
    }
#endif
  }
}

// max over the scores of chunks [c0, c1)
template <bool kDiag>
__device__ __noinline__ float chunks_max(const uint32_t s_addr, const uint32_t bias, const uint64_t sc2, const int climit,
                                            const int c0, const int c1) {
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll 1
  for (int c = c0; c < c1; ++c) {
    uint32_t raw[32];
    ptx::tmem_ld32(s_addr + c * 32, raw);
    ptx::tc_wait_ld_regs(raw);
    float t[32];
    score_chunk<kDiag>(raw, bias, c, sc2, climit, t);
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      mx0 = fmax3(mx0, t[i + 0], t[i + 1]);
      mx1 = fmax3(mx1, t[i + 2], t[i + 3]);
    }
  }
  return fmaxf(mx0, mx1);
}

// exact pass over chunks [c0, c1) against a known reference: P chunk c (16 packed columns) goes straight over S columns
// [16c, 16c+16), which belong to chunks already consumed. Returns the row sum.
template <typename T, bool kDiag>
__device__ __noinline__ float exact_pass(const uint32_t s_addr, const uint32_t bias, const uint64_t sc2, const int climit,
                                            const float m_ref, const int c0, const int c1) {
  const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
  const uint64_t negm = pk2(-m_use, -m_use);
  uint64_t la = pk2(0.f, 0.f), lb = pk2(0.f, 0.f);
#pragma unroll 1
  for (int c = c0; c < c1; ++c) {
    uint32_t raw[32];
    ptx::tmem_ld32(s_addr + c * 32, raw);
    ptx::tc_wait_ld_regs(raw);
    float t[32];
    score_chunk<kDiag>(raw, bias, c, sc2, climit, t);
    uint32_t pk[16];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      float e0, e1, e2, e3;
      upk2(fadd2(pk2(t[i + 0], t[i + 1]), negm), e0, e1);
      upk2(fadd2(pk2(t[i + 2], t[i + 3]), negm), e2, e3);
      const float p0 = CRAM_EX2_0(e0), p1 = CRAM_EX2_1(e1), p2 = CRAM_EX2_2(e2), p3 = CRAM_EX2_3(e3);
      la = fadd2(la, pk2(p0, p1));
      lb = fadd2(lb, pk2(p2, p3));
      pk[i / 2 + 0] = pack2<T>(p0, p1);
      pk[i / 2 + 1] = pack2<T>(p2, p3);
    }
    ptx::tmem_st16(s_addr + c * 16, pk);
  }
  float l0, l1, l2, l3;
  upk2(la, l0, l1);
  upk2(lb, l2, l3);
  return (l0 + l1) + (l2 + l3);
}

// O *= f, l *= f (f = 2^(m_old - m_new) per row): the correction step of the online softmax
// (l goes in and comes back by value: taken by reference it lived in local memory, in the hot loop too)
template <int DH>
__device__ __noinline__ float rescale_o(const uint32_t o_addr, const float f, const float l) {
  const uint64_t f2 = pk2(f, f);
#pragma unroll 1
  for (int c = 0; c < DH / 32; ++c) {
    uint32_t o[32];
    ptx::tmem_ld32(o_addr + c * 32, o);
    ptx::tc_wait_ld_regs(o);
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float a, b;
      upk2(fmul2(pk2(__uint_as_float(o[i]), __uint_as_float(o[i + 1])), f2), a, b);
      o[i] = __float_as_uint(a);
      o[i + 1] = __float_as_uint(b);
    }
    ptx::tmem_st32(o_addr + c * 32, o);
  }
  return l * f;
}

// One speculative pass over the 64 keys of half h (chunks 2h, 2h+1): P packed into pk, returns the row sum.
// `between` runs after the first chunk (the main loop uses it to hand the previous half over to the MMA warp once its
// tcgen05.st has had a chunk's worth of time to land).
template <typename T, bool kDiag, typename Between>
__device__ __forceinline__ float spec_half(const uint32_t s_addr, const uint32_t bias, const uint64_t sc2, const int climit,
                                           const float m_ref, const int h, uint32_t (&pk)[32], Between&& between) {
  const float m_spec = (m_ref == -INFINITY) ? 0.f : m_ref;
  const uint64_t negm = pk2(-m_spec, -m_spec);
  uint64_t la = pk2(0.f, 0.f), lb = pk2(0.f, 0.f);
  uint32_t raw[2][32];
  // TMEM -> registers in 32-column chunks; the load of the second chunk is in flight while the first is processed.
  ptx::tmem_ld32(s_addr + (2 * h) * 32, raw[0]);
#pragma unroll
  for (int cc = 0; cc < 2; ++cc) {
    ptx::tc_wait_ld_regs(raw[cc]);
    if (cc == 0) ptx::tmem_ld32(s_addr + (2 * h + 1) * 32, raw[1]);
    float t[32];
    score_chunk<kDiag>(raw[cc], bias, 2 * h + cc, sc2, climit, t);
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      float e0, e1, e2, e3;
      upk2(fadd2(pk2(t[i + 0], t[i + 1]), negm), e0, e1);
      upk2(fadd2(pk2(t[i + 2], t[i + 3]), negm), e2, e3);
      const float p0 = CRAM_EX2_0(e0), p1 = CRAM_EX2_1(e1), p2 = CRAM_EX2_2(e2), p3 = CRAM_EX2_3(e3);
      la = fadd2(la, pk2(p0, p1));
      lb = fadd2(lb, pk2(p2, p3));
      pk[cc * 16 + i / 2 + 0] = pack2<T>(p0, p1);
      pk[cc * 16 + i / 2 + 1] = pack2<T>(p2, p3);
    }
    if (cc == 0) between();
  }
  float l0, l1, l2, l3;
  upk2(la, l0, l1);
  upk2(lb, l2, l3);
  return (l0 + l1) + (l2 + l3);
}

// ---- MMA issue helpers (compile-time K steps; see ptx::mma_ss_imm) ---------------------------------
constexpr uint32_t kDescHi = ptx::umma_desc_hi(1024);  // SBO = 8 rows x 128 B, version 1, SWIZZLE_128B
// byte offset (>> 4) of K-step kk inside a K-major 128B-swizzled [128 x DH] tile
constexpr int kmajor_off16(int kk, int slab_bytes) { return ((kk >> 2) * slab_bytes + (kk & 3) * 32) >> 4; }

template <int... KK>
__device__ __forceinline__ void issue_qk_steps(uint32_t d_tmem, uint32_t q_lo, uint32_t k_lo, uint32_t idesc,
                                               std::integer_sequence<int, KK...>) {
  (ptx::mma_ss_imm<kDescHi, 0, kmajor_off16(KK, kBlockM * 128), kmajor_off16(KK, kBlockN * 128)>(d_tmem, q_lo, k_lo, idesc,
                                                                                                    KK > 0 ? 1u : 0u),
   ...);
}
// O += P V for key steps KK (16 keys each): P step kk sits at TMEM column offset 8*kk (packed 16-bit),
// V step kk 16 rows x 128 B further down the MN-major tile.
template <int kFirst, int... KK>
__device__ __forceinline__ void issue_pv_steps(uint32_t o_tmem, uint32_t p_tmem, uint32_t v_lo, uint32_t idesc,
                                               bool acc_first, std::integer_sequence<int, KK...>) {
  (ptx::mma_ts_imm<kDescHi, 0, (kFirst + KK) * 8, (kFirst + KK) * 128>(o_tmem, p_tmem, v_lo, idesc,
                                                                       (acc_first || KK > 0) ? 1u : 0u),
   ...);
}

// ---- work items ------------------------------------------------------------------------------------------------
// One item = 256 query rows (a pair of 128-row tiles) of one (sequence, head). Item order: heads fastest (the 4 / 7 / 8
// query heads of a KV head run side by side and share its tiles in L2), then the row-block pairs of ONE sequence, heaviest
// (latest rows) first, then the next sequence: all readers of a sequence's K / V (8.4 MB at 2k tokens) run within ~2 waves,
// so the tiles come from DRAM once (pair-major order re-read them ~4x: 1079 MB per launch against 617 MB).
struct alignas(16) Item {
  int idx;  // position in the work list; >= n_items marks the end of the list
  int head, kv_head, seq, pair, q0, len_q, k0, len_k, diag, nt[2], n_max;
  // Dead-tile skipping: on a head that carries the bias, a tile whose keys all have zero credibility (log = -inf)
  // contributes P = 0 to every row -- it is never loaded, multiplied or exponentiated. nt / n_max then count LIVE tiles
  // and the roles walk them through LiveTiles (virtual index v -> key tile j); dense = 1: every tile is live, v == j.
  unsigned live_lo, live_hi;
  int dense;
};
static_assert(sizeof(Item) == 64, "one scheduler-ring slot");

struct LiveTiles {
  unsigned long long m;
  int v;
  bool dense;
  __device__ __forceinline__ explicit LiveTiles(const Item& w)
      : m(((unsigned long long)w.live_hi << 32) | w.live_lo), v(0), dense(w.dense != 0) {}
  // key tile of the next live tile (callers stop after nt / n_max calls)
  __device__ __forceinline__ int next() {
    if (dense) return v++;
    const int j = __ffsll((long long)m) - 1;
    m &= m - 1;
    ++v;
    return j;
  }
};
__device__ __forceinline__ unsigned long long low_bits(int n) { return n >= 64 ? ~0ull : ((1ull << n) - 1ull); }

__device__ __forceinline__ Item decode_item(const cram_attn_prefill_params& p, const int num_pairs, const int idx) {
  Item w;
  w.idx = idx;
  w.head = idx % p.n_heads;
  const int rest = idx / p.n_heads;
  w.pair = num_pairs - 1 - rest % num_pairs;
  w.seq = rest / num_pairs;
  w.q0 = p.cu_seqlens_q[w.seq];
  w.len_q = p.cu_seqlens_q[w.seq + 1] - w.q0;
  const int32_t* cuk = p.cu_seqlens_k ? p.cu_seqlens_k : p.cu_seqlens_q;
  w.k0 = cuk[w.seq];
  w.len_k = cuk[w.seq + 1] - w.k0;
  w.diag = w.len_k - w.len_q;  // query i sees keys <= i + diag
  w.kv_head = w.head / (p.n_heads / p.n_kv_heads);
#pragma unroll
  for (int s = 0; s < kStages; ++s) {  // number of 128-key tiles query stage s visits
    const int mt = w.pair * kStages + s;
    int kv_end = w.len_k;
    if (p.causal) kv_end = min(w.len_k, mt * kBlockM + kBlockM + w.diag);
    w.nt[s] = (mt * kBlockM < w.len_q && kv_end > 0) ? (kv_end + kBlockN - 1) / kBlockN : 0;
  }
  w.n_max = max(w.nt[0], w.nt[1]);
  w.live_lo = w.live_hi = 0u;
  w.dense = 1;
  return w;
}

// Pre-pass of dead-tile skipping: one warp per (sequence, 128-key tile) ORs the tile's bit into live[seq] when any of its
// keys is unbiased (>= bias_len) or has a log-credibility above -inf. 2 bytes per key read once per launch, against the
// K / V tile loads, MMAs and exponentials of every biased (head, row block) that would have visited a dead tile.
__global__ void live_tiles_kernel(const cram_attn_prefill_params p, unsigned long long* __restrict__ live) {
  const int tiles = (p.max_seqlen_k + kBlockN - 1) / kBlockN;
  const int wid = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  const int seq = wid / tiles, j = wid % tiles;
  if (seq >= p.n_seq) return;
  const int32_t* cuk = p.cu_seqlens_k ? p.cu_seqlens_k : p.cu_seqlens_q;
  const int k0 = cuk[seq], len_k = cuk[seq + 1] - k0;
  if (j * kBlockN >= len_k) return;
  const int bias_len = p.bias_len ? p.bias_len[seq] : len_k;
  bool any = false;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int key = j * kBlockN + lane * 4 + e;
    if (key < len_k) any |= key >= bias_len || !(load_logcred(p.logcred, (int64_t)k0 + key, p.logcred_dtype) == -INFINITY);
  }
  if (__any_sync(0xffffffffu, any) && lane == 0) atomicOr(&live[seq], 1ull << j);
}

// PERSISTENT kernel: one CTA per SM walks a dynamic queue of work items (atomic counter; the first item of a CTA is its
// block index). TMEM, the barriers and the K / V / bias rings live across items: the producer runs into the next item's
// Q / K / V while the softmax warps finish the current one, the MMA warp issues the next item's first Q K^T right behind
// the last P V of the current one, and the epilogue of item i overlaps the first tensor work of item i+1. All ring / phase counters are running counts over the CTA's stream of
// tiles; every role derives them from the same item sequence.
template <typename T, int DH>
__global__ void __launch_bounds__(kThreads, 1)
attn_prefill_sm100_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                          const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_o,
                          const cram_attn_prefill_params p,
                          const int num_pairs, const int n_items, unsigned* __restrict__ work_counter,
                          const unsigned long long* __restrict__ live_masks, const int mask_tiles) {
  using S = Smem<DH>;
  constexpr int kSlabs = DH / 64;
  constexpr uint32_t kAbFmt = DType<T>::code == CRAM_BF16 ? 1u : 0u;
  constexpr uint32_t kIdescQK = ptx::umma_idesc_f16(kAbFmt, kBlockM, kBlockN, 0);
  constexpr uint32_t kIdescPV = ptx::umma_idesc_f16(kAbFmt, kBlockM, DH, 1);

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((ptx::smem_u32(smem) & 1023u) != 0u) __trap();  // the 128B-swizzled TMA tiles need a 1024-byte aligned base
  Barriers* bar = reinterpret_cast<Barriers*>(smem + S::kBar);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + S::kTmemPtr);
  float* s_bias = reinterpret_cast<float*>(smem + S::kBias);
  Item* s_items = reinterpret_cast<Item*>(smem + S::kSched);
  // copy a ring slot to registers / from registers (4 x 16 bytes)
  auto read_slot = [&](int slot) {
    Item w;
    const uint4* src = reinterpret_cast<const uint4*>(&s_items[slot]);
    uint4* dst = reinterpret_cast<uint4*>(&w);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = src[i];
    return w;
  };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 8 && lane == 0) {
    ptx::prefetch_tmap(&tmap_q);
    ptx::prefetch_tmap(&tmap_k);
    ptx::prefetch_tmap(&tmap_v);
    ptx::prefetch_tmap(&tmap_o);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&bar->q_full[i], 1);
      ptx::mbar_init(&bar->q_empty[i], 1);
      ptx::mbar_init(&bar->o_final[i], 1);
      ptx::mbar_init(&bar->s_full[i], 1);
      ptx::mbar_init(&bar->p_full[i][0], kBlockM);
      ptx::mbar_init(&bar->p_full[i][1], kBlockM);
      ptx::mbar_init(&bar->o_half[i], 1);
    }
    for (int i = 0; i < kKV; ++i) {
      ptx::mbar_init(&bar->k_full[i], 1);
      ptx::mbar_init(&bar->k_empty[i], 1);
      ptx::mbar_init(&bar->v_full[i], 1);
      ptx::mbar_init(&bar->v_empty[i], 1);
    }
    for (int i = 0; i < kBiasRing; ++i) ptx::mbar_init(&bar->bias_full[i], 1);
    for (int i = 0; i < kSched; ++i) {
      ptx::mbar_init(&bar->item_full[i], 1);
      ptx::mbar_init(&bar->item_empty[i], kItemConsumers);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 9) ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // Every consumer role walks the same item sequence: entry `it` of the scheduler ring. The items arrive DECODED (the
  // producer did the cu_seqlens loads and the divisions while it was running ahead), one 64-byte broadcast read per warp.
  auto next_item = [&](int it) -> Item {
    const int slot = it % kSched;
    ptx::mbar_wait(&bar->item_full[slot], (it / kSched) & 1);
    const Item w = read_slot(slot);
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(&bar->item_empty[slot]);
    return w;
  };

  if (warp >= 8) {
    setmaxnreg_dec<104>();
    if (warp == 8) {
      // ============================== scheduler + producer =======================================
      // Per item: K[0], Q of both stages (once the previous item's last Q K^T of that stage has completed), K[1], then per
      // tile j: V[j], K[j+2]. gt = running tile count: ring slot gt % kKV, phase gt / kKV. A slot is re-filled as soon as
      // the last MMA that reads it has completed (k_empty / v_empty are tcgen05.commit arrivals).
      int gt = 0, qi[2] = {0, 0};
      // publish entry `it` of the scheduler ring
      auto publish = [&](int it, const Item& w) {
        const int slot = it % kSched;
        if (it >= kSched) ptx::mbar_wait(&bar->item_empty[slot], ((it / kSched) & 1) ^ 1);
        const uint4* src = reinterpret_cast<const uint4*>(&w);
        uint4* dst = reinterpret_cast<uint4*>(&s_items[slot]);
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = src[i];
        ptx::mbar_arrive(&bar->item_full[slot]);
      };
      // Next work item that has query rows (items of a ragged batch whose row block lies past the end of their sequence
      // are never published: every published entry is either real work or the end marker, which is what lets the MMA
      // warp read exactly one entry ahead). The first item of a CTA is its block index, the rest come from the queue.
      bool first = true;
#ifdef CRAM_ABL_STATIC_QUEUE
      int static_next = (int)blockIdx.x;
#endif
      auto take = [&]() -> Item {
        for (;;) {
#ifdef CRAM_ABL_STATIC_QUEUE
          static_next += first ? 0 : (int)gridDim.x;
          const int i = first ? (int)blockIdx.x : static_next;
#else
          const int i = first ? (int)blockIdx.x : (int)(atomicAdd(work_counter, 1u) + gridDim.x);
#endif
          first = false;
          if (i >= n_items) {
            Item e = {};
            e.idx = n_items;
            return e;
          }
          Item w = decode_item(p, num_pairs, i);
          if (w.pair * kStages * kBlockM >= w.len_q) continue;
          // (a sequence longer than the caller's max_seqlen_k has tiles the pre-pass never looked at: it runs dense)
          if (w.n_max <= mask_tiles && ((p.head_mask >> w.head) & 1ull)) {
            const unsigned long long m = live_masks[w.seq] & low_bits(w.n_max);
            if (m != low_bits(w.n_max)) {
              w.nt[0] = __popcll(m & low_bits(w.nt[0]));
              w.nt[1] = __popcll(m & low_bits(w.nt[1]));
              w.n_max = max(w.nt[0], w.nt[1]);
              w.live_lo = (unsigned)m;
              w.live_hi = (unsigned)(m >> 32);
              w.dense = 0;
            }
          }
          return w;
        }
      };
      // lane 0 runs the queue and the loads; the whole warp only meets at the end
      if (lane == 0) {
        CRAM_EV_DECL;
        Item w = take();
        publish(0, w);
        for (int it = 0; w.idx < n_items; ++it) {
          CRAM_EV(3, it, 0, 0);
          // The NEXT item is taken LATE: three tiles before the end of this item's loads (short items: behind their
          // first loads). Taking costs ~3800 cycles of dependent global latency (a contended atomic, cu_seqlens, the mask
          // word); there the producer is between two ring-slot waits with a tile of slack, while at the top of the loop
          // (v5.4) it sat in front of the next item's Q / K[0] at every item boundary (event log:
          // profiles/r2_attn_ablation_and_trace.md). It is also the latest point that still has the next item's record and first
          // loads out before the consumers get there -- and the later an item is claimed from the dynamic queue, the
          // better the balance at the tail (claiming one more item ahead cost 7 % at 16 x 2048 causal).
          Item w_next;
          bool taken = false;
          const int take_at = max(w.n_max - 3, 0);
          if (w.n_max > 0) {  // n_max = 0: rows that see no key at all (len_k < len_q); the softmax warps write zeros
          LiveTiles lk(w), lv(w);  // K runs two live tiles ahead of V
          auto load_k = [&](int vt) {
            const int j = lk.next();
            const int g = gt + vt, st = g % kKV;
            ptx::mbar_wait(&bar->k_empty[st], ((g / kKV) & 1) ^ 1);
            ptx::mbar_expect_tx(&bar->k_full[st], S::kTileBytes);
#pragma unroll
            for (int sl = 0; sl < kSlabs; ++sl)
              ptx::tma_load_2d_hint(smem + S::kK + st * S::kTileBytes + sl * S::kSlabBytes, &tmap_k, &bar->k_full[st],
                                    w.kv_head * DH + sl * 64, w.k0 + j * kBlockN, ptx::kL2EvictLast);
          };
          // K[0] goes out before the Q tiles: it only needs a free ring slot, while a Q tile has to wait for the previous
          // item's last Q K^T of its stage.
          load_k(0);
#pragma unroll
          for (int s = 0; s < kStages; ++s) {
            if (w.nt[s] == 0) continue;
            if (qi[s] > 0) ptx::mbar_wait(&bar->q_empty[s], (qi[s] - 1) & 1);
            ++qi[s];
            ptx::mbar_expect_tx(&bar->q_full[s], S::kQTileBytes);
#pragma unroll
            for (int sl = 0; sl < kSlabs; ++sl)
              ptx::tma_load_2d_hint(smem + S::kQ + s * S::kQTileBytes + sl * S::kQSlabBytes, &tmap_q, &bar->q_full[s],
                                    w.head * DH + sl * 64, w.q0 + (w.pair * kStages + s) * kBlockM, ptx::kL2EvictFirst);
          }
          if (w.n_max > 1) load_k(1);
          CRAM_EV(3, it, 0, 2);
          for (int vt = 0; vt < w.n_max; ++vt) {
            const int j = lv.next();
            const int g = gt + vt, st = g % kKV;
            ptx::mbar_wait(&bar->v_empty[st], ((g / kKV) & 1) ^ 1);
            ptx::mbar_expect_tx(&bar->v_full[st], S::kTileBytes);
#pragma unroll
            for (int sl = 0; sl < kSlabs; ++sl)
              ptx::tma_load_2d_hint(smem + S::kV + st * S::kTileBytes + sl * S::kSlabBytes, &tmap_v, &bar->v_full[st],
                                    w.kv_head * DH + sl * 64, w.k0 + j * kBlockN, ptx::kL2EvictLast);
            if (vt + 2 < w.n_max) load_k(vt + 2);
            if (vt == take_at) {
              CRAM_EV(3, it, vt, 1);
              w_next = take();
              publish(it + 1, w_next);
              taken = true;
            }
          }
          }
          if (!taken) {
            w_next = take();
            publish(it + 1, w_next);
          }
          CRAM_EV(3, it, w.n_max, 3);
          gt += w.n_max;
          w = w_next;
        }
      }
      __syncwarp();
    } else if (warp == 10) {
      // ============================== bias vectors ===============================================
      // Its own warp (on a different sub-partition than the TMA producer): the log-credibility loads are dependent
      // global loads, and behind the TMA issue they kept the producer from running ahead. The vector of tile g
      // (already x log2e, -inf past the key length) lives in a ring of 4 until every P V[g] is done; tile g is
      // written once the V slot of tile g-2 has been seen free (the producer's own condition for V[g], so the phase
      // parity cannot alias: that barrier's next completion needs this very vector) -- two tiles of slack.
      int gt = 0;
      for (int it = 0;; ++it) {
        const Item w = next_item(it);
        if (w.idx >= n_items) break;
        const bool biased = p.logcred != nullptr && ((p.head_mask >> w.head) & 1ull);
        const int bias_len = p.bias_len ? p.bias_len[w.seq] : w.len_k;
        LiveTiles lt(w);
        for (int vt = 0; vt < w.n_max; ++vt) {
          const int j = lt.next();
          const int g = gt + vt;
          if (g >= kKV) ptx::mbar_wait(&bar->v_empty[g % kKV], ((g / kKV) & 1) ^ 1);
          float4 b4;
          float* bp = reinterpret_cast<float*>(&b4);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int key = j * kBlockN + lane * 4 + e;
            float v = 0.f;
            if (key >= w.len_k) v = -INFINITY;
            else if (biased && key < bias_len) v = load_logcred(p.logcred, (int64_t)w.k0 + key, p.logcred_dtype) * kLog2e;
            bp[e] = v;
          }
          reinterpret_cast<float4*>(s_bias + (g % kBiasRing) * kBlockN)[lane] = b4;
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&bar->bias_full[g % kBiasRing]);
        }
        gt += w.n_max;
      }
    } else if (warp == 9 && ptx::elect_one_sync()) {
      // ============================== MMA issuer (one elected lane) ==============================
      // Per query stage s the chain is Q K^T(j) -> softmax(j) -> P V(j) -> Q K^T(j+1): S is single-buffered and P
      // aliases it, the tensor pipe executes in order, so Q K^T(j+1) issued behind P V(j) cannot overtake it. While
      // stage s waits for its softmax, the pipe runs the other stage's P V + Q K^T (1024 cycles at d = 128). Behind the
      // last P V of an item comes the first Q K^T of the NEXT item (both stages).
      const uint32_t q_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + S::kQ), 16);
      const uint32_t k_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + S::kK), 16);
      const uint32_t v_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + S::kV), S::kSlabBytes);
      constexpr uint32_t kQTile16 = S::kQTileBytes >> 4;
      constexpr uint32_t kTile16 = S::kTileBytes >> 4;
      int it = 0, tc[2] = {0, 0}, qc[2] = {0, 0};
      // next item of the sequence that has tiles (this lane is the warp's only consumer of the scheduler ring)
      auto fetch = [&](Item& w) -> bool {
        const int slot = it % kSched;
        ptx::mbar_wait(&bar->item_full[slot], (it / kSched) & 1);
        w = read_slot(slot);
        ptx::mbar_arrive(&bar->item_empty[slot]);
        ++it;
        return w.idx < n_items;  // an item without tiles (n_max = 0) simply passes through the loops below
      };
      // Q K^T of tile j of item w (whose first tile is global tile g0) for stage s
      auto issue_qk = [&](int s, const Item& w, int g0, int j) {
        const int g = g0 + j, st = g % kKV;
        if (j == 0) {
          ptx::mbar_wait(&bar->q_full[s], qc[s] & 1);
          ++qc[s];
        }
        ptx::mbar_wait(&bar->k_full[st], (g / kKV) & 1);
        ptx::tc_fence_after();
        issue_qk_steps(tmem_base + s * 128, q_lo + s * kQTile16, k_lo + st * kTile16, kIdescQK,
                       std::make_integer_sequence<int, DH / 16>{});
        ptx::tc_commit(&bar->s_full[s]);
        if (s == 1 || j >= w.nt[1]) ptx::tc_commit(&bar->k_empty[st]);  // last stage to read K[j]
        if (j == w.nt[s] - 1) ptx::tc_commit(&bar->q_empty[s]);         // last Q K^T that reads this item's Q tile
      };
      Item cur, nxt;
      CRAM_EV_DECL;
      int ev_item = 0;
      bool have = fetch(cur);
      int gt = 0;
      if (have) {
#pragma unroll
        for (int s = 0; s < kStages; ++s)
          if (cur.nt[s] > 0) issue_qk(s, cur, gt, 0);
      }
      while (have) {
        const int gt_next = gt + cur.n_max;
        for (int j = 0; j < cur.n_max; ++j) {
          const int g = gt + j, st = g % kKV;
#pragma unroll
          for (int s = 0; s < kStages; ++s) {
            const int nt = cur.nt[s];
            if (j >= nt) continue;
            ptx::mbar_wait(&bar->v_full[st], (g / kKV) & 1);
            ptx::mbar_wait(&bar->p_full[s][0], tc[s] & 1);
            ptx::tc_fence_after();
            issue_pv_steps<0>(tmem_base + 256 + s * 128, tmem_base + s * 128, v_lo + st * kTile16, kIdescPV, j > 0,
                              std::make_integer_sequence<int, kBlockN / 32>{});
            ptx::tc_commit(&bar->o_half[s]);
            ptx::mbar_wait(&bar->p_full[s][1], tc[s] & 1);
            ptx::tc_fence_after();
            issue_pv_steps<kBlockN / 32>(tmem_base + 256 + s * 128, tmem_base + s * 128, v_lo + st * kTile16, kIdescPV, true,
                                         std::make_integer_sequence<int, kBlockN / 32>{});
            if (j == nt - 1) ptx::tc_commit(&bar->o_final[s]);
            if (s == 1 || j >= cur.nt[1]) ptx::tc_commit(&bar->v_empty[st]);  // last stage to read V[j]
            ++tc[s];
            if (j + 1 < nt) issue_qk(s, cur, gt, j + 1);
            CRAM_EV(2, ev_item, j, 4 + s);
          }
        }
        // The next item's first Q K^T of both stages goes out HERE, behind the last P V of the whole item, not behind the
        // last P V of each stage: chaining stage 0's (it finishes a tile earlier in a causal item) put 512 tensor cycles
        // in front of stage 1's last P V + Q K^T, the critical path of the item's end. Measured at 16 x 2048 causal:
        // 802 TFLOP/s chained (blocking on the next Q / K[0]), 841 chained only when both had landed, 896-904 not chained.
        // (The next item is also only READ here: its entry carries the live-tile mask, a dependent global load of the
        // producer's, and nothing above needs it.)
        const bool have_next = fetch(nxt);
        CRAM_EV(2, ev_item, 0, 6);
        if (have_next) {
#pragma unroll
          for (int s = 0; s < kStages; ++s)
            if (nxt.nt[s] > 0) issue_qk(s, nxt, gt_next, 0);
        }
        CRAM_EV(2, ev_item, 0, 7);
#ifdef CRAM_ATTN_TRACE
        ++ev_item;
#endif
        gt = gt_next;
        cur = nxt;
        have = have_next;
      }
    }
    // warp 11 and the other lanes of warps 8 and 9 idle: kItemConsumers counts the MMA lane, the bias warp and the eight
    // softmax warps
  } else {
    // ============================== softmax / correction / epilogue ================================
    // 200 / 104 registers (softmax / the other four warps): 256 x (200 - 168) = 128 x (168 - 104). The softmax code does not
    // spill at 192; at 224 / 56 the producer's work items lived in local memory, and 200 / 104 measured 0.5-1 % faster
    // (profiles/ab_regs.txt).
    setmaxnreg_inc<200>();
    const int s = warp >> 2;                      // query stage of this warpgroup
    const int row = (warp & 3) * 32 + lane;       // row inside the tile == TMEM lane
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t s_base = tmem_base + lane_off + s * 128;
    const uint32_t o_addr = tmem_base + lane_off + 256 + s * 128;
    const float scale2 = p.scale * kLog2e;
    const uint32_t bias_addr = ptx::smem_u32(s_bias);
    const uint64_t sc2 = pk2(scale2, scale2);
    int gt = 0, tc = 0, ic = 0;  // running counts: tiles of the CTA, tiles of this stage, items of this stage

    // hand a half of P over to the MMA warp: its tcgen05.st must have completed
    auto release = [&](int h) {
      ptx::tc_wait_st();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&bar->p_full[s][h]);
    };

    CRAM_EV_DECL;
    const bool ev_on = ((warp & 3) == 0 || (warp & 3) == 3) && lane == 0;  // rows 0-31 and 96-127 of the stage's tile
    const int ev_slot = s + ((warp & 3) == 3 ? 4 : 0);
    for (int it = 0;; ++it) {
      const Item w = next_item(it);
      if (w.idx >= n_items) break;
      if (ev_on) CRAM_EV(ev_slot, it, w.nt[s], 8);
      const int mt = w.pair * kStages + s;
      const int qrow = mt * kBlockM + row;          // row inside the sequence
      const int nt = w.nt[s];
      float m_ref = -INFINITY, m_next = -INFINITY, l = 0.f;

      LiveTiles lt(w);
      for (int vt = 0; vt < nt; ++vt, ++tc) {
        const int j = lt.next();  // key tile (causal geometry); vt = position in this item's stream of live tiles
        const int g = gt + vt, bs = g % kBiasRing;
        ptx::mbar_wait(&bar->bias_full[bs], (g / kBiasRing) & 1);
        ptx::mbar_wait(&bar->s_full[s], tc & 1);
        ptx::tc_fence_after();
        if (ev_on) CRAM_EV(ev_slot, it, vt, 9);
        const bool diag_tile = p.causal && (j * kBlockN + kBlockN - 1 > mt * kBlockM + w.diag);
        const int climit = qrow + w.diag - j * kBlockN;  // columns <= climit are visible
        const uint32_t b_addr = bias_addr + bs * kBlockN * 4;

        auto tile = [&](auto diag_tag) {
          constexpr bool kDiag = decltype(diag_tag)::value;
          if (vt == 0) {
            // reference of a fresh row block: the max of its first 32 keys (of the first live tile; a row that sees none of them yet
            // keeps m_ref = -inf and takes the exact path at its first visible key)
            m_ref = chunks_max<kDiag>(s_base, b_addr, sc2, climit, 0, 1);
            m_next = m_ref;
          } else {
            // delayed lazy rescale: a row whose previous tile asked for a larger reference moves now. s_full(j) has been
            // observed and Q K^T(j) was issued behind P V(j-1), so O holds every product up to tile j-1.
            const bool pend = m_next > m_ref;
            if (__any_sync(0xffffffffu, pend)) {
              l = rescale_o<DH>(o_addr, pend ? ex2(m_ref - m_next) : 1.0f, l);
              m_ref = m_next;
            }
          }
          // (the diagonal instantiation runs once per item and is cold in the instruction cache every time: its two halves
          // share one copy of the code)
          constexpr int kUnrollH = kDiag ? 1 : 2;
#pragma unroll kUnrollH
          for (int h = 0; h < 2; ++h) {
            uint32_t pk[32];
            const float sum = spec_half<T, kDiag>(s_base, b_addr, sc2, climit, m_ref, h, pk, [&] {
              if (h == 1) release(0);  // P of the first half was stored one chunk ago
            });
            // a row with no reference yet (every key so far masked out or of zero credibility) that now sees one
            const bool exact = !(sum < ExactAbove<T>::value) || (m_ref == -INFINITY && sum > 0.f);
            if (__any_sync(0xffffffffu, exact)) {
              // exact path (rare). O must hold every product issued so far before it is rescaled.
              const float mx = chunks_max<kDiag>(s_base, b_addr, sc2, climit, 2 * h, 2 * h + 2);
              const float m_new = exact ? fmaxf(mx, m_ref) : m_ref;
              if (h == 1) {
                ptx::mbar_wait(&bar->o_half[s], tc & 1);
                ptx::tc_fence_after();
              }
              if (vt > 0 || h == 1) l = rescale_o<DH>(o_addr, (m_new > m_ref) ? ex2(m_ref - m_new) : 1.0f, l);
              m_ref = m_new;
              m_next = fmaxf(m_next, m_ref);
              l += exact_pass<T, kDiag>(s_base, b_addr, sc2, climit, m_ref, 2 * h, 2 * h + 2);
            } else {
              ptx::tmem_st32(s_base + h * 32, pk);
              l += sum;
              if (!(sum < kSoftAbove)) m_next = fmaxf(m_next, m_ref + __log2f(sum));  // >= the true max of this half
            }
            if (ev_on) CRAM_EV(ev_slot, it, vt, 15 - h);  // 15: half 0 done, 14: half 1 done (before its hand-over)
            if (h == 1) release(1);
          }
        };
        if (diag_tile) tile(std::true_type{});
        else tile(std::false_type{});
        if (ev_on) CRAM_EV(ev_slot, it, vt, 10);
      }
      gt += w.n_max;

      // ---- epilogue: O / l -> global ---------------------------------------------------------------
      const bool row_ok = qrow < w.len_q;
      T* out = reinterpret_cast<T*>(p.out) + (int64_t)(w.q0 + qrow) * p.out_row_stride + (int64_t)w.head * DH;
      if (nt > 0) {
        ptx::mbar_wait(&bar->o_final[s], ic & 1);  // committed after the last P V of this stage and item
        ++ic;
        ptx::tc_fence_after();
        if (ev_on) CRAM_EV(ev_slot, it, 0, 11);
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        // Full 32-row slices (every warp except at the ragged end of a sequence) leave through shared memory and a TMA
        // store: one thread per row writing 16 bytes at a time straight to global memory is 32 separate 128-byte lines
        // per store instruction -- 2048 L1 wavefronts per stage and item, 2200-3100 cycles in the event log, in the LSU
        // path the other stage's bias loads need. Staged, it is 64 conflict-free shared-memory wavefronts per warp and
        // the copy engine does the rest while the warp goes on. (Boxes of 32 columns, double-buffered, measured the same.)
        const int slice_row0 = mt * kBlockM + (warp & 3) * 32;
        if (slice_row0 + 32 <= w.len_q) {
          // rounds of 64 columns: one 4 KB box per warp (32 rows x 128 bytes, 128B swizzle), one TMA store per round
          uint8_t* box = smem + S::kOut + warp * 2 * S::kOutBoxBytes;
          const uint32_t box_row = ptx::smem_u32(box) + lane * 128;
          const uint64_t inv2 = pk2(inv, inv);
          // (not unrolled, packed multiplies: this code runs once per item and is cold in the instruction cache -- its size
          // costs more than the TMEM latency a prefetch across rounds would hide)
#pragma unroll 1
          for (int r = 0; r < DH / 64; ++r) {
            uint32_t oa[32], ob[32];
            ptx::tmem_ld32(o_addr + r * 64, oa);
            ptx::tmem_ld32(o_addr + r * 64 + 32, ob);
            ptx::tc_wait_ld_regs(oa);
            ptx::tc_wait_ld_regs(ob);
            if (ev_on) CRAM_EV(ev_slot, it, r, 13);
            uint4 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint32_t(&o)[32] = i < 4 ? oa : ob;
              const int k = (i & 3) * 8;
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; e += 2)
                upk2(fmul2(pk2(__uint_as_float(o[k + e]), __uint_as_float(o[k + e + 1])), inv2), f[e], f[e + 1]);
              v[i].x = pack2<T>(f[0], f[1]);
              v[i].y = pack2<T>(f[2], f[3]);
              v[i].z = pack2<T>(f[4], f[5]);
              v[i].w = pack2<T>(f[6], f[7]);
            }
            // the box is free once the previous store has finished READING it (the issuing lane knows)
            if (lane == 0) ptx::tma_store_wait_read<0>();
            __syncwarp();
            if (ev_on) CRAM_EV(ev_slot, it, r, 14);
#pragma unroll
            for (int i = 0; i < 8; ++i)  // 128B swizzle: 16-byte chunk index ^= row & 7
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(box_row + ((i ^ (lane & 7)) << 4)), "r"(v[i].x),
                           "r"(v[i].y), "r"(v[i].z), "r"(v[i].w)
                           : "memory");
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              ptx::tma_store_2d(&tmap_o, box, w.head * DH + r * 64, w.q0 + slice_row0);
              ptx::tma_store_commit();
            }
          }
        } else {
#pragma unroll
        for (int c = 0; c < DH / 32; ++c) {
          uint32_t o[32];
          ptx::tmem_ld32(o_addr + c * 32, o);
          ptx::tc_wait_ld();
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              uint4 v;
              v.x = pack2<T>(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
              v.y = pack2<T>(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
              v.z = pack2<T>(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
              v.w = pack2<T>(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
              __stcs(reinterpret_cast<uint4*>(out + c * 32 + i), v);  // written once, read by the next kernel
            }
          }
        }
        }
        if (p.lse && row_ok)
          p.lse[(int64_t)w.head * p.total_q + w.q0 + qrow] = l > 0.f ? (m_ref + log2f(l)) * 0.6931471805599453f : -INFINITY;
      } else if (row_ok) {
#pragma unroll
        for (int c = 0; c < DH; c += 8) *reinterpret_cast<uint4*>(out + c) = make_uint4(0, 0, 0, 0);
        if (p.lse) p.lse[(int64_t)w.head * p.total_q + w.q0 + qrow] = -INFINITY;
      }
      if (ev_on) CRAM_EV(ev_slot, it, 0, 12);
    }
  }

  if (warp < 8 && lane == 0) ptx::tma_store_wait<0>();  // this lane's output boxes have reached global memory
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 9) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  return fn;
}

// rows x cols view of packed token rows; box = box_rows rows x 64 elements, 128-byte swizzle.
int make_tmap(CUtensorMap* m, const void* base, int dtype, int64_t cols, int64_t rows, int64_t row_stride_elems,
              int box_rows, int box_cols = 64, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return CRAM_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)row_stride_elems * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, dtype == CRAM_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (base %p cols %lld rows %lld stride %lld)", (int)r, base,
              (long long)cols, (long long)rows, (long long)row_stride_elems);
    return CRAM_ERR_CUDA;
  }
  return CRAM_OK;
}

template <typename T, int DH>
int launch_t(const cram_attn_prefill_params& p, cudaStream_t stream) {
  CUtensorMap tq, tk, tv, to;
  int rc;
  if ((rc = make_tmap(&tq, p.q, p.dtype, (int64_t)p.n_heads * DH, p.total_q, p.q_row_stride, kBlockM)) != CRAM_OK) return rc;
  if ((rc = make_tmap(&tk, p.k, p.dtype, (int64_t)p.n_kv_heads * DH, p.total_k, p.k_row_stride, kBlockN)) != CRAM_OK) return rc;
  if ((rc = make_tmap(&tv, p.v, p.dtype, (int64_t)p.n_kv_heads * DH, p.total_k, p.v_row_stride, kBlockN)) != CRAM_OK) return rc;
  if ((rc = make_tmap(&to, p.out, p.dtype, (int64_t)p.n_heads * DH, p.total_q, p.out_row_stride, 32)) != CRAM_OK) return rc;
  const int num_pairs = (p.max_seqlen_q + kStages * kBlockM - 1) / (kStages * kBlockM);
  const long long n_items = (long long)num_pairs * p.n_seq * p.n_heads;
  if (n_items == 0) return CRAM_OK;
  CRAM_REQUIRE(n_items < (1ll << 30), "cram_attn_prefill: too many work items");
  auto kern = attn_prefill_sm100_kernel<T, DH>;
  static PerDeviceOnce attr;  // per template instantiation, per device
  if (!attr.flag()) {
    CRAM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<DH>::kRequest));
    attr.flag() = true;
  }
  // per-launch scratch (work-queue counter + live-tile masks) out of a ring of 64 module globals, zeroed on the launching
  // stream
  static std::atomic<unsigned> launch_seq{0};
  LaunchScratch* ring = nullptr;
  CRAM_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ring), g_scratch));
  LaunchScratch* scratch = ring + (launch_seq.fetch_add(1) & 63u);
  // Dead-tile skipping needs a bias, a biased head, and a batch that fits the mask layout (<= 1024 sequences of <= 64
  // tiles = 8192 keys); anything else runs every tile. CRAM_B200_NO_DEAD_TILES=1 turns it off (A/B timing, tests).
  static const bool skip_off = [] {
    const char* e = getenv("CRAM_B200_NO_DEAD_TILES");
    return e != nullptr && e[0] != '\0' && e[0] != '0';
  }();
  const int max_tiles = (p.max_seqlen_k + kBlockN - 1) / kBlockN;
  const bool masks = !skip_off && p.logcred != nullptr && p.head_mask != 0 && p.n_seq <= kMaxMaskSeqs && max_tiles > 0 &&
                     max_tiles <= kMaxMaskTiles;
  CRAM_CUDA(cudaMemsetAsync(scratch, 0, masks ? 8 + sizeof(unsigned long long) * p.n_seq : 8, stream));
  if (masks) {
    const long long warps = (long long)p.n_seq * max_tiles;
    live_tiles_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, stream>>>(p, scratch->live);
    if ((rc = check_launch("live_tiles_kernel")) != CRAM_OK) return rc;
  }
  const int grid = (int)(n_items < sm_count() ? n_items : sm_count());
  kern<<<grid, kThreads, Smem<DH>::kRequest, stream>>>(tq, tk, tv, to, p, num_pairs, (int)n_items, &scratch->counter,
                                                       scratch->live, masks ? max_tiles : 0);
  return check_launch("attn_prefill_sm100_kernel");
}

}  // namespace

bool attn_sm100_supported(const cram_attn_prefill_params& p, const char** why) {
  auto no = [&](const char* w) {
    if (why) *why = w;
    return false;
  };
  if (p.dtype != CRAM_BF16 && p.dtype != CRAM_F16) return no("dtype must be bf16 or fp16");
  if (p.head_dim != 64 && p.head_dim != 128) return no("head_dim must be 64 or 128");
  if (p.q_row_stride % 8 || p.k_row_stride % 8 || p.v_row_stride % 8 || p.out_row_stride % 8)
    return no("row strides must be multiples of 8 elements");
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(p.q) || mis(p.k) || mis(p.v) || mis(p.out)) return no("q/k/v/out must be 16-byte aligned");
  return true;
}

int launch_attn_sm100(const cram_attn_prefill_params& p, cudaStream_t stream) {
  if (p.dtype == CRAM_BF16) {
    return p.head_dim == 128 ? launch_t<__nv_bfloat16, 128>(p, stream) : launch_t<__nv_bfloat16, 64>(p, stream);
  }
  return p.head_dim == 128 ? launch_t<__half, 128>(p, stream) : launch_t<__half, 64>(p, stream);
}

}  // namespace cram

#ifdef CRAM_ATTN_TRACE
extern "C" int cram_debug_read_trace(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, cram::g_trace, sizeof(cram::g_trace)) == cudaSuccess ? 0 : -1;
}
extern "C" int cram_debug_read_evlog(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, cram::g_evlog, sizeof(cram::g_evlog)) == cudaSuccess ? 0 : -1;
}
extern "C" int cram_debug_read_trace_fine(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, cram::g_trace_fine, sizeof(cram::g_trace_fine)) == cudaSuccess ? 0 : -1;
}
#endif
