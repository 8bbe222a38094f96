// Kernel (1), decode step, tensor-core variant (bf16 / fp16): one new query token per sequence over
// a paged KV cache, split-KV + combine. HBM-bound: every cached K/V byte is read exactly once per step.
//
// CTA = (split, kv head, sequence), 4 warps. The GQA group's query heads (<= 16) are the M rows of
// mma.sync.m16n8k16 tiles, so the whole group is evaluated against each staged K/V tile and the
// cache is streamed once per group. K/V tiles of 64 keys arrive through a 3-deep cp.async ring
// (16-byte chunks, XOR-swizzled rows so ldmatrix is conflict-free); warp w owns keys 16w..16w+15 of
// every tile and keeps its own online-softmax state, the four states are merged once at the end.
// The credibility bias is indexed by absolute key position and only applies to keys < bias_len
// (the prompt), as the reference's hook slices [..., :weight_len] (utils/re_weighting.py:57):
//     t = S*scale*log2e + (head selected ? logcred[key]*log2e : 0)
// Legacy mma.sync is deliberate here: the contraction is 16 x 64 x 128 per tile and the kernel is
// bound by the cache stream, not by the tensor pipe.
#include <cstdlib>

#include "common.cuh"

namespace cram {

namespace {

constexpr int kMmaThreads = 128;
constexpr int kMmaWarps = 4;
constexpr int kTileKeys = 64;
constexpr int kRing = 3;        // 3 x 32 KB per CTA, two CTAs per SM: 128 KB of cache reads in flight per SM
constexpr int kMaxSplitPages = 1024;  // page ids of one split staged in shared memory

__device__ __forceinline__ void cp_async16_zfill(uint32_t smem, const void* gmem, bool valid) {
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem), "l"(gmem), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
template <typename T>
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<__half>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <typename T>
__device__ __forceinline__ uint32_t pack_pair(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack_pair<__nv_bfloat16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <>
__device__ __forceinline__ uint32_t pack_pair<__half>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float fast_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// partial results: [n_seq, n_heads, n_splits, DH + 2] floats (o[DH] unnormalised, m, l)
template <typename T, int DH, bool kPagePow2>
__global__ void __launch_bounds__(kMmaThreads, 2)
attn_decode_mma_kernel(const cram_attn_decode_params p, const int n_splits, float* __restrict__ part) {
  constexpr int CHUNKS = DH / 8;                 // 16-byte chunks per K/V row
  constexpr int ROW_BYTES = DH * 2;
  constexpr int TILE_BYTES = kTileKeys * ROW_BYTES;
  constexpr int STAGE_BYTES = 2 * TILE_BYTES;    // K then V
  constexpr int KSTEPS = DH / 16;                // k-steps of Q K^T
  constexpr int NT_O = DH / 8;                   // n-tiles of the O accumulator
  constexpr int ROWS_PER_PASS = kMmaThreads / CHUNKS;
  extern __shared__ __align__(128) uint8_t dsm[];
  const uint32_t smem0 = (uint32_t)__cvta_generic_to_shared(dsm);

  // programmatic dependent launch (common.cuh): resident early, but nothing is read before the predecessor is done
  pdl_launch_dependents();
  pdl_wait();
  const int split = blockIdx.x, kvh = blockIdx.y, seq = blockIdx.z;
  const int n_rep = p.n_heads / p.n_kv_heads;
  const int h_base = kvh * n_rep;
  const int len = p.seqlens[seq];
  const int per_split = ((len + n_splits - 1) / n_splits + kTileKeys - 1) / kTileKeys * kTileKeys;
  const int k_begin = split * per_split, k_end = min(len, k_begin + per_split);
  const int n_tiles = k_end > k_begin ? (k_end - k_begin + kTileKeys - 1) / kTileKeys : 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g0 = lane >> 2, qd = lane & 3;       // accumulator rows g0 and g0 + 8, column pair qd
  const int bias_len = p.bias_len ? p.bias_len[seq] : len;
  constexpr float kLog2e = 1.4426950408889634f;
  const float scale2 = p.scale * kLog2e;
  const bool any_bias = p.logcred != nullptr && ((p.head_mask >> h_base) & ((1ull << n_rep) - 1ull)) != 0ull;
  const bool hb0 = p.logcred != nullptr && g0 < n_rep && ((p.head_mask >> (h_base + g0)) & 1ull);
  const bool hb1 = p.logcred != nullptr && g0 + 8 < n_rep && ((p.head_mask >> (h_base + g0 + 8)) & 1ull);

  // ---- Q as A fragments (rows = heads of the group, zero beyond n_rep) --------------------------------
  uint32_t qa[KSTEPS][4];
  {
    const T* qrow = reinterpret_cast<const T*>(p.q) + (int64_t)seq * p.q_row_stride + (int64_t)h_base * DH;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      const int c = ks * 16 + qd * 2;
      qa[ks][0] = g0 < n_rep ? *reinterpret_cast<const uint32_t*>(qrow + g0 * DH + c) : 0u;
      qa[ks][1] = g0 + 8 < n_rep ? *reinterpret_cast<const uint32_t*>(qrow + (g0 + 8) * DH + c) : 0u;
      qa[ks][2] = g0 < n_rep ? *reinterpret_cast<const uint32_t*>(qrow + g0 * DH + c + 8) : 0u;
      qa[ks][3] = g0 + 8 < n_rep ? *reinterpret_cast<const uint32_t*>(qrow + (g0 + 8) * DH + c + 8) : 0u;
    }
  }

  // page ids of this split -> shared memory (the block-table load would otherwise sit on the critical
  // path of every cp.async address)
  int32_t* s_pages = reinterpret_cast<int32_t*>(dsm + kRing * STAGE_BYTES);
  const int page0 = k_begin / p.page_size;
  {
    const int32_t* bt = p.block_table + (int64_t)seq * p.block_table_stride;
    const int n_pages = k_end > k_begin ? (k_end - 1) / p.page_size - page0 + 1 : 0;
    for (int i = threadIdx.x; i < n_pages; i += kMmaThreads) s_pages[i] = bt[page0 + i];
  }
  if (any_bias) {
    // pull this split's slice of the bias row towards L2 now: under a saturated HBM a first-touch miss
    // costs more than one tile iteration, which is all the register prefetch below can hide
    const int esz = p.logcred_dtype == CRAM_F32 ? 4 : 2;
    const uint8_t* b0 = reinterpret_cast<const uint8_t*>(p.logcred) + ((int64_t)seq * p.logcred_stride + k_begin) * esz;
    const int nbytes = (min(k_end, bias_len) - k_begin) * esz;
    for (int off = threadIdx.x * 128; off < nbytes; off += kMmaThreads * 128)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(b0 + off));
  }
  __syncthreads();
  const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache);
  const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache);
  const int ld_chunk = threadIdx.x % CHUNKS, ld_row0 = threadIdx.x / CHUNKS;

  const int page_shift = kPagePow2 ? 31 - __clz(p.page_size) : 0;
  // this thread's 4 bias values of tile t, fetched one tile ahead of their use as RAW bits (the
  // conversion and the x log2e happen at the use, so the loads stay in flight across the iteration)
  auto load_bias = [&](int t, uint32_t (&b)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int pos = k_begin + t * kTileKeys + warp * 16 + (i >> 1) * 8 + qd * 2 + (i & 1);
      uint32_t raw = 0u;  // +0.0 in every dtype
      if (any_bias && pos < bias_len && pos < k_end) {
        const int64_t idx = (int64_t)seq * p.logcred_stride + pos;
        raw = p.logcred_dtype == CRAM_F32 ? reinterpret_cast<const uint32_t*>(p.logcred)[idx]
                                          : (uint32_t) reinterpret_cast<const uint16_t*>(p.logcred)[idx];
      }
      b[i] = raw;
    }
  };
  auto bias_value = [&](uint32_t raw) -> float {
    float f;
    if (p.logcred_dtype == CRAM_BF16) f = __uint_as_float(raw << 16);
    else if (p.logcred_dtype == CRAM_F16) f = __half2float(__ushort_as_half((unsigned short)raw));
    else f = __uint_as_float(raw);
    return f * kLog2e;
  };
  auto load_tile = [&](int t) {
    const uint32_t st = smem0 + (t % kRing) * STAGE_BYTES;
    const int kb = k_begin + t * kTileKeys;
#pragma unroll
    for (int r = ld_row0; r < kTileKeys; r += ROWS_PER_PASS) {
      const int pos = kb + r;
      const bool ok = pos < k_end;
      const int ps = ok ? pos : k_begin;
      const int pg = kPagePow2 ? (ps >> page_shift) : ps / p.page_size;
      const int64_t page = s_pages[pg - page0];
      const int64_t off = (page * p.page_stride + (int64_t)(ps - pg * p.page_size) * p.token_stride + (int64_t)kvh * DH) * 2 +
                          ld_chunk * 16;
      const uint32_t dst = st + r * ROW_BYTES + ((ld_chunk ^ (r & 7)) << 4);
      cp_async16_zfill(dst, kc + off, ok);
      cp_async16_zfill(dst + TILE_BYTES, vc + off, ok);
    }
  };

  float o[NT_O][4];
#pragma unroll
  for (int i = 0; i < NT_O; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

#pragma unroll
  for (int t = 0; t < kRing - 1; ++t) {
    if (t < n_tiles) load_tile(t);
    cp_async_commit();
  }
  uint32_t bnext[4], bnext2[4];
  load_bias(0, bnext);
  load_bias(1, bnext2);

  for (int t = 0; t < n_tiles; ++t) {
    cp_async_wait<kRing - 2>();
    __syncthreads();  // tile t has landed for every thread; everyone is done with tile t-1's slot
    if (t + kRing - 1 < n_tiles) load_tile(t + kRing - 1);
    cp_async_commit();
    uint32_t bcur[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      bcur[i] = bnext[i];
      bnext[i] = bnext2[i];
    }
    load_bias(t + 2, bnext2);

    const uint32_t sK = smem0 + (t % kRing) * STAGE_BYTES, sV = sK + TILE_BYTES;
    const int key0 = warp * 16;                       // this warp's keys inside the tile
    const int pos0 = k_begin + t * kTileKeys + key0;  // absolute position of the first of them
    if (pos0 >= k_end) continue;                      // (uniform per warp) nothing visible: no state change

    // ---- S = Q K^T for 16 keys: 2 n-tiles x KSTEPS ---------------------------------------------------
    float s[2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
    {
      const int krow = key0 + ((lane >> 4) << 3) + (lane & 7);
      const int csel = (lane >> 3) & 1;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(sK + krow * ROW_BYTES + (((ks * 2 + csel) ^ (krow & 7)) << 4), b0, b1, b2, b3);
        mma16816<T>(s[0], qa[ks], b0, b1);
        mma16816<T>(s[1], qa[ks], b2, b3);
      }
    }

    // ---- bias, length mask, online softmax (rows g0 and g0+8; 4 keys per thread) ----------------------
    float t0[4], t1[4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int pos = pos0 + nt * 8 + qd * 2 + e;
        const float bias = bias_value(bcur[nt * 2 + e]);
        const bool vis = pos < k_end;
        t0[nt * 2 + e] = vis ? fmaf(s[nt][e], scale2, hb0 ? bias : 0.f) : -INFINITY;
        t1[nt * 2 + e] = vis ? fmaf(s[nt][2 + e], scale2, hb1 ? bias : 0.f) : -INFINITY;
      }
    }
    float mx0 = fmaxf(fmaxf(t0[0], t0[1]), fmaxf(t0[2], t0[3]));
    float mx1 = fmaxf(fmaxf(t1[0], t1[1]), fmaxf(t1[2], t1[3]));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float c0 = (m0 == -INFINITY) ? 0.f : fast_ex2(m0 - mn0);  // mn = -inf only if m = -inf too
    const float c1 = (m1 == -INFINITY) ? 0.f : fast_ex2(m1 - mn1);
    const float u0 = (mn0 == -INFINITY) ? 0.f : mn0, u1 = (mn1 == -INFINITY) ? 0.f : mn1;
    m0 = mn0;
    m1 = mn1;
    float p0[4], p1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      p0[i] = fast_ex2(t0[i] - u0);  // exp2(-inf) = 0
      p1[i] = fast_ex2(t1[i] - u1);
    }
    l0 = l0 * c0 + (p0[0] + p0[1]) + (p0[2] + p0[3]);  // per-thread partial; quad-reduced at the end
    l1 = l1 * c1 + (p1[0] + p1[1]) + (p1[2] + p1[3]);
#pragma unroll
    for (int i = 0; i < NT_O; ++i) {
      o[i][0] *= c0;
      o[i][1] *= c0;
      o[i][2] *= c1;
      o[i][3] *= c1;
    }
    uint32_t pa[4];
    pa[0] = pack_pair<T>(p0[0], p0[1]);  // row g0,   keys 0-7 of the warp
    pa[1] = pack_pair<T>(p1[0], p1[1]);  // row g0+8
    pa[2] = pack_pair<T>(p0[2], p0[3]);  // row g0,   keys 8-15
    pa[3] = pack_pair<T>(p1[2], p1[3]);

    // ---- O += P V : NT_O n-tiles, one k-step of 16 keys ----------------------------------------------
    {
      const int vrow = key0 + (((lane >> 3) & 1) << 3) + (lane & 7);
      const int csel = lane >> 4;
#pragma unroll
      for (int j = 0; j < NT_O; j += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(sV + vrow * ROW_BYTES + (((j + csel) ^ (vrow & 7)) << 4), b0, b1, b2, b3);
        mma16816<T>(o[j], pa, b0, b1);
        mma16816<T>(o[j + 1], pa, b2, b3);
      }
    }
  }
  cp_async_wait<0>();

  // ---- merge the 4 warps through shared memory (the ring is free now) ------------------------------------
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  __syncthreads();
  float* red = reinterpret_cast<float*>(dsm);  // [warp][16 rows][DH + 2]
  {
    float* r0 = red + ((warp * 16 + g0) * (DH + 2));
    float* r1 = red + ((warp * 16 + g0 + 8) * (DH + 2));
#pragma unroll
    for (int i = 0; i < NT_O; ++i) {
      r0[i * 8 + qd * 2] = o[i][0];
      r0[i * 8 + qd * 2 + 1] = o[i][1];
      r1[i * 8 + qd * 2] = o[i][2];
      r1[i * 8 + qd * 2 + 1] = o[i][3];
    }
    if (qd == 0) {
      r0[DH] = m0;
      r0[DH + 1] = l0;
      r1[DH] = m1;
      r1[DH + 1] = l1;
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < n_rep * DH; idx += kMmaThreads) {
    const int g = idx / DH, d = idx % DH;
    float mm = -INFINITY;
#pragma unroll
    for (int w = 0; w < kMmaWarps; ++w) mm = fmaxf(mm, red[(w * 16 + g) * (DH + 2) + DH]);
    float acc = 0.f, ll = 0.f;
#pragma unroll
    for (int w = 0; w < kMmaWarps; ++w) {
      const float* src = red + (w * 16 + g) * (DH + 2);
      const float f = src[DH] == -INFINITY ? 0.f : exp2f(src[DH] - mm);
      acc += src[d] * f;
      ll += src[DH + 1] * f;
    }
    const int head = h_base + g;
    if (n_splits == 1) {
      T* out = reinterpret_cast<T*>(p.out) + (int64_t)seq * p.out_row_stride + (int64_t)head * DH;
      out[d] = DType<T>::from_f(ll > 0.f ? acc / ll : 0.f);
    } else {
      float* dst = part + (((int64_t)seq * p.n_heads + head) * n_splits + split) * (DH + 2);
      dst[d] = acc;
      if (d == 0) {
        dst[DH] = mm;
        dst[DH + 1] = ll;
      }
    }
  }
}

template <typename T, int DH>
int launch_mma_t(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream) {
  constexpr size_t ring = (size_t)kRing * 2 * kTileKeys * DH * 2 + kMaxSplitPages * sizeof(int32_t);
  constexpr size_t red = (size_t)kMmaWarps * 16 * (DH + 2) * sizeof(float);
  constexpr size_t smem = ring > red ? ring : red;
  const bool pow2 = (p.page_size & (p.page_size - 1)) == 0;
  auto kern = pow2 ? attn_decode_mma_kernel<T, DH, true> : attn_decode_mma_kernel<T, DH, false>;
  static PerDeviceOnce attr[2];  // per template instantiation, per device
  if (!attr[pow2].flag()) {
    CRAM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr[pow2].flag() = true;
  }
  dim3 grid(n_splits, p.n_kv_heads, p.n_seq);
  CRAM_CUDA(launch_pdl(true, kern, grid, dim3(kMmaThreads), smem, stream, p, n_splits, reinterpret_cast<float*>(p.workspace)));
  return check_launch("attn_decode_mma_kernel");
}

}  // namespace

bool attn_decode_mma_supported(const cram_attn_decode_params& p) {
  return (p.dtype == CRAM_BF16 || p.dtype == CRAM_F16) && (p.head_dim == 64 || p.head_dim == 128) &&
         p.n_heads / p.n_kv_heads <= 16 && p.q_row_stride % 2 == 0 && (reinterpret_cast<uintptr_t>(p.q) & 3) == 0;
}

// splits: fill the resident CTA slots (2 per SM) once; never fewer than the page table staging allows
int attn_decode_mma_splits(const cram_attn_decode_params& p) {
  static const int forced = [] {  // A/B runs only: CRAM_B200_DECODE_SPLITS=<n> overrides the choice below
    const char* e = getenv("CRAM_B200_DECODE_SPLITS");
    return e ? atoi(e) : 0;
  }();
  if (forced > 0) return forced;
  const int base = p.n_seq * p.n_kv_heads;
  int want = (2 * sm_count()) / (base > 0 ? base : 1);
  const int max_by_len = (p.max_seqlen + 4 * kTileKeys - 1) / (4 * kTileKeys);  // >= 256 keys per split
  if (want > max_by_len) want = max_by_len;
  if (want > 64) want = 64;
  const int min_by_pages = (int)(((int64_t)p.max_seqlen + (int64_t)kMaxSplitPages * p.page_size - 1) /
                                 ((int64_t)kMaxSplitPages * p.page_size));
  if (want < min_by_pages) want = min_by_pages;
  return want < 1 ? 1 : want;
}

int launch_attn_decode_mma(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream) {
  {
    // one split may cover at most kMaxSplitPages pages (+1 tile of slack for the rounding of per_split)
    const int64_t per_split = (((int64_t)p.max_seqlen + n_splits - 1) / n_splits + kTileKeys - 1) / kTileKeys * kTileKeys;
    if (per_split / p.page_size + 2 > kMaxSplitPages) {
      set_error("cram_attn_decode: %d splits leave %lld keys per split; at most %d pages of %d fit", n_splits,
                (long long)per_split, kMaxSplitPages - 2, p.page_size);
      return CRAM_ERR_INVALID;
    }
  }
  if (p.dtype == CRAM_BF16)
    return p.head_dim == 128 ? launch_mma_t<__nv_bfloat16, 128>(p, n_splits, stream)
                             : launch_mma_t<__nv_bfloat16, 64>(p, n_splits, stream);
  return p.head_dim == 128 ? launch_mma_t<__half, 128>(p, n_splits, stream) : launch_mma_t<__half, 64>(p, n_splits, stream);
}

}  // namespace cram
