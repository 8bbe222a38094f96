// Kernel (1), decode step: one new query token per sequence over a paged KV cache, split-KV + combine.
// HBM-bound (every cached K/V byte is read once per step): CTA = (split, kv head, sequence); all the
// query heads of the GQA group are evaluated against each staged K/V tile so the cache is streamed
// once per group, 16-byte cp.async into shared memory, warp-shuffle dot-product reductions.
// The credibility bias is indexed by absolute key position and only applies to keys < bias_len
// (the prompt), as the reference's hook slices [..., :weight_len] (utils/re_weighting.py:57).
#include <mutex>

#include "common.cuh"

namespace cram {

constexpr int kDecThreads = 256;
constexpr int kDecWarps = 8;
constexpr int kDecKeys = 64;                       // keys per staged tile
constexpr int kKeysPerWarp = kDecKeys / kDecWarps;  // 8

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// partial results: [n_seq, n_heads, n_splits, DH + 2] floats (o[DH] unnormalised, m, l)
template <typename T, int DH, int G>
__global__ void __launch_bounds__(kDecThreads)
attn_decode_kernel(const cram_attn_decode_params p, const int n_splits, const int group_chunks, float* __restrict__ part) {
  constexpr int EPL = DH / 32;                   // elements per lane
  constexpr int ROW_BYTES = DH * (int)sizeof(T);
  constexpr int CHUNKS_PER_ROW = ROW_BYTES / 16;
  extern __shared__ __align__(16) uint8_t dsm[];
  T* sK = reinterpret_cast<T*>(dsm);
  T* sV = sK + kDecKeys * DH;
  float* s_red = reinterpret_cast<float*>(sV + kDecKeys * DH);  // [warps][G][DH+2]

  const int split = blockIdx.x;
  const int kvh = blockIdx.y / group_chunks, chunk = blockIdx.y % group_chunks;
  const int seq = blockIdx.z;
  const int n_rep = p.n_heads / p.n_kv_heads;
  const int h_base = kvh * n_rep + chunk * G;
  const int g_count = min(G, n_rep - chunk * G);
  const int len = p.seqlens[seq];
  const int per_split = ((len + n_splits - 1) / n_splits + kDecKeys - 1) / kDecKeys * kDecKeys;
  const int k_begin = split * per_split, k_end = min(len, k_begin + per_split);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bias_len = p.bias_len ? p.bias_len[seq] : len;
  const float kLog2e = 1.4426950408889634f;
  const float scale2 = p.scale * kLog2e;

  float q[G][EPL], o[G][EPL], m[G], l[G];
  bool hb[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
    hb[g] = false;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      o[g][e] = 0.f;
      q[g][e] = 0.f;
    }
    if (g < g_count) {
      const T* qp = reinterpret_cast<const T*>(p.q) + (int64_t)seq * p.q_row_stride + (int64_t)(h_base + g) * DH + lane * EPL;
#pragma unroll
      for (int e = 0; e < EPL; ++e) q[g][e] = DType<T>::to_f(qp[e]) * scale2;
      hb[g] = p.logcred != nullptr && ((p.head_mask >> (h_base + g)) & 1ull);
    }
  }

  const int32_t* bt = p.block_table + (int64_t)seq * p.block_table_stride;
  const T* kc = reinterpret_cast<const T*>(p.k_cache);
  const T* vc = reinterpret_cast<const T*>(p.v_cache);

  for (int kb = k_begin; kb < k_end; kb += kDecKeys) {
    __syncthreads();
    for (int c = threadIdx.x; c < kDecKeys * CHUNKS_PER_ROW; c += kDecThreads) {
      const int j = c / CHUNKS_PER_ROW, part_i = c % CHUNKS_PER_ROW;
      const int pos = kb + j;
      if (pos < k_end) {
        const int64_t page = bt[pos / p.page_size];
        const int64_t off = page * p.page_stride + (int64_t)(pos % p.page_size) * p.token_stride + (int64_t)kvh * DH;
        cp_async16(reinterpret_cast<uint8_t*>(sK + j * DH) + part_i * 16, reinterpret_cast<const uint8_t*>(kc + off) + part_i * 16);
        cp_async16(reinterpret_cast<uint8_t*>(sV + j * DH) + part_i * 16, reinterpret_cast<const uint8_t*>(vc + off) + part_i * 16);
      }
    }
    cp_async_wait_all();
    __syncthreads();

    // scores for this warp's 8 keys x G heads (every lane ends with the full sums)
    float s[kKeysPerWarp][G];
#pragma unroll
    for (int kk = 0; kk < kKeysPerWarp; ++kk) {
      const int j = warp * kKeysPerWarp + kk;
      const int pos = kb + j;
      float kf[EPL];
#pragma unroll
      for (int e = 0; e < EPL; ++e) kf[e] = pos < k_end ? DType<T>::to_f(sK[j * DH + lane * EPL + e]) : 0.f;
      float bias = 0.f;
      if (pos < k_end && pos < bias_len && p.logcred != nullptr)
        bias = load_logcred(p.logcred, (int64_t)seq * p.logcred_stride + pos, p.logcred_dtype) * kLog2e;
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float acc = 0.f;
#pragma unroll
        for (int e = 0; e < EPL; ++e) acc = fmaf(q[g][e], kf[e], acc);
        acc = warp_sum(acc);
        s[kk][g] = pos < k_end ? (hb[g] ? acc + bias : acc) : -INFINITY;
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      float mx = m[g];
#pragma unroll
      for (int kk = 0; kk < kKeysPerWarp; ++kk) mx = fmaxf(mx, s[kk][g]);
      if (mx == -INFINITY) continue;
      const float corr = m[g] == -INFINITY ? 0.f : exp2f(m[g] - mx);
      m[g] = mx;
      l[g] *= corr;
#pragma unroll
      for (int e = 0; e < EPL; ++e) o[g][e] *= corr;
#pragma unroll
      for (int kk = 0; kk < kKeysPerWarp; ++kk) {
        const float pj = s[kk][g] == -INFINITY ? 0.f : exp2f(s[kk][g] - mx);
        l[g] += pj;
        const int j = warp * kKeysPerWarp + kk;
        if (kb + j < k_end) {  // slots past the split hold stale shared memory
#pragma unroll
          for (int e = 0; e < EPL; ++e) o[g][e] = fmaf(pj, DType<T>::to_f(sV[j * DH + lane * EPL + e]), o[g][e]);
        }
      }
    }
  }

  // combine the 8 warps through shared memory
  __syncthreads();
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float* dst = s_red + ((warp * G + g) * (DH + 2));
#pragma unroll
    for (int e = 0; e < EPL; ++e) dst[lane * EPL + e] = o[g][e];
    if (lane == 0) {
      dst[DH] = m[g];
      dst[DH + 1] = l[g];
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < g_count * DH; idx += kDecThreads) {
    const int g = idx / DH, d = idx % DH;
    float mm = -INFINITY;
    for (int w = 0; w < kDecWarps; ++w) mm = fmaxf(mm, s_red[(w * G + g) * (DH + 2) + DH]);
    float acc = 0.f, ll = 0.f;
    for (int w = 0; w < kDecWarps; ++w) {
      const float* src = s_red + (w * G + g) * (DH + 2);
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
__global__ void attn_decode_combine_kernel(const float* __restrict__ part, int n_splits, int n_heads, T* __restrict__ out,
                                           int64_t out_row_stride) {
  const int head = blockIdx.x, seq = blockIdx.y;
  const float* src = part + ((int64_t)seq * n_heads + head) * n_splits * (DH + 2);
  float mm = -INFINITY;
  for (int s = 0; s < n_splits; ++s) mm = fmaxf(mm, src[s * (DH + 2) + DH]);
  for (int d = threadIdx.x; d < DH; d += blockDim.x) {
    float acc = 0.f, ll = 0.f;
    for (int s = 0; s < n_splits; ++s) {
      const float ms = src[s * (DH + 2) + DH];
      const float f = ms == -INFINITY ? 0.f : exp2f(ms - mm);
      acc += src[s * (DH + 2) + d] * f;
      ll += src[s * (DH + 2) + DH + 1] * f;
    }
    out[(int64_t)seq * out_row_stride + (int64_t)head * DH + d] = DType<T>::from_f(ll > 0.f ? acc / ll : 0.f);
  }
}

static int choose_splits(int n_seq, int n_kv_heads, int max_seqlen) {
  const int base = n_seq * n_kv_heads;
  int want = (2 * sm_count() + base - 1) / base;  // aim for >= 2 CTAs per SM
  const int max_by_len = (max_seqlen + 4 * kDecKeys - 1) / (4 * kDecKeys);  // >= 256 keys per split
  if (want > max_by_len) want = max_by_len;
  if (want < 1) want = 1;
  if (want > 64) want = 64;
  return want;
}

template <typename T, int DH, int G>
static int launch_decode_g(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream) {
  const int n_rep = p.n_heads / p.n_kv_heads;
  const int group_chunks = (n_rep + G - 1) / G;
  const size_t smem = (size_t)2 * kDecKeys * DH * sizeof(T) + (size_t)kDecWarps * G * (DH + 2) * sizeof(float);
  auto kern = attn_decode_kernel<T, DH, G>;
  CRAM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(n_splits, p.n_kv_heads * group_chunks, p.n_seq);
  kern<<<grid, kDecThreads, smem, stream>>>(p, n_splits, group_chunks, reinterpret_cast<float*>(p.workspace));
  int rc = check_launch("attn_decode_kernel");
  if (rc != CRAM_OK) return rc;
  if (n_splits > 1) {
    attn_decode_combine_kernel<T, DH><<<dim3(p.n_heads, p.n_seq), DH <= 128 ? DH : 128, 0, stream>>>(
        reinterpret_cast<const float*>(p.workspace), n_splits, p.n_heads, reinterpret_cast<T*>(p.out), p.out_row_stride);
    rc = check_launch("attn_decode_combine_kernel");
  }
  return rc;
}

static int launch_decode_combine(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream) {
  const float* part = reinterpret_cast<const float*>(p.workspace);
  const dim3 grid(p.n_heads, p.n_seq);
  if (p.dtype == CRAM_BF16) {
    if (p.head_dim == 128)
      attn_decode_combine_kernel<__nv_bfloat16, 128><<<grid, 128, 0, stream>>>(part, n_splits, p.n_heads, reinterpret_cast<__nv_bfloat16*>(p.out), p.out_row_stride);
    else
      attn_decode_combine_kernel<__nv_bfloat16, 64><<<grid, 64, 0, stream>>>(part, n_splits, p.n_heads, reinterpret_cast<__nv_bfloat16*>(p.out), p.out_row_stride);
  } else {
    if (p.head_dim == 128)
      attn_decode_combine_kernel<__half, 128><<<grid, 128, 0, stream>>>(part, n_splits, p.n_heads, reinterpret_cast<__half*>(p.out), p.out_row_stride);
    else
      attn_decode_combine_kernel<__half, 64><<<grid, 64, 0, stream>>>(part, n_splits, p.n_heads, reinterpret_cast<__half*>(p.out), p.out_row_stride);
  }
  return check_launch("attn_decode_combine_kernel");
}

template <typename T, int DH>
static int launch_decode_dh(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream) {
  const int n_rep = p.n_heads / p.n_kv_heads;
  if (n_rep == 1) return launch_decode_g<T, DH, 1>(p, n_splits, stream);
  if (n_rep == 2) return launch_decode_g<T, DH, 2>(p, n_splits, stream);
  if (n_rep <= 4) return launch_decode_g<T, DH, 4>(p, n_splits, stream);
  return launch_decode_g<T, DH, 8>(p, n_splits, stream);
}

template <typename T>
static int launch_decode_t(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream) {
  switch (p.head_dim) {
    case 64: return launch_decode_dh<T, 64>(p, n_splits, stream);
    case 128: return launch_decode_dh<T, 128>(p, n_splits, stream);
    default: set_error("cram_attn_decode: head_dim %d unsupported (64, 128)", p.head_dim); return CRAM_ERR_UNSUPPORTED;
  }
}

// tensor-core variant (attn_decode_mma.cu): bf16 / fp16, GQA group <= 16
bool attn_decode_mma_supported(const cram_attn_decode_params& p);
int launch_attn_decode_mma(const cram_attn_decode_params& p, int n_splits, cudaStream_t stream);
int attn_decode_mma_splits(const cram_attn_decode_params& p);

}  // namespace cram

using namespace cram;

extern "C" size_t cram_attn_decode_workspace_bytes(int n_seq, int n_heads, int head_dim, int max_seqlen, int n_splits) {
  if (n_splits <= 0) n_splits = 64;  // upper bound of the automatic choice
  (void)max_seqlen;
  return (size_t)n_seq * n_heads * n_splits * (head_dim + 2) * sizeof(float);
}

extern "C" int cram_attn_decode(const cram_attn_decode_params* pp, cram_stream_t stream) {
  CRAM_REQUIRE(pp != nullptr, "cram_attn_decode: null params");
  const cram_attn_decode_params& p = *pp;
  CRAM_REQUIRE(p.n_seq >= 0 && p.n_heads > 0 && p.n_kv_heads > 0 && p.n_heads % p.n_kv_heads == 0 && p.n_heads <= 64,
               "cram_attn_decode: bad head counts (%d / %d)", p.n_heads, p.n_kv_heads);
  CRAM_REQUIRE(p.page_size > 0 && p.max_seqlen >= 0, "cram_attn_decode: bad page_size / max_seqlen");
  if (p.n_seq == 0) return CRAM_OK;
  CRAM_REQUIRE(p.q && p.out && p.k_cache && p.v_cache && p.block_table && p.seqlens, "cram_attn_decode: null pointer");
  const int esz = (int)dtype_size(p.dtype);
  CRAM_REQUIRE((p.page_stride * esz) % 16 == 0 && (p.token_stride * esz) % 16 == 0 && (p.head_dim * esz) % 16 == 0,
               "cram_attn_decode: cache strides must be 16-byte multiples");
  CRAM_REQUIRE(((uintptr_t)p.k_cache & 15) == 0 && ((uintptr_t)p.v_cache & 15) == 0, "cram_attn_decode: cache must be 16-byte aligned");
  const bool use_mma = attn_decode_mma_supported(p);
  int n_splits = p.n_splits > 0 ? p.n_splits
                                : (use_mma ? attn_decode_mma_splits(p) : choose_splits(p.n_seq, p.n_kv_heads, p.max_seqlen));
  if (n_splits > 1) {
    const size_t need = (size_t)p.n_seq * p.n_heads * n_splits * (p.head_dim + 2) * sizeof(float);
    if (p.workspace == nullptr || p.workspace_bytes < need) {
      set_error("cram_attn_decode: workspace %zu bytes < %zu needed for %d splits", p.workspace_bytes, need, n_splits);
      return CRAM_ERR_WORKSPACE;
    }
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (use_mma) {
    const int rc = launch_attn_decode_mma(p, n_splits, st);
    if (rc != CRAM_OK || n_splits == 1) return rc;
    return launch_decode_combine(p, n_splits, st);
  }
  switch (p.dtype) {
    case CRAM_F32: return launch_decode_t<float>(p, n_splits, st);
    case CRAM_F16: return launch_decode_t<__half>(p, n_splits, st);
    case CRAM_BF16: return launch_decode_t<__nv_bfloat16>(p, n_splits, st);
    default: set_error("cram_attn_decode: bad dtype %d", p.dtype); return CRAM_ERR_INVALID;
  }
}

// ---- allocation-free variant -------------------------------------------------------------------------------------
// cram_attn_decode_static: cram_attn_decode for callers that may not allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc). The split-KV partials (n_seq x n_heads x n_splits x (head_dim + 2) floats) go into
// a file-scope device array instead of a caller workspace. Every GPU that loads the library holds the array until the
// process exits; every call on a device shares it, so calls that overlap on the GPU (different streams) race -- run
// them on one stream. The kernels write every partial the combine kernel reads, so nothing has to start at zero and the
// previous call's contents are simply overwritten.
namespace cram {
namespace {
constexpr size_t kStaticScratchBytes = (size_t)CRAM_ATTN_DECODE_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_decode_static_scratch[kStaticScratchBytes];

// The array's name is a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress, asked
// once per device with that device current.
int decode_static_scratch(void** out) {
  static void* ptrs[64] = {};
  static std::mutex mu;
  int dev = 0;
  CRAM_CUDA(cudaGetDevice(&dev));
  CRAM_REQUIRE(dev >= 0 && dev < 64, "cram_attn_decode_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* ptr = nullptr;
    CRAM_CUDA(cudaGetSymbolAddress((void**)&ptr, g_decode_static_scratch));
    CRAM_REQUIRE(ptr != nullptr, "cram_attn_decode_static: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = ptr;
  }
  *out = ptrs[dev];
  return CRAM_OK;
}
}  // namespace
}  // namespace cram

extern "C" int cram_attn_decode_static(const cram_attn_decode_params* pp, cram_stream_t stream) {
  CRAM_REQUIRE(pp != nullptr, "cram_attn_decode_static: null params");
  cram_attn_decode_params p = *pp;
  CRAM_REQUIRE(p.n_seq >= 0 && p.n_heads > 0 && p.n_kv_heads > 0 && p.n_heads % p.n_kv_heads == 0 && p.n_heads <= 64,
               "cram_attn_decode_static: bad head counts (%d / %d)", p.n_heads, p.n_kv_heads);
  CRAM_REQUIRE(p.page_size > 0 && p.max_seqlen >= 0 && p.head_dim > 0, "cram_attn_decode_static: bad page_size / max_seqlen / head_dim");
  if (p.n_seq == 0) return CRAM_OK;
  // fix the split count here (the same choice cram_attn_decode makes) so that the size check is about the launch that runs
  if (p.n_splits <= 0)
    p.n_splits = attn_decode_mma_supported(p) ? attn_decode_mma_splits(p) : choose_splits(p.n_seq, p.n_kv_heads, p.max_seqlen);
  const size_t need = p.n_splits > 1 ? (size_t)p.n_seq * p.n_heads * p.n_splits * (p.head_dim + 2) * sizeof(float) : 0;
  if (need > kStaticScratchBytes) {
    set_error("cram_attn_decode_static: %zu bytes of split-KV scratch needed (%d sequences x %d heads x %d splits), the static "
              "array holds %zu: pass a smaller n_splits / batch, or use cram_attn_decode with a caller workspace",
              need, p.n_seq, p.n_heads, p.n_splits, kStaticScratchBytes);
    return CRAM_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  const int rc = decode_static_scratch(&scratch);
  if (rc != CRAM_OK) return rc;
  p.workspace = scratch;
  p.workspace_bytes = kStaticScratchBytes;
  return cram_attn_decode(&p, stream);
}
