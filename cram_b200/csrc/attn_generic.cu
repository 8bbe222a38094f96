// Generic (CUDA-core, fp32-math) credibility-biased attention for packed variable-length rows.
// Covers what the tcgen05 kernel does not: fp32 tensors (the tiny parity model of BASELINE.json
// configs[0]), head dims other than 64/128, and it is the in-library cross-check of the fast kernel.
// Same semantics as attn_prefill_sm100.cu; see include/cram_b200.h (cram_attn_prefill).
//
// One warp per query row, 16 rows per CTA, keys in chunks of 32 (lane j owns key j for QK^T, lanes own
// strided output columns for PV), online softmax with warp-shuffle reductions.
#include "attn.cuh"

namespace cram {

constexpr int kGenWarps = 16;
constexpr int kGenThreads = kGenWarps * 32;
constexpr int kGenKeys = 32;

template <typename T>
__global__ void __launch_bounds__(kGenThreads)
attn_generic_kernel(const cram_attn_prefill_params p) {
  extern __shared__ __align__(16) float smem_f[];
  const int dh = p.head_dim;
  const int kstride = dh + 4;
  float* sK = smem_f;                          // [32][dh+4]
  float* sV = sK + kGenKeys * kstride;         // [32][dh]
  float* sQ = sV + kGenKeys * dh;              // [16][dh]

  const int seq = blockIdx.z, head = blockIdx.y;
  const int q0 = p.cu_seqlens_q[seq], len_q = p.cu_seqlens_q[seq + 1] - q0;
  const int32_t* cuk = p.cu_seqlens_k ? p.cu_seqlens_k : p.cu_seqlens_q;
  const int k0 = cuk[seq], len_k = cuk[seq + 1] - k0;
  const int row0 = blockIdx.x * kGenWarps;
  if (row0 >= len_q) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = row0 + warp;
  const bool row_ok = row < len_q;
  const int kv_head = head / (p.n_heads / p.n_kv_heads);
  const int diag = len_k - len_q;  // query i sees keys <= i + diag
  const bool biased = p.logcred != nullptr && ((p.head_mask >> head) & 1ull);
  const int bias_len = p.bias_len ? p.bias_len[seq] : len_k;

  const T* Q = reinterpret_cast<const T*>(p.q);
  const T* K = reinterpret_cast<const T*>(p.k);
  const T* V = reinterpret_cast<const T*>(p.v);

  // stage this CTA's query rows (fp32)
  for (int i = threadIdx.x; i < kGenWarps * dh; i += kGenThreads) {
    const int r = i / dh, d = i - r * dh;
    float v = 0.f;
    if (row0 + r < len_q) v = DType<T>::to_f(Q[(int64_t)(q0 + row0 + r) * p.q_row_stride + (int64_t)head * dh + d]);
    sQ[r * dh + d] = v;
  }

  const int last_row = min(row0 + kGenWarps, len_q) - 1;
  const int k_end = p.causal ? min(len_k, last_row + diag + 1) : len_k;
  const int my_limit = p.causal ? min(len_k, row + diag + 1) : len_k;  // keys < my_limit are visible

  float m = -INFINITY, l = 0.f;
  float o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = 0.f;
  const float4* q4 = reinterpret_cast<const float4*>(sQ + warp * dh);

  for (int kb = 0; kb < k_end; kb += kGenKeys) {
    __syncthreads();
    for (int i = threadIdx.x; i < kGenKeys * dh; i += kGenThreads) {
      const int j = i / dh, d = i - j * dh;
      float kv = 0.f, vv = 0.f;
      if (kb + j < len_k) {
        kv = DType<T>::to_f(K[(int64_t)(k0 + kb + j) * p.k_row_stride + (int64_t)kv_head * dh + d]);
        vv = DType<T>::to_f(V[(int64_t)(k0 + kb + j) * p.v_row_stride + (int64_t)kv_head * dh + d]);
      }
      sK[j * kstride + d] = kv;
      sV[j * dh + d] = vv;
    }
    __syncthreads();
    if (!row_ok) continue;

    const int key = kb + lane;
    float s = -INFINITY;
    if (key < my_limit) {
      const float4* k4 = reinterpret_cast<const float4*>(sK + lane * kstride);
      float acc = 0.f;
      for (int d = 0; d < dh / 4; ++d) {
        const float4 a = q4[d], b = k4[d];
        acc = fmaf(a.x, b.x, acc);
        acc = fmaf(a.y, b.y, acc);
        acc = fmaf(a.z, b.z, acc);
        acc = fmaf(a.w, b.w, acc);
      }
      s = acc * p.scale;
      if (biased && key < bias_len) s += load_logcred(p.logcred, k0 + key, p.logcred_dtype);
    }
    const float m_new = fmaxf(m, warp_max(s));
    if (m_new == -INFINITY) continue;  // nothing visible yet
    const float pj = (s == -INFINITY) ? 0.f : expf(s - m_new);
    const float corr = (m == -INFINITY) ? 0.f : expf(m - m_new);
    l = l * corr + warp_sum(pj);
    m = m_new;
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] *= corr;
    for (int j = 0; j < kGenKeys; ++j) {
      const float pb = __shfl_sync(0xffffffffu, pj, j);
      if (pb != 0.f) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int d = lane + 32 * i;
          if (d < dh) o[i] = fmaf(pb, sV[j * dh + d], o[i]);
        }
      }
    }
  }

  if (!row_ok) return;
  const float inv = l > 0.f ? 1.f / l : 0.f;
  T* out = reinterpret_cast<T*>(p.out) + (int64_t)(q0 + row) * p.out_row_stride + (int64_t)head * dh;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int d = lane + 32 * i;
    if (d < dh) out[d] = DType<T>::from_f(o[i] * inv);
  }
  if (p.lse && lane == 0) p.lse[(int64_t)head * p.total_q + q0 + row] = l > 0.f ? m + logf(l) : -INFINITY;
}

int launch_attn_generic(const cram_attn_prefill_params& p, cudaStream_t stream) {
  CRAM_REQUIRE(p.head_dim % 8 == 0 && p.head_dim <= 256, "cram_attn_prefill(generic): head_dim %d must be a multiple of 8, <= 256", p.head_dim);
  const size_t smem = sizeof(float) * ((size_t)kGenKeys * (p.head_dim + 4) + (size_t)kGenKeys * p.head_dim + (size_t)kGenWarps * p.head_dim);
  dim3 grid((p.max_seqlen_q + kGenWarps - 1) / kGenWarps, p.n_heads, p.n_seq);
  if (grid.x == 0) return CRAM_OK;
#define LAUNCH_GEN(T)                                                                                       \
  do {                                                                                                      \
    CRAM_CUDA(cudaFuncSetAttribute(attn_generic_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    attn_generic_kernel<T><<<grid, kGenThreads, smem, stream>>>(p);                                         \
  } while (0)
  switch (p.dtype) {
    case CRAM_F32: LAUNCH_GEN(float); break;
    case CRAM_F16: LAUNCH_GEN(__half); break;
    case CRAM_BF16: LAUNCH_GEN(__nv_bfloat16); break;
    default: set_error("cram_attn_prefill: bad dtype %d", p.dtype); return CRAM_ERR_INVALID;
  }
#undef LAUNCH_GEN
  return check_launch("attn_generic_kernel");
}

}  // namespace cram
