// HBM-bound layer glue around the library GEMMs, plus the sweep's splice / answer-logit kernels and
// the paged-KV append. Every kernel rounds where HF's eager modules round (transformers
// models/llama/modeling_llama.py:52-70, 138-176, 178-186) so that the host forward tracks the hooked
// reference within bf16 noise. All are streaming kernels: 16-byte accesses, one pass where possible.
#include <mutex>

#include "common.cuh"

namespace cram {

template <typename T>
struct Vec {  // 16 bytes of T
  static constexpr int N = 16 / sizeof(T);
  T v[N];
};
template <typename T>
__device__ __forceinline__ Vec<T> ldvec(const T* p) {
  Vec<T> r;
  *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(p);
  return r;
}
template <typename T>
__device__ __forceinline__ void stvec(T* p, const Vec<T>& r) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&r);
}

// ------------------------------------------------------------------------------------------------
// RMSNorm (+ fused residual add). One CTA per row.
// ------------------------------------------------------------------------------------------------
constexpr int kNormThreads = 256;

template <typename T>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_kernel(T* __restrict__ x, const T* __restrict__ residual, const T* __restrict__ w, T* __restrict__ out,
               int hidden, float eps) {
  constexpr int N = Vec<T>::N;
  __shared__ float s_part[kNormThreads / 32];
  __shared__ float s_inv;
  pdl_launch_dependents();
  pdl_wait();
  T* xr = x + (int64_t)blockIdx.x * hidden;
  const T* rr = residual ? residual + (int64_t)blockIdx.x * hidden : nullptr;
  T* orow = out + (int64_t)blockIdx.x * hidden;
  float ss = 0.f;
  for (int i = threadIdx.x * N; i < hidden; i += kNormThreads * N) {
    Vec<T> a = ldvec(xr + i);
    if (rr) {
      Vec<T> b = ldvec(rr + i);
#pragma unroll
      for (int j = 0; j < N; ++j) a.v[j] = DType<T>::from_f(DType<T>::to_f(b.v[j]) + DType<T>::to_f(a.v[j]));
      stvec(xr + i, a);
    }
#pragma unroll
    for (int j = 0; j < N; ++j) {
      float f = DType<T>::to_f(a.v[j]);
      ss += f * f;
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < kNormThreads / 32 ? s_part[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) s_inv = rsqrtf(v / (float)hidden + eps);
  }
  __syncthreads();
  const float inv = s_inv;
  for (int i = threadIdx.x * N; i < hidden; i += kNormThreads * N) {
    Vec<T> a = ldvec(xr + i);  // same thread wrote these elements above; L1-resident
    Vec<T> ww = ldvec(w + i);
    Vec<T> o;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      float n = round_to<T>(DType<T>::to_f(a.v[j]) * inv);
      o.v[j] = DType<T>::from_f(DType<T>::to_f(ww.v[j]) * n);
    }
    stvec(orow + i, o);
  }
}

// ------------------------------------------------------------------------------------------------
// Rotary embedding in place on the first n_rot_heads heads of each row (Q heads then K heads of a
// fused QKV buffer). out = x*cos + rotate_half(x)*sin with HF's three roundings.
// ------------------------------------------------------------------------------------------------
// I: index type of the flattened (row, head, vector) space -- int32 whenever it fits (a 64-bit division per item
// costs more than the item's arithmetic)
template <typename T, typename I>
__global__ void rope_kernel(T* __restrict__ qk, int64_t row_stride, int64_t n_rows, int n_rot_heads, int head_dim,
                            const T* __restrict__ cos_t, const T* __restrict__ sin_t,
                            const int32_t* __restrict__ positions) {
  constexpr int N = Vec<T>::N;
  const int half = head_dim / 2;
  const int vec_per_head = half / N;
  const I per_row = (I)n_rot_heads * vec_per_head;
  const I total = (I)n_rows * per_row;
  for (I idx = blockIdx.x * (I)blockDim.x + threadIdx.x; idx < total; idx += (I)gridDim.x * blockDim.x) {
    const I r = idx / per_row;
    const int rem = (int)(idx - r * per_row);
    const int h = rem / vec_per_head;
    const int j = (rem - h * vec_per_head) * N;
    const int pos = positions[r];
    T* base = qk + (int64_t)r * row_stride + (int64_t)h * head_dim;
    Vec<T> x1 = ldvec(base + j), x2 = ldvec(base + j + half);
    Vec<T> c1 = ldvec(cos_t + (int64_t)pos * head_dim + j), c2 = ldvec(cos_t + (int64_t)pos * head_dim + j + half);
    Vec<T> s1 = ldvec(sin_t + (int64_t)pos * head_dim + j), s2 = ldvec(sin_t + (int64_t)pos * head_dim + j + half);
    Vec<T> o1, o2;
#pragma unroll
    for (int e = 0; e < N; ++e) {
      const float a = DType<T>::to_f(x1.v[e]), b = DType<T>::to_f(x2.v[e]);
      // explicit _rn intrinsics: no FMA contraction, so fp32 also matches torch's mul, mul, add
      o1.v[e] = DType<T>::from_f(__fadd_rn(round_to<T>(__fmul_rn(a, DType<T>::to_f(c1.v[e]))),
                                           round_to<T>(__fmul_rn(-b, DType<T>::to_f(s1.v[e])))));
      o2.v[e] = DType<T>::from_f(__fadd_rn(round_to<T>(__fmul_rn(b, DType<T>::to_f(c2.v[e]))),
                                           round_to<T>(__fmul_rn(a, DType<T>::to_f(s2.v[e])))));
    }
    stvec(base + j, o1);
    stvec(base + j + half, o2);
  }
}

// ------------------------------------------------------------------------------------------------
// SwiGLU gate: out = dtype(dtype(silu(gate)) * up), gate/up the two halves of a fused projection row.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void silu_mul_kernel(const T* __restrict__ gu, T* __restrict__ out, int64_t n_rows, int inter) {
  constexpr int N = Vec<T>::N;
  const int vec_per_row = inter / N;
  const int64_t total = n_rows * vec_per_row;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / vec_per_row;
    const int j = (int)(idx - r * vec_per_row) * N;
    Vec<T> g = ldvec(gu + r * 2 * inter + j), u = ldvec(gu + r * 2 * inter + inter + j), o;
#pragma unroll
    for (int e = 0; e < N; ++e) {
      // exact expf + IEEE divide: 599 us at [32852 x 2 x 14336] against 427 us for the bare stream; deciding the rounding
      // from ex2.approx / rcp.approx with an exact fallback near ties was bit-identical but slower (637-733 us: branches)
      const float x = DType<T>::to_f(g.v[e]);
      const float s = round_to<T>(x / (1.0f + expf(-x)));
      o.v[e] = DType<T>::from_f(s * DType<T>::to_f(u.v[e]));
    }
    stvec(out + r * inter + j, o);
  }
}

// ------------------------------------------------------------------------------------------------
// SwiGLU gate for 16-bit dtypes through a table. T(silu(g)) depends on the 16 bits of g alone, so the exact expression
// above is evaluated ONCE per bit pattern (65 536 entries, 128 KB, a module global per dtype) and the streaming kernel
// looks it up from shared memory: bit-identical to silu_mul_kernel by construction, ~6 instructions per element instead
// of ~25 (expf + IEEE divide made the plain kernel issue-bound at 0.70 of the copy bandwidth).
// ------------------------------------------------------------------------------------------------
__device__ uint16_t g_silu_table[2][65536];  // [0] fp16, [1] bf16

template <typename T>
__global__ void silu_table_kernel(uint16_t* __restrict__ table) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 65536) return;
  const uint16_t bits = (uint16_t)i;
  const float x = DType<T>::to_f(*reinterpret_cast<const T*>(&bits));
  const T s = DType<T>::from_f(x / (1.0f + expf(-x)));
  table[i] = *reinterpret_cast<const uint16_t*>(&s);
}

constexpr int kSiluTabThreads = 1024;

template <typename T>
__global__ void __launch_bounds__(kSiluTabThreads, 1)
silu_mul_table_kernel(const T* __restrict__ gu, T* __restrict__ out, int64_t n_rows, int inter,
                      const uint16_t* __restrict__ table) {
  extern __shared__ uint16_t s_tab[];
  for (int i = threadIdx.x; i < 65536 / 8; i += kSiluTabThreads)
    reinterpret_cast<uint4*>(s_tab)[i] = reinterpret_cast<const uint4*>(table)[i];
  __syncthreads();
  constexpr int N = Vec<T>::N;  // 8
  const int vec_per_row = inter / N;
  const int64_t total = n_rows * vec_per_row;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / vec_per_row;
    const int j = (int)(idx - r * vec_per_row) * N;
    Vec<T> g = ldvec(gu + r * 2 * inter + j), u = ldvec(gu + r * 2 * inter + inter + j), o;
#pragma unroll
    for (int e = 0; e < N; ++e) {
      const uint16_t sb = s_tab[*reinterpret_cast<const uint16_t*>(&g.v[e])];
      const float sv = DType<T>::to_f(*reinterpret_cast<const T*>(&sb));
      o.v[e] = DType<T>::from_f(sv * DType<T>::to_f(u.v[e]));
    }
    stvec(out + r * inter + j, o);
  }
}

// ------------------------------------------------------------------------------------------------
// Sweep: build n_var single-head variants of an attention output.
// ------------------------------------------------------------------------------------------------
__global__ void splice_heads_kernel(const uint4* __restrict__ clean, const uint4* __restrict__ biased,
                                    const int32_t* __restrict__ heads, int n_var, int64_t n_rows, int vec_per_row,
                                    int vec_per_head, uint4* __restrict__ out) {
  const int64_t per_var = n_rows * vec_per_row;
  const int64_t total = per_var * n_var;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx / per_var);
    const int64_t within = idx - (int64_t)v * per_var;
    const int col = (int)(within % vec_per_row);
    const bool take = (col / vec_per_head) == heads[v];
    out[idx] = take ? biased[within] : clean[within];
  }
}

// ------------------------------------------------------------------------------------------------
// Sweep epilogue: final norm + lm_head rows of the answer tokens + sum + IE, one CTA per variant.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
answer_ie_kernel(const T* __restrict__ x, int64_t T_len, int hidden, const T* __restrict__ norm_w, float eps,
                 const T* __restrict__ lm_head, const int64_t* __restrict__ answer_ids, int n_answer,
                 const float* __restrict__ ori, float* __restrict__ cur, float* __restrict__ ie) {
  __shared__ float s_part[8];
  __shared__ float s_bcast;
  const int v = blockIdx.x;
  float acc_sum = 0.f;  // thread 0 only
  for (int k = 0; k < n_answer; ++k) {
    const T* row = x + ((int64_t)v * T_len + (T_len - n_answer + k)) * hidden;
    const T* wrow = lm_head + answer_ids[k] * (int64_t)hidden;
    float ss = 0.f;
    for (int i = threadIdx.x; i < hidden; i += 256) {
      const float f = DType<T>::to_f(row[i]);
      ss += f * f;
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < 32) {
      float t = threadIdx.x < 8 ? s_part[threadIdx.x] : 0.f;
      t = warp_sum(t);
      if (threadIdx.x == 0) s_bcast = rsqrtf(t / (float)hidden + eps);
    }
    __syncthreads();
    const float inv = s_bcast;
    float dot = 0.f;
    for (int i = threadIdx.x; i < hidden; i += 256) {
      const float n = round_to<T>(DType<T>::to_f(row[i]) * inv);
      const float h = round_to<T>(DType<T>::to_f(norm_w[i]) * n);
      dot += h * DType<T>::to_f(wrow[i]);
    }
    dot = warp_sum(dot);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = dot;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < 8; ++i) t += s_part[i];
      acc_sum += round_to<T>(t);  // one logit, stored in model dtype
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float c = round_to<T>(acc_sum);  // .sum() of a model-dtype tensor: fp32 accumulate, one rounding
    if (cur) cur[v] = c;
    if (ie && ori) ie[v] = round_to<T>(ori[0] - c);
  }
}

// ------------------------------------------------------------------------------------------------
// Paged KV append.
// ------------------------------------------------------------------------------------------------
__global__ void kv_append_kernel(const uint4* __restrict__ k, const uint4* __restrict__ v, int64_t k_stride_v,
                                 int64_t v_stride_v, int64_t n_rows, int vec_per_row,
                                 const int64_t* __restrict__ slot_mapping, uint4* __restrict__ kc,
                                 uint4* __restrict__ vc, int64_t page_stride_v, int64_t token_stride_v, int page_size) {
  const int64_t total = n_rows * vec_per_row;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / vec_per_row;
    const int c = (int)(idx - r * vec_per_row);
    const int64_t slot = slot_mapping[r];
    if (slot < 0) continue;
    const int64_t dst = (slot / page_size) * page_stride_v + (slot % page_size) * token_stride_v + c;
    kc[dst] = k[r * k_stride_v + c];
    vc[dst] = v[r * v_stride_v + c];
  }
}

// Rotary embedding + paged KV append in one pass over a fused QKV row (decode: one launch per layer instead of
// two). Work item = (row, head, 16-byte vector j of the first half of the head): Q heads are rotated in place, K heads
// are rotated in place AND written to the cache slot, V heads are copied to the cache slot. Same arithmetic as
// rope_kernel (products rounded to T before the add, as torch does).
template <typename T, typename I>
__global__ void rope_kv_append_kernel(T* __restrict__ qkv, int64_t row_stride, int64_t n_rows, int n_heads, int n_kv_heads,
                                      int head_dim, const T* __restrict__ cos_t, const T* __restrict__ sin_t,
                                      const int32_t* __restrict__ positions, const int64_t* __restrict__ slot_mapping,
                                      T* __restrict__ kc, T* __restrict__ vc, int64_t page_stride, int64_t token_stride,
                                      int page_size) {
  constexpr int N = Vec<T>::N;
  const int half = head_dim / 2;
  const int vec_per_head = half / N;
  const int heads_all = n_heads + 2 * n_kv_heads;
  pdl_launch_dependents();
  pdl_wait();
  const I per_row = (I)heads_all * vec_per_head;
  const I total = (I)n_rows * per_row;
  for (I idx = blockIdx.x * (I)blockDim.x + threadIdx.x; idx < total; idx += (I)gridDim.x * blockDim.x) {
    const I r = idx / per_row;
    const int rem = (int)(idx - r * per_row);
    const int h = rem / vec_per_head;
    const int j = (rem - h * vec_per_head) * N;
    T* base = qkv + (int64_t)r * row_stride + (int64_t)h * head_dim;
    Vec<T> x1 = ldvec(base + j), x2 = ldvec(base + j + half);
    const int64_t slot = slot_mapping[r];
    if (h >= n_heads + n_kv_heads) {  // V head: copy only
      if (slot >= 0) {
        T* dst = vc + (slot / page_size) * page_stride + (slot % page_size) * token_stride +
                 (int64_t)(h - n_heads - n_kv_heads) * head_dim;
        stvec(dst + j, x1);
        stvec(dst + j + half, x2);
      }
      continue;
    }
    const int pos = positions[r];
    Vec<T> c1 = ldvec(cos_t + (int64_t)pos * head_dim + j), c2 = ldvec(cos_t + (int64_t)pos * head_dim + j + half);
    Vec<T> s1 = ldvec(sin_t + (int64_t)pos * head_dim + j), s2 = ldvec(sin_t + (int64_t)pos * head_dim + j + half);
    Vec<T> o1, o2;
#pragma unroll
    for (int e = 0; e < N; ++e) {
      const float a = DType<T>::to_f(x1.v[e]), b = DType<T>::to_f(x2.v[e]);
      o1.v[e] = DType<T>::from_f(__fadd_rn(round_to<T>(__fmul_rn(a, DType<T>::to_f(c1.v[e]))),
                                           round_to<T>(__fmul_rn(-b, DType<T>::to_f(s1.v[e])))));
      o2.v[e] = DType<T>::from_f(__fadd_rn(round_to<T>(__fmul_rn(b, DType<T>::to_f(c2.v[e]))),
                                           round_to<T>(__fmul_rn(a, DType<T>::to_f(s2.v[e])))));
    }
    stvec(base + j, o1);
    stvec(base + j + half, o2);
    if (h >= n_heads && slot >= 0) {
      T* dst = kc + (slot / page_size) * page_stride + (slot % page_size) * token_stride + (int64_t)(h - n_heads) * head_dim;
      stvec(dst + j, o1);
      stvec(dst + j + half, o2);
    }
  }
}

static inline int stream_grid(int64_t work_items, int threads) {
  int64_t b = (work_items + threads - 1) / threads;
  int64_t cap = (int64_t)sm_count() * 16;  // multiple of the SM count; grid-stride covers the rest
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace cram

using namespace cram;

#define DISPATCH_DTYPE(dtype, ...)                          \
  switch (dtype) {                                          \
    case CRAM_F32: { using T = float; __VA_ARGS__; break; } \
    case CRAM_F16: { using T = __half; __VA_ARGS__; break; } \
    case CRAM_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; } \
    default: set_error("bad dtype %d", dtype); return CRAM_ERR_INVALID; \
  }

extern "C" int cram_rmsnorm(void* x, const void* residual, const void* weight, void* out, int64_t n_rows, int hidden,
                            float eps, int dtype, cram_stream_t stream) {
  CRAM_REQUIRE(n_rows >= 0 && hidden > 0, "cram_rmsnorm: bad shape");
  if (n_rows == 0) return CRAM_OK;
  const int vec = 16 / (int)dtype_size(dtype);
  CRAM_REQUIRE(hidden % vec == 0, "cram_rmsnorm: hidden %d not a multiple of %d", hidden, vec);
  CRAM_REQUIRE(aligned16(x) && aligned16(weight) && aligned16(out) && (!residual || aligned16(residual)),
               "cram_rmsnorm: pointers must be 16-byte aligned");
  CRAM_REQUIRE(n_rows < (1ll << 31), "cram_rmsnorm: too many rows");
  DISPATCH_DTYPE(dtype, (launch_pdl(n_rows <= 256, rmsnorm_kernel<T>, dim3((unsigned)n_rows), dim3(kNormThreads), 0, (cudaStream_t)stream,
                                    (T*)x, (const T*)residual, (const T*)weight, (T*)out, hidden, eps)));
  return check_launch("rmsnorm_kernel");
}

extern "C" int cram_rope(void* qk, int64_t row_stride, int64_t n_rows, int n_rot_heads, int head_dim,
                         const void* cos_table, const void* sin_table, const int32_t* positions, int dtype,
                         cram_stream_t stream) {
  CRAM_REQUIRE(n_rows >= 0 && n_rot_heads > 0 && head_dim > 0, "cram_rope: bad shape");
  if (n_rows == 0) return CRAM_OK;
  const int vec = 16 / (int)dtype_size(dtype);
  CRAM_REQUIRE((head_dim / 2) % vec == 0 && row_stride % vec == 0, "cram_rope: head_dim/2 and row_stride must be multiples of %d", vec);
  CRAM_REQUIRE(aligned16(qk) && aligned16(cos_table) && aligned16(sin_table), "cram_rope: pointers must be 16-byte aligned");
  const int64_t items = n_rows * n_rot_heads * (head_dim / 2 / vec);
  if (items + (int64_t)sm_count() * 16 * 256 < (1ll << 31)) {
    DISPATCH_DTYPE(dtype, (rope_kernel<T, int><<<stream_grid(items, 256), 256, 0, (cudaStream_t)stream>>>(
                              (T*)qk, row_stride, n_rows, n_rot_heads, head_dim, (const T*)cos_table,
                              (const T*)sin_table, positions)));
  } else {
    DISPATCH_DTYPE(dtype, (rope_kernel<T, int64_t><<<stream_grid(items, 256), 256, 0, (cudaStream_t)stream>>>(
                              (T*)qk, row_stride, n_rows, n_rot_heads, head_dim, (const T*)cos_table,
                              (const T*)sin_table, positions)));
  }
  return check_launch("rope_kernel");
}

extern "C" int cram_rope_kv_append(void* qkv, int64_t row_stride, int64_t n_rows, int n_heads, int n_kv_heads, int head_dim,
                                   const void* cos_table, const void* sin_table, const int32_t* positions,
                                   const int64_t* slot_mapping, void* k_cache, void* v_cache, int64_t page_stride,
                                   int64_t token_stride, int page_size, int dtype, cram_stream_t stream) {
  CRAM_REQUIRE(n_rows >= 0 && n_heads > 0 && n_kv_heads > 0 && head_dim > 0 && page_size > 0, "cram_rope_kv_append: bad shape");
  if (n_rows == 0) return CRAM_OK;
  const int vec = 16 / (int)dtype_size(dtype);
  CRAM_REQUIRE((head_dim / 2) % vec == 0 && row_stride % vec == 0 && page_stride % vec == 0 && token_stride % vec == 0,
               "cram_rope_kv_append: head_dim/2 and strides must be multiples of %d", vec);
  CRAM_REQUIRE(aligned16(qkv) && aligned16(cos_table) && aligned16(sin_table) && aligned16(k_cache) && aligned16(v_cache),
               "cram_rope_kv_append: pointers must be 16-byte aligned");
  const int64_t items = n_rows * (n_heads + 2 * n_kv_heads) * (head_dim / 2 / vec);
  if (items + (int64_t)sm_count() * 16 * 256 < (1ll << 31)) {
    DISPATCH_DTYPE(dtype, (launch_pdl(n_rows <= 256, rope_kv_append_kernel<T, int>, dim3(stream_grid(items, 256)), dim3(256), 0,
                                      (cudaStream_t)stream, (T*)qkv, row_stride, n_rows, n_heads, n_kv_heads, head_dim,
                                      (const T*)cos_table, (const T*)sin_table, positions, slot_mapping, (T*)k_cache,
                                      (T*)v_cache, page_stride, token_stride, page_size)));
  } else {
    DISPATCH_DTYPE(dtype, (launch_pdl(false, rope_kv_append_kernel<T, int64_t>, dim3(stream_grid(items, 256)), dim3(256), 0,
                                      (cudaStream_t)stream, (T*)qkv, row_stride, n_rows, n_heads, n_kv_heads, head_dim,
                                      (const T*)cos_table, (const T*)sin_table, positions, slot_mapping, (T*)k_cache,
                                      (T*)v_cache, page_stride, token_stride, page_size)));
  }
  return check_launch("rope_kv_append_kernel");
}

// The 16-bit tables are filled on first use per device (a launch on the caller's stream; later callers on other streams
// wait for that launch through an event).
struct SiluTableState {
  cudaEvent_t ready[64][2] = {};
  bool launched[64][2] = {};
  std::mutex mu;
};
static SiluTableState g_silu_state;

static int silu_table_for(int dtype, cudaStream_t stream, const uint16_t** table) {
  const int which = dtype == CRAM_BF16 ? 1 : 0;
  int dev = 0;
  CRAM_CUDA(cudaGetDevice(&dev));
  uint16_t* base = nullptr;
  CRAM_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&base), g_silu_table));
  uint16_t* t = base + which * 65536;
  std::lock_guard<std::mutex> lock(g_silu_state.mu);
  cudaEvent_t& ev = g_silu_state.ready[dev & 63][which];
  if (!g_silu_state.launched[dev & 63][which]) {
    CRAM_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    if (which) silu_table_kernel<__nv_bfloat16><<<64, 1024, 0, stream>>>(t);
    else silu_table_kernel<__half><<<64, 1024, 0, stream>>>(t);
    int rc = check_launch("silu_table_kernel");
    if (rc != CRAM_OK) return rc;
    CRAM_CUDA(cudaEventRecord(ev, stream));
    CRAM_CUDA(cudaFuncSetAttribute(silu_mul_table_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    CRAM_CUDA(cudaFuncSetAttribute(silu_mul_table_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    g_silu_state.launched[dev & 63][which] = true;
  } else if (cudaEventQuery(ev) != cudaSuccess) {
    (void)cudaGetLastError();  // cudaErrorNotReady is not an error
    CRAM_CUDA(cudaStreamWaitEvent(stream, ev, 0));
  }
  *table = t;
  return CRAM_OK;
}

extern "C" int cram_silu_mul(const void* gate_up, void* out, int64_t n_rows, int inter, int dtype,
                             cram_stream_t stream) {
  CRAM_REQUIRE(n_rows >= 0 && inter > 0, "cram_silu_mul: bad shape");
  if (n_rows == 0) return CRAM_OK;
  const int vec = 16 / (int)dtype_size(dtype);
  CRAM_REQUIRE(inter % vec == 0, "cram_silu_mul: inter must be a multiple of %d", vec);
  CRAM_REQUIRE(aligned16(gate_up) && aligned16(out), "cram_silu_mul: pointers must be 16-byte aligned");
  const int64_t items = n_rows * (inter / vec);
  // table path: 16-bit dtypes, enough work to amortise the 128 KB table load of every (persistent) CTA
  if (dtype != CRAM_F32 && items >= (int64_t)sm_count() * kSiluTabThreads * 8) {
    const uint16_t* table = nullptr;
    int rc = silu_table_for(dtype, (cudaStream_t)stream, &table);
    if (rc != CRAM_OK) return rc;
    const int grid = sm_count();
    if (dtype == CRAM_BF16)
      silu_mul_table_kernel<__nv_bfloat16><<<grid, kSiluTabThreads, 131072, (cudaStream_t)stream>>>(
          (const __nv_bfloat16*)gate_up, (__nv_bfloat16*)out, n_rows, inter, table);
    else
      silu_mul_table_kernel<__half><<<grid, kSiluTabThreads, 131072, (cudaStream_t)stream>>>(
          (const __half*)gate_up, (__half*)out, n_rows, inter, table);
    return check_launch("silu_mul_table_kernel");
  }
  DISPATCH_DTYPE(dtype, (silu_mul_kernel<T><<<stream_grid(items, 256), 256, 0, (cudaStream_t)stream>>>(
                            (const T*)gate_up, (T*)out, n_rows, inter)));
  return check_launch("silu_mul_kernel");
}

extern "C" int cram_sweep_splice_heads(const void* clean, const void* biased, const int32_t* heads, int n_var,
                                       int64_t n_rows, int n_heads, int head_dim, int dtype, void* out,
                                       cram_stream_t stream) {
  CRAM_REQUIRE(n_var >= 0 && n_rows >= 0 && n_heads > 0 && head_dim > 0, "cram_sweep_splice_heads: bad shape");
  if (n_var == 0 || n_rows == 0) return CRAM_OK;
  const int vec = 16 / (int)dtype_size(dtype);
  CRAM_REQUIRE(head_dim % vec == 0, "cram_sweep_splice_heads: head_dim must be a multiple of %d", vec);
  CRAM_REQUIRE(aligned16(clean) && aligned16(biased) && aligned16(out), "cram_sweep_splice_heads: 16-byte alignment");
  const int vec_per_head = head_dim / vec, vec_per_row = n_heads * vec_per_head;
  const int64_t items = (int64_t)n_var * n_rows * vec_per_row;
  splice_heads_kernel<<<stream_grid(items, 256), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)clean, (const uint4*)biased, heads, n_var, n_rows, vec_per_row, vec_per_head, (uint4*)out);
  return check_launch("splice_heads_kernel");
}

extern "C" int cram_sweep_answer_ie(const void* x, int n_var, int64_t seq_len, int hidden, const void* norm_weight,
                                    float rms_eps, const void* lm_head, const int64_t* answer_ids, int n_answer,
                                    int dtype, const float* ori, float* cur, float* ie, cram_stream_t stream) {
  CRAM_REQUIRE(n_var >= 0 && hidden > 0 && n_answer > 0 && n_answer <= seq_len, "cram_sweep_answer_ie: bad shape");
  if (n_var == 0) return CRAM_OK;
  DISPATCH_DTYPE(dtype, (answer_ie_kernel<T><<<n_var, 256, 0, (cudaStream_t)stream>>>(
                            (const T*)x, seq_len, hidden, (const T*)norm_weight, rms_eps, (const T*)lm_head, answer_ids,
                            n_answer, ori, cur, ie)));
  return check_launch("answer_ie_kernel");
}

extern "C" int cram_kv_append(const void* k, const void* v, int64_t k_row_stride, int64_t v_row_stride, int64_t n_rows,
                              int n_kv_heads, int head_dim, int dtype, const int64_t* slot_mapping, void* k_cache,
                              void* v_cache, int64_t page_stride, int64_t token_stride, int page_size,
                              cram_stream_t stream) {
  CRAM_REQUIRE(n_rows >= 0 && n_kv_heads > 0 && head_dim > 0 && page_size > 0, "cram_kv_append: bad shape");
  if (n_rows == 0) return CRAM_OK;
  const int vec = 16 / (int)dtype_size(dtype);
  CRAM_REQUIRE((n_kv_heads * head_dim) % vec == 0 && k_row_stride % vec == 0 && v_row_stride % vec == 0 &&
                   page_stride % vec == 0 && token_stride % vec == 0,
               "cram_kv_append: strides must be multiples of %d elements", vec);
  CRAM_REQUIRE(aligned16(k) && aligned16(v) && aligned16(k_cache) && aligned16(v_cache), "cram_kv_append: 16-byte alignment");
  const int vec_per_row = n_kv_heads * head_dim / vec;
  kv_append_kernel<<<stream_grid(n_rows * vec_per_row, 256), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)k, (const uint4*)v, k_row_stride / vec, v_row_stride / vec, n_rows, vec_per_row, slot_mapping,
      (uint4*)k_cache, (uint4*)v_cache, page_stride / vec, token_stride / vec, page_size);
  return check_launch("kv_append_kernel");
}
