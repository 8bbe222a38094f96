// Kernel (2): token -> document span index, credibility vector and its log, bit-exact with
// Re_Weighting_Strategy.decode_with_special_attention (reference utils/re_weighting.py:99-111)
// followed by torch.log on the fp16 vector (:122/:166) and the cast to the model dtype (:56).
//
// One CTA per packed sequence. Integer work, latency bound at prompt sizes (T*8 B in, T*8 B out);
// the reference spends T*D*2 device syncs on it, this is one launch for a whole batch.
#include "common.cuh"

namespace cram {

constexpr int kSpanThreads = 512;
constexpr int kMaxDocs = 256;

// log of an fp16 value the way torch evaluates it: float log, rounded to fp16. Using the correctly
// rounded float log (via double) reproduces torch's result for every non-negative fp16 input
// (tests/golden/log_chain.npz holds all 65 536 cases).
__device__ __forceinline__ uint16_t log_fp16_bits(uint16_t cred_bits) {
  float x = __half2float(__ushort_as_half(cred_bits));
  float r = static_cast<float>(log(static_cast<double>(x)));
  return __half_as_ushort(__float2half_rn(r));
}

__device__ __forceinline__ void store_logcred(void* out, int64_t i, uint16_t log16, int dtype) {
  __half h = __ushort_as_half(log16);
  if (dtype == CRAM_F16) {
    reinterpret_cast<__half*>(out)[i] = h;
  } else if (dtype == CRAM_BF16) {
    reinterpret_cast<__nv_bfloat16*>(out)[i] = __float2bfloat16_rn(__half2float(h));
  } else {
    reinterpret_cast<float*>(out)[i] = __half2float(h);
  }
}

__global__ void __launch_bounds__(kSpanThreads)
span_index_kernel(const int64_t* __restrict__ offset_starts, const int32_t* __restrict__ cu_seqlens,
                  const int32_t* __restrict__ span_begin, const int32_t* __restrict__ span_end,
                  const float* __restrict__ scores, const int32_t* __restrict__ cu_docs, int logcred_dtype,
                  int32_t* __restrict__ doc_id, uint16_t* __restrict__ cred_f16, void* __restrict__ logcred,
                  int32_t* __restrict__ err_flags) {
  __shared__ int s_start[kMaxDocs];
  __shared__ int s_end[kMaxDocs];
  __shared__ uint16_t s_cred[kMaxDocs];
  __shared__ uint16_t s_log[kMaxDocs];
  __shared__ int s_err;

  const int seq = blockIdx.x;
  const int t0 = cu_seqlens[seq];
  const int T = cu_seqlens[seq + 1] - t0;
  const int d0 = cu_docs[seq];
  const int D = min(cu_docs[seq + 1] - d0, kMaxDocs);
  const int64_t* off = offset_starts + t0;

  for (int d = threadIdx.x; d < D; d += kSpanThreads) {
    s_start[d] = -1;
    s_end[d] = -1;
    uint16_t c = __half_as_ushort(__float2half_rn(scores[d0 + d]));
    s_cred[d] = c;
    s_log[d] = log_fp16_bits(c);
  }
  if (threadIdx.x == 0) s_err = 0;
  __syncthreads();

  // phase 1: per document, the LAST token index whose offset start is <= the span begin / end.
  for (int d = 0; d < D; ++d) {
    const int64_t sb = span_begin[d0 + d];
    const int64_t se = span_end[d0 + d];
    int best_s = -1, best_e = -1;
    for (int t = threadIdx.x; t < T; t += kSpanThreads) {
      const int64_t o = __ldg(off + t);
      if (sb >= o) best_s = t;  // t increases per thread, so the last hit is the max
      if (se >= o) best_e = t;
    }
    best_s = warp_max_i(best_s);
    best_e = warp_max_i(best_e);
    if ((threadIdx.x & 31) == 0) {
      if (best_s >= 0) atomicMax(&s_start[d], best_s);
      if (best_e >= 0) atomicMax(&s_end[d], best_e);
    }
  }
  __syncthreads();

  if (threadIdx.x < D) {
    const int a = s_start[threadIdx.x], b = s_end[threadIdx.x];
    int e = 0;
    if (a < 0 || b < 0) e |= 1;
    else if (b + 1 < a) e |= 2;
    if (e) atomicOr(&s_err, e);
  }

  // phase 2: the last document whose token range covers t wins (later writes overwrite earlier).
  const uint16_t one_bits = 0x3C00u;  // fp16 1.0
  for (int t = threadIdx.x; t < T; t += kSpanThreads) {
    int who = -1;
    for (int d = D - 1; d >= 0; --d) {
      const int a = s_start[d], b = s_end[d];
      if (a >= 0 && b >= 0 && t >= a && t <= b) {
        who = d;
        break;
      }
    }
    const uint16_t c = who >= 0 ? s_cred[who] : one_bits;
    if (doc_id) doc_id[t0 + t] = who;
    if (cred_f16) cred_f16[t0 + t] = c;
    if (logcred) store_logcred(logcred, t0 + t, who >= 0 ? s_log[who] : (uint16_t)0, logcred_dtype);
  }
  __syncthreads();
  if (threadIdx.x == 0 && err_flags) {
    int e = s_err;
    if (cu_docs[seq + 1] - d0 > kMaxDocs) e |= 4;
    err_flags[seq] = e;
  }
}

__global__ void log_cred_kernel(const uint16_t* __restrict__ cred, int64_t n, int dtype, void* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    store_logcred(out, i, log_fp16_bits(cred[i]), dtype);
}

}  // namespace cram

extern "C" int cram_span_index(const int64_t* offset_starts, const int32_t* cu_seqlens, int n_seq,
                               const int32_t* span_begin, const int32_t* span_end, const float* scores,
                               const int32_t* cu_docs, int logcred_dtype, int32_t* doc_id, uint16_t* cred_f16,
                               void* logcred, int32_t* err_flags, cram_stream_t stream) {
  using namespace cram;
  CRAM_REQUIRE(n_seq >= 0, "cram_span_index: n_seq < 0");
  CRAM_REQUIRE(logcred_dtype >= CRAM_F32 && logcred_dtype <= CRAM_BF16, "cram_span_index: bad logcred_dtype %d", logcred_dtype);
  if (n_seq == 0) return CRAM_OK;
  CRAM_REQUIRE(offset_starts && cu_seqlens && cu_docs, "cram_span_index: null input");
  span_index_kernel<<<n_seq, kSpanThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      offset_starts, cu_seqlens, span_begin, span_end, scores, cu_docs, logcred_dtype, doc_id, cred_f16, logcred,
      err_flags);
  return check_launch("span_index_kernel");
}

extern "C" int cram_log_cred(const uint16_t* cred_f16, int64_t n, int logcred_dtype, void* logcred,
                             cram_stream_t stream) {
  using namespace cram;
  CRAM_REQUIRE(n >= 0, "cram_log_cred: n < 0");
  CRAM_REQUIRE(logcred_dtype >= CRAM_F32 && logcred_dtype <= CRAM_BF16, "cram_log_cred: bad dtype");
  if (n == 0) return CRAM_OK;
  int blocks = (int)((n + 255) / 256);
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  log_cred_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(cred_f16, n, logcred_dtype, logcred);
  return check_launch("log_cred_kernel");
}
