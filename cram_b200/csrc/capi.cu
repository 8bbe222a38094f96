// C-ABI plumbing: error string, device queries, attention dispatch, host-side head ranking.
#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "attn.cuh"

namespace cram {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int g_pdl = 1;

int sm_count() {
  static int n[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;  // B200
  int& slot = n[dev & 63];
  if (slot == 0) {
    int v = 0;
    slot = (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) ? v : 148;
  }
  return slot;
}

}  // namespace cram

using namespace cram;

extern "C" int cram_abi_version(void) { return CRAM_B200_ABI_VERSION; }
extern "C" const char* cram_last_error(void) { return g_err; }

extern "C" int cram_attn_prefill(const cram_attn_prefill_params* pp, cram_stream_t stream) {
  CRAM_REQUIRE(pp != nullptr, "cram_attn_prefill: null params");
  const cram_attn_prefill_params& p = *pp;
  CRAM_REQUIRE(p.n_seq >= 0 && p.total_q >= 0 && p.total_k >= 0, "cram_attn_prefill: negative sizes");
  CRAM_REQUIRE(p.n_heads > 0 && p.n_heads <= 64 && p.n_kv_heads > 0 && p.n_heads % p.n_kv_heads == 0,
               "cram_attn_prefill: bad head counts (%d query / %d kv; at most 64 query heads)", p.n_heads, p.n_kv_heads);
  CRAM_REQUIRE(p.head_dim > 0, "cram_attn_prefill: bad head_dim");
  CRAM_REQUIRE(p.dtype >= CRAM_F32 && p.dtype <= CRAM_BF16, "cram_attn_prefill: bad dtype %d", p.dtype);
  CRAM_REQUIRE(p.logcred == nullptr || (p.logcred_dtype >= CRAM_F32 && p.logcred_dtype <= CRAM_BF16),
               "cram_attn_prefill: bad logcred_dtype %d", p.logcred_dtype);
  if (p.n_seq == 0 || p.total_q == 0) return CRAM_OK;
  CRAM_REQUIRE(p.q && p.k && p.v && p.out && p.cu_seqlens_q, "cram_attn_prefill: null tensor pointer");
  CRAM_REQUIRE(p.max_seqlen_q > 0, "cram_attn_prefill: max_seqlen_q must be > 0");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const char* why = "";
  const bool fast_ok = attn_sm100_supported(p, &why);
  if (p.kernel == 2) {
    if (!fast_ok) {
      set_error("cram_attn_prefill: tcgen05 kernel requested but %s", why);
      return CRAM_ERR_UNSUPPORTED;
    }
    return launch_attn_sm100(p, st);
  }
  if (p.kernel == 1 || !fast_ok) return launch_attn_generic(p, st);
  return launch_attn_sm100(p, st);
}

// utils/find_best_heads.py:113-127: float64 mean over samples, descending; ties by flat index ascending
// (the reference's np.argsort()[::-1] leaves tie order to the numpy build -- SURVEY.md section 4).
extern "C" int cram_rank_heads(const double* scores, int n_samples, int n_layers, int n_heads, int32_t* order_out,
                               double* mean_out) {
  CRAM_REQUIRE(scores && order_out && n_samples > 0 && n_layers > 0 && n_heads > 0, "cram_rank_heads: bad arguments");
  const int n = n_layers * n_heads;
  std::vector<double> mean(n, 0.0);
  // numpy's mean over axis 0 adds the samples in order (pairwise summation only applies along the
  // contiguous reduction axis), then divides once.
  for (int s = 0; s < n_samples; ++s)
    for (int i = 0; i < n; ++i) mean[i] += scores[(size_t)s * n + i];
  for (int i = 0; i < n; ++i) mean[i] /= (double)n_samples;
  std::vector<int32_t> idx(n);
  std::iota(idx.begin(), idx.end(), 0);
  std::stable_sort(idx.begin(), idx.end(), [&](int32_t a, int32_t b) { return mean[a] > mean[b]; });
  for (int i = 0; i < n; ++i) {
    order_out[i] = idx[i];
    if (mean_out) mean_out[i] = mean[idx[i]];
  }
  return CRAM_OK;
}

extern "C" void cram_debug_pdl(int on) { cram::g_pdl = on; }
