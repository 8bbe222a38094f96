/*
 * cram_b200.h -- C ABI of libcram_b200.so: the B200 (sm_100a) kernels behind CrAM's
 * credibility-aware attention path.
 *
 * The reference (yeahjack/CrAM) is pure Python and has no FFI of its own; each entry point below
 * names the reference code it replaces (paths relative to the reference checkout). INTEGRATION.md
 * shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller unless marked "host";
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), allocates nothing,
 *     and returns 0 on success or a negative cram_status; cram_last_error() gives the message
 *     (thread-local);
 *   - dtypes are cram_dtype codes; 16-bit values travel as uint16_t bit patterns;
 *   - there is no CPU fallback: without a CUDA device every compute call returns CRAM_ERR_CUDA.
 */
#ifndef CRAM_B200_H_
#define CRAM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRAM_B200_ABI_VERSION 1

typedef void* cram_stream_t; /* cudaStream_t */

enum cram_dtype { CRAM_F32 = 0, CRAM_F16 = 1, CRAM_BF16 = 2 };

enum cram_status {
  CRAM_OK = 0,
  CRAM_ERR_INVALID = -1,     /* bad argument (shape, alignment, dtype) */
  CRAM_ERR_UNSUPPORTED = -2, /* valid request this build has no kernel for */
  CRAM_ERR_CUDA = -3,        /* CUDA runtime / driver error */
  CRAM_ERR_WORKSPACE = -4    /* caller-provided workspace too small */
};

int cram_abi_version(void);
const char* cram_last_error(void);

/* ------------------------------------------------------------------------------------------------
 * (2) token -> document span index and credibility vector.
 * Replaces the Python double loop of Re_Weighting_Strategy.decode_with_special_attention
 * (utils/re_weighting.py:99-111) and the torch.log / .to(dtype) that follow it (:122, :56, :166).
 *
 * Batched over n_seq packed sequences. Sequence s owns tokens [cu_seqlens[s], cu_seqlens[s+1]) and
 * documents [cu_docs[s], cu_docs[s+1]). For document d with inclusive character span
 * [span_begin[d], span_end[d]]:
 *     start_tok = max{ t : span_begin[d] >= offset_starts[t] },  end_tok likewise for span_end[d]
 * (token indices local to the sequence, no monotonicity assumed), and tokens start_tok..end_tok get
 * doc_id = d - cu_docs[s] and cred = fp16(scores[d]); later documents overwrite earlier ones.
 * Tokens of no document get doc_id -1, cred 1. logcred = log(cred) evaluated as the reference does
 * (fp16 in, fp16 out) and then cast to `logcred_dtype` (bit-exact with torch for every cred >= 0).
 *
 * err_flags[s] (optional): bit 0 = a document found no start/end token (the reference raises
 * TypeError at :111), bit 1 = end_tok + 1 < start_tok (the reference raises in torch.full).
 * Outputs doc_id / cred_f16 / logcred may each be NULL.
 */
int cram_span_index(const int64_t* offset_starts, const int32_t* cu_seqlens, int n_seq,
                    const int32_t* span_begin, const int32_t* span_end, const float* scores,
                    const int32_t* cu_docs, int logcred_dtype, int32_t* doc_id, uint16_t* cred_f16,
                    void* logcred, int32_t* err_flags, cram_stream_t stream);

/* log chain alone: logcred[i] = cast(log_fp16(cred_f16[i])) (utils/re_weighting.py:122 then :56). */
int cram_log_cred(const uint16_t* cred_f16, int64_t n, int logcred_dtype, void* logcred,
                  cram_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (1) fused credibility-biased attention, prefill / chunked prefill.
 * Replaces edit_attention_mask (utils/re_weighting.py:36-61) + HF eager_attention_forward
 * (transformers models/llama/modeling_llama.py:199-221) for one layer:
 *     S[q,k] = scale * <Q[q], K[k]>  (+ logcred[k] if head in head_mask, k < bias_len, k visible)
 *     P = softmax over visible k (causal: k <= q + (len_k - len_q)),  O = P V
 * which equals re-weighting the clean probabilities by cred[k] and renormalising (README.md:21).
 * Rows with no visible key of finite logit produce 0 (the reference never creates such rows).
 *
 * Layout: packed variable-length rows. Token t / head h lives at base + t*row_stride + h*head_dim
 * (strides in elements), so Q, K, V can alias one fused QKV projection buffer. GQA: query head h
 * reads KV head h / (n_heads / n_kv_heads).
 */
typedef struct cram_attn_prefill_params {
  const void* q;
  const void* k;
  const void* v;
  void* out;                  /* [total_q, n_heads*head_dim], row stride out_row_stride */
  float* lse;                 /* optional [n_heads, total_q], natural-log sum-exp of the biased logits */
  int64_t q_row_stride, k_row_stride, v_row_stride, out_row_stride;
  const int32_t* cu_seqlens_q; /* [n_seq+1] */
  const int32_t* cu_seqlens_k; /* [n_seq+1] or NULL (= cu_seqlens_q) */
  int32_t n_seq, total_q, total_k, max_seqlen_q, max_seqlen_k;
  int32_t n_heads, n_kv_heads, head_dim;
  int32_t dtype;              /* of q, k, v, out */
  float scale;
  const void* logcred;        /* packed like K rows: [total_k], or NULL for clean attention */
  int32_t logcred_dtype;
  uint64_t head_mask;         /* bit h set => query head h carries the bias */
  const int32_t* bias_len;    /* [n_seq] keys < bias_len[s] are biased; NULL => every key */
  int32_t causal;
  int32_t kernel;             /* 0 = auto, 1 = generic CUDA-core kernel, 2 = tcgen05 kernel */
} cram_attn_prefill_params;

int cram_attn_prefill(const cram_attn_prefill_params* p, cram_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (1) decode step over a paged KV cache (one new query token per sequence), split-KV + combine.
 * Same arithmetic as above with len_q = 1; the bias is indexed by absolute key position, only
 * keys < bias_len[s] (the prompt) carry it -- utils/re_weighting.py:57 slices [..., :weight_len].
 * Page p, slot j, kv head g at cache + p*page_stride + j*token_stride + g*head_dim.
 */
typedef struct cram_attn_decode_params {
  const void* q;              /* [n_seq, n_heads*head_dim] */
  void* out;                  /* [n_seq, n_heads*head_dim] */
  int64_t q_row_stride, out_row_stride;
  const void* k_cache;
  const void* v_cache;
  int64_t page_stride, token_stride;
  const int32_t* block_table; /* [n_seq, block_table_stride] page ids */
  int32_t block_table_stride;
  const int32_t* seqlens;     /* [n_seq] keys visible to the query (includes the new token) */
  int32_t n_seq, page_size, max_seqlen;
  int32_t n_heads, n_kv_heads, head_dim;
  int32_t dtype;
  float scale;
  const void* logcred;        /* [n_seq, logcred_stride] or NULL */
  int64_t logcred_stride;
  int32_t logcred_dtype;
  uint64_t head_mask;
  const int32_t* bias_len;    /* [n_seq] */
  int32_t n_splits;           /* 0 = choose */
  void* workspace;            /* >= cram_attn_decode_workspace_bytes(...) */
  size_t workspace_bytes;
} cram_attn_decode_params;

size_t cram_attn_decode_workspace_bytes(int n_seq, int n_heads, int head_dim, int max_seqlen, int n_splits);
int cram_attn_decode(const cram_attn_decode_params* p, cram_stream_t stream);

/* Allocation-free variant of cram_attn_decode. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same inputs, outputs and arithmetic as cram_attn_decode; p->workspace / p->workspace_bytes are ignored and
 * the split-KV partials go into a static device array of CRAM_ATTN_DECODE_STATIC_SCRATCH_BYTES inside the library (every
 * GPU that loads the library holds it until the process exits). It calls no cudaMalloc / cudaFree / allocating routine.
 * A launch that needs more (n_seq * n_heads * n_splits * (head_dim + 2) * 4 bytes, n_splits as given or as chosen
 * automatically) is rejected with CRAM_ERR_WORKSPACE and a message in cram_last_error(). All calls on a device share the
 * array: calls that overlap on the GPU (different streams, no ordering between them) race -- keep them on one stream.
 * The array keeps the previous call's contents; the kernels overwrite every partial they later read, nothing needs
 * clearing. */
#define CRAM_ATTN_DECODE_STATIC_SCRATCH_BYTES (32u << 20)
int cram_attn_decode_static(const cram_attn_decode_params* p, cram_stream_t stream);

/* scatter freshly projected K/V rows into the paged cache: row r goes to slot_mapping[r]
 * (= page*page_size + slot; negative = skip). */
int cram_kv_append(const void* k, const void* v, int64_t k_row_stride, int64_t v_row_stride,
                   int64_t n_rows, int n_kv_heads, int head_dim, int dtype, const int64_t* slot_mapping,
                   void* k_cache, void* v_cache, int64_t page_stride, int64_t token_stride, int page_size,
                   cram_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (3) head-intervention sweep kernels. Replace the body of Find_Best_Heads.cal_logits
 * (utils/re_weighting.py:152-175): instead of L*H full forwards, layer l's H single-head variants
 * are built from ONE all-heads-biased attention launch and the cached clean attention output, then
 * run as a batch through the suffix layers.
 *
 * cram_sweep_splice_heads: out[v] = clean, except head heads[v]'s head_dim columns taken from
 * biased. clean/biased are [n_rows, n_heads*head_dim]; out is [n_var, n_rows, n_heads*head_dim].
 */
int cram_sweep_splice_heads(const void* clean, const void* biased, const int32_t* heads, int n_var,
                            int64_t n_rows, int n_heads, int head_dim, int dtype, void* out,
                            cram_stream_t stream);

/* cram_sweep_answer_ie: the fused epilogue of a swept forward (utils/re_weighting.py:160,170-171).
 * For variant v, rows T-n..T-1 of its final residual stream x[v] ([n_var, T, hidden]): final RMSNorm,
 * dot with the lm_head rows of answer_ids[k], round to model dtype, sum over k in model dtype
 * (cur[v]), and ie[v] = model_dtype(ori - cur[v]). `ori` (device scalar, f32 holding a model-dtype
 * value) may be NULL, then only cur is produced (that is how the clean pass makes `ori`). */
int cram_sweep_answer_ie(const void* x, int n_var, int64_t T, int hidden, const void* norm_weight,
                         float rms_eps, const void* lm_head, const int64_t* answer_ids, int n_answer,
                         int dtype, const float* ori, float* cur, float* ie, cram_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Layer glue (HBM-bound elementwise work around the library GEMMs), written to round exactly where
 * HF's eager modules round (transformers models/llama/modeling_llama.py:52-70 rmsnorm, :138-176
 * rotary, :178-186 MLP act) so the host forward matches the hooked reference within bf16 noise.
 */
/* out = weight * dtype(x * rsqrt(mean(x^2) + eps));  if `residual` != NULL first x = x + residual
 * (written back to x, rounded to dtype) -- the decoder layer's residual add fused with the norm. */
int cram_rmsnorm(void* x, const void* residual, const void* weight, void* out, int64_t n_rows, int hidden,
                 float eps, int dtype, cram_stream_t stream);
/* in-place rotary embedding on the Q and K heads of a fused QKV buffer; cos/sin are [n_pos, head_dim]
 * tables in `dtype`, positions[r] indexes them. */
int cram_rope(void* qk, int64_t row_stride, int64_t n_rows, int n_rot_heads, int head_dim,
              const void* cos_table, const void* sin_table, const int32_t* positions, int dtype,
              cram_stream_t stream);
/* cram_rope on the Q and K heads of a fused [.., (n_heads + 2 n_kv_heads) * head_dim] row AND cram_kv_append of
 * the rotated K and of V, in one launch (the decode step's per-layer glue). */
int cram_rope_kv_append(void* qkv, int64_t row_stride, int64_t n_rows, int n_heads, int n_kv_heads, int head_dim,
                        const void* cos_table, const void* sin_table, const int32_t* positions,
                        const int64_t* slot_mapping, void* k_cache, void* v_cache, int64_t page_stride,
                        int64_t token_stride, int page_size, int dtype, cram_stream_t stream);
/* Decode-step projection (<= 64 token rows) with the step's glue fused in; streams w once (HBM-bound).
 *   mode 0: y[m, n] = sum_k x[m,k] w[n,k] (+ bias[n])        w = [n_out, k]
 *   mode 1: y[m, n] += T(sum_k x[m,k] w[n,k])                 the residual add of o_proj / down_proj, in place
 *   mode 2: y[m, n] = T(T(silu(T(g))) * T(u)), g = x . w[n], u = x . w[n_out + n]     w = [2 n_out, k] (gate, up):
 *           LlamaMLP act_fn(gate_proj(x)) * up_proj(x) (transformers modeling_llama.py LlamaMLP.forward)
 * bf16 / fp16; k and the row strides multiples of 8. Split-K runs inside a thread-block cluster (reduction in
 * distributed shared memory): no workspace, results independent of timing. */
int cram_linear_decode(const void* x, int64_t ldx, int m, const void* w, int64_t ldw, int n_out, int k,
                       const void* bias, void* y, int64_t ldy, int mode, int dtype, cram_stream_t stream);
/* out[r, j] = silu(gate_up[r, j]) * gate_up[r, inter + j] */
int cram_silu_mul(const void* gate_up, void* out, int64_t n_rows, int inter, int dtype, cram_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Host-side ranking (utils/find_best_heads.py:113-150). HOST pointers. scores is [n_samples, L, H]
 * float64; order_out receives flat indices l*H+h sorted by mean descending, ties by flat index
 * ascending; mean_out the float64 means in that order. */
int cram_rank_heads(const double* scores, int n_samples, int n_layers, int n_heads, int32_t* order_out,
                    double* mean_out);

#ifdef __cplusplus
}
#endif
#endif /* CRAM_B200_H_ */
