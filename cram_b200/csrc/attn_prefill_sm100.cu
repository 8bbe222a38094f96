// Kernel (1): fused credibility-biased FlashAttention-style prefill for sm_100a.
//
//   S = Q K^T      tcgen05.mma (SS), 128x128 fp32 accumulator in TMEM, operands TMA-staged in
//                  128B-swizzled shared memory
//   online softmax one thread per query row straight out of TMEM (tcgen05.ld 32x32b); the per-key
//                  log-credibility bias (already x log2e, with the key-length mask folded in as
//                  -inf) is one FFMA: t = S*scale*log2e + bias[k]
//   O += P V       tcgen05.mma (TS): P is written back to TMEM as bf16/fp16 over the S columns and
//                  consumed as the A operand; V is the MN-major B operand from shared memory
//
// This replaces the reference's dense [b,H,q,k] mask rewrite (utils/re_weighting.py:36-61) plus HF
// eager attention (modeling_llama.py:199-221): the bias is O(T) instead of O(H*T^2) and S/P never
// touch HBM. Biasing head h by logcred[k] before the softmax equals multiplying its probabilities
// by cred[k] and renormalising (README.md:21).
//
// CTA = 128 query rows of one (sequence, head); 6 warps:
//   warps 0-3  softmax + correction + epilogue (TMEM lanes 32w..32w+31)
//   warp  4    producer: TMA loads of Q, K[j], V[j]; builds the per-tile bias vector
//   warp  5    MMA issuer (one elected lane); owns the TMEM allocation
// Pipelines (all mbarrier based): K ring x2 and V ring x2 (freed by tcgen05.commit of the MMA that
// read them), S/P double buffer in TMEM so QK^T of tile j+1 overlaps the softmax of tile j, and a
// lazy O rescale (only when a row's max grows by more than 2^8) so the correction rarely touches TMEM.
#include <mutex>

#include "attn.cuh"
#include "sm100_ptx.cuh"

namespace cram {

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockN = 128;
constexpr int kThreads = 192;
constexpr int kTmemCols = 512;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.0f;  // log2 units

struct Barriers {
  uint64_t q_full;
  uint64_t k_full[2], k_empty[2];
  uint64_t v_full[2], v_empty[2];
  uint64_t bias_full[2];
  uint64_t s_full[2], p_full[2];
  uint64_t o_done;
};

template <int DH>
struct Smem {
  static constexpr int kTileBytes = kBlockN * DH * 2;  // one Q / K / V tile
  static constexpr int kSlabBytes = kBlockN * 128;     // 128 rows x 64 elements (one 128B swizzle slab)
  static constexpr int kQ = 0;
  static constexpr int kK = kQ + kTileBytes;
  static constexpr int kV = kK + 2 * kTileBytes;
  static constexpr int kBias = kV + 2 * kTileBytes;
  static constexpr int kBar = kBias + 2 * kBlockN * 4;
  static constexpr int kTmemPtr = kBar + (int)sizeof(Barriers);
  static constexpr int kEnd = kTmemPtr + 16;
  // dynamic smem request: + 1024 for manual alignment; padded above half an SM so that exactly one
  // CTA (and therefore one 512-column TMEM allocation) is resident per SM.
  static constexpr int kRequest = (kEnd + 1024) > 120 * 1024 ? (kEnd + 1024) : 120 * 1024;
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

template <typename T, int DH>
__global__ void __launch_bounds__(kThreads, 1)
attn_prefill_sm100_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                          const __grid_constant__ CUtensorMap tmap_v, const cram_attn_prefill_params p,
                          const int num_m_tiles) {
  using S = Smem<DH>;
  constexpr int kSlabs = DH / 64;
  constexpr uint32_t kAbFmt = DType<T>::code == CRAM_BF16 ? 1u : 0u;
  constexpr uint32_t kIdescQK = ptx::umma_idesc_f16(kAbFmt, kBlockM, kBlockN, 0);
  constexpr uint32_t kIdescPV = ptx::umma_idesc_f16(kAbFmt, kBlockM, DH, 1);

  // ---- which tile ------------------------------------------------------------------------------
  const int head = blockIdx.x % p.n_heads;
  const int rest = blockIdx.x / p.n_heads;
  const int seq = rest % p.n_seq;
  const int m_tile = num_m_tiles - 1 - rest / p.n_seq;  // heaviest (latest) query tiles first
  const int q0 = p.cu_seqlens_q[seq], len_q = p.cu_seqlens_q[seq + 1] - q0;
  if (m_tile * kBlockM >= len_q) return;
  const int32_t* cuk = p.cu_seqlens_k ? p.cu_seqlens_k : p.cu_seqlens_q;
  const int k0 = cuk[seq], len_k = cuk[seq + 1] - k0;
  const int diag = len_k - len_q;  // query i sees keys <= i + diag
  const int kv_head = head / (p.n_heads / p.n_kv_heads);
  int kv_end = len_k;
  if (p.causal) kv_end = min(len_k, m_tile * kBlockM + kBlockM + diag);
  const int n_tiles = kv_end > 0 ? (kv_end + kBlockN - 1) / kBlockN : 0;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Barriers* bar = reinterpret_cast<Barriers*>(smem + S::kBar);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + S::kTmemPtr);
  float* s_bias = reinterpret_cast<float*>(smem + S::kBias);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 4 && lane == 0) {
    ptx::prefetch_tmap(&tmap_q);
    ptx::prefetch_tmap(&tmap_k);
    ptx::prefetch_tmap(&tmap_v);
    ptx::mbar_init(&bar->q_full, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&bar->k_full[i], 1);
      ptx::mbar_init(&bar->k_empty[i], 1);
      ptx::mbar_init(&bar->v_full[i], 1);
      ptx::mbar_init(&bar->v_empty[i], 1);
      ptx::mbar_init(&bar->bias_full[i], 1);
      ptx::mbar_init(&bar->s_full[i], 1);
      ptx::mbar_init(&bar->p_full[i], 128);
    }
    ptx::mbar_init(&bar->o_done, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 5) ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_o = tmem_base + 2 * kBlockN;

  if (warp == 4) {
    // ============================== producer ====================================================
    if (n_tiles > 0) {
      const bool biased = p.logcred != nullptr && ((p.head_mask >> head) & 1ull);
      const int bias_len = p.bias_len ? p.bias_len[seq] : len_k;
      if (lane == 0) {
        ptx::mbar_expect_tx(&bar->q_full, S::kTileBytes);
#pragma unroll
        for (int s = 0; s < kSlabs; ++s)
          ptx::tma_load_2d(smem + S::kQ + s * S::kSlabBytes, &tmap_q, &bar->q_full, head * DH + s * 64,
                           q0 + m_tile * kBlockM);
      }
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        if (lane == 0) {
          ptx::mbar_wait(&bar->k_empty[st], ph ^ 1);
          ptx::mbar_expect_tx(&bar->k_full[st], S::kTileBytes);
#pragma unroll
          for (int s = 0; s < kSlabs; ++s)
            ptx::tma_load_2d(smem + S::kK + st * S::kTileBytes + s * S::kSlabBytes, &tmap_k, &bar->k_full[st],
                             kv_head * DH + s * 64, k0 + j * kBlockN);
        }
        // V slot and bias slot st are free once P[j-2] V[j-2] has completed.
        ptx::mbar_wait(&bar->v_empty[st], ph ^ 1);
        {
          float4 b4;
          float* bp = reinterpret_cast<float*>(&b4);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int key = j * kBlockN + lane * 4 + e;
            float v = 0.f;
            if (key >= len_k) v = -INFINITY;
            else if (biased && key < bias_len) v = load_logcred(p.logcred, (int64_t)k0 + key, p.logcred_dtype) * kLog2e;
            bp[e] = v;
          }
          reinterpret_cast<float4*>(s_bias + st * kBlockN)[lane] = b4;
        }
        __syncwarp();
        if (lane == 0) {
          ptx::mbar_arrive(&bar->bias_full[st]);
          ptx::mbar_expect_tx(&bar->v_full[st], S::kTileBytes);
#pragma unroll
          for (int s = 0; s < kSlabs; ++s)
            ptx::tma_load_2d(smem + S::kV + st * S::kTileBytes + s * S::kSlabBytes, &tmap_v, &bar->v_full[st],
                             kv_head * DH + s * 64, k0 + j * kBlockN);
        }
      }
    }
  } else if (warp == 5) {
    // ============================== MMA issuer ==================================================
    if (lane == 0 && n_tiles > 0) {
      const uint32_t q_addr = ptx::smem_u32(smem + S::kQ);
      const uint32_t k_addr = ptx::smem_u32(smem + S::kK);
      const uint32_t v_addr = ptx::smem_u32(smem + S::kV);
      auto issue_qk = [&](int j) {
        const int st = j & 1;
        ptx::mbar_wait(&bar->k_full[st], (j >> 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < DH / 16; ++kk) {
          const uint32_t off = (kk >> 2) * S::kSlabBytes + (kk & 3) * 32;
          ptx::mma_ss(tmem_base + st * kBlockN, ptx::umma_desc_sw128(q_addr + off, 16, 1024),
                      ptx::umma_desc_sw128(k_addr + st * S::kTileBytes + off, 16, 1024), kIdescQK, kk > 0);
        }
        ptx::tc_commit(&bar->k_empty[st]);
        ptx::tc_commit(&bar->s_full[st]);
      };
      ptx::mbar_wait(&bar->q_full, 0);
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        if (j + 1 < n_tiles) issue_qk(j + 1);
        ptx::mbar_wait(&bar->p_full[st], ph);
        ptx::mbar_wait(&bar->v_full[st], ph);
        ptx::tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kBlockN / 16; ++kk) {
          ptx::mma_ts(tmem_o, tmem_base + st * kBlockN + kk * 8,
                      ptx::umma_desc_sw128(v_addr + st * S::kTileBytes + kk * 2048, S::kSlabBytes, 1024), kIdescPV,
                      (j > 0 || kk > 0) ? 1u : 0u);
        }
        ptx::tc_commit(&bar->v_empty[st]);
        ptx::tc_commit(&bar->o_done);
      }
    }
  } else {
    // ============================== softmax / correction / epilogue ================================
    const int row = warp * 32 + lane;             // row inside the tile == TMEM lane
    const int qrow = m_tile * kBlockM + row;      // row inside the sequence
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    const float scale2 = p.scale * kLog2e;
    float m_ref = -INFINITY, l = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      ptx::mbar_wait(&bar->bias_full[st], ph);
      ptx::mbar_wait(&bar->s_full[st], ph);
      ptx::tc_fence_after();
      uint32_t s[4][32];
      const uint32_t s_addr = tmem_base + lane_off + st * kBlockN;
#pragma unroll
      for (int c = 0; c < 4; ++c) ptx::tmem_ld32(s_addr + c * 32, s[c]);
      ptx::tc_wait_ld();

      const float4* bias4 = reinterpret_cast<const float4*>(s_bias + st * kBlockN);
      const bool diag_tile = p.causal && (j * kBlockN + kBlockN - 1 > m_tile * kBlockM + diag);
      const int climit = qrow + diag - j * kBlockN;  // columns <= climit are visible
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 b = bias4[c * 8 + i / 4];
          float t0 = fmaf(__uint_as_float(s[c][i + 0]), scale2, b.x);
          float t1 = fmaf(__uint_as_float(s[c][i + 1]), scale2, b.y);
          float t2 = fmaf(__uint_as_float(s[c][i + 2]), scale2, b.z);
          float t3 = fmaf(__uint_as_float(s[c][i + 3]), scale2, b.w);
          if (diag_tile) {
            const int col = c * 32 + i;
            t0 = (col + 0 <= climit) ? t0 : -INFINITY;
            t1 = (col + 1 <= climit) ? t1 : -INFINITY;
            t2 = (col + 2 <= climit) ? t2 : -INFINITY;
            t3 = (col + 3 <= climit) ? t3 : -INFINITY;
          }
          s[c][i + 0] = __float_as_uint(t0);
          s[c][i + 1] = __float_as_uint(t1);
          s[c][i + 2] = __float_as_uint(t2);
          s[c][i + 3] = __float_as_uint(t3);
          mx0 = fmaxf(mx0, t0);
          mx1 = fmaxf(mx1, t1);
          mx2 = fmaxf(mx2, t2);
          mx3 = fmaxf(mx3, t3);
        }
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));

      // lazy rescale: move the reference max only when this tile beats it by more than 2^8
      const bool need = mx > m_ref + kRescaleThreshold;
      const float m_new = need ? mx : m_ref;
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        const float f = need ? ex2(m_ref - m_new) : 1.0f;  // m_ref = -inf gives 0
        ptx::mbar_wait(&bar->o_done, (j - 1) & 1);         // P[j-1] V[j-1] has landed in O
        ptx::tc_fence_after();
#pragma unroll
        for (int c = 0; c < DH / 32; ++c) {
          uint32_t o[32];
          ptx::tmem_ld32(tmem_o + lane_off + c * 32, o);
          ptx::tc_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
          ptx::tmem_st32(tmem_o + lane_off + c * 32, o);
        }
        l *= f;
      }
      m_ref = m_new;
      const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;

      float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t pk[32];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = half * 2 + cc;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float p0 = ex2(__uint_as_float(s[c][i + 0]) - m_use);
            const float p1 = ex2(__uint_as_float(s[c][i + 1]) - m_use);
            const float p2 = ex2(__uint_as_float(s[c][i + 2]) - m_use);
            const float p3 = ex2(__uint_as_float(s[c][i + 3]) - m_use);
            l0 += p0;
            l1 += p1;
            l2 += p2;
            l3 += p3;
            pk[cc * 16 + i / 2 + 0] = pack2<T>(p0, p1);
            pk[cc * 16 + i / 2 + 1] = pack2<T>(p2, p3);
          }
        }
        ptx::tmem_st32(s_addr + half * 32, pk);  // P over the S columns: 64 keys -> 32 packed columns
      }
      l += (l0 + l1) + (l2 + l3);
      ptx::tc_wait_st();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&bar->p_full[st]);
    }

    // ---- epilogue: O / l -> global ---------------------------------------------------------------
    const bool row_ok = qrow < len_q;
    T* out = reinterpret_cast<T*>(p.out) + (int64_t)(q0 + qrow) * p.out_row_stride + (int64_t)head * DH;
    if (n_tiles > 0) {
      ptx::mbar_wait(&bar->o_done, (n_tiles - 1) & 1);
      ptx::tc_fence_after();
      const float inv = l > 0.f ? 1.0f / l : 0.f;
#pragma unroll
      for (int c = 0; c < DH / 32; ++c) {
        uint32_t o[32];
        ptx::tmem_ld32(tmem_o + lane_off + c * 32, o);
        ptx::tc_wait_ld();
        if (row_ok) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 v;
            v.x = pack2<T>(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
            v.y = pack2<T>(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
            v.z = pack2<T>(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
            v.w = pack2<T>(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
            *reinterpret_cast<uint4*>(out + c * 32 + i) = v;
          }
        }
      }
      if (p.lse && row_ok)
        p.lse[(int64_t)head * p.total_q + q0 + qrow] = l > 0.f ? (m_ref + log2f(l)) * 0.6931471805599453f : -INFINITY;
    } else if (row_ok) {
#pragma unroll
      for (int c = 0; c < DH; c += 8) *reinterpret_cast<uint4*>(out + c) = make_uint4(0, 0, 0, 0);
      if (p.lse) p.lse[(int64_t)head * p.total_q + q0 + qrow] = -INFINITY;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 5) ptx::tmem_dealloc(tmem_base, kTmemCols);
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

// rows x cols view of packed token rows; box = 128 rows x 64 elements, 128-byte swizzle.
int make_tmap(CUtensorMap* m, const void* base, int dtype, int64_t cols, int64_t rows, int64_t row_stride_elems) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return CRAM_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)row_stride_elems * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)kBlockN};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, dtype == CRAM_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (base %p cols %lld rows %lld stride %lld)", (int)r, base,
              (long long)cols, (long long)rows, (long long)row_stride_elems);
    return CRAM_ERR_CUDA;
  }
  return CRAM_OK;
}

template <typename T, int DH>
int launch_t(const cram_attn_prefill_params& p, cudaStream_t stream) {
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = make_tmap(&tq, p.q, p.dtype, (int64_t)p.n_heads * DH, p.total_q, p.q_row_stride)) != CRAM_OK) return rc;
  if ((rc = make_tmap(&tk, p.k, p.dtype, (int64_t)p.n_kv_heads * DH, p.total_k, p.k_row_stride)) != CRAM_OK) return rc;
  if ((rc = make_tmap(&tv, p.v, p.dtype, (int64_t)p.n_kv_heads * DH, p.total_k, p.v_row_stride)) != CRAM_OK) return rc;
  const int num_m_tiles = (p.max_seqlen_q + kBlockM - 1) / kBlockM;
  const long long grid = (long long)num_m_tiles * p.n_seq * p.n_heads;
  if (grid == 0) return CRAM_OK;
  CRAM_REQUIRE(grid < (1ll << 31), "cram_attn_prefill: grid too large");
  auto kern = attn_prefill_sm100_kernel<T, DH>;
  static bool attr_set = false;  // per template instantiation
  if (!attr_set) {
    CRAM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<DH>::kRequest));
    attr_set = true;
  }
  kern<<<(unsigned)grid, kThreads, Smem<DH>::kRequest, stream>>>(tq, tk, tv, p, num_m_tiles);
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
