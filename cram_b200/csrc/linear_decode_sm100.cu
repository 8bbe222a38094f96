// Decode-step projections for sm_100a: y[m, n] = sum_k x[m, k] W[n, k] for m <= 64 rows (one new token per
// sequence), with the step's element-wise glue fused into the epilogue. This is the weight-streaming half of the
// greedy decode loop behind run_RAG_with_attention_weighting (utils/re_weighting.py:127-135 -> HF generate ->
// modeling_llama.py LlamaAttention q/k/v/o_proj and LlamaMLP down_proj(act_fn(gate_proj(x)) * up_proj(x))).
//
// HBM-bound (every weight byte is read once per token step, 2 x 64 FLOP per byte), so the design is a streaming one:
//   * swap-AB: the 128 weight rows of a tile are the UMMA M dimension, the <= 64 token rows the N dimension:
//     tcgen05.mma.kind::f16 M128 N64 K16, both operands K-major in 128B-swizzled shared memory, fp32 accumulator
//     in 64 TMEM columns;
//   * one producer lane feeds a 4-stage ring with TMA (16 KB of weights + 8 KB of x per 64-wide K block; x rows
//     beyond m are zero-filled by the tensor map), one elected lane issues the MMAs, tcgen05.commit frees ring slots;
//   * 2 CTAs per SM (2 x 96 KB of ring = 128 KB of weights in flight per SM); grid = weight tiles x split-K, the
//     split chosen on the host so that the grid fills 2 x #SM slots; split-K partials go through an fp32
//     workspace and the LAST CTA of a tile (atomic ticket) sums them in split order -- deterministic;
//   * epilogues: 0 = (+ bias) -> y; 1 = y += T(acc) (the residual add of o_proj / down_proj, in place);
//     2 = SwiGLU: the tile holds 64 gate rows and their 64 up rows, out = T(T(silu(T(gate))) * T(up)) -- the
//     roundings of cram_silu_mul / HF eager.
#include <mutex>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace cram {

namespace {

constexpr int kTileN = 128;   // weight rows per tile (UMMA M, TMEM lanes)
constexpr int kRows = 64;     // token rows (UMMA N, TMEM columns)
constexpr int kBlockK = 64;   // one 128-byte swizzle slab of bf16/fp16
constexpr int kRing = 4;
constexpr int kWBytes = kTileN * kBlockK * 2;  // 16 KB
constexpr int kXBytes = kRows * kBlockK * 2;   // 8 KB
constexpr int kStageBytes = kWBytes + kXBytes;
constexpr int kThreads = 192;  // warps 0-3 epilogue, warp 4 producer, warp 5 MMA issuer / TMEM owner
constexpr int kTmemCols = 64;
constexpr int kMaxTiles = 4096;
constexpr uint32_t kDescHi = ptx::umma_desc_hi(1024);

struct Bars {
  uint64_t full[kRing], empty[kRing];
  uint64_t acc_full;
  uint32_t tmem_ptr;
  int is_last;
};
constexpr int kSmemRequest = kRing * kStageBytes + (int)sizeof(Bars) + 1024;

struct Params {
  void* y;
  int64_t ldy;
  const void* bias;
  int m, n_out, k_blocks, split;
  int up_row0;  // mode 2: first up row (= n_out)
  float* partials;
  int* counters;
};

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <typename T>
__device__ __forceinline__ float rnd(float v) {
  return DType<T>::to_f(DType<T>::from_f(v));
}

template <typename T, int kMode>
__global__ void __launch_bounds__(kThreads, 2)
linear_decode_sm100_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                           const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Bars* bar = reinterpret_cast<Bars*>(smem + kRing * kStageBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x / p.split, sp = blockIdx.x - tile * p.split;
  // K blocks [kb0, kb1) of this split: the first `rem` splits take one more
  const int per = p.k_blocks / p.split, rem = p.k_blocks - per * p.split;
  const int kb0 = sp * per + min(sp, rem), n_kb = per + (sp < rem ? 1 : 0);
  // weight row of TMEM lane 0 / of lane 64 (mode 2: gate rows then their up rows)
  const int row_lo = kMode == 2 ? tile * 64 : tile * kTileN;
  const int row_hi = kMode == 2 ? p.up_row0 + tile * 64 : tile * kTileN + 64;

  if (warp == 4 && lane == 0) {
    ptx::prefetch_tmap(&tmap_w);
    ptx::prefetch_tmap(&tmap_x);
    for (int i = 0; i < kRing; ++i) {
      ptx::mbar_init(&bar->full[i], 1);
      ptx::mbar_init(&bar->empty[i], 1);
    }
    ptx::mbar_init(&bar->acc_full, 1);
    ptx::fence_mbar_init();
    // The weights do not depend on the previous kernel of the stream: the first ring of W tiles goes out before
    // the grid dependency resolves (and before the TMEM allocation / CTA sync), so under programmatic dependent
    // launch the HBM stream of this kernel starts while the previous kernel is still draining.
    ptx::griddep_launch_dependents();
    for (int i = 0; i < n_kb && i < kRing; ++i) {
      ptx::mbar_expect_tx(&bar->full[i], kStageBytes);
      uint8_t* w_dst = smem + i * kStageBytes;
      const int kc = (kb0 + i) * kBlockK;
      ptx::tma_load_2d(w_dst, &tmap_w, &bar->full[i], kc, row_lo);               // TMEM lanes 0-63
      ptx::tma_load_2d(w_dst + kWBytes / 2, &tmap_w, &bar->full[i], kc, row_hi);  // TMEM lanes 64-127
    }
  }
  if (warp == 5) ptx::tmem_alloc(&bar->tmem_ptr, kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = bar->tmem_ptr;
  ptx::griddep_wait();  // x, y, the split-K workspace: everything below may depend on the previous kernel

  if (warp == 4) {
    if (lane == 0) {
      for (int i = 0; i < n_kb && i < kRing; ++i)
        ptx::tma_load_2d(smem + i * kStageBytes + kWBytes, &tmap_x, &bar->full[i], (kb0 + i) * kBlockK, 0);
      for (int i = kRing; i < n_kb; ++i) {
        const int st = i % kRing;
        ptx::mbar_wait(&bar->empty[st], ((i / kRing) & 1) ^ 1);
        ptx::mbar_expect_tx(&bar->full[st], kStageBytes);
        uint8_t* w_dst = smem + st * kStageBytes;
        const int kc = (kb0 + i) * kBlockK;
        ptx::tma_load_2d(w_dst, &tmap_w, &bar->full[st], kc, row_lo);
        ptx::tma_load_2d(w_dst + kWBytes / 2, &tmap_w, &bar->full[st], kc, row_hi);
        ptx::tma_load_2d(w_dst + kWBytes, &tmap_x, &bar->full[st], kc, 0);
      }
    }
  } else if (warp == 5) {
    if (ptx::elect_one_sync()) {
      constexpr uint32_t kAbFmt = DType<T>::code == CRAM_BF16 ? 1u : 0u;
      constexpr uint32_t kIdesc = ptx::umma_idesc_f16(kAbFmt, kTileN, kRows, 0);
      const uint32_t w_lo = ptx::umma_desc_lo(ptx::smem_u32(smem), 16);
      const uint32_t x_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + kWBytes), 16);
      for (int i = 0; i < n_kb; ++i) {
        const int st = i % kRing;
        ptx::mbar_wait(&bar->full[st], (i / kRing) & 1);
        ptx::tc_fence_after();
        const uint32_t a = w_lo + st * (kStageBytes >> 4), b = x_lo + st * (kStageBytes >> 4);
        ptx::mma_ss_imm<kDescHi, 0, 0, 0>(tmem, a, b, kIdesc, i > 0 ? 1u : 0u);
        ptx::mma_ss_imm<kDescHi, 0, 2, 2>(tmem, a, b, kIdesc, 1u);
        ptx::mma_ss_imm<kDescHi, 0, 4, 4>(tmem, a, b, kIdesc, 1u);
        ptx::mma_ss_imm<kDescHi, 0, 6, 6>(tmem, a, b, kIdesc, 1u);
        ptx::tc_commit(&bar->empty[st]);
      }
      ptx::tc_commit(&bar->acc_full);
    }
  } else {
    // ============================== epilogue (warps 0-3, TMEM lanes 32w..32w+31) ======================
    const int nl = warp * 32 + lane;  // TMEM lane = weight row inside the tile
    float acc[kRows];
    if (n_kb > 0) {
      ptx::mbar_wait(&bar->acc_full, 0);
      ptx::tc_fence_after();
      uint32_t r0[32], r1[32];
      const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
      ptx::tmem_ld32(taddr, r0);
      ptx::tmem_ld32(taddr + 32, r1);
      ptx::tc_wait_ld();
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        acc[i] = __uint_as_float(r0[i]);
        acc[32 + i] = __uint_as_float(r1[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < kRows; ++i) acc[i] = 0.f;
    }
    bool finish = true;
    if (p.split > 1) {
      // partial tile layout [m / 4][weight row][4]: consecutive lanes write consecutive 16-byte pieces
      float4* mine = reinterpret_cast<float4*>(p.partials + (int64_t)blockIdx.x * kTileN * kRows) + nl;
#pragma unroll
      for (int i = 0; i < kRows / 4; ++i)
        __stcg(mine + i * kTileN, make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]));
      __threadfence();
      epi_bar();
      if (threadIdx.x == 0) {
        const int ticket = atomicAdd(p.counters + tile, 1);
        bar->is_last = ticket == p.split - 1;
        if (ticket == p.split - 1) p.counters[tile] = 0;  // ready for the next launch (stream order)
      }
      epi_bar();
      finish = bar->is_last != 0;
      if (finish) {
        __threadfence();
#pragma unroll
        for (int i = 0; i < kRows; ++i) acc[i] = 0.f;
        for (int s = 0; s < p.split; ++s) {  // fixed order: the sum does not depend on which CTA came last
          const float4* src =
              reinterpret_cast<const float4*>(p.partials + ((int64_t)tile * p.split + s) * kTileN * kRows) + nl;
#pragma unroll
          for (int i = 0; i < kRows / 4; ++i) {
            const float4 v = __ldcg(src + i * kTileN);
            acc[4 * i] += v.x;
            acc[4 * i + 1] += v.y;
            acc[4 * i + 2] += v.z;
            acc[4 * i + 3] += v.w;
          }
        }
      }
    }
    if (finish) {
      T* y = reinterpret_cast<T*>(p.y);
      if (kMode == 2) {
        // lanes 64-127 hold the up rows of the gate rows in lanes 0-63: hand them over through the (now idle) ring
        T* s_up = reinterpret_cast<T*>(smem);  // [m][64]
        if (nl >= 64) {
#pragma unroll
          for (int m = 0; m < kRows; ++m) s_up[m * 64 + (nl - 64)] = DType<T>::from_f(acc[m]);
        }
        epi_bar();
        const int n = tile * 64 + nl;
        if (nl < 64 && n < p.n_out) {
#pragma unroll
          for (int m = 0; m < kRows; ++m) {
            if (m < p.m) {
              const float g = rnd<T>(acc[m]);
              const float s = rnd<T>(g / (1.0f + expf(-g)));
              y[(int64_t)m * p.ldy + n] = DType<T>::from_f(s * DType<T>::to_f(s_up[m * 64 + nl]));
            }
          }
        }
      } else {
        const int n = tile * kTileN + nl;
        if (n < p.n_out) {
          const float b = (kMode == 0 && p.bias) ? DType<T>::to_f(reinterpret_cast<const T*>(p.bias)[n]) : 0.f;
          if (kMode == 1) {
#pragma unroll
            for (int m0 = 0; m0 < kRows; m0 += 16) {  // 16 loads in flight, then their 16 stores
              float res[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) res[i] = m0 + i < p.m ? DType<T>::to_f(y[(int64_t)(m0 + i) * p.ldy + n]) : 0.f;
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (m0 + i < p.m) y[(int64_t)(m0 + i) * p.ldy + n] = DType<T>::from_f(res[i] + rnd<T>(acc[m0 + i]));
            }
          } else {
#pragma unroll
            for (int m = 0; m < kRows; ++m)
              if (m < p.m) y[(int64_t)m * p.ldy + n] = DType<T>::from_f(acc[m] + b);
          }
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 5) ptx::tmem_dealloc(tmem, kTmemCols);
}

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

// [rows x cols] K-contiguous matrix, box = 64 rows x 64 elements (one 128-byte swizzle slab), zero fill outside.
int tmap_rows_by_k(CUtensorMap* m, const void* base, int dtype, int64_t rows, int64_t cols, int64_t ld) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return CRAM_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {kBlockK, 64};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, dtype == CRAM_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (base %p rows %lld cols %lld ld %lld)", (int)r, base,
              (long long)rows, (long long)cols, (long long)ld);
    return CRAM_ERR_CUDA;
  }
  return CRAM_OK;
}

inline int tiles_of(int n_out, int mode) { return mode == 2 ? (n_out + 63) / 64 : (n_out + kTileN - 1) / kTileN; }

// split-K factor: fill the 2 x #SM CTA slots in as few waves as possible; a CTA costs ~4 K-blocks of start-up /
// epilogue latency and the reducing CTA reads `split` partial tiles.
int g_pdl = 1;          // programmatic dependent launch on (cram_debug_linear_pdl switches it off for A/B timing)
int g_force_split = 0;  // scripts/profile_linear.py only (cram_debug_linear_split)
int choose_split(int n_tiles, int k_blocks) {
  if (g_force_split > 0) return g_force_split < k_blocks ? g_force_split : k_blocks;
  const int slots = 2 * sm_count();
  int best = 1;
  double best_cost = 1e30;
  for (int s = 1; s <= k_blocks && s <= 32; ++s) {
    const int waves = (n_tiles * s + slots - 1) / slots;
    const double cost = waves * ((k_blocks + s - 1) / s + 4.0) + 0.25 * s;
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
}

constexpr size_t kCounterBytes = kMaxTiles * sizeof(int);

template <typename T, int kMode>
int launch_t(const CUtensorMap& tw, const CUtensorMap& tx, const Params& p, int n_tiles, cudaStream_t stream) {
  auto kern = linear_decode_sm100_kernel<T, kMode>;
  static bool attr_set = false;  // per template instantiation
  if (!attr_set) {
    CRAM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemRequest));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(n_tiles * p.split));
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemRequest;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see griddep_wait in the kernel
  attr[0].val.programmaticStreamSerializationAllowed = g_pdl ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CRAM_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, p));
  return check_launch("linear_decode_sm100_kernel");
}

}  // namespace

}  // namespace cram

using namespace cram;

extern "C" void cram_debug_linear_split(int split) { g_force_split = split; }
extern "C" void cram_debug_linear_pdl(int on) { g_pdl = on; }

extern "C" size_t cram_linear_decode_workspace_bytes(int n_out, int k, int mode) {
  if (n_out <= 0 || k <= 0) return 0;
  const int n_tiles = tiles_of(n_out, mode);
  const int split = choose_split(n_tiles, (k + kBlockK - 1) / kBlockK);
  return kCounterBytes + (split > 1 ? (size_t)n_tiles * split * kTileN * kRows * sizeof(float) : 0);
}

extern "C" int cram_linear_decode(const void* x, int64_t ldx, int m, const void* w, int64_t ldw, int n_out, int k,
                                  const void* bias, void* y, int64_t ldy, int mode, int dtype, void* workspace,
                                  size_t workspace_bytes, cram_stream_t stream) {
  CRAM_REQUIRE(dtype == CRAM_BF16 || dtype == CRAM_F16, "cram_linear_decode: dtype must be bf16 or fp16");
  CRAM_REQUIRE(mode >= 0 && mode <= 2, "cram_linear_decode: mode must be 0 (bias), 1 (residual add) or 2 (SwiGLU)");
  CRAM_REQUIRE(m >= 0 && m <= kRows, "cram_linear_decode: at most %d rows (got %d)", kRows, m);
  CRAM_REQUIRE(n_out > 0 && k > 0, "cram_linear_decode: bad shape");
  if (m == 0) return CRAM_OK;
  CRAM_REQUIRE(k % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "cram_linear_decode: k and the x / w row strides must be multiples of 8");
  auto mis = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
  CRAM_REQUIRE(!mis(x) && !mis(w), "cram_linear_decode: x and w must be 16-byte aligned");
  CRAM_REQUIRE(workspace && !mis(workspace), "cram_linear_decode: workspace missing");
  const int n_tiles = tiles_of(n_out, mode);
  CRAM_REQUIRE(n_tiles <= kMaxTiles, "cram_linear_decode: n_out too large");
  Params p{};
  p.y = y;
  p.ldy = ldy;
  p.bias = mode == 0 ? bias : nullptr;
  p.m = m;
  p.n_out = n_out;
  p.k_blocks = (k + kBlockK - 1) / kBlockK;
  p.split = choose_split(n_tiles, p.k_blocks);
  p.up_row0 = n_out;
  const size_t need = kCounterBytes + (p.split > 1 ? (size_t)n_tiles * p.split * kTileN * kRows * sizeof(float) : 0);
  CRAM_REQUIRE(workspace_bytes >= need, "cram_linear_decode: workspace of %zu bytes, need %zu", workspace_bytes, need);
  p.counters = reinterpret_cast<int*>(workspace);  // zero before the first use; every launch leaves them zero
  p.partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kCounterBytes);
  CUtensorMap tw, tx;
  int rc;
  if ((rc = tmap_rows_by_k(&tw, w, dtype, mode == 2 ? 2 * (int64_t)n_out : n_out, k, ldw)) != CRAM_OK) return rc;
  if ((rc = tmap_rows_by_k(&tx, x, dtype, m, k, ldx)) != CRAM_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == CRAM_BF16) {
    if (mode == 0) return launch_t<__nv_bfloat16, 0>(tw, tx, p, n_tiles, s);
    if (mode == 1) return launch_t<__nv_bfloat16, 1>(tw, tx, p, n_tiles, s);
    return launch_t<__nv_bfloat16, 2>(tw, tx, p, n_tiles, s);
  }
  if (mode == 0) return launch_t<__half, 0>(tw, tx, p, n_tiles, s);
  if (mode == 1) return launch_t<__half, 1>(tw, tx, p, n_tiles, s);
  return launch_t<__half, 2>(tw, tx, p, n_tiles, s);
}
