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
//   * 2 CTAs per SM (2 x 96 KB of ring = 128 KB of weights in flight per SM);
//   * split-K across a thread-block CLUSTER (2, 4 or 8 CTAs share a weight tile, each streams a K range): the
//     partial accumulators are reduce-scattered through distributed shared memory -- CTA j receives everybody's
//     token rows [j*64/split, (j+1)*64/split), sums them in rank order (deterministic) and runs the epilogue for
//     those rows. No global workspace, no atomics, no second kernel;
//   * programmatic dependent launch: the weights do not depend on the previous kernel, so the first ring of W
//     tiles is requested before griddepcontrol.wait; only x (and y) wait for the dependency;
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
constexpr int kRingShared = 4;  // two CTAs per SM: grids that fill the machine
constexpr int kRingDeep = 8;    // one CTA per SM: small projections (<= #SM CTAs), twice the weights in flight per CTA
constexpr int kWBytes = kTileN * kBlockK * 2;  // 16 KB
constexpr int kXBytes = kRows * kBlockK * 2;   // 8 KB
constexpr int kStageBytes = kWBytes + kXBytes;
constexpr int kThreads = 192;  // warps 0-3 epilogue, warp 4 producer, warp 5 MMA issuer / TMEM owner
constexpr int kTmemCols = 64;
constexpr uint32_t kDescHi = ptx::umma_desc_hi(1024);

template <int kRing>
struct Bars {
  uint64_t full[kRing], empty[kRing];
  uint64_t acc_full;
  uint32_t tmem_ptr;
};
template <int kRing>
constexpr int smem_request() {
  return kRing * kStageBytes + (int)sizeof(Bars<kRing>) + 1024;
}
static_assert(kTileN * kRows * 4 <= kRingShared * kStageBytes, "the reduce-scatter buffer reuses the drained ring");

// Debug timeline (scripts/profile_linear.py trace builds a separate library with -DCRAM_LINEAR_TRACE; the product
// build compiles none of this): %globaltimer stamps of every CTA of the last 32 launches.
#ifdef CRAM_LINEAR_TRACE
__device__ unsigned long long g_lin_trace[32][320][8];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define LIN_TRACE(e)                                                                         \
  do {                                                                                       \
    if (blockIdx.x < 320) g_lin_trace[p.trace_slot & 31][blockIdx.x][e] = gtime();           \
  } while (0)
#else
#define LIN_TRACE(e) \
  do {               \
  } while (0)
#endif

struct Params {
  int trace_slot;
  void* y;
  int64_t ldy;
  const void* bias;
  int m, n_out, k_blocks, split;
  int up_row0;  // mode 2: first up row (= n_out)
};

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
// all threads of all CTAs of the cluster
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `local_addr` (a shared::cta address of this CTA) inside CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <typename T>
__device__ __forceinline__ float rnd(float v) {
  return DType<T>::to_f(DType<T>::from_f(v));
}

// What the epilogue of weight row n needs from global memory besides the accumulator -- the bias (mode 0) or the
// residual values it adds to (mode 1) -- requested early, so that the latency hides behind the main loop.
template <typename T, int kMode, int kCount>
__device__ __forceinline__ void load_pre(const Params& p, int n, int m0, float (&pre)[kCount]) {
  if (n >= p.n_out) return;
  if (kMode == 0) {
    pre[0] = p.bias ? DType<T>::to_f(reinterpret_cast<const T*>(p.bias)[n]) : 0.f;
  } else if (kMode == 1) {
    const T* y = reinterpret_cast<const T*>(p.y);
#pragma unroll
    for (int i = 0; i < kCount; ++i) pre[i] = m0 + i < p.m ? DType<T>::to_f(y[(int64_t)(m0 + i) * p.ldy + n]) : 0.f;
  }
}

// Epilogue of one weight row n for kCount token rows starting at m0 (v[i]: fp32 sums; up[i]: mode 2 only).
template <typename T, int kMode, int kCount>
__device__ __forceinline__ void store_rows(const Params& p, int n, int m0, const float (&v)[kCount], const float (&up)[kCount],
                                           const float (&pre)[kCount]) {
  if (n >= p.n_out) return;
  T* y = reinterpret_cast<T*>(p.y);
#pragma unroll
  for (int i = 0; i < kCount; ++i) {
    if (m0 + i < p.m) {
      float r;
      if (kMode == 0) {
        r = v[i] + pre[0];
      } else if (kMode == 1) {
        r = pre[i] + rnd<T>(v[i]);
      } else {
        const float g = rnd<T>(v[i]);
        r = rnd<T>(g / (1.0f + expf(-g))) * rnd<T>(up[i]);
      }
      y[(int64_t)(m0 + i) * p.ldy + n] = DType<T>::from_f(r);
    }
  }
}

// Split-K over a cluster of kSplit CTAs: reduce-scatter of the [128 x 64] fp32 partial tiles through distributed
// shared memory. `recv` is this CTA's drained ring, as float4 [source rank][kMS / 4][weight row] (lane-contiguous).
template <typename T, int kMode, int kSplit>
__device__ __forceinline__ void cluster_reduce_and_store(const Params& p, const float (&acc)[kRows], uint8_t* recv, int tile,
                                                         int nl, bool epilogue_thread) {
  constexpr int kMS = kRows / kSplit;  // token rows per CTA after the scatter
  constexpr int kQ = kMS / 4;
  const uint32_t rank = cluster_ctarank();
  const int n = kMode == 2 ? tile * 64 + nl : tile * kTileN + nl;
  const bool owner = epilogue_thread && (kMode != 2 || nl < 64);
  float pre[kMS];
  if (owner) load_pre<T, kMode, kMS>(p, n, (int)rank * kMS, pre);
  cluster_sync();  // every CTA of the cluster has drained its ring: the rings may now be overwritten remotely
  if (epilogue_thread) {
    const uint32_t mine = ptx::smem_u32(recv) + ((rank * kQ) * kTileN + nl) * 16;
#pragma unroll
    for (int j = 0; j < kSplit; ++j) {
      const uint32_t dst = map_to_cta(mine, (uint32_t)j);
#pragma unroll
      for (int q = 0; q < kQ; ++q)
        st_cluster_v4(dst + q * kTileN * 16, acc[j * kMS + 4 * q], acc[j * kMS + 4 * q + 1], acc[j * kMS + 4 * q + 2],
                      acc[j * kMS + 4 * q + 3]);
    }
  }
  cluster_sync();  // all partial rows of this CTA's token slice have landed
#ifdef CRAM_LINEAR_TRACE
  if (threadIdx.x == 0) LIN_TRACE(6);
#endif
  if (!owner) return;
  const float4* r = reinterpret_cast<const float4*>(recv);
  auto sum_row = [&](int row, float (&out)[kMS]) {
#pragma unroll
    for (int e = 0; e < kMS; ++e) out[e] = 0.f;
#pragma unroll
    for (int s = 0; s < kSplit; ++s) {  // rank order: the sum does not depend on timing
#pragma unroll
      for (int q = 0; q < kQ; ++q) {
        const float4 v = r[(s * kQ + q) * kTileN + row];
        out[4 * q] += v.x;
        out[4 * q + 1] += v.y;
        out[4 * q + 2] += v.z;
        out[4 * q + 3] += v.w;
      }
    }
  };
  float v[kMS], up[kMS];
  sum_row(nl, v);
  if (kMode == 2) sum_row(nl + 64, up);
  store_rows<T, kMode, kMS>(p, n, (int)rank * kMS, v, kMode == 2 ? up : v, pre);
}

template <typename T, int kMode, int kRing>
__global__ void __launch_bounds__(kThreads, kRing == kRingShared ? 2 : 1)
linear_decode_sm100_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                           const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Bars<kRing>* bar = reinterpret_cast<Bars<kRing>*>(smem + kRing * kStageBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // grid = tiles x split, the `split` CTAs of a tile are consecutive (= one cluster)
  const int tile = blockIdx.x / p.split, sp = blockIdx.x - tile * p.split;
  // K blocks [kb0, kb0 + n_kb) of this split: the first `rem` splits take one more
  const int per = p.k_blocks / p.split, rem = p.k_blocks - per * p.split;
  const int kb0 = sp * per + min(sp, rem), n_kb = per + (sp < rem ? 1 : 0);
  // weight row of TMEM lane 0 / of lane 64 (mode 2: gate rows, then their up rows)
  const int row_lo = kMode == 2 ? tile * 64 : tile * kTileN;
  const int row_hi = kMode == 2 ? p.up_row0 + tile * 64 : tile * kTileN + 64;

  if (warp == 4 && lane == 0) {
    LIN_TRACE(0);
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
  if (threadIdx.x == 0) LIN_TRACE(1);
  ptx::griddep_wait();  // x and y: everything below may depend on the previous kernel
  if (threadIdx.x == 0) LIN_TRACE(2);

  float acc[kRows];
  float pre[kRows];  // split == 1: bias / residual values requested before the main loop ends
  const int nl = (warp & 3) * 32 + lane;  // epilogue warps: TMEM lane = weight row inside the tile
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
    __syncwarp();
  } else if (warp == 5) {
    if (ptx::elect_one_sync()) {
      constexpr uint32_t kAbFmt = DType<T>::code == CRAM_BF16 ? 1u : 0u;
      constexpr uint32_t kIdesc = ptx::umma_idesc_f16(kAbFmt, kTileN, kRows, 0);
      const uint32_t w_lo = ptx::umma_desc_lo(ptx::smem_u32(smem), 16);
      const uint32_t x_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + kWBytes), 16);
      for (int i = 0; i < n_kb; ++i) {
        const int st = i % kRing;
        ptx::mbar_wait(&bar->full[st], (i / kRing) & 1);
        if (i == 0) LIN_TRACE(3);
        ptx::tc_fence_after();
        const uint32_t a = w_lo + st * (kStageBytes >> 4), b = x_lo + st * (kStageBytes >> 4);
        ptx::mma_ss_imm<kDescHi, 0, 0, 0>(tmem, a, b, kIdesc, i > 0 ? 1u : 0u);
        ptx::mma_ss_imm<kDescHi, 0, 2, 2>(tmem, a, b, kIdesc, 1u);
        ptx::mma_ss_imm<kDescHi, 0, 4, 4>(tmem, a, b, kIdesc, 1u);
        ptx::mma_ss_imm<kDescHi, 0, 6, 6>(tmem, a, b, kIdesc, 1u);
        ptx::tc_commit(&bar->empty[st]);
      }
      ptx::tc_commit(&bar->acc_full);
      LIN_TRACE(4);
    }
    __syncwarp();
  } else {
    // ============================== accumulator -> registers (warps 0-3, TMEM lanes 32w..32w+31) =======
    if (kMode != 2 && p.split == 1) load_pre<T, kMode, kRows>(p, tile * kTileN + nl, 0, pre);
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
      if (threadIdx.x == 0) LIN_TRACE(5);
    } else {
#pragma unroll
      for (int i = 0; i < kRows; ++i) acc[i] = 0.f;
    }
  }

  if (p.split > 1) {
    // every thread of the cluster takes part in the two cluster barriers inside
    if (p.split == 2) cluster_reduce_and_store<T, kMode, 2>(p, acc, smem, tile, nl, warp < 4);
    else if (p.split == 4) cluster_reduce_and_store<T, kMode, 4>(p, acc, smem, tile, nl, warp < 4);
    else cluster_reduce_and_store<T, kMode, 8>(p, acc, smem, tile, nl, warp < 4);
  } else if (warp < 4) {
    if (kMode == 2) {
      // lanes 0-63 hold gate rows, lanes 64-127 their up rows. Gate threads finish token rows 0-31 of their weight
      // row, up threads rows 32-63: each side hands the other the half it needs through the drained ring.
      float* s_up = reinterpret_cast<float*>(smem);            // [32 token rows][64]: up values, rows 0-31
      float* s_gate = reinterpret_cast<float*>(smem) + 32 * 64;  // [32 token rows][64]: gate values, rows 32-63
      const int row = nl & 63;
      if (nl >= 64) {
#pragma unroll
        for (int m = 0; m < 32; ++m) s_up[m * 64 + row] = acc[m];
      } else {
#pragma unroll
        for (int m = 0; m < 32; ++m) s_gate[m * 64 + row] = acc[32 + m];
      }
      epi_bar();
      float g[32], u[32];
      if (nl >= 64) {
#pragma unroll
        for (int m = 0; m < 32; ++m) {
          g[m] = s_gate[m * 64 + row];
          u[m] = acc[32 + m];
        }
      } else {
#pragma unroll
        for (int m = 0; m < 32; ++m) {
          g[m] = acc[m];
          u[m] = s_up[m * 64 + row];
        }
      }
      store_rows<T, kMode, 32>(p, tile * 64 + row, nl >= 64 ? 32 : 0, g, u, g);
    } else {
      store_rows<T, kMode, kRows>(p, tile * kTileN + nl, 0, acc, acc, pre);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) LIN_TRACE(7);
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

int g_force_split = 0;  // scripts/profile_linear.py only (cram_debug_linear_split)

int g_force_ring = 0;   // 0 = by grid size, 4 / 8 = forced (scripts/profile_linear.py only)

void fill_config(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int grid, int split, int smem, cudaStream_t stream) {
  cfg->gridDim = dim3((unsigned)grid);
  cfg->blockDim = dim3(kThreads);
  cfg->dynamicSmemBytes = smem;
  cfg->stream = stream;
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see griddep_wait in the kernel
  attr[0].val.programmaticStreamSerializationAllowed = g_pdl ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = (unsigned)split;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg->attrs = attr;
  cfg->numAttrs = 2;
}

// How many clusters of `split` CTAs are resident at once with 2 CTAs per SM. Clusters of 8 have to fit inside one
// GPC: measured on B200, a 224-CTA grid of 8-clusters runs in one wave while 288 CTAs do not (scripts/
// profile_linear.py splits), hence ~30 of them. (cudaOccupancyMaxActiveClusters answers 148 / 74 / 33 / 15 here,
// i.e. for one CTA per SM, which is not what the launches do.)
int resident_clusters(int split) {
  const int sms = sm_count();
  return split == 8 ? 30 * sms / 148 : 2 * sms / split;
}

// A grid of at most #SM CTAs runs one CTA per SM and takes the 8-stage ring (128 KB of weights in flight per CTA:
// under programmatic dependent launch most of a small projection is on chip before its input exists).
bool deep_ring(int n_tiles, int split) {
  if (g_force_ring) return g_force_ring == kRingDeep;
  return n_tiles * split <= sm_count();
}

// split-K factor (cluster size 1, 2, 4 or 8). Streaming model: a CTA keeps its ring (4 or 8 K-blocks of 16 KB of
// weights) in flight, i.e. ~2.5 or ~5 blocks/us at loaded HBM latency; the device moves ~390 blocks/us (6.4 TB/s);
// the reduce-scatter costs a few hundred ns per doubling. Pick the smallest split that reaches the best modelled time.
int choose_split(int n_tiles, int k_blocks) {
  if (g_force_split > 0) return g_force_split;
  int best = 1;
  double best_t = 1e30;
  const double device_rate = 390.0 * sm_count() / 148.0;
  for (int s = 1; s <= 8 && s <= k_blocks; s *= 2) {
    const int cap = resident_clusters(s);
    const double conc = (double)(n_tiles < cap ? n_tiles : cap) * s;
    const double per_cta = deep_ring(n_tiles, s) ? 5.0 : 2.5;
    const double rate = conc * per_cta < device_rate ? conc * per_cta : device_rate;
    const double t = (double)n_tiles * k_blocks / rate + (s > 1 ? 1.0 + 0.4 * s : 0.0);
    if (t < best_t * 0.97) {
      best_t = t;
      best = s;
    }
  }
  return best;
}

template <typename T, int kMode, int kRing>
int launch_r(const CUtensorMap& tw, const CUtensorMap& tx, const Params& p, int n_tiles, cudaStream_t stream) {
  auto kern = linear_decode_sm100_kernel<T, kMode, kRing>;
  static PerDeviceOnce smem_attr;  // per template instantiation, per device
  if (!smem_attr.flag()) {
    CRAM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_request<kRing>()));
    smem_attr.flag() = true;
  }
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[2];
  fill_config(&cfg, attr, n_tiles * p.split, p.split, smem_request<kRing>(), stream);
  CRAM_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, p));
  return check_launch("linear_decode_sm100_kernel");
}

template <typename T, int kMode>
int launch_t(const CUtensorMap& tw, const CUtensorMap& tx, Params p, int n_tiles, cudaStream_t stream) {
  p.split = choose_split(n_tiles, p.k_blocks);
  static int launches = 0;
  p.trace_slot = launches++;
  return deep_ring(n_tiles, p.split) ? launch_r<T, kMode, kRingDeep>(tw, tx, p, n_tiles, stream)
                                     : launch_r<T, kMode, kRingShared>(tw, tx, p, n_tiles, stream);
}

}  // namespace

}  // namespace cram

using namespace cram;

#ifdef CRAM_LINEAR_TRACE
extern "C" int cram_debug_linear_trace(unsigned long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, cram::g_lin_trace, sizeof(cram::g_lin_trace)) == cudaSuccess ? 0 : -1;
}
#endif
extern "C" void cram_debug_linear_split(int split) { g_force_split = split; }
extern "C" void cram_debug_linear_ring(int ring) { g_force_ring = ring; }

extern "C" int cram_linear_decode(const void* x, int64_t ldx, int m, const void* w, int64_t ldw, int n_out, int k,
                                  const void* bias, void* y, int64_t ldy, int mode, int dtype, cram_stream_t stream) {
  CRAM_REQUIRE(dtype == CRAM_BF16 || dtype == CRAM_F16, "cram_linear_decode: dtype must be bf16 or fp16");
  CRAM_REQUIRE(mode >= 0 && mode <= 2, "cram_linear_decode: mode must be 0 (bias), 1 (residual add) or 2 (SwiGLU)");
  CRAM_REQUIRE(m >= 0 && m <= kRows, "cram_linear_decode: at most %d rows (got %d)", kRows, m);
  CRAM_REQUIRE(n_out > 0 && k > 0, "cram_linear_decode: bad shape");
  if (m == 0) return CRAM_OK;
  CRAM_REQUIRE(k % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "cram_linear_decode: k and the x / w row strides must be multiples of 8");
  auto mis = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
  CRAM_REQUIRE(!mis(x) && !mis(w), "cram_linear_decode: x and w must be 16-byte aligned");
  const int n_tiles = tiles_of(n_out, mode);
  Params p{};
  p.y = y;
  p.ldy = ldy;
  p.bias = mode == 0 ? bias : nullptr;
  p.m = m;
  p.n_out = n_out;
  p.k_blocks = (k + kBlockK - 1) / kBlockK;
  p.split = 1;
  p.up_row0 = n_out;
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
