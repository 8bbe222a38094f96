// Shared helpers for libcram_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/cram_b200.h"

namespace cram {

void set_error(const char* fmt, ...);

#define CRAM_REQUIRE(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      ::cram::set_error(__VA_ARGS__);    \
      return CRAM_ERR_INVALID;           \
    }                                    \
  } while (0)

#define CRAM_CUDA(expr)                                                                       \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      ::cram::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return CRAM_ERR_CUDA;                                                                   \
    }                                                                                         \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
    return CRAM_ERR_CUDA;
  }
  return CRAM_OK;
}

// B200: 148 SMs. Queried once per DEVICE (a process may drive several GPUs); used to size persistent grids.
int sm_count();

// One-time-per-device flag: cudaFuncSetAttribute is per device, so a `static bool` per template instantiation would
// skip the attribute on the second GPU a process touches and the launch would fail there.
struct PerDeviceOnce {
  bool done[64] = {};
  bool& flag() {
    int d = 0;
    cudaGetDevice(&d);
    return done[d & 63];
  }
};

// Programmatic dependent launch for the decode step's chain of short kernels: a kernel launched this way may become
// resident while its predecessor in the stream is still running; it must execute pdl_wait() before touching
// anything the predecessor wrote, and calls pdl_launch_dependents() so that ITS successor can do the same. Both are
// no-ops for a normally launched kernel. `early` is set for decode-sized launches only (a prefill-sized grid gains
// nothing from it). cram_debug_pdl(0) switches the attribute off (A/B timing).
extern int g_pdl;
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(bool early, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = (g_pdl && early) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

template <typename T>
struct DType;
template <>
struct DType<float> {
  static constexpr int code = CRAM_F32;
  __device__ static float to_f(float x) { return x; }
  __device__ static float from_f(float x) { return x; }
};
template <>
struct DType<__half> {
  static constexpr int code = CRAM_F16;
  __device__ static float to_f(__half x) { return __half2float(x); }
  __device__ static __half from_f(float x) { return __float2half_rn(x); }
};
template <>
struct DType<__nv_bfloat16> {
  static constexpr int code = CRAM_BF16;
  __device__ static float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }
  __device__ static __nv_bfloat16 from_f(float x) { return __float2bfloat16_rn(x); }
};

// round a float through the model dtype (what a torch op on that dtype stores)
template <typename T>
__device__ __forceinline__ float round_to(float x) {
  return DType<T>::to_f(DType<T>::from_f(x));
}

// load a log-credibility value of runtime dtype as float
__device__ __forceinline__ float load_logcred(const void* p, int64_t i, int dtype) {
  if (dtype == CRAM_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
  if (dtype == CRAM_F16) return __half2float(reinterpret_cast<const __half*>(p)[i]);
  return reinterpret_cast<const float*>(p)[i];
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_max_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

inline size_t dtype_size(int dtype) { return dtype == CRAM_F32 ? 4 : 2; }

}  // namespace cram
