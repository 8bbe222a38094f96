// Microbenchmark: tcgen05.ld (TMEM -> registers) throughput per SM as a function of how many warps load at once,
// and tcgen05.st. One CTA per SM, 512 columns allocated; every warp loops over 32x32b.x32 loads of its own lane quadrant.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};" ::"r"(r[0]),
      "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
      "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]),
      "r"(r[29]), "r"(r[30]), "r"(r[31]), "r"(taddr) : "memory");
}
// mode 0: back-to-back loads (4 in flight before each wait); mode 1: one load, wait (latency); mode 2: stores
__global__ void k(int mode, int iters, long long* cyc, uint32_t* out) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tptr + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;  // each warp its own columns
  uint32_t r[32], acc = 0;
  for (int i = 0; i < 32; ++i) r[i] = threadIdx.x + i;
  st32(base, r); st32(base + 32, r);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (mode == 0) {
      ld32(base, r); acc += r[3];
      ld32(base + 32, r); acc += r[5];
      ld32(base, r); acc += r[7];
      ld32(base + 32, r); acc += r[9];
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    } else if (mode == 1) {
      ld32(base, r);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      acc += r[3];
    } else {
      st32(base, r); st32(base + 32, r); st32(base, r); st32(base + 32, r);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tptr) : "memory");
}
int main() {
  long long* cyc; uint32_t* out;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&out, 148 * 1024 * 4);
  const char* names[] = {"ld x32, 4 per wait", "ld x32 latency (1 per wait)", "st x32, 4 per wait"};
  for (int mode = 0; mode < 3; ++mode)
    for (int warps : {1, 4, 8, 16}) {
      const int iters = 2000;
      k<<<148, warps * 32>>>(mode, iters, cyc, out);
      cudaError_t e = cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      const double per_iter = (double)h / iters;
      const int n_per_iter = mode == 1 ? 1 : 4;
      printf("%-28s warps=%2d  cycles/iter %.1f  -> %.1f cycles per x32 op per warp, SM-wide %.1f B/clk  (%s)\n", names[mode], warps,
             per_iter, per_iter / n_per_iter, warps * n_per_iter * 4096.0 / per_iter, cudaGetErrorString(e));
    }
  return 0;
}
