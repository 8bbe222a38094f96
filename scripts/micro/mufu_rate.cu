// Microbenchmark: MUFU.EX2 issue rate per SM sub-partition for f32 / f16x2 / bf16x2 variants, and the
// FFMA2 / FMNMX3 / F2FP rates, to size the softmax pipeline of the attention kernel.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
template <int MODE>
__global__ void k(uint32_t* out, long long* cyc, int iters) {
  uint32_t x[16];
  uint64_t y[8];
  for (int i = 0; i < 8; ++i) y[i] = 0x3f8000003f800000ull + threadIdx.x + i;
  for (int i = 0; i < 16; ++i) x[i] = 0x3c003c00u + threadIdx.x + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(x[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(x[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(x[i]));
      if (MODE == 3) { if (i < 8) asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(y[i])); }
      if (MODE == 8) { if (i < 8) asm volatile("add.rn.f32x2 %0, %0, %0;" : "+l"(y[i])); }
      if (MODE == 9) asm volatile("add.rn.f32 %0, %0, %0;" : "+r"(x[i]));
      if (MODE == 10) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(x[(i + 1) & 15]), "r"(x[(i + 2) & 15]));
      if (MODE == 11) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(x[(i + 1) & 15]), "r"(x[(i + 2) & 15]));
      if (MODE == 4) asm volatile("max.f32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(x[(i + 1) & 15]), "r"(x[(i + 2) & 15]));
      if (MODE == 5) asm volatile("cvt.rn.bf16x2.f32 %0, %0, %1;" : "+r"(x[i]) : "r"(x[(i + 1) & 15]));
      if (MODE == 6) asm volatile("{.reg .b16 a, b; mov.b32 {a, b}, %0; ex2.approx.f16 a, a; mov.b32 %0, {a, b};}" : "+r"(x[i]));
      if (MODE == 7) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+r"(x[i]));
      if (MODE == 12) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(x[i]) : "r"(x[(i + 1) & 15]));
      if (MODE == 13) asm volatile("cvt.rn.f16x2.f32 %0, %0, %1;" : "+r"(x[i]) : "r"(x[(i + 1) & 15]));
      if (MODE == 14) asm volatile("{.reg .b16 a, b; mov.b32 {a, b}, %0; cvt.f32.f16 %0, a;}" : "+r"(x[i]));
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
  for (int i = 0; i < 16; ++i) s ^= x[i];
  for (int i = 0; i < 8; ++i) s ^= (uint32_t)(y[i] >> 13);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE>
void run(const char* name) {
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  for (int warps : {4, 8, 16}) {
    int iters = 2000;
    k<MODE><<<148, warps * 32>>>(out, cyc, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = (double)h[0] / (iters * 16.0);   // cycles per instruction per warp
    printf("%-12s warps/SM=%2d  cycles per warp-instr (per warp) %.2f  -> per SMSP one instr every %.2f cycles\n", name, warps, c, c / (warps / 4.0));
  }
  cudaFree(out); cudaFree(cyc);
}
int main() {
  run<0>("ex2.f32"); run<1>("ex2.f16x2"); run<2>("ex2.bf16x2"); run<12>("hadd2.f16x2"); run<13>("cvt.f16x2.f32"); run<14>("cvt.f32.f16");
  run<3>("ffma2 (x8/16)"); run<8>("fadd2 (x8/16)"); run<7>("ffma"); run<9>("fadd"); run<4>("fmnmx3"); run<5>("cvt.bf16x2"); run<10>("lop3"); run<11>("imad");
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
}
