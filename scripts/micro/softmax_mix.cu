// Microbenchmark: the per-element instruction mix of the attention kernel's softmax (FFMA2 score, FADD2 minus reference,
// MUFU.EX2, FADD2 row sum, F2FP pack) with 1 / 2 / 4 warps per SM sub-partition and no TMEM / barrier traffic at all:
// how many cycles per element does the mix cost when the pipes are free, and do MUFU and the FMA pipe overlap?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o softmax_mix softmax_mix.cu && ./softmax_mix
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ uint64_t pk2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(uint64_t r, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r)); }
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) { uint64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack(float lo, float hi) { uint32_t r; asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }

// MODE 0: full mix; 1: no MUFU (identity); 2: MUFU + pack only (no FFMA2 / FADD2); 3: full mix + row max (FMNMX3)
template <int MODE>
__global__ void k(const float* __restrict__ in, uint32_t* out, long long* cyc, int iters) {
  __shared__ float bias[128];
  if (threadIdx.x < 128) bias[threadIdx.x] = in[threadIdx.x] * 0.01f;
  float s[32];
  for (int i = 0; i < 32; ++i) s[i] = in[(threadIdx.x * 32 + i) & 1023];
  __syncthreads();
  const uint64_t sc2 = pk2(0.127f, 0.127f);
  uint64_t negm = pk2(-1.5f, -1.5f);
  uint64_t la = pk2(0.f, 0.f), lb = pk2(0.f, 0.f);
  uint32_t acc = 0;
  float mx0 = -1e30f, mx1 = -1e30f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    uint32_t pk[16];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 b = *reinterpret_cast<const float4*>(&bias[(i + 32 * (it & 3))]);
      uint64_t t01, t23;
      if (MODE == 2) {
        t01 = pk2(s[i], s[i + 1]);
        t23 = pk2(s[i + 2], s[i + 3]);
      } else {
        t01 = ffma2(pk2(s[i], s[i + 1]), sc2, pk2(b.x, b.y));
        t23 = ffma2(pk2(s[i + 2], s[i + 3]), sc2, pk2(b.z, b.w));
      }
      float e0, e1, e2, e3;
      if (MODE == 3) {
        float a0, a1, a2, a3;
        upk2(t01, a0, a1);
        upk2(t23, a2, a3);
        asm("max.f32 %0, %0, %1, %2;" : "+f"(mx0) : "f"(a0), "f"(a1));
        asm("max.f32 %0, %0, %1, %2;" : "+f"(mx1) : "f"(a2), "f"(a3));
      }
      if (MODE == 2) {
        upk2(t01, e0, e1);
        upk2(t23, e2, e3);
      } else {
        upk2(fadd2(t01, negm), e0, e1);
        upk2(fadd2(t23, negm), e2, e3);
      }
      float p0, p1, p2, p3;
      if (MODE == 1) { p0 = e0; p1 = e1; p2 = e2; p3 = e3; }
      else { p0 = ex2(e0); p1 = ex2(e1); p2 = ex2(e2); p3 = ex2(e3); }
      if (MODE != 2) {
        la = fadd2(la, pk2(p0, p1));
        lb = fadd2(lb, pk2(p2, p3));
      }
      pk[i / 2] = pack(p0, p1);
      pk[i / 2 + 1] = pack(p2, p3);
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) acc ^= pk[i];
    s[it & 31] += 1e-3f;  // keep the loop body from being hoisted
  }
  long long t1 = clock64();
  float l0, l1, l2, l3;
  upk2(la, l0, l1);
  upk2(lb, l2, l3);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc ^ __float_as_uint(l0 + l1 + l2 + l3 + mx0 + mx1);
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE>
void run(const char* name, const float* in) {
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  for (int warps : {4, 8, 16}) {
    int iters = 4000;
    k<MODE><<<148, warps * 32>>>(in, out, cyc, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double per_elem_warp = (double)h[0] / (iters * 32.0);
    printf("%-28s warps/SMSP=%d  cycles per element: per warp %.2f, per sub-partition %.2f\n", name, warps / 4, per_elem_warp,
           per_elem_warp / (warps / 4.0));
  }
  cudaFree(out); cudaFree(cyc);
}
int main() {
  float* in; cudaMalloc(&in, 4096); 
  float h[1024]; for (int i = 0; i < 1024; ++i) h[i] = (i % 37) * 0.1f - 1.7f;
  cudaMemcpy(in, h, 4096, cudaMemcpyHostToDevice);
  run<0>("full mix", in); run<3>("full mix + FMNMX3", in); run<1>("mix without MUFU", in); run<2>("MUFU + F2FP + LDS only", in);
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
}
