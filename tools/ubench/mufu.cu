// Micro-benchmark: sustained ex2 throughput per SM on B200 under the instruction mixes of the attention softmax.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/mufu tools/ubench/mufu.cu && ./tools/ubench/mufu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack(float a, float b) { __nv_bfloat162 v = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&v); }
__device__ __forceinline__ float fmax3(float a, float b, float c) { float d; asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

// MODE 0: MUFU only. 1: FFMA + MUFU. 2: FFMA + MUFU + F2FP (pack pairs). 3: + FMNMX3 per pair. 4: as 2 but PRMT truncation pack.
// 5: as 3 + one 16-byte st.shared per 8 values
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out, int iters, float sc, float off) {
  __shared__ uint4 sm[512];
  float x[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) x[i] = threadIdx.x * 1e-3f + i * 0.01f;
  float acc = 0.f, mx = -1e30f;
  uint32_t pacc = 0;
  for (int it = 0; it < iters; ++it) {
    float y[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float t = x[i];
      if (MODE >= 1) t = fmaf(t, sc, off);
      y[i] = ex2(t);
    }
    if (MODE == 3 || MODE == 5) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) mx = fmax3(mx, x[i], x[i + 1]);
    }
    if (MODE == 2 || MODE == 3 || MODE == 5) {
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 32; i += 2) pk[i >> 1] = pack(y[i], y[i + 1]);
      if (MODE == 5) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sm[(threadIdx.x + q * 7) & 511] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) pacc ^= pk[i];
      }
    } else if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) pacc ^= __byte_perm(__float_as_uint(y[i]), __float_as_uint(y[i + 1]), 0x7632);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) acc += y[i];
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] += 1e-6f * (MODE == 0 ? 0.f : 1.f) + (MODE == 0 ? y[i] * 1e-9f : 0.f);
  }
  if (acc + mx + pacc == 12345.678f) out[0] = acc + sm[0].x;
}

template <int MODE>
void run(const char* name, int warps) {
  float* d;
  cudaMalloc(&d, 4);
  const int iters = 2000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, warps * 32>>>(d, 10, 1.0001f, -0.5f);
  cudaEventRecord(e0);
  k<MODE><<<148, warps * 32>>>(d, iters, 1.0001f, -0.5f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  int clk;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double ex = 148.0 * warps * 32 * 32.0 * iters;
  printf("%-44s warps/SM %2d: %7.3f ms  %6.2f ex2/ns/SM  (= %5.2f per clk at 1.9 GHz)\n", name, warps, ms, ex / 148 / (ms * 1e6),
         ex / 148 / (ms * 1e6) / 1.9);
  cudaFree(d);
}

int main() {
  for (int w : {4, 8, 16}) {
    if (w == 4) { run<0>("MUFU only", 4); run<1>("FFMA+MUFU", 4); run<2>("FFMA+MUFU+F2FP", 4); run<3>("FFMA+MUFU+F2FP+FMNMX3", 4); run<4>("FFMA+MUFU+PRMT", 4); run<5>("FFMA+MUFU+F2FP+FMNMX3+STS", 4); }
    if (w == 8) { run<0>("MUFU only", 8); run<1>("FFMA+MUFU", 8); run<2>("FFMA+MUFU+F2FP", 8); run<3>("FFMA+MUFU+F2FP+FMNMX3", 8); run<4>("FFMA+MUFU+PRMT", 8); run<5>("FFMA+MUFU+F2FP+FMNMX3+STS", 8); }
    if (w == 16) { run<0>("MUFU only", 16); run<1>("FFMA+MUFU", 16); run<2>("FFMA+MUFU+F2FP", 16); run<3>("FFMA+MUFU+F2FP+FMNMX3", 16); run<4>("FFMA+MUFU+PRMT", 16); run<5>("FFMA+MUFU+F2FP+FMNMX3+STS", 16); }
  }
  return 0;
}
