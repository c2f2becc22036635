// Micro-benchmark: tcgen05.ld / tcgen05.st throughput per SM on B200 (how fast can the softmax warps read S?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/tmem tools/ubench/tmem.cu && ./tools/ubench/tmem
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void ld32(uint32_t a, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(a)
      : "memory");
}
__device__ __forceinline__ void st32(uint32_t a, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      :
      : "r"(a), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
        "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
        "r"(r[30]), "r"(r[31])
      : "memory");
}

// MODE 0: loads only (2 x32 loads then wait, like the softmax). 1: stores only. 2: load + 64 FFMA of work per 64 values
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(uint32_t* out, int iters, long long* cycles) {
  __shared__ uint32_t tm;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tm)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tm + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;   // each warp its own 64 columns
  uint32_t r[32], q[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) { r[i] = threadIdx.x + i; q[i] = i; }
  st32(base, r);
  st32(base + 32, q);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  __syncthreads();
  const long long t0 = clock64();
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 2) {
      ld32(base, r);
      ld32(base + 32, q);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (MODE == 2) {
#pragma unroll
        for (int i = 0; i < 32; ++i) acc = fmaf(__uint_as_float(r[i]), 1.0001f, acc) + __uint_as_float(q[i]);
      } else {
        acc += __uint_as_float(r[it & 31]) + __uint_as_float(q[it & 31]);
      }
    } else {
      r[it & 31] += it;
      st32(base, r);
      st32(base + 32, r);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
  if (acc == 123.456f) out[0] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

template <int MODE>
void run(const char* name, int warps) {
  uint32_t* d;
  long long* c;
  cudaMalloc(&d, 4);
  cudaMalloc(&c, 8);
  const int iters = 4000;
  k<MODE><<<148, warps * 32>>>(d, 100, c);
  k<MODE><<<148, warps * 32>>>(d, iters, c);
  cudaError_t e = cudaDeviceSynchronize();
  long long cyc = 0;
  cudaMemcpy(&cyc, c, 8, cudaMemcpyDeviceToHost);
  const double bytes = (double)warps * 32 * 64 * 4 * iters;   // per SM
  printf("%-28s warps/SM %2d: %9lld clk  %7.1f B/clk/SM  (%s)\n", name, warps, cyc, bytes / cyc, cudaGetErrorString(e));
  cudaFree(d);
  cudaFree(c);
}

int main() {
  for (int w : {4, 8, 16}) {
    run<0>("tcgen05.ld 2 x (32x32b.x32)", w);
    run<2>("  + 64 FFMA", w);
    run<1>("tcgen05.st 2 x (32x32b.x32)", w);
  }
  return 0;
}
