// Throughput of fp32 -> bf16x2 conversion forms on sm_100a: 16 independent dependency chains per thread, the chain
// runs only through the conversion itself (the packed result is re-read as a float).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cvt_bench cvt_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(uint32_t* out, long long* cyc, float seed, int iters) {
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = seed * (threadIdx.x + i) + 1.0f;
  const float y = seed;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      uint32_t r;
      if (MODE == 0) asm volatile("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y), "f"(x[i]));
      else if (MODE == 1) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y), "f"(x[i]));
      else if (MODE == 2) asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(__float_as_uint(x[i])), "r"(__float_as_uint(y)));
      else if (MODE == 3) asm volatile("fma.rn.f32 %0, %1, %2, %2;" : "=r"(r) : "f"(x[i]), "f"(y));
      else asm volatile("ex2.approx.f32 %0, %1;" : "=r"(r) : "f"(x[i]));
      x[i] = __uint_as_float(r);
    }
  }
  const long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc ^= __float_as_uint(x[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  const int iters = 4000;
  const char* names[5] = {"cvt.rn.relu.bf16x2.f32", "cvt.rn.bf16x2.f32", "prmt.b32", "fma.rn.f32", "ex2.approx.f32 (MUFU)"};
  for (int threads = 256; threads <= 1024; threads *= 4)
    for (int m = 0; m < 5; ++m) {
      for (int rep = 0; rep < 2; ++rep) {
        if (m == 0) k<0><<<148, threads>>>(out, cyc, 0.37f, iters);
        if (m == 1) k<1><<<148, threads>>>(out, cyc, 0.37f, iters);
        if (m == 2) k<2><<<148, threads>>>(out, cyc, 0.37f, iters);
        if (m == 3) k<3><<<148, threads>>>(out, cyc, 0.37f, iters);
        if (m == 4) k<4><<<148, threads>>>(out, cyc, 0.37f, iters);
      }
      cudaDeviceSynchronize();
      long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      const double ops = double(threads) * iters * 16;
      printf("threads %4d  %-26s %9lld cycles  %6.1f thread-ops/clk/SM\n", threads, names[m], h[0], ops / h[0]);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
