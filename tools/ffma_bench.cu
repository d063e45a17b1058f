// FP32 FMA issue-rate microbenchmark for sm_100a (dev tool): scalar FFMA vs packed FFMA2 (fma.rn.f32x2), with
// and without interleaved integer ALU work, at 4 and 8 warps per scheduler.  Prints FMA lanes per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ffma_bench.bin tools/ffma_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
  float a[16];
  u64 p[8];
  unsigned z[4] = {threadIdx.x, threadIdx.x * 3u, 7u, 11u};
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = seed + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) p[i] = ((u64)__float_as_uint(seed + i) << 32) | __float_as_uint(seed - i);
  const u64 m = ((u64)__float_as_uint(0.999f) << 32) | __float_as_uint(1.001f);
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {  // 16 scalar FFMA
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], 0.999f, seed);
    } else if (MODE == 1) {  // 8 FFMA2 = 16 FMA lanes-ops per thread
#pragma unroll
      for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], m, m);
    } else if (MODE == 2) {  // 8 FFMA2 + 8 integer LOP3/SHF
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        p[i] = fma2(p[i], m, m);
        z[i & 3] = (z[i & 3] << 16) ^ (z[(i + 1) & 3] & 0xffff0000u);
      }
    } else {  // 16 FFMA + 8 integer
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        a[i] = fmaf(a[i], 0.999f, seed);
        if (i & 1) z[(i >> 1) & 3] = (z[(i >> 1) & 3] << 16) ^ (z[((i >> 1) + 1) & 3] & 0xffff0000u);
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s += __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + z[0] + z[1] + z[2] + z[3];
}
template <int MODE>
void run(const char* name, int ctas_per_sm, float* out) {
  int dev, sms, khz;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const int iters = 20000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<MODE><<<sms * ctas_per_sm, 256>>>(out, 100, 1.0f);
  cudaEventRecord(e0);
  k<MODE><<<sms * ctas_per_sm, 256>>>(out, iters, 1.0f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double fma = (double)sms * ctas_per_sm * 256 * 16.0 * iters;
  printf("%-28s %2d warps/sched: %.3f ms, %.1f FMA/clk/SM (at %d MHz nominal)\n", name, ctas_per_sm * 2, ms,
         fma / (ms * 1e-3) / sms / (khz * 1e3), khz / 1000);
}
int main() {
  float* out;
  cudaMalloc(&out, 148 * 8 * 256 * 4);
  for (int c = 2; c <= 4; c += 2) {
    run<0>("FFMA x16", c, out);
    run<1>("FFMA2 x8", c, out);
    run<2>("FFMA2 x8 + 8 int", c, out);
    run<3>("FFMA x16 + 8 int", c, out);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
