// Squeeze-excite MLP on the pooled sums produced by the depthwise kernel:
//   gate[n,c] = sigmoid( W2 . silu( W1 . (pool[n,:]/hw) + b1 ) + b2 )
// Replaces timm `se.conv_reduce` -> SiLU -> `se.conv_expand` -> sigmoid;
// the multiply by the gate itself is fused into the A-operand path of the project GEMM.
// Latency-bound micro-kernel: one block per image, everything fp32 (accurate expf).
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int kSeThreads = 1024;
constexpr int kSeG = 2;  // images per block (128 blocks of 1024 threads at n = 256: one wave)

// Latency-bound micro-kernel: the cost is chains of L2-latency loads, so the work is spread over 32 warps
// per block and every loop keeps many independent loads in flight (unrolled).  One block computes the gates
// of kSeG images; blocks start on different W1 rows so they do not all hit the same L2 lines at once.
// Everything fp32, accurate expf.
__global__ void __launch_bounds__(kSeThreads, 1)
se_kernel(const pool_t* __restrict__ pool, int n, int C, int RD, float inv_hw, const float* __restrict__ w1,
          const float* __restrict__ b1, const float* __restrict__ w2t, const float* __restrict__ b2,
          float* __restrict__ gate) {
  extern __shared__ __align__(16) float sm[];
  float* mean = sm;                 // [kSeG][C]
  float* hid = sm + kSeG * C;       // [kSeG][RD]
  const int n0 = blockIdx.x * kSeG;
  const int ng = min(kSeG, n - n0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  griddep_wait();
  griddep_launch_dependents();
  for (int i = tid; i < kSeG * C; i += kSeThreads) {
    const int g = i / C, c = i - g * C;
    mean[i] = g < ng ? pool_value(pool[(size_t)(n0 + g) * C + c]) * inv_hw : 0.f;
  }
  __syncthreads();
  const int C4 = C >> 2;
  const int j0 = (blockIdx.x * 5) % RD;  // stagger
  for (int jj = warp; jj < RD; jj += kSeThreads / 32) {
    int j = jj + j0;
    if (j >= RD) j -= RD;
    const float4* wr = reinterpret_cast<const float4*>(w1 + (size_t)j * C);
    float s[kSeG];
#pragma unroll
    for (int g = 0; g < kSeG; ++g) s[g] = 0.f;
#pragma unroll 4
    for (int c4 = lane; c4 < C4; c4 += 32) {
      const float4 w = __ldg(wr + c4);
#pragma unroll
      for (int g = 0; g < kSeG; ++g) {
        const float4 m = *reinterpret_cast<const float4*>(mean + g * C + 4 * c4);
        s[g] = fmaf(w.x, m.x, fmaf(w.y, m.y, fmaf(w.z, m.z, fmaf(w.w, m.w, s[g]))));
      }
    }
#pragma unroll
    for (int g = 0; g < kSeG; ++g) {
      const float t = warp_sum(s[g]);
      if (lane == 0) hid[g * RD + j] = silu_acc(t + b1[j]);
    }
  }
  __syncthreads();
  for (int c = tid; c < C; c += kSeThreads) {
    float acc[kSeG];
    const float bb = b2[c];
#pragma unroll
    for (int g = 0; g < kSeG; ++g) acc[g] = bb;
#pragma unroll 16
    for (int j = 0; j < RD; ++j) {
      const float w = __ldg(w2t + (size_t)j * C + c);
#pragma unroll
      for (int g = 0; g < kSeG; ++g) acc[g] = fmaf(w, hid[g * RD + j], acc[g]);
    }
#pragma unroll
    for (int g = 0; g < kSeG; ++g)
      if (g < ng) gate[(size_t)(n0 + g) * C + c] = sigmoid_acc(acc[g]);
  }
}

cudaError_t launch_se(const pool_t* pool, int n, int c, int rd, float inv_hw, const float* w1, const float* b1,
                      const float* w2t, const float* b2, float* gate, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c % 4) return cudaErrorInvalidValue;
  const size_t smem = (size_t)kSeG * (c + rd) * sizeof(float);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  count_launch();
  return launch_pdl(se_kernel, dim3((n + kSeG - 1) / kSeG), dim3(kSeThreads), smem, s, pool, n, c, rd, inv_hw, w1, b1,
                    w2t, b2, gate);
}

}  // namespace b200df
