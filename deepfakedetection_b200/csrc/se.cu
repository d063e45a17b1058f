// Squeeze-excite MLP on the pooled sums produced by the depthwise kernel:
//   gate[n,c] = sigmoid( W2 . silu( W1 . (pool[n,:]/hw) + b1 ) + b2 )
// Replaces timm `se.conv_reduce` -> SiLU -> `se.conv_expand` -> sigmoid;
// the multiply by the gate itself is fused into the A-operand path of the project GEMM.
// Latency-bound micro-kernel: one block per image, everything fp32 (accurate expf).
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int kSeThreads = 256;

__global__ void __launch_bounds__(kSeThreads)
se_kernel(const pool_t* __restrict__ pool, int C, int RD, float inv_hw, const float* __restrict__ w1,
          const float* __restrict__ b1, const float* __restrict__ w2t, const float* __restrict__ b2,
          float* __restrict__ gate) {
  extern __shared__ float sm[];
  float* mean = sm;       // [C]
  float* hid = sm + C;    // [RD]
  const int n = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int c = tid; c < C; c += kSeThreads) mean[c] = pool_value(pool[(size_t)n * C + c]) * inv_hw;
  __syncthreads();
  for (int j = warp; j < RD; j += kSeThreads / 32) {
    const float* wr = w1 + (size_t)j * C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(__ldg(wr + c), mean[c], s);
    s = warp_sum(s);
    if (lane == 0) hid[j] = silu_acc(s + b1[j]);
  }
  __syncthreads();
  for (int c = tid; c < C; c += kSeThreads) {
    float s = b2[c];
    for (int j = 0; j < RD; ++j) s = fmaf(__ldg(w2t + (size_t)j * C + c), hid[j], s);
    gate[(size_t)n * C + c] = sigmoid_acc(s);
  }
}

cudaError_t launch_se(const pool_t* pool, int n, int c, int rd, float inv_hw, const float* w1, const float* b1,
                      const float* w2t, const float* b2, float* gate, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  se_kernel<<<n, kSeThreads, (c + rd) * sizeof(float), s>>>(pool, c, rd, inv_hw, w1, b1, w2t, b2, gate);
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
