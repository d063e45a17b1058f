// Pointwise (1x1) conv as a CUDA-core FFMA GEMM -- the *strict fp32* path
// (B200DF_PREC_FP32: every multiply-add in fp32, max |dlogit| <= 1e-3 vs the oracle) and the
// cross-check for the tcgen05 kernel in pwconv_tc.cu.  Not the throughput path.
//   out[m, co] = act( sum_k x[m,k] * gate[m / rows_per_img, k] * w[co,k] + bias[co] ) (+ residual[m,co])
// Replaces timm `conv_pw`/`conv_pwl`/`conv_head` (+ folded BN) -- SURVEY.md 8a rows E1-E8.
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int SBM = 64, SBN = 64, SBK = 16, SThreads = 256;


template <typename T>
__global__ void __launch_bounds__(SThreads)
pwconv_simt_kernel(const T* __restrict__ x, long long M, int K, int N, const T* __restrict__ w,
                   const float* __restrict__ bias, const float* __restrict__ gate, int rows_per_img,
                   const T* __restrict__ residual, int act, T* __restrict__ out) {
  __shared__ float As[SBK][SBM + 1];
  __shared__ float Bs[SBK][SBN + 1];
  const long long m0 = (long long)blockIdx.x * SBM;
  const int n0 = blockIdx.y * SBN;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 4x4 outputs each
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += SBK) {
    // A tile: 64 rows x 16 k
    for (int i = tid; i < SBM * SBK; i += SThreads) {
      int r = i / SBK, kk = i % SBK;
      long long m = m0 + r;
      int k = k0 + kk;
      float v = 0.f;
      if (m < M && k < K) {
        v = to_f<T>(x[m * K + k]);
        if (gate) {
          v *= gate[(m / rows_per_img) * K + k];
          // the tensor-core path rounds the gated operand to bf16; mirror that for bf16 storage
          if constexpr (is_bf16<T>::value) v = __bfloat162float(__float2bfloat16_rn(v));
        }
      }
      As[kk][r] = v;
    }
    for (int i = tid; i < SBN * SBK; i += SThreads) {
      int c = i / SBK, kk = i % SBK;
      int nn = n0 + c, k = k0 + kk;
      Bs[kk][c] = (nn < N && k < K) ? to_f<T>(w[(size_t)nn * K + k]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SBK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int nn = n0 + tx * 4 + j;
      if (nn >= N) continue;
      float v = acc[i][j] + bias[nn];
      v = apply_act<is_bf16<T>::value>(v, act);
      if (residual) v += to_f<T>(residual[m * N + nn]);
      out[m * N + nn] = from_f<T>(v);
    }
  }
}

cudaError_t launch_pwconv_simt(const void* x, long long m, int cin, int cout, const void* wgt, const float* bias,
                               const float* gate, int rows_per_img, const void* residual, int act, void* out,
                               int dtype, cudaStream_t s) {
  if (m <= 0) return cudaSuccess;
  dim3 grid((unsigned)((m + SBM - 1) / SBM), (cout + SBN - 1) / SBN);
  if (dtype == kBF16) {
    pwconv_simt_kernel<bf16><<<grid, SThreads, 0, s>>>(
        reinterpret_cast<const bf16*>(x), m, cin, cout, reinterpret_cast<const bf16*>(wgt), bias, gate,
        rows_per_img, reinterpret_cast<const bf16*>(residual), act, reinterpret_cast<bf16*>(out));
  } else {
    pwconv_simt_kernel<float><<<grid, SThreads, 0, s>>>(
        reinterpret_cast<const float*>(x), m, cin, cout, reinterpret_cast<const float*>(wgt), bias, gate,
        rows_per_img, reinterpret_cast<const float*>(residual), act, reinterpret_cast<float*>(out));
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
