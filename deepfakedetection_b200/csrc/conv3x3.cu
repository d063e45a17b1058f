// Dense 3x3 conv (stride 1|2, pad 1) as a CUDA-core implicit GEMM, NHWC, any dtype -- the *strict fp32* path
// and the cross-check / fallback for the tcgen05 kernel in conv3x3_tc.cu.
//   out[m, co] = act( sum_{tap,c} x[img, oy*s + ky - 1, ox*s + kx - 1, c] * w[co, tap*Cin + c] + bias[co] ) (+ res[m, co])
// Replaces (inside the reference's `model(images)`, /root/reference/orchestration/orchestrator.py:590):
//   timm EfficientFormerV2 `stem.conv2` and `stages.N.downsample.conv` (conv bias + BN folded),
//   fastervit `patch_embed.conv_down.3`, `levels.{0,1}.blocks.N.conv{1,2}` (+BN, GELU / residual) and
//   `levels.N.downsample.reduction.0`.
// M = n*Ho*Wo output pixels, N = Cout, K = 9*Cin with k = (ky*3 + kx)*Cin + c, so that with NHWC activations every
// 8-channel chunk of a K tile is one contiguous 16-byte (bf16) / 32-byte (fp32) load; zero padding = skipped loads.
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int CBM = 64, CBN = 64, CBK = 16, CThreads = 256;

template <typename T>
__global__ void __launch_bounds__(CThreads)
conv3x3_simt_kernel(const T* __restrict__ x, int n, int H, int W, int Cin, int Cout, int stride,
                    const T* __restrict__ w, const float* __restrict__ bias, const T* __restrict__ residual, int act,
                    T* __restrict__ out, int Ho, int Wo, int cinp) {
  __shared__ float As[CBK][CBM + 1];
  __shared__ float Bs[CBK][CBN + 1];
  __shared__ int s_img[CBM], s_iy[CBM], s_ix[CBM];
  const long long M = (long long)n * Ho * Wo;
  const long long m0 = (long long)blockIdx.x * CBM;
  const int n0 = blockIdx.y * CBN;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  if (tid < CBM) {
    const long long m = m0 + tid;
    int img = -1, iy = 0, ix = 0;
    if (m < M) {
      img = (int)(m / (Ho * Wo));
      const int r = (int)(m - (long long)img * Ho * Wo);
      const int oy = r / Wo, ox = r - oy * Wo;
      iy = oy * stride - 1;
      ix = ox * stride - 1;
    }
    s_img[tid] = img;
    s_iy[tid] = iy;
    s_ix[tid] = ix;
  }
  __syncthreads();
  const int K = 9 * Cin;
  const size_t wrow = (size_t)9 * cinp;   // weight row pitch: every tap holds cinp (>= Cin) channels
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += CBK) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if (tid < 2 * CBM) {  // A: thread = (pixel row, 8-channel chunk)
      const int r = tid >> 1, ch = tid & 1, k8 = k0 + ch * 8;
      if (k8 < K && s_img[r] >= 0) {
        const int tap = k8 / Cin, c = k8 - tap * Cin;
        const int ky = tap / 3, kx = tap - ky * 3;
        const int iy = s_iy[r] + ky, ix = s_ix[r] + kx;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) load8<T>(x + (((size_t)s_img[r] * H + iy) * W + ix) * Cin + c, v);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) As[ch * 8 + j][r] = v[j];
    } else {  // B: thread = (output channel, 8-element chunk of its K row)
      const int t = tid - 2 * CBM, cidx = t >> 1, ch = t & 1, k8 = k0 + ch * 8, co = n0 + cidx;
      if (k8 < K && co < Cout) {
        const int tap = k8 / Cin, c = k8 - tap * Cin;
        load8<T>(w + (size_t)co * wrow + (size_t)tap * cinp + c, v);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) Bs[ch * 8 + j][cidx] = v[j];
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < CBK; ++kk) {
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
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int nn = n0 + tx * 4 + j;
      if (nn >= Cout) continue;
      float r = apply_act<is_bf16<T>::value>(acc[i][j] + bias[nn], act);
      if (residual) r += to_f<T>(residual[m * Cout + nn]);
      out[m * Cout + nn] = from_f<T>(r);
    }
  }
}

cudaError_t launch_conv3x3_simt(const void* x, int n, int h, int w, int cin, int cout, int stride, const void* wgt,
                                const float* bias, const void* residual, int act, void* out, int dtype,
                                cudaStream_t s, int cinp) {
  if (n <= 0) return cudaSuccess;
  if (cinp <= 0) cinp = cin;
  if (cin % 8 || cinp % 8 || cinp < cin || (stride != 1 && stride != 2)) return cudaErrorInvalidValue;
  const int Ho = (h - 1) / stride + 1, Wo = (w - 1) / stride + 1;  // pad 1, k 3
  const long long M = (long long)n * Ho * Wo;
  dim3 grid((unsigned)((M + CBM - 1) / CBM), (cout + CBN - 1) / CBN);
  if (dtype == kBF16) {
    conv3x3_simt_kernel<bf16><<<grid, CThreads, 0, s>>>(
        reinterpret_cast<const bf16*>(x), n, h, w, cin, cout, stride, reinterpret_cast<const bf16*>(wgt), bias,
        reinterpret_cast<const bf16*>(residual), act, reinterpret_cast<bf16*>(out), Ho, Wo, cinp);
  } else {
    conv3x3_simt_kernel<float><<<grid, CThreads, 0, s>>>(
        reinterpret_cast<const float*>(x), n, h, w, cin, cout, stride, reinterpret_cast<const float*>(wgt), bias,
        reinterpret_cast<const float*>(residual), act, reinterpret_cast<float*>(out), Ho, Wo, cinp);
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
