// Depthwise kxk conv (k = 3|5, stride 1|2, pad (k-1)/2 or an explicit pad-before for TF-"same" layers) + folded-BN bias + SiLU, NHWC,
// with the squeeze-excite global pooling fused in (per-image per-channel sums).
// Replaces timm `conv_dw` + `bn1/bn2` and the `se` mean_hw, i.e. the
// depthwise leg of every MBConv inside the reference's `model(images)`
// (/root/reference/orchestration/orchestrator.py:590).
//
// HBM-bound: each thread owns 8 channels (one 128-bit vector) of a 1 x TW output strip and slides the
// k x (TW*s + k - s) input window through registers, so every input vector is loaded once per
// thread and re-used by up to k taps; neighbouring strips hit L1.  Accumulation is fp32.
// Pool sums: Q24 fixed-point integer atomics (shared per block, then one global atomic per (block, channel)),
// so the sums are independent of scheduling order (bit-reproducible).
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int kDwThreads = 256;

template <typename T, int K, int S, int TW>
__global__ void __launch_bounds__(kDwThreads)
dwconv_kernel(const T* __restrict__ x, T* __restrict__ out, const float* __restrict__ wgt,
              const float* __restrict__ bias, pool_t* __restrict__ pool, int H, int W, int C, int Ho, int Wo, int PAD,
              int act) {
  extern __shared__ unsigned long long spool[];  // [C] Q24
  constexpr int IW = (TW - 1) * S + K;  // input columns a strip touches
  constexpr bool kFast = is_bf16<T>::value;
  const int n = blockIdx.y;
  const int CG = C >> 3;
  const int WT = (Wo + TW - 1) / TW;
  const int total = Ho * WT * CG;
  const bool do_pool = pool != nullptr;
  if (do_pool) {
    for (int i = threadIdx.x; i < C; i += kDwThreads) spool[i] = 0ull;
    __syncthreads();
  }
  const T* xn = x + (size_t)n * H * W * C;
  T* on = out + (size_t)n * Ho * Wo * C;

  for (int idx = blockIdx.x * kDwThreads + threadIdx.x; idx < total; idx += gridDim.x * kDwThreads) {
    const int cg = idx % CG;
    const int rest = idx / CG;
    const int wt = rest % WT;
    const int oh = rest / WT;
    const int c0 = cg << 3;
    const int ow0 = wt * TW;
    float acc[TW][8];
    {
      float b[8];
      load8<float>(bias + c0, b);
#pragma unroll
      for (int j = 0; j < TW; ++j)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[j][e] = b[e];
    }
    const int iw0 = ow0 * S - PAD;
#pragma unroll
    for (int ky = 0; ky < K; ++ky) {
      const int ih = oh * S - PAD + ky;
      if (ih < 0 || ih >= H) continue;
      float wr[K][8];
#pragma unroll
      for (int kx = 0; kx < K; ++kx) load8<float>(wgt + (size_t)(ky * K + kx) * C + c0, wr[kx]);
      const T* xrow = xn + ((size_t)ih * W) * C + c0;
#pragma unroll
      for (int ix = 0; ix < IW; ++ix) {
        const int iw = iw0 + ix;
        if (iw < 0 || iw >= W) continue;
        float v[8];
        load8<T>(xrow + (size_t)iw * C, v);
#pragma unroll
        for (int j = 0; j < TW; ++j) {
          constexpr int dummy = 0;
          (void)dummy;
          const int kx = ix - j * S;
          if (kx >= 0 && kx < K) {
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[j][e] = fmaf(v[e], wr[kx][e], acc[j][e]);
          }
        }
      }
    }
    float psum[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) psum[e] = 0.f;
#pragma unroll
    for (int j = 0; j < TW; ++j) {
      const int ow = ow0 + j;
      if (ow < Wo) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          f[e] = apply_act<kFast>(acc[j][e], act);
        }
        if constexpr (kFast) {
          // pool what the consumer will actually read (bf16-rounded values)
          uint4 pk = pack_bf16x8(f);
          *reinterpret_cast<uint4*>(on + ((size_t)oh * Wo + ow) * C + c0) = pk;
          float r[8];
          unpack_bf16x8(pk, r);
#pragma unroll
          for (int e = 0; e < 8; ++e) psum[e] += r[e];
        } else {
          store8<T>(on + ((size_t)oh * Wo + ow) * C + c0, f);
#pragma unroll
          for (int e = 0; e < 8; ++e) psum[e] += f[e];
        }
      }
    }
    if (do_pool) {
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(&spool[c0 + e], (unsigned long long)__float2ll_rn(psum[e] * kPoolScale));
    }
  }
  if (do_pool) {
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += kDwThreads)
      atomicAdd(reinterpret_cast<unsigned long long*>(pool) + (size_t)n * C + i, spool[i]);
  }
}

template <typename T, int K, int S, int TW>
static cudaError_t dw_launch(const void* x, int n, int h, int w, int c, const float* wgt, const float* bias, void* out,
                             pool_t* pool, int pad_lo, int act, cudaStream_t s) {
  const int PAD = pad_lo < 0 ? (K - 1) / 2 : pad_lo;
  if (PAD > K - 1) return cudaErrorInvalidValue;
  const int Ho = (h - 1) / S + 1, Wo = (w - 1) / S + 1;  // total padding is always k - 1
  const int WT = (Wo + TW - 1) / TW;
  const long long total = (long long)Ho * WT * (c / 8);
  int bx = (int)((total + kDwThreads - 1) / kDwThreads);
  // keep >= ~2 waves over the chip but bound the number of per-block pool flushes
  int max_bx = (4 * kNumSMs * 8 + n - 1) / n;
  if (max_bx < 1) max_bx = 1;
  if (bx > max_bx) bx = max_bx;
  dim3 grid(bx, n);
  dwconv_kernel<T, K, S, TW><<<grid, kDwThreads, c * sizeof(unsigned long long), s>>>(
      reinterpret_cast<const T*>(x), reinterpret_cast<T*>(out), wgt, bias, pool, h, w, c, Ho, Wo, PAD, act);
  count_launch();
  return cudaGetLastError();
}

template <typename T>
static cudaError_t dw_dispatch(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                               const float* bias, void* out, pool_t* pool, int pad_lo, int act, cudaStream_t s) {
  if (k == 3 && stride == 1) return dw_launch<T, 3, 1, 4>(x, n, h, w, c, wgt, bias, out, pool, pad_lo, act, s);
  if (k == 3 && stride == 2) return dw_launch<T, 3, 2, 2>(x, n, h, w, c, wgt, bias, out, pool, pad_lo, act, s);
  if (k == 5 && stride == 1) return dw_launch<T, 5, 1, 4>(x, n, h, w, c, wgt, bias, out, pool, pad_lo, act, s);
  if (k == 5 && stride == 2) return dw_launch<T, 5, 2, 2>(x, n, h, w, c, wgt, bias, out, pool, pad_lo, act, s);
  return cudaErrorInvalidValue;
}

cudaError_t launch_dwconv(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                          const float* bias, void* out, pool_t* pool, int dtype, cudaStream_t s, int pad_lo, int act) {
  if (n <= 0) return cudaSuccess;
  if (c % 8 != 0 || c > 6144) return cudaErrorInvalidValue;
  if (dtype == kBF16) {
    DwPlan plan;  // bf16: TMA-staged kernel (dwconv_tma.cu) whenever the shape allows it
    if (plan_dwconv_tma(plan, x, n, h, w, c, k, stride, wgt, bias, out, pool, pad_lo, act) == 0)
      return launch_dwconv_tma(plan, s);
    return dw_dispatch<bf16>(x, n, h, w, c, k, stride, wgt, bias, out, pool, pad_lo, act, s);
  }
  return dw_dispatch<float>(x, n, h, w, c, k, stride, wgt, bias, out, pool, pad_lo, act, s);
}

}  // namespace b200df
