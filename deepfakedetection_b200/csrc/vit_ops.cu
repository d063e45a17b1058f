// Memory-bound glue ops of the two transformer-style backbones (EfficientFormerV2-S1, FasterViT-0/2), NHWC / token-major,
// any engine dtype (bf16 | fp32), fp32 math.  Each replaces a piece of the reference's `model(images)` call
// (/root/reference/orchestration/orchestrator.py:590) whose arithmetic lives in timm / fastervit:
//   layernorm          : nn.LayerNorm / LayerNorm2d over the channel (innermost) dim        (fastervit HAT norm1/2, downsample.norm)
//   upsample2x_act     : nn.Upsample(scale 2, bilinear, align_corners=False) + GELU          (timm Attention2d, stride 2)
//   win_part / win_rev : window_partition / window_reverse, carrier-token slots left free    (fastervit FasterViTLayer.forward)
//   ct_pool            : AvgPool2d(5, 3) of the pos-embedded map -> 4x4 carrier tokens, windowed order (TokenInitializer)
//   add_pos            : x[:, rows] += precomputed PosEmbMLPSwinv1D table                     (HAT.pos_embed / hat_pos_embed)
//   copy_rows          : carrier tokens <-> window sequences (torch.cat / split in HAT.forward)
//   pool_affine_fc     : [residual + scale*A + shift] -> BN -> global mean -> Linear          (both nets' head)
//   cam_linear, lin_head(_backward): Grad-CAM when everything after the target activation is affine
#include "common.cuh"
#include "kernels.h"

#include <algorithm>

namespace b200df {

// ---- LayerNorm over the innermost dim ------------------------------------------------------------------------
// One warp per R rows; the rows (C = 8 * 32 * NV channels at most) stay in registers between the mean, variance and
// normalise passes, and all R * NV 16-byte loads of a warp are issued before the first reduction (HBM-bound op: what
// matters is bytes in flight per SM).
template <typename T, int NV, int R>
__global__ void __launch_bounds__(256)
layernorm_kernel(const T* __restrict__ x, T* __restrict__ y, const float* __restrict__ gamma,
                 const float* __restrict__ beta, long long rows, int C, float eps) {
  const int lane = threadIdx.x & 31;
  const long long row0 = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * R;
  griddep_wait();
  griddep_launch_dependents();
  if (row0 >= rows) return;
  const int nv = C >> 3;
  float v[R][NV][8];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (row0 + r < rows && vi < nv) {
        load8<T>(x + (row0 + r) * C + vi * 8, v[r][i]);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[r][i][e] = 0.f;
      }
    }
  float g[NV][8], b[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = lane + 32 * i;
    if (vi < nv) {
      load8<float>(gamma + vi * 8, g[i]);
      load8<float>(beta + vi * 8, b[i]);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (row0 + r >= rows) break;
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) sum += v[r][i][e];   // lanes past the row hold zeros
    const float mean = warp_sum(sum) / (float)C;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (lane + 32 * i < nv) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float d = v[r][i][e] - mean;
          sq = fmaf(d, d, sq);
        }
      }
    }
    const float rstd = 1.0f / sqrtf(warp_sum(sq) / (float)C + eps);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nv) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf((v[r][i][e] - mean) * rstd, g[i][e], b[i][e]);
        store8<T>(y + (row0 + r) * C + vi * 8, o);
      }
    }
  }
}

// C <= 128 (FasterViT's level-0 / level-1 downsample norms, C = 64 .. 192 / 2): a row is only C / 8 <= 16 vectors, so a warp
// takes 32 / LPR rows at once, LPR lanes each (the full-warp kernel above left 16-24 lanes idle: 183 us for the 56 x 56 x 64
// norm of FasterViT-0 against 25 us of HBM time).  R such row groups per warp; reductions stay inside the LPR-lane group.
template <typename T, int LPR, int R>
__global__ void __launch_bounds__(256)
layernorm_small_kernel(const T* __restrict__ x, T* __restrict__ y, const float* __restrict__ gamma,
                       const float* __restrict__ beta, long long rows, int C, float eps) {
  constexpr int RPW = 32 / LPR;                     // rows per warp pass
  const int lane = threadIdx.x & 31, sub = lane / LPR, vi = lane % LPR;
  const long long warp_id = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long row0 = warp_id * (RPW * R) + sub;
  const int nv = C >> 3;
  const bool act = vi < nv;
  griddep_wait();
  griddep_launch_dependents();
  float v[R][8];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const long long row = row0 + (long long)r * RPW;
    if (act && row < rows) {
      load8<T>(x + row * C + vi * 8, v[r]);
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[r][e] = 0.f;
    }
  }
  float g[8], b[8];
  if (act) {
    load8<float>(gamma + vi * 8, g);
    load8<float>(beta + vi * 8, b);
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const long long row = row0 + (long long)r * RPW;
    float sum = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) sum += v[r][e];
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum / (float)C;
    float sq = 0.f;
    if (act) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = v[r][e] - mean;
        sq = fmaf(d, d, sq);
      }
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = 1.0f / sqrtf(sq / (float)C + eps);
    if (act && row < rows) {
      float o8[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o8[e] = fmaf((v[r][e] - mean) * rstd, g[e], b[e]);
      store8<T>(y + row * C + vi * 8, o8);
    }
  }
}

template <typename T, int LPR, int R>
static cudaError_t ln_small_launch(const void* x, void* y, const float* gamma, const float* beta, long long rows, int c, float eps,
                                   cudaStream_t s) {
  const long long per_warp = (32 / LPR) * R;
  const long long warps = (rows + per_warp - 1) / per_warp;
  const unsigned grid = (unsigned)((warps + 7) / 8);
  return launch_pdl(layernorm_small_kernel<T, LPR, R>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const T*>(x),
                    reinterpret_cast<T*>(y), gamma, beta, rows, c, eps);
}

template <typename T, int NV, int R>
static cudaError_t ln_launch(const void* x, void* y, const float* gamma, const float* beta, long long rows, int c, float eps,
                             cudaStream_t s) {
  const long long warps = (rows + R - 1) / R;
  const unsigned grid = (unsigned)((warps + 7) / 8);
  return launch_pdl(layernorm_kernel<T, NV, R>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const T*>(x),
                    reinterpret_cast<T*>(y), gamma, beta, rows, c, eps);
}

cudaError_t launch_layernorm(const void* x, void* y, const float* gamma, const float* beta, long long rows, int c,
                             float eps, int dtype, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  if (c % 8 || c > 1024) return cudaErrorInvalidValue;
  const int nvl = (c / 8 + 31) / 32;   // 16-byte vectors per lane
  count_launch();
  if (c <= 128) {      // several rows per warp
    if (dtype == kBF16)
      return c <= 64 ? ln_small_launch<bf16, 8, 4>(x, y, gamma, beta, rows, c, eps, s)
                     : ln_small_launch<bf16, 16, 4>(x, y, gamma, beta, rows, c, eps, s);
    return c <= 64 ? ln_small_launch<float, 8, 4>(x, y, gamma, beta, rows, c, eps, s)
                   : ln_small_launch<float, 16, 4>(x, y, gamma, beta, rows, c, eps, s);
  }
#define LN_CASE(T_)                                                                         \
  if (nvl == 1) return ln_launch<T_, 1, 4>(x, y, gamma, beta, rows, c, eps, s);             \
  if (nvl == 2) return ln_launch<T_, 2, 2>(x, y, gamma, beta, rows, c, eps, s);             \
  return ln_launch<T_, 4, 1>(x, y, gamma, beta, rows, c, eps, s)
  if (dtype == kBF16) { LN_CASE(bf16); }
  LN_CASE(float);
#undef LN_CASE
}

// ---- bilinear x2 upsample (align_corners = False) + activation -------------------------------------------------
// dst index d -> src = max((d + 0.5) / 2 - 0.5, 0); i0 = floor(src), i1 = min(i0 + 1, side - 1), frac = src - i0.
// One CTA per (image, output row): the row's two source rows and the y weight are fixed, all index math is 32-bit
// (the flat 64-bit index with three runtime divisions per 8 outputs made the first version issue-bound at 85 us).
template <typename T>
__global__ void __launch_bounds__(256)
upsample2x_act_kernel(const T* __restrict__ x, T* __restrict__ y, int side, int C, int act) {
  const int C8 = C >> 3, os = side * 2;
  const int oy = blockIdx.x % os, img = blockIdx.x / os;
  const float sy = fmaxf((oy + 0.5f) * 0.5f - 0.5f, 0.f);
  const int y0 = (int)sy, y1 = min(y0 + 1, side - 1);
  const float fy = sy - (float)y0;
  const T* r0 = x + ((size_t)img * side + y0) * side * C;
  const T* r1 = x + ((size_t)img * side + y1) * side * C;
  T* out = y + ((size_t)img * os + oy) * os * C;
  griddep_wait();
  griddep_launch_dependents();
  for (int i = threadIdx.x; i < os * C8; i += blockDim.x) {
    const int ox = i / C8, c8 = i - ox * C8;
    const float sx = fmaxf((ox + 0.5f) * 0.5f - 0.5f, 0.f);
    const int x0 = (int)sx, x1 = min(x0 + 1, side - 1);
    const float fx = sx - (float)x0;
    float a[8], b[8], c[8], d[8], o[8];
    load8<T>(r0 + x0 * C + c8 * 8, a);
    load8<T>(r0 + x1 * C + c8 * 8, b);
    load8<T>(r1 + x0 * C + c8 * 8, c);
    load8<T>(r1 + x1 * C + c8 * 8, d);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      // torch's order: interpolate along x within each row, then along y
      const float top = a[e] * (1.f - fx) + b[e] * fx;
      const float bot = c[e] * (1.f - fx) + d[e] * fx;
      o[e] = apply_act<is_bf16<T>::value>(top * (1.f - fy) + bot * fy, act);
    }
    store8<T>(out + (size_t)i * 8, o);
  }
}

cudaError_t launch_upsample2x_act(const void* x, void* y, int n, int side, int c, int act, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c % 8) return cudaErrorInvalidValue;
  const unsigned grid = (unsigned)(n * side * 2);
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(upsample2x_act_kernel<bf16>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const bf16*>(x),
                      reinterpret_cast<bf16*>(y), side, c, act);
  return launch_pdl(upsample2x_act_kernel<float>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const float*>(x),
                    reinterpret_cast<float*>(y), side, c, act);
}

// ---- window partition / reverse ----------------------------------------------------------------------------------
// NHWC [n, side, side, C] <-> sequences [n * nw^2, ct + ws^2, C] (window (wy, wx) -> sequence wy*nw + wx; token
// (iy, ix) -> row ct + iy*ws + ix; rows [0, ct) belong to the carrier tokens and are not touched).
// `pos` (nullable, [ws^2, C] fp32) is added on the way in (HAT.pos_embed of the level's first block).
template <typename T, bool kReverse>
__global__ void __launch_bounds__(256)
window_kernel(const T* __restrict__ src, T* __restrict__ dst, int n, int side, int ws, int ct, int C,
              const float* __restrict__ pos) {
  const int C8 = C >> 3, nw = side / ws, seq_rows = ct + ws * ws;
  const long long total = (long long)n * side * side * C8;
  griddep_wait();
  griddep_launch_dependents();
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    long long r = idx / C8;
    const int x = (int)(r % side);
    r /= side;
    const int y = (int)(r % side);
    const int img = (int)(r / side);
    const int wy = y / ws, wx = x / ws, iy = y - wy * ws, ix = x - wx * ws;
    const size_t tok = ((size_t)(img * nw + wy) * nw + wx) * seq_rows + ct + iy * ws + ix;
    const size_t pix = ((size_t)img * side + y) * side + x;
    float f[8];
    if constexpr (!kReverse) {
      load8<T>(src + pix * C + c8 * 8, f);
      if (pos) {
        float p[8];
        load8<float>(pos + (size_t)(iy * ws + ix) * C + c8 * 8, p);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] += p[e];
      }
      store8<T>(dst + tok * C + c8 * 8, f);
    } else {
      load8<T>(src + tok * C + c8 * 8, f);
      store8<T>(dst + pix * C + c8 * 8, f);
    }
  }
}

cudaError_t launch_window(const void* src, void* dst, int n, int side, int ws, int ct, int c, const float* pos,
                          bool reverse, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c % 8 || side % ws) return cudaErrorInvalidValue;
  const long long total = (long long)n * side * side * (c / 8);
  const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, (long long)kNumSMs * 16);
  count_launch();
#define WIN_LAUNCH(T_, R_)                                                                                        \
  return launch_pdl(window_kernel<T_, R_>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const T_*>(src),          \
                    reinterpret_cast<T_*>(dst), n, side, ws, ct, c, pos)
  if (dtype == kBF16) {
    if (reverse) WIN_LAUNCH(bf16, true);
    WIN_LAUNCH(bf16, false);
  }
  if (reverse) WIN_LAUNCH(float, true);
  WIN_LAUNCH(float, false);
#undef WIN_LAUNCH
}

// ---- carrier-token initialiser: AvgPool2d(kernel, stride) of [n, side, side, C] -> [n, g*g, C] in WINDOWED order ----
// g = (side - kernel) / stride + 1 pooled positions per side (4); token index = ((wy*nwin + wx)*cs + iy)*cs + ix for
// pooled position (wy*cs + iy, wx*cs + ix), cs = carrier tokens per window side (2).
template <typename T>
__global__ void __launch_bounds__(256)
ct_pool_kernel(const T* __restrict__ x, T* __restrict__ ct, int n, int side, int C, int kernel, int stride, int g,
               int cs) {
  const int C8 = C >> 3, nwin = g / cs;
  const long long total = (long long)n * g * g * C8;
  const float inv = 1.0f / (float)(kernel * kernel);
  griddep_wait();
  griddep_launch_dependents();
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    long long r = idx / C8;
    const int px = (int)(r % g);
    r /= g;
    const int py = (int)(r % g);
    const int img = (int)(r / g);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    for (int dy = 0; dy < kernel; ++dy)
      for (int dx = 0; dx < kernel; ++dx) {
        float f[8];
        load8<T>(x + (((size_t)img * side + py * stride + dy) * side + px * stride + dx) * C + c8 * 8, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] *= inv;
    const int wy = py / cs, iy = py - wy * cs, wx = px / cs, ix = px - wx * cs;
    const int tok = ((wy * nwin + wx) * cs + iy) * cs + ix;
    store8<T>(ct + ((size_t)img * g * g + tok) * C + c8 * 8, acc);
  }
}

cudaError_t launch_ct_pool(const void* x, void* ct, int n, int side, int c, int kernel, int stride, int cs, int dtype,
                           cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int g = (side - kernel) / stride + 1;
  if (c % 8 || g % cs) return cudaErrorInvalidValue;
  const long long total = (long long)n * g * g * (c / 8);
  const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, (long long)kNumSMs * 16);
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(ct_pool_kernel<bf16>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const bf16*>(x),
                      reinterpret_cast<bf16*>(ct), n, side, c, kernel, stride, g, cs);
  return launch_pdl(ct_pool_kernel<float>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const float*>(x),
                    reinterpret_cast<float*>(ct), n, side, c, kernel, stride, g, cs);
}

// ---- x[seq, row_off + r, :] += pos[r, :]  for r < nrows ----------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
add_pos_kernel(T* __restrict__ x, long long nseq, int seq_rows, int row_off, int nrows, int C,
               const float* __restrict__ pos) {
  const int C8 = C >> 3;
  const long long total = nseq * nrows * C8;
  griddep_wait();
  griddep_launch_dependents();
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    long long r = idx / C8;
    const int row = (int)(r % nrows);
    const long long seq = r / nrows;
    T* p = x + ((size_t)seq * seq_rows + row_off + row) * C + c8 * 8;
    float f[8], q[8];
    load8<T>(p, f);
    load8<float>(pos + (size_t)row * C + c8 * 8, q);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] += q[e];
    store8<T>(p, f);
  }
}

cudaError_t launch_add_pos(void* x, long long nseq, int seq_rows, int row_off, int nrows, int c, const float* pos,
                           int dtype, cudaStream_t s) {
  if (nseq <= 0 || nrows <= 0) return cudaSuccess;
  if (c % 8) return cudaErrorInvalidValue;
  const long long total = nseq * nrows * (c / 8);
  const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, (long long)kNumSMs * 16);
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(add_pos_kernel<bf16>, dim3(grid), dim3(256), 0, s, reinterpret_cast<bf16*>(x), nseq, seq_rows,
                      row_off, nrows, c, pos);
  return launch_pdl(add_pos_kernel<float>, dim3(grid), dim3(256), 0, s, reinterpret_cast<float*>(x), nseq, seq_rows,
                    row_off, nrows, c, pos);
}

// ---- dst[seq, dst_off + r, :] = src[seq, src_off + r, :]  for r < nrows -------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
copy_rows_kernel(const T* __restrict__ src, T* __restrict__ dst, long long nseq, int src_rows, int src_off,
                 int dst_rows, int dst_off, int nrows, int C) {
  const int C8 = C >> 3;
  const long long total = nseq * nrows * C8;
  griddep_wait();
  griddep_launch_dependents();
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    long long r = idx / C8;
    const int row = (int)(r % nrows);
    const long long seq = r / nrows;
    const uint4* sp = reinterpret_cast<const uint4*>(src + ((size_t)seq * src_rows + src_off + row) * C + c8 * 8);
    uint4* dp = reinterpret_cast<uint4*>(dst + ((size_t)seq * dst_rows + dst_off + row) * C + c8 * 8);
    dp[0] = sp[0];
    if constexpr (sizeof(T) == 4) dp[1] = sp[1];
  }
}

cudaError_t launch_copy_rows(const void* src, void* dst, long long nseq, int src_rows, int src_off, int dst_rows,
                             int dst_off, int nrows, int c, int dtype, cudaStream_t s) {
  if (nseq <= 0 || nrows <= 0) return cudaSuccess;
  if (c % 8) return cudaErrorInvalidValue;
  const long long total = nseq * nrows * (c / 8);
  const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, (long long)kNumSMs * 16);
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(copy_rows_kernel<bf16>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const bf16*>(src),
                      reinterpret_cast<bf16*>(dst), nseq, src_rows, src_off, dst_rows, dst_off, nrows, c);
  return launch_pdl(copy_rows_kernel<float>, dim3(grid), dim3(256), 0, s, reinterpret_cast<const float*>(src),
                    reinterpret_cast<float*>(dst), nseq, src_rows, src_off, dst_rows, dst_off, nrows, c);
}

// ---- head: [resid + a_scale*A + a_shift] -> (x n_scale + n_shift) -> mean_hw -> Linear -------------------------------------
// One CTA per image; thread = channel (coalesced over the NHWC rows), fixed summation order (bit-reproducible).
template <typename T>
__global__ void __launch_bounds__(256)
pool_affine_fc_kernel(const T* __restrict__ a, const T* __restrict__ resid, int HW, int C,
                      const float* __restrict__ a_scale, const float* __restrict__ a_shift,
                      const float* __restrict__ n_scale, const float* __restrict__ n_shift,
                      const float* __restrict__ w, const float* __restrict__ b, int NC, float* __restrict__ logits) {
  extern __shared__ float pooled[];  // [C]
  const int img = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* an = a + (size_t)img * HW * C;
  const T* rn = resid ? resid + (size_t)img * HW * C : nullptr;
  griddep_wait();
  griddep_launch_dependents();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float as = a_scale ? a_scale[c] : 1.f, ah = a_shift ? a_shift[c] : 0.f;
    float sum = 0.f;
    for (int p = 0; p < HW; ++p) {
      float v = fmaf(to_f<T>(an[(size_t)p * C + c]), as, ah);
      if (rn) v += to_f<T>(rn[(size_t)p * C + c]);
      sum += v;
    }
    pooled[c] = fmaf(sum / (float)HW, n_scale[c], n_shift[c]);
  }
  __syncthreads();
  for (int j = warp; j < NC; j += blockDim.x >> 5) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(__ldg(w + (size_t)j * C + c), pooled[c], s);
    s = warp_sum(s);
    if (lane == 0) logits[(size_t)img * NC + j] = s + b[j];
  }
}

cudaError_t launch_pool_affine_fc(const void* a, const void* resid, int n, int hw, int c, const float* a_scale,
                                  const float* a_shift, const float* n_scale, const float* n_shift, const float* w,
                                  const float* b, int num_classes, float* logits, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(pool_affine_fc_kernel<bf16>, dim3(n), dim3(256), c * sizeof(float), s,
                      reinterpret_cast<const bf16*>(a), reinterpret_cast<const bf16*>(resid), hw, c, a_scale, a_shift,
                      n_scale, n_shift, w, b, num_classes, logits);
  return launch_pdl(pool_affine_fc_kernel<float>, dim3(n), dim3(256), c * sizeof(float), s,
                    reinterpret_cast<const float*>(a), reinterpret_cast<const float*>(resid), hw, c, a_scale, a_shift,
                    n_scale, n_shift, w, b, num_classes, logits);
}

// ---- Grad-CAM with a spatially constant gradient: cam[n, p] = relu(sum_k wcam[c_n, k] A[n, p, k]) ------------------------------
// wcam[c, k] = d logit_c / d A[k, p] (the same for every p).  target_cls == nullptr -> argmax of `logits`.
// An out-of-range class writes NaN (the C ABI validates explicit targets on the host before the launch).
template <typename T>
__global__ void __launch_bounds__(256)
cam_linear_kernel(const T* __restrict__ a, int HW, int C, const float* __restrict__ wcam, int NC,
                  const int32_t* __restrict__ target_cls, const float* __restrict__ logits, float* __restrict__ cam) {
  const int img = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  griddep_wait();
  griddep_launch_dependents();
  int cls;
  if (target_cls) {
    cls = target_cls[img];
  } else {
    cls = 0;
    float best = logits[(size_t)img * NC];
    for (int j = 1; j < NC; ++j) {
      const float v = logits[(size_t)img * NC + j];
      if (v > best) { best = v; cls = j; }
    }
  }
  const bool ok = cls >= 0 && cls < NC;
  const float* wk = wcam + (size_t)(ok ? cls : 0) * C;
  const T* an = a + (size_t)img * HW * C;
  for (int p = warp; p < HW; p += blockDim.x >> 5) {
    float s = 0.f;
    for (int k = lane; k < C; k += 32) s = fmaf(__ldg(wk + k), to_f<T>(an[(size_t)p * C + k]), s);
    s = warp_sum(s);
    if (lane == 0) cam[(size_t)img * HW + p] = ok ? fmaxf(s, 0.f) : __int_as_float(0x7fc00000);
  }
}

cudaError_t launch_cam_linear(const void* a, int n, int hw, int c, const float* wcam, int num_classes,
                              const int32_t* target_cls, const float* logits, float* cam_lowres, int dtype,
                              cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(cam_linear_kernel<bf16>, dim3(n), dim3(256), 0, s, reinterpret_cast<const bf16*>(a), hw, c, wcam,
                      num_classes, target_cls, logits, cam_lowres);
  return launch_pdl(cam_linear_kernel<float>, dim3(n), dim3(256), 0, s, reinterpret_cast<const float*>(a), hw, c, wcam,
                    num_classes, target_cls, logits, cam_lowres);
}

// ---- split (hook-protocol) head for the affine case ---------------------------------------------------------------------
// logits[n, j] = aux[n, j] + sum_{k,p} wcam[j, k] A[n, k, p]   (A fp32 NCHW, as handed to pytorch-grad-cam's hooks)
// mode 0: logits = aux + lin(A);  mode 1: aux_out = logits_in - lin(A)  (computed by the trunk entry point)
__global__ void __launch_bounds__(256)
lin_head_kernel(const float* __restrict__ a_nchw, const float* __restrict__ aux, int HW, int C,
                const float* __restrict__ wcam, int NC, int mode, float* __restrict__ out) {
  __shared__ float red[8];
  const int img = blockIdx.x, j = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* an = a_nchw + (size_t)img * C * HW;
  float s = 0.f;
  for (int k = warp; k < C; k += 8) {
    float t = 0.f;
    for (int p = lane; p < HW; p += 32) t += an[(size_t)k * HW + p];
    s = fmaf(__ldg(wcam + (size_t)j * C + k), t, s);
  }
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    const float base = aux[(size_t)img * NC + j];
    out[(size_t)img * NC + j] = mode == 0 ? base + t : base - t;
  }
}

cudaError_t launch_lin_head(const float* a_nchw, const float* aux, int n, int hw, int c, const float* wcam,
                            int num_classes, int mode, float* out, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  lin_head_kernel<<<dim3(n, num_classes), 256, 0, s>>>(a_nchw, aux, hw, c, wcam, num_classes, mode, out);
  count_launch();
  return cudaGetLastError();
}

// grad_A[n, k, p] = sum_j grad_logits[n, j] wcam[j, k]
__global__ void __launch_bounds__(256)
lin_head_backward_kernel(const float* __restrict__ grad_logits, int HW, int C, const float* __restrict__ wcam, int NC,
                         float* __restrict__ grad_a) {
  const int img = blockIdx.x;
  for (int idx = threadIdx.x; idx < C * HW; idx += blockDim.x) {
    const int k = idx / HW;
    float g = 0.f;
    for (int j = 0; j < NC; ++j) g = fmaf(grad_logits[(size_t)img * NC + j], __ldg(wcam + (size_t)j * C + k), g);
    grad_a[(size_t)img * C * HW + idx] = g;
  }
}

cudaError_t launch_lin_head_backward(const float* grad_logits, int n, int hw, int c, const float* wcam,
                                     int num_classes, float* grad_a, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  lin_head_backward_kernel<<<n, 256, 0, s>>>(grad_logits, hw, c, wcam, num_classes, grad_a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
