// Head and Grad-CAM kernels (everything downstream of the raw `conv_head` output A).
//   head_pool   : pooled[n,c] = mean_hw silu(A*scale + shift)          (timm bn2 + SiLU + global_pool)
//   classifier  : logits = pooled . W^T + b                            (timm `classifier`)
//   cam         : fused Grad-CAM -- channel weights (closed-form gradient through BN->SiLU->mean->Linear,
//                 reduced over hw) * A, summed over channels, ReLU      (pytorch-grad-cam's GradCAM)
//   cam_upsample: min-max -> bilinear x32 -> relu -> min-max            (pytorch-grad-cam scale_cam_image)
// Reference call sites: /root/reference/orchestration/orchestrator.py:590-592 (forward),
// /root/reference/web_ui.py:96-114 (target layer rule), web_ui.py:275-279 (GradCAM call).
// All HBM/latency-bound; one block per image, 128-bit NHWC loads, warp-shuffle reductions.
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int kTailThreads = 256;

// ---- head pool ----------------------------------------------------------------------------------
// One CTA per image.  A thread owns (8-channel group, pixel phase q of Q): its pixels q, q+Q, ... are independent
// 16-byte loads (all in flight at once), the Q partial sums meet in shared memory and are added in fixed order
// (bit-reproducible).  Q is chosen so that ~1024 threads are busy whatever C is.
template <typename T>
__global__ void __launch_bounds__(1024)
head_pool_kernel(const T* __restrict__ a, int HW, int C, int Q, const float* __restrict__ scale,
                 const float* __restrict__ shift, float* __restrict__ pooled) {
  extern __shared__ float part[];  // [Q][C]
  const int n = blockIdx.x;
  const int CG = C >> 3;
  const float inv = 1.0f / (float)HW;
  const T* an = a + (size_t)n * HW * C;
  griddep_wait();
  griddep_launch_dependents();
  for (int idx = threadIdx.x; idx < CG * Q; idx += blockDim.x) {
    const int q = idx / CG, cg = idx - q * CG;
    float sc[8], sh[8], acc[8];
    load8<float>(scale + cg * 8, sc);
    load8<float>(shift + cg * 8, sh);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    const T* ap = an + cg * 8;
#pragma unroll 4
    for (int p = q; p < HW; p += Q) {
      float v[8];
      load8<T>(ap + (size_t)p * C, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += silu_acc(fmaf(v[e], sc[e], sh[e]));
    }
    store8<float>(part + (size_t)q * C + cg * 8, acc);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float v = 0.f;
    for (int q = 0; q < Q; ++q) v += part[(size_t)q * C + c];
    pooled[(size_t)n * C + c] = v * inv;
  }
}

cudaError_t launch_head_pool(const void* a, int n, int hw, int c, const float* scale, const float* shift,
                             float* pooled, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int cg = c >> 3;
  if (cg < 1) return cudaErrorInvalidValue;
  int q = 1024 / cg;
  q = q < 1 ? 1 : (q > 8 ? 8 : q);
  if (q > hw) q = hw > 0 ? hw : 1;
  while (q > 1 && (size_t)q * c * sizeof(float) > 48 * 1024) --q;
  if ((size_t)q * c * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  int threads = (cg * q + 31) / 32 * 32;
  if (threads > 1024) threads = 1024;
  const size_t smem = (size_t)q * c * sizeof(float);
  count_launch();
  if (dtype == kBF16)
    return launch_pdl(head_pool_kernel<bf16>, dim3(n), dim3(threads), smem, s, reinterpret_cast<const bf16*>(a), hw, c, q,
                      scale, shift, pooled);
  return launch_pdl(head_pool_kernel<float>, dim3(n), dim3(threads), smem, s, reinterpret_cast<const float*>(a), hw, c, q,
                    scale, shift, pooled);
}

// ---- classifier ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
classifier_kernel(const float* __restrict__ pooled, int C, const float* __restrict__ w, const float* __restrict__ b,
                  int NC, float* __restrict__ logits) {
  const int n = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* pn = pooled + (size_t)n * C;
  griddep_wait();
  griddep_launch_dependents();
  for (int j = warp; j < NC; j += 4) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(__ldg(w + (size_t)j * C + c), pn[c], s);
    s = warp_sum(s);
    if (lane == 0) logits[(size_t)n * NC + j] = s + b[j];
  }
}

cudaError_t launch_classifier(const float* pooled, int n, int c, const float* w, const float* b, int num_classes,
                              float* logits, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  return launch_pdl(classifier_kernel, dim3(n), dim3(128), 0, s, pooled, c, w, b, num_classes, logits);
}

// ---- fused Grad-CAM -------------------------------------------------------------------------------
// w[n,k] = Wfc[c_n,k]/HW * scale_k * mean_hw silu'(A*scale_k + shift_k);  cam[n,p] = relu(sum_k w[n,k] A[n,p,k])
template <typename T>
__global__ void __launch_bounds__(1024)
cam_kernel(const T* __restrict__ a, int HW, int C, int Q, const float* __restrict__ scale, const float* __restrict__ shift,
           const float* __restrict__ wfc, int NC, const int32_t* __restrict__ target, const float* __restrict__ logits,
           float* __restrict__ cam) {
  extern __shared__ float wk[];  // [C] weights + [Q][C] partial sums
  const int n = blockIdx.x;
  const int CG = C >> 3;
  const T* an = a + (size_t)n * HW * C;
  int cls;
  if (target != nullptr) {
    cls = target[n];
  } else {  // argmax of logits (first max wins, like torch.argmax)
    cls = 0;
    float best = logits[(size_t)n * NC];
    for (int j = 1; j < NC; ++j) {
      float v = logits[(size_t)n * NC + j];
      if (v > best) { best = v; cls = j; }
    }
  }
  cls = min(max(cls, 0), NC - 1);
  const float inv = 1.0f / (float)HW;
  // mean_hw silu'(.) per channel: (8-channel group, pixel phase q of Q) per thread, Q partial sums combined in
  // fixed order (same scheme as head_pool_kernel)
  float* part = wk + C;  // [Q][C]
  for (int idx = threadIdx.x; idx < CG * Q; idx += blockDim.x) {
    const int q = idx / CG, cg = idx - q * CG;
    float sc[8], sh[8], acc[8];
    load8<float>(scale + cg * 8, sc);
    load8<float>(shift + cg * 8, sh);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll 4
    for (int p = q; p < HW; p += Q) {
      float v[8];
      load8<T>(an + (size_t)p * C + cg * 8, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float z = fmaf(v[e], sc[e], sh[e]);
        float sg = sigmoid_acc(z);
        acc[e] += sg * fmaf(z, 1.0f - sg, 1.0f);  // silu'(z)
      }
    }
    store8<float>(part + (size_t)q * C + cg * 8, acc);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f;
    for (int q = 0; q < Q; ++q) a += part[(size_t)q * C + c];
    wk[c] = __ldg(wfc + (size_t)cls * C + c) * inv * __ldg(scale + c) * (a * inv);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int p = warp; p < HW; p += nw) {
    float s = 0.f;
    for (int cg = lane; cg < CG; cg += 32) {
      float v[8];
      load8<T>(an + (size_t)p * C + cg * 8, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) s = fmaf(v[e], wk[cg * 8 + e], s);
    }
    s = warp_sum(s);
    if (lane == 0) cam[(size_t)n * HW + p] = fmaxf(s, 0.f);
  }
}

cudaError_t launch_cam(const void* a, int n, int hw, int c, const float* scale, const float* shift, const float* wfc,
                       int num_classes, const int32_t* target_cls, const float* logits, float* cam_lowres, int dtype,
                       cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (target_cls == nullptr && logits == nullptr) return cudaErrorInvalidValue;
  const int cg = c >> 3;
  if (cg < 1) return cudaErrorInvalidValue;
  int q = 1024 / cg;
  q = q < 1 ? 1 : (q > 8 ? 8 : q);
  if (q > hw) q = hw > 0 ? hw : 1;
  while (q > 1 && (size_t)(q + 1) * c * sizeof(float) > 48 * 1024) --q;
  const size_t smem = (size_t)(q + 1) * c * sizeof(float);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  int threads = (cg * q + 31) / 32 * 32;
  threads = threads > 1024 ? 1024 : (threads < 256 ? 256 : threads);
  if (dtype == kBF16)
    cam_kernel<bf16><<<n, threads, smem, s>>>(reinterpret_cast<const bf16*>(a), hw, c, q, scale, shift, wfc, num_classes,
                                              target_cls, logits, cam_lowres);
  else
    cam_kernel<float><<<n, threads, smem, s>>>(reinterpret_cast<const float*>(a), hw, c, q, scale, shift, wfc,
                                               num_classes, target_cls, logits, cam_lowres);
  count_launch();
  return cudaGetLastError();
}

// ---- CAM post-processing ----------------------------------------------------------------------------
__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < kTailThreads / 32; ++i) r = fmaxf(r, red[i]);
  __syncthreads();
  return r;
}
__device__ __forceinline__ float block_reduce_min(float v, float* red) {
  v = warp_min(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < kTailThreads / 32; ++i) r = fminf(r, red[i]);
  __syncthreads();
  return r;
}

__device__ __forceinline__ float bilerp(const float* src, int side, int oy, int ox, float rscale) {
  // cv2.resize / F.interpolate(align_corners=False): half-pixel centres, clamped
  float sy = fmaxf(((float)oy + 0.5f) * rscale - 0.5f, 0.f);
  float sx = fmaxf(((float)ox + 0.5f) * rscale - 0.5f, 0.f);
  int y0 = min((int)sy, side - 1), x0 = min((int)sx, side - 1);
  int y1 = min(y0 + 1, side - 1), x1 = min(x0 + 1, side - 1);
  float fy = sy - (float)y0, fx = sx - (float)x0;
  float v00 = src[y0 * side + x0], v01 = src[y0 * side + x1];
  float v10 = src[y1 * side + x0], v11 = src[y1 * side + x1];
  float top = v00 + (v01 - v00) * fx, bot = v10 + (v11 - v10) * fx;
  return top + (bot - top) * fy;
}

// One CTA (1024 threads) per image.  The bilinear sample positions are separable: per-axis (i0, i1, frac) tables
// in shared memory, a warp per output row, lanes along the row (coalesced stores, no divisions).  Same arithmetic
// as bilerp() above, evaluated twice per pixel (min/max pass, write pass) instead of three times.
constexpr int kUpThreads = 1024;
__global__ void __launch_bounds__(kUpThreads)
cam_upsample_kernel(const float* __restrict__ cam_low, int side, int out_side, float* __restrict__ cam) {
  extern __shared__ float sm[];  // [side*side] + 64 reduction scratch + 3 tables of out_side
  float* src = sm;
  float* red = sm + side * side;
  int* t0 = reinterpret_cast<int*>(red + 64);
  int* t1 = t0 + out_side;
  float* tf = reinterpret_cast<float*>(t1 + out_side);
  const int n = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = kUpThreads / 32;
  const int S2 = side * side;
  auto reduce2 = [&](float& lo, float& hi) {  // block-wide min and max
    lo = warp_min(lo);
    hi = warp_max(hi);
    __syncthreads();
    if (lane == 0) { red[warp] = lo; red[32 + warp] = hi; }
    __syncthreads();
    lo = red[0]; hi = red[32];
    for (int i = 1; i < nw; ++i) { lo = fminf(lo, red[i]); hi = fmaxf(hi, red[32 + i]); }
  };
  float lo = INFINITY, hi = -INFINITY;
  for (int i = tid; i < S2; i += kUpThreads) {
    const float v = cam_low[(size_t)n * S2 + i];
    src[i] = v;
    lo = fminf(lo, v);
    hi = fmaxf(hi, v);
  }
  const float rscale = (float)side / (float)out_side;
  for (int o = tid; o < out_side; o += kUpThreads) {
    const float sp = fmaxf(((float)o + 0.5f) * rscale - 0.5f, 0.f);
    const int i0 = min((int)sp, side - 1);
    t0[o] = i0;
    t1[o] = min(i0 + 1, side - 1);
    tf[o] = sp - (float)i0;
  }
  reduce2(lo, hi);
  hi = hi - lo;  // == max over (v - lo): rounding is monotonic
  const float inv = 1.0f / (hi + 1e-7f);
  __syncthreads();
  for (int i = tid; i < S2; i += kUpThreads) src[i] = (src[i] - lo) * inv;
  __syncthreads();
  auto sample = [&](int oy, int ox) {
    const int y0 = t0[oy], y1 = t1[oy], x0 = t0[ox], x1 = t1[ox];
    const float fy = tf[oy], fx = tf[ox];
    const float v00 = src[y0 * side + x0], v01 = src[y0 * side + x1];
    const float v10 = src[y1 * side + x0], v11 = src[y1 * side + x1];
    const float top = v00 + (v01 - v00) * fx, bot = v10 + (v11 - v10) * fx;
    return fmaxf(top + (bot - top) * fy, 0.f);
  };
  float lo2 = INFINITY, hi2 = -INFINITY;
  for (int oy = warp; oy < out_side; oy += nw)
    for (int ox = lane; ox < out_side; ox += 32) {
      const float v = sample(oy, ox);
      lo2 = fminf(lo2, v);
      hi2 = fmaxf(hi2, v);
    }
  reduce2(lo2, hi2);
  hi2 = hi2 - lo2;
  const float inv2 = 1.0f / (hi2 + 1e-7f);
  float* on = cam + (size_t)n * out_side * out_side;
  for (int oy = warp; oy < out_side; oy += nw)
    for (int ox = lane; ox < out_side; ox += 32) on[oy * out_side + ox] = (sample(oy, ox) - lo2) * inv2;
}

cudaError_t launch_cam_upsample(const float* cam_lowres, int n, int side, int out_side, float* cam, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const size_t smem = (size_t)(side * side + 64 + 3 * out_side) * sizeof(float);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  cam_upsample_kernel<<<n, kUpThreads, smem, s>>>(cam_lowres, side, out_side, cam);
  count_launch();
  return cudaGetLastError();
}

// ---- split (hook-protocol) helpers ----------------------------------------------------------------
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ a, int HW, int C, float* __restrict__ out) {
  // tile transpose through smem: 32 (hw) x 32 (c)
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const T* an = a + (size_t)n * HW * C;
  float* on = out + (size_t)n * HW * C;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    int p = p0 + i, c = c0 + threadIdx.x;
    if (p < HW && c < C) {
      if constexpr (is_bf16<T>::value) tile[i][threadIdx.x] = __bfloat162float(an[(size_t)p * C + c]);
      else tile[i][threadIdx.x] = an[(size_t)p * C + c];
    }
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    int c = c0 + i, p = p0 + threadIdx.x;
    if (p < HW && c < C) on[(size_t)c * HW + p] = tile[threadIdx.x][i];
  }
}

cudaError_t launch_nhwc_to_nchw_f32(const void* a, int n, int hw, int c, float* out, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  dim3 grid((hw + 31) / 32, (c + 31) / 32, n), block(32, 8);
  if (dtype == kBF16) nhwc_to_nchw_kernel<bf16><<<grid, block, 0, s>>>(reinterpret_cast<const bf16*>(a), hw, c, out);
  else nhwc_to_nchw_kernel<float><<<grid, block, 0, s>>>(reinterpret_cast<const float*>(a), hw, c, out);
  count_launch();
  return cudaGetLastError();
}

// pooled[n,c] = mean_hw silu(a_nchw*scale+shift): one warp per (n,c) row of HW contiguous floats
__global__ void __launch_bounds__(kTailThreads)
head_from_nchw_kernel(const float* __restrict__ a, int total_rows, int HW, int C, const float* __restrict__ scale,
                      const float* __restrict__ shift, float* __restrict__ pooled) {
  const int row = blockIdx.x * (kTailThreads / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= total_rows) return;
  const int c = row % C;
  const float sc = scale[c], sh = shift[c];
  float s = 0.f;
  for (int p = lane; p < HW; p += 32) s += silu_acc(fmaf(a[(size_t)row * HW + p], sc, sh));
  s = warp_sum(s);
  if (lane == 0) pooled[row] = s / (float)HW;
}

cudaError_t launch_head_from_nchw(const float* a, int n, int hw, int c, const float* scale, const float* shift,
                                  float* pooled, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  int rows = n * c;
  head_from_nchw_kernel<<<(rows + 7) / 8, kTailThreads, 0, s>>>(a, rows, hw, c, scale, shift, pooled);
  count_launch();
  return cudaGetLastError();
}

// grad_a[n,k,p] = (sum_j grad_logits[n,j] Wfc[j,k]) / HW * scale_k * silu'(a*scale_k + shift_k)
__global__ void __launch_bounds__(kTailThreads)
head_backward_kernel(const float* __restrict__ a, const float* __restrict__ gl, int total_rows, int HW, int C,
                     const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ wfc,
                     int NC, float* __restrict__ ga) {
  const int row = blockIdx.x * (kTailThreads / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= total_rows) return;
  const int c = row % C, n = row / C;
  float g = 0.f;
  for (int j = 0; j < NC; ++j) g = fmaf(gl[(size_t)n * NC + j], wfc[(size_t)j * C + c], g);
  const float sc = scale[c], sh = shift[c];
  g = g / (float)HW * sc;
  for (int p = lane; p < HW; p += 32) {
    float z = fmaf(a[(size_t)row * HW + p], sc, sh);
    float sg = sigmoid_acc(z);
    ga[(size_t)row * HW + p] = g * sg * fmaf(z, 1.0f - sg, 1.0f);
  }
}

cudaError_t launch_head_backward(const float* a, const float* grad_logits, int n, int hw, int c, const float* scale,
                                 const float* shift, const float* wfc, int num_classes, float* grad_a,
                                 cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  int rows = n * c;
  head_backward_kernel<<<(rows + 7) / 8, kTailThreads, 0, s>>>(a, grad_logits, rows, hw, c, scale, shift, wfc,
                                                                num_classes, grad_a);
  count_launch();
  return cudaGetLastError();
}

// ---- softmax + argmax packed records for the all-gather ------------------------------------------------
__global__ void softmax_pack_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, int n,
                                    int NC, float* __restrict__ packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  griddep_wait();
  griddep_launch_dependents();
  if (i >= n) return;
  const float* l = logits + (size_t)i * NC;
  float mx = l[0];
  int am = 0;
  for (int j = 1; j < NC; ++j)
    if (l[j] > mx) { mx = l[j]; am = j; }
  float den = 0.f;
  for (int j = 0; j < NC; ++j) den += expf(l[j] - mx);
  float* o = packed + (size_t)i * (NC + 2);
  for (int j = 0; j < NC; ++j) o[j] = expf(l[j] - mx) / den;
  o[NC] = (float)am;
  o[NC + 1] = labels ? (float)labels[i] : -1.0f;
}

cudaError_t launch_softmax_pack(const float* logits, const int64_t* labels, int n, int num_classes, float* packed,
                                cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  return launch_pdl(softmax_pack_kernel, dim3((n + 127) / 128), dim3(128), 0, s, logits, labels, n, num_classes, packed);
}

}  // namespace b200df
