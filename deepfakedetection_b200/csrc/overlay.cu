// Grad-CAM overlay on the device: `_tensor_to_rgb` + pytorch-grad-cam `show_cam_on_image(rgb, cam, use_rgb=True)`
// (/root/reference/web_ui.py:72-93 and :281-282), one CTA per image:
//   rgb  = clamp(x * std + mean, 0, 1)                       (only when the transform normalised the input)
//   heat = JET[uint8(255 * clip(cam, 0, 1))] / 255            (OpenCV COLORMAP_JET, RGB order)
//   o    = (1 - w) * heat + w * rgb ;  o /= max(o) ;  out = uint8(255 * o)
// Every product / sum / quotient is a separately rounded fp32 operation (no FMA contraction, IEEE division),
// like the numpy expressions it replaces, so the bytes are identical to the host path.  Two passes over the
// image: the global max, then the write (the blend is recomputed: 10 flops against 15 bytes per pixel).
#include "common.cuh"
#include "kernels.h"

namespace b200df {

__constant__ unsigned char kJetRGB[256 * 3] = {
#include "jet_lut.inc"
};

constexpr int kOvThreads = 1024;

__device__ __forceinline__ void overlay_pixel(const float* __restrict__ x, const float* __restrict__ cam, size_t plane,
                                              size_t i, bool denorm, const float (&mean)[3], const float (&sd)[3],
                                              float w_heat, float w_img, float (&o)[3]) {
  const float m = fminf(fmaxf(cam[i], 0.f), 1.f);
  const int idx = (int)__fmul_rn(255.f, m);  // np.uint8(255 * mask): truncation
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float v = x[c * plane + i];
    if (denorm) v = __fadd_rn(__fmul_rn(v, sd[c]), mean[c]);
    v = fminf(fmaxf(v, 0.f), 1.f);
    const float heat = __fdiv_rn((float)kJetRGB[idx * 3 + c], 255.f);
    o[c] = __fadd_rn(__fmul_rn(w_heat, heat), __fmul_rn(w_img, v));
  }
}

__global__ void __launch_bounds__(kOvThreads)
cam_overlay_kernel(const float* __restrict__ x_nchw, const float* __restrict__ cam, int hw, int denorm, float m0,
                   float m1, float m2, float s0, float s1, float s2, float w_img, float w_heat,
                   unsigned char* __restrict__ out) {
  __shared__ float red[32];
  const int n = blockIdx.x, tid = threadIdx.x;
  const float mean[3] = {m0, m1, m2}, sd[3] = {s0, s1, s2};
  const float* x = x_nchw + (size_t)n * 3 * hw;
  const float* cm = cam + (size_t)n * hw;
  float mx = 0.f;  // every blended value is >= 0
  for (int i = tid; i < hw; i += kOvThreads) {
    float o[3];
    overlay_pixel(x, cm, (size_t)hw, (size_t)i, denorm != 0, mean, sd, w_heat, w_img, o);
    mx = fmaxf(mx, fmaxf(o[0], fmaxf(o[1], o[2])));
  }
  mx = warp_max(mx);
  if ((tid & 31) == 0) red[tid >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < kOvThreads / 32; ++i) mx = fmaxf(mx, red[i]);
  unsigned char* on = out + (size_t)n * hw * 3;
  for (int i = tid; i < hw; i += kOvThreads) {
    float o[3];
    overlay_pixel(x, cm, (size_t)hw, (size_t)i, denorm != 0, mean, sd, w_heat, w_img, o);
#pragma unroll
    for (int c = 0; c < 3; ++c) on[(size_t)i * 3 + c] = (unsigned char)(int)__fmul_rn(255.f, __fdiv_rn(o[c], mx));
  }
}

cudaError_t launch_cam_overlay(const float* x_nchw, const float* cam, int n, int h, int w, int denorm, const float* mean,
                               const float* stdv, float image_weight, unsigned char* out, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  // numpy evaluates (1 - image_weight) in double and rounds the result to fp32 once
  const float w_heat = (float)(1.0 - (double)image_weight);
  cam_overlay_kernel<<<n, kOvThreads, 0, s>>>(x_nchw, cam, h * w, denorm, mean[0], mean[1], mean[2], stdv[0], stdv[1],
                                              stdv[2], image_weight, w_heat, out);
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
