// Shared device helpers for the b200df kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "act.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "b200df kernels are written for sm_100a (Blackwell) only"
#endif

namespace b200df {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

typedef __nv_bfloat16 bf16;

// ---- activations --------------------------------------------------------------------------
// Accurate forms (fp32 strict path) and fast forms (bf16 path: one MUFU.TANH per element).
__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }
// Programmatic dependent launch (kernels.h: launch_pdl): block until every prerequisite grid has completed and
// its writes are visible; then let the next grid in the stream start placing its CTAs.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

__device__ __forceinline__ float silu_acc(float x) { return x / (1.0f + expf(-x)); }

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// silu(x) = x*sigmoid(x) = h + h*tanh(h), h = x/2
__device__ __forceinline__ float silu_fast(float x) {
  float h = 0.5f * x;
  return fmaf(h, tanh_fast(h), h);
}
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }

template <bool kFast>
__device__ __forceinline__ float silu(float x) {
  if constexpr (kFast) return silu_fast(x);
  else return silu_acc(x);
}

// GELU (nn.GELU default = exact erf form).  Accurate: erff.  Fast (bf16 path): the tanh form
//   gelu(x) = h + h*tanh(h*(1.5957691216 + 0.2854191951*h^2)), h = x/2
// (|tanh form - erf form| <= 4.8e-4 absolute, an order of magnitude below one bf16 ulp of the stored result).
__device__ __forceinline__ float gelu_acc(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_fast(float x) {
  const float h = 0.5f * x;
  const float u = h * fmaf(0.2854191951f, h * h, 1.5957691216f);
  return fmaf(h, tanh_fast(u), h);
}

template <bool kFast>
__device__ __forceinline__ float apply_act(float x, int act) {
  switch (act) {
    case kActSiLU: return silu<kFast>(x);
    case kActGELU: return kFast ? gelu_fast(x) : gelu_acc(x);
    case kActReLU: return fmaxf(x, 0.f);
    default: return x;
  }
}
template <int ACT, bool kFast>
__device__ __forceinline__ float act_fn(float x) {
  if constexpr (ACT == kActSiLU) return silu<kFast>(x);
  else if constexpr (ACT == kActGELU) return kFast ? gelu_fast(x) : gelu_acc(x);
  else if constexpr (ACT == kActReLU) return fmaxf(x, 0.f);
  else return x;
}

// ---- 8-channel vector access (NHWC, C % 8 == 0) ---------------------------------------------
// bf16: one 128-bit access; fp32: two 128-bit accesses.
__device__ __forceinline__ void unpack_bf16x8(const uint4& v, float (&f)[8]) {
  f[0] = __uint_as_float(v.x << 16);
  f[1] = __uint_as_float(v.x & 0xffff0000u);
  f[2] = __uint_as_float(v.y << 16);
  f[3] = __uint_as_float(v.y & 0xffff0000u);
  f[4] = __uint_as_float(v.z << 16);
  f[5] = __uint_as_float(v.z & 0xffff0000u);
  f[6] = __uint_as_float(v.w << 16);
  f[7] = __uint_as_float(v.w & 0xffff0000u);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);  // .x = lo (low 16 bits)
  return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ uint4 pack_bf16x8(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16x2(f[0], f[1]);
  v.y = pack_bf16x2(f[2], f[3]);
  v.z = pack_bf16x2(f[4], f[5]);
  v.w = pack_bf16x2(f[6], f[7]);
  return v;
}

template <typename T>
__device__ __forceinline__ void load8(const T* __restrict__ p, float (&f)[8]);

template <>
__device__ __forceinline__ void load8<bf16>(const bf16* __restrict__ p, float (&f)[8]) {
  uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  unpack_bf16x8(v, f);
}
template <>
__device__ __forceinline__ void load8<float>(const float* __restrict__ p, float (&f)[8]) {
  float4 a = __ldg(reinterpret_cast<const float4*>(p));
  float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

template <typename T>
__device__ __forceinline__ void store8(T* __restrict__ p, const float (&f)[8]);

template <>
__device__ __forceinline__ void store8<bf16>(bf16* __restrict__ p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = pack_bf16x8(f);
}
template <>
__device__ __forceinline__ void store8<float>(float* __restrict__ p, const float (&f)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}

template <typename T> struct is_bf16 { static constexpr bool value = false; };
template <> struct is_bf16<bf16> { static constexpr bool value = true; };

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// ---- deterministic pooled sums ----------------------------------------------------------------
// SE pool sums are accumulated across CTAs with 64-bit INTEGER atomics on a Q24 fixed-point value
// (resolution 2^-24, range +-2^39): integer addition is associative, so the result does not depend on
// the order in which tiles retire and a forward pass is bit-reproducible run to run.
typedef long long pool_t;
constexpr float kPoolScale = 16777216.0f;         // 2^24
constexpr float kPoolInvScale = 1.0f / 16777216.0f;
__device__ __forceinline__ void pool_add(pool_t* dst, float v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(dst), (unsigned long long)__float2ll_rn(v * kPoolScale));
}
__device__ __forceinline__ float pool_value(pool_t v) { return (float)((double)v * (double)kPoolInvScale); }

// ---- warp reductions ----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace b200df
