// Small-sequence multi-head attention, one CTA per sequence (talking heads) or per (sequence, head):
//   S = (q k^T) * scale + bias;  [talking_head1: 1x1 mix across heads];  P = softmax_j(S);  [talking_head2];
//   out = P v (+ v_local) ; [GELU]
// Replaces, inside the reference's `model(images)` (/root/reference/orchestration/orchestrator.py:590):
//   timm EfficientFormerV2 `Attention2d` (49 tokens, 8 heads, d_k 32, d_v 128, learned bias table gathered at load
//   time, talking heads, + `v_local`) and `Attention2dDownsample` (49 queries x 196 keys, d_k 16, d_v 64)
//   -- SURVEY.md 8a rows F2/F3; fastervit `WindowAttention` (53 = 4 carrier + 49 window tokens, 16 carrier tokens,
//   or 49 tokens; head dim 32 / 48; relative-position bias precomputed at load time) -- SURVEY.md 8a rows V3/V4.
// Sequences are at most 196 keys long, so scores, probabilities and the head's q/k/v slices live in shared memory
// (fp32) and nothing but q, k, v is read and nothing but `out` is written: the op is bound by its 2-8 KB/token of
// I/O, not by the 0.6-5 M MAC per sequence.  q/k/v are addressed as row pitch + column offset, so the fused q|k or
// q|k|v GEMM outputs are consumed in place (no split / permute kernels).
#include "common.cuh"
#include "kernels.h"

#include <stdlib.h>

namespace b200df {

constexpr int kAttnThreads = 256;
constexpr int kAttnMaxHeads = 16;

// float offset of the V slice inside the CTA's shared memory (rounded up: its rows are read as float4)
__host__ __device__ inline int attn_vs_offset(int heads_in_smem, int nq, int nk, int dk) {
  const int f = heads_in_smem * nq * (nk + 1) + nq * dk + nk * (dk + 1);
  return (f + 3) & ~3;
}

template <typename T, bool TH>
__global__ void __launch_bounds__(kAttnThreads)
attn_kernel(const AttnArgs a) {
  extern __shared__ float sm[];
  constexpr bool kFast = is_bf16<T>::value;
  const int seq = blockIdx.x;
  const int H = a.heads, nq = a.nq, nk = a.nk, dk = a.dk, dv = a.dv;
  const int sp = nk + 1;
  const int dkp = dk + 1;
  float* S = sm;                                // TH: [H][nq][sp], else [nq][sp]
  float* qs = S + (TH ? H : 1) * nq * sp;       // [nq][dk]
  float* ks = qs + nq * dk;                     // [nk][dk + 1]
  float* vs = sm + attn_vs_offset(TH ? H : 1, nq, nk, dk);  // [nk][dv], 16-byte aligned
  const T* q = reinterpret_cast<const T*>(a.q);
  const T* k = reinterpret_cast<const T*>(a.k);
  const T* v = reinterpret_cast<const T*>(a.v);
  const T* vl = reinterpret_cast<const T*>(a.vlocal);
  T* out = reinterpret_cast<T*>(a.out);
  const int tid = threadIdx.x;
  const int h_begin = TH ? 0 : (int)blockIdx.y, h_end = TH ? H : (int)blockIdx.y + 1;
  const int dk8 = dk >> 3, dv8 = dv >> 3;
  griddep_wait();
  griddep_launch_dependents();

  // ---- scores ----
  for (int h = h_begin; h < h_end; ++h) {
    float* Sh = S + (TH ? h * nq * sp : 0);
    for (int idx = tid; idx < nq * dk8; idx += kAttnThreads) {
      const int i = idx / dk8, d8 = idx - i * dk8;
      float f[8];
      load8<T>(q + ((size_t)seq * nq + i) * a.q_pitch + a.q_off + h * dk + d8 * 8, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) qs[i * dk + d8 * 8 + e] = f[e];
    }
    for (int idx = tid; idx < nk * dk8; idx += kAttnThreads) {
      const int j = idx / dk8, d8 = idx - j * dk8;
      float f[8];
      load8<T>(k + ((size_t)seq * nk + j) * a.k_pitch + a.k_off + h * dk + d8 * 8, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ks[j * dkp + d8 * 8 + e] = f[e];
    }
    __syncthreads();
    for (int idx = tid; idx < nq * nk; idx += kAttnThreads) {
      const int i = idx / nk, j = idx - i * nk;
      const float* qi = qs + i * dk;
      const float* kj = ks + j * dkp;
      float d = 0.f;
      for (int e = 0; e < dk; ++e) d = fmaf(qi[e], kj[e], d);
      d *= a.scale;
      if (a.bias) d += __ldg(a.bias + ((size_t)h * nq + i) * nk + j);
      Sh[i * sp + j] = d;
    }
    __syncthreads();
  }

  // ---- talking head 1 (1x1 conv across the head axis, EfficientFormerV2) ----
  if constexpr (TH) {
    for (int idx = tid; idx < nq * nk; idx += kAttnThreads) {
      const int i = idx / nk, j = idx - i * nk;
      float t[kAttnMaxHeads];
      for (int h = 0; h < H; ++h) t[h] = S[(h * nq + i) * sp + j];
      for (int ho = 0; ho < H; ++ho) {
        float o = __ldg(a.th1_b + ho);
        for (int h = 0; h < H; ++h) o = fmaf(__ldg(a.th1_w + ho * H + h), t[h], o);
        S[(ho * nq + i) * sp + j] = o;
      }
    }
    __syncthreads();
  }

  // ---- softmax over keys: one warp per (head, query) row ----
  {
    const int rows = (TH ? H : 1) * nq, lane = tid & 31, warp = tid >> 5;
    for (int r = warp; r < rows; r += kAttnThreads / 32) {
      float* row = S + r * sp;
      float mx = -INFINITY;
      for (int j = lane; j < nk; j += 32) mx = fmaxf(mx, row[j]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int j = lane; j < nk; j += 32) {
        const float e = kFast ? __expf(row[j] - mx) : expf(row[j] - mx);
        row[j] = e;
        sum += e;
      }
      sum = warp_sum(sum);
      const float inv = 1.0f / sum;
      for (int j = lane; j < nk; j += 32) row[j] *= inv;
    }
    __syncthreads();
  }

  // ---- talking head 2 ----
  if constexpr (TH) {
    for (int idx = tid; idx < nq * nk; idx += kAttnThreads) {
      const int i = idx / nk, j = idx - i * nk;
      float t[kAttnMaxHeads];
      for (int h = 0; h < H; ++h) t[h] = S[(h * nq + i) * sp + j];
      for (int ho = 0; ho < H; ++ho) {
        float o = __ldg(a.th2_b + ho);
        for (int h = 0; h < H; ++h) o = fmaf(__ldg(a.th2_w + ho * H + h), t[h], o);
        S[(ho * nq + i) * sp + j] = o;
      }
    }
    __syncthreads();
  }

  // ---- out = P v (+ v_local) (GELU) ----
  const int oc = H * dv;
  for (int h = h_begin; h < h_end; ++h) {
    const float* Sh = S + (TH ? h * nq * sp : 0);
    for (int idx = tid; idx < nk * dv8; idx += kAttnThreads) {
      const int j = idx / dv8, d8 = idx - j * dv8;
      float f[8];
      load8<T>(v + ((size_t)seq * nk + j) * a.v_pitch + a.v_off + h * dv + d8 * 8, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) vs[j * dv + d8 * 8 + e] = f[e];
    }
    __syncthreads();
    for (int idx = tid; idx < nq * dv8; idx += kAttnThreads) {
      const int i = idx / dv8, d8 = idx - i * dv8;
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = 0.f;
      const float* pi = Sh + i * sp;
      for (int j = 0; j < nk; ++j) {
        const float p = pi[j];
        const float4 v0 = *reinterpret_cast<const float4*>(vs + j * dv + d8 * 8);
        const float4 v1 = *reinterpret_cast<const float4*>(vs + j * dv + d8 * 8 + 4);
        o[0] = fmaf(p, v0.x, o[0]); o[1] = fmaf(p, v0.y, o[1]); o[2] = fmaf(p, v0.z, o[2]); o[3] = fmaf(p, v0.w, o[3]);
        o[4] = fmaf(p, v1.x, o[4]); o[5] = fmaf(p, v1.y, o[5]); o[6] = fmaf(p, v1.z, o[6]); o[7] = fmaf(p, v1.w, o[7]);
      }
      const size_t off = ((size_t)seq * nq + i) * oc + h * dv + d8 * 8;
      if (vl) {
        float l[8];
        load8<T>(vl + off, l);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += l[e];
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = apply_act<kFast>(o[e], a.act);
      store8<T>(out + off, o);
    }
    __syncthreads();
  }
}

static size_t attn_smem(const AttnArgs& a, bool th) {
  return ((size_t)attn_vs_offset(th ? a.heads : 1, a.nq, a.nk, a.dk) + (size_t)a.nk * a.dv) * sizeof(float);
}

template <typename T, bool TH>
static cudaError_t attn_launch(const AttnArgs& a, cudaStream_t s) {
  const size_t smem = attn_smem(a, TH);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(attn_kernel<T, TH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
  }
  count_launch();
  return launch_pdl(attn_kernel<T, TH>, dim3(a.nseq, TH ? 1 : a.heads), dim3(kAttnThreads), smem, s, a);
}

cudaError_t launch_attention(const AttnArgs& a, int dtype, cudaStream_t s) {
  if (a.nseq <= 0) return cudaSuccess;
  if (dtype == kBF16) {   // tensor-core kernel (attn_mma.cu) for the shapes it is built for; B200DF_GEMM=simt / B200DF_ATTN=simt: this one
    const char* e1 = getenv("B200DF_ATTN");
    const char* e2 = getenv("B200DF_GEMM");
    if (!((e1 && e1[0] == 's') || (e2 && e2[0] == 's'))) {
      const cudaError_t e = launch_attention_mma(a, s);
      if (e != cudaErrorNotSupported) return e;
    }
  }
  const bool th = a.th1_w != nullptr;
  if (a.dk % 8 || a.dv % 8 || a.heads > kAttnMaxHeads || a.heads < 1 || a.nq < 1 || a.nk < 1) return cudaErrorInvalidValue;
  if (a.q_pitch % 8 || a.k_pitch % 8 || a.v_pitch % 8 || a.q_off % 8 || a.k_off % 8 || a.v_off % 8) return cudaErrorInvalidValue;
  if (th && (!a.th1_b || !a.th2_w || !a.th2_b)) return cudaErrorInvalidValue;
  if (dtype == kBF16) return th ? attn_launch<bf16, true>(a, s) : attn_launch<bf16, false>(a, s);
  return th ? attn_launch<float, true>(a, s) : attn_launch<float, false>(a, s);
}

}  // namespace b200df
