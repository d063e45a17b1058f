// Backward-data kernels for Grad-CAM on FasterViT.
//
// The reference's target rule (/root/reference/web_ui.py:94-112: the last nn.Conv2d in modules() order) picks
// `levels.2.global_tokenizer.pos_embed` for fastervit: the depthwise 3x3 whose pooled output becomes the carrier tokens.
// Between that activation and the logits lie every HAT block of level 2, the level-2 downsample, level 3 and the head, so
// pytorch-grad-cam's `loss.backward()` (web_ui.py:275-279) is a real backward pass through windowed attention, LayerNorm and
// the GELU MLPs.  Only gradients with respect to ACTIVATIONS are needed (no weight gradients), so the GEMMs of the
// backward pass are the forward tcgen05 / CUDA-core GEMM kernels on transposed packed weights (nets.inc); this file holds
// the rest, any engine dtype (bf16 | fp32) for tensors, fp32 math:
//   layernorm_bwd      : dx of nn.LayerNorm given x, gamma, dy (+ the residual branch's gradient)
//   gelu_bwd           : dpre = dh * gelu'(pre), erf form
//   attention_bwd      : dq, dk, dv of softmax(q k^T * scale + bias) v, one CTA per (sequence, head)
//   conv3x3_s2_bwd     : backward-data of the level-2 downsample conv (3x3, stride 2, pad 1)
//   move_rows          : backward of the carrier-token cat / split copies (the destination rows were overwritten)
//   tail_bwd           : d logits[target] / d (level-3 output) of BatchNorm2d -> mean -> Linear (general d loss / d logits
//                        for the hook protocol: tail_bwd_general); ct_pool_bwd: the AvgPool2d's backward as fp32 NCHW
//   cam_from_ct_grad   : channel weights = spatial mean of dA (through the AvgPool2d in closed form), cam = relu(sum_k w_k A_k)
#include "common.cuh"
#include "kernels.h"

namespace b200df {

namespace {

// ---- LayerNorm backward ------------------------------------------------------------------------------------------
// y = (x - mean) * rstd * gamma + beta  ->  dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma
// out = res + dx (res nullable).  One warp per row, three passes over a row that stays in L1.
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const T* __restrict__ dy,
                     const T* __restrict__ res, T* __restrict__ out, long long rows, int C, float eps) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const T* xr = x + row * C;
  const T* dr = dy + row * C;
  float sum = 0.f;
  for (int c = lane; c < C; c += 32) sum += to_f(xr[c]);
  const float mean = warp_sum(sum) / (float)C;
  float sq = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float d = to_f(xr[c]) - mean;
    sq = fmaf(d, d, sq);
  }
  const float rstd = 1.0f / sqrtf(warp_sum(sq) / (float)C + eps);
  float sg = 0.f, sgx = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float g = to_f(dr[c]) * __ldg(gamma + c);
    sg += g;
    sgx = fmaf(g, (to_f(xr[c]) - mean) * rstd, sgx);
  }
  sg = warp_sum(sg) / (float)C;
  sgx = warp_sum(sgx) / (float)C;
  for (int c = lane; c < C; c += 32) {
    const float g = to_f(dr[c]) * __ldg(gamma + c);
    const float xh = (to_f(xr[c]) - mean) * rstd;
    float d = rstd * (g - sg - xh * sgx);
    if (res) d += to_f(res[row * C + c]);
    out[row * C + c] = from_f<T>(d);
  }
}

// ---- GELU backward (erf form: Phi(x) + x * phi(x)) ---------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
gelu_bwd_kernel(const T* __restrict__ pre, T* __restrict__ g, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const float x = to_f(pre[i]);
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  const float pdf = 0.3989422804014327f * expf(-0.5f * x * x);
  g[i] = from_f<T>(to_f(g[i]) * fmaf(x, pdf, cdf));
}

// ---- move rows: dst[seq, dst_off + r, :] = src[seq, src_off + r, :]; src rows <- 0 -----------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
move_rows_kernel(T* __restrict__ src, T* __restrict__ dst, long long nseq, int src_rows, int src_off, int dst_rows,
                 int dst_off, int nrows, int C) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = nseq * nrows * C;
  if (i >= total) return;
  const int c = (int)(i % C);
  const long long t = i / C;
  const int r = (int)(t % nrows);
  const long long sq = t / nrows;
  T* sp = src + (sq * src_rows + src_off + r) * C + c;
  dst[(sq * dst_rows + dst_off + r) * C + c] = *sp;
  *sp = from_f<T>(0.f);
}

// ---- head backward: logits = W (n_scale * mean_p(A) + n_shift) + b  ->  dA[p, c] = W[cls, c] * n_scale[c] / HW ------------------------
template <typename T>
__global__ void __launch_bounds__(256)
tail_bwd_kernel(const float* __restrict__ n_scale, const float* __restrict__ w, int HW, int C, int NC,
                const int32_t* __restrict__ target_cls, const float* __restrict__ logits, T* __restrict__ g) {
  const int img = blockIdx.x;
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
  T* gn = g + (size_t)img * HW * C;
  const float inv = 1.0f / (float)HW;
  for (int i = threadIdx.x; i < HW * C; i += blockDim.x) {
    const int c = i % C;
    gn[i] = from_f<T>(ok ? __ldg(w + (size_t)cls * C + c) * __ldg(n_scale + c) * inv : __int_as_float(0x7fc00000));
  }
}

// General form for the hook protocol (pytorch-grad-cam's loss.backward() with any d loss / d logits):
// dA[p, c] = (sum_j gl[n, j] W[j, c]) * n_scale[c] / HW
template <typename T>
__global__ void __launch_bounds__(256)
tail_bwd_general_kernel(const float* __restrict__ n_scale, const float* __restrict__ w, int HW, int C, int NC,
                        const float* __restrict__ gl, T* __restrict__ g) {
  extern __shared__ float wsum[];   // [C]
  const int img = blockIdx.x;
  const float inv = 1.0f / (float)HW;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int j = 0; j < NC; ++j) s = fmaf(gl[(size_t)img * NC + j], __ldg(w + (size_t)j * C + c), s);
    wsum[c] = s * __ldg(n_scale + c) * inv;
  }
  __syncthreads();
  T* gn = g + (size_t)img * HW * C;
  for (int i = threadIdx.x; i < HW * C; i += blockDim.x) gn[i] = from_f<T>(wsum[i % C]);
}

// backward of the carrier-token initialiser's AvgPool2d(kernel, stride) (vit_ops.cu: ct_pool, windowed token order):
// dA[n, c, y, x] = sum over pooled positions (py, px) whose window holds (y, x) of dct[n, tok(py, px), c] / kernel^2,
// written as fp32 NCHW -- the tensor pytorch-grad-cam's gradient hook on the target layer receives.
template <typename T>
__global__ void __launch_bounds__(256)
ct_pool_bwd_kernel(const T* __restrict__ gct, float* __restrict__ ga, int side, int C, int kernel, int stride, int g, int cs) {
  const int img = blockIdx.y, nwin = g / cs;
  const float inv = 1.0f / (float)(kernel * kernel);
  const int total = C * side * side;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int x = i % side, y = (i / side) % side, c = i / (side * side);
    float s = 0.f;
    for (int py = 0; py < g; ++py) {
      if (y < py * stride || y >= py * stride + kernel) continue;
      for (int px = 0; px < g; ++px) {
        if (x < px * stride || x >= px * stride + kernel) continue;
        const int wy = py / cs, iy = py - wy * cs, wx = px / cs, ix = px - wx * cs;
        const int tok = ((wy * nwin + wx) * cs + iy) * cs + ix;
        s += to_f(gct[((size_t)img * g * g + tok) * C + c]);
      }
    }
    ga[(size_t)img * total + i] = s * inv;
  }
}

// ---- Grad-CAM from the carrier-token gradient ----------------------------------------------------------------------------
// A = pos_embed output [HW, C]; ct = AvgPool2d(kernel, stride)(A) in windowed order.  Every pooled token averages
// kernel^2 positions with weight 1 / kernel^2, so sum_p dA[p, c] = sum_t dct[t, c]: pytorch-grad-cam's channel weights
// (the spatial MEAN of dA) are  w_c = sum_t dct[t, c] / HW  without materialising dA.  cam[p] = relu(sum_c w_c A[p, c]).
template <typename T>
__global__ void __launch_bounds__(256)
cam_from_ct_grad_kernel(const T* __restrict__ a, const T* __restrict__ gct, int HW, int C, int NT, float* __restrict__ cam) {
  extern __shared__ float wch[];   // [C]
  const int img = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* gn = gct + (size_t)img * NT * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int t = 0; t < NT; ++t) s += to_f(gn[(size_t)t * C + c]);
    wch[c] = s / (float)HW;
  }
  __syncthreads();
  const T* an = a + (size_t)img * HW * C;
  for (int p = warp; p < HW; p += blockDim.x >> 5) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(wch[c], to_f(an[(size_t)p * C + c]), s);
    s = warp_sum(s);
    if (lane == 0) cam[(size_t)img * HW + p] = (s != s) ? s : fmaxf(s, 0.f);   // a NaN gradient (invalid class) stays visible
  }
}

// ---- 3x3 stride-2 pad-1 conv, backward-data ---------------------------------------------------------------------------------
// y[oy, ox, co] = sum_{ky,kx,ci} x[2 oy + ky - 1, 2 ox + kx - 1, ci] w[co, ky*3+kx, ci]
//   => dx[iy, ix, ci] = sum over taps with (iy + 1 - ky) and (ix + 1 - kx) even and in range, sum_co dy[oy, ox, co] w[co, tap, ci].
// One CTA per input pixel and group of kImgs images (the weight slab of a tap is read once per group); threads over ci.
constexpr int kCbImgs = 8;
template <typename T>
__global__ void __launch_bounds__(256)
conv3x3_s2_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ w, T* __restrict__ dx, int n, int hin, int cin, int cinp,
                      int cout) {
  extern __shared__ float sdy[];   // [kCbImgs][cout]
  const int ho = (hin - 1) / 2 + 1;
  const int ix = blockIdx.x % hin, iy = blockIdx.x / hin;
  const int img0 = blockIdx.y * kCbImgs, ni = min(kCbImgs, n - img0);
  float acc[kCbImgs][2];
#pragma unroll
  for (int i = 0; i < kCbImgs; ++i) acc[i][0] = acc[i][1] = 0.f;
  for (int ky = 0; ky < 3; ++ky) {
    const int ty = iy + 1 - ky;
    if (ty < 0 || (ty & 1) || (ty >> 1) >= ho) continue;
    for (int kx = 0; kx < 3; ++kx) {
      const int tx = ix + 1 - kx;
      if (tx < 0 || (tx & 1) || (tx >> 1) >= ho) continue;
      const int oy = ty >> 1, ox = tx >> 1, tap = ky * 3 + kx;
      __syncthreads();
      for (int i = threadIdx.x; i < ni * cout; i += blockDim.x) {
        const int im = i / cout, co = i - im * cout;
        sdy[im * cout + co] = to_f(dy[(((size_t)(img0 + im) * ho + oy) * ho + ox) * cout + co]);
      }
      __syncthreads();
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int ci = threadIdx.x + r * 256;
        if (ci >= cin) continue;
        for (int co = 0; co < cout; ++co) {
          const float wv = to_f(w[((size_t)co * 9 + tap) * cinp + ci]);
#pragma unroll
          for (int im = 0; im < kCbImgs; ++im) acc[im][r] = fmaf(sdy[im * cout + co], wv, acc[im][r]);
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int ci = threadIdx.x + r * 256;
    if (ci >= cin) continue;
#pragma unroll
    for (int im = 0; im < kCbImgs; ++im)
      if (im < ni) dx[(((size_t)(img0 + im) * hin + iy) * hin + ix) * cin + ci] = from_f<T>(acc[im][r]);
  }
}

// ---- attention backward -------------------------------------------------------------------------------------------------------
// Forward (attn.cu / attn_mma.cu, no talking heads): S = q k^T * scale + bias, P = softmax_j(S), o = P v.
// Given do:  dv = P^T do;  dP = do v^T;  dS = P * (dP - rowsum(dP * P));  dq = scale * dS k;  dk = scale * dS^T q.
// One CTA per (sequence, head); q, k, v, do of the head and the two score matrices live in shared memory in fp32.
template <typename T>
__global__ void __launch_bounds__(256)
attention_bwd_kernel(const AttnArgs a, const T* __restrict__ dout, T* __restrict__ dq, T* __restrict__ dk, T* __restrict__ dv) {
  extern __shared__ float sm[];
  const int seq = blockIdx.x, h = blockIdx.y;
  const int nq = a.nq, nk = a.nk, dkk = a.dk, dvv = a.dv, H = a.heads;
  const int sp = nk + 1, qp = dkk + 1, vp = dvv + 1;
  float* qs = sm;                  // [nq][dk + 1]
  float* ks = qs + nq * qp;        // [nk][dk + 1]
  float* vs = ks + nk * qp;        // [nk][dv + 1]
  float* os = vs + nk * vp;        // [nq][dv + 1]   do
  float* P = os + nq * vp;         // [nq][nk + 1]
  float* D = P + nq * sp;          // [nq][nk + 1]   dP, then dS
  const T* q = reinterpret_cast<const T*>(a.q);
  const T* k = reinterpret_cast<const T*>(a.k);
  const T* v = reinterpret_cast<const T*>(a.v);
  const int tid = threadIdx.x, nt = blockDim.x;
  const int oc = H * dvv;
  for (int i = tid; i < nq * dkk; i += nt) {
    const int r = i / dkk, d = i - r * dkk;
    qs[r * qp + d] = to_f(q[((size_t)seq * nq + r) * a.q_pitch + a.q_off + h * dkk + d]);
  }
  for (int i = tid; i < nk * dkk; i += nt) {
    const int r = i / dkk, d = i - r * dkk;
    ks[r * qp + d] = to_f(k[((size_t)seq * nk + r) * a.k_pitch + a.k_off + h * dkk + d]);
  }
  for (int i = tid; i < nk * dvv; i += nt) {
    const int r = i / dvv, d = i - r * dvv;
    vs[r * vp + d] = to_f(v[((size_t)seq * nk + r) * a.v_pitch + a.v_off + h * dvv + d]);
  }
  for (int i = tid; i < nq * dvv; i += nt) {
    const int r = i / dvv, d = i - r * dvv;
    os[r * vp + d] = to_f(dout[((size_t)seq * nq + r) * oc + h * dvv + d]);
  }
  __syncthreads();
  for (int idx = tid; idx < nq * nk; idx += nt) {
    const int i = idx / nk, j = idx - i * nk;
    float s = 0.f, dp = 0.f;
    for (int e = 0; e < dkk; ++e) s = fmaf(qs[i * qp + e], ks[j * qp + e], s);
    for (int e = 0; e < dvv; ++e) dp = fmaf(os[i * vp + e], vs[j * vp + e], dp);
    s *= a.scale;
    if (a.bias) s += __ldg(a.bias + ((size_t)h * nq + i) * nk + j);
    P[i * sp + j] = s;
    D[i * sp + j] = dp;
  }
  __syncthreads();
  {   // softmax rows, then dS = P * (dP - sum_j dP P): one warp per row
    const int lane = tid & 31, warp = tid >> 5;
    for (int i = warp; i < nq; i += nt >> 5) {
      float* pr = P + i * sp;
      float* dr = D + i * sp;
      float mx = -INFINITY;
      for (int j = lane; j < nk; j += 32) mx = fmaxf(mx, pr[j]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int j = lane; j < nk; j += 32) {
        const float e = expf(pr[j] - mx);
        pr[j] = e;
        sum += e;
      }
      const float inv = 1.0f / warp_sum(sum);
      float dot = 0.f;
      for (int j = lane; j < nk; j += 32) {
        pr[j] *= inv;
        dot = fmaf(dr[j], pr[j], dot);
      }
      dot = warp_sum(dot);
      for (int j = lane; j < nk; j += 32) dr[j] = pr[j] * (dr[j] - dot);
    }
  }
  __syncthreads();
  for (int idx = tid; idx < nk * dvv; idx += nt) {   // dv[j, d] = sum_i P[i, j] do[i, d]
    const int j = idx / dvv, d = idx - j * dvv;
    float s = 0.f;
    for (int i = 0; i < nq; ++i) s = fmaf(P[i * sp + j], os[i * vp + d], s);
    dv[((size_t)seq * nk + j) * a.v_pitch + a.v_off + h * dvv + d] = from_f<T>(s);
  }
  for (int idx = tid; idx < nq * dkk; idx += nt) {   // dq[i, d] = scale * sum_j dS[i, j] k[j, d]
    const int i = idx / dkk, d = idx - i * dkk;
    float s = 0.f;
    for (int j = 0; j < nk; ++j) s = fmaf(D[i * sp + j], ks[j * qp + d], s);
    dq[((size_t)seq * nq + i) * a.q_pitch + a.q_off + h * dkk + d] = from_f<T>(s * a.scale);
  }
  for (int idx = tid; idx < nk * dkk; idx += nt) {   // dk[j, d] = scale * sum_i dS[i, j] q[i, d]
    const int j = idx / dkk, d = idx - j * dkk;
    float s = 0.f;
    for (int i = 0; i < nq; ++i) s = fmaf(D[i * sp + j], qs[i * qp + d], s);
    dk[((size_t)seq * nk + j) * a.k_pitch + a.k_off + h * dkk + d] = from_f<T>(s * a.scale);
  }
}

}  // namespace

// ---- launchers (plain stream-ordered launches: none of these kernels reads its inputs before its predecessor is done) ------
template <typename T> static const T* cp(const void* p) { return reinterpret_cast<const T*>(p); }
template <typename T> static T* mp(void* p) { return reinterpret_cast<T*>(p); }

cudaError_t launch_layernorm_bwd(const void* x, const float* gamma, const void* dy, const void* res, void* out,
                                 long long rows, int c, float eps, int dtype, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  const unsigned grid = (unsigned)((rows + 7) / 8);
  count_launch();
  if (dtype == kBF16)
    layernorm_bwd_kernel<bf16><<<grid, 256, 0, s>>>(cp<bf16>(x), gamma, cp<bf16>(dy), cp<bf16>(res), mp<bf16>(out), rows, c, eps);
  else
    layernorm_bwd_kernel<float><<<grid, 256, 0, s>>>(cp<float>(x), gamma, cp<float>(dy), cp<float>(res), mp<float>(out), rows, c, eps);
  return cudaGetLastError();
}

cudaError_t launch_gelu_bwd(const void* pre, void* g, long long total, int dtype, cudaStream_t s) {
  if (total <= 0) return cudaSuccess;
  const unsigned grid = (unsigned)((total + 255) / 256);
  count_launch();
  if (dtype == kBF16) gelu_bwd_kernel<bf16><<<grid, 256, 0, s>>>(cp<bf16>(pre), mp<bf16>(g), total);
  else gelu_bwd_kernel<float><<<grid, 256, 0, s>>>(cp<float>(pre), mp<float>(g), total);
  return cudaGetLastError();
}

cudaError_t launch_move_rows(void* src, void* dst, long long nseq, int src_rows, int src_off, int dst_rows, int dst_off,
                             int nrows, int c, int dtype, cudaStream_t s) {
  const long long total = nseq * nrows * c;
  if (total <= 0) return cudaSuccess;
  const unsigned grid = (unsigned)((total + 255) / 256);
  count_launch();
  if (dtype == kBF16)
    move_rows_kernel<bf16><<<grid, 256, 0, s>>>(mp<bf16>(src), mp<bf16>(dst), nseq, src_rows, src_off, dst_rows, dst_off, nrows, c);
  else
    move_rows_kernel<float><<<grid, 256, 0, s>>>(mp<float>(src), mp<float>(dst), nseq, src_rows, src_off, dst_rows, dst_off, nrows, c);
  return cudaGetLastError();
}

cudaError_t launch_tail_bwd(const float* n_scale, const float* w, int n, int hw, int c, int num_classes,
                            const int32_t* target_cls, const float* logits, void* g, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  if (dtype == kBF16) tail_bwd_kernel<bf16><<<n, 256, 0, s>>>(n_scale, w, hw, c, num_classes, target_cls, logits, mp<bf16>(g));
  else tail_bwd_kernel<float><<<n, 256, 0, s>>>(n_scale, w, hw, c, num_classes, target_cls, logits, mp<float>(g));
  return cudaGetLastError();
}

cudaError_t launch_tail_bwd_general(const float* n_scale, const float* w, int n, int hw, int c, int num_classes,
                                    const float* grad_logits, void* g, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  count_launch();
  if (dtype == kBF16)
    tail_bwd_general_kernel<bf16><<<n, 256, c * sizeof(float), s>>>(n_scale, w, hw, c, num_classes, grad_logits, mp<bf16>(g));
  else
    tail_bwd_general_kernel<float><<<n, 256, c * sizeof(float), s>>>(n_scale, w, hw, c, num_classes, grad_logits, mp<float>(g));
  return cudaGetLastError();
}

cudaError_t launch_ct_pool_bwd(const void* gct, float* ga_nchw, int n, int side, int c, int kernel, int stride, int cs,
                               int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int g = (side - kernel) / stride + 1;
  if (g % cs) return cudaErrorInvalidValue;
  const dim3 grid((c * side * side + 255) / 256, n);
  count_launch();
  if (dtype == kBF16) ct_pool_bwd_kernel<bf16><<<grid, 256, 0, s>>>(cp<bf16>(gct), ga_nchw, side, c, kernel, stride, g, cs);
  else ct_pool_bwd_kernel<float><<<grid, 256, 0, s>>>(cp<float>(gct), ga_nchw, side, c, kernel, stride, g, cs);
  return cudaGetLastError();
}

cudaError_t launch_cam_from_ct_grad(const void* a, const void* gct, int n, int hw, int c, int ntok, float* cam_lowres,
                                    int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  count_launch();
  if (dtype == kBF16)
    cam_from_ct_grad_kernel<bf16><<<n, 256, c * sizeof(float), s>>>(cp<bf16>(a), cp<bf16>(gct), hw, c, ntok, cam_lowres);
  else
    cam_from_ct_grad_kernel<float><<<n, 256, c * sizeof(float), s>>>(cp<float>(a), cp<float>(gct), hw, c, ntok, cam_lowres);
  return cudaGetLastError();
}

cudaError_t launch_conv3x3_s2_bwd(const void* dy, const void* w, void* dx, int n, int hin, int cin, int cinp, int cout,
                                  int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (cin > 512 || (size_t)kCbImgs * cout * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  const dim3 grid(hin * hin, (n + kCbImgs - 1) / kCbImgs);
  const size_t smem = (size_t)kCbImgs * cout * sizeof(float);
  count_launch();
  if (dtype == kBF16)
    conv3x3_s2_bwd_kernel<bf16><<<grid, 256, smem, s>>>(cp<bf16>(dy), cp<bf16>(w), mp<bf16>(dx), n, hin, cin, cinp, cout);
  else
    conv3x3_s2_bwd_kernel<float><<<grid, 256, smem, s>>>(cp<float>(dy), cp<float>(w), mp<float>(dx), n, hin, cin, cinp, cout);
  return cudaGetLastError();
}

// dq / dk / dv are written with q's / k's / v's pitch and column offset (the fused q|k|v layout of the forward GEMM).
cudaError_t launch_attention_bwd(const AttnArgs& a, const void* dout, void* dq, void* dk, void* dv, int dtype, cudaStream_t s) {
  if (a.nseq <= 0) return cudaSuccess;
  if (a.th1_w || a.vlocal || a.act != kActNone) return cudaErrorNotSupported;
  const size_t smem = ((size_t)(a.nq + a.nk) * (a.dk + 1) + (size_t)(a.nq + a.nk) * (a.dv + 1) + 2 * (size_t)a.nq * (a.nk + 1)) *
                      sizeof(float);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  static PerDeviceOnce once_b, once_f;
  count_launch();
  const dim3 grid(a.nseq, a.heads);
  if (dtype == kBF16) {
    if (once_b.need()) {
      cudaError_t e = cudaFuncSetAttribute(attention_bwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return e;
    }
    attention_bwd_kernel<bf16><<<grid, 256, smem, s>>>(a, cp<bf16>(dout), mp<bf16>(dq), mp<bf16>(dk), mp<bf16>(dv));
  } else {
    if (once_f.need()) {
      cudaError_t e = cudaFuncSetAttribute(attention_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return e;
    }
    attention_bwd_kernel<float><<<grid, 256, smem, s>>>(a, cp<float>(dout), mp<float>(dq), mp<float>(dk), mp<float>(dv));
  }
  return cudaGetLastError();
}

}  // namespace b200df
