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
// out = res + dx (res nullable).  One warp per row.
// C % 8 == 0, C <= 1024: a lane keeps its (up to four) 16-byte vectors of x and dy in registers across the passes.
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const T* __restrict__ dy,
                     const T* __restrict__ res, T* __restrict__ out, long long rows, int C, float eps) {
  griddep_wait();
  griddep_launch_dependents();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nv = C >> 3;
  float xv[4][8], gv[4][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int vi = lane + 32 * i;
    if (vi < nv) {
      float d8[8], g8[8];
      load8<T>(x + row * C + vi * 8, xv[i]);
      load8<T>(dy + row * C + vi * 8, d8);
      load8<float>(gamma + vi * 8, g8);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        gv[i][e] = d8[e] * g8[e];
        sum += xv[i][e];
      }
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) xv[i][e] = gv[i][e] = 0.f;
    }
  }
  const float mean = warp_sum(sum) / (float)C;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (lane + 32 * i < nv) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = xv[i][e] - mean;
        sq = fmaf(d, d, sq);
      }
    }
  const float rstd = 1.0f / sqrtf(warp_sum(sq) / (float)C + eps);
  float sg = 0.f, sgx = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (lane + 32 * i < nv) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        xv[i][e] = (xv[i][e] - mean) * rstd;     // xhat
        sg += gv[i][e];
        sgx = fmaf(gv[i][e], xv[i][e], sgx);
      }
    }
  sg = warp_sum(sg) / (float)C;
  sgx = warp_sum(sgx) / (float)C;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int vi = lane + 32 * i;
    if (vi < nv) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = rstd * (gv[i][e] - sg - xv[i][e] * sgx);
      if (res) {
        float r8[8];
        load8<T>(res + row * C + vi * 8, r8);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += r8[e];
      }
      store8<T>(out + row * C + vi * 8, o);
    }
  }
}

// ---- GELU backward (erf form: Phi(x) + x * phi(x)), eight elements per thread -------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
gelu_bwd_kernel(const T* __restrict__ pre, T* __restrict__ g, long long total8) {
  griddep_wait();
  griddep_launch_dependents();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total8) return;
  float x[8], d[8];
  load8<T>(pre + i * 8, x);
  load8<T>(g + i * 8, d);
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float cdf = 0.5f * (1.0f + erff(x[e] * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * expf(-0.5f * x[e] * x[e]);
    d[e] *= fmaf(x[e], pdf, cdf);
  }
  store8<T>(g + i * 8, d);
}

// ---- move rows: dst[seq, dst_off + r, :] = src[seq, src_off + r, :]; src rows <- 0 -----------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
move_rows_kernel(T* __restrict__ src, T* __restrict__ dst, long long nseq, int src_rows, int src_off, int dst_rows,
                 int dst_off, int nrows, int C) {
  griddep_wait();
  griddep_launch_dependents();
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
  griddep_wait();
  griddep_launch_dependents();
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
  griddep_wait();
  griddep_launch_dependents();
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
  griddep_wait();
  griddep_launch_dependents();
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
  griddep_wait();
  griddep_launch_dependents();
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
  griddep_wait();
  griddep_launch_dependents();
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
// One CTA per (sequence, head); everything of the head lives in shared memory in fp32.  The five small products run on
// 4 x 4 register tiles per thread -- the first version (one output per thread, two shared-memory loads per FMA) was
// bound by the shared-memory pipe: 155 us per launch.  K and V are also kept TRANSPOSED ([d][key], keys padded to a
// multiple of 4 with zeros) so that four consecutive keys are one 16-byte load.
__host__ __device__ inline int abw_r4(int x) { return (x + 3) & ~3; }
// float offsets of the shared-memory arrays
struct AbwLayout {
  int qp, vp, kp, sp;            // pitches: q / k rows (dk + 4), do rows (dv + 4), transposed rows / score rows (keys4 + 4)
  int qs, ks, kT, vT, os, P, D, total;
  __host__ __device__ AbwLayout(int nq, int nk, int dk, int dv) {
    qp = dk + 4; vp = dv + 4; kp = abw_r4(nk) + 4; sp = kp;
    qs = 0;
    ks = qs + nq * qp;
    kT = ks + nk * qp;
    vT = kT + dk * kp;
    os = vT + dv * kp;
    P = os + nq * vp;
    D = P + nq * sp;
    total = D + nq * sp;
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
attention_bwd_kernel(const AttnArgs a, const T* __restrict__ dout, T* __restrict__ dq, T* __restrict__ dk, T* __restrict__ dv) {
  griddep_wait();
  griddep_launch_dependents();
  extern __shared__ __align__(16) float sm[];
  const int seq = blockIdx.x, h = blockIdx.y;
  const int nq = a.nq, nk = a.nk, dkk = a.dk, dvv = a.dv, H = a.heads;
  const AbwLayout L(nq, nk, dkk, dvv);
  float* qs = sm + L.qs;   // [nq][qp]
  float* ks = sm + L.ks;   // [nk][qp]
  float* kT = sm + L.kT;   // [dk][kp]   keys >= nk: 0
  float* vT = sm + L.vT;   // [dv][kp]
  float* os = sm + L.os;   // [nq][vp]   do
  float* P = sm + L.P;     // [nq][sp]
  float* D = sm + L.D;     // [nq][sp]   dP, then dS
  const T* q = reinterpret_cast<const T*>(a.q);
  const T* k = reinterpret_cast<const T*>(a.k);
  const T* v = reinterpret_cast<const T*>(a.v);
  const int tid = threadIdx.x, nt = blockDim.x;
  const int oc = H * dvv, nk4 = abw_r4(nk) >> 2, nq4 = abw_r4(nq) >> 2;
  for (int i = tid; i < (dkk + dvv) * L.kp; i += nt) kT[i] = 0.f;   // kT and vT are adjacent: zero incl. the key padding
  __syncthreads();
  for (int i = tid; i < nq * dkk; i += nt) {
    const int r = i / dkk, d = i - r * dkk;
    qs[r * L.qp + d] = to_f(q[((size_t)seq * nq + r) * a.q_pitch + a.q_off + h * dkk + d]);
  }
  for (int i = tid; i < nk * dkk; i += nt) {
    const int r = i / dkk, d = i - r * dkk;
    const float x = to_f(k[((size_t)seq * nk + r) * a.k_pitch + a.k_off + h * dkk + d]);
    ks[r * L.qp + d] = x;
    kT[d * L.kp + r] = x;
  }
  for (int i = tid; i < nk * dvv; i += nt) {
    const int r = i / dvv, d = i - r * dvv;
    vT[d * L.kp + r] = to_f(v[((size_t)seq * nk + r) * a.v_pitch + a.v_off + h * dvv + d]);
  }
  for (int i = tid; i < nq * dvv; i += nt) {
    const int r = i / dvv, d = i - r * dvv;
    os[r * L.vp + d] = to_f(dout[((size_t)seq * nq + r) * oc + h * dvv + d]);
  }
  __syncthreads();
  // S = scale q k^T + bias and dP = do v^T on 4 (queries) x 4 (keys) tiles
  for (int item = tid; item < nq4 * nk4; item += nt) {
    const int i0 = (item / nk4) * 4, j0 = (item - (item / nk4) * nk4) * 4;
    int ir[4];
#pragma unroll
    for (int ii = 0; ii < 4; ++ii) ir[ii] = min(i0 + ii, nq - 1);
    float sacc[4][4], pacc[4][4];
#pragma unroll
    for (int ii = 0; ii < 4; ++ii)
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) sacc[ii][jj] = pacc[ii][jj] = 0.f;
    for (int e = 0; e < dkk; ++e) {
      const float4 kk = *reinterpret_cast<const float4*>(kT + e * L.kp + j0);
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) {
        const float qv = qs[ir[ii] * L.qp + e];
        sacc[ii][0] = fmaf(qv, kk.x, sacc[ii][0]); sacc[ii][1] = fmaf(qv, kk.y, sacc[ii][1]);
        sacc[ii][2] = fmaf(qv, kk.z, sacc[ii][2]); sacc[ii][3] = fmaf(qv, kk.w, sacc[ii][3]);
      }
    }
    for (int e = 0; e < dvv; ++e) {
      const float4 vv = *reinterpret_cast<const float4*>(vT + e * L.kp + j0);
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) {
        const float ov = os[ir[ii] * L.vp + e];
        pacc[ii][0] = fmaf(ov, vv.x, pacc[ii][0]); pacc[ii][1] = fmaf(ov, vv.y, pacc[ii][1]);
        pacc[ii][2] = fmaf(ov, vv.z, pacc[ii][2]); pacc[ii][3] = fmaf(ov, vv.w, pacc[ii][3]);
      }
    }
#pragma unroll
    for (int ii = 0; ii < 4; ++ii) {
      const int i = i0 + ii;
      if (i >= nq) continue;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int jx = j0 + jj;
        float sv = -INFINITY;   // key padding: P = 0
        if (jx < nk) {
          sv = sacc[ii][jj] * a.scale;
          if (a.bias) sv += __ldg(a.bias + ((size_t)h * nq + i) * nk + jx);
        }
        P[i * L.sp + jx] = sv;
        D[i * L.sp + jx] = jx < nk ? pacc[ii][jj] : 0.f;
      }
    }
  }
  __syncthreads();
  {   // softmax rows, then dS = P * (dP - sum_j dP P): one warp per row (padded keys: P = exp(-inf) = 0)
    const int lane = tid & 31, warp = tid >> 5, nkp = nk4 * 4;
    for (int i = warp; i < nq; i += nt >> 5) {
      float* pr = P + i * L.sp;
      float* dr = D + i * L.sp;
      float mx = -INFINITY;
      for (int jx = lane; jx < nkp; jx += 32) mx = fmaxf(mx, pr[jx]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int jx = lane; jx < nkp; jx += 32) {
        const float e = expf(pr[jx] - mx);
        pr[jx] = e;
        sum += e;
      }
      const float inv = 1.0f / warp_sum(sum);
      float dot = 0.f;
      for (int jx = lane; jx < nkp; jx += 32) {
        pr[jx] *= inv;
        dot = fmaf(dr[jx], pr[jx], dot);
      }
      dot = warp_sum(dot);
      for (int jx = lane; jx < nkp; jx += 32) dr[jx] = pr[jx] * (dr[jx] - dot);
    }
  }
  __syncthreads();
  // dv[j, d] = sum_i P[i, j] do[i, d]  and  dk[j, d] = scale sum_i dS[i, j] q[i, d]  on 4 (keys) x 4 (channels) tiles
  {
    const int dv4 = dvv >> 2, dk4 = dkk >> 2;
    for (int item = tid; item < nk4 * (dv4 + dk4); item += nt) {
      const int j0 = (item / (dv4 + dk4)) * 4, c = item - (item / (dv4 + dk4)) * (dv4 + dk4);
      const bool is_v = c < dv4;
      const int d0 = (is_v ? c : c - dv4) * 4;
      const float* lhs = is_v ? P : D;                       // [i][j]
      const float* rhs = is_v ? os + d0 : qs + d0;           // [i][d]
      const int rp = is_v ? L.vp : L.qp;
      float acc[4][4];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) acc[jj][dd] = 0.f;
      for (int i = 0; i < nq; ++i) {
        const float4 l4 = *reinterpret_cast<const float4*>(lhs + i * L.sp + j0);
        const float4 r4 = *reinterpret_cast<const float4*>(rhs + i * rp);
        const float lv[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          acc[jj][0] = fmaf(lv[jj], r4.x, acc[jj][0]); acc[jj][1] = fmaf(lv[jj], r4.y, acc[jj][1]);
          acc[jj][2] = fmaf(lv[jj], r4.z, acc[jj][2]); acc[jj][3] = fmaf(lv[jj], r4.w, acc[jj][3]);
        }
      }
      const float sc = is_v ? 1.0f : a.scale;
      T* dst = is_v ? dv : dk;
      const int pitch = is_v ? a.v_pitch : a.k_pitch, off = is_v ? a.v_off + h * dvv : a.k_off + h * dkk;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        if (j0 + jj >= nk) continue;
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) dst[((size_t)seq * nk + j0 + jj) * pitch + off + d0 + dd] = from_f<T>(acc[jj][dd] * sc);
      }
    }
  }
  // dq[i, d] = scale sum_j dS[i, j] k[j, d]  on 4 (queries) x 4 (channels) tiles
  {
    const int dk4 = dkk >> 2;
    for (int item = tid; item < nq4 * dk4; item += nt) {
      const int i0 = (item / dk4) * 4, d0 = (item - (item / dk4) * dk4) * 4;
      int ir[4];
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) ir[ii] = min(i0 + ii, nq - 1);
      float acc[4][4];
#pragma unroll
      for (int ii = 0; ii < 4; ++ii)
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) acc[ii][dd] = 0.f;
      for (int jx = 0; jx < nk; ++jx) {
        const float4 k4 = *reinterpret_cast<const float4*>(ks + jx * L.qp + d0);
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          const float dsv = D[ir[ii] * L.sp + jx];
          acc[ii][0] = fmaf(dsv, k4.x, acc[ii][0]); acc[ii][1] = fmaf(dsv, k4.y, acc[ii][1]);
          acc[ii][2] = fmaf(dsv, k4.z, acc[ii][2]); acc[ii][3] = fmaf(dsv, k4.w, acc[ii][3]);
        }
      }
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) {
        if (i0 + ii >= nq) continue;
#pragma unroll
        for (int dd = 0; dd < 4; ++dd)
          dq[((size_t)seq * nq + i0 + ii) * a.q_pitch + a.q_off + h * dkk + d0 + dd] = from_f<T>(acc[ii][dd] * a.scale);
      }
    }
  }
}

}  // namespace

// ---- launchers (programmatic dependent launch like the forward kernels: every kernel starts with griddepcontrol.wait) ------
template <typename T> static const T* cp(const void* p) { return reinterpret_cast<const T*>(p); }
template <typename T> static T* mp(void* p) { return reinterpret_cast<T*>(p); }

cudaError_t launch_layernorm_bwd(const void* x, const float* gamma, const void* dy, const void* res, void* out,
                                 long long rows, int c, float eps, int dtype, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  if (c % 8 || c > 1024) return cudaErrorInvalidValue;
  const unsigned grid = (unsigned)((rows + 7) / 8);
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16)
    e_ = launch_pdl(layernorm_bwd_kernel<bf16>, dim3(grid), dim3(256), 0, s, cp<bf16>(x), gamma, cp<bf16>(dy), cp<bf16>(res), mp<bf16>(out), rows, c, eps);
  else
    e_ = launch_pdl(layernorm_bwd_kernel<float>, dim3(grid), dim3(256), 0, s, cp<float>(x), gamma, cp<float>(dy), cp<float>(res), mp<float>(out), rows, c, eps);
  return e_;
}

cudaError_t launch_gelu_bwd(const void* pre, void* g, long long total, int dtype, cudaStream_t s) {
  if (total <= 0) return cudaSuccess;
  if (total % 8) return cudaErrorInvalidValue;
  const long long total8 = total / 8;
  const unsigned grid = (unsigned)((total8 + 255) / 256);
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16) e_ = launch_pdl(gelu_bwd_kernel<bf16>, dim3(grid), dim3(256), 0, s, cp<bf16>(pre), mp<bf16>(g), total8);
  else e_ = launch_pdl(gelu_bwd_kernel<float>, dim3(grid), dim3(256), 0, s, cp<float>(pre), mp<float>(g), total8);
  return e_;
}

cudaError_t launch_move_rows(void* src, void* dst, long long nseq, int src_rows, int src_off, int dst_rows, int dst_off,
                             int nrows, int c, int dtype, cudaStream_t s) {
  const long long total = nseq * nrows * c;
  if (total <= 0) return cudaSuccess;
  const unsigned grid = (unsigned)((total + 255) / 256);
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16)
    e_ = launch_pdl(move_rows_kernel<bf16>, dim3(grid), dim3(256), 0, s, mp<bf16>(src), mp<bf16>(dst), nseq, src_rows, src_off, dst_rows, dst_off, nrows, c);
  else
    e_ = launch_pdl(move_rows_kernel<float>, dim3(grid), dim3(256), 0, s, mp<float>(src), mp<float>(dst), nseq, src_rows, src_off, dst_rows, dst_off, nrows, c);
  return e_;
}

cudaError_t launch_tail_bwd(const float* n_scale, const float* w, int n, int hw, int c, int num_classes,
                            const int32_t* target_cls, const float* logits, void* g, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16) e_ = launch_pdl(tail_bwd_kernel<bf16>, dim3(n), dim3(256), 0, s, n_scale, w, hw, c, num_classes, target_cls, logits, mp<bf16>(g));
  else e_ = launch_pdl(tail_bwd_kernel<float>, dim3(n), dim3(256), 0, s, n_scale, w, hw, c, num_classes, target_cls, logits, mp<float>(g));
  return e_;
}

cudaError_t launch_tail_bwd_general(const float* n_scale, const float* w, int n, int hw, int c, int num_classes,
                                    const float* grad_logits, void* g, int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16)
    e_ = launch_pdl(tail_bwd_general_kernel<bf16>, dim3(n), dim3(256), c * sizeof(float), s, n_scale, w, hw, c, num_classes, grad_logits, mp<bf16>(g));
  else
    e_ = launch_pdl(tail_bwd_general_kernel<float>, dim3(n), dim3(256), c * sizeof(float), s, n_scale, w, hw, c, num_classes, grad_logits, mp<float>(g));
  return e_;
}

cudaError_t launch_ct_pool_bwd(const void* gct, float* ga_nchw, int n, int side, int c, int kernel, int stride, int cs,
                               int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int g = (side - kernel) / stride + 1;
  if (g % cs) return cudaErrorInvalidValue;
  const dim3 grid((c * side * side + 255) / 256, n);
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16) e_ = launch_pdl(ct_pool_bwd_kernel<bf16>, dim3(grid), dim3(256), 0, s, cp<bf16>(gct), ga_nchw, side, c, kernel, stride, g, cs);
  else e_ = launch_pdl(ct_pool_bwd_kernel<float>, dim3(grid), dim3(256), 0, s, cp<float>(gct), ga_nchw, side, c, kernel, stride, g, cs);
  return e_;
}

cudaError_t launch_cam_from_ct_grad(const void* a, const void* gct, int n, int hw, int c, int ntok, float* cam_lowres,
                                    int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (c * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16)
    e_ = launch_pdl(cam_from_ct_grad_kernel<bf16>, dim3(n), dim3(256), c * sizeof(float), s, cp<bf16>(a), cp<bf16>(gct), hw, c, ntok, cam_lowres);
  else
    e_ = launch_pdl(cam_from_ct_grad_kernel<float>, dim3(n), dim3(256), c * sizeof(float), s, cp<float>(a), cp<float>(gct), hw, c, ntok, cam_lowres);
  return e_;
}

cudaError_t launch_conv3x3_s2_bwd(const void* dy, const void* w, void* dx, int n, int hin, int cin, int cinp, int cout,
                                  int dtype, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (cin > 512 || (size_t)kCbImgs * cout * sizeof(float) > 48 * 1024) return cudaErrorInvalidValue;
  const dim3 grid(hin * hin, (n + kCbImgs - 1) / kCbImgs);
  const size_t smem = (size_t)kCbImgs * cout * sizeof(float);
  count_launch();
  cudaError_t e_ = cudaSuccess;
  if (dtype == kBF16)
    e_ = launch_pdl(conv3x3_s2_bwd_kernel<bf16>, dim3(grid), dim3(256), smem, s, cp<bf16>(dy), cp<bf16>(w), mp<bf16>(dx), n, hin, cin, cinp, cout);
  else
    e_ = launch_pdl(conv3x3_s2_bwd_kernel<float>, dim3(grid), dim3(256), smem, s, cp<float>(dy), cp<float>(w), mp<float>(dx), n, hin, cin, cinp, cout);
  return e_;
}

// dq / dk / dv are written with q's / k's / v's pitch and column offset (the fused q|k|v layout of the forward GEMM).
cudaError_t launch_attention_bwd(const AttnArgs& a, const void* dout, void* dq, void* dk, void* dv, int dtype, cudaStream_t s) {
  if (a.nseq <= 0) return cudaSuccess;
  if (a.th1_w || a.vlocal || a.act != kActNone) return cudaErrorNotSupported;
  if (a.dk % 4 || a.dv % 4) return cudaErrorInvalidValue;
  const size_t smem = (size_t)AbwLayout(a.nq, a.nk, a.dk, a.dv).total * sizeof(float);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  static PerDeviceOnce once_b, once_f;
  count_launch();
  cudaError_t e_ = cudaSuccess;
  const dim3 grid(a.nseq, a.heads);
  if (dtype == kBF16) {
    if (once_b.need()) {
      cudaError_t e = cudaFuncSetAttribute(attention_bwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return e;
    }
    e_ = launch_pdl(attention_bwd_kernel<bf16>, dim3(grid), dim3(256), smem, s, a, cp<bf16>(dout), mp<bf16>(dq), mp<bf16>(dk), mp<bf16>(dv));
  } else {
    if (once_f.need()) {
      cudaError_t e = cudaFuncSetAttribute(attention_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return e;
    }
    e_ = launch_pdl(attention_bwd_kernel<float>, dim3(grid), dim3(256), smem, s, a, cp<float>(dout), mp<float>(dq), mp<float>(dk), mp<float>(dv));
  }
  return e_;
}

}  // namespace b200df
