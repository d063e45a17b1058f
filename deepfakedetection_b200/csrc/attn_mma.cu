// Small-sequence multi-head attention on the tensor cores (bf16 path), one CTA per sequence, eight warps.
//   S = (q k^T) * scale + bias;  [talking_head1];  P = softmax_j(S);  [talking_head2];  out = P v (+ v_local); [GELU]
// Same contract as attn.cu (the fp32 / any-shape CUDA-core kernel): replaces timm EfficientFormerV2 `Attention2d` /
// `Attention2dDownsample` and fastervit `WindowAttention` inside the reference's `model(images)`
// (/root/reference/orchestration/orchestrator.py:590).
//
// Why warp-level MMA (mma.sync m16n8k16, bf16 x bf16 -> fp32) and not a tcgen05 tile: a sequence is 16-53 queries by
// 16-196 keys per head with d = 16-128 -- one 128-row UMMA tile would hold 2.4-8 sequences of DIFFERENT keys, i.e. mostly
// masked cross terms, and the probabilities would have to leave TMEM, be re-laid out as a shared-memory operand and come
// back between the two products.  With register-resident fragments the whole (head, 16-query block) unit is one warp's
// work, flash-attention style: the S accumulators of two adjacent 8-key tiles ARE the A fragment of the P.V product, the
// softmax is two quad shuffles per row, and nothing but q, k, v is read and nothing but `out` is written.
//   * staging (per pass of `hp` heads): K and V rows as they are ([key][d], pitch d+8, 16-byte copies): the q.k^T B
//     fragment is a conflict-free 32-bit load, the P.V B fragment (keys x d, i.e. V transposed) comes out of
//     `ldmatrix.trans`, so no transposition is ever written; key pads are zeroed;
//   * Q fragments come straight from global memory (L2-hot GEMM output), the bias table from L2;
//   * talking heads (EfficientFormerV2) couple the heads between the two products: S of all heads goes to shared memory in
//     fp32, the two 8x8 mixes and the softmax run there (as in attn.cu), and the P.V A fragments are built from it.
#include "common.cuh"
#include "kernels.h"

#include <algorithm>

namespace b200df {

namespace {

constexpr int kMmaThreads = 256;
constexpr int kMmaWarps = 8;

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int DK, int DV, int NKT>
struct AttnGeom {
  static constexpr int NK8 = NKT * 8;                 // keys padded to whole 8-key S tiles
  static constexpr int NKP = (NK8 + 15) / 16 * 16;    // keys padded to whole 16-key P.V steps
  static constexpr int KP = DK + 8;                   // K row pitch (bf16 elements)
  static constexpr int VP = DV + 8;                   // V row pitch (bf16 elements; 16-byte aligned rows for ldmatrix)
  static constexpr int SP = NKP + 8;                  // fp32 score row pitch (talking-head path)
  static constexpr size_t k_bytes(int heads, int nk) { return (size_t)heads * ((nk + 7) / 8 * 8) * KP * 2; }
  static constexpr size_t v_bytes(int heads) { return (size_t)heads * NKP * VP * 2; }
};

// ---- staging ---------------------------------------------------------------------------------------------------
template <int DK, int DV, int NKT>
__device__ __forceinline__ void stage_k(const AttnArgs& a, int seq, int h0, int nh, bf16* Ks) {
  using G = AttnGeom<DK, DV, NKT>;
  const bf16* k = reinterpret_cast<const bf16*>(a.k);
  const int rows = G::NK8;
  for (int idx = threadIdx.x; idx < nh * rows * (DK / 8); idx += kMmaThreads) {
    // 8 consecutive lanes take 8 consecutive rows of ONE 16-byte column: with the 16-byte-odd row pitch the 128-bit
    // shared stores of a quarter warp land in 8 different bank groups (row-major lane order measured 8-way conflicts)
    int r = idx >> 3;
    const int c8 = r % (DK / 8);
    r /= (DK / 8);
    const int j = (r % (rows / 8)) * 8 + (idx & 7), hl = r / (rows / 8);
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (j < a.nk)
      v = __ldg(reinterpret_cast<const uint4*>(k + ((size_t)seq * a.nk + j) * a.k_pitch + a.k_off + (h0 + hl) * DK + c8 * 8));
    *reinterpret_cast<uint4*>(Ks + ((size_t)hl * rows + j) * G::KP + c8 * 8) = v;
  }
}

template <int DK, int DV, int NKT>
__device__ __forceinline__ void stage_vt(const AttnArgs& a, int seq, int h0, int nh, bf16* Vs) {
  using G = AttnGeom<DK, DV, NKT>;
  const bf16* v = reinterpret_cast<const bf16*>(a.v);
  for (int idx = threadIdx.x; idx < nh * G::NKP * (DV / 8); idx += kMmaThreads) {
    int r = idx >> 3;
    const int c8 = r % (DV / 8);
    r /= (DV / 8);
    const int j = (r % (G::NKP / 8)) * 8 + (idx & 7), hl = r / (G::NKP / 8);
    uint4 raw = make_uint4(0u, 0u, 0u, 0u);
    if (j < a.nk)
      raw = __ldg(reinterpret_cast<const uint4*>(v + ((size_t)seq * a.nk + j) * a.v_pitch + a.v_off + (h0 + hl) * DV + c8 * 8));
    *reinterpret_cast<uint4*>(Vs + ((size_t)hl * G::NKP + j) * G::VP + c8 * 8) = raw;
  }
}

// B fragment of the P.V product for keys [k0, k0 + 16) x channels [d0, d0 + 8): V is staged [key][d], the fragment wants
// (key pair, channel) per thread = the transposed 8x8 tiles -> ldmatrix.trans (lanes 0-15 supply the 16 key-row addresses)
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t& b0, uint32_t& b1, const bf16* row_ptr) {
  const uint32_t addr = static_cast<uint32_t>(__cvta_generic_to_shared(row_ptr));
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(addr));
}

// ---- S = q k^T for one (head, 16-query block): s[NKT][4] in the mma C layout -------------------------------------------
template <int DK, int DV, int NKT>
__device__ __forceinline__ void qk_unit(const AttnArgs& a, int seq, int h, int r0, const bf16* Kh, float (&s)[NKT][4]) {
  using G = AttnGeom<DK, DV, NKT>;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const bf16* q = reinterpret_cast<const bf16*>(a.q);
#pragma unroll
  for (int jt = 0; jt < NKT; ++jt)
#pragma unroll
    for (int e = 0; e < 4; ++e) s[jt][e] = 0.f;
  const int i0 = r0 + g, i1 = r0 + g + 8;
  const bf16* q0 = q + ((size_t)seq * a.nq + min(i0, a.nq - 1)) * a.q_pitch + a.q_off + h * DK;
  const bf16* q1 = q + ((size_t)seq * a.nq + min(i1, a.nq - 1)) * a.q_pitch + a.q_off + h * DK;
#pragma unroll
  for (int ks = 0; ks < DK / 16; ++ks) {
    uint32_t af[4];
    af[0] = i0 < a.nq ? __ldg(reinterpret_cast<const uint32_t*>(q0 + ks * 16 + 2 * t)) : 0u;
    af[1] = i1 < a.nq ? __ldg(reinterpret_cast<const uint32_t*>(q1 + ks * 16 + 2 * t)) : 0u;
    af[2] = i0 < a.nq ? __ldg(reinterpret_cast<const uint32_t*>(q0 + ks * 16 + 2 * t + 8)) : 0u;
    af[3] = i1 < a.nq ? __ldg(reinterpret_cast<const uint32_t*>(q1 + ks * 16 + 2 * t + 8)) : 0u;
#pragma unroll
    for (int jt = 0; jt < NKT; ++jt) {
      const bf16* kr = Kh + (size_t)(jt * 8 + g) * G::KP + ks * 16 + 2 * t;
      mma_bf16_16816(s[jt], af, *reinterpret_cast<const uint32_t*>(kr), *reinterpret_cast<const uint32_t*>(kr + 8));
    }
  }
}

// ---- out = P v for one (head, 16-query block); P given as A fragments per 16-key step ------------------------------------
template <int DK, int DV, int NKT, typename AFrag>
__device__ __forceinline__ void pv_unit(const AttnArgs& a, int seq, int h, int r0, const bf16* Vh, AFrag afrag, float inv0,
                                        float inv1) {
  using G = AttnGeom<DK, DV, NKT>;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  float o[DV / 8][4];
#pragma unroll
  for (int dt = 0; dt < DV / 8; ++dt)
#pragma unroll
    for (int e = 0; e < 4; ++e) o[dt][e] = 0.f;
#pragma unroll
  for (int kt = 0; kt < G::NKP / 16; ++kt) {
    uint32_t af[4];
    afrag(kt, af);
#pragma unroll
    for (int dt = 0; dt < DV / 8; ++dt) {
      uint32_t b0, b1;
      ldmatrix_x2_trans(b0, b1, Vh + (size_t)(kt * 16 + (lane & 15)) * G::VP + dt * 8);
      mma_bf16_16816(o[dt], af, b0, b1);
    }
  }
  const int oc = a.heads * DV;
  bf16* out = reinterpret_cast<bf16*>(a.out);
  const bf16* vl = reinterpret_cast<const bf16*>(a.vlocal);
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int i = r0 + g + 8 * half;
    if (i >= a.nq) continue;
    const float inv = half ? inv1 : inv0;
    const size_t base = ((size_t)seq * a.nq + i) * oc + h * DV + 2 * t;
#pragma unroll
    for (int dt = 0; dt < DV / 8; ++dt) {
      float x0 = o[dt][2 * half] * inv, x1 = o[dt][2 * half + 1] * inv;
      if (vl) {
        const uint32_t l = __ldg(reinterpret_cast<const uint32_t*>(vl + base + dt * 8));
        x0 += __uint_as_float(l << 16);
        x1 += __uint_as_float(l & 0xffff0000u);
      }
      x0 = apply_act<true>(x0, a.act);
      x1 = apply_act<true>(x1, a.act);
      *reinterpret_cast<uint32_t*>(out + base + dt * 8) = pack_bf16x2(x0, x1);
    }
  }
}

// ---- no talking heads: everything between the two products stays in registers -----------------------------------------------
template <int DK, int DV, int NKT>
__global__ void __launch_bounds__(kMmaThreads)
attn_mma_kernel(const AttnArgs a, int hp) {
  using G = AttnGeom<DK, DV, NKT>;
  extern __shared__ __align__(16) uint8_t sm[];
  bf16* Ks = reinterpret_cast<bf16*>(sm);
  bf16* Vt = reinterpret_cast<bf16*>(sm + G::k_bytes(hp, G::NK8));
  const int seq = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int RB = (a.nq + 15) / 16;
  griddep_wait();
  griddep_launch_dependents();
  for (int h0 = 0; h0 < a.heads; h0 += hp) {
    const int nh = min(hp, a.heads - h0);
    __syncthreads();   // the previous pass is done with the staged tiles
    stage_k<DK, DV, NKT>(a, seq, h0, nh, Ks);
    stage_vt<DK, DV, NKT>(a, seq, h0, nh, Vt);
    __syncthreads();
    for (int u = warp; u < nh * RB; u += kMmaWarps) {
      const int hl = u / RB, r0 = (u - hl * RB) * 16, h = h0 + hl;
      float s[NKT][4];
      qk_unit<DK, DV, NKT>(a, seq, h, r0, Ks + (size_t)hl * G::NK8 * G::KP, s);
      // scale + bias, key mask, row max / sum (two rows per thread: g and g + 8; a row lives in one quad)
      float mx[2] = {-INFINITY, -INFINITY};
      if (a.bias_frag) {
        // bias in the accumulator-fragment order (one float4 per thread and key tile; masked keys hold -inf, rows past
        // the sequence 0): NKT independent 16-byte loads instead of 4*NKT predicated scalar ones
        const float4* bf = reinterpret_cast<const float4*>(a.bias_frag) + ((size_t)(h * RB + r0 / 16) * NKT) * 32 + lane;
        float4 b4[NKT];
#pragma unroll
        for (int jt = 0; jt < NKT; ++jt) b4[jt] = __ldg(bf + jt * 32);
#pragma unroll
        for (int jt = 0; jt < NKT; ++jt) {
          s[jt][0] = fmaf(s[jt][0], a.scale, b4[jt].x);
          s[jt][1] = fmaf(s[jt][1], a.scale, b4[jt].y);
          s[jt][2] = fmaf(s[jt][2], a.scale, b4[jt].z);
          s[jt][3] = fmaf(s[jt][3], a.scale, b4[jt].w);
          mx[0] = fmaxf(mx[0], fmaxf(s[jt][0], s[jt][1]));
          mx[1] = fmaxf(mx[1], fmaxf(s[jt][2], s[jt][3]));
        }
      } else {
#pragma unroll
        for (int jt = 0; jt < NKT; ++jt)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i = r0 + g + 8 * (e >> 1), j = jt * 8 + 2 * t + (e & 1);
            float v = -INFINITY;
            if (j < a.nk) {
              v = s[jt][e] * a.scale;
              if (a.bias && i < a.nq) v += __ldg(a.bias + ((size_t)h * a.nq + i) * a.nk + j);
            }
            s[jt][e] = v;
            mx[e >> 1] = fmaxf(mx[e >> 1], v);
          }
      }
      float sum[2] = {0.f, 0.f};
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      }
#pragma unroll
      for (int jt = 0; jt < NKT; ++jt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float p = __expf(s[jt][e] - mx[e >> 1]);   // exp(-inf) = 0 for the masked keys
          s[jt][e] = p;
          sum[e >> 1] += p;
        }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 1);
        sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 2);
      }
      auto afrag = [&](int kt, uint32_t (&af)[4]) {   // C fragments of S tiles 2kt, 2kt+1 = A fragment of key step kt
        af[0] = pack_bf16x2(s[2 * kt][0], s[2 * kt][1]);
        af[1] = pack_bf16x2(s[2 * kt][2], s[2 * kt][3]);
        if (2 * kt + 1 < NKT) {
          af[2] = pack_bf16x2(s[2 * kt + 1][0], s[2 * kt + 1][1]);
          af[3] = pack_bf16x2(s[2 * kt + 1][2], s[2 * kt + 1][3]);
        } else {
          af[2] = 0u;
          af[3] = 0u;
        }
      };
      pv_unit<DK, DV, NKT>(a, seq, h, r0, Vt + (size_t)hl * G::NKP * G::VP, afrag, 1.0f / sum[0], 1.0f / sum[1]);
    }
  }
}

// ---- talking heads (EfficientFormerV2 Attention2d): scores of ALL heads meet in shared memory between the products ---------
template <int DK, int DV, int NKT>
__global__ void __launch_bounds__(kMmaThreads)
attn_mma_th_kernel(const AttnArgs a, int hp) {
  using G = AttnGeom<DK, DV, NKT>;
  extern __shared__ __align__(16) uint8_t sm[];
  const int H = a.heads, nq = a.nq, nk = a.nk;
  float* S = reinterpret_cast<float*>(sm);                                        // [H][nq][SP] fp32
  uint8_t* region = sm + (((size_t)H * nq * G::SP * sizeof(float) + 15) & ~size_t(15));
  bf16* Ks = reinterpret_cast<bf16*>(region);                                      // all heads (scores phase)
  bf16* Vt = reinterpret_cast<bf16*>(region);                                      // hp heads at a time (output phase)
  const int seq = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int RB = (nq + 15) / 16;
  __shared__ __align__(16) float thw[2 * (16 * 16 + 16)];
  for (int i = threadIdx.x; i < 2 * (H * H + H); i += kMmaThreads) {   // weights only: before the dependency wait
    const int pass = i / (H * H + H), r = i - pass * (H * H + H);
    const float* w = pass == 0 ? a.th1_w : a.th2_w;
    const float* b = pass == 0 ? a.th1_b : a.th2_b;
    thw[pass * (16 * 16 + 16) + (r < H * H ? r : 16 * 16 + r - H * H)] = r < H * H ? __ldg(w + r) : __ldg(b + r - H * H);
  }
  griddep_wait();
  griddep_launch_dependents();
  // ---- scores ----
  stage_k<DK, DV, NKT>(a, seq, 0, H, Ks);
  __syncthreads();
  for (int u = warp; u < H * RB; u += kMmaWarps) {
    const int h = u / RB, r0 = (u - h * RB) * 16;
    float s[NKT][4];
    qk_unit<DK, DV, NKT>(a, seq, h, r0, Ks + (size_t)h * G::NK8 * G::KP, s);
    const float4* bf = a.bias_frag ? reinterpret_cast<const float4*>(a.bias_frag) + ((size_t)(h * RB + r0 / 16) * NKT) * 32 + lane
                                   : nullptr;
    float4 b4[NKT];
#pragma unroll
    for (int jt = 0; jt < NKT; ++jt) b4[jt] = bf ? __ldg(bf + jt * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int jt = 0; jt < NKT; ++jt) {
      const float bb[4] = {b4[jt].x, b4[jt].y, b4[jt].z, b4[jt].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = r0 + g + 8 * (e >> 1), j = jt * 8 + 2 * t + (e & 1);
        if (i < nq && j < nk) {
          float v = fmaf(s[jt][e], a.scale, bb[e]);
          if (!bf && a.bias) v += __ldg(a.bias + ((size_t)h * nq + i) * nk + j);
          S[((size_t)h * nq + i) * G::SP + j] = v;
        }
      }
    }
  }
  // zero the key padding of every row once: it is read as P = 0 by the 16-key steps of the second product
  for (int idx = threadIdx.x; idx < H * nq * (G::SP - nk); idx += kMmaThreads) {
    const int r = idx / (G::SP - nk), c = idx - r * (G::SP - nk);
    S[(size_t)r * G::SP + nk + c] = 0.f;
  }
  __syncthreads();
  // ---- talking head 1 -> softmax over keys -> talking head 2 ----
  if (NKT == 7 && H == 8) {
    // EfficientFormerV2's shape: ONE pass over S.  A quarter warp owns a query row for ALL heads (8 lanes x 7 keys each,
    // 8 heads: 56 values per lane in registers), so the two 8x8 head mixes are register-local and the softmax
    // reductions are three shuffles inside the 8-lane group; four rows per warp are in flight at once.
    const float* w1 = thw, *b1 = thw + 256, *w2 = thw + 272, *b2 = thw + 272 + 256;
    const int q4 = lane >> 3, l8 = lane & 7;
    const size_t hs = (size_t)nq * G::SP;
    for (int base = warp * 4; base < nq; base += kMmaWarps * 4) {
      const int i = base + q4;
      const bool row_ok = i < nq;
      float* sp = S + (size_t)min(i, nq - 1) * G::SP + l8;
      float v[8][7], o[8][7];
#pragma unroll
      for (int h = 0; h < 8; ++h)
#pragma unroll
        for (int k = 0; k < 7; ++k) v[h][k] = sp[h * hs + 8 * k];     // j = l8 + 8k < 56 <= SP: pads hold zeros
#pragma unroll
      for (int ho = 0; ho < 8; ++ho) {
        const float4 wa = *reinterpret_cast<const float4*>(w1 + ho * 8), wb = *reinterpret_cast<const float4*>(w1 + ho * 8 + 4);
        const float bb = b1[ho];
        float mx = -INFINITY;
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          const float x = bb + wa.x * v[0][k] + wa.y * v[1][k] + wa.z * v[2][k] + wa.w * v[3][k] + wb.x * v[4][k] +
                          wb.y * v[5][k] + wb.z * v[6][k] + wb.w * v[7][k];
          o[ho][k] = l8 + 8 * k < nk ? x : -INFINITY;
          mx = fmaxf(mx, o[ho][k]);
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          o[ho][k] = __expf(o[ho][k] - mx);
          sum += o[ho][k];
        }
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        sum += __shfl_xor_sync(0xffffffffu, sum, 4);
        const float inv = 1.0f / sum;
#pragma unroll
        for (int k = 0; k < 7; ++k) o[ho][k] *= inv;
      }
#pragma unroll
      for (int ho = 0; ho < 8; ++ho) {
        const float4 wa = *reinterpret_cast<const float4*>(w2 + ho * 8), wb = *reinterpret_cast<const float4*>(w2 + ho * 8 + 4);
        const float bb = b2[ho];
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          const float x = bb + wa.x * o[0][k] + wa.y * o[1][k] + wa.z * o[2][k] + wa.w * o[3][k] + wb.x * o[4][k] +
                          wb.y * o[5][k] + wb.z * o[6][k] + wb.w * o[7][k];
          if (row_ok && l8 + 8 * k < nk) sp[ho * hs + 8 * k] = x;
        }
      }
    }
    __syncthreads();
  } else {
    for (int pass = 0; pass < 2; ++pass) {
      const float* w = thw + pass * (16 * 16 + 16);   // [H][H] mix + [H] bias, staged in shared memory at kernel start
      const float* b = w + 16 * 16;
      for (int idx = threadIdx.x; idx < nq * nk; idx += kMmaThreads) {
        const int i = idx / nk, j = idx - i * nk;
        float* sp = S + (size_t)i * G::SP + j;
        const size_t hs = (size_t)nq * G::SP;
        if (H == 8) {
          float v[8], o[8];
  #pragma unroll
          for (int h = 0; h < 8; ++h) v[h] = sp[h * hs];
  #pragma unroll
          for (int ho = 0; ho < 8; ++ho) {
            const float4 w0 = *reinterpret_cast<const float4*>(w + ho * 8), w1 = *reinterpret_cast<const float4*>(w + ho * 8 + 4);
            o[ho] = b[ho] + w0.x * v[0] + w0.y * v[1] + w0.z * v[2] + w0.w * v[3] + w1.x * v[4] + w1.y * v[5] + w1.z * v[6] + w1.w * v[7];
          }
  #pragma unroll
          for (int h = 0; h < 8; ++h) sp[h * hs] = o[h];
        } else {
          float v[16];
          for (int h = 0; h < H; ++h) v[h] = sp[h * hs];
          for (int ho = 0; ho < H; ++ho) {
            float o = b[ho];
            for (int h = 0; h < H; ++h) o = fmaf(w[ho * H + h], v[h], o);
            sp[ho * hs] = o;
          }
        }
      }
      __syncthreads();
      if (pass == 0) {
        for (int r = warp; r < H * nq; r += kMmaWarps) {
          float* row = S + (size_t)r * G::SP;
          float mx = -INFINITY;
          for (int j = lane; j < nk; j += 32) mx = fmaxf(mx, row[j]);
          mx = warp_max(mx);
          float sum = 0.f;
          for (int j = lane; j < nk; j += 32) {
            const float e = __expf(row[j] - mx);
            row[j] = e;
            sum += e;
          }
          const float inv = 1.0f / warp_sum(sum);
          for (int j = lane; j < nk; j += 32) row[j] *= inv;
        }
        __syncthreads();
      }
    }
  }
  // ---- out = P v (+ v_local), hp heads per pass ----
  for (int h0 = 0; h0 < H; h0 += hp) {
    const int nh = min(hp, H - h0);
    __syncthreads();
    stage_vt<DK, DV, NKT>(a, seq, h0, nh, Vt);
    __syncthreads();
    for (int u = warp; u < nh * RB; u += kMmaWarps) {
      const int hl = u / RB, r0 = (u - hl * RB) * 16, h = h0 + hl;
      const float* s0 = S + ((size_t)h * nq + min(r0 + g, nq - 1)) * G::SP;
      const float* s1 = S + ((size_t)h * nq + min(r0 + g + 8, nq - 1)) * G::SP;
      const bool ok0 = r0 + g < nq, ok1 = r0 + g + 8 < nq;
      auto afrag = [&](int kt, uint32_t (&af)[4]) {
        const float2 p00 = *reinterpret_cast<const float2*>(s0 + kt * 16 + 2 * t);
        const float2 p10 = *reinterpret_cast<const float2*>(s1 + kt * 16 + 2 * t);
        const float2 p01 = *reinterpret_cast<const float2*>(s0 + kt * 16 + 2 * t + 8);
        const float2 p11 = *reinterpret_cast<const float2*>(s1 + kt * 16 + 2 * t + 8);
        af[0] = ok0 ? pack_bf16x2(p00.x, p00.y) : 0u;
        af[1] = ok1 ? pack_bf16x2(p10.x, p10.y) : 0u;
        af[2] = ok0 ? pack_bf16x2(p01.x, p01.y) : 0u;
        af[3] = ok1 ? pack_bf16x2(p11.x, p11.y) : 0u;
      };
      pv_unit<DK, DV, NKT>(a, seq, h, r0, Vt + (size_t)hl * G::NKP * G::VP, afrag, 1.0f, 1.0f);
    }
  }
}

template <int DK, int DV, int NKT>
cudaError_t mma_launch(const AttnArgs& a, bool th, cudaStream_t s) {
  using G = AttnGeom<DK, DV, NKT>;
  const size_t budget = th ? 200 * 1024 : 100 * 1024;  // no talking heads: two CTAs per SM
  const size_t s_bytes = th ? (((size_t)a.heads * a.nq * G::SP * sizeof(float) + 15) & ~size_t(15)) : 0;
  int hp = a.heads;
  auto need = [&](int h) {
    return th ? s_bytes + std::max(G::k_bytes(a.heads, G::NK8), G::v_bytes(h)) : G::k_bytes(h, G::NK8) + G::v_bytes(h);
  };
  while (hp > 1 && need(hp) > budget) hp = (hp + 1) / 2;
  const size_t smem = need(hp);
  if (smem > 220 * 1024) return cudaErrorInvalidValue;
  count_launch();
  if (th) {
    static PerDeviceOnce once;
    if (once.need()) {
      cudaError_t e = cudaFuncSetAttribute(attn_mma_th_kernel<DK, DV, NKT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
      if (e != cudaSuccess) return e;
    }
    return launch_pdl(attn_mma_th_kernel<DK, DV, NKT>, dim3(a.nseq), dim3(kMmaThreads), smem, s, a, hp);
  }
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(attn_mma_kernel<DK, DV, NKT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e != cudaSuccess) return e;
  }
  return launch_pdl(attn_mma_kernel<DK, DV, NKT>, dim3(a.nseq), dim3(kMmaThreads), smem, s, a, hp);
}

}  // namespace

// ---- bias table in accumulator-fragment order ---------------------------------------------------------------------------
// element (h, rb, jt, lane, e): query i = rb*16 + (lane>>2) + 8*(e>>1), key j = jt*8 + 2*(lane&3) + (e&1);
// value = bias[h][i][j], 0 for i >= nq, -inf for j >= nk (the key mask comes for free with the bias add).
size_t attention_bias_frag_elems(int heads, int nq, int nk) {
  return (size_t)heads * ((nq + 15) / 16) * ((nk + 7) / 8) * 32 * 4;
}

void attention_bias_frag_host(const float* bias, int heads, int nq, int nk, float* out) {
  const int RB = (nq + 15) / 16, NKT = (nk + 7) / 8;
  for (int h = 0; h < heads; ++h)
    for (int rb = 0; rb < RB; ++rb)
      for (int jt = 0; jt < NKT; ++jt)
        for (int lane = 0; lane < 32; ++lane)
          for (int e = 0; e < 4; ++e) {
            const int i = rb * 16 + (lane >> 2) + 8 * (e >> 1), j = jt * 8 + 2 * (lane & 3) + (e & 1);
            const float v = j >= nk ? -INFINITY : (i >= nq ? 0.f : (bias ? bias[((size_t)h * nq + i) * nk + j] : 0.f));
            out[((((size_t)h * RB + rb) * NKT + jt) * 32 + lane) * 4 + e] = v;
          }
}

namespace {
__global__ void bias_frag_kernel(const float* __restrict__ bias, int heads, int nq, int nk, float* __restrict__ out, size_t total) {
  const int RB = (nq + 15) / 16, NKT = (nk + 7) / 8;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int e = (int)(idx & 3), lane = (int)((idx >> 2) & 31);
    size_t r = idx >> 7;
    const int jt = (int)(r % NKT);
    r /= NKT;
    const int rb = (int)(r % RB), h = (int)(r / RB);
    const int i = rb * 16 + (lane >> 2) + 8 * (e >> 1), j = jt * 8 + 2 * (lane & 3) + (e & 1);
    out[idx] = j >= nk ? -INFINITY : (i >= nq ? 0.f : (bias ? bias[((size_t)h * nq + i) * nk + j] : 0.f));
  }
}
}  // namespace

cudaError_t launch_attention_bias_frag(const float* bias, int heads, int nq, int nk, float* out, cudaStream_t s) {
  const size_t total = attention_bias_frag_elems(heads, nq, nk);
  bias_frag_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(bias, heads, nq, nk, out, total);
  return cudaGetLastError();
}

// Returns cudaErrorNotSupported for shapes this kernel is not instantiated for (the caller falls back to attn.cu).
cudaError_t launch_attention_mma(const AttnArgs& a, cudaStream_t s) {
  if (a.nseq <= 0) return cudaSuccess;
  const bool th = a.th1_w != nullptr;
  if (a.heads > 16 || a.nq > 64 || (a.q_pitch | a.k_pitch | a.v_pitch | a.q_off | a.k_off | a.v_off) % 8) return cudaErrorNotSupported;
  const int nkt = (a.nk + 7) / 8;
#define MMA_CASE(DK_, DV_, NKT_, TH_) \
  if (a.dk == DK_ && a.dv == DV_ && nkt == NKT_ && th == TH_) return mma_launch<DK_, DV_, NKT_>(a, th, s)
  MMA_CASE(32, 32, 7, false);    // FasterViT-0 window (53 tokens) and level-3 (49 tokens) attention
  MMA_CASE(32, 32, 2, false);    // FasterViT-0 carrier-token attention (16 tokens)
  MMA_CASE(48, 48, 7, false);    // FasterViT-2
  MMA_CASE(48, 48, 2, false);
  MMA_CASE(32, 128, 7, true);    // EfficientFormerV2 Attention2d (49 tokens, talking heads)
  MMA_CASE(16, 64, 25, false);   // EfficientFormerV2 Attention2dDownsample (49 queries x 196 keys)
#undef MMA_CASE
  return cudaErrorNotSupported;
}

}  // namespace b200df
