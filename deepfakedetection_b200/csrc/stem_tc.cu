// Stem as an implicit GEMM on the 5th-gen tensor cores (bf16 path): 3x3 stride-2 conv 3->32 (pad 1) or 3->40
// (TF-"same" pad (0, 1), the efficientnet_pytorch B3 stem)
// (+folded BN bias, SiLU).  Replaces timm `conv_stem` + `bn1` at the head of the
// reference's `model(images)` call (/root/reference/orchestration/orchestrator.py:590).
//
//   out[p, oc] = silu( sum_k A[p, k] * W[oc, k] + bias[oc] ),  k = (c, ky, kx) in [0, 27) padded to 32
//
// HBM-bound by intent (0.6 MB in + 0.8 MB out per image): the SIMT version spends ~860 FMA issue slots per
// pixel; here a pixel costs 27 gathers, 16 packs and 4 shared stores, and the 27x32 contraction is two
// tcgen05.mma (M=128 pixels, N=32, K=16 each):
//   * a CTA (128 threads) owns tiles of 128 consecutive output pixels; thread t builds im2col row t straight
//     from the caller's fp32 NCHW image (or uint8 HWC with ToTensor+Normalize fused, orchestrator.py:340-345),
//     rounds to bf16 and writes it into the 128B-swizzled K-major A tile; zero padding = predicated loads;
//   * thread 0 issues the MMAs into a 32-column TMEM accumulator and commits to an mbarrier;
//   * every warp reads its 32 TMEM lanes (tcgen05.ld 32x32b.x32), adds bias, SiLU (one MUFU.TANH), packs
//     bf16 and writes its pixel's 32 channels (64 contiguous bytes; a warp writes 2 KB contiguous);
//   * latency is hidden across CTAs: 6 persistent CTAs per SM (TMEM 6 x 32 columns, 21 KB smem each).
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

namespace b200df {

using namespace tc;

constexpr int kStT = 128;       // pixels per tile = threads per CTA
constexpr int kStCtasPerSm = 6;

// OC: output channels (32 timm B0 | 40 efficientnet_pytorch B3; the MMA's N is OC rounded up to 16).
// PAD: zero rows/columns before the image (1 symmetric | 0 = TF-"same" (0, 1), Conv2dStaticSamePadding).
// ACT: common.cuh Act applied after the folded bias (SiLU EfficientNet | GELU EfficientFormerV2 `stem.conv1` |
// ReLU FasterViT `patch_embed.conv_down.0`).
template <bool kU8, int OC, int PAD, int ACT>
__global__ void __launch_bounds__(kStT, kStCtasPerSm)
stem_tc_kernel(const void* __restrict__ xin, int n, int H, int W, const float* __restrict__ wgt,
               const float* __restrict__ bias, bf16* __restrict__ out, float sc0, float sc1, float sc2, float sh0,
               float sh1, float sh2) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;                  // [128 rows x 128 B], chunks 0-3 of every row used (K = 32)
  constexpr int NB = (OC + 15) / 16 * 16;     // MMA N / rows of the weight tile (rows >= OC are zero)
  constexpr uint32_t kTmemCols = NB <= 32 ? 32 : 64;
  uint8_t* sB = smem + kStT * 128;     // [NB rows x 128 B]
  float* sbias = reinterpret_cast<float*>(sB + NB * 128);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sbias + NB);
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int Ho = H >> 1, Wo = W >> 1;
  const long long npix = (long long)n * Ho * Wo;
  const int tiles = (int)((npix + kStT - 1) / kStT);

  // weights [OC][27] fp32 -> bf16 K-major tile, k padded to 32 and rows padded to NB with zeros
  for (int i = tid; i < NB * 4; i += kStT) {
    const int oc = i >> 2, ch = i & 3;
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = ch * 8 + j;
      f[j] = (k < 27 && oc < OC) ? __ldg(wgt + oc * 27 + k) : 0.f;
    }
    *reinterpret_cast<uint4*>(sB + oc * 128 + ((ch ^ (oc & 7)) << 4)) = pack_bf16x8(f);
  }
  if (tid < NB) sbias[tid] = tid < OC ? __ldg(bias + tid) : 0.f;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr, kTmemCols);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  griddep_wait();  // `out` may still be read by the previous forward's kernels
  griddep_launch_dependents();
  const uint32_t idesc = make_idesc_bf16_m128((uint32_t)NB);
  const uint64_t da = make_sw128_kmajor_desc(smem_u32(sA));
  const uint64_t db = make_sw128_kmajor_desc(smem_u32(sB));
  uint32_t phase = 0;

  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long p = (long long)t * kStT + tid;
    const bool valid = p < npix;
    int img = 0, oh = 0, ow = 0;
    if (valid) {
      img = (int)(p / (Ho * Wo));
      const int r = (int)(p - (long long)img * Ho * Wo);
      oh = r / Wo;
      ow = r - oh * Wo;
    }
    // ---- im2col row of this pixel: k = c*9 + ky*3 + kx ----
    float f[32];
#pragma unroll
    for (int k = 27; k < 32; ++k) f[k] = 0.f;
    const int ih0 = 2 * oh - PAD, iw0 = 2 * ow - PAD;
    if constexpr (!kU8) {
      // fp32 NCHW: taps kx = 1, 2 of a pixel are the aligned pair (2 ow, 2 ow + 1) -> one 8-byte load; tap kx = 0
      // (2 ow - 1) is the second half of the LEFT neighbour's pair -> a shuffle from lane - 1.  Lanes whose left
      // neighbour sits in another warp / row (lane 0, ow == 0) load it themselves (ow == 0: zero padding).
      const float* x = reinterpret_cast<const float*>(xin);
      const int lane = tid & 31;
      // PAD == 0 (TF-"same"): taps kx = 0, 1 are the aligned pair and tap kx = 2 (2 ow + 2) is the first half of the
      // RIGHT neighbour's pair -> a shuffle from lane + 1 (lane 31 / ow == Wo - 1: own load / zero padding).
      const bool own_side = PAD == 1 ? (lane == 0 || ow == 0) : (lane == 31 || ow == Wo - 1);
      const bool side_in = PAD == 1 ? ow > 0 : ow < Wo - 1;   // the side tap lies inside the image
      float2 pr[9];
      float own[9];
      // all 18 loads first (predicated, no branches: they must all be in flight together), shuffles afterwards
#pragma unroll
      for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int ih = ih0 + ky;
          const bool row_ok = valid && ih >= 0 && ih < H;
          const float* rowp = x + (((size_t)img * 3 + c) * H + (row_ok ? ih : 0)) * W;
          pr[c * 3 + ky] = make_float2(0.f, 0.f);
          own[c * 3 + ky] = 0.f;
          if (row_ok) pr[c * 3 + ky] = __ldg(reinterpret_cast<const float2*>(rowp + 2 * ow));
          if (row_ok && own_side && side_in) own[c * 3 + ky] = __ldg(rowp + (PAD == 1 ? 2 * ow - 1 : 2 * ow + 2));
        }
      }
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        if constexpr (PAD == 1) {
          const float up = __shfl_up_sync(0xffffffffu, pr[i].y, 1);
          f[i * 3 + 0] = own_side ? own[i] : up;
          f[i * 3 + 1] = pr[i].x;
          f[i * 3 + 2] = pr[i].y;
        } else {
          const float dn = __shfl_down_sync(0xffffffffu, pr[i].x, 1);
          f[i * 3 + 0] = pr[i].x;
          f[i * 3 + 1] = pr[i].y;
          f[i * 3 + 2] = own_side ? own[i] : dn;
        }
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int ih = ih0 + ky;
          const bool row_ok = valid && ih >= 0 && ih < H;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const int iw = iw0 + kx;
            const bool ok = row_ok && iw >= 0 && iw < W;
            float v = 0.f;
            const uint8_t* x = reinterpret_cast<const uint8_t*>(xin);
            if (ok) {
              const float u = (float)__ldg(x + (((size_t)img * H + ih) * W + iw) * 3 + c);
              v = fmaf(u, c == 0 ? sc0 : (c == 1 ? sc1 : sc2), c == 0 ? sh0 : (c == 1 ? sh1 : sh2));
            }
            f[c * 9 + ky * 3 + kx] = v;
          }
        }
      }
    }
#pragma unroll
    for (int ch = 0; ch < 4; ++ch) {
      uint4 pk;
      pk.x = pack_bf16x2(f[ch * 8 + 0], f[ch * 8 + 1]);
      pk.y = pack_bf16x2(f[ch * 8 + 2], f[ch * 8 + 3]);
      pk.z = pack_bf16x2(f[ch * 8 + 4], f[ch * 8 + 5]);
      pk.w = pack_bf16x2(f[ch * 8 + 6], f[ch * 8 + 7]);
      *reinterpret_cast<uint4*>(sA + tid * 128 + ((ch ^ (tid & 7)) << 4)) = pk;
    }
    fence_proxy_async_smem();
    tcgen05_fence_before();  // this thread's TMEM reads of the previous tile are ordered before the barrier
    __syncthreads();
    if (tid == 0) {
      tcgen05_fence_after();
      umma_bf16(tmem_base, da, db, idesc, 0u);
      umma_bf16(tmem_base, da + 2ull, db + 2ull, idesc, 1u);  // +32 B along K inside the swizzle atom
      umma_commit(bar);
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    tcgen05_fence_after();
    uint32_t r[32];
    uint32_t r2[32];
    tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16), r);
    if constexpr (OC > 48) tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + 32u, r2);
    else if constexpr (OC > 32) tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + 32u, reinterpret_cast<uint32_t(&)[16]>(r2));
    tmem_ld_wait();
    // ---- epilogue: bias + SiLU + bf16, OC * 2 contiguous bytes per pixel ----
    uint4 o[(OC + 7) / 8];
#pragma unroll
    for (int ch = 0; ch < (OC < 32 ? OC / 8 : 4); ++ch) {
      float g[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = act_fn<ACT, true>(__uint_as_float(r[ch * 8 + j]) + sbias[ch * 8 + j]);
      o[ch] = pack_bf16x8(g);
    }
    if constexpr (OC > 32) {
#pragma unroll
      for (int ch = 4; ch < OC / 8; ++ch) {
        float g[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = act_fn<ACT, true>(__uint_as_float(r2[(ch - 4) * 8 + j]) + sbias[ch * 8 + j]);
        o[ch] = pack_bf16x8(g);
      }
    }
    if (valid) {
      uint4* dst = reinterpret_cast<uint4*>(out + (size_t)p * OC);
#pragma unroll
      for (int ch = 0; ch < OC / 8; ++ch) dst[ch] = o[ch];
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <bool kU8, int OC, int PAD, int ACT>
static cudaError_t stem_tc_launch(const void* x, int n, int h, int w, const float* wgt, const float* bias, void* out,
                                  float sc0, float sc1, float sc2, float sh0, float sh1, float sh2, cudaStream_t s) {
  constexpr int NB = (OC + 15) / 16 * 16;
  if (n <= 0) return cudaSuccess;
  if ((h & 1) || (w & 1)) return cudaErrorInvalidValue;
  const long long npix = (long long)n * (h / 2) * (w / 2);
  const long long tiles = (npix + kStT - 1) / kStT;
  const size_t smem = 1024 + kStT * 128 + NB * 128 + NB * sizeof(float) + 32;
  const int grid = (int)(tiles < (long long)kNumSMs * kStCtasPerSm ? tiles : (long long)kNumSMs * kStCtasPerSm);
  count_launch();
  return launch_pdl(stem_tc_kernel<kU8, OC, PAD, ACT>, dim3(grid), dim3(kStT), smem, s, x, n, h, w, wgt, bias,
                    reinterpret_cast<bf16*>(out), sc0, sc1, sc2, sh0, sh1, sh2);
}

bool stem_tc_supported(int oc, int pad, int act) {
  return (oc == 32 && pad == 1 && act == kActSiLU) || (oc == 40 && pad == 0 && act == kActSiLU) ||
         (oc == 16 && pad == 1 && act == kActGELU) || (oc == 64 && pad == 1 && act == kActReLU);
}

cudaError_t launch_stem_tc(const float* x, int n, int h, int w, const float* wgt, const float* bias, void* out,
                           cudaStream_t s, int oc, int pad, int act) {
  if (oc == 32 && pad == 1 && act == kActSiLU)
    return stem_tc_launch<false, 32, 1, kActSiLU>(x, n, h, w, wgt, bias, out, 0, 0, 0, 0, 0, 0, s);
  if (oc == 40 && pad == 0 && act == kActSiLU)
    return stem_tc_launch<false, 40, 0, kActSiLU>(x, n, h, w, wgt, bias, out, 0, 0, 0, 0, 0, 0, s);
  if (oc == 16 && pad == 1 && act == kActGELU)
    return stem_tc_launch<false, 16, 1, kActGELU>(x, n, h, w, wgt, bias, out, 0, 0, 0, 0, 0, 0, s);
  if (oc == 64 && pad == 1 && act == kActReLU)
    return stem_tc_launch<false, 64, 1, kActReLU>(x, n, h, w, wgt, bias, out, 0, 0, 0, 0, 0, 0, s);
  return cudaErrorInvalidValue;
}

cudaError_t launch_stem_tc_u8(const uint8_t* x, int n, int h, int w, const float* wgt, const float* bias, float sc0,
                              float sc1, float sc2, float sh0, float sh1, float sh2, void* out, cudaStream_t s, int oc,
                              int pad, int act) {
  if (oc == 32 && pad == 1 && act == kActSiLU)
    return stem_tc_launch<true, 32, 1, kActSiLU>(x, n, h, w, wgt, bias, out, sc0, sc1, sc2, sh0, sh1, sh2, s);
  if (oc == 40 && pad == 0 && act == kActSiLU)
    return stem_tc_launch<true, 40, 0, kActSiLU>(x, n, h, w, wgt, bias, out, sc0, sc1, sc2, sh0, sh1, sh2, s);
  if (oc == 16 && pad == 1 && act == kActGELU)
    return stem_tc_launch<true, 16, 1, kActGELU>(x, n, h, w, wgt, bias, out, sc0, sc1, sc2, sh0, sh1, sh2, s);
  if (oc == 64 && pad == 1 && act == kActReLU)
    return stem_tc_launch<true, 64, 1, kActReLU>(x, n, h, w, wgt, bias, out, sc0, sc1, sc2, sh0, sh1, sh2, s);
  return cudaErrorInvalidValue;
}

}  // namespace b200df
