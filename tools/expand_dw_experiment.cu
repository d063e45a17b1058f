// EXPERIMENT, NOT PART OF THE LIBRARY BUILD (measured slower than the two separate kernels; kept as the record of why).
//
// Fused MBConv front for EfficientNet-B0 `blocks.1.0`: 1x1 expand 16 -> 96 (+BN +SiLU) -> depthwise 3x3 stride 2 (+BN +SiLU)
// -> SE pool sums in ONE kernel, so that the network's largest tensor (112 x 112 x 96 per image: 563 MB written + 624 MB
// read per 256 images, 1.19 of the step's 7.14 GB) never reaches HBM.  The expand is a K = 16 contraction: one
// `mma.sync.m16n8k16` per 16 pixels x 8 channels from registers, written as bf16 into the shared-memory tile the depthwise
// loop of dwconv_tma.cu reads.  Correct (unit test against torch and against pwconv_tc + dwconv_tma: max 1 bf16 ulp).
//
// Result on a B200, N = 256 (r02_c, `profiles/r02_c_expand_dw_experiment.txt`): 320 us against 126 + 132 = 258 us for
// the separate kernels (first version 371 us; row-ordered m-tiles + fragment prefetch 320; one strip per tile with three
// CTAs per SM 351).  ncu: 179 M warp instructions, issue slots 54 % busy, MUFU pipe 34 %, DRAM 4 % -- the kernel is
// INSTRUCTION-bound: the bias + tanh + fma + bf16 pack of every expanded element (x 1.07-1.13 for the recomputed halo)
// costs ~110 M issue slots on the CUDA cores, more than the whole separate depthwise kernel, and at 100 % issue the
// kernel would still take 162 us.  In the separate version that epilogue work hides under the expand GEMM's 563 MB HBM
// write; fusing removes the bytes but serialises the two instruction streams.  What would make the fusion pay is an
// activation that costs fewer issue slots per element, not a different tiling.
//
// Build (needs the csrc headers):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Ideepfakedetection_b200/csrc
//                                       -c tools/expand_dw_experiment.cu
// The ExpandDwParams / ExpandDwPlan structs it expects in kernels.h:
//   struct ExpandDwParams { DwTmaParams d; const __nv_bfloat16* x; int H, W, cin; const __nv_bfloat16* ew; const float* eb; };
//   struct ExpandDwPlan { ExpandDwParams p; int grid, block, n; size_t smem; const void* x; const void* out; };
//
// A CTA owns (image, 32-channel block, oth x otw output tile):
//   phase A  every pixel of the input box (tile + halo; out-of-image pixels are the conv's ZERO padding of the EXPANDED
//            tensor, so they are written as 0, not as silu(bias)): 16 input channels -> this block's 32 channels, + bias,
//            SiLU, bf16, into the tile layout of dwconv_tma.cu's stride-2 path ([row][column parity][column pair][32 ch]);
//   phase B  the depthwise loop of dwconv_tma.cu (register weights, sliding fp32x2 accumulators, SiLU, pool partials).
// No TMA / double buffering: the producer is the CTA itself; two CTAs per SM overlap each other's phases.
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

#include <stdlib.h>

#include <algorithm>

namespace b200df {

using namespace tc;

namespace {

constexpr int kCB = 32;   // channels per CTA tile

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 bf2_to_f2(uint32_t raw) {
  return pk2(__uint_as_float(raw << 16), __uint_as_float(raw & 0xffff0000u));
}
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// bf16 pair scaled by 1/2 (exact: a power of two) -- the accumulators then hold h = x/2, silu(x) = h + h tanh(h)
__device__ __forceinline__ uint32_t half_bf16x2(uint32_t raw) {
  const float lo = __uint_as_float(raw << 16) * 0.5f, hi = __uint_as_float(raw & 0xffff0000u) * 0.5f;
  return pack_bf16x2(lo, hi);
}

}  // namespace

// K = 3, S = 2, pad 1 (the box starts one row above and TWO columns left of the tile: an even column, as in dwconv_tma.cu)
template <int TH, int MINB>
__global__ void __launch_bounds__(224, MINB)
expand_dw_s2_kernel(const ExpandDwParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* tile = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int K = 3, S = 2, PAD = 1, PADE = 2, XOFF = PADE - PAD;
  constexpr int RIN = (TH - 1) * S + K;
  const DwTmaParams& d = p.d;
  const int nwarps = blockDim.x >> 5;
  float* spool = reinterpret_cast<float*>(tile + d.stage_stride);   // [warp][channel]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cp = lane & 15, xs = lane >> 4;
  const int g8 = lane >> 2, t4 = lane & 3;
  const int total = d.tiles_x * d.tiles_y * d.ncb * d.n;
  const int row_bytes = d.bw * 128;
  const size_t out_row = (size_t)d.Wo * d.C;
  const int col_pairs = d.otw / 2;
  const int items = col_pairs * (d.oth / TH);

  for (int i = tid; i < nwarps * kCB; i += blockDim.x) spool[i] = 0.f;

  int cb_loaded = -1;
  uint32_t bw_frag[4][2];     // expand weights of this channel block, B fragments of 4 n8 tiles (K = 16), pre-scaled by 1/2
  float eb[4][2];             // expand bias / 2 at columns nt*8 + 2*t4 + {0, 1}
  f32x2 w[K * K];             // depthwise weights / 2 of this thread's channel pair
  f32x2 bias2 = 0ull;
  griddep_wait();
  griddep_launch_dependents();
  __syncthreads();

  for (int t = blockIdx.x; t < total; t += gridDim.x) {
    const int cb = t % d.ncb;
    int r = t / d.ncb;
    const int tx = r % d.tiles_x;
    r /= d.tiles_x;
    const int ty = r % d.tiles_y;
    const int img = r / d.tiles_y;
    const int c0 = cb * kCB + cp * 2;
    const bool c_ok = c0 < d.C;
    if (cb != cb_loaded) {
      cb_loaded = cb;
      // n-tile nt, column j  <->  channel (j / 2) * 8 + nt * 2 + (j % 2): after the four MMAs a thread (columns 2*t4, 2*t4+1
      // of every n-tile) holds the 8 CONSECUTIVE channels t4*8 .. t4*8+7 of its two pixels -> one 16-byte store each
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const int co = cb * kCB + (g8 >> 1) * 8 + nt * 2 + (g8 & 1);
        const bf16* wr = p.ew + (size_t)co * p.cin;
        const bool ok = co < d.C;
        bw_frag[nt][0] = ok ? half_bf16x2(__ldg(reinterpret_cast<const uint32_t*>(wr + 2 * t4))) : 0u;
        bw_frag[nt][1] = ok ? half_bf16x2(__ldg(reinterpret_cast<const uint32_t*>(wr + 2 * t4 + 8))) : 0u;
        const int cc = cb * kCB + t4 * 8 + nt * 2;
        eb[nt][0] = cc < d.C ? 0.5f * __ldg(p.eb + cc) : 0.f;
        eb[nt][1] = cc + 1 < d.C ? 0.5f * __ldg(p.eb + cc + 1) : 0.f;
      }
#pragma unroll
      for (int i = 0; i < K * K; ++i) {
        const float2 v = c_ok ? __ldg(reinterpret_cast<const float2*>(d.wgt + (size_t)i * d.C + c0)) : make_float2(0.f, 0.f);
        w[i] = pk2(v.x * 0.5f, v.y * 0.5f);
      }
      const float2 bb = c_ok ? __ldg(reinterpret_cast<const float2*>(d.bias + c0)) : make_float2(0.f, 0.f);
      bias2 = pk2(bb.x * 0.5f, bb.y * 0.5f);
    }

    // ---------------- phase A: expand the input box into the tile ----------------
    const int iy0 = ty * d.oth * S - PAD, ix0 = tx * d.otw * S - PADE;
    const bf16* ximg = p.x + (size_t)img * p.H * p.W * p.cin;
    // m-tiles follow the tile's rows: (input row, column parity) = rp holds bw column pairs = two m-tiles of 16 pixels (the
    // tail of the second one is masked) -- the row test and all addresses are then per-warp constants plus a lane offset, no
    // divisions.  The A fragment of the NEXT m-tile is loaded before the current one is multiplied (global-load latency is
    // what this phase would otherwise wait for: 4 MMAs and 16 tanh per 4 loads).
    const int n_mt = d.bh * 4;
    auto frag_addr = [&](int mt, const bf16*(&src)[2], bool (&ok)[2], uint32_t (&soff)[2]) {
      const int rp = mt >> 1, hf = mt & 1;
      const int row = rp >> 1, par = rp & 1;
      const int gy = iy0 + row;
      const bool row_ok = gy >= 0 && gy < p.H;
#pragma unroll
      for (int hrow = 0; hrow < 2; ++hrow) {
        const int pair = hf * 16 + g8 + hrow * 8;
        const int gx = ix0 + pair * 2 + par;
        ok[hrow] = row_ok && pair < d.bw && gx >= 0 && gx < p.W;
        src[hrow] = ximg + ((size_t)(row_ok ? gy : 0) * p.W + (ok[hrow] ? gx : 0)) * p.cin;
        soff[hrow] = pair < d.bw ? (uint32_t)(rp * d.bw + pair) * 64u : 0xffffffffu;
      }
    };
    auto frag_load = [&](const bf16* const (&src)[2], const bool (&ok)[2], uint32_t (&a)[4]) {
      a[0] = ok[0] ? __ldg(reinterpret_cast<const uint32_t*>(src[0] + 2 * t4)) : 0u;
      a[1] = ok[1] ? __ldg(reinterpret_cast<const uint32_t*>(src[1] + 2 * t4)) : 0u;
      a[2] = ok[0] ? __ldg(reinterpret_cast<const uint32_t*>(src[0] + 2 * t4 + 8)) : 0u;
      a[3] = ok[1] ? __ldg(reinterpret_cast<const uint32_t*>(src[1] + 2 * t4 + 8)) : 0u;
    };
    uint32_t a[4], soff[2];
    bool ok[2];
    {
      const bf16* src[2];
      if (warp < n_mt) {
        frag_addr(warp, src, ok, soff);
        frag_load(src, ok, a);
      }
    }
    for (int mt = warp; mt < n_mt; mt += nwarps) {
      uint32_t a_n[4] = {0u, 0u, 0u, 0u}, soff_n[2] = {0xffffffffu, 0xffffffffu};
      bool ok_n[2] = {false, false};
      if (mt + nwarps < n_mt) {
        const bf16* src_n[2];
        frag_addr(mt + nwarps, src_n, ok_n, soff_n);
        frag_load(src_n, ok_n, a_n);
      }
      uint32_t v0[4], v1[4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        float acc[4] = {eb[nt][0], eb[nt][1], eb[nt][0], eb[nt][1]};
        mma_16816(acc, a, bw_frag[nt][0], bw_frag[nt][1]);
        // pixel g8: acc[0..1], pixel g8 + 8: acc[2..3]; channels t4*8 + nt*2 + {0, 1}
        v0[nt] = pack_bf16x2(fmaf(acc[0], tanh_fast(acc[0]), acc[0]), fmaf(acc[1], tanh_fast(acc[1]), acc[1]));
        v1[nt] = pack_bf16x2(fmaf(acc[2], tanh_fast(acc[2]), acc[2]), fmaf(acc[3], tanh_fast(acc[3]), acc[3]));
      }
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);   // outside the image: the depthwise conv's zero padding
      if (soff[0] != 0xffffffffu)
        *reinterpret_cast<uint4*>(tile + soff[0] + t4 * 16) = ok[0] ? make_uint4(v0[0], v0[1], v0[2], v0[3]) : z;
      if (soff[1] != 0xffffffffu)
        *reinterpret_cast<uint4*>(tile + soff[1] + t4 * 16) = ok[1] ? make_uint4(v1[0], v1[1], v1[2], v1[3]) : z;
#pragma unroll
      for (int e = 0; e < 4; ++e) a[e] = a_n[e];
      soff[0] = soff_n[0]; soff[1] = soff_n[1];
      ok[0] = ok_n[0]; ok[1] = ok_n[1];
    }
    __syncthreads();

    // ---------------- phase B: depthwise 3x3 stride 2 on the tile (the loop of dwconv_tma.cu, NC = 1) ----------------
    f32x2 ps = 0ull;
    for (int item = warp; item < items; item += nwarps) {
      const int st = (item * d.inv_cp) >> 16;
      const int x = (item - st * col_pairs) * 2 + xs;
      const int y0 = st * TH;
      f32x2 acc[TH];
#pragma unroll
      for (int j = 0; j < TH; ++j) acc[j] = bias2;
      const uint8_t* base = tile + (y0 * S * row_bytes + x * 64 + cp * 4);
#pragma unroll
      for (int rr = 0; rr < RIN; ++rr) {
        f32x2 v[K];
#pragma unroll
        for (int kx = 0; kx < K; ++kx) {
          const int ic = kx + XOFF;
          v[kx] = bf2_to_f2(*reinterpret_cast<const uint32_t*>(base + rr * row_bytes + (ic & 1) * (d.bw * 64) + (ic >> 1) * 64));
        }
#pragma unroll
        for (int ky = 0; ky < K; ++ky) {
          if ((rr - ky) >= 0 && (rr - ky) % S == 0 && (rr - ky) / S < TH) {
            const int j = (rr - ky) / S;
#pragma unroll
            for (int kx = 0; kx < K; ++kx) acc[j] = fma2(v[kx], w[ky * K + kx], acc[j]);
          }
        }
      }
      const int oy0 = ty * d.oth + y0;
      const int ox = tx * d.otw + x;
      if (c_ok && ox < d.Wo) {
        bf16* op = reinterpret_cast<bf16*>(d.out) + (((size_t)img * d.Ho + oy0) * d.Wo + ox) * d.C + c0;
#pragma unroll
        for (int j = 0; j < TH; ++j) {
          float h0, h1;
          upk2(acc[j], h0, h1);
          const f32x2 f = fma2(acc[j], pk2(tanh_fast(h0), tanh_fast(h1)), acc[j]);
          float f0, f1;
          upk2(f, f0, f1);
          if (oy0 + j < d.Ho) {
            *reinterpret_cast<uint32_t*>(op) = pack_bf16x2(f0, f1);
            ps = add2(ps, f);
          }
          op += out_row;
        }
      }
    }
    if (d.pool != nullptr) {
      float s0, s1;
      upk2(ps, s0, s1);
      s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
      s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
      if (xs == 0) *reinterpret_cast<float2*>(spool + warp * kCB + cp * 2) = make_float2(s0, s1);
    }
    __syncthreads();   // the tile is free again; the pool slots of this tile are complete
    if (d.pool != nullptr && tid < kCB) {
      float v = 0.f;
      for (int wq = 0; wq < nwarps; ++wq) v += spool[wq * kCB + tid];   // fixed order -> deterministic
      const int c = cb * kCB + tid;
      if (c < d.C) pool_add(d.pool + (size_t)img * d.C + c, v);
    }
    // (the next tile's phase A does not touch spool; its phase B rewrites every slot before the next barrier pair)
  }
}

int plan_expand_dw(ExpandDwPlan& plan, const void* x, int n, int h, int w, int cin, int c, int k, int stride, const void* ew,
                   const float* eb, const float* wgt, const float* bias, void* out, pool_t* pool) {
  // the one shape this kernel is built for (B0 `blocks.1.0`): K = 16 expand, 3x3 stride-2 pad-1 depthwise, even sizes
  if (n <= 0 || cin != 16 || k != 3 || stride != 2 || (h & 1) || (w & 1) || c % 8) return -1;
  const int Ho = h / 2, Wo = w / 2;
  constexpr int TH = 7;
  if (Ho % TH) return -1;
  // tile: as large as two CTAs per SM allow (halo recompute 1.07 at 14 x 28), warps = items / 4
  int otw = 28, m = 2;
  if (const char* e = getenv("B200DF_XDW_M")) m = atoi(e) == 1 ? 1 : 2;   // sweep aid: one strip per tile (3 CTAs per SM)
  while (otw > 2 && Wo % otw) otw -= 2;
  if ((Ho / TH) % m) m = 1;
  ExpandDwParams& p = plan.p;
  DwTmaParams& d = p.d;
  d.C = c; d.Ho = Ho; d.Wo = Wo;
  d.otw = otw; d.oth = m * TH;
  d.tiles_x = Wo / otw; d.tiles_y = Ho / d.oth;
  d.ncb = (c + kCB - 1) / kCB;
  d.n = n;
  d.bh = (d.oth - 1) * 2 + 3;
  d.bw = ((otw - 1) * 2 + 3 + 1 + 1) / 2;      // column pairs, incl. the even-start column
  d.gimg = 1;
  d.img_bytes = d.bw * 2 * d.bh * 64;
  const int col_groups = otw / 2;
  d.inv_cp = (65536 + col_groups - 1) / col_groups;
  d.inv_per_img = 0;
  d.box_bytes = (uint32_t)d.img_bytes;
  d.stage_stride = (int)((d.box_bytes + 1023) & ~1023u);
  d.cb_fixed = 0;
  d.wgt = wgt; d.bias = bias; d.out = out; d.pool = pool;
  p.x = reinterpret_cast<const bf16*>(x);
  p.H = h; p.W = w; p.cin = cin;
  p.ew = reinterpret_cast<const bf16*>(ew);
  p.eb = eb;
  plan.block = 224;
  plan.smem = 1024 + (size_t)d.stage_stride + (size_t)(plan.block / 32) * kCB * sizeof(float);
  if (plan.smem > 113 * 1024) return -1;
  const long long total = (long long)d.tiles_x * d.tiles_y * d.ncb * n;
  long long grid = std::min<long long>(total, (long long)kNumSMs * (plan.smem <= 74 * 1024 ? 3 : 2));
  if (grid >= d.ncb) grid = grid / d.ncb * d.ncb;   // a CTA then keeps one channel block: weights are loaded once
  plan.grid = (int)grid;
  plan.x = x;
  plan.out = out;
  plan.n = n;
  return 0;
}

cudaError_t launch_expand_dw(const ExpandDwPlan& plan, cudaStream_t s) {
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(expand_dw_s2_kernel<7, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(expand_dw_s2_kernel<7, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 74 * 1024);
    if (e != cudaSuccess) return e;
  }
  count_launch();
  if (plan.smem <= 74 * 1024)
    return launch_pdl(expand_dw_s2_kernel<7, 3>, dim3(plan.grid), dim3(plan.block), plan.smem, s, plan.p);
  return launch_pdl(expand_dw_s2_kernel<7, 2>, dim3(plan.grid), dim3(plan.block), plan.smem, s, plan.p);
}

}  // namespace b200df
