// Depthwise kxk conv (k = 3|5, stride 1|2, pad (k-1)/2) + folded-BN bias + SiLU + SE pool sums,
// bf16 NHWC, for sm_100a: TMA halo staging + register-resident weights + sliding accumulators.
// Replaces timm `conv_dw` + `bn1/bn2`(+SiLU) and the `se` mean_hw, the
// depthwise leg of every MBConv inside the reference's `model(images)`
// (/root/reference/orchestration/orchestrator.py:590).
//
// Shape of the problem: 12.2 MB of HBM traffic and 34.5 M MAC per image over 16 layers -- HBM-bound,
// but only if the instruction stream stays near 25-40 issue slots per output element, so:
//   * a CTA owns (image, 32-channel block, OTW x OTH output tile); the input tile INCLUDING its halo
//     arrives with ONE cp.async.bulk.tensor (TMA) per tile: out-of-image halo pixels and channels
//     past C are zero-filled by the TMA unit, so the compute loop has no bounds checks at all.
//     Stride-2 layers use a 5-D tensor map that splits columns into (pair, parity), which lands even
//     and odd input columns in separate contiguous planes -> bank-conflict-free reads at stride 2;
//   * a thread owns 2 channels x 1 output column x TH output rows; the k*k*2 folded weights live in
//     registers for the whole tile; input rows stream through once (k LDS.32 per row) and every loaded
//     value feeds up to k*k FMAs on the TH sliding fp32 accumulators;
//   * 16 lanes cover one pixel's 32 channels (64 B), the two half-warps take adjacent columns, so each
//     warp-wide LDS.32 reads 128 contiguous bytes and each store writes two full 64-B segments;
//   * tiles are double-buffered: the TMA for tile i+1 is in flight while tile i is computed;
//   * SE pooling, order-independent (bit-reproducible): per-thread sums -> shuffle -> a shared slot owned by
//     one thread per (warp, image, channel) -> fixed-order sum over warps -> ONE Q24 fixed-point integer
//     atomic per (tile, channel) in global memory.
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

namespace b200df {

using namespace tc;

constexpr int kDwCB = 32;       // channels per CTA tile (64 B of every pixel)
constexpr int kDwMaxGroup = 8;  // images per tile for the small-spatial layers


// ---- packed fp32x2 helpers (sm_100: FFMA2 / FADD2 / FMUL2 -- two fp32 lanes per issue slot) ----
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
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
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// two bf16 channels (one 32-bit word) -> packed fp32 pair
__device__ __forceinline__ f32x2 bf16x2_to_f32x2(uint32_t raw) {
  return pack2(__uint_as_float(raw << 16), __uint_as_float(raw & 0xffff0000u));
}

struct DwTile { int cb, tx, ty, ng; };

// (Three CTAs per SM for the 3x3 single-column variants -- 80 registers, <= 74 KB plans -- measured slower.)
// PL: zero rows/columns before the image.  (K-1)/2 everywhere except the TF-"same" 3x3 stride-2 layers of the
// efficientnet_pytorch layout (pad (0, 1): PL = 0, the box then starts on the even column 2*ox itself).
// ACT: activation after the folded-BN bias (common.cuh Act): SiLU (EfficientNet), GELU (EfficientFormerV2 `mlp.mid`),
// none (EfficientFormerV2 `stride_conv` / `v_local` / `q.local`, FasterViT `global_tokenizer.pos_embed`).
template <int K, int S, int TH, int NC, int PL = (K - 1) / 2, int ACT = kActSiLU>
__global__ void __launch_bounds__(256, 2)
dwconv_tma_kernel(const __grid_constant__ CUtensorMap tmap, const DwTmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-align by OFFSET (not by pointer round-trip) so the compiler keeps the shared address space -> LDS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full_bar[2];
  const int nwarps_ = blockDim.x >> 5;
  // pool slots [stage][warp][image in group][channel], after the two operand stages
  float* spool = reinterpret_cast<float*>(smem + 2 * p.stage_stride);
  const int spool_stage = nwarps_ * p.gimg * kDwCB;

  constexpr int PAD = PL;
  constexpr int PADE = (S == 2) ? (PL == 0 ? 0 : 2) : PAD;   // stride 2: the box starts on an even column
  constexpr int XOFF = PADE - PAD;           // extra left columns inside the box
  constexpr int RIN = (TH - 1) * S + K;      // input rows one strip touches

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int cp = lane & 15;   // channel pair inside the 32-channel block
  const int xs = lane >> 4;   // which of the warp's two adjacent output columns
  const int ngroups = (p.n + p.gimg - 1) / p.gimg;
  const int total = p.tiles_x * p.tiles_y * p.ncb * ngroups;

  if (tid == 0) {
    tma_prefetch_desc(&tmap);
    mbar_init(&full_bar[0], 1);
    mbar_init(&full_bar[1], 1);
    fence_barrier_init();
  }
  for (int i = tid; i < 2 * spool_stage; i += blockDim.x) spool[i] = 0.f;
  __syncthreads();

  // tile index -> (cb fastest, then tile x, tile y, image group): neighbouring CTAs read the same pixels'
  // other channel blocks at the same time (full DRAM lines, halos shared through L2).  Decoded once,
  // then advanced by gridDim.x in mixed radix (no divisions in the tile loop).
  auto decode = [&](int t) {
    DwTile d;
    d.cb = t % p.ncb;
    int r = t / p.ncb;
    d.tx = r % p.tiles_x;
    r /= p.tiles_x;
    d.ty = r % p.tiles_y;
    d.ng = r / p.tiles_y;
    return d;
  };
  const DwTile step = decode((int)gridDim.x);
  auto advance = [&](DwTile d) {
    d.cb += step.cb;
    int c = d.cb >= p.ncb;
    d.cb -= c ? p.ncb : 0;
    d.tx += step.tx + c;
    c = d.tx >= p.tiles_x;
    d.tx -= c ? p.tiles_x : 0;
    d.ty += step.ty + c;
    c = d.ty >= p.tiles_y;
    d.ty -= c ? p.tiles_y : 0;
    d.ng += step.ng + c;
    return d;
  };
  auto issue = [&](const DwTile& d, int stage) {
    mbar_arrive_expect_tx(&full_bar[stage], p.box_bytes);
    uint8_t* dst = smem + stage * p.stage_stride;
    if constexpr (S == 1) {
      tma_load_4d(dst, &tmap, d.cb * kDwCB, d.tx * p.otw - PAD, d.ty * p.oth - PAD, d.ng * p.gimg, &full_bar[stage]);
    } else {
      // dims: (c, column pair, column parity, row, image)
      tma_load_5d(dst, &tmap, d.cb * kDwCB, d.tx * p.otw - PADE / 2, 0, d.ty * p.oth * 2 - PAD, d.ng * p.gimg,
                  &full_bar[stage]);
    }
  };

  DwTile cur = decode((int)blockIdx.x);

  // a warp covers 2*NC output columns: half-warp xs takes columns x and (NC == 2) x + 2, so that the two
  // half-warps still read adjacent 64-B pixels (bank-conflict free) while a thread's two columns share K - 2
  // of the K + 2 inputs it loads per row
  constexpr int KV = K + 2 * (NC - 1);
  const int col_pairs = p.otw / (2 * NC);
  const int per_img = col_pairs * (p.oth / TH);
  const int items = per_img * p.gimg;
  const int row_bytes = (S == 1) ? p.bw * 64 : p.bw * 128;  // S=2: two parity planes of bw pairs per row
  const size_t out_row = (size_t)p.Wo * p.C;                // elements between output rows

  // folded weights + bias for this thread's channel pair, pre-scaled by 1/2 so the accumulators hold
  // h = x/2 directly (silu(x) = h + h*tanh(h)).  When gridDim.x is a multiple of ncb (the planner
  // arranges that) every tile of this CTA has the same channel block and they are loaded once.
  int c0 = cur.cb * kDwCB + cp * 2;
  bool c_ok = c0 < p.C;
  f32x2 w[K * K];
  f32x2 bias2 = 0ull;
  auto load_weights = [&]() {
    constexpr float kPre = ACT == kActNone ? 1.0f : 0.5f;   // SiLU / GELU work on h = x/2
    const f32x2 half2 = pack2(kPre, kPre);
#pragma unroll
    for (int i = 0; i < K * K; ++i) {
      float2 v = c_ok ? __ldg(reinterpret_cast<const float2*>(p.wgt + (size_t)i * p.C + c0)) : make_float2(0.f, 0.f);
      w[i] = mul2(pack2(v.x, v.y), half2);
    }
    float2 bb = c_ok ? __ldg(reinterpret_cast<const float2*>(p.bias + c0)) : make_float2(0.f, 0.f);
    bias2 = mul2(pack2(bb.x, bb.y), half2);
  };
  load_weights();
  griddep_wait();  // weights, barriers and pool slots are ready; activations may only be touched from here on
  griddep_launch_dependents();
  if (tid == 0 && (int)blockIdx.x < total) issue(cur, 0);

  int it = 0;
  for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
    const int stage = it & 1;
    const DwTile nxt = advance(cur);
    if (tid == 0 && t + (int)gridDim.x < total) issue(nxt, stage ^ 1);

    if (!p.cb_fixed) {  // cb changes from tile to tile: reload this channel pair's weights
      c0 = cur.cb * kDwCB + cp * 2;
      c_ok = c0 < p.C;
      load_weights();
    }

    mbar_wait(&full_bar[stage], (uint32_t)((it >> 1) & 1));
    const uint8_t* tile = smem + stage * p.stage_stride;

    for (int item = warp; item < items; item += nwarps) {
      const int g = (item * p.inv_per_img) >> 16;           // image inside the group
      if (cur.ng * p.gimg + g >= p.n) continue;             // ragged last group (the plan never depends on n,
                                                            // so a sample's result never depends on its batch)
      const int rem = item - g * per_img;
      const int st = (rem * p.inv_cp) >> 16;                // row strip
      const int x = (rem - st * col_pairs) * (2 * NC) + xs; // first output column inside the tile
      const int y0 = st * TH;                               // first output row inside the tile
      f32x2 acc[NC][TH];
#pragma unroll
      for (int q = 0; q < NC; ++q)
#pragma unroll
        for (int j = 0; j < TH; ++j) acc[q][j] = bias2;

      // (input row 0 of the strip, this thread's first column, its channel pair)
      const uint8_t* base = tile + (g * p.img_bytes + y0 * S * row_bytes + x * 64 + cp * 4);
#pragma unroll
      for (int rr = 0; rr < RIN; ++rr) {
        f32x2 v[KV];
#pragma unroll
        for (int kx = 0; kx < KV; ++kx) {
          uint32_t raw;
          if constexpr (S == 1) {
            raw = *reinterpret_cast<const uint32_t*>(base + rr * row_bytes + kx * 64);
          } else {
            const int ic = kx + XOFF;  // column offset inside the box relative to 2*x
            raw = *reinterpret_cast<const uint32_t*>(base + rr * row_bytes + (ic & 1) * (p.bw * 64) + (ic >> 1) * 64);
          }
          v[kx] = bf16x2_to_f32x2(raw);
        }
#pragma unroll
        for (int ky = 0; ky < K; ++ky) {
          // input row rr feeds output row (rr - ky) / S when that is an integer in [0, TH)
          if ((rr - ky) >= 0 && (rr - ky) % S == 0 && (rr - ky) / S < TH) {
            const int j = (rr - ky) / S;
#pragma unroll
            for (int q = 0; q < NC; ++q)
#pragma unroll
              for (int kx = 0; kx < K; ++kx) acc[q][j] = fma2(v[kx + 2 * q], w[ky * K + kx], acc[q][j]);
          }
        }
      }
      // epilogue: SiLU (h + h*tanh h), bf16 pack, store, pool partials
      const int oy0 = cur.ty * p.oth + y0;
      const int img = cur.ng * p.gimg + g;
      f32x2 ps = 0ull;
#pragma unroll
      for (int q = 0; q < NC; ++q) {
        const int ox = cur.tx * p.otw + x + 2 * q;
        if (c_ok && ox < p.Wo && img < p.n) {
          bf16* op = reinterpret_cast<bf16*>(p.out) + (((size_t)img * p.Ho + oy0) * p.Wo + ox) * p.C + c0;
          const bool rows_full = oy0 + TH <= p.Ho;
#pragma unroll
          for (int j = 0; j < TH; ++j) {
            float h0, h1;
            unpack2(acc[q][j], h0, h1);
            f32x2 f;
            if constexpr (ACT == kActSiLU) {
              f = fma2(acc[q][j], pack2(tanh_fast(h0), tanh_fast(h1)), acc[q][j]);
            } else if constexpr (ACT == kActGELU) {  // h + h tanh(h (c1 + c3 h^2)), common.cuh gelu_fast
              const f32x2 u = mul2(acc[q][j], fma2(mul2(acc[q][j], acc[q][j]), pack2(0.2854191951f, 0.2854191951f),
                                                   pack2(1.5957691216f, 1.5957691216f)));
              float u0, u1;
              unpack2(u, u0, u1);
              f = fma2(acc[q][j], pack2(tanh_fast(u0), tanh_fast(u1)), acc[q][j]);
            } else {
              f = acc[q][j];
            }
            float f0, f1;
            unpack2(f, f0, f1);
            if (rows_full || oy0 + j < p.Ho) {
              *reinterpret_cast<uint32_t*>(op) = pack_bf16x2(f0, f1);
              ps = add2(ps, f);
            }
            op += out_row;
          }
        }
      }
      if (p.pool != nullptr) {
        float s0, s1;
        unpack2(ps, s0, s1);
        s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
        s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
        if (xs == 0) {  // this thread is the only writer of its (warp, *, cp) slots: plain read-modify-write
          float2* slot = reinterpret_cast<float2*>(spool + stage * spool_stage + (warp * p.gimg + g) * kDwCB + cp * 2);
          float2 o = *slot;
          o.x += s0;
          o.y += s1;
          *slot = o;
        }
      }
    }
    __syncthreads();  // everyone is done with smem stage `stage`; spool[stage] is complete
    if (p.pool != nullptr) {
      for (int i = tid; i < p.gimg * kDwCB; i += blockDim.x) {
        const int g = i >> 5, cc = i & 31;
        float* sl = spool + stage * spool_stage + g * kDwCB + cc;
        float v = 0.f;
        for (int wq = 0; wq < nwarps_; ++wq) {  // fixed order -> deterministic
          v += sl[wq * p.gimg * kDwCB];
          sl[wq * p.gimg * kDwCB] = 0.f;
        }
        const int c = cur.cb * kDwCB + cc;
        const int img = cur.ng * p.gimg + g;
        if (c < p.C && img < p.n) pool_add(p.pool + (size_t)img * p.C + c, v);
      }
    }
    cur = nxt;
  }
}

// ---- host side ------------------------------------------------------------------------------------

typedef CUresult (*PFN_encodeTiledNd)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_tmap_nd_bf16(void* tmap_out, const void* base, int rank, const unsigned long long* dims,
                        const unsigned long long* strides_bytes, const unsigned* box, int swizzle,
                        const unsigned* elem_strides) {
  static PFN_encodeTiledNd enc = nullptr;
  if (!enc) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return -1;
    enc = reinterpret_cast<PFN_encodeTiledNd>(fn);
  }
  if (rank < 2 || rank > 5 || (reinterpret_cast<uintptr_t>(base) & 15)) return -2;
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = elem_strides ? elem_strides[i] : 1;
    if (i > 0) {
      gstr[i - 1] = strides_bytes[i];
      if (strides_bytes[i] % 16) return -2;
    }
  }
  CUtensorMapSwizzle sw = swizzle == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                          : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(reinterpret_cast<CUtensorMap*>(tmap_out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                   const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

namespace {

template <int K, int S, int TH, int NC, int PL = (K - 1) / 2, int ACT = kActSiLU>
cudaError_t dw_tma_launch(const DwPlan& pl, cudaStream_t s) {
  auto kern = dwconv_tma_kernel<K, S, TH, NC, PL, ACT>;
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
  }
  count_launch();
  return launch_pdl(kern, dim3(pl.grid), dim3(pl.block), pl.smem, s, *reinterpret_cast<const CUtensorMap*>(pl.tmap), pl.p);
}

template <int K, int S, int TH, int NC, int PL = (K - 1) / 2>
int dw_tma_occupancy(int block, size_t smem) {
  auto kern = dwconv_tma_kernel<K, S, TH, NC, PL>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) return 1;
  int nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, block, smem) != cudaSuccess || nb < 1) nb = 1;
  return nb;
}

int dw_tma_occupancy_dispatch(int k, int stride, int th, int nc, int pad_lo, int block, size_t smem) {
  if (k == 3 && stride == 2 && nc == 1 && pad_lo == 0)
    return th == 8 ? dw_tma_occupancy<3, 2, 8, 1, 0>(block, smem) : dw_tma_occupancy<3, 2, 7, 1, 0>(block, smem);
#define DW_OCC(K_, S_, TH_, NC_) \
  if (k == K_ && stride == S_ && th == TH_ && nc == NC_) return dw_tma_occupancy<K_, S_, TH_, NC_>(block, smem)
  DW_OCC(3, 1, 8, 1); DW_OCC(3, 1, 7, 1); DW_OCC(3, 2, 8, 1); DW_OCC(3, 2, 7, 1);
  DW_OCC(5, 1, 8, 1); DW_OCC(5, 1, 7, 1); DW_OCC(5, 2, 8, 1); DW_OCC(5, 2, 7, 1);
  DW_OCC(3, 1, 8, 2); DW_OCC(3, 1, 7, 2); DW_OCC(5, 1, 7, 2);
#undef DW_OCC
  return 1;
}

constexpr size_t kDwSmemLimitDefault = 110 * 1024;  // per CTA (2 operand stages + pool slots) -> 2 CTAs per SM
static size_t dw_smem_limit() {  // B200DF_DW_SMEM_KB: sweep aid (e.g. 74 -> plans that allow 3 CTAs per SM)
  static size_t v = 0;
  if (!v) {
    const char* e = getenv("B200DF_DW_SMEM_KB");
    v = e ? (size_t)atoi(e) * 1024 : kDwSmemLimitDefault;
  }
  return v;
}
#define kDwSmemLimit dw_smem_limit()

}  // namespace

int plan_dwconv_tma(DwPlan& plan, const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                    const float* bias, void* out, pool_t* pool, int pad_lo, int act) {
  if (n <= 0 || c % 8 || (k != 3 && k != 5) || (stride != 1 && stride != 2)) return -1;
  if (act != kActSiLU && !((act == kActNone || act == kActGELU) && k == 3 && pad_lo < 0)) return -1;
  plan.act = act;
  if (stride == 2 && (w & 1)) return -1;
  const int pad = pad_lo < 0 ? (k - 1) / 2 : pad_lo;
  // an explicit pad-before other than (k-1)/2 exists only for the TF-"same" 3x3 stride-2 layers: pad (0, 1)
  if (pad != (k - 1) / 2 && !(pad == 0 && k == 3 && stride == 2 && !(h & 1))) return -1;
  const int pade = stride == 2 ? (pad == 0 ? 0 : 2) : pad;  // the stride-2 box starts on an even column
  const int Ho = (h - 1) / stride + 1, Wo = (w - 1) / stride + 1;  // total padding is k - 1 either way
  // choose the strip height TH and the output tile: even width wasting at most one column, height a multiple
  // of TH dividing the strip count; maximise useful/loaded pixels subject to the per-stage smem limit, prefer
  // >= 8 warp items.  A few EfficientNet-B0 shapes have measured-better plans (tools/dw_sweep.py on a B200);
  // B200DF_DW_PLAN="th,otw,m,g" pins the choice for such sweeps.
  int f_th = 0, f_otw = 0, f_m = 0, f_g = 0, nc = 1;
  if (const char* e = getenv("B200DF_DW_PLAN")) {
    const int got = sscanf(e, "%d,%d,%d,%d,%d", &f_th, &f_otw, &f_m, &f_g, &nc);
    if (got < 4) f_th = f_otw = f_m = f_g = 0;
    if (got < 5) nc = 1;
  } else {
    // measured on a B200 at N = 256 (tools/dw_sweep.py); nc = 2: two output columns per thread
    struct Tuned { int h, k, stride, th, otw, m, g, nc; };
    static const Tuned tuned[] = {
        {112, 3, 1, 7, 16, 4, 1, 2},  // 3x3 s1 112^2, C=32  : 0.110 -> 0.100 ms
        {112, 3, 2, 7, 14, 1, 1, 1},  // 3x3 s2 112->56, C=96: 0.130 -> 0.125 ms
        {56, 3, 1, 7, 8, 8, 1, 2},    // 3x3 s1 56^2, C=144  : 0.139 -> 0.125 ms
        {56, 5, 2, 7, 4, 4, 1, 1},    // 5x5 s2 56->28       : 0.094 -> 0.082 ms
        {28, 5, 1, 7, 4, 4, 1, 2},    // 5x5 s1 28^2, C=240  : 0.106 -> 0.098 ms
        {14, 5, 2, 7, 8, 1, 1, 1},    // 5x5 s2 14->7        : 0.031 -> 0.030 ms
        {7, 5, 1, 7, 8, 1, 4, 2},     // 5x5 s1 7^2, C=1152  : 0.040 -> 0.035 ms
        {7, 3, 1, 7, 8, 1, 4, 2},     // 3x3 s1 7^2, C=1152  : 0.026 -> 0.022 ms
    };
    for (const Tuned& t : tuned)
      if (t.h == h && h == w && t.k == k && t.stride == stride) {
        f_th = t.th; f_otw = t.otw; f_m = t.m; f_g = t.g; nc = t.nc;
      }
    // EfficientNet-B3's wider 7^2 layers (C = 1392 / 2304) prefer other image-group sizes (DW_SWEEP_MODEL=b3
    // tools/dw_sweep.py, profiles/r01_h_dw_sweep_b3.txt): 5x5 g = 2 (0.048 -> 0.043 ms), 3x3 C=1392 g = 1, C=2304 g = 8
    if (h == 7 && w == 7 && stride == 1 && c > 1152 && f_th) f_g = k == 5 ? 2 : (c >= 2304 ? 8 : 1);
  }
  if (nc != 1 && (nc != 2 || stride != 1 || (k == 5 && f_th == 8))) return -1;
  double best_score = -1.0;
  int b_otw = 0, b_m = 0, b_bw = 0, b_bh = 0, b_g = 1, th = 8, strips_total = 0;
  for (int th_c = 8; th_c >= 7; --th_c) {
    if (f_th ? th_c != f_th : th_c != ((Ho % 8 == 0) ? 8 : 7)) continue;
    const int strips_c = (Ho + th_c - 1) / th_c;
    for (int otw = 2 * nc; otw <= 32; otw += 2 * nc) {
      if (f_otw && otw != f_otw) continue;
      const int tiles_x = (Wo + otw - 1) / otw;
      if (!f_otw && tiles_x * otw - Wo > 1) continue;
      for (int m = 1; m <= strips_c; ++m) {
        if (strips_c % m) continue;
        if (f_m && m != f_m) continue;
        const int oth = m * th_c;
        int bw, bh = (oth - 1) * stride + k;
        size_t bytes;
        if (stride == 1) {
          bw = otw + k - 1;
          bytes = (size_t)bw * bh * 64;
        } else {
          const int cols = (otw - 1) * 2 + k + (pade - pad);  // incl. the even-start offset
          bw = (cols + 1) / 2;                              // column pairs
          bytes = (size_t)bw * 2 * bh * 64;
        }
        if (bw > 256 || bh > 256) continue;
        const bool whole = tiles_x == 1 && m == strips_c;  // the tile is the whole image
        for (int g = 1; g <= (whole ? kDwMaxGroup : 1); g *= 2) {
          if (f_g && g != f_g) continue;
          if (2 * ((bytes * g + 1023) & ~size_t(1023)) + 2 * 8 * (size_t)g * kDwCB * 4 + 1024 > kDwSmemLimit) break;
          const int items = (otw / (2 * nc)) * m * g;
          double eff = (double)std::min(otw, Wo) * std::min(oth, Ho) * stride * stride / ((double)bytes / 64.0);
          double score = eff * std::min(1.0, 0.5 + items / 32.0);
          if (score > best_score) {
            best_score = score;
            b_otw = otw; b_m = m; b_bw = bw; b_bh = bh; b_g = g;
            th = th_c; strips_total = strips_c;
          }
        }
      }
    }
  }
  if (best_score < 0) return -1;
  DwTmaParams& p = plan.p;
  p.C = c; p.Ho = Ho; p.Wo = Wo;
  p.otw = b_otw; p.oth = b_m * th;
  p.tiles_x = (Wo + b_otw - 1) / b_otw;
  p.tiles_y = strips_total / b_m;
  p.ncb = (c + kDwCB - 1) / kDwCB;
  p.n = n;
  p.bw = b_bw; p.bh = b_bh;
  p.gimg = b_g;
  p.img_bytes = (int)((stride == 1 ? (size_t)b_bw : (size_t)b_bw * 2) * b_bh * 64);
  const int col_groups = b_otw / (2 * nc);
  p.inv_cp = (65536 + col_groups - 1) / col_groups;
  p.inv_per_img = (65536 + col_groups * b_m - 1) / (col_groups * b_m);
  p.box_bytes = (uint32_t)p.img_bytes * (uint32_t)b_g;
  p.stage_stride = (int)((p.box_bytes + 1023) & ~1023u);
  p.wgt = wgt; p.bias = bias; p.out = out; p.pool = pool;
  int rc;
  if (stride == 1) {
    unsigned long long dims[4] = {(unsigned long long)c, (unsigned long long)w, (unsigned long long)h, (unsigned long long)n};
    unsigned long long str[4] = {2, (unsigned long long)c * 2, (unsigned long long)w * c * 2, (unsigned long long)h * w * c * 2};
    unsigned box[4] = {kDwCB, (unsigned)b_bw, (unsigned)b_bh, (unsigned)b_g};
    rc = encode_tmap_nd_bf16(plan.tmap, x, 4, dims, str, box, 0);
  } else {
    unsigned long long dims[5] = {(unsigned long long)c, (unsigned long long)(w / 2), 2, (unsigned long long)h, (unsigned long long)n};
    unsigned long long str[5] = {2, (unsigned long long)c * 4, (unsigned long long)c * 2, (unsigned long long)w * c * 2,
                                 (unsigned long long)h * w * c * 2};
    unsigned box[5] = {kDwCB, (unsigned)b_bw, 2, (unsigned)b_bh, (unsigned)b_g};
    rc = encode_tmap_nd_bf16(plan.tmap, x, 5, dims, str, box, 0);
  }
  if (rc) return rc;
  plan.k = k; plan.stride = stride; plan.th = th; plan.nc = nc; plan.pad_lo = pad;
  const int items = col_groups * b_m * b_g;
  plan.block = items >= 8 ? 256 : (items >= 4 ? 128 : 64);
  plan.smem = 1024 + 2 * (size_t)p.stage_stride + 2 * (size_t)(plan.block / 32) * b_g * kDwCB * sizeof(float);
  const long long total = (long long)p.tiles_x * p.tiles_y * p.ncb * ((n + b_g - 1) / b_g);
  const int per_sm = dw_tma_occupancy_dispatch(k, stride, th, nc, pad, plan.block, plan.smem);
  long long grid = std::min<long long>(total, (long long)kNumSMs * per_sm);
  p.cb_fixed = 0;
  if (grid >= p.ncb) {  // a grid that is a multiple of ncb keeps every CTA on one channel block
    grid = grid / p.ncb * p.ncb;
    p.cb_fixed = 1;
  }
  plan.grid = (int)grid;
  plan.x = x;
  return 0;
}

cudaError_t launch_dwconv_tma(const DwPlan& pl, cudaStream_t s) {
  if (pl.act != kActSiLU) {  // 3x3 only (plan_dwconv_tma refuses the rest)
#define DW_CASE_ACT(S_, TH_, NC_)                                                                                       \
  if (pl.k == 3 && pl.stride == S_ && pl.th == TH_ && pl.nc == NC_)                                                      \
    return pl.act == kActGELU ? dw_tma_launch<3, S_, TH_, NC_, 1, kActGELU>(pl, s)                                      \
                              : dw_tma_launch<3, S_, TH_, NC_, 1, kActNone>(pl, s)
    DW_CASE_ACT(1, 8, 1); DW_CASE_ACT(1, 7, 1); DW_CASE_ACT(2, 8, 1); DW_CASE_ACT(2, 7, 1);
    DW_CASE_ACT(1, 8, 2); DW_CASE_ACT(1, 7, 2);
#undef DW_CASE_ACT
    return cudaErrorInvalidValue;
  }
  if (pl.k == 3 && pl.stride == 2 && pl.nc == 1 && pl.pad_lo == 0)
    return pl.th == 8 ? dw_tma_launch<3, 2, 8, 1, 0>(pl, s) : dw_tma_launch<3, 2, 7, 1, 0>(pl, s);
#define DW_CASE(K_, S_, TH_, NC_) \
  if (pl.k == K_ && pl.stride == S_ && pl.th == TH_ && pl.nc == NC_) return dw_tma_launch<K_, S_, TH_, NC_>(pl, s)
  DW_CASE(3, 1, 8, 1); DW_CASE(3, 1, 7, 1); DW_CASE(3, 2, 8, 1); DW_CASE(3, 2, 7, 1);
  DW_CASE(5, 1, 8, 1); DW_CASE(5, 1, 7, 1); DW_CASE(5, 2, 8, 1); DW_CASE(5, 2, 7, 1);
  DW_CASE(3, 1, 8, 2); DW_CASE(3, 1, 7, 2); DW_CASE(5, 1, 7, 2);
#undef DW_CASE
  return cudaErrorInvalidValue;
}

}  // namespace b200df
