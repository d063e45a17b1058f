// Dense 3x3 conv (pad 1, stride 1|2) as an implicit GEMM on the 5th-gen tensor cores (bf16 NHWC), for sm_100a.
//   out[img, oy, ox, co] = act( sum_{ky,kx,c} x[img, oy*s + ky - 1, ox*s + kx - 1, c] * w[co, ky, kx, c] + bias[co] ) (+ res)
// Replaces, inside the reference's `model(images)` (/root/reference/orchestration/orchestrator.py:590): fastervit
// `patch_embed.conv_down.3`, `levels.{0,1}.blocks.N.conv{1,2}` (+BN, GELU / residual) and
// `levels.N.downsample.reduction.0` (SURVEY.md 8a rows V0-V2 -- the FLOP bulk of FasterViT, K = 576 / 1152), and timm
// EfficientFormerV2 `stem.conv2` / `stages.N.downsample.conv` (rows F0, F1, F3).
//
// GEMM view: M = output pixels, N = Cout, K = 9 taps x Cin (each tap padded to whole 64-channel chunks).  No im2col
// buffer exists anywhere: the M tile is a RECTANGLE of output pixels (8 wide x TH tall x TI images, TH*TI = 16 -> 128
// rows) and the A operand of K chunk (tap, channel chunk) is ONE TMA box load from the NHWC activation tensor,
//     box [64 ch, 8, TH, TI] at (c0, ox0*s + kx - 1, oy0*s + ky - 1, img0), traversal stride s in W and H,
// which lands in shared memory as exactly the 128-row x 128-byte, 128B-swizzled K-major tile tcgen05.mma wants; the
// conv's zero padding, ragged image edges, ragged batches and channel tails are the TMA unit's out-of-bounds zero fill.
// Stride-1 layers (the FasterViT ConvBlocks) run in "vertical halo" mode: the tile is 8 wide x 16 tall in ONE image and a
// pipeline stage is (channel chunk, kernel column kx): ONE A box of 18 rows feeds the three taps ky = 0..2 -- a shift by one
// image row of an 8-pixel-wide tile is exactly one 1024-byte swizzle atom, i.e. a legal start address for the UMMA
// descriptor -- so the activation operand crosses L2 -> shared memory 3 times instead of 9 (the kernel is bound by that
// traffic: 216 KB per 128 x 64 x 576 tile at C = 64 is 27 KB/clk chip-wide at full tensor rate against ~6 KB/clk of L2).
// When all weights fit (cout_pad == bn and 9 * cpt tiles <= 80 KB: C = 64, 96) they are loaded ONCE per CTA and stay.
// The pipeline is the one of pwconv_tc.cu: a TMA producer warp, a multi-stage mbarrier ring, one MMA-issuing lane
// (M=128, N<=256, K=16 per instruction) into a double-buffered TMEM accumulator, and two groups of epilogue warps
// (tcgen05.ld -> bias -> GELU/ReLU -> +residual -> bf16 -> swizzled staging -> 4-D TMA store, edges clipped).
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

#include <stdlib.h>
#include <string.h>

namespace b200df {

using namespace tc;

namespace {

constexpr int kCBM = 128;
constexpr int kCBK = 64;
constexpr int kCABytes = kCBM * kCBK * 2;   // 16 KB
constexpr int kCMaxStages = 8;
constexpr int kCEpiWarps = 16;              // two accumulator groups x two column-half teams x 4 TMEM lane quarters
constexpr int kCEpiStageBytes = 32 * 128;   // per warp: two [32 rows x 32 cols] staging buffers
constexpr int kCMaxBias = 1024;
constexpr int kCThreads = 32 * (4 + kCEpiWarps);

typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t c2_pack(float lo, float hi) {
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void c2_unpack(f32x2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2_t c2_add(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

struct TileCoord { int nt, tx, ty, g; };

// Four K = 16 steps of one 64-wide K chunk in ONE asm block: the two descriptors differ from step to step only in the start
// address (low word, +32 bytes = +2), so the issuing thread spends 3 instructions per MMA instead of the ~14 of the generic
// 64-bit descriptor arithmetic.  With N <= 128 an MMA takes 32-64 clocks: the single issuing thread's instruction latency
// -- not the tensor pipe, not L2 -- was what bounded the small-N kernels (ncu, profiles/r02_d_conv3x3.txt).  a_lo / b_lo: low words of
// the K-major SW128 descriptors (make_sw128_kmajor_desc), hi: their common high word.  acc0: accumulate flag of step 0.
__device__ __forceinline__ void umma_bf16_k4(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                             uint32_t acc0) {
  asm volatile(
      "{\n"
      ".reg .pred p, t;\n"
      ".reg .b32 a, b;\n"
      ".reg .b64 da, db;\n"
      "setp.ne.b32 p, %5, 0;\n"
      "setp.eq.b32 t, 0, 0;\n"
      "mov.b64 da, {%1, %3};\n"
      "mov.b64 db, {%2, %3};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
      "add.u32 a, %1, 2;\n"
      "add.u32 b, %2, 2;\n"
      "mov.b64 da, {a, %3};\n"
      "mov.b64 db, {b, %3};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, t;\n"
      "add.u32 a, %1, 4;\n"
      "add.u32 b, %2, 4;\n"
      "mov.b64 da, {a, %3};\n"
      "mov.b64 db, {b, %3};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, t;\n"
      "add.u32 a, %1, 6;\n"
      "add.u32 b, %2, 6;\n"
      "mov.b64 da, {a, %3};\n"
      "mov.b64 db, {b, %3};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, t;\n"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(acc0)
      : "memory");
}

}  // namespace

template <int kAct, bool kRes>
__global__ void __launch_bounds__(kCThreads, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                  const __grid_constant__ CUtensorMap tmap_out, const ConvTcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int b_bytes = p.bn * kCBK * 2;
  const int a_bytes = p.vhalo ? (p.th + 2) * 1024 : kCABytes;
  const int res_bytes = p.b_res ? p.k_chunks * b_bytes : 0;                       // resident weights in front of the ring
  // classic schedule: one ring of (A box + weight tile) stages.  vhalo: an A ring of 18-row boxes and, unless the weights are
  // resident, a separate ring of weight tiles (three per A box) -- more, smaller stages in flight for the same shared memory
  const int stage_bytes = p.vhalo ? a_bytes : a_bytes + (p.b_res ? 0 : b_bytes);
  uint8_t* ring = smem + res_bytes;
  uint8_t* bring = ring + (size_t)p.stages * stage_bytes;
  uint8_t* epi_smem = bring + (size_t)p.b_stages * b_bytes;
  const int epi_warps_alloc = p.bn > 64 ? kCEpiWarps : kCEpiWarps / 2;   // bn <= 64: the second column-half team never stores
  float* sbias = reinterpret_cast<float*>(epi_smem + epi_warps_alloc * kCEpiStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sbias + kCMaxBias);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kCMaxStages;
  uint64_t* bfull_bar = bars + 2 * kCMaxStages;
  uint64_t* bempty_bar = bars + 3 * kCMaxStages;
  uint64_t* tmem_full = bars + 4 * kCMaxStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* bres_bar = tmem_empty + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bres_bar + 1);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int t_begin = blockIdx.x, t_end = p.num_tiles, t_step = gridDim.x;
  auto decode = [&](int t) {
    TileCoord c;
    c.nt = t % p.n_tiles;
    int r = t / p.n_tiles;
    c.tx = r % p.tiles_x;
    r /= p.tiles_x;
    c.ty = r % p.tiles_y;
    c.g = r / p.tiles_y;
    return c;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_out);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < p.b_stages; ++s) {
      mbar_init(&bfull_bar[s], 1);
      mbar_init(&bempty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 256);
    }
    mbar_init(bres_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, p.tmem_cols);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < p.cout_pad; i += blockDim.x) sbias[i] = __ldg(p.bias + i);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_ptr, 0);
  griddep_wait();               // everything above touched only weights / shared memory / TMEM
  griddep_launch_dependents();

  if (warp == 0) {
    // ================= TMA producer =================
    int s = 0;
    uint32_t ph = 0;
    if (p.b_res && t_begin < t_end) {   // all weight tiles, once (weights only: no dependence on the previous kernel's output)
      if (elect_one()) {
        mbar_arrive_expect_tx(bres_bar, (uint32_t)res_bytes);
        for (int kc = 0; kc < p.k_chunks; ++kc) tma_load_2d(smem + (size_t)kc * b_bytes, &tmap_w, kc * kCBK, 0, bres_bar);
      }
      __syncwarp();
    }
    int sb = 0;
    uint32_t phb = 0;
    for (int t = t_begin; p.vhalo && t < t_end; t += t_step) {
      const TileCoord tc_ = decode(t);
      const int ix0 = tc_.tx * 8 - 1, iy0 = tc_.ty * p.th - 1;
      for (int cc = 0; cc < p.cpt; ++cc)
        for (int kx = 0; kx < 3; ++kx) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(&full_bar[s], (uint32_t)a_bytes);
            tma_load_4d(ring + (size_t)s * a_bytes, &tmap_x, cc * kCBK, ix0 + kx, iy0, tc_.g, &full_bar[s]);   // th + 2 rows: ky = 0, 1, 2
          }
          __syncwarp();
          if (++s == p.stages) { s = 0; ph ^= 1; }
          for (int ky = 0; ky < 3 && !p.b_res; ++ky) {
            mbar_wait(&bempty_bar[sb], phb ^ 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&bfull_bar[sb], (uint32_t)b_bytes);
              tma_load_2d(bring + (size_t)sb * b_bytes, &tmap_w, ((ky * 3 + kx) * p.cpt + cc) * kCBK, tc_.nt * p.bn, &bfull_bar[sb]);
            }
            __syncwarp();
            if (++sb == p.b_stages) { sb = 0; phb ^= 1; }
          }
        }
    }
    for (int t = t_begin; !p.vhalo && t < t_end; t += t_step) {
      const TileCoord tc_ = decode(t);
      const int ix0 = tc_.tx * 8 * p.stride - 1, iy0 = tc_.ty * p.th * p.stride - 1, img0 = tc_.g * p.ti;
      for (int kc = 0; kc < p.k_chunks; ++kc) {
        const int tap = kc / p.cpt, cc = kc - tap * p.cpt;
        const int ky = tap / 3, kx = tap - ky * 3;
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* sa = ring + (size_t)s * stage_bytes;
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[s], (uint32_t)stage_bytes);
          tma_load_4d(sa, &tmap_x, cc * kCBK, ix0 + kx, iy0 + ky, img0, &full_bar[s]);
          if (!p.b_res) tma_load_2d(sa + kCABytes, &tmap_w, kc * kCBK, tc_.nt * p.bn, &full_bar[s]);
        }
        __syncwarp();
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = make_idesc_bf16_m128((uint32_t)p.bn);
    const uint64_t desc_res = make_sw128_kmajor_desc(smem_u32(smem));
    const uint64_t desc0 = make_sw128_kmajor_desc(smem_u32(ring));
    const uint64_t desc_b0 = make_sw128_kmajor_desc(smem_u32(bring));
    const int stage16 = stage_bytes >> 4;
    const uint32_t desc_hi = (uint32_t)(desc0 >> 32);
    const uint32_t lo0 = (uint32_t)desc0, lo_res = (uint32_t)desc_res, lo_b0 = (uint32_t)desc_b0;
    int sb = 0;
    uint32_t phb = 0;
    int s = 0;
    uint32_t ph = 0;
    int lt = 0;
    if (p.b_res && t_begin < t_end) {
      mbar_wait(bres_bar, 0);
      tcgen05_fence_after();
    }
    for (int t = t_begin; p.vhalo && t < t_end; t += t_step, ++lt) {
      const int acc = lt & 1;
      const uint32_t acc_ph = (lt >> 1) & 1;
      mbar_wait(&tmem_empty[acc], acc_ph ^ 1);
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.acc_stride);
      for (int cc = 0; cc < p.cpt; ++cc) {
        const int crem = p.cin - cc * kCBK;
        const int nk = crem >= kCBK ? 4 : ((crem + 15) >> 4);
        for (int kx = 0; kx < 3; ++kx) {
          mbar_wait(&full_bar[s], ph);
          tcgen05_fence_after();
          const uint32_t a0 = lo0 + (uint32_t)(s * stage16);
          for (int ky = 0; ky < 3; ++ky) {
            const uint32_t a_lo = a0 + (uint32_t)(ky * (1024 >> 4));          // one image row of the tile = one swizzle atom
            uint32_t b_lo;
            if (p.b_res) {
              b_lo = lo_res + (uint32_t)((((ky * 3 + kx) * p.cpt + cc) * b_bytes) >> 4);
            } else {
              mbar_wait(&bfull_bar[sb], phb);
              tcgen05_fence_after();
              b_lo = lo_b0 + (uint32_t)((sb * b_bytes) >> 4);
            }
            if (elect_one()) {
              const uint32_t first = (cc | kx | ky) ? 1u : 0u;
              if (nk == 4) {
                umma_bf16_k4(d_tmem, a_lo, b_lo, desc_hi, idesc, first);
              } else {
                for (int k = 0; k < nk; ++k)
                  umma_bf16(d_tmem, ((uint64_t)desc_hi << 32) | (a_lo + 2 * k), ((uint64_t)desc_hi << 32) | (b_lo + 2 * k), idesc,
                            (first | k) ? 1u : 0u);
              }
              if (!p.b_res) umma_commit(&bempty_bar[sb]);
              if (ky == 2) {
                umma_commit(&empty_bar[s]);
                if (cc == p.cpt - 1 && kx == 2) umma_commit(&tmem_full[acc]);
              }
            }
            __syncwarp();
            if (!p.b_res && ++sb == p.b_stages) { sb = 0; phb ^= 1; }
          }
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
    for (int t = t_begin; !p.vhalo && t < t_end; t += t_step, ++lt) {
      const int acc = lt & 1;
      const uint32_t acc_ph = (lt >> 1) & 1;
      mbar_wait(&tmem_empty[acc], acc_ph ^ 1);
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.acc_stride);
      for (int kc = 0; kc < p.k_chunks; ++kc) {
        mbar_wait(&full_bar[s], ph);
        tcgen05_fence_after();
        const uint64_t da = desc0 + (uint64_t)(uint32_t)(s * stage16);
        const uint64_t db = p.b_res ? desc_res + (uint64_t)(uint32_t)((kc * b_bytes) >> 4) : da + (uint64_t)(kCABytes >> 4);
        const int cc = kc % p.cpt;
        const int crem = p.cin - cc * kCBK;                       // valid channels of this chunk
        const int nk = crem >= kCBK ? 4 : ((crem + 15) >> 4);     // K = 16 steps that hold any non-zero column
        if (elect_one()) {
          if (nk == 4) {
            umma_bf16_k4(d_tmem, (uint32_t)da, (uint32_t)db, desc_hi, idesc, kc ? 1u : 0u);
          } else {
            for (int k = 0; k < nk; ++k)
              umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kc | k) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (kc == p.k_chunks - 1) umma_commit(&tmem_full[acc]);
        }
        __syncwarp();
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    const int ew = warp - 4;
    const int grp = ew >> 3;          // accumulator this warp drains
    const int half = (ew >> 2) & 1;   // column-half team
    const int q = warp & 3;           // TMEM lane quarter = tile rows 32q .. 32q+31 = 4 image rows of 8 pixels
    uint8_t* stg_base = epi_smem + (p.bn > 64 ? ew : grp * 4 + q) * kCEpiStageBytes;   // (bn <= 64: half-1 warps never store)
    const int img_in_tile = (4 * q) / p.th, y0 = (4 * q) % p.th;
    const int lx = lane & 7, ly = lane >> 3;
    int stores = 0, lt = 0;
    for (int t = t_begin + grp * t_step; t < t_end; t += 2 * t_step, lt += 2) {
      const TileCoord tc_ = decode(t);
      const uint32_t acc_ph = (lt >> 1) & 1;
      mbar_wait(&tmem_full[grp], acc_ph);
      tcgen05_fence_after();
      const int ox = tc_.tx * 8 + lx, oy = tc_.ty * p.th + y0 + ly, img = tc_.g * p.ti + img_in_tile;
      const bool row_ok = ox < p.Wo && oy < p.Ho && img < p.n_img;
      const size_t row = ((size_t)img * p.Ho + oy) * p.Wo + ox;
      const int ncol0 = tc_.nt * p.bn;
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * p.acc_stride);
      if (half * 64 >= p.bn) {  // this team has no column block in this tile
        tcgen05_fence_before();
        mbar_arrive(&tmem_empty[grp]);
      }
      for (int cb = half * 64; cb < p.bn; cb += 128) {
        const int ncols = min(64, p.bn - cb);
#pragma unroll
        for (int hb = 0; hb < 2; ++hb) {
          if (hb * 32 < ncols) {
            uint8_t* stg = stg_base + (stores & 1) * (kCEpiStageBytes / 2);
            if (stores >= 2) {
              if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              __syncwarp();
            }
            const int col = ncol0 + cb + hb * 32;
            uint32_t r[32];
            tmem_ld32(t_row + (uint32_t)(cb + hb * 32), r);
            uint4 res[4];
            if constexpr (kRes) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                res[c] = make_uint4(0u, 0u, 0u, 0u);
                if (row_ok && col + 8 * c < p.cout)
                  res[c] = __ldg(reinterpret_cast<const uint4*>(p.residual + row * p.cout + col + 8 * c));
              }
            }
            tmem_ld_wait();
            if (cb + 128 >= p.bn && hb * 32 + 32 >= ncols) {  // this warp's last read of the accumulator
              tcgen05_fence_before();
              mbar_arrive(&tmem_empty[grp]);
            }
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = act_fn<kAct, true>(__uint_as_float(r[j]) + sbias[col + j]);
            if constexpr (kRes) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const uint32_t rw[4] = {res[c].x, res[c].y, res[c].z, res[c].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  v[8 * c + 2 * j] += __uint_as_float(rw[j] << 16);
                  v[8 * c + 2 * j + 1] += __uint_as_float(rw[j] & 0xffff0000u);
                }
              }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint4 pk;
              pk.x = pack_bf16x2(v[8 * c + 0], v[8 * c + 1]);
              pk.y = pack_bf16x2(v[8 * c + 2], v[8 * c + 3]);
              pk.z = pack_bf16x2(v[8 * c + 4], v[8 * c + 5]);
              pk.w = pack_bf16x2(v[8 * c + 6], v[8 * c + 7]);
              // [32 rows x 64 B] staging tile in the 64-B swizzle of the store's tensor map
              *reinterpret_cast<uint4*>(stg + lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4)) = pk;
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              // box [32 cols, 8 px, 4 rows, 1 image]: pixels outside the image / images past the batch are clipped
              asm volatile(
                  "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                      reinterpret_cast<uint64_t>(&tmap_out)),
                  "r"(smem_u32(stg)), "r"(col), "r"(tc_.tx * 8), "r"(tc_.ty * p.th + y0), "r"(tc_.g * p.ti + img_in_tile)
                  : "memory");
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            ++stores;
          }
        }
      }
    }
    if (stores > 0 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------

static size_t conv_tc_smem(int bn, int stages, size_t stage_bytes = 0, size_t res_bytes = 0, int b_stages = 0) {
  if (!stage_bytes) stage_bytes = kCABytes + (size_t)bn * kCBK * 2;
  const int epi_warps = bn > 64 ? kCEpiWarps : kCEpiWarps / 2;
  return 1024 + res_bytes + (size_t)stages * stage_bytes + (size_t)b_stages * bn * kCBK * 2 + epi_warps * kCEpiStageBytes +
         kCMaxBias * sizeof(float) + (4 * kCMaxStages + 5) * sizeof(uint64_t) + 16;
}

static bool conv_vhalo_enabled() {   // B200DF_CONV_VHALO=0: the one-box-per-tap schedule everywhere (A/B aid)
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200DF_CONV_VHALO");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

int conv3x3_tc_cin_pad(int cin) { return (cin + kCBK - 1) / kCBK * kCBK; }

int plan_conv3x3_tc(Conv3Plan& plan, const void* x, int n, int h, int w, int cin, int cout, int cout_pad, int bn,
                    int stride, const void* wgt_tc, const float* bias, const void* residual, int act, void* out) {
  if (n <= 0 || cin % 8 || cout % 8 || (stride != 1 && stride != 2) || cout_pad % bn || bn % 16 || bn > 256 ||
      cout_pad > kCMaxBias || (cout_pad != bn && bn % 64) || (act != kActNone && act != kActGELU && act != kActReLU))
    return -1;
  ConvTcParams& p = plan.p;
  memset(&p, 0, sizeof(p));
  const int Ho = (h - 1) / stride + 1, Wo = (w - 1) / stride + 1;
  // tile height: least wasted rows; ties -> the taller tile (fewer images per tile)
  int best_th = 16;
  double best_eff = -1.0;
  for (int th = 16; th >= 4; th /= 2) {
    const int ti = 16 / th;
    const double eff = (double)Ho / ((Ho + th - 1) / th * th) * (double)n / ((n + ti - 1) / ti * ti);
    if (eff > best_eff + 1e-9) { best_eff = eff; best_th = th; }
  }
  // stride 1: vertical-halo schedule (16-row tiles of ONE image; a few wasted rows at 56 / 28 buy 2.7x less A traffic) when
  // at least two of its stages (18 KB of A + three weight tiles, or A alone with resident weights) fit
  p.vhalo = 0;
  if (stride == 1 && conv_vhalo_enabled()) {
    const size_t bb = (size_t)bn * kCBK * 2, kch = 9 * (size_t)(conv3x3_tc_cin_pad(cin) / kCBK);
    const bool res_ok = cout_pad == bn && kch * bb <= 80 * 1024;
    if (conv_tc_smem(bn, 3, 18 * 1024, res_ok ? kch * bb : 0, res_ok ? 0 : 3) <= 220 * 1024) p.vhalo = 1;
  }
  if (p.vhalo) best_th = 16;
  p.cin = cin; p.cout = cout; p.cout_pad = cout_pad; p.bn = bn;
  p.Ho = Ho; p.Wo = Wo; p.stride = stride; p.n_img = n;
  p.th = best_th; p.ti = 16 / best_th;
  p.tiles_x = (Wo + 7) / 8;
  p.tiles_y = (Ho + p.th - 1) / p.th;
  p.groups = (n + p.ti - 1) / p.ti;
  p.cpt = conv3x3_tc_cin_pad(cin) / kCBK;
  p.k_chunks = 9 * p.cpt;
  p.n_tiles = cout_pad / bn;
  const long long tiles = (long long)p.tiles_x * p.tiles_y * p.groups * p.n_tiles;
  if (tiles >= (1ll << 31)) return -1;
  p.num_tiles = (int)tiles;
  p.acc_stride = (bn + 31) / 32 * 32;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * p.acc_stride)) cols <<= 1;
  p.tmem_cols = cols;
  const size_t b_bytes = (size_t)bn * kCBK * 2;
  p.b_res = (p.vhalo && p.n_tiles == 1 && p.k_chunks * b_bytes <= 80 * 1024) ? 1 : 0;
  const size_t res_bytes = p.b_res ? p.k_chunks * b_bytes : 0;
  const size_t stage_bytes = p.vhalo ? (size_t)(p.th + 2) * 1024 : kCABytes + b_bytes;
  int stages = kCMaxStages, b_stages = 0;
  if (p.vhalo && !p.b_res) {   // three A boxes in flight, the rest of the budget for weight tiles (three per A box)
    stages = 3;
    b_stages = kCMaxStages;
    while (b_stages > 3 && conv_tc_smem(bn, stages, stage_bytes, 0, b_stages) > 220 * 1024) --b_stages;
  } else {
    while (stages > 2 && conv_tc_smem(bn, stages, stage_bytes, res_bytes) > 220 * 1024) --stages;
  }
  if (conv_tc_smem(bn, stages, stage_bytes, res_bytes, b_stages) > 220 * 1024) return -1;
  p.stages = stages;
  p.b_stages = b_stages;
  p.bias = bias;
  p.residual = reinterpret_cast<const bf16*>(residual);
  plan.act = act;
  plan.smem = conv_tc_smem(bn, stages, stage_bytes, res_bytes, b_stages);
  plan.grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  // A: NHWC activations [cin, w, h, n], box [64, 8 s, th s, ti] traversed with stride s in W and H, 128B swizzle
  {
    unsigned long long dims[4] = {(unsigned long long)cin, (unsigned long long)w, (unsigned long long)h, (unsigned long long)n};
    unsigned long long str[4] = {2, (unsigned long long)cin * 2, (unsigned long long)w * cin * 2, (unsigned long long)h * w * cin * 2};
    unsigned box[4] = {(unsigned)kCBK, (unsigned)(8 * stride), (unsigned)(p.vhalo ? p.th + 2 : p.th * stride), (unsigned)p.ti};
    unsigned es[4] = {1, (unsigned)stride, (unsigned)stride, 1};
    if (encode_tmap_nd_bf16(plan.tmap_x, x, 4, dims, str, box, 128, es)) return -3;
  }
  // B: weights [cout_pad, 9 * cin_pad], box [bn, 64]
  if (encode_tmap_2d_bf16(plan.tmap_w, wgt_tc, cout_pad, 9 * conv3x3_tc_cin_pad(cin), bn)) return -3;
  // out: NHWC [cout, Wo, Ho, n], box [32, 8, 4, 1], 64B swizzle
  {
    unsigned long long dims[4] = {(unsigned long long)cout, (unsigned long long)Wo, (unsigned long long)Ho, (unsigned long long)n};
    unsigned long long str[4] = {2, (unsigned long long)cout * 2, (unsigned long long)Wo * cout * 2, (unsigned long long)Ho * Wo * cout * 2};
    unsigned box[4] = {32, 8, 4, 1};
    if (encode_tmap_nd_bf16(plan.tmap_out, out, 4, dims, str, box, 64)) return -3;
  }
  plan.x = x;
  plan.out = out;
  plan.n = n;
  return 0;
}

cudaError_t launch_conv3x3_tc(const Conv3Plan& pl, cudaStream_t s) {
  const CUtensorMap* tx = reinterpret_cast<const CUtensorMap*>(pl.tmap_x);
  const CUtensorMap* tw = reinterpret_cast<const CUtensorMap*>(pl.tmap_w);
  const CUtensorMap* to = reinterpret_cast<const CUtensorMap*>(pl.tmap_out);
  const bool res = pl.p.residual != nullptr;
#define CONV_LAUNCH(A_, R_)                                                                                              \
  if (pl.act == A_ && res == R_) {                                                                                       \
    static PerDeviceOnce once;                                                                                           \
    if (once.need()) {                                                                                                   \
      cudaError_t e = cudaFuncSetAttribute(conv3x3_tc_kernel<A_, R_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); \
      if (e != cudaSuccess) return e;                                                                                    \
    }                                                                                                                    \
    count_launch();                                                                                                      \
    return launch_pdl(conv3x3_tc_kernel<A_, R_>, dim3(pl.grid), dim3(kCThreads), pl.smem, s, *tx, *tw, *to, pl.p);       \
  }
  CONV_LAUNCH(kActNone, false)
  CONV_LAUNCH(kActNone, true)
  CONV_LAUNCH(kActGELU, false)
  CONV_LAUNCH(kActReLU, false)
  CONV_LAUNCH(kActGELU, true)
  CONV_LAUNCH(kActReLU, true)
#undef CONV_LAUNCH
  return cudaErrorInvalidValue;
}

}  // namespace b200df
