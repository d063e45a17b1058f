// EXPERIMENT (off by default, B200DF_DW_TC=1; correct but measured ~2x slower than dwconv_tma.cu -- each tap
// re-streams the whole pixel tile from shared memory through the tensor core, ~80 cycles per N=16 MMA):
// Depthwise kxk stride-1 conv (k = 3|5, pad (k-1)/2) + folded-BN bias + SiLU + SE pool sums on the
// 5th-gen TENSOR cores (bf16 NHWC), for sm_100a.  Replaces timm `conv_dw` + `bn`(+SiLU) and the `se` mean_hw
//  inside the reference's `model(images)`
// (/root/reference/orchestration/orchestrator.py:590).
//
// Why tensor cores for a depthwise conv: on the CUDA cores this layer class costs ~25 issue slots per output
// and is issue-bound at ~45 % of the HBM roofline.  Written as
//     out[p, c] = sum_tap  in[p + tap, c] * w[tap, c]  =  sum_tap ( A_tap  x  diag(w_tap) )[p, c]
// every tap is a GEMM of a SHIFTED VIEW of the input halo tile (already in shared memory, K-major, 128B
// swizzle, exactly what TMA delivers) with a diagonal weight matrix.  A diagonal 64x64 matrix only couples
// channel c with itself, so per tap the 64-channel block needs four tcgen05.mma of M=128 pixels, N=16, K=16
// (one per 16-channel group) -- 97 % of the multiplies hit zeros, but the tensor pipe has ~60x the CUDA-core
// rate and the instruction stream per output drops by an order of magnitude:
//   * warp 0: TMA producer -- one 4-D box (64 ch, 16 x (16+k-1) pixels, OOB zero fill = conv padding) per
//     tile of 8 x 16 output pixels, multi-stage mbarrier ring;
//   * warp 1: MMA issuer -- k*k*4 MMAs per tile; the A descriptor of tap (ky,kx) is the stage base advanced
//     by (ky*16 + kx) pixel rows (128 B each; the swizzle is address-based, no base offset needed), the
//     8-row core groups are the 8 output columns of one output row (group stride = one halo row);
//     B = per-tap [16 x 64] tile holding the four 16x16 diagonal blocks side by side along K;
//   * warps 2-9: two epilogue groups (one per TMEM accumulator): tcgen05.ld, +bias, SiLU (one MUFU), bf16,
//     swizzled staging tile -> ONE 4-D TMA store per warp (out-of-image pixels / channels clipped);
//     SE pool sums: column sums of the staging tile, fixed-order combine, Q24 integer atomics.
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

#include <stdlib.h>

#include <algorithm>

namespace b200df {

using namespace tc;

constexpr int kTcCB = 64;   // channels per tile (128-B pixel rows)
constexpr int kTcOW = 8;    // output tile width  (= one 8-row core-matrix group)
constexpr int kTcOH = 16;   // output tile height (16 groups -> M = 128)
constexpr int kTcHW = 16;   // halo tile width in pixels (>= 8 + k - 1, multiple of 8: group stride % 1024 == 0)
constexpr int kTcMaxStages = 4;
constexpr int kTcThreads = 320;

// K-major SW128 descriptor with an explicit 8-row-group stride and swizzle base offset
__device__ __forceinline__ uint64_t make_sw128_desc_sbo(uint32_t smem_addr, uint32_t sbo_bytes, uint32_t base_off) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(base_off & 7u) << 49;
  d |= (uint64_t)2 << 61;
  return d;
}

struct SpIter {  // (tile x, tile y, image), advanced by a fixed step in mixed radix (no divisions per tile)
  int tx, ty, img, dx, dy, di, nx, ny;
  __device__ __forceinline__ SpIter(int i0, int step, int nx_, int ny_) : nx(nx_), ny(ny_) {
    tx = i0 % nx;
    int r = i0 / nx;
    ty = r % ny;
    img = r / ny;
    dx = step % nx;
    r = step / nx;
    dy = r % ny;
    di = r / ny;
  }
  __device__ __forceinline__ void next() {
    tx += dx;
    ty += dy;
    img += di;
    if (tx >= nx) { tx -= nx; ++ty; }
    if (ty >= ny) { ty -= ny; ++img; }
  }
};

template <int K>
__global__ void __launch_bounds__(kTcThreads, 1)
dwconv_tc_kernel(const __grid_constant__ CUtensorMap tmap_in, const __grid_constant__ CUtensorMap tmap_out,
                 const DwTcParams p) {
  constexpr int KK = K * K;
  constexpr int PAD = (K - 1) / 2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sB = smem + (size_t)p.stages * p.stage_stride;        // [KK][16 rows x 128 B]
  uint8_t* stg_all = sB + KK * 2048;                              // [8 warps][32 rows x 128 B]
  float* sbias = reinterpret_cast<float*>(stg_all + 8 * 4096);    // [64]
  float* slots = sbias + kTcCB;                                   // [2 parity][2 groups][4 warps][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(slots + 2 * 2 * 4 * kTcCB);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kTcMaxStages;
  uint64_t* tmem_full = bars + 2 * kTcMaxStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cb = blockIdx.x % p.ncb;            // this CTA's channel block (fixed)
  const int sp0 = blockIdx.x / p.ncb;           // first spatial tile
  const int sp_step = gridDim.x / p.ncb;
  const int sp_total = p.tiles_x * p.tiles_y * p.n;
  const int c0 = cb * kTcCB;

  // ---- per-CTA setup: diagonal weight tiles, bias, barriers, TMEM ----
  for (int i = tid; i < KK * 2048 / 16; i += kTcThreads) reinterpret_cast<uint4*>(sB)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (tid < kTcCB) sbias[tid] = (c0 + tid < p.C) ? __ldg(p.bias + c0 + tid) : 0.f;
  __syncthreads();
  for (int i = tid; i < KK * kTcCB; i += kTcThreads) {
    const int tap = i / kTcCB, c = i - tap * kTcCB;
    const int nn = c & 15, j = c >> 4;          // row of the [16 x 64] tile, 16-channel group
    const int k = 16 * j + nn;                  // column: group j's diagonal sits in K columns 16j .. 16j+15
    const int chunk = k >> 3;
    const float w = (c0 + c < p.C) ? __ldg(p.wgt + (size_t)tap * p.C + c0 + c) : 0.f;
    *reinterpret_cast<bf16*>(sB + tap * 2048 + nn * 128 + ((chunk ^ (nn & 7)) << 4) + (k & 7) * 2) =
        __float2bfloat16_rn(w);
  }
  if (tid == 0) {
    tma_prefetch_desc(&tmap_in);
    tma_prefetch_desc(&tmap_out);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr, 128);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      SpIter it(sp0, sp_step, p.tiles_x, p.tiles_y);
      int s = 0;
      uint32_t ph = 0;
      for (int sp = sp0; sp < sp_total; sp += sp_step, it.next()) {
        mbar_wait(&empty_bar[s], ph ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], p.box_bytes);
        tma_load_4d(smem + (size_t)s * p.stage_stride, &tmap_in, c0, it.tx * kTcOW - PAD, it.ty * kTcOH - PAD, it.img,
                    &full_bar[s]);
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = make_idesc_bf16_m128(16u);
    const uint64_t b_desc0 = make_sw128_kmajor_desc(smem_u32(sB));
    int s = 0;
    uint32_t ph = 0;
    int lt = 0;
    for (int sp = sp0; sp < sp_total; sp += sp_step, ++lt) {
      const int acc = lt & 1;
      mbar_wait(&tmem_empty[acc], ((lt >> 1) & 1) ^ 1);
      mbar_wait(&full_bar[s], ph);
      tcgen05_fence_after();
      if (lane == 0) {
        const uint32_t d0 = tmem_base + (uint32_t)(acc * kTcCB);
        const uint64_t a_desc0 = make_sw128_desc_sbo(smem_u32(smem + (size_t)s * p.stage_stride), kTcHW * 128, 0);
#pragma unroll
        for (int ky = 0; ky < K; ++ky) {
#pragma unroll
          for (int kx = 0; kx < K; ++kx) {
            // shifted view: +(ky*16 + kx) pixel rows of 128 B.  The 128B swizzle is a function of the absolute
            // shared-memory address bits for both the TMA write and the MMA read, so a start address that is
            // not 1024-B aligned needs NO descriptor base offset (measured: base_offset = kx gives wrong data)
            const uint64_t a_desc = a_desc0 + (uint64_t)(((ky * kTcHW + kx) * 128) >> 4);
            const uint64_t b_desc = b_desc0 + (uint64_t)((ky * K + kx) * (2048 >> 4));
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_bf16(d0 + 16 * j, a_desc + (uint64_t)(2 * j), b_desc + (uint64_t)(2 * j), idesc, (ky | kx) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);
        umma_commit(&tmem_full[acc]);
      }
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1; }
    }
  } else {
    // ================= epilogue (group = accumulator) =================
    const int ew = warp - 2;       // 0..7
    const int grp = ew >> 2;
    const int q = warp & 3;        // TMEM lane quarter of this warp
    uint8_t* stg = stg_all + ew * 4096;
    const int yy = 4 * q + (lane >> 3), xx = lane & 7;  // this lane's pixel inside the 8 x 16 tile (m = 32q + lane)
    SpIter it(sp0 + grp * sp_step, 2 * sp_step, p.tiles_x, p.tiles_y);
    int lt = 0;  // local tile count of this group
    bool pending = false;
    for (int sp = sp0 + grp * sp_step; sp < sp_total; sp += 2 * sp_step, ++lt, it.next()) {
      mbar_wait(&tmem_full[grp], lt & 1);
      tcgen05_fence_after();
      const int oy = it.ty * kTcOH + yy, ox = it.tx * kTcOW + xx;
      const bool valid = oy < p.H && ox < p.W;
      if (pending) {  // previous TMA store must have finished reading the staging tile
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
      }
#pragma unroll
      for (int hb = 0; hb < 2; ++hb) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * kTcCB + hb * 32), r);
        tmem_ld_wait();
        if (hb == 1) {
          tcgen05_fence_before();
          mbar_arrive(&tmem_empty[grp]);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float v = silu_fast(__uint_as_float(r[c * 8 + j]) + sbias[hb * 32 + c * 8 + j]);
            f[j] = valid ? v : 0.f;  // out-of-image pixels must not reach the pool sums
          }
          const int chunk = hb * 4 + c;
          *reinterpret_cast<uint4*>(stg + lane * 128 + ((chunk ^ (lane & 7)) << 4)) = pack_bf16x8(f);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        asm volatile(
            "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                reinterpret_cast<uint64_t>(&tmap_out)),
            "r"(smem_u32(stg)), "r"(c0), "r"(it.tx * kTcOW), "r"(it.ty * kTcOH + 4 * q), "r"(it.img)
            : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      pending = true;
      if (p.pool != nullptr) {
        // column sums of this warp's [32 rows x 64 ch] staging tile: lane owns channels 2*lane, 2*lane+1
        float s0 = 0.f, s1 = 0.f;
        const int chunk = lane >> 2, within = (lane & 3) * 4;
#pragma unroll
        for (int rr = 0; rr < 32; ++rr) {
          const uint32_t wv = *reinterpret_cast<const uint32_t*>(stg + rr * 128 + ((chunk ^ (rr & 7)) << 4) + within);
          s0 += __uint_as_float(wv << 16);
          s1 += __uint_as_float(wv & 0xffff0000u);
        }
        float* slot = slots + (((lt & 1) * 2 + grp) * 4 + q) * kTcCB;
        slot[2 * lane] = s0;
        slot[2 * lane + 1] = s1;
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        const int t128 = q * 32 + lane;  // any bijection of the group's 128 threads works
        if (t128 < kTcCB && c0 + t128 < p.C) {
          const float* sl = slots + ((lt & 1) * 2 + grp) * 4 * kTcCB + t128;
          const float v = ((sl[0] + sl[kTcCB]) + sl[2 * kTcCB]) + sl[3 * kTcCB];  // fixed order
          pool_add(p.pool + (size_t)it.img * p.C + c0 + t128, v);
        }
      }
    }
    if (pending && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// ---- host side ------------------------------------------------------------------------------------

static size_t dw_tc_smem(int k, int stages, int stage_stride) {
  return 1024 + (size_t)stages * stage_stride + (size_t)k * k * 2048 + 8 * 4096 + kTcCB * 4 + 2 * 2 * 4 * kTcCB * 4 +
         (2 * kTcMaxStages + 4) * 8 + 16;
}

int plan_dwconv_tc(DwPlan& plan, const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                   const float* bias, void* out, pool_t* pool) {
  if (n <= 0 || stride != 1 || (k != 3 && k != 5) || c % 8) return -1;
  DwTcParams& q = plan.q;
  q.C = c; q.H = h; q.W = w; q.n = n;
  q.tiles_x = (w + kTcOW - 1) / kTcOW;
  q.tiles_y = (h + kTcOH - 1) / kTcOH;
  q.ncb = (c + kTcCB - 1) / kTcCB;
  if (q.ncb > kNumSMs) return -1;
  const int hr = kTcOH + k - 1;
  q.box_bytes = (uint32_t)(kTcCB * 2 * kTcHW * hr);
  q.stage_stride = (int)((q.box_bytes + 1023) & ~1023u);
  int stages = kTcMaxStages;
  while (stages > 2 && dw_tc_smem(k, stages, q.stage_stride) > 220 * 1024) --stages;
  q.stages = stages;
  q.wgt = wgt; q.bias = bias; q.pool = pool;

  unsigned long long dims[4] = {(unsigned long long)c, (unsigned long long)w, (unsigned long long)h, (unsigned long long)n};
  unsigned long long str[4] = {2, (unsigned long long)c * 2, (unsigned long long)w * c * 2, (unsigned long long)h * w * c * 2};
  unsigned box_in[4] = {kTcCB, kTcHW, (unsigned)hr, 1};
  unsigned box_out[4] = {kTcCB, kTcOW, 4, 1};
  if (encode_tmap_nd_bf16(plan.tmap, x, 4, dims, str, box_in, 128)) return -3;
  if (encode_tmap_nd_bf16(plan.tmap_out, out, 4, dims, str, box_out, 128)) return -3;
  plan.k = k; plan.stride = 1; plan.tc = 1;
  const long long sp_total = (long long)q.tiles_x * q.tiles_y * n;
  long long per_cb = kNumSMs / q.ncb;               // CTAs per channel block
  if (per_cb > sp_total) per_cb = sp_total;
  plan.grid = (int)(per_cb * q.ncb);
  plan.block = kTcThreads;
  plan.smem = dw_tc_smem(k, stages, q.stage_stride);
  plan.x = x;
  return 0;
}

template <int K>
static cudaError_t dw_tc_launch(const DwPlan& pl, cudaStream_t s) {
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(dwconv_tc_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  dwconv_tc_kernel<K><<<pl.grid, pl.block, pl.smem, s>>>(*reinterpret_cast<const CUtensorMap*>(pl.tmap),
                                                          *reinterpret_cast<const CUtensorMap*>(pl.tmap_out), pl.q);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_dwconv_tc(const DwPlan& pl, cudaStream_t s) {
  if (pl.k == 3) return dw_tc_launch<3>(pl, s);
  if (pl.k == 5) return dw_tc_launch<5>(pl, s);
  return cudaErrorInvalidValue;
}

}  // namespace b200df
