// Pointwise (1x1) conv as a tcgen05 GEMM for sm_100a -- the bf16 throughput path.
//   out[m, co] = act( sum_k (x[m,k] * gate[m / rows_per_img, k]) * w[co,k] + bias[co] ) (+ residual[m,co])
// Replaces timm `conv_pw` (+bn1+SiLU), `conv_pwl` (+bn3, SE gate, skip add) and `conv_head`
//  inside the reference's `model(images)`
// (/root/reference/orchestration/orchestrator.py:590).
//
// Shape of the problem: M = n*H*W rows (up to 3.2 M), K = Cin in [16, 1152], N = Cout in [16, 1280].
// Every one of these GEMMs except conv_head is HBM-bound, so the kernel is built to stream:
//   * persistent CTAs, static round-robin over (m_tile, n_tile) tiles of 128 x BN;
//   * warp 0: TMA producer -- A tile [128 x 64] and B tile [BN x 64] per K chunk, 128B swizzle,
//     multi-stage mbarrier ring; K / M / N tails are zero-filled by TMA out-of-bounds handling;
//   * warp 1: one lane issues tcgen05.mma (M=128, N=BN, K=16) into a double-buffered TMEM
//     accumulator and commits to the ring's `empty` barriers and to `tmem_full`;
//   * warps 4-7: epilogue -- tcgen05.ld 32x32b, +bias, SiLU, +residual (fp32), bf16 pack, transpose
//     through a warp-private swizzled smem tile so global stores are 128 B contiguous per row;
//   * (gated variant) warps 8-11 multiply the landed A tile in place by the per-image SE gate and
//     hand it to the MMA warp through a second barrier (fence.proxy.async before the arrive).
// A [m, cin] and W [cout_pad, cin] are bf16 row-major (K-major operands, no transposes anywhere).
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

#include <mutex>

namespace b200df {

using namespace tc;

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kABytes = kBM * kBK * 2;  // 16 KB
constexpr int kMaxStages = 8;
constexpr int kEpiStageBytes = 32 * 128;  // per epilogue warp: 32 rows x 64 bf16

struct TcParams {
  int m, cin, cout, cout_pad, bn;
  int m_tiles, n_tiles, k_chunks, stages;
  uint32_t tmem_cols;
  const float* bias;
  const float* gate;
  int rows_per_img;
  const bf16* residual;
  int act;
  bf16* out;
};

template <bool kGate>
__global__ void __launch_bounds__(kGate ? 384 : 256, 1)
pwconv_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-B align the operand ring (128B swizzle atoms are 1024 B)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int b_bytes = p.bn * kBK * 2;
  const int stage_bytes = kABytes + b_bytes;
  uint8_t* epi_smem = smem + (size_t)p.stages * stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + 4 * kEpiStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kMaxStages;
  uint64_t* scaled_bar = bars + 2 * kMaxStages;
  uint64_t* tmem_full = bars + 3 * kMaxStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.m_tiles * p.n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&scaled_bar[s], 128);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, p.tmem_cols);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles;
        for (int kc = 0; kc < p.k_chunks; ++kc) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + (size_t)s * stage_bytes;
          mbar_arrive_expect_tx(&full_bar[s], (uint32_t)stage_bytes);
          tma_load_2d(sa, &tmap_a, kc * kBK, mt * kBM, &full_bar[s]);
          tma_load_2d(sa + kABytes, &tmap_b, kc * kBK, nt * p.bn, &full_bar[s]);
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = make_idesc_bf16_m128((uint32_t)p.bn);
    int s = 0;
    uint32_t ph = 0;
    int lt = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++lt) {
      const int acc = lt & 1;
      const uint32_t acc_ph = (lt >> 1) & 1;
      mbar_wait(&tmem_empty[acc], acc_ph ^ 1);
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.bn);
      for (int kc = 0; kc < p.k_chunks; ++kc) {
        mbar_wait(kGate ? &scaled_bar[s] : &full_bar[s], ph);
        tcgen05_fence_after();
        if (lane == 0) {
          const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
          const uint64_t da = make_sw128_kmajor_desc(a_addr);
          const uint64_t db = make_sw128_kmajor_desc(a_addr + kABytes);
          int krem = p.cin - kc * kBK;
          int nk = krem >= kBK ? kBK / 16 : (krem + 15) / 16;
          for (int k = 0; k < nk; ++k) {
            // +32 bytes per K=16 step inside the swizzle atom -> +2 in the (addr >> 4) field
            umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kc | k) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (kc == p.k_chunks - 1) umma_commit(&tmem_full[acc]);
        }
        __syncwarp();
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ================= epilogue =================
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    uint8_t* stg = epi_smem + q * kEpiStageBytes;
    const uint32_t stg_u32 = smem_u32(stg);
    int lt = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++lt) {
      const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles;
      const int acc = lt & 1;
      const uint32_t acc_ph = (lt >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_ph);
      tcgen05_fence_after();
      const int row0 = mt * kBM + q * 32;  // first row of this warp
      const int row = row0 + lane;
      const int ncol0 = nt * p.bn;
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * p.bn);
      for (int cb = 0; cb < p.bn; cb += 64) {
        const int ncols = min(64, p.bn - cb);  // multiple of 16
        uint32_t r[4][16];
#pragma unroll
        for (int g = 0; g < 4; ++g)
          if (g * 16 < ncols) tmem_ld16(t_row + (uint32_t)(cb + g * 16), r[g]);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if (g * 16 < ncols) {
            const int col = ncol0 + cb + g * 16;  // global output column of r[g][0]
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[g][j]) + __ldg(p.bias + col + j);
            if (p.act == 1) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = silu_fast(v[j]);
            }
            if (p.residual != nullptr && row < p.m) {
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                if (col + 8 * h < p.cout) {
                  float rr[8];
                  load8<bf16>(p.residual + (size_t)row * p.cout + col + 8 * h, rr);
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[8 * h + j] += rr[j];
                }
              }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint4 pk;
              pk.x = pack_bf16x2(v[8 * h + 0], v[8 * h + 1]);
              pk.y = pack_bf16x2(v[8 * h + 2], v[8 * h + 3]);
              pk.z = pack_bf16x2(v[8 * h + 4], v[8 * h + 5]);
              pk.w = pack_bf16x2(v[8 * h + 6], v[8 * h + 7]);
              const int chunk = g * 2 + h;
              *reinterpret_cast<uint4*>(stg + lane * 128 + ((chunk ^ (lane & 7)) << 4)) = pk;
            }
          }
        }
        __syncwarp();
        // coalesced copy-out: 8 lanes cover one 128-B row segment, 4 rows per instruction
        const int j = lane & 7;
        const int gcol = ncol0 + cb + j * 8;
        const bool col_ok = (j * 8 < ncols) && (gcol < p.cout);
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) {
          const int rr = ps * 4 + (lane >> 3);
          const int grow = row0 + rr;
          if (col_ok && grow < p.m) {
            uint4 pk;
            const uint32_t a = stg_u32 + rr * 128 + ((j ^ (rr & 7)) << 4);
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(pk.x), "=r"(pk.y), "=r"(pk.z), "=r"(pk.w)
                         : "r"(a));
            *reinterpret_cast<uint4*>(p.out + (size_t)grow * p.cout + gcol) = pk;
          }
        }
        __syncwarp();
      }
      tcgen05_fence_before();
      mbar_arrive(&tmem_empty[acc]);
    }
  } else if (kGate && warp >= 8) {
    // ================= SE-gate scaler: A[m,k] *= gate[img(m), k] in shared memory =================
    const int tr = threadIdx.x - 256;  // row of the A tile owned by this thread
    int s = 0;
    uint32_t ph = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int mt = t / p.n_tiles;
      const int row = mt * kBM + tr;
      const bool row_ok = row < p.m;
      const float* grow = p.gate + (size_t)(row_ok ? row / p.rows_per_img : 0) * p.cin;
      for (int kc = 0; kc < p.k_chunks; ++kc) {
        mbar_wait(&full_bar[s], ph);
        if (row_ok) {
          uint8_t* arow = smem + (size_t)s * stage_bytes + tr * 128;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int pc = (i + tr) & 7;       // physical 16-B chunk (staggered: no bank conflicts)
            const int lc = pc ^ (tr & 7);      // logical chunk under the 128B swizzle
            const int k0 = kc * kBK + lc * 8;
            if (k0 < p.cin) {
              uint4* ptr = reinterpret_cast<uint4*>(arow + (pc << 4));
              float a[8], g[8];
              unpack_bf16x8(*ptr, a);
              load8<float>(grow + k0, g);
#pragma unroll
              for (int e = 0; e < 8; ++e) a[e] *= g[e];
              *ptr = pack_bf16x8(a);
            }
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(&scaled_bar[s]);
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- host side --------------------------------------------------------------------------------

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

int encode_tmap_2d_bf16(void* tmap_out, const void* base, long long rows, int cols, int box_rows) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return -1;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (cols % 8) || box_rows > 256 || rows <= 0) return -2;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(reinterpret_cast<CUtensorMap*>(tmap_out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                   const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

static size_t tc_smem_bytes(int bn, int stages) {
  return 1024 /*align slack*/ + (size_t)stages * (kABytes + bn * kBK * 2) + 4 * kEpiStageBytes +
         (3 * kMaxStages + 4) * sizeof(uint64_t) + 16;
}

cudaError_t launch_pwconv_tc(const TcGemmArgs& a, cudaStream_t s) {
  if (a.m <= 0) return cudaSuccess;
  if (a.m >= (1ll << 31) - kBM || a.bn % 16 || a.bn < 16 || a.bn > 256 || a.cout_pad % a.bn || a.cin % 8 ||
      a.cout % 8)
    return cudaErrorInvalidValue;
  TcParams p;
  p.m = (int)a.m;
  p.cin = a.cin;
  p.cout = a.cout;
  p.cout_pad = a.cout_pad;
  p.bn = a.bn;
  p.m_tiles = (int)((a.m + kBM - 1) / kBM);
  p.n_tiles = a.cout_pad / a.bn;
  p.k_chunks = (a.cin + kBK - 1) / kBK;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * a.bn)) cols <<= 1;
  p.tmem_cols = cols;
  const size_t budget = 200 * 1024;
  int stages = kMaxStages;
  while (stages > 2 && tc_smem_bytes(a.bn, stages) > budget) --stages;
  p.stages = stages;
  p.bias = a.bias;
  p.gate = a.gate;
  p.rows_per_img = a.rows_per_img > 0 ? a.rows_per_img : 1;
  p.residual = reinterpret_cast<const bf16*>(a.residual);
  p.act = a.act;
  p.out = reinterpret_cast<bf16*>(a.out);
  const size_t smem = tc_smem_bytes(a.bn, stages);
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  const CUtensorMap* ta = reinterpret_cast<const CUtensorMap*>(a.tmap_a);
  const CUtensorMap* tb = reinterpret_cast<const CUtensorMap*>(a.tmap_b);
  cudaError_t e;
  if (a.gate) {
    static bool attr_done = false;
    if (!attr_done) {
      e = cudaFuncSetAttribute(pwconv_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e != cudaSuccess) return e;
      attr_done = true;
    }
    pwconv_tc_kernel<true><<<grid, 384, smem, s>>>(*ta, *tb, p);
  } else {
    static bool attr_done = false;
    if (!attr_done) {
      e = cudaFuncSetAttribute(pwconv_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e != cudaSuccess) return e;
      attr_done = true;
    }
    pwconv_tc_kernel<false><<<grid, 256, smem, s>>>(*ta, *tb, p);
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
