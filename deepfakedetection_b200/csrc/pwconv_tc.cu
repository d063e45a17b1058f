// Pointwise (1x1) conv as a tcgen05 GEMM for sm_100a -- the bf16 throughput path.
//   out[m, co] = act( sum_k (x[m,k] * gate[m / rows_per_img, k]) * w[co,k] + bias[co] ) (+ residual[m,co])
// Replaces timm `conv_pw` (+bn1+SiLU), `conv_pwl` (+bn3, SE gate, skip add) and `conv_head`
//  inside the reference's `model(images)`
// (/root/reference/orchestration/orchestrator.py:590).
//
// Shape of the problem: M = n*H*W rows (up to 3.2 M), K = Cin in [16, 1152], N = Cout in [16, 1280].
// Every one of these GEMMs except conv_head is HBM/L2-bound with a tiny K x N, so the kernel is a
// streaming pipeline whose cost is the EPILOGUE, not the MMA:
//   * persistent CTAs, static round-robin over (m_tile, n_tile) tiles of 128 x BN;
//   * warps 0, 2, 3: TMA producers (tiles dealt round-robin) -- A tile [128 x 64] and B tile [BN x 64] per
//     K chunk, 128B swizzle, multi-stage mbarrier ring; K / M / N tails zero-filled by TMA OOB handling;
//   * warp 1: one lane issues tcgen05.mma (M=128, N=BN, K=16) into a double-buffered TMEM
//     accumulator and commits to the ring's `empty` barriers and to `tmem_full`;
//   * warps 4-7 and 8-11: TWO epilogue groups, one per TMEM accumulator, alternating tiles --
//     tcgen05.ld 32x32b, bias from shared memory, packed f32x2 math, SiLU = h + h*tanh(h) (one MUFU),
//     +residual, cvt.rn.bf16x2, swizzled staging tile, then ONE TMA store (cp.async.bulk.tensor
//     shared -> global) per 32 x 64 block: no per-element address math, M / N tails clipped by the
//     tensor map;
//   * (gated variant) warps 12-15 multiply the landed A tile in place by the per-image fp32 SE gate
//     (packed f32x2 multiplies, product rounded once to bf16) and hand it to the MMA warp
//     through a second barrier (fence.proxy.async before the arrive).
// A [m, cin] and W [cout_pad, cin] are bf16 row-major (K-major operands, no transposes anywhere).
#include "common.cuh"
#include "kernels.h"
#include "tc.cuh"

#include <stdio.h>
#include <stdlib.h>

#include <mutex>

namespace b200df {

using namespace tc;

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kABytes = kBM * kBK * 2;  // 16 KB
constexpr int kMaxStages = 8;
constexpr int kScalWarps = 8;             // gated kernels: SE-gate scaler warps (the bottleneck role of the K-heavy 14^2 / 7^2 layers)
constexpr int kEpiWarps = 8;              // gated kernels: two groups of four (one group per TMEM accumulator)
constexpr int kEpiWarpsWide = 16;         // ungated kernels (expand, conv_head: wide N, SiLU): two column halves per group
constexpr int kEpiStageBytes = 32 * 128;  // one staging tile: 32 rows x 64 bf16; TWO per epilogue warp
constexpr int kMaxBiasFloats = 3072;  // widest layer: FasterViT-2 level-3 MLP fc1 768 -> 3072 (EfficientNet-B3: 2304)

struct TcParams {
  int m, cin, cout, cout_pad, bn;
  int rpi;            // rows per image (ungated: m, one "image")
  int tiles_per_img;  // ceil(rpi / 128): M tiles never straddle images
  int acc_stride;     // TMEM columns between the two accumulators (bn rounded up to 32)
  int b_res;          // scale_b only: the pristine weight matrix (all K chunks) stays resident in shared memory
  int prof_slot;      // TC_PROF builds: row of g_tc_prof this launch writes
  int np;             // TMA producer lanes (1..3); stages % np == 0
  int sg;             // gated: scaler groups (1, 2 or 4); stages % sg == 0
  int b_once;         // weights fit one (k chunk, n tile): load the B tile once per ring slot, not per tile
  int debug;          // B200DF_TC_DEBUG bits (experiments only): 1 = skip the output stores
  int scale_b;        // gated only: 1 = fold the gate into the B (weight) tile, 0 = into the A tile
  int m_tiles, n_tiles, k_chunks, stages;
  uint32_t tmem_cols;
  const float* bias;
  const float* gate;  // [m / rows_per_img, cin] fp32 or nullptr
  const bf16* residual;
};

typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t f2_pack(float lo, float hi) {
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(f32x2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2_t f2_add(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t f2_mul(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t f2_fma(f32x2_t a, f32x2_t b, f32x2_t c) {
  f32x2_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// two bf16 (one word) times two fp32 gates, product rounded once to bf16
__device__ __forceinline__ uint32_t scale_bf16x2(uint32_t a, float g0, float g1) {
  const f32x2_t pr = f2_mul(f2_pack(__uint_as_float(a << 16), __uint_as_float(a & 0xffff0000u)), f2_pack(g0, g1));
  float lo, hi;
  f2_unpack(pr, lo, hi);
  return pack_bf16x2(lo, hi);
}

// (n tile, M sub-tile, image) of a CTA's current tile, advanced by the grid stride in mixed radix: every role
// is a lone warp per scheduler (~0.2 IPC), so the per-tile bookkeeping has no divisions.
// Role-level timing (dev builds only: `make EXTRA=-DTC_PROF`): lane 0 of one warp per role accumulates the
// cycles it spends inside each mbarrier wait; tools/tc_prof.py prints the table.
#ifdef TC_PROF
constexpr int kProfLaunches = 64;
__device__ unsigned long long g_tc_prof[kProfLaunches * 148 * 32];
#define PROF_DECL unsigned long long prof_w[7] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull}; const long long prof_t0 = clock64();
#define PROF_WAIT(i, ...) { const long long prof_c0 = clock64(); __VA_ARGS__; prof_w[i] += (unsigned long long)(clock64() - prof_c0); }
#define PROF_BEGIN long long prof_b0 = clock64();
#define PROF_LAP(i) { const long long prof_b1 = clock64(); prof_w[i] += (unsigned long long)(prof_b1 - prof_b0); prof_b0 = prof_b1; }
#define PROF_END(role)                                                                          \
  if (lane == 0) {                                                                              \
    unsigned long long* g = g_tc_prof + ((size_t)p.prof_slot * 148 + blockIdx.x) * 32 + (role) * 8; \
    g[0] = (unsigned long long)(clock64() - prof_t0);                                           \
    for (int i_ = 0; i_ < 7; ++i_) g[1 + i_] = prof_w[i_];                                       \
  }
#else
#define PROF_DECL
#define PROF_WAIT(i, ...) { __VA_ARGS__; }
#define PROF_BEGIN
#define PROF_LAP(i)
#define PROF_END(role)
#endif

struct TileIter {
  int nt, sub, img, d_nt, d_sub, d_img, n_tiles, tiles_per_img;
  __device__ __forceinline__ TileIter(int t0, int step, int n_tiles_, int tiles_per_img_)
      : n_tiles(n_tiles_), tiles_per_img(tiles_per_img_) {
    nt = t0 % n_tiles;
    int mt = t0 / n_tiles;
    img = mt / tiles_per_img;
    sub = mt - img * tiles_per_img;
    d_nt = step % n_tiles;
    mt = step / n_tiles;
    d_img = mt / tiles_per_img;
    d_sub = mt - d_img * tiles_per_img;
  }
  __device__ __forceinline__ void next() {
    nt += d_nt;
    sub += d_sub;
    img += d_img;
    if (nt >= n_tiles) { nt -= n_tiles; ++sub; }
    if (sub >= tiles_per_img) { sub -= tiles_per_img; ++img; }
  }
};

template <bool kGate, int kAct, bool kRes>
__global__ void __launch_bounds__(640, 1)
pwconv_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_out, const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-B align the operand ring (128B swizzle atoms are 1024 B)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int b_bytes = p.bn * kBK * 2;
  const int stage_bytes = kABytes + b_bytes;
  uint8_t* bres = smem + (size_t)p.stages * stage_bytes;      // 1024-aligned: stage_bytes % 1024 == 0
  uint8_t* epi_smem = bres + (p.b_res ? (size_t)p.k_chunks * b_bytes : 0);
  float* sbias = reinterpret_cast<float*>(epi_smem + (kGate ? 1 : 2) * kEpiWarps * kEpiStageBytes);
  float* sgate = sbias + kMaxBiasFloats;  // [stages][64] fp32: the SE gate slice of every ring slot's (image, K chunk)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sgate + kMaxStages * kBK);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kMaxStages;
  uint64_t* scaled_bar = bars + 2 * kMaxStages;
  uint64_t* tmem_full = bars + 3 * kMaxStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* bres_bar = tmem_empty + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bres_bar + 1);

  // warp index through a shuffle: tells ptxas it is warp-uniform, so the role loops run on uniform registers
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int num_tiles = p.m_tiles * p.n_tiles;
  // static round-robin tile order: at any moment the 148 CTAs stream through one contiguous window of the
  // activation tensor (DRAM-page friendly; measured faster than a contiguous range per CTA)
  const int t_begin = blockIdx.x, t_end = num_tiles, t_step = gridDim.x;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_out);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&scaled_bar[s], kScalWarps / p.sg);  // one arrival per warp of the group that owns the chunk
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], kGate ? 128 : 256);
    }
    mbar_init(bres_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, p.tmem_cols);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < p.cout_pad; i += blockDim.x) sbias[i] = __ldg(p.bias + i);
  // K tails that are not a multiple of 16 (cin = 40, 24: the first EfficientNet-B3 blocks): the last MMA step reads
  // 8 columns past cin, where the operand tiles are TMA-zero-filled but the gate slice holds whatever shared memory
  // held (the bulk copy writes only cin - c_k floats) -- 0 x NaN garbage = NaN.  Start from zeros: entries that go
  // stale later are finite gates.
  if constexpr (kGate)
    for (int i = threadIdx.x; i < kMaxStages * kBK; i += blockDim.x) sgate[i] = 0.f;
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_ptr, 0);
  griddep_wait();               // everything above touched only weights / shared memory / TMEM
  griddep_launch_dependents();

  if (warp == 0 || warp == 2 || warp == 3) {
    // ================= TMA producers (up to three warps, one lane each) =================
    // The flattened (tile, K chunk) sequence is dealt round-robin to p.np producer lanes.  p.stages is a multiple
    // of p.np, hence every ring slot has exactly ONE producer and the parity waits stay unambiguous (a producer
    // two ring revolutions ahead of another on the same slot would alias mbarrier phases).  Measured: with the
    // warp-uniform loop one lane already keeps the ring full (it waits on `empty` 60-90 % of the time); the
    // extra lanes only matter for the thinnest tiles.
    const int pw = warp == 0 ? 0 : warp - 1;
    if (pw < p.np) {
      if (p.b_res && pw == 0 && elect_one()) {  // whole weight matrix, once per CTA (148 CTAs re-reading the
                                                // same few KB per tile serialise on one L2 slice)
        mbar_arrive_expect_tx(bres_bar, (uint32_t)(p.k_chunks * b_bytes));
        for (int kc = 0; kc < p.k_chunks; ++kc) tma_load_2d(bres + (size_t)kc * b_bytes, &tmap_b, kc * kBK, 0, bres_bar);
      }
      __syncwarp();
      PROF_DECL
      int it = pw;               // flattened (tile, K chunk) index of this producer's next load
      int s = pw;                // = it % stages (pw < np <= stages)
      uint32_t ph = 0;
      int kc = pw, t = t_begin;
      TileIter ti(t_begin, t_step, p.n_tiles, p.tiles_per_img);
      while (kc >= p.k_chunks) { kc -= p.k_chunks; t += t_step; ti.next(); }
      while (t < t_end) {
        PROF_WAIT(0, mbar_wait(&empty_bar[s], ph ^ 1))
        uint8_t* sa = smem + (size_t)s * stage_bytes;
        const bool load_b = !p.b_res && (!p.b_once || it < p.stages);  // b_once: every slot keeps the B tile of its first use
        const int c_k = kc * kBK, c_row = ti.sub * kBM, c_img = ti.img, c_n = ti.nt * p.bn;
        // gated: the image's gate slice for this K chunk rides the same barrier (the scaler warps then read it
        // from shared memory instead of stalling on an L2 round trip per chunk)
        const uint32_t g_bytes = kGate ? (uint32_t)min(kBK, p.cin - c_k) * 4u : 0u;
        const uint32_t tx_bytes = (uint32_t)(load_b ? stage_bytes : kABytes) + g_bytes;
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[s], tx_bytes);
          tma_load_3d(sa, &tmap_a, c_k, c_row, c_img, &full_bar[s]);
          if constexpr (kGate) bulk_load_1d(sgate + s * kBK, p.gate + (size_t)c_img * p.cin + c_k, g_bytes, &full_bar[s]);
          if (load_b) tma_load_2d(sa + kABytes, &tmap_b, c_k, c_n, &full_bar[s]);
        }
        __syncwarp();
        it += p.np;
        s += p.np;
        if (s >= p.stages) { s -= p.stages; ph ^= 1; }
        kc += p.np;
        while (kc >= p.k_chunks) { kc -= p.k_chunks; t += t_step; ti.next(); }
      }
      if (pw == 0) { PROF_END(0) }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = make_idesc_bf16_m128((uint32_t)p.bn);
    const uint64_t desc0 = make_sw128_kmajor_desc(smem_u32(smem));  // A tile of ring slot 0
    const int stage16 = stage_bytes >> 4;
    int s = 0;
    uint32_t ph = 0;
    int lt = 0;
    PROF_DECL
    for (int t = t_begin; t < t_end; t += t_step, ++lt) {
      const int acc = lt & 1;
      const uint32_t acc_ph = (lt >> 1) & 1;
      PROF_WAIT(0, mbar_wait(&tmem_empty[acc], acc_ph ^ 1))
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.acc_stride);
      for (int kc = 0; kc < p.k_chunks; ++kc) {
        PROF_WAIT(1, mbar_wait(kGate ? &scaled_bar[s] : &full_bar[s], ph))
        tcgen05_fence_after();
        PROF_BEGIN
        const uint64_t da = desc0 + (uint64_t)(uint32_t)(s * stage16);  // (addr >> 4) field; smem < 256 KB: no carry
        const uint64_t db = da + (uint64_t)(kABytes >> 4);
        const int krem = p.cin - kc * kBK;
        const uint32_t acc0 = kc ? 1u : 0u;
        if (elect_one()) {
          // +32 bytes per K=16 step inside the swizzle atom -> +2 in the (addr >> 4) field
          if (krem >= kBK) {
            umma_bf16(d_tmem, da, db, idesc, acc0);
            umma_bf16(d_tmem, da + 2ull, db + 2ull, idesc, 1u);
            umma_bf16(d_tmem, da + 4ull, db + 4ull, idesc, 1u);
            umma_bf16(d_tmem, da + 6ull, db + 6ull, idesc, 1u);
          } else {
            const int nk = (krem + 15) >> 4;
            for (int k = 0; k < nk; ++k)
              umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kc | k) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (kc == p.k_chunks - 1) umma_commit(&tmem_full[acc]);
        }
        __syncwarp();
        PROF_LAP(2)
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
    PROF_END(1)
  } else if (warp >= 4 && warp < 4 + (kGate ? kEpiWarps : kEpiWarpsWide)) {
    // ================= epilogue (group = accumulator; ungated: two column-half teams per group) =================
    constexpr int kHalves = kGate ? 1 : 2;   // teams of 4 warps per accumulator, alternating 64-column blocks
    constexpr int kBufs = 1;                 // one 4-KB staging area per warp (gated: 32 KB, ungated: 64 KB)
    constexpr int kNB = 2 * kBufs;           // ... used as 2-KB buffers of one [32 x 32-column] store box each
    const int ew = warp - 4;
    const int grp = ew / (4 * kHalves);
    const int half = (ew >> 2) % kHalves;
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    uint8_t* stg_base = epi_smem + ew * kBufs * kEpiStageBytes;
    const f32x2_t half2 = f2_pack(0.5f, 0.5f);
    int stores = 0;  // TMA stores issued by this warp (staging buffer = stores & 1)
    int lt = 0;
    PROF_DECL
    // this group's tiles: every second tile of the CTA, starting at t_begin + grp * t_step
    TileIter ti(t_begin + grp * t_step, 2 * t_step, p.n_tiles, p.tiles_per_img);
    for (int t = t_begin + grp * t_step; t < t_end; t += 2 * t_step, lt += 2, ti.next()) {
      const int nt = ti.nt, img = ti.img;
      const uint32_t acc_ph = (lt >> 1) & 1;
      PROF_WAIT(0, mbar_wait(&tmem_full[grp], acc_ph))
      tcgen05_fence_after();
      const int lrow0 = ti.sub * kBM + q * 32;  // first row of this warp inside the image
      const bool row_ok = lrow0 + lane < p.rpi;
      const size_t row = (size_t)img * p.rpi + lrow0 + lane;         // global row of this lane
      const int ncol0 = nt * p.bn;
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * p.acc_stride);
      if (half * 64 >= p.bn) {  // this team has no column block in this tile: release the accumulator at once
        tcgen05_fence_before();
        mbar_arrive(&tmem_empty[grp]);
      }
      for (int cb = half * 64; cb < p.bn; cb += 64 * kHalves) {
        const int ncols = min(64, p.bn - cb);  // multiple of 16
#pragma unroll
        for (int hb = 0; hb < 2; ++hb) {  // two 32-column halves, each with its own staging buffer and TMA store
          if (hb * 32 < ncols) {
            uint8_t* stg = stg_base + (stores % kNB) * (kEpiStageBytes / 2);
            if (stores >= kNB) {  // the store that last used this staging buffer must have finished READING it
              PROF_WAIT(1, if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kNB - 1) : "memory");
                        __syncwarp())
            }
            const int col = ncol0 + cb + hb * 32;  // global output column of r[0]
            uint32_t r[32];
            tmem_ld32(t_row + (uint32_t)(cb + hb * 32), r);
            // residual for this lane's row: issued before the TMEM wait so the two latencies overlap
            uint4 res[4];
            if constexpr (kRes) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                res[c] = make_uint4(0u, 0u, 0u, 0u);
                if (row_ok && col + 8 * c < p.cout)
                  res[c] = __ldg(reinterpret_cast<const uint4*>(p.residual + row * p.cout + col + 8 * c));
              }
            }
            PROF_WAIT(2, tmem_ld_wait())
            if (cb + 64 * kHalves >= p.bn && hb * 32 + 32 >= ncols) {  // this warp's last read of the accumulator
              tcgen05_fence_before();
              mbar_arrive(&tmem_empty[grp]);
            }
            f32x2_t v[16];
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
              const float4 b4 = *reinterpret_cast<const float4*>(sbias + col + 4 * j4);  // smem broadcast
              v[2 * j4] = f2_add(f2_pack(__uint_as_float(r[4 * j4]), __uint_as_float(r[4 * j4 + 1])),
                                 f2_pack(b4.x, b4.y));
              v[2 * j4 + 1] = f2_add(f2_pack(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])),
                                     f2_pack(b4.z, b4.w));
            }
            if constexpr (kAct == 1) {  // SiLU(x) = h + h tanh(h), h = x/2: all MUFUs first, then the FMAs
              float th[32];
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                v[j] = f2_mul(v[j], half2);
                float h0, h1;
                f2_unpack(v[j], h0, h1);
                th[2 * j] = tanh_fast(h0);
                th[2 * j + 1] = tanh_fast(h1);
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = f2_fma(v[j], f2_pack(th[2 * j], th[2 * j + 1]), v[j]);
            }
            if constexpr (kAct == 2) {  // GELU, tanh form: h + h tanh(h (c1 + c3 h^2)), h = x/2 (common.cuh gelu_fast)
              float th[32];
              const f32x2_t c1 = f2_pack(1.5957691216f, 1.5957691216f), c3 = f2_pack(0.2854191951f, 0.2854191951f);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                v[j] = f2_mul(v[j], half2);
                const f32x2_t u = f2_mul(v[j], f2_fma(f2_mul(v[j], v[j]), c3, c1));
                float u0, u1;
                f2_unpack(u, u0, u1);
                th[2 * j] = tanh_fast(u0);
                th[2 * j + 1] = tanh_fast(u1);
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = f2_fma(v[j], f2_pack(th[2 * j], th[2 * j + 1]), v[j]);
            }
            if constexpr (kAct == 3) {  // ReLU
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                float a0, a1;
                f2_unpack(v[j], a0, a1);
                v[j] = f2_pack(fmaxf(a0, 0.f), fmaxf(a1, 0.f));
              }
            }
            if constexpr (kRes) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const uint32_t rw[4] = {res[c].x, res[c].y, res[c].z, res[c].w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  v[4 * c + j] = f2_add(v[4 * c + j], f2_pack(__uint_as_float(rw[j] << 16),
                                                              __uint_as_float(rw[j] & 0xffff0000u)));
              }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint4 pk;
              float a0, a1;
              f2_unpack(v[4 * c + 0], a0, a1); pk.x = pack_bf16x2(a0, a1);
              f2_unpack(v[4 * c + 1], a0, a1); pk.y = pack_bf16x2(a0, a1);
              f2_unpack(v[4 * c + 2], a0, a1); pk.z = pack_bf16x2(a0, a1);
              f2_unpack(v[4 * c + 3], a0, a1); pk.w = pack_bf16x2(a0, a1);
              // [32 rows x 64 B] staging tile in the 64-B swizzle of the store's tensor map
              *reinterpret_cast<uint4*>(stg + lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4)) = pk;
            }
            // staging buffer -> global: one TMA store of the [32 rows x 32 cols] box (rows past the image and
            // columns >= cout are clipped by the tensor map; columns past `ncols` only exist when they are
            // >= cout, see choose_bn)
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && !(p.debug & 1)) {
              asm volatile(
                  "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                      reinterpret_cast<uint64_t>(&tmap_out)),
                  "r"(smem_u32(stg)), "r"(col), "r"(lrow0), "r"(img)
                  : "memory");
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            ++stores;
          }
        }
      }
    }
    if (ew == 0) { PROF_END(2) }
    if (stores > 0 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  } else if (kGate && warp >= 4 + kEpiWarps) {
    // ================= SE-gate scaler =================
    // M tiles never straddle images, so ONE gate vector applies to a whole tile and the gate can be folded
    // into whichever operand tile is smaller:  sum_k (a_k g_k) w_k  ==  sum_k a_k (g_k w_k).
    //   scale_b: B tile [BN x 64] (weights) *= gate   -- the big-spatial layers, BN = 16..96 rows
    //   else   : A tile [rows x 64] (activations) *= gate -- only the image's valid rows
    // A thread owns ONE 16-byte K chunk (8 channels): its 8 fp32 gate values come from the ring slot's gate
    // slice (bulk-copied by the producer on the same barrier as the tile) and stay in registers for all of
    // its rows (rows ts/8, ts/8 + 32, ...).
    // The warps form kScalGroups groups that take the (tile, K chunk) sequence in turn: a chunk's path through
    // a scaler (barrier wait -> gate + tile loads -> multiply -> stores -> proxy fence -> arrive) is one
    // dependent chain of ~900 cycles whatever the row count, so what matters is how many chunks are in
    // flight, not how many threads share one chunk.
    const int kScalGroups = p.sg;                        // every ring slot belongs to ONE group (stages % sg == 0)
    const int kTG = 32 * kScalWarps / kScalGroups;       // threads per group
    const int kRowsPass = kTG / 8;                       // rows covered by one pass of a group (16 or 32)
    const int tsa = threadIdx.x - 32 * (4 + kEpiWarps);
    const int sgrp = tsa / kTG;
    const int ts = tsa - sgrp * kTG;
    const int lc = ts & 7;          // logical 16-B chunk inside the 128-B swizzled row
    const int rg = ts >> 3;         // first row; rows rg + kRowsPass * j
    const int pc = lc ^ (rg & 7);   // physical chunk (kRowsPass % 8 == 0: every row of this thread has rg & 7)
    PROF_DECL
    PROF_WAIT(1, if (p.b_res) mbar_wait(bres_bar, 0))
    // (n tile, M sub-tile, image) of the current tile, advanced by the grid stride in mixed radix (no divisions)
    TileIter ti(t_begin, t_step, p.n_tiles, p.tiles_per_img);
    int s = sgrp;                   // ring slot of this group's next chunk (kScalGroups <= stages)
    uint32_t ph = 0;
    int kc = sgrp, t = t_begin;
    while (kc >= p.k_chunks) { kc -= p.k_chunks; t += t_step; ti.next(); }
    while (t < t_end) {
      const int nrows = p.scale_b ? p.bn : min(kBM, p.rpi - ti.sub * kBM);  // rows of the tile to scale
      // every 16-B chunk the MMA will read: K tails are issued in steps of 16 columns, so for cin % 16 == 8 the
      // last step reads one chunk past cin.  That chunk is zero in the source (TMA fill) and must be WRITTEN when
      // the scaled copy goes to a ring slot no TMA load ever touched (b_res); its gate entries are the zeros the
      // prologue put there.
      const int krem = p.cin - kc * kBK;
      const bool k_ok = lc < (krem >= kBK ? 8 : 2 * ((krem + 15) >> 4));
      PROF_WAIT(0, mbar_wait(&full_bar[s], ph))
      PROF_BEGIN
      if (k_ok) {
        const float4* sg = reinterpret_cast<const float4*>(sgate + s * kBK) + 2 * lc;
        const float4 g0 = sg[0], g1 = sg[1];
        uint8_t* base = smem + (size_t)s * stage_bytes + (p.scale_b ? kABytes : 0) + rg * 128 + (pc << 4);
        // source: the tile itself (scaled in place) or the resident pristine weights
        const uint8_t* src = p.b_res ? bres + (size_t)kc * b_bytes + rg * 128 + (pc << 4) : base;
        for (int r0 = rg; r0 < nrows; r0 += 4 * kRowsPass, base += 4 * kRowsPass * 128, src += 4 * kRowsPass * 128) {
          uint4 a[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (r0 + kRowsPass * j < nrows) a[j] = *reinterpret_cast<const uint4*>(src + j * kRowsPass * 128);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (r0 + kRowsPass * j < nrows) {
              a[j].x = scale_bf16x2(a[j].x, g0.x, g0.y);
              a[j].y = scale_bf16x2(a[j].y, g0.z, g0.w);
              a[j].z = scale_bf16x2(a[j].z, g1.x, g1.y);
              a[j].w = scale_bf16x2(a[j].w, g1.z, g1.w);
              *reinterpret_cast<uint4*>(base + j * kRowsPass * 128) = a[j];
            }
          }
        }
        PROF_LAP(2)
        fence_proxy_async_smem();
        PROF_LAP(3)
      }
      __syncwarp();  // every lane's writes + proxy fence precede the warp's single arrival
      if (lane == 0) mbar_arrive(&scaled_bar[s]);
      PROF_LAP(4)
      s += kScalGroups;
      if (s >= p.stages) { s -= p.stages; ph ^= 1; }
      kc += kScalGroups;
      while (kc >= p.k_chunks) { kc -= p.k_chunks; t += t_step; ti.next(); }
    }
    if (tsa < 32) { PROF_END(3) }
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

int encode_tmap_act_bf16(void* tmap_out, const void* base, long long m, int cols, int rows_per_img, int box_rows) {
  const long long rpi = rows_per_img > 0 ? rows_per_img : m;
  if (m <= 0 || m % rpi || (cols % 8) || box_rows > 256) return -2;
  unsigned long long dims[3] = {(unsigned long long)cols, (unsigned long long)rpi, (unsigned long long)(m / rpi)};
  unsigned long long str[3] = {2, (unsigned long long)cols * 2, (unsigned long long)rpi * cols * 2};
  // box_rows == 128: operand tiles, 64 columns (128-B rows, 128-B swizzle)
  // box_rows == 32 : epilogue store boxes, 32 columns (64-B rows, 64-B swizzle) -- half a 64-column block, so
  //                  that a warp's two staging buffers alternate inside one block
  const bool store_box = box_rows == 32;
  unsigned box[3] = {(unsigned)(store_box ? 32 : kBK), (unsigned)box_rows, 1};
  return encode_tmap_nd_bf16(tmap_out, base, 3, dims, str, box, store_box ? 64 : 128);
}

static size_t tc_smem_bytes(int bn, int stages, size_t bres_bytes, bool gated) {
  return 1024 /*align slack*/ + (size_t)stages * (kABytes + bn * kBK * 2) + bres_bytes +
         (gated ? 1 : 2) * kEpiWarps * kEpiStageBytes + (kMaxBiasFloats + kMaxStages * kBK) * sizeof(float) +
         (3 * kMaxStages + 5) * sizeof(uint64_t) + 16;
}

cudaError_t launch_pwconv_tc(const TcGemmArgs& a, cudaStream_t s) {
  if (a.m <= 0) return cudaSuccess;
  if (a.m >= (1ll << 31) - kBM || a.bn % 16 || a.bn < 16 || a.bn > 256 || a.cout_pad % a.bn || a.cin % 8 ||
      a.cout % 8 || a.cout_pad > kMaxBiasFloats || (a.cout_pad != a.bn && a.bn % 64))
    return cudaErrorInvalidValue;
  if (a.gate && (reinterpret_cast<uintptr_t>(a.gate) & 15)) return cudaErrorInvalidValue;  // bulk-copied in 16-B units
  TcParams p;
  p.m = (int)a.m;
  p.cin = a.cin;
  p.cout = a.cout;
  p.cout_pad = a.cout_pad;
  p.bn = a.bn;
  // M tiles are image-aligned (ungated: one image of m rows)
  const long long rpi = (a.gate && a.rows_per_img > 0) ? a.rows_per_img : a.m;
  if (a.m % rpi) return cudaErrorInvalidValue;
  p.rpi = (int)rpi;
  p.tiles_per_img = (int)((rpi + kBM - 1) / kBM);
  p.m_tiles = (int)(a.m / rpi) * p.tiles_per_img;
  // fold the gate into the smaller operand tile: B has bn rows, A has min(128, rpi) valid rows
  p.scale_b = (a.gate && a.bn <= 96 && a.bn < (rpi < kBM ? rpi : kBM)) ? 1 : 0;
  p.n_tiles = a.cout_pad / a.bn;
  static int dbg = -1;
  if (dbg < 0) {
    const char* e = getenv("B200DF_TC_DEBUG");
    dbg = e ? atoi(e) : 0;
  }
  p.debug = dbg;
  p.b_once = (p.n_tiles == 1 && a.cin <= kBK && !p.scale_b && !(dbg & 2)) ? 1 : 0;
  p.k_chunks = (a.cin + kBK - 1) / kBK;
  p.acc_stride = (a.bn + 31) / 32 * 32;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * p.acc_stride)) cols <<= 1;
  p.tmem_cols = cols;
  const size_t budget = 220 * 1024;
  size_t bres_bytes = (size_t)p.k_chunks * a.bn * kBK * 2;
  p.b_res = (p.scale_b && bres_bytes <= 40 * 1024 && !(dbg & 4)) ? 1 : 0;
  if (!p.b_res) bres_bytes = 0;
  int stages = kMaxStages;
  const bool gated = a.gate != nullptr;
  while (stages > 2 && tc_smem_bytes(a.bn, stages, bres_bytes, gated) > budget) --stages;
  if (p.b_res && p.k_chunks <= stages) stages = stages / p.k_chunks * p.k_chunks;  // slot <-> K chunk mapping fixed
  // producer lanes: every ring slot must belong to one producer (stages % np == 0)
  static int np_max = -1;
  if (np_max < 0) {
    const char* e = getenv("B200DF_TC_NP");
    np_max = e ? atoi(e) : 3;
    if (np_max < 1 || np_max > 3) np_max = 3;
  }
  if (np_max >= 2 && !p.b_res && stages >= 5 && stages % 2 && stages % 3) --stages;  // 5 -> 4, 7 -> 6
  // gated: two scaler groups alternate over the K chunks when every ring slot can belong to one group
  if (gated && stages % 2 && stages >= 5 && !(p.b_res && p.k_chunks <= stages && (stages - 1) % p.k_chunks)) --stages;
  p.sg = (gated && stages % 2 == 0) ? 2 : 1;
  {
    static int sg_max = -1;
    if (sg_max < 0) {
      const char* e = getenv("B200DF_TC_SG");
      sg_max = e ? atoi(e) : 2;
    }
    if (gated && sg_max >= 4 && stages % 4 == 0) p.sg = 4;
    if (sg_max == 1) p.sg = 1;
  }
  p.np = (np_max >= 3 && stages % 3 == 0) ? 3 : ((np_max >= 2 && stages % 2 == 0) ? 2 : 1);
  p.stages = stages;
#ifdef TC_PROF
  static int prof_launch = 0;
  p.prof_slot = prof_launch++ % kProfLaunches;
  if (getenv("B200DF_TC_PROF_PRINT"))
    fprintf(stderr, "tc_prof slot %d: m=%lld cin=%d cout=%d bn=%d stages=%d np=%d gate=%d scale_b=%d b_res=%d b_once=%d\n",
            p.prof_slot, (long long)a.m, a.cin, a.cout, a.bn, stages, p.np, a.gate != nullptr, p.scale_b, p.b_res, p.b_once);
#endif
  p.bias = a.bias;
  p.gate = a.gate;
  p.residual = reinterpret_cast<const bf16*>(a.residual);
  const size_t smem = tc_smem_bytes(a.bn, stages, bres_bytes, gated);
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  const CUtensorMap* ta = reinterpret_cast<const CUtensorMap*>(a.tmap_a);
  const CUtensorMap* tb = reinterpret_cast<const CUtensorMap*>(a.tmap_b);
  const CUtensorMap* to = reinterpret_cast<const CUtensorMap*>(a.tmap_out);
  const bool res = a.residual != nullptr;
  const int threads = 640;  // gated: 4 + 8 epilogue + 8 scaler warps; ungated: 4 + 16 epilogue warps
#define TC_LAUNCH(G_, A_, R_)                                                                                  \
  if ((a.gate != nullptr) == G_ && a.act == A_ && res == R_) {                                                 \
    static PerDeviceOnce once;                                                                                 \
    if (once.need()) {                                                                                         \
      cudaError_t e = cudaFuncSetAttribute(pwconv_tc_kernel<G_, A_, R_>,                                       \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);           \
      if (e != cudaSuccess) return e;                                                                          \
    }                                                                                                          \
    count_launch();                                                                                            \
    return launch_pdl(pwconv_tc_kernel<G_, A_, R_>, dim3(grid), dim3(threads), smem, s, *ta, *tb, *to, p);     \
  }
  TC_LAUNCH(false, 1, false)  // expand
  TC_LAUNCH(false, 0, false)  // conv_head
  TC_LAUNCH(true, 0, false)   // project
  TC_LAUNCH(true, 0, true)    // project + skip
  TC_LAUNCH(false, 0, true)
  TC_LAUNCH(false, 1, true)
  TC_LAUNCH(true, 1, false)
  TC_LAUNCH(true, 1, true)
  TC_LAUNCH(false, 2, false)  // EfficientFormerV2 / FasterViT fc1 (GELU)
  TC_LAUNCH(false, 2, true)
  TC_LAUNCH(false, 3, false)
#undef TC_LAUNCH
  return cudaErrorInvalidValue;
}

}  // namespace b200df

#ifdef TC_PROF
// dev builds only (not part of include/b200df.h): copies the role-timing table to the host
extern "C" __attribute__((visibility("default"))) int b200df_tc_prof_read(unsigned long long* out) {
  return (int)cudaMemcpyFromSymbol(out, b200df::g_tc_prof, sizeof(b200df::g_tc_prof));
}
#endif
