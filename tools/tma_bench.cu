// Microbenchmark (dev tool): per-SM TMA load / store rate as a function of the box row width.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_bench tools/tma_bench.cu -lcuda
// Each persistent CTA streams boxes [rows x row_bytes] of a 2-D bf16 tensor [R, C] (C*2 = row_bytes... or wider)
// through a 4-deep smem ring with no compute.  Reports GB/s and cycles per box row.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t su32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(su32(b)), "r"(c));
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) {
  asm volatile(
      "{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(su32(b)),
      "r"(ph)
      : "memory");
}

constexpr int kStages = 4;

// mode 0: loads, mode 1: stores
__global__ void __launch_bounds__(128) tma_kernel(const __grid_constant__ CUtensorMap tmap, int box_rows, int box_bytes,
                                                  int boxes_per_cta, int col_boxes, int mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (su32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full[kStages];
  const int stage_bytes = (box_rows * box_bytes + 1023) & ~1023;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int first = blockIdx.x * boxes_per_cta;
  if (mode == 0) {
    for (int i = 0; i < boxes_per_cta + kStages; ++i) {
      if (i >= kStages) mbar_wait(&full[(i - kStages) % kStages], ((i - kStages) / kStages) & 1);
      if (i < boxes_per_cta) {
        const int b = first + i, s = i % kStages;
        const int x = (b % col_boxes) * (box_bytes / 2), y = (b / col_boxes) * box_rows;
        mbar_expect(&full[s], box_rows * box_bytes);
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                su32(smem + s * stage_bytes)),
            "l"((uint64_t)&tmap), "r"(su32(&full[s])), "r"(x), "r"(y)
            : "memory");
      }
    }
  } else {
    for (int i = 0; i < boxes_per_cta; ++i) {
      const int b = first + i, s = i % kStages;
      const int x = (b % col_boxes) * (box_bytes / 2), y = (b / col_boxes) * box_rows;
      if (i >= kStages) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kStages - 1) : "memory");
      asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"((uint64_t)&tmap),
                   "r"(su32(smem + s * stage_bytes)), "r"(x), "r"(y)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                            const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  PFN_enc enc = (PFN_enc)fn;
  const size_t total = 1ull << 30;  // 1 GiB tensor
  void* buf;
  cudaMalloc(&buf, total);
  cudaMemset(buf, 1, total);
  cudaFuncSetAttribute(tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  // (tensor row bytes, box row bytes, box rows)
  const int cfgs[][3] = {{32, 32, 128},  {64, 64, 128},  {128, 128, 128}, {128, 128, 32}, {192, 128, 32}, {192, 64, 32},
                         {256, 128, 64}, {256, 256, 64}, {64, 64, 256},   {2304, 128, 128}, {2304, 64, 128}, {64, 64, 32},
                         {64, 128, 128}, {32, 128, 128}, {48, 128, 128}, {80, 128, 128}, {48, 64, 128}};
  for (int mode = 0; mode < 2; ++mode) {
    for (auto& c : cfgs) {
      const int row_bytes = c[0], box_bytes = c[1], box_rows = c[2];
      const cuuint64_t cols = row_bytes / 2, rows = total / row_bytes;
      CUtensorMap tm;
      cuuint64_t gdim[2] = {cols, rows};
      cuuint64_t gstr[1] = {(cuuint64_t)row_bytes};
      cuuint32_t box[2] = {(cuuint32_t)(box_bytes / 2), (cuuint32_t)box_rows};
      cuuint32_t es[2] = {1, 1};
      CUtensorMapSwizzle sw = box_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                              : box_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                              : box_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) {
        printf("encode failed %d for %d/%d/%d\n", (int)r, row_bytes, box_bytes, box_rows);
        continue;
      }
      const int col_boxes = row_bytes / box_bytes > 0 ? row_bytes / box_bytes : 1;
      const long long nboxes = (long long)(rows / box_rows) * col_boxes;
      const int grid = 148;
      long long per = nboxes / grid;
      if (per > 4000) per = 4000;
      const size_t smem = 1024 + (size_t)kStages * ((box_rows * box_bytes + 1023) & ~1023);
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        tma_kernel<<<grid, 128, smem>>>(tm, box_rows, box_bytes, (int)per, col_boxes, mode);
        cudaEventRecord(e1);
        cudaError_t ce = cudaDeviceSynchronize();
        if (ce != cudaSuccess) {
          printf("kernel failed: %s\n", cudaGetErrorString(ce));
          return 1;
        }
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double bytes = (double)per * grid * box_rows * (box_bytes < row_bytes ? box_bytes : row_bytes);
      const double rows_per_sm = (double)per * box_rows;
      printf("%s tensor_row=%4dB box=%3dB x %3d rows: %7.1f GB/s  %6.2f cycles/row (at 1.9 GHz)  %.3f ms\n",
             mode ? "STORE" : "LOAD ", row_bytes, box_bytes, box_rows, bytes / ms / 1e6, ms * 1e-3 * 1.9e9 / rows_per_sm, ms);
    }
  }
  return 0;
}
