// Host-side launcher declarations shared by the engine and the C-ABI per-kernel entry points.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "act.h"

namespace b200df {

enum DType : int { kBF16 = 0, kFP32 = 1 };
inline size_t dtype_size(int dt) { return dt == kBF16 ? 2 : 4; }

// thread-local launch counter (bench.py's gpu_launches)
void count_launch(int n = 1);

// Programmatic dependent launch (PDL): the hot-path kernels run `griddepcontrol.wait` before they touch any
// activation and trigger `griddepcontrol.launch_dependents` right after it, so the NEXT kernel's CTAs are placed
// as SMs free up and its prologue (barrier init, TMEM alloc, weight/bias staging, tensor-map prefetch) overlaps
// this kernel's tail.  Only kernels that contain the wait are launched through launch_pdl.  B200DF_PDL=0 -> off.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE attribute: a launcher that opts a kernel into
// more than 48 KB of dynamic shared memory must do so once per (kernel instantiation, device), not once per process.
struct PerDeviceOnce {
  bool done[64] = {};
  bool need() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
    if (done[d]) return false;
    done[d] = true;
    return true;
  }
};

// error plumbing: launchers return cudaError_t of the launch (cudaGetLastError()).
// oc: stem channels (32 | 40).  pad: zero rows/columns BEFORE the image (1 symmetric, 0 = TF-"same" (0, 1)).
// act: common.cuh Act (1 SiLU | 2 GELU | 3 ReLU) after the folded bias.  oc: 32 | 40 | 16 (EfficientFormerV2) | 64 (FasterViT).
cudaError_t launch_stem(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias,
                        void* out, int dtype, cudaStream_t s, int oc = 32, int pad = 1, int act = 1);

// u8 HWC input, fused ToTensor(/255)+Normalize: in_scale[c] = 1/(255*std[c]), in_shift[c] = -mean[c]/std[c]
cudaError_t launch_stem_u8(const uint8_t* x_hwc, int n, int h, int w, const float* wgt, const float* bias,
                           float sc0, float sc1, float sc2, float sh0, float sh1, float sh2, void* out,
                           int dtype, cudaStream_t s, int oc = 32, int pad = 1, int act = 1);

// Pooled sums are Q24 fixed-point int64 (common.cuh: pool_t), accumulated with integer atomics.
typedef long long pool_t;
// bf16 output only: implicit GEMM on tcgen05 (stem_tc.cu); launch_stem / launch_stem_u8 dispatch to these
bool stem_tc_supported(int oc, int pad, int act);
cudaError_t launch_stem_tc(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias, void* out,
                           cudaStream_t s, int oc = 32, int pad = 1, int act = 1);
cudaError_t launch_stem_tc_u8(const uint8_t* x_hwc, int n, int h, int w, const float* wgt, const float* bias,
                              float sc0, float sc1, float sc2, float sh0, float sh1, float sh2, void* out,
                              cudaStream_t s, int oc = 32, int pad = 1, int act = 1);

// pad_lo: zero rows/columns BEFORE the image; -1 = symmetric (k-1)/2.  The pad after is k - 1 - pad_lo, so the
// output size (h - 1) / stride + 1 does not depend on it (efficientnet_pytorch Conv2dStaticSamePadding).
cudaError_t launch_dwconv(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                          const float* bias, void* out, pool_t* pool, int dtype, cudaStream_t s, int pad_lo = -1,
                          int act = 1);

struct DwTmaParams {
  int C, Ho, Wo;
  int otw, oth;          // output tile (otw even, oth multiple of TH)
  int tiles_x, tiles_y, ncb, n;
  int bw;                // box width in pixels (S=1) or column PAIRS (S=2)
  int bh;                // box height in input rows
  int cb_fixed;          // gridDim.x % ncb == 0: a CTA's tiles all share one channel block
  int gimg;              // images per tile (the box spans gimg images; small-spatial layers)
  int img_bytes;         // bytes of one image's box inside a stage
  int inv_cp, inv_per_img;  // ceil(65536 / col_groups), ceil(65536 / (col_groups * strips)); col_groups = otw / (2 NC)
  uint32_t box_bytes;
  int stage_stride;      // bytes between the two smem stages (1024-aligned)
  const float* wgt;      // [k*k, C] fp32, BN folded
  const float* bias;     // [C]
  void* out;               // [n, Ho, Wo, C]
  pool_t* pool;          // [n, C] Q24 sums or nullptr
};


// A planned depthwise launch: tensor map + tile geometry for one (layer, n, buffers) combination.
struct DwPlan {
  alignas(64) unsigned char tmap[128];
  DwTmaParams p;
  int act = 1;           // common.cuh Act (SiLU | GELU | none; the last two: 3x3 only)
  int k = 0, stride = 0, th = 0;
  int pad_lo = 1;        // zero rows/columns before the image: (k-1)/2, or 0 for the TF-"same" 3x3 stride-2 layers
  int nc = 1;            // output columns per thread (2: big stride-1 layers, columns x and x+2 share loaded inputs)
  int grid = 0, block = 0;
  size_t smem = 0;
  const void* x = nullptr;  // cache key
};
// bf16 only.  Returns 0 on success (plan filled), <0 if this shape is not supported by the TMA kernel.
int plan_dwconv_tma(DwPlan& plan, const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                    const float* bias, void* out, pool_t* pool, int pad_lo = -1, int act = 1);
cudaError_t launch_dwconv_tma(const DwPlan& plan, cudaStream_t s);
// Generic N-d bf16 tensor map (dims/strides innermost first; strides[0] is implicit).  0 on success.
// elem_strides (nullable): traversal stride per dim -- a box of B elements then loads ceil(B / stride) elements.
int encode_tmap_nd_bf16(void* tmap_out, const void* base, int rank, const unsigned long long* dims,
                        const unsigned long long* strides_bytes, const unsigned* box, int swizzle,
                        const unsigned* elem_strides = nullptr);

cudaError_t launch_se(const pool_t* pool, int n, int c, int rd, float inv_hw, const float* w1, const float* b1,
                      const float* w2, const float* b2, float* gate, cudaStream_t s);

cudaError_t launch_pwconv_simt(const void* x, long long m, int cin, int cout, const void* wgt, const float* bias,
                               const float* gate, int rows_per_img, const void* residual, int act, void* out,
                               int dtype, cudaStream_t s);

// tcgen05 GEMM.  Tensor maps are encoded by the caller (engine caches them per plan);
// see pwconv_tc.cu for the layout contract.
struct TcGemmArgs {
  // CUtensorMap is 128 bytes, 64-byte aligned; kept opaque here.
  alignas(64) unsigned char tmap_a[128];
  alignas(64) unsigned char tmap_b[128];
  alignas(64) unsigned char tmap_out[128];  // out as [cout, rows_per_img, images], box [64, 32, 1] (TMA store)
  long long m;        // rows of A / out
  int cin;            // K
  int cout;           // valid output columns (row pitch of out / residual)
  int cout_pad;       // rows of the padded weight matrix (multiple of bn)
  int bn;             // N tile (multiple of 16, <= 256), divides cout_pad
  const float* bias;  // [cout_pad]
  const float* gate;  // fp32 [m / rows_per_img, cin] or nullptr
  int rows_per_img;
  const void* residual;  // bf16 [m, cout] or nullptr
  int act;               // common.cuh Act: 0 none, 1 SiLU, 2 GELU, 3 ReLU
  void* out;             // bf16 [m, cout]
};
// Encodes a 2D bf16 row-major [rows, cols] tensor map with box [box_rows, 64] and 128B swizzle (weights).
int encode_tmap_2d_bf16(void* tmap_out, const void* base, long long rows, int cols, int box_rows);
// Activation / output operand [m, cols] viewed as [images, rows_per_img, cols] (3-D map, box [1, box_rows, 64],
// 128B swizzle) so that M tiles never straddle two images: rows past an image's end are zero-filled on load
// and clipped on store.  rows_per_img <= 0 (ungated GEMM) means one "image" of m rows.  m % rows_per_img == 0.
int encode_tmap_act_bf16(void* tmap_out, const void* base, long long m, int cols, int rows_per_img, int box_rows);
cudaError_t launch_pwconv_tc(const TcGemmArgs& a, cudaStream_t s);

cudaError_t launch_head_pool(const void* a, int n, int hw, int c, const float* scale, const float* shift,
                             float* pooled, int dtype, cudaStream_t s);
cudaError_t launch_classifier(const float* pooled, int n, int c, const float* w, const float* b, int num_classes,
                              float* logits, cudaStream_t s);
cudaError_t launch_cam(const void* a, int n, int hw, int c, const float* scale, const float* shift, const float* wfc,
                       int num_classes, const int32_t* target_cls, const float* logits, float* cam_lowres,
                       int dtype, cudaStream_t s);
cudaError_t launch_cam_upsample(const float* cam_lowres, int n, int side, int out_side, float* cam, cudaStream_t s);

// layout helpers for the split (hook-protocol) API
cudaError_t launch_nhwc_to_nchw_f32(const void* a_nhwc, int n, int hw, int c, float* out_nchw, int dtype,
                                    cudaStream_t s);
cudaError_t launch_head_from_nchw(const float* a_nchw, int n, int hw, int c, const float* scale, const float* shift,
                                  float* pooled, cudaStream_t s);
cudaError_t launch_head_backward(const float* a_nchw, const float* grad_logits, int n, int hw, int c,
                                 const float* scale, const float* shift, const float* wfc, int num_classes,
                                 float* grad_a, cudaStream_t s);
cudaError_t launch_softmax_pack(const float* logits, const int64_t* labels, int n, int num_classes, float* packed,
                                cudaStream_t s);

// ---- dense 3x3 conv (conv3x3.cu: CUDA-core implicit GEMM, any dtype; conv3x3_tc.cu: tcgen05, bf16) ----------------
// x NHWC [n, h, w, cin]; wgt [cout, 9*cin] with k = (ky*3 + kx)*cin + c in the engine dtype; pad 1; stride 1 | 2.
// cinp: channel pitch of one tap inside a weight row (>= cin; the tcgen05 layout pads every tap to 64 channels).
cudaError_t launch_conv3x3_simt(const void* x, int n, int h, int w, int cin, int cout, int stride, const void* wgt,
                                const float* bias, const void* residual, int act, void* out, int dtype,
                                cudaStream_t s, int cinp = 0);

// tcgen05 implicit GEMM (conv3x3_tc.cu), bf16.  wgt_tc: [cout_pad, 9, cin_pad] bf16, cin_pad = conv3x3_tc_cin_pad(cin),
// rows >= cout and channels >= cin zero; bias [cout_pad]; (cout_pad, bn) as chosen by the engine (choose_bn).
struct ConvTcParams {
  int cin, cout, cout_pad, bn;
  int Ho, Wo, stride, n_img;
  int th, ti;                     // M tile: 8 wide x th tall x ti images (th * ti = 16)
  int tiles_x, tiles_y, groups;   // spatial tiles per image, image groups
  int cpt, k_chunks;              // 64-channel chunks per tap, 9 * cpt
  int n_tiles, num_tiles, stages, acc_stride;
  uint32_t tmem_cols;
  // vhalo (stride 1, th 16, ti 1): one A box of th + 2 rows serves the three taps of a kernel COLUMN (a row shift of the
  // 8-pixel-wide tile is one 1024-byte swizzle atom); b_res: all 9 * cpt weight tiles stay in shared memory (cout_pad == bn)
  int vhalo, b_res;
  int b_stages;                   // vhalo without resident weights: the weight tiles have their own ring (stages = A ring)
  const float* bias;
  const __nv_bfloat16* residual;
};
struct Conv3Plan {
  alignas(64) unsigned char tmap_x[128];
  alignas(64) unsigned char tmap_w[128];
  alignas(64) unsigned char tmap_out[128];
  ConvTcParams p;
  int act = 0, grid = 0, n = 0;
  size_t smem = 0;
  const void* x = nullptr;   // cache key
  void* out = nullptr;
};
int conv3x3_tc_cin_pad(int cin);
int plan_conv3x3_tc(Conv3Plan& plan, const void* x, int n, int h, int w, int cin, int cout, int cout_pad, int bn,
                    int stride, const void* wgt_tc, const float* bias, const void* residual, int act, void* out);
cudaError_t launch_conv3x3_tc(const Conv3Plan& plan, cudaStream_t s);

// ---- small-sequence attention (attn.cu) ---------------------------------------------------------------------------
struct AttnArgs {
  const void* q; const void* k; const void* v;   // engine dtype; token row = pitch elements, head h at off + h*d
  int q_pitch, k_pitch, v_pitch;
  int q_off, k_off, v_off;
  int nseq, nq, nk, dk, dv, heads;
  float scale;
  const float* bias;                             // [heads, nq, nk] fp32 or nullptr
  const float* bias_frag;                        // the same table in mma accumulator-fragment order (attn_mma.cu), or nullptr
  const float* th1_w; const float* th1_b;        // talking heads: [heads, heads], [heads] (nullptr = none)
  const float* th2_w; const float* th2_b;
  const void* vlocal;                            // [nseq*nq, heads*dv] added to the output, or nullptr
  int act;                                       // common.cuh Act on the output
  void* out;                                     // [nseq*nq, heads*dv]
};
cudaError_t launch_attention(const AttnArgs& a, int dtype, cudaStream_t s);
// bf16, warp-level tensor-core MMA (attn_mma.cu); cudaErrorNotSupported for shapes it is not instantiated for
cudaError_t launch_attention_mma(const AttnArgs& a, cudaStream_t s);
size_t attention_bias_frag_elems(int heads, int nq, int nk);
void attention_bias_frag_host(const float* bias, int heads, int nq, int nk, float* out);
cudaError_t launch_attention_bias_frag(const float* bias, int heads, int nq, int nk, float* out, cudaStream_t s);

// ---- transformer glue (vit_ops.cu) ----------------------------------------------------------------------------------
cudaError_t launch_layernorm(const void* x, void* y, const float* gamma, const float* beta, long long rows, int c,
                             float eps, int dtype, cudaStream_t s);
cudaError_t launch_upsample2x_act(const void* x, void* y, int n, int side, int c, int act, int dtype, cudaStream_t s);
cudaError_t launch_window(const void* src, void* dst, int n, int side, int ws, int ct, int c, const float* pos,
                          bool reverse, int dtype, cudaStream_t s);
cudaError_t launch_ct_pool(const void* x, void* ct, int n, int side, int c, int kernel, int stride, int cs, int dtype,
                           cudaStream_t s);
cudaError_t launch_add_pos(void* x, long long nseq, int seq_rows, int row_off, int nrows, int c, const float* pos,
                           int dtype, cudaStream_t s);
cudaError_t launch_copy_rows(const void* src, void* dst, long long nseq, int src_rows, int src_off, int dst_rows,
                             int dst_off, int nrows, int c, int dtype, cudaStream_t s);
cudaError_t launch_pool_affine_fc(const void* a, const void* resid, int n, int hw, int c, const float* a_scale,
                                  const float* a_shift, const float* n_scale, const float* n_shift, const float* w,
                                  const float* b, int num_classes, float* logits, int dtype, cudaStream_t s);
cudaError_t launch_cam_linear(const void* a, int n, int hw, int c, const float* wcam, int num_classes,
                              const int32_t* target_cls, const float* logits, float* cam_lowres, int dtype,
                              cudaStream_t s);
cudaError_t launch_lin_head(const float* a_nchw, const float* aux, int n, int hw, int c, const float* wcam,
                            int num_classes, int mode, float* out, cudaStream_t s);
cudaError_t launch_lin_head_backward(const float* grad_logits, int n, int hw, int c, const float* wcam,
                                     int num_classes, float* grad_a, cudaStream_t s);

// Grad-CAM backward pass of the FasterViT layer program (cam_bwd.cu); tensors in the engine dtype, fp32 math
cudaError_t launch_layernorm_bwd(const void* x, const float* gamma, const void* dy, const void* res, void* out,
                                 long long rows, int c, float eps, int dtype, cudaStream_t s);
cudaError_t launch_gelu_bwd(const void* pre, void* g, long long total, int dtype, cudaStream_t s);
cudaError_t launch_move_rows(void* src, void* dst, long long nseq, int src_rows, int src_off, int dst_rows, int dst_off,
                             int nrows, int c, int dtype, cudaStream_t s);
cudaError_t launch_tail_bwd(const float* n_scale, const float* w, int n, int hw, int c, int num_classes,
                            const int32_t* target_cls, const float* logits, void* g, int dtype, cudaStream_t s);
cudaError_t launch_tail_bwd_general(const float* n_scale, const float* w, int n, int hw, int c, int num_classes,
                                    const float* grad_logits, void* g, int dtype, cudaStream_t s);
cudaError_t launch_ct_pool_bwd(const void* gct, float* ga_nchw, int n, int side, int c, int kernel, int stride, int cs,
                               int dtype, cudaStream_t s);
cudaError_t launch_cam_from_ct_grad(const void* a, const void* gct, int n, int hw, int c, int ntok, float* cam_lowres,
                                    int dtype, cudaStream_t s);
cudaError_t launch_conv3x3_s2_bwd(const void* dy, const void* w, void* dx, int n, int hin, int cin, int cinp, int cout,
                                  int dtype, cudaStream_t s);
cudaError_t launch_attention_bwd(const AttnArgs& a, const void* dout, void* dq, void* dk, void* dv, int dtype, cudaStream_t s);

// on-device metrics tail (metrics.cu)
// Grad-CAM overlay (overlay.cu): de-normalise + JET heat map + blend + max-normalise -> uint8 HWC
cudaError_t launch_cam_overlay(const float* x_nchw, const float* cam, int n, int h, int w, int denorm, const float* mean,
                               const float* stdv, float image_weight, unsigned char* out, cudaStream_t s);

cudaError_t launch_threshold_hist(const float* scores, const int64_t* labels, int n, const double* thr, int nthr,
                                  unsigned long long* hist, cudaStream_t s);
cudaError_t launch_confusion(const int64_t* targets, const int64_t* preds, int n, int nc, unsigned long long* cm,
                             cudaStream_t s);
cudaError_t launch_auc_count(const float* pos, int np, const float* neg_sorted, int nn, unsigned long long* out,
                             cudaStream_t s);

}  // namespace b200df
