// Engine: host-side model description, weight folding/packing and the forward schedule.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "kernels.h"

namespace b200df {

struct TensorSlot {
  std::string name;
  size_t numel = 0;
  std::vector<float> host;  // raw fp32 copy as provided by set_tensor
  bool set = false;
};

// One MBConv block (timm DepthwiseSeparableConv when !expand, InvertedResidual otherwise).
struct BlockCfg {
  std::string prefix;  // "blocks.S.B."
  bool expand = false;
  bool skip = false;
  int cin = 0, mid = 0, cout = 0, k = 3, stride = 1, rd = 0;
  int hin = 0, hout = 0;  // square spatial sizes
  int pad_lo = -1;        // depthwise zero rows/columns before the image; -1 = symmetric (k-1)/2
  // state_dict names of this block's tensors ("" = absent); BN names are prefixes (.weight/.bias/.running_*)
  std::string n_expand_w, n_expand_bn, n_dw_w, n_dw_bn, n_se_rw, n_se_rb, n_se_ew, n_se_eb, n_proj_w, n_proj_bn;
};

// A packed pointwise-conv (GEMM) layer on the device.
struct PwLayer {
  int cin = 0, cout = 0, cout_pad = 0, bn = 0;
  void* w = nullptr;       // [cout_pad, cin] in the engine dtype (BN scale folded in)
  float* bias = nullptr;   // [cout_pad] fp32
};

struct BlockDev {
  PwLayer expand, project;
  float* dw_w = nullptr;   // [k*k, mid] fp32, BN folded
  float* dw_b = nullptr;   // [mid]
  float* se_w1 = nullptr;  // [rd, mid]
  float* se_b1 = nullptr;  // [rd]
  float* se_w2t = nullptr; // [rd, mid] (conv_expand transposed)
  float* se_b2 = nullptr;  // [mid]
  size_t pool_off = 0;     // float offset of this block's [n, mid] pool/gate slabs, per image
};

struct Plan {  // cached tensor maps for one (n, workspace) pair; the engine keeps a few (LRU)
  int n = -1;
  void* ws = nullptr;
  unsigned long long stamp = 0;   // last use
  std::vector<TcGemmArgs> gemms;  // in schedule order
  std::vector<DwPlan> dws;        // per block (bf16 path); x == nullptr -> not planned yet
  std::vector<Conv3Plan> convs;   // per op of a layer program (tcgen05 3x3 convs)
};

// Per-launch profiling record (b200df_set_profile): CUDA events around one kernel launch plus
// the ALGORITHMIC bytes / flops of that launch.
struct ProfRec {
  char name[48];
  double bytes = 0, flops = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
};

struct Workspace {
  size_t t0, t1, e, d, a, pool, gate, pooled, logits, cam_low, total;
};

// ---- layer programs (nets.inc): EfficientFormerV2 / FasterViT ---------------------------------------------------
enum OpKind { kOpStem = 0, kOpConv3, kOpPw, kOpDw, kOpAttn, kOpLn, kOpUp, kOpWin, kOpCtPool, kOpAddPos, kOpCopyRows, kOpTail,
              // Grad-CAM: activation checkpoints taken during the forward pass, and the ops of the backward program
              kOpSave, kOpZero, kOpLnBwd, kOpGeluBwd, kOpAttnBwd, kOpMoveRows, kOpTailBwd, kOpConv3Bwd, kOpCamCt };

struct NetOp {
  int kind = 0;
  char cls[24] = "";            // profiling class
  bool cam_only = false;        // forward-program op that only runs when a Grad-CAM backward pass follows (kOpSave)
  int in = -1, in2 = -1, in3 = -1, out = -1;   // activation buffer ids
  int i[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  float f[2] = {0.f, 0.f};
  const void* p[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  PwLayer pw;
  AttnArgs attn;
  double bytes = 0, wbytes = 0, flops = 0;     // algorithmic: activations per image, weights per launch, flops per image
};

struct NetWs { std::vector<size_t> off; size_t logits = 0, cam_low = 0, aux = 0, total = 0; };

// Debug checkpoint: after op `op`, buffer `buf` holds [n, rows, c] = the reference tensor `name` (tools/layer_diff.py)
struct NetMark { std::string name; int op, buf, rows, c, row_off, seq_rows; };

struct Net {
  std::vector<NetMark> marks;
  int stop_after = -1;             // debug: run_net returns after this op (-1 = run everything)
  std::vector<size_t> buf_elems;   // per image
  std::vector<int> buf_f32;
  std::vector<NetOp> ops;          // [0, fwd_end): forward program; [fwd_end, size): Grad-CAM backward program (may be empty)
  int fwd_end = -1;
  int first_cam_buf = -1;          // buffers >= this id exist only in the mode-1 (forward + CAM) workspace; -1 = none
  int cam_ntok = 0;                // carrier tokens per image (backward program's last op)
  int a_buf = -1, resid_buf = -1;  // Grad-CAM activation A (raw output of the reference's target conv) / its residual input
  int feat_c = 0, feat_side = 0;
  size_t wcam_off = ~size_t(0);    // arena offset of d logit / d A [num_classes, feat_c] (affine heads only)
  const float* wcam = nullptr;
};

}  // namespace b200df

struct b200df_engine {
  int arch = 0, num_classes = 2, device = 0;
  int precision = -1;  // -1 = not finalized
  int img = 224;
  // architecture constants (set by describe_*): stem/head widths, BN eps, stem pad-before, tensor names
  int stem_c = 32, head_c = 1280, stem_pad = 1;
  float bn_eps = 1e-5f;
  std::string n_stem_w, n_stem_bn, n_head_w, n_head_bn, n_fc_w, n_fc_b;
  std::vector<b200df::TensorSlot> slots;
  std::map<std::string, int> slot_index;
  std::vector<b200df::BlockCfg> blocks;
  // device weights
  void* arena = nullptr;
  size_t arena_bytes = 0;
  float* stem_w = nullptr;  // [32, 27]
  float* stem_b = nullptr;  // [32]
  std::vector<b200df::BlockDev> bdev;
  b200df::PwLayer head;      // conv_head, raw (no BN fold: its output is the Grad-CAM activation)
  float* head_scale = nullptr;  // bn2 folded scale/shift [1280]
  float* head_shift = nullptr;
  float* fc_w = nullptr;  // [num_classes, 1280]
  float* fc_b = nullptr;
  size_t pool_floats_per_img = 0;
  int gemm_impl = 0;  // 0 = tcgen05 (bf16), 1 = SIMT
  int dw_impl = 0;    // 0 = TMA halo kernel (bf16), 1 = plain kernel
  std::vector<b200df::Plan> plans;   // a few (n, workspace) plans, least recently used evicted
  unsigned long long plan_clock = 0;
  std::unique_ptr<b200df::Net> net;  // layer program (EfficientFormerV2 / FasterViT); null for the MBConv engines
  // hook protocol on a layer program with a backward program (FasterViT): b200df_trunk_ex leaves the activation checkpoints
  // in its workspace; b200df_head_backward runs the backward program on them
  void* cam_ws = nullptr;
  int cam_n = 0;
  float* head_scratch = nullptr;     // pooled [n, head_c] for the split (hook-protocol) head, grown outside captures
  size_t head_scratch_cap = 0;
  // profiling (off by default; never on inside a timed region)
  bool profile = false;
  std::vector<b200df::ProfRec> prof;   // event pool, reused across passes
  size_t prof_used = 0;
};
