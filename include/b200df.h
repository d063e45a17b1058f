/*
 * b200df.h -- C ABI of libb200df.so: the B200-native (sm_100a) engine for the
 * DeepfakeDetection eval hot path (batched backbone forward + Grad-CAM).
 *
 * Plain pointers and sizes only; no torch types.  All device pointers are raw
 * CUDA device addresses on the device the handle was created for; `stream` is a
 * cudaStream_t (CUstream) passed as void*.  Every call returns 0 on success or a
 * negative B200DF_E_* code; b200df_last_error() returns a thread-local message.
 * Nothing in here ever computes on the CPU: if no sm_100 device is present the
 * compute entry points fail with B200DF_E_NO_DEVICE (there is no fallback).
 *
 * What each entry point replaces in the reference (thourihan/DeepfakeDetection):
 *   b200df_create / set_tensor / finalize
 *        <- spec.builder(name, num_classes) + model.to(device) + model.eval() +
 *           model.load_state_dict(state, strict=False)
 *           (orchestration/model_registry.py:32-47, orchestration/orchestrator.py:350-377)
 *   b200df_forward
 *        <- `logits = model(images)` in the eval loops
 *           (orchestration/orchestrator.py:526-532 and :587-592; web_ui.py:266-268)
 *   b200df_forward_cam / b200df_cam
 *        <- GradCAM(model, [target_layer])(input_tensor, [ClassifierOutputTarget(c)])
 *           (web_ui.py:96-114 target rule, web_ui.py:275-279 call)
 *   b200df_head / b200df_head_backward
 *        <- the autograd leg pytorch-grad-cam drives through hooks on `_conv_head`
 *           (web_ui.py:108-114)
 *   b200df_softmax_pack
 *        <- probs = softmax(logits, 1); preds = argmax(probs, 1)
 *           (orchestration/orchestrator.py:591-592), packed for the all-gather
 */
#ifndef B200DF_H_
#define B200DF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define B200DF_ABI_VERSION 1

/* status codes */
#define B200DF_OK 0
#define B200DF_E_INVALID (-1)   /* bad argument / unknown tensor name / wrong shape */
#define B200DF_E_STATE (-2)     /* call out of order (e.g. forward before finalize)  */
#define B200DF_E_CUDA (-3)      /* a CUDA runtime/driver call failed                  */
#define B200DF_E_NO_DEVICE (-4) /* no sm_100 device: there is no CPU fallback         */
#define B200DF_E_WORKSPACE (-5) /* workspace too small or misaligned                  */

/* architectures */
#define B200DF_ARCH_EFFICIENTNET_B0 0 /* timm `efficientnet_b0` state_dict layout (conv_stem, blocks.S.R.*, ...)     */
#define B200DF_ARCH_EFFICIENTNET_B3 1 /* efficientnet_pytorch `efficientnet-b3` layout (_conv_stem, _blocks.N.*, _fc): */
                                      /* the reference's registry default, orchestration/model_registry.py:32-36,50-58 */
#define B200DF_ARCH_EFFICIENTFORMERV2_S1 2 /* timm `efficientformerv2_s1` layout (stem.conv1.conv, stages.S.blocks.B.*,  */
                                           /* norm, head, head_dist): model_registry.py:39-40, 61-67                     */
#define B200DF_ARCH_FASTERVIT_0 3 /* fastervit `faster_vit_0_224` layout (patch_embed.conv_down.N, levels.L.blocks.B.*,  */
                                  /* norm, head): model_registry.py:43-47, 68-74                                         */
#define B200DF_ARCH_FASTERVIT_2 4 /* fastervit `faster_vit_2_224`, the registry's prefix-table default (:68-74)           */

/* precisions (b200df_finalize) */
#define B200DF_PREC_BF16 0 /* bf16 NHWC activations/weights, fp32 accumulate, tcgen05 GEMMs */
#define B200DF_PREC_FP32 1 /* fp32 NHWC activations/weights, fp32 FFMA everywhere (strict)  */

typedef struct b200df_engine* b200df_handle;

const char* b200df_last_error(void);
int b200df_abi_version(void);

/* Number of CUDA devices with compute capability 10.x visible to this process (0 if none). */
int b200df_device_count(void);

/* ---- lifecycle ------------------------------------------------------------------------- */
int b200df_create(int arch, int num_classes, int device, b200df_handle* out);
int b200df_destroy(b200df_handle h);

/* Provide one fp32 tensor of the timm state_dict by name ("conv_stem.weight",
 * "blocks.3.1.bn2.running_var", ...).  `data` is host memory unless is_device != 0.
 * `numel` must match the architecture's shape for that name.  Unknown names return
 * B200DF_E_INVALID (the Python loader mirrors strict=False by skipping them loudly). */
int b200df_set_tensor(b200df_handle h, const char* name, const float* data, size_t numel, int is_device);

/* Number of tensors the architecture expects / name+numel of the i-th one. */
int b200df_num_tensors(b200df_handle h);
int b200df_tensor_info(b200df_handle h, int index, const char** name, size_t* numel);

/* Shape of the Grad-CAM activation A (raw conv_head / _conv_head output, web_ui.py:108-110) for one image:
 * channels (1280 for B0, 1536 for B3) and the square spatial side (7 at 224 x 224 input). */
int b200df_feature_shape(b200df_handle h, int* channels, int* side);

/* Fold BN into conv weights (fp32, on the host), convert to `precision`, lay out for the
 * kernels and upload.  May be called again after further set_tensor calls. */
int b200df_finalize(b200df_handle h, int precision);

/* ---- the hot path ---------------------------------------------------------------------- */
/* Bytes of device scratch needed for a batch of n images (mode: 0 = forward, 1 = forward+CAM). */
int b200df_workspace_bytes(b200df_handle h, int n, int mode, size_t* bytes);

/* images: float32 [n,3,224,224] NCHW contiguous (the reference's call convention,
 * orchestrator.py:588-590); logits: float32 [n,num_classes].  Asynchronous on `stream`. */
int b200df_forward(b200df_handle h, const float* images, int n, float* logits,
                   void* workspace, size_t workspace_bytes, void* stream);

/* Same forward, input given as uint8 [n,224,224,3] HWC crops; ToTensor (/255) and
 * Normalize(mean,std) (orchestrator.py:340-345) are fused into the stem.  mean/std: 3 floats
 * each (host), or NULL for "val_normalize: false". */
int b200df_forward_u8(b200df_handle h, const uint8_t* images_hwc, int n, const float* mean,
                      const float* stdv, float* logits, void* workspace, size_t workspace_bytes,
                      void* stream);

/* Forward + Grad-CAM in one pass.  target_cls: int32 [n] device (NULL = argmax of logits).
 * cam_lowres: float32 [n,s,s] (ReLU'd, unnormalised; s = b200df_feature_shape's side: 7, FasterViT 14) or NULL;
 * cam: float32 [n,224,224] in [0,1] (min-max, bilinear upsample, min-max) or NULL.  The workspace must be sized with
 * mode 1.  The target layer is the one the reference's rule picks (web_ui.py:94-112): conv_head / _conv_head
 * (EfficientNet, closed form), stages.3.blocks.5.mlp.fc2.conv (EfficientFormerV2, closed form),
 * levels.2.global_tokenizer.pos_embed (FasterViT: forward with activation checkpoints, then a backward pass through the
 * transformer blocks behind it -- csrc/cam_bwd.cu). */
int b200df_forward_cam(b200df_handle h, const float* images, int n, const int32_t* target_cls,
                       float* logits, float* cam_lowres, float* cam, void* workspace,
                       size_t workspace_bytes, void* stream);

/* Split form used by the autograd/hook protocol: trunk -> A (raw conv_head output,
 * float32 [n,1280,7,7] NCHW as a hook on `_conv_head` would see it), head: A -> logits,
 * head_backward: dL/dA given dL/dlogits. */
int b200df_trunk(b200df_handle h, const float* images, int n, float* a_nchw, void* workspace,
                 size_t workspace_bytes, void* stream);
int b200df_head(b200df_handle h, const float* a_nchw, int n, float* logits, void* stream);
int b200df_head_backward(b200df_handle h, const float* a_nchw, const float* grad_logits, int n,
                         float* grad_a_nchw, void* stream);
/* Same split with an auxiliary per-image term, aux float32 [n,num_classes]: EfficientFormerV2's Grad-CAM target
 * (the last nn.Conv2d, `stages.3.blocks.5.mlp.fc2.conv`, web_ui.py:96-105) is followed by an AFFINE map whose offset
 * depends on the block's residual input, not on A: trunk_ex writes aux = logits - lin(A), head_ex(A, aux) = aux + lin(A).
 * The EfficientNet engines write / ignore zeros.  FasterViT (target in front of the transformer blocks): trunk_ex runs the
 * whole forward with activation checkpoints, A = the pos_embed conv's output [n,C,14,14], aux = the logits; head_ex
 * returns aux; b200df_head_backward runs the backward program (csrc/cam_bwd.cu) from grad_logits on the checkpoints the
 * LAST trunk_ex call of this engine left in its workspace -- keep that workspace alive and untouched in between; another
 * batch size is refused with B200DF_E_STATE. */
int b200df_trunk_ex(b200df_handle h, const float* images, int n, float* a_nchw, float* aux, void* workspace,
                    size_t workspace_bytes, void* stream);
int b200df_head_ex(b200df_handle h, const float* a_nchw, const float* aux, int n, float* logits, void* stream);

/* Allocation-free variant of b200df_head for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  It has NO counterpart in the reference.
 * b200df_head keeps its pooled [n, head_c] scratch in a per-engine buffer that cudaMalloc grows on demand;
 * this entry point keeps it in a file-scope `static __device__` array of B200DF_HEAD_STATIC_MAX_FLOATS floats (1.5 MiB,
 * e.g. 256 images of EfficientNet-B3's 1536 or 307 of B0's 1280 channels) that every GPU loading the library holds until
 * exit.  It calls no cudaMalloc / cudaFree / allocating library routine; n * head_c beyond the maximum is rejected with
 * B200DF_E_INVALID, the layer-program architectures (whose split head, b200df_head_ex, uses no scratch) with
 * B200DF_E_STATE, and a current device other than the engine's with B200DF_E_STATE.  Same results as b200df_head.
 * ALL calls on a device share the one array, which keeps the last call's contents (every element a call reads it has
 * written first, so nothing needs clearing): calls that overlap on the GPU -- two streams, two engines, two host
 * threads -- RACE; serialise them on one stream or use b200df_head. */
#define B200DF_HEAD_STATIC_MAX_FLOATS (256 * 1536)
int b200df_head_static(b200df_handle h, const float* a_nchw, int n, float* logits, void* stream);

/* softmax + argmax of [n,num_classes] logits packed as records {p[0..C), pred, label} of
 * (C+2) floats per image -- the buffer the multi-GPU harness all-gathers. labels may be NULL. */
int b200df_softmax_pack(const float* logits, const int64_t* labels, int n, int num_classes,
                        float* packed, void* stream);

/* Grad-CAM overlay on the device: `_tensor_to_rgb(tensor, normalize)` (web_ui.py:72-93) followed by
 * pytorch-grad-cam `show_cam_on_image(rgb, cam, use_rgb=True, image_weight)` (web_ui.py:281-282).
 * images: float32 [n,3,h,w] NCHW (the tensor the model saw); cam: float32 [n,h,w] in [0,1];
 * denorm != 0: rgb = images * std + mean first (mean/std: 3 host floats each); out: uint8 [n,h,w,3] RGB.
 * Same separately-rounded fp32 operations as the numpy code: byte-identical to the host path. */
int b200df_cam_overlay(const float* images, const float* cam, int n, int h, int w, int denorm,
                       const float* mean, const float* stdv, float image_weight, unsigned char* out,
                       void* stream);

/* ---- on-device metrics tail (the counting work behind orchestrator.py:536-546 and :605-631) ---- */
/* Threshold sweep: hist[cls*(nthr+1) + j] += #{samples of class cls (label==1 -> 1, else 0) whose float32
 * score, promoted to double, is >= exactly j of the ascending `thresholds`} -- so for threshold k,
 * TP_k = sum_{j>k} hist[1][j], FP_k = sum_{j>k} hist[0][j] (same `score >= thr` decisions as the reference's
 * 501 balanced_accuracy_score calls).  hist: int64 [2*(nthr+1)] device, zeroed by the caller; nthr <= 1024. */
int b200df_threshold_hist(const float* scores, const int64_t* labels, int n, const double* thresholds,
                          int nthr, long long* hist, void* stream);
/* cm[t*num_classes + p] += 1 for every (target t, prediction p); int64 device, zeroed by the caller
 * (sklearn.metrics.confusion_matrix, orchestrator.py:631). */
int b200df_confusion_matrix(const int64_t* targets, const int64_t* preds, int n, int num_classes,
                            long long* cm, void* stream);
/* ROC-AUC counts (roc_auc_score, orchestrator.py:619-622): less_equal[0] += #{(p,q): neg[q] < pos[p]},
 * less_equal[1] += #{(p,q): neg[q] == pos[p]}; neg_scores_sorted ascending.  AUC = (less + equal/2)/(P*N). */
int b200df_auc_counts(const float* pos_scores, int n_pos, const float* neg_scores_sorted, int n_neg,
                      long long* less_equal, void* stream);

/* Count of kernel launches issued by this library on the calling thread since the last
 * reset (bench.py reports it as gpu_launches). */
long long b200df_launch_count(void);
void b200df_reset_launch_count(void);

/* ---- per-launch profiling (bench.py's per-kernel roofline) -------------------------------- */
/* When enabled, every kernel launch of forward / forward_cam is bracketed by two CUDA events on
 * the launching stream and tagged "<class>[.<block>]" with its ALGORITHMIC bytes and flops
 * (DESIGN.md section 4).  Records accumulate until b200df_profile_reset; read them after a
 * stream synchronize.  Keep it off inside timed regions. */
int b200df_set_profile(b200df_handle h, int enable);
int b200df_profile_count(b200df_handle h);
int b200df_profile_get(b200df_handle h, int index, const char** name, float* ms, double* bytes,
                       double* flops);
int b200df_profile_reset(b200df_handle h);

/* ---- debug checkpoints (layer programs: EfficientFormerV2 / FasterViT) --------------------- */
/* A mark says: after op `op` of the forward, the workspace at `ws_offset` (for batch n) holds sequences of
 * seqs_rows_c[0] rows x seqs_rows_c[3] channels (engine dtype), of which rows [seqs_rows_c[1], +seqs_rows_c[2]) equal the
 * reference module output `name` (token-major / NHWC).  b200df_debug_stop_after(h, op) makes forward return after that
 * op (-1 = run everything) so the buffer can be read before later layers reuse it.  Test / tooling use only. */
int b200df_debug_num_marks(b200df_handle h);
int b200df_debug_mark(b200df_handle h, int index, int n, const char** name, int* op, int* seqs_rows_c, size_t* ws_offset);
int b200df_debug_stop_after(b200df_handle h, int op);

/* ---- per-kernel entry points (unit tests, ncu) ------------------------------------------ */
/* dtype: 0 = bf16, 1 = fp32 for activations/weights; all layouts NHWC; bias fp32. */

/* Stem: fp32 NCHW [n,3,H,W] -> act [n,H/2,W/2,32]; w: [32,27] fp32 (BN folded), bias [32]; SiLU. */
int b200df_k_stem(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias,
                  void* out, int dtype, void* stream);

/* Same with `oc` output channels (32 | 40) and `pad` zero rows/columns BEFORE the image: 1 = symmetric,
 * 0 = the static TF-"same" split (0, 1) of efficientnet_pytorch's Conv2dStaticSamePadding (its stem at the
 * 300-pixel native size; /root/reference/orchestration/model_registry.py:32-36).  out: [n,H/2,W/2,oc]. */
int b200df_k_stem_ex(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias,
                     void* out, int dtype, int oc, int pad, void* stream);

/* Depthwise kxk (k=3|5, stride 1|2, pad (k-1)/2) + bias + SiLU; optional pooled sums
 * pool[n,c] += sum_hw(out) (must be zeroed by the caller) as Q24 fixed-point int64 (value * 2^24),
 * accumulated with integer atomics so the sums are order-independent and bit-reproducible.
 * wgt: [k*k, c] fp32 (tap-major). */
int b200df_k_dwconv(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                    const float* bias, void* out, long long* pool, int dtype, void* stream);

/* Same with an explicit pad-before `pad_lo` (pad after = k - 1 - pad_lo; -1 = symmetric): the TF-"same" 3x3
 * stride-2 layers of the efficientnet_pytorch layout use pad_lo = 0.  Output size (h - 1) / stride + 1. */
int b200df_k_dwconv_ex(const void* x, int n, int h, int w, int c, int k, int stride, int pad_lo, const float* wgt,
                       const float* bias, void* out, long long* pool, int dtype, void* stream);

/* Squeeze-excite MLP on pooled sums: gate[n,c] = sigmoid(W2 silu(W1 (pool*inv_hw) + b1) + b2).
 * w1: conv_reduce [rd,c]; w2t: conv_expand TRANSPOSED to [rd,c]. */
int b200df_k_se(const long long* pool, int n, int c, int rd, float inv_hw, const float* w1,
                const float* b1, const float* w2t, const float* b2, float* gate, void* stream);

/* Pointwise (1x1) conv as GEMM: out[m, cout] = act( (x[m,:] * gate[m / rows_per_img, :]) . w^T + bias )
 * (+ residual).  x: [m,cin]; w: [cout,cin] in `dtype`; gate fp32 [m/rows_per_img, cin] or NULL;
 * residual [m,cout] or NULL; act: 0 none, 1 SiLU.  impl: 0 = tcgen05 (bf16 only), 1 = SIMT. */
int b200df_k_pwconv(const void* x, long long m, int cin, int cout, const void* wgt, const float* bias,
                    const float* gate, int rows_per_img, const void* residual, int act, void* out,
                    int dtype, int impl, void* stream);

/* Head pooling: pooled[n,c] = mean_hw silu(a[n,hw,c]*scale[c] + shift[c]); a in `dtype` NHWC. */
int b200df_k_head_pool(const void* a, int n, int hw, int c, const float* scale, const float* shift,
                       float* pooled, int dtype, void* stream);

/* Fused Grad-CAM on the raw conv_head output a [n,hw,c] (dtype): closed-form channel weights
 * (BN -> SiLU -> mean -> Linear), weighted sum, ReLU -> cam_lowres [n,hw] fp32. */
int b200df_k_cam(const void* a, int n, int hw, int c, const float* scale, const float* shift,
                 const float* wfc, int num_classes, const int32_t* target_cls, float* cam_lowres,
                 int dtype, void* stream);

/* min-max -> bilinear (half-pixel, edge clamp) upsample to out_hw x out_hw -> relu -> min-max. */
int b200df_k_cam_upsample(const float* cam_lowres, int n, int side, int out_side, float* cam,
                          void* stream);

/* ---- kernels of the transformer-style backbones (EfficientFormerV2 / FasterViT) ---- */
/* act everywhere: 0 none, 1 SiLU, 2 GELU (nn.GELU, erf form; the bf16 path evaluates the tanh form), 3 ReLU. */

/* Stem with `oc` in {16, 32, 40, 64} output channels and activation `act` (timm efficientformerv2 `stem.conv1`:
 * oc 16 GELU; fastervit `patch_embed.conv_down.0`: oc 64 ReLU). */
int b200df_k_stem_act(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias, void* out,
                      int dtype, int oc, int pad, int act, void* stream);

/* Depthwise kxk + bias + `act` (no pooling): EfficientFormerV2 `mlp.mid` (GELU), `stride_conv` / `v_local` (none). */
int b200df_k_dwconv_act(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                        const float* bias, void* out, int act, int dtype, void* stream);

/* Dense 3x3 conv, pad 1, stride 1|2: out[n,ho,wo,cout] = act(conv(x) + bias) (+ residual).  x NHWC [n,h,w,cin];
 * wgt [cout, 9*cin] with k = (ky*3 + kx)*cin + c, in `dtype`.  impl: 0 = tcgen05 implicit GEMM (bf16), 1 = SIMT. */
int b200df_k_conv3x3(const void* x, int n, int h, int w, int cin, int cout, int stride, const void* wgt,
                     const float* bias, const void* residual, int act, void* out, int dtype, int impl, void* stream);

/* Small-sequence attention: S = q k^T * scale + bias [heads,nq,nk]; optional talking-head 1x1 mixes across heads
 * before / after the softmax; out = P v (+ vlocal) then `act`.  Token rows: pitch elements, head h at off + h*d. */
int b200df_k_attention(const void* q, const void* k, const void* v, int q_pitch, int k_pitch, int v_pitch, int q_off,
                       int k_off, int v_off, int nseq, int nq, int nk, int dk, int dv, int heads, float scale,
                       const float* bias, const float* th1_w, const float* th1_b, const float* th2_w,
                       const float* th2_b, const void* vlocal, int act, void* out, int dtype, void* stream);

/* LayerNorm over the innermost dim of [rows, c]. */
int b200df_k_layernorm(const void* x, void* y, const float* gamma, const float* beta, long long rows, int c, float eps,
                       int dtype, void* stream);

/* Backward-DATA kernels of the FasterViT Grad-CAM program (csrc/cam_bwd.cu; pytorch-grad-cam's loss.backward() through the
 * blocks behind the target layer, web_ui.py:275-279).  Tensors in `dtype`, fp32 math.
 *   attention_bwd : qkv / dqkv [nseq * n_tok, pitch] fused q|k|v rows (head h at off + h*d); dout [nseq * n_tok, heads*d]
 *   layernorm_bwd : out = res + dLayerNorm(x; gamma)(dy)   (res may be NULL)
 *   gelu_bwd      : g *= gelu'(pre), erf form, in place
 *   conv3x3_s2_bwd: dx [n,hin,hin,cin] of a 3x3 stride-2 pad-1 conv given dy [n,hin/2,hin/2,cout], wgt [cout, 9, cin]
 *   ct_pool_bwd   : fp32 NCHW gradient of the carrier-token AvgPool2d(kernel, stride) given dct [n, g*g, c] (windowed order) */
int b200df_k_attention_bwd(const void* qkv, int pitch, int q_off, int k_off, int v_off, int nseq, int n_tok, int d, int heads,
                           float scale, const float* bias, const void* dout, void* dqkv, int dtype, void* stream);
int b200df_k_layernorm_bwd(const void* x, const float* gamma, const void* dy, const void* res, void* out, long long rows,
                           int c, float eps, int dtype, void* stream);
int b200df_k_gelu_bwd(const void* pre, void* g, long long total, int dtype, void* stream);
int b200df_k_conv3x3_s2_bwd(const void* dy, const void* wgt, void* dx, int n, int hin, int cin, int cout, int dtype,
                            void* stream);
int b200df_k_ct_pool_bwd(const void* gct, float* ga_nchw, int n, int side, int c, int kernel, int stride, int cs, int dtype,
                         void* stream);

/* nn.Upsample(scale_factor=2, mode="bilinear") + act on NHWC [n,side,side,c] -> [n,2 side,2 side,c]. */
int b200df_k_upsample2x_act(const void* x, void* y, int n, int side, int c, int act, int dtype, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* B200DF_H_ */
