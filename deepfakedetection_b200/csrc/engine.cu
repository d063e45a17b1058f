// libb200df.so C-ABI implementation (see include/b200df.h for the contract and the reference
// interfaces each entry point replaces).
#include "engine.h"

#include <cuda_bf16.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "../../include/b200df.h"

using namespace b200df;

// ---- error / launch-count plumbing --------------------------------------------------------------
static thread_local char g_err[512] = "";
static thread_local long long g_launches = 0;

namespace b200df {
void count_launch(int n) { g_launches += n; }
bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("B200DF_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}
}  // namespace b200df

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define CU_OK(expr)                                                                         \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) return fail(B200DF_E_CUDA, "%s failed: %s (%s:%d)", #expr,       \
                                       cudaGetErrorString(_e), __FILE__, __LINE__);         \
  } while (0)

extern "C" const char* b200df_last_error(void) { return g_err; }
extern "C" int b200df_abi_version(void) { return B200DF_ABI_VERSION; }
extern "C" long long b200df_launch_count(void) { return g_launches; }
extern "C" void b200df_reset_launch_count(void) { g_launches = 0; }

extern "C" int b200df_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  int ok = 0;
  for (int i = 0; i < n; ++i) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ++ok;
  }
  return ok;
}

static int require_sm100(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(B200DF_E_NO_DEVICE, "no CUDA device visible: libb200df has no CPU fallback");
  }
  if (device < 0 || device >= n) return fail(B200DF_E_INVALID, "device %d out of range (%d visible)", device, n);
  int major = 0;
  CU_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  if (major != 10)
    return fail(B200DF_E_NO_DEVICE, "device %d has compute capability %d.x; kernels are sm_100a only", device, major);
  return B200DF_OK;
}

// layer programs (nets.inc, included further down)
namespace {
bool is_net_arch(int arch);
void describe_net(b200df_engine* e);
int finalize_net(b200df_engine* h, int precision);
NetWs layout_net_workspace(const b200df_engine* h, int n, int mode = 0);
}  // namespace

// ---- EfficientNet-B0 description (timm layout) ---------------------------
static void add_slot(b200df_engine* e, const std::string& name, size_t numel) {
  TensorSlot s;
  s.name = name;
  s.numel = numel;
  e->slot_index[name] = (int)e->slots.size();
  e->slots.push_back(std::move(s));
}
static void add_bn(b200df_engine* e, const std::string& p, int c) {
  add_slot(e, p + ".weight", c);
  add_slot(e, p + ".bias", c);
  add_slot(e, p + ".running_mean", c);
  add_slot(e, p + ".running_var", c);
}

struct StageCfg { int k, s, exp, cout, reps; };

// Shared by both layouts: walks the stage table, registers every tensor under the layout's names.
// timm_layout: blocks.S.R.{conv_pw,bn1,conv_dw,bn2,se.conv_reduce,se.conv_expand,conv_pwl,bn3} (the expand-less
// first stage is a DepthwiseSeparableConv: conv_dw,bn1,se,conv_pw,bn2).  Otherwise efficientnet_pytorch:
// _blocks.N.{_expand_conv,_bn0,_depthwise_conv,_bn1,_se_reduce,_se_expand,_project_conv,_bn2}, flat block index.
static void describe_blocks(b200df_engine* e, const StageCfg* stages, int nstages, bool timm_layout, bool tf_same) {
  int cin = e->stem_c, h = e->img / 2, flat = 0;
  for (int s = 0; s < nstages; ++s) {
    for (int r = 0; r < stages[s].reps; ++r, ++flat) {
      BlockCfg b;
      b.expand = stages[s].exp != 1;
      b.cin = cin;
      b.cout = stages[s].cout;
      b.mid = cin * stages[s].exp;
      b.k = stages[s].k;
      b.stride = r == 0 ? stages[s].s : 1;
      b.skip = (b.stride == 1 && b.cin == b.cout);
      b.rd = cin / 4 > 0 ? cin / 4 : 1;  // timm: make_divisible off for se (rd = cin/4); e_p: max(1, int(cin*0.25))
      b.hin = h;
      b.hout = (h - 1) / b.stride + 1;   // total padding k-1 in both layouts
      // Conv2dStaticSamePadding at the 300-pixel native size: only the 3x3 stride-2 layers are asymmetric, (0, 1)
      b.pad_lo = (tf_same && b.k == 3 && b.stride == 2) ? 0 : -1;
      if (timm_layout) {
        const std::string p = "blocks." + std::to_string(s) + "." + std::to_string(r) + ".";
        b.prefix = p;
        if (b.expand) {
          b.n_expand_w = p + "conv_pw.weight"; b.n_expand_bn = p + "bn1";
          b.n_dw_w = p + "conv_dw.weight";     b.n_dw_bn = p + "bn2";
          b.n_proj_w = p + "conv_pwl.weight";  b.n_proj_bn = p + "bn3";
        } else {
          b.n_dw_w = p + "conv_dw.weight";     b.n_dw_bn = p + "bn1";
          b.n_proj_w = p + "conv_pw.weight";   b.n_proj_bn = p + "bn2";
        }
        b.n_se_rw = p + "se.conv_reduce.weight"; b.n_se_rb = p + "se.conv_reduce.bias";
        b.n_se_ew = p + "se.conv_expand.weight"; b.n_se_eb = p + "se.conv_expand.bias";
      } else {
        const std::string p = "_blocks." + std::to_string(flat) + ".";
        b.prefix = p;
        if (b.expand) { b.n_expand_w = p + "_expand_conv.weight"; b.n_expand_bn = p + "_bn0"; }
        b.n_dw_w = p + "_depthwise_conv.weight"; b.n_dw_bn = p + "_bn1";
        b.n_se_rw = p + "_se_reduce.weight";     b.n_se_rb = p + "_se_reduce.bias";
        b.n_se_ew = p + "_se_expand.weight";     b.n_se_eb = p + "_se_expand.bias";
        b.n_proj_w = p + "_project_conv.weight"; b.n_proj_bn = p + "_bn2";
      }
      // registration order = state_dict order of the layout
      if (b.expand) {
        add_slot(e, b.n_expand_w, (size_t)b.mid * b.cin);
        add_bn(e, b.n_expand_bn, b.mid);
      }
      add_slot(e, b.n_dw_w, (size_t)b.mid * b.k * b.k);
      add_bn(e, b.n_dw_bn, b.mid);
      add_slot(e, b.n_se_rw, (size_t)b.rd * b.mid);
      add_slot(e, b.n_se_rb, b.rd);
      add_slot(e, b.n_se_ew, (size_t)b.mid * b.rd);
      add_slot(e, b.n_se_eb, b.mid);
      add_slot(e, b.n_proj_w, (size_t)b.cout * b.mid);
      add_bn(e, b.n_proj_bn, b.cout);
      e->blocks.push_back(b);
      cin = b.cout;
      h = b.hout;
    }
  }
}

static void describe_tail(b200df_engine* e) {
  add_slot(e, e->n_head_w, (size_t)e->head_c * e->blocks.back().cout);
  add_bn(e, e->n_head_bn, e->head_c);
  add_slot(e, e->n_fc_w, (size_t)e->num_classes * e->head_c);
  add_slot(e, e->n_fc_b, e->num_classes);
}

static void describe_effnet_b0(b200df_engine* e) {
  static const StageCfg stages[7] = {{3, 1, 1, 16, 1},  {3, 2, 6, 24, 2},  {5, 2, 6, 40, 2}, {3, 2, 6, 80, 3},
                                     {5, 1, 6, 112, 3}, {5, 2, 6, 192, 4}, {3, 1, 6, 320, 1}};
  e->stem_c = 32; e->head_c = 1280; e->stem_pad = 1; e->bn_eps = 1e-5f;
  e->n_stem_w = "conv_stem.weight"; e->n_stem_bn = "bn1";
  e->n_head_w = "conv_head.weight"; e->n_head_bn = "bn2";
  e->n_fc_w = "classifier.weight";  e->n_fc_b = "classifier.bias";
  add_slot(e, e->n_stem_w, (size_t)e->stem_c * 27);
  add_bn(e, e->n_stem_bn, e->stem_c);
  describe_blocks(e, stages, 7, true, false);
  describe_tail(e);
}

// EfficientNet-B3 as built by efficientnet_pytorch `EfficientNet.from_name("efficientnet-b3")`
// (/root/reference/orchestration/model_registry.py:32-36, 50-58): width x1.2 / depth x1.4 of
// the base table (stem 40, stages 24/32/48/96/136/232/384 x 2/3/3/5/5/6/2, head 1536), BN eps 1e-3, SE squeeze
// int(cin/4), static TF-"same" padding computed for 300 pixels and applied to the 224-pixel crops.
static void describe_effnet_b3(b200df_engine* e) {
  static const StageCfg stages[7] = {{3, 1, 1, 24, 2},  {3, 2, 6, 32, 3},  {5, 2, 6, 48, 3}, {3, 2, 6, 96, 5},
                                     {5, 1, 6, 136, 5}, {5, 2, 6, 232, 6}, {3, 1, 6, 384, 2}};
  e->stem_c = 40; e->head_c = 1536; e->stem_pad = 0; e->bn_eps = 1e-3f;
  e->n_stem_w = "_conv_stem.weight"; e->n_stem_bn = "_bn0";
  e->n_head_w = "_conv_head.weight"; e->n_head_bn = "_bn1";
  e->n_fc_w = "_fc.weight";          e->n_fc_b = "_fc.bias";
  add_slot(e, e->n_stem_w, (size_t)e->stem_c * 27);
  add_bn(e, e->n_stem_bn, e->stem_c);
  describe_blocks(e, stages, 7, false, true);
  describe_tail(e);
}

// ---- lifecycle ----------------------------------------------------------------------------------------
extern "C" int b200df_create(int arch, int num_classes, int device, b200df_handle* out) {
  if (!out) return fail(B200DF_E_INVALID, "out is NULL");
  *out = nullptr;
  if (arch != B200DF_ARCH_EFFICIENTNET_B0 && arch != B200DF_ARCH_EFFICIENTNET_B3 && !is_net_arch(arch))
    return fail(B200DF_E_INVALID, "unknown arch id %d", arch);
  if (num_classes < 1 || num_classes > 4096) return fail(B200DF_E_INVALID, "num_classes %d out of range", num_classes);
  // No device is needed to describe a model (tests list the expected tensors on CPU boxes);
  // finalize() and every compute entry point require an sm_100 device and never fall back.
  b200df_engine* e = new b200df_engine();
  e->arch = arch;
  e->num_classes = num_classes;
  e->device = device;
  if (is_net_arch(arch)) describe_net(e);
  else if (arch == B200DF_ARCH_EFFICIENTNET_B3) describe_effnet_b3(e);
  else describe_effnet_b0(e);
  *out = e;
  return B200DF_OK;
}

static void free_device(b200df_engine* e) {
  if (e->arena) {
    cudaFree(e->arena);
    e->arena = nullptr;
  }
  if (e->head_scratch) {
    cudaFree(e->head_scratch);
    e->head_scratch = nullptr;
    e->head_scratch_cap = 0;
  }
  e->bdev.clear();
  e->plans.clear();
  e->precision = -1;
}

extern "C" int b200df_destroy(b200df_handle h) {
  if (!h) return B200DF_OK;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(h->device);
  free_device(h);
  for (auto& r : h->prof) {
    if (r.e0) cudaEventDestroy(r.e0);
    if (r.e1) cudaEventDestroy(r.e1);
  }
  cudaSetDevice(prev);
  delete h;
  return B200DF_OK;
}

extern "C" int b200df_num_tensors(b200df_handle h) { return h ? (int)h->slots.size() : 0; }

extern "C" int b200df_tensor_info(b200df_handle h, int index, const char** name, size_t* numel) {
  if (!h || index < 0 || index >= (int)h->slots.size()) return fail(B200DF_E_INVALID, "tensor index out of range");
  if (name) *name = h->slots[index].name.c_str();
  if (numel) *numel = h->slots[index].numel;
  return B200DF_OK;
}

extern "C" int b200df_feature_shape(b200df_handle h, int* channels, int* side) {
  if (h && h->net) {
    if (channels) *channels = h->net->feat_c;
    if (side) *side = h->net->feat_side;
    return B200DF_OK;
  }
  if (!h || h->blocks.empty()) return fail(B200DF_E_INVALID, "NULL handle");
  if (channels) *channels = h->head_c;
  if (side) *side = h->blocks.back().hout;
  return B200DF_OK;
}

extern "C" int b200df_set_tensor(b200df_handle h, const char* name, const float* data, size_t numel, int is_device) {
  if (!h || !name || !data) return fail(B200DF_E_INVALID, "NULL argument");
  auto it = h->slot_index.find(name);
  if (it == h->slot_index.end()) return fail(B200DF_E_INVALID, "unknown tensor '%s'", name);
  TensorSlot& s = h->slots[it->second];
  if (numel != s.numel)
    return fail(B200DF_E_INVALID, "tensor '%s': got %zu elements, expected %zu", name, numel, s.numel);
  s.host.resize(numel);
  if (is_device) {
    CU_OK(cudaMemcpy(s.host.data(), data, numel * sizeof(float), cudaMemcpyDeviceToHost));
  } else {
    memcpy(s.host.data(), data, numel * sizeof(float));
  }
  s.set = true;
  return B200DF_OK;
}

// ---- finalize: fold BN (fp32, host), convert, lay out, upload -----------------------------------------
namespace {

struct Packer {
  std::vector<uint8_t> buf;
  size_t alloc(size_t bytes) {
    size_t off = (buf.size() + 1023) & ~size_t(1023);
    buf.resize(off + bytes, 0);
    return off;
  }
};

inline uint16_t f2bf(float f) {  // round-to-nearest-even, like __float2bfloat16_rn
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

const std::vector<float>& T(const b200df_engine* e, const std::string& name) {
  return e->slots[e->slot_index.at(name)].host;
}

// BN(eval): y = (x - mean)/sqrt(var + eps)*gamma + beta  ->  scale, shift
void bn_fold(const b200df_engine* e, const std::string& p, int c, std::vector<float>& scale, std::vector<float>& shift) {
  const auto& g = T(e, p + ".weight");
  const auto& b = T(e, p + ".bias");
  const auto& m = T(e, p + ".running_mean");
  const auto& v = T(e, p + ".running_var");
  scale.resize(c);
  shift.resize(c);
  for (int i = 0; i < c; ++i) {
    float s = g[i] / sqrtf(v[i] + e->bn_eps);
    scale[i] = s;
    shift[i] = b[i] - m[i] * s;
  }
}

int choose_bn(int cout, int* cout_pad) {
  int pad = (cout + 15) / 16 * 16;
  if (pad <= 256) {
    *cout_pad = pad;
    return pad;
  }
  // several N tiles: the epilogue's 64-column TMA-store boxes must not reach into the next tile's
  // columns, so the N tile is a multiple of 64 (padded weight rows are zero, padded columns are clipped)
  // Among 256 / 192 / 128 pick the tile with the least padded columns (ties: the wider one): 1152 -> 6 x 192
  // exactly instead of 5 x 256 = 1280 (10 % dead columns and a worse wave count at 98 M tiles).
  int best_bn = 256, best_pad = (pad + 255) / 256 * 256;
  for (int bn = 192; bn >= 128; bn -= 64) {
    const int p2 = (pad + bn - 1) / bn * bn;
    if (p2 < best_pad) { best_pad = p2; best_bn = bn; }
  }
  *cout_pad = best_pad;
  return best_bn;
}

struct PwOff { size_t w, bias; };

// w_src: [cout, cin] fp32; scale may be empty (no fold); shift -> bias (empty -> zeros)
PwOff pack_pw(Packer& pk, PwLayer& L, const std::vector<float>& w_src, int cout, int cin,
              const std::vector<float>& scale, const std::vector<float>& shift, int precision) {
  L.cin = cin;
  L.cout = cout;
  L.bn = choose_bn(cout, &L.cout_pad);
  const size_t es = precision == B200DF_PREC_BF16 ? 2 : 4;
  PwOff o;
  o.w = pk.alloc((size_t)L.cout_pad * cin * es);
  o.bias = pk.alloc((size_t)L.cout_pad * sizeof(float));
  for (int oc = 0; oc < cout; ++oc) {
    const float s = scale.empty() ? 1.0f : scale[oc];
    for (int k = 0; k < cin; ++k) {
      float v = w_src[(size_t)oc * cin + k] * s;
      if (es == 2) reinterpret_cast<uint16_t*>(pk.buf.data() + o.w)[(size_t)oc * cin + k] = f2bf(v);
      else reinterpret_cast<float*>(pk.buf.data() + o.w)[(size_t)oc * cin + k] = v;
    }
    reinterpret_cast<float*>(pk.buf.data() + o.bias)[oc] = shift.empty() ? 0.0f : shift[oc];
  }
  return o;
}

size_t pack_f32(Packer& pk, const std::vector<float>& v) {
  size_t off = pk.alloc(v.size() * sizeof(float));
  memcpy(pk.buf.data() + off, v.data(), v.size() * sizeof(float));
  return off;
}

}  // namespace

extern "C" int b200df_finalize(b200df_handle h, int precision) {
  if (!h) return fail(B200DF_E_INVALID, "NULL handle");
  if (precision != B200DF_PREC_BF16 && precision != B200DF_PREC_FP32)
    return fail(B200DF_E_INVALID, "unknown precision %d", precision);
  for (const auto& s : h->slots)
    if (!s.set) return fail(B200DF_E_STATE, "tensor '%s' was never provided (set_tensor)", s.name.c_str());
  int rc = require_sm100(h->device);
  if (rc != B200DF_OK) return rc;
  CU_OK(cudaSetDevice(h->device));
  free_device(h);
  h->gemm_impl = precision == B200DF_PREC_BF16 ? 0 : 1;
  if (const char* env = getenv("B200DF_GEMM")) {  // debugging aid: force the SIMT GEMM / conv on the bf16 path
    if (strcmp(env, "simt") == 0) h->gemm_impl = 1;
  }
  h->dw_impl = 0;
  if (h->net) {
    rc = finalize_net(h, precision);
    if (rc == B200DF_OK) h->precision = precision;
    return rc;
  }

  Packer pk;
  std::vector<float> sc, sh, tmp;
  // stem
  bn_fold(h, h->n_stem_bn, h->stem_c, sc, sh);
  tmp = T(h, h->n_stem_w);
  for (int oc = 0; oc < h->stem_c; ++oc)
    for (int t = 0; t < 27; ++t) tmp[oc * 27 + t] *= sc[oc];
  const size_t stem_w = pack_f32(pk, tmp), stem_b = pack_f32(pk, sh);

  struct BOff { PwOff ex, pr; size_t dw_w, dw_b, w1, b1, w2t, b2; };
  std::vector<BOff> boff(h->blocks.size());
  h->bdev.assign(h->blocks.size(), BlockDev());
  size_t pool_off = 0;
  for (size_t i = 0; i < h->blocks.size(); ++i) {
    const BlockCfg& b = h->blocks[i];
    BlockDev& d = h->bdev[i];
    if (b.expand) {
      bn_fold(h, b.n_expand_bn, b.mid, sc, sh);
      boff[i].ex = pack_pw(pk, d.expand, T(h, b.n_expand_w), b.mid, b.cin, sc, sh, precision);
    }
    // depthwise: [mid,1,k,k] -> tap-major [k*k, mid], BN folded
    bn_fold(h, b.n_dw_bn, b.mid, sc, sh);
    {
      const auto& w = T(h, b.n_dw_w);
      const int kk = b.k * b.k;
      tmp.assign((size_t)kk * b.mid, 0.f);
      for (int c = 0; c < b.mid; ++c)
        for (int t = 0; t < kk; ++t) tmp[(size_t)t * b.mid + c] = w[(size_t)c * kk + t] * sc[c];
      boff[i].dw_w = pack_f32(pk, tmp);
      boff[i].dw_b = pack_f32(pk, sh);
    }
    // squeeze-excite
    boff[i].w1 = pack_f32(pk, T(h, b.n_se_rw));
    boff[i].b1 = pack_f32(pk, T(h, b.n_se_rb));
    {
      const auto& w2 = T(h, b.n_se_ew);  // [mid, rd]
      tmp.assign((size_t)b.rd * b.mid, 0.f);
      for (int c = 0; c < b.mid; ++c)
        for (int j = 0; j < b.rd; ++j) tmp[(size_t)j * b.mid + c] = w2[(size_t)c * b.rd + j];
      boff[i].w2t = pack_f32(pk, tmp);
    }
    boff[i].b2 = pack_f32(pk, T(h, b.n_se_eb));
    // project
    bn_fold(h, b.n_proj_bn, b.cout, sc, sh);
    boff[i].pr = pack_pw(pk, d.project, T(h, b.n_proj_w), b.cout, b.mid, sc, sh, precision);
    d.pool_off = pool_off;
    pool_off += (size_t)b.mid;
  }
  h->pool_floats_per_img = pool_off;
  // head: conv_head stays un-folded (its raw output is the Grad-CAM activation); bn2 -> scale/shift
  const int cin_head = h->blocks.back().cout;
  PwOff head_off = pack_pw(pk, h->head, T(h, h->n_head_w), h->head_c, cin_head, {}, {}, precision);
  bn_fold(h, h->n_head_bn, h->head_c, sc, sh);
  const size_t hs = pack_f32(pk, sc), hb = pack_f32(pk, sh);
  const size_t fw = pack_f32(pk, T(h, h->n_fc_w)), fb = pack_f32(pk, T(h, h->n_fc_b));

  h->arena_bytes = (pk.buf.size() + 1023) & ~size_t(1023);
  CU_OK(cudaMalloc(&h->arena, h->arena_bytes));
  CU_OK(cudaMemcpy(h->arena, pk.buf.data(), pk.buf.size(), cudaMemcpyHostToDevice));
  uint8_t* base = reinterpret_cast<uint8_t*>(h->arena);
  auto F = [&](size_t off) { return reinterpret_cast<float*>(base + off); };
  h->stem_w = F(stem_w);
  h->stem_b = F(stem_b);
  for (size_t i = 0; i < h->blocks.size(); ++i) {
    BlockDev& d = h->bdev[i];
    if (h->blocks[i].expand) {
      d.expand.w = base + boff[i].ex.w;
      d.expand.bias = F(boff[i].ex.bias);
    }
    d.project.w = base + boff[i].pr.w;
    d.project.bias = F(boff[i].pr.bias);
    d.dw_w = F(boff[i].dw_w);
    d.dw_b = F(boff[i].dw_b);
    d.se_w1 = F(boff[i].w1);
    d.se_b1 = F(boff[i].b1);
    d.se_w2t = F(boff[i].w2t);
    d.se_b2 = F(boff[i].b2);
  }
  h->head.w = base + head_off.w;
  h->head.bias = F(head_off.bias);
  h->head_scale = F(hs);
  h->head_shift = F(hb);
  h->fc_w = F(fw);
  h->fc_b = F(fb);
  h->precision = precision;
  if (const char* env = getenv("B200DF_DW")) {  // debugging aid: force the plain (non-TMA) depthwise kernel
    if (strcmp(env, "plain") == 0) h->dw_impl = 1;
  }
  return B200DF_OK;
}

// ---- workspace ------------------------------------------------------------------------------------------
static Workspace layout_workspace(const b200df_engine* h, int n) {
  const size_t es = h->precision == B200DF_PREC_BF16 ? 2 : 4;
  size_t trunk = (size_t)h->stem_c * (h->img / 2) * (h->img / 2);  // stem output
  size_t e = 0, d = 0;
  for (const auto& b : h->blocks) {
    trunk = std::max(trunk, (size_t)b.cout * b.hout * b.hout);
    if (b.expand) e = std::max(e, (size_t)b.mid * b.hin * b.hin);
    d = std::max(d, (size_t)b.mid * b.hout * b.hout);
  }
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  auto al = [](size_t x) { return (x + 1023) & ~size_t(1023); };
  Workspace w;
  size_t off = 0;
  w.t0 = off; off += al(trunk * n * es);
  w.t1 = off; off += al(trunk * n * es);
  w.e = off; off += al(e * n * es);
  w.d = off; off += al(d * n * es);
  w.a = off; off += al((size_t)h->head_c * hw * n * es);
  w.pool = off; off += al(h->pool_floats_per_img * n * sizeof(pool_t));
  w.gate = off; off += al(h->pool_floats_per_img * n * sizeof(float));
  w.pooled = off; off += al((size_t)h->head_c * n * sizeof(float));
  w.logits = off; off += al((size_t)h->num_classes * n * sizeof(float));
  w.cam_low = off; off += al((size_t)hw * n * sizeof(float));
  w.total = off;
  return w;
}

extern "C" int b200df_workspace_bytes(b200df_handle h, int n, int mode, size_t* bytes) {
  if (!h || !bytes || n < 0) return fail(B200DF_E_INVALID, "bad argument");
  if (h->precision < 0) return fail(B200DF_E_STATE, "b200df_finalize has not been called");
  *bytes = h->net ? layout_net_workspace(h, n > 0 ? n : 1, mode).total : layout_workspace(h, n > 0 ? n : 1).total;
  return B200DF_OK;
}

// ---- forward schedule ---------------------------------------------------------------------------------------
namespace {

// Profiling: when h->profile is set, every launch of the schedule is bracketed by two CUDA events
// on the launching stream and tagged with its algorithmic bytes / flops.
ProfRec* prof_begin(b200df_engine* h, cudaStream_t s, const char* cls, int layer, double bytes, double flops) {
  if (!h->profile) return nullptr;
  if (h->prof_used == h->prof.size()) {
    ProfRec r;
    if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return nullptr;
    h->prof.push_back(r);
  }
  ProfRec* r = &h->prof[h->prof_used++];
  if (layer >= 0) snprintf(r->name, sizeof(r->name), "%s.%d", cls, layer);
  else snprintf(r->name, sizeof(r->name), "%s", cls);
  r->bytes = bytes;
  r->flops = flops;
  cudaEventRecord(r->e0, s);
  return r;
}
void prof_end(ProfRec* r, cudaStream_t s) {
  if (r) cudaEventRecord(r->e1, s);
}

struct GemmCall {  // one pointwise conv in schedule order
  const PwLayer* L;
  const void* x;
  long long m;
  const float* gate;
  int rows_per_img;
  const void* residual;
  int act;
  void* out;
};

// The plan (tensor maps, depthwise tile plans) of this (n, workspace) pair: the engine keeps a few of them, so a
// loader that alternates full batches / a ragged tail / val and test splits does not re-encode ~80 tensor maps
// per switch.  Least recently used is evicted.
Plan& plan_for(b200df_engine* h, int n, void* ws) {
  constexpr size_t kMaxPlans = 6;
  ++h->plan_clock;
  for (auto& p : h->plans)
    if (p.n == n && p.ws == ws) {
      p.stamp = h->plan_clock;
      return p;
    }
  if (h->plans.size() < kMaxPlans) {
    h->plans.reserve(kMaxPlans);   // references handed out earlier stay valid
    h->plans.emplace_back();
  } else {
    size_t lru = 0;
    for (size_t i = 1; i < h->plans.size(); ++i)
      if (h->plans[i].stamp < h->plans[lru].stamp) lru = i;
    std::swap(h->plans[lru], h->plans.back());
    h->plans.back() = Plan();
  }
  Plan& p = h->plans.back();
  p.n = n;
  p.ws = ws;
  p.stamp = h->plan_clock;
  return p;
}

int run_gemm(b200df_engine* h, Plan& plan, const GemmCall& g, int index, cudaStream_t s) {
  const int dt = h->precision == B200DF_PREC_BF16 ? kBF16 : kFP32;
  if (h->gemm_impl == 1) {
    CU_OK(launch_pwconv_simt(g.x, g.m, g.L->cin, g.L->cout, g.L->w, g.L->bias, g.gate, g.rows_per_img, g.residual,
                             g.act, g.out, dt, s));
    return B200DF_OK;
  }
  if ((int)plan.gemms.size() <= index) plan.gemms.resize(index + 1);
  TcGemmArgs& a = plan.gemms[index];
  if (a.out != g.out || a.m != g.m) {  // (re)encode tensor maps for this (n, workspace)
    memset(&a, 0, sizeof(a));
    const int rpi = g.gate ? g.rows_per_img : 0;
    int r1 = encode_tmap_act_bf16(a.tmap_a, g.x, g.m, g.L->cin, rpi, 128);
    int r2 = encode_tmap_2d_bf16(a.tmap_b, g.L->w, g.L->cout_pad, g.L->cin, g.L->bn);
    int r3 = encode_tmap_act_bf16(a.tmap_out, g.out, g.m, g.L->cout, rpi, 32);
    if (r1 || r2 || r3)
      return fail(B200DF_E_CUDA, "cuTensorMapEncodeTiled failed (%d, %d, %d) for gemm %d", r1, r2, r3, index);
    a.m = g.m;
    a.cin = g.L->cin;
    a.cout = g.L->cout;
    a.cout_pad = g.L->cout_pad;
    a.bn = g.L->bn;
    a.bias = g.L->bias;
    a.gate = g.gate;
    a.rows_per_img = g.rows_per_img;
    a.residual = g.residual;
    a.act = g.act;
    a.out = g.out;
  }
  CU_OK(launch_pwconv_tc(a, s));
  return B200DF_OK;
}

enum InputKind { kInF32NCHW = 0, kInU8HWC = 1 };

// dense 3x3 conv of a layer program: tcgen05 implicit GEMM on the bf16 path, CUDA-core implicit GEMM otherwise
// (B200DF_GEMM=simt, or B200DF_CONV_TC_S2=0 for the stride-2 layers only: debugging aids)
int run_conv3(b200df_engine* h, Plan& plan, int index, const void* x, int n, int hin, int cin, int cout, int stride,
              const void* w, const float* bias, const void* residual, int act, void* out, int dt, cudaStream_t s,
              int cinp, int cout_pad, int bn) {
  static int tc_s2 = -1;
  if (tc_s2 < 0) {
    const char* e = getenv("B200DF_CONV_TC_S2");
    tc_s2 = (e && e[0] == '0') ? 0 : 1;
  }
  if (dt == kBF16 && h->gemm_impl == 0 && (stride == 1 || tc_s2)) {
    Conv3Plan& cp = plan.convs[index];
    if (cp.x != x || cp.out != out || cp.n != n) {
      cp.x = nullptr;
      if (plan_conv3x3_tc(cp, x, n, hin, hin, cin, cout, cout_pad, bn, stride, w, bias, residual, act, out) != 0)
        return fail(B200DF_E_CUDA, "tcgen05 3x3 conv: cannot plan op %d (cin %d cout %d stride %d)", index, cin, cout, stride);
    }
    CU_OK(launch_conv3x3_tc(cp, s));
    return B200DF_OK;
  }
  CU_OK(launch_conv3x3_simt(x, n, hin, hin, cin, cout, stride, w, bias, residual, act, out, dt, s, cinp));
  return B200DF_OK;
}

// Runs stem -> 16 MBConv blocks -> conv_head.  Leaves A (raw conv_head output, NHWC) in ws.a.
int run_trunk(b200df_engine* h, const void* images, int in_kind, const float* in_sc, const float* in_sh, int n,
              uint8_t* ws, const Workspace& W, cudaStream_t s) {
  const int dt = h->precision == B200DF_PREC_BF16 ? kBF16 : kFP32;
  Plan& plan = plan_for(h, n, ws);
  pool_t* pool = reinterpret_cast<pool_t*>(ws + W.pool);
  float* gate = reinterpret_cast<float*>(ws + W.gate);
  CU_OK(cudaMemsetAsync(pool, 0, h->pool_floats_per_img * n * sizeof(pool_t), s));
  void* cur = ws + W.t0;
  void* nxt = ws + W.t1;
  const double es = h->precision == B200DF_PREC_BF16 ? 2.0 : 4.0;
  const double ho0 = h->img / 2;
  ProfRec* pr = prof_begin(h, s, "stem", -1,
                           (double)n * (3.0 * h->img * h->img * (in_kind == kInF32NCHW ? 4.0 : 1.0) + ho0 * ho0 * h->stem_c * es),
                           2.0 * n * ho0 * ho0 * h->stem_c * 27);
  if (in_kind == kInF32NCHW) {
    CU_OK(launch_stem(reinterpret_cast<const float*>(images), n, h->img, h->img, h->stem_w, h->stem_b, cur, dt, s,
                      h->stem_c, h->stem_pad));
  } else {
    CU_OK(launch_stem_u8(reinterpret_cast<const uint8_t*>(images), n, h->img, h->img, h->stem_w, h->stem_b, in_sc[0],
                         in_sc[1], in_sc[2], in_sh[0], in_sh[1], in_sh[2], cur, dt, s, h->stem_c, h->stem_pad));
  }
  prof_end(pr, s);
  int gi = 0;
  for (size_t i = 0; i < h->blocks.size(); ++i) {
    const BlockCfg& b = h->blocks[i];
    const BlockDev& d = h->bdev[i];
    const long long m_in = (long long)n * b.hin * b.hin, m_out = (long long)n * b.hout * b.hout;
    const void* dw_in = cur;
    const int li = (int)i;
    if (b.expand) {
      pr = prof_begin(h, s, "expand", li, (double)m_in * (b.cin + b.mid) * es + (double)b.mid * b.cin * es,
                      2.0 * m_in * b.cin * b.mid);
      GemmCall g{&d.expand, cur, m_in, nullptr, 1, nullptr, 1, ws + W.e};
      int rc = run_gemm(h, plan, g, gi++, s);
      if (rc) return rc;
      prof_end(pr, s);
      dw_in = ws + W.e;
    }
    pool_t* pool_b = pool + d.pool_off * n;
    float* gate_b = gate + d.pool_off * n;
    pr = prof_begin(h, s, "dw", li, ((double)m_in + (double)m_out) * b.mid * es, 2.0 * m_out * b.mid * b.k * b.k);
    {
      bool done = false;
      if (dt == kBF16 && h->dw_impl == 0) {
        if (plan.dws.size() != h->blocks.size()) plan.dws.assign(h->blocks.size(), DwPlan());
        DwPlan& dp = plan.dws[i];
        if (dp.x != dw_in || dp.p.n != n || dp.p.pool != pool_b) {
          dp.x = nullptr;
          if (plan_dwconv_tma(dp, dw_in, n, b.hin, b.hin, b.mid, b.k, b.stride, d.dw_w, d.dw_b, ws + W.d, pool_b,
                              b.pad_lo) != 0)
            dp.x = nullptr;
        }
        if (dp.x != nullptr) {
          CU_OK(launch_dwconv_tma(dp, s));
          done = true;
        }
      }
      if (!done)
        CU_OK(launch_dwconv(dw_in, n, b.hin, b.hin, b.mid, b.k, b.stride, d.dw_w, d.dw_b, ws + W.d, pool_b, dt, s,
                            b.pad_lo));
    }
    prof_end(pr, s);
    pr = prof_begin(h, s, "se", li, (double)n * b.mid * 8.0 + 2.0 * b.mid * b.rd * 4.0, 4.0 * n * b.mid * b.rd);
    CU_OK(launch_se(pool_b, n, b.mid, b.rd, 1.0f / (float)(b.hout * b.hout), d.se_w1, d.se_b1, d.se_w2t, d.se_b2,
                    gate_b, s));
    prof_end(pr, s);
    pr = prof_begin(h, s, "project", li,
                    (double)m_out * (b.mid + b.cout * (b.skip ? 2 : 1)) * es + (double)b.mid * b.cout * es,
                    2.0 * m_out * b.mid * b.cout);
    GemmCall g{&d.project, ws + W.d, m_out, gate_b, b.hout * b.hout, b.skip ? cur : nullptr, 0, nxt};
    int rc = run_gemm(h, plan, g, gi++, s);
    if (rc) return rc;
    prof_end(pr, s);
    std::swap(cur, nxt);
  }
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  const int cin_head = h->blocks.back().cout;
  pr = prof_begin(h, s, "conv_head", -1, (double)n * hw * (cin_head + h->head_c) * es + (double)h->head_c * cin_head * es,
                  2.0 * n * hw * cin_head * h->head_c);
  GemmCall g{&h->head, cur, (long long)n * hw, nullptr, 1, nullptr, 0, ws + W.a};
  int rc = run_gemm(h, plan, g, gi++, s);
  prof_end(pr, s);
  return rc;
}

int check_ws(int n, const void* ws, size_t ws_bytes, size_t need) {
  if (n > 0) {
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 1023))
      return fail(B200DF_E_WORKSPACE, "workspace must be non-NULL and 1024-byte aligned");
    if (ws_bytes < need)
      return fail(B200DF_E_WORKSPACE, "workspace too small: %zu < %zu bytes for n=%d", ws_bytes, need, n);
  }
  return B200DF_OK;
}

int check_call(b200df_engine* h, int n, const void* ws, size_t ws_bytes, Workspace* W) {
  if (!h) return fail(B200DF_E_INVALID, "NULL handle");
  if (h->precision < 0) return fail(B200DF_E_STATE, "b200df_finalize has not been called");
  if (n < 0) return fail(B200DF_E_INVALID, "negative batch");
  if (h->net) return B200DF_OK;   // layer programs lay out their own workspace (check_net_call)
  *W = layout_workspace(h, n > 0 ? n : 1);
  return check_ws(n, ws, ws_bytes, W->total);
}

// Explicit Grad-CAM targets are validated on the host (pytorch-grad-cam's ClassifierOutputTarget raises IndexError on a
// class the model does not have, /root/reference/web_ui.py:278): one small D2H copy + a stream synchronize, only on
// the CAM entry points and only when the caller passes targets.
int check_targets(const int32_t* target_cls, int n, int num_classes, cudaStream_t s) {
  if (!target_cls || n <= 0) return B200DF_OK;
  std::vector<int32_t> t(n);
  CU_OK(cudaMemcpyAsync(t.data(), target_cls, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  CU_OK(cudaStreamSynchronize(s));
  for (int i = 0; i < n; ++i)
    if (t[i] < 0 || t[i] >= num_classes)
      return fail(B200DF_E_INVALID, "Grad-CAM target class %d (image %d) out of range [0, %d)", t[i], i, num_classes);
  return B200DF_OK;
}

int run_head(b200df_engine* h, int n, uint8_t* ws, const Workspace& W, float* logits, cudaStream_t s) {
  const int dt = h->precision == B200DF_PREC_BF16 ? kBF16 : kFP32;
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  float* pooled = reinterpret_cast<float*>(ws + W.pooled);
  const double es = h->precision == B200DF_PREC_BF16 ? 2.0 : 4.0;
  ProfRec* pr = prof_begin(h, s, "head_pool", -1, (double)n * hw * h->head_c * es + (double)n * h->head_c * 4, 8.0 * n * hw * h->head_c);
  CU_OK(launch_head_pool(ws + W.a, n, hw, h->head_c, h->head_scale, h->head_shift, pooled, dt, s));
  prof_end(pr, s);
  pr = prof_begin(h, s, "classifier", -1, (double)n * h->head_c * 4 + (double)n * h->num_classes * 4,
                  2.0 * n * h->head_c * h->num_classes);
  CU_OK(launch_classifier(pooled, n, h->head_c, h->fc_w, h->fc_b, h->num_classes, logits, s));
  prof_end(pr, s);
  return B200DF_OK;
}

}  // namespace

#include "nets.inc"

namespace {
int check_net_call(b200df_engine* h, int n, const void* ws, size_t ws_bytes, NetWs* W, int mode = 0) {
  *W = layout_net_workspace(h, n > 0 ? n : 1, mode);
  return check_ws(n, ws, ws_bytes, W->total);
}
}  // namespace

extern "C" int b200df_forward(b200df_handle h, const float* images, int n, float* logits, void* workspace,
                              size_t workspace_bytes, void* stream) {
  Workspace W;
  int rc = check_call(h, n, workspace, workspace_bytes, &W);
  if (rc || n == 0) return rc;
  if (!images || !logits) return fail(B200DF_E_INVALID, "NULL images/logits");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  if (h->net) {
    NetWs NW;
    rc = check_net_call(h, n, workspace, workspace_bytes, &NW);
    if (rc) return rc;
    return run_net(h, images, kInF32NCHW, nullptr, nullptr, n, ws, NW, logits, s);
  }
  rc = run_trunk(h, images, kInF32NCHW, nullptr, nullptr, n, ws, W, s);
  if (rc) return rc;
  return run_head(h, n, ws, W, logits, s);
}

extern "C" int b200df_forward_u8(b200df_handle h, const uint8_t* images_hwc, int n, const float* mean,
                                 const float* stdv, float* logits, void* workspace, size_t workspace_bytes,
                                 void* stream) {
  Workspace W;
  int rc = check_call(h, n, workspace, workspace_bytes, &W);
  if (rc || n == 0) return rc;
  if (!images_hwc || !logits) return fail(B200DF_E_INVALID, "NULL images/logits");
  float sc[3], sh[3];
  for (int c = 0; c < 3; ++c) {  // ToTensor: /255; Normalize: (x - mean)/std  (orchestrator.py:340-345)
    float sd = stdv ? stdv[c] : 1.0f, mu = mean ? mean[c] : 0.0f;
    sc[c] = 1.0f / (255.0f * sd);
    sh[c] = -mu / sd;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  if (h->net) {
    NetWs NW;
    rc = check_net_call(h, n, workspace, workspace_bytes, &NW);
    if (rc) return rc;
    return run_net(h, images_hwc, kInU8HWC, sc, sh, n, ws, NW, logits, s);
  }
  rc = run_trunk(h, images_hwc, kInU8HWC, sc, sh, n, ws, W, s);
  if (rc) return rc;
  return run_head(h, n, ws, W, logits, s);
}

extern "C" int b200df_forward_cam(b200df_handle h, const float* images, int n, const int32_t* target_cls,
                                  float* logits, float* cam_lowres, float* cam, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  Workspace W;
  int rc = check_call(h, n, workspace, workspace_bytes, &W);
  if (rc || n == 0) return rc;
  if (!images) return fail(B200DF_E_INVALID, "NULL images");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  const int dt = h->precision == B200DF_PREC_BF16 ? kBF16 : kFP32;
  rc = check_targets(target_cls, n, h->num_classes, s);
  if (rc) return rc;
  if (h->net) {
    Net* net = h->net.get();
    const bool has_bwd = net->fwd_end >= 0 && (size_t)net->fwd_end < net->ops.size();
    if (!net->wcam && !has_bwd) return fail(B200DF_E_STATE, "Grad-CAM is not built for this architecture");
    NetWs NW;
    if (!net->wcam) {
      // FasterViT: the target activation (global_tokenizer.pos_embed, web_ui.py:94-112) sits in front of the transformer
      // blocks -> forward program with activation checkpoints, then the backward program (nets.inc, cam_bwd.cu)
      rc = check_net_call(h, n, workspace, workspace_bytes, &NW, 1);
      if (rc) return rc;
      float* lg = logits ? logits : reinterpret_cast<float*>(ws + NW.logits);
      float* low = cam_lowres ? cam_lowres : reinterpret_cast<float*>(ws + NW.cam_low);
      CamRun cr;
      cr.target_cls = target_cls; cr.logits = lg; cr.cam_low = low;
      rc = run_net(h, images, kInF32NCHW, nullptr, nullptr, n, ws, NW, lg, s, 0, &cr);
      if (rc) return rc;
      rc = run_net(h, nullptr, kInF32NCHW, nullptr, nullptr, n, ws, NW, lg, s, 1, &cr);
      if (rc) return rc;
      if (cam) {
        ProfRec* pr = prof_begin(h, s, "cam_upsample", -1, (double)n * (net->feat_side * net->feat_side + (double)h->img * h->img) * 4,
                                 8.0 * n * h->img * h->img);
        CU_OK(launch_cam_upsample(low, n, net->feat_side, h->img, cam, s));
        prof_end(pr, s);
      }
      return B200DF_OK;
    }
    rc = check_net_call(h, n, workspace, workspace_bytes, &NW);
    if (rc) return rc;
    float* lg = logits ? logits : reinterpret_cast<float*>(ws + NW.logits);
    rc = run_net(h, images, kInF32NCHW, nullptr, nullptr, n, ws, NW, lg, s);
    if (rc) return rc;
    const int fs = net->feat_side, fhw = fs * fs;
    float* low = cam_lowres ? cam_lowres : reinterpret_cast<float*>(ws + NW.cam_low);
    ProfRec* pr = prof_begin(h, s, "cam", -1, (double)n * fhw * net->feat_c * (dt == kBF16 ? 2.0 : 4.0) + (double)n * fhw * 4,
                             2.0 * n * fhw * net->feat_c);
    CU_OK(launch_cam_linear(ws + NW.off[net->a_buf], n, fhw, net->feat_c, net->wcam, h->num_classes, target_cls, lg, low, dt, s));
    prof_end(pr, s);
    if (cam) {
      pr = prof_begin(h, s, "cam_upsample", -1, (double)n * (fhw + (double)h->img * h->img) * 4, 8.0 * n * h->img * h->img);
      CU_OK(launch_cam_upsample(low, n, fs, h->img, cam, s));
      prof_end(pr, s);
    }
    return B200DF_OK;
  }
  const int side = h->blocks.back().hout, hw = side * side;
  rc = run_trunk(h, images, kInF32NCHW, nullptr, nullptr, n, ws, W, s);
  if (rc) return rc;
  float* lg = logits ? logits : reinterpret_cast<float*>(ws + W.logits);
  rc = run_head(h, n, ws, W, lg, s);
  if (rc) return rc;
  float* low = cam_lowres ? cam_lowres : reinterpret_cast<float*>(ws + W.cam_low);
  const double es = dt == kBF16 ? 2.0 : 4.0;
  ProfRec* pr = prof_begin(h, s, "cam", -1, (double)n * hw * h->head_c * es + (double)n * hw * 4, 12.0 * n * hw * h->head_c);
  CU_OK(launch_cam(ws + W.a, n, hw, h->head_c, h->head_scale, h->head_shift, h->fc_w, h->num_classes, target_cls, lg, low,
                   dt, s));
  prof_end(pr, s);
  if (cam) {
    pr = prof_begin(h, s, "cam_upsample", -1, (double)n * (hw + (double)h->img * h->img) * 4, 8.0 * n * h->img * h->img);
    CU_OK(launch_cam_upsample(low, n, side, h->img, cam, s));
    prof_end(pr, s);
  }
  return B200DF_OK;
}

// aux: float32 [n, num_classes] (nullable).  For the affine-head architectures (EfficientFormerV2: everything after
// the Grad-CAM activation A is affine, but part of it -- the block's residual input -- is not a function of A) it
// receives logits - lin(A), so that head_ex(A, aux) reproduces the logits; the MBConv engines write zeros.
static int trunk_impl(b200df_handle h, const float* images, int n, float* a_nchw, float* aux, void* workspace,
                      size_t workspace_bytes, void* stream) {
  Workspace W;
  int rc = check_call(h, n, workspace, workspace_bytes, &W);
  if (rc || n == 0) return rc;
  if (!images || !a_nchw) return fail(B200DF_E_INVALID, "NULL images/output");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  const int dt = h->precision == B200DF_PREC_BF16 ? kBF16 : kFP32;
  if (h->net) {
    Net* net = h->net.get();
    const bool has_bwd = net->fwd_end >= 0 && (size_t)net->fwd_end < net->ops.size();
    if (!net->wcam && !has_bwd)
      return fail(B200DF_E_STATE, "the split trunk/head (Grad-CAM hook) path is not built for this architecture");
    if (!aux) return fail(B200DF_E_INVALID, "this architecture's split head needs the aux output (b200df_trunk_ex)");
    NetWs NW;
    rc = check_net_call(h, n, workspace, workspace_bytes, &NW, net->wcam ? 0 : 1);
    if (rc) return rc;
    const int fhw = net->feat_side * net->feat_side;
    if (!net->wcam) {
      // FasterViT: the whole forward with activation checkpoints; aux = the logits themselves (head_ex returns them,
      // head_backward runs the backward program on the checkpoints this call leaves in `workspace`)
      CamRun cr;
      rc = run_net(h, images, kInF32NCHW, nullptr, nullptr, n, ws, NW, aux, s, 0, &cr);
      if (rc) return rc;
      CU_OK(launch_nhwc_to_nchw_f32(ws + NW.off[net->a_buf], n, fhw, net->feat_c, a_nchw, dt, s));
      h->cam_ws = workspace;
      h->cam_n = n;
      return B200DF_OK;
    }
    float* lg = reinterpret_cast<float*>(ws + NW.logits);
    rc = run_net(h, images, kInF32NCHW, nullptr, nullptr, n, ws, NW, lg, s);
    if (rc) return rc;
    CU_OK(launch_nhwc_to_nchw_f32(ws + NW.off[net->a_buf], n, fhw, net->feat_c, a_nchw, dt, s));
    CU_OK(launch_lin_head(a_nchw, lg, n, fhw, net->feat_c, net->wcam, h->num_classes, 1, aux, s));
    return B200DF_OK;
  }
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  rc = run_trunk(h, images, kInF32NCHW, nullptr, nullptr, n, ws, W, s);
  if (rc) return rc;
  CU_OK(launch_nhwc_to_nchw_f32(ws + W.a, n, hw, h->head_c, a_nchw, dt, s));
  if (aux) CU_OK(cudaMemsetAsync(aux, 0, (size_t)n * h->num_classes * sizeof(float), s));
  return B200DF_OK;
}

extern "C" int b200df_trunk(b200df_handle h, const float* images, int n, float* a_nchw, void* workspace,
                            size_t workspace_bytes, void* stream) {
  return trunk_impl(h, images, n, a_nchw, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int b200df_trunk_ex(b200df_handle h, const float* images, int n, float* a_nchw, float* aux, void* workspace,
                               size_t workspace_bytes, void* stream) {
  return trunk_impl(h, images, n, a_nchw, aux, workspace, workspace_bytes, stream);
}

// Scratch for the split head: pooled [n, head_c] lives in a per-ENGINE buffer on the engine's device, grown on
// demand (outside stream capture), so head()/head_backward() need no caller workspace.
static int head_scratch(b200df_engine* h, int n, float** out) {
  const size_t need = (size_t)n * h->head_c * sizeof(float);
  if (need > h->head_scratch_cap) {
    if (h->head_scratch) CU_OK(cudaFree(h->head_scratch));
    h->head_scratch = nullptr;
    h->head_scratch_cap = 0;
    CU_OK(cudaMalloc(&h->head_scratch, need));
    h->head_scratch_cap = need;
  }
  *out = h->head_scratch;
  return B200DF_OK;
}

static int head_impl(b200df_handle h, const float* a_nchw, const float* aux, int n, float* logits, void* stream) {
  if (!h || h->precision < 0) return fail(B200DF_E_STATE, "engine not finalized");
  if (n <= 0) return n == 0 ? B200DF_OK : fail(B200DF_E_INVALID, "negative batch");
  if (!a_nchw || !logits) return fail(B200DF_E_INVALID, "NULL argument");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (h->net) {
    Net* net = h->net.get();
    const bool has_bwd = net->fwd_end >= 0 && (size_t)net->fwd_end < net->ops.size();
    if (!net->wcam && !has_bwd)
      return fail(B200DF_E_STATE, "the split trunk/head (Grad-CAM hook) path is not built for this architecture");
    if (!aux) return fail(B200DF_E_INVALID, "this architecture's split head needs the aux input (b200df_head_ex)");
    if (!net->wcam) {   // FasterViT: aux holds the logits of the trunk call (the path from A to them is not a function of A alone)
      if (logits != aux)
        CU_OK(cudaMemcpyAsync(logits, aux, (size_t)n * h->num_classes * sizeof(float), cudaMemcpyDeviceToDevice, s));
      return B200DF_OK;
    }
    CU_OK(launch_lin_head(a_nchw, aux, n, net->feat_side * net->feat_side, net->feat_c, net->wcam, h->num_classes, 0, logits, s));
    return B200DF_OK;
  }
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  float* pooled = nullptr;
  int rc = head_scratch(h, n, &pooled);
  if (rc) return rc;
  CU_OK(launch_head_from_nchw(a_nchw, n, hw, h->head_c, h->head_scale, h->head_shift, pooled, s));
  CU_OK(launch_classifier(pooled, n, h->head_c, h->fc_w, h->fc_b, h->num_classes, logits, s));
  return B200DF_OK;
}

extern "C" int b200df_head(b200df_handle h, const float* a_nchw, int n, float* logits, void* stream) {
  return head_impl(h, a_nchw, nullptr, n, logits, stream);
}

extern "C" int b200df_head_ex(b200df_handle h, const float* a_nchw, const float* aux, int n, float* logits, void* stream) {
  return head_impl(h, a_nchw, aux, n, logits, stream);
}

// ---- allocation-free split head (include/b200df.h: b200df_head_static) -----------------------------------------------
// Same two launches as head_impl's MBConv branch, with the pooled [n, head_c] scratch in a file-scope device array
// instead of the engine's cudaMalloc'd buffer.  The array's NAME is a host-side stand-in in host code: its device
// address is looked up with cudaGetSymbolAddress, once per device, with that device current, and cached.
static __device__ __align__(256) float g_head_static_scratch[B200DF_HEAD_STATIC_MAX_FLOATS];
static float* g_head_static_ptr[64] = {};

static int head_static_scratch(int device, float** out) {
  if (device < 0 || device >= 64) return fail(B200DF_E_INVALID, "b200df_head_static: device ordinal %d out of range", device);
  if (!g_head_static_ptr[device]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_head_static_scratch);
    if (e != cudaSuccess || !p)
      return fail(B200DF_E_CUDA, "b200df_head_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    g_head_static_ptr[device] = p;
  }
  *out = g_head_static_ptr[device];
  return B200DF_OK;
}

extern "C" int b200df_head_static(b200df_handle h, const float* a_nchw, int n, float* logits, void* stream) {
  if (!h || h->precision < 0) return fail(B200DF_E_STATE, "engine not finalized");
  if (n <= 0) return n == 0 ? B200DF_OK : fail(B200DF_E_INVALID, "negative batch");
  if (!a_nchw || !logits) return fail(B200DF_E_INVALID, "NULL argument");
  if (h->net)
    return fail(B200DF_E_STATE, "b200df_head_static: this architecture's split head (b200df_head_ex) uses no scratch memory");
  if ((size_t)n * h->head_c > (size_t)B200DF_HEAD_STATIC_MAX_FLOATS)
    return fail(B200DF_E_INVALID, "b200df_head_static: n * head_c = %d * %d exceeds the static scratch of %d floats",
                n, h->head_c, (int)B200DF_HEAD_STATIC_MAX_FLOATS);
  int dev = -1;
  CU_OK(cudaGetDevice(&dev));
  if (dev != h->device)
    return fail(B200DF_E_STATE, "b200df_head_static: current device %d is not the engine's device %d", dev, h->device);
  float* pooled = nullptr;
  int rc = head_static_scratch(dev, &pooled);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  // every pooled[n * head_c] element is overwritten before it is read: nothing has to start at zero
  CU_OK(launch_head_from_nchw(a_nchw, n, hw, h->head_c, h->head_scale, h->head_shift, pooled, s));
  CU_OK(launch_classifier(pooled, n, h->head_c, h->fc_w, h->fc_b, h->num_classes, logits, s));
  return B200DF_OK;
}

extern "C" int b200df_head_backward(b200df_handle h, const float* a_nchw, const float* grad_logits, int n,
                                    float* grad_a_nchw, void* stream) {
  if (!h || h->precision < 0) return fail(B200DF_E_STATE, "engine not finalized");
  if (n <= 0) return n == 0 ? B200DF_OK : fail(B200DF_E_INVALID, "negative batch");
  if (!a_nchw || !grad_logits || !grad_a_nchw) return fail(B200DF_E_INVALID, "NULL argument");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (h->net) {
    Net* net = h->net.get();
    const bool has_bwd = net->fwd_end >= 0 && (size_t)net->fwd_end < net->ops.size();
    if (!net->wcam && !has_bwd)
      return fail(B200DF_E_STATE, "the split trunk/head (Grad-CAM hook) path is not built for this architecture");
    if (!net->wcam) {
      // FasterViT: the backward program on the checkpoints of the LAST b200df_trunk_ex call of this engine (its workspace
      // must still be alive and untouched; same batch size)
      if (!h->cam_ws || h->cam_n != n)
        return fail(B200DF_E_STATE, "b200df_head_backward: no matching b200df_trunk_ex call (batch %d, last trunk batch %d)", n,
                    h->cam_n);
      const NetWs NW = layout_net_workspace(h, n, 1);
      CamRun cr;
      cr.grad_logits = grad_logits;
      cr.grad_a = grad_a_nchw;
      return run_net(h, nullptr, kInF32NCHW, nullptr, nullptr, n, reinterpret_cast<uint8_t*>(h->cam_ws), NW, nullptr, s, 1, &cr);
    }
    CU_OK(launch_lin_head_backward(grad_logits, n, net->feat_side * net->feat_side, net->feat_c, net->wcam, h->num_classes,
                                   grad_a_nchw, s));
    return B200DF_OK;
  }
  const int hw = h->blocks.back().hout * h->blocks.back().hout;
  CU_OK(launch_head_backward(a_nchw, grad_logits, n, hw, h->head_c, h->head_scale, h->head_shift, h->fc_w, h->num_classes,
                             grad_a_nchw, s));
  return B200DF_OK;
}

extern "C" int b200df_softmax_pack(const float* logits, const int64_t* labels, int n, int num_classes, float* packed,
                                   void* stream) {
  if (n < 0 || num_classes < 1 || (n > 0 && (!logits || !packed))) return fail(B200DF_E_INVALID, "bad argument");
  CU_OK(launch_softmax_pack(logits, labels, n, num_classes, packed, reinterpret_cast<cudaStream_t>(stream)));
  return B200DF_OK;
}

// ---- Grad-CAM overlay (web_ui.py:72-93, 281-282) ---------------------------------------------------------------
extern "C" int b200df_cam_overlay(const float* images, const float* cam, int n, int h, int w, int denorm,
                                  const float* mean, const float* stdv, float image_weight, unsigned char* out,
                                  void* stream) {
  if (n < 0 || h < 1 || w < 1 || (long long)h * w > (1 << 28) || !mean || !stdv || (n > 0 && (!images || !cam || !out)))
    return fail(B200DF_E_INVALID, "bad argument");
  if (!(image_weight >= 0.f && image_weight <= 1.f))
    return fail(B200DF_E_INVALID, "image_weight should be in the range [0, 1]");
  CU_OK(launch_cam_overlay(images, cam, n, h, w, denorm, mean, stdv, image_weight, out,
                           reinterpret_cast<cudaStream_t>(stream)));
  return B200DF_OK;
}

// ---- on-device metrics tail (orchestrator.py:536-546, 605-631) ------------------------------------------------
extern "C" int b200df_threshold_hist(const float* scores, const int64_t* labels, int n, const double* thresholds,
                                     int nthr, long long* hist, void* stream) {
  if (n < 0 || nthr < 1 || !thresholds || !hist || (n > 0 && (!scores || !labels)))
    return fail(B200DF_E_INVALID, "bad argument");
  CU_OK(launch_threshold_hist(scores, labels, n, thresholds, nthr, reinterpret_cast<unsigned long long*>(hist),
                              reinterpret_cast<cudaStream_t>(stream)));
  return B200DF_OK;
}

extern "C" int b200df_confusion_matrix(const int64_t* targets, const int64_t* preds, int n, int num_classes,
                                       long long* cm, void* stream) {
  if (n < 0 || num_classes < 1 || !cm || (n > 0 && (!targets || !preds))) return fail(B200DF_E_INVALID, "bad argument");
  CU_OK(launch_confusion(targets, preds, n, num_classes, reinterpret_cast<unsigned long long*>(cm),
                         reinterpret_cast<cudaStream_t>(stream)));
  return B200DF_OK;
}

extern "C" int b200df_auc_counts(const float* pos_scores, int n_pos, const float* neg_scores_sorted, int n_neg,
                                 long long* less_equal, void* stream) {
  if (n_pos < 0 || n_neg < 0 || !less_equal) return fail(B200DF_E_INVALID, "bad argument");
  CU_OK(launch_auc_count(pos_scores, n_pos, neg_scores_sorted, n_neg, reinterpret_cast<unsigned long long*>(less_equal),
                         reinterpret_cast<cudaStream_t>(stream)));
  return B200DF_OK;
}

// ---- debug checkpoints of the layer programs (tools/layer_diff.py, tests) ---------------------------------------
extern "C" int b200df_debug_num_marks(b200df_handle h) { return (h && h->net) ? (int)h->net->marks.size() : 0; }

extern "C" int b200df_debug_mark(b200df_handle h, int index, int n, const char** name, int* op, int* seqs_rows_c,
                                 size_t* ws_offset) {
  if (!h || !h->net || index < 0 || index >= (int)h->net->marks.size()) return fail(B200DF_E_INVALID, "mark index out of range");
  if (h->precision < 0) return fail(B200DF_E_STATE, "b200df_finalize has not been called");
  const NetMark& m = h->net->marks[index];
  if (name) *name = m.name.c_str();
  if (op) *op = m.op;
  if (seqs_rows_c) { seqs_rows_c[0] = m.seq_rows; seqs_rows_c[1] = m.row_off; seqs_rows_c[2] = m.rows; seqs_rows_c[3] = m.c; }
  if (ws_offset) *ws_offset = layout_net_workspace(h, n > 0 ? n : 1).off[m.buf];
  return B200DF_OK;
}

extern "C" int b200df_debug_stop_after(b200df_handle h, int op) {
  if (!h || !h->net) return fail(B200DF_E_INVALID, "not a layer-program engine");
  h->net->stop_after = op;
  return B200DF_OK;
}

// ---- profiling (bench.py's per-kernel roofline; never enabled inside a timed region) ---------------------------
extern "C" int b200df_set_profile(b200df_handle h, int enable) {
  if (!h) return fail(B200DF_E_INVALID, "NULL handle");
  h->profile = enable != 0;
  h->prof_used = 0;
  return B200DF_OK;
}

extern "C" int b200df_profile_count(b200df_handle h) { return h ? (int)h->prof_used : 0; }

extern "C" int b200df_profile_get(b200df_handle h, int index, const char** name, float* ms, double* bytes,
                                  double* flops) {
  if (!h || index < 0 || index >= (int)h->prof_used) return fail(B200DF_E_INVALID, "profile index out of range");
  ProfRec& r = h->prof[index];
  CU_OK(cudaEventSynchronize(r.e1));
  float t = 0.f;
  CU_OK(cudaEventElapsedTime(&t, r.e0, r.e1));
  if (name) *name = r.name;
  if (ms) *ms = t;
  if (bytes) *bytes = r.bytes;
  if (flops) *flops = r.flops;
  return B200DF_OK;
}

extern "C" int b200df_profile_reset(b200df_handle h) {
  if (!h) return fail(B200DF_E_INVALID, "NULL handle");
  h->prof_used = 0;
  return B200DF_OK;
}

// ---- per-kernel entry points (unit tests / ncu) ---------------------------------------------------------------
static cudaStream_t S(void* stream) { return reinterpret_cast<cudaStream_t>(stream); }

extern "C" int b200df_k_stem(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias, void* out,
                             int dtype, void* stream) {
  CU_OK(launch_stem(x_nchw, n, h, w, wgt, bias, out, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_stem_ex(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias, void* out,
                                int dtype, int oc, int pad, void* stream) {
  CU_OK(launch_stem(x_nchw, n, h, w, wgt, bias, out, dtype, S(stream), oc, pad));
  return B200DF_OK;
}

extern "C" int b200df_k_dwconv_ex(const void* x, int n, int h, int w, int c, int k, int stride, int pad_lo,
                                  const float* wgt, const float* bias, void* out, long long* pool, int dtype,
                                  void* stream) {
  CU_OK(launch_dwconv(x, n, h, w, c, k, stride, wgt, bias, out, pool, dtype, S(stream), pad_lo));
  return B200DF_OK;
}

extern "C" int b200df_k_dwconv(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                               const float* bias, void* out, long long* pool, int dtype, void* stream) {
  CU_OK(launch_dwconv(x, n, h, w, c, k, stride, wgt, bias, out, pool, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_se(const long long* pool, int n, int c, int rd, float inv_hw, const float* w1, const float* b1,
                           const float* w2t, const float* b2, float* gate, void* stream) {
  CU_OK(launch_se(pool, n, c, rd, inv_hw, w1, b1, w2t, b2, gate, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_pwconv(const void* x, long long m, int cin, int cout, const void* wgt, const float* bias,
                               const float* gate, int rows_per_img, const void* residual, int act, void* out,
                               int dtype, int impl, void* stream) {
  if (impl == 1) {
    CU_OK(launch_pwconv_simt(x, m, cin, cout, wgt, bias, gate, rows_per_img, residual, act, out, dtype, S(stream)));
    return B200DF_OK;
  }
  if (dtype != kBF16) return fail(B200DF_E_INVALID, "the tcgen05 GEMM is bf16-only");
  if (cin % 8 || cout % 8) return fail(B200DF_E_INVALID, "cin and cout must be multiples of 8");
  // test-only convenience: pad weights/bias to the N tile on the fly (the engine does this once in finalize)
  int cout_pad = 0;
  const int bn = choose_bn(cout, &cout_pad);
  void* wpad = nullptr;
  float* bpad = nullptr;
  CU_OK(cudaMalloc(&wpad, (size_t)cout_pad * cin * 2));
  CU_OK(cudaMalloc(&bpad, (size_t)cout_pad * sizeof(float)));
  cudaStream_t s = S(stream);

  CU_OK(cudaMemsetAsync(wpad, 0, (size_t)cout_pad * cin * 2, s));
  CU_OK(cudaMemsetAsync(bpad, 0, (size_t)cout_pad * sizeof(float), s));
  CU_OK(cudaMemcpyAsync(wpad, wgt, (size_t)cout * cin * 2, cudaMemcpyDeviceToDevice, s));
  CU_OK(cudaMemcpyAsync(bpad, bias, (size_t)cout * sizeof(float), cudaMemcpyDeviceToDevice, s));
  TcGemmArgs a;
  memset(&a, 0, sizeof(a));
  if (gate && (rows_per_img <= 0 || m % rows_per_img)) {
    cudaFree(wpad);
    cudaFree(bpad);
    return fail(B200DF_E_INVALID, "gated tcgen05 GEMM needs m to be a multiple of rows_per_img");
  }
  int r1 = encode_tmap_act_bf16(a.tmap_a, x, m, cin, gate ? rows_per_img : 0, 128);
  int r2 = encode_tmap_2d_bf16(a.tmap_b, wpad, cout_pad, cin, bn);
  int r3 = encode_tmap_act_bf16(a.tmap_out, out, m, cout, gate ? rows_per_img : 0, 32);
  int rc = B200DF_OK;
  if (r1 || r2 || r3) {
    rc = fail(B200DF_E_CUDA, "cuTensorMapEncodeTiled failed (%d, %d, %d)", r1, r2, r3);
  } else {
    a.m = m; a.cin = cin; a.cout = cout; a.cout_pad = cout_pad; a.bn = bn;
    a.bias = bpad; a.gate = gate; a.rows_per_img = rows_per_img; a.residual = residual; a.act = act; a.out = out;
    cudaError_t e = launch_pwconv_tc(a, s);
    if (e != cudaSuccess) rc = fail(B200DF_E_CUDA, "launch_pwconv_tc: %s", cudaGetErrorString(e));
  }
  cudaError_t es = cudaStreamSynchronize(s);
  cudaFree(wpad);
  cudaFree(bpad);
  if (rc == B200DF_OK && es != cudaSuccess) rc = fail(B200DF_E_CUDA, "pwconv_tc execution: %s", cudaGetErrorString(es));
  return rc;
}

extern "C" int b200df_k_head_pool(const void* a, int n, int hw, int c, const float* scale, const float* shift,
                                  float* pooled, int dtype, void* stream) {
  CU_OK(launch_head_pool(a, n, hw, c, scale, shift, pooled, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_cam(const void* a, int n, int hw, int c, const float* scale, const float* shift,
                            const float* wfc, int num_classes, const int32_t* target_cls, float* cam_lowres,
                            int dtype, void* stream) {
  if (!target_cls) return fail(B200DF_E_INVALID, "target_cls is required for the standalone CAM kernel");
  CU_OK(launch_cam(a, n, hw, c, scale, shift, wfc, num_classes, target_cls, nullptr, cam_lowres, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_cam_upsample(const float* cam_lowres, int n, int side, int out_side, float* cam,
                                     void* stream) {
  CU_OK(launch_cam_upsample(cam_lowres, n, side, out_side, cam, S(stream)));
  return B200DF_OK;
}

// ---- per-kernel entry points of the transformer-style backbones ------------------------------------------------------
extern "C" int b200df_k_stem_act(const float* x_nchw, int n, int h, int w, const float* wgt, const float* bias, void* out,
                                 int dtype, int oc, int pad, int act, void* stream) {
  CU_OK(launch_stem(x_nchw, n, h, w, wgt, bias, out, dtype, S(stream), oc, pad, act));
  return B200DF_OK;
}

extern "C" int b200df_k_dwconv_act(const void* x, int n, int h, int w, int c, int k, int stride, const float* wgt,
                                   const float* bias, void* out, int act, int dtype, void* stream) {
  CU_OK(launch_dwconv(x, n, h, w, c, k, stride, wgt, bias, out, nullptr, dtype, S(stream), -1, act));
  return B200DF_OK;
}

extern "C" int b200df_k_conv3x3(const void* x, int n, int h, int w, int cin, int cout, int stride, const void* wgt,
                                const float* bias, const void* residual, int act, void* out, int dtype, int impl,
                                void* stream) {
  if (impl == 1 || dtype != kBF16) {
    CU_OK(launch_conv3x3_simt(x, n, h, w, cin, cout, stride, wgt, bias, residual, act, out, dtype, S(stream)));
    return B200DF_OK;
  }
  if (cin % 8 || cout % 8) return fail(B200DF_E_INVALID, "cin and cout must be multiples of 8");
  // test-only convenience: re-lay the [cout, 9*cin] weights out as [cout_pad, 9, cin_pad] on the fly (the engine does
  // this once in finalize)
  int cout_pad = 0;
  const int bn = choose_bn(cout, &cout_pad), cinp = conv3x3_tc_cin_pad(cin);
  cudaStream_t s = S(stream);
  void* wpad = nullptr;
  float* bpad = nullptr;
  CU_OK(cudaMalloc(&wpad, (size_t)cout_pad * 9 * cinp * 2));
  CU_OK(cudaMalloc(&bpad, (size_t)cout_pad * sizeof(float)));
  CU_OK(cudaMemsetAsync(wpad, 0, (size_t)cout_pad * 9 * cinp * 2, s));
  CU_OK(cudaMemsetAsync(bpad, 0, (size_t)cout_pad * sizeof(float), s));
  CU_OK(cudaMemcpy2DAsync(wpad, (size_t)cinp * 2, wgt, (size_t)cin * 2, (size_t)cin * 2, (size_t)cout * 9,
                          cudaMemcpyDeviceToDevice, s));
  CU_OK(cudaMemcpyAsync(bpad, bias, (size_t)cout * sizeof(float), cudaMemcpyDeviceToDevice, s));
  Conv3Plan plan;
  int rc = B200DF_OK;
  if (plan_conv3x3_tc(plan, x, n, h, w, cin, cout, cout_pad, bn, stride, wpad, bpad, residual, act, out) != 0) {
    rc = fail(B200DF_E_INVALID, "shape not supported by the tcgen05 3x3 conv (cin %d cout %d stride %d act %d)", cin, cout,
              stride, act);
  } else {
    cudaError_t e = launch_conv3x3_tc(plan, s);
    if (e != cudaSuccess) rc = fail(B200DF_E_CUDA, "launch_conv3x3_tc: %s", cudaGetErrorString(e));
  }
  cudaError_t es = cudaStreamSynchronize(s);
  cudaFree(wpad);
  cudaFree(bpad);
  if (rc == B200DF_OK && es != cudaSuccess) rc = fail(B200DF_E_CUDA, "conv3x3_tc execution: %s", cudaGetErrorString(es));
  return rc;
}

extern "C" int b200df_k_attention(const void* q, const void* k, const void* v, int q_pitch, int k_pitch, int v_pitch,
                                  int q_off, int k_off, int v_off, int nseq, int nq, int nk, int dk, int dv, int heads,
                                  float scale, const float* bias, const float* th1_w, const float* th1_b,
                                  const float* th2_w, const float* th2_b, const void* vlocal, int act, void* out,
                                  int dtype, void* stream) {
  AttnArgs a;
  memset(&a, 0, sizeof(a));
  a.q = q; a.k = k; a.v = v;
  a.q_pitch = q_pitch; a.k_pitch = k_pitch; a.v_pitch = v_pitch;
  a.q_off = q_off; a.k_off = k_off; a.v_off = v_off;
  a.nseq = nseq; a.nq = nq; a.nk = nk; a.dk = dk; a.dv = dv; a.heads = heads; a.scale = scale;
  a.bias = bias; a.th1_w = th1_w; a.th1_b = th1_b; a.th2_w = th2_w; a.th2_b = th2_b;
  a.vlocal = vlocal; a.act = act; a.out = out;
  float* frag = nullptr;
  if (dtype == kBF16 && bias) {   // test-only convenience: the engine lays the table out once at finalize
    CU_OK(cudaMalloc(&frag, attention_bias_frag_elems(heads, nq, nk) * sizeof(float)));
    CU_OK(launch_attention_bias_frag(bias, heads, nq, nk, frag, S(stream)));
    a.bias_frag = frag;
  }
  cudaError_t e = launch_attention(a, dtype, S(stream));
  if (frag) {
    cudaStreamSynchronize(S(stream));
    cudaFree(frag);
  }
  CU_OK(e);
  return B200DF_OK;
}

extern "C" int b200df_k_layernorm(const void* x, void* y, const float* gamma, const float* beta, long long rows, int c,
                                  float eps, int dtype, void* stream) {
  CU_OK(launch_layernorm(x, y, gamma, beta, rows, c, eps, dtype, S(stream)));
  return B200DF_OK;
}

// ---- backward-data kernels of the FasterViT Grad-CAM program (cam_bwd.cu), exposed for direct tests ----
extern "C" int b200df_k_attention_bwd(const void* qkv, int pitch, int q_off, int k_off, int v_off, int nseq, int n_tok, int d,
                                      int heads, float scale, const float* bias, const void* dout, void* dqkv, int dtype,
                                      void* stream) {
  AttnArgs a;
  memset(&a, 0, sizeof(a));
  a.q = a.k = a.v = qkv;
  a.q_pitch = a.k_pitch = a.v_pitch = pitch;
  a.q_off = q_off; a.k_off = k_off; a.v_off = v_off;
  a.nseq = nseq; a.nq = a.nk = n_tok; a.dk = a.dv = d; a.heads = heads; a.scale = scale; a.bias = bias;
  CU_OK(launch_attention_bwd(a, dout, dqkv, dqkv, dqkv, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_layernorm_bwd(const void* x, const float* gamma, const void* dy, const void* res, void* out,
                                      long long rows, int c, float eps, int dtype, void* stream) {
  CU_OK(launch_layernorm_bwd(x, gamma, dy, res, out, rows, c, eps, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_gelu_bwd(const void* pre, void* g, long long total, int dtype, void* stream) {
  CU_OK(launch_gelu_bwd(pre, g, total, dtype, S(stream)));
  return B200DF_OK;
}

// wgt: [cout, 9, cin] in `dtype` (tap-major, the layout the forward kernels read with cin_pad == cin)
extern "C" int b200df_k_conv3x3_s2_bwd(const void* dy, const void* wgt, void* dx, int n, int hin, int cin, int cout, int dtype,
                                       void* stream) {
  CU_OK(launch_conv3x3_s2_bwd(dy, wgt, dx, n, hin, cin, cin, cout, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_ct_pool_bwd(const void* gct, float* ga_nchw, int n, int side, int c, int kernel, int stride, int cs,
                                    int dtype, void* stream) {
  CU_OK(launch_ct_pool_bwd(gct, ga_nchw, n, side, c, kernel, stride, cs, dtype, S(stream)));
  return B200DF_OK;
}

extern "C" int b200df_k_upsample2x_act(const void* x, void* y, int n, int side, int c, int act, int dtype, void* stream) {
  CU_OK(launch_upsample2x_act(x, y, n, side, c, act, dtype, S(stream)));
  return B200DF_OK;
}

