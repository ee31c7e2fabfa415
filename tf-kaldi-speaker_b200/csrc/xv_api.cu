// libxvector.so C-ABI, part 1: context life cycle, config -> layer plan, weight store, BN folding.
// Mirrors what Trainer.__init__/build("predict")/load() set up in the reference (model/trainer.py:87-219,
// 277-295, 325-338) and the layer table of model/tdnn.py:36-179 + model/pooling.py:96-229.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>

#include "xv_ctx.h"

namespace xv {

static thread_local std::string g_create_err;

int fail(xv_ctx* c, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) c->err = buf; else g_create_err = buf;
  return code;
}

int cuda_fail(xv_ctx* c, cudaError_t e, const char* what) {
  return fail(c, XV_ECUDA, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}

int ensure(xv_ctx* c, DevBuf* b, size_t bytes, bool zero) {
  if (bytes == 0) bytes = 16;
  if (b->bytes < bytes) {
    if (b->p) {
      cudaError_t e = cudaStreamSynchronize(c->stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "sync before buffer growth");
      cudaFree(b->p);
      b->p = nullptr;
      b->bytes = 0;
    }
    const size_t want = (bytes + 255) & ~size_t(255);
    cudaError_t e = cudaMalloc(&b->p, want);
    if (e != cudaSuccess) {
      b->p = nullptr;
      fail(c, XV_ENOMEM, "cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e));
      return XV_ENOMEM;
    }
    b->bytes = want;
    if (zero) {
      e = cudaMemsetAsync(b->p, 0, want, c->stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "cudaMemsetAsync");
    }
  }
  return XV_OK;
}

void release(DevBuf* b) {
  if (b->p) cudaFree(b->p);
  b->p = nullptr;
  b->bytes = 0;
}

void count_launch(xv_ctx* c, const char* family) {
  c->launches++;
  c->launch_by_family[family]++;
}

ProfScope::ProfScope(xv_ctx* ctx, const char* family) : c(ctx), idx(-1) {
  count_launch(c, family);
  if (!c->profiling) return;
  int fam = -1;
  for (size_t i = 0; i < c->prof_family_names.size(); ++i)
    if (c->prof_family_names[i] == family) fam = static_cast<int>(i);
  if (fam < 0) {
    c->prof_family_names.push_back(family);
    fam = static_cast<int>(c->prof_family_names.size()) - 1;
  }
  ProfileSlot s;
  if (cudaEventCreate(&s.a) != cudaSuccess || cudaEventCreate(&s.b) != cudaSuccess) return;
  s.family = fam;
  cudaEventRecord(s.a, c->stream);
  c->prof.push_back(s);
  idx = static_cast<int>(c->prof.size()) - 1;
}
ProfScope::~ProfScope() {
  if (idx >= 0) cudaEventRecord(c->prof[idx].b, c->stream);
}

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);   // NaN
  const uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return static_cast<uint16_t>(u >> 16);
}

static bool jbool(const JValue& cfg, const char* key, bool dflt) {
  const JValue* v = cfg.get(key);
  return v ? v->truthy() : dflt;
}
static int jint(const JValue& cfg, const char* key, int dflt) {
  const JValue* v = cfg.get(key);
  return (v && v->type == JValue::Num) ? static_cast<int>(v->num) : dflt;
}
static std::string jstr(const JValue& cfg, const char* key, const char* dflt) {
  const JValue* v = cfg.get(key);
  return (v && v->type == JValue::Str) ? v->str : std::string(dflt);
}

// ---- layer plan -------------------------------------------------------------------------------
static Layer make_layer(const std::string& scope, const std::string& name, bool conv, bool has_bn, int act, int taps,
                        int cin, int cout, int in_buf, bool nested_scope) {
  Layer l;
  l.name = name;
  // tdnn(): variables live at tdnn/<name>_{conv|dense|bn|relu}; common.dense_*(): at <scope>/<name>/<name>_{dense|bn|relu}
  const std::string base = nested_scope ? scope + "/" + name + "/" + name : scope + "/" + name;
  const std::string aff = conv ? "_conv" : "_dense";
  l.kernel_var = base + aff + "/kernel";
  l.bias_var = base + aff + "/bias";
  l.bn_scope = base + "_bn";
  l.alpha_var = base + "_relu/alpha";
  l.has_bn = has_bn;
  l.act = act;
  l.taps = taps;
  l.dil = 1;
  l.cin = cin;
  l.cout = cout;
  l.in_buf = in_buf;
  l.ep[kStageDense] = name + aff;
  l.ep[kStageBn] = has_bn ? name + "_bn" : "";
  l.ep[kStageAct] = act == kActRelu ? name + "_relu" : (act == kActTanh ? name + "_tanh" : "");
  return l;
}

static int last_stage(const Layer& l) {
  if (l.act != kActNone) return kStageAct;
  return l.has_bn ? kStageBn : kStageDense;
}

static int build_side_net(xv_ctx* c, const char* prefix, const JValue* nodes, int last_type, int in_buf, int in_dim,
                          std::vector<int>* out_layers, int* out_dim) {
  const std::string att_scope = c->scope + "/attention";
  *out_dim = in_dim;
  if (!nodes || nodes->type != JValue::Arr) return XV_OK;
  const int n = static_cast<int>(nodes->arr.size());
  for (int i = 0; i < n; ++i) {
    const int width = static_cast<int>(nodes->arr[i].num);
    const bool last = i == n - 1;
    bool has_bn = true;
    int act = kActRelu;
    if (last) {
      // model/pooling.py:104-117: 0 affine, 1 affine+relu, 2 affine+bn+relu, 3 affine+tanh
      if (last_type == 0) { has_bn = false; act = kActNone; }
      else if (last_type == 1) { has_bn = false; act = kActRelu; }
      else if (last_type == 2) { has_bn = true; act = kActRelu; }
      else if (last_type == 3) { has_bn = false; act = kActTanh; }
      else return fail(c, XV_EINVAL, "%s network type %d is not one of 0..3", prefix, last_type);
    }
    char name[64];
    snprintf(name, sizeof(name), "%s%d", prefix, i);
    Layer l = make_layer(att_scope, name, false, has_bn, act, 1, *out_dim, width, in_buf, true);
    c->frame_layers.push_back(l);
    in_buf = static_cast<int>(c->frame_layers.size()) - 1;
    out_layers->push_back(in_buf);
    *out_dim = width;
  }
  return XV_OK;
}

static int build_plan(xv_ctx* c) {
  const JValue& cfg = c->cfg;
  const std::string net = jstr(cfg, "network_type", "");
  // model/trainer.py:101-110: tdnn | tdnn-s (dead: quit() at model/tdnn.py:201-202) | extended_tdnn | resnet_18
  if (net != "tdnn" && net != "extended_tdnn" && net != "resnet_18")
    return fail(c, XV_EUNSUPPORTED, "Not implement %s network", net.c_str());   // model/trainer.py:109-110
  c->resnet = net == "resnet_18";
  c->scope = net == "tdnn" ? "tdnn" : (net == "extended_tdnn" ? "etdnn" : "resnet_18");
  c->relu_type = cfg.has("network_relu_type") ? jstr(cfg, "network_relu_type", "") : "";
  if (c->relu_type != "prelu" && c->relu_type != "lrelu") c->relu_type = "";
  c->pooling_type = jstr(cfg, "pooling_type", "");
  c->embedding_node = jstr(cfg, "embedding_node", net == "extended_tdnn" ? "tdnn12_dense" : "tdnn6_dense");
  c->pool_nodes = jint(cfg, "num_nodes_pooling_layer", 1500);      // model/tdnn.py:114-116
  c->last_nodes = jint(cfg, "num_nodes_last_layer", 512);          // model/tdnn.py:152-154
  c->last_layer_no_bn = jbool(cfg, "last_layer_no_bn", false);     // :163-164
  c->last_layer_linear = jbool(cfg, "last_layer_linear", false);   // :173-174
  c->feature_norm = jbool(cfg, "feature_norm", false);             // model/trainer.py:400-403
  if (c->feature_norm) {
    const JValue* f = cfg.get("feature_scaling_factor");
    if (!f || f->type != JValue::Num)
      return fail(c, XV_EINVAL, "If feature normalization is applied, scaling factor is necessary.");
    c->feature_scaling_factor = static_cast<float>(f->num);
  }
  const int hidden = 512;
  std::vector<int> ctx_of;
  const char* seg_names[2] = {"tdnn6", "tdnn7"};
  if (c->resnet) {
    // model/resnet.py:152-351.  The 2-D convolution stack (conv0_1 .. conv4b) is planned in xv_resnet.cu; its
    // output [B, T', F', 512] viewed as rows of F'*512 values feeds conv5 (a (1,F') VALID conv = one GEMM),
    // dense1 and dense2, which run on the same spliced-GEMM kernel as the TDNN frame layers.
    c->resnet_pool_nodes = 1500;     // model/resnet.py:281-282 hard-codes 1500 (num_nodes_pooling_layer is not read)
    c->pool_nodes = c->resnet_pool_nodes;
    int rc = resnet_build_plan(c);
    if (rc != XV_OK) return rc;
    const int fq = c->resnet_f_out;
    auto named = [&](const char* name, int cin, int cout, int in_buf) {
      Layer l = make_layer("resnet_18", name, false, true, kActRelu, 1, cin, cout, in_buf, false);
      l.kernel_var = std::string("resnet_18/") + name + "/kernel";
      l.bias_var = std::string("resnet_18/") + name + "/bias";
      l.ep[0] = l.ep[1] = l.ep[2] = "";      // resnet_18 publishes no frame-level endpoints (model/resnet.py:303-349)
      return l;
    };
    c->frame_layers.push_back(named("conv5", fq * 512, hidden, -1));
    c->frame_layers.push_back(named("dense1", hidden, hidden, 0));
    c->frame_layers.push_back(named("dense2", hidden, c->pool_nodes, 1));
    c->n_main_frame = 3;
    c->total_context = 0;
    ctx_of.assign(3, 0);
  } else {
    // (name, conv kernel width | 0 = dense) per frame layer: model/tdnn.py:38-130 and :376-518
    struct Row { const char* name; int taps; };
    static const Row tdnn_rows[] = {{"tdnn1", 5}, {"tdnn2", 5}, {"tdnn3", 7}, {"tdnn4", 0}, {"tdnn5", 0}};
    static const Row etdnn_rows[] = {{"tdnn1", 5}, {"tdnn2", 0}, {"tdnn3", 5}, {"tdnn4", 0}, {"tdnn5", 7},
                                     {"tdnn6", 0}, {"tdnn7", 9}, {"tdnn8", 0}, {"tdnn9", 0}, {"tdnn10", 0}};
    const bool ext = net == "extended_tdnn";
    const Row* rows = ext ? etdnn_rows : tdnn_rows;
    const int n_rows = ext ? 10 : 5;
    if (ext) { seg_names[0] = "tdnn12"; seg_names[1] = "tdnn13"; }
    int ctx_sum = 0;
    int cin = c->dim;
    for (int i = 0; i < n_rows; ++i) {
      const bool conv = rows[i].taps > 0;
      const int cout = i == n_rows - 1 ? c->pool_nodes : hidden;
      c->frame_layers.push_back(make_layer(c->scope, rows[i].name, conv, true, kActRelu, conv ? rows[i].taps : 1, cin,
                                           cout, i - 1, false));
      cin = cout;
      ctx_sum += c->frame_layers[i].taps - 1;
      ctx_of.push_back(ctx_sum);
    }
    c->n_main_frame = n_rows;
    c->total_context = ctx_sum;
  }
  for (int i = 0; i < c->n_main_frame; ++i) {
    const Layer& l = c->frame_layers[i];
    for (int s = 0; s < 3; ++s) {
      if (l.ep[s].empty()) continue;
      NodeInfo ni;
      ni.level = 0; ni.layer = i; ni.stage = s; ni.width = l.cout; ni.context = ctx_of[i];
      c->nodes[l.ep[s]] = ni;
    }
  }
  // ---- pooling ----
  if (c->pooling_type == "statistics_pooling") {
    c->pooled_dim = 2 * c->pool_nodes;
  } else if (c->pooling_type == "self_attention") {
    const char* need[] = {"att_key_input", "att_key_num_nodes", "att_key_network_type", "att_value_input",
                          "att_value_num_nodes", "att_value_network_type", "att_apply_nonlinear", "att_use_scale",
                          "att_num_heads", "att_split_value", "att_split_key"};
    for (const char* k : need)
      if (!cfg.has(k)) return fail(c, XV_EINVAL, "self_attention pooling needs config key '%s'", k);
    c->att_heads = jint(cfg, "att_num_heads", 1);
    c->att_split_value = jbool(cfg, "att_split_value", true);
    c->att_split_key = jbool(cfg, "att_split_key", true);
    c->att_use_scale = jbool(cfg, "att_use_scale", true);
    c->att_apply_nonlinear = jbool(cfg, "att_apply_nonlinear", false);
    auto find_input = [&](const char* key, int* buf, int* dim) -> int {
      const std::string ep = jstr(cfg, key, "");
      auto it = c->nodes.find(ep);
      if (it == c->nodes.end()) return fail(c, XV_ENOTFOUND, "%s = '%s' is not an endpoint of the network", key, ep.c_str());
      const NodeInfo& ni = it->second;
      if (ni.level != 0 || ni.stage != last_stage(c->frame_layers[ni.layer]) || ni.context != c->total_context)
        return fail(c, XV_EUNSUPPORTED,
                    "%s = '%s': only the final activation of a full-context frame layer (tdnn3/4/5_relu) is supported",
                    key, ep.c_str());
      *buf = ni.layer;
      *dim = ni.width;
      return XV_OK;
    };
    int kin = 0, vin = 0, rc;
    if ((rc = find_input("att_key_input", &c->att_key_buf, &kin)) != XV_OK) return rc;
    if ((rc = find_input("att_value_input", &c->att_value_buf, &vin)) != XV_OK) return rc;
    rc = build_side_net(c, "att_key", cfg.get("att_key_num_nodes"), jint(cfg, "att_key_network_type", 0),
                        c->att_key_buf, kin, &c->key_layers, &c->att_key_dim);
    if (rc != XV_OK) return rc;
    if (c->key_layers.empty())
      return fail(c, XV_EINVAL, "att_key_num_nodes must not be empty (model/pooling.py:104-117 always builds the last key layer)");
    rc = build_side_net(c, "att_value", cfg.get("att_value_num_nodes"), jint(cfg, "att_value_network_type", 0),
                        c->att_value_buf, vin, &c->value_layers, &c->att_value_dim);
    if (rc != XV_OK) return rc;
    const int h = c->att_heads;
    if (h < 1) return fail(c, XV_EINVAL, "att_num_heads must be >= 1");
    if (c->att_split_value && c->att_value_dim % h != 0)
      return fail(c, XV_EINVAL, "The dim of the value must be divided by the num of heads.");   // model/pooling.py:142
    if (c->att_split_key && c->att_key_dim % h != 0)
      return fail(c, XV_EINVAL, "The dim of the key must be divided by the num of heads.");
    c->att_dk = c->att_split_key ? c->att_key_dim / h : c->att_key_dim;
    c->att_dv = c->att_split_value ? c->att_value_dim / h : c->att_value_dim;
    c->pooled_dim = 2 * h * c->att_dv;
    for (size_t i = c->n_main_frame; i < c->frame_layers.size(); ++i) {
      const Layer& l = c->frame_layers[i];
      for (int s = 0; s < 3; ++s) {
        if (l.ep[s].empty()) continue;
        NodeInfo ni;
        ni.level = 0; ni.layer = static_cast<int>(i); ni.stage = s; ni.width = l.cout; ni.context = c->total_context;
        c->nodes[l.ep[s]] = ni;
      }
    }
    NodeInfo w;
    w.level = 2; w.width = h; w.context = c->total_context;
    c->nodes["attention_weights"] = w;
    NodeInfo b;
    b.level = 1; b.kind = 2; b.width = c->pooled_dim;
    c->nodes["att_output_before_nonlinear"] = b;
    if (c->att_apply_nonlinear) {
      c->has_att_post = true;
      b.kind = 3; c->nodes["att_post_bn"] = b;
      b.kind = 4; c->nodes["att_post_relu"] = b;
    }
  } else {
    return fail(c, XV_EUNSUPPORTED, "Not implement %s pooling", c->pooling_type.c_str());   // model/pooling.py:22-23
  }
  NodeInfo p;
  p.level = 1; p.kind = 1; p.width = c->pooled_dim;
  c->nodes["pooling"] = p;
  // ---- segment level ----
  c->seg_layers.push_back(make_layer(c->scope, seg_names[0], false, true, kActRelu, 1, c->pooled_dim, 512, -1, false));
  c->seg_layers.push_back(make_layer(c->scope, seg_names[1], false, !c->last_layer_no_bn,
                                     c->last_layer_linear ? kActNone : kActRelu, 1, 512, c->last_nodes, -1, false));
  for (int i = 0; i < 2; ++i) {
    const Layer& l = c->seg_layers[i];
    for (int s = 0; s < 3; ++s) {
      if (l.ep[s].empty()) continue;
      NodeInfo ni;
      ni.level = 1; ni.layer = i; ni.stage = s; ni.width = l.cout;
      c->nodes[l.ep[s]] = ni;
    }
  }
  NodeInfo o;
  o.level = 1; o.kind = 5; o.layer = 1; o.stage = last_stage(c->seg_layers[1]); o.width = c->last_nodes;
  c->nodes["output"] = o;
  // bf16 operand geometry
  for (size_t i = 0; i < c->frame_layers.size(); ++i) {
    Layer& l = c->frame_layers[i];
    l.n_alloc = round_up(l.cout, kBN);
    if (l.in_buf < 0 && !c->resnet && l.taps > 1) {
      // first TDNN layer: the features are cast to bf16 rows padded to a multiple of 64 channels and the context is
      // spliced by the TMA row coordinate like every other layer (K = taps * 64 for the 23..64-dim front ends); the
      // im2col rows are never materialised
      l.a_cin = round_up(l.cin, kBK);
      l.a_taps = l.taps;
      l.kpad = l.a_taps * l.a_cin;
    } else if (l.in_buf < 0) {        // reads rows prepared by the caller (ResNet tail: conv5 over flattened F' x C)
      l.kpad = round_up(l.taps * l.cin, kBK);
      l.a_cin = l.kpad;
      l.a_taps = 1;
    } else {
      l.a_cin = round_up(c->frame_layers[l.in_buf].cout, kBN);
      l.a_taps = l.taps;
      l.kpad = l.a_taps * l.a_cin;
    }
  }
  for (Layer& l : c->seg_layers) {
    l.n_alloc = round_up(l.cout, kBN);
    l.kpad = round_up(l.cin, kBK);
    l.a_cin = l.kpad;
    l.a_taps = 1;
  }
  c->frame_out.resize(c->frame_layers.size());
  return XV_OK;
}

// ---- weight folding ---------------------------------------------------------------------------
const HostTensor* find_weight(xv_ctx* c, const std::string& name) {
  auto it = c->weights.find(name);
  return it == c->weights.end() ? nullptr : &it->second;
}

int upload(xv_ctx* c, DevBuf* b, const void* host, size_t bytes) {
  int rc = ensure(c, b, bytes);
  if (rc != XV_OK) return rc;
  cudaError_t e = cudaMemcpy(b->p, host, bytes, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaMemcpy(weights)");
  return XV_OK;
}

static int need_vec(xv_ctx* c, const std::string& name, int n, const HostTensor** out) {
  const HostTensor* t = find_weight(c, name);
  if (!t) return fail(c, XV_ENOTFOUND, "missing variable '%s'", name.c_str());
  if (static_cast<int>(t->data.size()) != n)
    return fail(c, XV_EINVAL, "variable '%s' has %zu elements, expected %d", name.c_str(), t->data.size(), n);
  *out = t;
  return XV_OK;
}

// s = gamma / sqrt(var + eps), h = beta - s * mean   (inference BN, eps = 1e-3)
int fold_bn(xv_ctx* c, const std::string& scope, int n, std::vector<double>* s, std::vector<double>* h) {
  const HostTensor *g, *b, *m, *v;
  int rc;
  if ((rc = need_vec(c, scope + "/gamma", n, &g)) != XV_OK) return rc;
  if ((rc = need_vec(c, scope + "/beta", n, &b)) != XV_OK) return rc;
  if ((rc = need_vec(c, scope + "/moving_mean", n, &m)) != XV_OK) return rc;
  if ((rc = need_vec(c, scope + "/moving_variance", n, &v)) != XV_OK) return rc;
  s->resize(n);
  h->resize(n);
  for (int i = 0; i < n; ++i) {
    (*s)[i] = static_cast<double>(g->data[i]) / sqrt(static_cast<double>(v->data[i]) + static_cast<double>(kBnEpsilon));
    (*h)[i] = static_cast<double>(b->data[i]) - (*s)[i] * static_cast<double>(m->data[i]);
  }
  return XV_OK;
}

static int act_slopes(xv_ctx* c, const Layer& l, int n, const std::string& alpha_var, std::vector<float>* slope) {
  slope->assign(n, 1.f);
  if (l.act != kActRelu) return XV_OK;
  if (c->relu_type == "prelu") {           // model/common.py:27-42
    const HostTensor* a;
    int rc = need_vec(c, alpha_var, n, &a);
    if (rc != XV_OK) return rc;
    for (int i = 0; i < n; ++i) (*slope)[i] = a->data[i];
  } else if (c->relu_type == "lrelu") {
    slope->assign(n, kLreluAlpha);
  } else {
    slope->assign(n, 0.f);
  }
  return XV_OK;
}

static int finalize_layer(xv_ctx* c, Layer& l) {
  const HostTensor* k = find_weight(c, l.kernel_var);
  if (!k) return fail(c, XV_ENOTFOUND, "missing variable '%s'", l.kernel_var.c_str());
  const int K = l.taps * l.cin;
  bool shape_ok;
  if (k->shape.size() == 4)        // conv2d (1,k): [1,k,Cin,Cout]; resnet conv5 (1,F') over [F',512] rows: [1,F',512,Cout]
    shape_ok = k->shape[0] == 1 && k->shape[1] * k->shape[2] == static_cast<int64_t>(l.taps) * l.cin && k->shape[3] == l.cout &&
               (k->shape[1] == l.taps || l.taps == 1);
  else if (k->shape.size() == 3)   // conv1d: [k,Cin,Cout] (model/tdnn.py:378-383)
    shape_ok = k->shape[0] == l.taps && k->shape[1] == l.cin && k->shape[2] == l.cout;
  else
    shape_ok = k->shape.size() == 2 && l.taps == 1 && k->shape[0] == l.cin && k->shape[1] == l.cout;
  if (!shape_ok)
    return fail(c, XV_EINVAL, "variable '%s' has the wrong shape for a %d-tap %d->%d layer", l.kernel_var.c_str(),
                l.taps, l.cin, l.cout);
  const HostTensor* b;
  int rc;
  if ((rc = need_vec(c, l.bias_var, l.cout, &b)) != XV_OK) return rc;
  std::vector<double> s(l.cout, 1.0), h(l.cout, 0.0);
  if (l.has_bn && (rc = fold_bn(c, l.bn_scope, l.cout, &s, &h)) != XV_OK) return rc;
  std::vector<float> slope;
  if ((rc = act_slopes(c, l, l.cout, l.alpha_var, &slope)) != XV_OK) return rc;
  l.plain_relu = l.act == kActRelu;
  for (float s0 : slope) l.plain_relu = l.plain_relu && s0 == 0.f;

  // validation-path parameters (fp32, unpadded)
  if ((rc = upload(c, &l.w32, k->data.data(), sizeof(float) * K * l.cout)) != XV_OK) return rc;
  if ((rc = upload(c, &l.bias, b->data.data(), sizeof(float) * l.cout)) != XV_OK) return rc;
  std::vector<float> sf(l.cout), hf(l.cout);
  for (int i = 0; i < l.cout; ++i) { sf[i] = static_cast<float>(s[i]); hf[i] = static_cast<float>(h[i]); }
  if ((rc = upload(c, &l.bn_scale, sf.data(), sizeof(float) * l.cout)) != XV_OK) return rc;
  if ((rc = upload(c, &l.bn_shift, hf.data(), sizeof(float) * l.cout)) != XV_OK) return rc;
  if ((rc = upload(c, &l.slope_raw, slope.data(), sizeof(float) * l.cout)) != XV_OK) return rc;

  // tcgen05-path parameters: per-stage folded epilogue + bf16 K-major weights
  for (int st = 0; st < 3; ++st) {
    std::vector<float> fs(l.n_alloc, 0.f), fh(l.n_alloc, 0.f), fl(l.n_alloc, 1.f);
    for (int i = 0; i < l.cout; ++i) {
      const bool bn = l.has_bn && st >= kStageBn;
      fs[i] = bn ? static_cast<float>(s[i]) : 1.f;
      fh[i] = bn ? static_cast<float>(s[i] * static_cast<double>(b->data[i]) + h[i]) : b->data[i];
      fl[i] = st == kStageAct ? slope[i] : 1.f;
    }
    if ((rc = upload(c, &l.f_scale[st], fs.data(), sizeof(float) * l.n_alloc)) != XV_OK) return rc;
    if ((rc = upload(c, &l.f_shift[st], fh.data(), sizeof(float) * l.n_alloc)) != XV_OK) return rc;
    if ((rc = upload(c, &l.f_slope[st], fl.data(), sizeof(float) * l.n_alloc)) != XV_OK) return rc;
  }
  if (c->bf16) {
    std::vector<uint16_t> wt(static_cast<size_t>(l.n_alloc) * l.kpad, 0);
    for (int j = 0; j < l.taps; ++j)
      for (int ci = 0; ci < l.cin; ++ci) {
        const int kk = l.a_taps == 1 ? j * l.cin + ci : j * l.a_cin + ci;
        const float* src = k->data.data() + (static_cast<size_t>(j) * l.cin + ci) * l.cout;
        for (int n = 0; n < l.cout; ++n) wt[static_cast<size_t>(n) * l.kpad + kk] = f32_to_bf16_rne(src[n]);
      }
    if ((rc = upload(c, &l.wt, wt.data(), wt.size() * 2)) != XV_OK) return rc;
    const char* e = make_tmap_bf16_2d(&l.tm_b, l.wt.p, l.kpad, l.n_alloc, static_cast<uint64_t>(l.kpad) * 2, kBK, kBLoadRows);
    if (e) return fail(c, XV_ECUDA, "%s", e);
  }
  return XV_OK;
}

}  // namespace xv

using namespace xv;

extern "C" {

const char* xv_version(void) { return "xvector-b200 0.1 (sm_100a)"; }

const char* xv_last_error(xv_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int xv_create(int device, const char* config_json, xv_ctx** out) {
  if (!out || !config_json) return fail(nullptr, XV_EINVAL, "xv_create: null argument");
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(nullptr, XV_ECUDA, "no CUDA device available (%s); libxvector has no CPU fallback",
                cudaGetErrorString(e));
  if (device < 0 || device >= n_dev) return fail(nullptr, XV_EINVAL, "device %d out of range [0,%d)", device, n_dev);
  if ((e = cudaSetDevice(device)) != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
  if (prop.major != 10)
    return fail(nullptr, XV_EUNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
                prop.major, prop.minor);
  xv_ctx* c = new xv_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  std::string perr;
  JParser parser(config_json);
  if (!parser.parse(&c->cfg, &perr) || c->cfg.type != JValue::Obj) {
    fail(nullptr, XV_EINVAL, "config JSON: %s", perr.empty() ? "top level must be an object" : perr.c_str());
    delete c;
    return XV_EINVAL;
  }
  c->dim = jint(c->cfg, "feature_dim", 0);
  if (c->dim <= 0) {
    fail(nullptr, XV_EINVAL, "config needs a positive \"feature_dim\"");
    delete c;
    return XV_EINVAL;
  }
  const std::string prec = jstr(c->cfg, "precision", "bf16");
  if (prec != "bf16" && prec != "fp32") {
    fail(nullptr, XV_EINVAL, "precision must be \"bf16\" or \"fp32\", got \"%s\"", prec.c_str());
    delete c;
    return XV_EINVAL;
  }
  c->bf16 = prec == "bf16";
  c->batch_frames = jint(c->cfg, "batch_frames", 262144);
  if (c->batch_frames < 1024) c->batch_frames = 1024;
  int rc = build_plan(c);
  if (rc != XV_OK) {
    g_create_err = c->err;
    delete c;
    return rc;
  }
  if (c->resnet && !c->cfg.has("batch_frames")) c->batch_frames = 16384;   // keeps a stage's activations L2-sized
  const char* ge = gemm_init();
  if (!ge && c->resnet) ge = conv_gemm_init();
  if (!ge && c->resnet) ge = conv_halo_init();
  if (ge) {
    fail(nullptr, XV_ECUDA, "%s", ge);
    delete c;
    return XV_ECUDA;
  }
  bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; ok && i < kNumEvents; ++i) ok = cudaEventCreate(&c->events[i]) == cudaSuccess;
  for (int i = 0; ok && i < 2; ++i)
    ok = cudaEventCreateWithFlags(&c->h2d_done[i], cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->buf_free[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    fail(nullptr, XV_ECUDA, "stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
    xv_destroy(c);
    return XV_ECUDA;
  }
  *out = c;
  return XV_OK;
}

// Parses a config exactly as xv_create does and builds the layer plan on the host only (no device is touched): the
// check a caller can run anywhere (tests/test_host.py feeds it the reference's 42 nnet_conf JSONs).
int xv_config_check(const char* config_json, char* buf, size_t buf_size) {
  if (!config_json) return fail(nullptr, XV_EINVAL, "xv_config_check: null argument");
  xv_ctx* c = new xv_ctx();
  std::string perr;
  JParser parser(config_json);
  if (!parser.parse(&c->cfg, &perr) || c->cfg.type != JValue::Obj) {
    fail(nullptr, XV_EINVAL, "config JSON: %s", perr.empty() ? "top level must be an object" : perr.c_str());
    delete c;
    return XV_EINVAL;
  }
  c->dim = jint(c->cfg, "feature_dim", 0);
  if (c->dim <= 0) {
    fail(nullptr, XV_EINVAL, "config needs a positive \"feature_dim\"");
    delete c;
    return XV_EINVAL;
  }
  const int rc = build_plan(c);
  if (rc != XV_OK) {
    g_create_err = c->err;
    delete c;
    return rc;
  }
  if (buf && buf_size > 0)
    snprintf(buf, buf_size, "scope=%s;pooling=%s;frame_layers=%zu;seg_layers=%zu;convs=%zu;nodes=%zu;context=%d;pooled_dim=%d;embedding_node=%s",
             c->scope.c_str(), c->pooling_type.c_str(), c->frame_layers.size(), c->seg_layers.size(), c->convs.size(),
             c->nodes.size(), c->total_context, c->pooled_dim, c->embedding_node.c_str());
  delete c;
  return XV_OK;
}

void xv_destroy(xv_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
  auto rel_layer = [](Layer& l) {
    release(&l.wt); release(&l.w32); release(&l.bias); release(&l.bn_scale); release(&l.bn_shift);
    release(&l.slope_raw);
    for (int s = 0; s < 3; ++s) { release(&l.f_scale[s]); release(&l.f_shift[s]); release(&l.f_slope[s]); }
  };
  for (Layer& l : c->frame_layers) rel_layer(l);
  for (Layer& l : c->seg_layers) rel_layer(l);
  train_release(c);
  resnet_release(c);
  for (DevBuf& b : c->frame_out) release(&b);
  DevBuf* bufs[] = {&c->att_query, &c->att_post_scale, &c->att_post_shift, &c->att_post_slope, &c->d_feats,
                    &c->d_feat_ring[0], &c->d_feat_ring[1], &c->a1, &c->scores, &c->scores_acc, &c->att_query_pad, &c->pool_acc, &c->pool_ws,
                    &c->pooled_f32, &c->pooled_bf16, &c->pooled_pre_f32, &c->seg_out[0], &c->seg_out[1],
                    &c->seg_out_bf16, &c->d_cu, &c->d_segs, &c->d_tile_seg_ptr, &c->d_gather_meta, &c->d_gather_out,
                    &c->flush_buf, &c->d_view_src, &c->d_feat_gather};
  for (DevBuf* b : bufs) release(b);
  for (int i = 0; i < 2; ++i)
    if (c->h_pinned[i]) cudaFreeHost(c->h_pinned[i]);
  if (c->h_tables) cudaFreeHost(c->h_tables);
  if (c->tables_ev) cudaEventDestroy(c->tables_ev);
  for (auto& s : c->prof) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  for (int i = 0; i < kNumEvents; ++i)
    if (c->events[i]) cudaEventDestroy(c->events[i]);
  for (int i = 0; i < 2; ++i) {
    if (c->h2d_done[i]) cudaEventDestroy(c->h2d_done[i]);
    if (c->buf_free[i]) cudaEventDestroy(c->buf_free[i]);
  }
  if (c->stream) cudaStreamDestroy(c->stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  delete c;
}

int xv_set_weight(xv_ctx* c, const char* tf_name, const float* host, const int64_t* shape, int rank) {
  if (!c) return XV_EINVAL;
  if (!tf_name || !host || !shape || rank < 1 || rank > 4) return fail(c, XV_EINVAL, "xv_set_weight: bad argument");
  HostTensor t;
  size_t n = 1;
  for (int i = 0; i < rank; ++i) {
    if (shape[i] <= 0) return fail(c, XV_EINVAL, "xv_set_weight('%s'): non-positive dimension", tf_name);
    t.shape.push_back(shape[i]);
    n *= static_cast<size_t>(shape[i]);
  }
  t.data.assign(host, host + n);
  c->weights[tf_name] = std::move(t);
  c->finalized = false;
  return XV_OK;
}

int xv_finalize_weights(xv_ctx* c) {
  if (!c) return XV_EINVAL;
  cudaSetDevice(c->device);
  int rc;
  if (c->resnet && (rc = resnet_finalize(c)) != XV_OK) return rc;
  for (Layer& l : c->frame_layers)
    if ((rc = finalize_layer(c, l)) != XV_OK) return rc;
  for (Layer& l : c->seg_layers)
    if ((rc = finalize_layer(c, l)) != XV_OK) return rc;
  if (c->pooling_type == "self_attention") {
    const std::string qname = c->scope + "/attention/query";
    const HostTensor* q = find_weight(c, qname);
    if (!q) return fail(c, XV_ENOTFOUND, "missing variable '%s'", qname.c_str());
    if (q->shape.size() != 2 || q->shape[0] != c->att_heads || q->shape[1] != c->att_dk)
      return fail(c, XV_EINVAL, "variable '%s' must be [%d, %d]", qname.c_str(), c->att_heads, c->att_dk);
    if ((rc = upload(c, &c->att_query, q->data.data(), q->data.size() * sizeof(float))) != XV_OK) return rc;
    if (c->has_att_post) {
      std::vector<double> s, h;
      if ((rc = fold_bn(c, c->scope + "/attention/att_post_bn", c->pooled_dim, &s, &h)) != XV_OK) return rc;
      std::vector<float> sf(c->pooled_dim), hf(c->pooled_dim), sl;
      for (int i = 0; i < c->pooled_dim; ++i) { sf[i] = static_cast<float>(s[i]); hf[i] = static_cast<float>(h[i]); }
      Layer fake;
      fake.act = kActRelu;
      if ((rc = act_slopes(c, fake, c->pooled_dim, c->scope + "/attention/att_post_relu/alpha", &sl)) != XV_OK) return rc;
      if ((rc = upload(c, &c->att_post_scale, sf.data(), sf.size() * 4)) != XV_OK) return rc;
      if ((rc = upload(c, &c->att_post_shift, hf.data(), hf.size() * 4)) != XV_OK) return rc;
      if ((rc = upload(c, &c->att_post_slope, sl.data(), sl.size() * 4)) != XV_OK) return rc;
    }
  }
  c->finalized = true;
  return XV_OK;
}

int xv_node_info(xv_ctx* c, const char* node, int32_t* width, int32_t* frame_level, int32_t* context) {
  if (!c) return XV_EINVAL;
  const std::string name = (node && node[0]) ? node : c->embedding_node;
  auto it = c->nodes.find(name);
  if (it == c->nodes.end()) return fail(c, XV_ENOTFOUND, "unknown node '%s'", name.c_str());
  if (width) *width = it->second.width;
  if (frame_level) *frame_level = it->second.level == 1 ? 0 : 1;
  if (context) *context = it->second.context;
  return XV_OK;
}

}  // extern "C"
