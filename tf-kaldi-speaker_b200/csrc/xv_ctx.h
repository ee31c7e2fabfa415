// The xv_ctx object: model plan, weights, device buffers, staged job.  Internal to libxvector.so.
#pragma once
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/xvector.h"
#include "xv_internal.h"
#include "xv_json.h"

namespace xv {

constexpr float kBnEpsilon = 1e-3f;     // tf.layers.batch_normalization default (TF 1.x)
constexpr float kLreluAlpha = 0.2f;     // tf.nn.leaky_relu default
constexpr int kNumEvents = 16;

enum Stage { kStageDense = 0, kStageBn = 1, kStageAct = 2 };
enum ActKind { kActNone = 0, kActRelu = 1, kActTanh = 2 };   // kActRelu = relu / lrelu / prelu per network_relu_type

struct HostTensor {
  std::vector<int64_t> shape;
  std::vector<float> data;
};

// Grow-only device buffer.
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

// One affine layer: y = act(bn(x.W + b)).  Frame level (applied to every packed row, optionally with a spliced
// context of `taps` rows) or segment level (one row per utterance).
struct Layer {
  std::string name;          // "tdnn1" ... "att_key0"
  std::string kernel_var, bias_var, bn_scope, alpha_var;
  bool has_bn = false;
  bool plain_relu = false;   // activation is max(x, 0) on every channel (fast epilogue path)
  int act = kActNone;
  int taps = 1, dil = 1, cin = 0, cout = 0;
  int in_buf = -1;           // index of the frame buffer this layer reads (-1: the features)
  std::string ep[3];         // endpoint name per stage ("" = stage does not exist)
  // device-side parameters
  int kpad = 0;              // padded K of the bf16 operand (multiple of 64)
  int a_cin = 0;             // channel count of the bf16 A buffer this layer reads (taps*a_cin == kpad)
  int a_taps = 1;            // taps used by the bf16 kernel (1 when the input is pre-spliced)
  int n_alloc = 0;           // cout rounded up to a multiple of 256
  DevBuf wt;                 // bf16 [n_alloc, kpad], K-major
  DevBuf w32;                // fp32 [taps*cin, cout]
  DevBuf bias, bn_scale, bn_shift, slope_raw;    // fp32 [cout] (validation path)
  DevBuf f_scale[3], f_shift[3], f_slope[3];     // folded per-stage epilogue params, fp32 [n_alloc]
  CUtensorMap tm_b;
};

struct NodeInfo {
  int level = 0;     // 0 frame level, 1 utterance level, 2 attention weights
  int layer = -1;    // index into frame_layers / seg_layers; -1 for pooling nodes
  int stage = kStageAct;
  int width = 0;
  int context = 0;
  int kind = 0;      // utterance level: 0 layer output, 1 "pooling", 2 "att_output_before_nonlinear", 3 att_post_bn, 4 att_post_relu, 5 "output"
};

// One 2-D convolution of the ResNet-18 front end (model/resnet.py:9-149, 217-251): conv (no bias) -> BN
// [-> + residual] [-> activation], NHWC.  Activations live in "slots" (ActSlot) that rotate over four buffers.
struct ActSlot {
  int buf = 0;            // index into xv_ctx::res_buf
  int f = 0, c = 0;       // frequency bins, channels
  int tshift = 0;         // number of stride-2 time reductions applied so far
  bool pad_even = false;  // allocate an even number of time rows (zero pad row): read by a time-strided conv
};
struct ConvLayer {
  std::string name, kernel_var, bn_scope, alpha_var;
  int kh = 3, kw = 3, st = 1, sf = 1;
  int cin = 0, cout = 0;
  bool act = true;        // activation after (BN + residual)
  int in_slot = -1, out_slot = -1, res_slot = -1;
  bool plain_relu = false;
  int bn_tile = 64;       // UMMA N of the tile: 64 | 128 | 256
  DevBuf wt;              // bf16 [cout][kh*kw*cin], K-major
  DevBuf w32;             // fp32 [kh*kw*cin][cout]
  DevBuf scale, shift, slope;   // fp32 [cout]
  CUtensorMap tm_b;
};

struct Batch {
  int32_t utt0 = 0, utt1 = 0;   // utterances [utt0, utt1)
  int64_t row0 = 0;             // first packed row (rows of the frame-level tail: T' per utterance)
  int64_t rows = 0;
  int64_t in_row0 = 0;          // first packed INPUT frame and number of input frames (differ from row0/rows only
  int64_t in_rows = 0;          // for resnet_18 with time stride)
  int32_t t_in = 0;             // resnet_18: the common utterance length of the batch
  int32_t seg0 = 0;             // first entry of this batch in the tile_seg_ptr / segs tables
  int32_t tile0 = 0;
  int32_t m_tiles = 0;
};

struct TrainState;   // xv_train.cu

struct ProfileSlot {
  cudaEvent_t a, b;
  int family;
};

}  // namespace xv

struct xv_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t events[xv::kNumEvents] = {};
  cudaEvent_t h2d_done[2] = {}, buf_free[2] = {};
  std::string err;

  // ---- configuration (reference nnet_conf JSON + extensions) ----
  xv::JValue cfg;
  int dim = 0;
  bool bf16 = true;
  int64_t batch_frames = 262144;
  bool resnet = false;
  std::string scope = "tdnn";    // variable scope: "tdnn" | "etdnn" | "resnet_18"
  std::string relu_type;         // "", "prelu", "lrelu"
  std::string pooling_type;
  std::string embedding_node;
  int pool_nodes = 1500, last_nodes = 512;
  bool last_layer_no_bn = false, last_layer_linear = false;
  bool feature_norm = false;
  float feature_scaling_factor = 1.f;
  // attention
  int att_heads = 1;
  bool att_split_value = true, att_split_key = true, att_use_scale = true, att_apply_nonlinear = false;
  int att_key_buf = -1, att_value_buf = -1;   // frame buffer indices feeding the key / value nets
  int att_key_dim = 0, att_value_dim = 0, att_dk = 0, att_dv = 0;
  int pooled_dim = 0;

  // ---- model plan ----
  std::vector<xv::Layer> frame_layers;    // tdnn1..5 then att_key*, att_value*
  int n_main_frame = 0;                    // tdnn1..tdnn5
  std::vector<int> key_layers, value_layers;
  std::vector<xv::Layer> seg_layers;      // tdnn6, tdnn7
  std::map<std::string, xv::NodeInfo> nodes;
  int total_context = 14;
  bool has_att_post = false;
  xv::DevBuf att_query, att_post_scale, att_post_shift, att_post_slope;

  // ---- resnet_18 front end (xv_resnet.cu) ----
  bool resnet_time_stride = false, resnet_maxpool = false;
  int resnet_blocks[4] = {2, 2, 2, 2};
  int resnet_pool_nodes = 1500;
  int resnet_f_out = 0;                    // F' = dim / 8
  int resnet_out_slot = -1;                // slot holding the stage-4 output
  int resnet_conv0_slot = -1, resnet_pool_slot = -1;
  std::vector<xv::ActSlot> slots;
  std::vector<xv::ConvLayer> convs;        // conv1a_conv0 ... conv4b_conv1 (conv0_1 is the SIMT stem)
  xv::DevBuf conv0_w, conv0_scale, conv0_shift, conv0_slope;   // stem: fp32 [9][64] + folded BN
  xv::DevBuf res_buf[4];
  std::vector<int32_t> cu_in;              // [n+1] packed input-frame offsets (== cu unless time stride)

  xv::TrainState* train = nullptr;         // training step state (xv_train_init)

  std::map<std::string, xv::HostTensor> weights;
  bool finalized = false;

  // ---- device buffers ----
  xv::DevBuf d_feats;                      // staged features (all utterances) fp32 [rows, dim]
  xv::DevBuf d_feat_ring[2];               // pipelined H2D staging for xv_forward
  xv::DevBuf a1;                           // spliced bf16 [batch rows, k1pad]
  std::vector<xv::DevBuf> frame_out;       // per frame layer output (bf16 or fp32) [batch rows, n_alloc | cout]
  xv::DevBuf scores;                       // attention scores/weights fp32 [H][batch rows]
  xv::DevBuf scores_acc;                   // fused key-layer epilogue: fp64 score accumulators [H][batch rows]
  xv::DevBuf att_query_pad;                // query flattened + zero padded to the key layer's n_alloc
  xv::DevBuf pool_acc;                     // fused pooling accumulators
  xv::DevBuf pool_ws;                      // launch_pool workspace
  xv::DevBuf pooled_f32, pooled_bf16, pooled_pre_f32;
  xv::DevBuf seg_out[2], seg_out_bf16;
  xv::DevBuf d_cu, d_segs, d_tile_seg_ptr;
  xv::DevBuf d_gather_meta, d_gather_out;
  xv::DevBuf flush_buf;
  void* h_pinned[2] = {nullptr, nullptr};
  size_t h_pinned_bytes[2] = {0, 0};

  // ---- staged job ----
  bool staged = false;
  int32_t n_utts = 0;
  int64_t total_rows = 0;
  std::vector<int32_t> lengths;
  std::vector<int32_t> cu;                 // [n+1] packed row offsets
  std::vector<xv::Batch> batches;
  std::vector<xv::PoolSeg> segs;
  std::vector<int32_t> tile_seg_ptr;
  // utterances given as VIEWS into the feature buffer (xv_forward_views: the extractor's overlapping chunks)
  bool views = false;
  bool pipeline_hint = false;              // the features of the job being planned will be streamed from the host
  std::vector<int64_t> view_row0;          // [n] first source row of every view (ascending)
  xv::DevBuf d_view_src;                   // int64 [n]: source row of a view relative to its batch's first source row
  xv::DevBuf d_feat_gather;                // fp32 [batch rows, dim]: the views of a batch packed back to back
  // pinned staging for the per-call tables (no stream synchronisation in plan_job)
  void* h_tables = nullptr;
  size_t h_tables_bytes = 0;
  cudaEvent_t tables_ev = nullptr;
  const float* host_feats = nullptr;       // xv_forward pipelined path
  bool host_pinned = false;          // xv_forward source is pinned host memory
  // result of the last run
  const float* result = nullptr;
  int result_ld = 0, result_width = 0;
  bool ran = false;

  // ---- accounting ----
  int64_t launches = 0;
  std::map<std::string, int64_t> launch_by_family;
  bool profiling = false;
  std::vector<xv::ProfileSlot> prof;
  std::vector<std::string> prof_family_names;
};

namespace xv {
// helpers implemented in xv_api.cu
int fail(xv_ctx* c, int code, const char* fmt, ...);
int cuda_fail(xv_ctx* c, cudaError_t e, const char* what);
int ensure(xv_ctx* c, DevBuf* b, size_t bytes, bool zero = false);
void release(DevBuf* b);
void count_launch(xv_ctx* c, const char* family);
struct ProfScope {
  xv_ctx* c;
  int idx;
  ProfScope(xv_ctx* ctx, const char* family);
  ~ProfScope();
};
// resnet_18 front end (xv_resnet.cu)
int resnet_build_plan(xv_ctx* c);
int resnet_finalize(xv_ctx* c);
int resnet_out_len(const xv_ctx* c, int t);      // frames after the (optional) three time strides
// runs conv0_1 .. conv4b on `nb` utterances of `t` frames each (feats fp32 [nb*t, dim] on the device); *out points at
// the stage-4 activations [nb*T', F'*512] (bf16, or fp32 in validation mode) inside a ctx-owned buffer.
int resnet_run_batch(xv_ctx* c, const float* feats, int nb, int t, const void** out);
void resnet_release(xv_ctx* c);
void train_release(xv_ctx* c);                   // xv_train.cu
// shared with xv_api.cu
const HostTensor* find_weight(xv_ctx* c, const std::string& name);
int upload(xv_ctx* c, DevBuf* b, const void* host, size_t bytes);
int fold_bn(xv_ctx* c, const std::string& scope, int n, std::vector<double>* s, std::vector<double>* h);
uint16_t f32_to_bf16_rne(float f);
// forward pieces (xv_forward.cu)
int plan_job(xv_ctx* c, const int32_t* lengths, int32_t n_utts, const int64_t* view_row0 = nullptr, int64_t view_total_rows = 0);
int run_job(xv_ctx* c, const char* node, bool pipelined_h2d, float* frames_out, const int64_t* frames_off,
            int64_t frames_cap);
}  // namespace xv
