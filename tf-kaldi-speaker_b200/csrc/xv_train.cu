// One data-parallel training step of the TDNN x-vector network (BASELINE.json config 5).
//
// Replaces sess.run(train_op) of Trainer.train / TrainerMGPU (model/trainer.py:590-673, 1133-1248):
//   forward in TRAINING mode (tf.layers.batch_normalization with batch statistics, model/tdnn.py:36-181),
//   statistics pooling (model/pooling.py:40-50), additive_margin_softmax / softmax (model/loss.py:9-48, 201-286),
//   total loss = loss + L2 regularisation of every kernel (model/trainer.py:472-499),
//   hand-derived backward pass, gradient averaging over ranks = ncclAllReduce(sum) * 1/N in place of the
//   reference's per-variable add_n/N (model/trainer.py:1225-1248), SGD / momentum / Adam update (:225-246),
//   BN moving averages taken from rank 0 (= tower 0, :1183-1190).
//
// Layout: the B utterances of T frames are packed back to back, R = B*T rows; layer l keeps the row index and
// its valid rows are (r mod T) < T - context_l.  Invalid rows of every activation / gradient are written as 0,
// so the spliced GEMMs need no masks.  All parameters live in ONE flat fp32 buffer and all gradients in a
// second flat buffer of the same layout (+ a tail carrying the moving statistics and the loss).
//
// Exchange step: the gradient buffer is cut into three buckets in the order the backward pass retires them
// (segment layers + softmax + tail | tdnn5..tdnn3 | tdnn2..tdnn1); each bucket's ncclAllReduce is launched on a side
// stream as soon as its last gradient is written, so all but the last (small) bucket overlap the rest of the backward
// pass.  ncclCommGetAsyncError is polled while the host waits for a step; a failing rank aborts the communicator.
//
// Streams of a step (bf16 path): the compute stream carries the critical chain (forward, loss, data gradients, batch-norm
// backward); a side stream carries everything that only produces parameter gradients (weight-gradient GEMMs of all layers,
// the loss head's weight gradient), the work with slack at the step boundaries (gradient zeros, AM-softmax column norms,
// the recast of the data-gradient weight layouts) and the early update of bucket 0; the communication stream carries the
// bucketed all-reduces (and bucket 0's update behind its all-reduce); the copy stream carries the next batch of a
// pipelined step (xv_train_submit / xv_train_collect).  Events order every buffer reuse across them.
//
// Two precisions: "bf16" (product path: the 14 frame-layer GEMMs on tcgen05 with bf16 operands, fp32 masters) and
// "fp32" (validation path: SIMT FFMA GEMMs); both share every other kernel (xv_train_kernels.cuh).
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <cmath>

#include <algorithm>
#include <chrono>
#include <thread>

#include "xv_ctx.h"
#include "xv_train_kernels.cuh"
#include "xv_train_stream.cuh"

// ================================================================================================ host side
namespace xv {

struct TrLayer {
  std::string name;
  bool frame = true, has_bn = true, act = true, fused_bn = false;
  int taps = 1, cin = 0, cout = 0, context = 0;
  int64_t off_kernel = -1, off_bias = -1, off_gamma = -1, off_beta = -1, off_alpha = -1, off_mm = -1, off_mv = -1;
  DevBuf z, y, mean, rstd;
  // mixed precision (bf16 tensor-core GEMMs, fp32 master weights): frame layers only
  int a_cin = 0, a_taps = 1;       // channels / taps of the bf16 A operand (layer 0 reads pre-spliced features)
  int ktotal = 0;                  // a_taps * a_cin
  int n_alloc = 0, c_alloc = 0;    // cout, a_cin rounded up to the 256-wide N tile
  int ldb = 0;                     // row stride of this layer's bf16 y / dz copies (cout rounded up to 64)
  DevBuf wt_fwd, wt_dgrad, yb;
  CUtensorMap tm_wf, tm_wd;        // weight operands (forward / data gradient)
  CUtensorMap tm_af, tm_adz;       // activation operands of the current (B, T): this layer's input | its dZ
};

struct UidBlob { char bytes[128]; };     // ncclUniqueId is a 128-byte POD passed by value

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, UidBlob, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*CommAbort)(void*) = nullptr;
  int (*CommGetAsyncError)(void*, int*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNumBuckets = 3;
struct TrainState {
  std::vector<TrLayer> layers;           // tdnn1..5 (frame), tdnn6, tdnn7 (segment)
  int n_frame = 5;
  int64_t n_train = 0, n_total = 0;      // trainable floats; + moving statistics; (+1 loss slot in the gradient buffer)
  std::map<std::string, std::pair<int64_t, int64_t>> index;   // TF variable name -> (offset, size)
  std::vector<ParamSeg> segs;
  std::vector<UpdChunk> chunks;
  DevBuf params, grads, accum, d_segs, d_chunks;
  int64_t off_w = -1, off_b = -1;
  int embed = 512, speakers = 0;
  bool amsoftmax = true;
  float margin = 0.f, lam_min = 0.f, lam_base = 0.f, lam_gamma = 0.f, lam_power = 0.f;
  float wd = 0.f, wd_out = 0.f, bn_momentum = 0.99f, opt_momentum = 0.f;
  int optimizer = 0, nesterov = 0;       // 0 sgd, 1 momentum, 2 adam
  int64_t adam_t = 0;                    // Adam steps taken (beta powers; model/trainer.py:241-243)
  float clip_norm = 0.f;                 // > 0: per-variable tf.clip_by_norm (clip_gradient / clip_gradient_norm)
  DevBuf sqnorms;
  std::string relu_type;
  DevBuf const_slope, feats, labels, pooled, logits, gr, coef, rnorm, dbuf[2], scratch_stats;
  // fp64 sums of a step, zeroed by ONE memset: [0..8) scalars (0 loss, 1 regularisation), then per layer
  // [2][sum_ld] forward statistics (sum z, sum z^2) and [3][sum_ld] backward sums
  DevBuf sum_area;
  int sum_ld = 0;
  int cap_b = 0, cap_t = 0;
  bool mixed = false;                    // context precision bf16: frame-layer GEMMs on tcgen05, everything else fp32
  bool fused_stats = true;               // batch statistics from the forward GEMM's epilogue (XV_TRAIN_FUSED_STATS=0: separate pass)
  DevBuf xb0, dzb, ones_pad, zeros_pad;
  RecastJobs recast;                     // forward operand layouts of the frame layers
  RecastJobs recast_d;                   // data-gradient layouts ([Cin][taps*Cout]): first needed by the NEXT step's backward pass
  int chunk_split = 0;                   // first update chunk of the early bucket (segment layers, loss head, tail)
  bool early_update = true;              // XV_TRAIN_EARLY_UPDATE=0: one update launch at the end of the step
  // data parallel
  NcclApi nccl;
  void* comm = nullptr;
  int rank = 0, world = 1;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_ready[kNumBuckets] = {}, ev_done = nullptr;
  int64_t bucket_lo[kNumBuckets] = {}, bucket_hi[kNumBuckets] = {};   // float ranges of the gradient buffer
  bool overlap = true;                   // XV_TRAIN_OVERLAP=0: one allreduce after the backward pass
  // gradients cross NVLink as bf16 (fp32 masters, fp32 moving statistics / loss in the tail): default for the bf16 path,
  // config "grad_allreduce_dtype": "fp32" | "bf16" or XV_TRAIN_COMM=fp32|bf16 override; needs the bucketed exchange
  bool comm_bf16 = false;
  DevBuf gcomm;                          // bf16 [n_train]: the trainable range of the gradient buffer as it travels
  bool poisoned = false;                 // the communicator was aborted after a failed step
  // weight-gradient GEMMs of the bf16 path on a side stream: dW_l needs only dZ_l and the stored forward activations, so it runs
  // beside the critical chain dgrad_l -> BN backward_{l-1} (fills the idle second wave of the dgrad GEMM and the latency-bound
  // row-streaming kernels).  dZ is double-buffered by layer parity; XV_TRAIN_WGRAD_STREAM=0 keeps everything on one stream.
  bool wg_overlap = true, wg_overlap_head = true;
  // XV_TRAIN_TIMELINE=<global step>: that step records an event in front of every launch of the compute stream and prints the
  // time between consecutive marks per kernel family (the in-step critical path, warm caches; side-stream work shows as ~0)
  int64_t timeline_step = -1;
  std::vector<std::pair<const char*, cudaEvent_t>> marks;
  bool marking = false;
  bool bulk_rows = false;                 // XV_TRAIN_BULK=0: per-thread-load versions of the row-streaming kernels (xv_train_stream.cuh)
  cudaStream_t wg_stream = nullptr;
  cudaEvent_t ev_dz[2] = {}, ev_wg[2] = {}, ev_wg_mark = nullptr, ev_fork = nullptr, ev_side_rd[2] = {}, ev_rnorm = nullptr;
  size_t dzb_half = 0;                   // elements between the two dZ buffers inside dzb (0: single buffer)
  // pipelined steps (xv_train_submit / xv_train_collect): the batch of step k+1 is copied on the copy stream into the
  // alternate feature / label buffers while step k computes; every step leaves its loss in a pinned slot behind an event
  DevBuf feats_alt, labels_alt;
  struct PipeSlot {
    cudaEvent_t copied = nullptr, end = nullptr;   // H2D of the batch done | last command of the step (after the D2H of the loss)
    float* h_loss = nullptr;                       // pinned: [0] tower-loss sum as float, then two doubles at +2 floats
    bool used = false;
  } pipe[2];
  int pipe_next = 0;                     // slot of the next submit
  int pipe_inflight = 0;                 // submitted and not yet collected (0..2); the oldest is slot (pipe_next - inflight) & 1

  double* scalars() const { return static_cast<double*>(sum_area.p); }
  double* fwd_sums(int layer) const { return static_cast<double*>(sum_area.p) + 8 + static_cast<size_t>(layer) * 5 * sum_ld; }
  double* bwd_sums(int layer) const { return fwd_sums(layer) + 2 * static_cast<size_t>(sum_ld); }
  size_t sum_area_doubles() const { return 8 + layers.size() * 5 * static_cast<size_t>(sum_ld); }
};

namespace {

int tr_fail_nccl(xv_ctx* c, TrainState* t, int rc, const char* what) {
  return fail(c, XV_ECUDA, "%s: NCCL error %d (%s)", what, rc, t->nccl.GetErrorString ? t->nccl.GetErrorString(rc) : "?");
}

int load_nccl(xv_ctx* c, NcclApi* api) {
  if (api->lib) return XV_OK;
  // Prefer a libnccl that is already mapped into the process (torch's), then the system one.
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(c, XV_EUNSUPPORTED, "cannot load libnccl.so.2: %s", dlerror());
  api->lib = h;
  api->GetUniqueId = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclGetUniqueId"));
  api->CommInitRank = reinterpret_cast<int (*)(void**, int, UidBlob, int)>(dlsym(h, "ncclCommInitRank"));
  api->AllReduce = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(dlsym(h, "ncclAllReduce"));
  api->CommDestroy = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommDestroy"));
  api->CommAbort = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommAbort"));
  api->CommGetAsyncError = reinterpret_cast<int (*)(void*, int*)>(dlsym(h, "ncclCommGetAsyncError"));
  api->GetErrorString = reinterpret_cast<const char* (*)(int)>(dlsym(h, "ncclGetErrorString"));
  if (!api->GetUniqueId || !api->CommInitRank || !api->AllReduce || !api->CommDestroy)
    return fail(c, XV_EUNSUPPORTED, "libnccl.so.2 lacks a required symbol");
  return XV_OK;
}

double jnum(const JValue& cfg, const char* key, double dflt) {
  const JValue* v = cfg.get(key);
  if (!v) return dflt;
  if (v->type == JValue::Num) return v->num;
  if (v->type == JValue::Str) return atof(v->str.c_str());
  return dflt;
}

int64_t add_var(xv_ctx* c, TrainState* t, std::vector<float>* host, const std::string& name, int64_t expect, float wd,
                bool moving, int* rc) {
  const HostTensor* h = find_weight(c, name);
  if (!h) { *rc = fail(c, XV_ENOTFOUND, "missing variable '%s'", name.c_str()); return -1; }
  if (static_cast<int64_t>(h->data.size()) != expect) {
    *rc = fail(c, XV_EINVAL, "variable '%s' has %zu elements, expected %lld", name.c_str(), h->data.size(), (long long)expect);
    return -1;
  }
  const int64_t off = static_cast<int64_t>(host->size());
  host->insert(host->end(), h->data.begin(), h->data.end());
  t->index[name] = std::make_pair(off, expect);
  ParamSeg s;
  s.off = off; s.size = expect; s.wd = wd; s.moving = moving ? 1 : 0;
  t->segs.push_back(s);
  return off;
}

}  // namespace

void train_release(xv_ctx* c) {
  TrainState* t = c->train;
  if (!t) return;
  if (t->comm && t->nccl.CommDestroy) t->nccl.CommDestroy(t->comm);
  if (t->comm_stream) cudaStreamDestroy(t->comm_stream);
  for (int i = 0; i < kNumBuckets; ++i)
    if (t->ev_ready[i]) cudaEventDestroy(t->ev_ready[i]);
  if (t->ev_done) cudaEventDestroy(t->ev_done);
  if (t->wg_stream) cudaStreamDestroy(t->wg_stream);
  for (int i = 0; i < 2; ++i) {
    if (t->ev_dz[i]) cudaEventDestroy(t->ev_dz[i]);
    if (t->ev_wg[i]) cudaEventDestroy(t->ev_wg[i]);
  }
  if (t->ev_wg_mark) cudaEventDestroy(t->ev_wg_mark);
  for (int i = 0; i < 2; ++i) {
    if (t->pipe[i].copied) cudaEventDestroy(t->pipe[i].copied);
    if (t->pipe[i].end) cudaEventDestroy(t->pipe[i].end);
    if (t->pipe[i].h_loss) cudaFreeHost(t->pipe[i].h_loss);
  }
  if (t->ev_fork) cudaEventDestroy(t->ev_fork);
  if (t->ev_rnorm) cudaEventDestroy(t->ev_rnorm);
  for (int i = 0; i < 2; ++i)
    if (t->ev_side_rd[i]) cudaEventDestroy(t->ev_side_rd[i]);
  for (TrLayer& l : t->layers) {
    release(&l.z); release(&l.y); release(&l.mean); release(&l.rstd); release(&l.wt_fwd); release(&l.wt_dgrad); release(&l.yb);
  }
  DevBuf* bufs[] = {&t->params, &t->grads, &t->accum, &t->d_segs, &t->d_chunks, &t->const_slope, &t->feats, &t->labels,
                    &t->pooled, &t->logits, &t->gr, &t->coef, &t->rnorm, &t->dbuf[0], &t->dbuf[1], &t->scratch_stats,
                    &t->sum_area, &t->sqnorms, &t->xb0, &t->dzb, &t->ones_pad, &t->zeros_pad, &t->gcomm, &t->feats_alt, &t->labels_alt};
  for (DevBuf* b : bufs) release(b);
  delete t;
  c->train = nullptr;
}

namespace {
// (Re)builds the bf16 operand layouts of every frame layer from the fp32 masters: one launch.
void tr_count(xv_ctx* c, TrainState* t, const char* fam) {
  count_launch(c, fam);
  if (!t->marking) return;
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, c->stream);
  t->marks.push_back(std::make_pair(fam, e));
}

void tr_timeline_report(xv_ctx* c, TrainState* t) {
  tr_count(c, t, "end");
  t->marking = false;
  cudaStreamSynchronize(c->stream);
  std::map<std::string, std::pair<double, int>> fam;
  double total = 0;
  fprintf(stderr, "[xv_train timeline]");
  for (size_t i = 0; i + 1 < t->marks.size(); ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, t->marks[i].second, t->marks[i + 1].second);
    fam[t->marks[i].first].first += ms * 1e3;
    fam[t->marks[i].first].second++;
    total += ms * 1e3;
    fprintf(stderr, " %s=%.1f", t->marks[i].first + 6, ms * 1e3);
  }
  fprintf(stderr, "\n[xv_train timeline] total %.1f us:", total);
  for (const auto& kv : fam) fprintf(stderr, " %s x%d = %.1f;", kv.first.c_str() + 6, kv.second.second, kv.second.first);
  fprintf(stderr, "\n");
  for (auto& m : t->marks) cudaEventDestroy(m.second);
  t->marks.clear();
}

int launch_recast(xv_ctx* c, TrainState* t, const RecastJobs& jobs, cudaStream_t stream) {
  if (!t->mixed || jobs.n == 0) return XV_OK;
  tr_count(c, t, "train.cast_w");
  const int grid = std::min(jobs.total_tiles, c->sm_count * 16);
  recast_weights_kernel<<<grid, 256, 0, stream>>>(jobs);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(c, e, "weight recast launch");
  return XV_OK;
}
}  // namespace

static int train_init(xv_ctx* c, int32_t num_speakers) {
  if (c->resnet || c->scope != "tdnn" || c->pooling_type != "statistics_pooling")
    return fail(c, XV_EUNSUPPORTED, "xv_train_init: the training step covers network_type tdnn with statistics_pooling");
  if (c->feature_norm) return fail(c, XV_EUNSUPPORTED, "xv_train_init: feature_norm is not covered");
  train_release(c);
  TrainState* t = new TrainState();
  c->train = t;
  t->mixed = c->bf16;
  t->fused_stats = !(getenv("XV_TRAIN_FUSED_STATS") && atoi(getenv("XV_TRAIN_FUSED_STATS")) == 0);
  t->overlap = !(getenv("XV_TRAIN_OVERLAP") && atoi(getenv("XV_TRAIN_OVERLAP")) == 0);
  t->wg_overlap = !(getenv("XV_TRAIN_WGRAD_STREAM") && atoi(getenv("XV_TRAIN_WGRAD_STREAM")) == 0);
  if (getenv("XV_TRAIN_TIMELINE")) t->timeline_step = atoll(getenv("XV_TRAIN_TIMELINE"));
  t->early_update = !(getenv("XV_TRAIN_EARLY_UPDATE") && atoi(getenv("XV_TRAIN_EARLY_UPDATE")) == 0);
  t->bulk_rows = getenv("XV_TRAIN_BULK") && atoi(getenv("XV_TRAIN_BULK")) == 1;     // measured no faster than the per-thread-load kernels (r02w): opt-in
  t->wg_overlap_head = !(getenv("XV_TRAIN_WGRAD_STREAM") && atoi(getenv("XV_TRAIN_WGRAD_STREAM")) == 1);   // 1: frame layers only
  {
    const JValue* gd = c->cfg.get("grad_allreduce_dtype");
    std::string want = (gd && gd->type == JValue::Str) ? gd->str : (c->bf16 ? "bf16" : "fp32");
    if (getenv("XV_TRAIN_COMM")) want = getenv("XV_TRAIN_COMM");
    if (want != "bf16" && want != "fp32") return fail(c, XV_EINVAL, "grad_allreduce_dtype must be \"bf16\" or \"fp32\", got \"%s\"", want.c_str());
    t->comm_bf16 = want == "bf16" && t->overlap;
  }
  const JValue& cfg = c->cfg;
  const JValue* lf = cfg.get("loss_func");
  const std::string loss = (lf && lf->type == JValue::Str) ? lf->str : "softmax";
  if (loss != "softmax" && loss != "additive_margin_softmax")
    return fail(c, XV_EUNSUPPORTED, "Not implement %s loss", loss.c_str());       // model/trainer.py:111-129
  t->amsoftmax = loss == "additive_margin_softmax";
  if (t->amsoftmax) {
    const char* need[] = {"amsoftmax_m", "amsoftmax_lambda_min", "amsoftmax_lambda_base", "amsoftmax_lambda_gamma", "amsoftmax_lambda_power"};
    for (const char* k : need)
      if (!cfg.has(k)) return fail(c, XV_EINVAL, "additive_margin_softmax needs config key '%s'", k);
    t->margin = static_cast<float>(jnum(cfg, "amsoftmax_m", 0));
    t->lam_min = static_cast<float>(jnum(cfg, "amsoftmax_lambda_min", 0));
    t->lam_base = static_cast<float>(jnum(cfg, "amsoftmax_lambda_base", 0));
    t->lam_gamma = static_cast<float>(jnum(cfg, "amsoftmax_lambda_gamma", 0));
    t->lam_power = static_cast<float>(jnum(cfg, "amsoftmax_lambda_power", 0));
  }
  if (!cfg.has("weight_l2_regularizer") || !cfg.has("batchnorm_momentum"))
    return fail(c, XV_EINVAL, "training needs config keys 'weight_l2_regularizer' and 'batchnorm_momentum'");
  t->wd = static_cast<float>(jnum(cfg, "weight_l2_regularizer", 0));
  t->wd_out = static_cast<float>(jnum(cfg, "output_weight_l2_regularizer", t->wd));   // model/loss.py:233-235
  t->bn_momentum = static_cast<float>(jnum(cfg, "batchnorm_momentum", 0.99));
  const JValue* ov = cfg.get("optimizer");                                             // model/trainer.py:225-246
  const std::string opt = (ov && ov->type == JValue::Str) ? ov->str : "sgd";
  if (opt == "sgd") {
    if (cfg.has("momentum")) return fail(c, XV_EINVAL, "Using sgd as the optimizer and you should not specify the momentum.");
    t->optimizer = 0;
  } else if (opt == "momentum") {
    t->optimizer = 1;
    t->opt_momentum = static_cast<float>(jnum(cfg, "momentum", 0.9));
    const JValue* nv = cfg.get("use_nesterov");
    t->nesterov = (nv && nv->truthy()) ? 1 : 0;
  } else if (opt == "adam") {
    t->optimizer = 2;
  } else {
    return fail(c, XV_EUNSUPPORTED, "Optimizer %s is not supported.", opt.c_str());
  }
  const JValue* cg = cfg.get("clip_gradient");                                        // model/trainer.py:372-373, 529-533
  if (cg && cg->truthy()) {
    if (!cfg.has("clip_gradient_norm")) return fail(c, XV_EINVAL, "clip_gradient needs the config key 'clip_gradient_norm'");
    t->clip_norm = static_cast<float>(jnum(cfg, "clip_gradient_norm", 0));
    if (!(t->clip_norm > 0.f)) return fail(c, XV_EINVAL, "clip_gradient_norm must be positive");
  }
  t->relu_type = c->relu_type;
  t->speakers = num_speakers;
  t->embed = c->last_nodes;
  if (num_speakers < 2) return fail(c, XV_EINVAL, "xv_train_init: num_speakers must be >= 2");

  // ---- flat parameter buffer: trainables first, then the moving statistics ----
  std::vector<float> host;
  int rc = XV_OK;
  int ctx_sum = 0;
  auto add_layer = [&](const Layer& L, bool frame) {
    TrLayer l;
    l.name = L.name; l.frame = frame; l.has_bn = L.has_bn; l.act = L.act != kActNone; l.taps = L.taps; l.cin = L.cin; l.cout = L.cout;
    l.fused_bn = frame && L.taps > 1;                      // rank-4 conv inputs use fused batch norm (unbiased moving variance)
    if (frame) ctx_sum += L.taps - 1;
    l.context = frame ? ctx_sum : 0;
    l.off_kernel = add_var(c, t, &host, L.kernel_var, static_cast<int64_t>(L.taps) * L.cin * L.cout, t->wd, false, &rc);
    if (rc == XV_OK) l.off_bias = add_var(c, t, &host, L.bias_var, L.cout, 0.f, false, &rc);
    if (rc == XV_OK && L.has_bn) {
      l.off_gamma = add_var(c, t, &host, L.bn_scope + "/gamma", L.cout, 0.f, false, &rc);
      if (rc == XV_OK) l.off_beta = add_var(c, t, &host, L.bn_scope + "/beta", L.cout, 0.f, false, &rc);
    }
    if (rc == XV_OK && l.act && t->relu_type == "prelu") l.off_alpha = add_var(c, t, &host, L.alpha_var, L.cout, 0.f, false, &rc);
    t->layers.push_back(l);
  };
  for (int i = 0; i < c->n_main_frame && rc == XV_OK; ++i) add_layer(c->frame_layers[i], true);
  for (size_t i = 0; i < c->seg_layers.size() && rc == XV_OK; ++i) add_layer(c->seg_layers[i], false);
  if (rc != XV_OK) return rc;
  t->n_frame = c->n_main_frame;
  for (int i = 0; i < t->n_frame; ++i) {
    const TrLayer& l = t->layers[i];
    if (!l.has_bn) return fail(c, XV_EUNSUPPORTED, "xv_train_init: frame layer %s has no batch norm", l.name.c_str());
    if (l.cout % 4 != 0) return fail(c, XV_EUNSUPPORTED, "training needs frame-layer widths that are multiples of 4 (%s has %d)", l.name.c_str(), l.cout);
  }
  t->off_w = add_var(c, t, &host, "softmax/output/kernel", static_cast<int64_t>(t->embed) * num_speakers, t->wd_out, false, &rc);
  if (rc == XV_OK && !t->amsoftmax) t->off_b = add_var(c, t, &host, "softmax/output/bias", num_speakers, 0.f, false, &rc);
  if (rc != XV_OK) return rc;
  t->n_train = static_cast<int64_t>(host.size());
  for (size_t i = 0; i < t->layers.size() && rc == XV_OK; ++i) {
    TrLayer& l = t->layers[i];
    if (!l.has_bn) continue;
    const Layer& L = i < static_cast<size_t>(t->n_frame) ? c->frame_layers[i] : c->seg_layers[i - t->n_frame];
    l.off_mm = add_var(c, t, &host, L.bn_scope + "/moving_mean", l.cout, 0.f, true, &rc);
    if (rc == XV_OK) l.off_mv = add_var(c, t, &host, L.bn_scope + "/moving_variance", l.cout, 0.f, true, &rc);
  }
  if (rc != XV_OK) return rc;
  t->n_total = static_cast<int64_t>(host.size());
  if ((rc = upload(c, &t->params, host.data(), sizeof(float) * t->n_total)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->grads, sizeof(float) * (t->n_total + 1), true)) != XV_OK) return rc;
  if (t->optimizer != 0) {               // momentum accumulator | Adam first and second moments
    const size_t nacc = static_cast<size_t>(t->n_total) * (t->optimizer == 2 ? 2 : 1);
    if ((rc = ensure(c, &t->accum, sizeof(float) * nacc)) != XV_OK) return rc;
    cudaMemsetAsync(t->accum.p, 0, sizeof(float) * nacc, c->stream);
  }
  if (t->clip_norm > 0.f && (rc = ensure(c, &t->sqnorms, sizeof(double) * t->segs.size())) != XV_OK) return rc;
  if ((rc = upload(c, &t->d_segs, t->segs.data(), sizeof(ParamSeg) * t->segs.size())) != XV_OK) return rc;
  for (size_t si = 0; si < t->segs.size(); ++si)
    for (int64_t o = 0; o < t->segs[si].size; o += kUpdChunk) {
      UpdChunk ch;
      ch.off = t->segs[si].off + o;
      ch.n = static_cast<int32_t>(std::min<int64_t>(kUpdChunk, t->segs[si].size - o));
      ch.seg = static_cast<int32_t>(si);
      t->chunks.push_back(ch);
    }
  if ((rc = upload(c, &t->d_chunks, t->chunks.data(), sizeof(UpdChunk) * t->chunks.size())) != XV_OK) return rc;
  // gradient buckets in the order the backward pass retires them (see the file header)
  {
    const int64_t seg0 = t->layers[t->n_frame].off_kernel;
    const int64_t mid = t->layers[t->n_frame / 2].off_kernel;
    t->bucket_lo[0] = seg0; t->bucket_hi[0] = t->n_total + 1;
    t->bucket_lo[1] = mid;  t->bucket_hi[1] = seg0;
    t->bucket_lo[2] = 0;    t->bucket_hi[2] = mid;
    // the update of bucket 0 (segment layers, loss head, moving statistics) can run as soon as its gradients are final:
    // chunks [chunk_split, end) -- valid only if the chunk table is ordered the way the flat buffer is
    const int n_chunks = static_cast<int>(t->chunks.size());
    t->chunk_split = n_chunks;
    for (int i = 0; i < n_chunks; ++i)
      if (t->chunks[i].off >= seg0) { t->chunk_split = i; break; }
    for (int i = t->chunk_split; i < n_chunks; ++i)
      if (t->chunks[i].off < seg0) t->early_update = false;
  }
  int maxc = 4;
  for (const TrLayer& l : t->layers) maxc = std::max(maxc, l.cout);
  t->sum_ld = (maxc + kBN - 1) / kBN * kBN;
  if ((rc = ensure(c, &t->sum_area, sizeof(double) * t->sum_area_doubles(), true)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->scratch_stats, sizeof(float) * 2 * maxc)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->const_slope, sizeof(float) * 2 * maxc)) != XV_OK) return rc;
  const float slope = t->relu_type == "lrelu" ? kLreluAlpha : 0.f;
  fill_f32_kernel<<<64, 256, 0, c->stream>>>(static_cast<float*>(t->const_slope.p), maxc, slope);
  fill_f32_kernel<<<64, 256, 0, c->stream>>>(static_cast<float*>(t->const_slope.p) + maxc, maxc, 1.f);   // linear layers
  if ((rc = ensure(c, &t->rnorm, sizeof(float) * num_speakers)) != XV_OK) return rc;
  for (TrLayer& l : t->layers) {
    if ((rc = ensure(c, &l.mean, sizeof(float) * l.cout)) != XV_OK) return rc;
    if ((rc = ensure(c, &l.rstd, sizeof(float) * l.cout)) != XV_OK) return rc;
  }
  t->recast.n = 0;
  t->recast.total_tiles = 0;
  t->recast_d.n = 0;
  t->recast_d.total_tiles = 0;
  if (t->mixed) {
    const char* we = wgrad_init();
    if (we) return fail(c, XV_ECUDA, "%s", we);
    int max_alloc = 256;
    int prev_ldb = 0;
    auto add_job = [&](RecastJob j, int fast, int64_t slow) {
      RecastJobs& set = j.kind == 0 ? t->recast : t->recast_d;
      const int tile_fast = j.kind == 0 ? 64 : 32;        // forward layout: 64 (k) x 32 (n) tiles, bf16 pairs along k
      j.tiles_x = (fast + tile_fast - 1) / tile_fast;
      j.tile0 = set.total_tiles;
      set.total_tiles += j.tiles_x * static_cast<int32_t>((slow + 31) / 32);
      set.j[set.n++] = j;
    };
    if (t->n_frame > kMaxRecastJobs) return fail(c, XV_EUNSUPPORTED, "too many frame layers for the recast table");
    for (int i = 0; i < t->n_frame; ++i) {
      TrLayer& l = t->layers[i];
      if (i == 0) { l.a_taps = 1; l.a_cin = (l.taps * l.cin + kBK - 1) / kBK * kBK; }
      else { l.a_taps = l.taps; l.a_cin = prev_ldb; }
      l.ktotal = l.a_taps * l.a_cin;
      l.n_alloc = (l.cout + kBN - 1) / kBN * kBN;
      l.c_alloc = (l.a_cin + kBN - 1) / kBN * kBN;
      l.ldb = (l.cout + kBK - 1) / kBK * kBK;
      prev_ldb = l.ldb;
      max_alloc = std::max(max_alloc, std::max(l.n_alloc, l.c_alloc));
      if ((rc = ensure(c, &l.wt_fwd, sizeof(__nv_bfloat16) * static_cast<size_t>(l.n_alloc) * l.ktotal)) != XV_OK) return rc;
      const char* e = make_tmap_bf16_2d(&l.tm_wf, l.wt_fwd.p, l.ktotal, l.n_alloc, static_cast<uint64_t>(l.ktotal) * 2, kBK, kBLoadRows);
      RecastJob jf = {};
      jf.w = static_cast<const float*>(t->params.p) + l.off_kernel; jf.dst = static_cast<__nv_bfloat16*>(l.wt_fwd.p);
      jf.kind = 0; jf.taps = l.taps; jf.cin = l.cin; jf.cout = l.cout; jf.cstride = i == 0 ? l.cin : l.a_cin;
      jf.ktotal = l.ktotal; jf.n_alloc = l.n_alloc;
      add_job(jf, l.ktotal, l.n_alloc);
      if (!e && i > 0) {
        const size_t kd = static_cast<size_t>(l.taps) * l.ldb;
        if ((rc = ensure(c, &l.wt_dgrad, sizeof(__nv_bfloat16) * l.c_alloc * kd)) != XV_OK) return rc;
        e = make_tmap_bf16_2d(&l.tm_wd, l.wt_dgrad.p, kd, l.c_alloc, kd * 2, kBK, kBLoadRows);
        RecastJob jd = {};
        jd.w = jf.w; jd.dst = static_cast<__nv_bfloat16*>(l.wt_dgrad.p);
        jd.kind = 1; jd.taps = l.taps; jd.cin = l.cin; jd.cout = l.cout; jd.ldz = l.ldb; jd.c_alloc = l.c_alloc;
        add_job(jd, l.ldb, static_cast<int64_t>(l.c_alloc) * l.taps);
      }
      if (e) return fail(c, XV_ECUDA, "%s", e);
    }
    if ((rc = ensure(c, &t->ones_pad, sizeof(float) * max_alloc)) != XV_OK) return rc;
    if ((rc = ensure(c, &t->zeros_pad, sizeof(float) * max_alloc)) != XV_OK) return rc;
    fill_f32_kernel<<<64, 256, 0, c->stream>>>(static_cast<float*>(t->ones_pad.p), max_alloc, 1.f);
    cudaMemsetAsync(t->zeros_pad.p, 0, sizeof(float) * max_alloc, c->stream);
    if ((rc = launch_recast(c, t, t->recast, c->stream)) != XV_OK) return rc;
    if ((rc = launch_recast(c, t, t->recast_d, c->stream)) != XV_OK) return rc;
  }
  cudaError_t e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "xv_train_init");
  return XV_OK;
}

}  // namespace xv

// ================================================================================================ the step
namespace xv {
namespace {

#define TR_CUDA(call, what)                                   \
  do {                                                        \
    cudaError_t e__ = (call);                                 \
    if (e__ != cudaSuccess) return cuda_fail(c, e__, what);   \
  } while (0)

int ew_grid(const xv_ctx* c) { return c->sm_count * 8; }

int tr_buffers(xv_ctx* c, TrainState* t, int b, int tt) {
  if (t->cap_b == b && t->cap_t == tt) return XV_OK;
  const int64_t rows = static_cast<int64_t>(b) * tt;
  int rc;
  int maxc = 0;
  int max_ldb = 0;
  for (size_t i = 0; i < t->layers.size(); ++i) {
    TrLayer& l = t->layers[i];
    const int64_t r = l.frame ? rows : b;
    const bool last_frame = static_cast<int>(i) == t->n_frame - 1;
    if ((rc = ensure(c, &l.z, sizeof(float) * r * l.cout)) != XV_OK) return rc;
    // frame layers: the fp32 path keeps y (SIMT GEMM operand), the bf16 path a bf16 copy; the last frame layer's
    // output is never stored (pooling is fused into its normalisation pass)
    if (!l.frame || (!t->mixed && !last_frame)) { if ((rc = ensure(c, &l.y, sizeof(float) * r * l.cout)) != XV_OK) return rc; }
    if (l.frame && t->mixed && !last_frame) { if ((rc = ensure(c, &l.yb, sizeof(__nv_bfloat16) * r * l.ldb)) != XV_OK) return rc; }
    maxc = std::max(maxc, l.frame ? l.cout : 0);
    if (l.frame) max_ldb = std::max(max_ldb, l.ldb);
  }
  if (t->mixed) {
    if ((rc = ensure(c, &t->xb0, sizeof(__nv_bfloat16) * rows * t->layers[0].a_cin)) != XV_OK) return rc;
    const bool two_dz = t->wg_overlap;
    t->dzb_half = two_dz ? static_cast<size_t>(rows) * max_ldb : 0;
    if ((rc = ensure(c, &t->dzb, sizeof(__nv_bfloat16) * rows * max_ldb * (two_dz ? 2 : 1))) != XV_OK) return rc;
    if (two_dz && !t->wg_stream) {
      cudaError_t e = cudaStreamCreateWithFlags(&t->wg_stream, cudaStreamNonBlocking);
      for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&t->ev_dz[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t->ev_wg[i], cudaEventDisableTiming);
      }
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t->ev_wg_mark, cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t->ev_fork, cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t->ev_rnorm, cudaEventDisableTiming);
      for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&t->ev_side_rd[i], cudaEventDisableTiming);
      if (e != cudaSuccess) return cuda_fail(c, e, "weight-gradient stream / events");
    }
    for (int i = 0; i < t->n_frame; ++i) {
      TrLayer& l = t->layers[i];
      const void* a = i == 0 ? t->xb0.p : t->layers[i - 1].yb.p;
      const char* err = make_tmap_bf16_2d(&l.tm_af, a, l.a_cin, rows, static_cast<uint64_t>(l.a_cin) * 2, kBK, kBM);
      if (!err) err = make_tmap_bf16_2d(&l.tm_adz, static_cast<__nv_bfloat16*>(t->dzb.p) + (i & 1) * t->dzb_half, l.ldb, rows, static_cast<uint64_t>(l.ldb) * 2, kBK, kBM);
      if (err) return fail(c, XV_ECUDA, "%s", err);
    }
  }
  const int pool_c = t->layers[t->n_frame - 1].cout;
  if ((rc = ensure(c, &t->feats, sizeof(float) * rows * c->dim)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->labels, sizeof(int32_t) * b)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->feats_alt, sizeof(float) * rows * c->dim)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->labels_alt, sizeof(int32_t) * b)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->pooled, sizeof(float) * b * 2 * pool_c)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->logits, sizeof(float) * static_cast<size_t>(b) * t->speakers)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->gr, sizeof(float) * static_cast<size_t>(b) * t->speakers)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->coef, sizeof(float) * b)) != XV_OK) return rc;
  const size_t dbytes = sizeof(float) * std::max<int64_t>(rows * maxc, static_cast<int64_t>(b) * std::max(2 * pool_c, t->embed));
  for (int i = 0; i < 2; ++i)
    if ((rc = ensure(c, &t->dbuf[i], dbytes)) != XV_OK) return rc;
  t->cap_b = b;
  t->cap_t = tt;
  return XV_OK;
}

// One frame-layer GEMM of the mixed-precision step on the tcgen05 kernel of the embedding path (xv_gemm_sm100.cu):
// out[m, n] = sum_{j, k} a[m + j*dil, k] * wt[n][j*a_cin + k] + shift[n], fp32 stores masked to n < cols; optional
// fused batch statistics (stat != null: column sums of out and out^2 over the rows with (m mod stat_t) < stat_valid).
int tr_gemm_tc(xv_ctx* c, TrainState* t, const CUtensorMap& tm_a, const CUtensorMap& tm_b, int64_t rows, int a_cin, int taps,
               int dil, int n_alloc, const float* shift, float* out, int ldo, int cols, double* stat, int stat_t,
               int stat_valid, const char* fam) {
  GemmPlan g;
  g.tm_a = tm_a;
  g.tm_b = tm_b;
  g.tm_out = tm_a;       // never dereferenced (no bf16 output)
  GemmParams& p = g.p;
  p.m_rows = static_cast<int32_t>(rows);
  p.m_tiles = static_cast<int32_t>((rows + kBM - 1) / kBM);
  p.n_tiles = n_alloc / kBN;
  p.taps = taps; p.dilation = dil; p.kb_per_tap = a_cin / kBK; p.cin = a_cin;
  p.scale = static_cast<const float*>(t->ones_pad.p); p.shift = shift; p.slope = static_cast<const float*>(t->ones_pad.p);
  p.act_tanh = 0; p.relu_only = 0;
  p.out_bf16 = nullptr; p.out_f32 = out; p.ldo = ldo; p.f32_cols = cols;
  p.pool_acc = nullptr; p.pool_ld = 0; p.tile_seg_ptr = nullptr; p.segs = nullptr; p.row_weight = nullptr;
  p.heads = 1; p.head_width = n_alloc;
  p.stat_out = stat; p.stat_ld = t->sum_ld; p.stat_t = stat_t; p.stat_valid = stat_valid;
  g.pool = false;
  tr_count(c, t, fam);
  cudaError_t e = launch_gemm(g, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "tcgen05 training GEMM launch");
  return XV_OK;
}

// Host wait for the stream of a data-parallel step: polls the communicator for asynchronous errors so that a peer's
// failure surfaces as an error here instead of a hang inside the collective.
int dp_wait(xv_ctx* c, TrainState* t, cudaEvent_t ev = nullptr) {     // ev: wait for this event instead of the whole stream
  if (t->world <= 1 || !t->comm || !t->nccl.CommGetAsyncError) {
    cudaError_t e = ev ? cudaEventSynchronize(ev) : cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
    return XV_OK;
  }
  int spins = 0;
  for (;;) {
    cudaError_t q = ev ? cudaEventQuery(ev) : cudaStreamQuery(c->stream);
    if (q == cudaSuccess) break;
    if (q != cudaErrorNotReady) return cuda_fail(c, q, "stream query");
    if ((++spins & 63) == 0) {
      int async_err = 0;
      const int nrc = t->nccl.CommGetAsyncError(t->comm, &async_err);
      if (nrc != 0 || async_err != 0) {
        if (t->nccl.CommAbort) t->nccl.CommAbort(t->comm);
        t->comm = nullptr;
        t->poisoned = true;
        return tr_fail_nccl(c, t, nrc != 0 ? nrc : async_err, "asynchronous NCCL error during the training step (communicator aborted)");
      }
      std::this_thread::sleep_for(std::chrono::microseconds(20));
    }
  }
  int async_err = 0;
  if (t->nccl.CommGetAsyncError(t->comm, &async_err) == 0 && async_err != 0) {
    if (t->nccl.CommAbort) t->nccl.CommAbort(t->comm);
    t->comm = nullptr;
    t->poisoned = true;
    return tr_fail_nccl(c, t, async_err, "asynchronous NCCL error during the training step (communicator aborted)");
  }
  return XV_OK;
}

}  // namespace

// One bucket of the exchange on the communication stream.  fp32: the bucket's range of the gradient buffer in place.
// bf16: the trainable part is cast to bf16 (on the communication stream, off the compute stream's critical path),
// all-reduced as bf16, and read back by the update kernel; the tail (moving statistics + loss slot) stays fp32.
static int exchange_bucket(TrainState* t, int k) {
  float* G = static_cast<float*>(t->grads.p);
  const int64_t lo = t->bucket_lo[k], hi = t->bucket_hi[k];
  if (!t->comm_bf16)
    return t->nccl.AllReduce(G + lo, G + lo, static_cast<size_t>(hi - lo), /*ncclFloat32*/ 7, /*ncclSum*/ 0, t->comm, t->comm_stream);
  const int64_t hi_t = std::min(hi, t->n_train);
  int nrc = 0;
  if (hi_t > lo) {
    __nv_bfloat16* gb = static_cast<__nv_bfloat16*>(t->gcomm.p);
    cudaError_t e = launch_f32_to_bf16(G + lo, gb + lo, static_cast<size_t>(hi_t - lo), t->comm_stream);
    if (e != cudaSuccess) return 1;       // ncclUnhandledCudaError
    nrc = t->nccl.AllReduce(gb + lo, gb + lo, static_cast<size_t>(hi_t - lo), /*ncclBfloat16*/ 9, /*ncclSum*/ 0, t->comm, t->comm_stream);
  }
  if (nrc == 0 && hi > t->n_train) {
    const int64_t a = std::max(lo, t->n_train);
    nrc = t->nccl.AllReduce(G + a, G + a, static_cast<size_t>(hi - a), /*ncclFloat32*/ 7, /*ncclSum*/ 0, t->comm, t->comm_stream);
  }
  return nrc;
}

static int train_step_body(xv_ctx* c, TrainState* t, const float* feats, const int32_t* labels, int32_t b, int32_t tt, float lr,
                           int64_t global_step, float* loss_out, int pipe_slot = -1) {
  int rc = tr_buffers(c, t, b, tt);
  if (rc != XV_OK) return rc;
  cudaStream_t s = c->stream;
  const int64_t rows = static_cast<int64_t>(b) * tt;
  float* P = static_cast<float*>(t->params.p);
  float* G = static_cast<float*>(t->grads.p);
  double* scalars = t->scalars();
  int maxc = 4;
  for (const TrLayer& l : t->layers) maxc = std::max(maxc, l.cout);
  const float* slope_const = static_cast<const float*>(t->const_slope.p);
  const float* slope_linear = slope_const + maxc;
  const int nl = static_cast<int>(t->layers.size());
  t->marking = t->timeline_step >= 0 && global_step == t->timeline_step;
  if (t->marking) tr_count(c, t, "train.h2d+memsets");
  if (feats != nullptr) {          // null: the batch was staged into t->feats / t->labels by the caller (pipelined step)
    TR_CUDA(cudaMemcpyAsync(t->feats.p, feats, sizeof(float) * rows * c->dim, cudaMemcpyDefault, s), "H2D features");
    TR_CUDA(cudaMemcpyAsync(t->labels.p, labels, sizeof(int32_t) * b, cudaMemcpyDefault, s), "H2D labels");
  }
  TR_CUDA(cudaMemsetAsync(t->sum_area.p, 0, sizeof(double) * t->sum_area_doubles(), s), "memset sums");
  // gradient buffer: the frame-layer range is accumulated into (tensor-core wgrad = split-K reductions; biases in front
  // of a batch norm keep their exact zero); the tail carries rank 0's moving statistics + the loss slot
  if (t->clip_norm > 0.f) TR_CUDA(cudaMemsetAsync(t->sqnorms.p, 0, sizeof(double) * t->segs.size(), s), "memset norms");
  // (bf16 path: only the side-stream weight-gradient GEMMs need these zeros, and the side stream clears them itself below)
  const bool side0 = t->mixed && t->wg_overlap && t->wg_stream != nullptr;
  if (!side0) TR_CUDA(cudaMemsetAsync(G, 0, sizeof(float) * t->layers[t->n_frame].off_kernel, s), "memset frame gradients");
  TR_CUDA(cudaMemsetAsync(G + t->n_train, 0, sizeof(float) * (t->n_total - t->n_train + 1), s), "memset tail");
  auto slope_of = [&](const TrLayer& l) -> const float* {
    if (!l.act) return slope_linear;
    return l.off_alpha >= 0 ? P + l.off_alpha : slope_const;
  };
  auto finalize_of = [&](const TrLayer& l) {
    BnFinalize f;
    f.mean = static_cast<float*>(l.mean.p); f.rstd = static_cast<float*>(l.rstd.p);
    f.mov_mean = l.has_bn ? P + l.off_mm : nullptr; f.mov_var = l.has_bn ? P + l.off_mv : nullptr;
    // new moving statistics go to the tail of the gradient buffer on rank 0 (tower 0), to a scratch area elsewhere
    f.new_mm = (t->rank == 0 && l.has_bn) ? G + l.off_mm : static_cast<float*>(t->scratch_stats.p);
    f.new_mv = (t->rank == 0 && l.has_bn) ? G + l.off_mv : static_cast<float*>(t->scratch_stats.p) + maxc;
    f.momentum = t->bn_momentum; f.eps = kBnEpsilon; f.unbiased = l.fused_bn ? 1 : 0;
    return f;
  };
  const PoolGrad no_pg = {nullptr, nullptr, 0.f, 0};

  // 1/|w_s| of the AM-softmax weight columns depends on the weights only: it runs on the side stream beside the forward pass
  const bool rnorm_early = t->amsoftmax && side0 && t->wg_overlap_head;
  if (side0) {
    TR_CUDA(cudaEventRecord(t->ev_fork, s), "cudaEventRecord");                   // behind the previous step's update
    TR_CUDA(cudaStreamWaitEvent(t->wg_stream, t->ev_fork, 0), "cudaStreamWaitEvent");
    TR_CUDA(cudaMemsetAsync(G, 0, sizeof(float) * t->layers[t->n_frame].off_kernel, t->wg_stream), "memset frame gradients");
  }
  if (rnorm_early) {
    tr_count(c, t, "train.col_rnorm");
    col_rnorm_kernel<<<(t->speakers + 31) / 32, 256, 0, t->wg_stream>>>(P + t->off_w, t->embed, t->speakers, static_cast<float*>(t->rnorm.p));
    TR_CUDA(cudaGetLastError(), "column norm launch");
  }
  if (side0) TR_CUDA(cudaEventRecord(t->ev_rnorm, t->wg_stream), "cudaEventRecord");

  // ------------------------------------------------------------------ forward
  const float* in = static_cast<const float*>(t->feats.p);
  int in_c = c->dim;
  for (int i = 0; i < nl; ++i) {
    TrLayer& l = t->layers[i];
    const int64_t r = l.frame ? rows : b;
    const RowMask mask = l.frame ? RowMask{tt, tt - l.context} : RowMask{1, 1};
    const double n = l.frame ? static_cast<double>(b) * (tt - l.context) : static_cast<double>(b);
    if (i == t->n_frame) {
      in = static_cast<const float*>(t->pooled.p);
      in_c = 2 * t->layers[i - 1].cout;
    }
    if (in_c != l.cin) return fail(c, XV_ESTATE, "training plan: layer %s expects %d inputs, got %d", l.name.c_str(), l.cin, in_c);
    const bool tc = t->mixed && l.frame;
    bool have_stats = false;
    if (tc) {
      if (i == 0) {
        tr_count(c, t, "train.splice_cast");
        TR_CUDA(launch_splice_cast(static_cast<const float*>(t->feats.p), rows, c->dim, l.taps, 1,
                                   static_cast<__nv_bfloat16*>(t->xb0.p), l.a_cin, s), "splice launch");
      }
      // the bias array doubles as the padded shift vector: columns >= cout are masked (f32_cols) and their statistics
      // never read, so the few floats the epilogue reads past the bias (the next variable of the flat buffer) are inert
      have_stats = t->fused_stats;
      if ((rc = tr_gemm_tc(c, t, l.tm_af, l.tm_wf, rows, l.a_cin, l.a_taps, 1, l.n_alloc, P + l.off_bias,
                           static_cast<float*>(l.z.p), l.cout, l.cout, have_stats ? t->fwd_sums(i) : nullptr, tt,
                           tt - l.context, "train.tc_fwd")) != XV_OK) return rc;
    } else if (!l.frame) {
      SkinnyArgs sa;
      sa.a = in; sa.lda = l.cin; sa.w = P + l.off_kernel; sa.ldw = l.cout; sa.out = static_cast<float*>(l.z.p); sa.ldo = l.cout;
      sa.col_scale = nullptr; sa.bias = P + l.off_bias; sa.m = b; sa.n = l.cout; sa.k = l.cin; sa.ks = 0;
      tr_count(c, t, "train.skinny_fwd");
      TR_CUDA(launch_skinny_gemm<false>(sa, c->sm_count, s), "segment-level forward GEMM launch");
    } else {
      SimtGemmArgs ga;
      ga.a = in; ga.a_rows = r; ga.cin = l.cin; ga.taps = l.taps; ga.dilation = 1;
      ga.w = P + l.off_kernel; ga.n = l.cout; ga.bias = P + l.off_bias;
      ga.bn_scale = nullptr; ga.bn_shift = nullptr; ga.slope = nullptr; ga.act_tanh = 0;
      ga.out_dense = static_cast<float*>(l.z.p); ga.out_bn = nullptr; ga.out_act = nullptr; ga.m_rows = r;
      tr_count(c, t, "train.gemm_fwd");
      TR_CUDA(launch_simt_gemm(ga, s), "forward GEMM launch");
    }
    if (!l.frame) {
      SegFwdArgs fa;
      fa.z = static_cast<const float*>(l.z.p); fa.rows = b; fa.c_total = l.cout; fa.has_bn = l.has_bn ? 1 : 0;
      fa.fin = finalize_of(l);
      fa.gamma = l.has_bn ? P + l.off_gamma : nullptr; fa.beta = l.has_bn ? P + l.off_beta : nullptr; fa.slope = slope_of(l);
      fa.y = static_cast<float*>(l.y.p);
      tr_count(c, t, "train.seg_bn_act");
      seg_bn_act_fwd_kernel<<<(l.cout + 31) / 32, dim3(32, 8), 0, s>>>(fa);
      TR_CUDA(cudaGetLastError(), "segment BN/activation launch");
      in = static_cast<const float*>(l.y.p);
      in_c = l.cout;
      continue;
    }
    if (!have_stats) {
      tr_count(c, t, "train.bn_stats");
      TR_CUDA(launch_col_reduce4<0>(nullptr, static_cast<const float*>(l.z.p), mask, r, l.cout, nullptr, nullptr, nullptr, nullptr,
                                    nullptr, no_pg, t->fwd_sums(i), t->sum_ld, c->sm_count, s), "BN statistics launch");
    }
    BnFwdArgs fa;
    fa.z = static_cast<const float*>(l.z.p); fa.rows = r; fa.c_total = l.cout; fa.mask = mask;
    fa.st.sums = t->fwd_sums(i); fa.st.ld = t->sum_ld; fa.st.n = n;
    fa.fin = finalize_of(l);
    fa.gamma = P + l.off_gamma; fa.beta = P + l.off_beta; fa.slope = slope_of(l);
    fa.y = nullptr; fa.yb = nullptr; fa.ldb = l.ldb;
    if (i == t->n_frame - 1) {
      // normalise + activate + statistics pooling (model/pooling.py:40-50) in one pass; y is never stored
      tr_count(c, t, "train.bn_act_pool");
      bn_act_pool_fwd_kernel<<<dim3((l.cout / 4 + 31) / 32, b), dim3(32, 8), 0, s>>>(fa, static_cast<float*>(t->pooled.p));
      TR_CUDA(cudaGetLastError(), "BN/activation/pooling launch");
    } else {
      if (tc) fa.yb = static_cast<__nv_bfloat16*>(l.yb.p); else fa.y = static_cast<float*>(l.y.p);
      tr_count(c, t, "train.bn_act");
      BsArgs sa = {};
      sa.z = fa.z; sa.dy = nullptr; sa.rows = r; sa.c_total = l.cout; sa.mask = mask;
      sa.st = fa.st; sa.fin = fa.fin; sa.y = fa.y; sa.yb = fa.yb; sa.ldb = fa.ldb;
      sa.gamma = fa.gamma; sa.beta = fa.beta; sa.slope = fa.slope;
      dim3 sg, sb;
      if (t->bulk_rows && bs_plan(&sa, c->sm_count, &sg, &sb)) {
        TR_CUDA(bs_launch<kBsFwd>(sa, sg, sb, s), "BN/activation launch");
      } else {
        const RowGrid rg = row_grid(l.cout / 4, c->sm_count);
        bn_act_fwd_kernel<<<rg.grid, rg.block, 0, s>>>(fa);
        TR_CUDA(cudaGetLastError(), "BN/activation launch");
      }
      in = static_cast<const float*>(l.y.p);
    }
    in_c = l.cout;
  }

  // Side stream for everything that only PRODUCES parameter gradients (nothing on the critical chain reads them before the
  // exchange / update): the loss head's weight gradient, the segment layers' and the frame layers' weight gradients.
  const bool wg_side = t->mixed && t->wg_overlap && t->wg_stream != nullptr;
  const bool head_side = wg_side && t->wg_overlap_head;
  bool wg_launched = false;
  bool wg_reads[2] = {false, false};            // a side-stream wgrad of this step reads dZ buffer [parity]
  bool side_rd[2] = {false, false};             // a side-stream kernel of this step reads dbuf[i]
  // side stream <- everything enqueued on the main stream so far
  auto side_fork = [&]() -> int {
    TR_CUDA(cudaEventRecord(t->ev_fork, s), "cudaEventRecord");
    TR_CUDA(cudaStreamWaitEvent(t->wg_stream, t->ev_fork, 0), "cudaStreamWaitEvent");
    wg_launched = true;
    return XV_OK;
  };
  // the side-stream work enqueued so far reads dbuf[i]: the main stream must not overwrite it earlier
  auto side_reads_dbuf = [&](int i) -> int {
    TR_CUDA(cudaEventRecord(t->ev_side_rd[i], t->wg_stream), "cudaEventRecord");
    side_rd[i] = true;
    return XV_OK;
  };
  auto before_dbuf_write = [&](int i) -> int {
    if (!side_rd[i]) return XV_OK;
    TR_CUDA(cudaStreamWaitEvent(s, t->ev_side_rd[i], 0), "cudaStreamWaitEvent");
    side_rd[i] = false;
    return XV_OK;
  };

  // ------------------------------------------------------------------ loss (model/loss.py:9-48, 201-286)
  const TrLayer& last = t->layers.back();
  const float* fvec = static_cast<const float*>(last.y.p);
  const int E = t->embed, S = t->speakers;
  float fa = 0.f;
  // the side stream's work of the step start (gradient zeros, AM-softmax column norms) is complete before anything below
  // reads the norms or stores a parameter gradient
  if (side0) TR_CUDA(cudaStreamWaitEvent(s, t->ev_rnorm, 0), "cudaStreamWaitEvent");
  if (t->amsoftmax) {
    const double lam = std::max(static_cast<double>(t->lam_min),
                                static_cast<double>(t->lam_base) * pow(1.0 + static_cast<double>(t->lam_gamma) * static_cast<double>(global_step),
                                                                        -static_cast<double>(t->lam_power)));
    fa = static_cast<float>(1.0 / (1.0 + lam));
    if (!rnorm_early) {
      tr_count(c, t, "train.col_rnorm");
      col_rnorm_kernel<<<(S + 31) / 32, 256, 0, s>>>(P + t->off_w, E, S, static_cast<float*>(t->rnorm.p));
      TR_CUDA(cudaGetLastError(), "column norm launch");
    }
  }
  {
    // logits = f . W (softmax: + bias; AM-softmax: columns scaled by 1/|w_s|, model/loss.py:241-243)
    SkinnyArgs sa;
    sa.a = fvec; sa.lda = E; sa.w = P + t->off_w; sa.ldw = S; sa.out = static_cast<float*>(t->logits.p); sa.ldo = S;
    sa.col_scale = t->amsoftmax ? static_cast<const float*>(t->rnorm.p) : nullptr;
    sa.bias = t->amsoftmax ? nullptr : P + t->off_b;
    sa.m = b; sa.n = S; sa.k = E; sa.ks = 0;
    tr_count(c, t, "train.skinny_fwd");
    TR_CUDA(launch_skinny_gemm<false>(sa, c->sm_count, s), "logits GEMM launch");
  }
  tr_count(c, t, "train.loss");
  softmax_loss_kernel<<<b, kLossThreads, 0, s>>>(static_cast<float*>(t->logits.p), S, fvec, E, static_cast<const int32_t*>(t->labels.p), b,
                                        t->amsoftmax ? static_cast<const float*>(t->rnorm.p) : nullptr, t->amsoftmax ? 1 : 0,
                                        t->margin, fa, static_cast<float*>(t->gr.p), static_cast<float*>(t->coef.p), scalars);
  TR_CUDA(cudaGetLastError(), "loss launch");
  // dW = f^T G (+ the l2_normalize Jacobian for AM-softmax), db = column sums of G; the tower loss goes to the tail slot
  if (E <= kHeadMaxE && E % 32 == 0) {
    cudaStream_t hs = s;
    if (head_side) {
      if ((rc = side_fork()) != XV_OK) return rc;
      hs = t->wg_stream;
    }
    tr_count(c, t, "train.head_wgrad");
    head_wgrad_kernel<<<(S + 31) / 32, 256, 0, hs>>>(fvec, static_cast<const float*>(t->logits.p), b, E, S, P + t->off_w,
                                                    t->amsoftmax ? static_cast<const float*>(t->rnorm.p) : nullptr, G + t->off_w,
                                                    t->amsoftmax ? nullptr : G + t->off_b, scalars, G + t->n_total);
    TR_CUDA(cudaGetLastError(), "softmax weight-gradient launch");
  } else {
    RowsTArgs wa;
    wa.a = fvec; wa.lda = E; wa.dz = static_cast<const float*>(t->logits.p); wa.ldz = S; wa.out = G + t->off_w; wa.ldo = S;
    wa.rows = b; wa.k = E; wa.n = S;
    tr_count(c, t, "train.gemm_wgrad");
    TR_CUDA(launch_rowsT_gemm(wa, s), "softmax wgrad launch");
    if (t->amsoftmax) amsoftmax_dw_kernel<<<(S + 31) / 32, 256, 0, s>>>(G + t->off_w, P + t->off_w, static_cast<const float*>(t->rnorm.p), E, S);
    else col_sum_small_kernel<<<(S + 255) / 256, 256, 0, s>>>(static_cast<const float*>(t->logits.p), b, S, G + t->off_b);
    loss_to_tail_kernel<<<1, 1, 0, s>>>(scalars, G + t->n_total);
    TR_CUDA(cudaGetLastError(), "softmax gradient launch");
  }
  int cur = 0;
  float* dy = static_cast<float*>(t->dbuf[cur].p);
  {
    SkinnyArgs sa;
    sa.a = static_cast<const float*>(t->gr.p); sa.lda = S; sa.w = P + t->off_w; sa.ldw = S; sa.out = dy; sa.ldo = E;
    sa.col_scale = nullptr; sa.bias = nullptr; sa.m = b; sa.n = E; sa.k = S; sa.ks = 0;
    tr_count(c, t, "train.skinny_dgrad");
    TR_CUDA(launch_skinny_gemm<true>(sa, c->sm_count, s), "softmax dgrad launch");
    if (t->amsoftmax) {
      tr_count(c, t, "train.axpy");
      axpy_rows_kernel<<<ew_grid(c), 256, 0, s>>>(dy, fvec, static_cast<const float*>(t->coef.p), b, E);
      TR_CUDA(cudaGetLastError(), "axpy launch");
    }
  }

  // ------------------------------------------------------------------ exchange plumbing
  const bool dp = t->world > 1;
  const bool overlap = dp && t->overlap;
  // ---- update of a chunk range on a stream (clipping norms + optimizer), see the end of the step
  const int n_chunks = static_cast<int>(t->chunks.size());
  // bf16 exchange: the summed gradients of the trainable range come back in the bf16 travel buffer
  const __nv_bfloat16* gsum_bf16 = (dp && overlap && t->comm_bf16) ? static_cast<const __nv_bfloat16*>(t->gcomm.p) : nullptr;
  float lr_eff = lr;
  if (t->optimizer == 2) {
    ++t->adam_t;
    lr_eff = static_cast<float>(lr * std::sqrt(1.0 - std::pow(0.999, static_cast<double>(t->adam_t))) /
                                (1.0 - std::pow(0.9, static_cast<double>(t->adam_t))));
  }
  auto launch_update = [&](int c_lo, int c_hi, cudaStream_t us) -> int {
    const int nc = c_hi - c_lo;
    if (nc <= 0) return XV_OK;
    const UpdChunk* chunks = static_cast<const UpdChunk*>(t->d_chunks.p) + c_lo;
    const double* sqn = nullptr;
    if (t->clip_norm > 0.f) {
      // clipping acts on the AVERAGED gradient (model/trainer.py:1244-1246), per variable; the norms were zeroed at the step start
      double* d_sqn = static_cast<double*>(t->sqnorms.p);
      tr_count(c, t, "train.grad_norms");
      grad_sqnorm_kernel<<<nc, 256, 0, us>>>(P, G, gsum_bf16, static_cast<const ParamSeg*>(t->d_segs.p), chunks, 1.f / t->world, d_sqn);
      TR_CUDA(cudaGetLastError(), "gradient norm launch");
      sqn = d_sqn;
    }
    tr_count(c, t, "train.update");
    apply_update_kernel<<<nc, 256, 0, us>>>(P, G, gsum_bf16, static_cast<float*>(t->accum.p), static_cast<const ParamSeg*>(t->d_segs.p),
                                            chunks, t->n_total, 1.f / t->world, lr_eff, t->optimizer, t->opt_momentum, t->nesterov,
                                            sqn, t->clip_norm, scalars + 1);
    TR_CUDA(cudaGetLastError(), "update launch");
    return XV_OK;
  };
  // Bucket 0 (segment layers, loss head, moving statistics: half of the parameters) is updated while the frame layers are still
  // in their backward pass -- nothing there reads those variables.  Data parallel: on the communication stream behind the
  // bucket's all-reduce; one GPU: on the side stream behind the weight gradients it holds.
  const bool early_dp = t->early_update && overlap && t->chunk_split < n_chunks;
  const bool early_side = t->early_update && !dp && wg_side && head_side && t->chunk_split < n_chunks;
  const int late_hi = (early_dp || early_side) ? t->chunk_split : n_chunks;
  auto bucket_ready = [&](int k) -> int {       // every gradient of bucket k has been written on the main stream
    if (!overlap || t->bucket_hi[k] <= t->bucket_lo[k]) return XV_OK;
    TR_CUDA(cudaEventRecord(t->ev_ready[k], s), "cudaEventRecord");
    TR_CUDA(cudaStreamWaitEvent(t->comm_stream, t->ev_ready[k], 0), "cudaStreamWaitEvent");
    if (wg_launched) {                          // ... and the weight gradients that were sent to the side stream
      TR_CUDA(cudaEventRecord(t->ev_wg_mark, t->wg_stream), "cudaEventRecord");
      TR_CUDA(cudaStreamWaitEvent(t->comm_stream, t->ev_wg_mark, 0), "cudaStreamWaitEvent");
    }
    int nrc = exchange_bucket(t, k);
    if (nrc != 0) return tr_fail_nccl(c, t, nrc, "ncclAllReduce");
    tr_count(c, t, "train.allreduce");
    if (k == 0 && early_dp) return launch_update(t->chunk_split, n_chunks, t->comm_stream);
    return XV_OK;
  };

  // ------------------------------------------------------------------ backward
  PoolGrad pg = no_pg;
  for (int i = nl - 1; i >= 0; --i) {
    TrLayer& l = t->layers[i];
    const int64_t r = l.frame ? rows : b;
    const RowMask mask = l.frame ? RowMask{tt, tt - l.context} : RowMask{1, 1};
    const double n = l.frame ? static_cast<double>(b) * (tt - l.context) : static_cast<double>(b);
    const bool tc = t->mixed && l.frame;
    const float *mean = nullptr, *rstd = nullptr, *gamma = nullptr, *beta = nullptr;
    if (l.has_bn) {
      mean = static_cast<const float*>(l.mean.p); rstd = static_cast<const float*>(l.rstd.p);
      gamma = P + l.off_gamma; beta = P + l.off_beta;
    }
    if (!l.frame) {
      // ---- segment-level layer: reduce + apply + parameter gradients in one kernel, dy -> dz in place
      SegBwdArgs ba;
      ba.dy = dy; ba.z = static_cast<const float*>(l.z.p); ba.rows = b; ba.c_total = l.cout;
      ba.mean = mean; ba.rstd = rstd; ba.gamma = gamma; ba.beta = beta; ba.slope = slope_of(l);
      ba.g_gamma = l.has_bn ? G + l.off_gamma : nullptr; ba.g_beta = l.has_bn ? G + l.off_beta : nullptr;
      ba.g_alpha = l.off_alpha >= 0 ? G + l.off_alpha : nullptr; ba.g_bias = G + l.off_bias;
      tr_count(c, t, "train.seg_bn_bwd");
      seg_bn_bwd_kernel<<<(l.cout + 31) / 32, dim3(32, 8), 0, s>>>(ba);
      TR_CUDA(cudaGetLastError(), "segment BN backward launch");
      RowsTArgs wa;
      wa.a = i == t->n_frame ? static_cast<const float*>(t->pooled.p) : static_cast<const float*>(t->layers[i - 1].y.p);
      wa.lda = l.cin; wa.dz = dy; wa.ldz = l.cout; wa.out = G + l.off_kernel; wa.ldo = l.cout; wa.rows = b; wa.k = l.cin; wa.n = l.cout;
      tr_count(c, t, "train.seg_wgrad");
      if (head_side) {
        if ((rc = side_fork()) != XV_OK) return rc;
        TR_CUDA(launch_rowsT_gemm(wa, t->wg_stream), "segment wgrad launch");
        if ((rc = side_reads_dbuf(cur)) != XV_OK) return rc;           // dz lives in dbuf[cur]
      } else {
        TR_CUDA(launch_rowsT_gemm(wa, s), "segment wgrad launch");
      }
      if ((rc = before_dbuf_write(cur ^ 1)) != XV_OK) return rc;
      float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
      SkinnyArgs sa;
      sa.a = dy; sa.lda = l.cout; sa.w = P + l.off_kernel; sa.ldw = l.cout; sa.out = dx; sa.ldo = l.cin;
      sa.col_scale = nullptr; sa.bias = nullptr; sa.m = b; sa.n = l.cin; sa.k = l.cout; sa.ks = 0;
      tr_count(c, t, "train.skinny_dgrad");
      TR_CUDA(launch_skinny_gemm<true>(sa, c->sm_count, s), "segment-level dgrad launch");
      cur ^= 1;
      dy = dx;
      if (i == t->n_frame) {
        // dy now holds d(pooled) [B, 2C]: the frame-level kernels below rebuild the pooling gradient from it
        const TrLayer& pl = t->layers[i - 1];
        pg.pooled = static_cast<const float*>(t->pooled.p); pg.dpooled = dy;
        pg.inv_l = 1.f / static_cast<float>(tt - pl.context); pg.enabled = 1;
        if ((rc = bucket_ready(0)) != XV_OK) return rc;
        if (early_side) {
          if ((rc = side_fork()) != XV_OK) return rc;
          if ((rc = launch_update(t->chunk_split, n_chunks, t->wg_stream)) != XV_OK) return rc;
        }
      }
      continue;
    }
    // ---- frame-level layer
    const bool from_pool = i == t->n_frame - 1;
    const PoolGrad& lpg = from_pool ? pg : no_pg;
    double* bs = t->bwd_sums(i);
    tr_count(c, t, "train.bn_bwd");
    BsArgs sa = {};
    sa.z = static_cast<const float*>(l.z.p); sa.dy = from_pool ? nullptr : dy; sa.rows = r; sa.c_total = l.cout; sa.mask = mask;
    sa.mean = mean; sa.rstd = rstd; sa.gamma = gamma; sa.beta = beta; sa.slope = slope_of(l); sa.pg = lpg;
    sa.sums_out = bs; sa.sums_ld = t->sum_ld; sa.n = n; sa.ldb = l.ldb;
    dim3 sg, sb;
    const bool staged = t->bulk_rows && bs_plan(&sa, c->sm_count, &sg, &sb);
    if (staged) {
      TR_CUDA(bs_launch<kBsReduce>(sa, sg, sb, s), "BN backward reduce launch");
    } else {
      TR_CUDA(launch_col_reduce4<1>(from_pool ? nullptr : dy, static_cast<const float*>(l.z.p), mask, r, l.cout, mean, rstd, gamma, beta,
                                    slope_of(l), lpg, bs, t->sum_ld, c->sm_count, s), "BN backward reduce launch");
    }
    BnBwdArgs ba;
    ba.dy = from_pool ? nullptr : dy; ba.z = static_cast<const float*>(l.z.p); ba.rows = r; ba.c_total = l.cout; ba.mask = mask;
    ba.mean = mean; ba.rstd = rstd; ba.gamma = gamma; ba.beta = beta; ba.slope = slope_of(l);
    ba.sums = bs; ba.sums_ld = t->sum_ld; ba.n = n; ba.pg = lpg;
    ba.g_gamma = G + l.off_gamma; ba.g_beta = G + l.off_beta; ba.g_alpha = l.off_alpha >= 0 ? G + l.off_alpha : nullptr;
    ba.ldb = l.ldb;
    float* dz32 = nullptr;
    __nv_bfloat16* dzb_l = static_cast<__nv_bfloat16*>(t->dzb.p) + (i & 1) * t->dzb_half;
    if (tc) {
      ba.dz32 = nullptr; ba.dzb = dzb_l;
      // layer i+2's weight gradient may still be reading this dZ buffer on the side stream
      if (wg_side && wg_reads[i & 1]) TR_CUDA(cudaStreamWaitEvent(s, t->ev_wg[i & 1], 0), "cudaStreamWaitEvent");
    } else {
      dz32 = from_pool ? static_cast<float*>(t->dbuf[cur ^ 1].p) : dy;      // in place unless the gradient was virtual
      ba.dz32 = dz32; ba.dzb = nullptr;
    }
    if (staged) {
      sa.dz32 = ba.dz32; sa.dzb = ba.dzb;
      sa.g_gamma = ba.g_gamma; sa.g_beta = ba.g_beta; sa.g_alpha = ba.g_alpha;
      TR_CUDA(bs_launch<kBsApply>(sa, sg, sb, s), "BN backward apply launch");
    } else {
      const RowGrid rg = row_grid(l.cout / 4, c->sm_count);
      bn_bwd_apply_kernel<<<rg.grid, rg.block, 0, s>>>(ba);
      TR_CUDA(cudaGetLastError(), "BN backward apply launch");
    }
    if (!tc && from_pool) { cur ^= 1; dy = dz32; }
    if (tc) {
      // The data gradient is the critical path (the next layer's BN backward waits for it) and is enqueued first; dW on the
      // MN-major tcgen05 kernel (xv_wgrad_sm100.cu) follows on the side stream: layer 0 contracts the pre-spliced features
      // (one tap of width taps*dim), the others shift the previous layer's bf16 activations by the tap index
      if (wg_side) TR_CUDA(cudaEventRecord(t->ev_dz[i & 1], s), "cudaEventRecord");
      if (i > 0) {
        if ((rc = before_dbuf_write(cur ^ 1)) != XV_OK) return rc;
        float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
        // dX[m, c] = sum_j sum_n dZ[m - j, n] W[j, c, n]: the forward kernel with tap step -1 (rows < 0 are TMA zero fill)
        if ((rc = tr_gemm_tc(c, t, l.tm_adz, l.tm_wd, rows, l.ldb, l.taps, -1, l.c_alloc, static_cast<const float*>(t->zeros_pad.p),
                             dx, l.cin, l.cin, nullptr, 1, 1, "train.tc_dgrad")) != XV_OK) return rc;
        cur ^= 1;
        dy = dx;
      }
      cudaStream_t ws = s;
      if (wg_side) {
        TR_CUDA(cudaStreamWaitEvent(t->wg_stream, t->ev_dz[i & 1], 0), "cudaStreamWaitEvent");
        ws = t->wg_stream;
      }
      const void* a_bf16 = i == 0 ? t->xb0.p : t->layers[i - 1].yb.p;
      cudaError_t ce;
      tr_count(c, t, "train.tc_wgrad");
      const char* we = launch_wgrad(a_bf16, rows, l.a_cin, dzb_l, l.ldb, rows, i == 0 ? 1 : l.taps, i == 0 ? l.taps * l.cin : l.cin,
                                    l.cout, G + l.off_kernel, l.cout, c->sm_count, ws, &ce);
      if (we) return fail(c, XV_ECUDA, "%s", we);
      TR_CUDA(ce, "tcgen05 wgrad launch");
      if (wg_side) {
        TR_CUDA(cudaEventRecord(t->ev_wg[i & 1], t->wg_stream), "cudaEventRecord");
        wg_reads[i & 1] = true;
        wg_launched = true;
      }
    } else {
      const float* a_in = i == 0 ? static_cast<const float*>(t->feats.p) : static_cast<const float*>(t->layers[i - 1].y.p);
      TrGemmArgs wa;
      wa.a = a_in; wa.lda = i == 0 ? c->dim : l.cin; wa.dz = dy; wa.w = nullptr; wa.out = G + l.off_kernel; wa.rows = r; wa.cin = l.cin;
      wa.nz = l.cout; wa.taps = l.taps;
      tr_count(c, t, "train.gemm_wgrad");
      TR_CUDA(launch_tr_gemm<kWgrad>(wa, s), "wgrad launch");
      if (i > 0) {
        float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
        TrGemmArgs da;
        da.a = nullptr; da.lda = 0; da.dz = dy; da.w = P + l.off_kernel; da.out = dx; da.rows = r; da.cin = l.cin; da.nz = l.cout;
        da.taps = l.taps;
        tr_count(c, t, "train.gemm_dgrad");
        TR_CUDA(launch_tr_gemm<kDgrad>(da, s), "dgrad launch");
        cur ^= 1;
        dy = dx;
      }
    }
    if (i == t->n_frame / 2 && (rc = bucket_ready(1)) != XV_OK) return rc;
    if (i == 0 && (rc = bucket_ready(2)) != XV_OK) return rc;
  }

  // ------------------------------------------------------------------ exchange, update
  if (wg_launched) {                              // the update (and the next step's forward) follow every weight gradient
    TR_CUDA(cudaEventRecord(t->ev_wg_mark, t->wg_stream), "cudaEventRecord");
    TR_CUDA(cudaStreamWaitEvent(s, t->ev_wg_mark, 0), "cudaStreamWaitEvent");
  }
  if (dp && overlap) {
    TR_CUDA(cudaEventRecord(t->ev_done, t->comm_stream), "cudaEventRecord");
    TR_CUDA(cudaStreamWaitEvent(s, t->ev_done, 0), "cudaStreamWaitEvent");
  } else if (dp) {
    // ONE collective: gradients of every variable + rank 0's moving statistics + the tower loss
    const int nrc = t->nccl.AllReduce(G, G, static_cast<size_t>(t->n_total + 1), /*ncclFloat32*/ 7, /*ncclSum*/ 0, t->comm, s);
    if (nrc != 0) return tr_fail_nccl(c, t, nrc, "ncclAllReduce");
    tr_count(c, t, "train.allreduce");
  }
  if ((rc = launch_update(0, late_hi, s)) != XV_OK) return rc;
  // bf16 operand copies of the NEW weights for the next step: the forward layouts now, the data-gradient layouts (first read by
  // the next step's backward pass) on the side stream when there is one -- the main stream meets it again before that step's loss
  if ((rc = launch_recast(c, t, t->recast, s)) != XV_OK) return rc;
  if (side0) {
    TR_CUDA(cudaEventRecord(t->ev_fork, s), "cudaEventRecord");
    TR_CUDA(cudaStreamWaitEvent(t->wg_stream, t->ev_fork, 0), "cudaStreamWaitEvent");
    if ((rc = launch_recast(c, t, t->recast_d, t->wg_stream)) != XV_OK) return rc;
  } else {
    if ((rc = launch_recast(c, t, t->recast_d, s)) != XV_OK) return rc;
  }
  if (t->marking) tr_timeline_report(c, t);
  if (loss_out) {
    float loss_sum = 0.f;
    double sc[2] = {0, 0};
    TR_CUDA(cudaMemcpyAsync(&loss_sum, G + t->n_total, sizeof(float), cudaMemcpyDeviceToHost, s), "D2H loss");
    TR_CUDA(cudaMemcpyAsync(sc, scalars, sizeof(double) * 2, cudaMemcpyDeviceToHost, s), "D2H scalars");
    if ((rc = dp_wait(c, t)) != XV_OK) return rc;
    loss_out[0] = loss_sum / t->world;                 // mean of the tower losses (model/trainer.py:1218-1219)
    loss_out[1] = loss_out[0] + static_cast<float>(sc[1]);
  } else if (pipe_slot >= 0) {
    TrainState::PipeSlot& ps = t->pipe[pipe_slot];
    TR_CUDA(cudaMemcpyAsync(ps.h_loss, G + t->n_total, sizeof(float), cudaMemcpyDeviceToHost, s), "D2H loss");
    TR_CUDA(cudaMemcpyAsync(ps.h_loss + 2, scalars, sizeof(double) * 2, cudaMemcpyDeviceToHost, s), "D2H scalars");
    TR_CUDA(cudaEventRecord(ps.end, s), "cudaEventRecord");
  }
  return XV_OK;
}

static bool is_device_pointer(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Pipelined step, first half: the batch goes to the alternate feature / label buffers on the COPY stream (it overlaps the step
// that is still running on the compute stream), then the whole step is enqueued behind it; nothing is waited for.
static int train_submit_body(xv_ctx* c, TrainState* t, const float* feats, const int32_t* labels, int32_t b, int32_t tt, float lr,
                             int64_t global_step) {
  int rc = tr_buffers(c, t, b, tt);
  if (rc != XV_OK) return rc;
  const int slot = t->pipe_next;
  TrainState::PipeSlot& ps = t->pipe[slot];
  if (!ps.end) {
    TR_CUDA(cudaEventCreateWithFlags(&ps.copied, cudaEventDisableTiming), "cudaEventCreate");
    TR_CUDA(cudaEventCreateWithFlags(&ps.end, cudaEventDisableTiming), "cudaEventCreate");
    TR_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ps.h_loss), 32, cudaHostAllocDefault), "cudaHostAlloc");
  }
  // the buffers about to be overwritten were read by the step two submits ago (same slot): it has been collected, hence
  // finished, unless the caller never collected it -- the event makes the order explicit either way
  std::swap(t->feats, t->feats_alt);
  std::swap(t->labels, t->labels_alt);
  if (ps.used) TR_CUDA(cudaStreamWaitEvent(c->copy_stream, ps.end, 0), "cudaStreamWaitEvent");
  const int64_t rows = static_cast<int64_t>(b) * tt;
  TR_CUDA(cudaMemcpyAsync(t->feats.p, feats, sizeof(float) * rows * c->dim, cudaMemcpyDefault, c->copy_stream), "H2D features");
  TR_CUDA(cudaMemcpyAsync(t->labels.p, labels, sizeof(int32_t) * b, cudaMemcpyDefault, c->copy_stream), "H2D labels");
  TR_CUDA(cudaEventRecord(ps.copied, c->copy_stream), "cudaEventRecord");
  TR_CUDA(cudaStreamWaitEvent(c->stream, ps.copied, 0), "cudaStreamWaitEvent");
  rc = train_step_body(c, t, nullptr, nullptr, b, tt, lr, global_step, nullptr, slot);
  if (rc != XV_OK) return rc;
  ps.used = true;
  t->pipe_next ^= 1;
  t->pipe_inflight++;
  return XV_OK;
}

static int train_step(xv_ctx* c, const float* feats, const int32_t* labels, int32_t b, int32_t tt, float lr,
                      int64_t global_step, float* loss_out, bool pipelined = false) {
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "xv_train_step: call xv_train_init first");
  if (t->poisoned) return fail(c, XV_ESTATE, "xv_train_step: the communicator was aborted by an earlier failed step; re-create the context");
  if (t->world > 1 && !t->comm) return fail(c, XV_ESTATE, "xv_train_step: world > 1 but no communicator (xv_comm_init)");
  if (!pipelined && t->pipe_inflight > 0)
    return fail(c, XV_ESTATE, "xv_train_step: %d submitted step(s) have not been collected (xv_train_collect)", t->pipe_inflight);
  if (pipelined && t->pipe_inflight >= 2)
    return fail(c, XV_ESTATE, "xv_train_submit: two steps are in flight; collect one first (xv_train_collect)");
  // Everything that can fail for reasons local to this rank is checked BEFORE any work (and any collective) is enqueued.
  int rc = XV_OK;
  if (b < 1 || tt <= c->total_context) rc = fail(c, XV_EINVAL, "xv_train_step: need B >= 1 and T > %d", c->total_context);
  else if (static_cast<int64_t>(b) * tt >= (1LL << 31)) rc = fail(c, XV_EINVAL, "xv_train_step: B*T must stay below 2^31 rows");
  else if (!is_device_pointer(labels))       // labels resident on the device are the caller's responsibility
    for (int u = 0; u < b; ++u)
      if (labels[u] < 0 || labels[u] >= t->speakers) {
        rc = fail(c, XV_EINVAL, "label %d of utterance %d is outside [0,%d)", labels[u], u, t->speakers);
        break;
      }
  if (rc == XV_OK)
    rc = pipelined ? train_submit_body(c, t, feats, labels, b, tt, lr, global_step)
                   : train_step_body(c, t, feats, labels, b, tt, lr, global_step, loss_out);
  if (rc != XV_OK && t->world > 1 && t->comm) {
    // The peers are (or will be) inside this step's collectives: abort the communicator so that they get an
    // asynchronous error (dp_wait) instead of waiting forever.  The context cannot train any further.
    const std::string msg = c->err;
    if (t->nccl.CommAbort) t->nccl.CommAbort(t->comm);
    t->comm = nullptr;
    t->poisoned = true;
    fail(c, rc, "%s (data-parallel step failed on rank %d: communicator aborted)", msg.c_str(), t->rank);
  }
  return rc;
}

}  // namespace xv

using namespace xv;

extern "C" {

int xv_train_init(xv_ctx* c, int32_t num_speakers) {
  if (!c) return XV_EINVAL;
  cudaSetDevice(c->device);
  int rc = train_init(c, num_speakers);
  if (rc != XV_OK && c->train) train_release(c);
  return rc;
}

int xv_train_step(xv_ctx* c, const float* feats, const int32_t* labels, int32_t batch, int32_t frames, float learning_rate,
                  int64_t global_step, float* loss_out) {
  if (!c) return XV_EINVAL;
  if (!feats || !labels) return fail(c, XV_EINVAL, "xv_train_step: null argument");
  cudaSetDevice(c->device);
  return train_step(c, feats, labels, batch, frames, learning_rate, global_step, loss_out);
}

int xv_train_submit(xv_ctx* c, const float* feats, const int32_t* labels, int32_t batch, int32_t frames, float learning_rate,
                    int64_t global_step) {
  if (!c) return XV_EINVAL;
  if (!feats || !labels) return fail(c, XV_EINVAL, "xv_train_submit: null argument");
  cudaSetDevice(c->device);
  return train_step(c, feats, labels, batch, frames, learning_rate, global_step, nullptr, true);
}

int xv_train_collect(xv_ctx* c, float* loss_out) {
  if (!c) return XV_EINVAL;
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "xv_train_collect: call xv_train_init first");
  if (t->pipe_inflight <= 0) return fail(c, XV_ESTATE, "xv_train_collect: no submitted step is waiting");
  cudaSetDevice(c->device);
  const int slot = (t->pipe_next - t->pipe_inflight) & 1;
  TrainState::PipeSlot& ps = t->pipe[slot];
  t->pipe_inflight--;
  if (t->poisoned) return fail(c, XV_ESTATE, "xv_train_collect: the communicator was aborted by an earlier failed step");
  int rc = dp_wait(c, t, ps.end);
  if (rc != XV_OK) return rc;
  if (loss_out) {
    double sc[2];
    memcpy(sc, ps.h_loss + 2, sizeof(sc));
    loss_out[0] = ps.h_loss[0] / t->world;
    loss_out[1] = loss_out[0] + static_cast<float>(sc[1]);
  }
  return XV_OK;
}

static int train_read(xv_ctx* c, const char* name, float* host, int64_t cap, bool grad) {
  if (!c) return XV_EINVAL;
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "no training state (call xv_train_init)");
  if (!name || !host) return fail(c, XV_EINVAL, "null argument");
  auto it = t->index.find(name);
  if (it == t->index.end()) return fail(c, XV_ENOTFOUND, "unknown variable '%s'", name);
  if (cap < it->second.second) return fail(c, XV_ECAPACITY, "buffer holds %lld floats, need %lld", (long long)cap, (long long)it->second.second);
  cudaSetDevice(c->device);
  const float* src = static_cast<const float*>(grad ? t->grads.p : t->params.p) + it->second.first;
  cudaError_t e = cudaMemcpyAsync(host, src, sizeof(float) * it->second.second, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "D2H of a variable");
  return XV_OK;
}

int xv_train_get_weight(xv_ctx* c, const char* tf_name, float* host, int64_t capacity) {
  return train_read(c, tf_name, host, capacity, false);
}
int xv_train_get_grad(xv_ctx* c, const char* tf_name, float* host, int64_t capacity) {
  return train_read(c, tf_name, host, capacity, true);
}

int xv_train_num_variables(xv_ctx* c) {
  if (!c || !c->train) return -1;
  return static_cast<int>(c->train->index.size());
}

int xv_train_variable_name(xv_ctx* c, int32_t i, char* buf, size_t buf_size, int64_t* size) {
  if (!c) return XV_EINVAL;
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "no training state (call xv_train_init)");
  if (i < 0 || i >= static_cast<int32_t>(t->index.size()) || !buf || buf_size == 0) return fail(c, XV_EINVAL, "xv_train_variable_name: bad argument");
  auto it = t->index.begin();
  std::advance(it, i);
  snprintf(buf, buf_size, "%s", it->first.c_str());
  if (size) *size = it->second.second;
  return XV_OK;
}

int xv_comm_unique_id(xv_ctx* c, void* out128) {
  if (!c || !out128) return XV_EINVAL;
  NcclApi api;
  int rc = load_nccl(c, &api);
  if (rc != XV_OK) return rc;
  const int nrc = api.GetUniqueId(out128);
  if (nrc != 0) return fail(c, XV_ECUDA, "ncclGetUniqueId failed (%d)", nrc);
  return XV_OK;
}

int xv_comm_init(xv_ctx* c, const void* nccl_unique_id, int rank, int world) {
  if (!c) return XV_EINVAL;
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "xv_comm_init: call xv_train_init first");
  if (!nccl_unique_id || world < 1 || rank < 0 || rank >= world) return fail(c, XV_EINVAL, "xv_comm_init: bad argument");
  cudaSetDevice(c->device);
  t->rank = rank;
  t->world = world;
  if (world == 1) return XV_OK;
  int rc = load_nccl(c, &t->nccl);
  if (rc != XV_OK) return rc;
  UidBlob id;
  memcpy(id.bytes, nccl_unique_id, sizeof(id.bytes));
  const int nrc = t->nccl.CommInitRank(&t->comm, world, id, rank);
  if (nrc != 0) return tr_fail_nccl(c, t, nrc, "ncclCommInitRank");
  if (t->comm_bf16 && (rc = ensure(c, &t->gcomm, sizeof(__nv_bfloat16) * static_cast<size_t>(t->n_train))) != XV_OK) return rc;
  if (!t->comm_stream) {
    cudaError_t e = cudaStreamCreateWithFlags(&t->comm_stream, cudaStreamNonBlocking);
    for (int i = 0; i < kNumBuckets && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&t->ev_ready[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t->ev_done, cudaEventDisableTiming);
    if (e != cudaSuccess) return cuda_fail(c, e, "xv_comm_init: stream / event creation");
  }
  return XV_OK;
}

int xv_comm_bench(xv_ctx* c, int32_t iters, float* ms_per_iter) {
  if (!c) return XV_EINVAL;
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "xv_comm_bench: call xv_train_init first");
  if (iters < 1 || !ms_per_iter) return fail(c, XV_EINVAL, "xv_comm_bench: bad argument");
  if (t->world <= 1 || !t->comm) return fail(c, XV_ESTATE, "xv_comm_bench: no communicator (xv_comm_init with world > 1)");
  cudaSetDevice(c->device);
  // the gradient buffer is scratch between steps (every step rewrites it): reduce it in place, bucket by bucket
  cudaError_t e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
  cudaEvent_t a = nullptr, b = nullptr;
  if ((e = cudaEventCreate(&a)) != cudaSuccess || (e = cudaEventCreate(&b)) != cudaSuccess) return cuda_fail(c, e, "cudaEventCreate");
  float* G = static_cast<float*>(t->grads.p);
  cudaMemsetAsync(G, 0, sizeof(float) * (t->n_total + 1), t->comm_stream);
  cudaEventRecord(a, t->comm_stream);
  int nrc = 0;
  for (int it = 0; it < iters && nrc == 0; ++it) {
    if (t->overlap) {
      for (int k = 0; k < kNumBuckets && nrc == 0; ++k)
        if (t->bucket_hi[k] > t->bucket_lo[k]) nrc = exchange_bucket(t, k);
    } else {
      nrc = t->nccl.AllReduce(G, G, static_cast<size_t>(t->n_total + 1), 7, 0, t->comm, t->comm_stream);
    }
  }
  cudaEventRecord(b, t->comm_stream);
  int rc = XV_OK;
  if (nrc != 0) rc = tr_fail_nccl(c, t, nrc, "ncclAllReduce");
  e = cudaStreamSynchronize(t->comm_stream);
  if (rc == XV_OK && e != cudaSuccess) rc = cuda_fail(c, e, "stream synchronize");
  float ms = 0.f;
  if (rc == XV_OK && (e = cudaEventElapsedTime(&ms, a, b)) != cudaSuccess) rc = cuda_fail(c, e, "cudaEventElapsedTime");
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  *ms_per_iter = ms / iters;
  return rc;
}

int xv_comm_dtype(xv_ctx* c) {
  if (!c || !c->train) return XV_EINVAL;
  return (c->train->comm_bf16 && c->train->world > 1) ? XV_DTYPE_BF16 : XV_DTYPE_F32;
}

int xv_comm_info(xv_ctx* c, int32_t* rank, int32_t* world, int64_t* grad_elems, int32_t* buckets) {
  if (!c) return XV_EINVAL;
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "xv_comm_info: call xv_train_init first");
  if (rank) *rank = t->rank;
  if (world) *world = t->world;
  if (grad_elems) *grad_elems = t->n_total + 1;
  if (buckets) *buckets = (t->world > 1 && t->overlap) ? kNumBuckets : 1;
  return XV_OK;
}

}  // extern "C"
