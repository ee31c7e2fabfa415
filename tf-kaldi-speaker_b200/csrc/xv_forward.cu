// libxvector.so C-ABI, part 2: the forward pass.  Replaces sess.run(self.embeddings, ...) of
// Trainer.predict (model/trainer.py:886-913) for a ragged batch of utterances packed back to back.
//
// Data layout in HBM (bf16 mode):  features fp32 [rows, D] -> spliced bf16 [rows, K1pad] -> per-layer
// activations bf16 [rows, N_alloc] (row = packed frame index, same index in every layer; an utterance keeps
// its row offset and its valid length shrinks by the layer context: T-4, T-8, T-14).  The pre-pooling layer
// is never stored when statistics pooling is fused into its epilogue.  Utterance-level tensors are
// [n_utts, width] fp32 (+ a bf16 copy of the pooled vector as the tdnn6 GEMM operand).
#include <algorithm>

#include "xv_ctx.h"

namespace xv {

static int64_t round_up64(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

int plan_job(xv_ctx* c, const int32_t* lengths, int32_t n_utts, const int64_t* view_row0, int64_t view_total_rows) {
  if (n_utts < 0) return fail(c, XV_EINVAL, "n_utts must be >= 0");
  c->staged = false;
  c->ran = false;
  c->n_utts = n_utts;
  c->views = view_row0 != nullptr;
  c->view_row0.clear();
  if (c->views) {
    if (c->resnet && c->resnet_time_stride) return fail(c, XV_EUNSUPPORTED, "views are not supported with resnet_time_stride");
    for (int u = 0; u < n_utts; ++u) {
      if (view_row0[u] < 0 || view_row0[u] + lengths[u] > view_total_rows)
        return fail(c, XV_EINVAL, "view %d (rows %lld..%lld) lies outside the feature buffer of %lld rows", u, (long long)view_row0[u],
                    (long long)(view_row0[u] + lengths[u]), (long long)view_total_rows);
      if (u > 0 && view_row0[u] < view_row0[u - 1]) return fail(c, XV_EINVAL, "views must be given in ascending source-row order (view %d)", u);
    }
    c->view_row0.assign(view_row0, view_row0 + n_utts);
  }
  c->lengths.assign(lengths, lengths + n_utts);
  c->cu.assign(n_utts + 1, 0);
  c->cu_in.assign(n_utts + 1, 0);
  int64_t rows = 0, tail_rows = 0;
  for (int u = 0; u < n_utts; ++u) {
    if (lengths[u] <= c->total_context || lengths[u] < 1)
      return fail(c, XV_EINVAL, "utterance %d has %d frames; the network context needs more than %d", u, lengths[u],
                  c->total_context);
    rows += lengths[u];
    tail_rows += resnet_out_len(c, lengths[u]);     // == lengths[u] unless resnet_18 with time stride
    if (rows >= (int64_t(1) << 31) - 4096) return fail(c, XV_EINVAL, "more than 2^31 frames in one call");
    c->cu_in[u + 1] = static_cast<int32_t>(rows);
    c->cu[u + 1] = static_cast<int32_t>(tail_rows);
  }
  c->total_rows = c->views ? view_total_rows : rows;      // rows of the caller's feature buffer
  // Features that come from the host are copied batch by batch on a second stream while the previous batch computes: a
  // job of only one or two full batches (BASELINE configs[0]: 300 k frames) is cut into four so that the copy hides
  int64_t batch_frames = c->batch_frames;
  if (c->pipeline_hint && !c->resnet && rows < 4 * batch_frames)
    batch_frames = std::max<int64_t>(32768, round_up64((rows + 3) / 4, kBM));
  c->batches.clear();
  c->segs.assign(n_utts, PoolSeg{0, 0, 0, 0});
  c->tile_seg_ptr.clear();
  int u = 0;
  while (u < n_utts) {
    Batch b;
    b.utt0 = u;
    b.row0 = c->cu[u];
    b.in_row0 = c->cu_in[u];
    b.t_in = lengths[u];
    int64_t r = 0;
    // resnet_18: a batch is a run of EQUAL-length utterances (one 5-D tensor map describes [B,T,F,C])
    while (u < n_utts && (r == 0 || (r + lengths[u] <= batch_frames && (!c->resnet || lengths[u] == b.t_in)))) r += lengths[u++];
    b.utt1 = u;
    b.in_rows = r;
    if (c->views) {                      // source rows this batch's views cover: [first view's start, furthest end)
      int64_t hi = 0;
      for (int v = b.utt0; v < b.utt1; ++v) hi = std::max<int64_t>(hi, c->view_row0[v] + lengths[v]);
      b.in_row0 = c->view_row0[b.utt0];
      b.in_rows = hi - b.in_row0;
    }
    b.rows = c->cu[u] - b.row0;
    b.m_tiles = static_cast<int32_t>((b.rows + kBM - 1) / kBM);
    b.tile0 = static_cast<int32_t>(c->tile_seg_ptr.size() / 2);
    // segment table: one valid-row range per utterance (batch-local rows); per tile the [lo,hi) range of
    // utterances whose valid rows intersect the tile.
    for (int v = b.utt0; v < b.utt1; ++v) {
      PoolSeg s;
      s.utt = v;
      s.r0 = static_cast<int32_t>(c->cu[v] - b.row0);
      s.r1 = s.r0 + (c->cu[v + 1] - c->cu[v]) - c->total_context;
      s.pad = 0;
      c->segs[v] = s;
    }
    int lo = b.utt0;
    for (int t = 0; t < b.m_tiles; ++t) {
      const int t0 = t * kBM, t1 = t0 + kBM;
      while (lo < b.utt1 && c->segs[lo].r1 <= t0) ++lo;
      int hi = lo;
      while (hi < b.utt1 && c->segs[hi].r0 < t1) ++hi;
      c->tile_seg_ptr.push_back(lo);
      c->tile_seg_ptr.push_back(hi);
    }
    c->batches.push_back(b);
  }
  // upload tables through a pinned staging area (no stream synchronisation: the previous call's copies are awaited
  // only if they are still in flight when the staging area is about to be overwritten)
  int rc;
  std::vector<int32_t> cu_all(c->cu);       // d_cu = [global cu (n+1)] [batch-local cu, one (nb+1) run per batch]
  for (const Batch& b : c->batches)
    for (int v = b.utt0; v <= b.utt1; ++v) cu_all.push_back(static_cast<int32_t>(c->cu[v] - b.row0));
  std::vector<int64_t> view_src;
  if (c->views) {
    view_src.resize(n_utts);
    for (const Batch& b : c->batches)
      for (int v = b.utt0; v < b.utt1; ++v) view_src[v] = c->view_row0[v] - b.in_row0;
  }
  const size_t cu_bytes = cu_all.size() * sizeof(int32_t);
  const size_t seg_bytes = sizeof(PoolSeg) * static_cast<size_t>(n_utts);
  const size_t tsp_bytes = sizeof(int32_t) * c->tile_seg_ptr.size();
  const size_t vs_bytes = sizeof(int64_t) * view_src.size();
  auto up16 = [](size_t v) { return (v + 15) / 16 * 16; };
  const size_t off_seg = up16(cu_bytes), off_tsp = off_seg + up16(seg_bytes), off_vs = off_tsp + up16(tsp_bytes);
  const size_t total = off_vs + up16(vs_bytes) + 16;
  if ((rc = ensure(c, &c->d_cu, std::max<size_t>(cu_bytes, 16))) != XV_OK) return rc;
  if ((rc = ensure(c, &c->d_segs, std::max<size_t>(seg_bytes, sizeof(PoolSeg)))) != XV_OK) return rc;
  if ((rc = ensure(c, &c->d_tile_seg_ptr, std::max<size_t>(tsp_bytes, 8))) != XV_OK) return rc;
  if (c->views && (rc = ensure(c, &c->d_view_src, std::max<size_t>(vs_bytes, 8))) != XV_OK) return rc;
  cudaError_t e = cudaSuccess;
  if (c->tables_ev == nullptr) e = cudaEventCreateWithFlags(&c->tables_ev, cudaEventDisableTiming);
  else e = cudaEventSynchronize(c->tables_ev);
  if (e != cudaSuccess) return cuda_fail(c, e, "table staging event");
  if (c->h_tables_bytes < total) {
    if (c->h_tables) cudaFreeHost(c->h_tables);
    c->h_tables = nullptr;
    c->h_tables_bytes = 0;
    if ((e = cudaMallocHost(&c->h_tables, total * 2)) != cudaSuccess) return cuda_fail(c, e, "cudaMallocHost(tables)");
    c->h_tables_bytes = total * 2;
  }
  char* h = static_cast<char*>(c->h_tables);
  memcpy(h, cu_all.data(), cu_bytes);
  if (seg_bytes) memcpy(h + off_seg, c->segs.data(), seg_bytes);
  if (tsp_bytes) memcpy(h + off_tsp, c->tile_seg_ptr.data(), tsp_bytes);
  if (vs_bytes) memcpy(h + off_vs, view_src.data(), vs_bytes);
  e = cudaMemcpyAsync(c->d_cu.p, h, cu_bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && seg_bytes) e = cudaMemcpyAsync(c->d_segs.p, h + off_seg, seg_bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && tsp_bytes) e = cudaMemcpyAsync(c->d_tile_seg_ptr.p, h + off_tsp, tsp_bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && vs_bytes) e = cudaMemcpyAsync(c->d_view_src.p, h + off_vs, vs_bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaEventRecord(c->tables_ev, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "upload of length tables");
  return XV_OK;
}

// Attention fusions of a layer launch: softmax-weighted pooled sums (value layer) or score accumulation (last key layer).
struct GemmExtras {
  const float* row_weight = nullptr;
  int heads = 1, head_width = 0, weight_ld = 0;
  const float* dot_q = nullptr;
  double* dot_out = nullptr;
  int dot_ld = 0;
  float dot_scale = 1.f;
};

// ---- one tcgen05 layer launch -------------------------------------------------------------------
static int gemm_layer(xv_ctx* c, const Layer& l, const void* a_base, int64_t a_rows_alloc, int stage,
                      __nv_bfloat16* out_bf16, float* out_f32, int64_t m_rows, bool pool, double* pool_acc,
                      const int32_t* tile_seg, const PoolSeg* segs, const GemmExtras* ex = nullptr) {
  if (m_rows <= 0) return XV_OK;
  GemmPlan g;
  const char* err = make_tmap_bf16_2d(&g.tm_a, a_base, l.a_cin, a_rows_alloc, static_cast<uint64_t>(l.a_cin) * 2, kBK, kBM);
  if (!err && out_bf16 != nullptr)
    err = make_tmap_bf16_2d(&g.tm_out, out_bf16, l.n_alloc, m_rows, static_cast<uint64_t>(l.n_alloc) * 2, 32, 32, 64);
  if (err) return fail(c, XV_ECUDA, "%s", err);
  if (out_bf16 == nullptr) g.tm_out = g.tm_a;   // never dereferenced
  g.tm_b = l.tm_b;
  GemmParams& p = g.p;
  p.m_rows = static_cast<int32_t>(m_rows);
  p.m_tiles = static_cast<int32_t>((m_rows + kBM - 1) / kBM);
  p.n_tiles = l.n_alloc / kBN;
  p.taps = l.a_taps;
  p.dilation = l.dil;
  p.kb_per_tap = l.a_cin / kBK;
  p.cin = l.a_cin;
  p.scale = static_cast<const float*>(l.f_scale[stage].p);
  p.shift = static_cast<const float*>(l.f_shift[stage].p);
  p.slope = static_cast<const float*>(l.f_slope[stage].p);
  p.act_tanh = (l.act == kActTanh && stage == kStageAct) ? 1 : 0;
  p.relu_only = (stage == kStageAct && l.plain_relu) ? 1 : 0;
  p.out_bf16 = out_bf16;
  p.out_f32 = out_f32;
  p.ldo = l.n_alloc;
  p.pool_acc = pool_acc;
  p.pool_ld = l.n_alloc;
  p.tile_seg_ptr = tile_seg;
  p.segs = segs;
  p.row_weight = nullptr;
  p.heads = 1;
  p.head_width = l.n_alloc;
  if (ex != nullptr) {
    p.row_weight = ex->row_weight; p.heads = ex->heads; p.head_width = ex->head_width; p.weight_ld = ex->weight_ld;
    p.dot_q = ex->dot_q; p.dot_out = ex->dot_out; p.dot_ld = ex->dot_ld; p.dot_scale = ex->dot_scale;
  }
  g.pool = pool;
  const std::string fam = std::string(pool ? "gemm_pool." : "gemm.") + l.name;
  ProfScope ps(c, fam.c_str());
  cudaError_t e = launch_gemm(g, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "tcgen05 GEMM launch");
  return XV_OK;
}

static int simt_layer(xv_ctx* c, const Layer& l, const float* a, int64_t a_rows, int stage, float* out, int64_t m_rows) {
  SimtGemmArgs s;
  s.a = a; s.a_rows = a_rows; s.cin = l.cin; s.taps = l.taps; s.dilation = l.dil;
  s.w = static_cast<const float*>(l.w32.p); s.n = l.cout;
  s.bias = static_cast<const float*>(l.bias.p);
  s.bn_scale = (l.has_bn && stage >= kStageBn) ? static_cast<const float*>(l.bn_scale.p) : nullptr;
  s.bn_shift = static_cast<const float*>(l.bn_shift.p);
  s.slope = (stage == kStageAct && l.act == kActRelu) ? static_cast<const float*>(l.slope_raw.p) : nullptr;
  s.act_tanh = (stage == kStageAct && l.act == kActTanh) ? 1 : 0;
  s.out_dense = nullptr; s.out_bn = nullptr; s.out_act = out;   // exactly one stage is materialised
  s.m_rows = m_rows;
  ProfScope ps(c, "simt_gemm");
  cudaError_t e = launch_simt_gemm(s, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "SIMT GEMM launch");
  return XV_OK;
}

static int final_stage(const Layer& l) { return l.act != kActNone ? kStageAct : (l.has_bn ? kStageBn : kStageDense); }

// Copies utterance-level rows to the caller.
static int fetch_result(xv_ctx* c, float* out, int64_t cap) {
  if (!c->ran || !c->result) return fail(c, XV_ESTATE, "xv_fetch: no utterance-level result (call xv_run first)");
  const int64_t need = static_cast<int64_t>(c->n_utts) * c->result_width;
  if (cap < need) return fail(c, XV_ECAPACITY, "output buffer holds %lld floats, need %lld", (long long)cap, (long long)need);
  if (c->n_utts > 0) {
    cudaError_t e = cudaMemcpy2DAsync(out, sizeof(float) * c->result_width, c->result, sizeof(float) * c->result_ld,
                                      sizeof(float) * c->result_width, c->n_utts, cudaMemcpyDeviceToHost, c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "D2H of the result");
  }
  cudaError_t e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
  return XV_OK;
}

int run_job(xv_ctx* c, const char* node_name, bool pipelined_h2d, float* frames_out, const int64_t* frames_off,
            int64_t frames_cap) {
  if (!c->finalized) return fail(c, XV_ESTATE, "weights are not finalized (call xv_finalize_weights)");
  const std::string name = (node_name && node_name[0]) ? node_name : c->embedding_node;
  auto it = c->nodes.find(name);
  if (it == c->nodes.end()) return fail(c, XV_ENOTFOUND, "unknown node '%s'", name.c_str());
  const NodeInfo ni = it->second;
  const bool utt_level = ni.level == 1;
  if (utt_level && frames_out) return fail(c, XV_EINVAL, "node '%s' is utterance level; use xv_forward", name.c_str());
  if (!utt_level && !frames_out) return fail(c, XV_EINVAL, "node '%s' is frame level; use xv_forward_frames", name.c_str());
  const bool attention = c->pooling_type == "self_attention";
  const int n = c->n_utts;
  c->ran = false;
  c->result = nullptr;
  int rc;

  // ---- which frame layers run, and at which stage ----
  const int nl = static_cast<int>(c->frame_layers.size());
  std::vector<int> stage_of(nl, -1);
  auto need_chain = [&](int layer, int stage) {
    int l = layer, s = stage;
    while (l >= 0) {
      if (stage_of[l] < 0 || l == layer) stage_of[l] = s;
      l = c->frame_layers[l].in_buf;
      if (l >= 0) s = final_stage(c->frame_layers[l]);
    }
  };
  int target_layer = -1;
  if (ni.level == 0) {
    target_layer = ni.layer;
    need_chain(ni.layer, ni.stage);
  } else {
    need_chain(c->n_main_frame - 1, kStageAct);
    if (attention) {
      need_chain(c->key_layers.back(), final_stage(c->frame_layers[c->key_layers.back()]));
      if (!c->value_layers.empty())
        need_chain(c->value_layers.back(), final_stage(c->frame_layers[c->value_layers.back()]));
      need_chain(c->att_value_buf, kStageAct);
    }
  }
  // A layer that feeds another layer must be at its final stage; the target layer is a leaf unless it also
  // feeds something (only possible for frame-level requests, where it is the last of its chain).
  const int pool_layer = c->n_main_frame - 1;
  // Attention pooling fuses into the value layer's epilogue when the values are the last main layer itself (no value
  // network), every head owns a slice of its columns (or there is one head), and the keys do not read that layer:
  // tdnn1-4 -> key layers -> scores -> softmax -> tdnn5 with softmax-weighted sums in the epilogue, never stored.
  bool att_fusable = attention && c->value_layers.empty() && c->att_value_buf == pool_layer &&
                     (c->att_split_value || c->att_heads == 1) && c->att_value_dim / c->att_heads >= 32;
  if (att_fusable) {
    static const bool no_fuse = getenv("XV_ATT_POOL_IMPL") != nullptr && getenv("XV_ATT_POOL_IMPL")[0] == 's';   // 's'tandalone
    int l = c->key_layers.front();
    while (l >= 0 && l != pool_layer) l = c->frame_layers[l].in_buf;
    if (l == pool_layer || no_fuse) att_fusable = false;
  }
  const bool fuse_pool = c->bf16 && utt_level && (!attention || att_fusable);    // tdnn5 never stored
  const bool fuse_att = fuse_pool && attention;
  // ... and the scores fuse into the last key layer's epilogue (its output is then never stored) when every head owns a
  // slice of the key columns (or there is one head) that is at least half a tile wide.
  const int key_last = attention ? c->key_layers.back() : -1;
  const bool fuse_scores = fuse_att && (c->att_split_key || c->att_heads == 1) && c->att_dk >= 128 &&
                           c->att_heads * c->att_dk <= c->frame_layers[key_last].n_alloc &&
                           !(getenv("XV_ATT_SCORES_IMPL") != nullptr && getenv("XV_ATT_SCORES_IMPL")[0] == 's');

  // ---- buffers ----
  int64_t max_rows = 0, max_in_rows = 0;
  for (const Batch& b : c->batches) {
    max_rows = std::max(max_rows, b.rows);
    max_in_rows = std::max(max_in_rows, b.in_rows);
  }
  const int64_t rows_alloc = round_up64(std::max<int64_t>(max_rows, 1), kBM);
  const size_t esz = c->bf16 ? 2 : 4;
  for (int l = 0; l < nl; ++l) {
    if (stage_of[l] < 0) continue;
    if (fuse_pool && l == pool_layer) continue;
    if (fuse_scores && l == key_last) continue;
    const int w = c->bf16 ? c->frame_layers[l].n_alloc : c->frame_layers[l].cout;
    if ((rc = ensure(c, &c->frame_out[l], static_cast<size_t>(rows_alloc) * w * esz, true)) != XV_OK) return rc;
  }
  const Layer& l0 = c->frame_layers[0];
  if (c->bf16 && !c->resnet && (rc = ensure(c, &c->a1, static_cast<size_t>(rows_alloc) * l0.a_cin * 2, true)) != XV_OK) return rc;
  if (pipelined_h2d)
    for (int s = 0; s < 2; ++s)
      if ((rc = ensure(c, &c->d_feat_ring[s], static_cast<size_t>(max_in_rows) * c->dim * 4)) != XV_OK) return rc;
  const Layer& l6 = c->seg_layers[0];
  const Layer& l7 = c->seg_layers[1];
  const int64_t n_alloc_rows = round_up64(std::max(n, 1), kBM);
  if (utt_level || ni.level == 2) {
    if ((rc = ensure(c, &c->pooled_f32, static_cast<size_t>(n_alloc_rows) * c->pooled_dim * 4)) != XV_OK) return rc;
    if ((rc = ensure(c, &c->pooled_pre_f32, static_cast<size_t>(n_alloc_rows) * c->pooled_dim * 4)) != XV_OK) return rc;
    if ((rc = ensure(c, &c->pooled_bf16, static_cast<size_t>(n_alloc_rows) * l6.kpad * 2, true)) != XV_OK) return rc;
    if ((rc = ensure(c, &c->seg_out[0], static_cast<size_t>(n_alloc_rows) * l6.n_alloc * 4)) != XV_OK) return rc;
    if ((rc = ensure(c, &c->seg_out[1], static_cast<size_t>(n_alloc_rows) * l7.n_alloc * 4)) != XV_OK) return rc;
    if ((rc = ensure(c, &c->seg_out_bf16, static_cast<size_t>(n_alloc_rows) * l7.kpad * 2, true)) != XV_OK) return rc;
    if (fuse_pool) {
      const size_t bytes = static_cast<size_t>(std::max(n, 1)) * 2 * c->frame_layers[pool_layer].n_alloc * sizeof(double);
      if ((rc = ensure(c, &c->pool_acc, bytes)) != XV_OK) return rc;
      cudaError_t e = cudaMemsetAsync(c->pool_acc.p, 0, bytes, c->stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "cudaMemsetAsync(pool_acc)");
    } else {
      const int hout = (attention && !c->att_split_value) ? c->att_heads : 1;
      const int cval = attention ? c->att_value_dim : c->pool_nodes;
      int max_nb = 1;
      for (const Batch& b : c->batches) max_nb = std::max(max_nb, b.utt1 - b.utt0);
      if ((rc = ensure(c, &c->pool_ws, pool_workspace_bytes(max_nb, hout, cval))) != XV_OK) return rc;
    }
  }
  if (attention && (utt_level || ni.level == 2))
    if ((rc = ensure(c, &c->scores, static_cast<size_t>(c->att_heads) * rows_alloc * 4)) != XV_OK) return rc;

  if (fuse_scores) {
    const int kn = c->frame_layers[key_last].n_alloc;
    if ((rc = ensure(c, &c->scores_acc, static_cast<size_t>(c->att_heads) * rows_alloc * sizeof(double))) != XV_OK) return rc;
    if ((rc = ensure(c, &c->att_query_pad, sizeof(float) * kn)) != XV_OK) return rc;
    cudaError_t qe = cudaMemsetAsync(c->att_query_pad.p, 0, sizeof(float) * kn, c->stream);   // refreshed per call: weights may change
    if (qe == cudaSuccess)
      qe = cudaMemcpyAsync(c->att_query_pad.p, c->att_query.p, sizeof(float) * c->att_heads * c->att_dk, cudaMemcpyDeviceToDevice,
                           c->stream);
    if (qe != cudaSuccess) return cuda_fail(c, qe, "query copy");
  }
  const int32_t* d_cu_global = static_cast<const int32_t*>(c->d_cu.p);
  const int32_t* d_cu_local = d_cu_global + (n + 1);
  size_t cu_local_off = 0;
  cudaError_t e;

  for (size_t bi = 0; bi < c->batches.size(); ++bi) {
    const Batch& b = c->batches[bi];
    const int nb = b.utt1 - b.utt0;
    const int32_t* cu_b = d_cu_local + cu_local_off;
    cu_local_off += nb + 1;
    // ---- features of this batch ----
    const float* feats_b;
    const int slot = static_cast<int>(bi & 1);
    if (pipelined_h2d) {
      const float* src = c->host_feats + b.in_row0 * c->dim;
      const size_t bytes = static_cast<size_t>(b.in_rows) * c->dim * 4;
      if (bi >= 2 && (e = cudaStreamWaitEvent(c->copy_stream, c->buf_free[slot], 0)) != cudaSuccess)
        return cuda_fail(c, e, "cudaStreamWaitEvent");
      const void* from = src;
      if (!c->host_pinned) {   // pageable host memory: bounce through a pinned slot
        if (c->h_pinned_bytes[slot] < bytes) {
          if (c->h_pinned[slot]) { cudaEventSynchronize(c->h2d_done[slot]); cudaFreeHost(c->h_pinned[slot]); }
          c->h_pinned[slot] = nullptr;
          c->h_pinned_bytes[slot] = 0;
          if ((e = cudaMallocHost(&c->h_pinned[slot], bytes)) != cudaSuccess) return cuda_fail(c, e, "cudaMallocHost");
          c->h_pinned_bytes[slot] = bytes;
        } else if (bi >= 2) {
          if ((e = cudaEventSynchronize(c->h2d_done[slot])) != cudaSuccess) return cuda_fail(c, e, "cudaEventSynchronize");
        }
        memcpy(c->h_pinned[slot], src, bytes);
        from = c->h_pinned[slot];
      }
      if ((e = cudaMemcpyAsync(c->d_feat_ring[slot].p, from, bytes, cudaMemcpyHostToDevice, c->copy_stream)) != cudaSuccess)
        return cuda_fail(c, e, "H2D of features");
      if ((e = cudaEventRecord(c->h2d_done[slot], c->copy_stream)) != cudaSuccess) return cuda_fail(c, e, "cudaEventRecord");
      if ((e = cudaStreamWaitEvent(c->stream, c->h2d_done[slot], 0)) != cudaSuccess) return cuda_fail(c, e, "cudaStreamWaitEvent");
      feats_b = static_cast<const float*>(c->d_feat_ring[slot].p);
    } else {
      feats_b = static_cast<const float*>(c->d_feats.p) + b.in_row0 * c->dim;
    }
    if (c->views) {
      // the batch's views (possibly overlapping source rows) packed back to back: one gather pass over the features
      if ((rc = ensure(c, &c->d_feat_gather, static_cast<size_t>(std::max<int64_t>(max_rows, 1)) * c->dim * 4)) != XV_OK) return rc;
      count_launch(c, "gather_views");
      e = launch_gather_views(feats_b, cu_b, static_cast<const int64_t*>(c->d_view_src.p) + b.utt0, nb, b.rows, c->dim,
                              static_cast<float*>(c->d_feat_gather.p), c->stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "gather_views launch");
      feats_b = static_cast<const float*>(c->d_feat_gather.p);
    }
    // ---- frame layers ----
    const void* a_first = nullptr;          // operand of the first frame layer
    if (c->resnet) {
      // conv0_1 .. conv4b: [nb, T, F] -> [nb*T', F'*512]; conv5 then reads these rows as a plain GEMM operand
      if ((rc = resnet_run_batch(c, feats_b, nb, b.t_in, &a_first)) != XV_OK) return rc;
    } else if (c->bf16) {
      ProfScope ps(c, "splice_cast");
      // fp32 -> bf16 rows padded to a_cin channels (taps > 1: the context splice happens in tdnn1's TMA loads)
      if (l0.a_taps > 1)
        e = launch_splice_cast(feats_b, b.rows, c->dim, 1, 1, static_cast<__nv_bfloat16*>(c->a1.p), l0.a_cin, c->stream);
      else
        e = launch_splice_cast(feats_b, b.rows, c->dim, l0.taps, l0.dil, static_cast<__nv_bfloat16*>(c->a1.p), l0.kpad, c->stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "splice_cast launch");
      a_first = c->a1.p;
    } else {
      a_first = feats_b;
    }
    for (int l = 0; l < nl; ++l) {
      if (stage_of[l] < 0) continue;
      if (fuse_att && l == pool_layer) continue;             // runs after the attention weights exist (below)
      const Layer& L = c->frame_layers[l];
      const bool pool_here = fuse_pool && l == pool_layer;
      if (fuse_scores && l == key_last) {
        GemmExtras ex;
        ex.heads = c->att_heads; ex.head_width = c->att_dk;
        ex.dot_q = static_cast<const float*>(c->att_query_pad.p); ex.dot_out = static_cast<double*>(c->scores_acc.p);
        ex.dot_ld = static_cast<int>(b.rows);
        ex.dot_scale = c->att_use_scale ? 1.0f / sqrtf(static_cast<float>(c->att_dk)) : 1.0f;
        if ((e = cudaMemsetAsync(c->scores_acc.p, 0, sizeof(double) * c->att_heads * b.rows, c->stream)) != cudaSuccess)
          return cuda_fail(c, e, "cudaMemsetAsync(scores)");
        rc = gemm_layer(c, L, c->frame_out[L.in_buf].p, rows_alloc, stage_of[l], nullptr, nullptr, b.rows, false, nullptr, nullptr,
                        nullptr, &ex);
      } else if (c->bf16) {
        const void* a = L.in_buf < 0 ? a_first : c->frame_out[L.in_buf].p;
        const int64_t a_rows = (L.in_buf < 0 && c->resnet) ? b.rows : rows_alloc;
        rc = gemm_layer(c, L, a, a_rows, stage_of[l], pool_here ? nullptr : static_cast<__nv_bfloat16*>(c->frame_out[l].p),
                        nullptr, b.rows, pool_here, static_cast<double*>(c->pool_acc.p),
                        static_cast<const int32_t*>(c->d_tile_seg_ptr.p) + 2 * b.tile0, static_cast<const PoolSeg*>(c->d_segs.p));
      } else {
        const float* a = L.in_buf < 0 ? static_cast<const float*>(a_first) : static_cast<const float*>(c->frame_out[L.in_buf].p);
        rc = simt_layer(c, L, a, b.rows, stage_of[l], static_cast<float*>(c->frame_out[l].p), b.rows);
      }
      if (rc != XV_OK) return rc;
    }
    // ---- pooling of this batch ----
    const bool need_pool = utt_level || ni.level == 2;
    if (fuse_att) {
      const int kl = c->key_layers.back();
      const Layer& KL = c->frame_layers[kl];
      const Layer& VL = c->frame_layers[pool_layer];
      const float scale = c->att_use_scale ? 1.0f / sqrtf(static_cast<float>(c->att_dk)) : 1.0f;
      if (!fuse_scores) {
        ProfScope ps(c, "att_scores");
        e = launch_att_scores(c->frame_out[kl].p, XV_DTYPE_BF16, KL.n_alloc, b.rows, static_cast<const float*>(c->att_query.p),
                              c->att_heads, c->att_dk, c->att_split_key ? 1 : 0, scale, static_cast<float*>(c->scores.p), c->stream);
        if (e != cudaSuccess) return cuda_fail(c, e, "att_scores launch");
      }
      {
        ProfScope ps(c, "att_softmax");
        if (fuse_scores)
          e = launch_att_softmax_f64(static_cast<const double*>(c->scores_acc.p), static_cast<float*>(c->scores.p), b.rows, cu_b,
                                     c->total_context, nb, c->att_heads, c->stream);
        else
          e = launch_att_softmax(static_cast<float*>(c->scores.p), b.rows, cu_b, c->total_context, nb, c->att_heads, c->stream);
        if (e != cudaSuccess) return cuda_fail(c, e, "att_softmax launch");
      }
      const void* a = VL.in_buf < 0 ? a_first : c->frame_out[VL.in_buf].p;
      GemmExtras ex;
      ex.row_weight = static_cast<const float*>(c->scores.p); ex.heads = c->att_heads;
      ex.head_width = c->att_value_dim / c->att_heads; ex.weight_ld = static_cast<int>(b.rows);
      rc = gemm_layer(c, VL, a, rows_alloc, stage_of[pool_layer], nullptr, nullptr, b.rows, true, static_cast<double*>(c->pool_acc.p),
                      static_cast<const int32_t*>(c->d_tile_seg_ptr.p) + 2 * b.tile0, static_cast<const PoolSeg*>(c->d_segs.p), &ex);
      if (rc != XV_OK) return rc;
    }
    if (need_pool && !fuse_pool) {
      const int dtype = c->bf16 ? XV_DTYPE_BF16 : XV_DTYPE_F32;
      PoolArgs pa;
      pa.dtype = dtype;
      pa.rows = b.rows; pa.cu = cu_b; pa.n = nb; pa.context = c->total_context;
      pa.acc = static_cast<float*>(c->pool_ws.p);
      float* dst = static_cast<float*>(c->has_att_post ? c->pooled_pre_f32.p : c->pooled_f32.p);
      pa.out = dst + static_cast<size_t>(b.utt0) * c->pooled_dim;
      pa.ldout = c->pooled_dim;
      pa.out_bf16 = (c->bf16 && !c->has_att_post) ? static_cast<__nv_bfloat16*>(c->pooled_bf16.p) + static_cast<size_t>(b.utt0) * l6.kpad : nullptr;
      pa.ldout_bf16 = l6.kpad;
      if (!attention) {
        const Layer& P = c->frame_layers[pool_layer];
        pa.x = c->frame_out[pool_layer].p; pa.c = P.cout; pa.ldx = c->bf16 ? P.n_alloc : P.cout;
        pa.weights = nullptr; pa.heads = 1; pa.split_value = 1;
      } else {
        const int kl = c->key_layers.back();
        const Layer& KL = c->frame_layers[kl];
        const int vl = c->value_layers.empty() ? c->att_value_buf : c->value_layers.back();
        const Layer& VL = c->frame_layers[vl];
        const float scale = c->att_use_scale ? 1.0f / sqrtf(static_cast<float>(c->att_dk)) : 1.0f;
        {
          ProfScope ps(c, "att_scores");
          e = launch_att_scores(c->frame_out[kl].p, dtype, c->bf16 ? KL.n_alloc : KL.cout, b.rows,
                                static_cast<const float*>(c->att_query.p), c->att_heads, c->att_dk, c->att_split_key ? 1 : 0,
                                scale, static_cast<float*>(c->scores.p), c->stream);
          if (e != cudaSuccess) return cuda_fail(c, e, "att_scores launch");
        }
        {
          ProfScope ps(c, "att_softmax");
          e = launch_att_softmax(static_cast<float*>(c->scores.p), b.rows, cu_b, c->total_context, nb, c->att_heads, c->stream);
          if (e != cudaSuccess) return cuda_fail(c, e, "att_softmax launch");
        }
        pa.x = c->frame_out[vl].p; pa.c = VL.cout; pa.ldx = c->bf16 ? VL.n_alloc : VL.cout;
        pa.weights = static_cast<const float*>(c->scores.p); pa.heads = c->att_heads; pa.split_value = c->att_split_value ? 1 : 0;
      }
      if (utt_level) {
        ProfScope ps(c, "pool");
        count_launch(c, "pool_finalize");
        e = launch_pool(pa, c->sm_count, c->stream);
        if (e != cudaSuccess) return cuda_fail(c, e, "pooling launch");
      }
    }
    // ---- frame-level outputs of this batch ----
    if (!utt_level) {
      const int width = ni.level == 2 ? c->att_heads : ni.width;
      std::vector<int64_t> meta(3 * static_cast<size_t>(nb));   // src_row0 | dst_off | n_rows (int32 packed in int64 slots)
      std::vector<int32_t> nrows(nb);
      int64_t total = 0;
      int max_valid = 0;
      for (int v = 0; v < nb; ++v) {
        const int u = b.utt0 + v;
        const int valid = c->lengths[u] - ni.context;
        meta[v] = c->cu[u] - b.row0;
        meta[nb + v] = total;
        nrows[v] = valid;
        total += static_cast<int64_t>(valid) * width;
        max_valid = std::max(max_valid, valid);
      }
      if ((rc = ensure(c, &c->d_gather_meta, sizeof(int64_t) * 2 * nb + sizeof(int32_t) * nb)) != XV_OK) return rc;
      if ((rc = ensure(c, &c->d_gather_out, sizeof(float) * static_cast<size_t>(total))) != XV_OK) return rc;
      int64_t* d_meta = static_cast<int64_t*>(c->d_gather_meta.p);
      int32_t* d_nrows = reinterpret_cast<int32_t*>(d_meta + 2 * nb);
      e = cudaMemcpyAsync(d_meta, meta.data(), sizeof(int64_t) * 2 * nb, cudaMemcpyHostToDevice, c->stream);
      if (e == cudaSuccess) e = cudaMemcpyAsync(d_nrows, nrows.data(), sizeof(int32_t) * nb, cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "gather metadata upload");
      if (ni.level == 2) {
        count_launch(c, "gather");
        e = launch_gather_att_weights(static_cast<const float*>(c->scores.p), b.rows, cu_b, c->total_context, c->att_heads, nb,
                                      d_meta + nb, static_cast<float*>(c->d_gather_out.p), c->stream);
      } else {
        const Layer& T = c->frame_layers[target_layer];
        count_launch(c, "gather");
        e = launch_gather_rows(c->frame_out[target_layer].p, c->bf16 ? XV_DTYPE_BF16 : XV_DTYPE_F32,
                               c->bf16 ? T.n_alloc : T.cout, T.cout, d_meta, d_nrows, d_meta + nb, nb, max_valid,
                               static_cast<float*>(c->d_gather_out.p), c->stream);
      }
      if (e != cudaSuccess) return cuda_fail(c, e, "gather launch");
      // the metadata vectors are pageable: make sure the async copies have consumed them
      if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
      for (int v = 0; v < nb; ++v) {
        const int u = b.utt0 + v;
        const int64_t cnt = static_cast<int64_t>(nrows[v]) * width;
        if (frames_off[u] < 0 || frames_off[u] + cnt > frames_cap)
          return fail(c, XV_ECAPACITY, "frame output of utterance %d does not fit the output buffer", u);
        e = cudaMemcpyAsync(frames_out + frames_off[u], static_cast<const float*>(c->d_gather_out.p) + meta[nb + v],
                            sizeof(float) * cnt, cudaMemcpyDeviceToHost, c->stream);
        if (e != cudaSuccess) return cuda_fail(c, e, "D2H of frame outputs");
      }
      if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
    }
    if (pipelined_h2d && (e = cudaEventRecord(c->buf_free[slot], c->stream)) != cudaSuccess)
      return cuda_fail(c, e, "cudaEventRecord");
  }
  if (!utt_level) {
    c->ran = true;
    return XV_OK;
  }

  // ---- utterance level: finalize pooling, segment affines ----
  if (fuse_pool) {
    const Layer& P = c->frame_layers[pool_layer];
    count_launch(c, "pool_finalize");
    e = launch_pool_finalize(static_cast<const double*>(c->pool_acc.p), P.n_alloc, d_cu_global, c->total_context, n, P.cout,
                             static_cast<float*>(c->has_att_post ? c->pooled_pre_f32.p : c->pooled_f32.p), c->pooled_dim,
                             c->has_att_post ? nullptr : static_cast<__nv_bfloat16*>(c->pooled_bf16.p), l6.kpad, fuse_att ? 1 : 0,
                             c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "pool finalize launch");
  }
  if (c->has_att_post && ni.kind != 2 && ni.kind != 3) {
    count_launch(c, "affine");
    e = launch_affine_rows(static_cast<const float*>(c->pooled_pre_f32.p), c->pooled_dim, n, c->pooled_dim,
                           static_cast<const float*>(c->att_post_scale.p), static_cast<const float*>(c->att_post_shift.p),
                           static_cast<const float*>(c->att_post_slope.p), static_cast<float*>(c->pooled_f32.p), c->pooled_dim,
                           c->bf16 ? static_cast<__nv_bfloat16*>(c->pooled_bf16.p) : nullptr, l6.kpad, c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "att_post affine launch");
  }
  auto set_result = [&](const void* p, int ld, int width) {
    c->result = static_cast<const float*>(p);
    c->result_ld = ld;
    c->result_width = width;
    c->ran = true;
  };
  if (ni.kind == 1 || ni.kind == 4) { set_result(c->pooled_f32.p, c->pooled_dim, c->pooled_dim); return XV_OK; }
  if (ni.kind == 2) { set_result(c->has_att_post ? c->pooled_pre_f32.p : c->pooled_f32.p, c->pooled_dim, c->pooled_dim); return XV_OK; }
  if (ni.kind == 3) {
    count_launch(c, "affine");
    e = launch_affine_rows(static_cast<const float*>(c->pooled_pre_f32.p), c->pooled_dim, n, c->pooled_dim,
                           static_cast<const float*>(c->att_post_scale.p), static_cast<const float*>(c->att_post_shift.p),
                           nullptr, static_cast<float*>(c->seg_out[0].p), c->pooled_dim, nullptr, 0, c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "att_post_bn launch");
    set_result(c->seg_out[0].p, c->pooled_dim, c->pooled_dim);
    return XV_OK;
  }
  // layer outputs: tdnn6 (layer 0) or tdnn7 (layer 1; needs tdnn6 at its activation)
  const int want_layer = ni.layer;
  const int want_stage = ni.stage;
  if (c->bf16) {
    if (want_layer == 0) {
      rc = gemm_layer(c, l6, c->pooled_bf16.p, n_alloc_rows, want_stage, nullptr, static_cast<float*>(c->seg_out[0].p), n, false,
                      nullptr, nullptr, nullptr);
      if (rc != XV_OK) return rc;
      set_result(c->seg_out[0].p, l6.n_alloc, l6.cout);
    } else {
      rc = gemm_layer(c, l6, c->pooled_bf16.p, n_alloc_rows, kStageAct, static_cast<__nv_bfloat16*>(c->seg_out_bf16.p), nullptr, n,
                      false, nullptr, nullptr, nullptr);
      if (rc != XV_OK) return rc;
      rc = gemm_layer(c, l7, c->seg_out_bf16.p, n_alloc_rows, want_stage, nullptr, static_cast<float*>(c->seg_out[1].p), n, false,
                      nullptr, nullptr, nullptr);
      if (rc != XV_OK) return rc;
      set_result(c->seg_out[1].p, l7.n_alloc, l7.cout);
    }
  } else {
    if (want_layer == 0) {
      rc = simt_layer(c, l6, static_cast<const float*>(c->pooled_f32.p), n, want_stage, static_cast<float*>(c->seg_out[0].p), n);
      if (rc != XV_OK) return rc;
      set_result(c->seg_out[0].p, l6.cout, l6.cout);
    } else {
      rc = simt_layer(c, l6, static_cast<const float*>(c->pooled_f32.p), n, kStageAct, static_cast<float*>(c->seg_out[0].p), n);
      if (rc != XV_OK) return rc;
      rc = simt_layer(c, l7, static_cast<const float*>(c->seg_out[0].p), n, want_stage, static_cast<float*>(c->seg_out[1].p), n);
      if (rc != XV_OK) return rc;
      set_result(c->seg_out[1].p, l7.cout, l7.cout);
    }
  }
  if (ni.kind == 5 && c->feature_norm) {
    count_launch(c, "l2_scale");
    e = launch_l2_scale_rows(const_cast<float*>(c->result), c->result_ld, n, c->result_width, c->feature_scaling_factor, 1e-12f, c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "l2_scaling launch");
  }
  return XV_OK;
}

static int classify_pointer(const void* p) {   // 0 pageable host, 1 pinned host, 2 device/managed
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return 2;
  if (a.type == cudaMemoryTypeHost) return 1;
  return 0;
}

}  // namespace xv

using namespace xv;

extern "C" {

static int stage_common(xv_ctx* c, const float* feats, const int32_t* lengths, int32_t n_utts, const int64_t* view_row0,
                        int64_t total_rows) {
  if (!c) return XV_EINVAL;
  if (!feats || !lengths) return fail(c, XV_EINVAL, "xv_stage: null argument");
  cudaSetDevice(c->device);
  int rc = plan_job(c, lengths, n_utts, view_row0, total_rows);
  if (rc != XV_OK) return rc;
  const size_t bytes = static_cast<size_t>(c->total_rows) * c->dim * 4;
  if ((rc = ensure(c, &c->d_feats, bytes)) != XV_OK) return rc;
  const int kind = classify_pointer(feats);
  cudaError_t e = cudaMemcpyAsync(c->d_feats.p, feats, bytes, kind == 2 ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                  c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "H2D of features");
  c->host_feats = nullptr;
  c->staged = true;
  return XV_OK;
}

int xv_stage(xv_ctx* c, const float* feats, const int32_t* lengths, int32_t n_utts) {
  return stage_common(c, feats, lengths, n_utts, nullptr, 0);
}

int xv_stage_views(xv_ctx* c, const float* feats, int64_t total_rows, const int64_t* view_row0, const int32_t* lengths,
                   int32_t n_views) {
  if (c && !view_row0) return fail(c, XV_EINVAL, "xv_stage_views: null argument");
  return stage_common(c, feats, lengths, n_views, view_row0, total_rows);
}

int xv_run(xv_ctx* c, const char* node) {
  if (!c) return XV_EINVAL;
  if (!c->staged) return fail(c, XV_ESTATE, "xv_run: nothing staged (call xv_stage first)");
  cudaSetDevice(c->device);
  return run_job(c, node, false, nullptr, nullptr, 0);
}

int xv_fetch(xv_ctx* c, float* out, int64_t out_capacity) {
  if (!c) return XV_EINVAL;
  if (!out) return fail(c, XV_EINVAL, "xv_fetch: null output");
  cudaSetDevice(c->device);
  return fetch_result(c, out, out_capacity);
}

int xv_sync(xv_ctx* c) {
  if (!c) return XV_EINVAL;
  cudaSetDevice(c->device);
  cudaError_t e = cudaStreamSynchronize(c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->copy_stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
  return XV_OK;
}

static int forward_common(xv_ctx* c, const float* feats, const int32_t* lengths, int32_t n_utts, const char* node,
                          float* out, const int64_t* offs, int64_t cap, bool frames, const int64_t* view_row0 = nullptr,
                          int64_t view_total_rows = 0) {
  if (!c) return XV_EINVAL;
  if (!feats || !lengths || !out || (frames && !offs)) return fail(c, XV_EINVAL, "xv_forward: null argument");
  cudaSetDevice(c->device);
  const int kind = classify_pointer(feats);
  c->pipeline_hint = kind != 2;
  int rc = plan_job(c, lengths, n_utts, view_row0, view_total_rows);
  c->pipeline_hint = false;
  if (rc != XV_OK) return rc;
  if (kind == 2) {
    // device-resident features: no staging copy at all
    const size_t bytes = static_cast<size_t>(c->total_rows) * c->dim * 4;
    if ((rc = ensure(c, &c->d_feats, bytes)) != XV_OK) return rc;
    cudaError_t e = cudaMemcpyAsync(c->d_feats.p, feats, bytes, cudaMemcpyDeviceToDevice, c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "D2D of features");
    c->staged = true;
    rc = run_job(c, node, false, frames ? out : nullptr, offs, cap);
  } else {
    c->host_feats = feats;
    c->host_pinned = kind == 1;   // pinned: async copies straight from the caller's buffer
    c->staged = false;
    rc = run_job(c, node, true, frames ? out : nullptr, offs, cap);
    c->host_feats = nullptr;
    if (rc != XV_OK) {
      cudaStreamSynchronize(c->copy_stream);
      cudaStreamSynchronize(c->stream);
      return rc;
    }
  }
  if (rc != XV_OK) return rc;
  if (!frames) return fetch_result(c, out, cap);
  return xv_sync(c);
}

int xv_forward(xv_ctx* c, const float* feats, const int32_t* lengths, int32_t n_utts, const char* node, float* out,
               int64_t out_capacity) {
  return forward_common(c, feats, lengths, n_utts, node, out, nullptr, out_capacity, false);
}

int xv_forward_views(xv_ctx* c, const float* feats, int64_t total_rows, const int64_t* view_row0, const int32_t* lengths,
                     int32_t n_views, const char* node, float* out, int64_t out_capacity) {
  if (c && !view_row0) return fail(c, XV_EINVAL, "xv_forward_views: null argument");
  return forward_common(c, feats, lengths, n_views, node, out, nullptr, out_capacity, false, view_row0, total_rows);
}

int xv_forward_frames(xv_ctx* c, const float* feats, const int32_t* lengths, int32_t n_utts, const char* node,
                      float* out, const int64_t* out_offsets, int64_t out_capacity) {
  return forward_common(c, feats, lengths, n_utts, node, out, out_offsets, out_capacity, true);
}

}  // extern "C"
