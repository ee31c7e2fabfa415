/* libxvector.so -- C-ABI of the B200-native x-vector embedding hot path.
 *
 * The reference (entn-at/tf-kaldi-speaker) has no FFI/plugin interface: the path is Python
 * that builds TensorFlow graph nodes.  This library sits one level below those Python
 * signatures; each entry point names the reference code it replaces (paths relative to the
 * reference root).  Plain pointers and sizes only; no torch / numpy types.
 *
 * Ownership : the caller owns every buffer it passes; the library copies weights at
 *             xv_set_weight and owns all device memory; results are written into
 *             caller-allocated buffers.
 * Errors    : 0 on success, negative XV_E* otherwise; xv_last_error() gives the message.
 * Threading : an xv_ctx is bound to one device and one CUDA stream and is NOT re-entrant;
 *             multi-GPU = one ctx (one process or thread) per GPU.
 * There is no CPU fallback: every compute entry point fails with XV_ECUDA without a GPU.
 */
#ifndef XVECTOR_H_
#define XVECTOR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XV_OK            0
#define XV_EINVAL       -1   /* bad argument / malformed config / utterance shorter than context */
#define XV_ECUDA        -2   /* CUDA runtime or driver error (message has the CUDA string) */
#define XV_ENOMEM       -3
#define XV_ESTATE       -4   /* call order violated (e.g. forward before finalize_weights) */
#define XV_ENOTFOUND    -5   /* unknown node / missing weight */
#define XV_EUNSUPPORTED -6   /* network / pooling type not implemented (reference raises NotImplementedError:
                                model/trainer.py:109-110, model/pooling.py:22-23) */
#define XV_ECAPACITY    -7   /* output buffer too small */

#define XV_DTYPE_F32  0
#define XV_DTYPE_BF16 1

typedef struct xv_ctx xv_ctx;

/* ---- life cycle ------------------------------------------------------------------------
 * Replaces Trainer.__init__ + Trainer.build("predict") (model/trainer.py:87-219, 325-338):
 * config_json is the text of the reference nnet_conf JSON (misc/utils.py:13-41) extended with
 *   "feature_dim": D            (nnet/feature_dim, extract.py:54-55)
 *   "precision": "bf16"|"fp32"  (bf16 = tcgen05 path; fp32 = SIMT validation mode)
 *   "batch_frames": int         (optional frame budget of one packed launch sequence)
 * Lazily-injected defaults (num_nodes_pooling_layer=1500 model/tdnn.py:114-116, ...) are applied. */
int  xv_create(int device, const char* config_json, xv_ctx** out);
void xv_destroy(xv_ctx* ctx);
/* Parses config_json and builds the layer plan exactly as xv_create does, on the host only (no device needed): XV_OK
 * and a one-line summary ("scope=tdnn;pooling=...;frame_layers=5;...") in buf, or the error xv_create would give
 * (message through xv_last_error(NULL)).  Every reference nnet_conf JSON must get through (misc/utils.py:13-41). */
int  xv_config_check(const char* config_json, char* buf, size_t buf_size);
/* ctx may be NULL: returns the calling thread's last error from xv_create. */
const char* xv_last_error(xv_ctx* ctx);

/* ---- weights ---------------------------------------------------------------------------
 * Replaces Trainer.load() / tf.train.Saver.restore (model/trainer.py:277-295).  tf_name is the
 * TensorFlow variable name ("tdnn/tdnn1_conv/kernel" [1,k,Cin,Cout], "tdnn/tdnn4_dense/kernel"
 * [in,out], "tdnn/tdnn1_bn/moving_variance", "tdnn/attention/query" [H,dk], ...).  fp32, C order. */
int  xv_set_weight(xv_ctx* ctx, const char* tf_name, const float* host, const int64_t* shape, int rank);
/* Folds BN (eps 1e-3) + bias into per-channel scale/shift, pads K/N, casts to bf16 K-major,
 * builds TMA descriptors.  Fails with XV_ENOTFOUND naming the first missing variable. */
int  xv_finalize_weights(xv_ctx* ctx);

/* ---- endpoints -------------------------------------------------------------------------
 * endpoints[node] of model/tdnn.py:35-179 / model/pooling.py:24,198,220.  width = last-axis size;
 * frame_level = 1 for [B,T',width] nodes; context = T - T' (frames lost to VALID convolutions). */
int  xv_node_info(xv_ctx* ctx, const char* node, int32_t* width, int32_t* frame_level, int32_t* context);

/* ---- the hot path ----------------------------------------------------------------------
 * Replaces sess.run(self.embeddings, {features: x, is_training: False}) (model/trainer.py:909)
 * for a ragged batch: feats = n_utts utterances packed back to back, [sum(lengths), D] fp32
 * (host, pinned host or device memory), lengths[u] frames each.  node = "" -> the config's
 * embedding_node.  Utterance-level nodes: out[n_utts, width] fp32 (host).  Each utterance is
 * computed exactly as the reference computes a batch-of-one (no masking approximations). */
int  xv_forward(xv_ctx* ctx, const float* feats, const int32_t* lengths, int32_t n_utts,
                const char* node, float* out, int64_t out_capacity);
/* Frame-level nodes (extract_frame.py / extract_attention.py readers): utterance u's
 * [lengths[u]-context, width] block (attention_weights: [H, lengths[u]-context]) is written at
 * out + out_offsets[u] (offsets in floats). */
int  xv_forward_frames(xv_ctx* ctx, const float* feats, const int32_t* lengths, int32_t n_utts,
                       const char* node, float* out, const int64_t* out_offsets, int64_t out_capacity);

/* As xv_forward, with the utterances given as VIEWS into one feature buffer: view i = rows
 * [view_row0[i], view_row0[i] + lengths[i]) of feats [total_rows, D]; views may overlap and must come in ascending
 * view_row0 order.  This is the extractor's chunk-and-average (egs/voxceleb/v1/nnet/lib/extract.py:68-86: chunks of
 * chunk_size frames every chunk_size/2 frames) without materialising the overlapping copies on the host: the features
 * cross the bus once, a gather pass packs each batch's views on the device.  out[n_views, width]. */
int  xv_forward_views(xv_ctx* ctx, const float* feats, int64_t total_rows, const int64_t* view_row0,
                      const int32_t* lengths, int32_t n_views, const char* node, float* out, int64_t out_capacity);

/* Same path split in three so a caller (bench.py) can time device-resident work:
 * xv_stage  : H2D of features + length tables (kept on the device until the next xv_stage);
 * xv_run    : launches every kernel of the forward pass asynchronously on the ctx stream;
 * xv_fetch  : D2H of the utterance-level result of the last xv_run + stream sync. */
int  xv_stage(xv_ctx* ctx, const float* feats, const int32_t* lengths, int32_t n_utts);
int  xv_stage_views(xv_ctx* ctx, const float* feats, int64_t total_rows, const int64_t* view_row0,
                    const int32_t* lengths, int32_t n_views);
int  xv_run(xv_ctx* ctx, const char* node);
int  xv_fetch(xv_ctx* ctx, float* out, int64_t out_capacity);
int  xv_sync(xv_ctx* ctx);

/* ---- standalone pooling ----------------------------------------------------------------
 * statistics_pooling(features,...) (model/pooling.py:27-52) for direct callers: x is a DEVICE
 * (or host; then copied) buffer [cu_lens[n], C] of dtype XV_DTYPE_*, utterance u = rows
 * cu_lens[u]..cu_lens[u+1]; out (host) [n, 2C] = [mean | std], population variance, floor 1e-12. */
int  xv_stat_pooling(xv_ctx* ctx, const void* x, const int32_t* cu_lens, int32_t n, int32_t C,
                     int dtype, float* out);
/* As xv_stat_pooling but leaves the result on the device (out_dev [n,2C] fp32) and does not
 * synchronise: the bandwidth-roofline entry point. cu_lens_dev is a device array [n+1]. */
int  xv_stat_pooling_dev(xv_ctx* ctx, const void* x_dev, const int32_t* cu_lens_dev, int32_t n, int32_t C,
                         int dtype, int32_t max_len, float* out_dev);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * xv_stat_pooling_dev (same kernels) but the [n,2C] fp32 sum / sum-of-squares workspace lives in a file-scope
 * __device__ array of XV_NOALLOC_POOL_MAX_FLOATS floats (2 MiB, held by every GPU that loads the library until
 * exit): nothing is allocated or freed, on the device or through any allocating library routine, and nothing is
 * synchronised.  x_dev [total_rows, C] of dtype XV_DTYPE_*, cu_lens_dev [n+1] and out_dev [n,2C] fp32 are DEVICE
 * buffers of the caller; total_rows = cu_lens[n] is passed by value so the call needs no read-back.
 * n*2*C > XV_NOALLOC_POOL_MAX_FLOATS is rejected with XV_EINVAL and a message in xv_last_error().
 * Every call on a device shares the one array, which keeps the previous call's contents: each call clears the
 * part it uses on the ctx stream first, so calls ordered on the GPU are independent, but calls that OVERLAP on
 * the GPU (two ctx of one device inside this entry point at once) race on it: the caller serialises them. */
#define XV_NOALLOC_POOL_MAX_FLOATS (512 * 1024)
int  xv_stat_pooling_noalloc(xv_ctx* ctx, const void* x_dev, const int32_t* cu_lens_dev, int32_t n, int32_t C,
                             int dtype, int64_t total_rows, float* out_dev);
/* The weighted-statistics core of self_attention (model/pooling.py:186-218): value [cu[n],Cv],
 * key [cu[n],Ck] (dtype as above, device or host), query [H,dk] fp32 host.  out (host)
 * [n, 2*H*dv]; weights_out (host, may be NULL) packed [H, L_u] per utterance at H*cu_lens[u]. */
int  xv_attention_pooling(xv_ctx* ctx, const void* value, const void* key, const float* query,
                          const int32_t* cu_lens, int32_t n, int32_t Cv, int32_t Ck, int32_t H,
                          int split_value, int split_key, int use_scale, int dtype,
                          float* out, float* weights_out);

/* One affine (+BN +activation) layer of model/common.py:113-223 / model/tdnn.py:42-55 on fp32 HOST rows through
 * the SIMT validation kernel, for direct array callers of dense_bn_relu()/dense()/dense_relu()/dense_tanh():
 * out[r, n] = act(bn_scale[n] * (sum_{j<taps,c} x[r+j, c] * w[j*cin+c, n] + bias[n]) + bn_shift[n]); out has
 * rows-taps+1 rows.  bn_scale/bn_shift (already folded with eps) and slope (negative-side slope per channel)
 * may be NULL (= no BN / no activation). */
int  xv_dense_rows_f32(xv_ctx* ctx, const float* x, int64_t rows, int32_t cin, int32_t taps, const float* w,
                       int32_t n, const float* bias, const float* bn_scale, const float* bn_shift,
                       const float* slope, int act_tanh, float* out);

/* prelu() (model/common.py:27-42): out = relu(x) + alpha*(x-|x|)/2 per channel, x/out HOST [rows,width] fp32,
 * alpha [width] (or [1] when shared != 0). */
int  xv_prelu_f32(xv_ctx* ctx, const float* x, int64_t rows, int32_t width, const float* alpha, int shared, float* out);
/* l2_scaling() (model/common.py:45-58): out = x * rsqrt(max(sum_row x^2, eps)) * factor, HOST [rows,width] fp32. */
int  xv_l2_scaling_f32(xv_ctx* ctx, const float* x, int64_t rows, int32_t width, float factor, float eps, float* out);

/* ---- measurement / plumbing --------------------------------------------------------------
 * CUDA events on the ctx stream (the stream every kernel of this library is launched on). */
int  xv_event_record(xv_ctx* ctx, int slot);                 /* slot in [0,16) */
int  xv_event_elapsed_ms(xv_ctx* ctx, int slot_a, int slot_b, float* ms);
/* Number of kernels of THIS library launched by the ctx so far, and a per-kernel-family
 * breakdown written as "name=count;..." into buf. */
int64_t xv_launch_count(xv_ctx* ctx);
int  xv_launch_summary(xv_ctx* ctx, char* buf, size_t buf_size);
/* Accumulated device time (ms) of a kernel family between xv_profile_begin/end (events around
 * every launch of that family; serialises nothing else).  family: "gemm", "pool", "cast", ... */
int  xv_profile_enable(xv_ctx* ctx, int on);
int  xv_profile_query(xv_ctx* ctx, const char* family, float* total_ms, int64_t* launches);
/* Pinned host memory for callers that want true async H2D. */
void* xv_host_alloc(size_t bytes);
void  xv_host_free(void* p);
/* Raw device buffers (for the standalone pooling benches and tests). */
void* xv_device_alloc(xv_ctx* ctx, size_t bytes);
void  xv_device_free(xv_ctx* ctx, void* p);
int   xv_memcpy_h2d(xv_ctx* ctx, void* dst_dev, const void* src_host, size_t bytes);
int   xv_memcpy_d2h(xv_ctx* ctx, void* dst_host, const void* src_dev, size_t bytes);
/* Converts fp32 host data to a bf16 device buffer (round-to-nearest-even). */
int   xv_upload_bf16(xv_ctx* ctx, void* dst_dev, const float* src_host, size_t count);
/* Writes `bytes` of a scratch buffer (> L2) to flush the L2 between timed iterations. */
int   xv_flush_l2(xv_ctx* ctx);

/* Generic C = act(scale*(A.W)+shift) on the tcgen05 path, for kernel-level tests and the GEMM
 * roofline bench: A [M, taps-spliced view of Cin] bf16 device, Wt [N,K] bf16 device (K-major),
 * out bf16 device [M, ldo].  See DESIGN.md "GEMM kernel". */
int   xv_gemm_bf16(xv_ctx* ctx, const void* a_dev, int64_t a_rows, int32_t cin, int32_t taps, int32_t dilation,
                   const void* wt_dev, int32_t n, const float* scale_dev, const float* shift_dev,
                   float neg_slope, void* out_dev, int32_t ldo, int64_t m_rows);

/* Bench variant of xv_gemm_bf16: caller supplies the per-channel negative-slope array, ldo == n, and
 * nothing is synchronised (time it with xv_event_record on the ctx stream). */
int   xv_gemm_bf16_async(xv_ctx* ctx, const void* a_dev, int64_t a_rows, int32_t cin, int32_t taps, int32_t dilation,
                         const void* wt_dev, int32_t n, const float* scale_dev, const float* shift_dev,
                         const float* slope_dev, void* out_dev, int64_t m_rows);

/* ---- training step (BASELINE.json config 5) ------------------------------------------------------------
 * Replaces sess.run(train_op, {features, labels, is_training: True, global_step, learning_rate}) of
 * Trainer.train / TrainerMGPU (model/trainer.py:590-673, 1133-1248) for network_type tdnn + statistics_pooling with
 * loss_func softmax | additive_margin_softmax (model/loss.py:9-48, 201-286), optimizer sgd | momentum | adam
 * (model/trainer.py:225-246), clip_gradient / clip_gradient_norm = per-variable tf.clip_by_norm (model/trainer.py:529-533).
 * The context's "precision" picks the arithmetic: "bf16" (product path) runs the frame-layer GEMMs (forward, data and
 * weight gradients) on tcgen05 with bf16 operands, fp32 accumulation and fp32 master weights; "fp32" is the
 * validation-grade path (SIMT FFMA GEMMs).  Both share every other kernel.  Every variable the reference's graph would
 * hold (tdnn/..., softmax/output/kernel [512, num_speakers] (+ /bias for softmax), BN moving statistics) is first given
 * with xv_set_weight; xv_train_init copies them into the flat parameter buffer that the step updates in place. */
int  xv_train_init(xv_ctx* ctx, int32_t num_speakers);
/* One step on this rank's slice of the batch: feats [batch, frames, D] fp32 (host or device), labels [batch]
 * int32 (host).  Forward in training mode (BN batch statistics of THIS rank, model/trainer.py:1183-1190), loss +
 * L2 regularisation, backward, gradient average over ranks (one ncclAllReduce when xv_comm_init was called with
 * world > 1; == add_n/N of model/trainer.py:1240-1244), optimizer update, BN moving averages from rank 0.
 * loss_out (may be NULL): [0] = mean tower loss, [1] = [0] + regularisation loss.  With loss_out the call returns when the
 * step has finished; with NULL it only enqueues the step.  feats / labels may also be DEVICE buffers of the caller
 * (device labels are not range-checked). */
int  xv_train_step(xv_ctx* ctx, const float* feats, const int32_t* labels, int32_t batch, int32_t frames,
                   float learning_rate, int64_t global_step, float* loss_out);
/* The same step as a two-stage pipeline, in place of the reference's queue runners + sess.run(train_op) loop
 * (model/trainer.py:629-660, dataset/data_loader.py:369-495: there the input queue hides the host work; here the copy of
 * batch k+1 overlaps the kernels of step k and the host never waits for the step it has just enqueued).
 * xv_train_submit copies the batch on the copy stream into the alternate batch buffers and enqueues the step behind it
 * without waiting; xv_train_collect waits for the OLDEST submitted step and returns its loss (loss_out as above, may be
 * NULL).  At most two steps may be in flight (XV_ESTATE beyond that); the usual loop is submit(0); for k: submit(k+1),
 * collect(k).  The host buffers of a submit must stay untouched until that step has been collected (pinned memory is
 * copied asynchronously).  Steps, updates and the data-parallel exchange are ordered exactly as with xv_train_step. */
int  xv_train_submit(xv_ctx* ctx, const float* feats, const int32_t* labels, int32_t batch, int32_t frames,
                     float learning_rate, int64_t global_step);
int  xv_train_collect(xv_ctx* ctx, float* loss_out);
/* Current value / last averaged gradient (including the wd*W regularisation term) of a variable, by TF name. */
int  xv_train_get_weight(xv_ctx* ctx, const char* tf_name, float* host, int64_t capacity);
int  xv_train_get_grad(xv_ctx* ctx, const char* tf_name, float* host, int64_t capacity);
/* The variables the training state holds (what tf.train.Saver would write, model/trainer.py:583): count (-1 without a
 * training state), then name i (in name order) and its element count. */
int  xv_train_num_variables(xv_ctx* ctx);
int  xv_train_variable_name(xv_ctx* ctx, int32_t i, char* buf, size_t buf_size, int64_t* size);
/* Data parallelism: rank 0 creates a 128-byte ncclUniqueId (xv_comm_unique_id), the launcher hands it to every
 * rank, each rank calls xv_comm_init(ctx, id, rank, world) after xv_train_init.  NCCL is loaded at run time
 * (libnccl.so.2, the copy already mapped by the process if there is one).  world == 1 needs no NCCL. */
int  xv_comm_unique_id(xv_ctx* ctx, void* out128);
int  xv_comm_init(xv_ctx* ctx, const void* nccl_unique_id, int rank, int world);
/* The exchange step of xv_train_step as configured: this rank, the world size, the fp32 elements every step all-reduces
 * (gradients + moving statistics + loss slot) and the number of buckets they travel in (3, launched on a side stream
 * as the backward pass retires them; 1 when world == 1 or XV_TRAIN_OVERLAP=0).  A step that fails on one rank aborts
 * the communicator (ncclCommAbort); the peers see an asynchronous NCCL error instead of a hang (ncclCommGetAsyncError is
 * polled while a step is waited for) and every context of the job must be re-created. */
int  xv_comm_info(xv_ctx* ctx, int32_t* rank, int32_t* world, int64_t* grad_elems, int32_t* buckets);
/* XV_DTYPE_BF16 when the gradients of the trainable variables cross NVLink as bf16 (default of the bf16 precision with
 * world > 1: fp32 master weights and fp32 sums on arrival, half the bytes in flight; the tail -- moving statistics and the
 * loss -- always travels as fp32), XV_DTYPE_F32 otherwise.  Config key "grad_allreduce_dtype": "bf16" | "fp32". */
int  xv_comm_dtype(xv_ctx* ctx);
/* The exchange step alone (measurement): all-reduces the step's gradient buckets `iters` times back to back on the
 * communication stream and reports the mean time of one full exchange (CUDA events on that stream).  Collective: every
 * rank must call it with the same iters.  The gradient buffer is scratch between steps. */
int  xv_comm_bench(xv_ctx* ctx, int32_t iters, float* ms_per_iter);

/* One 2-D convolution of the ResNet-18 front end on the implicit-GEMM tcgen05 path (kernel-level tests and the
 * conv roofline bench): replaces tf.layers.conv2d(padding='same', use_bias=False) + batch_normalization
 * [+ shortcut] + activation of model/resnet.py:25-86.  x [nb,t,f,cin] bf16 NHWC (device), wt [cout][kh*kw*cin]
 * bf16 K-major (device; TF kernel [kh,kw,cin,cout] transposed), strides st/sf in {1,2} (TensorFlow SAME padding),
 * out [nb,ceil(t/st),ceil(f/sf),cout] bf16 = act(scale*conv + shift + residual); residual may be NULL.
 * cin % 64 == 0, cout % 64 == 0.  Asynchronous on the ctx stream. */
int   xv_conv2d_bf16_async(xv_ctx* ctx, const void* x_dev, int32_t nb, int32_t t, int32_t f, int32_t cin,
                           const void* wt_dev, int32_t kh, int32_t kw, int32_t st, int32_t sf, int32_t cout,
                           const float* scale_dev, const float* shift_dev, const float* slope_dev,
                           const void* residual_dev, void* out_dev);

/* Weight gradient of a spliced (1 x taps) layer on the tcgen05 path (kernel-level tests; used by xv_train_step in
 * bf16 mode): dW[j*cin + c, n] = sum_{m<rows} a[m + j, c] * dz[m, n] -- what tf.gradients produces for the kernels of
 * model/tdnn.py:42-103 under model/trainer.py:501-527.  a [a_rows, lda] bf16, dz [rows, ldz] bf16, out fp32
 * [taps*cin, ldo] (zeroed by the call), all device pointers; lda, ldz multiples of 8, ldo a multiple of 4. */
int   xv_wgrad_bf16_async(xv_ctx* ctx, const void* a_dev, int64_t a_rows, int32_t lda, const void* dz_dev, int32_t ldz,
                          int64_t rows, int32_t taps, int32_t cin, int32_t n_valid, float* out_dev, int32_t ldo);

/* ---- post-extraction vector operations -------------------------------------------------
 * What egs/voxceleb/v1/nnet/run_extract_embeddings.sh:80-103 runs on the x-vectors through Kaldi binaries, and the
 * cosine trial scores of misc/utils.py:307-346.  Host pointers in and out; synchronous.
 *
 * ivector-normalize-length: out_i = x_i / |x_i| (scaleup = 0, the recipe's --scaleup=false) or x_i * sqrt(dim) / |x_i|
 * (scaleup = 1); a zero vector is copied unchanged. */
int   xv_vec_normalize_length(xv_ctx* ctx, const float* x, int64_t n, int32_t dim, int scaleup, float* out);
/* ivector-mean over spk2utt: out[g] = mean of the vectors with group[i] == g (group[i] < 0: not used), summed in input
 * order; counts[g] (may be NULL) = number of members (num_utts.ark); an empty group gives a zero vector. */
int   xv_vec_group_mean(xv_ctx* ctx, const float* x, int64_t n, int32_t dim, const int32_t* group, int32_t n_groups,
                        float* out, int32_t* counts);
/* out[i, j] = a_i . b_j / (sqrt(|a_i|^2 + eps) * sqrt(|b_j|^2 + eps)); compute_cos_pairwise_eer uses eps = 1e-12 and
 * a == b (misc/utils.py:318,326). */
int   xv_cosine_scores(xv_ctx* ctx, const float* a, int64_t na, const float* b, int64_t nb, int32_t dim, float eps,
                       float* out);

const char* xv_version(void);

#ifdef __cplusplus
}
#endif
#endif /* XVECTOR_H_ */
