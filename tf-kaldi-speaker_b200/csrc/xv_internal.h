// Internal declarations shared by the kernels and the C-ABI layer of libxvector.so.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace xv {

// ---- tcgen05 GEMM tile configuration (see DESIGN.md "GEMM kernel") ----
constexpr int kBM = 128;        // rows (frames) per tile  = UMMA M
constexpr int kBN = 256;        // output channels per tile = UMMA N
constexpr int kBK = 64;         // K per pipeline stage = one 128-byte swizzle atom of bf16
constexpr int kBLoadRows = 128; // rows of one weight TMA box: a 256-column tile is two boxes (one per CTA of a cluster pair)
constexpr int kUK = 16;         // UMMA K for 16-bit inputs
constexpr int kStages = 4;      // TMA->MMA smem ring depth
constexpr int kAccStages = 2;   // TMEM accumulator double buffer (2 x 256 columns = all 512)
constexpr int kGemmThreads = 320;  // warp0 TMA, warp1 MMA+TMEM alloc, warps2-9 epilogue

// One valid-row range of one utterance inside the packed row space: rows [r0, r1) belong to
// utterance `utt` and are valid outputs of the pooled layer.
struct PoolSeg {
  int32_t utt;
  int32_t r0;
  int32_t r1;
  int32_t pad;
};

struct GemmParams {
  int32_t m_rows;      // rows of output to produce (stores are masked to row < m_rows)
  int32_t m_tiles;     // ceil(m_rows / kBM)
  int32_t n_tiles;     // N_alloc / kBN
  int32_t taps;        // spliced context width (conv kernel width); 1 for dense layers
  int32_t dilation;    // row step between taps (reference uses 1 everywhere)
  int32_t kb_per_tap;  // Cin / kBK
  int32_t cin;         // channels of the A buffer (multiple of kBK)
  // epilogue: y = s*acc + h ; y = y>=0 ? y : slope*y ; optional tanh
  const float* scale;  // [N_alloc]
  const float* shift;  // [N_alloc]
  const float* slope;  // [N_alloc] negative-side slope (0 relu, 0.2 lrelu, alpha prelu, 1 linear)
  int32_t act_tanh;
  int32_t relu_only;   // every slope is 0 (plain ReLU) and no tanh: fast epilogue path
  __nv_bfloat16* out_bf16;  // [m_rows, ldo] or null
  float* out_f32;           // [m_rows, ldo] or null
  int32_t ldo;
  int32_t f32_cols = 0x7fffffff;   // fp32 stores are masked to columns < f32_cols (a multiple of 4); training: cout = 1500
  // fused statistics-pooling partial sums (tdnn5): acc[utt][0|1][pool_ld] += sum x | sum x^2
  double* pool_acc;    // fp64: atomics become order-independent at fp32 precision
  int32_t pool_ld;
  int32_t col_stationary = 0;   // set by launch_gemm: grid = cpn x n_tiles CTAs, each keeps one column tile (xv_gemm_sm100.cu)
  const int32_t* tile_seg_ptr;  // [2 * m_tiles]: (lo, hi) index range into segs per tile
  const PoolSeg* segs;
  // optional per-row weights (attention pooling): weighted sums  sum w x | sum w x^2 ; [heads, m_rows]
  const float* row_weight;   // [heads, weight_ld] softmax weights of the rows (null: plain sums)
  int32_t heads;        // number of heads; column n belongs to head n / head_width
  int32_t head_width;
  int32_t weight_ld = 0;
  // fused attention scores (last key layer): dot_out[h][row] += dot_scale * sum_{n in head h} y[row, n] * dot_q[n]
  // (fp64 atomics: order-independent at fp32 precision); heads / head_width above give the column -> head map
  const float* dot_q = nullptr;   // [N_alloc], zero beyond the key width
  double* dot_out = nullptr;      // [heads][dot_ld], zeroed by the caller
  int32_t dot_ld = 0;
  float dot_scale = 1.f;
  // fused batch-norm batch statistics (training forward, storing kernel): stat_out[0][n] += sum_r y[r, n],
  // stat_out[1][n] += sum_r y[r, n]^2 over the rows with (r mod stat_t) < stat_valid; [2][stat_ld] fp64, zeroed by the caller
  double* stat_out = nullptr;
  int32_t stat_ld = 0;
  int32_t stat_t = 1, stat_valid = 1;
};

struct GemmPlan {
  CUtensorMap tm_a;
  CUtensorMap tm_b;
  CUtensorMap tm_out;  // bf16 output [m_rows, ldo], box 32x32, SWIZZLE_64B (unused for fp32 / pooled outputs)
  GemmParams p;
  bool pool;
};

// Host-side helpers (xv_gemm_sm100.cu)
const char* gemm_init();   // resolves cuTensorMapEncodeTiled, sets smem attributes; returns error text or null
const char* make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows,
                              uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_rows, int swizzle_bytes = 128);
cudaError_t launch_gemm(const GemmPlan& plan, int sm_count, cudaStream_t stream);

// ---- implicit-GEMM 2-D convolution (xv_conv_sm100.cu), ResNet-18 front end ----
// Per-tap coordinate offsets into the 5-D input map (sf*C, F/sf, st, T/st, B): see xv_conv_sm100.cu.
struct ConvTap {
  int32_t c0, c1, c2, c3;
};
struct ConvGemmParams {
  int32_t n_utts, t_out, f_out;
  int32_t fb, tb;            // output tile = tb time steps x fb frequency bins (fb*tb <= 128)
  int32_t tiles_f, tiles_t;  // f_out / fb, ceil(t_out / tb)
  int32_t tile_rows;         // fb * tb
  int32_t m_tiles, n_tiles;  // n_utts * tiles_t * tiles_f, cout / BN
  int32_t taps, kb_per_tap, cin;
  ConvTap tap[9];
  const float* scale;        // folded BN: y = scale*acc + shift
  const float* shift;
  const float* slope;        // negative-side slope per channel (1 = no activation)
  int32_t relu_only;
  const __nv_bfloat16* residual;   // [n_utts][t_alloc_res][f_out][cout] added before the activation, or null
  int32_t t_alloc_res;
  int32_t cout;
  // halo mode (xv_conv_halo_sm100.cu; 3x3 stride-1 layers with 64 / 128 input channels): the tile is tb time steps x
  // fbh = fb + 2 columns INCLUDING the frequency halo, so that every tap is the same shared-memory box shifted by a
  // constant number of rows; columns fb, fb+1 of a tile row are junk and are dropped by the epilogue
  int32_t fbh = 0;
  int32_t a_stages = 0, a_stage_bytes = 0;   // halo ring: stages (2..8) of round_up((tb+2)*fbh*128, 1024) bytes
};
struct ConvGemmPlan {
  CUtensorMap tm_a, tm_b, tm_out;
  ConvGemmParams p;
  int bn;                    // 64 | 128 | 256
  uint32_t halo_smem = 0;    // dynamic shared memory of the halo kernel (conv_halo_plan)
};
const char* conv_gemm_init();
const char* conv_halo_init();
bool conv_halo_plan(ConvGemmParams* p, uint32_t* smem_bytes);   // needs fb, tb, fbh, kb_per_tap
cudaError_t launch_conv_halo(const ConvGemmPlan& plan, int sm_count, cudaStream_t stream);   // plan.tm_a = halo box map
cudaError_t launch_conv_gemm(const ConvGemmPlan& plan, int sm_count, cudaStream_t stream);
// rank-N bf16 tensor map (dims/strides innermost first; strides[0] is implied = 2 bytes).
const char* make_tmap_bf16_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                              const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

// ---- tensor-core weight gradient (xv_wgrad_sm100.cu), training step ----
struct WgradParams {
  int64_t rows;              // packed frames contracted over
  int64_t rows_per_split;    // multiple of 64
  int32_t splits;
  int32_t taps, cin;         // dW rows = taps * cin
  int32_t n_valid;           // dW columns written
  int32_t m_tiles_per_tap, m_tiles, n_tiles;
  int32_t ldo;
  float* out;
};
const char* wgrad_init();
// dW[j*cin + c, n] += sum_m a[m + j, c] * dz[m, n]; a [a_rows, lda] bf16, dz [rows, ldz] bf16, out fp32 pre-zeroed.
const char* launch_wgrad(const void* a, int64_t a_rows, int lda, const void* dz, int ldz, int64_t rows, int taps, int cin,
                         int n_valid, float* out, int ldo, int sm_count, cudaStream_t stream, cudaError_t* cuda_err);

// ---- SIMT / elementwise kernels (xv_kernels.cu) ----
// features fp32 [rows, D] -> spliced bf16 [rows, kpad]: out[r, j*D + c] = in[r + j*dil, c] (0 beyond rows / kpad)
cudaError_t launch_splice_cast(const float* feats, int64_t rows, int d, int taps, int dil, __nv_bfloat16* out,
                               int kpad, cudaStream_t stream);
// views -> packed rows: dst[r, :] = src[src0[v] + (r - cu[v]), :] for the view v with cu[v] <= r < cu[v+1] (fp32, dim floats per row)
cudaError_t launch_gather_views(const float* src, const int32_t* cu, const int64_t* src0, int32_t n_views, int64_t rows, int dim,
                                float* dst, cudaStream_t stream);
// fp32 validation GEMM with the same tap splice: out = act(scale*(A.W)+shift), optional 3 stage outputs.
struct SimtGemmArgs {
  const float* a;       // [a_rows, cin]
  int64_t a_rows;
  int32_t cin, taps, dilation;
  const float* w;       // [taps*cin, n] row-major (TF kernel layout flattened)
  int32_t n;
  const float* bias;    // [n]
  const float* bn_scale;  // [n] or null (no BN)
  const float* bn_shift;  // [n]
  const float* slope;   // [n] or null (no activation)
  int32_t act_tanh;
  float* out_dense;     // [m_rows, n] or null
  float* out_bn;        // or null
  float* out_act;       // or null
  int64_t m_rows;
};
cudaError_t launch_simt_gemm(const SimtGemmArgs& a, cudaStream_t stream);

// Ragged (optionally attention-weighted) mean/std pooling over packed rows (xv_pool.cu).
// x [rows, ldx] (bf16 or fp32); utterance u = rows [cu[u], cu[u+1]); its last `context` rows are ignored.
struct PoolArgs {
  const void* x;
  int dtype;             // XV_DTYPE_F32 (0) | XV_DTYPE_BF16 (1)
  int32_t c;             // channels
  int32_t ldx;           // row stride (elements)
  int64_t rows;          // total packed rows
  const int32_t* cu;     // [n+1] device
  int32_t n;
  int32_t context;
  const float* weights;  // null -> statistics pooling; else softmax weights [heads][rows]
  int32_t heads;
  int32_t split_value;   // 1: column c belongs to head c/(C/heads); 0: every head pools all columns
  float* acc;            // workspace, pool_workspace_bytes(n, hout, c)
  float* out;            // [n, ldout] fp32: mean block | std block (head-major)
  int32_t ldout;
  __nv_bfloat16* out_bf16;  // optional bf16 copy
  int32_t ldout_bf16;
};
size_t pool_workspace_bytes(int32_t n, int32_t hout, int32_t c);
cudaError_t launch_pool(const PoolArgs& a, int sm_count, cudaStream_t stream);

// Finalize the fused (GEMM-epilogue) pooling accumulators: acc [n][2][acc_ld] (sum, sumsq) -> mean|std.
cudaError_t launch_pool_finalize(const double* acc, int acc_ld, const int32_t* cu, int context, int n, int c,
                                 float* out_f32, int ldo_f32, __nv_bfloat16* out_bf16, int ldo_bf16, int weighted,
                                 cudaStream_t stream);

// Attention: scores[h, r] = scale * sum_d key[r, h*dk + d] * q[h, d]  (split_key) or key[r, d] (shared key)
cudaError_t launch_att_scores(const void* key, int dtype, int ldk, int64_t rows, const float* query, int heads,
                              int dk, int split_key, float scale, float* scores, cudaStream_t stream);
// Per (utt, head): softmax over the valid rows, in place -> weights.
cudaError_t launch_att_softmax(float* scores, int64_t rows, const int32_t* cu, int context, int n, int heads,
                               cudaStream_t stream);
// Same from fp64 score accumulators (fused key-layer epilogue) into the fp32 weight buffer.
cudaError_t launch_att_softmax_f64(const double* acc, float* weights, int64_t rows, const int32_t* cu, int context, int n,
                                   int heads, cudaStream_t stream);
cudaError_t launch_gather_att_weights(const float* w, int64_t rows, const int32_t* cu, int context, int heads, int n,
                                      const int64_t* dst_off, float* dst, cudaStream_t stream);

// Misc
cudaError_t launch_gather_rows(const void* src, int dtype, int lds, int width, const int64_t* src_row0,
                               const int32_t* n_rows, const int64_t* dst_off, int n_utts, int max_rows, float* dst,
                               cudaStream_t stream);
cudaError_t launch_f32_to_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t stream);
cudaError_t launch_fill_u32(uint32_t* dst, size_t n, uint32_t v, cudaStream_t stream);
cudaError_t launch_affine_rows(const float* in, int ldi, int n_rows, int width, const float* scale,
                               const float* shift, const float* slope, float* out_f32, int ldo,
                               __nv_bfloat16* out_bf16, int ldob, cudaStream_t stream);
cudaError_t launch_l2_scale_rows(float* x, int ld, int n_rows, int width, float factor, float eps,
                                 cudaStream_t stream);

}  // namespace xv
