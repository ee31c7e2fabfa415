// Device kernels of the training step (xv_train.cu): everything that is not a frame-layer tensor-core GEMM.
//
// Replaces, op by op, what tf.gradients + tf.layers.batch_normalization(training=True) emit for the graph of
// model/tdnn.py:36-181 under model/trainer.py:472-588: batch-norm forward/backward in batch-statistics mode, the
// activation derivative, statistics pooling forward/backward (model/pooling.py:40-50), the loss head
// (model/loss.py:9-48, 201-286) and the optimizer update (model/trainer.py:225-246).
//
// Round-2 layout: the step is built from FEW fat kernels instead of many thin ones --
//   * batch statistics come out of the forward GEMM's epilogue (xv_gemm_sm100.cu, GemmParams::stat_out) or, in the
//     fp32 validation path, out of col_stats4_kernel; mean / rstd / moving averages are finalised inside the
//     normalise+activate pass that needs them anyway (no bn_finalize launch);
//   * the last frame layer's normalise+activate pass IS the statistics pooling (its output is never stored), and its
//     backward pass rebuilds the pooling gradient on the fly from (pooled, d pooled) instead of reading a [rows, C]
//     gradient tensor a separate pool_bwd pass would have written;
//   * the segment-level layers (B rows) do statistics + normalise (forward) and reduce + apply (backward) in one
//     kernel each;
//   * parameter gradients are written by the kernels that own the sums (no fp64 -> fp32 copy launches);
//   * the softmax weight gradient, the l2_normalize Jacobian of AM-softmax and the bias gradient are one kernel;
//   * the regularisation loss rides in the update kernel; the bf16 operand copies of the master weights are refreshed
//     by ONE launch after the update instead of one launch per layer and layout every step.
#pragma once
#include <math.h>

#include "xv_internal.h"

namespace xv {
namespace {

// ------------------------------------------------------------------------------------------------ fp32 GEMMs
// (validation-grade path and the skinny segment-level products)
// DGRAD:  out[m, c]        = sum_{j<taps} sum_{n<nz} dz[(m - j), n] * w[(j*cin + c), n]          (rows m-j < 0 -> 0)
// WGRAD:  out[j*cin + c, n] = sum_{m<rows}            a[(m + j), c] * dz[m, n]                    (rows m+j >= rows -> 0)
struct TrGemmArgs {
  const float* a;     // WGRAD: activations [rows, lda]
  int32_t lda;
  const float* dz;    // [rows, nz]
  const float* w;     // DGRAD: [taps*cin, nz]
  float* out;         // DGRAD: [rows, cin]; WGRAD: [taps*cin, nz]
  int64_t rows;
  int32_t cin, nz, taps;
};

enum { kDgrad = 0, kWgrad = 1 };

template <int MODE>
__global__ void __launch_bounds__(256) tr_gemm_kernel(const TrGemmArgs g) {
  __shared__ float sa[16][65];
  __shared__ float sb[16][65];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t m0 = static_cast<int64_t>(blockIdx.x) * 64;
  const int n0 = blockIdx.y * 64;
  const int64_t M = MODE == kDgrad ? g.rows : static_cast<int64_t>(g.taps) * g.cin;
  const int N = MODE == kDgrad ? g.cin : g.nz;
  const int64_t K = MODE == kDgrad ? static_cast<int64_t>(g.taps) * g.nz : g.rows;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int64_t k0 = 0; k0 < K; k0 += 16) {
    if (MODE == kDgrad) {
      // A(m,k) = dz[m - j, n], k = j*nz + n: threads run along k (contiguous n)
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        const int kk = i & 15, mm = i >> 4;
        const int64_t m = m0 + mm, k = k0 + kk;
        float v = 0.f;
        if (m < M && k < K) {
          const int j = static_cast<int>(k / g.nz), n = static_cast<int>(k - static_cast<int64_t>(j) * g.nz);
          if (m - j >= 0) v = g.dz[(m - j) * g.nz + n];
        }
        sa[kk][mm] = v;
      }
      // B(k,c) = w[(j*cin + c), n]: threads along k
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        const int kk = i & 15, nn = i >> 4;
        const int64_t k = k0 + kk;
        const int c = n0 + nn;
        float v = 0.f;
        if (c < N && k < K) {
          const int j = static_cast<int>(k / g.nz), n = static_cast<int>(k - static_cast<int64_t>(j) * g.nz);
          v = g.w[(static_cast<int64_t>(j) * g.cin + c) * g.nz + n];
        }
        sb[kk][nn] = v;
      }
    } else {
      // A(mm,k) = a[(k + j), c], mm = j*cin + c: threads along mm (contiguous c)
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        const int mm = i & 63, kk = i >> 6;
        const int64_t m = m0 + mm, k = k0 + kk;
        float v = 0.f;
        if (m < M && k < K) {
          const int j = static_cast<int>(m / g.cin), c = static_cast<int>(m - static_cast<int64_t>(j) * g.cin);
          if (k + j < g.rows) v = g.a[(k + j) * g.lda + c];
        }
        sa[kk][mm] = v;
      }
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        const int nn = i & 63, kk = i >> 6;
        const int64_t k = k0 + kk;
        const int n = n0 + nn;
        sb[kk][nn] = (n < N && k < K) ? g.dz[k * g.nz + n] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = sa[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = sb[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < N) g.out[m * N + n] = acc[i][j];
    }
  }
}

template <int MODE>
cudaError_t launch_tr_gemm(const TrGemmArgs& g, cudaStream_t s) {
  const int64_t M = MODE == kDgrad ? g.rows : static_cast<int64_t>(g.taps) * g.cin;
  const int N = MODE == kDgrad ? g.cin : g.nz;
  if (M <= 0 || N <= 0) return cudaSuccess;
  dim3 grid(static_cast<unsigned>((M + 63) / 64), static_cast<unsigned>((N + 63) / 64));
  tr_gemm_kernel<MODE><<<grid, 256, 0, s>>>(g);
  return cudaGetLastError();
}

// Segment-level weight gradient: out[k, n] = sum_{b < rows} a[b, k] * dz[b, n]  (rows = utterances of the step).
// Both operands are read along their contiguous axis; 64 x 64 output tile per block, 4 x 4 per thread.
struct RowsTArgs {
  const float* a; int32_t lda;     // [rows, K]
  const float* dz; int32_t ldz;    // [rows, N]
  float* out; int32_t ldo;         // [K, N]
  int32_t rows, k, n;
};
__global__ void __launch_bounds__(256) rowsT_gemm_kernel(const RowsTArgs g) {
  __shared__ __align__(16) float As[32][64];
  __shared__ __align__(16) float Ds[32][64];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int k0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int b0 = 0; b0 < g.rows; b0 += 32) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int idx = i * 256 + threadIdx.x;
      const int col = idx & 63, bb = idx >> 6;
      const int b = b0 + bb;
      As[bb][col] = (b < g.rows && k0 + col < g.k) ? g.a[static_cast<int64_t>(b) * g.lda + k0 + col] : 0.f;
      Ds[bb][col] = (b < g.rows && n0 + col < g.n) ? g.dz[static_cast<int64_t>(b) * g.ldz + n0 + col] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int bb = 0; bb < 32; ++bb) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[bb][ty * 4]);
      const float4 d4 = *reinterpret_cast<const float4*>(&Ds[bb][tx * 4]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w};
      const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], dv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = k0 + ty * 4 + i;
    if (k >= g.k) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < g.n) g.out[static_cast<int64_t>(k) * g.ldo + n] = acc[i][j];
    }
  }
}
inline cudaError_t launch_rowsT_gemm(const RowsTArgs& g, cudaStream_t s) {
  if (g.k <= 0 || g.n <= 0) return cudaSuccess;
  rowsT_gemm_kernel<<<dim3((g.k + 63) / 64, (g.n + 63) / 64), 256, 0, s>>>(g);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ shared pieces
// Valid rows of a frame-level tensor: (r mod T) < valid_len.
struct RowMask {
  int32_t t, valid_len;   // utterance length in rows and number of valid rows per utterance
  __device__ __forceinline__ bool ok(int64_t r) const {     // rows of a step fit 32 bits (checked by xv_train_step)
    return static_cast<int32_t>(static_cast<uint32_t>(r) % static_cast<uint32_t>(t)) < valid_len;
  }
};

// A strided row walk (r, r + step, ...) that knows its utterance and its position inside it without dividing per row (the
// division -- 64-bit for the utterance index -- was a quarter of the row-streaming kernels' instructions; ncu source view r02t).
struct RowWalk {
  uint32_t t, valid_len, rem;
  int32_t ut;
  __device__ __forceinline__ RowWalk(const RowMask& m, int64_t first) {
    t = static_cast<uint32_t>(m.t); valid_len = static_cast<uint32_t>(m.valid_len);
    const uint32_t f = static_cast<uint32_t>(first);
    rem = f % t; ut = static_cast<int32_t>(f / t);
  }
  __device__ __forceinline__ bool ok() const { return rem < valid_len; }
  __device__ __forceinline__ void advance(uint32_t step) {
    rem += step;
    while (rem >= t) { rem -= t; ++ut; }
  }
};

// Batch statistics of one layer as fp64 column sums: sums[c] = sum z, sums[ld + c] = sum z^2 over n rows.
struct BnStat {
  const double* sums;     // null: the layer has no batch norm
  int32_t ld;
  double n;
};
__device__ __forceinline__ void bn_mean_rstd(const BnStat& b, int c, float eps, float* mean, float* rstd, double* var_out) {
  const double inv_n = 1.0 / b.n;                   // uniform: one divide per thread, hoisted by the compiler
  const double m = b.sums[c] * inv_n;
  double var = b.sums[b.ld + c] * inv_n - m * m;    // fp64 where the cancellation is
  if (var < 0.0) var = 0.0;
  *mean = static_cast<float>(m);
  *rstd = rsqrtf(static_cast<float>(var) + eps);    // 1 ulp in fp32: every consumer of rstd is fp32 anyway
  *var_out = var;
}
// What the pass that first needs mean / rstd leaves behind for the backward pass and the optimizer:
// mean, rstd [C]; the new moving statistics (momentum update, model/tdnn.py:49-52) go to new_mm / new_mv -- the tail of
// the gradient buffer on rank 0 (tower 0, model/trainer.py:1183-1190), a scratch area on the other ranks.
struct BnFinalize {
  float* mean; float* rstd;
  const float* mov_mean; const float* mov_var;
  float* new_mm; float* new_mv;
  float momentum, eps;
  int32_t unbiased;       // fused batch norm (rank-4 conv inputs) feeds the UNBIASED batch variance to the moving average
};
__device__ __forceinline__ void bn_finalize_channel(const BnStat& b, const BnFinalize& f, int c, float mean, float rstd, double var) {
  f.mean[c] = mean;
  f.rstd[c] = rstd;
  const double m = b.sums[c] / b.n;
  const double vm = f.unbiased ? var * (b.n / fmax(b.n - 1.0, 1.0)) : var;
  f.new_mm[c] = static_cast<float>(static_cast<double>(f.mov_mean[c]) * f.momentum + m * (1.0 - f.momentum));
  f.new_mv[c] = static_cast<float>(static_cast<double>(f.mov_var[c]) * f.momentum + vm * (1.0 - f.momentum));
}

// Per-channel constants of y = act(gamma * (z - mean) * rstd + beta), folded so that a row loop pays ONE FMA for the normalised
// value and ONE (independent) FMA for the pre-activation: xhat = z * xa + xb, pre = z * pa + pb.  Every kernel of the step --
// forward, pooled forward, backward reduce and apply -- uses this form, so they agree on the sign of `pre` bit for bit.
// (Against the textbook order the rounding differs by ~6e-8 * |mean| * rstd: the row kernels are bound by dependent-instruction
// latency, and this removes a third of their arithmetic.)  No batch norm: pre = z, xhat = 0.
struct BnFold { float xa, xb, pa, pb; };
__device__ __forceinline__ BnFold bn_fold(bool has_bn, float mu, float rs, float ga, float be) {
  BnFold f;
  if (!has_bn) { f.xa = 0.f; f.xb = 0.f; f.pa = 1.f; f.pb = 0.f; return f; }
  f.xa = rs;
  f.xb = -mu * rs;
  f.pa = ga * rs;
  f.pb = fmaf(ga, f.xb, be);
  return f;
}

template <int Q> struct VecF;
template <> struct VecF<4> {
  static __device__ __forceinline__ void load(const float* p, float* v) { const float4 t = *reinterpret_cast<const float4*>(p); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
  static __device__ __forceinline__ void store(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
  static __device__ __forceinline__ void store_bf16(__nv_bfloat16* p, const float* v) {
    const __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
    uint2 u;
    u.x = *reinterpret_cast<const uint32_t*>(&lo);
    u.y = *reinterpret_cast<const uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = u;
  }
};

// Gradient of statistics pooling w.r.t. its input, rebuilt on the fly (model/pooling.py:40-50):
// pooled = [mean | std] per utterance; d x[r, c] = dmean/L + dstd * (x - mean) / (L * std); no gradient through the
// variance floor.  enabled == 0: the incoming gradient is read from memory instead.
struct PoolGrad {
  const float* pooled;    // [B, 2C]
  const float* dpooled;   // [B, 2C]
  float inv_l;
  int32_t enabled;
};
__device__ __forceinline__ float pool_grad(float y, float mu, float sd, float dm, float ds, float inv_l) {
  float out = dm * inv_l;
  if (sd * sd > 1.0000001e-12f) out += ds * (y - mu) * inv_l / sd;
  return out;
}
// The same per (utterance, channel) as two coefficients, so that the row loop pays one subtract and one FMA per element
// instead of a divide: d x = ca + cb * (y - mean), ca = dmean / L, cb = dstd / (L * std) (0 under the variance floor).
struct PoolCoef { float mu[4], ca[4], cb[4]; };
__device__ __forceinline__ void pool_coef_load(const PoolGrad& pg, int64_t u, int c_total, int c0, PoolCoef* k) {
  float sd[4], dm[4], ds[4];
  const float* pp = pg.pooled + u * 2 * c_total + c0;
  const float* dp = pg.dpooled + u * 2 * c_total + c0;
  const float4 a = *reinterpret_cast<const float4*>(pp), b = *reinterpret_cast<const float4*>(pp + c_total);
  const float4 c = *reinterpret_cast<const float4*>(dp), d = *reinterpret_cast<const float4*>(dp + c_total);
  k->mu[0] = a.x; k->mu[1] = a.y; k->mu[2] = a.z; k->mu[3] = a.w;
  sd[0] = b.x; sd[1] = b.y; sd[2] = b.z; sd[3] = b.w;
  dm[0] = c.x; dm[1] = c.y; dm[2] = c.z; dm[3] = c.w;
  ds[0] = d.x; ds[1] = d.y; ds[2] = d.z; ds[3] = d.w;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    k->ca[j] = dm[j] * pg.inv_l;
    k->cb[j] = sd[j] * sd[j] > 1.0000001e-12f ? ds[j] * pg.inv_l / sd[j] : 0.f;
  }
}

// ------------------------------------------------------------------------------------------------ frame-level reductions
// Column sums over the valid rows of a [rows, C] matrix, four channels per thread (16-byte loads), fp64 accumulators.
// KIND 0 (fp32 validation path; the bf16 path takes these sums from the GEMM epilogue): v0 = sum z, v1 = sum z^2.
// KIND 1 (backward through activation + batch norm): g = dy * act'(pre); v0 = sum g, v1 = sum g*xhat, v2 = sum dy*pre [pre<0].
// `out` ([NV][out_ld]) must be zero on entry.
template <int KIND>
__global__ void __launch_bounds__(256, 2) col_reduce4_kernel(const float* __restrict__ dy, const float* __restrict__ z,
                                                          const RowMask mask, int64_t rows, int c_total, int64_t rows_per_block,
                                                          const float* __restrict__ mean, const float* __restrict__ rstd,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          const float* __restrict__ slope, const PoolGrad pg,
                                                          double* __restrict__ out, int out_ld) {
  constexpr int NV = KIND == 0 ? 2 : 3;
  constexpr int U = 4;                                 // rows in flight per thread: loads first, math after (2 rows and 3 thinner blocks per SM: 7 % slower step, r02x)
  __shared__ double red[NV][8][129];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + cx) * 4;
  const int64_t r0 = static_cast<int64_t>(blockIdx.y) * rows_per_block;
  const int64_t r1 = min(r0 + rows_per_block, rows);
  double v[NV][4];
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) v[k][j] = 0.0;
  if (c0 < c_total) {
    float sl[4];
    BnFold bf[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      sl[j] = 0.f;
      bf[j] = bn_fold(false, 0.f, 1.f, 1.f, 0.f);
      if (KIND == 1) {
        sl[j] = slope[c0 + j];
        if (mean != nullptr) bf[j] = bn_fold(true, mean[c0 + j], rstd[c0 + j], gamma[c0 + j], beta[c0 + j]);
      }
    }
    int64_t cur_u = -1;
    PoolCoef pc;
#pragma unroll
    for (int j = 0; j < 4; ++j) { pc.mu[j] = 0.f; pc.ca[j] = 0.f; pc.cb[j] = 0.f; }
    const bool read_dy = KIND == 1 && !pg.enabled;
    RowWalk walk(mask, r0 + ry);
    for (int64_t rb = r0 + ry; rb < r1; rb += 8 * U) {
      float zz[U][4], dd[U][4];
      bool ok[U];
      int32_t uts[U];
      // invalid rows hold finite values (zeros / products of real data): loading them unconditionally keeps U loads in flight
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t r = rb + 8 * u;
        ok[u] = r < r1 && walk.ok();
        uts[u] = walk.ut;
        walk.advance(8);
        if (r < r1) {
          VecF<4>::load(z + r * c_total + c0, zz[u]);
          if (read_dy) VecF<4>::load(dy + r * c_total + c0, dd[u]);
        }
      }
      // the U rows of a group are summed in fp32 (at most U terms), the running sums stay fp64
      float p0[4] = {0.f, 0.f, 0.f, 0.f}, p1[4] = {0.f, 0.f, 0.f, 0.f}, p2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!ok[u]) continue;
        if (KIND == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            p0[j] += zz[u][j];
            p1[j] = fmaf(zz[u][j], zz[u][j], p1[j]);
          }
        } else {
          if (pg.enabled) {
            const int64_t ut = uts[u];
            if (ut != cur_u) {                        // per-utterance pooling coefficients: reloaded when the row walk enters a new utterance
              cur_u = ut;
              pool_coef_load(pg, ut, c_total, c0, &pc);
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float xh = fmaf(zz[u][j], bf[j].xa, bf[j].xb), pre = fmaf(zz[u][j], bf[j].pa, bf[j].pb);
            float d = dd[u][j];
            if (pg.enabled) d = fmaf(pc.cb[j], (pre >= 0.f ? pre : pre * sl[j]) - pc.mu[j], pc.ca[j]);
            const float g = pre > 0.f ? d : d * sl[j];
            p0[j] += g;
            p1[j] = fmaf(g, xh, p1[j]);
            if (pre < 0.f) p2[j] = fmaf(d, pre, p2[j]);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[0][j] += p0[j]; v[1][j] += p1[j];
        if (KIND == 1) v[2][j] += p2[j];
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) red[k][ry][cx * 4 + j] = v[k][j];
  __syncthreads();
  for (int i = threadIdx.x; i < NV * 128; i += 256) {
    const int k = i >> 7, cc = i & 127;
    const int c = blockIdx.x * 128 + cc;
    if (c >= c_total) continue;
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) sum += red[k][j][cc];
    atomicAdd(out + static_cast<int64_t>(k) * out_ld + c, sum);
  }
}

template <int KIND>
cudaError_t launch_col_reduce4(const float* dy, const float* z, RowMask mask, int64_t rows, int c_total, const float* mean,
                               const float* rstd, const float* gamma, const float* beta, const float* slope, const PoolGrad& pg,
                               double* out, int out_ld, int sm_count, cudaStream_t s) {
  const int bx = (c_total + 127) / 128;
  int64_t chunks = (2 * sm_count + bx - 1) / bx;       // ~2 blocks per SM, each thread walks >= 16 rows in groups of 4
  int64_t rpb = (rows + chunks - 1) / chunks;
  if (rpb < 32) rpb = 32;
  dim3 grid(bx, static_cast<unsigned>((rows + rpb - 1) / rpb));
  col_reduce4_kernel<KIND><<<grid, 256, 0, s>>>(dy, z, mask, rows, c_total, rpb, mean, rstd, gamma, beta, slope, pg, out, out_ld);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ frame-level elementwise
// y = act(BN(z)) on valid rows, 0 elsewhere; optional fp32 output, optional bf16 operand copy (pad columns zeroed).
// The batch statistics arrive as fp64 sums; every thread derives mean / rstd of its four channels itself, block 0 also
// leaves them (and the new moving averages) behind for the backward pass / the optimizer.
struct BnFwdArgs {
  const float* z; int64_t rows; int32_t c_total; RowMask mask;
  BnStat st; BnFinalize fin;
  const float* gamma; const float* beta; const float* slope;
  float* y; __nv_bfloat16* yb; int32_t ldb;
};
__global__ void __launch_bounds__(256, 2) bn_act_fwd_kernel(const BnFwdArgs a) {
  constexpr int U = 4;
  const int nq = a.c_total / 4;
  // block b owns the rows [b*rpb, (b+1)*rpb): few blocks (2 per SM), many rows per thread -- the per-channel set-up (fp64
  // divide + square root) is paid once per block, not once per handful of rows
  const int64_t rpb = (a.rows + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rpb, r1 = min(r0 + rpb, a.rows);
  for (int q = threadIdx.x; q < nq; q += blockDim.x) {
    const int c0 = q * 4;
    float sl[4];
    BnFold bf[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      bf[k] = bn_fold(false, 0.f, 1.f, 1.f, 0.f);
      if (a.st.sums != nullptr) {
        double var;
        float mu, rs;
        bn_mean_rstd(a.st, c0 + k, a.fin.eps, &mu, &rs, &var);
        bf[k] = bn_fold(true, mu, rs, a.gamma[c0 + k], a.beta[c0 + k]);
        if (blockIdx.x == 0 && threadIdx.y == 0) bn_finalize_channel(a.st, a.fin, c0 + k, mu, rs, var);
      }
      sl[k] = a.slope[c0 + k];
    }
    RowWalk walk(a.mask, r0 + threadIdx.y);
    for (int64_t rb = r0 + threadIdx.y; rb < r1; rb += static_cast<int64_t>(blockDim.y) * U) {
      float v[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t r = rb + static_cast<int64_t>(u) * blockDim.y;
        if (r < r1) VecF<4>::load(a.z + r * a.c_total + c0, v[u]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t r = rb + static_cast<int64_t>(u) * blockDim.y;
        const bool ok = walk.ok();
        walk.advance(blockDim.y);
        if (r >= r1) continue;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float t = fmaf(v[u][k], bf[k].pa, bf[k].pb);                          // same expression in every backward kernel
          v[u][k] = ok ? (t >= 0.f ? t : t * sl[k]) : 0.f;
        }
        if (a.y != nullptr) VecF<4>::store(a.y + r * a.c_total + c0, v[u]);
        if (a.yb != nullptr) {
          VecF<4>::store_bf16(a.yb + r * a.ldb + c0, v[u]);
          if (q == nq - 1)
            for (int k = a.c_total; k < a.ldb; ++k) a.yb[r * a.ldb + k] = __float2bfloat16(0.f);
        }
      }
    }
  }
}

// Last frame layer: normalise + activate + statistics pooling in one pass over z; y is never stored.
// grid (ceil(C/128), B), block (32, 8): thread (tx, ty) owns four channels and the rows ty, ty+8, ... of utterance
// blockIdx.y.  Sums are taken about the utterance's first frame (pivot) in fp64: mean = p + S1/L, var = S2/L - (S1/L)^2.
__global__ void __launch_bounds__(256) bn_act_pool_fwd_kernel(const BnFwdArgs a, float* __restrict__ pooled) {
  __shared__ double red[2][8][129];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c0 = (blockIdx.x * 32 + tx) * 4;
  const int u = blockIdx.y;
  const int64_t base = static_cast<int64_t>(u) * a.mask.t;
  double s1[4] = {0.0, 0.0, 0.0, 0.0}, s2[4] = {0.0, 0.0, 0.0, 0.0};
  float piv[4] = {0.f, 0.f, 0.f, 0.f};
  if (c0 < a.c_total) {
    float sl[4];
    BnFold bf[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      bf[k] = bn_fold(false, 0.f, 1.f, 1.f, 0.f);
      if (a.st.sums != nullptr) {
        double var;
        float mu, rs;
        bn_mean_rstd(a.st, c0 + k, a.fin.eps, &mu, &rs, &var);
        bf[k] = bn_fold(true, mu, rs, a.gamma[c0 + k], a.beta[c0 + k]);
        if (u == 0 && ty == 0) bn_finalize_channel(a.st, a.fin, c0 + k, mu, rs, var);
      }
      sl[k] = a.slope[c0 + k];
    }
    auto act_row = [&](int64_t r, float* v) {
      VecF<4>::load(a.z + r * a.c_total + c0, v);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float t = fmaf(v[k], bf[k].pa, bf[k].pb);
        v[k] = t >= 0.f ? t : t * sl[k];
      }
    };
    act_row(base, piv);
    for (int i0 = ty; i0 < a.mask.valid_len; i0 += 8 * 4) {
      float v[4][4];
#pragma unroll
      for (int u = 0; u < 4; ++u)                       // four independent 16-byte loads, then the math
        if (i0 + 8 * u < a.mask.valid_len) VecF<4>::load(a.z + (base + i0 + 8 * u) * a.c_total + c0, v[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (i0 + 8 * u >= a.mask.valid_len) continue;
        float p1[4], p2[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float t = fmaf(v[u][k], bf[k].pa, bf[k].pb);
          t = t >= 0.f ? t : t * sl[k];
          const float d = t - piv[k];
          p1[k] = d; p2[k] = d * d;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { s1[k] += p1[k]; s2[k] += p2[k]; }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) { red[0][ty][tx * 4 + k] = s1[k]; red[1][ty][tx * 4 + k] = s2[k]; }
  __syncthreads();
  if (ty == 0 && c0 < a.c_total) {
    const double inv = 1.0 / static_cast<double>(a.mask.valid_len);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      double t1 = 0.0, t2 = 0.0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { t1 += red[0][j][tx * 4 + k]; t2 += red[1][j][tx * 4 + k]; }
      const double m = t1 * inv;
      float var = static_cast<float>(t2 * inv - m * m);
      var = var <= 1e-12f ? 1e-12f : var;               // VAR2STD_EPSILON floor (model/pooling.py:46-47)
      pooled[static_cast<int64_t>(u) * 2 * a.c_total + c0 + k] = static_cast<float>(static_cast<double>(piv[k]) + m);
      pooled[static_cast<int64_t>(u) * 2 * a.c_total + a.c_total + c0 + k] = sqrtf(var);
    }
  }
}

// dz = gamma*rstd*(g - sum_g/n - xhat*sum_gx/n) on valid rows, 0 elsewhere.  The incoming gradient is read from `dy` or
// rebuilt from the pooling gradient (pg.enabled).  Outputs: dz32 (fp32, may alias dy) and / or dzb (bf16 GEMM operand).
// Block 0 also writes the parameter gradients the sums hold: d beta = sum g, d gamma = sum g*xhat, d alpha (PReLU).
struct BnBwdArgs {
  const float* dy; const float* z; int64_t rows; int32_t c_total; RowMask mask;
  const float* mean; const float* rstd; const float* gamma; const float* beta; const float* slope;   // mean == null: no BN
  const double* sums; int32_t sums_ld; double n;
  PoolGrad pg;
  float* dz32; __nv_bfloat16* dzb; int32_t ldb;
  float* g_gamma; float* g_beta; float* g_alpha;     // may be null
};
__global__ void __launch_bounds__(256, 2) bn_bwd_apply_kernel(const BnBwdArgs a) {
  constexpr int U = 4;
  const int nq = a.c_total / 4;
  const int64_t rpb = (a.rows + gridDim.x - 1) / gridDim.x;        // contiguous row range per block (see bn_act_fwd_kernel)
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rpb, r1 = min(r0 + rpb, a.rows);
  const double inv_n = 1.0 / a.n;
  for (int q = threadIdx.x; q < nq; q += blockDim.x) {
    const int c0 = q * 4;
    float sl[4];
    BnFold bf[4];
    // dz = gr * (g - k1 - xhat * k2) = g * gr + (e0 + xhat * e1): the fp64 sums only enter through the three constants
    float grf[4], e0[4], e1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      sl[k] = a.slope[c0 + k];
      bf[k] = bn_fold(false, 0.f, 1.f, 1.f, 0.f);
      grf[k] = 1.f; e0[k] = 0.f; e1[k] = 0.f;
      if (a.mean != nullptr) {
        const float ga = a.gamma[c0 + k], rs = a.rstd[c0 + k];
        bf[k] = bn_fold(true, a.mean[c0 + k], rs, ga, a.beta[c0 + k]);
        const double gr = static_cast<double>(ga) * rs;
        grf[k] = static_cast<float>(gr);
        e0[k] = static_cast<float>(-gr * (a.sums[c0 + k] * inv_n));
        e1[k] = static_cast<float>(-gr * (a.sums[a.sums_ld + c0 + k] * inv_n));
      }
      if (blockIdx.x == 0 && threadIdx.y == 0) {
        if (a.g_beta != nullptr) a.g_beta[c0 + k] = static_cast<float>(a.sums[c0 + k]);
        if (a.g_gamma != nullptr) a.g_gamma[c0 + k] = static_cast<float>(a.sums[a.sums_ld + c0 + k]);
        if (a.g_alpha != nullptr) a.g_alpha[c0 + k] = static_cast<float>(a.sums[2 * a.sums_ld + c0 + k]);
      }
    }
    int64_t cur_u = -1;
    PoolCoef pc;
#pragma unroll
    for (int k = 0; k < 4; ++k) { pc.mu[k] = 0.f; pc.ca[k] = 0.f; pc.cb[k] = 0.f; }
    RowWalk walk(a.mask, r0 + threadIdx.y);
    for (int64_t rb = r0 + threadIdx.y; rb < r1; rb += static_cast<int64_t>(blockDim.y) * U) {
      float zz[U][4], d[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t r = rb + static_cast<int64_t>(u) * blockDim.y;
        if (r < r1) {
          VecF<4>::load(a.z + r * a.c_total + c0, zz[u]);
          if (!a.pg.enabled) VecF<4>::load(a.dy + r * a.c_total + c0, d[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t r = rb + static_cast<int64_t>(u) * blockDim.y;
        const bool row_ok = walk.ok();
        const int64_t ut = walk.ut;
        walk.advance(blockDim.y);
        if (r >= r1) continue;
        float out[4] = {0.f, 0.f, 0.f, 0.f};
        if (row_ok) {
          if (a.pg.enabled) {
            if (ut != cur_u) {
              cur_u = ut;
              pool_coef_load(a.pg, ut, a.c_total, c0, &pc);
            }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float xh = fmaf(zz[u][k], bf[k].xa, bf[k].xb), pre = fmaf(zz[u][k], bf[k].pa, bf[k].pb);
            float dk = d[u][k];
            if (a.pg.enabled) dk = fmaf(pc.cb[k], (pre >= 0.f ? pre : pre * sl[k]) - pc.mu[k], pc.ca[k]);
            const float g = pre > 0.f ? dk : dk * sl[k];
            out[k] = fmaf(g, grf[k], fmaf(xh, e1[k], e0[k]));          // no batch norm: grf = 1, e0 = e1 = 0
          }
        }
        if (a.dz32 != nullptr) VecF<4>::store(a.dz32 + r * a.c_total + c0, out);
        if (a.dzb != nullptr) {
          VecF<4>::store_bf16(a.dzb + r * a.ldb + c0, out);
          if (q == nq - 1)
            for (int k = a.c_total; k < a.ldb; ++k) a.dzb[r * a.ldb + k] = __float2bfloat16(0.f);
        }
      }
    }
  }
}

// Launch geometry of the row-streaming kernels: x covers the channel groups, y the rows.
struct RowGrid { dim3 grid, block; };
inline RowGrid row_grid(int groups, int sm_count, int blocks_per_sm = 2) {
  int bx = 32;
  while (bx < groups && bx < 256) bx <<= 1;
  RowGrid g;
  g.block = dim3(bx, 256 / bx);
  g.grid = dim3(sm_count * blocks_per_sm);       // few fat blocks: see bn_act_fwd_kernel
  return g;
}

// ------------------------------------------------------------------------------------------------ segment-level layers
// (rows = utterances of the step, a few dozen: one kernel does statistics AND normalisation)
// Forward: y = act(BN_batch(z)); one block per 32 channels, block (32, 8).
struct SegFwdArgs {
  const float* z; int32_t rows, c_total;
  int32_t has_bn;
  BnFinalize fin;
  const float* gamma; const float* beta; const float* slope;
  float* y;
};
__global__ void __launch_bounds__(256) seg_bn_act_fwd_kernel(const SegFwdArgs a) {
  __shared__ double red[2][8][33];
  __shared__ float sh_mu[32], sh_rs[32];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
  const bool ok = c < a.c_total;
  if (a.has_bn) {
    double v0 = 0.0, v1 = 0.0;
    if (ok)
      for (int r = ty; r < a.rows; r += 8) {
        const double x = a.z[static_cast<int64_t>(r) * a.c_total + c];
        v0 += x; v1 += x * x;
      }
    red[0][ty][tx] = v0; red[1][ty][tx] = v1;
    __syncthreads();
    if (ty == 0 && ok) {
      double t0 = 0.0, t1 = 0.0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { t0 += red[0][j][tx]; t1 += red[1][j][tx]; }
      red[0][0][tx] = t0; red[1][0][tx] = t1;
      BnStat st;
      st.sums = &red[0][0][0]; st.ld = static_cast<int32_t>(&red[1][0][0] - &red[0][0][0]); st.n = static_cast<double>(a.rows);
      float mu, rs;
      double var;
      bn_mean_rstd(st, tx, a.fin.eps, &mu, &rs, &var);
      sh_mu[tx] = mu; sh_rs[tx] = rs;
      // bn_finalize_channel with the block-local sums (index tx) and the global channel index c
      a.fin.mean[c] = mu;
      a.fin.rstd[c] = rs;
      const double m = t0 / st.n;
      const double vm = a.fin.unbiased ? var * (st.n / fmax(st.n - 1.0, 1.0)) : var;
      a.fin.new_mm[c] = static_cast<float>(static_cast<double>(a.fin.mov_mean[c]) * a.fin.momentum + m * (1.0 - a.fin.momentum));
      a.fin.new_mv[c] = static_cast<float>(static_cast<double>(a.fin.mov_var[c]) * a.fin.momentum + vm * (1.0 - a.fin.momentum));
    }
    __syncthreads();
  }
  if (!ok) return;
  const float mu = a.has_bn ? sh_mu[tx] : 0.f, rs = a.has_bn ? sh_rs[tx] : 1.f;
  const float ga = a.has_bn ? a.gamma[c] : 1.f, be = a.has_bn ? a.beta[c] : 0.f, sl = a.slope[c];
  for (int r = ty; r < a.rows; r += 8) {
    float t = a.z[static_cast<int64_t>(r) * a.c_total + c];
    if (a.has_bn) t = fmaf(ga, (t - mu) * rs, be);
    a.y[static_cast<int64_t>(r) * a.c_total + c] = t >= 0.f ? t : t * sl;
  }
}

// Backward: dz = BN'(act'(dy)) in place over dy, parameter gradients written directly.
struct SegBwdArgs {
  float* dy; const float* z; int32_t rows, c_total;
  const float* mean; const float* rstd; const float* gamma; const float* beta; const float* slope;   // mean == null: no BN
  float* g_gamma; float* g_beta; float* g_alpha; float* g_bias;     // g_gamma/g_beta/g_alpha may be null
};
__global__ void __launch_bounds__(256) seg_bn_bwd_kernel(const SegBwdArgs a) {
  __shared__ double red[3][8][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
  const bool ok = c < a.c_total;
  const bool bn = a.mean != nullptr;
  float mu = 0.f, rs = 1.f, ga = 1.f, be = 0.f, sl = 0.f;
  if (ok) {
    sl = a.slope[c];
    if (bn) { mu = a.mean[c]; rs = a.rstd[c]; ga = a.gamma[c]; be = a.beta[c]; }
  }
  double v0 = 0.0, v1 = 0.0, v2 = 0.0;
  if (ok)
    for (int r = ty; r < a.rows; r += 8) {
      const float zz = a.z[static_cast<int64_t>(r) * a.c_total + c];
      const float d = a.dy[static_cast<int64_t>(r) * a.c_total + c];
      float xh = 0.f, pre = zz;
      if (bn) { xh = (zz - mu) * rs; pre = fmaf(ga, xh, be); }
      const float g = pre > 0.f ? d : d * sl;
      v0 += g; v1 += static_cast<double>(g) * xh;
      if (pre < 0.f) v2 += static_cast<double>(d) * pre;
    }
  red[0][ty][tx] = v0; red[1][ty][tx] = v1; red[2][ty][tx] = v2;
  __syncthreads();
  if (!ok) return;
  double t0 = 0.0, t1 = 0.0, t2 = 0.0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { t0 += red[0][j][tx]; t1 += red[1][j][tx]; t2 += red[2][j][tx]; }
  if (ty == 0) {
    if (a.g_beta != nullptr) a.g_beta[c] = static_cast<float>(t0);
    if (a.g_gamma != nullptr) a.g_gamma[c] = static_cast<float>(t1);
    if (a.g_alpha != nullptr) a.g_alpha[c] = static_cast<float>(t2);
    // a bias in front of a batch norm: sum_rows dz = -gamma*rstd*(sum xhat)*(sum g*xhat)/n with sum xhat == 0, i.e. the
    // gradient is identically zero (TF evaluates the same sum and gets rounding residue of the order 1e-9); without
    // batch norm dz = g and the bias gradient is its column sum
    a.g_bias[c] = bn ? 0.f : static_cast<float>(t0);
  }
  const double n = static_cast<double>(a.rows);
  const double k1 = t0 / n, k2 = t1 / n, gr = static_cast<double>(ga) * rs;
  for (int r = ty; r < a.rows; r += 8) {
    const float zz = a.z[static_cast<int64_t>(r) * a.c_total + c];
    const float d = a.dy[static_cast<int64_t>(r) * a.c_total + c];
    float out;
    if (bn) {
      const float xh = (zz - mu) * rs;
      const float pre = fmaf(ga, xh, be);
      const float g = pre > 0.f ? d : d * sl;
      out = static_cast<float>(gr * (static_cast<double>(g) - k1 - static_cast<double>(xh) * k2));
    } else {
      out = zz > 0.f ? d : d * sl;
    }
    a.dy[static_cast<int64_t>(r) * a.c_total + c] = out;
  }
}

// ------------------------------------------------------------------------------------------------ bf16 operand copies
// Master weights stay fp32 in the flat buffer; after every update ONE launch refreshes the bf16 layouts the
// tensor-core GEMMs read.  A job is one (layer, layout); its 32 x 32 tiles are numbered after the previous job's.
//   kind 0, forward operand: wt[n][k] (K-major), k = j*cstride + c  ->  W[j, c, n]; zero for n >= cout, c >= cin, j >= taps.
//   kind 1, data-gradient operand: wd[c][j*ldz + n] = W[j, c, n]; zero for c >= cin, n >= cout.
struct RecastJob {
  const float* w; __nv_bfloat16* dst;
  int32_t kind, taps, cin, cout, cstride, ktotal, n_alloc, ldz, c_alloc;
  int32_t tiles_x, tile0;          // tiles along the fast axis; first global tile index of this job
};
constexpr int kMaxRecastJobs = 16;
struct RecastJobs { RecastJob j[kMaxRecastJobs]; int32_t n, total_tiles; };

__global__ void __launch_bounds__(256) recast_weights_kernel(const RecastJobs jobs) {
  __shared__ float tile[64][33];
  for (int t = blockIdx.x; t < jobs.total_tiles; t += gridDim.x) {
    int ji = 0;
    while (ji + 1 < jobs.n && jobs.j[ji + 1].tile0 <= t) ++ji;
    const RecastJob& J = jobs.j[ji];
    const int lt = t - J.tile0;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    if (J.kind == 0) {
      // 64 (k) x 32 (n) tile: 8 independent loads per thread along n (coalesced), bf16 PAIRS along k on the way out
      const int k0 = (lt % J.tiles_x) * 64, n0 = (lt / J.tiles_x) * 32;
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int k = k0 + ty + 8 * u, n = n0 + tx;
        const int j = k / J.cstride, c = k - j * J.cstride;
        v[u] = (k < J.ktotal && j < J.taps && c < J.cin && n < J.cout) ? J.w[(static_cast<int64_t>(j) * J.cin + c) * J.cout + n] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) tile[ty + 8 * u][tx] = v[u];
      __syncthreads();
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int n = n0 + ty + 8 * u, k = k0 + 2 * tx;
        if (n < J.n_alloc && k < J.ktotal)        // ktotal is a multiple of 64: the pair never straddles the end
          *reinterpret_cast<__nv_bfloat162*>(J.dst + static_cast<int64_t>(n) * J.ktotal + k) =
              __floats2bfloat162_rn(tile[2 * tx][ty + 8 * u], tile[2 * tx + 1][ty + 8 * u]);
      }
      __syncthreads();
    } else {
      // rows = (c, j) pairs, columns = n: contiguous on both sides, no transpose
      const int row_len = J.ldz;
      const int n0 = (lt % J.tiles_x) * 32;
      const int64_t row0 = static_cast<int64_t>(lt / J.tiles_x) * 32;
      const int64_t n_rows = static_cast<int64_t>(J.c_alloc) * J.taps;
      for (int i = ty; i < 32; i += 8) {
        const int64_t q = row0 + i;
        const int n = n0 + tx;
        if (q >= n_rows || n >= row_len) continue;
        const int j = static_cast<int>(q % J.taps);
        const int c = static_cast<int>(q / J.taps);
        float v = 0.f;
        if (c < J.cin && n < J.cout) v = J.w[(static_cast<int64_t>(j) * J.cin + c) * J.cout + n];
        J.dst[q * row_len + n] = __float2bfloat16(v);
      }
    }
  }
}

__global__ void axpy_rows_kernel(float* __restrict__ y, const float* __restrict__ x, const float* __restrict__ coef,
                                 int64_t rows, int width) {
  const int64_t total = rows * width;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    y[i] = fmaf(coef[i / width], x[i], y[i]);
}

// ------------------------------------------------------------------------------------------------ loss
// r[s] = rsqrt(max(sum_e w[e,s]^2, 1e-12))  (tf.nn.l2_normalize(w, dim=0), model/loss.py:241)
__global__ void __launch_bounds__(256) col_rnorm_kernel(const float* __restrict__ w, int e_dim, int s_dim, float* __restrict__ r) {
  __shared__ double red[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int s = blockIdx.x * 32 + cx;
  double ss = 0.0;
  if (s < s_dim)
    for (int e = ry; e < e_dim; e += 8) {
      const double v = w[static_cast<int64_t>(e) * s_dim + s];
      ss += v * v;
    }
  red[ry][cx] = ss;
  __syncthreads();
  if (ry == 0 && s < s_dim) {
    double t = 0.0;
    for (int j = 0; j < 8; ++j) t += red[j][cx];
    r[s] = static_cast<float>(1.0 / sqrt(fmax(t, 1e-12)));
  }
}

// sh: one slot per warp of the block (<= 32)
__device__ __forceinline__ double block_reduce(double v, double* sh, bool is_max) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double other = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmax(v, other) : v + other;
  }
  __syncthreads();
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  v = sh[0];
  for (int k = 1; k < static_cast<int>(blockDim.x >> 5); ++k) v = is_max ? fmax(v, sh[k]) : v + sh[k];
  return v;
}

// One block per utterance.  logits[b, :] = f.wn (AM-softmax) or f.W + bias (softmax) on entry; on exit it holds
// G = d(loss)/d(base logits); gr = G * r[s] (AM-softmax) for the feature gradient; coef[b] = (dloss/d||f||)/||f||.
constexpr int kLossThreads = 1024;      // one block per utterance: the fp64 exp / log passes over S logits are spread over 32 warps
__global__ void __launch_bounds__(kLossThreads) softmax_loss_kernel(float* __restrict__ logits, int s_dim, const float* __restrict__ f,
                                                           int e_dim, const int32_t* __restrict__ labels, int batch,
                                                           const float* __restrict__ r, int amsoftmax, float margin, float fa,
                                                           float* __restrict__ gr, float* __restrict__ coef,
                                                           double* __restrict__ loss_sum) {
  __shared__ double sh[32];
  const int b = blockIdx.x;
  float* row = logits + static_cast<int64_t>(b) * s_dim;
  const int lab = labels[b];
  double sel = 0.0, fn_raw = 0.0, fn = 0.0, cosc = 0.0, new_label = 0.0;
  bool clipped = false;
  if (amsoftmax) {
    double ss = 0.0;
    for (int e = threadIdx.x; e < e_dim; e += blockDim.x) {
      const double v = f[static_cast<int64_t>(b) * e_dim + e];
      ss += v * v;
    }
    ss = block_reduce(ss, sh, false);
    const double eps = 1e-12;
    fn_raw = sqrt(ss);
    fn = fmax(fn_raw, eps);
    sel = row[lab];
    const double cs = sel / fn;
    clipped = cs < -1.0 + eps || cs > 1.0 - eps;
    cosc = fmin(fmax(cs, -1.0 + eps), 1.0 - eps);
    new_label = sel + fa * ((cosc - margin) * fn - sel);       // fs*logits + fa*logits_amsoftmax at the label
    __syncthreads();
    if (threadIdx.x == 0) row[lab] = static_cast<float>(new_label);
    __syncthreads();
  } else {
    new_label = row[lab];
  }
  double mx = -1e300;
  for (int s = threadIdx.x; s < s_dim; s += blockDim.x) mx = fmax(mx, static_cast<double>(row[s]));
  mx = block_reduce(mx, sh, true);
  double z = 0.0;
  for (int s = threadIdx.x; s < s_dim; s += blockDim.x) z += exp(static_cast<double>(row[s]) - mx);
  z = block_reduce(z, sh, false);
  const double lse = mx + log(z);
  const double inv_b = 1.0 / batch;
  double g_label = 0.0;
  for (int s = threadIdx.x; s < s_dim; s += blockDim.x) {
    double g = exp(static_cast<double>(row[s]) - lse);
    if (s == lab) {
      g -= 1.0;
      g *= inv_b;
      g_label = g;
      if (amsoftmax && clipped) g *= (1.0 - fa);
    } else {
      g *= inv_b;
    }
    row[s] = static_cast<float>(g);
    if (gr != nullptr) gr[static_cast<int64_t>(b) * s_dim + s] = static_cast<float>(g * (r != nullptr ? r[s] : 1.f));
  }
  g_label = block_reduce(g_label, sh, false);                   // only the label's owner contributed
  if (threadIdx.x == 0) {
    atomicAdd(loss_sum, (lse - new_label) * inv_b);
    double cf = 0.0;
    if (amsoftmax && fn_raw > 1e-12) cf = (clipped ? g_label * fa * (cosc - margin) : -g_label * fa * margin) / fn_raw;
    coef[b] = static_cast<float>(cf);
  }
}

// Softmax-layer weight gradient in one pass.  Block = a strip of 32 speaker columns x all E rows of W [E, S]:
//   dwn[e, s] = sum_b f[b, e] * G[b, s]                                   (G = d loss / d logits, [B, S])
//   AM-softmax (r != null): wn = W * r_s;  dW = r_s * (dwn - wn * sum_e' dwn[e', s] * wn[e', s])   (tf.nn.l2_normalize
//     Jacobian, model/loss.py:241; flat where sum w^2 <= 1e-12);  plain softmax: dW = dwn, db[s] = sum_b G[b, s].
// Thread (cx, ry) keeps the E/8 accumulators of rows ry*E/8 .. of column s0 + cx in registers; f is staged through
// shared memory 16 utterances at a time and read as broadcast float4.  Block 0 also copies the tower loss into the
// tail slot of the gradient buffer (it rides in the same allreduce).
constexpr int kHeadMaxE = 512;
__global__ void __launch_bounds__(256) head_wgrad_kernel(const float* __restrict__ f, const float* __restrict__ G, int batch,
                                                         int e_dim, int s_dim, const float* __restrict__ w,
                                                         const float* __restrict__ r, float* __restrict__ dw,
                                                         float* __restrict__ db, const double* __restrict__ loss_src,
                                                         float* __restrict__ loss_slot) {
  __shared__ __align__(16) float fs[16 * kHeadMaxE];
  __shared__ float gs[16][32];
  __shared__ double red[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int s0 = blockIdx.x * 32;
  const int s = s0 + cx;
  const bool ok = s < s_dim;
  const int ept = e_dim >> 3;                       // rows per thread (multiple of 4, <= 64)
  const int e_base = ry * ept;
  float acc[64];
#pragma unroll
  for (int k = 0; k < 64; ++k) acc[k] = 0.f;
  double gsum = 0.0;
  if (blockIdx.x == 0 && threadIdx.x == 0 && loss_slot != nullptr) loss_slot[0] = static_cast<float>(loss_src[0]);
  for (int b0 = 0; b0 < batch; b0 += 16) {
    const int nb = min(16, batch - b0);
    __syncthreads();
    // stage 16 utterances of f and of this strip of G: the inner loop then touches shared memory only
    {
      // the nb rows of f are one contiguous run of nb*e_dim floats: every thread issues its (up to 8) 16-byte loads
      // back to back, then stores -- a load -> store loop with run-time bounds exposes one L2 round trip per element
      const float4* src = reinterpret_cast<const float4*>(f + static_cast<int64_t>(b0) * e_dim);
      const int n4 = nb * e_dim / 4;
      float4 tmp[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i4 = threadIdx.x + 256 * j;
        if (i4 < n4) tmp[j] = src[i4];
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i4 = threadIdx.x + 256 * j;
        if (i4 < n4) {
          const int gi = i4 * 4;
          const int bb = gi / e_dim, e = gi - bb * e_dim;
          *reinterpret_cast<float4*>(fs + bb * kHeadMaxE + e) = tmp[j];
        }
      }
    }
    for (int i = threadIdx.x; i < nb * 32; i += 256) {
      const int bb = i >> 5, c = i & 31;
      gs[bb][c] = s0 + c < s_dim ? G[static_cast<int64_t>(b0 + bb) * s_dim + s0 + c] : 0.f;
    }
    __syncthreads();
    for (int bb = 0; bb < nb; ++bb) {
      const float g = gs[bb][cx];
      gsum += g;
      const float4* fp = reinterpret_cast<const float4*>(fs + bb * kHeadMaxE + e_base);
#pragma unroll
      for (int k4 = 0; k4 < 16; ++k4) {
        if (k4 * 4 < ept) {
          const float4 f4 = fp[k4];
          acc[k4 * 4 + 0] = fmaf(f4.x, g, acc[k4 * 4 + 0]);
          acc[k4 * 4 + 1] = fmaf(f4.y, g, acc[k4 * 4 + 1]);
          acc[k4 * 4 + 2] = fmaf(f4.z, g, acc[k4 * 4 + 2]);
          acc[k4 * 4 + 3] = fmaf(f4.w, g, acc[k4 * 4 + 3]);
        }
      }
    }
  }
  if (r == nullptr) {
    if (ok) {
#pragma unroll
      for (int k = 0; k < 64; ++k)
        if (k < ept) dw[static_cast<int64_t>(e_base + k) * s_dim + s] = acc[k];
      if (ry == 0 && db != nullptr) db[s] = static_cast<float>(gsum);
    }
    return;
  }
  const double rs = ok ? static_cast<double>(r[s]) : 0.0;
  const float* wcol = w + static_cast<int64_t>(e_base) * s_dim + s;
  double dot = 0.0;
  if (ok) {
#pragma unroll
    for (int k0 = 0; k0 < 64; k0 += 16) {           // 16 independent loads in flight, then the fp64 chain
      float wv[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) wv[j] = k0 + j < ept ? wcol[static_cast<int64_t>(k0 + j) * s_dim] : 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) dot += static_cast<double>(acc[k0 + j]) * (static_cast<double>(wv[j]) * rs);
    }
  }
  red[ry][cx] = dot;
  __syncthreads();
  if (!ok) return;
  dot = 0.0;
#pragma unroll
  for (int j = 0; j < 8; ++j) dot += red[j][cx];
  const bool floored = rs >= 0.999999e6;                         // sum w^2 <= 1e-12: l2_normalize's max() is flat
  float* dcol = dw + static_cast<int64_t>(e_base) * s_dim + s;
#pragma unroll
  for (int k0 = 0; k0 < 64; k0 += 16) {
    float wv[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) wv[j] = k0 + j < ept ? wcol[static_cast<int64_t>(k0 + j) * s_dim] : 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (k0 + j < ept) {
        const double gval = floored ? rs * acc[k0 + j]
                                    : rs * (static_cast<double>(acc[k0 + j]) - static_cast<double>(wv[j]) * rs * dot);
        dcol[static_cast<int64_t>(k0 + j) * s_dim] = static_cast<float>(gval);
      }
    }
  }
}

// Fallback for embedding widths the strip kernel does not take: the l2_normalize Jacobian in place over dwn.
__global__ void __launch_bounds__(256) amsoftmax_dw_kernel(float* __restrict__ dwn, const float* __restrict__ w,
                                                           const float* __restrict__ r, int e_dim, int s_dim) {
  __shared__ double red[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int s = blockIdx.x * 32 + cx;
  const double rs = s < s_dim ? r[s] : 0.0;
  double dot = 0.0;
  if (s < s_dim)
    for (int e = ry; e < e_dim; e += 8) {
      const int64_t i = static_cast<int64_t>(e) * s_dim + s;
      dot += static_cast<double>(dwn[i]) * (static_cast<double>(w[i]) * rs);
    }
  red[ry][cx] = dot;
  __syncthreads();
  dot = 0.0;
  for (int j = 0; j < 8; ++j) dot += red[j][cx];
  if (s >= s_dim) return;
  const bool floored = rs >= 0.999999e6;
  for (int e = ry; e < e_dim; e += 8) {
    const int64_t i = static_cast<int64_t>(e) * s_dim + s;
    const double g = floored ? rs * dwn[i] : rs * (static_cast<double>(dwn[i]) - static_cast<double>(w[i]) * rs * dot);
    dwn[i] = static_cast<float>(g);
  }
}
// column sums of a [rows, n] matrix -> out[n] (softmax bias gradient in the fallback path)
__global__ void __launch_bounds__(256) col_sum_small_kernel(const float* __restrict__ x, int rows, int n, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  double s = 0.0;
  for (int r = 0; r < rows; ++r) s += x[static_cast<int64_t>(r) * n + c];
  out[c] = static_cast<float>(s);
}
__global__ void loss_to_tail_kernel(const double* __restrict__ scalars, float* __restrict__ slot) { slot[0] = static_cast<float>(scalars[0]); }

// ------------------------------------------------------------------------------------------------ skinny GEMMs
// The segment-level layers and the softmax layer multiply a [B, K] matrix with B = the utterances of a step (64): too
// few rows for a 128-row tensor-core tile to pay, and fp32 keeps the loss head exact.  One block = 64 rows x 64 columns
// x one K chunk; the K chunks of a tile are spread over blocks (split-K) and added with fp32 atomics into a zeroed
// output, so that ~2 blocks per SM stream the weight matrix once at full bandwidth.
//   TB = false: out[m, n] += cs[n] * sum_k a[m, k] * w[k, n]      (forward:  w = kernel [K, N])
//   TB = true : out[m, n] += cs[n] * sum_k a[m, k] * w[n, k]      (backward: w = kernel [N, K], contraction over its columns)
struct SkinnyArgs {
  const float* a; int32_t lda;
  const float* w; int32_t ldw;
  float* out; int32_t ldo;
  const float* col_scale;    // [n] or null
  const float* bias;         // [n] or null, added by the first K chunk
  int32_t m, n, k, ks;       // ks: K chunk per block (multiple of 64)
};

template <bool TB>
__global__ void __launch_bounds__(256) skinny_gemm_kernel(const SkinnyArgs g) {
  __shared__ float As[64][65];
  __shared__ __align__(16) float Bs[64][68];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;       // 16 x 16 threads, 4 x 4 outputs each
  const int n0 = blockIdx.x * 64, m0 = blockIdx.z * 64;
  const int kbeg = blockIdx.y * g.ks, kend = min(kbeg + g.ks, g.k);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = kbeg; k0 < kend; k0 += 64) {
    // every thread first issues its 32 loads of this K step, then stores: a load -> store loop pays one memory round trip
    // per unrolled group, and these GEMMs are nothing but latency (64 rows against a weight matrix streamed once)
    float av[16], bv[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = i * 256 + threadIdx.x;
      const int mm = idx >> 6, kk = idx & 63;
      const int m = m0 + mm, k = k0 + kk;
      av[i] = (m < g.m && k < kend) ? g.a[static_cast<int64_t>(m) * g.lda + k] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = i * 256 + threadIdx.x;
      if (TB) {
        const int nn = idx >> 6, kk = idx & 63;
        const int n = n0 + nn, k = k0 + kk;
        bv[i] = (n < g.n && k < kend) ? g.w[static_cast<int64_t>(n) * g.ldw + k] : 0.f;
      } else {
        const int kk = idx >> 6, nn = idx & 63;
        const int n = n0 + nn, k = k0 + kk;
        bv[i] = (n < g.n && k < kend) ? g.w[static_cast<int64_t>(k) * g.ldw + n] : 0.f;
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = i * 256 + threadIdx.x;
      As[idx & 63][idx >> 6] = av[i];
      if (TB) Bs[idx & 63][idx >> 6] = bv[i];
      else Bs[idx >> 6][idx & 63] = bv[i];
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < 64; ++kk) {
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      float a[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty + 16 * i];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[i][0] = fmaf(a[i], b.x, acc[i][0]);
        acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
        acc[i][2] = fmaf(a[i], b.z, acc[i][2]);
        acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
      }
    }
    __syncthreads();
  }
  // split-K partial sums meet in global memory: one vector reduction per row and thread where the layout allows it (four
  // adjacent columns = 16 bytes, two when only the row pitch is even) instead of four scalar atomics
  const int nq = n0 + tx * 4;
  float cs[4] = {1.f, 1.f, 1.f, 1.f}, bi[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (nq + j < g.n) {
      if (g.col_scale != nullptr) cs[j] = g.col_scale[nq + j];
      if (g.bias != nullptr && blockIdx.y == 0) bi[j] = g.bias[nq + j];
    }
  const bool base_ok = (reinterpret_cast<uintptr_t>(g.out) & 15) == 0;
  const bool vec4 = base_ok && (g.ldo & 3) == 0 && nq + 3 < g.n;
  const bool vec2 = base_ok && (g.ldo & 1) == 0 && nq + 3 < g.n;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty + 16 * i;
    if (m >= g.m) continue;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = fmaf(acc[i][j], cs[j], bi[j]);
    float* dst = g.out + static_cast<int64_t>(m) * g.ldo + nq;
    if (vec4) {
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
    } else if (vec2) {
      asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst), "f"(v[0]), "f"(v[1]) : "memory");
      asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst + 2), "f"(v[2]), "f"(v[3]) : "memory");
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (nq + j < g.n) atomicAdd(dst + j, v[j]);
    }
  }
}

template <bool TB>
cudaError_t launch_skinny_gemm(SkinnyArgs g, int sm_count, cudaStream_t s) {
  cudaError_t e = g.ldo == g.n ? cudaMemsetAsync(g.out, 0, sizeof(float) * g.n * g.m, s)
                              : cudaMemset2DAsync(g.out, sizeof(float) * g.ldo, 0, sizeof(float) * g.n, g.m, s);
  if (e != cudaSuccess) return e;
  const int nb = (g.n + 63) / 64, mb = (g.m + 63) / 64, kb = (g.k + 63) / 64;
  // one 64-wide K step per block when the row count is a single tile (every product of the training step): one memory round
  // trip per block and ~kb x nb blocks in flight; otherwise ~2 blocks per SM
  int splits = mb == 1 ? kb : (2 * sm_count + nb * mb - 1) / (nb * mb);
  if (splits > kb) splits = kb;
  if (splits < 1) splits = 1;
  g.ks = (kb + splits - 1) / splits * 64;
  splits = (g.k + g.ks - 1) / g.ks;
  skinny_gemm_kernel<TB><<<dim3(nb, splits, mb), 256, 0, s>>>(g);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ optimizer
struct ParamSeg {
  int64_t off, size;
  float wd;          // L2 regulariser scale (kernels only)
  int32_t moving;    // 1: moving statistic (copied from the tail of the gradient buffer, no gradient)
};
// The flat buffer cut into chunks of at most kUpdChunk elements that never straddle a variable: one block per chunk,
// no per-element search for the owning variable.
constexpr int kUpdChunk = 4096;
struct UpdChunk { int64_t off; int32_t n; int32_t seg; };

// Per-variable squared L2 norm of the final gradient g = gsum/world + wd*p (tf.clip_by_norm works per tensor,
// model/trainer.py:529-533 and :1245-1246): norms[seg] += sum g^2.
// gb (may be null): the summed gradients of the trainable range as they came back from a bf16 exchange
__device__ __forceinline__ float grad_sum(const float* g, const __nv_bfloat16* gb, int64_t i) {
  return gb != nullptr ? __bfloat162float(gb[i]) : g[i];
}
__global__ void __launch_bounds__(256) grad_sqnorm_kernel(const float* __restrict__ p, const float* __restrict__ g,
                                                          const __nv_bfloat16* __restrict__ gb,
                                                          const ParamSeg* __restrict__ segs, const UpdChunk* __restrict__ chunks,
                                                          float inv_world, double* __restrict__ norms) {
  __shared__ double sh[8];
  const UpdChunk ch = chunks[blockIdx.x];
  const ParamSeg sg = segs[ch.seg];
  double s = 0.0;
  if (!sg.moving)
    for (int i = threadIdx.x; i < ch.n; i += 256) {
      const double v = fmaf(sg.wd, p[ch.off + i], grad_sum(g, gb, ch.off + i) * inv_world);
      s += v * v;
    }
  s = block_reduce(s, sh, false);
  if (threadIdx.x == 0 && s != 0.0) atomicAdd(norms + ch.seg, s);
}

// g = gsum/world + wd*p, optionally clipped per variable to clip_norm (g * clip / max(|g|, clip)); stored back for
// inspection.  sgd: p -= lr*g; momentum: a = m*a + g; p -= lr*(nesterov ? g + m*a : a);
// adam (tf.train.AdamOptimizer defaults b1 = 0.9, b2 = 0.999, eps = 1e-8; lr already holds lr*sqrt(1-b2^t)/(1-b1^t)):
//   m = b1*m + (1-b1)*g; v = b2*v + (1-b2)*g^2; p -= lr*m/(sqrt(v)+eps), m = accum[0..total), v = accum[total..2 total)
// The regularisation loss sum_kernels (wd/2)*|W|^2 (tf.contrib.layers.l2_regularizer, model/trainer.py:472-499) of the
// PRE-update weights is accumulated on the way (reg_out += ...).
__global__ void __launch_bounds__(256) apply_update_kernel(float* __restrict__ p, float* __restrict__ g,
                                                           const __nv_bfloat16* __restrict__ gb, float* __restrict__ accum,
                                                           const ParamSeg* __restrict__ segs, const UpdChunk* __restrict__ chunks,
                                                           int64_t total, float inv_world, float lr, int optimizer, float momentum,
                                                           int nesterov, const double* __restrict__ sqnorms, float clip_norm,
                                                           double* __restrict__ reg_out) {
  __shared__ double sh[8];
  const UpdChunk ch = chunks[blockIdx.x];
  const ParamSeg sg = segs[ch.seg];
  if (sg.moving) {
    for (int i = threadIdx.x; i < ch.n; i += 256) p[ch.off + i] = g[ch.off + i];   // rank 0's new moving statistic (other ranks sent zeros)
    return;
  }
  float clip = 1.f;
  if (sqnorms != nullptr) clip = clip_norm / fmaxf(static_cast<float>(sqrt(sqnorms[ch.seg])), clip_norm);
  double reg = 0.0;
  for (int k = threadIdx.x; k < ch.n; k += 256) {
    const int64_t i = ch.off + k;
    const float pv = p[i];
    reg += static_cast<double>(pv) * pv;
    float grad = fmaf(sg.wd, pv, grad_sum(g, gb, i) * inv_world);
    if (sqnorms != nullptr) grad *= clip;
    g[i] = grad;
    if (optimizer == 0) {
      p[i] = pv - lr * grad;
    } else if (optimizer == 1) {
      const float a = fmaf(momentum, accum[i], grad);
      accum[i] = a;
      p[i] = pv - lr * (nesterov ? fmaf(momentum, a, grad) : a);
    } else {
      const float m = fmaf(0.9f, accum[i], 0.1f * grad);
      const float v = fmaf(0.999f, accum[total + i], 0.001f * grad * grad);
      accum[i] = m;
      accum[total + i] = v;
      p[i] = pv - lr * m / (sqrtf(v) + 1e-8f);
    }
  }
  if (sg.wd != 0.f) {                                 // uniform per block: every thread takes the reduction's barriers
    reg = block_reduce(reg, sh, false);
    if (threadIdx.x == 0 && reg != 0.0) atomicAdd(reg_out, 0.5 * sg.wd * reg);
  }
}

__global__ void fill_f32_kernel(float* __restrict__ dst, int64_t n, float v) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    dst[i] = v;
}

}  // namespace
}  // namespace xv
