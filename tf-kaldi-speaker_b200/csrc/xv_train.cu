// One data-parallel training step of the TDNN x-vector network (BASELINE.json config 5).
//
// Replaces sess.run(train_op) of Trainer.train / TrainerMGPU (model/trainer.py:590-673, 1133-1248):
//   forward in TRAINING mode (tf.layers.batch_normalization with batch statistics, model/tdnn.py:36-181),
//   statistics pooling (model/pooling.py:40-50), additive_margin_softmax / softmax (model/loss.py:9-48, 201-286),
//   total loss = loss + L2 regularisation of every kernel (model/trainer.py:472-499),
//   hand-derived backward pass, gradient averaging over ranks = ncclAllReduce(sum) * 1/N in place of the
//   reference's per-variable add_n/N (model/trainer.py:1225-1248), SGD / momentum update (:225-246),
//   BN moving averages taken from rank 0 (= tower 0, :1183-1190).
//
// Layout: the B utterances of T frames are packed back to back, R = B*T rows; layer l keeps the row index and
// its valid rows are (r mod T) < T - context_l.  Invalid rows of every activation / gradient are written as 0,
// so the spliced GEMMs need no masks.  All parameters live in ONE flat fp32 buffer and all gradients in a
// second flat buffer of the same layout (+ a tail carrying the moving statistics and the loss), so the
// exchange step is a single collective.
//
// This file is the fp32 path (SIMT FFMA GEMMs, fp64 reductions): the validation-grade implementation whose
// gradients match the fp64 oracle to <= 1e-4.
#include <dlfcn.h>
#include <math.h>
#include <cmath>

#include <algorithm>

#include "xv_ctx.h"

namespace xv {

namespace {

// ------------------------------------------------------------------------------------------------ GEMMs
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

}  // namespace
}  // namespace xv

namespace xv {
namespace {

// ------------------------------------------------------------------------------------------------ reductions
// Per-column sums over the VALID rows of a [rows, C] matrix ((r mod T) < valid_len), NV fp64 sums per column.
struct RowMask {
  int32_t t, valid_len;   // utterance length in rows and number of valid rows per utterance
  __device__ __forceinline__ bool ok(int64_t r) const {     // rows of a step fit 32 bits (checked by xv_train_step)
    return static_cast<int32_t>(static_cast<uint32_t>(r) % static_cast<uint32_t>(t)) < valid_len;
  }
};

struct StatsF {            // v0 = sum z, v1 = sum z^2
  const float* z; int32_t ld;
  static constexpr int NV = 2;
  __device__ __forceinline__ void operator()(int64_t r, int c, double* v) const {
    const double x = z[r * ld + c];
    v[0] += x; v[1] += x * x;
  }
};
struct SumF {              // v0 = sum x
  const float* x; int32_t ld;
  static constexpr int NV = 1;
  __device__ __forceinline__ void operator()(int64_t r, int c, double* v) const { v[0] += x[r * ld + c]; }
};
// Backward through activation + batch norm: g = dy * act'(pre); v0 = sum g, v1 = sum g*xhat, v2 = sum dy*pre [pre<0]
struct BnBwdF {
  const float* dy; const float* z; int32_t ld;
  const float* mean; const float* rstd; const float* gamma; const float* beta; const float* slope;   // mean==null: no BN
  static constexpr int NV = 3;
  __device__ __forceinline__ void operator()(int64_t r, int c, double* v) const {
    const float zz = z[r * ld + c];
    float xh = 0.f, pre = zz;
    if (mean != nullptr) { xh = (zz - mean[c]) * rstd[c]; pre = fmaf(gamma[c], xh, beta[c]); }
    const float d = dy[r * ld + c];
    const float g = pre > 0.f ? d : d * slope[c];
    v[0] += g; v[1] += static_cast<double>(g) * xh;
    if (pre < 0.f) v[2] += static_cast<double>(d) * pre;
  }
};

template <class F>
__global__ void __launch_bounds__(256) col_reduce_kernel(const F f, const RowMask mask, int64_t rows, int c_total,
                                                         int64_t rows_per_block, double* __restrict__ out) {
  __shared__ double red[F::NV][8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  const int64_t r0 = static_cast<int64_t>(blockIdx.y) * rows_per_block;
  const int64_t r1 = min(r0 + rows_per_block, rows);
  double v[F::NV];
#pragma unroll
  for (int k = 0; k < F::NV; ++k) v[k] = 0.0;
  if (c < c_total) {
#pragma unroll 4
    for (int64_t r = r0 + ry; r < r1; r += 8)
      if (mask.ok(r)) f(r, c, v);
  }
#pragma unroll
  for (int k = 0; k < F::NV; ++k) red[k][ry][cx] = v[k];
  __syncthreads();
  if (ry == 0 && c < c_total) {
#pragma unroll
    for (int k = 0; k < F::NV; ++k) {
      double s = 0.0;
      for (int j = 0; j < 8; ++j) s += red[k][j][cx];
      atomicAdd(out + static_cast<int64_t>(k) * c_total + c, s);
    }
  }
}

template <class F>
cudaError_t launch_col_reduce(const F& f, RowMask mask, int64_t rows, int c_total, double* out, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(double) * F::NV * c_total, s);
  if (e != cudaSuccess) return e;
  int64_t rpb = rows / 64;
  if (rpb < 64) rpb = 64;
  dim3 grid((c_total + 31) / 32, static_cast<unsigned>((rows + rpb - 1) / rpb));
  col_reduce_kernel<F><<<grid, 256, 0, s>>>(f, mask, rows, c_total, rpb, out);
  return cudaGetLastError();
}

// Four-channel variants of the two hot reductions (widths that are multiples of 4): one 16-byte load per tensor and row,
// per-channel constants in registers, fp64 accumulators.  KIND 0: v0 = sum z, v1 = sum z^2.  KIND 1: the BnBwdF sums.
template <int KIND>
__global__ void __launch_bounds__(256) col_reduce4_kernel(const float* __restrict__ dy, const float* __restrict__ z,
                                                          const RowMask mask, int64_t rows, int c_total, int64_t rows_per_block,
                                                          const float* __restrict__ mean, const float* __restrict__ rstd,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          const float* __restrict__ slope, double* __restrict__ out) {
  constexpr int NV = KIND == 0 ? 2 : 3;
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
    float mu[4], rs[4], ga[4], be[4], sl[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mu[j] = 0.f; rs[j] = 1.f; ga[j] = 1.f; be[j] = 0.f; sl[j] = 0.f;
      if (KIND == 1) {
        sl[j] = slope[c0 + j];
        if (mean != nullptr) { mu[j] = mean[c0 + j]; rs[j] = rstd[c0 + j]; ga[j] = gamma[c0 + j]; be[j] = beta[c0 + j]; }
      }
    }
#pragma unroll 2
    for (int64_t r = r0 + ry; r < r1; r += 8) {
      if (!mask.ok(r)) continue;
      const float4 z4 = *reinterpret_cast<const float4*>(z + r * c_total + c0);
      const float zz[4] = {z4.x, z4.y, z4.z, z4.w};
      if (KIND == 0) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const double x = zz[j];
          v[0][j] += x; v[1][j] += x * x;
        }
      } else {
        const float4 d4 = *reinterpret_cast<const float4*>(dy + r * c_total + c0);
        const float dd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float xh = 0.f, pre = zz[j];
          if (mean != nullptr) { xh = (zz[j] - mu[j]) * rs[j]; pre = fmaf(ga[j], xh, be[j]); }
          const float g = pre > 0.f ? dd[j] : dd[j] * sl[j];
          v[0][j] += g; v[1][j] += static_cast<double>(g) * xh;
          if (pre < 0.f) v[2][j] += static_cast<double>(dd[j]) * pre;
        }
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
    atomicAdd(out + static_cast<int64_t>(k) * c_total + c, sum);
  }
}

template <int KIND>
cudaError_t launch_col_reduce4(const float* dy, const float* z, RowMask mask, int64_t rows, int c_total, const float* mean,
                               const float* rstd, const float* gamma, const float* beta, const float* slope, double* out,
                               int sm_count, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(double) * (KIND == 0 ? 2 : 3) * c_total, s);
  if (e != cudaSuccess) return e;
  const int bx = (c_total + 127) / 128;
  int64_t chunks = (4 * sm_count + bx - 1) / bx;
  int64_t rpb = (rows + chunks - 1) / chunks;
  if (rpb < 64) rpb = 64;
  dim3 grid(bx, static_cast<unsigned>((rows + rpb - 1) / rpb));
  col_reduce4_kernel<KIND><<<grid, 256, 0, s>>>(dy, z, mask, rows, c_total, rpb, mean, rstd, gamma, beta, slope, out);
  return cudaGetLastError();
}

// mean / biased variance / rstd from the fp64 sums; moving-average targets for the flat-buffer tail.
__global__ void bn_finalize_kernel(const double* __restrict__ sums, int c_total, double n, float eps, int unbiased_moving,
                                   float* __restrict__ mean, float* __restrict__ rstd, const float* __restrict__ mov_mean,
                                   const float* __restrict__ mov_var, float momentum, float* __restrict__ new_mov_mean,
                                   float* __restrict__ new_mov_var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= c_total) return;
  const double m = sums[c] / n;
  double var = sums[c_total + c] / n - m * m;
  if (var < 0.0) var = 0.0;
  mean[c] = static_cast<float>(m);
  rstd[c] = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
  const double vm = unbiased_moving ? var * (n / fmax(n - 1.0, 1.0)) : var;
  new_mov_mean[c] = static_cast<float>(static_cast<double>(mov_mean[c]) * momentum + m * (1.0 - momentum));
  new_mov_var[c] = static_cast<float>(static_cast<double>(mov_var[c]) * momentum + vm * (1.0 - momentum));
}

// ------------------------------------------------------------------------------------------------ elementwise
// The three row-streaming kernels below give every thread FOUR consecutive channels (one 16-byte access per tensor) and
// keep the per-channel constants in registers while the thread walks down the rows; widths that are not a multiple
// of 4 take the scalar path (Q = 1).
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
template <> struct VecF<1> {
  static __device__ __forceinline__ void load(const float* p, float* v) { v[0] = *p; }
  static __device__ __forceinline__ void store(float* p, const float* v) { *p = v[0]; }
  static __device__ __forceinline__ void store_bf16(__nv_bfloat16* p, const float* v) { *p = __float2bfloat16(v[0]); }
};

// y = act(BN(z)) on valid rows, 0 elsewhere; optional fp32 output, optional bf16 operand copy (pad columns zeroed).
template <int Q>
__global__ void __launch_bounds__(256) bn_act_fwd_kernel(const float* __restrict__ z, int64_t rows, int c_total, RowMask mask,
                                                         const float* __restrict__ mean, const float* __restrict__ rstd,
                                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                                         const float* __restrict__ slope, float* __restrict__ y,
                                                         __nv_bfloat16* __restrict__ yb, int ldb) {
  const int nq = c_total / Q;
  const int64_t rstep = static_cast<int64_t>(gridDim.x) * blockDim.y;
  for (int q = threadIdx.x; q < nq; q += blockDim.x) {
    const int c0 = q * Q;
    float mu[Q], rs[Q], ga[Q], be[Q], sl[Q];
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      mu[k] = 0.f; rs[k] = 1.f; ga[k] = 1.f; be[k] = 0.f;
      if (mean != nullptr) { mu[k] = mean[c0 + k]; rs[k] = rstd[c0 + k]; ga[k] = gamma[c0 + k]; be[k] = beta[c0 + k]; }
      sl[k] = slope[c0 + k];
    }
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.y + threadIdx.y; r < rows; r += rstep) {
      float v[Q];
#pragma unroll
      for (int k = 0; k < Q; ++k) v[k] = 0.f;
      if (mask.ok(r)) {
        VecF<Q>::load(z + r * c_total + c0, v);
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          float t = v[k];
          if (mean != nullptr) t = fmaf(ga[k], (t - mu[k]) * rs[k], be[k]);      // same expression as BnBwdF / the apply kernel
          v[k] = t >= 0.f ? t : t * sl[k];
        }
      }
      if (y != nullptr) VecF<Q>::store(y + r * c_total + c0, v);
      if (yb != nullptr) {
        VecF<Q>::store_bf16(yb + r * ldb + c0, v);
        if (q == nq - 1)
          for (int k = c_total; k < ldb; ++k) yb[r * ldb + k] = __float2bfloat16(0.f);
      }
    }
  }
}

// dz = gamma*rstd*(g - sum_g/n - xhat*sum_gx/n) on valid rows, 0 elsewhere (in place over dy); optional bf16 copy.
template <int Q>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(float* __restrict__ dy, const float* __restrict__ z, int64_t rows,
                                                           int c_total, RowMask mask, const float* __restrict__ mean,
                                                           const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, const float* __restrict__ slope,
                                                           const double* __restrict__ sums, double n,
                                                           __nv_bfloat16* __restrict__ dzb, int ldb) {
  const int nq = c_total / Q;
  const int64_t rstep = static_cast<int64_t>(gridDim.x) * blockDim.y;
  for (int q = threadIdx.x; q < nq; q += blockDim.x) {
    const int c0 = q * Q;
    float mu[Q], rs[Q], ga[Q], be[Q], sl[Q];
    double k1[Q], k2[Q], gr[Q];
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      sl[k] = slope[c0 + k];
      mu[k] = 0.f; rs[k] = 1.f; ga[k] = 1.f; be[k] = 0.f; k1[k] = 0.0; k2[k] = 0.0; gr[k] = 1.0;
      if (mean != nullptr) {
        mu[k] = mean[c0 + k]; rs[k] = rstd[c0 + k]; ga[k] = gamma[c0 + k]; be[k] = beta[c0 + k];
        k1[k] = sums[c0 + k] / n; k2[k] = sums[c_total + c0 + k] / n;
        gr[k] = static_cast<double>(ga[k]) * rs[k];
      }
    }
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.y + threadIdx.y; r < rows; r += rstep) {
      float out[Q];
#pragma unroll
      for (int k = 0; k < Q; ++k) out[k] = 0.f;
      if (mask.ok(r)) {
        float zz[Q], d[Q];
        VecF<Q>::load(z + r * c_total + c0, zz);
        VecF<Q>::load(dy + r * c_total + c0, d);
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          if (mean != nullptr) {
            const float xh = (zz[k] - mu[k]) * rs[k];
            const float pre = fmaf(ga[k], xh, be[k]);
            const float g = pre > 0.f ? d[k] : d[k] * sl[k];
            const double t = static_cast<double>(g) - k1[k] - static_cast<double>(xh) * k2[k];
            out[k] = static_cast<float>(gr[k] * t);
          } else {
            out[k] = zz[k] > 0.f ? d[k] : d[k] * sl[k];
          }
        }
      }
      VecF<Q>::store(dy + r * c_total + c0, out);
      if (dzb != nullptr) {
        VecF<Q>::store_bf16(dzb + r * ldb + c0, out);
        if (q == nq - 1)
          for (int k = c_total; k < ldb; ++k) dzb[r * ldb + k] = __float2bfloat16(0.f);
      }
    }
  }
}

// Statistics-pooling backward (model/pooling.py:40-50): pooled = [mean | std]; dx = dmean/L + dstd*(x-mean)/(L*std).
template <int Q>
__global__ void __launch_bounds__(256) pool_bwd_kernel(const float* __restrict__ x, int64_t rows, int c_total, RowMask mask,
                                                       const float* __restrict__ pooled, const float* __restrict__ dpooled,
                                                       float* __restrict__ dx) {
  const int nq = c_total / Q;
  const float inv_l = 1.f / static_cast<float>(mask.valid_len);
  const int64_t rstep = static_cast<int64_t>(gridDim.x) * blockDim.y;
  for (int q = threadIdx.x; q < nq; q += blockDim.x) {
    const int c0 = q * Q;
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.y + threadIdx.y; r < rows; r += rstep) {
      float out[Q];
#pragma unroll
      for (int k = 0; k < Q; ++k) out[k] = 0.f;
      if (mask.ok(r)) {
        const int64_t u = r / mask.t;
        float xv[Q], mu[Q], sd[Q], dm[Q], ds[Q];
        VecF<Q>::load(x + r * c_total + c0, xv);
        VecF<Q>::load(pooled + u * 2 * c_total + c0, mu);
        VecF<Q>::load(pooled + u * 2 * c_total + c_total + c0, sd);
        VecF<Q>::load(dpooled + u * 2 * c_total + c0, dm);
        VecF<Q>::load(dpooled + u * 2 * c_total + c_total + c0, ds);
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          out[k] = dm[k] * inv_l;
          if (sd[k] * sd[k] > 1.0000001e-12f) out[k] += ds[k] * (xv[k] - mu[k]) * inv_l / sd[k];   // variance floor: no gradient
        }
      }
      VecF<Q>::store(dx + r * c_total + c0, out);
    }
  }
}

// Launch geometry of the row-streaming kernels: x covers the channel groups, y the rows.
struct RowGrid { dim3 grid, block; };
inline RowGrid row_grid(int groups, int sm_count) {
  int bx = 32;
  while (bx < groups && bx < 256) bx <<= 1;
  RowGrid g;
  g.block = dim3(bx, 256 / bx);
  g.grid = dim3(sm_count * 8);
  return g;
}

// ------------------------------------------------------------------------------------------------ bf16 operand copies
// Master weights stay fp32 in the flat buffer; every step re-casts them into the two bf16 layouts the GEMMs read.
// Forward operand: wt[n][k] (K-major), k = j*cstride + c  ->  W[j, c, n]; zero for n >= cout, c >= cin, j >= taps.
__global__ void cast_w_fwd_kernel(const float* __restrict__ w, int taps, int cin, int cout, int cstride, int ktotal,
                                  int n_alloc, __nv_bfloat16* __restrict__ wt) {
  __shared__ float tile[32][33];
  const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int k = k0 + i, n = n0 + threadIdx.x;
    const int j = k / cstride, c = k - j * cstride;
    float v = 0.f;
    if (k < ktotal && j < taps && c < cin && n < cout) v = w[(static_cast<int64_t>(j) * cin + c) * cout + n];
    tile[i][threadIdx.x] = v;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int n = n0 + i, k = k0 + threadIdx.x;
    if (n < n_alloc && k < ktotal) wt[static_cast<int64_t>(n) * ktotal + k] = __float2bfloat16(tile[threadIdx.x][i]);
  }
}
// Data-gradient operand: wd[c][j*ldz + n] = W[j, c, n]; zero for c >= cin, n >= cout.
__global__ void cast_w_dgrad_kernel(const float* __restrict__ w, int taps, int cin, int cout, int ldz, int c_alloc,
                                    __nv_bfloat16* __restrict__ wd) {
  const int64_t total = static_cast<int64_t>(c_alloc) * taps * ldz;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int n = static_cast<int>(i % ldz);
    const int64_t q = i / ldz;
    const int j = static_cast<int>(q % taps);
    const int c = static_cast<int>(q / taps);
    float v = 0.f;
    if (c < cin && n < cout) v = w[(static_cast<int64_t>(j) * cin + c) * cout + n];
    wd[i] = __float2bfloat16(v);
  }
}
__global__ void pad_copy_kernel(const float* __restrict__ src, int n, float* __restrict__ dst, int n_alloc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_alloc) dst[i] = i < n ? src[i] : 0.f;
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
__global__ void __launch_bounds__(256) softmax_loss_kernel(float* __restrict__ logits, int s_dim, const float* __restrict__ f,
                                                           int e_dim, const int32_t* __restrict__ labels, int batch,
                                                           const float* __restrict__ r, int amsoftmax, float margin, float fa,
                                                           float* __restrict__ gr, float* __restrict__ coef,
                                                           double* __restrict__ loss_sum) {
  __shared__ double sh[8];
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

// dW[e,s] = r_s * (dwn[e,s] - wn[e,s] * sum_e' dwn[e',s]*wn[e',s]),  wn = W * r_s   (in place over dwn)
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
  const bool floored = rs >= 0.999999e6;                         // sum w^2 <= 1e-12: l2_normalize's max() is flat
  for (int e = ry; e < e_dim; e += 8) {
    const int64_t i = static_cast<int64_t>(e) * s_dim + s;
    const double g = floored ? rs * dwn[i] : rs * (static_cast<double>(dwn[i]) - static_cast<double>(w[i]) * rs * dot);
    dwn[i] = static_cast<float>(g);
  }
}

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
#pragma unroll 4
    for (int i = 0; i < 16; ++i) {
      const int idx = i * 256 + threadIdx.x;
      const int mm = idx >> 6, kk = idx & 63;
      const int m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < g.m && k < kend) ? g.a[static_cast<int64_t>(m) * g.lda + k] : 0.f;
    }
#pragma unroll 4
    for (int i = 0; i < 16; ++i) {
      const int idx = i * 256 + threadIdx.x;
      if (TB) {
        const int nn = idx >> 6, kk = idx & 63;
        const int n = n0 + nn, k = k0 + kk;
        Bs[kk][nn] = (n < g.n && k < kend) ? g.w[static_cast<int64_t>(n) * g.ldw + k] : 0.f;
      } else {
        const int kk = idx >> 6, nn = idx & 63;
        const int n = n0 + nn, k = k0 + kk;
        Bs[kk][nn] = (n < g.n && k < kend) ? g.w[static_cast<int64_t>(k) * g.ldw + n] : 0.f;
      }
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
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty + 16 * i;
    if (m >= g.m) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= g.n) continue;
      float v = acc[i][j];
      if (g.col_scale != nullptr) v *= g.col_scale[n];
      if (g.bias != nullptr && blockIdx.y == 0) v += g.bias[n];
      atomicAdd(g.out + static_cast<int64_t>(m) * g.ldo + n, v);
    }
  }
}

template <bool TB>
cudaError_t launch_skinny_gemm(SkinnyArgs g, int sm_count, cudaStream_t s) {
  cudaError_t e = cudaMemset2DAsync(g.out, sizeof(float) * g.ldo, 0, sizeof(float) * g.n, g.m, s);
  if (e != cudaSuccess) return e;
  const int nb = (g.n + 63) / 64, mb = (g.m + 63) / 64, kb = (g.k + 63) / 64;
  int splits = (2 * sm_count + nb * mb - 1) / (nb * mb);
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

__global__ void reg_loss_kernel(const float* __restrict__ p, const ParamSeg* __restrict__ segs, int n_segs,
                                double* __restrict__ out) {
  __shared__ double sh[8];
  const ParamSeg sg = segs[blockIdx.y];
  double s = 0.0;
  if (sg.wd != 0.f)
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < sg.size;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
      const double v = p[sg.off + i];
      s += v * v;
    }
  s = block_reduce(s, sh, false);
  if (threadIdx.x == 0 && s != 0.0) atomicAdd(out, 0.5 * sg.wd * s);
}

// Per-variable squared L2 norm of the final gradient g = gsum/world + wd*p (tf.clip_by_norm works per tensor,
// model/trainer.py:529-533 and :1245-1246): norms[seg] += sum g^2.  grid (x, n_segs).
__global__ void grad_sqnorm_kernel(const float* __restrict__ p, const float* __restrict__ g, const ParamSeg* __restrict__ segs,
                                   float inv_world, double* __restrict__ norms) {
  __shared__ double sh[8];
  const ParamSeg sg = segs[blockIdx.y];
  double s = 0.0;
  if (!sg.moving)
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < sg.size;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
      const double v = fmaf(sg.wd, p[sg.off + i], g[sg.off + i] * inv_world);
      s += v * v;
    }
  s = block_reduce(s, sh, false);
  if (threadIdx.x == 0 && s != 0.0) atomicAdd(norms + blockIdx.y, s);
}

// g = gsum/world + wd*p, optionally clipped per variable to clip_norm (g * clip / max(|g|, clip)); stored back for
// inspection.  sgd: p -= lr*g; momentum: a = m*a + g; p -= lr*(nesterov ? g + m*a : a);
// adam (tf.train.AdamOptimizer defaults b1 = 0.9, b2 = 0.999, eps = 1e-8; lr already holds lr*sqrt(1-b2^t)/(1-b1^t)):
//   m = b1*m + (1-b1)*g; v = b2*v + (1-b2)*g^2; p -= lr*m/(sqrt(v)+eps), m = accum[0..total), v = accum[total..2 total)
__global__ void apply_update_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ accum,
                                    const ParamSeg* __restrict__ segs, int n_segs, int64_t total, float inv_world, float lr,
                                    int optimizer, float momentum, int nesterov, const double* __restrict__ sqnorms,
                                    float clip_norm) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int lo = 0, hi = n_segs - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (segs[mid].off <= i) lo = mid; else hi = mid - 1;
    }
    const ParamSeg sg = segs[lo];
    if (sg.moving) {
      p[i] = g[i];                       // rank 0's new moving statistic (other ranks contributed zeros)
      continue;
    }
    float grad = fmaf(sg.wd, p[i], g[i] * inv_world);
    if (sqnorms != nullptr) grad *= clip_norm / fmaxf(static_cast<float>(sqrt(sqnorms[lo])), clip_norm);
    g[i] = grad;
    if (optimizer == 0) {
      p[i] -= lr * grad;
    } else if (optimizer == 1) {
      const float a = fmaf(momentum, accum[i], grad);
      accum[i] = a;
      p[i] -= lr * (nesterov ? fmaf(momentum, a, grad) : a);
    } else {
      const float m = fmaf(0.9f, accum[i], 0.1f * grad);
      const float v = fmaf(0.999f, accum[total + i], 0.001f * grad * grad);
      accum[i] = m;
      accum[total + i] = v;
      p[i] -= lr * m / (sqrtf(v) + 1e-8f);
    }
  }
}

}  // namespace
}  // namespace xv

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
  DevBuf wt_fwd, wt_dgrad, bias_pad, yb;
  CUtensorMap tm_wf, tm_wd;
};

struct UidBlob { char bytes[128]; };     // ncclUniqueId is a 128-byte POD passed by value

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, UidBlob, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
struct TrainState {
  std::vector<TrLayer> layers;           // tdnn1..5 (frame), tdnn6, tdnn7 (segment)
  int n_frame = 5;
  int64_t n_train = 0, n_total = 0;      // trainable floats; + moving statistics; (+1 loss slot in the gradient buffer)
  std::map<std::string, std::pair<int64_t, int64_t>> index;   // TF variable name -> (offset, size)
  std::vector<ParamSeg> segs;
  DevBuf params, grads, accum, d_segs;
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
  DevBuf const_slope, zeros, feats, labels, d_cu, pooled, logits, gr, coef, rnorm, dbuf[2], sums, scalars, pool_ws, scratch;
  int cap_b = 0, cap_t = 0;
  bool mixed = false;                    // context precision bf16: frame-layer GEMMs on tcgen05, everything else fp32
  DevBuf xb0, dzb, ones_pad, zeros_pad;
  // data parallel
  NcclApi nccl;
  void* comm = nullptr;
  int rank = 0, world = 1;
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

__global__ void d2f_kernel(const double* __restrict__ src, float* __restrict__ dst, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = static_cast<float>(src[i]);
}
__global__ void fill_f32_kernel(float* __restrict__ dst, int64_t n, float v) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    dst[i] = v;
}
__global__ void loss_to_tail_kernel(const double* __restrict__ scalars, float* __restrict__ slot) { slot[0] = static_cast<float>(scalars[0]); }

}  // namespace

void train_release(xv_ctx* c) {
  TrainState* t = c->train;
  if (!t) return;
  if (t->comm && t->nccl.CommDestroy) t->nccl.CommDestroy(t->comm);
  for (TrLayer& l : t->layers) { release(&l.z); release(&l.y); release(&l.mean); release(&l.rstd); }
  DevBuf* bufs[] = {&t->params, &t->grads, &t->accum, &t->d_segs, &t->const_slope, &t->zeros, &t->feats, &t->labels, &t->d_cu,
                    &t->pooled, &t->logits, &t->gr, &t->coef, &t->rnorm, &t->dbuf[0], &t->dbuf[1], &t->sums, &t->scalars,
                    &t->pool_ws, &t->scratch, &t->sqnorms};
  for (DevBuf* b : bufs) release(b);
  delete t;
  c->train = nullptr;
}

static int train_init(xv_ctx* c, int32_t num_speakers) {
  if (c->resnet || c->scope != "tdnn" || c->pooling_type != "statistics_pooling")
    return fail(c, XV_EUNSUPPORTED, "xv_train_init: the training step covers network_type tdnn with statistics_pooling");
  if (c->feature_norm) return fail(c, XV_EUNSUPPORTED, "xv_train_init: feature_norm is not covered");
  train_release(c);
  TrainState* t = new TrainState();
  c->train = t;
  t->mixed = c->bf16;
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
  int maxc = std::max(num_speakers, 1);
  for (const TrLayer& l : t->layers) maxc = std::max(maxc, l.cout);
  if ((rc = ensure(c, &t->const_slope, sizeof(float) * 2 * maxc)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->zeros, sizeof(float) * maxc)) != XV_OK) return rc;
  const float slope = t->relu_type == "lrelu" ? kLreluAlpha : 0.f;
  fill_f32_kernel<<<64, 256, 0, c->stream>>>(static_cast<float*>(t->const_slope.p), maxc, slope);
  fill_f32_kernel<<<64, 256, 0, c->stream>>>(static_cast<float*>(t->const_slope.p) + maxc, maxc, 1.f);   // linear layers
  cudaMemsetAsync(t->zeros.p, 0, sizeof(float) * maxc, c->stream);
  if ((rc = ensure(c, &t->sums, sizeof(double) * 3 * maxc)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->scalars, sizeof(double) * 4)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->rnorm, sizeof(float) * num_speakers)) != XV_OK) return rc;
  for (TrLayer& l : t->layers) {
    if ((rc = ensure(c, &l.mean, sizeof(float) * l.cout)) != XV_OK) return rc;
    if ((rc = ensure(c, &l.rstd, sizeof(float) * l.cout)) != XV_OK) return rc;
  }
  if (t->mixed) {
    const char* we = wgrad_init();
    if (we) return fail(c, XV_ECUDA, "%s", we);
    int max_alloc = 256;
    int prev_ldb = 0;
    for (int i = 0; i < t->n_frame; ++i) {
      TrLayer& l = t->layers[i];
      if (l.cout % 4 != 0) return fail(c, XV_EUNSUPPORTED, "bf16 training needs layer widths that are multiples of 4 (%s has %d)", l.name.c_str(), l.cout);
      if (i == 0) { l.a_taps = 1; l.a_cin = (l.taps * l.cin + kBK - 1) / kBK * kBK; }
      else { l.a_taps = l.taps; l.a_cin = prev_ldb; }
      l.ktotal = l.a_taps * l.a_cin;
      l.n_alloc = (l.cout + kBN - 1) / kBN * kBN;
      l.c_alloc = (l.a_cin + kBN - 1) / kBN * kBN;
      l.ldb = (l.cout + kBK - 1) / kBK * kBK;
      prev_ldb = l.ldb;
      max_alloc = std::max(max_alloc, std::max(l.n_alloc, l.c_alloc));
      if ((rc = ensure(c, &l.wt_fwd, sizeof(__nv_bfloat16) * static_cast<size_t>(l.n_alloc) * l.ktotal)) != XV_OK) return rc;
      if ((rc = ensure(c, &l.bias_pad, sizeof(float) * l.n_alloc)) != XV_OK) return rc;
      const char* e = make_tmap_bf16_2d(&l.tm_wf, l.wt_fwd.p, l.ktotal, l.n_alloc, static_cast<uint64_t>(l.ktotal) * 2, kBK, kBLoadRows);
      if (!e && i > 0) {
        const size_t kd = static_cast<size_t>(l.taps) * l.ldb;
        if ((rc = ensure(c, &l.wt_dgrad, sizeof(__nv_bfloat16) * l.c_alloc * kd)) != XV_OK) return rc;
        e = make_tmap_bf16_2d(&l.tm_wd, l.wt_dgrad.p, kd, l.c_alloc, kd * 2, kBK, kBLoadRows);
      }
      if (e) return fail(c, XV_ECUDA, "%s", e);
    }
    if ((rc = ensure(c, &t->ones_pad, sizeof(float) * max_alloc)) != XV_OK) return rc;
    if ((rc = ensure(c, &t->zeros_pad, sizeof(float) * max_alloc)) != XV_OK) return rc;
    fill_f32_kernel<<<64, 256, 0, c->stream>>>(static_cast<float*>(t->ones_pad.p), max_alloc, 1.f);
    cudaMemsetAsync(t->zeros_pad.p, 0, sizeof(float) * max_alloc, c->stream);
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
    const bool bf16_only = t->mixed && static_cast<int>(i) < t->n_frame - 1;   // y feeds tensor-core GEMMs only
    if ((rc = ensure(c, &l.z, sizeof(float) * r * l.cout)) != XV_OK) return rc;
    if (!bf16_only && (rc = ensure(c, &l.y, sizeof(float) * r * l.cout)) != XV_OK) return rc;
    if (bf16_only && (rc = ensure(c, &l.yb, sizeof(__nv_bfloat16) * r * l.ldb)) != XV_OK) return rc;
    maxc = std::max(maxc, l.frame ? l.cout : 0);
    if (l.frame) max_ldb = std::max(max_ldb, l.ldb);
  }
  if (t->mixed) {
    if ((rc = ensure(c, &t->xb0, sizeof(__nv_bfloat16) * rows * t->layers[0].a_cin)) != XV_OK) return rc;
    if ((rc = ensure(c, &t->dzb, sizeof(__nv_bfloat16) * rows * max_ldb)) != XV_OK) return rc;
  }
  const int pool_c = t->layers[t->n_frame - 1].cout;
  if ((rc = ensure(c, &t->feats, sizeof(float) * rows * c->dim)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->labels, sizeof(int32_t) * b)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->d_cu, sizeof(int32_t) * (b + 1))) != XV_OK) return rc;
  if ((rc = ensure(c, &t->pooled, sizeof(float) * b * 2 * pool_c)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->pool_ws, pool_workspace_bytes(b, 1, pool_c))) != XV_OK) return rc;
  if ((rc = ensure(c, &t->logits, sizeof(float) * static_cast<size_t>(b) * t->speakers)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->gr, sizeof(float) * static_cast<size_t>(b) * t->speakers)) != XV_OK) return rc;
  if ((rc = ensure(c, &t->coef, sizeof(float) * b)) != XV_OK) return rc;
  const size_t dbytes = sizeof(float) * std::max<int64_t>(rows * maxc, static_cast<int64_t>(b) * std::max(2 * pool_c, t->embed));
  for (int i = 0; i < 2; ++i)
    if ((rc = ensure(c, &t->dbuf[i], dbytes)) != XV_OK) return rc;
  std::vector<int32_t> cu(b + 1);
  for (int u = 0; u <= b; ++u) cu[u] = u * tt;
  TR_CUDA(cudaMemcpyAsync(t->d_cu.p, cu.data(), sizeof(int32_t) * (b + 1), cudaMemcpyHostToDevice, c->stream), "H2D cu");
  TR_CUDA(cudaStreamSynchronize(c->stream), "sync");
  t->cap_b = b;
  t->cap_t = tt;
  return XV_OK;
}

// One frame-layer GEMM of the mixed-precision step on the tcgen05 kernel of the embedding path (xv_gemm_sm100.cu):
// out[m, n] = sum_{j, k} a[m + j*dil, k] * wt[n][j*a_cin + k] + shift[n], fp32 stores masked to n < cols.
int tr_gemm_tc(xv_ctx* c, TrainState* t, const CUtensorMap& tm_b, const void* a, int64_t rows, int a_cin, int taps, int dil,
               int n_alloc, const float* shift, float* out, int ldo, int cols, const char* fam) {
  GemmPlan g;
  const char* err = make_tmap_bf16_2d(&g.tm_a, a, a_cin, rows, static_cast<uint64_t>(a_cin) * 2, kBK, kBM);
  if (err) return fail(c, XV_ECUDA, "%s", err);
  g.tm_b = tm_b;
  g.tm_out = g.tm_a;     // never dereferenced (no bf16 output)
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
  g.pool = false;
  count_launch(c, fam);
  cudaError_t e = launch_gemm(g, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "tcgen05 training GEMM launch");
  return XV_OK;
}

}  // namespace

static int train_step(xv_ctx* c, const float* feats, const int32_t* labels, int32_t b, int32_t tt, float lr,
                      int64_t global_step, float* loss_out) {
  TrainState* t = c->train;
  if (!t) return fail(c, XV_ESTATE, "xv_train_step: call xv_train_init first");
  if (b < 1 || tt <= c->total_context) return fail(c, XV_EINVAL, "xv_train_step: need B >= 1 and T > %d", c->total_context);
  if (static_cast<int64_t>(b) * tt >= (1LL << 31)) return fail(c, XV_EINVAL, "xv_train_step: B*T must stay below 2^31 rows");
  int rc = tr_buffers(c, t, b, tt);
  if (rc != XV_OK) return rc;
  cudaStream_t s = c->stream;
  const int64_t rows = static_cast<int64_t>(b) * tt;
  float* P = static_cast<float*>(t->params.p);
  float* G = static_cast<float*>(t->grads.p);
  double* sums = static_cast<double*>(t->sums.p);
  double* scalars = static_cast<double*>(t->scalars.p);
  int maxc = t->speakers;
  for (const TrLayer& l : t->layers) maxc = std::max(maxc, l.cout);
  const float* slope_const = static_cast<const float*>(t->const_slope.p);
  const float* slope_linear = slope_const + maxc;
  for (int u = 0; u < b; ++u)
    if (labels[u] < 0 || labels[u] >= t->speakers) return fail(c, XV_EINVAL, "label %d of utterance %d is outside [0,%d)", labels[u], u, t->speakers);
  TR_CUDA(cudaMemcpyAsync(t->feats.p, feats, sizeof(float) * rows * c->dim, cudaMemcpyDefault, s), "H2D features");
  TR_CUDA(cudaMemcpyAsync(t->labels.p, labels, sizeof(int32_t) * b, cudaMemcpyHostToDevice, s), "H2D labels");
  TR_CUDA(cudaMemsetAsync(scalars, 0, sizeof(double) * 4, s), "memset");
  // the tail of the gradient buffer: moving statistics (rank 0 only) + loss slot
  TR_CUDA(cudaMemsetAsync(G + t->n_train, 0, sizeof(float) * (t->n_total - t->n_train + 1), s), "memset");
  auto slope_of = [&](const TrLayer& l) -> const float* {
    if (!l.act) return slope_linear;
    return l.off_alpha >= 0 ? P + l.off_alpha : slope_const;
  };

  // ------------------------------------------------------------------ forward
  const float* in = static_cast<const float*>(t->feats.p);
  int in_c = c->dim;
  for (size_t i = 0; i < t->layers.size(); ++i) {
    TrLayer& l = t->layers[i];
    const int64_t r = l.frame ? rows : b;
    const RowMask mask = l.frame ? RowMask{tt, tt - l.context} : RowMask{1, 1};
    const double n = l.frame ? static_cast<double>(b) * (tt - l.context) : static_cast<double>(b);
    if (static_cast<int>(i) == t->n_frame) {
      // statistics pooling of the last frame layer (model/pooling.py:40-50)
      const TrLayer& pl = t->layers[i - 1];
      PoolArgs pa;
      pa.x = pl.y.p; pa.dtype = XV_DTYPE_F32; pa.c = pl.cout; pa.ldx = pl.cout; pa.rows = rows;
      pa.cu = static_cast<const int32_t*>(t->d_cu.p); pa.n = b; pa.context = pl.context;
      pa.weights = nullptr; pa.heads = 1; pa.split_value = 1; pa.acc = static_cast<float*>(t->pool_ws.p);
      pa.out = static_cast<float*>(t->pooled.p); pa.ldout = 2 * pl.cout; pa.out_bf16 = nullptr; pa.ldout_bf16 = 0;
      count_launch(c, "train.pool");
      TR_CUDA(launch_pool(pa, c->sm_count, s), "pooling launch");
      in = static_cast<const float*>(t->pooled.p);
      in_c = 2 * pl.cout;
    }
    if (in_c != l.cin) return fail(c, XV_ESTATE, "training plan: layer %s expects %d inputs, got %d", l.name.c_str(), l.cin, in_c);
    const bool tc = t->mixed && l.frame;
    if (tc) {
      const void* a_bf16;
      if (i == 0) {
        count_launch(c, "train.splice_cast");
        TR_CUDA(launch_splice_cast(static_cast<const float*>(t->feats.p), rows, c->dim, l.taps, 1,
                                   static_cast<__nv_bfloat16*>(t->xb0.p), l.a_cin, s), "splice launch");
        a_bf16 = t->xb0.p;
      } else {
        a_bf16 = t->layers[i - 1].yb.p;
      }
      count_launch(c, "train.cast_w");
      cast_w_fwd_kernel<<<dim3(l.ktotal / 32, l.n_alloc / 32), dim3(32, 8), 0, s>>>(
          P + l.off_kernel, l.taps, l.cin, l.cout, i == 0 ? l.cin : l.a_cin, l.ktotal, l.n_alloc,
          static_cast<__nv_bfloat16*>(l.wt_fwd.p));
      pad_copy_kernel<<<(l.n_alloc + 255) / 256, 256, 0, s>>>(P + l.off_bias, l.cout, static_cast<float*>(l.bias_pad.p), l.n_alloc);
      TR_CUDA(cudaGetLastError(), "weight cast launch");
      if ((rc = tr_gemm_tc(c, t, l.tm_wf, a_bf16, rows, l.a_cin, l.a_taps, 1, l.n_alloc, static_cast<const float*>(l.bias_pad.p),
                           static_cast<float*>(l.z.p), l.cout, l.cout, "train.tc_fwd")) != XV_OK) return rc;
    } else if (!l.frame) {
      SkinnyArgs sa;
      sa.a = in; sa.lda = l.cin; sa.w = P + l.off_kernel; sa.ldw = l.cout; sa.out = static_cast<float*>(l.z.p); sa.ldo = l.cout;
      sa.col_scale = nullptr; sa.bias = P + l.off_bias; sa.m = b; sa.n = l.cout; sa.k = l.cin; sa.ks = 0;
      count_launch(c, "train.skinny_fwd");
      TR_CUDA(launch_skinny_gemm<false>(sa, c->sm_count, s), "segment-level forward GEMM launch");
    } else {
      SimtGemmArgs ga;
      ga.a = in; ga.a_rows = r; ga.cin = l.cin; ga.taps = l.taps; ga.dilation = 1;
      ga.w = P + l.off_kernel; ga.n = l.cout; ga.bias = P + l.off_bias;
      ga.bn_scale = nullptr; ga.bn_shift = nullptr; ga.slope = nullptr; ga.act_tanh = 0;
      ga.out_dense = static_cast<float*>(l.z.p); ga.out_bn = nullptr; ga.out_act = nullptr; ga.m_rows = r;
      count_launch(c, "train.gemm_fwd");
      TR_CUDA(launch_simt_gemm(ga, s), "forward GEMM launch");
    }
    const float *mean = nullptr, *rstd = nullptr, *gamma = nullptr, *beta = nullptr;
    if (l.has_bn) {
      count_launch(c, "train.bn_stats");
      if (l.cout % 4 == 0)
        TR_CUDA(launch_col_reduce4<0>(nullptr, static_cast<const float*>(l.z.p), mask, r, l.cout, nullptr, nullptr, nullptr, nullptr,
                                      nullptr, sums, c->sm_count, s), "BN statistics launch");
      else
        TR_CUDA(launch_col_reduce(StatsF{static_cast<const float*>(l.z.p), l.cout}, mask, r, l.cout, sums, s), "BN statistics launch");
      // new moving statistics go to the tail of the gradient buffer on rank 0 (tower 0), zeros elsewhere
      float* nm = t->rank == 0 ? G + l.off_mm : static_cast<float*>(t->dbuf[0].p);
      float* nv = t->rank == 0 ? G + l.off_mv : static_cast<float*>(t->dbuf[0].p) + l.cout;
      bn_finalize_kernel<<<(l.cout + 127) / 128, 128, 0, s>>>(sums, l.cout, n, kBnEpsilon, l.fused_bn ? 1 : 0,
                                                             static_cast<float*>(l.mean.p), static_cast<float*>(l.rstd.p),
                                                             P + l.off_mm, P + l.off_mv, t->bn_momentum, nm, nv);
      TR_CUDA(cudaGetLastError(), "BN finalize launch");
      mean = static_cast<const float*>(l.mean.p); rstd = static_cast<const float*>(l.rstd.p);
      gamma = P + l.off_gamma; beta = P + l.off_beta;
    }
    count_launch(c, "train.bn_act");
    const bool bf16_only = tc && static_cast<int>(i) < t->n_frame - 1;
    {
      float* y32 = bf16_only ? nullptr : static_cast<float*>(l.y.p);
      __nv_bfloat16* y16 = bf16_only ? static_cast<__nv_bfloat16*>(l.yb.p) : nullptr;
      if (l.cout % 4 == 0) {
        const RowGrid rg = row_grid(l.cout / 4, c->sm_count);
        bn_act_fwd_kernel<4><<<rg.grid, rg.block, 0, s>>>(static_cast<const float*>(l.z.p), r, l.cout, mask, mean, rstd, gamma, beta,
                                                          slope_of(l), y32, y16, l.ldb);
      } else {
        const RowGrid rg = row_grid(l.cout, c->sm_count);
        bn_act_fwd_kernel<1><<<rg.grid, rg.block, 0, s>>>(static_cast<const float*>(l.z.p), r, l.cout, mask, mean, rstd, gamma, beta,
                                                          slope_of(l), y32, y16, l.ldb);
      }
    }
    TR_CUDA(cudaGetLastError(), "BN/activation launch");
    in = static_cast<const float*>(l.y.p);
    in_c = l.cout;
  }

  // ------------------------------------------------------------------ loss (model/loss.py:9-48, 201-286)
  const TrLayer& last = t->layers.back();
  const float* fvec = static_cast<const float*>(last.y.p);
  const int E = t->embed, S = t->speakers;
  float fa = 0.f;
  if (t->amsoftmax) {
    const double lam = std::max(static_cast<double>(t->lam_min),
                                static_cast<double>(t->lam_base) * pow(1.0 + static_cast<double>(t->lam_gamma) * static_cast<double>(global_step),
                                                                        -static_cast<double>(t->lam_power)));
    fa = static_cast<float>(1.0 / (1.0 + lam));
    col_rnorm_kernel<<<(S + 31) / 32, 256, 0, s>>>(P + t->off_w, E, S, static_cast<float*>(t->rnorm.p));
    TR_CUDA(cudaGetLastError(), "column norm launch");
  }
  {
    // logits = f . W (softmax: + bias; AM-softmax: columns scaled by 1/|w_s|, model/loss.py:241-243)
    SkinnyArgs sa;
    sa.a = fvec; sa.lda = E; sa.w = P + t->off_w; sa.ldw = S; sa.out = static_cast<float*>(t->logits.p); sa.ldo = S;
    sa.col_scale = t->amsoftmax ? static_cast<const float*>(t->rnorm.p) : nullptr;
    sa.bias = t->amsoftmax ? nullptr : P + t->off_b;
    sa.m = b; sa.n = S; sa.k = E; sa.ks = 0;
    count_launch(c, "train.skinny_fwd");
    TR_CUDA(launch_skinny_gemm<false>(sa, c->sm_count, s), "logits GEMM launch");
  }
  count_launch(c, "train.loss");
  softmax_loss_kernel<<<b, 256, 0, s>>>(static_cast<float*>(t->logits.p), S, fvec, E, static_cast<const int32_t*>(t->labels.p), b,
                                        t->amsoftmax ? static_cast<const float*>(t->rnorm.p) : nullptr, t->amsoftmax ? 1 : 0,
                                        t->margin, fa, static_cast<float*>(t->gr.p), static_cast<float*>(t->coef.p), scalars);
  TR_CUDA(cudaGetLastError(), "loss launch");
  // dW = f^T G  (+ the l2_normalize Jacobian for AM-softmax), db = column sums of G
  {
    TrGemmArgs wa;
    wa.a = fvec; wa.lda = E; wa.dz = static_cast<const float*>(t->logits.p); wa.w = nullptr; wa.out = G + t->off_w;
    wa.rows = b; wa.cin = E; wa.nz = S; wa.taps = 1;
    count_launch(c, "train.gemm_wgrad");
    TR_CUDA(launch_tr_gemm<kWgrad>(wa, s), "softmax wgrad launch");
    if (t->amsoftmax) {
      amsoftmax_dw_kernel<<<(S + 31) / 32, 256, 0, s>>>(G + t->off_w, P + t->off_w, static_cast<const float*>(t->rnorm.p), E, S);
      TR_CUDA(cudaGetLastError(), "AM-softmax dW launch");
    } else {
      TR_CUDA(launch_col_reduce(SumF{static_cast<const float*>(t->logits.p), S}, RowMask{1, 1}, b, S, sums, s), "bias grad launch");
      d2f_kernel<<<(S + 255) / 256, 256, 0, s>>>(sums, G + t->off_b, S);
      TR_CUDA(cudaGetLastError(), "d2f launch");
    }
  }
  int cur = 0;
  float* dy = static_cast<float*>(t->dbuf[cur].p);
  {
    SkinnyArgs sa;
    sa.a = static_cast<const float*>(t->gr.p); sa.lda = S; sa.w = P + t->off_w; sa.ldw = S; sa.out = dy; sa.ldo = E;
    sa.col_scale = nullptr; sa.bias = nullptr; sa.m = b; sa.n = E; sa.k = S; sa.ks = 0;
    count_launch(c, "train.skinny_dgrad");
    TR_CUDA(launch_skinny_gemm<true>(sa, c->sm_count, s), "softmax dgrad launch");
    if (t->amsoftmax) {
      axpy_rows_kernel<<<ew_grid(c), 256, 0, s>>>(dy, fvec, static_cast<const float*>(t->coef.p), b, E);
      TR_CUDA(cudaGetLastError(), "axpy launch");
    }
  }

  // ------------------------------------------------------------------ backward
  for (int i = static_cast<int>(t->layers.size()) - 1; i >= 0; --i) {
    TrLayer& l = t->layers[i];
    const int64_t r = l.frame ? rows : b;
    const RowMask mask = l.frame ? RowMask{tt, tt - l.context} : RowMask{1, 1};
    const double n = l.frame ? static_cast<double>(b) * (tt - l.context) : static_cast<double>(b);
    if (i == t->n_frame - 1) {
      // dy currently holds d(pooled) [B, 2C]: statistics-pooling backward into [rows, C]
      float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
      count_launch(c, "train.pool_bwd");
      if (l.cout % 4 == 0) {
        const RowGrid rg = row_grid(l.cout / 4, c->sm_count);
        pool_bwd_kernel<4><<<rg.grid, rg.block, 0, s>>>(static_cast<const float*>(l.y.p), rows, l.cout, mask,
                                                        static_cast<const float*>(t->pooled.p), dy, dx);
      } else {
        const RowGrid rg = row_grid(l.cout, c->sm_count);
        pool_bwd_kernel<1><<<rg.grid, rg.block, 0, s>>>(static_cast<const float*>(l.y.p), rows, l.cout, mask,
                                                        static_cast<const float*>(t->pooled.p), dy, dx);
      }
      TR_CUDA(cudaGetLastError(), "pooling backward launch");
      cur ^= 1;
      dy = dx;
    }
    const bool tc = t->mixed && l.frame;
    const float *mean = nullptr, *rstd = nullptr, *gamma = nullptr, *beta = nullptr;
    if (l.has_bn) {
      mean = static_cast<const float*>(l.mean.p); rstd = static_cast<const float*>(l.rstd.p);
      gamma = P + l.off_gamma; beta = P + l.off_beta;
    }
    count_launch(c, "train.bn_bwd");
    if (l.cout % 4 == 0)
      TR_CUDA(launch_col_reduce4<1>(dy, static_cast<const float*>(l.z.p), mask, r, l.cout, mean, rstd, gamma, beta, slope_of(l), sums,
                                    c->sm_count, s), "BN backward reduce launch");
    else
      TR_CUDA(launch_col_reduce(BnBwdF{dy, static_cast<const float*>(l.z.p), l.cout, mean, rstd, gamma, beta, slope_of(l)}, mask, r,
                              l.cout, sums, s), "BN backward reduce launch");
    if (l.has_bn) {
      d2f_kernel<<<(l.cout + 255) / 256, 256, 0, s>>>(sums, G + l.off_beta, l.cout);
      d2f_kernel<<<(l.cout + 255) / 256, 256, 0, s>>>(sums + l.cout, G + l.off_gamma, l.cout);
    }
    if (l.off_alpha >= 0) d2f_kernel<<<(l.cout + 255) / 256, 256, 0, s>>>(sums + 2 * l.cout, G + l.off_alpha, l.cout);
    TR_CUDA(cudaGetLastError(), "d2f launch");
    {
      __nv_bfloat16* dz16 = tc ? static_cast<__nv_bfloat16*>(t->dzb.p) : nullptr;
      if (l.cout % 4 == 0) {
        const RowGrid rg = row_grid(l.cout / 4, c->sm_count);
        bn_bwd_apply_kernel<4><<<rg.grid, rg.block, 0, s>>>(dy, static_cast<const float*>(l.z.p), r, l.cout, mask, mean, rstd, gamma,
                                                            beta, slope_of(l), sums, n, dz16, l.ldb);
      } else {
        const RowGrid rg = row_grid(l.cout, c->sm_count);
        bn_bwd_apply_kernel<1><<<rg.grid, rg.block, 0, s>>>(dy, static_cast<const float*>(l.z.p), r, l.cout, mask, mean, rstd, gamma,
                                                            beta, slope_of(l), sums, n, dz16, l.ldb);
      }
    }
    TR_CUDA(cudaGetLastError(), "BN backward apply launch");
    // now dy == dz of this layer
    if (l.has_bn) {
      // a bias in front of a batch norm: sum_rows dz = -gamma*rstd*(sum xhat)*(sum g*xhat)/n with sum xhat == 0, i.e. the
      // gradient is identically zero (TF evaluates the same sum and gets rounding residue of the order 1e-9)
      TR_CUDA(cudaMemsetAsync(G + l.off_bias, 0, sizeof(float) * l.cout, s), "memset bias grad");
    } else {
      double* bsum = sums;     // reuse after the apply kernel has consumed the BN sums (same stream: ordered)
      TR_CUDA(launch_col_reduce(SumF{dy, l.cout}, mask, r, l.cout, bsum, s), "bias grad launch");
      d2f_kernel<<<(l.cout + 255) / 256, 256, 0, s>>>(bsum, G + l.off_bias, l.cout);
      TR_CUDA(cudaGetLastError(), "d2f launch");
    }
    const float* a_in;
    int lda;
    if (i == 0) { a_in = static_cast<const float*>(t->feats.p); lda = c->dim; }
    else if (i == t->n_frame) { a_in = static_cast<const float*>(t->pooled.p); lda = l.cin; }
    else { a_in = static_cast<const float*>(t->layers[i - 1].y.p); lda = l.cin; }
    if (tc) {
      // dW on the MN-major tcgen05 kernel (xv_wgrad_sm100.cu): layer 0 contracts the pre-spliced features (one tap of width
      // taps*dim), the others shift the previous layer's bf16 activations by the tap index
      const void* a_bf16 = i == 0 ? t->xb0.p : t->layers[i - 1].yb.p;
      TR_CUDA(cudaMemsetAsync(G + l.off_kernel, 0, sizeof(float) * static_cast<size_t>(l.taps) * l.cin * l.cout, s), "memset dW");
      cudaError_t ce;
      count_launch(c, "train.tc_wgrad");
      const char* we = launch_wgrad(a_bf16, rows, l.a_cin, t->dzb.p, l.ldb, rows, i == 0 ? 1 : l.taps, i == 0 ? l.taps * l.cin : l.cin,
                                    l.cout, G + l.off_kernel, l.cout, c->sm_count, s, &ce);
      if (we) return fail(c, XV_ECUDA, "%s", we);
      TR_CUDA(ce, "tcgen05 wgrad launch");
      if (i > 0) {
        float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
        count_launch(c, "train.cast_w");
        cast_w_dgrad_kernel<<<ew_grid(c), 256, 0, s>>>(P + l.off_kernel, l.taps, l.cin, l.cout, l.ldb, l.c_alloc,
                                                       static_cast<__nv_bfloat16*>(l.wt_dgrad.p));
        TR_CUDA(cudaGetLastError(), "dgrad weight cast launch");
        // dX[m, c] = sum_j sum_n dZ[m - j, n] W[j, c, n]: the forward kernel with tap step -1 (rows < 0 are TMA zero fill)
        if ((rc = tr_gemm_tc(c, t, l.tm_wd, t->dzb.p, rows, l.ldb, l.taps, -1, l.c_alloc, static_cast<const float*>(t->zeros_pad.p),
                             dx, l.cin, l.cin, "train.tc_dgrad")) != XV_OK) return rc;
        cur ^= 1;
        dy = dx;
      }
    } else {
      TrGemmArgs wa;
      wa.a = a_in; wa.lda = lda; wa.dz = dy; wa.w = nullptr; wa.out = G + l.off_kernel; wa.rows = r; wa.cin = l.cin; wa.nz = l.cout;
      wa.taps = l.taps;
      count_launch(c, "train.gemm_wgrad");
      TR_CUDA(launch_tr_gemm<kWgrad>(wa, s), "wgrad launch");
      if (i > 0 && !l.frame) {
        float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
        SkinnyArgs sa;
        sa.a = dy; sa.lda = l.cout; sa.w = P + l.off_kernel; sa.ldw = l.cout; sa.out = dx; sa.ldo = l.cin;
        sa.col_scale = nullptr; sa.bias = nullptr; sa.m = b; sa.n = l.cin; sa.k = l.cout; sa.ks = 0;
        count_launch(c, "train.skinny_dgrad");
        TR_CUDA(launch_skinny_gemm<true>(sa, c->sm_count, s), "segment-level dgrad launch");
        cur ^= 1;
        dy = dx;
      } else if (i > 0) {
        float* dx = static_cast<float*>(t->dbuf[cur ^ 1].p);
        TrGemmArgs da;
        da.a = nullptr; da.lda = 0; da.dz = dy; da.w = P + l.off_kernel; da.out = dx; da.rows = r; da.cin = l.cin; da.nz = l.cout;
        da.taps = l.taps;
        count_launch(c, "train.gemm_dgrad");
        TR_CUDA(launch_tr_gemm<kDgrad>(da, s), "dgrad launch");
        cur ^= 1;
        dy = dx;
      }
    }
  }

  // ------------------------------------------------------------------ regularisation, exchange, update
  reg_loss_kernel<<<dim3(128, static_cast<unsigned>(t->segs.size())), 256, 0, s>>>(P, static_cast<const ParamSeg*>(t->d_segs.p),
                                                                                 static_cast<int>(t->segs.size()), scalars + 1);
  TR_CUDA(cudaGetLastError(), "regularisation launch");
  loss_to_tail_kernel<<<1, 1, 0, s>>>(scalars, G + t->n_total);
  TR_CUDA(cudaGetLastError(), "loss slot launch");
  if (t->world > 1) {
    // ONE collective: gradients of every variable + rank 0's moving statistics + the tower loss
    const int nrc = t->nccl.AllReduce(G, G, static_cast<size_t>(t->n_total + 1), /*ncclFloat32*/ 7, /*ncclSum*/ 0, t->comm, s);
    if (nrc != 0) return tr_fail_nccl(c, t, nrc, "ncclAllReduce");
    count_launch(c, "train.allreduce");
  }
  const double* sqn = nullptr;
  if (t->clip_norm > 0.f) {
    // clipping acts on the AVERAGED gradient (model/trainer.py:1244-1246), per variable
    double* d_sqn = static_cast<double*>(t->sqnorms.p);
    TR_CUDA(cudaMemsetAsync(d_sqn, 0, sizeof(double) * t->segs.size(), s), "memset norms");
    count_launch(c, "train.grad_norms");
    grad_sqnorm_kernel<<<dim3(64, static_cast<unsigned>(t->segs.size())), 256, 0, s>>>(
        P, G, static_cast<const ParamSeg*>(t->d_segs.p), 1.f / t->world, d_sqn);
    TR_CUDA(cudaGetLastError(), "gradient norm launch");
    sqn = d_sqn;
  }
  float lr_eff = lr;
  if (t->optimizer == 2) {
    ++t->adam_t;
    lr_eff = static_cast<float>(lr * std::sqrt(1.0 - std::pow(0.999, static_cast<double>(t->adam_t))) /
                                (1.0 - std::pow(0.9, static_cast<double>(t->adam_t))));
  }
  count_launch(c, "train.update");
  apply_update_kernel<<<ew_grid(c), 256, 0, s>>>(P, G, static_cast<float*>(t->accum.p), static_cast<const ParamSeg*>(t->d_segs.p),
                                                 static_cast<int>(t->segs.size()), t->n_total, 1.f / t->world, lr_eff,
                                                 t->optimizer, t->opt_momentum, t->nesterov, sqn, t->clip_norm);
  TR_CUDA(cudaGetLastError(), "update launch");
  if (loss_out) {
    float loss_sum = 0.f;
    double sc[2] = {0, 0};
    TR_CUDA(cudaMemcpyAsync(&loss_sum, G + t->n_total, sizeof(float), cudaMemcpyDeviceToHost, s), "D2H loss");
    TR_CUDA(cudaMemcpyAsync(sc, scalars, sizeof(double) * 2, cudaMemcpyDeviceToHost, s), "D2H scalars");
    TR_CUDA(cudaStreamSynchronize(s), "stream synchronize");
    loss_out[0] = loss_sum / t->world;                 // mean of the tower losses (model/trainer.py:1218-1219)
    loss_out[1] = loss_out[0] + static_cast<float>(sc[1]);
  }
  return XV_OK;
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
  return XV_OK;
}

}  // extern "C"
