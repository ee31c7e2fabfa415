// Pooling kernels: ragged (optionally attention-weighted) mean / std over packed utterances.
//
// statistics_pooling (model/pooling.py:40-50) and the weighted statistics of self_attention
// (model/pooling.py:186-218) are one streaming reduction over a [rows, C] matrix:
//   mean[c] = sum_l w_l x[l,c],  var[c] = sum_l w_l (x[l,c]-mean[c])^2,  w_l = 1/L or softmax weights.
// The reference makes two passes; here ONE pass reads x exactly once (HBM-bound: C*sizeof(T) bytes per frame)
// and stays accurate by accumulating about a per-(utterance,column) pivot p = x[first row]:
//   S1 = sum w (x-p), S2 = sum w (x-p)^2  ->  mean = p + S1/W,  var = S2/W - (S1/W)^2,  W = sum w,
// which has no catastrophic cancellation because p lies inside the data range.
//
// Work decomposition: the packed row space is cut into fixed-size row chunks (grid.x) so the load is
// balanced regardless of utterance lengths; a chunk walks the utterance segments it intersects and
// flushes one fp32 atomicAdd per (segment, column).  Threads own V consecutive columns (one 16/8-byte
// vector load per row) and `rl` row lanes stride the rows; loads are unrolled x4 for memory-level parallelism.
#include <float.h>

#include <stdlib.h>

#include "xv_internal.h"
#include "xv_ptx.cuh"

namespace xv {

namespace {

template <typename T, int V>
struct VecLoad;
template <>
struct VecLoad<float, 4> {
  static __device__ __forceinline__ void ld(const float* p, float (&o)[4]) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
};
template <>
struct VecLoad<float, 2> {
  static __device__ __forceinline__ void ld(const float* p, float (&o)[2]) {
    const float2 v = __ldg(reinterpret_cast<const float2*>(p));
    o[0] = v.x; o[1] = v.y;
  }
};
template <>
struct VecLoad<float, 1> {
  static __device__ __forceinline__ void ld(const float* p, float (&o)[1]) { o[0] = __ldg(p); }
};
__device__ __forceinline__ void unpack_bf16x2(uint32_t u, float& a, float& b) {
  a = __uint_as_float(u << 16);
  b = __uint_as_float(u & 0xffff0000u);
}
template <>
struct VecLoad<__nv_bfloat16, 8> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    unpack_bf16x2(v.x, o[0], o[1]); unpack_bf16x2(v.y, o[2], o[3]);
    unpack_bf16x2(v.z, o[4], o[5]); unpack_bf16x2(v.w, o[6], o[7]);
  }
};
template <>
struct VecLoad<__nv_bfloat16, 4> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[4]) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    unpack_bf16x2(v.x, o[0], o[1]); unpack_bf16x2(v.y, o[2], o[3]);
  }
};
template <>
struct VecLoad<__nv_bfloat16, 2> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[2]) {
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(p));
    unpack_bf16x2(v, o[0], o[1]);
  }
};
template <>
struct VecLoad<__nv_bfloat16, 1> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[1]) {
    o[0] = __bfloat162float(p[0]);
  }
};

struct PoolAccumParams {
  const void* x;
  int64_t rows;
  int32_t c, ldx;
  const int32_t* cu;
  int32_t n, context;
  const float* weights;   // [heads][rows] or null
  int32_t head_width;     // split_value: columns per head (weights head = col / head_width); else 0 -> head = blockIdx.z
  float* acc;             // [n][hout][2][c]
  int32_t hout;
  int32_t rows_per_chunk, threads_x, rl;
  // bulk (TMA-staged) variant
  int32_t rows_per_stage;   // even; stage = rows_per_stage * ldx * sizeof(T) bytes <= kPoolStageBytes
  int32_t span;             // rows per CTA (multiple of rows_per_stage)
  int64_t bulk_end;         // rows [0, bulk_end) may be fetched with 16-byte-granular bulk copies
};

constexpr int kPoolStages = 3;
constexpr uint32_t kPoolStageBytes = 32768;
constexpr uint32_t kPoolBulkSmem = kPoolStages * kPoolStageBytes + 64;

template <typename T, int V>
struct SmemLoad;
template <>
struct SmemLoad<float, 4> {
  static __device__ __forceinline__ void ld(const float* p, float (&o)[4]) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
};
template <>
struct SmemLoad<float, 2> {
  static __device__ __forceinline__ void ld(const float* p, float (&o)[2]) {
    const float2 v = *reinterpret_cast<const float2*>(p);
    o[0] = v.x; o[1] = v.y;
  }
};
template <>
struct SmemLoad<__nv_bfloat16, 8> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[8]) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    unpack_bf16x2(v.x, o[0], o[1]); unpack_bf16x2(v.y, o[2], o[3]);
    unpack_bf16x2(v.z, o[4], o[5]); unpack_bf16x2(v.w, o[6], o[7]);
  }
};
template <>
struct SmemLoad<__nv_bfloat16, 4> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[4]) {
    const uint2 v = *reinterpret_cast<const uint2*>(p);
    unpack_bf16x2(v.x, o[0], o[1]); unpack_bf16x2(v.y, o[2], o[3]);
  }
};

template <typename T, int V>
__global__ void __launch_bounds__(512) pool_accum_kernel(const PoolAccumParams a) {
  __shared__ float red[512 * 2 * V > 4096 ? 4096 : 512 * 2 * V];   // [blockDim][2V] capped at 16 KB
  const T* __restrict__ x = reinterpret_cast<const T*>(a.x);
  const int cvi = threadIdx.x % a.threads_x;
  const int rli = threadIdx.x / a.threads_x;
  const int col0 = (blockIdx.y * a.threads_x + cvi) * V;
  const bool col_ok = col0 < a.c;
  const int hz = blockIdx.z;
  const int64_t R0 = static_cast<int64_t>(blockIdx.x) * a.rows_per_chunk;
  const int64_t R1 = min(R0 + a.rows_per_chunk, a.rows);
  const float* wrow = nullptr;
  if (a.weights != nullptr) {
    const int h = a.head_width > 0 ? (col_ok ? col0 / a.head_width : 0) : hz;
    wrow = a.weights + static_cast<int64_t>(h) * a.rows;
  }
  // first utterance whose end lies beyond R0
  int lo = 0, hi = a.n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (static_cast<int64_t>(__ldg(a.cu + mid + 1)) > R0) hi = mid; else lo = mid + 1;
  }
  for (int u = lo; u < a.n; ++u) {
    const int64_t cs = __ldg(a.cu + u);
    if (cs >= R1) break;
    const int64_t ve = static_cast<int64_t>(__ldg(a.cu + u + 1)) - a.context;
    const int64_t s0 = cs > R0 ? cs : R0;
    const int64_t s1 = ve < R1 ? ve : R1;
    if (s0 >= s1) continue;
    float s[V], q[V], piv[V];
#pragma unroll
    for (int e = 0; e < V; ++e) { s[e] = 0.f; q[e] = 0.f; piv[e] = 0.f; }
    if (col_ok) {
      VecLoad<T, V>::ld(x + cs * a.ldx + col0, piv);
      int64_t r = s0 + rli;
      const int64_t step = a.rl;
      for (; r + 3 * step < s1; r += 4 * step) {
        float v0[V], v1[V], v2[V], v3[V];
        VecLoad<T, V>::ld(x + r * a.ldx + col0, v0);
        VecLoad<T, V>::ld(x + (r + step) * a.ldx + col0, v1);
        VecLoad<T, V>::ld(x + (r + 2 * step) * a.ldx + col0, v2);
        VecLoad<T, V>::ld(x + (r + 3 * step) * a.ldx + col0, v3);
        float w0 = 1.f, w1 = 1.f, w2 = 1.f, w3 = 1.f;
        if (wrow) { w0 = __ldg(wrow + r); w1 = __ldg(wrow + r + step); w2 = __ldg(wrow + r + 2 * step); w3 = __ldg(wrow + r + 3 * step); }
#pragma unroll
        for (int e = 0; e < V; ++e) {
          const float d0 = v0[e] - piv[e], d1 = v1[e] - piv[e], d2 = v2[e] - piv[e], d3 = v3[e] - piv[e];
          s[e] += w0 * d0; q[e] = fmaf(w0 * d0, d0, q[e]);
          s[e] += w1 * d1; q[e] = fmaf(w1 * d1, d1, q[e]);
          s[e] += w2 * d2; q[e] = fmaf(w2 * d2, d2, q[e]);
          s[e] += w3 * d3; q[e] = fmaf(w3 * d3, d3, q[e]);
        }
      }
      for (; r < s1; r += step) {
        float v0[V];
        VecLoad<T, V>::ld(x + r * a.ldx + col0, v0);
        const float w0 = wrow ? __ldg(wrow + r) : 1.f;
#pragma unroll
        for (int e = 0; e < V; ++e) {
          const float d0 = v0[e] - piv[e];
          s[e] += w0 * d0; q[e] = fmaf(w0 * d0, d0, q[e]);
        }
      }
    }
    if (a.rl > 1) {
      float* mine = red + threadIdx.x * (2 * V);
#pragma unroll
      for (int e = 0; e < V; ++e) { mine[e] = s[e]; mine[V + e] = q[e]; }
      __syncthreads();
      if (rli == 0) {
        for (int k = 1; k < a.rl; ++k) {
          const float* o = red + (k * a.threads_x + cvi) * (2 * V);
#pragma unroll
          for (int e = 0; e < V; ++e) { s[e] += o[e]; q[e] += o[V + e]; }
        }
      }
      __syncthreads();
    }
    if (rli == 0 && col_ok) {
      float* dst = a.acc + ((static_cast<int64_t>(u) * a.hout + hz) * 2) * a.c + col0;
#pragma unroll
      for (int e = 0; e < V; ++e) {
        atomicAdd(dst + e, s[e]);
        atomicAdd(dst + a.c + e, q[e]);
      }
    }
  }
}

// Bulk-staged variant of pool_accum_kernel for the HBM roofline: a persistent CTA owns a contiguous span of
// packed rows; warp 0 streams it through a 3-stage shared-memory ring with 1-D bulk copies (cp.async.bulk,
// 30 KB per request, completion on an mbarrier) so ~90 KB per CTA (two CTAs per SM) are always in flight
// without any per-thread load instruction; the consumer warps (one thread per V columns) reduce each stage
// from shared memory with conflict-free vector loads, walk the utterance segments the span crosses and flush
// one fp32 atomicAdd per (segment, column).  Same pivot-shifted sums as pool_accum_kernel.
template <typename T, int V>
__global__ void __launch_bounds__(544) pool_accum_bulk_kernel(const PoolAccumParams a) {
  extern __shared__ __align__(128) uint8_t pool_smem[];
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(pool_smem + kPoolStages * kPoolStageBytes);
  uint64_t* empty_bar = full_bar + kPoolStages;
  const T* __restrict__ x = reinterpret_cast<const T*>(a.x);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_cons_warps = (blockDim.x >> 5) - 1;
  const int64_t R0 = static_cast<int64_t>(blockIdx.x) * a.span;
  const int64_t R1 = min(R0 + a.span, a.rows);
  if (R0 >= R1) return;
  const int n_chunks = static_cast<int>((R1 - R0 + a.rows_per_stage - 1) / a.rows_per_stage);
  if (threadIdx.x == 0) {
    for (int s = 0; s < kPoolStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], n_cons_warps);
    }
    ptx::fence_mbar_init();
  }
  __syncthreads();
  const uint32_t row_bytes = static_cast<uint32_t>(a.ldx) * sizeof(T);
  if (warp == 0) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      for (int i = 0; i < n_chunks; ++i) {
        const int64_t c0 = R0 + static_cast<int64_t>(i) * a.rows_per_stage;
        int64_t c1 = min(c0 + a.rows_per_stage, R1);
        c1 = min(c1, a.bulk_end);                       // a trailing odd row is read straight from global memory
        ptx::mbar_wait(&empty_bar[st], ph ^ 1);
        if (c1 > c0) {
          const uint32_t bytes = static_cast<uint32_t>(c1 - c0) * row_bytes;
          ptx::mbar_arrive_expect_tx(&full_bar[st], bytes);
          ptx::bulk_load_1d(pool_smem + st * kPoolStageBytes, x + c0 * a.ldx, bytes, &full_bar[st]);
        } else {
          ptx::mbar_arrive(&full_bar[st]);
        }
        if (++st == kPoolStages) { st = 0; ph ^= 1; }
      }
    }
    return;
  }
  const int cvi = threadIdx.x - 32;
  const int col0 = cvi * V;
  const bool col_ok = col0 < a.c;
  const int hz = blockIdx.y;
  const float* wrow = nullptr;
  if (a.weights != nullptr) {
    const int h = a.head_width > 0 ? (col_ok ? col0 / a.head_width : 0) : hz;
    wrow = a.weights + static_cast<int64_t>(h) * a.rows;
  }
  // first utterance whose end lies beyond R0
  int u = 0;
  {
    int lo = 0, hi = a.n;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (static_cast<int64_t>(__ldg(a.cu + mid + 1)) > R0) hi = mid; else lo = mid + 1;
    }
    u = lo;
  }
  float s[V], q[V], piv[V];
  int64_t cs = 0, ve = 0, ue = 0;            // utterance start, valid end, end
  bool dirty = false;
  auto open_utt = [&]() {
#pragma unroll
    for (int e = 0; e < V; ++e) { s[e] = 0.f; q[e] = 0.f; piv[e] = 0.f; }
    dirty = false;
    if (u < a.n) {
      cs = __ldg(a.cu + u);
      ue = __ldg(a.cu + u + 1);
      ve = ue - a.context;
      if (col_ok && cs < ve) VecLoad<T, V>::ld(x + cs * a.ldx + col0, piv);
    } else {
      cs = ve = ue = a.rows + 1;
    }
  };
  auto flush = [&]() {
    if (dirty && col_ok) {
      float* dst = a.acc + ((static_cast<int64_t>(u) * a.hout + hz) * 2) * a.c + col0;
#pragma unroll
      for (int e = 0; e < V; ++e) {
        atomicAdd(dst + e, s[e]);
        atomicAdd(dst + a.c + e, q[e]);
      }
    }
  };
  open_utt();
  uint32_t st = 0, ph = 0;
  for (int i = 0; i < n_chunks; ++i) {
    const int64_t c0 = R0 + static_cast<int64_t>(i) * a.rows_per_stage;
    const int64_t c1 = min(c0 + a.rows_per_stage, R1);
    ptx::mbar_wait(&full_bar[st], ph);
    const T* sbase = reinterpret_cast<const T*>(pool_smem + st * kPoolStageBytes) + col0;
    if (c0 >= cs && c1 <= ve && c1 <= a.bulk_end) {
      // fast path: the whole stage lies inside the current utterance's valid rows
      if (col_ok) {
        const int nr = static_cast<int>(c1 - c0);
        if (wrow == nullptr) {
#pragma unroll 5
          for (int r = 0; r < nr; ++r) {
            float v[V];
            SmemLoad<T, V>::ld(sbase + r * a.ldx, v);
#pragma unroll
            for (int e = 0; e < V; ++e) {
              const float d = v[e] - piv[e];
              s[e] += d;
              q[e] = fmaf(d, d, q[e]);
            }
          }
        } else {
          for (int r = 0; r < nr; ++r) {
            float v[V];
            SmemLoad<T, V>::ld(sbase + r * a.ldx, v);
            const float w = __ldg(wrow + c0 + r);
#pragma unroll
            for (int e = 0; e < V; ++e) {
              const float d = v[e] - piv[e];
              s[e] = fmaf(w, d, s[e]);
              q[e] = fmaf(w * d, d, q[e]);
            }
          }
        }
        dirty = true;
      }
    } else {
      for (int64_t r = c0; r < c1; ++r) {
        while (r >= ue && u < a.n) { flush(); ++u; open_utt(); }
        if (r < cs || r >= ve || !col_ok) continue;
        float v[V];
        if (r < a.bulk_end) SmemLoad<T, V>::ld(sbase + (r - c0) * a.ldx, v);
        else VecLoad<T, V>::ld(x + r * a.ldx + col0, v);
        const float w = wrow ? __ldg(wrow + r) : 1.f;
#pragma unroll
        for (int e = 0; e < V; ++e) {
          const float d = v[e] - piv[e];
          s[e] = fmaf(w, d, s[e]);
          q[e] = fmaf(w * d, d, q[e]);
        }
        dirty = true;
      }
    }
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(&empty_bar[st]);
    if (++st == kPoolStages) { st = 0; ph ^= 1; }
  }
  flush();
}

template <typename AccT>
struct PoolFinalizeParams {
  const AccT* acc;       // [n][hout][2][acc_ld]
  int32_t acc_ld, n, hout, c;
  const int32_t* cu;
  int32_t context;
  int32_t weighted;      // 1: weights sum to one (no division by the count)
  const void* x;         // pivot source (row cu[u]) or null (unshifted sums)
  int32_t dtype, ldx;
  float* out; int32_t ldout;
  __nv_bfloat16* out_bf16; int32_t ldout_bf16;
};

template <typename AccT>
__global__ void pool_finalize_kernel(const PoolFinalizeParams<AccT> a) {
  const int64_t total = static_cast<int64_t>(a.n) * a.hout * a.c;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % a.c);
    const int64_t t = i / a.c;
    const int hh = static_cast<int>(t % a.hout);
    const int u = static_cast<int>(t / a.hout);
    const AccT* p = a.acc + ((static_cast<int64_t>(u) * a.hout + hh) * 2) * a.acc_ld + c;
    const int32_t cs = a.cu[u];
    const int32_t len = a.cu[u + 1] - cs - a.context;
    const AccT inv = a.weighted ? AccT(1) : AccT(1) / static_cast<AccT>(len);
    const AccT m_acc = p[0] * inv;
    const float m = static_cast<float>(m_acc);
    float var = static_cast<float>(p[a.acc_ld] * inv - m_acc * m_acc);
    float piv = 0.f;
    if (a.x != nullptr) {
      piv = a.dtype == 1 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(a.x)[static_cast<int64_t>(cs) * a.ldx + c])
                         : reinterpret_cast<const float*>(a.x)[static_cast<int64_t>(cs) * a.ldx + c];
    }
    const float mean = piv + m;
    var = var <= 1e-12f ? 1e-12f : var;          // VAR2STD_EPSILON floor (model/pooling.py:46-47)
    const float sd = sqrtf(var);
    const int64_t om = static_cast<int64_t>(hh) * a.c + c;
    const int64_t os = static_cast<int64_t>(a.hout) * a.c + om;
    if (a.out) {
      a.out[static_cast<int64_t>(u) * a.ldout + om] = mean;
      a.out[static_cast<int64_t>(u) * a.ldout + os] = sd;
    }
    if (a.out_bf16) {
      a.out_bf16[static_cast<int64_t>(u) * a.ldout_bf16 + om] = __float2bfloat16_rn(mean);
      a.out_bf16[static_cast<int64_t>(u) * a.ldout_bf16 + os] = __float2bfloat16_rn(sd);
    }
  }
}

// scores[h, r] = scale * sum_d key[r, koff(h) + d] * q[h, d]; one warp per row (model/pooling.py:186-194).
template <typename T>
__global__ void att_scores_kernel(const T* __restrict__ key, int ldk, int64_t rows, const float* __restrict__ query,
                                  int heads, int dk, int split_key, float scale, float* __restrict__ scores) {
  const int64_t r = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= rows) return;
  const T* krow = key + r * ldk;
  for (int h = 0; h < heads; ++h) {
    const T* kp = krow + (split_key ? h * dk : 0);
    const float* qp = query + h * dk;
    float s = 0.f;
    for (int d = lane; d < dk; d += 32) s = fmaf(static_cast<float>(kp[d]), __ldg(qp + d), s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) scores[static_cast<int64_t>(h) * rows + r] = s * scale;
  }
}

// Softmax over the valid rows of each (utterance, head) (model/pooling.py:197): fp32 in place, or from the fp64 score
// accumulators of the fused key-layer epilogue (src != null) into the fp32 weight buffer.
template <typename S>
__global__ void att_softmax_kernel(const S* __restrict__ src, float* __restrict__ scores, int64_t rows,
                                   const int32_t* __restrict__ cu, int context) {
  __shared__ float sh[32];
  const int u = blockIdx.x, h = blockIdx.y;
  const int32_t cs = cu[u];
  const int32_t len = cu[u + 1] - cs - context;
  const S* in = src + static_cast<int64_t>(h) * rows + cs;
  float* s = scores + static_cast<int64_t>(h) * rows + cs;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float m = -FLT_MAX;
  for (int i = threadIdx.x; i < len; i += blockDim.x) m = fmaxf(m, static_cast<float>(in[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) sh[wid] = m;
  __syncthreads();
  m = sh[0];
  for (int k = 1; k < nw; ++k) m = fmaxf(m, sh[k]);
  __syncthreads();
  float z = 0.f;
  for (int i = threadIdx.x; i < len; i += blockDim.x) {
    const float e = expf(static_cast<float>(in[i]) - m);
    s[i] = e;
    z += e;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
  if (lane == 0) sh[wid] = z;
  __syncthreads();
  z = 0.f;
  for (int k = 0; k < nw; ++k) z += sh[k];
  const float inv = 1.f / z;
  for (int i = threadIdx.x; i < len; i += blockDim.x) s[i] *= inv;
}

// dst + dst_off[u] <- fp32 copy of rows [src_row0[u], +n_rows[u]) x width of src (bf16 or fp32).
template <typename T>
__global__ void gather_rows_kernel(const T* __restrict__ src, int lds, int width, const int64_t* __restrict__ src_row0,
                                   const int32_t* __restrict__ n_rows, const int64_t* __restrict__ dst_off,
                                   float* __restrict__ dst) {
  const int u = blockIdx.y;
  const int64_t r0 = src_row0[u];
  const int nr = n_rows[u];
  float* d = dst + dst_off[u];
  for (int r = blockIdx.x; r < nr; r += gridDim.x) {
    const T* s = src + (r0 + r) * lds;
    for (int c = threadIdx.x; c < width; c += blockDim.x) d[static_cast<int64_t>(r) * width + c] = static_cast<float>(s[c]);
  }
}

// attention_weights endpoint: dst + dst_off[u] + h*len + l <- w[h][cu[u] + l]
__global__ void gather_att_weights_kernel(const float* __restrict__ w, int64_t rows, const int32_t* __restrict__ cu,
                                          int context, int heads, const int64_t* __restrict__ dst_off,
                                          float* __restrict__ dst) {
  const int u = blockIdx.x;
  const int32_t cs = cu[u];
  const int32_t len = cu[u + 1] - cs - context;
  float* d = dst + dst_off[u];
  for (int h = 0; h < heads; ++h)
    for (int i = threadIdx.x; i < len; i += blockDim.x) d[static_cast<int64_t>(h) * len + i] = w[static_cast<int64_t>(h) * rows + cs + i];
}

template <typename T>
int pick_vec(const void* base, int ldx, int c, int head_width, int vmax) {
  int v = vmax;
  while (v > 1) {
    const bool ok = (ldx % v == 0) && (c % v == 0) && (reinterpret_cast<uintptr_t>(base) % (v * sizeof(T)) == 0) &&
                    (head_width <= 0 || head_width % v == 0);
    if (ok) break;
    v >>= 1;
  }
  return v;
}

template <typename T, int V>
cudaError_t launch_accum_v(const PoolAccumParams& p, dim3 grid, int block, cudaStream_t stream) {
  pool_accum_kernel<T, V><<<grid, block, 0, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace

static cudaError_t launch_pool_finalize_f32(const PoolArgs& a, int hout, int sm_count, cudaStream_t stream) {
  PoolFinalizeParams<float> f;
  f.acc = a.acc; f.acc_ld = a.c; f.n = a.n; f.hout = hout; f.c = a.c; f.cu = a.cu; f.context = a.context;
  f.weighted = a.weights != nullptr; f.x = a.x; f.dtype = a.dtype; f.ldx = a.ldx;
  f.out = a.out; f.ldout = a.ldout; f.out_bf16 = a.out_bf16; f.ldout_bf16 = a.ldout_bf16;
  const int64_t total = static_cast<int64_t>(a.n) * hout * a.c;
  int64_t blocks = (total + 255) / 256;
  if (blocks > sm_count * 8) blocks = sm_count * 8;
  pool_finalize_kernel<float><<<static_cast<int>(blocks), 256, 0, stream>>>(f);
  return cudaGetLastError();
}

size_t pool_workspace_bytes(int32_t n, int32_t hout, int32_t c) {
  return static_cast<size_t>(n) * hout * 2 * c * sizeof(float);
}

cudaError_t launch_pool(const PoolArgs& a, int sm_count, cudaStream_t stream) {
  if (a.n <= 0 || a.rows <= 0) return cudaSuccess;
  const int hout = (a.weights != nullptr && !a.split_value) ? a.heads : 1;
  cudaError_t e = cudaMemsetAsync(a.acc, 0, pool_workspace_bytes(a.n, hout, a.c), stream);
  if (e != cudaSuccess) return e;
  const int head_width = (a.weights != nullptr && a.split_value) ? a.c / a.heads : 0;
  PoolAccumParams p;
  p.x = a.x; p.rows = a.rows; p.c = a.c; p.ldx = a.ldx; p.cu = a.cu; p.n = a.n; p.context = a.context;
  p.weights = a.weights; p.head_width = head_width; p.acc = a.acc; p.hout = hout;
  int v;
  if (a.dtype == 1) v = pick_vec<__nv_bfloat16>(a.x, a.ldx, a.c, head_width, 8);
  else v = pick_vec<float>(a.x, a.ldx, a.c, head_width, 4);
  const int nv = (a.c + v - 1) / v;
  // ---- bulk-staged kernel whenever the layout allows 16-byte-granular bulk copies ----
  {
    static const bool force_ldg = getenv("XV_POOL_IMPL") != nullptr && getenv("XV_POOL_IMPL")[0] == 'l';
    const size_t esz = a.dtype == 1 ? 2 : 4;
    const size_t row_bytes = static_cast<size_t>(a.ldx) * esz;
    const bool v_ok = a.dtype == 1 ? (v == 8 || v == 4) : (v == 4 || v == 2);
    if (!force_ldg && v_ok && nv <= 512 && row_bytes % 8 == 0 && 2 * row_bytes <= kPoolStageBytes &&
        reinterpret_cast<uintptr_t>(a.x) % 16 == 0) {
      int rps = static_cast<int>(kPoolStageBytes / row_bytes) & ~1;
      if (rps > 64) rps = 64;
      const int ctas = sm_count * 2;
      int64_t span = (a.rows + ctas - 1) / ctas;
      span = ((span + rps - 1) / rps) * rps;
      p.rows_per_stage = rps;
      p.span = static_cast<int32_t>(span);
      p.bulk_end = row_bytes % 16 == 0 ? a.rows : (a.rows & ~int64_t(1));
      p.threads_x = 0; p.rl = 1; p.rows_per_chunk = 0;
      dim3 grid(static_cast<unsigned>((a.rows + span - 1) / span), static_cast<unsigned>(hout));
      const int block = 32 + ((nv + 31) / 32) * 32;
      // function attributes are per device: one ctx per GPU may live in one process (multi-GPU extractor, TrainerMGPU)
      static bool attr_done_dev[64] = {};
      int cur_dev = 0;
      cudaGetDevice(&cur_dev);
      bool& attr_done = attr_done_dev[cur_dev & 63];
      if (!attr_done) {
        cudaFuncSetAttribute(pool_accum_bulk_kernel<__nv_bfloat16, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPoolBulkSmem);
        cudaFuncSetAttribute(pool_accum_bulk_kernel<__nv_bfloat16, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPoolBulkSmem);
        cudaFuncSetAttribute(pool_accum_bulk_kernel<float, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPoolBulkSmem);
        cudaFuncSetAttribute(pool_accum_bulk_kernel<float, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPoolBulkSmem);
        attr_done = true;
      }
      if (a.dtype == 1 && v == 8) pool_accum_bulk_kernel<__nv_bfloat16, 8><<<grid, block, kPoolBulkSmem, stream>>>(p);
      else if (a.dtype == 1) pool_accum_bulk_kernel<__nv_bfloat16, 4><<<grid, block, kPoolBulkSmem, stream>>>(p);
      else if (v == 4) pool_accum_bulk_kernel<float, 4><<<grid, block, kPoolBulkSmem, stream>>>(p);
      else pool_accum_bulk_kernel<float, 2><<<grid, block, kPoolBulkSmem, stream>>>(p);
      e = cudaGetLastError();
      if (e != cudaSuccess) return e;
      return launch_pool_finalize_f32(a, hout, sm_count, stream);
    }
  }
  int tx = nv < 512 ? ((nv + 31) / 32) * 32 : 512;
  int rl = tx >= 256 ? 1 : 256 / tx;
  // keep the cross-row-lane reduction buffer within its 16 KB cap
  while (rl > 1 && tx * rl * 2 * v > 4096) rl >>= 1;
  p.threads_x = tx; p.rl = rl;
  int64_t rpc = a.rows / (static_cast<int64_t>(sm_count) * 8);
  if (rpc < 64) rpc = 64;
  if (rpc > 512) rpc = 512;
  rpc = ((rpc + 4 * rl - 1) / (4 * rl)) * (4 * rl);
  p.rows_per_chunk = static_cast<int32_t>(rpc);
  dim3 grid(static_cast<unsigned>((a.rows + rpc - 1) / rpc), static_cast<unsigned>((nv + tx - 1) / tx),
            static_cast<unsigned>(hout));
  const int block = tx * rl;
  if (a.dtype == 1) {
    switch (v) {
      case 8: e = launch_accum_v<__nv_bfloat16, 8>(p, grid, block, stream); break;
      case 4: e = launch_accum_v<__nv_bfloat16, 4>(p, grid, block, stream); break;
      case 2: e = launch_accum_v<__nv_bfloat16, 2>(p, grid, block, stream); break;
      default: e = launch_accum_v<__nv_bfloat16, 1>(p, grid, block, stream); break;
    }
  } else {
    switch (v) {
      case 4: e = launch_accum_v<float, 4>(p, grid, block, stream); break;
      case 2: e = launch_accum_v<float, 2>(p, grid, block, stream); break;
      default: e = launch_accum_v<float, 1>(p, grid, block, stream); break;
    }
  }
  if (e != cudaSuccess) return e;
  return launch_pool_finalize_f32(a, hout, sm_count, stream);
}

cudaError_t launch_pool_finalize(const double* acc, int acc_ld, const int32_t* cu, int context, int n, int c,
                                 float* out_f32, int ldo_f32, __nv_bfloat16* out_bf16, int ldo_bf16, int weighted,
                                 cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  PoolFinalizeParams<double> f;
  f.acc = acc; f.acc_ld = acc_ld; f.n = n; f.hout = 1; f.c = c; f.cu = cu; f.context = context;
  f.weighted = weighted; f.x = nullptr; f.dtype = 0; f.ldx = 0;
  f.out = out_f32; f.ldout = ldo_f32; f.out_bf16 = out_bf16; f.ldout_bf16 = ldo_bf16;
  const int64_t total = static_cast<int64_t>(n) * c;
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  pool_finalize_kernel<double><<<static_cast<int>(blocks), 256, 0, stream>>>(f);
  return cudaGetLastError();
}

cudaError_t launch_att_scores(const void* key, int dtype, int ldk, int64_t rows, const float* query, int heads,
                              int dk, int split_key, float scale, float* scores, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  const unsigned blocks = static_cast<unsigned>((rows + 7) / 8);
  if (dtype == 1)
    att_scores_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(key), ldk, rows,
                                                                  query, heads, dk, split_key, scale, scores);
  else
    att_scores_kernel<float><<<blocks, 256, 0, stream>>>(reinterpret_cast<const float*>(key), ldk, rows, query, heads,
                                                          dk, split_key, scale, scores);
  return cudaGetLastError();
}

cudaError_t launch_att_softmax(float* scores, int64_t rows, const int32_t* cu, int context, int n, int heads,
                               cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  att_softmax_kernel<float><<<dim3(n, heads), 256, 0, stream>>>(scores, scores, rows, cu, context);
  return cudaGetLastError();
}

cudaError_t launch_att_softmax_f64(const double* acc, float* weights, int64_t rows, const int32_t* cu, int context, int n,
                                   int heads, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  att_softmax_kernel<double><<<dim3(n, heads), 256, 0, stream>>>(acc, weights, rows, cu, context);
  return cudaGetLastError();
}

cudaError_t launch_gather_rows(const void* src, int dtype, int lds, int width, const int64_t* src_row0,
                               const int32_t* n_rows, const int64_t* dst_off, int n_utts, int max_rows, float* dst,
                               cudaStream_t stream) {
  if (n_utts <= 0 || max_rows <= 0) return cudaSuccess;
  int bx = max_rows < 64 ? max_rows : 64;
  dim3 grid(bx, n_utts);
  const int block = width >= 256 ? 256 : (width >= 128 ? 128 : 64);
  if (dtype == 1)
    gather_rows_kernel<__nv_bfloat16><<<grid, block, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(src), lds, width,
                                                                   src_row0, n_rows, dst_off, dst);
  else
    gather_rows_kernel<float><<<grid, block, 0, stream>>>(reinterpret_cast<const float*>(src), lds, width, src_row0,
                                                           n_rows, dst_off, dst);
  return cudaGetLastError();
}

cudaError_t launch_gather_att_weights(const float* w, int64_t rows, const int32_t* cu, int context, int heads, int n,
                                      const int64_t* dst_off, float* dst, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  gather_att_weights_kernel<<<n, 256, 0, stream>>>(w, rows, cu, context, heads, dst_off, dst);
  return cudaGetLastError();
}

}  // namespace xv
