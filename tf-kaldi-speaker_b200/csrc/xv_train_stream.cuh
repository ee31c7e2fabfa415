// Bulk-staged versions of the three row-streaming kernels of the training step (xv_train_kernels.cuh: bn_act_fwd_kernel,
// col_reduce4_kernel<1>, bn_bwd_apply_kernel).  Same arithmetic, different data movement:
//
// The per-thread-load kernels keep 16 warps per SM with 4-8 16-byte loads each in flight and stall at ~2.7 TB/s (ncu:
// 28-34 % of DRAM peak, r02b / r02o).  Here one persistent CTA per SM owns a contiguous range of rows -- ONE contiguous
// byte range of each [rows, C] fp32 tensor -- and a producer lane streams it through a 4-stage shared-memory ring with 1-D
// bulk copies (cp.async.bulk, ~24 KB per tensor per request, completion on an mbarrier): ~190 KB per SM are in flight with
// no load instruction in the consumer threads, which read the staged rows with conflict-free 16-byte shared loads (one
// thread per four channels, the per-channel constants in registers) and write their results straight to global memory.
// The pattern is pool_accum_bulk_kernel's (xv_pool.cu), which reaches 94 % of the HBM copy bandwidth.
//
// References: model/tdnn.py:49-58 (batch norm + activation, training mode), tf.gradients of the same under
// model/trainer.py:501-527; model/pooling.py:40-50 for the pooling gradient rebuilt on the fly (PoolGrad).
#pragma once

#include "xv_ptx.cuh"
#include "xv_train_kernels.cuh"

namespace xv {
namespace {

constexpr int kBsStages = 4;
constexpr uint32_t kBsStageBytes = 24576;            // per tensor and stage: 12 rows of 512 floats, 4 rows of 1500
constexpr int kBsMaxThreads = 544;                   // 1 producer warp + up to 512 consumers
constexpr uint32_t kBsSmemBytes = kBsStages * 2 * kBsStageBytes + 128;   // two tensors (z, dy) + barriers

enum BsMode { kBsFwd = 0, kBsReduce = 1, kBsApply = 2 };

struct BsArgs {
  // geometry
  const float* z; const float* dy;       // [rows, c_total] fp32, dense; dy null (forward, or the pooling gradient is virtual)
  int64_t rows; int32_t c_total; RowMask mask;
  int32_t rows_per_stage;                // kBsStageBytes / (4 c_total)
  int32_t qpad;                          // channel quads rounded up to a warp multiple: consumer thread t -> quad t % qpad, row lane t / qpad
  int64_t span;                          // rows per CTA
  // forward
  BnStat st; BnFinalize fin;
  float* y; __nv_bfloat16* yb; int32_t ldb;
  // backward
  const float* mean; const float* rstd;  // null: no batch norm
  const float* gamma; const float* beta; const float* slope;
  PoolGrad pg;
  double* sums_out; int32_t sums_ld;     // reduce: [3][sums_ld] (zero on entry); apply: the finished sums
  double n;
  float* dz32; __nv_bfloat16* dzb;       // apply outputs (ldb as above)
  float* g_gamma; float* g_beta; float* g_alpha;
};

__device__ __forceinline__ void bs_named_barrier(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(kBsMaxThreads) bn_stream_kernel(const BsArgs a) {
  extern __shared__ __align__(128) uint8_t bs_smem[];
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bs_smem + kBsStages * 2 * kBsStageBytes);
  uint64_t* empty_bar = full_bar + kBsStages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_cons = blockDim.x - 32;
  const int64_t R0 = static_cast<int64_t>(blockIdx.x) * a.span;
  const int64_t R1 = min(R0 + a.span, a.rows);
  const bool two = MODE != kBsFwd && a.dy != nullptr;           // a second tensor travels through the ring
  const int n_chunks = R1 > R0 ? static_cast<int>((R1 - R0 + a.rows_per_stage - 1) / a.rows_per_stage) : 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kBsStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], n_cons >> 5);
    }
    ptx::fence_mbar_init();
  }
  __syncthreads();
  const uint32_t row_bytes = static_cast<uint32_t>(a.c_total) * 4u;
  if (warp == 0) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      for (int i = 0; i < n_chunks; ++i) {
        const int64_t c0 = R0 + static_cast<int64_t>(i) * a.rows_per_stage;
        const int64_t c1 = min(c0 + a.rows_per_stage, R1);
        const uint32_t bytes = static_cast<uint32_t>(c1 - c0) * row_bytes;
        ptx::mbar_wait(&empty_bar[st], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&full_bar[st], two ? 2 * bytes : bytes);
        ptx::bulk_load_1d(bs_smem + (st * 2) * kBsStageBytes, a.z + c0 * a.c_total, bytes, &full_bar[st]);
        if (two) ptx::bulk_load_1d(bs_smem + (st * 2 + 1) * kBsStageBytes, a.dy + c0 * a.c_total, bytes, &full_bar[st]);
        if (++st == kBsStages) { st = 0; ph ^= 1; }
      }
    }
    return;
  }
  // ------------------------------------------------------------------------------------------- consumers
  const int tc = threadIdx.x - 32;
  const int q = tc % a.qpad, rl = tc / a.qpad;
  const int n_lanes = n_cons / a.qpad;
  const int nq = a.c_total >> 2;
  const bool active = q < nq && rl < n_lanes;
  const int c0 = q * 4;
  float sl[4], grf[4], e0[4], e1[4];         // apply: dz = g * grf + (e0 + xhat * e1), see bn_bwd_apply_kernel
  BnFold bf[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    sl[k] = 0.f; grf[k] = 1.f; e0[k] = 0.f; e1[k] = 0.f;
    bf[k] = bn_fold(false, 0.f, 1.f, 1.f, 0.f);
  }
  const bool has_bn = MODE == kBsFwd ? a.st.sums != nullptr : a.mean != nullptr;
  if (active) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      sl[k] = a.slope[c0 + k];
      if (MODE == kBsFwd) {
        if (has_bn) {
          double var;
          float mu, rs;
          bn_mean_rstd(a.st, c0 + k, a.fin.eps, &mu, &rs, &var);
          bf[k] = bn_fold(true, mu, rs, a.gamma[c0 + k], a.beta[c0 + k]);
          if (blockIdx.x == 0 && rl == 0) bn_finalize_channel(a.st, a.fin, c0 + k, mu, rs, var);
        }
      } else if (has_bn) {
        const float ga = a.gamma[c0 + k], rs = a.rstd[c0 + k];
        bf[k] = bn_fold(true, a.mean[c0 + k], rs, ga, a.beta[c0 + k]);
        if (MODE == kBsApply) {
          const double inv_n = 1.0 / a.n;
          const double gr = static_cast<double>(ga) * rs;
          grf[k] = static_cast<float>(gr);
          e0[k] = static_cast<float>(-gr * (a.sums_out[c0 + k] * inv_n));
          e1[k] = static_cast<float>(-gr * (a.sums_out[a.sums_ld + c0 + k] * inv_n));
        }
      }
      if (MODE == kBsApply && blockIdx.x == 0 && rl == 0) {
        if (a.g_beta != nullptr) a.g_beta[c0 + k] = static_cast<float>(a.sums_out[c0 + k]);
        if (a.g_gamma != nullptr) a.g_gamma[c0 + k] = static_cast<float>(a.sums_out[a.sums_ld + c0 + k]);
        if (a.g_alpha != nullptr) a.g_alpha[c0 + k] = static_cast<float>(a.sums_out[2 * a.sums_ld + c0 + k]);
      }
    }
  }
  // reduce: a thread sees at most span / n_lanes rows (22 of a 512-wide layer, 87 of the 1500-wide one at 64 x 200): its own
  // running sums stay fp32, everything they are combined with (row lanes, CTAs) is fp64
  float acc[3][4];
#pragma unroll
  for (int v = 0; v < 3; ++v)
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[v][k] = 0.f;
  int64_t cur_u = -1;
  PoolCoef pc;
#pragma unroll
  for (int k = 0; k < 4; ++k) { pc.mu[k] = 0.f; pc.ca[k] = 0.f; pc.cb[k] = 0.f; }
  const bool virt = MODE != kBsFwd && a.pg.enabled;

  // A thread's rows are R0 + rl, R0 + rl + n_lanes, ... (rows_per_stage is a multiple of n_lanes, so the walk continues
  // uniformly across stages): the position inside the utterance and the utterance index advance by addition -- a division per
  // row was a quarter of this kernel's instructions (ncu source view, r02t).  Everything is addressed with 32-bit offsets
  // from per-thread base pointers.
  const uint32_t t_len = static_cast<uint32_t>(a.mask.t), valid_len = static_cast<uint32_t>(a.mask.valid_len);
  const uint32_t first = static_cast<uint32_t>(R0) + static_cast<uint32_t>(rl);
  uint32_t rem = first % t_len;
  int32_t ut = static_cast<int32_t>(first / t_len);
  const uint32_t step_rows = static_cast<uint32_t>(n_lanes);
  const uint32_t row_f = static_cast<uint32_t>(a.c_total);                      // floats per staged row
  const size_t out_row0 = static_cast<size_t>(first);
  float* y_p = (MODE == kBsFwd && a.y != nullptr) ? a.y + out_row0 * a.c_total + c0 : nullptr;
  float* dz_p = (MODE == kBsApply && a.dz32 != nullptr) ? a.dz32 + out_row0 * a.c_total + c0 : nullptr;
  __nv_bfloat16* ob_p = nullptr;
  if (MODE == kBsFwd && a.yb != nullptr) ob_p = a.yb + out_row0 * a.ldb + c0;
  if (MODE == kBsApply && a.dzb != nullptr) ob_p = a.dzb + out_row0 * a.ldb + c0;
  const uint32_t f32_step = step_rows * row_f, b16_step = step_rows * static_cast<uint32_t>(a.ldb);
  const bool pad_cols = q == nq - 1 && a.ldb > a.c_total;

  uint32_t st = 0, ph = 0;
  for (int i = 0; i < n_chunks; ++i) {
    const int nr = min(a.rows_per_stage, static_cast<int>(R1 - R0) - i * a.rows_per_stage);
    ptx::mbar_wait(&full_bar[st], ph);
    if (active) {
      const float* zs = reinterpret_cast<const float*>(bs_smem + (st * 2) * kBsStageBytes) + c0 + rl * row_f;
      const float* ds = reinterpret_cast<const float*>(bs_smem + (st * 2 + 1) * kBsStageBytes) + c0 + rl * row_f;
#pragma unroll 2
      for (int rr = rl; rr < nr; rr += n_lanes) {
        const bool ok = rem < valid_len;
        float zz[4];
        VecF<4>::load(zs, zz);
        if (MODE == kBsFwd) {
          float v[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float t = fmaf(zz[k], bf[k].pa, bf[k].pb);                  // same expression in every backward kernel
            v[k] = ok ? (t >= 0.f ? t : t * sl[k]) : 0.f;
          }
          if (y_p != nullptr) { VecF<4>::store(y_p, v); y_p += f32_step; }
          if (ob_p != nullptr) {
            VecF<4>::store_bf16(ob_p, v);
            if (pad_cols)
              for (int k = a.c_total - c0; k < a.ldb - c0; ++k) ob_p[k] = __float2bfloat16(0.f);
            ob_p += b16_step;
          }
        } else {
          float dd[4] = {0.f, 0.f, 0.f, 0.f};
          if (two) VecF<4>::load(ds, dd);
          float out[4] = {0.f, 0.f, 0.f, 0.f};
          if (ok) {
            if (virt && ut != static_cast<int32_t>(cur_u)) {   // per-utterance pooling coefficients: reloaded when the row walk enters a new utterance
              cur_u = ut;
              pool_coef_load(a.pg, ut, a.c_total, c0, &pc);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float xh = fmaf(zz[k], bf[k].xa, bf[k].xb), pre = fmaf(zz[k], bf[k].pa, bf[k].pb);
              float d = dd[k];
              if (virt) d = fmaf(pc.cb[k], (pre >= 0.f ? pre : pre * sl[k]) - pc.mu[k], pc.ca[k]);
              const float g = pre > 0.f ? d : d * sl[k];
              if (MODE == kBsReduce) {
                acc[0][k] += g;
                acc[1][k] = fmaf(g, xh, acc[1][k]);
                if (pre < 0.f) acc[2][k] = fmaf(d, pre, acc[2][k]);
              } else {
                out[k] = fmaf(g, grf[k], fmaf(xh, e1[k], e0[k]));
              }
            }
          }
          if (MODE == kBsApply) {
            if (dz_p != nullptr) { VecF<4>::store(dz_p, out); dz_p += f32_step; }
            if (ob_p != nullptr) {
              VecF<4>::store_bf16(ob_p, out);
              if (pad_cols)
                for (int k = a.c_total - c0; k < a.ldb - c0; ++k) ob_p[k] = __float2bfloat16(0.f);
              ob_p += b16_step;
            }
          }
        }
        zs += f32_step;
        ds += f32_step;
        rem += step_rows;
        while (rem >= t_len) { rem -= t_len; ++ut; }
      }
    }
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(&empty_bar[st]);
    if (++st == kBsStages) { st = 0; ph ^= 1; }
  }
  if (MODE == kBsReduce) {
    // row lanes of a channel quad are combined through the (now idle) ring, lane 0 issues the fp64 atomics
    double* red = reinterpret_cast<double*>(bs_smem);           // [n_lanes][3][c_total]: <= 4 * 3 * 512 * 8 = 48 KB
    if (n_lanes > 1) {
      bs_named_barrier(1, n_cons);                               // every consumer has left the last stage
      if (active && rl > 0) {
#pragma unroll
        for (int v = 0; v < 3; ++v)
#pragma unroll
          for (int k = 0; k < 4; ++k) red[(static_cast<size_t>(rl) * 3 + v) * a.c_total + c0 + k] = static_cast<double>(acc[v][k]);
      }
      bs_named_barrier(1, n_cons);
    }
    if (active && rl == 0 && n_chunks > 0) {
#pragma unroll
      for (int v = 0; v < 3; ++v)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          double s = static_cast<double>(acc[v][k]);
          for (int l = 1; l < n_lanes; ++l) s += red[(static_cast<size_t>(l) * 3 + v) * a.c_total + c0 + k];
          atomicAdd(a.sums_out + static_cast<int64_t>(v) * a.sums_ld + c0 + k, s);
        }
    }
  }
}

// Geometry shared by the three modes.  False: the layout cannot be staged (the caller keeps the per-thread-load kernels).
inline bool bs_plan(BsArgs* a, int sm_count, dim3* grid, dim3* block) {
  if (a->c_total < 4 || a->c_total % 4 != 0) return false;
  const uint32_t row_bytes = static_cast<uint32_t>(a->c_total) * 4u;
  if (row_bytes > kBsStageBytes) return false;
  if (reinterpret_cast<uintptr_t>(a->z) % 16 != 0 || (a->dy != nullptr && reinterpret_cast<uintptr_t>(a->dy) % 16 != 0)) return false;
  const int nq = a->c_total / 4;
  const int qpad = (nq + 31) / 32 * 32;
  if (qpad > kBsMaxThreads - 32) return false;
  int lanes = (kBsMaxThreads - 32) / qpad;
  a->rows_per_stage = static_cast<int32_t>(kBsStageBytes / row_bytes);
  if (lanes > a->rows_per_stage) lanes = a->rows_per_stage;
  a->rows_per_stage -= a->rows_per_stage % lanes;          // the consumers' row walk continues uniformly across stages
  if (a->rows >= (1LL << 31)) return false;                // 32-bit row arithmetic in the kernel
  a->qpad = qpad;
  int ctas = sm_count;
  if (static_cast<int64_t>(ctas) * a->rows_per_stage > a->rows) ctas = static_cast<int>((a->rows + a->rows_per_stage - 1) / a->rows_per_stage);
  if (ctas < 1) ctas = 1;
  a->span = (a->rows + ctas - 1) / ctas;
  *grid = dim3(static_cast<unsigned>((a->rows + a->span - 1) / a->span));
  *block = dim3(32 + qpad * lanes);
  return true;
}

template <int MODE>
cudaError_t bs_launch(const BsArgs& a, dim3 grid, dim3 block, cudaStream_t s) {
  // function attributes are per device: one ctx per GPU may live in one process (TrainerMGPU)
  static bool attr_done_dev[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done_dev[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(bn_stream_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBsSmemBytes);
    if (e != cudaSuccess) return e;
    attr_done_dev[dev & 63] = true;
  }
  bn_stream_kernel<MODE><<<grid, block, kBsSmemBytes, s>>>(a);
  return cudaGetLastError();
}

}  // namespace
}  // namespace xv
