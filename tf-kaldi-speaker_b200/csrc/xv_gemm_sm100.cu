// TDNN layer as ONE kernel on sm_100a: frame splicing folded into TMA operand loads, bf16 tcgen05.mma
// with fp32 accumulators in TMEM, and an epilogue that fuses bias + batch-norm (folded affine) + ReLU /
// leaky / PReLU (+ tanh) and, for the pre-pooling layer, the per-utterance running sum and
// sum-of-squares of statistics pooling.
//
// Replaces, per layer, tf.layers.conv2d(1xk)/tf.layers.dense -> tf.layers.batch_normalization ->
// relu of the reference (model/tdnn.py:42-130) and, when pooling is fused, the reduce_mean /
// squared_difference passes of statistics_pooling (model/pooling.py:40-43).
//
// Math:  out[m, n] = act( scale[n] * sum_{j<taps} sum_{c<cin} A[m + j*dil, c] * Wt[n, j*cin + c] + shift[n] )
//
// A is the PACKED activation matrix [rows, cin] (all utterances back to back).  The "im2col" row of
// output frame m is the contiguous slab A[m .. m+taps-1, :], so K-block (tap j, channel block cb) of the
// A tile is simply the TMA box at coordinates (cb*64, m0 + j*dil): no im2col buffer exists.
// Rows past the end of the buffer are zero-filled by TMA.  Rows whose window straddles an utterance
// boundary are computed but never consumed by valid rows (every layer is VALID) nor by pooling.
//
// Warp roles (192 threads, 1 CTA/SM, persistent over tiles):
//   warp 0 : TMA producer (one lane)          -- fills the 4-stage smem ring (A 16 KB + B 32 KB per stage)
//   warp 1 : TMEM allocator + MMA issuer      -- 4 x tcgen05.mma (128x256x16) per stage, commit -> mbarrier
//   warps 2-5 : epilogue                      -- tcgen05.ld 32 lanes x 32 columns, affine+act, store/pool
// TMEM holds two 128x256 fp32 accumulators so the epilogue of tile i overlaps the MMAs of tile i+1.
#include <stdio.h>

#include "xv_internal.h"
#include "xv_ptx.cuh"

namespace xv {

namespace {

constexpr uint32_t kABytes = kBM * kBK * 2;                  // 16384
constexpr uint32_t kBBytes = kBN * kBK * 2;                  // 32768
constexpr uint32_t kStageBytes = kABytes + kBBytes;          // 49152
constexpr uint32_t kRingBytes = kStages * kStageBytes;       // 196608
constexpr uint32_t kEpiParamBytes = kAccStages * kBN * 12;   // (scale,shift) float2 + slope float
constexpr uint32_t kTransposeBytes = 4 * 32 * 33 * 4;        // per-epilogue-warp 32x33 fp32 (pooling only)
constexpr uint32_t kBarrierBytes = 256;
constexpr uint32_t kTmemCols = 512;

__host__ __device__ constexpr uint32_t gemm_smem_bytes(bool pool) {
  return 1024 /*alignment slack*/ + kRingBytes + kEpiParamBytes + (pool ? kTransposeBytes : 0) + kBarrierBytes;
}

template <bool kPool>
__global__ void __launch_bounds__(kGemmThreads, 1)
xv_tdnn_gemm_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                    const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 128-byte swizzle atoms need 1024-byte aligned stage buffers.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = smem;
  float2* ep_ss = reinterpret_cast<float2*>(smem + kRingBytes);                       // [2][256] scale,shift
  float* ep_sl = reinterpret_cast<float*>(smem + kRingBytes + kAccStages * kBN * 8);  // [2][256] slope
  float* tr_base = reinterpret_cast<float*>(smem + kRingBytes + kEpiParamBytes);      // [4][32*33]
  uint64_t* bars =
      reinterpret_cast<uint64_t*>(smem + kRingBytes + kEpiParamBytes + (kPool ? kTransposeBytes : 0));
  uint64_t* full_bar = bars;                       // [kStages]   TMA -> MMA
  uint64_t* empty_bar = bars + kStages;            // [kStages]   MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * kStages;        // [kAccStages] MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + kAccStages;   // [kAccStages] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tm_a);
    ptx::prefetch_tensormap(&tm_b);
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < kAccStages; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], 4);   // one arrive per epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = p.m_tiles * p.n_tiles;
  const int num_kb = p.taps * p.kb_per_tap;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int m0 = (t / p.n_tiles) * kBM;
        const int n0 = (t % p.n_tiles) * kBN;
        for (int tap = 0; tap < p.taps; ++tap) {
          const int row = m0 + tap * p.dilation;
          for (int cb = 0; cb < p.kb_per_tap; ++cb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * kStageBytes;
            uint8_t* sb = sa + kABytes;
            ptx::mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
            ptx::tma_load_2d(sa, &tm_a, &full_bar[stage], cb * kBK, row);
            ptx::tma_load_2d(sb, &tm_b, &full_bar[stage], tap * p.cin + cb * kBK, n0);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(kBM, kBN);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);   // epilogue has drained this accumulator
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * kBN;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);     // TMA bytes have landed
          ptx::tc_fence_after();
          const uint32_t a_addr = ptx::smem_u32(ring + stage * kStageBytes);
          const uint64_t adesc = ptx::make_kmajor_sw128_desc(a_addr);
          const uint64_t bdesc = ptx::make_kmajor_sw128_desc(a_addr + kABytes);
#pragma unroll
          for (int k = 0; k < kBK / kUK; ++k) {
            // advance 16 elements (32 bytes) along K inside the 128-byte swizzle atom: +2 in the address field
            ptx::umma_bf16(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(&empty_bar[stage]);          // frees the smem slot when these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        ptx::umma_commit(&tfull_bar[as]);               // accumulator complete -> epilogue
        if (++as == kAccStages) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const uint32_t q = warp & 3;                 // TMEM lane quarter this warp may access
    const uint32_t etid = threadIdx.x - 64;      // 0..127
    float* tr = tr_base + q * (32 * 33);
    uint32_t as = 0, aphase = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int mt = t / p.n_tiles;
      const int m0 = mt * kBM;
      const int n0 = (t % p.n_tiles) * kBN;
      // per-tile epilogue parameters -> smem (double buffered by accumulator stage)
      float2* ss = ep_ss + as * kBN;
      float* sl = ep_sl + as * kBN;
      for (int i = etid; i < kBN; i += 128) {
        ss[i] = make_float2(__ldg(p.scale + n0 + i), __ldg(p.shift + n0 + i));
        sl[i] = __ldg(p.slope + n0 + i);
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();

      const int mw = m0 + q * 32;                // first row of this warp
      const int row = mw + lane;
      const bool row_ok = row < p.m_rows;
      int seg_lo = 0, seg_hi = 0;
      if (kPool) {
        seg_lo = __ldg(p.tile_seg_ptr + 2 * mt);       // [lo, hi) utterance range intersecting this tile
        seg_hi = __ldg(p.tile_seg_ptr + 2 * mt + 1);
      }
#pragma unroll 1
      for (int c = 0; c < kBN / 32; ++c) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(tmem_base + ((q * 32) << 16) + as * kBN + c * 32, r);
        ptx::tmem_ld_wait();
        float y[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float2 sh = ss[c * 32 + j];
          float v = fmaf(__uint_as_float(r[j]), sh.x, sh.y);
          v = v >= 0.f ? v : v * sl[c * 32 + j];
          y[j] = v;
        }
        if (p.act_tanh) {
#pragma unroll
          for (int j = 0; j < 32; ++j) y[j] = tanhf(y[j]);
        }
        if (p.out_bf16 != nullptr && row_ok) {
          uint4* dst = reinterpret_cast<uint4*>(p.out_bf16 + static_cast<size_t>(row) * p.ldo + n0 + c * 32);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            __nv_bfloat162 b0 = __floats2bfloat162_rn(y[g * 8 + 0], y[g * 8 + 1]);
            __nv_bfloat162 b1 = __floats2bfloat162_rn(y[g * 8 + 2], y[g * 8 + 3]);
            __nv_bfloat162 b2 = __floats2bfloat162_rn(y[g * 8 + 4], y[g * 8 + 5]);
            __nv_bfloat162 b3 = __floats2bfloat162_rn(y[g * 8 + 6], y[g * 8 + 7]);
            uint4 v;
            v.x = *reinterpret_cast<uint32_t*>(&b0);
            v.y = *reinterpret_cast<uint32_t*>(&b1);
            v.z = *reinterpret_cast<uint32_t*>(&b2);
            v.w = *reinterpret_cast<uint32_t*>(&b3);
            dst[g] = v;
          }
        }
        if (p.out_f32 != nullptr && row_ok) {
          float4* dst = reinterpret_cast<float4*>(p.out_f32 + static_cast<size_t>(row) * p.ldo + n0 + c * 32);
#pragma unroll
          for (int g = 0; g < 8; ++g) dst[g] = make_float4(y[g * 4], y[g * 4 + 1], y[g * 4 + 2], y[g * 4 + 3]);
        }
        if (kPool) {
          // Transpose the warp's 32 rows x 32 columns through padded smem, then lane = column sums the
          // valid rows of every utterance segment that intersects this warp's rows.
          const int col = n0 + c * 32 + lane;
          const int head = p.row_weight != nullptr ? col / p.head_width : 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) tr[lane * 33 + j] = y[j];
          __syncwarp();
          for (int s = seg_lo; s < seg_hi; ++s) {
            const PoolSeg sg = p.segs[s];
            const int lo = max(sg.r0, mw) - mw;
            const int hi = min(sg.r1, mw + 32) - mw;
            if (lo < hi) {
              float s1 = 0.f, s2 = 0.f;
              if (p.row_weight == nullptr) {
                for (int rr = lo; rr < hi; ++rr) {
                  const float v = tr[rr * 33 + lane];
                  s1 += v;
                  s2 = fmaf(v, v, s2);
                }
              } else {
                const float* wrow = p.row_weight + static_cast<size_t>(min(head, p.heads - 1)) * p.m_rows + mw;
                for (int rr = lo; rr < hi; ++rr) {
                  const float v = tr[rr * 33 + lane];
                  const float w = __ldg(wrow + rr);
                  s1 = fmaf(w, v, s1);
                  s2 = fmaf(w * v, v, s2);
                }
              }
              float* acc = p.pool_acc + static_cast<size_t>(sg.utt) * 2 * p.pool_ld + col;
              atomicAdd(acc, s1);
              atomicAdd(acc + p.pool_ld, s2);
            }
          }
          __syncwarp();
        }
      }
      // release the accumulator stage back to the MMA warp
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
char g_err[256];

}  // namespace

const char* gemm_init() {
  if (g_encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
      snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled not available from the driver: %s",
               cudaGetErrorString(e));
      return g_err;
    }
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  cudaError_t e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       gemm_smem_bytes(false));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             gemm_smem_bytes(true));
  if (e != cudaSuccess) {
    snprintf(g_err, sizeof(g_err), "cudaFuncSetAttribute(max dynamic smem) failed: %s", cudaGetErrorString(e));
    return g_err;
  }
  return nullptr;
}

const char* make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows,
                              uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_rows) {
  if (g_encode == nullptr) return "gemm_init() not called";
  cuuint64_t dims[2] = {inner, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_err, sizeof(g_err),
             "cuTensorMapEncodeTiled failed (CUresult %d): inner=%llu rows=%llu stride=%llu box=%ux%u base=%p",
             static_cast<int>(r), (unsigned long long)inner, (unsigned long long)rows,
             (unsigned long long)row_stride_bytes, box_inner, box_rows, base);
    return g_err;
  }
  return nullptr;
}

cudaError_t launch_gemm(const GemmPlan& plan, int sm_count, cudaStream_t stream) {
  const int tiles = plan.p.m_tiles * plan.p.n_tiles;
  if (tiles <= 0) return cudaSuccess;
  const int grid = tiles < sm_count ? tiles : sm_count;
  if (plan.pool) {
    xv_tdnn_gemm_kernel<true><<<grid, kGemmThreads, gemm_smem_bytes(true), stream>>>(plan.tm_a, plan.tm_b, plan.p);
  } else {
    xv_tdnn_gemm_kernel<false><<<grid, kGemmThreads, gemm_smem_bytes(false), stream>>>(plan.tm_a, plan.tm_b, plan.p);
  }
  return cudaGetLastError();
}

}  // namespace xv
