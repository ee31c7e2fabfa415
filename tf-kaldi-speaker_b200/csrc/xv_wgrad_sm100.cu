// Weight gradient of a spliced (1 x k) TDNN layer on the sm_100a tensor cores:
//
//     dW[j*cin + c, n] = sum_{m < rows} A[m + j, c] * dZ[m, n]          (rows m + j >= rows contribute 0)
//
// = the backward of tf.layers.conv2d(1xk) / tf.layers.dense w.r.t. its kernel (what tf.gradients emits for
// model/tdnn.py:42-47,99-103 in the training graph of model/trainer.py:501-527).
//
// The contraction runs over the packed frame index m, which is the OUTER (row) index of both row-major operands
// A [rows, cin] and dZ [rows, cout]: both are "MN-major" UMMA operands.  A K-block of 64 frames is fetched as
// TMA boxes {64 channels, 64 rows} (128-byte swizzle): 2 boxes for the 128 channels of the M tile, 4 boxes for the
// 256 columns of the N tile; the shared-memory descriptors use the MN-major canonical layout (8 rows of 128 bytes
// per swizzle atom, SBO = 1024 bytes between 8-frame groups, LBO = 8192 bytes between 64-wide channel groups)
// and the instruction descriptor sets a_major = b_major = MN.  Tap j of the kernel simply shifts the A row
// coordinate by j.  The frame dimension is split over CTAs (split-K) and the fp32 partial tiles are added to dW
// with vector reductions (red.global.add.v4.f32), so the output must be zeroed first.
#include <stdio.h>

#include "xv_internal.h"
#include "xv_ptx.cuh"

namespace xv {

namespace {

constexpr int kWgThreads = 192;                        // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
constexpr uint32_t kWgABytes = 128 * 64 * 2;           // 16384: 2 boxes of {64 ch, 64 rows}
constexpr uint32_t kWgBBytes = 256 * 64 * 2;           // 32768: 4 boxes
constexpr uint32_t kWgStage = kWgABytes + kWgBBytes;
constexpr uint32_t kWgSmem = 1024 + kStages * kWgStage + 256;

// MN-major operand, 128-byte swizzle: start address, LBO (between 64-element MN groups), SBO (between 8-row K groups)
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
  d |= static_cast<uint64_t>(sbo_bytes >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(kWgThreads, 1)
xv_wgrad_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, const WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* ring = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kWgStage);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tfull_bar = bars + 2 * kStages;
  uint64_t* tempty_bar = tfull_bar + kAccStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tm_a);
    ptx::prefetch_tensormap(&tm_b);
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < kAccStages; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], 4);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int mn_tiles = p.m_tiles * p.n_tiles;
  const int total = mn_tiles * p.splits;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int split = t / mn_tiles;
        const int rem = t - split * mn_tiles;
        const int mt = rem / p.n_tiles, nt = rem - mt * p.n_tiles;
        const int tap = mt / p.m_tiles_per_tap;
        const int c0 = (mt - tap * p.m_tiles_per_tap) * 128;
        const int n0 = nt * 256;
        const int64_t r_begin = static_cast<int64_t>(split) * p.rows_per_split;
        const int64_t r_end = min(r_begin + p.rows_per_split, p.rows);
        for (int64_t r = r_begin; r < r_end; r += 64) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = ring + stage * kWgStage;
          uint8_t* sb = sa + kWgABytes;
          ptx::mbar_arrive_expect_tx(&full_bar[stage], kWgStage);
          ptx::tma_load_2d(sa, &tm_a, &full_bar[stage], c0, static_cast<int32_t>(r) + tap);
          ptx::tma_load_2d(sa + 8192, &tm_a, &full_bar[stage], c0 + 64, static_cast<int32_t>(r) + tap);
#pragma unroll
          for (int i = 0; i < 4; ++i) ptx::tma_load_2d(sb + i * 8192, &tm_b, &full_bar[stage], n0 + 64 * i, static_cast<int32_t>(r));
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // warp-uniform issue loop, one elected lane issues (see xv_gemm_sm100.cu)
    constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 256) | (1u << 15) | (1u << 16);   // both operands MN-major
    uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int split = t / mn_tiles;
      const int64_t r_begin = static_cast<int64_t>(split) * p.rows_per_split;
      const int64_t r_end = min(r_begin + p.rows_per_split, p.rows);
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * 256;
      uint32_t first = 1;
      for (int64_t r = r_begin; r < r_end; r += 64) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t a_addr = ptx::smem_u32(ring + stage * kWgStage);
          const uint32_t b_addr = a_addr + kWgABytes;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // 16 frames per instruction = two 8-row swizzle atoms = 2048 bytes further into every 64-wide group
            const uint64_t adesc = make_mnmajor_sw128_desc(a_addr + k * 2048, 8192, 1024);
            const uint64_t bdesc = make_mnmajor_sw128_desc(b_addr + k * 2048, 8192, 1024);
            ptx::umma_bf16(tmem_d, adesc, bdesc, idesc, (first && k == 0) ? 0u : 1u);
          }
          ptx::umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        first = 0;
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      if (ptx::elect_one()) ptx::umma_commit(&tfull_bar[as]);
      __syncwarp();
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
  } else {
    const uint32_t q = warp & 3;                     // TMEM lane quarter of this warp
    uint32_t as = 0, aphase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int rem = t % mn_tiles;
      const int mt = rem / p.n_tiles, nt = rem - mt * p.n_tiles;
      const int tap = mt / p.m_tiles_per_tap;
      const int c = (mt - tap * p.m_tiles_per_tap) * 128 + q * 32 + lane;     // channel of this thread's accumulator row
      const int n0 = nt * 256;
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const bool row_ok = c < p.cin;
      float* dst_row = p.out + (static_cast<int64_t>(tap) * p.cin + c) * p.ldo + n0;
#pragma unroll 1
      for (int ch = 0; ch < 8; ++ch) {
        uint32_t acc[32];
        ptx::tmem_ld_32x32(tmem_base + ((q * 32) << 16) + as * 256 + ch * 32, acc);
        ptx::tmem_ld_wait();
        if (ch == 7) {
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
        }
        if (row_ok) {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int n = n0 + ch * 32 + g * 4;
            if (n + 3 < p.n_valid) {
              red_add_v4(dst_row + ch * 32 + g * 4, __uint_as_float(acc[g * 4]), __uint_as_float(acc[g * 4 + 1]),
                         __uint_as_float(acc[g * 4 + 2]), __uint_as_float(acc[g * 4 + 3]));
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (n + e < p.n_valid) atomicAdd(dst_row + ch * 32 + g * 4 + e, __uint_as_float(acc[g * 4 + e]));
            }
          }
        }
      }
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

const char* wgrad_init() {
  static char err[160];
  cudaError_t e = cudaFuncSetAttribute(xv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmem);
  if (e != cudaSuccess) {
    snprintf(err, sizeof(err), "cudaFuncSetAttribute(wgrad smem) failed: %s", cudaGetErrorString(e));
    return err;
  }
  return nullptr;
}

// a [rows, lda] bf16, dz [rows, ldz] bf16 (device), out fp32 [taps*cin, ldo] (zeroed by the caller), n_valid <= ldo.
const char* launch_wgrad(const void* a, int64_t a_rows, int lda, const void* dz, int ldz, int64_t rows, int taps, int cin,
                         int n_valid, float* out, int ldo, int sm_count, cudaStream_t stream, cudaError_t* cuda_err) {
  *cuda_err = cudaSuccess;
  if (rows <= 0 || cin <= 0 || n_valid <= 0) return nullptr;
  CUtensorMap tm_a, tm_b;
  const char* e = make_tmap_bf16_2d(&tm_a, a, lda, a_rows, static_cast<uint64_t>(lda) * 2, 64, 64);
  if (!e) e = make_tmap_bf16_2d(&tm_b, dz, ldz, rows, static_cast<uint64_t>(ldz) * 2, 64, 64);
  if (e) return e;
  WgradParams p;
  p.rows = rows; p.taps = taps; p.cin = cin; p.n_valid = n_valid; p.out = out; p.ldo = ldo;
  p.m_tiles_per_tap = (cin + 127) / 128;
  p.m_tiles = taps * p.m_tiles_per_tap;
  p.n_tiles = (n_valid + 255) / 256;
  const int mn = p.m_tiles * p.n_tiles;
  const int64_t kblocks = (rows + 63) / 64;
  // Split of the contraction (frame) axis: the persistent grid works through mn * splits units of kb k-blocks each in
  // ceil(units / SMs) rounds, and every unit ends with a 128 x 256 fp32 reduction into `out` (~6 k-blocks' worth of time).
  // Taking ~2 units per SM regardless of how they divide into rounds left the last round mostly empty (tdnn3 of the training
  // step: 336 units = 3 rounds of 34 k-blocks where 280 units = 2 rounds of 40 do; tdnn5: 312 units = 3 rounds where 144 fit
  // one): pick the split with the shortest schedule.
  int64_t splits = 1;
  {
    const int64_t max_splits = kblocks / 4 > 1 ? kblocks / 4 : 1;       // at least 4 k-blocks per unit
    int64_t best_cost = -1;
    for (int64_t sp = 1; sp <= max_splits; ++sp) {
      const int64_t kb = (kblocks + sp - 1) / sp;
      const int64_t sp_eff = (kblocks + kb - 1) / kb;
      const int64_t rounds = (mn * sp_eff + sm_count - 1) / sm_count;
      const int64_t cost = rounds * (kb + 6);
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; splits = sp; }
    }
  }
  const int64_t kb_per_split = (kblocks + splits - 1) / splits;
  p.rows_per_split = kb_per_split * 64;
  p.splits = static_cast<int32_t>((kblocks + kb_per_split - 1) / kb_per_split);
  const int total = mn * p.splits;
  const int grid = total < sm_count ? total : sm_count;
  xv_wgrad_kernel<<<grid, kWgThreads, kWgSmem, stream>>>(tm_a, tm_b, p);
  *cuda_err = cudaGetLastError();
  return nullptr;
}

}  // namespace xv

#include "xv_ctx.h"

using namespace xv;

extern "C" int xv_wgrad_bf16_async(xv_ctx* c, const void* a_dev, int64_t a_rows, int32_t lda, const void* dz_dev, int32_t ldz,
                                   int64_t rows, int32_t taps, int32_t cin, int32_t n_valid, float* out_dev, int32_t ldo) {
  if (!c) return XV_EINVAL;
  if (!a_dev || !dz_dev || !out_dev || rows <= 0 || a_rows < rows || taps <= 0 || cin <= 0 || cin > lda || n_valid <= 0 ||
      n_valid > ldz || n_valid > ldo || lda % 8 != 0 || ldz % 8 != 0 || ldo % 4 != 0)
    return fail(c, XV_EINVAL, "xv_wgrad_bf16_async: bad argument (lda, ldz multiples of 8; ldo multiple of 4)");
  cudaSetDevice(c->device);
  const char* e = wgrad_init();
  if (e) return fail(c, XV_ECUDA, "%s", e);
  cudaError_t ce = cudaMemsetAsync(out_dev, 0, static_cast<size_t>(taps) * cin * ldo * sizeof(float), c->stream);
  if (ce != cudaSuccess) return fail(c, XV_ECUDA, "xv_wgrad_bf16_async: memset: %s", cudaGetErrorString(ce));
  e = launch_wgrad(a_dev, a_rows, lda, dz_dev, ldz, rows, taps, cin, n_valid, out_dev, ldo, c->sm_count, c->stream, &ce);
  if (e) return fail(c, XV_ECUDA, "%s", e);
  if (ce != cudaSuccess) return fail(c, XV_ECUDA, "xv_wgrad_bf16_async: launch: %s", cudaGetErrorString(ce));
  return XV_OK;
}
