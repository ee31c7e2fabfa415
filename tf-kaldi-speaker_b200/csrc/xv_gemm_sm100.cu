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
// Warp roles (320 threads, 1 CTA/SM, persistent over tiles):
//   warp 0    : TMA producer (one lane)        -- fills the 4-stage smem ring (A 16 KB + B 32 KB per stage)
//   warp 1    : TMEM allocator + MMA issuer    -- 4 x tcgen05.mma (128x256x16) per stage, commit -> mbarrier
//   warps 2-9 : epilogue, 2 warps per TMEM lane quarter (each takes 128 of the tile's 256 columns):
//               tcgen05.ld 32 lanes x 32 columns -> affine + activation -> bf16 tile staged in 64B-swizzled
//               smem -> TMA store (full-line writes), or fp32 direct store, or pooling partial sums.
// TMEM holds two 128x256 fp32 accumulators so the epilogue of tile i overlaps the MMAs of tile i+1; the
// accumulator is handed back to the MMA warp as soon as the last tcgen05.ld of the tile has completed.
#include <stdio.h>
#include <stdlib.h>

#include <type_traits>

#include "xv_internal.h"
#include "xv_ptx.cuh"

namespace xv {

namespace {

// Epilogue warps: 4 TMEM lane quarters x kParts column parts.  The storing kernel runs 8 (each needs a 2 KB TMA store
// staging buffer and shared memory is full).  The pooled kernel stores nothing and can run 12 or 16
// (-DXV_POOL_EPI_WARPS=...).  Measured on tdnn5: 12 warps = 8 warps (890 TFLOP/s either way), 16 warps slower
// (11.6 -> 14.3 ms: the register cap drops to 96 and every warp repeats the per-tile segment search) -- the pooled
// epilogue is bound by instruction throughput, not by the number of warps hiding latencies.
#ifndef XV_POOL_EPI_WARPS
#define XV_POOL_EPI_WARPS 8
#endif
constexpr int kEpiWarpsStore = 8;
template <bool kPool>
struct EpiCfg {
  static constexpr int kEpiWarps = kPool ? XV_POOL_EPI_WARPS : kEpiWarpsStore;
  static constexpr int kParts = kEpiWarps / 4;
  static constexpr int kChunksPerWarp = (8 + kParts - 1) / kParts;   // 32-column chunks of the 256-column tile per warp (the last part may get fewer)
  static constexpr int kThreads = 64 + kEpiWarps * 32;
};
constexpr uint32_t kABytes = kBM * kBK * 2;                  // 16384
constexpr uint32_t kBBytes = kBN * kBK * 2;                  // 32768
constexpr uint32_t kBHalfBytes = kBLoadRows * kBK * 2;       // 16384: one TMA box of the weight tile (xv_internal.h)
constexpr uint32_t kStageBytes = kABytes + kBBytes;          // 49152
constexpr uint32_t kRingBytes = kStages * kStageBytes;       // 196608
constexpr int kStages2Sm = 6;                                // 2-SM pairs: A 16 KB + this CTA's weight half 16 KB per stage
constexpr uint32_t kStageBytes2Sm = kABytes + kBHalfBytes;   // 32768
static_assert(kStages2Sm * kStageBytes2Sm == kRingBytes, "both ring layouts fill the same shared memory");
constexpr uint32_t kEpiParamBytes = kAccStages * kBN * 16;   // (scale,shift) float2 + slope float + query float = 8192
constexpr uint32_t kStoreStageBytes = 32 * 64;               // per warp: 32 rows x 32 bf16, SWIZZLE_64B
constexpr uint32_t kEpiScratchBytes = kEpiWarpsStore * kStoreStageBytes;
constexpr uint32_t kBarrierBytes = 256;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kSmemBytes = 1024 /*alignment slack (normally unused: the array is declared __align__(1024))*/ + kRingBytes + kEpiParamBytes + kEpiScratchBytes + kBarrierBytes;
static_assert(kSmemBytes <= 232448, "exceeds the 227 KB per-CTA shared memory limit");
static_assert((kRingBytes + kEpiParamBytes) % 1024 == 0, "store staging must stay 1024-byte aligned");

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Pooled epilogue, last step: v[0..7] = this thread's partial column sums (column slot 2i + e = column 8i + 2(lane&3) + e
// of a 32-column chunk), v[8..15] the partial sums of squares; the 8 lanes that share (lane & 3) hold different rows.
// A 14-shuffle butterfly over lane bits 4, 3, 2 halves the live values each step; afterwards every lane owns two adjacent
// columns of either the sum (lane bit 4 = 0) or the sum of squares (bit 4 = 1) and issues two fp64 atomics
// (order-independent at fp32 precision; no cancellation noise in E[x^2] - mean^2).  acc_row = pool_acc + utt*2*pool_ld.
__device__ __forceinline__ void pool_reduce_atomics(float (&v)[16], uint32_t lane, double* acc_row, int pool_ld, int col0) {
  {
    const bool up = (lane & 16) != 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float send = up ? v[j] : v[j + 8];
      const float keep = up ? v[j + 8] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  {
    const bool up = (lane & 8) != 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float send = up ? v[j] : v[j + 4];
      const float keep = up ? v[j + 4] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  {
    const bool up = (lane & 4) != 0;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float send = up ? v[j] : v[j + 2];
      const float keep = up ? v[j + 2] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
  }
  const int kind = (lane >> 4) & 1;                      // 0: sum, 1: sum of squares
  const int ii = ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
  double* dst = acc_row + static_cast<size_t>(kind) * pool_ld + col0 + 8 * ii + (lane & 3) * 2;
  atomicAdd(dst, static_cast<double>(v[0]));
  atomicAdd(dst + 1, static_cast<double>(v[1]));
}

// k2Sm: the CTA pair of a 2-CTA cluster runs ONE 256 x 256 tcgen05.mma.cta_group::2 per K step.  Each CTA stages its own
// 128 activation rows and its own 128-row half of the weight tile (32 KB per k-block and SM instead of 48 KB, six ring
// stages instead of four), the leader (cluster rank 0) issues every MMA, commits multicast to both CTAs' barriers, both
// epilogues read their own TMEM and hand the accumulator back on the LEADER's barrier.  Halves the shared-memory operand
// reads and the L2 -> SM weight traffic per flop, which is what the power-capped tensor pipe is short of.
// kWPool: the pooled epilogue with softmax row weights (attention pooling).  A separate instantiation, so that the code of the
// weighted sums does not enter the register allocation of the unweighted kernel (the headline path's tdnn5) and vice versa.
template <bool kPool, bool k2Sm, bool kWPool = false>
__global__ void __launch_bounds__(EpiCfg<kPool>::kThreads, 1)
xv_tdnn_gemm_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                    const __grid_constant__ CUtensorMap tm_out, const GemmParams p) {
  // 128-byte swizzle atoms need 1024-byte aligned stage buffers.  The pad is added as an OFFSET to the
  // __shared__ symbol (not via an integer round trip) so the compiler keeps the shared address space and emits
  // LDS/STS rather than generic loads in the epilogue.
  constexpr int kEpiWarps = EpiCfg<kPool>::kEpiWarps;
  constexpr int kParts = EpiCfg<kPool>::kParts;
  constexpr int kChunksPerWarp = EpiCfg<kPool>::kChunksPerWarp;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* ring = smem;
  float2* ep_ss = reinterpret_cast<float2*>(smem + kRingBytes);                       // [2][256] scale,shift
  float* ep_sl = reinterpret_cast<float*>(smem + kRingBytes + kAccStages * kBN * 8);  // [2][256] slope
  float* ep_sq = reinterpret_cast<float*>(smem + kRingBytes + kAccStages * kBN * 12); // [2][256] attention query (fused scores)
  uint8_t* scratch = smem + kRingBytes + kEpiParamBytes;                              // per-warp staging / transpose
  uint64_t* bars = reinterpret_cast<uint64_t*>(scratch + kEpiScratchBytes);
  constexpr int kNumStages = k2Sm ? kStages2Sm : kStages;
  constexpr uint32_t kStageB = k2Sm ? kStageBytes2Sm : kStageBytes;
  uint64_t* full_bar = bars;                       // [kNumStages]   TMA -> MMA
  uint64_t* empty_bar = bars + kNumStages;         // [kNumStages]   MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * kNumStages;     // [kAccStages] MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + kAccStages;   // [kAccStages] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  // Thread-block cluster of 1 (plain launch) or 2 CTAs.  A pair works on two row tiles of the SAME column tile: each
  // CTA fetches one 128-row half of every weight tile and multicasts it into both shared memories, so the weights
  // cross the L2 -> SM fabric once per pair (48 -> 32 KB per k-block and SM; the K = 512 layers are bound by that feed).
  const uint32_t crank = ptx::cluster_ctarank();
  const uint32_t csize = ptx::cluster_nctarank();
  const uint16_t cmask = static_cast<uint16_t>((1u << csize) - 1u);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tm_a);
    ptx::prefetch_tensormap(&tm_b);
    if (!kPool && p.out_bf16 != nullptr) ptx::prefetch_tensormap(&tm_out);
    for (int s = 0; s < kNumStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], k2Sm ? 1 : csize);   // multicast pairs: every CTA of the cluster retires its own MMAs on this stage
    }
    for (int s = 0; s < kAccStages; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], k2Sm ? 2 * kEpiWarps : kEpiWarps);   // one arrive per epilogue warp (of both CTAs: 2-SM leader)
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    if (k2Sm) {
      ptx::tmem_alloc_2sm(tmem_slot, kTmemCols);
      ptx::tmem_relinquish_2sm();
    } else {
      ptx::tmem_alloc(tmem_slot, kTmemCols);
      ptx::tmem_relinquish();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (csize > 1) ptx::cluster_sync();               // the peer's barriers are initialised before anything targets them
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work units: (group of csize consecutive row tiles, column tile); CTA `crank` takes row tile group*csize + crank.
  // A row tile past the end (odd tile count) is a phantom: its loads are zero-filled, its epilogue only hands the
  // accumulator back.
  // Column-stationary schedule (pooled layer, p.col_stationary): the grid is cpn x n_tiles CTAs, CTA b keeps column tile
  // b % n_tiles for its whole life and walks a contiguous range of row tiles, so the pooling sums of an utterance stay in
  // registers from tile to tile (see the epilogue); the n_tiles CTAs of one row range run in lock step and share every
  // activation tile through the L2.
  const bool col_stat = kPool && p.col_stationary != 0;
  int unit0 = blockIdx.x / csize;
  int unit_step = gridDim.x / csize;
  int n_units = ((p.m_tiles + static_cast<int>(csize) - 1) / static_cast<int>(csize)) * p.n_tiles;
  int nt_fixed = -1;
  if (col_stat) {                                   // units = groups of csize consecutive row tiles of ONE column tile
    const int cl = blockIdx.x / csize, n_cl = gridDim.x / csize;
    const int cpn = n_cl / p.n_tiles;
    const int j = cl / p.n_tiles;
    const int groups = (p.m_tiles + static_cast<int>(csize) - 1) / static_cast<int>(csize);
    nt_fixed = cl % p.n_tiles;
    unit0 = static_cast<int>(static_cast<int64_t>(j) * groups / cpn);
    n_units = static_cast<int>(static_cast<int64_t>(j + 1) * groups / cpn);
    unit_step = 1;
  }
  // 2-SM pairs: the leader's barriers as shared::cluster addresses
  const uint32_t lead_full0 = k2Sm ? ptx::mapa_u32(ptx::smem_u32(&full_bar[0]), 0) : 0u;
  const uint32_t lead_tempty0 = k2Sm ? ptx::mapa_u32(ptx::smem_u32(&tempty_bar[0]), 0) : 0u;
  // epilogue warp -> MMA warp: this accumulator stage is drained (lane 0 of the warp calls)
  auto release_acc = [&](uint32_t as) {
    if (k2Sm) ptx::mbar_arrive_cluster(lead_tempty0 + as * 8);
    else ptx::mbar_arrive(&tempty_bar[as]);
  };
  const int num_kb = p.taps * p.kb_per_tap;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int u = unit0; u < n_units; u += unit_step) {
        const int m0 = ((col_stat ? u : u / p.n_tiles) * static_cast<int>(csize) + static_cast<int>(crank)) * kBM;
        const int n0 = (col_stat ? nt_fixed : u % p.n_tiles) * kBN;
        for (int tap = 0; tap < p.taps; ++tap) {
          const int row = m0 + tap * p.dilation;
          for (int cb = 0; cb < p.kb_per_tap; ++cb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * kStageB;
            uint8_t* sb = sa + kABytes;
            const int kc = tap * p.cin + cb * kBK;
            if (k2Sm) {
              // both CTAs' boxes complete on the leader's barrier, which expects the bytes of the pair
              if (crank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * kStageBytes2Sm);
              ptx::tma_load_2d_2sm(sa, &tm_a, lead_full0 + stage * 8, cb * kBK, row);
              ptx::tma_load_2d_2sm(sb, &tm_b, lead_full0 + stage * 8, kc, n0 + static_cast<int>(crank) * kBLoadRows);
              if (++stage == kNumStages) { stage = 0; phase ^= 1; }
              continue;
            }
            ptx::mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
            ptx::tma_load_2d(sa, &tm_a, &full_bar[stage], cb * kBK, row);
            if (csize > 1) {
              ptx::tma_load_2d_multicast(sb + crank * kBHalfBytes, &tm_b, &full_bar[stage], kc, n0 + static_cast<int>(crank) * kBLoadRows, cmask);
            } else {
              ptx::tma_load_2d(sb, &tm_b, &full_bar[stage], kc, n0);
              ptx::tma_load_2d(sb + kBHalfBytes, &tm_b, &full_bar[stage], kc, n0 + kBLoadRows);
            }
            if (++stage == kNumStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp walks the loop (barrier waits included) so that every operand is warp-uniform and the compiler
    // issues UTCHMMA straight from uniform registers; one elected lane issues.  A lone lane under `if (lane == 0)`
    // costs ~450 clk per (wait, 4 MMAs, commit) round in its own instruction stream -- 88% of the 512 clk the four
    // MMAs execute in (tools/debug/umma_issue.cu) -- which leaves no slack to absorb a late TMA arrival.
    constexpr uint32_t idesc = ptx::make_idesc_bf16(k2Sm ? 2 * kBM : kBM, kBN);
    const uint32_t ring_lo = ptx::sw128_desc_lo(ptx::smem_u32(ring));
    uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
    if (k2Sm && crank != 0) n_units = unit0;        // the peer's MMA warp only owns the TMEM allocation
    for (int u = unit0; u < n_units; u += unit_step) {
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);   // epilogue has drained this accumulator
      ptx::tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * kBN;
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);     // TMA bytes have landed
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t a_lo = ring_lo + stage * (kStageB >> 4);
          const uint32_t b_lo = a_lo + (kABytes >> 4);
#pragma unroll
          for (int k = 0; k < kBK / kUK; ++k) {
            // advance 16 elements (32 bytes) along K inside the 128-byte swizzle atom: +2 in the address field
            if (k2Sm) ptx::umma_bf16_lo_2sm(tmem_d, a_lo + 2 * k, b_lo + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            else ptx::umma_bf16_lo(tmem_d, a_lo + 2 * k, b_lo + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          // frees the smem slot (in every CTA of the cluster: the peer multicasts into it) when these MMAs retire
          if (k2Sm) ptx::umma_commit_2sm(&empty_bar[stage], 3);
          else if (csize > 1) ptx::umma_commit_multicast(&empty_bar[stage], cmask);
          else ptx::umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == kNumStages) { stage = 0; phase ^= 1; }
      }
      if (ptx::elect_one()) {                                   // accumulator complete -> epilogue (of both CTAs)
        if (k2Sm) ptx::umma_commit_2sm(&tfull_bar[as], 3);
        else ptx::umma_commit(&tfull_bar[as]);
      }
      __syncwarp();
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const uint32_t e = warp - 2;                 // 0..7
    const uint32_t q = warp & 3;                 // TMEM lane quarter this warp may access
    const uint32_t half = e >> 2;                // which column part of the tile's 256 columns
    const uint32_t etid = threadIdx.x - 64;      // 0..255
    uint8_t* stage_buf = scratch + (kPool ? 0 : e) * kStoreStageBytes;         // bf16 store staging (non-pool)
    const bool relu_only = p.relu_only != 0;
    const bool store_bf16 = !kPool && p.out_bf16 != nullptr;
    // plain ReLU layer that only stores bf16 (every TDNN frame layer of the extraction path): BN on packed fp32 pairs and
    // ReLU folded into the bf16 conversion (cvt.rn.relu.bf16x2.f32) -- 1 issue slot per element instead of 2.5.  Measured
    // neutral on the step time (tdnn1 / tdnn4 are not bound by epilogue instructions: with the TMA stores skipped
    // altogether tdnn4 takes the same 3.74 ms -- it waits on its activation loads, L2 hit rate 64 %); kept for the issue
    // slots and the power it saves under the cap.
    const bool fast_store = store_bf16 && relu_only && p.dot_out == nullptr && p.out_f32 == nullptr;
    uint32_t as = 0, aphase = 0;
    // epilogue parameters of one column tile -> shared memory.  Pooled kernel: (scale, shift) are stored per COLUMN PAIR
    // as (s0, s1, h0, h1) so that one LDS.128 yields the two packed-fp32 operands of an FFMA2.
    auto fill_params = [&](float2* ss, float* sl, float* sq, int n0) {
      if (etid < kBN) {
        const float sc = __ldg(p.scale + n0 + etid), sh = __ldg(p.shift + n0 + etid);
        if (kPool || fast_store) {
          float* ssf = reinterpret_cast<float*>(ss);
          ssf[(etid >> 1) * 4 + (etid & 1)] = sc;
          ssf[(etid >> 1) * 4 + 2 + (etid & 1)] = sh;
        } else {
          ss[etid] = make_float2(sc, sh);
        }
        sl[etid] = __ldg(p.slope + n0 + etid);
        if (!kPool && p.dot_out != nullptr) sq[etid] = __ldg(p.dot_q + n0 + etid);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
    };
    if (col_stat) fill_params(ep_ss, ep_sl, ep_sq, nt_fixed * kBN);       // the column tile never changes
    int slot_n0_a = -1, slot_n0_b = -1;          // column tile whose parameters each accumulator stage's slot holds
    bool bar_prev_tile = true;                   // (nothing to protect before the first tile)
    static_assert(kAccStages == 2, "one parameter slot per accumulator stage");
    // pooling sums carried in registers from tile to tile (packed pairs: [chunk][i] = columns 8i + 2(lane&3) + {0,1})
    uint64_t acc_s[kPool ? kChunksPerWarp : 1][4], acc_q[kPool ? kChunksPerWarp : 1][4];
#pragma unroll
    for (int ci = 0; ci < (kPool ? kChunksPerWarp : 1); ++ci)
#pragma unroll
      for (int i = 0; i < 4; ++i) { acc_s[ci][i] = 0ull; acc_q[ci][i] = 0ull; }
    int held_utt = -1, held_n0 = 0, held_tiles = 0;
    PoolSeg held_sg = {0, 0, 0, 0};                 // valid-row range of utterance held_utt
    auto flush_held = [&]() {
      if (kPool && held_utt >= 0) {
        double* acc_row = p.pool_acc + static_cast<size_t>(held_utt) * 2 * p.pool_ld;
#pragma unroll
        for (int ci = 0; ci < kChunksPerWarp; ++ci) {
          const int c = half * kChunksPerWarp + ci;
          if (c < 8) {
            float v[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              ptx::f2_unpack(acc_s[ci][i], v[2 * i], v[2 * i + 1]);
              ptx::f2_unpack(acc_q[ci][i], v[8 + 2 * i], v[8 + 2 * i + 1]);
              acc_s[ci][i] = 0ull;
              acc_q[ci][i] = 0ull;
            }
            pool_reduce_atomics(v, lane, acc_row, p.pool_ld, held_n0 + c * 32);
          }
        }
      }
      held_utt = -1;
      held_tiles = 0;
    };
    for (int u = unit0; u < n_units; u += unit_step) {
      const int mt = (col_stat ? u : u / p.n_tiles) * static_cast<int>(csize) + static_cast<int>(crank);
      const int m0 = mt * kBM;
      const int n0 = (col_stat ? nt_fixed : u % p.n_tiles) * kBN;
      if (mt >= p.m_tiles) {                       // phantom row tile of an odd pair: just recycle the accumulator
        ptx::mbar_wait(&tfull_bar[as], aphase);
        ptx::tc_fence_after();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(as);
        if (++as == kAccStages) { as = 0; aphase ^= 1; }
        bar_prev_tile = false;
        continue;
      }
      // per-tile epilogue parameters -> smem (double buffered by accumulator stage)
      float2* ss = ep_ss + (col_stat ? 0 : as * kBN);
      float* sl = ep_sl + (col_stat ? 0 : as * kBN);
      float* sq = ep_sq + (col_stat ? 0 : as * kBN);
      // The slot of this accumulator stage still holds the parameters of column tile slot_n0_{a,b}: a CTA whose units keep
      // one column tile (n_tiles divides the unit stride, e.g. 2 column tiles over 74 pairs) fills each slot once and
      // never meets the 256-thread barrier again (8 % of the storing epilogue's stall samples).  A slot is rewritten only
      // after a barrier every epilogue warp has passed since its last use: the previous tile's, else one taken here.
      if (!col_stat && (as ? slot_n0_b : slot_n0_a) != n0) {
        if (!bar_prev_tile) asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
        bar_prev_tile = true;
        fill_params(ss, sl, sq, n0);
        if (as) slot_n0_b = n0; else slot_n0_a = n0;
      } else {
        bar_prev_tile = false;
      }
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();

      const int mw = m0 + q * 32;                // first row of this warp
      const int row = mw + lane;
      const bool row_ok = row < p.m_rows;
      // segments (utterance valid-row ranges) that intersect this warp's 32 rows: [w_first, w_last]
      int w_first = 0, w_last = -1;
      bool w_full = false;
      PoolSeg w_sg = {0, 0, 0, 0};
      if (kPool && held_utt >= 0 && held_sg.r0 <= mw && held_sg.r1 >= mw + 32) {
        // still inside the utterance whose sums this warp holds (segments are disjoint): no table lookups
        w_first = w_last = 0;
        w_sg = held_sg;
        w_full = true;
      } else if (kPool) {
        const int seg_lo = __ldg(p.tile_seg_ptr + 2 * mt);       // [lo, hi) utterance range intersecting this tile
        const int seg_hi = __ldg(p.tile_seg_ptr + 2 * mt + 1);
        for (int s = seg_lo; s < seg_hi; ++s) {
          const PoolSeg sg = p.segs[s];
          if (max(sg.r0, mw) < min(sg.r1, mw + 32)) {
            if (w_last < w_first) { w_first = s; w_sg = sg; }
            w_last = s;
          }
        }
        w_full = w_first == w_last && w_sg.r0 <= mw && w_sg.r1 >= mw + 32 && !kWPool;
      }
      // attention pooling, common case (one utterance under this warp's 32 rows, heads at least a tile wide): the masked
      // softmax weights of the thread's four rows for the (at most two) heads of this tile, loaded once per tile
      float tw_a[4] = {0.f, 0.f, 0.f, 0.f}, tw_b[4] = {0.f, 0.f, 0.f, 0.f};
      int tile_bound = 0;
      const bool tile_weights = kPool && kWPool && w_first == w_last && p.head_width >= kBN;
      if (tile_weights) {
        const int ha = n0 / p.head_width;
        tile_bound = (ha + 1) * p.head_width - n0;
        const int lo = max(w_sg.r0, mw) - mw, hi = min(w_sg.r1, mw + 32) - mw;
        const int rq0 = lane >> 2;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int rr = rq0 + 8 * k;
          const bool in = rr >= lo && rr < hi;
          const int r = min(mw + rr, p.m_rows - 1);
          tw_a[k] = (in && ha < p.heads) ? __ldg(p.row_weight + static_cast<size_t>(ha) * p.weight_ld + r) : 0.f;
          tw_b[k] = (in && tile_bound < kBN && ha + 1 < p.heads) ? __ldg(p.row_weight + static_cast<size_t>(ha + 1) * p.weight_ld + r) : 0.f;
        }
      }
      if (kPool) {
        // ---- pooled layer: nothing is stored; per-utterance column sums of y and y^2 ----
        // tcgen05.ld.16x256b gives each thread 4 rows x 8 columns of the warp's 32x32 block (rows rq+{0,8,16,24},
        // columns 8i + 2(lane&3) + {0,1}), so the row reduction is 3 in-thread adds per column followed by a
        // 14-shuffle butterfly over the 8 lanes that share (lane & 3): no shared-memory traffic at all (the smem
        // pipe is what the UMMA operand reads and TMA fills need).  After the butterfly every lane owns two
        // adjacent columns of either the sum (lane bit 4 = 0) or the sum of squares (bit 4 = 1) and issues two fp64
        // atomics (order-independent at fp32 precision; no cancellation noise in E[x^2] - mean^2).
        const int rq = lane >> 2, cq = (lane & 3) * 2;
        const int c_end = min(8, static_cast<int>(half + 1) * kChunksPerWarp);
        {
          // ---- statistics pooling (unweighted), and attention pooling when one utterance lies under the warp's rows.
          // The thread's 4 rows are added into register accumulators that live
          // across tiles (same utterance, same columns); the butterfly and the atomics run once per utterance and warp
          // instead of once per tile.  All math on packed fp32 pairs.  A warp whose 32 rows straddle an utterance edge
          // makes one masked pass per utterance over the same accumulator tile (re-read from TMEM), flushing in between.
          auto accumulate = [&](auto masked_c, auto weighted_c, int lo, int hi, bool release) {
            constexpr bool kMasked = decltype(masked_c)::value;
            constexpr bool kWeighted = decltype(weighted_c)::value;     // attention pooling: sum w*y and sum w*y^2
            uint64_t mk[4];
            if (kMasked) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float m = (rq + 8 * k >= lo && rq + 8 * k < hi) ? 1.f : 0.f;
                mk[k] = ptx::f2_pack(m, m);
              }
            }
            // 16-column half chunks, software-pipelined: the tcgen05.ld of half chunk h+1 is in flight while h is reduced
            uint32_t r0[2][8], r1[2][8];               // [buffer]: rows q*32 + rq + {0, 8} | the same + 16
            const uint32_t tbase = tmem_base + ((q * 32) << 16) + as * kBN + half * kChunksPerWarp * 32;
            const int n_half = 2 * (c_end - static_cast<int>(half) * kChunksPerWarp);
            ptx::tmem_ld_16x256b_x2(tbase, r0[0]);
            ptx::tmem_ld_16x256b_x2(tbase + (16u << 16), r1[0]);
#pragma unroll
            for (int h = 0; h < 2 * kChunksPerWarp; ++h) {
              if (h < n_half) {
                ptx::tmem_ld_wait();
                if (h + 1 < n_half) {
                  ptx::tmem_ld_16x256b_x2(tbase + (h + 1) * 16, r0[(h + 1) & 1]);
                  ptx::tmem_ld_16x256b_x2(tbase + (16u << 16) + (h + 1) * 16, r1[(h + 1) & 1]);
                } else if (release) {
                  ptx::tc_fence_before();              // last TMEM read of the tile: hand the accumulator back
                  __syncwarp();
                  if (lane == 0) release_acc(as);
                }
                const int ci = h >> 1;
                const int col = (half * kChunksPerWarp + ci) * 32 + (h & 1) * 16;
#pragma unroll
                for (int ii = 0; ii < 2; ++ii) {
                  const int i = (h & 1) * 2 + ii;
                  const ulonglong2 sh = *reinterpret_cast<const ulonglong2*>(ss + col + 8 * ii + cq);   // (s0,s1) | (h0,h1)
                  float2 ng = make_float2(0.f, 0.f);
                  if (!relu_only) ng = *reinterpret_cast<const float2*>(sl + col + 8 * ii + cq);
                  // weighted: a tile touches at most two heads (head_width >= the tile width), the second one from
                  // column tile_bound on; only the 16-column half chunk that holds the boundary selects per column
                  const int rel = tile_bound - col;                  // warp-uniform
                  const bool w_mixed = kWeighted && rel > 0 && rel < 16;
                  const bool sec0 = 8 * ii + cq >= rel, sec1 = 8 * ii + cq + 1 >= rel;
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    const uint32_t* src = k < 2 ? r0[h & 1] : r1[h & 1];
                    uint64_t v = ptx::f2_fma(ptx::f2_pack_u32(src[4 * ii + 2 * (k & 1)], src[4 * ii + 2 * (k & 1) + 1]), sh.x, sh.y);
                    float v0, v1;
                    ptx::f2_unpack(v, v0, v1);
                    if (relu_only) {
                      v0 = fmaxf(v0, 0.f);
                      v1 = fmaxf(v1, 0.f);
                    } else {
                      v0 = v0 >= 0.f ? v0 : v0 * ng.x;
                      v1 = v1 >= 0.f ? v1 : v1 * ng.y;
                      if (p.act_tanh) { v0 = tanhf(v0); v1 = tanhf(v1); }
                    }
                    v = ptx::f2_pack(v0, v1);
                    if (kWeighted) {
                      const float wu = rel <= 0 ? tw_b[k] : tw_a[k];          // softmax weights: 0 outside the utterance's valid rows
                      const uint64_t w = w_mixed ? ptx::f2_pack(sec0 ? tw_b[k] : tw_a[k], sec1 ? tw_b[k] : tw_a[k]) : ptx::f2_pack(wu, wu);
                      const uint64_t t = ptx::f2_mul(v, w);
                      acc_s[ci][i] = ptx::f2_add(acc_s[ci][i], t);
                      acc_q[ci][i] = ptx::f2_fma(t, v, acc_q[ci][i]);
                    } else {
                      if (kMasked) v = ptx::f2_mul(v, mk[k]);
                      acc_s[ci][i] = ptx::f2_add(acc_s[ci][i], v);
                      acc_q[ci][i] = ptx::f2_fma(v, v, acc_q[ci][i]);
                    }
                  }
                }
              }
            }
          };
          if (!kWPool || tile_weights) {
          bool released = false;
          for (int s = w_first; s <= w_last; ++s) {
            const PoolSeg sg = s == w_first ? w_sg : p.segs[s];
            const int lo = max(sg.r0, mw) - mw;
            const int hi = min(sg.r1, mw + 32) - mw;
            if (lo >= hi) continue;
            if (held_utt != sg.utt || held_n0 != n0 || held_tiles >= 64) {
              flush_held();
              held_utt = sg.utt;
              held_n0 = n0;
              held_sg = sg;
            }
            ++held_tiles;
            released = s == w_last;
            if (kWPool) accumulate(std::false_type{}, std::true_type{}, lo, hi, released);   // the weights carry the row mask
            else if (lo == 0 && hi == 32) accumulate(std::false_type{}, std::false_type{}, lo, hi, released);
            else accumulate(std::true_type{}, std::false_type{}, lo, hi, released);
          }
          if (!released) {                           // no valid row under this warp: just hand the accumulator back
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) release_acc(as);
          }
          } else {
#pragma unroll 1
        for (int c = half * kChunksPerWarp; c < c_end; ++c) {
          uint32_t ra[16], rb[16];
          const uint32_t tcol = as * kBN + c * 32;
          ptx::tmem_ld_16x256b_x4(tmem_base + ((q * 32) << 16) + tcol, ra);
          ptx::tmem_ld_16x256b_x4(tmem_base + ((q * 32 + 16) << 16) + tcol, rb);
          ptx::tmem_ld_wait();
          if (c == c_end - 1) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) release_acc(as);
          }
          float y[4][8];                           // [row k: rq + 8k][column slot 2i + e]
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 sh = *reinterpret_cast<const float4*>(ss + c * 32 + 8 * i + cq);
            const float2 ng = *reinterpret_cast<const float2*>(sl + c * 32 + 8 * i + cq);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t* src = k < 2 ? ra : rb;
              float v0 = fmaf(__uint_as_float(src[4 * i + 2 * (k & 1)]), sh.x, sh.z);       // pair layout (s0, s1, h0, h1)
              float v1 = fmaf(__uint_as_float(src[4 * i + 2 * (k & 1) + 1]), sh.y, sh.w);
              if (relu_only) {
                v0 = fmaxf(v0, 0.f);
                v1 = fmaxf(v1, 0.f);
              } else {
                v0 = v0 >= 0.f ? v0 : v0 * ng.x;
                v1 = v1 >= 0.f ? v1 : v1 * ng.y;
                if (p.act_tanh) { v0 = tanhf(v0); v1 = tanhf(v1); }
              }
              y[k][2 * i] = v0;
              y[k][2 * i + 1] = v1;
            }
          }
          for (int s = w_first; s <= w_last; ++s) {
            const PoolSeg sg = s == w_first ? w_sg : p.segs[s];
            float v[16];
            {
              const int lo = max(sg.r0, mw) - mw;
              const int hi = min(sg.r1, mw + 32) - mw;
              if (lo >= hi) continue;
              float m[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) m[k] = (rq + 8 * k >= lo && rq + 8 * k < hi) ? 1.f : 0.f;
              // attention pooling: the row mask becomes the softmax weight of (head of the column, row); a 32-column
              // chunk touches at most two heads (head_width >= 32), the second one from column `bound` on
              float mb[4] = {0.f, 0.f, 0.f, 0.f};
              int bound = 32;
              if (tile_weights) {
                bound = tile_bound - c * 32;
#pragma unroll
                for (int k = 0; k < 4; ++k) { m[k] = tw_a[k]; mb[k] = tw_b[k]; }
                if (bound <= 0) {                       // the whole chunk belongs to the second head
#pragma unroll
                  for (int k = 0; k < 4; ++k) m[k] = tw_b[k];
                  bound = 32;
                }
              } else if (kWPool) {
                const int nbase = n0 + c * 32;
                const int ha = nbase / p.head_width;
                bound = (ha + 1) * p.head_width - nbase;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const int r = min(mw + rq + 8 * k, p.m_rows - 1);
                  const float wa = ha < p.heads ? __ldg(p.row_weight + static_cast<size_t>(ha) * p.weight_ld + r) : 0.f;
                  const float wb = (bound < 32 && ha + 1 < p.heads) ? __ldg(p.row_weight + static_cast<size_t>(ha + 1) * p.weight_ld + r) : 0.f;
                  mb[k] = m[k] * wb;
                  m[k] *= wa;
                }
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const bool second = 8 * (j >> 1) + cq + (j & 1) >= bound;
                float a1 = 0.f, a2 = 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float t = (second ? mb[k] : m[k]) * y[k][j];
                  a1 += t;
                  a2 = fmaf(t, y[k][j], a2);
                }
                v[j] = a1;
                v[8 + j] = a2;
              }
            }
            pool_reduce_atomics(v, lane, p.pool_acc + static_cast<size_t>(sg.utt) * 2 * p.pool_ld, p.pool_ld, n0 + c * 32);
          }
        }
          }
        }
      } else if (!store_bf16 && p.out_f32 != nullptr && p.dot_out == nullptr) {
      // ---- fp32-only output (training forward / data gradient, segment layers): tcgen05.ld.16x256b gives each thread
      // 4 rows x 2 adjacent columns per 8-column group, so a quad of lanes writes one full 32-byte sector of a row
      // (st.global.v2: 8 sectors per instruction instead of 32 half-written ones with one row per lane), and the batch-norm
      // column statistics need the 14-shuffle butterfly of the pooled epilogue instead of a 62-shuffle transpose.
      const int rq = lane >> 2, cq = (lane & 3) * 2;
      float mk[4];
      bool rok[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int rr = mw + rq + 8 * k;
        rok[k] = rr < p.m_rows;
        mk[k] = (rok[k] && p.stat_out != nullptr &&
                 static_cast<int32_t>(static_cast<uint32_t>(rr) % static_cast<uint32_t>(p.stat_t)) < p.stat_valid) ? 1.f : 0.f;
      }
#pragma unroll 1
      for (int ci = 0; ci < kChunksPerWarp; ++ci) {
        const int c = half * kChunksPerWarp + ci;
        uint32_t ra[16], rb[16];
        const uint32_t tcol = as * kBN + c * 32;
        ptx::tmem_ld_16x256b_x4(tmem_base + ((q * 32) << 16) + tcol, ra);
        ptx::tmem_ld_16x256b_x4(tmem_base + ((q * 32 + 16) << 16) + tcol, rb);
        ptx::tmem_ld_wait();
        if (ci == kChunksPerWarp - 1) {
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(as);
        }
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int col = c * 32 + 8 * i + cq;                      // column inside the tile
          const float4 sh = *reinterpret_cast<const float4*>(ss + col);      // (s0, h0, s1, h1)
          const float2 ng = *reinterpret_cast<const float2*>(sl + col);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t* src = k < 2 ? ra : rb;
            float v0 = fmaf(__uint_as_float(src[4 * i + 2 * (k & 1)]), sh.x, sh.y);
            float v1 = fmaf(__uint_as_float(src[4 * i + 2 * (k & 1) + 1]), sh.z, sh.w);
            if (relu_only) {
              v0 = fmaxf(v0, 0.f);
              v1 = fmaxf(v1, 0.f);
            } else {
              v0 = v0 >= 0.f ? v0 : v0 * ng.x;
              v1 = v1 >= 0.f ? v1 : v1 * ng.y;
              if (p.act_tanh) { v0 = tanhf(v0); v1 = tanhf(v1); }
            }
            if (rok[k] && n0 + col < p.f32_cols)
              *reinterpret_cast<float2*>(p.out_f32 + static_cast<size_t>(mw + rq + 8 * k) * p.ldo + n0 + col) = make_float2(v0, v1);
            const float m0v = v0 * mk[k], m1v = v1 * mk[k];
            v[2 * i] += m0v;
            v[2 * i + 1] += m1v;
            v[8 + 2 * i] = fmaf(m0v, m0v, v[8 + 2 * i]);
            v[8 + 2 * i + 1] = fmaf(m1v, m1v, v[8 + 2 * i + 1]);
          }
        }
        if (p.stat_out != nullptr) pool_reduce_atomics(v, lane, p.stat_out, p.stat_ld, n0 + c * 32);
      }
      } else {
      // fused attention scores: this warp's 128 columns lie in at most two heads (head_width >= 128)
      float dot_a = 0.f, dot_b = 0.f;
      int dot_ha = 0, dot_bound = 0;
      const bool dots = p.dot_out != nullptr;
      if (dots) {
        const int half_base = n0 + half * (kBN / kParts);
        dot_ha = half_base / p.head_width;
        dot_bound = (dot_ha + 1) * p.head_width - half_base;
      }
#pragma unroll 1
      for (int ci = 0; ci < kChunksPerWarp; ++ci) {
        const int c = half * kChunksPerWarp + ci;   // 32-column chunk of the tile
        uint32_t r[32];
        ptx::tmem_ld_32x32(tmem_base + ((q * 32) << 16) + as * kBN + c * 32, r);
        ptx::tmem_ld_wait();
        if (ci == kChunksPerWarp - 1) {
          // last TMEM read of this tile by this warp: hand the accumulator back to the MMA warp now
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(as);
        }
        if (fast_store) {
          const ulonglong2* sp = reinterpret_cast<const ulonglong2*>(ss + c * 32);      // per column pair: (s0,s1) | (h0,h1)
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const ulonglong2 sh = sp[j];
            w[j] = ptx::f2_relu_bf16x2(ptx::f2_fma(ptx::f2_pack_u32(r[2 * j], r[2 * j + 1]), sh.x, sh.y));
          }
          if (lane == 0) ptx::tma_store_wait_read<0>();        // previous store has finished reading the buffer
          __syncwarp();
          uint8_t* rowp = stage_buf + lane * 64;
          const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
          for (int g = 0; g < 4; ++g)
            *reinterpret_cast<uint4*>(rowp + ((g ^ sw) << 4)) = make_uint4(w[4 * g], w[4 * g + 1], w[4 * g + 2], w[4 * g + 3]);
          ptx::fence_proxy_async_smem();                        // generic-proxy writes -> visible to the TMA engine
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_2d(&tm_out, stage_buf, n0 + c * 32, mw);   // rows >= m_rows are clipped by the tensor map
            ptx::tma_store_commit();
          }
          continue;
        }
        float y[32];
        const float4* ss4 = reinterpret_cast<const float4*>(ss + c * 32);
        if (relu_only) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float4 sh = ss4[j];            // (scale, shift) of two columns in one broadcast LDS.128
            y[2 * j] = fmaxf(fmaf(__uint_as_float(r[2 * j]), sh.x, sh.y), 0.f);
            y[2 * j + 1] = fmaxf(fmaf(__uint_as_float(r[2 * j + 1]), sh.z, sh.w), 0.f);
          }
        } else {
          const float2* sl2 = reinterpret_cast<const float2*>(sl + c * 32);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float4 sh = ss4[j];
            const float2 ng = sl2[j];
            float v0 = fmaf(__uint_as_float(r[2 * j]), sh.x, sh.y);
            float v1 = fmaf(__uint_as_float(r[2 * j + 1]), sh.z, sh.w);
            y[2 * j] = v0 >= 0.f ? v0 : v0 * ng.x;
            y[2 * j + 1] = v1 >= 0.f ? v1 : v1 * ng.y;
          }
          if (p.act_tanh) {
#pragma unroll
            for (int j = 0; j < 32; ++j) y[j] = tanhf(y[j]);
          }
        }
        if (p.stat_out != nullptr) {
          // Batch-norm batch statistics of the training forward pass (tf.layers.batch_normalization(training=True),
          // model/tdnn.py:49-52): column sums of y and y^2 over this warp's VALID rows.  x[0..31] = masked y, x[32..63] =
          // its square; a halving butterfly over the five lane bits (62 shuffles) leaves every lane with two adjacent
          // columns of either the sum (lane bit 4 = 0) or the sum of squares (bit 4 = 1): two fp64 atomics per lane
          // (order-independent at fp32 precision).
          const float mk = (row_ok && static_cast<int32_t>(static_cast<uint32_t>(row) % static_cast<uint32_t>(p.stat_t)) < p.stat_valid) ? 1.f : 0.f;
          float x[64];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            x[j] = y[j] * mk;
            x[32 + j] = x[j] * x[j];
          }
#pragma unroll
          for (int h = 32; h >= 2; h >>= 1) {              // lane bit h/2 ... : masks 16, 8, 4, 2, 1
            const bool up = (lane & (h >> 1)) != 0;
#pragma unroll
            for (int j = 0; j < h; ++j) {
              const float send = up ? x[j] : x[j + h];
              const float keep = up ? x[j + h] : x[j];
              x[j] = keep + __shfl_xor_sync(0xffffffffu, send, h >> 1);
            }
          }
          const int kind = (lane >> 4) & 1;
          const int col = ((lane >> 3) & 1) * 16 + ((lane >> 2) & 1) * 8 + ((lane >> 1) & 1) * 4 + (lane & 1) * 2;
          double* dst = p.stat_out + static_cast<size_t>(kind) * p.stat_ld + n0 + c * 32 + col;
          atomicAdd(dst, static_cast<double>(x[0]));
          atomicAdd(dst + 1, static_cast<double>(x[1]));
        }
        if (dots) {
          const float4* q4 = reinterpret_cast<const float4*>(sq + c * 32);
          const int bnd = dot_bound - ci * 32;            // columns of this chunk from `bnd` on belong to the next head
          float acc_lo = 0.f, acc_hi = 0.f;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const float4 qv = q4[g];
            const float t0 = y[4 * g] * qv.x, t1 = y[4 * g + 1] * qv.y, t2 = y[4 * g + 2] * qv.z, t3 = y[4 * g + 3] * qv.w;
            if (4 * g + 3 < bnd) {
              acc_lo += (t0 + t1) + (t2 + t3);
            } else if (4 * g >= bnd) {
              acc_hi += (t0 + t1) + (t2 + t3);
            } else {
              acc_lo += (4 * g < bnd ? t0 : 0.f) + (4 * g + 1 < bnd ? t1 : 0.f) + (4 * g + 2 < bnd ? t2 : 0.f);
              acc_hi += (4 * g < bnd ? 0.f : t0) + (4 * g + 1 < bnd ? 0.f : t1) + (4 * g + 2 < bnd ? 0.f : t2) + t3;
            }
          }
          dot_a += acc_lo;
          dot_b += acc_hi;
        }
        if (store_bf16) {
          // stage this warp's 32 rows x 32 bf16 in the TMA SWIZZLE_64B layout (16-byte chunk index XOR
          // ((row >> 1) & 3)): conflict-free 16-byte shared stores, then one TMA store writes full lines.
          if (lane == 0) ptx::tma_store_wait_read<0>();        // previous store has finished reading the buffer
          __syncwarp();
          uint8_t* rowp = stage_buf + lane * 64;
          const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 v;
            v.x = pack_bf16(y[g * 8 + 0], y[g * 8 + 1]);
            v.y = pack_bf16(y[g * 8 + 2], y[g * 8 + 3]);
            v.z = pack_bf16(y[g * 8 + 4], y[g * 8 + 5]);
            v.w = pack_bf16(y[g * 8 + 6], y[g * 8 + 7]);
            *reinterpret_cast<uint4*>(rowp + ((g ^ sw) << 4)) = v;
          }
          ptx::fence_proxy_async_smem();                        // generic-proxy writes -> visible to the TMA engine
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_2d(&tm_out, stage_buf, n0 + c * 32, mw);   // rows >= m_rows are clipped by the tensor map
            ptx::tma_store_commit();
          }
        }
        if (!kPool && p.out_f32 != nullptr && row_ok) {
          float4* dst = reinterpret_cast<float4*>(p.out_f32 + static_cast<size_t>(row) * p.ldo + n0 + c * 32);
#pragma unroll
          for (int g = 0; g < 8; ++g)
            if (n0 + c * 32 + g * 4 < p.f32_cols) dst[g] = make_float4(y[g * 4], y[g * 4 + 1], y[g * 4 + 2], y[g * 4 + 3]);
        }
      }
      if (dots && row_ok) {
        if (dot_ha < p.heads) atomicAdd(p.dot_out + static_cast<size_t>(dot_ha) * p.dot_ld + row, static_cast<double>(dot_a * p.dot_scale));
        if (dot_bound < kBN / kParts && dot_ha + 1 < p.heads)
          atomicAdd(p.dot_out + static_cast<size_t>(dot_ha + 1) * p.dot_ld + row, static_cast<double>(dot_b * p.dot_scale));
      }
      }
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
    flush_held();
    if (store_bf16 && lane == 0) ptx::tma_store_wait<0>();      // all bulk stores of this warp have completed
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (csize > 1) ptx::cluster_sync();               // no CTA leaves while the peer may still signal its barriers
  if (warp == 1) {
    ptx::tc_fence_after();
    if (k2Sm) ptx::tmem_dealloc_2sm(tmem_base, kTmemCols);
    else ptx::tmem_dealloc(tmem_base, kTmemCols);
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
  cudaError_t e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(xv_tdnn_gemm_kernel<true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) {
    snprintf(g_err, sizeof(g_err), "cudaFuncSetAttribute(max dynamic smem) failed: %s", cudaGetErrorString(e));
    return g_err;
  }
  return nullptr;
}

const char* make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows,
                              uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_rows, int swizzle_bytes) {
  if (g_encode == nullptr) return "gemm_init() not called";
  cuuint64_t dims[2] = {inner, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
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

const char* make_tmap_bf16_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                              const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
  if (g_encode == nullptr) return "gemm_init() not called";
  cuuint64_t d[5];
  cuuint64_t st[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 1; i < rank; ++i) st[i - 1] = strides_bytes[i];
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), d, st, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled(rank %d) failed (CUresult %d): dims=%llu,%llu,%llu box=%u,%u,%u base=%p",
             rank, static_cast<int>(r), (unsigned long long)dims[0], (unsigned long long)dims[1],
             (unsigned long long)(rank > 2 ? dims[2] : 0), box[0], box[1], rank > 2 ? box[2] : 0, base);
    return g_err;
  }
  return nullptr;
}

namespace {
// Co-resident 2-CTA clusters of one kernel variant (one CTA per SM; a pair shares a TPC), -1 = not asked yet.
template <bool kPool, bool k2Sm>
int max_cluster_pairs() {
  static int cached = -1;
  if (cached >= 0) return cached;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2, 1, 1);
  cfg.blockDim = dim3(EpiCfg<kPool>::kThreads, 1, 1);
  cfg.dynamicSmemBytes = kSmemBytes;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, xv_tdnn_gemm_kernel<kPool, k2Sm>, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
  cached = n;
  if (getenv("XV_GEMM_TRACE")) fprintf(stderr, "xv gemm: %d co-resident 2-CTA clusters (pool=%d, 2sm=%d)\n", n, int(kPool), int(k2Sm));
  return cached;
}

template <bool kPool, bool k2Sm>
cudaError_t launch_kernel(const cudaLaunchConfig_t& cfg, const GemmPlan& plan, const GemmParams& p) {
  if constexpr (kPool) {
    if (p.row_weight != nullptr) return cudaLaunchKernelEx(&cfg, xv_tdnn_gemm_kernel<true, k2Sm, true>, plan.tm_a, plan.tm_b, plan.tm_out, p);
  }
  return cudaLaunchKernelEx(&cfg, xv_tdnn_gemm_kernel<kPool, k2Sm, false>, plan.tm_a, plan.tm_b, plan.tm_out, p);
}

template <bool kPool>
cudaError_t launch_variant(const GemmPlan& plan, int sm_count, cudaStream_t stream) {
  GemmParams p = plan.p;
  // XV_GEMM_2SM=0: no cta_group::2 pairs (multicast pairs instead); XV_GEMM_CLUSTER=0: no multicast pairs either (A/B switches)
  static const int want_2sm = getenv("XV_GEMM_2SM") ? atoi(getenv("XV_GEMM_2SM")) : 1;
  static const int want = getenv("XV_GEMM_CLUSTER") ? atoi(getenv("XV_GEMM_CLUSTER")) : 1;
  static const int want_cs = getenv("XV_POOL_COLSTAT") ? atoi(getenv("XV_POOL_COLSTAT")) : 1;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(EpiCfg<kPool>::kThreads, 1, 1);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  // pooled layer: column-stationary schedule, floor(clusters / n_tiles) clusters per
  // column tile (144 of 148 SMs at 6 column tiles; the register-resident sums save more than the four idle SMs cost)
  // (attention pooling too since its sums live in registers across tiles; XV_POOL_COLSTAT=2 keeps it on the row-major schedule)
  const bool col_stat = kPool && want_cs && (p.row_weight == nullptr || want_cs != 2);
  const int groups2 = (p.m_tiles + 1) / 2;         // pairs of row tiles
  // cta_group::2 pairs for every layer with at least two row tiles.  Measured on one box, pairs vs multicast pairs:
  // tdnn2 K=2560 1341 -> 1465, tdnn3 K=3584 1393 -> 1508, pooled tdnn5 K=512 1074 -> 1216, tdnn4 K=512 966 -> 1084,
  // tdnn1 K=192 363 -> 381 TFLOP/s, attention key layer 0 5.82 -> 5.46 ms, weighted pooled layer 10.65 -> 9.67 ms.
  // (The short-K layers LOST while the epilogues handed the accumulator back with a cluster-scope RELEASE arrive: its
  // memory barrier sat on the MMA warp's critical path -- see ptx::mbar_arrive_cluster.)
  if (want_2sm && p.m_tiles >= 2) {
    const int pairs = max_cluster_pairs<kPool, true>();
    if (2 * pairs >= sm_count - 4) {
      int n = groups2 * p.n_tiles < pairs ? groups2 * p.n_tiles : pairs;
      if (col_stat && p.n_tiles <= pairs) {
        int cpn = pairs / p.n_tiles;
        if (cpn > groups2) cpn = groups2;
        n = cpn * p.n_tiles;
        p.col_stationary = 1;
      }
      cfg.gridDim = dim3(2 * n, 1, 1);
      cfg.attrs = &attr;
      cfg.numAttrs = 1;
      return launch_kernel<kPool, true>(cfg, plan, p);
    }
  }
  if (col_stat && p.n_tiles <= sm_count) {
    int cpn = sm_count / p.n_tiles;
    if (cpn > p.m_tiles) cpn = p.m_tiles;
    p.col_stationary = 1;
    cfg.gridDim = dim3(cpn * p.n_tiles, 1, 1);
    return launch_kernel<kPool, false>(cfg, plan, p);
  }
  // multicast pairs (1-SM MMAs, weight tile halves multicast) when the pairs cover (nearly) the whole chip
  // (the pooled kernel gained nothing from them: 880 vs 890 TFLOP/s)
  int pairs = 0;
  if (want && p.m_tiles >= 2 && (!kPool || want > 1)) {
    pairs = max_cluster_pairs<kPool, false>();
    if (2 * pairs < sm_count - 4) pairs = 0;
  }
  if (pairs > 0) {
    const int units = groups2 * p.n_tiles;
    const int n = units < pairs ? units : pairs;
    cfg.gridDim = dim3(2 * n, 1, 1);
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
  } else {
    const int tiles = p.m_tiles * p.n_tiles;
    cfg.gridDim = dim3(tiles < sm_count ? tiles : sm_count, 1, 1);
  }
  return launch_kernel<kPool, false>(cfg, plan, p);
}
}  // namespace

cudaError_t launch_gemm(const GemmPlan& plan, int sm_count, cudaStream_t stream) {
  if (plan.p.m_tiles * plan.p.n_tiles <= 0) return cudaSuccess;
  return plan.pool ? launch_variant<true>(plan, sm_count, stream) : launch_variant<false>(plan, sm_count, stream);
}

}  // namespace xv
