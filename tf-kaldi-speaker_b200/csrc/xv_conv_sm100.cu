// 2-D convolution of the ResNet-18 front end as an implicit GEMM on sm_100a: the im2col operand never exists.
//
// Replaces tf.layers.conv2d(padding='same', use_bias=False) -> tf.layers.batch_normalization [-> + shortcut]
// [-> relu] of conv_block / identity_block (model/resnet.py:9-149).
//
// Activations are NHWC bf16 [B, T, F, C] (H = time, W = frequency).  An output tile is a box of tb time steps x
// fb frequency bins of ONE utterance (fb*tb <= 128 rows of the UMMA M dimension); for kernel tap (dt, df) and
// channel block cb the A operand of that tile is the same box shifted by the tap, which is exactly one TMA box
// of a 5-D tensor map over the input viewed as [B][T/st][st][F/sf][sf*C] (st, sf = strides; the parity split
// turns a strided read into a dense box).  Coordinates that fall outside the utterance (TensorFlow 'SAME'
// padding, asymmetric for stride 2) are zero-filled by the TMA unit, so padding costs nothing and utterances
// never see each other's frames.  B = bf16 weights [Cout][kh*kw*Cin] (K-major), accumulators in TMEM.
//
// Warp roles as in xv_gemm_sm100.cu: warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue
// (tcgen05.ld -> folded BN affine -> + residual -> activation -> bf16).  The four warps that share a column
// half stage their 128 rows x 32 columns in one SWIZZLE_64B buffer and ONE 4-D TMA store writes the
// (32 ch, fb, tb) box; rows beyond the utterance end are clipped by the tensor map.
#include <stdio.h>
#include <stdlib.h>

#include "xv_internal.h"
#include "xv_ptx.cuh"

namespace xv {

namespace {

constexpr int kConvEpiWarps = 8;
constexpr uint32_t kConvABytes = kBM * kBK * 2;   // 16384

template <int BN>
struct ConvSmem {
  static constexpr uint32_t kBBytes = BN * kBK * 2;
  static constexpr uint32_t kStage = kConvABytes + kBBytes;
  static constexpr uint32_t kRing = kStages * kStage;
  static constexpr uint32_t kParams = kAccStages * BN * 12;
  static constexpr uint32_t kStaging = 2 * 128 * 64;          // one 128-row x 32-column bf16 buffer per column half
  // cta_group::2 pairs: each CTA stages its 128 rows and ITS half of the weight tile, six stages
  static constexpr int kStages2 = 6;
  static constexpr uint32_t kStage2 = kConvABytes + kBBytes / 2;
  static constexpr uint32_t kRing2 = kStages2 * kStage2;
  static constexpr uint32_t kRingMax = kRing > kRing2 ? kRing : kRing2;
  static constexpr uint32_t kTotal = 1024 + kRingMax + kParams + kStaging + 256;
};
static_assert(ConvSmem<64>::kStage2 % 1024 == 0 && ConvSmem<128>::kStage2 % 1024 == 0 && ConvSmem<256>::kStage2 % 1024 == 0,
              "swizzled stage buffers stay 1024-byte aligned");
static_assert(ConvSmem<256>::kTotal <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void unpack2(uint32_t u, float& a, float& b) {
  a = __uint_as_float(u << 16);
  b = __uint_as_float(u & 0xffff0000u);
}

// k2Sm: one 256 x BN tcgen05.mma.cta_group::2 per K step across the CTA pair (see xv_gemm_sm100.cu).
template <int BN, bool k2Sm>
__global__ void __launch_bounds__(kGemmThreads, 1)
xv_conv_gemm_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                    const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ ConvGemmParams p) {
  using S = ConvSmem<BN>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* ring = smem;
  float2* ep_ss = reinterpret_cast<float2*>(smem + S::kRingMax);
  float* ep_sl = reinterpret_cast<float*>(smem + S::kRingMax + kAccStages * BN * 8);
  uint8_t* staging = smem + S::kRingMax + S::kParams;
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + S::kStaging);
  constexpr int kNumStages = k2Sm ? S::kStages2 : kStages;
  constexpr uint32_t kStageB = k2Sm ? S::kStage2 : S::kStage;
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kNumStages;
  uint64_t* tfull_bar = bars + 2 * kNumStages;
  uint64_t* tempty_bar = tfull_bar + kAccStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;
  // cluster of 1 or 2 CTAs; a pair takes two row tiles of one column tile and multicasts the halves of every weight
  // tile to each other (see xv_gemm_sm100.cu)
  const uint32_t crank = ptx::cluster_ctarank();
  const uint32_t csize = ptx::cluster_nctarank();
  const uint16_t cmask = static_cast<uint16_t>((1u << csize) - 1u);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tm_a);
    ptx::prefetch_tensormap(&tm_b);
    ptx::prefetch_tensormap(&tm_out);
    for (int s = 0; s < kNumStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], k2Sm ? 1 : csize);
    }
    for (int s = 0; s < kAccStages; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], k2Sm ? 2 * kConvEpiWarps : kConvEpiWarps);   // 2-SM: both CTAs' epilogues arrive on the leader's
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
  if (csize > 1) ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int n_units = ((p.m_tiles + static_cast<int>(csize) - 1) / static_cast<int>(csize)) * p.n_tiles;
  const int unit0 = blockIdx.x / csize;
  const int unit_step = gridDim.x / csize;
  const int num_kb = p.taps * p.kb_per_tap;
  const int tiles_per_utt = p.tiles_t * p.tiles_f;
  const uint32_t a_bytes = static_cast<uint32_t>(p.tile_rows) * kBK * 2;
  const uint32_t lead_full0 = k2Sm ? ptx::mapa_u32(ptx::smem_u32(&full_bar[0]), 0) : 0u;
  const uint32_t lead_tempty0 = k2Sm ? ptx::mapa_u32(ptx::smem_u32(&tempty_bar[0]), 0) : 0u;
  auto release_acc = [&](uint32_t as) {              // epilogue warp (lane 0) -> MMA warp of the pair's leader
    if (k2Sm) ptx::mbar_arrive_cluster(lead_tempty0 + as * 8);
    else ptx::mbar_arrive(&tempty_bar[as]);
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int u = unit0; u < n_units; u += unit_step) {
        const int mt = (u / p.n_tiles) * static_cast<int>(csize) + static_cast<int>(crank);
        const int n0 = (u % p.n_tiles) * BN;
        const int b = mt / tiles_per_utt;
        const int rem = mt - b * tiles_per_utt;
        const int tt = rem / p.tiles_f;
        const int to0 = tt * p.tb;
        const int fo0 = (rem - tt * p.tiles_f) * p.fb;
        for (int tap = 0; tap < p.taps; ++tap) {
          const ConvTap tp = p.tap[tap];
          for (int cb = 0; cb < p.kb_per_tap; ++cb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * kStageB;
            uint8_t* sb = sa + kConvABytes;
            const int kc = tap * p.cin + cb * kBK;
            if (k2Sm) {
              if (crank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * (a_bytes + S::kBBytes / 2));
              ptx::tma_load_5d_2sm(sa, &tm_a, lead_full0 + stage * 8, tp.c0 + cb * kBK, fo0 + tp.c1, tp.c2, to0 + tp.c3, b);
              ptx::tma_load_2d_2sm(sb, &tm_b, lead_full0 + stage * 8, kc, n0 + static_cast<int>(crank) * (BN / 2));
              if (++stage == kNumStages) { stage = 0; phase ^= 1; }
              continue;
            }
            ptx::mbar_arrive_expect_tx(&full_bar[stage], a_bytes + S::kBBytes);
            ptx::tma_load_5d(sa, &tm_a, &full_bar[stage], tp.c0 + cb * kBK, fo0 + tp.c1, tp.c2, to0 + tp.c3, b);
            if (csize > 1) {
              ptx::tma_load_2d_multicast(sb + crank * (S::kBBytes / 2), &tm_b, &full_bar[stage], kc, n0 + static_cast<int>(crank) * (BN / 2), cmask);
            } else {
              ptx::tma_load_2d(sb, &tm_b, &full_bar[stage], kc, n0);
              ptx::tma_load_2d(sb + S::kBBytes / 2, &tm_b, &full_bar[stage], kc, n0 + BN / 2);
            }
            if (++stage == kNumStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // warp-uniform loop, one elected lane issues (see xv_gemm_sm100.cu: a lone lane spends ~450 clk per round in
    // its own instruction stream, more than the 270 clk four 128-column MMAs execute in)
    constexpr uint32_t idesc = ptx::make_idesc_bf16(k2Sm ? 2 * kBM : kBM, BN);
    const uint32_t ring_lo = ptx::sw128_desc_lo(ptx::smem_u32(ring));
    uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
    const int mma_units = (k2Sm && crank != 0) ? unit0 : n_units;   // the peer's MMA warp only owns the TMEM allocation
    for (int u = unit0; u < mma_units; u += unit_step) {
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t a_lo = ring_lo + stage * (kStageB >> 4);
          const uint32_t b_lo = a_lo + (kConvABytes >> 4);
#pragma unroll
          for (int k = 0; k < kBK / kUK; ++k) {
            if (k2Sm) ptx::umma_bf16_lo_2sm(tmem_d, a_lo + 2 * k, b_lo + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            else ptx::umma_bf16_lo(tmem_d, a_lo + 2 * k, b_lo + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          if (k2Sm) ptx::umma_commit_2sm(&empty_bar[stage], 3);
          else if (csize > 1) ptx::umma_commit_multicast(&empty_bar[stage], cmask);
          else ptx::umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == kNumStages) { stage = 0; phase ^= 1; }
      }
      if (ptx::elect_one()) {
        if (k2Sm) ptx::umma_commit_2sm(&tfull_bar[as], 3);
        else ptx::umma_commit(&tfull_bar[as]);
      }
      __syncwarp();
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const uint32_t e = warp - 2;
    const uint32_t q = warp & 3;                 // TMEM lane quarter
    const uint32_t half = e >> 2;                // column half
    const uint32_t etid = threadIdx.x - 64;
    constexpr int kChunks = BN / 64;             // 32-column chunks per warp
    uint8_t* stage_buf = staging + half * (128 * 64);
    const bool issuer = q == 0 && lane == 0;
    const bool relu_only = p.relu_only != 0;
    const int r = q * 32 + lane;                 // row of the tile = TMEM lane
    const int t_l = r / p.fb;
    const int f_l = r - t_l * p.fb;
    uint32_t as = 0, aphase = 0;
    int slot_n0_a = -1, slot_n0_b = -1;          // column tile whose parameters each accumulator stage's slot holds
    bool bar_prev_tile = true;
    for (int u = unit0; u < n_units; u += unit_step) {
      const int mt = (u / p.n_tiles) * static_cast<int>(csize) + static_cast<int>(crank);
      const int n0 = (u % p.n_tiles) * BN;
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
      const int b = mt / tiles_per_utt;
      const int rem = mt - b * tiles_per_utt;
      const int tt = rem / p.tiles_f;
      const int to0 = tt * p.tb;
      const int fo0 = (rem - tt * p.tiles_f) * p.fb;
      float2* ss = ep_ss + as * BN;
      float* sl = ep_sl + as * BN;
      if ((as ? slot_n0_b : slot_n0_a) != n0) {   // this stage's slot holds another column tile's parameters (see xv_gemm_sm100.cu)
        if (!bar_prev_tile) asm volatile("bar.sync 1, 256;" ::: "memory");
        bar_prev_tile = true;
        if (etid < BN) {
          ss[etid] = make_float2(__ldg(p.scale + n0 + etid), __ldg(p.shift + n0 + etid));
          sl[etid] = __ldg(p.slope + n0 + etid);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (as) slot_n0_b = n0; else slot_n0_a = n0;
      } else {
        bar_prev_tile = false;
      }
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const bool row_ok = r < p.tile_rows && (to0 + t_l) < p.t_out;
      const __nv_bfloat16* res_row = nullptr;
      if (p.residual != nullptr && row_ok)
        res_row = p.residual + ((static_cast<size_t>(b) * p.t_alloc_res + to0 + t_l) * p.f_out + fo0 + f_l) * p.cout + n0;
#pragma unroll 1
      for (int ci = 0; ci < kChunks; ++ci) {
        const int c = half * kChunks + ci;
        uint32_t acc[32];
        ptx::tmem_ld_32x32(tmem_base + ((q * 32) << 16) + as * BN + c * 32, acc);
        uint4 rv[4];
        if (res_row != nullptr) {
          const uint4* rp = reinterpret_cast<const uint4*>(res_row + c * 32);
#pragma unroll
          for (int g = 0; g < 4; ++g) rv[g] = __ldg(rp + g);
        } else {
#pragma unroll
          for (int g = 0; g < 4; ++g) rv[g] = make_uint4(0, 0, 0, 0);
        }
        ptx::tmem_ld_wait();
        if (ci == kChunks - 1) {
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(as);
        }
        float y[32];
        const float4* ss4 = reinterpret_cast<const float4*>(ss + c * 32);
        const float2* sl2 = reinterpret_cast<const float2*>(sl + c * 32);
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(rv);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float4 sh = ss4[j];
          float r0, r1;
          unpack2(rw[j], r0, r1);
          float v0 = fmaf(__uint_as_float(acc[2 * j]), sh.x, sh.y) + r0;
          float v1 = fmaf(__uint_as_float(acc[2 * j + 1]), sh.z, sh.w) + r1;
          if (relu_only) {
            v0 = fmaxf(v0, 0.f);
            v1 = fmaxf(v1, 0.f);
          } else {
            const float2 ng = sl2[j];
            v0 = v0 >= 0.f ? v0 : v0 * ng.x;
            v1 = v1 >= 0.f ? v1 : v1 * ng.y;
          }
          y[2 * j] = v0;
          y[2 * j + 1] = v1;
        }
        // the previous chunk's bulk store must have finished reading the staging buffer
        if (issuer) ptx::tma_store_wait_read<0>();
        asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
        uint8_t* rowp = stage_buf + r * 64;
        const uint32_t sw = (r >> 1) & 3;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 v;
          v.x = pack2(y[g * 8 + 0], y[g * 8 + 1]);
          v.y = pack2(y[g * 8 + 2], y[g * 8 + 3]);
          v.z = pack2(y[g * 8 + 4], y[g * 8 + 5]);
          v.w = pack2(y[g * 8 + 6], y[g * 8 + 7]);
          *reinterpret_cast<uint4*>(rowp + ((g ^ sw) << 4)) = v;
        }
        ptx::fence_proxy_async_smem();
        asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
        if (issuer) {
          ptx::tma_store_4d(&tm_out, stage_buf, n0 + c * 32, fo0, to0, b);
          ptx::tma_store_commit();
        }
      }
      if (++as == kAccStages) { as = 0; aphase ^= 1; }
    }
    if (issuer) ptx::tma_store_wait<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (csize > 1) ptx::cluster_sync();
  if (warp == 1) {
    ptx::tc_fence_after();
    if (k2Sm) ptx::tmem_dealloc_2sm(tmem_base, kTmemCols);
    else ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace

template <int BN, bool k2Sm>
static cudaError_t conv_set_smem() {
  return cudaFuncSetAttribute(xv_conv_gemm_kernel<BN, k2Sm>, cudaFuncAttributeMaxDynamicSharedMemorySize, ConvSmem<BN>::kTotal);
}

const char* conv_gemm_init() {
  static char err[160];
  cudaError_t e = conv_set_smem<64, false>();
  if (e == cudaSuccess) e = conv_set_smem<128, false>();
  if (e == cudaSuccess) e = conv_set_smem<256, false>();
  if (e == cudaSuccess) e = conv_set_smem<64, true>();
  if (e == cudaSuccess) e = conv_set_smem<128, true>();
  if (e == cudaSuccess) e = conv_set_smem<256, true>();
  if (e != cudaSuccess) {
    snprintf(err, sizeof(err), "cudaFuncSetAttribute(conv max dynamic smem) failed: %s", cudaGetErrorString(e));
    return err;
  }
  return nullptr;
}

namespace {
template <int BN, bool k2Sm>
int conv_cluster_pairs() {
  static int cached = -1;
  if (cached >= 0) return cached;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2, 1, 1);
  cfg.blockDim = dim3(kGemmThreads, 1, 1);
  cfg.dynamicSmemBytes = ConvSmem<BN>::kTotal;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, xv_conv_gemm_kernel<BN, k2Sm>, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
  cached = n;
  return cached;
}

template <int BN>
cudaError_t launch_conv_variant(const ConvGemmPlan& plan, int sm_count, cudaStream_t stream) {
  const ConvGemmParams& p = plan.p;
  static const int want = getenv("XV_CONV_CLUSTER") ? atoi(getenv("XV_CONV_CLUSTER")) : 1;
  // cta_group::2 pairs for the long-K convolutions (3x3 with >= 128 input channels: >= 18 k-blocks); the short ones keep
  // multicast pairs (see launch_gemm in xv_gemm_sm100.cu for the measurements behind the rule).  XV_CONV_2SM=0|1|2.
  // Measured on resnet18 (one box, 1650 MHz: not power-capped like the TDNN step): conv.c512 1460 -> 1465, conv.c256
  // 1225 -> 1229 TFLOP/s -- within noise; these layers are not bound by weight traffic or operand reads.
  static const int want_2sm = getenv("XV_CONV_2SM") ? atoi(getenv("XV_CONV_2SM")) : 1;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kGemmThreads, 1, 1);
  cfg.dynamicSmemBytes = ConvSmem<BN>::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  const int units = ((p.m_tiles + 1) / 2) * p.n_tiles;
  if (want_2sm && p.m_tiles >= 2 && (p.taps * p.kb_per_tap >= 16 || want_2sm > 1)) {
    const int pairs = conv_cluster_pairs<BN, true>();
    static bool traced = false;
    if (!traced && getenv("XV_GEMM_TRACE")) { traced = true; fprintf(stderr, "xv conv<%d>: %d co-resident cta_group::2 pairs\n", BN, pairs); }
    if (2 * pairs >= sm_count - 4) {
      const int n = units < pairs ? units : pairs;
      cfg.gridDim = dim3(2 * n, 1, 1);
      cfg.attrs = &attr;
      cfg.numAttrs = 1;
      return cudaLaunchKernelEx(&cfg, xv_conv_gemm_kernel<BN, true>, plan.tm_a, plan.tm_b, plan.tm_out, p);
    }
  }
  int pairs = 0;
  if (want && p.m_tiles >= 2) {
    pairs = conv_cluster_pairs<BN, false>();
    if (2 * pairs < sm_count - 4) pairs = 0;
  }
  if (pairs > 0) {
    const int n = units < pairs ? units : pairs;
    cfg.gridDim = dim3(2 * n, 1, 1);
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
  } else {
    const int tiles = p.m_tiles * p.n_tiles;
    cfg.gridDim = dim3(tiles < sm_count ? tiles : sm_count, 1, 1);
  }
  return cudaLaunchKernelEx(&cfg, xv_conv_gemm_kernel<BN, false>, plan.tm_a, plan.tm_b, plan.tm_out, p);
}
}  // namespace

// plan.tm_b must map the weights in boxes of bn / 2 rows (one per CTA of a cluster pair).
cudaError_t launch_conv_gemm(const ConvGemmPlan& plan, int sm_count, cudaStream_t stream) {
  if (plan.p.m_tiles * plan.p.n_tiles <= 0) return cudaSuccess;
  if (plan.bn == 64) return launch_conv_variant<64>(plan, sm_count, stream);
  if (plan.bn == 128) return launch_conv_variant<128>(plan, sm_count, stream);
  return launch_conv_variant<256>(plan, sm_count, stream);
}

}  // namespace xv
