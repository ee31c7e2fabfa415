// 3x3 stride-1 'SAME' convolutions with 64 or 128 input channels (ResNet-18 stages 1-2) on sm_100a: weights-stationary
// halo-tile variant of xv_conv_sm100.cu (same contract: model/resnet.py:9-149 conv2d + batch_normalization
// [+ shortcut] [+ relu]).
//
// Why a second kernel.  With K = 9*64 and 64..128 output channels a 128-row tile is only 1152..4608 tensor-pipe
// cycles of math, while the generic kernel streams a 16 KB activation box AND an 8..16 KB weight tile through L2 for
// every (tap, channel block): ~190 B/clk/SM wanted, ~45 B/clk/SM is what L2 delivers with all 148 SMs pulling
// (measured: stage 1 ran at 368 TFLOP/s, exactly that cap).  Two changes remove the traffic:
//
//  * weights stay in shared memory.  Each CTA owns ONE 64-column slab of output channels for its whole life and loads
//    the 9 x (Cin/64) weight tiles of that slab once (72 KB for Cin = 64, 144 KB for Cin = 128).
//  * the activation tile is loaded once per (tile, channel block) together with its halo: a box of (tb + 2) time steps
//    x (fb + 2) frequency bins x 64 channels = consecutive 128-byte rows in shared memory (SWIZZLE_128B).  The output
//    tile is laid out on the SAME (fb + 2)-wide grid (two junk columns per time step), so output row m reads, for tap
//    (dt, df), halo row m + dt*(fb+2) + df: every tap is the same box shifted by a constant number of rows, i.e. the
//    same UMMA descriptor with a different start address.  The 128-byte swizzle is a function of the absolute
//    shared-memory address (bits [4,7) ^= bits [7,10)) for TMA and for the tensor core alike, so a start address that
//    is a multiple of 128 B but not of 1024 B needs NO base_offset in the descriptor (measured: base_offset =
//    (address >> 7) & 7 gives wrong sums, 0 is exact).  Rows the shifted view reads beyond the box only feed junk
//    output rows (an accumulator row depends on its own A row alone).
//
// L2 traffic per 128x64 output tile drops from 9*(Cin/64) x 24 KB to (Cin/64) x <= 21 KB.  Out-of-range coordinates
// (frequency / time padding, utterance ends) are zero-filled by the TMA unit exactly as in the generic kernel.
// Roles: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM owner), warps 2..9 = epilogue (lane quarter = warp % 4,
// 32-column half = (warp - 2) / 4); four 64-column accumulators in TMEM decouple the epilogue from the MMA stream.
#include <stdio.h>

#include "xv_internal.h"
#include "xv_ptx.cuh"

namespace xv {

namespace {

constexpr int kHaloBN = 64;                      // output channels per CTA slab
constexpr int kHaloAcc = 4;                      // TMEM accumulator stages (4 x 64 columns)
constexpr int kHaloMaxAStages = 8;
constexpr uint32_t kHaloBTile = kHaloBN * kBK * 2;   // one (tap, channel block) weight tile: 8 KB
constexpr uint32_t kHaloStaging = 2 * 128 * 64;      // two 32-column halves x 128 rows x 64 B
constexpr uint32_t kHaloParams = 1024;               // (scale, shift) float2 + slope float for 64 columns
constexpr uint32_t kHaloBars = 256;
constexpr uint32_t kSmemLimit = 232448;              // 227 KB

__device__ __forceinline__ uint32_t hpack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void hunpack2(uint32_t u, float& a, float& b) {
  a = __uint_as_float(u << 16);
  b = __uint_as_float(u & 0xffff0000u);
}

__global__ void __launch_bounds__(kGemmThreads, 1)
xv_conv_halo_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                    const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ ConvGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t b_bytes = 9u * p.kb_per_tap * kHaloBTile;
  uint8_t* b_res = smem;                                     // resident weights [tap][cb][64 x 64]
  uint8_t* a_ring = smem + b_bytes;
  uint8_t* staging = a_ring + static_cast<uint32_t>(p.a_stages) * p.a_stage_bytes;
  float2* ep_ss = reinterpret_cast<float2*>(staging + kHaloStaging);
  float* ep_sl = reinterpret_cast<float*>(staging + kHaloStaging + kHaloBN * 8);
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + kHaloStaging + kHaloParams);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kHaloMaxAStages;
  uint64_t* b_full = a_empty + kHaloMaxAStages;
  uint64_t* tfull_bar = b_full + 1;
  uint64_t* tempty_bar = tfull_bar + kHaloAcc;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kHaloAcc);

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = kHaloAcc * kHaloBN;
  const int n_stages = p.a_stages;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tm_a);
    ptx::prefetch_tensormap(&tm_b);
    ptx::prefetch_tensormap(&tm_out);
    for (int s = 0; s < n_stages; ++s) { ptx::mbar_init(&a_full[s], 1); ptx::mbar_init(&a_empty[s], 1); }
    ptx::mbar_init(b_full, 1);
    for (int s = 0; s < kHaloAcc; ++s) { ptx::mbar_init(&tfull_bar[s], 1); ptx::mbar_init(&tempty_bar[s], 8); }
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

  // this CTA's slab of output channels and its share of the row tiles
  const int nt = blockIdx.x % p.n_tiles;
  const int n0 = nt * kHaloBN;
  const int m_first = blockIdx.x / p.n_tiles;
  const int m_step = gridDim.x / p.n_tiles;
  const int tiles_per_utt = p.tiles_t * p.tiles_f;
  const int fbh = p.fbh;
  const uint32_t a_bytes = static_cast<uint32_t>((p.tb + 2) * fbh) * kBK * 2;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      ptx::mbar_arrive_expect_tx(b_full, b_bytes);
      for (int tap = 0; tap < 9; ++tap)
        for (int cb = 0; cb < p.kb_per_tap; ++cb)
          ptx::tma_load_2d(b_res + (tap * p.kb_per_tap + cb) * kHaloBTile, &tm_b, b_full, tap * p.cin + cb * kBK, n0);
      uint32_t sa = 0, pa = 0;
      for (int mt = m_first; mt < p.m_tiles; mt += m_step) {
        const int b = mt / tiles_per_utt;
        const int rem = mt - b * tiles_per_utt;
        const int tt = rem / p.tiles_f;
        const int to0 = tt * p.tb;
        const int fo0 = (rem - tt * p.tiles_f) * p.fb;
        for (int cb = 0; cb < p.kb_per_tap; ++cb) {
          ptx::mbar_wait(&a_empty[sa], pa ^ 1);
          ptx::mbar_arrive_expect_tx(&a_full[sa], a_bytes);
          // halo box: channels [cb*64, +64), bins [fo0-1, fo0+fb+1), steps [to0-1, to0+tb+1) of utterance b
          ptx::tma_load_5d(a_ring + sa * p.a_stage_bytes, &tm_a, &a_full[sa], cb * kBK, fo0 - 1, 0, to0 - 1, b);
          if (++sa == static_cast<uint32_t>(n_stages)) { sa = 0; pa ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs the loop (waits included) so that every operand is warp-uniform; one elected lane issues.
    // A single thread sustains ~450 clk per (wait, 4 MMAs, commit) round in its own instruction stream while four
    // N = 64 MMAs execute in 192 clk (tools/debug/umma_issue.cu), so the loop is kept to a few instructions per MMA:
    // 32-bit descriptor arithmetic, per-tap offsets in registers, one commit per channel block.
    constexpr uint32_t idesc = ptx::make_idesc_bf16(kBM, kHaloBN);
    const uint32_t a_lo0 = ptx::sw128_desc_lo(ptx::smem_u32(a_ring));
    const uint32_t b_lo0 = ptx::sw128_desc_lo(ptx::smem_u32(b_res));
    const uint32_t a_stage_step = static_cast<uint32_t>(p.a_stage_bytes) >> 4;
    const uint32_t b_tap_step = static_cast<uint32_t>(p.kb_per_tap) * (kHaloBTile >> 4);
    uint32_t tap_off[9];
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) tap_off[tap] = static_cast<uint32_t>((tap / 3) * fbh + tap % 3) * 8u;   // rows * 128 B >> 4
    uint32_t sa = 0, pa = 0, as = 0, aphase = 0;
    ptx::mbar_wait(b_full, 0);
    for (int mt = m_first; mt < p.m_tiles; mt += m_step) {
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
      const uint32_t tmem_d = tmem_base + as * kHaloBN;
      for (int cb = 0; cb < p.kb_per_tap; ++cb) {
        ptx::mbar_wait(&a_full[sa], pa);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t a_lo = a_lo0 + sa * a_stage_step;
          uint32_t b_lo = b_lo0 + cb * (kHaloBTile >> 4);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
#pragma unroll
            for (int k = 0; k < kBK / kUK; ++k)
              ptx::umma_bf16_lo(tmem_d, a_lo + tap_off[tap] + 2 * k, b_lo + 2 * k, idesc, (cb | tap | k) ? 1u : 0u);
            b_lo += b_tap_step;
          }
          ptx::umma_commit(&a_empty[sa]);
        }
        __syncwarp();
        if (++sa == static_cast<uint32_t>(n_stages)) { sa = 0; pa ^= 1; }
      }
      if (ptx::elect_one()) ptx::umma_commit(&tfull_bar[as]);
      __syncwarp();
      if (++as == kHaloAcc) { as = 0; aphase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const uint32_t q = warp & 3;                 // TMEM lane quarter
    const uint32_t half = (warp - 2) >> 2;       // 32-column half of the slab
    const uint32_t etid = threadIdx.x - 64;
    uint8_t* stage_buf = staging + half * (128 * 64);
    const bool issuer = q == 0 && lane == 0;
    const bool relu_only = p.relu_only != 0;
    const int r = q * 32 + lane;                 // accumulator row = TMEM lane = position on the (tb x fbh) grid
    const int t_l = r / fbh;
    const int f_l = r - t_l * fbh;
    const bool col_ok = f_l < p.fb && t_l < p.tb;          // not a halo (junk) column, inside the tile
    const int rs = t_l * p.fb + f_l;                        // compacted row in the store staging buffer (fb x tb box)
    // the slab's folded-BN parameters never change: load them once
    if (etid < kHaloBN) {
      ep_ss[etid] = make_float2(__ldg(p.scale + n0 + etid), __ldg(p.shift + n0 + etid));
      ep_sl[etid] = __ldg(p.slope + n0 + etid);
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const float4* ss4 = reinterpret_cast<const float4*>(ep_ss + half * 32);
    const float2* sl2 = reinterpret_cast<const float2*>(ep_sl + half * 32);
    const int ncol = n0 + half * 32;
    uint32_t as = 0, aphase = 0;
    // shortcut rows of one tile for this thread (zeros when there is no shortcut / the row is junk or past the end)
    auto load_shortcut = [&](int mt, uint4 (&rv)[4]) {
      const int b = mt / tiles_per_utt;
      const int rem = mt - b * tiles_per_utt;
      const int tt = rem / p.tiles_f;
      const int to0 = tt * p.tb;
      const int fo0 = (rem - tt * p.tiles_f) * p.fb;
      if (p.residual != nullptr && mt < p.m_tiles && col_ok && (to0 + t_l) < p.t_out) {
        const uint4* rp = reinterpret_cast<const uint4*>(
            p.residual + ((static_cast<size_t>(b) * p.t_alloc_res + to0 + t_l) * p.f_out + fo0 + f_l) * p.cout + ncol);
#pragma unroll
        for (int g = 0; g < 4; ++g) rv[g] = __ldg(rp + g);
      } else {
#pragma unroll
        for (int g = 0; g < 4; ++g) rv[g] = make_uint4(0, 0, 0, 0);
      }
    };
    uint4 rv[4], rn[4];
    load_shortcut(m_first, rv);
    for (int mt = m_first; mt < p.m_tiles; mt += m_step) {
      const int b = mt / tiles_per_utt;
      const int rem = mt - b * tiles_per_utt;
      const int tt = rem / p.tiles_f;
      const int to0 = tt * p.tb;
      const int fo0 = (rem - tt * p.tiles_f) * p.fb;
      // the NEXT tile's shortcut rows are requested now: a full HBM round trip hides behind this tile's epilogue
      load_shortcut(mt + m_step, rn);
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      uint32_t acc[32];
      ptx::tmem_ld_32x32(tmem_base + ((q * 32) << 16) + as * kHaloBN + half * 32, acc);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
      if (++as == kHaloAcc) { as = 0; aphase ^= 1; }
      uint32_t o[16];
      const uint32_t* rw = reinterpret_cast<const uint32_t*>(rv);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float4 sh = ss4[j];
        float r0, r1;
        hunpack2(rw[j], r0, r1);
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
        o[j] = hpack2(v0, v1);
      }
      // the previous tile's bulk store must have finished reading the staging buffer
      if (issuer) ptx::tma_store_wait_read<0>();
      asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
      if (col_ok) {                              // junk columns are dropped: the staging rows are the dense fb x tb box
        uint8_t* rowp = stage_buf + rs * 64;
        const uint32_t sw = (rs >> 1) & 3;
#pragma unroll
        for (int g = 0; g < 4; ++g)
          *reinterpret_cast<uint4*>(rowp + ((g ^ sw) << 4)) = make_uint4(o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
      }
      ptx::fence_proxy_async_smem();
      asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
      if (issuer) {
        ptx::tma_store_4d(&tm_out, stage_buf, ncol, fo0, to0, b);
        ptx::tma_store_commit();
      }
#pragma unroll
      for (int g = 0; g < 4; ++g) rv[g] = rn[g];
    }
    if (issuer) ptx::tma_store_wait<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace

const char* conv_halo_init() {
  static char err[160];
  cudaError_t e = cudaFuncSetAttribute(xv_conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
  if (e != cudaSuccess) {
    snprintf(err, sizeof(err), "cudaFuncSetAttribute(conv halo max dynamic smem) failed: %s", cudaGetErrorString(e));
    return err;
  }
  return nullptr;
}

// Shared-memory plan of one launch: resident weights + as many halo stages as fit (2..8).  false = does not fit.
bool conv_halo_plan(ConvGemmParams* p, uint32_t* smem_bytes) {
  const uint32_t b_bytes = 9u * p->kb_per_tap * kHaloBTile;
  const uint32_t stage = ((static_cast<uint32_t>((p->tb + 2) * p->fbh) * kBK * 2) + 1023u) & ~1023u;
  const uint32_t fixed = 1024 + b_bytes + kHaloStaging + kHaloParams + kHaloBars;
  if (fixed + 2 * stage > kSmemLimit) return false;
  uint32_t n = (kSmemLimit - fixed) / stage;
  if (n > static_cast<uint32_t>(kHaloMaxAStages)) n = kHaloMaxAStages;
  p->a_stage_bytes = static_cast<int32_t>(stage);
  p->a_stages = static_cast<int32_t>(n);
  *smem_bytes = fixed + n * stage;
  return true;
}

cudaError_t launch_conv_halo(const ConvGemmPlan& plan, int sm_count, cudaStream_t stream) {
  const int tiles = plan.p.m_tiles * plan.p.n_tiles;
  if (tiles <= 0) return cudaSuccess;
  int grid = tiles < sm_count ? tiles : sm_count;
  grid -= grid % plan.p.n_tiles;               // every CTA owns one slab: whole groups of n_tiles CTAs
  if (grid < plan.p.n_tiles) grid = plan.p.n_tiles;
  xv_conv_halo_kernel<<<grid, kGemmThreads, plan.halo_smem, stream>>>(plan.tm_a, plan.tm_b, plan.tm_out, plan.p);
  return cudaGetLastError();
}

}  // namespace xv
