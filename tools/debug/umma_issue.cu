// Microbenchmark: cost of the single-thread tcgen05 issue loop (wait full barrier, 4 x UMMA 128xNx16, commit) in
// several code shapes.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I tf-kaldi-speaker_b200/csrc ...
#include <stdio.h>
#include <stdlib.h>

#include "xv_ptx.cuh"
using namespace xv;

// MODE 0: lane 0 alone: wait(r - depth) + 4 MMA + commit      (shape of the product kernels)
// MODE 1: lane 0 alone: 4 MMA + commit, no waits
// MODE 2: lane 0 alone: 4 MMA, no commit, no waits
// MODE 3: whole warp runs the loop, elect_one guards MMA + commit, wait by all lanes
// MODE 4: lane 0 alone, the try_wait of round r+1 is issued before the MMAs of round r
template <int N, int MODE>
__global__ void __launch_bounds__(128, 1) issue_kernel(int rounds, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar[8];
  __shared__ uint64_t done_bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < (192 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) ptx::mbar_init(&bar[i], 1);
    ptx::mbar_init(&done_bar, 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = slot;
  constexpr uint32_t idesc = ptx::make_idesc_bf16(128, N);
  constexpr int kDepth = 4;
  const uint32_t base = ptx::smem_u32(smem);
  if (threadIdx.x < 32) {
    const bool lead = MODE == 3 ? true : threadIdx.x == 0;
    long long t0 = clock64();
    if (lead) {
      uint32_t s = 0, ph = 0;
      bool ready = false;
      for (int r = 0; r < rounds; ++r) {
        if (MODE == 0 || MODE == 3) {
          if (r >= kDepth) ptx::mbar_wait(&bar[s], ph ^ 1);
        } else if (MODE == 4) {
          if (r >= kDepth && !ready) ptx::mbar_wait(&bar[s], ph ^ 1);
          const uint32_t s2 = s + 1 == kDepth ? 0 : s + 1;
          const uint32_t ph2 = s + 1 == kDepth ? ph ^ 1 : ph;
          ready = r + 1 >= kDepth ? ptx::mbar_try_wait(&bar[s2], ph2 ^ 1) : true;
        }
        const uint32_t a0 = base + s * 49152;
        const uint64_t ad = ptx::make_kmajor_sw128_desc(a0);
        const uint64_t bd = ptx::make_kmajor_sw128_desc(a0 + 16384);
        if (MODE == 3) {
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::umma_bf16(tm + (r & 1) * N, ad + 2 * k, bd + 2 * k, idesc, k ? 1u : 0u);
            ptx::umma_commit(&bar[s]);
          }
          __syncwarp();
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) ptx::umma_bf16(tm + (r & 1) * N, ad + 2 * k, bd + 2 * k, idesc, k ? 1u : 0u);
          if (MODE != 2) ptx::umma_commit(&bar[s]);
        }
        if (++s == kDepth) { s = 0; ph ^= 1; }
      }
      if (MODE == 3) {
        if (ptx::elect_one()) ptx::umma_commit(&done_bar);
        __syncwarp();
      } else {
        ptx::umma_commit(&done_bar);
      }
      ptx::mbar_wait(&done_bar, 0);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

template <int N, int MODE>
void run(int grid, int rounds) {
  long long* d;
  cudaMalloc(&d, grid * sizeof(long long));
  cudaFuncSetAttribute(issue_kernel<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  issue_kernel<N, MODE><<<grid, 128, 200 * 1024>>>(rounds, d);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  issue_kernel<N, MODE><<<grid, 128, 200 * 1024>>>(rounds, d);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[256];
  cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  printf("N=%3d mode=%d grid=%3d: %.1f clk/round (4 MMAs), %.0f TFLOP/s chip  [%s]\n", N, MODE, grid, double(h[0]) / rounds,
         grid * double(rounds) * 4 * 2.0 * 128 * N * 16 / (ms * 1e-3) / 1e12, cudaGetErrorString(err));
  cudaFree(d);
}

int main() {
  const int grid = 148, rounds = 20000;
  run<256, 0>(grid, rounds); run<256, 1>(grid, rounds); run<256, 2>(grid, rounds); run<256, 3>(grid, rounds); run<256, 4>(grid, rounds);
  run<64, 0>(grid, rounds); run<64, 1>(grid, rounds); run<64, 2>(grid, rounds); run<64, 3>(grid, rounds); run<64, 4>(grid, rounds);
  run<128, 0>(grid, rounds); run<128, 2>(grid, rounds);
  return 0;
}
