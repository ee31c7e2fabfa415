// Microbenchmark: issue rate of tcgen05.mma kind::f16 (bf16, M=128, K=16) as a function of N, operands in shared memory
// (SWIZZLE_128B K-major), no loads, one issuing thread per CTA.   nvcc -gencode arch=compute_100a,code=sm_100a -O3
// -I tf-kaldi-speaker_b200/csrc tools/debug/umma_rate.cu -o tools/debug/umma_rate
#include <stdio.h>
#include <stdlib.h>

#include "xv_ptx.cuh"
using namespace xv;

template <int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(int rounds, int per_commit, int a_shift_rows, long long* cycles, int depth, int lg) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar[16];
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { for (int i = 0; i < 16; ++i) ptx::mbar_init(&bar[i], 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = ptx::make_idesc_bf16(128, N);
    const uint32_t a0 = ptx::smem_u32(smem);
    const uint32_t b0 = a0 + 24 * 1024;
    long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
      if (r >= depth) ptx::mbar_wait(&bar[r & (depth - 1)], ((r >> lg) - 1) & 1);     // at most `depth` commits in flight
      for (int j = 0; j < per_commit; ++j) {
        const uint64_t ad = ptx::make_kmajor_sw128_desc(a0 + ((j * a_shift_rows) & 63) * 128u);
        const uint64_t bd = ptx::make_kmajor_sw128_desc(b0);
#pragma unroll
        for (int k = 0; k < 4; ++k) ptx::umma_bf16(tm + (r & 1) * N, ad + 2 * k, bd + 2 * k, idesc, (j | k) ? 1u : 0u);
      }
      ptx::umma_commit(&bar[r & (depth - 1)]);
    }
    for (int r = rounds > depth ? rounds - depth : 0; r < rounds; ++r) ptx::mbar_wait(&bar[r & (depth - 1)], (r >> lg) & 1);
    long long t1 = clock64();
    cycles[blockIdx.x] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

template <int N>
void run(int grid, int rounds, int per_commit, int shift, int lg = 2) {
  const int depth = 1 << lg;
  long long* d;
  cudaMalloc(&d, grid * sizeof(long long));
  cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  rate_kernel<N><<<grid, 128, 100 * 1024>>>(rounds, per_commit, shift, d, depth, lg);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  rate_kernel<N><<<grid, 128, 100 * 1024>>>(rounds, per_commit, shift, d, depth, lg);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[256];
  cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  const double n_mma = double(rounds) * per_commit * 4;
  printf("N=%3d grid=%3d depth=%2d per_commit=%2d shift=%d: %.1f clk/MMA (cta0), %.1f ns/MMA, %.0f TFLOP/s chip  [%s]\n", N, grid,
         depth, per_commit, shift, h[0] / n_mma, ms * 1e6 / n_mma, grid * n_mma * 2.0 * 128 * N * 16 / (ms * 1e-3) / 1e12,
         cudaGetErrorString(err));
  cudaFree(d);
}

int main() {
  const int grid = 148;
  for (int lg : {1, 2, 3, 4})
    for (int pc : {1, 2, 4}) {
      run<64>(grid, 20000 / pc, pc, 0, lg);
      run<256>(grid, 20000 / pc, pc, 0, lg);
    }
  return 0;
}
