// libxvector.so C-ABI, part 4: statistics pooling for callers that must not allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// xv_stat_pooling_dev keeps its [n, 2C] fp32 sum / sum-of-squares workspace in a ctx buffer that grows with
// cudaMalloc on demand.  The variant below keeps the same workspace in a file-scope __device__ array of
// XV_NOALLOC_POOL_MAX_FLOATS floats (2 MiB, resident on every GPU that loads the library until the process exits)
// and launches the very same kernels (launch_pool: memset of the workspace, pool_accum_*_kernel, pool_finalize_kernel).
//
// The array's name is only a host-side stand-in in host code: the device address is taken once per device with
// cudaGetSymbolAddress while that device is current, and that pointer is what the kernels receive.
//
// Every call on a device shares the one array, which still holds the previous call's sums when the next call
// starts: launch_pool clears the part it uses first (cudaMemsetAsync on the ctx stream), so calls that are ordered
// on the GPU (same stream, or ordered by events) are independent; calls that OVERLAP on the GPU (two ctx on one
// device running this entry point concurrently) race on the array and must be serialised by the caller.
#include <mutex>

#include "xv_ctx.h"

using namespace xv;

namespace {

constexpr int64_t kNoAllocPoolMaxFloats = XV_NOALLOC_POOL_MAX_FLOATS;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_pool_scratch[kNoAllocPoolMaxFloats];

std::mutex g_scratch_mu;
float* g_scratch_dev[kMaxDevices] = {};   // device address of g_noalloc_pool_scratch, per device ordinal

// Device address of the scratch array on `device` (which must be current).  Looked up once per device.
cudaError_t scratch_address(int device, float** out) {
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (g_scratch_dev[device] == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_pool_scratch);
    if (e != cudaSuccess) return e;
    if (p == nullptr) return cudaErrorInvalidSymbol;
    g_scratch_dev[device] = p;
  }
  *out = g_scratch_dev[device];
  return cudaSuccess;
}

}  // namespace

extern "C" int xv_stat_pooling_noalloc(xv_ctx* c, const void* x_dev, const int32_t* cu_lens_dev, int32_t n, int32_t C,
                                       int dtype, int64_t total_rows, float* out_dev) {
  if (!c) return XV_EINVAL;
  if (!x_dev || !cu_lens_dev || !out_dev || n < 0 || C <= 0 || total_rows < 0 ||
      (dtype != XV_DTYPE_F32 && dtype != XV_DTYPE_BF16))
    return fail(c, XV_EINVAL, "xv_stat_pooling_noalloc: bad argument");
  const int64_t need = static_cast<int64_t>(n) * 2 * C;
  if (need > kNoAllocPoolMaxFloats)
    return fail(c, XV_EINVAL,
                "xv_stat_pooling_noalloc: n*2*C = %lld floats exceeds the built-in scratch of %lld floats "
                "(XV_NOALLOC_POOL_MAX_FLOATS); use xv_stat_pooling_dev or split the call",
                static_cast<long long>(need), static_cast<long long>(kNoAllocPoolMaxFloats));
  if (total_rows > 0x7fffffffLL) return fail(c, XV_EINVAL, "xv_stat_pooling_noalloc: total_rows exceeds int32 offsets");
  cudaError_t e = cudaSetDevice(c->device);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaSetDevice");
  if (n == 0 || total_rows == 0) return XV_OK;
  float* scratch = nullptr;
  e = scratch_address(c->device, &scratch);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaGetSymbolAddress(noalloc pooling scratch)");
  PoolArgs a;
  a.x = x_dev; a.dtype = dtype; a.c = C; a.ldx = C; a.rows = total_rows; a.cu = cu_lens_dev; a.n = n; a.context = 0;
  a.weights = nullptr; a.heads = 1; a.split_value = 1; a.acc = scratch;   // cleared by launch_pool before use
  a.out = out_dev; a.ldout = 2 * C; a.out_bf16 = nullptr; a.ldout_bf16 = 0;
  e = launch_pool(a, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "pooling launch (noalloc)");
  return XV_OK;
}
