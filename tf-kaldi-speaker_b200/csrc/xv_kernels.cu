// SIMT kernels of libxvector.so: input splice+cast, the fp32 validation GEMM, statistics / attention
// pooling reductions, and small gather / elementwise helpers.  Hand-written for sm_100a; no library calls.
#include <float.h>

#include "xv_internal.h"

namespace xv {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// ------------------------------------------------------------------------------------------------
// Layer-1 operand: fp32 features [rows, D] -> bf16 spliced rows [rows, kpad],
// out[r, j*D + c] = in[r + j*dil, c].  Fuses the fp32->bf16 cast with the context splice of tdnn1
// (model/tdnn.py:38-47: expand_dims + conv2d (1,5)); K = 5*D is tiny (150) so this is the one layer whose
// im2col rows are materialised (384 B/frame), which removes the 60-byte row stride TMA cannot address.
__global__ void splice_cast_kernel(const float* __restrict__ in, int64_t rows, int d, int taps, int dil,
                                   __nv_bfloat16* __restrict__ out, int kpad) {
  const int vec_per_row = kpad >> 3;
  const int64_t total = rows * vec_per_row;
  const int k_real = taps * d;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / vec_per_row;
    const int k0 = static_cast<int>(i - r * vec_per_row) << 3;
    float v[8];
    if (dil == 1) {
      // rows are contiguous with stride d, so the spliced row is the contiguous slab in[r*d .. r*d + taps*d)
      const int64_t base = r * d + k0;
      const int64_t total_in = rows * d;
#pragma unroll
      for (int e = 0; e < 8; ++e)
        v[e] = (k0 + e < k_real && base + e < total_in) ? __ldg(in + base + e) : 0.f;
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int k = k0 + e;
        float x = 0.f;
        if (k < k_real) {
          const int j = k / d;
          const int c = k - j * d;
          const int64_t rr = r + static_cast<int64_t>(j) * dil;
          if (rr < rows) x = __ldg(in + rr * d + c);
        }
        v[e] = x;
      }
    }
    __nv_bfloat162 b0 = __floats2bfloat162_rn(v[0], v[1]);
    __nv_bfloat162 b1 = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 b2 = __floats2bfloat162_rn(v[4], v[5]);
    __nv_bfloat162 b3 = __floats2bfloat162_rn(v[6], v[7]);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t*>(&b0);
    o.y = *reinterpret_cast<uint32_t*>(&b1);
    o.z = *reinterpret_cast<uint32_t*>(&b2);
    o.w = *reinterpret_cast<uint32_t*>(&b3);
    *reinterpret_cast<uint4*>(out + r * kpad + k0) = o;
  }
}

// ------------------------------------------------------------------------------------------------
// fp32 validation GEMM (the "fp32 validation mode" of BASELINE.json): same tap-splice addressing as the
// tcgen05 kernel but plain FFMA, fp32 in / fp32 accumulate / fp32 out, with the three reference endpoints of a
// layer (x_dense|conv, x_bn, x_relu) written separately (model/tdnn.py:48-55).  Not throughput-judged.
constexpr int kSBM = 64, kSBN = 64, kSBK = 16;

__global__ void __launch_bounds__(256) simt_gemm_kernel(const SimtGemmArgs a) {
  __shared__ float As[kSBK][kSBM + 4];
  __shared__ float Ws[kSBK][kSBN + 4];
  const int tx = threadIdx.x & 15;
  const int ty = threadIdx.x >> 4;
  const int64_t m0 = static_cast<int64_t>(blockIdx.x) * kSBM;
  const int n0 = blockIdx.y * kSBN;
  const int K = a.taps * a.cin;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += kSBK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = threadIdx.x + i * 256;
      {
        const int kk = e & 15, mm = e >> 4;
        const int k = k0 + kk;
        float v = 0.f;
        if (k < K) {
          const int j = k / a.cin;
          const int c = k - j * a.cin;
          const int64_t r = m0 + mm + static_cast<int64_t>(j) * a.dilation;
          if (r < a.a_rows) v = __ldg(a.a + r * a.cin + c);
        }
        As[kk][mm] = v;
      }
      {
        const int nn = e & 63, kk = e >> 6;
        const int k = k0 + kk, n = n0 + nn;
        Ws[kk][nn] = (k < K && n < a.n) ? __ldg(a.w + static_cast<int64_t>(k) * a.n + n) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kSBK; ++kk) {
      float av[4], wv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) wv[j] = Ws[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= a.m_rows) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= a.n) continue;
      float v = acc[i][j] + (a.bias ? __ldg(a.bias + n) : 0.f);
      if (a.out_dense) a.out_dense[m * a.n + n] = v;
      if (a.bn_scale) v = fmaf(v, __ldg(a.bn_scale + n), __ldg(a.bn_shift + n));
      if (a.out_bn) a.out_bn[m * a.n + n] = v;
      if (a.slope) v = v >= 0.f ? v : v * __ldg(a.slope + n);
      if (a.act_tanh) v = tanhf(v);
      if (a.out_act) a.out_act[m * a.n + n] = v;
    }
  }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dst[i] = __float2bfloat16_rn(src[i]);
}

__global__ void fill_u32_kernel(uint32_t* __restrict__ dst, size_t n, uint32_t v) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dst[i] = v;
}

// out = act(scale*in + shift) on [n_rows, width] (att_post_bn + relu, model/pooling.py:222-229).
__global__ void affine_rows_kernel(const float* __restrict__ in, int ldi, int n_rows, int width,
                                   const float* __restrict__ scale, const float* __restrict__ shift,
                                   const float* __restrict__ slope, float* __restrict__ out_f32, int ldo,
                                   __nv_bfloat16* __restrict__ out_bf16, int ldob) {
  const int64_t total = static_cast<int64_t>(n_rows) * width;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / width);
    const int c = static_cast<int>(i - static_cast<int64_t>(r) * width);
    float v = in[static_cast<int64_t>(r) * ldi + c];
    if (scale) v = fmaf(v, scale[c], shift[c]);
    if (slope) v = v >= 0.f ? v : v * slope[c];
    if (out_f32) out_f32[static_cast<int64_t>(r) * ldo + c] = v;
    if (out_bf16) out_bf16[static_cast<int64_t>(r) * ldob + c] = __float2bfloat16_rn(v);
  }
}

// l2_scaling (model/common.py:45-58): x * rsqrt(max(sum x^2, eps)) * factor, one warp per row.
__global__ void l2_scale_rows_kernel(float* __restrict__ x, int ld, int n_rows, int width, float factor, float eps) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= n_rows) return;
  float* p = x + static_cast<int64_t>(row) * ld;
  float s = 0.f;
  for (int c = lane; c < width; c += 32) s = fmaf(p[c], p[c], s);
  s = warp_sum(s);
  const float inv = rsqrtf(fmaxf(s, eps)) * factor;
  for (int c = lane; c < width; c += 32) p[c] *= inv;
}

}  // namespace

// One warp per packed row: find the view the row belongs to (binary search over the batch-local offsets) and copy its
// `dim` floats from the source row.  Used by xv_forward_views (overlapping chunks of the extractor, extract.py:68-86).
static __global__ void gather_views_kernel(const float* __restrict__ src, const int32_t* __restrict__ cu,
                                           const int64_t* __restrict__ src0, int32_t n_views, int64_t rows, int dim,
                                           float* __restrict__ dst) {
  const int64_t r = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  int lo = 0, hi = n_views - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (cu[mid] <= r) lo = mid; else hi = mid - 1;
  }
  const float* s = src + (src0[lo] + (r - cu[lo])) * dim;
  float* d = dst + r * dim;
  for (int k = lane; k < dim; k += 32) d[k] = __ldg(s + k);
}

cudaError_t launch_gather_views(const float* src, const int32_t* cu, const int64_t* src0, int32_t n_views, int64_t rows, int dim,
                                float* dst, cudaStream_t stream) {
  if (rows <= 0 || n_views <= 0) return cudaSuccess;
  const int64_t blocks = (rows + 7) / 8;
  gather_views_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(src, cu, src0, n_views, rows, dim, dst);
  return cudaGetLastError();
}

cudaError_t launch_splice_cast(const float* feats, int64_t rows, int d, int taps, int dil, __nv_bfloat16* out,
                               int kpad, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  const int64_t total = rows * (kpad >> 3);
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  splice_cast_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(feats, rows, d, taps, dil, out, kpad);
  return cudaGetLastError();
}

cudaError_t launch_simt_gemm(const SimtGemmArgs& a, cudaStream_t stream) {
  if (a.m_rows <= 0) return cudaSuccess;
  dim3 grid(static_cast<unsigned>((a.m_rows + kSBM - 1) / kSBM), static_cast<unsigned>((a.n + kSBN - 1) / kSBN));
  simt_gemm_kernel<<<grid, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_f32_to_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  f32_to_bf16_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(src, dst, n);
  return cudaGetLastError();
}

cudaError_t launch_fill_u32(uint32_t* dst, size_t n, uint32_t v, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  fill_u32_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(dst, n, v);
  return cudaGetLastError();
}

cudaError_t launch_affine_rows(const float* in, int ldi, int n_rows, int width, const float* scale,
                               const float* shift, const float* slope, float* out_f32, int ldo,
                               __nv_bfloat16* out_bf16, int ldob, cudaStream_t stream) {
  const int64_t total = static_cast<int64_t>(n_rows) * width;
  if (total <= 0) return cudaSuccess;
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  affine_rows_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(in, ldi, n_rows, width, scale, shift, slope,
                                                                   out_f32, ldo, out_bf16, ldob);
  return cudaGetLastError();
}

cudaError_t launch_l2_scale_rows(float* x, int ld, int n_rows, int width, float factor, float eps,
                                 cudaStream_t stream) {
  if (n_rows <= 0) return cudaSuccess;
  l2_scale_rows_kernel<<<(n_rows + 7) / 8, 256, 0, stream>>>(x, ld, n_rows, width, factor, eps);
  return cudaGetLastError();
}

}  // namespace xv
