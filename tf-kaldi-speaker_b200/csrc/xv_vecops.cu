// Post-extraction vector operations on the device: what the recipe does to the x-vectors
// after extract.py -- egs/voxceleb/v1/nnet/run_extract_embeddings.sh:80-103 (ivector-normalize-length --scaleup=false,
// ivector-mean over spk2utt) and the cosine trial scores of misc/utils.py:307-346 (compute_cos_pairwise_eer).
// Tiny, bandwidth-bound fp32 kernels: one warp per vector for the norms, one block per speaker for the means
// (deterministic member order), a 64 x 64 tiled dot-product kernel for the score matrix.
#include <math.h>

#include <vector>

#include "xv_ctx.h"

namespace xv {
namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// y = x / ratio, ratio = |x| (scaleup = 0) or |x| / sqrt(dim) (scaleup = 1); a zero vector is left untouched
// (ivector-normalize-length warns "Zero iVector" and writes it unchanged).
__global__ void normalize_length_kernel(const float* __restrict__ x, int64_t n, int dim, int scaleup, float* __restrict__ y) {
  const int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= n) return;
  const float* p = x + row * dim;
  float ss = 0.f;
  for (int c = lane; c < dim; c += 32) ss = fmaf(p[c], p[c], ss);
  ss = warp_sum(ss);
  float ratio = sqrtf(ss);
  if (scaleup) ratio /= sqrtf(static_cast<float>(dim));
  const float inv = ratio == 0.f ? 1.f : 1.f / ratio;
  for (int c = lane; c < dim; c += 32) y[row * dim + c] = p[c] * inv;
}

// mean[g] = sum_{i in members(g)} x[i] / count(g); members in CSR form (start[g] .. start[g+1]) listing row indices.
__global__ void group_mean_kernel(const float* __restrict__ x, int dim, const int64_t* __restrict__ start,
                                  const int64_t* __restrict__ member, float* __restrict__ out) {
  const int g = blockIdx.x;
  const int64_t a = start[g], b = start[g + 1];
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {
    float s = 0.f;
    for (int64_t k = a; k < b; ++k) s += x[member[k] * dim + c];
    out[static_cast<int64_t>(g) * dim + c] = b > a ? s * (1.f / static_cast<float>(b - a)) : 0.f;
  }
}

// rnorm[i] = 1 / sqrt(|x_i|^2 + eps)   (misc/utils.py:318: embeddings /= sqrt(sum(e**2) + 1e-12))
__global__ void row_rnorm_kernel(const float* __restrict__ x, int64_t n, int dim, float eps, float* __restrict__ r) {
  const int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= n) return;
  float ss = 0.f;
  for (int c = lane; c < dim; c += 32) ss = fmaf(x[row * dim + c], x[row * dim + c], ss);
  ss = warp_sum(ss);
  if (lane == 0) r[row] = 1.f / sqrtf(ss + eps);
}

// s[i, j] = (a_i . b_j) * ra[i] * rb[j]
__global__ void __launch_bounds__(256) cosine_scores_kernel(const float* __restrict__ a, const float* __restrict__ ra, int64_t na,
                                                            const float* __restrict__ b, const float* __restrict__ rb, int64_t nb,
                                                            int dim, float* __restrict__ s) {
  __shared__ float As[32][65];
  __shared__ float Bs[32][65];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t i0 = static_cast<int64_t>(blockIdx.y) * 64, j0 = static_cast<int64_t>(blockIdx.x) * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < dim; k0 += 32) {
    for (int idx = threadIdx.x; idx < 64 * 32; idx += 256) {
      const int r = idx >> 5, k = idx & 31;
      As[k][r] = (i0 + r < na && k0 + k < dim) ? a[(i0 + r) * dim + k0 + k] : 0.f;
      Bs[k][r] = (j0 + r < nb && k0 + k < dim) ? b[(j0 + r) * dim + k0 + k] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = As[k][ty + 16 * i]; bv[i] = Bs[k][tx + 16 * i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = i0 + ty + 16 * i;
    if (r >= na) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t c = j0 + tx + 16 * j;
      if (c < nb) s[r * nb + c] = acc[i][j] * ra[r] * rb[c];
    }
  }
}

struct Scratch {
  xv_ctx* c;
  std::vector<DevBuf> bufs;
  explicit Scratch(xv_ctx* ctx) : c(ctx) {}
  ~Scratch() {
    cudaStreamSynchronize(c->stream);
    for (DevBuf& b : bufs) release(&b);
  }
  void* get(size_t bytes, int* rc) {
    bufs.emplace_back();
    *rc = ensure(c, &bufs.back(), bytes ? bytes : 16);
    return bufs.back().p;
  }
};

}  // namespace
}  // namespace xv

using namespace xv;

extern "C" {

int xv_vec_normalize_length(xv_ctx* c, const float* x, int64_t n, int32_t dim, int scaleup, float* out) {
  if (!c) return XV_EINVAL;
  if (!x || !out || n < 0 || dim <= 0) return fail(c, XV_EINVAL, "xv_vec_normalize_length: bad argument");
  if (n == 0) return XV_OK;
  cudaSetDevice(c->device);
  Scratch sc(c);
  int rc;
  float* dx = static_cast<float*>(sc.get(sizeof(float) * n * dim, &rc));
  if (rc != XV_OK) return rc;
  cudaError_t e = cudaMemcpyAsync(dx, x, sizeof(float) * n * dim, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) {
    count_launch(c, "vec.normalize_length");
    normalize_length_kernel<<<static_cast<unsigned>((n * 32 + 255) / 256), 256, 0, c->stream>>>(dx, n, dim, scaleup ? 1 : 0, dx);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, dx, sizeof(float) * n * dim, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "xv_vec_normalize_length");
  return XV_OK;
}

int xv_vec_group_mean(xv_ctx* c, const float* x, int64_t n, int32_t dim, const int32_t* group, int32_t n_groups, float* out,
                      int32_t* counts) {
  if (!c) return XV_EINVAL;
  if (!x || !group || !out || n < 0 || dim <= 0 || n_groups < 0) return fail(c, XV_EINVAL, "xv_vec_group_mean: bad argument");
  if (n_groups == 0) return XV_OK;
  std::vector<int64_t> start(static_cast<size_t>(n_groups) + 1, 0), member;
  for (int64_t i = 0; i < n; ++i) {
    if (group[i] >= n_groups) return fail(c, XV_EINVAL, "xv_vec_group_mean: group %d of vector %lld is outside [0,%d)", group[i], (long long)i, n_groups);
    if (group[i] >= 0) ++start[group[i] + 1];
  }
  for (int g = 0; g < n_groups; ++g) start[g + 1] += start[g];
  member.resize(static_cast<size_t>(start[n_groups]));
  std::vector<int64_t> fill(start.begin(), start.end() - 1);
  for (int64_t i = 0; i < n; ++i)
    if (group[i] >= 0) member[static_cast<size_t>(fill[group[i]]++)] = i;        // members keep the input order (ivector-mean sums in spk2utt order)
  if (counts)
    for (int g = 0; g < n_groups; ++g) counts[g] = static_cast<int32_t>(start[g + 1] - start[g]);
  cudaSetDevice(c->device);
  Scratch sc(c);
  int rc;
  float* dx = static_cast<float*>(sc.get(sizeof(float) * n * dim, &rc));
  if (rc != XV_OK) return rc;
  int64_t* ds = static_cast<int64_t*>(sc.get(sizeof(int64_t) * start.size(), &rc));
  if (rc != XV_OK) return rc;
  int64_t* dm = static_cast<int64_t*>(sc.get(sizeof(int64_t) * member.size(), &rc));
  if (rc != XV_OK) return rc;
  float* dout = static_cast<float*>(sc.get(sizeof(float) * static_cast<size_t>(n_groups) * dim, &rc));
  if (rc != XV_OK) return rc;
  cudaError_t e = cudaSuccess;
  if (n > 0) e = cudaMemcpyAsync(dx, x, sizeof(float) * n * dim, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ds, start.data(), sizeof(int64_t) * start.size(), cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && !member.empty())
    e = cudaMemcpyAsync(dm, member.data(), sizeof(int64_t) * member.size(), cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) {
    count_launch(c, "vec.group_mean");
    group_mean_kernel<<<n_groups, 128, 0, c->stream>>>(dx, dim, ds, dm, dout);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout, sizeof(float) * static_cast<size_t>(n_groups) * dim, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);       // start / member are host vectors: drain before they go
  if (e != cudaSuccess) return cuda_fail(c, e, "xv_vec_group_mean");
  return XV_OK;
}

int xv_cosine_scores(xv_ctx* c, const float* a, int64_t na, const float* b, int64_t nb, int32_t dim, float eps, float* out) {
  if (!c) return XV_EINVAL;
  if (!a || !b || !out || na < 0 || nb < 0 || dim <= 0) return fail(c, XV_EINVAL, "xv_cosine_scores: bad argument");
  if (na == 0 || nb == 0) return XV_OK;
  cudaSetDevice(c->device);
  Scratch sc(c);
  int rc;
  float* da = static_cast<float*>(sc.get(sizeof(float) * na * dim, &rc));
  if (rc != XV_OK) return rc;
  float* db = static_cast<float*>(sc.get(sizeof(float) * nb * dim, &rc));
  if (rc != XV_OK) return rc;
  float* ra = static_cast<float*>(sc.get(sizeof(float) * na, &rc));
  if (rc != XV_OK) return rc;
  float* rb = static_cast<float*>(sc.get(sizeof(float) * nb, &rc));
  if (rc != XV_OK) return rc;
  float* ds = static_cast<float*>(sc.get(sizeof(float) * na * nb, &rc));
  if (rc != XV_OK) return rc;
  cudaError_t e = cudaMemcpyAsync(da, a, sizeof(float) * na * dim, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(db, b, sizeof(float) * nb * dim, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) {
    count_launch(c, "vec.cosine_scores");
    row_rnorm_kernel<<<static_cast<unsigned>((na * 32 + 255) / 256), 256, 0, c->stream>>>(da, na, dim, eps, ra);
    row_rnorm_kernel<<<static_cast<unsigned>((nb * 32 + 255) / 256), 256, 0, c->stream>>>(db, nb, dim, eps, rb);
    const dim3 grid(static_cast<unsigned>((nb + 63) / 64), static_cast<unsigned>((na + 63) / 64));
    cosine_scores_kernel<<<grid, 256, 0, c->stream>>>(da, ra, na, db, rb, nb, dim, ds);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, ds, sizeof(float) * na * nb, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "xv_cosine_scores");
  return XV_OK;
}

}  // extern "C"
