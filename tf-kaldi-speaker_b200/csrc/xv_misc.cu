// libxvector.so C-ABI, part 3: standalone pooling entry points, the raw GEMM entry point used by kernel
// tests and the roofline bench, CUDA-event timing, launch accounting, memory helpers.
#include <math.h>
#include <stdio.h>

#include <algorithm>
#include <vector>

#include "xv_ctx.h"

using namespace xv;

namespace {

int classify(const void* p) {   // 0 host, 2 device
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? 2 : 0;
}

// Brings a [rows, C] matrix of `dtype` to the device if the caller passed host memory (fp32 host data is
// converted to bf16 on the device when dtype == BF16 is requested for a host pointer holding bf16 bits).
int to_device(xv_ctx* c, const void* x, size_t bytes, DevBuf* tmp, const void** out) {
  if (classify(x) == 2) {
    *out = x;
    return XV_OK;
  }
  int rc = ensure(c, tmp, bytes);
  if (rc != XV_OK) return rc;
  cudaError_t e = cudaMemcpyAsync(tmp->p, x, bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "H2D copy");
  *out = tmp->p;
  return XV_OK;
}

}  // namespace

extern "C" {

int xv_stat_pooling_dev(xv_ctx* c, const void* x_dev, const int32_t* cu_lens_dev, int32_t n, int32_t C, int dtype,
                        int32_t max_len, float* out_dev) {
  if (!c) return XV_EINVAL;
  if (!x_dev || !cu_lens_dev || !out_dev || n < 0 || C <= 0 || (dtype != XV_DTYPE_F32 && dtype != XV_DTYPE_BF16))
    return fail(c, XV_EINVAL, "xv_stat_pooling_dev: bad argument");
  (void)max_len;
  cudaSetDevice(c->device);
  if (n == 0) return XV_OK;
  int rc = ensure(c, &c->pool_ws, pool_workspace_bytes(n, 1, C));
  if (rc != XV_OK) return rc;
  // total rows = cu[n]; read it back once (tiny, and only on this bench/test entry point)
  int32_t rows = 0;
  cudaError_t e = cudaMemcpyAsync(&rows, cu_lens_dev + n, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "read of cu_lens[n]");
  PoolArgs a;
  a.x = x_dev; a.dtype = dtype; a.c = C; a.ldx = C; a.rows = rows; a.cu = cu_lens_dev; a.n = n; a.context = 0;
  a.weights = nullptr; a.heads = 1; a.split_value = 1; a.acc = static_cast<float*>(c->pool_ws.p);
  a.out = out_dev; a.ldout = 2 * C; a.out_bf16 = nullptr; a.ldout_bf16 = 0;
  ProfScope ps(c, "pool");
  count_launch(c, "pool_finalize");
  e = launch_pool(a, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "pooling launch");
  return XV_OK;
}

int xv_stat_pooling(xv_ctx* c, const void* x, const int32_t* cu_lens, int32_t n, int32_t C, int dtype, float* out) {
  if (!c) return XV_EINVAL;
  if (!x || !cu_lens || !out || n < 0 || C <= 0 || (dtype != XV_DTYPE_F32 && dtype != XV_DTYPE_BF16))
    return fail(c, XV_EINVAL, "xv_stat_pooling: bad argument");
  cudaSetDevice(c->device);
  if (n == 0) return XV_OK;
  for (int u = 0; u < n; ++u)
    if (cu_lens[u + 1] <= cu_lens[u]) return fail(c, XV_EINVAL, "xv_stat_pooling: utterance %d is empty", u);
  const int64_t rows = cu_lens[n];
  const size_t esz = dtype == XV_DTYPE_BF16 ? 2 : 4;
  DevBuf tmp_x, tmp_cu, tmp_out;
  const void* xd = nullptr;
  int rc = to_device(c, x, static_cast<size_t>(rows) * C * esz, &tmp_x, &xd);
  if (rc == XV_OK) rc = ensure(c, &tmp_cu, sizeof(int32_t) * (n + 1));
  if (rc == XV_OK) rc = ensure(c, &tmp_out, sizeof(float) * n * 2 * C);
  if (rc == XV_OK) {
    cudaError_t e = cudaMemcpyAsync(tmp_cu.p, cu_lens, sizeof(int32_t) * (n + 1), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) rc = cuda_fail(c, e, "H2D of cu_lens");
  }
  if (rc == XV_OK)
    rc = xv_stat_pooling_dev(c, xd, static_cast<const int32_t*>(tmp_cu.p), n, C, dtype, 0, static_cast<float*>(tmp_out.p));
  if (rc == XV_OK) {
    cudaError_t e = cudaMemcpyAsync(out, tmp_out.p, sizeof(float) * n * 2 * C, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) rc = cuda_fail(c, e, "D2H of pooling result");
  }
  cudaStreamSynchronize(c->stream);
  release(&tmp_x); release(&tmp_cu); release(&tmp_out);
  return rc;
}

int xv_attention_pooling(xv_ctx* c, const void* value, const void* key, const float* query, const int32_t* cu_lens,
                         int32_t n, int32_t Cv, int32_t Ck, int32_t H, int split_value, int split_key, int use_scale,
                         int dtype, float* out, float* weights_out) {
  if (!c) return XV_EINVAL;
  if (!value || !key || !query || !cu_lens || !out || n < 0 || Cv <= 0 || Ck <= 0 || H <= 0 ||
      (dtype != XV_DTYPE_F32 && dtype != XV_DTYPE_BF16))
    return fail(c, XV_EINVAL, "xv_attention_pooling: bad argument");
  if (split_value && Cv % H) return fail(c, XV_EINVAL, "The dim of the value must be divided by the num of heads.");
  if (split_key && Ck % H) return fail(c, XV_EINVAL, "The dim of the key must be divided by the num of heads.");
  cudaSetDevice(c->device);
  if (n == 0) return XV_OK;
  for (int u = 0; u < n; ++u)
    if (cu_lens[u + 1] <= cu_lens[u]) return fail(c, XV_EINVAL, "xv_attention_pooling: utterance %d is empty", u);
  const int64_t rows = cu_lens[n];
  const size_t esz = dtype == XV_DTYPE_BF16 ? 2 : 4;
  const int dk = split_key ? Ck / H : Ck;
  const int dv = split_value ? Cv / H : Cv;
  const int hout = split_value ? 1 : H;
  const int width = 2 * H * dv;
  DevBuf tv, tk, tq, tcu, tout, tsc, tacc, toff;
  const void *vd = nullptr, *kd = nullptr;
  int rc = to_device(c, value, static_cast<size_t>(rows) * Cv * esz, &tv, &vd);
  if (rc == XV_OK) rc = to_device(c, key, static_cast<size_t>(rows) * Ck * esz, &tk, &kd);
  if (rc == XV_OK) rc = ensure(c, &tq, sizeof(float) * H * dk);
  if (rc == XV_OK) rc = ensure(c, &tcu, sizeof(int32_t) * (n + 1));
  if (rc == XV_OK) rc = ensure(c, &tout, sizeof(float) * static_cast<size_t>(n) * width);
  if (rc == XV_OK) rc = ensure(c, &tsc, sizeof(float) * static_cast<size_t>(H) * rows);
  if (rc == XV_OK) rc = ensure(c, &tacc, pool_workspace_bytes(n, hout, Cv));
  cudaError_t e = cudaSuccess;
  if (rc == XV_OK) {
    e = cudaMemcpyAsync(tq.p, query, sizeof(float) * H * dk, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(tcu.p, cu_lens, sizeof(int32_t) * (n + 1), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) rc = cuda_fail(c, e, "H2D of attention inputs");
  }
  if (rc == XV_OK) {
    const float scale = use_scale ? 1.0f / sqrtf(static_cast<float>(dk)) : 1.0f;
    count_launch(c, "att_scores");
    e = launch_att_scores(kd, dtype, Ck, rows, static_cast<const float*>(tq.p), H, dk, split_key ? 1 : 0, scale,
                          static_cast<float*>(tsc.p), c->stream);
    if (e == cudaSuccess) {
      count_launch(c, "att_softmax");
      e = launch_att_softmax(static_cast<float*>(tsc.p), rows, static_cast<const int32_t*>(tcu.p), 0, n, H, c->stream);
    }
    if (e == cudaSuccess) {
      PoolArgs a;
      a.x = vd; a.dtype = dtype; a.c = Cv; a.ldx = Cv; a.rows = rows; a.cu = static_cast<const int32_t*>(tcu.p); a.n = n;
      a.context = 0; a.weights = static_cast<const float*>(tsc.p); a.heads = H; a.split_value = split_value ? 1 : 0;
      a.acc = static_cast<float*>(tacc.p); a.out = static_cast<float*>(tout.p); a.ldout = width; a.out_bf16 = nullptr;
      a.ldout_bf16 = 0;
      count_launch(c, "pool");
      count_launch(c, "pool_finalize");
      e = launch_pool(a, c->sm_count, c->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, tout.p, sizeof(float) * static_cast<size_t>(n) * width, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && weights_out) {
      // packed per utterance: [H, L_u] at H * cu_lens[u]
      std::vector<int64_t> off(n);
      for (int u = 0; u < n; ++u) off[u] = static_cast<int64_t>(H) * cu_lens[u];
      rc = ensure(c, &toff, sizeof(int64_t) * n);
      DevBuf tw;
      if (rc == XV_OK) rc = ensure(c, &tw, sizeof(float) * static_cast<size_t>(H) * rows);
      if (rc == XV_OK) {
        e = cudaMemcpyAsync(toff.p, off.data(), sizeof(int64_t) * n, cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess) {
          count_launch(c, "gather");
          e = launch_gather_att_weights(static_cast<const float*>(tsc.p), rows, static_cast<const int32_t*>(tcu.p), 0, H, n,
                                        static_cast<const int64_t*>(toff.p), static_cast<float*>(tw.p), c->stream);
        }
        if (e == cudaSuccess)
          e = cudaMemcpyAsync(weights_out, tw.p, sizeof(float) * static_cast<size_t>(H) * rows, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
      }
      release(&tw);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess && rc == XV_OK) rc = cuda_fail(c, e, "attention pooling");
  }
  cudaStreamSynchronize(c->stream);
  DevBuf* all[] = {&tv, &tk, &tq, &tcu, &tout, &tsc, &tacc, &toff};
  for (DevBuf* b : all) release(b);
  return rc;
}

int xv_gemm_bf16(xv_ctx* c, const void* a_dev, int64_t a_rows, int32_t cin, int32_t taps, int32_t dilation,
                 const void* wt_dev, int32_t n, const float* scale_dev, const float* shift_dev, float neg_slope,
                 void* out_dev, int32_t ldo, int64_t m_rows) {
  if (!c) return XV_EINVAL;
  if (!a_dev || !wt_dev || !scale_dev || !shift_dev || !out_dev) return fail(c, XV_EINVAL, "xv_gemm_bf16: null argument");
  if (cin % kBK || n % kBN || ldo != n || taps < 1 || dilation == 0 || a_rows < kBM || m_rows > a_rows)
    return fail(c, XV_EINVAL,
                "xv_gemm_bf16: needs cin %% 64 == 0, n %% 256 == 0, ldo == n, a_rows >= 128 (allocation), m_rows <= a_rows");
  cudaSetDevice(c->device);
  const int kpad = taps * cin;
  GemmPlan g;
  if (m_rows <= 0) return XV_OK;
  const char* err = make_tmap_bf16_2d(&g.tm_a, a_dev, cin, a_rows, static_cast<uint64_t>(cin) * 2, kBK, kBM);
  if (!err) err = make_tmap_bf16_2d(&g.tm_b, wt_dev, kpad, n, static_cast<uint64_t>(kpad) * 2, kBK, kBLoadRows);
  if (!err) err = make_tmap_bf16_2d(&g.tm_out, out_dev, n, m_rows, static_cast<uint64_t>(n) * 2, 32, 32, 64);
  if (err) return fail(c, XV_ECUDA, "%s", err);
  DevBuf slope;
  int rc = ensure(c, &slope, sizeof(float) * n);
  if (rc != XV_OK) return rc;
  uint32_t bits;
  memcpy(&bits, &neg_slope, 4);
  cudaError_t e = launch_fill_u32(static_cast<uint32_t*>(slope.p), n, bits, c->stream);
  if (e != cudaSuccess) { release(&slope); return cuda_fail(c, e, "fill launch"); }
  GemmParams& p = g.p;
  memset(&p, 0, sizeof(p));
  p.m_rows = static_cast<int32_t>(m_rows);
  p.m_tiles = static_cast<int32_t>((m_rows + kBM - 1) / kBM);
  p.n_tiles = n / kBN;
  p.taps = taps; p.dilation = dilation; p.kb_per_tap = cin / kBK; p.cin = cin;
  p.scale = scale_dev; p.shift = shift_dev; p.slope = static_cast<const float*>(slope.p);
  p.relu_only = neg_slope == 0.f ? 1 : 0;
  p.out_bf16 = static_cast<__nv_bfloat16*>(out_dev); p.ldo = ldo; p.heads = 1; p.head_width = n;
  g.pool = false;
  {
    ProfScope ps(c, "gemm");
    e = launch_gemm(g, c->sm_count, c->stream);
  }
  // the slope buffer must outlive the kernel: free it after the stream drains
  cudaError_t e2 = cudaStreamSynchronize(c->stream);
  release(&slope);
  if (e != cudaSuccess) return cuda_fail(c, e, "tcgen05 GEMM launch");
  if (e2 != cudaSuccess) return cuda_fail(c, e2, "tcgen05 GEMM execution");
  return XV_OK;
}

int xv_gemm_bf16_async(xv_ctx* c, const void* a_dev, int64_t a_rows, int32_t cin, int32_t taps, int32_t dilation,
                       const void* wt_dev, int32_t n, const float* scale_dev, const float* shift_dev,
                       const float* slope_dev, void* out_dev, int64_t m_rows) {
  // Bench variant: caller provides the slope array, nothing is synchronised.
  if (!c) return XV_EINVAL;
  if (!a_dev || !wt_dev || !scale_dev || !shift_dev || !slope_dev || !out_dev)
    return fail(c, XV_EINVAL, "xv_gemm_bf16_async: null argument");
  if (cin % kBK || n % kBN || taps < 1 || dilation == 0 || a_rows < kBM || m_rows > a_rows)
    return fail(c, XV_EINVAL, "xv_gemm_bf16_async: bad geometry");
  const int kpad = taps * cin;
  GemmPlan g;
  if (m_rows <= 0) return XV_OK;
  const char* err = make_tmap_bf16_2d(&g.tm_a, a_dev, cin, a_rows, static_cast<uint64_t>(cin) * 2, kBK, kBM);
  if (!err) err = make_tmap_bf16_2d(&g.tm_b, wt_dev, kpad, n, static_cast<uint64_t>(kpad) * 2, kBK, kBLoadRows);
  if (!err) err = make_tmap_bf16_2d(&g.tm_out, out_dev, n, m_rows, static_cast<uint64_t>(n) * 2, 32, 32, 64);
  if (err) return fail(c, XV_ECUDA, "%s", err);
  GemmParams& p = g.p;
  memset(&p, 0, sizeof(p));
  p.m_rows = static_cast<int32_t>(m_rows);
  p.m_tiles = static_cast<int32_t>((m_rows + kBM - 1) / kBM);
  p.n_tiles = n / kBN;
  p.taps = taps; p.dilation = dilation; p.kb_per_tap = cin / kBK; p.cin = cin;
  p.scale = scale_dev; p.shift = shift_dev; p.slope = slope_dev;
  p.out_bf16 = static_cast<__nv_bfloat16*>(out_dev); p.ldo = n; p.heads = 1; p.head_width = n;
  g.pool = false;
  ProfScope ps(c, "gemm");
  cudaError_t e = launch_gemm(g, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "tcgen05 GEMM launch");
  return XV_OK;
}

int xv_dense_rows_f32(xv_ctx* c, const float* x, int64_t rows, int32_t cin, int32_t taps, const float* w, int32_t n,
                      const float* bias, const float* bn_scale, const float* bn_shift, const float* slope, int act_tanh,
                      float* out) {
  if (!c) return XV_EINVAL;
  if (!x || !w || !bias || !out || rows < taps || cin <= 0 || taps < 1 || n <= 0 || (bn_scale && !bn_shift))
    return fail(c, XV_EINVAL, "xv_dense_rows_f32: bad argument");
  cudaSetDevice(c->device);
  const int64_t m_rows = rows - taps + 1;
  const size_t K = static_cast<size_t>(taps) * cin;
  DevBuf dx, dw, dv, dout;
  int rc = ensure(c, &dx, sizeof(float) * rows * cin);
  if (rc == XV_OK) rc = ensure(c, &dw, sizeof(float) * K * n);
  if (rc == XV_OK) rc = ensure(c, &dv, sizeof(float) * 4 * n);
  if (rc == XV_OK) rc = ensure(c, &dout, sizeof(float) * m_rows * n);
  cudaError_t e = cudaSuccess;
  if (rc == XV_OK) {
    float* v = static_cast<float*>(dv.p);
    e = cudaMemcpyAsync(dx.p, x, sizeof(float) * rows * cin, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dw.p, w, sizeof(float) * K * n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(v, bias, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && bn_scale) e = cudaMemcpyAsync(v + n, bn_scale, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && bn_scale) e = cudaMemcpyAsync(v + 2 * n, bn_shift, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && slope) e = cudaMemcpyAsync(v + 3 * n, slope, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
      SimtGemmArgs s;
      s.a = static_cast<const float*>(dx.p); s.a_rows = rows; s.cin = cin; s.taps = taps; s.dilation = 1;
      s.w = static_cast<const float*>(dw.p); s.n = n; s.bias = v;
      s.bn_scale = bn_scale ? v + n : nullptr; s.bn_shift = v + 2 * n; s.slope = slope ? v + 3 * n : nullptr;
      s.act_tanh = act_tanh; s.out_dense = nullptr; s.out_bn = nullptr; s.out_act = static_cast<float*>(dout.p);
      s.m_rows = m_rows;
      count_launch(c, "simt_gemm");
      e = launch_simt_gemm(s, c->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout.p, sizeof(float) * m_rows * n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) rc = cuda_fail(c, e, "xv_dense_rows_f32");
  }
  cudaStreamSynchronize(c->stream);
  release(&dx); release(&dw); release(&dv); release(&dout);
  return rc;
}

int xv_prelu_f32(xv_ctx* c, const float* x, int64_t rows, int32_t width, const float* alpha, int shared, float* out) {
  if (!c) return XV_EINVAL;
  if (!x || !alpha || !out || rows <= 0 || width <= 0 || rows > 0x7fffffff) return fail(c, XV_EINVAL, "xv_prelu_f32: bad argument");
  cudaSetDevice(c->device);
  DevBuf dx, da;
  int rc = ensure(c, &dx, sizeof(float) * rows * width);
  if (rc == XV_OK) rc = ensure(c, &da, sizeof(float) * width);
  if (rc == XV_OK) {
    std::vector<float> al(width);
    for (int i = 0; i < width; ++i) al[i] = shared ? alpha[0] : alpha[i];
    cudaError_t e = cudaMemcpyAsync(dx.p, x, sizeof(float) * rows * width, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(da.p, al.data(), sizeof(float) * width, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) {
      count_launch(c, "affine");
      e = launch_affine_rows(static_cast<const float*>(dx.p), width, static_cast<int>(rows), width, nullptr, nullptr,
                             static_cast<const float*>(da.p), static_cast<float*>(dx.p), width, nullptr, 0, c->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dx.p, sizeof(float) * rows * width, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) rc = cuda_fail(c, e, "xv_prelu_f32");
  }
  cudaStreamSynchronize(c->stream);
  release(&dx); release(&da);
  return rc;
}

int xv_l2_scaling_f32(xv_ctx* c, const float* x, int64_t rows, int32_t width, float factor, float eps, float* out) {
  if (!c) return XV_EINVAL;
  if (!x || !out || rows <= 0 || width <= 0 || rows > 0x7fffffff) return fail(c, XV_EINVAL, "xv_l2_scaling_f32: bad argument");
  cudaSetDevice(c->device);
  DevBuf dx;
  int rc = ensure(c, &dx, sizeof(float) * rows * width);
  if (rc == XV_OK) {
    cudaError_t e = cudaMemcpyAsync(dx.p, x, sizeof(float) * rows * width, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
      count_launch(c, "l2_scale");
      e = launch_l2_scale_rows(static_cast<float*>(dx.p), width, static_cast<int>(rows), width, factor, eps, c->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dx.p, sizeof(float) * rows * width, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) rc = cuda_fail(c, e, "xv_l2_scaling_f32");
  }
  cudaStreamSynchronize(c->stream);
  release(&dx);
  return rc;
}

// ---- measurement ---------------------------------------------------------------------------------
int xv_event_record(xv_ctx* c, int slot) {
  if (!c) return XV_EINVAL;
  if (slot < 0 || slot >= kNumEvents) return fail(c, XV_EINVAL, "event slot out of range");
  cudaError_t e = cudaEventRecord(c->events[slot], c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaEventRecord");
  return XV_OK;
}

int xv_event_elapsed_ms(xv_ctx* c, int a, int b, float* ms) {
  if (!c) return XV_EINVAL;
  if (a < 0 || a >= kNumEvents || b < 0 || b >= kNumEvents || !ms) return fail(c, XV_EINVAL, "bad event arguments");
  cudaError_t e = cudaEventSynchronize(c->events[b]);
  if (e == cudaSuccess) e = cudaEventElapsedTime(ms, c->events[a], c->events[b]);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaEventElapsedTime");
  return XV_OK;
}

int64_t xv_launch_count(xv_ctx* c) { return c ? c->launches : 0; }

int xv_launch_summary(xv_ctx* c, char* buf, size_t buf_size) {
  if (!c || !buf || buf_size == 0) return XV_EINVAL;
  std::string s;
  for (const auto& kv : c->launch_by_family) {
    char tmp[128];
    snprintf(tmp, sizeof(tmp), "%s=%lld;", kv.first.c_str(), static_cast<long long>(kv.second));
    s += tmp;
  }
  snprintf(buf, buf_size, "%s", s.c_str());
  return XV_OK;
}

int xv_profile_enable(xv_ctx* c, int on) {
  if (!c) return XV_EINVAL;
  cudaStreamSynchronize(c->stream);
  for (auto& s : c->prof) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  c->prof.clear();
  c->profiling = on != 0;
  return XV_OK;
}

int xv_profile_query(xv_ctx* c, const char* family, float* total_ms, int64_t* launches) {
  if (!c || !family) return XV_EINVAL;
  cudaError_t e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "stream synchronize");
  int fam = -1;
  for (size_t i = 0; i < c->prof_family_names.size(); ++i)
    if (c->prof_family_names[i] == family) fam = static_cast<int>(i);
  float tot = 0.f;
  int64_t cnt = 0;
  for (const auto& s : c->prof) {
    if (s.family != fam) continue;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) { tot += ms; ++cnt; }
  }
  if (total_ms) *total_ms = tot;
  if (launches) *launches = cnt;
  return XV_OK;
}

// ---- memory helpers -------------------------------------------------------------------------------
void* xv_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 16) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void xv_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
void* xv_device_alloc(xv_ctx* c, size_t bytes) {
  if (!c) return nullptr;
  cudaSetDevice(c->device);
  void* p = nullptr;
  if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) {
    fail(c, XV_ENOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return p;
}
void xv_device_free(xv_ctx* c, void* p) {
  if (!c || !p) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  cudaFree(p);
}
int xv_memcpy_h2d(xv_ctx* c, void* dst, const void* src, size_t bytes) {
  if (!c) return XV_EINVAL;
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaMemcpy H2D");
  return XV_OK;
}
int xv_memcpy_d2h(xv_ctx* c, void* dst, const void* src, size_t bytes) {
  if (!c) return XV_EINVAL;
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "cudaMemcpy D2H");
  return XV_OK;
}
int xv_upload_bf16(xv_ctx* c, void* dst_dev, const float* src_host, size_t count) {
  if (!c) return XV_EINVAL;
  DevBuf tmp;
  int rc = ensure(c, &tmp, count * 4);
  if (rc != XV_OK) return rc;
  cudaError_t e = cudaMemcpyAsync(tmp.p, src_host, count * 4, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) {
    count_launch(c, "cast");
    e = launch_f32_to_bf16(static_cast<const float*>(tmp.p), static_cast<__nv_bfloat16*>(dst_dev), count, c->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  release(&tmp);
  if (e != cudaSuccess) return cuda_fail(c, e, "xv_upload_bf16");
  return XV_OK;
}
int xv_flush_l2(xv_ctx* c) {
  if (!c) return XV_EINVAL;
  const size_t bytes = size_t(256) << 20;   // 256 MiB > 126 MB L2
  int rc = ensure(c, &c->flush_buf, bytes);
  if (rc != XV_OK) return rc;
  cudaError_t e = launch_fill_u32(static_cast<uint32_t*>(c->flush_buf.p), bytes / 4, 0u, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "L2 flush launch");
  return XV_OK;
}

}  // extern "C"
