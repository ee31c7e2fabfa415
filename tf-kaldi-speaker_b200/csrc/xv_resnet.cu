// resnet_18 front end (model/resnet.py:152-251): conv0_1 stem, optional 3x3 max pooling, four stages of
// conv_block / identity_block.  Plans the layer sequence from the nnet_conf JSON, folds BN, and runs one batch
// of equal-length utterances through the implicit-GEMM kernel of xv_conv_sm100.cu (bf16 mode) or a plain SIMT
// fp32 convolution (validation mode).  conv5 / dense1 / dense2 / pooling / tdnn6 / tdnn7 are ordinary frame and
// segment layers and live in xv_api.cu / xv_forward.cu.
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "xv_ctx.h"

namespace xv {

namespace {

// TensorFlow 'SAME': out = ceil(in/s); total = max((out-1)*s + k - in, 0); before = total/2
inline void same_pads(int n_in, int k, int s, int* n_out, int* before) {
  *n_out = (n_in + s - 1) / s;
  const int total = std::max((*n_out - 1) * s + k - n_in, 0);
  *before = total / 2;
}
inline int floor_div(int a, int b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }

template <typename T>
__device__ __forceinline__ T cvt_out(float v);
template <>
__device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// ---- stem: conv0_1 (3x3, Cin = 1, SAME, no bias) + BN + activation, fp32 features -> NHWC [B,T,F,64] ----
// One thread = one (b,t,f) position x 8 output channels; the 9 taps x 64 weights sit in shared memory.
template <typename T>
__global__ void __launch_bounds__(256) resnet_stem_kernel(const float* __restrict__ x, int nb, int t, int f,
                                                          const float* __restrict__ w /*[9][64]*/,
                                                          const float* __restrict__ scale, const float* __restrict__ shift,
                                                          const float* __restrict__ slope, T* __restrict__ out,
                                                          int64_t out_utt_stride) {
  __shared__ __align__(16) float sw[9 * 64];
  __shared__ __align__(16) float ss[3 * 64];
  for (int i = threadIdx.x; i < 9 * 64; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < 64; i += blockDim.x) { ss[i] = scale[i]; ss[64 + i] = shift[i]; ss[128 + i] = slope[i]; }
  __syncthreads();
  // 32-bit index arithmetic (a batch holds < 2^28 positions), weights and BN constants read as 16-byte shared loads:
  // the kernel is bound by instruction issue, not by its 128 bytes of output per position
  const uint32_t n_pos = static_cast<uint32_t>(nb) * static_cast<uint32_t>(t) * static_cast<uint32_t>(f);
  const uint32_t total = n_pos * 8u;
  const uint32_t cg = threadIdx.x & 7u;                       // blockDim.x and the grid stride are multiples of 8
  const float4* w4 = reinterpret_cast<const float4*>(sw) + cg * 2;
  const float4 sc0 = reinterpret_cast<const float4*>(ss)[cg * 2], sc1 = reinterpret_cast<const float4*>(ss)[cg * 2 + 1];
  const float4 sh0 = reinterpret_cast<const float4*>(ss + 64)[cg * 2], sh1 = reinterpret_cast<const float4*>(ss + 64)[cg * 2 + 1];
  const float4 sl0 = reinterpret_cast<const float4*>(ss + 128)[cg * 2], sl1 = reinterpret_cast<const float4*>(ss + 128)[cg * 2 + 1];
  const float scv[8] = {sc0.x, sc0.y, sc0.z, sc0.w, sc1.x, sc1.y, sc1.z, sc1.w};
  const float shv[8] = {sh0.x, sh0.y, sh0.z, sh0.w, sh1.x, sh1.y, sh1.z, sh1.w};
  const float slv[8] = {sl0.x, sl0.y, sl0.z, sl0.w, sl1.x, sl1.y, sl1.z, sl1.w};
  const uint32_t uf = static_cast<uint32_t>(f), ut = static_cast<uint32_t>(t);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t pos = i >> 3;
    const uint32_t bt = pos / uf;
    const int fi = static_cast<int>(pos - bt * uf);
    const uint32_t b = bt / ut;
    const int ti = static_cast<int>(bt - b * ut);
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    const float* xb = x + static_cast<int64_t>(b) * t * f;
#pragma unroll
    for (int dt = 0; dt < 3; ++dt) {
      const int tt = ti + dt - 1;
      const bool tok = tt >= 0 && tt < t;
#pragma unroll
      for (int df = 0; df < 3; ++df) {
        const int ff = fi + df - 1;
        const float v = (tok && ff >= 0 && ff < f) ? __ldg(xb + tt * f + ff) : 0.f;
        const float4 wa = w4[(dt * 3 + df) * 16], wb = w4[(dt * 3 + df) * 16 + 1];
        acc[0] = fmaf(v, wa.x, acc[0]); acc[1] = fmaf(v, wa.y, acc[1]); acc[2] = fmaf(v, wa.z, acc[2]); acc[3] = fmaf(v, wa.w, acc[3]);
        acc[4] = fmaf(v, wb.x, acc[4]); acc[5] = fmaf(v, wb.y, acc[5]); acc[6] = fmaf(v, wb.z, acc[6]); acc[7] = fmaf(v, wb.w, acc[7]);
      }
    }
    T* dst = out + static_cast<int64_t>(b) * out_utt_stride + (static_cast<int64_t>(ti) * f + fi) * 64 + cg * 8;
    alignas(16) T y[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v = fmaf(acc[k], scv[k], shv[k]);
      v = v >= 0.f ? v : v * slv[k];
      y[k] = cvt_out<T>(v);
    }
    // 8 channels = one 16-byte (bf16) or two 16-byte (fp32) stores; consecutive threads write consecutive chunks
    const uint4* src = reinterpret_cast<const uint4*>(y);
#pragma unroll
    for (int k = 0; k < static_cast<int>(sizeof(T) * 8 / 16); ++k) reinterpret_cast<uint4*>(dst)[k] = src[k];
  }
}

// ---- tf.layers.max_pooling2d(x, (3,3), (1,1), 'same') on NHWC (model/resnet.py:230-231) ----
template <typename T>
__global__ void __launch_bounds__(256) maxpool3x3_kernel(const T* __restrict__ x, int nb, int t, int f, int c,
                                                         int64_t in_utt_stride, T* __restrict__ out, int64_t out_utt_stride) {
  const int64_t total = static_cast<int64_t>(nb) * t * f * c;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int ch = static_cast<int>(i % c);
    const int64_t pos = i / c;
    const int fi = static_cast<int>(pos % f);
    const int64_t bt = pos / f;
    const int ti = static_cast<int>(bt % t);
    const int b = static_cast<int>(bt / t);
    float m = -INFINITY;
    for (int dt = -1; dt <= 1; ++dt) {
      const int tt = ti + dt;
      if (tt < 0 || tt >= t) continue;
      for (int df = -1; df <= 1; ++df) {
        const int ff = fi + df;
        if (ff < 0 || ff >= f) continue;
        m = fmaxf(m, static_cast<float>(x[static_cast<int64_t>(b) * in_utt_stride + (static_cast<int64_t>(tt) * f + ff) * c + ch]));
      }
    }
    out[static_cast<int64_t>(b) * out_utt_stride + (static_cast<int64_t>(ti) * f + fi) * c + ch] = cvt_out<T>(m);
  }
}

// ---- fp32 validation convolution: direct NHWC conv, 64 positions x 64 channels per block, K chunks of 16 ----
struct SimtConvArgs {
  const float* x; int nb, t_in, f_in, cin; int64_t in_utt_stride;
  const float* w;  // [kh*kw*cin][cout]
  int kh, kw, st, sf, pt, pf;
  int t_out, f_out, cout; int64_t out_utt_stride;
  const float* scale; const float* shift; const float* slope;
  const float* res; int64_t res_utt_stride;
  float* out;
};

__global__ void __launch_bounds__(256) simt_conv_kernel(const SimtConvArgs a) {
  __shared__ float sa[16][65];
  __shared__ float sb[16][64];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t m0 = static_cast<int64_t>(blockIdx.x) * 64;
  const int n0 = blockIdx.y * 64;
  const int64_t m_total = static_cast<int64_t>(a.nb) * a.t_out * a.f_out;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int k_total = a.kh * a.kw * a.cin;
  for (int k0 = 0; k0 < k_total; k0 += 16) {
    // A tile: 64 positions x 16 k
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      const int mm = i >> 4, kk = i & 15;
      const int64_t m = m0 + mm;
      const int k = k0 + kk;
      float v = 0.f;
      if (m < m_total && k < k_total) {
        const int tap = k / a.cin, ci = k - tap * a.cin;
        const int dt = tap / a.kw, df = tap - dt * a.kw;
        const int fo = static_cast<int>(m % a.f_out);
        const int64_t bt = m / a.f_out;
        const int to = static_cast<int>(bt % a.t_out);
        const int b = static_cast<int>(bt / a.t_out);
        const int ti = to * a.st + dt - a.pt, fi = fo * a.sf + df - a.pf;
        if (ti >= 0 && ti < a.t_in && fi >= 0 && fi < a.f_in)
          v = a.x[static_cast<int64_t>(b) * a.in_utt_stride + (static_cast<int64_t>(ti) * a.f_in + fi) * a.cin + ci];
      }
      sa[kk][mm] = v;
    }
    for (int i = threadIdx.x; i < 16 * 64; i += 256) {
      const int kk = i >> 6, nn = i & 63;
      const int k = k0 + kk, n = n0 + nn;
      sb[kk][nn] = (k < k_total && n < a.cout) ? a.w[static_cast<int64_t>(k) * a.cout + n] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = sa[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = sb[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= m_total) continue;
    const int fo = static_cast<int>(m % a.f_out);
    const int64_t bt = m / a.f_out;
    const int to = static_cast<int>(bt % a.t_out);
    const int b = static_cast<int>(bt / a.t_out);
    const int64_t pos = (static_cast<int64_t>(to) * a.f_out + fo) * a.cout;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= a.cout) continue;
      float v = fmaf(acc[i][j], a.scale[n], a.shift[n]);
      if (a.res) v += a.res[static_cast<int64_t>(b) * a.res_utt_stride + pos + n];
      v = v >= 0.f ? v : v * a.slope[n];
      a.out[static_cast<int64_t>(b) * a.out_utt_stride + pos + n] = v;
    }
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ plan
static bool cfg_bool(const JValue& cfg, const char* key, bool dflt) {
  const JValue* v = cfg.get(key);
  return v ? v->truthy() : dflt;
}

int resnet_out_len(const xv_ctx* c, int t) {
  if (!c->resnet || !c->resnet_time_stride) return t;
  for (int i = 0; i < 3; ++i) t = (t + 1) / 2;
  return t;
}

static int new_slot(xv_ctx* c, int buf, int f, int ch, int tshift) {
  ActSlot s;
  s.buf = buf; s.f = f; s.c = ch; s.tshift = tshift;
  c->slots.push_back(s);
  return static_cast<int>(c->slots.size()) - 1;
}

static int free_buf(const xv_ctx* c, int a, int b, int d) {
  for (int i = 0; i < 4; ++i) {
    const bool used = (a >= 0 && c->slots[a].buf == i) || (b >= 0 && c->slots[b].buf == i) || (d >= 0 && c->slots[d].buf == i);
    if (!used) return i;
  }
  return 3;
}

static ConvLayer make_conv(const std::string& name, const std::string& bn, const std::string& relu, int k, int st, int sf,
                           int cin, int cout, bool act, int in_slot, int out_slot, int res_slot) {
  ConvLayer l;
  l.name = name;
  l.kernel_var = "resnet_18/" + name + "/kernel";
  l.bn_scope = "resnet_18/" + bn;
  l.alpha_var = relu.empty() ? "" : "resnet_18/" + relu + "/alpha";
  l.kh = l.kw = k; l.st = st; l.sf = sf; l.cin = cin; l.cout = cout; l.act = act;
  l.in_slot = in_slot; l.out_slot = out_slot; l.res_slot = res_slot;
  l.bn_tile = cout >= 256 ? 256 : cout;
  return l;
}

int resnet_build_plan(xv_ctx* c) {
  const JValue& cfg = c->cfg;
  // model/resnet.py:189-190 asserts a 40-dim input; BASELINE.json config 4 also names 64-dim fbank, so any
  // multiple of 8 (three frequency halvings) is accepted here.  DESIGN.md records the relaxation.
  if (c->dim % 8 != 0 || c->dim < 8)
    return fail(c, XV_EINVAL, "resnet_18 needs a feature_dim that is a multiple of 8 (reference: 40), got %d", c->dim);
  if (!cfg.has("resnet_time_stride") || !cfg.has("resnet_maxpooling"))
    return fail(c, XV_EINVAL, "resnet_18 needs config keys 'resnet_time_stride' and 'resnet_maxpooling'");   // params.* read at model/resnet.py:187,230
  c->resnet_time_stride = cfg_bool(cfg, "resnet_time_stride", false);
  c->resnet_maxpool = cfg_bool(cfg, "resnet_maxpooling", false);
  const JValue* blocks = cfg.get("resnet_blocks");             // default [2,2,2,2] (model/resnet.py:203-204)
  if (blocks && blocks->type == JValue::Arr) {
    if (blocks->arr.size() != 4) return fail(c, XV_EINVAL, "resnet_blocks must have 4 entries");
    for (int i = 0; i < 4; ++i) {
      c->resnet_blocks[i] = static_cast<int>(blocks->arr[i].num);
      if (c->resnet_blocks[i] < 1) return fail(c, XV_EINVAL, "resnet_blocks entries must be >= 1");
    }
  }
  const int ts = c->resnet_time_stride ? 2 : 1;
  c->slots.clear();
  c->convs.clear();
  int f = c->dim, tshift = 0;
  c->resnet_conv0_slot = new_slot(c, 0, f, 64, 0);
  int cur = c->resnet_conv0_slot;
  c->resnet_pool_slot = -1;
  if (c->resnet_maxpool) {
    c->resnet_pool_slot = new_slot(c, 1, f, 64, 0);
    cur = c->resnet_pool_slot;
  }
  int cin = 64;
  const int filters[4] = {64, 128, 256, 512};
  for (int stage = 0; stage < 4; ++stage) {
    const int filt = filters[stage];
    const int st = stage == 0 ? 1 : ts, sf = stage == 0 ? 1 : 2;
    char nm[32];
    snprintf(nm, sizeof(nm), "conv%da", stage + 1);
    const std::string a(nm);
    if (st == 2) c->slots[cur].pad_even = true;
    const int f2 = f / sf, tshift2 = tshift + (st == 2 ? 1 : 0);
    // conv_block (model/resnet.py:9-87)
    const int mid = new_slot(c, free_buf(c, cur, -1, -1), f2, filt, tshift2);
    const int sc = new_slot(c, free_buf(c, cur, mid, -1), f2, filt, tshift2);
    const int out = new_slot(c, free_buf(c, cur, mid, sc), f2, filt, tshift2);
    c->convs.push_back(make_conv(a + "_conv0", a + "_bn0", a + "_relu0", 3, st, sf, cin, filt, true, cur, mid, -1));
    c->convs.push_back(make_conv(a + "_conv_short", a + "_bn_short", "", 1, st, sf, cin, filt, false, cur, sc, -1));
    c->convs.push_back(make_conv(a + "_conv1", a + "_bn1", a + "_relu_final", 3, 1, 1, filt, filt, true, mid, out, sc));
    cur = out; f = f2; tshift = tshift2; cin = filt;
    for (int i = 0; i < c->resnet_blocks[stage] - 1; ++i) {
      // identity_block (model/resnet.py:90-149)
      snprintf(nm, sizeof(nm), "conv%db_%d", stage + 1, i);
      const std::string bname(nm);
      const int mid2 = new_slot(c, free_buf(c, cur, -1, -1), f, filt, tshift);
      const int out2 = new_slot(c, free_buf(c, cur, mid2, -1), f, filt, tshift);
      c->convs.push_back(make_conv(bname + "_conv0", bname + "_bn0", bname + "_relu0", 3, 1, 1, filt, filt, true, cur, mid2, -1));
      c->convs.push_back(make_conv(bname + "_conv1", bname + "_bn1", bname + "_relu_final", 3, 1, 1, filt, filt, true, mid2, out2, cur));
      cur = out2;
    }
  }
  c->resnet_out_slot = cur;
  c->resnet_f_out = f;
  return XV_OK;
}

// ------------------------------------------------------------------------------------------------ weights
static int slopes_for(xv_ctx* c, bool act, const std::string& alpha_var, int n, std::vector<float>* slope, bool* plain) {
  slope->assign(n, 1.f);
  *plain = false;
  if (!act) return XV_OK;
  if (c->relu_type == "prelu") {
    const HostTensor* a = find_weight(c, alpha_var);
    if (!a) return fail(c, XV_ENOTFOUND, "missing variable '%s'", alpha_var.c_str());
    if (static_cast<int>(a->data.size()) != n) return fail(c, XV_EINVAL, "variable '%s' must have %d elements", alpha_var.c_str(), n);
    for (int i = 0; i < n; ++i) (*slope)[i] = a->data[i];
  } else if (c->relu_type == "lrelu") {
    slope->assign(n, kLreluAlpha);
  } else {
    slope->assign(n, 0.f);
    *plain = true;
  }
  return XV_OK;
}

int resnet_finalize(xv_ctx* c) {
  int rc;
  {
    // stem: conv0_1 [3,3,1,64] + conv0_bn + conv0_relu
    const HostTensor* k = find_weight(c, "resnet_18/conv0_1/kernel");
    if (!k) return fail(c, XV_ENOTFOUND, "missing variable 'resnet_18/conv0_1/kernel'");
    if (k->shape.size() != 4 || k->shape[0] != 3 || k->shape[1] != 3 || k->shape[2] != 1 || k->shape[3] != 64)
      return fail(c, XV_EINVAL, "variable 'resnet_18/conv0_1/kernel' must be [3,3,1,64]");
    std::vector<double> s, h;
    if ((rc = fold_bn(c, "resnet_18/conv0_bn", 64, &s, &h)) != XV_OK) return rc;
    std::vector<float> sf(64), hf(64), sl;
    for (int i = 0; i < 64; ++i) { sf[i] = static_cast<float>(s[i]); hf[i] = static_cast<float>(h[i]); }
    bool plain;
    if ((rc = slopes_for(c, true, "resnet_18/conv0_relu/alpha", 64, &sl, &plain)) != XV_OK) return rc;
    if ((rc = upload(c, &c->conv0_w, k->data.data(), sizeof(float) * 9 * 64)) != XV_OK) return rc;
    if ((rc = upload(c, &c->conv0_scale, sf.data(), 256)) != XV_OK) return rc;
    if ((rc = upload(c, &c->conv0_shift, hf.data(), 256)) != XV_OK) return rc;
    if ((rc = upload(c, &c->conv0_slope, sl.data(), 256)) != XV_OK) return rc;
  }
  for (ConvLayer& l : c->convs) {
    const HostTensor* k = find_weight(c, l.kernel_var);
    if (!k) return fail(c, XV_ENOTFOUND, "missing variable '%s'", l.kernel_var.c_str());
    if (k->shape.size() != 4 || k->shape[0] != l.kh || k->shape[1] != l.kw || k->shape[2] != l.cin || k->shape[3] != l.cout)
      return fail(c, XV_EINVAL, "variable '%s' must be [%d,%d,%d,%d]", l.kernel_var.c_str(), l.kh, l.kw, l.cin, l.cout);
    std::vector<double> s, h;
    if ((rc = fold_bn(c, l.bn_scope, l.cout, &s, &h)) != XV_OK) return rc;
    std::vector<float> sf(l.cout), hf(l.cout), sl;
    for (int i = 0; i < l.cout; ++i) { sf[i] = static_cast<float>(s[i]); hf[i] = static_cast<float>(h[i]); }
    if ((rc = slopes_for(c, l.act, l.alpha_var, l.cout, &sl, &l.plain_relu)) != XV_OK) return rc;
    if ((rc = upload(c, &l.scale, sf.data(), sizeof(float) * l.cout)) != XV_OK) return rc;
    if ((rc = upload(c, &l.shift, hf.data(), sizeof(float) * l.cout)) != XV_OK) return rc;
    if ((rc = upload(c, &l.slope, sl.data(), sizeof(float) * l.cout)) != XV_OK) return rc;
    const int K = l.kh * l.kw * l.cin;
    if (c->bf16) {
      std::vector<uint16_t> wt(static_cast<size_t>(l.cout) * K);
      for (int kk = 0; kk < K; ++kk) {
        const float* src = k->data.data() + static_cast<size_t>(kk) * l.cout;
        for (int n = 0; n < l.cout; ++n) wt[static_cast<size_t>(n) * K + kk] = f32_to_bf16_rne(src[n]);
      }
      if ((rc = upload(c, &l.wt, wt.data(), wt.size() * 2)) != XV_OK) return rc;
      const char* e = make_tmap_bf16_2d(&l.tm_b, l.wt.p, K, l.cout, static_cast<uint64_t>(K) * 2, kBK, l.bn_tile / 2);
      if (e) return fail(c, XV_ECUDA, "%s", e);
    } else {
      if ((rc = upload(c, &l.w32, k->data.data(), sizeof(float) * K * l.cout)) != XV_OK) return rc;
    }
  }
  return XV_OK;
}

void resnet_release(xv_ctx* c) {
  for (ConvLayer& l : c->convs) {
    release(&l.wt); release(&l.w32); release(&l.scale); release(&l.shift); release(&l.slope);
  }
  release(&c->conv0_w); release(&c->conv0_scale); release(&c->conv0_shift); release(&c->conv0_slope);
  for (int i = 0; i < 4; ++i) release(&c->res_buf[i]);
}

// ------------------------------------------------------------------------------------------------ run
namespace {
struct SlotGeom {
  int t = 0, t_alloc = 0;
};

// Output tile (fb frequency bins x tb time steps, fb | f_out, fb*tb <= 128) with the fewest tiles per utterance.
void pick_tile(int t_out, int f_out, int* fb_out, int* tb_out) {
  int best_fb = 1, best_tb = 128;
  int64_t best_tiles = -1;
  for (int fb = 1; fb <= f_out && fb <= 128; ++fb) {
    if (f_out % fb) continue;
    int tb = 128 / fb;
    if (tb > t_out) tb = t_out;
    const int64_t tiles = static_cast<int64_t>(f_out / fb) * ((t_out + tb - 1) / tb);
    if (best_tiles < 0 || tiles < best_tiles || (tiles == best_tiles && fb > best_fb)) {
      best_tiles = tiles; best_fb = fb; best_tb = tb;
    }
  }
  *fb_out = best_fb;
  *tb_out = best_tb;
}

// Halo-tile geometry (xv_conv_halo_sm100.cu): tile rows live on a (fb + 2)-wide grid, tb = 128 / (fb + 2) time steps;
// the pair with the most useful rows per 128-row MMA wins (ties: wider fb = fewer junk columns re-read).
bool pick_halo_tile(int t_out, int f_out, int* fb_out, int* tb_out) {
  int best = 0;
  for (int fb = 1; fb <= f_out && fb + 2 <= 128; ++fb) {
    if (f_out % fb) continue;
    const int tb = 128 / (fb + 2);
    if (tb < 1) continue;
    if (fb * tb > best || (fb * tb == best && fb > *fb_out)) { best = fb * tb; *fb_out = fb; *tb_out = tb; }
  }
  (void)t_out;
  return best >= 64;        // below half-filled MMAs the generic kernel is the better one
}

int halo_env(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}
}  // namespace

// One implicit-GEMM convolution launch: builds the 5-D input map, the 4-D output map and the tap table.
struct ConvBf16Args {
  const void* in; int nb, t_in, ta_in, f_in, cin;
  int kh, kw, st, sf, cout, bn_tile;
  const CUtensorMap* tm_b;
  const void* wt;             // the weights tm_b describes ([cout][kh*kw*cin] bf16): the halo kernel maps them in 64-row slabs
  const float *scale, *shift, *slope; bool plain_relu;
  const void* res; int ta_res;
  void* out; int ta_out;
};

static int conv_bf16_launch(xv_ctx* c, const ConvBf16Args& a, const char* name) {
  int t_out, f_out, pt, pf;
  same_pads(a.t_in, a.kh, a.st, &t_out, &pt);
  same_pads(a.f_in, a.kw, a.sf, &f_out, &pf);
  if (a.f_in % a.sf != 0 || (a.st == 2 && a.ta_in % 2 != 0) || a.cin % kBK != 0 || a.cout % a.bn_tile != 0 || a.kh * a.kw > 9)
    return fail(c, XV_EINVAL, "conv %s: unsupported geometry (F %d, stride %dx%d, Cin %d, Cout %d)", name, a.f_in, a.st, a.sf,
                a.cin, a.cout);
  ConvGemmPlan plan;
  plan.bn = a.bn_tile;
  ConvGemmParams& p = plan.p;
  p.n_utts = a.nb; p.t_out = t_out; p.f_out = f_out;
  // 3x3 stride-1 layers with <= 128-wide output tiles: activation tile + halo loaded once per channel block
  // (Cin = 128 fits as well, but with 64-column MMAs it only ties the generic kernel: XV_CONV_HALO=2 forces it)
  const int halo_mode = halo_env("XV_CONV_HALO", 1);
  bool halo = a.kh == 3 && a.kw == 3 && a.st == 1 && a.sf == 1 && a.cout % 64 == 0 &&
              ((halo_mode == 1 && a.cin == 64) || (halo_mode == 2 && a.cin <= 128));
  if (halo) halo = pick_halo_tile(t_out, f_out, &p.fb, &p.tb);
  if (halo) {
    p.fbh = p.fb + 2;
    p.kb_per_tap = a.cin / kBK;
    halo = conv_halo_plan(&p, &plan.halo_smem);
  }
  if (!halo) {
    p.fbh = 0;
    pick_tile(t_out, f_out, &p.fb, &p.tb);
  }
  p.tiles_f = f_out / p.fb;
  p.tiles_t = (t_out + p.tb - 1) / p.tb;
  p.tile_rows = p.fb * p.tb;
  p.m_tiles = a.nb * p.tiles_t * p.tiles_f;
  p.n_tiles = a.cout / (halo ? 64 : a.bn_tile);
  p.taps = a.kh * a.kw;
  p.kb_per_tap = a.cin / kBK;
  p.cin = a.cin;
  for (int dt = 0; dt < a.kh; ++dt)
    for (int df = 0; df < a.kw; ++df) {
      const int at = dt - pt, af = df - pf;
      ConvTap& tp = p.tap[dt * a.kw + df];
      tp.c3 = floor_div(at, a.st);
      tp.c2 = at - tp.c3 * a.st;
      tp.c1 = floor_div(af, a.sf);
      tp.c0 = (af - tp.c1 * a.sf) * a.cin;
    }
  p.scale = a.scale; p.shift = a.shift; p.slope = a.slope;
  p.relu_only = a.plain_relu ? 1 : 0;
  p.residual = static_cast<const __nv_bfloat16*>(a.res);
  p.t_alloc_res = a.ta_res;
  p.cout = a.cout;
  const uint64_t cb = static_cast<uint64_t>(a.cin) * 2;
  const uint64_t dims_a[5] = {static_cast<uint64_t>(a.sf) * a.cin, static_cast<uint64_t>(a.f_in / a.sf),
                              static_cast<uint64_t>(a.st), static_cast<uint64_t>(a.st == 2 ? a.ta_in / 2 : a.t_in),
                              static_cast<uint64_t>(a.nb)};
  const uint64_t str_a[5] = {2, a.sf * cb, a.f_in * cb, a.st * a.f_in * cb, static_cast<uint64_t>(a.ta_in) * a.f_in * cb};
  const uint32_t box_a[5] = {static_cast<uint32_t>(kBK), static_cast<uint32_t>(halo ? p.fbh : p.fb), 1,
                             static_cast<uint32_t>(halo ? p.tb + 2 : p.tb), 1};
  const char* err = make_tmap_bf16_nd(&plan.tm_a, a.in, 5, dims_a, str_a, box_a, 128);
  const uint64_t co = static_cast<uint64_t>(a.cout) * 2;
  const uint64_t dims_o[4] = {static_cast<uint64_t>(a.cout), static_cast<uint64_t>(f_out), static_cast<uint64_t>(t_out),
                              static_cast<uint64_t>(a.nb)};
  const uint64_t str_o[4] = {2, co, f_out * co, static_cast<uint64_t>(a.ta_out) * f_out * co};
  const uint32_t box_o[4] = {32, static_cast<uint32_t>(p.fb), static_cast<uint32_t>(p.tb), 1};
  if (!err) err = make_tmap_bf16_nd(&plan.tm_out, a.out, 4, dims_o, str_o, box_o, 64);
  if (err) return fail(c, XV_ECUDA, "%s (layer %s)", err, name);
  if (halo) {
    const int K = a.kh * a.kw * a.cin;
    const char* be = make_tmap_bf16_2d(&plan.tm_b, a.wt, K, a.cout, static_cast<uint64_t>(K) * 2, kBK, 64);
    if (be) return fail(c, XV_ECUDA, "%s (layer %s)", be, name);
  } else {
    plan.tm_b = *a.tm_b;
  }
  cudaError_t e = halo ? launch_conv_halo(plan, c->sm_count, c->stream) : launch_conv_gemm(plan, c->sm_count, c->stream);
  if (e != cudaSuccess) return cuda_fail(c, e, "conv GEMM launch");
  return XV_OK;
}

int resnet_run_batch(xv_ctx* c, const float* feats, int nb, int t, const void** out) {
  const size_t esz = c->bf16 ? 2 : 4;
  const int ns = static_cast<int>(c->slots.size());
  std::vector<SlotGeom> g(ns);
  size_t max_bytes = 0;
  for (int i = 0; i < ns; ++i) {
    int tt = t;
    for (int k = 0; k < c->slots[i].tshift; ++k) tt = (tt + 1) / 2;
    g[i].t = tt;
    g[i].t_alloc = c->slots[i].pad_even ? tt + (tt & 1) : tt;
    max_bytes = std::max(max_bytes, static_cast<size_t>(nb) * g[i].t_alloc * c->slots[i].f * c->slots[i].c * esz);
  }
  int rc;
  for (int i = 0; i < 4; ++i)
    if ((rc = ensure(c, &c->res_buf[i], max_bytes + 4096)) != XV_OK) return rc;
  cudaError_t e;
  auto slot_ptr = [&](int s) { return c->res_buf[c->slots[s].buf].p; };
  auto utt_stride = [&](int s) { return static_cast<int64_t>(g[s].t_alloc) * c->slots[s].f * c->slots[s].c; };
  auto zero_pad_rows = [&](int s) -> cudaError_t {
    if (g[s].t_alloc == g[s].t) return cudaSuccess;
    const size_t row = static_cast<size_t>(c->slots[s].f) * c->slots[s].c * esz;
    return cudaMemset2DAsync(static_cast<uint8_t*>(slot_ptr(s)) + row * g[s].t, row * g[s].t_alloc, 0, row, nb, c->stream);
  };
  const int grid_ew = c->sm_count * 8;
  // ---- stem ----
  {
    const int s0 = c->resnet_conv0_slot;
    ProfScope ps(c, "resnet_stem");
    if (c->bf16)
      resnet_stem_kernel<__nv_bfloat16><<<grid_ew, 256, 0, c->stream>>>(
          feats, nb, t, c->dim, static_cast<const float*>(c->conv0_w.p), static_cast<const float*>(c->conv0_scale.p),
          static_cast<const float*>(c->conv0_shift.p), static_cast<const float*>(c->conv0_slope.p),
          static_cast<__nv_bfloat16*>(slot_ptr(s0)), utt_stride(s0));
    else
      resnet_stem_kernel<float><<<grid_ew, 256, 0, c->stream>>>(
          feats, nb, t, c->dim, static_cast<const float*>(c->conv0_w.p), static_cast<const float*>(c->conv0_scale.p),
          static_cast<const float*>(c->conv0_shift.p), static_cast<const float*>(c->conv0_slope.p),
          static_cast<float*>(slot_ptr(s0)), utt_stride(s0));
    if ((e = cudaGetLastError()) != cudaSuccess) return cuda_fail(c, e, "resnet stem launch");
    if ((e = zero_pad_rows(s0)) != cudaSuccess) return cuda_fail(c, e, "cudaMemset2DAsync");
  }
  if (c->resnet_pool_slot >= 0) {
    const int s0 = c->resnet_conv0_slot, s1 = c->resnet_pool_slot;
    ProfScope ps(c, "resnet_maxpool");
    if (c->bf16)
      maxpool3x3_kernel<__nv_bfloat16><<<grid_ew, 256, 0, c->stream>>>(static_cast<const __nv_bfloat16*>(slot_ptr(s0)), nb, t,
                                                                      c->dim, 64, utt_stride(s0),
                                                                      static_cast<__nv_bfloat16*>(slot_ptr(s1)), utt_stride(s1));
    else
      maxpool3x3_kernel<float><<<grid_ew, 256, 0, c->stream>>>(static_cast<const float*>(slot_ptr(s0)), nb, t, c->dim, 64,
                                                              utt_stride(s0), static_cast<float*>(slot_ptr(s1)), utt_stride(s1));
    if ((e = cudaGetLastError()) != cudaSuccess) return cuda_fail(c, e, "max pooling launch");
    if ((e = zero_pad_rows(s1)) != cudaSuccess) return cuda_fail(c, e, "cudaMemset2DAsync");
  }
  // ---- conv blocks ----
  for (const ConvLayer& l : c->convs) {
    const ActSlot& si = c->slots[l.in_slot];
    const ActSlot& so = c->slots[l.out_slot];
    const int t_in = g[l.in_slot].t, f_in = si.f;
    int t_out, f_out, pt, pf;
    same_pads(t_in, l.kh, l.st, &t_out, &pt);
    same_pads(f_in, l.kw, l.sf, &f_out, &pf);
    if (t_out != g[l.out_slot].t || f_out != so.f)
      return fail(c, XV_ESTATE, "resnet plan: layer %s geometry mismatch (%dx%d vs %dx%d)", l.name.c_str(), t_out, f_out,
                  g[l.out_slot].t, so.f);
    char fam[32];
    snprintf(fam, sizeof(fam), "conv.c%d", l.cout);
    ProfScope ps(c, fam);
    if (c->bf16) {
      ConvBf16Args ca;
      ca.in = slot_ptr(l.in_slot); ca.nb = nb; ca.t_in = t_in; ca.ta_in = g[l.in_slot].t_alloc; ca.f_in = f_in; ca.cin = l.cin;
      ca.kh = l.kh; ca.kw = l.kw; ca.st = l.st; ca.sf = l.sf; ca.cout = l.cout; ca.bn_tile = l.bn_tile; ca.tm_b = &l.tm_b; ca.wt = l.wt.p;
      ca.scale = static_cast<const float*>(l.scale.p); ca.shift = static_cast<const float*>(l.shift.p);
      ca.slope = static_cast<const float*>(l.slope.p); ca.plain_relu = l.plain_relu;
      ca.res = l.res_slot >= 0 ? slot_ptr(l.res_slot) : nullptr;
      ca.ta_res = l.res_slot >= 0 ? g[l.res_slot].t_alloc : 0;
      ca.out = slot_ptr(l.out_slot); ca.ta_out = g[l.out_slot].t_alloc;
      if ((rc = conv_bf16_launch(c, ca, l.name.c_str())) != XV_OK) return rc;
    } else {
      SimtConvArgs a;
      a.x = static_cast<const float*>(slot_ptr(l.in_slot)); a.nb = nb; a.t_in = t_in; a.f_in = f_in; a.cin = l.cin;
      a.in_utt_stride = utt_stride(l.in_slot);
      a.w = static_cast<const float*>(l.w32.p);
      a.kh = l.kh; a.kw = l.kw; a.st = l.st; a.sf = l.sf; a.pt = pt; a.pf = pf;
      a.t_out = t_out; a.f_out = f_out; a.cout = l.cout; a.out_utt_stride = utt_stride(l.out_slot);
      a.scale = static_cast<const float*>(l.scale.p); a.shift = static_cast<const float*>(l.shift.p);
      a.slope = static_cast<const float*>(l.slope.p);
      a.res = l.res_slot >= 0 ? static_cast<const float*>(slot_ptr(l.res_slot)) : nullptr;
      a.res_utt_stride = l.res_slot >= 0 ? utt_stride(l.res_slot) : 0;
      a.out = static_cast<float*>(slot_ptr(l.out_slot));
      const int64_t m_total = static_cast<int64_t>(nb) * t_out * f_out;
      dim3 grid(static_cast<unsigned>((m_total + 63) / 64), static_cast<unsigned>((l.cout + 63) / 64));
      simt_conv_kernel<<<grid, 256, 0, c->stream>>>(a);
      if ((e = cudaGetLastError()) != cudaSuccess) return cuda_fail(c, e, "SIMT conv launch");
    }
    if ((e = zero_pad_rows(l.out_slot)) != cudaSuccess) return cuda_fail(c, e, "cudaMemset2DAsync");
  }
  *out = slot_ptr(c->resnet_out_slot);
  return XV_OK;
}

}  // namespace xv

using namespace xv;

extern "C" int xv_conv2d_bf16_async(xv_ctx* c, const void* x_dev, int32_t nb, int32_t t, int32_t f, int32_t cin,
                                    const void* wt_dev, int32_t kh, int32_t kw, int32_t st, int32_t sf, int32_t cout,
                                    const float* scale_dev, const float* shift_dev, const float* slope_dev,
                                    const void* residual_dev, void* out_dev) {
  if (!c) return XV_EINVAL;
  if (!x_dev || !wt_dev || !scale_dev || !shift_dev || !slope_dev || !out_dev || nb <= 0 || t <= 0 || f <= 0)
    return fail(c, XV_EINVAL, "xv_conv2d_bf16_async: bad argument");
  if ((st != 1 && st != 2) || (sf != 1 && sf != 2) || (st == 2 && t % 2 != 0))
    return fail(c, XV_EINVAL, "xv_conv2d_bf16_async: strides must be 1 or 2 (time stride 2 needs an even T)");
  cudaSetDevice(c->device);
  const char* ge = conv_gemm_init();
  if (!ge) ge = conv_halo_init();
  if (ge) return fail(c, XV_ECUDA, "%s", ge);
  const int bn_tile = cout % 256 == 0 ? 256 : (cout % 128 == 0 ? 128 : 64);
  const int K = kh * kw * cin;
  CUtensorMap tm_b;
  const char* e = make_tmap_bf16_2d(&tm_b, wt_dev, K, cout, static_cast<uint64_t>(K) * 2, kBK, bn_tile / 2);
  if (e) return fail(c, XV_ECUDA, "%s", e);
  ConvBf16Args a;
  a.in = x_dev; a.nb = nb; a.t_in = t; a.ta_in = t; a.f_in = f; a.cin = cin;
  a.kh = kh; a.kw = kw; a.st = st; a.sf = sf; a.cout = cout; a.bn_tile = bn_tile; a.tm_b = &tm_b; a.wt = wt_dev;
  a.scale = scale_dev; a.shift = shift_dev; a.slope = slope_dev; a.plain_relu = false;
  const int t_out = (t + st - 1) / st;
  a.res = residual_dev; a.ta_res = t_out; a.out = out_dev; a.ta_out = t_out;
  ProfScope ps(c, "conv2d");
  return conv_bf16_launch(c, a, "xv_conv2d_bf16_async");
}
