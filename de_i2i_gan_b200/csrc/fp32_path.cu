// FP32-parity path (SURVEY.md K1 "FP32-parity mode", north-star tolerance 1e-3 relative in FP32): the same operators on FP32 NHWC
// activations with exact FP32 arithmetic (CUDA-core FMA, FP32 accumulation), so that the mirrored modules can be compared with the
// reference - which trains in FP32 (defectgan_model.py:54, autocast commented out) - at FP32 tolerances, independent of the BF16
// storage format of the production path.  Selected per tensor dtype by the Python wrappers (DEI2I_KERNELS=fp32); every entry point
// below has the semantics of its BF16 sibling in include/dei2i.h.  These kernels are written for clarity and exactness, not for
// speed: 64 x 64 x 16 shared-memory tiles for the implicit GEMMs, one thread per element for the streaming ops, every reduction
// in a fixed order (deterministic).
#include <string.h>

#include "common.cuh"
#include "conv_common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {
void count_launch(int n = 1);
void trace_begin();
void trace_kernel(const char* fmt, ...);

// im2col gather over an FP32 NHWC tensor (the FP32 twin of conv_common.cuh's Gather)
struct GatherF {
  const float* src;
  int N, Hl, Wl, Hs, Ws, C;
  int Ho, Wo;
  int kh, kw, stride, pad, pad_mode;
  int mode;  // 0: forward, 1: dgrad
  int off;
  long long M;
};
DEVINL bool axis_f(const GatherF& g, int o, int r, int Ll, int Ls, int& i_out) {
  if (g.mode == 0) {
    int i = o * g.stride + r - g.pad;
    if (g.pad_mode == DEI2I_PAD_REFLECT) {
      i = i < 0 ? -i : i;
      i = i >= Ll ? 2 * Ll - 2 - i : i;
    } else if (i < 0 || i >= Ll) {
      return false;
    }
    if (Ls != Ll) i = (int)(((long long)i * Ls) / Ll);
    i_out = i;
    return true;
  }
  const int t = o + g.off - r;
  if (t < 0) return false;
  const int q = t / g.stride;
  if (q * g.stride != t || q >= Ls) return false;
  i_out = q;
  return true;
}
// 4 consecutive channels (k .. k+3 within one tap) of GEMM row m; zeros where the tap falls outside
DEVINL float4 gather4(const GatherF& g, long long m, int k) {
  if (m >= g.M) return make_float4(0, 0, 0, 0);
  const int tap = k / g.C, ci = k - tap * g.C;
  if (tap >= g.kh * g.kw) return make_float4(0, 0, 0, 0);
  const int r = tap / g.kw, s = tap - r * g.kw;
  const int HoWo = g.Ho * g.Wo;
  const int img = (int)(m / HoWo);
  const int rem = (int)(m - (long long)img * HoWo);
  const int oy = rem / g.Wo, ox = rem - oy * g.Wo;
  int iy, ix;
  if (!axis_f(g, oy, r, g.Hl, g.Hs, iy) || !axis_f(g, ox, s, g.Wl, g.Ws, ix)) return make_float4(0, 0, 0, 0);
  return __ldg(reinterpret_cast<const float4*>(g.src + (((long long)img * g.Hs + iy) * g.Ws + ix) * g.C + ci));
}

struct EpiF {
  const float* scale;
  const float* bias;
  const float* noise;
  const float* noise_weight;
  const float* residual;
  float* out;
  int act, ld, ncols;
};

constexpr int FM = 64, FN = 64, FK = 16;

// D[m, n] = sum_k A[m, k] * B[n, k]: A gathered (im2col), B = packed FP32 weights [rows, Kf]
__global__ void __launch_bounds__(256)
conv_f32_kernel(GatherF g, const float* __restrict__ wmat, int rows, int Kf, int K, EpiF e) {
  __shared__ float As[FK][FM + 4], Bs[FK][FN + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const long long m0 = (long long)blockIdx.x * FM;
  const int n0 = blockIdx.y * FN;
  const int lr = threadIdx.x >> 2, lk = (threadIdx.x & 3) * 4;  // loader: row lr, k chunk lk
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += FK) {
    const float4 a = gather4(g, m0 + lr, k0 + lk);
    float4 b = make_float4(0, 0, 0, 0);
    if (n0 + lr < rows && k0 + lk < Kf) b = __ldg(reinterpret_cast<const float4*>(wmat + (size_t)(n0 + lr) * Kf + k0 + lk));
    As[lk][lr] = a.x; As[lk + 1][lr] = a.y; As[lk + 2][lr] = a.z; As[lk + 3][lr] = a.w;
    Bs[lk][lr] = b.x; Bs[lk + 1][lr] = b.y; Bs[lk + 2][lr] = b.z; Bs[lk + 3][lr] = b.w;
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < FK; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
  const float nzw = e.noise != nullptr ? __ldg(e.noise_weight) : 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= g.M) continue;
    const float nz = e.noise != nullptr ? __ldg(e.noise + m) * nzw : 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= e.ld) continue;
      float v = 0.f;
      if (n < e.ncols) {
        v = acc[i][j] * osc;
        if (e.bias != nullptr) v += __ldg(e.bias + n);
        v = apply_act(v + nz, e.act);
        if (e.residual != nullptr) v += __ldg(e.residual + m * e.ld + n);
      }
      e.out[m * e.ld + n] = v;
    }
  }
}

// wgrad: P[co, k] = sum_{m in split} dY[m, co] * A[m, k]  -> partial [splits, Cout, Kf] (second pass shared with the BF16 path)
__global__ void __launch_bounds__(256)
wgrad_f32_kernel(GatherF g, const float* __restrict__ dy, int cout, int cout_ld, int Kf, int K, long long m_per_split,
                 float* __restrict__ partial) {
  __shared__ float Ds[FK][FM + 4], As[FK][FN + 4];  // [pixel][co], [pixel][k]
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int k0 = blockIdx.x * FN, co0 = blockIdx.y * FM, split = blockIdx.z;
  const long long ms = (long long)split * m_per_split, me = min(g.M, ms + m_per_split);
  const int lp = threadIdx.x >> 4, lc = (threadIdx.x & 15) * 4;  // loader: pixel lp (0..15), column chunk lc (0..60)
  float acc[4][4] = {};
  for (long long mb = ms; mb < me; mb += FK) {
    const long long m = mb + lp;
    float4 d = make_float4(0, 0, 0, 0), a = make_float4(0, 0, 0, 0);
    if (m < me) {
      if (co0 + lc < cout_ld) d = __ldg(reinterpret_cast<const float4*>(dy + m * cout_ld + co0 + lc));
      if (k0 + lc < K) a = gather4(g, m, k0 + lc);
    }
    *reinterpret_cast<float4*>(&Ds[lp][lc]) = d;
    *reinterpret_cast<float4*>(&As[lp][lc]) = a;
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < FK; ++kk) {
      float dv[4], av[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) dv[i] = Ds[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) av[j] = As[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(dv[i], av[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int co = co0 + ty * 4 + i;
    if (co >= cout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + tx * 4 + j;
      if (k < Kf) partial[((size_t)split * cout + co) * Kf + k] = acc[i][j];
    }
  }
}

__global__ void pack_f32_kernel(const float* __restrict__ w, int Cout, int Cin_real, int taps, int Cin, int Cout_ld, int Kf, int Kd,
                                float* __restrict__ w_fwd, float* __restrict__ w_dg) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (long long)Cout * Kf) {
    const int co = (int)(i / Kf), k = (int)(i - (long long)co * Kf);
    const int tap = k / Cin, ci = k - tap * Cin;
    w_fwd[i] = (tap < taps && ci < Cin_real) ? w[((long long)co * Cin_real + ci) * taps + tap] : 0.f;
  }
  if (w_dg != nullptr && i < (long long)Cin * Kd) {
    const int ci = (int)(i / Kd), k = (int)(i - (long long)ci * Kd);
    const int tap = k / Cout_ld, co = k - tap * Cout_ld;
    w_dg[i] = (tap < taps && co < Cout && ci < Cin_real) ? w[((long long)co * Cin_real + ci) * taps + tap] : 0.f;
  }
}

static int out_dim(int L, int k, int stride, int pad) { return (L + 2 * pad - k) / stride + 1; }
static int ru(int a, int b) { return (a + b - 1) / b * b; }
static int f32_wgrad_splits(long long M, int tiles) {
  int s = (148 * 4 + tiles - 1) / tiles;
  const long long maxs = (M + 255) / 256;
  if (s > maxs) s = (int)maxs;
  return s < 1 ? 1 : s;
}

}  // namespace dei2i

using namespace dei2i;
#define ST ((cudaStream_t)stream)
#define LAUNCH_OK(name) (count_launch(), check_cuda(cudaGetLastError(), name))

extern "C" int dei2i_pack_weight_f32(const float* w, int Cout, int Cin_real, int kh, int kw, int Cin, int Cout_ld, float* w_fwd,
                                     float* w_dgrad, dei2i_stream stream) {
  DEI2I_REQUIRE(w && w_fwd && Cout > 0 && Cin_real > 0 && Cin % 8 == 0 && Cin >= Cin_real && Cout_ld % 8 == 0, "pack_weight_f32: bad arguments");
  const int taps = kh * kw, Kf = ru(taps * Cin, 64), Kd = ru(taps * Cout_ld, 64);
  const long long n = (long long)Cout * Kf > (long long)Cin * Kd ? (long long)Cout * Kf : (long long)Cin * Kd;
  pack_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ST>>>(w, Cout, Cin_real, taps, Cin, Cout_ld, Kf, Kd, w_fwd, w_dgrad);
  return LAUNCH_OK("pack_f32_kernel");
}

static void fill_gather(GatherF& gg, const float* src, const dei2i_conv_geom* g, int Ho, int Wo) {
  memset(&gg, 0, sizeof(gg));
  gg.src = src; gg.N = g->N; gg.Hl = g->H; gg.Wl = g->W; gg.Hs = g->Hs; gg.Ws = g->Ws; gg.C = g->Cin; gg.Ho = Ho; gg.Wo = Wo;
  gg.kh = g->kh; gg.kw = g->kw; gg.stride = g->stride; gg.pad = g->pad; gg.pad_mode = g->pad_mode; gg.mode = 0;
  gg.M = (long long)g->N * Ho * Wo;
}

extern "C" int dei2i_conv2d_fwd_f32(const float* x, const float* w_fwd, float* y, const dei2i_conv_geom* g, const dei2i_conv_epilogue* e,
                                    dei2i_stream stream) {
  trace_begin();
  DEI2I_REQUIRE(x && w_fwd && y && g && e, "conv2d_fwd_f32: null pointer");
  DEI2I_REQUIRE(g->Cin % 8 == 0 && g->Cout_ld % 8 == 0 && g->Cout <= g->Cout_ld && e->stats_partial == nullptr && e->out_pad == 0,
                "conv2d_fwd_f32: bad geometry / unsupported epilogue field");
  const int Ho = out_dim(g->H, g->kh, g->stride, g->pad), Wo = out_dim(g->W, g->kw, g->stride, g->pad);
  DEI2I_REQUIRE(Ho > 0 && Wo > 0, "conv2d_fwd_f32: empty output");
  {  // DEI2I_KERNELS=tf32: kind::tf32 tcgen05 kernel where the TMA-im2col rules admit the geometry
    const int rc = conv_fwd_tf32(x, w_fwd, y, g, e, ST);
    if (rc != 1) return rc;
  }
  GatherF gg;
  fill_gather(gg, x, g, Ho, Wo);
  EpiF ep{e->scale, e->bias, e->noise, e->noise_weight, (const float*)e->residual, y, e->act, g->Cout_ld, g->Cout};
  const int K = g->kh * g->kw * g->Cin, Kf = ru(K, 64);
  dim3 grid((unsigned)((gg.M + FM - 1) / FM), (unsigned)((g->Cout_ld + FN - 1) / FN));
  conv_f32_kernel<<<grid, 256, 0, ST>>>(gg, w_fwd, g->Cout, Kf, K, ep);
  trace_kernel("conv_f32");
  return LAUNCH_OK("conv_f32_kernel");
}

extern "C" int dei2i_conv2d_dgrad_f32(const float* dy, const float* w_dgrad, float* dx, const dei2i_conv_geom* g, const float* scale,
                                      dei2i_stream stream) {
  trace_begin();
  DEI2I_REQUIRE(dy && w_dgrad && dx && g, "conv2d_dgrad_f32: null pointer");
  DEI2I_REQUIRE(g->Hs == g->H && g->Ws == g->W, "conv2d_dgrad_f32: resized inputs are not supported");
  const int Ho = out_dim(g->H, g->kh, g->stride, g->pad), Wo = out_dim(g->W, g->kw, g->stride, g->pad);
  {
    const int rc = conv_dgrad_tf32(dy, w_dgrad, dx, g, scale, ST);
    if (rc != 1) return rc;
  }
  const bool full = g->pad_mode == DEI2I_PAD_REFLECT;  // reflect: gradient of the PADDED input (the caller folds it)
  const int Hd = full ? g->H + 2 * g->pad : g->H, Wd = full ? g->W + 2 * g->pad : g->W;
  GatherF gg;
  memset(&gg, 0, sizeof(gg));
  gg.src = dy; gg.N = g->N; gg.Hl = Ho; gg.Wl = Wo; gg.Hs = Ho; gg.Ws = Wo; gg.C = g->Cout_ld; gg.Ho = Hd; gg.Wo = Wd;
  gg.kh = g->kh; gg.kw = g->kw; gg.stride = g->stride; gg.pad = g->pad; gg.pad_mode = g->pad_mode; gg.mode = 1;
  gg.off = full ? 0 : g->pad;
  gg.M = (long long)g->N * Hd * Wd;
  EpiF ep{scale, nullptr, nullptr, nullptr, nullptr, dx, 0, g->Cin, g->Cin};
  const int K = g->kh * g->kw * g->Cout_ld, Kd = ru(K, 64);
  dim3 grid((unsigned)((gg.M + FM - 1) / FM), (unsigned)((g->Cin + FN - 1) / FN));
  conv_f32_kernel<<<grid, 256, 0, ST>>>(gg, w_dgrad, g->Cin, Kd, K, ep);
  trace_kernel("conv_f32(dgrad)");
  return LAUNCH_OK("conv_f32_kernel");
}

extern "C" size_t dei2i_conv2d_wgrad_f32_workspace(const dei2i_conv_geom* g) {
  if (g == nullptr) return 0;
  const int Ho = out_dim(g->H, g->kh, g->stride, g->pad), Wo = out_dim(g->W, g->kw, g->stride, g->pad);
  const int K = g->kh * g->kw * g->Cin, Kf = ru(K, 64);
  const int tiles = ((K + FN - 1) / FN) * ((g->Cout + FM - 1) / FM);
  const int splits = f32_wgrad_splits((long long)g->N * Ho * Wo, tiles);
  return (size_t)splits * g->Cout * Kf * sizeof(float) + 1024 * sizeof(float) + 256;
}

extern "C" int dei2i_conv2d_wgrad_f32(const float* x, const float* dy, float* dw, int Cin_real, float beta, const dei2i_conv_geom* g,
                                      const float* sn_w_orig, const float* sn_u, const float* sn_v, const float* sn_inv_sigma,
                                      void* workspace, size_t workspace_bytes, dei2i_stream stream) {
  trace_begin();
  DEI2I_REQUIRE(x && dy && dw && g && workspace, "conv2d_wgrad_f32: null pointer");
  DEI2I_REQUIRE(workspace_bytes >= dei2i_conv2d_wgrad_f32_workspace(g), "conv2d_wgrad_f32: workspace too small");
  const int Ho = out_dim(g->H, g->kh, g->stride, g->pad), Wo = out_dim(g->W, g->kw, g->stride, g->pad);
  GatherF gg;
  fill_gather(gg, x, g, Ho, Wo);
  const int K = g->kh * g->kw * g->Cin, Kf = ru(K, 64);
  const int kt = (K + FN - 1) / FN, ct = (g->Cout + FM - 1) / FM;
  const int splits = f32_wgrad_splits(gg.M, kt * ct);
  long long mps = (gg.M + splits - 1) / splits;
  mps = (mps + FK - 1) / FK * FK;
  const int splits2 = (int)((gg.M + mps - 1) / mps);
  wgrad_f32_kernel<<<dim3(kt, ct, splits2), 256, 0, ST>>>(gg, dy, g->Cout, g->Cout_ld, Kf, K, mps, (float*)workspace);
  trace_kernel("wgrad_f32");
  if (int rc = LAUNCH_OK("wgrad_f32_kernel")) return rc;
  return wgrad_second_pass((float*)workspace, splits2, g, Cin_real, Kf, dw, beta, sn_w_orig, sn_u, sn_v, sn_inv_sigma, ST);
}

// ======================================================================================================================
// Streaming operators on FP32 NHWC tensors [N, H, W, C] (C = channel pitch, multiple of 8).  One thread per element / per channel;
// per-channel reductions by one block per 32 channels walking the pixels with 32 lanes, lanes combined in a fixed order.
namespace dei2i {

__global__ void nchw_to_nhwc_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int C, int HW, int C_ld) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * HW * C_ld) return;
  const int c = (int)(i % C_ld);
  const long long p = i / C_ld;
  const int n = (int)(p / HW), px = (int)(p - (long long)n * HW);
  dst[i] = c < C ? src[((long long)n * C + c) * HW + px] : 0.f;
}
__global__ void nhwc_to_nchw_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int C, int HW, int C_ld) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * C * HW) return;
  const int px = (int)(i % HW);
  const long long t = i / HW;
  const int c = (int)(t % C), n = (int)(t / C);
  dst[i] = src[((long long)n * HW + px) * C_ld + c];
}

// block = 32 channels x 32 lanes; returns (in lane 0 of each channel column) the fixed-order sum over lanes.  Every reduction of this
// path accumulates in FP64: gradients such as the scalar noise weight are sums of millions of signed terms, and FP32 summation noise
// alone would cost more than the 1e-3 this path exists to demonstrate.
DEVINL double lanes_sum(double v, double (*sh)[33]) {
  const int ch = threadIdx.x & 31, ln = threadIdx.x >> 5;
  __syncthreads();
  sh[ln][ch] = v;
  __syncthreads();
  double t = 0.0;
  if (ln == 0)
    for (int l = 0; l < 32; ++l) t += sh[l][ch];
  return t;
}

// (mean, biased variance) per channel of group g (per image or over the batch): exact two passes
__global__ void __launch_bounds__(1024)
channel_stats_f32_kernel(const float* __restrict__ x, long long pix_per_group, int C, float* __restrict__ stats) {
  __shared__ double sh[32][33];
  __shared__ double mean_s[32];
  const int ch = threadIdx.x & 31, ln = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + ch, g = blockIdx.y;
  const float* xg = x + (size_t)g * pix_per_group * C;
  double a = 0.0;
  if (c < C)
    for (long long p = ln; p < pix_per_group; p += 32) a += xg[p * C + c];
  a = lanes_sum(a, sh);
  if (ln == 0) mean_s[ch] = a / (double)pix_per_group;
  __syncthreads();
  const double mean = mean_s[ch];
  a = 0.0;
  if (c < C)
    for (long long p = ln; p < pix_per_group; p += 32) {
      const double d = (double)xg[p * C + c] - mean;
      a += d * d;
    }
  a = lanes_sum(a, sh);
  if (ln == 0 && c < C) {
    stats[((size_t)g * 2 + 0) * C + c] = (float)mean;
    stats[((size_t)g * 2 + 1) * C + c] = (float)(a / (double)pix_per_group);
  }
}

__global__ void bn_fwd_f32_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ residual,
                                  const float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                                  float* __restrict__ running_mean, float* __restrict__ running_var, float* __restrict__ save_mean,
                                  float* __restrict__ save_rstd, long long npix, int C, int training, float momentum, float eps, int act) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix * C) return;
  const int c = (int)(i % C);
  const float mean = training ? stats[c] : running_mean[c], var = training ? stats[C + c] : running_var[c];
  const float rstd = rsqrtf(var + eps);
  float v = apply_act((x[i] - mean) * rstd * gamma[c] + beta[c], act);
  if (residual != nullptr) v += residual[i];
  y[i] = v;
  if (i < C) {  // pixel 0: one thread per channel owns the statistics' side effects
    if (save_mean != nullptr) { save_mean[c] = mean; save_rstd[c] = rstd; }
  }
}
// running statistics in a kernel of their own (after the element pass has read them in eval mode / never in training mode)
__global__ void bn_running_f32_kernel(const float* __restrict__ stats, float* __restrict__ running_mean, float* __restrict__ running_var,
                                      long long npix, int C, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = stats[c], var = stats[C + c];
  running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
  const float unb = npix > 1 ? var * (float)npix / (float)(npix - 1) : var;
  running_var[c] = (1.f - momentum) * running_var[c] + momentum * unb;
}

DEVINL float act_grad_from_z(float z, int act) { return act == ACT_LRELU ? (z > 0.f ? 1.f : 0.2f) : (act == ACT_RELU ? (z > 0.f ? 1.f : 0.f) : 1.f); }

// red[c] = sum dz (= dbeta), red[C + c] = sum dz * xhat (= dgamma)
__global__ void __launch_bounds__(1024)
bn_bwd_reduce_f32_kernel(const float* __restrict__ x, const float* __restrict__ dy, const float* __restrict__ gamma,
                         const float* __restrict__ beta, const float* __restrict__ mean, const float* __restrict__ rstd, long long npix,
                         int C, int act, float* __restrict__ red) {
  __shared__ double sh[32][33];
  const int ch = threadIdx.x & 31, ln = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + ch;
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float m = mean[c], r = rstd[c], g = gamma[c], b = beta[c];
    for (long long p = ln; p < npix; p += 32) {
      const float xh = (x[p * C + c] - m) * r;
      const float dz = dy[p * C + c] * act_grad_from_z(xh * g + b, act);
      s += dz;
      q += (double)dz * xh;
    }
  }
  s = lanes_sum(s, sh);
  q = lanes_sum(q, sh);
  if (ln == 0 && c < C) { red[c] = (float)s; red[C + c] = (float)q; }
}
__global__ void bn_bwd_apply_f32_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx,
                                        const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                                        const float* __restrict__ rstd, const float* __restrict__ red, long long npix, int C, int training,
                                        int act) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix * C) return;
  const int c = (int)(i % C);
  const float r = rstd[c], g = gamma[c];
  const float xh = (x[i] - mean[c]) * r;
  const float dz = dy[i] * act_grad_from_z(xh * g + beta[c], act);
  const float k1 = training ? red[c] / (float)npix : 0.f, k2 = training ? red[C + c] / (float)npix : 0.f;
  dx[i] = g * r * (dz - k1 - xh * k2);
}

DEVINL int bcls_f(int i, int n) { return i < 2 ? i : (i >= n - 2 ? i - n + 5 : 2); }
DEVINL int refl_f(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i); }

// SPADE on the border-class map, FP32: out[n, hp, wp, c] = [relu](xhat * (1 + gamma5[cls]) + beta5[cls]), reflect padding written too
__global__ void spade_cls_fwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ gb5, float* __restrict__ out,
                                         const float* __restrict__ stats, float* __restrict__ mean_rstd, int N, int H, int W, int C, int up,
                                         int relu, float eps, int pad) {
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * Hp * Wp * C) return;
  const int c = (int)(i % C);
  long long p = i / C;
  const int wp = (int)(p % Wp);
  p /= Wp;
  const int hp = (int)(p % Hp), n = (int)(p / Hp);
  const int h = refl_f(hp - pad, H), w = refl_f(wp - pad, W);
  const float mean = stats[((size_t)n * 2 + 0) * C + c], rstd = rsqrtf(stats[((size_t)n * 2 + 1) * C + c] + eps);
  if (hp == 0 && wp == 0) { mean_rstd[((size_t)n * 2 + 0) * C + c] = mean; mean_rstd[((size_t)n * 2 + 1) * C + c] = rstd; }
  const int Hx = H / up, Wx = W / up;
  const float xh = (x[(((size_t)n * Hx + h / up) * Wx + w / up) * C + c] - mean) * rstd;
  const int cls = bcls_f(h, H) * 5 + bcls_f(w, W);
  const float* gb = gb5 + ((size_t)n * 25 + cls) * 2 * C;
  const float v = xh * (1.f + gb[c]) + gb[C + c];
  out[i] = relu ? fmaxf(v, 0.f) : v;
}
// folded gradient of the padded output at interior position (h, w)
DEVINL float folded_f(const float* __restrict__ d, size_t img, int h, int w, int H, int W, int pad, int C, int c) {
  const int Wp = W + 2 * pad;
  float v = d[(img + (size_t)(h + pad) * Wp + (w + pad)) * C + c];
  if (pad == 0) return v;
  const int h2 = (h >= 1 && h <= pad) ? pad - h : ((h <= H - 2 && h >= H - 1 - pad) ? 2 * H - 2 - h + pad : -1);
  const int w2 = (w >= 1 && w <= pad) ? pad - w : ((w <= W - 2 && w >= W - 1 - pad) ? 2 * W - 2 - w + pad : -1);
  if (h2 >= 0) v += d[(img + (size_t)h2 * Wp + (w + pad)) * C + c];
  if (w2 >= 0) v += d[(img + (size_t)(h + pad) * Wp + w2) * C + c];
  if (h2 >= 0 && w2 >= 0) v += d[(img + (size_t)h2 * Wp + w2) * C + c];
  return v;
}
// class sums: grid (25 classes, N, C / 32), block 32 ch x 32 lanes over the class's rectangle
__global__ void __launch_bounds__(1024)
spade_cls_bwd1_f32_kernel(const float* __restrict__ x, const float* __restrict__ gb5, const float* __restrict__ dout,
                          const float* __restrict__ mean_rstd, float* __restrict__ dgb5, int H, int W, int C, int up, int relu, int pad) {
  __shared__ double sh[32][33];
  const int ch = threadIdx.x & 31, ln = threadIdx.x >> 5;
  const int cls = blockIdx.x, n = blockIdx.y, c = blockIdx.z * 32 + ch;
  const int chh = cls / 5, cw = cls - chh * 5;
  const int h0 = chh < 2 ? chh : (chh == 2 ? 2 : H - 5 + chh), h1 = chh == 2 ? H - 2 : h0 + 1;
  const int w0 = cw < 2 ? cw : (cw == 2 ? 2 : W - 5 + cw), w1 = cw == 2 ? W - 2 : w0 + 1;
  const int nw = w1 - w0, npx = (h1 - h0) * nw;
  const int Hx = H / up, Wx = W / up;
  double ag = 0.0, ab = 0.0;
  if (c < C) {
    const float mean = mean_rstd[((size_t)n * 2 + 0) * C + c], rstd = mean_rstd[((size_t)n * 2 + 1) * C + c];
    const float g1 = 1.f + gb5[((size_t)n * 25 + cls) * 2 * C + c], b = gb5[((size_t)n * 25 + cls) * 2 * C + C + c];
    const size_t img = (size_t)n * (H + 2 * pad) * (W + 2 * pad);
    for (int i = ln; i < npx; i += 32) {
      const int r = i / nw, h = h0 + r, w = w0 + (i - r * nw);
      const float xh = (x[(((size_t)n * Hx + h / up) * Wx + w / up) * C + c] - mean) * rstd;
      const float d = folded_f(dout, img, h, w, H, W, pad, C, c);
      const float gm = (!relu || xh * g1 + b > 0.f) ? d : 0.f;
      ag += (double)gm * xh;
      ab += gm;
    }
  }
  ag = lanes_sum(ag, sh);
  ab = lanes_sum(ab, sh);
  if (ln == 0 && c < C) {
    dgb5[((size_t)n * 25 + cls) * 2 * C + c] = (float)ag;
    dgb5[((size_t)n * 25 + cls) * 2 * C + C + c] = (float)ab;
  }
}
__global__ void spade_cls_red_f32_kernel(const float* __restrict__ gb5, const float* __restrict__ dgb5, float* __restrict__ red, int C) {
  const int n = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    double s1 = 0.0, s2 = 0.0;
    for (int cls = 0; cls < 25; ++cls) {
      const double g1 = 1.0 + gb5[((size_t)n * 25 + cls) * 2 * C + c];
      s1 += g1 * dgb5[((size_t)n * 25 + cls) * 2 * C + C + c];
      s2 += g1 * dgb5[((size_t)n * 25 + cls) * 2 * C + c];
    }
    red[(size_t)n * 2 * C + c] = (float)s1;
    red[(size_t)n * 2 * C + C + c] = (float)s2;
  }
}
__global__ void spade_cls_bwd2_f32_kernel(const float* __restrict__ x, const float* __restrict__ gb5, const float* __restrict__ dout,
                                          float* __restrict__ dx, const float* __restrict__ mean_rstd, const float* __restrict__ red, int N,
                                          int H, int W, int C, int up, int relu, int pad) {
  const int Hx = H / up, Wx = W / up;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * Hx * Wx * C) return;
  const int c = (int)(i % C);
  long long p = i / C;
  const int wx = (int)(p % Wx);
  p /= Wx;
  const int hx = (int)(p % Hx), n = (int)(p / Hx);
  const float mean = mean_rstd[((size_t)n * 2 + 0) * C + c], rstd = mean_rstd[((size_t)n * 2 + 1) * C + c];
  const float xh = (x[i] - mean) * rstd;
  const size_t img = (size_t)n * (H + 2 * pad) * (W + 2 * pad);
  float acc = 0.f;
  for (int dy = 0; dy < up; ++dy)
    for (int dxx = 0; dxx < up; ++dxx) {
      const int h = hx * up + dy, w = wx * up + dxx;
      const int cls = bcls_f(h, H) * 5 + bcls_f(w, W);
      const float g1 = 1.f + gb5[((size_t)n * 25 + cls) * 2 * C + c], b = gb5[((size_t)n * 25 + cls) * 2 * C + C + c];
      const float d = folded_f(dout, img, h, w, H, W, pad, C, c);
      acc += ((!relu || xh * g1 + b > 0.f) ? d : 0.f) * g1;
    }
  const float cnt = (float)H * (float)W, uu = (float)(up * up);
  dx[i] = rstd * (acc - uu * red[(size_t)n * 2 * C + c] / cnt - xh * uu * red[(size_t)n * 2 * C + C + c] / cnt);
}
// dense SPADE forward (spatial label maps): out = [relu](xhat * (1 + gamma) + beta), gb fp32 [N, H, W, 2C]
__global__ void spade_fwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ gb, float* __restrict__ out,
                                     const float* __restrict__ stats, float* __restrict__ mean_rstd, int N, int H, int W, int C, int up,
                                     int relu, float eps) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * H * W * C) return;
  const int c = (int)(i % C);
  long long p = i / C;
  const int w = (int)(p % W);
  p /= W;
  const int h = (int)(p % H), n = (int)(p / H);
  const float mean = stats[((size_t)n * 2 + 0) * C + c], rstd = rsqrtf(stats[((size_t)n * 2 + 1) * C + c] + eps);
  if (h == 0 && w == 0) { mean_rstd[((size_t)n * 2 + 0) * C + c] = mean; mean_rstd[((size_t)n * 2 + 1) * C + c] = rstd; }
  const int Hx = H / up, Wx = W / up;
  const float xh = (x[(((size_t)n * Hx + h / up) * Wx + w / up) * C + c] - mean) * rstd;
  const float* g = gb + (((size_t)n * H + h) * W + w) * 2 * C;
  const float v = xh * (1.f + g[c]) + g[C + c];
  out[i] = relu ? fmaxf(v, 0.f) : v;
}

__global__ void reflect_pad_f32_kernel(const float* __restrict__ x, float* __restrict__ xp, int N, int H, int W, int C, int pad) {
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * Hp * Wp * C) return;
  const int c = (int)(i % C);
  long long p = i / C;
  const int wp = (int)(p % Wp);
  p /= Wp;
  const int hp = (int)(p % Hp), n = (int)(p / Hp);
  xp[i] = x[(((size_t)n * H + refl_f(hp - pad, H)) * W + refl_f(wp - pad, W)) * C + c];
}
__global__ void reflect_fold_f32_kernel(const float* __restrict__ dxp, float* __restrict__ dx, int N, int H, int W, int C, int pad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)N * H * W * C) return;
  const int c = (int)(i % C);
  long long p = i / C;
  const int w = (int)(p % W);
  p /= W;
  const int h = (int)(p % H), n = (int)(p / H);
  // generic adjoint: every padded position that mirrors onto (h, w)
  float v = 0.f;
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  for (int hp = 0; hp < Hp; ++hp) {
    if (refl_f(hp - pad, H) != h) continue;
    for (int wp = 0; wp < Wp; ++wp)
      if (refl_f(wp - pad, W) == w) v += dxp[(((size_t)n * Hp + hp) * Wp + wp) * C + c];
  }
  dx[i] = v;
}

__global__ void act_bwd_f32_kernel(const float* __restrict__ y, const float* __restrict__ dy, float* __restrict__ dx, long long n, int act) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float d = 1.f;
  if (act != ACT_NONE) {
    const float v = y[i];
    d = act == ACT_LRELU ? (v > 0.f ? 1.f : 0.2f) : act == ACT_RELU ? (v > 0.f ? 1.f : 0.f) : act == ACT_TANH ? 1.f - v * v : v * (1.f - v);
  }
  dx[i] = dy[i] * d;
}
// column sums of g [M, C] (bias gradient) / sum_{m, c < C_real} g[m, c] * noise[m] (noise weight gradient)
__global__ void __launch_bounds__(1024)
colsum_f32_kernel(const float* __restrict__ g, long long M, int C, float* __restrict__ out) {
  __shared__ double sh[32][33];
  const int ch = threadIdx.x & 31, ln = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + ch;
  double a = 0.0;
  if (c < C)
    for (long long p = ln; p < M; p += 32) a += g[p * C + c];
  a = lanes_sum(a, sh);
  if (ln == 0 && c < C) out[c] = (float)a;
}
__global__ void __launch_bounds__(1024)
noise_wgrad_f32_kernel(const float* __restrict__ dy, const float* __restrict__ noise, long long M, int C, int C_real, float beta,
                       float* __restrict__ dw) {
  __shared__ double shd[1024];
  double a = 0.0;
  for (long long m = threadIdx.x; m < M; m += blockDim.x) {
    double t = 0.0;
    for (int c = 0; c < C_real; ++c) t += dy[m * C + c];
    a += t * noise[m];
  }
  shd[threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < (int)blockDim.x; ++k) t += shd[k];
    dw[0] = (float)((beta != 0.f ? (double)beta * dw[0] : 0.0) + t);
  }
}
static unsigned nb(long long n) { return (unsigned)((n + 255) / 256); }
}  // namespace dei2i

extern "C" int dei2i_nchw_to_nhwc_f32(const float* src, float* dst, int N, int C, int H, int W, int C_ld, dei2i_stream stream) {
  DEI2I_REQUIRE(src && dst && C <= C_ld && C_ld % 8 == 0, "nchw_to_nhwc_f32: bad arguments");
  nchw_to_nhwc_f32_kernel<<<nb((long long)N * H * W * C_ld), 256, 0, ST>>>(src, dst, N, C, H * W, C_ld);
  return LAUNCH_OK("nchw_to_nhwc_f32_kernel");
}
extern "C" int dei2i_nhwc_to_nchw_f32(const float* src, float* dst, int N, int C, int H, int W, int C_ld, dei2i_stream stream) {
  DEI2I_REQUIRE(src && dst && C <= C_ld, "nhwc_to_nchw_f32: bad arguments");
  nhwc_to_nchw_f32_kernel<<<nb((long long)N * C * H * W), 256, 0, ST>>>(src, dst, N, C, H * W, C_ld);
  return LAUNCH_OK("nhwc_to_nchw_f32_kernel");
}
extern "C" int dei2i_channel_stats_f32(const float* x, int N, int HW, int C, int per_sample, float* stats, dei2i_stream stream) {
  DEI2I_REQUIRE(x && stats && C > 0, "channel_stats_f32: bad arguments");
  const int groups = per_sample ? N : 1;
  channel_stats_f32_kernel<<<dim3((C + 31) / 32, groups), 1024, 0, ST>>>(x, per_sample ? (long long)HW : (long long)N * HW, C, stats);
  return LAUNCH_OK("channel_stats_f32_kernel");
}
extern "C" int dei2i_bn_fwd_f32(const float* x, float* y, const float* residual, const float* stats, const float* gamma, const float* beta,
                                float* running_mean, float* running_var, float* save_mean, float* save_rstd, int N, int HW, int C,
                                int training, float momentum, float eps, int act, dei2i_stream stream) {
  DEI2I_REQUIRE(x && y && gamma && beta && running_mean && running_var && (!training || stats), "bn_fwd_f32: bad arguments");
  const long long npix = (long long)N * HW;
  bn_fwd_f32_kernel<<<nb(npix * C), 256, 0, ST>>>(x, y, residual, stats, gamma, beta, running_mean, running_var, save_mean, save_rstd, npix,
                                                  C, training, momentum, eps, act);
  if (training) {
    bn_running_f32_kernel<<<(C + 255) / 256, 256, 0, ST>>>(stats, running_mean, running_var, npix, C, momentum);
    count_launch();
  }
  return LAUNCH_OK("bn_fwd_f32_kernel");
}
extern "C" int dei2i_bn_bwd_f32(const float* x, const float* dy, float* dx, const float* gamma, const float* beta, const float* save_mean,
                                const float* save_rstd, float* dbeta_dgamma /*[2, C]*/, int N, int HW, int C, int training, int act,
                                dei2i_stream stream) {
  DEI2I_REQUIRE(x && dy && dx && gamma && beta && save_mean && save_rstd && dbeta_dgamma, "bn_bwd_f32: bad arguments");
  const long long npix = (long long)N * HW;
  bn_bwd_reduce_f32_kernel<<<(C + 31) / 32, 1024, 0, ST>>>(x, dy, gamma, beta, save_mean, save_rstd, npix, C, act, dbeta_dgamma);
  count_launch();
  bn_bwd_apply_f32_kernel<<<nb(npix * C), 256, 0, ST>>>(x, dy, dx, gamma, beta, save_mean, save_rstd, dbeta_dgamma, npix, C, training, act);
  return LAUNCH_OK("bn_bwd_apply_f32_kernel");
}
extern "C" int dei2i_spade_cls_fwd_f32(const float* x, const float* gb5, float* out, const float* stats, float* mean_rstd, int N, int H,
                                       int W, int C, int up, int relu, float eps, int out_pad, dei2i_stream stream) {
  DEI2I_REQUIRE(x && gb5 && out && stats && mean_rstd && up >= 1 && H % up == 0 && W % up == 0 && H >= 5 && W >= 5, "spade_cls_fwd_f32: bad arguments");
  spade_cls_fwd_f32_kernel<<<nb((long long)N * (H + 2 * out_pad) * (W + 2 * out_pad) * C), 256, 0, ST>>>(x, gb5, out, stats, mean_rstd, N, H, W,
                                                                                                       C, up, relu, eps, out_pad);
  return LAUNCH_OK("spade_cls_fwd_f32_kernel");
}
extern "C" int dei2i_spade_cls_bwd_f32(const float* x, const float* gb5, const float* dout, float* dgb5, float* dx, const float* mean_rstd,
                                       float* red /*[N, 2, C] scratch*/, int N, int H, int W, int C, int up, int relu, int out_pad,
                                       dei2i_stream stream) {
  DEI2I_REQUIRE(x && gb5 && dout && dgb5 && dx && mean_rstd && red && up >= 1 && H >= 5 && W >= 5, "spade_cls_bwd_f32: bad arguments");
  spade_cls_bwd1_f32_kernel<<<dim3(25, N, (C + 31) / 32), 1024, 0, ST>>>(x, gb5, dout, mean_rstd, dgb5, H, W, C, up, relu, out_pad);
  spade_cls_red_f32_kernel<<<N, 256, 0, ST>>>(gb5, dgb5, red, C);
  count_launch(2);
  spade_cls_bwd2_f32_kernel<<<nb((long long)N * (H / up) * (W / up) * C), 256, 0, ST>>>(x, gb5, dout, dx, mean_rstd, red, N, H, W, C, up, relu,
                                                                                      out_pad);
  return LAUNCH_OK("spade_cls_bwd2_f32_kernel");
}
extern "C" int dei2i_spade_fwd_f32(const float* x, const float* gb, float* out, const float* stats, float* mean_rstd, int N, int H, int W,
                                   int C, int up, int relu, float eps, dei2i_stream stream) {
  DEI2I_REQUIRE(x && gb && out && stats && mean_rstd && up >= 1, "spade_fwd_f32: bad arguments");
  spade_fwd_f32_kernel<<<nb((long long)N * H * W * C), 256, 0, ST>>>(x, gb, out, stats, mean_rstd, N, H, W, C, up, relu, eps);
  return LAUNCH_OK("spade_fwd_f32_kernel");
}
extern "C" int dei2i_reflect_pad_f32(const float* x, float* xp, int N, int H, int W, int C, int pad, dei2i_stream stream) {
  DEI2I_REQUIRE(x && xp && pad >= 0 && pad < H && pad < W, "reflect_pad_f32: bad arguments");
  reflect_pad_f32_kernel<<<nb((long long)N * (H + 2 * pad) * (W + 2 * pad) * C), 256, 0, ST>>>(x, xp, N, H, W, C, pad);
  return LAUNCH_OK("reflect_pad_f32_kernel");
}
extern "C" int dei2i_reflect_fold_f32(const float* dxp, float* dx, int N, int H, int W, int C, int pad, dei2i_stream stream) {
  DEI2I_REQUIRE(dxp && dx && pad >= 0 && pad < H && pad < W, "reflect_fold_f32: bad arguments");
  reflect_fold_f32_kernel<<<nb((long long)N * H * W * C), 256, 0, ST>>>(dxp, dx, N, H, W, C, pad);
  return LAUNCH_OK("reflect_fold_f32_kernel");
}
extern "C" int dei2i_act_bwd_f32(const float* y, const float* dy, float* dx, float* bias_grad, long long M, int C, int act,
                                 dei2i_stream stream) {
  DEI2I_REQUIRE(dy && (y || act == DEI2I_ACT_NONE) && (dx || bias_grad), "act_bwd_f32: bad arguments");
  const float* g = dy;
  if (dx != nullptr) {
    act_bwd_f32_kernel<<<nb(M * C), 256, 0, ST>>>(y, dy, dx, M * C, act);
    count_launch();
    g = dx;
  } else {
    DEI2I_REQUIRE(act == DEI2I_ACT_NONE, "act_bwd_f32: bias gradient without dx needs act == none");
  }
  if (bias_grad != nullptr) {
    colsum_f32_kernel<<<(C + 31) / 32, 1024, 0, ST>>>(g, M, C, bias_grad);
    count_launch();
  }
  return check_cuda(cudaGetLastError(), "act_bwd_f32_kernel");
}
extern "C" int dei2i_noise_wgrad_f32(const float* dy, const float* noise, float* dw, long long M, int C, int C_real, float beta,
                                     dei2i_stream stream) {
  DEI2I_REQUIRE(dy && noise && dw, "noise_wgrad_f32: bad arguments");
  noise_wgrad_f32_kernel<<<1, 1024, 0, ST>>>(dy, noise, M, C, C_real, beta, dw);
  return LAUNCH_OK("noise_wgrad_f32_kernel");
}
