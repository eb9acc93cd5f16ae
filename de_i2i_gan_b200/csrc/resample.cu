// Resampling kernels of the WGAN / StarGAN-v2 paths: nearest up-sampling fused with zero padding, max / average pooling,
// global average pooling.  All HBM-streaming: one 16-byte (8 x bf16) access per thread and element group, NHWC.
#include "common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {
void count_launch(int n = 1);

static inline unsigned nblocks(long long n, int per) {
  long long b = (n + per - 1) / per;
  if (b > 148ll * 32) b = 148ll * 32;
  return (unsigned)(b < 1 ? 1 : b);
}
DEVINL void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
  f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}
DEVINL uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// out[n, up*h + dy + pt, up*w + dx + pl, :] = x[n, h, w, :]; the pad frame is zero.  (nn.Upsample(scale_factor=up, 'nearest')
// followed by the zero padding of nn.Conv2d(padding='same'), which is asymmetric for even kernels.)
__global__ void upsample_pad_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int N, int H, int W, int C8, int up,
                                        int pt, int pl, int Ho, int Wo) {
  const long long total = (long long)N * Ho * Wo * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    long long t = fast_div(i, C8);
    const int wo = fast_mod(t, Wo);
    t = fast_div(t, Wo);
    const int ho = fast_mod(t, Ho);
    const int n = (int)fast_div(t, Ho);
    const int hu = ho - pt, wu = wo - pl;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (hu >= 0 && hu < H * up && wu >= 0 && wu < W * up) v = __ldg(x + (((long long)n * H + hu / up) * W + wu / up) * C8 + c8);
    out[i] = v;
  }
}
// adjoint: dx[n,h,w,:] = sum_{dy,dx < up} dout[n, up*h + dy + pt, up*w + dx + pl, :]
__global__ void upsample_pad_bwd_kernel(const uint4* __restrict__ dout, uint4* __restrict__ dx, int N, int H, int W, int C8, int up,
                                        int pt, int pl, int Ho, int Wo) {
  const long long total = (long long)N * H * W * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    long long t = fast_div(i, C8);
    const int w = fast_mod(t, W);
    t = fast_div(t, W);
    const int h = fast_mod(t, H);
    const int n = (int)fast_div(t, H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int dy = 0; dy < up; ++dy)
      for (int dxx = 0; dxx < up; ++dxx) {
        float f[8];
        unpack8(__ldg(dout + (((long long)n * Ho + up * h + dy + pt) * Wo + up * w + dxx + pl) * C8 + c8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
      }
    dx[i] = pack8(acc);
  }
}

// nn.MaxPool2d(k, s, p): out = max over the window (padding = -inf); idx = window offset (r*k + c) of the FIRST maximum in
// scan order, which is the element torch routes the gradient to.
__global__ void maxpool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, uint2* __restrict__ idx, int N, int H, int W,
                                   int C8, int k, int s, int p, int Ho, int Wo) {
  const long long total = (long long)N * Ho * Wo * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    long long t = fast_div(i, C8);
    const int wo = fast_mod(t, Wo);
    t = fast_div(t, Wo);
    const int ho = fast_mod(t, Ho);
    const int n = (int)fast_div(t, Ho);
    float best[8];
    unsigned bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      best[j] = -INFINITY;
      bi[j] = 0;
    }
    bool first = true;
    for (int r = 0; r < k; ++r) {
      const int h = ho * s - p + r;
      if (h < 0 || h >= H) continue;
      for (int c = 0; c < k; ++c) {
        const int w = wo * s - p + c;
        if (w < 0 || w >= W) continue;
        float f[8];
        unpack8(__ldg(x + (((long long)n * H + h) * W + w) * C8 + c8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (first || f[j] > best[j]) {
            best[j] = f[j];
            bi[j] = (unsigned)(r * k + c);
          }
        first = false;
      }
    }
    out[i] = pack8(best);
    idx[i] = make_uint2(bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24), bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24));
  }
}
// gather form of the backward: input pixel (h, w) collects dout of every window whose recorded argmax is (h, w)
__global__ void maxpool_bwd_kernel(const uint4* __restrict__ dout, const uint2* __restrict__ idx, uint4* __restrict__ dx, int N, int H,
                                   int W, int C8, int k, int s, int p, int Ho, int Wo) {
  const long long total = (long long)N * H * W * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    long long t = fast_div(i, C8);
    const int w = fast_mod(t, W);
    t = fast_div(t, W);
    const int h = fast_mod(t, H);
    const int n = (int)fast_div(t, H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < k; ++r) {
      const int hn = h + p - r;
      if (hn < 0 || hn % s != 0) continue;
      const int ho = hn / s;
      if (ho >= Ho) continue;
      for (int c = 0; c < k; ++c) {
        const int wn = w + p - c;
        if (wn < 0 || wn % s != 0) continue;
        const int wo = wn / s;
        if (wo >= Wo) continue;
        const long long o = (((long long)n * Ho + ho) * Wo + wo) * C8 + c8;
        const uint2 id = __ldg(idx + o);
        float f[8];
        unpack8(__ldg(dout + o), f);
        const unsigned me = (unsigned)(r * k + c);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const unsigned b = ((j < 4 ? id.x : id.y) >> (8 * (j & 3))) & 255u;
          if (b == me) acc[j] += f[j];
        }
      }
    }
    dx[i] = pack8(acc);
  }
}

// F.avg_pool2d(x, k) (stride k, no padding; H, W multiples of k) and its adjoint, with an optional output scale
// (StarGAN-v2 folds the residual 1/sqrt(2) here).
__global__ void avgpool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int N, int H, int W, int C8, int k, float scale) {
  const int Ho = H / k, Wo = W / k;
  const long long total = (long long)N * Ho * Wo * C8;
  const float m = scale / (float)(k * k);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    long long t = fast_div(i, C8);
    const int wo = fast_mod(t, Wo);
    t = fast_div(t, Wo);
    const int ho = fast_mod(t, Ho);
    const int n = (int)fast_div(t, Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < k; ++r)
      for (int c = 0; c < k; ++c) {
        float f[8];
        unpack8(__ldg(x + (((long long)n * H + ho * k + r) * W + wo * k + c) * C8 + c8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= m;
    out[i] = pack8(acc);
  }
}
__global__ void avgpool_bwd_kernel(const uint4* __restrict__ dout, uint4* __restrict__ dx, int N, int H, int W, int C8, int k, float scale) {
  const int Ho = H / k, Wo = W / k;
  const long long total = (long long)N * H * W * C8;
  const float m = scale / (float)(k * k);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    long long t = fast_div(i, C8);
    const int w = fast_mod(t, W);
    t = fast_div(t, W);
    const int h = fast_mod(t, H);
    const int n = (int)fast_div(t, H);
    float f[8];
    unpack8(__ldg(dout + (((long long)n * Ho + h / k) * Wo + w / k) * C8 + c8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= m;
    dx[i] = pack8(f);
  }
}

// nn.AdaptiveAvgPool2d(1): out fp32 [N, C] = mean over H*W; block per (image, 8-channel group) row of 32 lanes
__global__ void gap_fwd_kernel(const uint4* __restrict__ x, float* __restrict__ out, int HW, int C8) {
  const int n = blockIdx.y;
  const int c8 = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  if (c8 >= C8) return;
  const int lane = threadIdx.x & 31;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int p = lane; p < HW; p += 32) {
    float f[8];
    unpack8(__ldg(x + ((long long)n * HW + p) * C8 + c8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += f[j];
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float v = warp_sum(acc[j]);
    if (lane == 0) out[(long long)n * C8 * 8 + c8 * 8 + j] = v / (float)HW;
  }
}
__global__ void gap_bwd_kernel(const float* __restrict__ dout, uint4* __restrict__ dx, int N, int HW, int C8) {
  const long long total = (long long)N * HW * C8;
  const float m = 1.f / (float)HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = fast_mod(i, C8);
    const int n = (int)(i / ((long long)HW * C8));
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = __ldg(dout + (long long)n * C8 * 8 + c8 * 8 + j) * m;
    dx[i] = pack8(f);
  }
}

}  // namespace dei2i

using namespace dei2i;
#define ST ((cudaStream_t)stream)
#define LAUNCH_OK(name) (count_launch(), check_cuda(cudaGetLastError(), name))

extern "C" int dei2i_upsample_pad_fwd(const void* x, void* out, int N, int H, int W, int C, int up, int pad_top, int pad_left,
                                      int pad_bottom, int pad_right, dei2i_stream stream) {
  DEI2I_REQUIRE(x && out && C % 8 == 0 && up >= 1 && pad_top >= 0 && pad_left >= 0 && pad_bottom >= 0 && pad_right >= 0,
                "upsample_pad_fwd: bad arguments");
  const int Ho = H * up + pad_top + pad_bottom, Wo = W * up + pad_left + pad_right;
  upsample_pad_fwd_kernel<<<nblocks((long long)N * Ho * Wo * (C / 8), 256), 256, 0, ST>>>((const uint4*)x, (uint4*)out, N, H, W, C / 8, up,
                                                                                          pad_top, pad_left, Ho, Wo);
  return LAUNCH_OK("upsample_pad_fwd_kernel");
}
extern "C" int dei2i_upsample_pad_bwd(const void* dout, void* dx, int N, int H, int W, int C, int up, int pad_top, int pad_left,
                                      int pad_bottom, int pad_right, dei2i_stream stream) {
  DEI2I_REQUIRE(dout && dx && C % 8 == 0 && up >= 1 && pad_top >= 0 && pad_left >= 0 && pad_bottom >= 0 && pad_right >= 0,
                "upsample_pad_bwd: bad arguments");
  const int Ho = H * up + pad_top + pad_bottom, Wo = W * up + pad_left + pad_right;
  upsample_pad_bwd_kernel<<<nblocks((long long)N * H * W * (C / 8), 256), 256, 0, ST>>>((const uint4*)dout, (uint4*)dx, N, H, W, C / 8, up,
                                                                                        pad_top, pad_left, Ho, Wo);
  return LAUNCH_OK("upsample_pad_bwd_kernel");
}

extern "C" int dei2i_maxpool_fwd(const void* x, void* out, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                                 dei2i_stream stream) {
  DEI2I_REQUIRE(x && out && idx && C % 8 == 0 && k >= 1 && k <= 15 && stride >= 1 && pad >= 0 && 2 * pad <= k, "maxpool_fwd: bad arguments");
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  DEI2I_REQUIRE(Ho > 0 && Wo > 0, "maxpool_fwd: empty output");
  maxpool_fwd_kernel<<<nblocks((long long)N * Ho * Wo * (C / 8), 256), 256, 0, ST>>>((const uint4*)x, (uint4*)out, (uint2*)idx, N, H, W,
                                                                                     C / 8, k, stride, pad, Ho, Wo);
  return LAUNCH_OK("maxpool_fwd_kernel");
}
extern "C" int dei2i_maxpool_bwd(const void* dout, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride, int pad,
                                 dei2i_stream stream) {
  DEI2I_REQUIRE(dout && idx && dx && C % 8 == 0 && k >= 1 && k <= 15 && stride >= 1 && pad >= 0, "maxpool_bwd: bad arguments");
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  maxpool_bwd_kernel<<<nblocks((long long)N * H * W * (C / 8), 256), 256, 0, ST>>>((const uint4*)dout, (const uint2*)idx, (uint4*)dx, N, H,
                                                                                   W, C / 8, k, stride, pad, Ho, Wo);
  return LAUNCH_OK("maxpool_bwd_kernel");
}

extern "C" int dei2i_avgpool_fwd(const void* x, void* out, int N, int H, int W, int C, int k, float scale, dei2i_stream stream) {
  DEI2I_REQUIRE(x && out && C % 8 == 0 && k >= 1 && H % k == 0 && W % k == 0, "avgpool_fwd: bad arguments");
  avgpool_fwd_kernel<<<nblocks((long long)N * (H / k) * (W / k) * (C / 8), 256), 256, 0, ST>>>((const uint4*)x, (uint4*)out, N, H, W, C / 8,
                                                                                               k, scale);
  return LAUNCH_OK("avgpool_fwd_kernel");
}
extern "C" int dei2i_avgpool_bwd(const void* dout, void* dx, int N, int H, int W, int C, int k, float scale, dei2i_stream stream) {
  DEI2I_REQUIRE(dout && dx && C % 8 == 0 && k >= 1 && H % k == 0 && W % k == 0, "avgpool_bwd: bad arguments");
  avgpool_bwd_kernel<<<nblocks((long long)N * H * W * (C / 8), 256), 256, 0, ST>>>((const uint4*)dout, (uint4*)dx, N, H, W, C / 8, k, scale);
  return LAUNCH_OK("avgpool_bwd_kernel");
}

extern "C" int dei2i_gap_fwd(const void* x, float* out, int N, int HW, int C, dei2i_stream stream) {
  DEI2I_REQUIRE(x && out && C % 8 == 0 && HW >= 1, "gap_fwd: bad arguments");
  const int C8 = C / 8;
  gap_fwd_kernel<<<dim3((C8 + 7) / 8, N), 256, 0, ST>>>((const uint4*)x, out, HW, C8);
  return LAUNCH_OK("gap_fwd_kernel");
}
extern "C" int dei2i_gap_bwd(const float* dout, void* dx, int N, int HW, int C, dei2i_stream stream) {
  DEI2I_REQUIRE(dout && dx && C % 8 == 0 && HW >= 1, "gap_bwd: bad arguments");
  gap_bwd_kernel<<<nblocks((long long)N * HW * (C / 8), 256), 256, 0, ST>>>(dout, (uint4*)dx, N, HW, C / 8);
  return LAUNCH_OK("gap_bwd_kernel");
}
