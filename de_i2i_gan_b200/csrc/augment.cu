// DiffAugment (defectGAN/utils/diffaug.py:9-75) as fused streaming kernels on FP32 NCHW images - the tensors that sit between the
// generator output and the discriminator input (defectgan_model.py:200-203, 266-271).
//
//   color        x1 = x + (b - 0.5);  x2 = (x1 - mean_c(x1)) * 2 s + mean_c(x1);  x3 = (x2 - mean_chw(x2)) * (k + 0.5) + mean_chw(x2)
//   translation  y[i, j] = x3[i + tx, j + ty], zero outside the image (F.pad by one + clamped gather)
//   cutout       y[i, j] = 0 inside the (clamped) ch x cw window centred at (ox, oy)
//
// The colour ops are point-wise, so they commute with the translation; saturation keeps every pixel's channel mean and brightness
// adds a constant, hence mean_chw(x2) = mean_chw(x) + (b - 0.5): ONE reduction over the input, then ONE pass that reads x at the
// translated position and writes y.  Backward: g3 = adjoint(translation, cutout)(dy); g2 = K g3 + (1 - K) mean_chw(g3);
// g1 = S g2 + (1 - S) mean_c(g2) - again one reduction (over dy, restricted to the positions that survive the window and the
// frame) and one pass.  Random parameters are drawn on the host side with the reference's torch calls; reductions are block-partial
// sums finished in a fixed order (deterministic).
#include <string.h>

#include "common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {
void count_launch(int n = 1);

struct AugGeom {
  int N, C, H, W, policy;
  int ch, cw;  // cutout window
};

// params per image: [0] b - 0.5, [1] S = 2 s, [2] K = k + 0.5, [3] tx, [4] ty, [5] ox, [6] oy, [7] unused
DEVINL bool cut_out(const AugGeom& g, const float* __restrict__ p, int i, int j) {
  if (!(g.policy & DEI2I_AUG_CUTOUT)) return false;
  const int ox = (int)p[5], oy = (int)p[6];
  // rows clamp(ox - ch/2 + [0, ch), 0, H - 1): a contiguous range (clamping only shortens it)
  const int r0 = max(ox - g.ch / 2, 0), r1 = min(ox - g.ch / 2 + g.ch - 1, g.H - 1);
  const int c0 = max(oy - g.cw / 2, 0), c1 = min(oy - g.cw / 2 + g.cw - 1, g.W - 1);
  const int rr0 = min(r0, g.H - 1), rr1 = max(r1, 0), cc0 = min(c0, g.W - 1), cc1 = max(c1, 0);
  return i >= rr0 && i <= rr1 && j >= cc0 && j <= cc1;
}

// part[n][blk] = sum over this block's pixels of sum_c v[n, c, i, j] * keep(i, j)   (keep == 1 for the forward: plain image sum)
template <bool BWD>
__global__ void __launch_bounds__(256)
aug_sum_kernel(const float* __restrict__ v, const float* __restrict__ params, AugGeom g, float* __restrict__ part) {
  __shared__ float sh[32];
  const int n = blockIdx.y;
  const float* p = params + n * 8;
  const int HW = g.H * g.W;
  const int tx = (g.policy & DEI2I_AUG_TRANSLATION) ? (int)p[3] : 0, ty = (g.policy & DEI2I_AUG_TRANSLATION) ? (int)p[4] : 0;
  float acc = 0.f;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    bool keep = true;
    if (BWD) {  // output position (i, j) contributes iff it is outside the cutout window and its source lies inside the image
      const int i = px / g.W, j = px - i * g.W;
      const int si = i + tx, sj = j + ty;
      keep = !cut_out(g, p, i, j) && si >= 0 && si < g.H && sj >= 0 && sj < g.W;
    }
    if (keep)
      for (int c = 0; c < g.C; ++c) acc += v[((size_t)n * g.C + c) * HW + px];
  }
  acc = warp_sum(acc);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sh[w] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += sh[k];
    part[(size_t)n * gridDim.x + blockIdx.x] = t;
  }
}

__global__ void aug_sum_finish_kernel(const float* __restrict__ part, int nblk, float scale, float* __restrict__ out) {
  const int n = threadIdx.x;  // one thread per image, blocks summed in order
  float t = 0.f;
  for (int b = 0; b < nblk; ++b) t += part[(size_t)n * nblk + b];
  out[n] = t * scale;
}

constexpr int kAugMaxC = 4;

__global__ void __launch_bounds__(256)
aug_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ params, const float* __restrict__ mean,
               AugGeom g) {
  const int n = blockIdx.y;
  const float* p = params + n * 8;
  const int HW = g.H * g.W;
  const bool color = g.policy & DEI2I_AUG_COLOR;
  const float b = color ? p[0] : 0.f, S = color ? p[1] : 1.f, K = color ? p[2] : 1.f;
  const float M = color ? mean[n] + b : 0.f;
  const int tx = (g.policy & DEI2I_AUG_TRANSLATION) ? (int)p[3] : 0, ty = (g.policy & DEI2I_AUG_TRANSLATION) ? (int)p[4] : 0;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    const int i = px / g.W, j = px - i * g.W;
    const int si = i + tx, sj = j + ty;
    const bool live = !cut_out(g, p, i, j) && si >= 0 && si < g.H && sj >= 0 && sj < g.W;
    float v[kAugMaxC];
    float m = 0.f;
#pragma unroll
    for (int c = 0; c < kAugMaxC; ++c)
      if (c < g.C) {
        v[c] = live ? x[((size_t)n * g.C + c) * HW + si * g.W + sj] + b : 0.f;
        m += v[c];
      }
    m /= (float)g.C;
#pragma unroll
    for (int c = 0; c < kAugMaxC; ++c)
      if (c < g.C) {
        const float x2 = (v[c] - m) * S + m;
        y[((size_t)n * g.C + c) * HW + px] = live ? (x2 - M) * K + M : 0.f;
      }
  }
}

// dx[n, c, si, sj] for every SOURCE position: the gradient arrives from output (si - tx, sj - ty) if that lies in the image and outside
// the cutout window; the two mean terms reach every source position.
__global__ void __launch_bounds__(256)
aug_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, const float* __restrict__ params, const float* __restrict__ gmean,
               AugGeom g) {
  const int n = blockIdx.y;
  const float* p = params + n * 8;
  const int HW = g.H * g.W;
  const bool color = g.policy & DEI2I_AUG_COLOR;
  const float S = color ? p[1] : 1.f, K = color ? p[2] : 1.f;
  const float G = color ? gmean[n] : 0.f;  // mean over (c, h, w) of g3
  const int tx = (g.policy & DEI2I_AUG_TRANSLATION) ? (int)p[3] : 0, ty = (g.policy & DEI2I_AUG_TRANSLATION) ? (int)p[4] : 0;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    const int si = px / g.W, sj = px - si * g.W;
    const int i = si - tx, j = sj - ty;
    const bool live = i >= 0 && i < g.H && j >= 0 && j < g.W && !cut_out(g, p, i, j);
    float g2[kAugMaxC];
    float m = 0.f;
#pragma unroll
    for (int c = 0; c < kAugMaxC; ++c)
      if (c < g.C) {
        const float g3 = live ? dy[((size_t)n * g.C + c) * HW + i * g.W + j] : 0.f;
        g2[c] = K * g3 + (1.f - K) * G;
        m += g2[c];
      }
    m /= (float)g.C;
#pragma unroll
    for (int c = 0; c < kAugMaxC; ++c)
      if (c < g.C) dx[((size_t)n * g.C + c) * HW + px] = S * g2[c] + (1.f - S) * m;
  }
}

static int aug_geom(AugGeom& g, int N, int C, int H, int W, int policy) {
  DEI2I_REQUIRE(N > 0 && N <= 1024 && C > 0 && C <= kAugMaxC && H > 0 && W > 0, "diffaug: bad sizes (C <= %d, N <= 1024)", kAugMaxC);
  DEI2I_REQUIRE((policy & ~(DEI2I_AUG_COLOR | DEI2I_AUG_TRANSLATION | DEI2I_AUG_CUTOUT)) == 0, "diffaug: unknown policy bits");
  g.N = N; g.C = C; g.H = H; g.W = W; g.policy = policy;
  g.ch = (int)(H * 0.5 + 0.5); g.cw = (int)(W * 0.5 + 0.5);  // diffaug.py:56 (ratio 0.5)
  return 0;
}
static int aug_blocks(int H, int W, int N) {
  int b = (H * W + 1023) / 1024;
  const int cap = (148 * 8 + N - 1) / N;
  return b < 1 ? 1 : (b > cap ? cap : b);
}

}  // namespace dei2i

using namespace dei2i;
#define ST ((cudaStream_t)stream)

extern "C" size_t dei2i_diffaug_scratch(int N, int H, int W) { return (size_t)N * (aug_blocks(H, W, N) + 1); }

extern "C" int dei2i_diffaug_fwd(const float* x, float* y, const float* params, float* scratch, int N, int C, int H, int W, int policy,
                                 dei2i_stream stream) {
  DEI2I_REQUIRE(x && y && params && scratch, "diffaug_fwd: null pointer");
  AugGeom g;
  if (int rc = aug_geom(g, N, C, H, W, policy)) return rc;
  const int nb = aug_blocks(H, W, N);
  float* mean = scratch + (size_t)N * nb;
  if (policy & DEI2I_AUG_COLOR) {
    aug_sum_kernel<false><<<dim3(nb, N), 256, 0, ST>>>(x, params, g, scratch);
    aug_sum_finish_kernel<<<1, N, 0, ST>>>(scratch, nb, 1.f / ((float)C * H * W), mean);
    count_launch(2);
  }
  aug_fwd_kernel<<<dim3(nb, N), 256, 0, ST>>>(x, y, params, mean, g);
  count_launch();
  return check_cuda(cudaGetLastError(), "aug_fwd_kernel");
}

extern "C" int dei2i_diffaug_bwd(const float* dy, float* dx, const float* params, float* scratch, int N, int C, int H, int W, int policy,
                                 dei2i_stream stream) {
  DEI2I_REQUIRE(dy && dx && params && scratch, "diffaug_bwd: null pointer");
  AugGeom g;
  if (int rc = aug_geom(g, N, C, H, W, policy)) return rc;
  const int nb = aug_blocks(H, W, N);
  float* gmean = scratch + (size_t)N * nb;
  if (policy & DEI2I_AUG_COLOR) {
    aug_sum_kernel<true><<<dim3(nb, N), 256, 0, ST>>>(dy, params, g, scratch);
    aug_sum_finish_kernel<<<1, N, 0, ST>>>(scratch, nb, 1.f / ((float)C * H * W), gmean);
    count_launch(2);
  }
  aug_bwd_kernel<<<dim3(nb, N), 256, 0, ST>>>(dy, dx, params, gmean, g);
  count_launch();
  return check_cuda(cudaGetLastError(), "aug_bwd_kernel");
}

// ---------------------------------------------------------------------------------------------------------------------
// MAE pre-training input (defectgan_model.py:361-383): masked_imgs = imgs * m + token * (1 - m)  (MaskToken.forward,
// architecture.py:411-414) with m the shifted patch mask of generate_shifted_mask (utils/util.py:60-71): a coarse Bernoulli
// grid [N, gh, gw] nearest-upsampled by patch_size and cropped at (h_shift, w_shift), evaluated on the fly from the coarse
// grid - the full-resolution mask is written once as an output because the caller returns it.
// token: broadcastable [tn, tc, th, tw] with every extent 1 or full (zero / scalar / vector / position / full / per-image mean).
namespace dei2i {
struct MaskGeom {
  int N, C, H, W, patch, hs, ws, gh, gw;
  int tn, tc, th, tw;  // token extents
};
DEVINL float coarse_mask(const float* __restrict__ coarse, const MaskGeom& g, int n, int i, int j) {
  return coarse[((size_t)n * g.gh + (i + g.hs) / g.patch) * g.gw + (j + g.ws) / g.patch];
}
__global__ void __launch_bounds__(256)
mask_token_fwd_kernel(const float* __restrict__ img, const float* __restrict__ coarse, const float* __restrict__ token,
                      float* __restrict__ out, float* __restrict__ mask, MaskGeom g) {
  const int n = blockIdx.y;
  const int HW = g.H * g.W;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    const int i = px / g.W, j = px - i * g.W;
    const float m = coarse_mask(coarse, g, n, i, j);
    if (mask != nullptr) mask[(size_t)n * HW + px] = m;
    for (int c = 0; c < g.C; ++c) {
      float t = 0.f;
      if (token != nullptr)
        t = token[(((size_t)(g.tn > 1 ? n : 0) * g.tc + (g.tc > 1 ? c : 0)) * g.th + (g.th > 1 ? i : 0)) * g.tw + (g.tw > 1 ? j : 0)];
      const size_t o = ((size_t)n * g.C + c) * HW + px;
      out[o] = img[o] * m + t * (1.f - m);
    }
  }
}
// per-(n, c) partial sums of img * m over the block's pixels ('mean' token: masked_imgs.mean((2,3)) / mask_ratio, architecture.py:416-418)
__global__ void __launch_bounds__(256)
mask_mean_kernel(const float* __restrict__ img, const float* __restrict__ coarse, float* __restrict__ part, MaskGeom g) {
  __shared__ float sh[32];
  const int n = blockIdx.y, c = blockIdx.z;
  const int HW = g.H * g.W;
  float acc = 0.f;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    const int i = px / g.W, j = px - i * g.W;
    acc += img[((size_t)n * g.C + c) * HW + px] * coarse_mask(coarse, g, n, i, j);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += sh[k];
    part[((size_t)n * g.C + c) * gridDim.x + blockIdx.x] = t;
  }
}
// gradient of a learnable token: d_token[e] = sum over the broadcast dimensions of dy * (1 - m).  One thread per token element for
// the spatial tokens (position / full: loops over N [and C]); the scalar / vector tokens go through per-block partial sums.
__global__ void __launch_bounds__(256)
mask_token_bwd_spatial_kernel(const float* __restrict__ dy, const float* __restrict__ coarse, float* __restrict__ dtoken, MaskGeom g) {
  const int HW = g.H * g.W;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= g.tc * HW) return;
  const int tc_i = e / HW, px = e - tc_i * HW;
  const int i = px / g.W, j = px - i * g.W;
  float acc = 0.f;
  for (int n = 0; n < g.N; ++n) {
    const float w = 1.f - coarse_mask(coarse, g, n, i, j);
    if (w == 0.f) continue;
    if (g.tc > 1) acc += dy[((size_t)n * g.C + tc_i) * HW + px] * w;
    else
      for (int c = 0; c < g.C; ++c) acc += dy[((size_t)n * g.C + c) * HW + px] * w;
  }
  dtoken[e] = acc;
}
__global__ void __launch_bounds__(256)
mask_token_bwd_channel_kernel(const float* __restrict__ dy, const float* __restrict__ coarse, float* __restrict__ part, MaskGeom g) {
  __shared__ float sh[32];
  const int n = blockIdx.y, c = blockIdx.z;
  const int HW = g.H * g.W;
  float acc = 0.f;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    const int i = px / g.W, j = px - i * g.W;
    acc += dy[((size_t)n * g.C + c) * HW + px] * (1.f - coarse_mask(coarse, g, n, i, j));
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += sh[k];
    part[((size_t)c * g.N + n) * gridDim.x + blockIdx.x] = t;  // [C][N * blocks]: channel-major so that a channel's partials are contiguous
  }
}
// out[v] = scale * sum_b part[v * nblk + b]  (fixed order, one thread per output)
__global__ void rows_sum_kernel(const float* __restrict__ part, int V, int nblk, float scale, float* __restrict__ out) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  float t = 0.f;
  for (int b = 0; b < nblk; ++b) t += part[(size_t)v * nblk + b];
  out[v] = t * scale;
}
static int mask_geom(MaskGeom& g, int N, int C, int H, int W, int patch, int hs, int ws, int gh, int gw, const int* tdims) {
  DEI2I_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0 && patch > 0 && hs >= 0 && hs < patch && ws >= 0 && ws < patch, "mask_token: bad geometry");
  DEI2I_REQUIRE((H + hs + patch - 1) / patch <= gh && (W + ws + patch - 1) / patch <= gw, "mask_token: coarse grid %dx%d too small", gh, gw);
  g.N = N; g.C = C; g.H = H; g.W = W; g.patch = patch; g.hs = hs; g.ws = ws; g.gh = gh; g.gw = gw;
  g.tn = tdims ? tdims[0] : 1; g.tc = tdims ? tdims[1] : 1; g.th = tdims ? tdims[2] : 1; g.tw = tdims ? tdims[3] : 1;
  DEI2I_REQUIRE((g.tn == 1 || g.tn == N) && (g.tc == 1 || g.tc == C) && (g.th == 1 || g.th == H) && (g.tw == 1 || g.tw == W),
                "mask_token: token extents must be 1 or the image's");
  return 0;
}
}  // namespace dei2i

extern "C" size_t dei2i_mask_token_scratch(int N, int C, int H, int W) { return (size_t)N * C * (aug_blocks(H, W, N) + 1); }

extern "C" int dei2i_mask_token_fwd(const float* img, const float* coarse, const float* token, const int* token_dims, float* out,
                                    float* mask, int N, int C, int H, int W, int patch, int h_shift, int w_shift, int gh, int gw,
                                    dei2i_stream stream) {
  DEI2I_REQUIRE(img && coarse && out && (token == nullptr || token_dims != nullptr), "mask_token_fwd: null pointer");
  MaskGeom g;
  if (int rc = mask_geom(g, N, C, H, W, patch, h_shift, w_shift, gh, gw, token ? token_dims : nullptr)) return rc;
  mask_token_fwd_kernel<<<dim3(aug_blocks(H, W, N), N), 256, 0, ST>>>(img, coarse, token, out, mask, g);
  count_launch();
  return check_cuda(cudaGetLastError(), "mask_token_fwd_kernel");
}

/* token_mean[n, c] = mean_{h,w}(img * m) / mask_ratio */
extern "C" int dei2i_mask_mean(const float* img, const float* coarse, float* token_mean, float* scratch, float mask_ratio, int N, int C,
                               int H, int W, int patch, int h_shift, int w_shift, int gh, int gw, dei2i_stream stream) {
  DEI2I_REQUIRE(img && coarse && token_mean && scratch && mask_ratio > 0.f, "mask_mean: bad arguments");
  MaskGeom g;
  if (int rc = mask_geom(g, N, C, H, W, patch, h_shift, w_shift, gh, gw, nullptr)) return rc;
  const int nb = aug_blocks(H, W, N);
  mask_mean_kernel<<<dim3(nb, N, C), 256, 0, ST>>>(img, coarse, scratch, g);
  rows_sum_kernel<<<(N * C + 255) / 256, 256, 0, ST>>>(scratch, N * C, nb, 1.f / ((float)H * W * mask_ratio), token_mean);
  count_launch(2);
  return check_cuda(cudaGetLastError(), "mask_mean_kernel");
}

extern "C" int dei2i_mask_token_bwd(const float* dy, const float* coarse, const int* token_dims, float* dtoken, float* scratch, int N,
                                    int C, int H, int W, int patch, int h_shift, int w_shift, int gh, int gw, dei2i_stream stream) {
  DEI2I_REQUIRE(dy && coarse && token_dims && dtoken && scratch, "mask_token_bwd: null pointer");
  MaskGeom g;
  if (int rc = mask_geom(g, N, C, H, W, patch, h_shift, w_shift, gh, gw, token_dims)) return rc;
  DEI2I_REQUIRE(g.tn == 1 && ((g.th == 1 && g.tw == 1) || (g.th == H && g.tw == W)), "mask_token_bwd: learnable tokens are [1, 1|C, 1, 1] or [1, 1|C, H, W]");
  if (g.th > 1) {
    mask_token_bwd_spatial_kernel<<<(g.tc * H * W + 255) / 256, 256, 0, ST>>>(dy, coarse, dtoken, g);
    count_launch();
    return check_cuda(cudaGetLastError(), "mask_token_bwd_spatial_kernel");
  }
  const int nb = aug_blocks(H, W, N);
  mask_token_bwd_channel_kernel<<<dim3(nb, N, C), 256, 0, ST>>>(dy, coarse, scratch, g);
  if (g.tc > 1) rows_sum_kernel<<<1, 256, 0, ST>>>(scratch, C, N * nb, 1.f, dtoken);
  else rows_sum_kernel<<<1, 32, 0, ST>>>(scratch, 1, C * N * nb, 1.f, dtoken);
  count_launch(2);
  return check_cuda(cudaGetLastError(), "mask_token_bwd_channel_kernel");
}

// ---------------------------------------------------------------------------------------------------------------------
// Training input pipeline on the device (train_defectgan.py:57-65): Resize(1.5 S) -> RandomResizedCrop(S, scale (0.6, 1)) ->
// random horizontal / vertical flip -> ColorJitter(brightness, saturation, contrast 0.2) -> ToTensor -> Normalize(0.5, 0.5),
// as ONE kernel from the decoded uint8 HWC image to the FP32 NCHW batch the trainer consumes.  The two resampling stages
// compose to one affine map from output pixel to source pixel; it is sampled with the triangle (bilinear, antialiased) filter
// PIL / torchvision use, support = max(1, source pixels per output pixel).  Random parameters come from the host
// (data/pipeline.py): params fp32 [N, 12] = crop box in SOURCE pixels (y0, x0, h, w), flip_h, flip_v, jitter factors
// (brightness, contrast, saturation) and their order (3 slots, op ids 0 / 1 / 2).
namespace dei2i {
DEVINL float tri(float t) { t = fabsf(t); return t < 1.f ? 1.f - t : 0.f; }
__global__ void __launch_bounds__(256)
augment_train_kernel(const uint8_t* __restrict__ src, const long long* __restrict__ src_off, const int* __restrict__ src_hw,
                     const float* __restrict__ params, float* __restrict__ out, int N, int S) {
  const int n = blockIdx.y;
  const float* p = params + n * 12;
  const int H = src_hw[2 * n], W = src_hw[2 * n + 1];
  const uint8_t* img = src + src_off[n];
  const float y0 = p[0], x0 = p[1], ch = p[2], cw = p[3];
  const bool fh = p[4] > 0.5f, fv = p[5] > 0.5f;
  const float sy = ch / (float)S, sx = cw / (float)S;           // source pixels per output pixel
  const float ry = fmaxf(sy, 1.f), rx = fmaxf(sx, 1.f);         // filter support (antialias when down-sampling)
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < S * S; px += gridDim.x * blockDim.x) {
    int oy = px / S, ox = px - oy * S;
    const int ty = fv ? S - 1 - oy : oy, tx = fh ? S - 1 - ox : ox;  // flips act on the cropped image
    const float cy = y0 + (ty + 0.5f) * sy, cx = x0 + (tx + 0.5f) * sx;  // centre in source coordinates (pixel centres at +0.5)
    // taps stay inside the crop box (the crop happens before the final resize, so its edge is the image edge for the filter)
    const int ya = max((int)floorf(cy - ry), max((int)floorf(y0), 0)), yb = min((int)ceilf(cy + ry), min((int)ceilf(y0 + ch), H));
    const int xa = max((int)floorf(cx - rx), max((int)floorf(x0), 0)), xb = min((int)ceilf(cx + rx), min((int)ceilf(x0 + cw), W));
    float acc[3] = {0.f, 0.f, 0.f}, wsum = 0.f;
    for (int y = ya; y < yb; ++y) {
      const float wy = tri((y + 0.5f - cy) / ry);
      if (wy == 0.f) continue;
      for (int x = xa; x < xb; ++x) {
        const float w = wy * tri((x + 0.5f - cx) / rx);
        if (w == 0.f) continue;
        const uint8_t* q = img + ((size_t)y * W + x) * 3;
        acc[0] += w * q[0]; acc[1] += w * q[1]; acc[2] += w * q[2];
        wsum += w;
      }
    }
    float v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) v[c] = wsum > 0.f ? acc[c] / (wsum * 255.f) : 0.f;
    out[((size_t)n * 3 + 0) * S * S + px] = v[0];
    out[((size_t)n * 3 + 1) * S * S + px] = v[1];
    out[((size_t)n * 3 + 2) * S * S + px] = v[2];
  }
}
// ColorJitter + Normalize on the resampled [0, 1] batch, in place.  order: 3 op ids (0 brightness, 1 contrast, 2 saturation);
// contrast blends with the mean grey level of the CURRENT image, so each contrast op is preceded by a per-image reduction.
__global__ void __launch_bounds__(256)
jitter_gray_sum_kernel(const float* __restrict__ x, float* __restrict__ part, int S2) {
  __shared__ float sh[32];
  const int n = blockIdx.y;
  float acc = 0.f;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < S2; px += gridDim.x * blockDim.x) {
    const float* q = x + (size_t)n * 3 * S2 + px;
    acc += 0.299f * q[0] + 0.587f * q[S2] + 0.114f * q[2 * S2];
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += sh[k];
    part[(size_t)n * gridDim.x + blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(256)
jitter_apply_kernel(float* __restrict__ x, const float* __restrict__ params, const float* __restrict__ gray_mean, int S2, int op_slot,
                    int normalize) {
  const int n = blockIdx.y;
  const float* p = params + n * 12;
  const int op = op_slot >= 0 ? (int)p[9 + op_slot] : -1;
  const float f = op >= 0 ? p[6 + op] : 1.f;
  const float gm = op == 1 ? gray_mean[n] : 0.f;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < S2; px += gridDim.x * blockDim.x) {
    float* q = x + (size_t)n * 3 * S2 + px;
    float v[3] = {q[0], q[S2], q[2 * S2]};
    if (op == 0) {  // brightness: blend with black
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = fminf(fmaxf(v[c] * f, 0.f), 1.f);
    } else if (op == 1) {  // contrast: blend with the mean grey level
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = fminf(fmaxf(f * v[c] + (1.f - f) * gm, 0.f), 1.f);
    } else if (op == 2) {  // saturation: blend with the pixel's grey level
      const float g = 0.299f * v[0] + 0.587f * v[1] + 0.114f * v[2];
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = fminf(fmaxf(f * v[c] + (1.f - f) * g, 0.f), 1.f);
    }
    if (normalize) {
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = (v[c] - 0.5f) * 2.f;
    }
    q[0] = v[0]; q[S2] = v[1]; q[2 * S2] = v[2];
  }
}
}  // namespace dei2i

extern "C" size_t dei2i_augment_scratch(int N, int S) { return (size_t)N * (aug_blocks(S, S, N) + 1); }

extern "C" int dei2i_augment_train(const void* src_u8, const long long* src_offsets, const int* src_hw, const float* params, float* out,
                                   float* scratch, int N, int S, int jitter, dei2i_stream stream) {
  DEI2I_REQUIRE(src_u8 && src_offsets && src_hw && params && out && scratch && N > 0 && N <= 1024 && S > 0, "augment_train: bad arguments");
  const int nb = aug_blocks(S, S, N);
  augment_train_kernel<<<dim3(nb, N), 256, 0, ST>>>((const uint8_t*)src_u8, src_offsets, src_hw, params, out, N, S);
  count_launch();
  float* gm = scratch + (size_t)N * nb;
  for (int slot = 0; slot < (jitter ? 3 : 0); ++slot) {
    // the grey mean is only consumed by a contrast op; computing it for every slot keeps the launch sequence static (graph friendly)
    jitter_gray_sum_kernel<<<dim3(nb, N), 256, 0, ST>>>(out, scratch, S * S);
    aug_sum_finish_kernel<<<1, N, 0, ST>>>(scratch, nb, 1.f / ((float)S * S), gm);
    jitter_apply_kernel<<<dim3(nb, N), 256, 0, ST>>>(out, params, gm, S * S, slot, 0);
    count_launch(3);
  }
  jitter_apply_kernel<<<dim3(nb, N), 256, 0, ST>>>(out, params, gm, S * S, -1, 1);  // Normalize(0.5, 0.5)
  count_launch();
  return check_cuda(cudaGetLastError(), "augment_train_kernel");
}
