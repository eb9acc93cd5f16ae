// Kernels for the "skinny" convolutions of the DefectGAN path - the ones that never reach the tensor roofline because one
// GEMM dimension is tiny, so they are bounded by HBM (or L2) traffic, not by tcgen05 throughput:
//
//   wgrad_narrow_kernel  : dW of convolutions with Cout_ld <= 32 (generator heads 64->4, discriminator classifiers 2048->1,
//                          2048->6, reduced-width test networks).  One thread per pair of K elements, FP32 CUDA-core FMAs,
//                          the pixel loop split across blockIdx.z; fixed summation order (no atomics).
//   conv_c8_fwd_kernel   : forward / dgrad-as-convolution over an 8-channel input (SPADE mlp_shared 6->128 on the label map,
//                          stem 3->64 7x7, discriminator conv0 3->64, dgrad of the heads 4->64): warp-level mma.sync
//                          m16n8k16 BF16 with the im2col gather done straight from global memory (16 B per tap), weights
//                          resident in shared memory, epilogue staged through shared memory for full-line stores.
//   wgrad_c8_kernel      : dW of the same 8-channel-input convolutions: dY and im2col tiles staged in shared memory with
//                          cp.async, transposed into mma.sync fragments with ldmatrix.trans, FP32 partial tiles per CTA.
//
// All of them write into / read from the same buffers and packed-weight layouts as the tcgen05 kernels, so the C ABI
// (dei2i_conv2d_fwd / _dgrad / _wgrad) just dispatches by geometry.
#include <string.h>

#include "common.cuh"
#include "conv_common.cuh"

namespace dei2i {

void count_launch(int n = 1);

// ======================================================================================================================
// narrow-Cout wgrad:  partial[split, co, k] = sum_{m in split} dY[m, co] * A[m, k],  co in [co0, co0 + 8)
//   thread = two adjacent input channels (k, k+1) of one tap; all threads of a block share the pixel walk.
__global__ void __launch_bounds__(256)
wgrad_narrow_kernel(const Gather g, const __nv_bfloat16* __restrict__ dy, int Cout, int Cout_ld, int Kf,
                    float* __restrict__ partial, long long pix_per_split) {
  const int Kreal = g.ntaps * g.cin8 * 8;
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
  const int co0 = blockIdx.y * 8;
  const int split = blockIdx.z;
  if (k >= Kreal) return;
  const int tap = k / g.C;
  const int ci = k - tap * g.C;
  const int r = tap / g.kw, s = tap - r * g.kw;
  const long long m_begin = (long long)split * pix_per_split;
  const long long m_end = min(g.M, m_begin + pix_per_split);
  const int HoWo = g.Ho * g.Wo;
  int img = (int)(m_begin / HoWo);
  int rem = (int)(m_begin - (long long)img * HoWo);
  int oy = rem / g.Wo, ox = rem - oy * g.Wo;
  float a0[8], a1[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) a0[c] = a1[c] = 0.f;
  const __nv_bfloat16* dyp = dy + m_begin * Cout_ld + co0;
  for (long long m = m_begin; m < m_end; ++m, dyp += Cout_ld) {
    int iy, ix;
    const bool ok = resolve_axis(g, oy, r, g.Hl, g.Hs, iy) && resolve_axis(g, ox, s, g.Wl, g.Ws, ix);
    if (ok) {
      const uint32_t xv = __ldg(reinterpret_cast<const uint32_t*>(g.src + (((long long)img * g.Hs + iy) * g.Ws + ix) * g.C + ci));
      const float x0 = bf16lo(xv), x1 = bf16hi(xv);
      float d[8];
      unpack8_bf16(__ldg(reinterpret_cast<const uint4*>(dyp)), d);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        a0[c] = fmaf(x0, d[c], a0[c]);
        a1[c] = fmaf(x1, d[c], a1[c]);
      }
    }
    if (++ox == g.Wo) {
      ox = 0;
      if (++oy == g.Ho) { oy = 0; ++img; }
    }
  }
  for (int c = 0; c < 8 && co0 + c < Cout; ++c) {
    float* o = partial + ((long long)split * Cout + co0 + c) * Kf + k;
    *reinterpret_cast<float2*>(o) = make_float2(a0[c], a1[c]);
  }
}

// splits chosen so that the grid fills the machine; shared with the workspace query
int wgrad_narrow_splits(int Kreal, int Cout_ld, long long M) {
  const int bx = (Kreal / 2 + 255) / 256;
  const int by = Cout_ld / 8;
  long long splits = (148 * 4 + bx * by - 1) / (bx * by);
  const long long max_splits = (M + 63) / 64;  // at least 64 pixels per split
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  if (splits > 1024) splits = 1024;
  return (int)splits;
}

int run_wgrad_narrow(const Gather& g, const void* dy, int Cout, int Cout_ld, int Kf, float* partial, int splits, cudaStream_t st) {
  const int Kreal = g.ntaps * g.cin8 * 8;
  const long long pps = (g.M + splits - 1) / splits;
  dim3 grid((unsigned)((Kreal / 2 + 255) / 256), (unsigned)(Cout_ld / 8), (unsigned)splits);
  // columns [Kreal, Kf) of the partial buffer are never read by wgrad_reduce_kernel (it walks real taps / channels only)
  wgrad_narrow_kernel<<<grid, 256, 0, st>>>(g, (const __nv_bfloat16*)dy, Cout, Cout_ld, Kf, partial, pps);
  count_launch();
  return check_cuda(cudaGetLastError(), "wgrad_narrow_kernel launch");
}

}  // namespace dei2i
