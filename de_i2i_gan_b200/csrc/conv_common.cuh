// Shared by the convolution kernels (conv_igemm.cu: cp.async-gather kernels; conv_tma.cu: TMA-im2col kernels).
#pragma once
#include "common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {

DEVINL void unpack8_bf16(const uint4& v, float (&f)[8]) {
  f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
  f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}

// im2col gather geometry shared by the cp.async-gather kernels (conv_igemm.cu) and the skinny kernels (conv_skinny.cu)
struct Gather {
  const __nv_bfloat16* src;
  int N, Hl, Wl, Hs, Ws, C;  // logical (Hl, Wl) nearest-mapped onto stored (Hs, Ws); C = channel pitch
  int Ho, Wo;                // row domain: m -> (img, oy, ox)
  int kh, kw, stride, pad, pad_mode;
  int mode;                  // 0: forward (iy = oy*stride + r - pad), 1: dgrad (iy = (oy + off - r) / stride)
  int off;
  int cin8;                  // 16-byte chunks per tap along K (= K-channels / 8)
  int ntaps;
  long long M;               // N*Ho*Wo
};

DEVINL bool resolve_axis(const Gather& g, int o, int r, int Ll, int Ls, int& i_out) {
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
  } else {
    int t = o + g.off - r;
    if (t < 0) return false;
    int q = t / g.stride;
    if (q * g.stride != t || q >= Ls) return false;
    i_out = q;
    return true;
  }
}

// address of the 8-channel chunk c8 of tap (r, s) for row (img, oy, ox); nullptr => zero fill
DEVINL const __nv_bfloat16* gather_ptr(const Gather& g, int img, int oy, int ox, int r, int s, int c8) {
  int iy, ix;
  if (!resolve_axis(g, oy, r, g.Hl, g.Hs, iy)) return nullptr;
  if (!resolve_axis(g, ox, s, g.Wl, g.Ws, ix)) return nullptr;
  return g.src + (((long long)img * g.Hs + iy) * g.Ws + ix) * g.C + c8 * 8;
}

struct Epilogue {
  const float* scale;
  const float* bias;
  const float* noise;
  const float* noise_weight;
  const __nv_bfloat16* residual;
  void* out;
  int act;
  int out_fp32;
  int ld;     // output pitch (elements)
  int ncols;  // real columns that carry bias (Cout)
  float* stats;  // optional [ceil(M / 128)][2][ld] per-tile (mean, M2) of the stored output (TMA kernels, bf16 output only)
  int res_fp32;  // residual is fp32 [M, ld] (FP32-storage TF32 kernel; needs out_fp32)
};


// Where output row m = (img, p, q) of the GEMM lands: out[img, p*osy + oy0, q*osx + ox0, :] of an [*, OH, OW, ld] tensor.
// linear = 1 means plain row-major rows (offset m * ld).
struct OutMap {
  int linear;
  int Ho, Wo;  // traversal domain per image
  int OH, OW, osy, osx, oy0, ox0;
};

struct TmaConvParams {
  int Ho, Wo;      // traversal (GEMM row) domain per image
  long long M;     // N * Ho * Wo
  int stride;      // traversal stride of the im2col descriptor
  int lower;       // bounding-box lower corner in H
  int lower_w;     // ... and in W (differs from `lower` for non-square filters)
  int kh, kw;      // filter size (tap t = r * kw + s)
  int flip;        // 1: TMA filter offsets are (kh-1-r, kw-1-s) - dgrad as a convolution over dY with the un-flipped weight pack
  int cin_blocks;  // input channels / 64
  // weight-matrix column block of tap (r, s): ((b_py + b_sub*r) * b_kw + (b_px + b_sub*s)) * cin_blocks + cb.
  // Plain convolutions: b_sub = 1, b_py = b_px = 0, b_kw = kw.  Strided dgrad = stride^2 stride-1 sub-convolutions over dY,
  // one per output parity (py, px), each using every stride-th tap of the full filter.
  int b_sub, b_py, b_px, b_kw;
  // npar > 1: ONE launch covers all b_sub^2 parity sub-convolutions - work tile = (m tile, n tile, parity), parity fastest so that
  // the CTAs of a wave share the dY tiles (identical for the four parities) in L2; the tile's (py, px) replace b_py / b_px and are
  // added to om.oy0 / om.ox0.  Four launches of 28 .. 64 CTAs each become one of 112 .. 148.
  int npar;
  int issue_mode;  // set by run_conv_tma from the "mma_issue" tuning knob
  int num_kb;      // taps * cin_blocks
  int m_tiles, n_tiles;
  OutMap om;
  Epilogue e;
};

struct TmaWgradParams {
  int Ho, Wo;        // output (dY) domain per image = traversal domain of the im2col descriptor over x
  long long M;       // N * Ho * Wo
  int stride, lower, lower_w, kw;  // lower: bounding-box lower corner in H, lower_w: in W
  int cin_blocks;    // Cin / 64
  int num_boxes;     // taps * cin_blocks  (64-wide K slices of the packed weight layout)
  int k_tiles;       // ceil(num_boxes / 4)
  int pair;            // 1: planned for wgrad_pair_kernel (CTA pairs, [256 co x boxes_per_tile <= 8 boxes] per pair)
  int boxes_per_tile;  // ceil(num_boxes / k_tiles) <= 4: 9 boxes run as 3 + 3 + 3, not 4 + 4 + 1 (every k tile gets the same number of CTAs)
  int co_groups;     // ceil(Cout / (128 * CO_T))
  int splits, pb_per_split, num_pb;  // 64-pixel blocks
  float* partial;    // [splits, Cout, Kf] fp32
  int cout, Kf;
};

void wgrad_tma_plan(TmaWgradParams& p);
bool wgrad_ring_applicable(const TmaWgradParams& p, int kh, int C, int W, int cout_ld);
int wgrad_ring_splits(const TmaWgradParams& p);
int run_wgrad_ring(const void* x, int N, int H, int W, int C, const void* dy, int cout_ld, TmaWgradParams& p, int kh, int upper, int upper_w,
                   cudaStream_t st);
int run_wgrad_tma(const void* x, int N, int H, int W, int C, const void* dy, int cout_ld, TmaWgradParams& p, int upper,
                  int upper_w, cudaStream_t st);

// conv_igemm.cu
int wgrad_second_pass(float* partial, int splits, const dei2i_conv_geom* g, int Cin_real, int Kf, float* dw, float beta,
                      const float* sn_w_orig, const float* sn_u, const float* sn_v, const float* sn_inv_sigma, cudaStream_t st);

// TF32 tensor-core variants of the FP32-parity path; return 1 when the geometry is not eligible
int conv_fwd_tf32(const float* x, const float* w_fwd, float* y, const dei2i_conv_geom* g, const dei2i_conv_epilogue* e, cudaStream_t st);
int conv_dgrad_tf32(const float* dy, const float* w_dgrad, float* dx, const dei2i_conv_geom* g, const float* scale, cudaStream_t st);

// conv_skinny.cu
int wgrad_narrow_splits(int Kreal, int Cout_ld, long long M);
int run_wgrad_narrow(const Gather& g, const void* dy, int Cout, int Cout_ld, int Kf, float* partial, int splits, cudaStream_t st);

bool wgrad_c8_applicable(const Gather& g, int Cout_ld);  // needs g.C, g.ntaps, g.M only
int wgrad_c8_splits(const Gather& g, int Cout_ld);
int run_wgrad_c8(const Gather& g, const void* dy, int Cout, int Cout_ld, int Kf, float* partial, cudaStream_t st);
bool conv_narrow_fwd_applicable(const Gather& g, int rows_ld);
int run_conv_narrow_fwd(const Gather& g, const Epilogue& e, const void* wmat, int rows, int Kf, cudaStream_t st);
bool conv_heads_fwd_applicable(const Gather& g, int rows_ld);
int run_conv_heads_fwd(const Gather& g, const Epilogue& e, const void* wmat, int rows, int Kf, cudaStream_t st);
bool wgrad_heads_applicable(const Gather& g, int Cout, int Cout_ld);
int wgrad_heads_splits();
int run_wgrad_heads(const Gather& g, const void* dy, int Cout, int Cout_ld, int Kf, float* partial, cudaStream_t st);
bool conv_c8_applicable(const Gather& g, int rows_ld);
int run_conv_c8(const Gather& g, const Epilogue& e, const void* wmat, int rows, int Kf, cudaStream_t st);

// returns 0 on success; the caller has checked applicability (see tma_conv_applicable)
int run_conv_tma(const void* x, int N, int H, int W, int C, const void* wmat, int rows, int K, const TmaConvParams& p,
                 int upper, int upper_w, cudaStream_t st, bool tf32 = false);

}  // namespace dei2i
