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
void trace_kernel(const char* fmt, ...);
extern int g_tune_conv_path;  // 2: CUDA-core heads wgrad kernel instead of the mma.sync one (A/B testing)

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
  trace_kernel("wgrad_narrow");
  return check_cuda(cudaGetLastError(), "wgrad_narrow_kernel launch");
}


// ======================================================================================================================
// 8-channel-input convolution (forward, or dgrad expressed through Gather mode 1):  D[m, n] = sum_k A[m, k] * W[n, k]
//   k = tap * 8 + ci.  One warp = 16 output pixels x all (<= NT * 8) output channels; K step = 2 taps (mma m16n8k16).
//   A fragments are loaded straight from global memory: a (pixel, tap) is 16 contiguous bytes, the 4 lanes of a fragment
//   group read 4 bytes each, the 8 groups read 8 neighbouring pixels -> 128-byte segments; tap re-use is served by L1.
DEVINL void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// like resolve_axis (conv_common.cuh) with the nearest-resize index in 32-bit arithmetic
DEVINL bool c8_axis(const Gather& g, int o, int r, int Ll, int Ls, int& i_out) {
  if (g.mode == 0) {
    int i = o * g.stride + r - g.pad;
    if (g.pad_mode == DEI2I_PAD_REFLECT) {
      i = i < 0 ? -i : i;
      i = i >= Ll ? 2 * Ll - 2 - i : i;
    } else if (i < 0 || i >= Ll) {
      return false;
    }
    if (Ls != Ll) i = Ls == 1 ? 0 : (int)(((unsigned)i * (unsigned)Ls) / (unsigned)Ll);
    i_out = i;
    return true;
  }
  return resolve_axis(g, o, r, Ll, Ls, i_out);
}

struct C8Params {
  Gather g;  // g.C == 8
  Epilogue e;
  const __nv_bfloat16* w;  // [rows, Kf] packed weights, k = tap * 8 + ci
  int rows;                // real rows of w (output channels)
  int Kf;                  // row pitch of w
  int ksteps;              // ceil(ntaps / 2)
  int m_tiles;             // ceil(M / 128)
  // affine tap walk (FAST kernels): input row of tap (r, s) for output (oy, ox) is y0 + dir * r with
  // y0 = oy * ystep + yoff, and the element offset of the tap relative to (y0, x0) is dir * (r * rowp + s * colp)
  int dir, ystep, yoff, rowp, colp;
};

// When the gather is affine in the tap index (no reflect padding, no general nearest resize, dgrad only at stride 1) the
// per-tap address is base(pixel) + table[tap] and the bounds test two unsigned compares; returns false otherwise.
static bool c8_affine(const Gather& g, int& dir, int& ystep, int& yoff, int& rowp, int& colp) {
  if (g.pad_mode == DEI2I_PAD_REFLECT && g.pad > 0) return false;
  if (g.mode == 0) {
    if (!((g.Hs == g.Hl || g.Hs == 1) && (g.Ws == g.Wl || g.Ws == 1))) return false;
    dir = 1; ystep = g.stride; yoff = -g.pad;
  } else {
    if (g.stride != 1 || g.Hs != g.Hl || g.Ws != g.Wl) return false;
    dir = -1; ystep = 1; yoff = g.off;
  }
  rowp = g.Hs == g.Hl ? g.Ws * 8 : 0;  // a 1 x 1 stored map broadcasts: every in-bounds tap reads the same pixel
  colp = g.Ws == g.Wl ? 8 : 0;
  return true;
}

constexpr int kC8Threads = 256;

DEVINL void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

template <int NT, bool FAST>
__global__ void __launch_bounds__(kC8Threads, 2)
conv_c8_fwd_kernel(const C8Params p) {
  extern __shared__ __align__(16) uint8_t c8_smem[];
  __shared__ int4 tap_tab[64];  // (dir * r, dir * s, element offset, 0); taps >= ntaps fail every bounds test
  __shared__ __align__(8) float bias_s[NT * 8];
  const Gather& g = p.g;
  const Epilogue& e = p.e;
  const int KP = p.ksteps * 16 + 8;  // weight row pitch in smem (elements): an odd number of 16-byte chunks -> conflict-free ldmatrix
  __nv_bfloat16* w_s = reinterpret_cast<__nv_bfloat16*>(c8_smem);
  constexpr int kStagePitch = NT * 16 + 16;  // bytes per staged output row (bf16), +16: conflict-free fragment writes
  uint8_t* stage_base = c8_smem + ((NT * 8 * KP * 2 + 15) & ~15);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tq = lane & 3;

  // weights -> smem (zero rows / columns beyond the real ones), tap table, bias
  {
    const int kreal = g.ntaps * 8;
    const int kw8 = KP / 8;
    for (int i = threadIdx.x; i < NT * 8 * kw8; i += kC8Threads) {
      const int n = i / kw8, k8 = i - n * kw8;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (n < p.rows && k8 * 8 < kreal) v = __ldg(reinterpret_cast<const uint4*>(p.w + (long long)n * p.Kf + k8 * 8));
      *reinterpret_cast<uint4*>(w_s + n * KP + k8 * 8) = v;
    }
    if (threadIdx.x < 64) {
      const int t = threadIdx.x;
      const int r = t / g.kw, s = t - r * g.kw;
      tap_tab[t] = t < g.ntaps ? make_int4(p.dir * r, p.dir * s, p.dir * (r * p.rowp + s * p.colp), 0) : make_int4(-(1 << 28), 0, 0, 0);
    }
    if (threadIdx.x < NT * 8) bias_s[threadIdx.x] = (e.bias != nullptr && (int)threadIdx.x < e.ncols) ? __ldg(e.bias + threadIdx.x) : 0.f;
  }
  __syncthreads();

  const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
  const float nzw = e.noise != nullptr ? __ldg(e.noise_weight) : 0.f;
  // none / leaky-relu / relu are max(v, slope * v); tanh and sigmoid take the (warp-uniform) slow branch
  const float slope = e.act == ACT_LRELU ? 0.2f : (e.act == ACT_RELU ? 0.f : 1.f);
  const bool slow_act = e.act == ACT_TANH || e.act == ACT_SIGMOID;
  const unsigned M = (unsigned)g.M, HoWo = (unsigned)(g.Ho * g.Wo), Wo = (unsigned)g.Wo;
  uint8_t* stage = stage_base + warp * (16 * kStagePitch);
  const uint32_t* w_words = reinterpret_cast<const uint32_t*>(w_s);
  const int KPW = KP / 2;
  // ldmatrix.x4 of B: matrices (n-tile j, k 0-7) (j, k 8-15) (j+1, k 0-7) (j+1, k 8-15); this lane supplies row (lane % 8) of matrix (lane / 8)
  const uint32_t b_lane = smem_u32(w_s) + ((((lane >> 4) * 8 + (lane & 7)) * KP + ((lane >> 3) & 1) * 8) * 2);
  // role in the coalesced copy of the staged tile: 16-byte chunk cp_ch of rows cp_row0, cp_row0 + cp_rows, ...
  const int chunks = e.ld / 8;
  const int cp_rows = 32 / chunks;
  const int cp_row0 = lane / chunks, cp_ch = lane - cp_row0 * chunks;

  for (int tile = blockIdx.x; tile < p.m_tiles; tile += gridDim.x) {
    const unsigned m_w = (unsigned)tile * 128u + (unsigned)warp * 16u;
    if (m_w >= M) continue;
    int r_img[2], r_oy[2], r_ox[2];
    bool r_ok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const unsigned m = m_w + gid + 8 * h;
      r_ok[h] = m < M;
      const unsigned mm = r_ok[h] ? m : 0u;
      const unsigned img = mm / HoWo, rem = mm - img * HoWo;
      const unsigned oy = rem / Wo;
      r_img[h] = (int)img; r_oy[h] = (int)oy; r_ox[h] = (int)(rem - oy * Wo);
    }
    float acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;

    // FAST: per-row origin (y0, x0) and base pointer; rows beyond M fail every bounds test
    int y0[2], x0[2];
    const __nv_bfloat16* base[2];
    if (FAST) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        y0[h] = r_ok[h] ? r_oy[h] * p.ystep + p.yoff : -(1 << 28);
        x0[h] = r_ox[h] * p.ystep + p.yoff;
        base[h] = g.src + (long long)r_img[h] * g.Hs * g.Ws * 8 + (long long)y0[h] * p.rowp + (long long)x0[h] * p.colp + 2 * tq;
      }
    }
    auto load_a = [&](int ks, uint32_t (&a)[4]) {
      if (FAST) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int4 t = tap_tab[2 * ks + j];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t v = 0;
            if ((unsigned)(y0[h] + t.x) < (unsigned)g.Hl && (unsigned)(x0[h] + t.y) < (unsigned)g.Wl)
              v = __ldg(reinterpret_cast<const uint32_t*>(base[h] + t.z));
            a[j * 2 + h] = v;
          }
        }
        return;
      }
      int tap = 2 * ks;
      int r = tap / g.kw, s = tap - r * g.kw;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t v = 0;
          int iy, ix;
          if (tap < g.ntaps && r_ok[h] && c8_axis(g, r_oy[h], r, g.Hl, g.Hs, iy) && c8_axis(g, r_ox[h], s, g.Wl, g.Ws, ix))
            v = __ldg(reinterpret_cast<const uint32_t*>(g.src + (((long long)r_img[h] * g.Hs + iy) * g.Ws + ix) * 8 + 2 * tq));
          a[j * 2 + h] = v;
        }
        ++tap;
        if (++s == g.kw) { s = 0; ++r; }
      }
    };

    uint32_t a_cur[4], a_nxt[4];
    load_a(0, a_cur);
    for (int ks = 0; ks < p.ksteps; ++ks) {
      if (ks + 1 < p.ksteps) load_a(ks + 1, a_nxt);
      const uint32_t baddr = b_lane + ks * 32;
#pragma unroll
      for (int j = 0; j < NT; j += 4) {
        uint32_t b0[4], b1[4];
        ldsm_x4(b0, baddr + j * 8 * KP * 2);
        ldsm_x4(b1, baddr + (j + 2) * 8 * KP * 2);
        mma_bf16_16816(acc[j], a_cur, b0[0], b0[1]);
        mma_bf16_16816(acc[j + 1], a_cur, b0[2], b0[3]);
        mma_bf16_16816(acc[j + 2], a_cur, b1[0], b1[1]);
        mma_bf16_16816(acc[j + 3], a_cur, b1[2], b1[3]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) a_cur[i] = a_nxt[i];
    }
    (void)w_words; (void)KPW;

    // ---- epilogue: rows gid / gid + 8, columns j * 8 + 2 * tq (+1)
    float nz[2] = {0.f, 0.f};
    if (e.noise != nullptr) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (r_ok[h]) nz[h] = __ldg(e.noise + m_w + gid + 8 * h) * nzw;
    }
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int col = j * 8 + 2 * tq;
      const float2 bb = *reinterpret_cast<const float2*>(bias_s + col);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float v0 = fmaf(acc[j][2 * h], osc, bb.x + nz[h]);
        float v1 = fmaf(acc[j][2 * h + 1], osc, bb.y + nz[h]);
        if (slow_act) {
          v0 = apply_act(v0, e.act);
          v1 = apply_act(v1, e.act);
        } else {
          v0 = fmaxf(v0, v0 * slope);
          v1 = fmaxf(v1, v1 * slope);
        }
        const long long m = (long long)m_w + gid + 8 * h;
        if (e.residual != nullptr && r_ok[h] && col < e.ld) {
          const uint32_t rv = __ldg(reinterpret_cast<const uint32_t*>(e.residual + m * e.ld + col));
          v0 += bf16lo(rv);
          v1 += bf16hi(rv);
        }
        if (e.out_fp32) {
          if (r_ok[h] && col < e.ld) *reinterpret_cast<float2*>(reinterpret_cast<float*>(e.out) + m * e.ld + col) = make_float2(v0, v1);
        } else {
          *reinterpret_cast<uint32_t*>(stage + (gid + 8 * h) * kStagePitch + col * 2) = pack_bf16(v0, v1);
        }
      }
    }
    if (!e.out_fp32) {
      __syncwarp();
      if (cp_row0 < cp_rows) {
        for (int row = cp_row0; row < 16; row += cp_rows) {
          const unsigned m = m_w + row;
          if (m < M)
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(e.out) + (long long)m * e.ld + cp_ch * 8) =
                *reinterpret_cast<const uint4*>(stage + row * kStagePitch + cp_ch * 16);
        }
      }
      __syncwarp();
    }
  }
}

template <int NT, bool FAST>
static int launch_c8(const C8Params& p, cudaStream_t st) {
  const int KP = p.ksteps * 16 + 8;
  const size_t smem = ((size_t)(NT * 8 * KP * 2 + 15) & ~(size_t)15) + (size_t)(kC8Threads / 32) * 16 * (NT * 16 + 16);
  if (smem > 200 * 1024) {
    set_error("conv_c8: filter too large for the resident-weight kernel (%zu bytes of shared memory)", smem);
    return DEI2I_ERR_INVALID;
  }
  static size_t attr = 0;
  if (smem > attr) {
    DEI2I_CUDA(cudaFuncSetAttribute(conv_c8_fwd_kernel<NT, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = smem;
  }
  const int grid = p.m_tiles < 148 * 2 ? p.m_tiles : 148 * 2;
  conv_c8_fwd_kernel<NT, FAST><<<grid, kC8Threads, smem, st>>>(p);
  count_launch();
  trace_kernel("conv_c8_fwd<%d,%d>", NT, (int)FAST);
  return check_cuda(cudaGetLastError(), "conv_c8_fwd_kernel launch");
}

bool conv_c8_applicable(const Gather& g, int rows_ld) {
  if (g.C != 8 || rows_ld > 128 || g.M >= (1ll << 31) - 256) return false;
  const int ksteps = (g.ntaps + 1) / 2;
  const int nt = rows_ld <= 64 ? 8 : 16;
  return (size_t)nt * 8 * (ksteps * 16 + 8) * 2 <= 150 * 1024;
}

int run_conv_c8(const Gather& g, const Epilogue& e, const void* wmat, int rows, int Kf, cudaStream_t st) {
  C8Params p;
  memset(&p, 0, sizeof(p));
  p.g = g; p.e = e; p.w = (const __nv_bfloat16*)wmat; p.rows = rows; p.Kf = Kf;
  p.ksteps = (g.ntaps + 1) / 2;
  p.m_tiles = (int)((g.M + 127) / 128);
  if (g.ntaps <= 62 && c8_affine(g, p.dir, p.ystep, p.yoff, p.rowp, p.colp))
    return e.ld <= 64 ? launch_c8<8, true>(p, st) : launch_c8<16, true>(p, st);
  return e.ld <= 64 ? launch_c8<8, false>(p, st) : launch_c8<16, false>(p, st);
}


// ======================================================================================================================
// wgrad of an 8-channel-input convolution:  partial[cta, co, tap * 8 + ci] = sum_{m in cta range} dY[m, co] * A[m, tap * 8 + ci]
//   mma.sync m16n8k16 with M = 16 output channels, N = the 8 channels of one tap, K = 16 pixels.  Both operands have the
//   reduction index (pixel) as the slow one in memory, so the tiles ([P pixels][Cout] of dY, [P pixels][taps * 8] of the
//   im2col matrix, 16-byte cp.async with zero fill) are transposed on the way to the registers by ldmatrix.trans.
//   Warp w owns m-tile (w % warps_m) and taps [ (w / warps_m) * tpw, ... + tpw ).
DEVINL void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
// 16-byte cp.async through L1 (.ca): neighbouring taps / pixels re-read the same lines, and a 1 x 1 label map is ONE line that
// every CTA reads for every tap - with .cg all of those requests would queue on a single L2 slice
DEVINL void cp_async_16_ca(uint32_t dst_smem, const void* src, bool valid) {
  uint32_t sz = valid ? 16u : 0u;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(sz) : "memory");
}
DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
DEVINL void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct WgC8Params {
  Gather g;  // forward gather geometry of x, g.C == 8
  const __nv_bfloat16* dy;
  float* partial;  // [grid, Cout, Kf]
  int Cout, Cout_ld, Kf;
  int warps_m;     // power of two >= ceil(Cout_ld / 16)
  int tpw;         // taps per warp
  int chunks_per_cta, num_chunks;  // chunks of P pixels
};

constexpr int kWgC8Stages = 3;

template <int NT, int P>
__global__ void __launch_bounds__(256, NT <= 9 ? 2 : 1)
wgrad_c8_kernel(const WgC8Params p) {
  extern __shared__ __align__(16) uint8_t wg_smem[];
  __shared__ int8_t tap_r[64], tap_s[64];
  const Gather& g = p.g;
  const int pitch_dy = ((p.Cout_ld / 8) | 1) * 16;
  const int pitch_a = (g.ntaps | 1) * 16;
  const int stage_bytes = P * (pitch_dy + pitch_a);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tq = lane & 3;
  if (threadIdx.x < g.ntaps) {
    tap_r[threadIdx.x] = (int8_t)(threadIdx.x / g.kw);
    tap_s[threadIdx.x] = (int8_t)(threadIdx.x % g.kw);
  }
  __syncthreads();

  const int c_begin = blockIdx.x * p.chunks_per_cta;
  const int c_end = min(p.num_chunks, c_begin + p.chunks_per_cta);
  const int HoWo = g.Ho * g.Wo;
  const uint32_t smem0 = smem_u32(wg_smem);

  // producer role of every thread: pixel slot (tid % P), taps / dY chunks strided by 256 / P
  constexpr int TPP = 256 / P;  // threads per pixel
  const int my_pix = threadIdx.x % P, my_sub = threadIdx.x / P;
  const int dy_chunks = p.Cout_ld / 8;
  auto issue = [&](int chunk, int stage) {
    const uint32_t s_dy = smem0 + stage * stage_bytes;
    const uint32_t s_a = s_dy + P * pitch_dy;
    const long long m = (long long)chunk * P + my_pix;
    const bool ok = m < g.M;
    const long long mm = ok ? m : 0;
    const int img = (int)(mm / HoWo);
    const int rem = (int)(mm - (long long)img * HoWo);
    const int oy = rem / g.Wo, ox = rem - oy * g.Wo;
    const __nv_bfloat16* dsrc = p.dy + mm * p.Cout_ld;
    for (int c = my_sub; c < dy_chunks; c += TPP) cp_async_16(s_dy + my_pix * pitch_dy + c * 16, dsrc + c * 8, ok);
    for (int t = my_sub; t < g.ntaps; t += TPP) {
      int iy, ix;
      const bool v = ok && c8_axis(g, oy, tap_r[t], g.Hl, g.Hs, iy) && c8_axis(g, ox, tap_s[t], g.Wl, g.Ws, ix);
      const __nv_bfloat16* src = v ? g.src + (((long long)img * g.Hs + iy) * g.Ws + ix) * 8 : g.src;
      cp_async_16_ca(s_a + my_pix * pitch_a + t * 16, src, v);
    }
  };

  const int m_tile = warp & (p.warps_m - 1);
  const int tap0 = (warp / p.warps_m) * p.tpw;
  const int ntw = max(0, min(p.tpw, g.ntaps - tap0));  // taps of this warp
  const bool m_active = m_tile * 16 < p.Cout_ld;
  float acc[NT][4];
#pragma unroll
  for (int j = 0; j < NT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;

  // prologue
#pragma unroll
  for (int s = 0; s < kWgC8Stages - 1; ++s) {
    if (c_begin + s < c_end) issue(c_begin + s, s);
    cp_async_commit();
  }
  for (int c = c_begin; c < c_end; ++c) {
    const int it = c - c_begin;
    cp_async_wait_group<kWgC8Stages - 2>();
    __syncthreads();  // chunk c has landed for everyone; everyone is done with the stage about to be refilled
    if (c + kWgC8Stages - 1 < c_end) issue(c + kWgC8Stages - 1, (it + kWgC8Stages - 1) % kWgC8Stages);
    cp_async_commit();
    if (m_active && ntw > 0) {
      const uint32_t s_dy = smem0 + (it % kWgC8Stages) * stage_bytes;
      const uint32_t s_a = s_dy + P * pitch_dy;
#pragma unroll
      for (int pp = 0; pp < P; pp += 32) {
        uint32_t a0[4], a1[4];
        // matrices: (pix lo, co lo) (pix lo, co hi) (pix hi, co lo) (pix hi, co hi); lane supplies row (lane % 8) of matrix (lane / 8)
        const int mi = lane >> 3, mr = lane & 7;
        const uint32_t a_addr = s_dy + (pp + (mi >> 1) * 8 + mr) * pitch_dy + (m_tile * 16 + (mi & 1) * 8) * 2;
        ldsm_x4_trans(a0, a_addr);
        ldsm_x4_trans(a1, a_addr + 16 * pitch_dy);
        // B: four pixel octets of one tap -> (b0, b1) of k-step 0 and of k-step 1
        const uint32_t b_addr = s_a + (pp + mi * 8 + mr) * pitch_a + tap0 * 16;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
          if (j < ntw) {
            uint32_t b[4];
            ldsm_x4_trans(b, b_addr + j * 16);
            mma_bf16_16816(acc[j], a0, b[0], b[1]);
            mma_bf16_16816(acc[j], a1, b[2], b[3]);
          }
        }
      }
    }
  }
  cp_async_wait_group<0>();
  if (m_active) {
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      if (j < ntw) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int co = m_tile * 16 + gid + 8 * h;
          if (co < p.Cout)
            *reinterpret_cast<float2*>(p.partial + ((long long)blockIdx.x * p.Cout + co) * p.Kf + (tap0 + j) * 8 + 2 * tq) =
                make_float2(acc[j][2 * h], acc[j][2 * h + 1]);
        }
      }
    }
  }
}

static void wgrad_c8_plan(const Gather& g, int Cout_ld, int& warps_m, int& tpw, int& P, int& grid, int& cpc, int& nchunks) {
  const int m_tiles = (Cout_ld + 15) / 16;
  warps_m = 1;
  while (warps_m < m_tiles) warps_m *= 2;
  const int warps_n = 8 / warps_m;
  tpw = (g.ntaps + warps_n - 1) / warps_n;
  P = tpw <= 9 ? 64 : 32;
  nchunks = (int)((g.M + P - 1) / P);
  grid = nchunks < 148 * 2 ? nchunks : 148 * 2;
  if (tpw > 9 && grid > 148) grid = 148;
  cpc = (nchunks + grid - 1) / grid;
  grid = (nchunks + cpc - 1) / cpc;
}

bool wgrad_c8_applicable(const Gather& g, int Cout_ld) {
  if (g.C != 8 || Cout_ld < 16 || Cout_ld > 128 || g.ntaps > 64) return false;
  int warps_m, tpw, P, grid, cpc, nchunks;
  wgrad_c8_plan(g, Cout_ld, warps_m, tpw, P, grid, cpc, nchunks);
  return tpw <= 25;
}

int wgrad_c8_splits(const Gather& g, int Cout_ld) {
  int warps_m, tpw, P, grid, cpc, nchunks;
  wgrad_c8_plan(g, Cout_ld, warps_m, tpw, P, grid, cpc, nchunks);
  return grid;
}

template <int NT, int P>
static int launch_wgrad_c8(const WgC8Params& p, int grid, cudaStream_t st) {
  const int pitch_dy = ((p.Cout_ld / 8) | 1) * 16, pitch_a = (p.g.ntaps | 1) * 16;
  const size_t smem = (size_t)kWgC8Stages * P * (pitch_dy + pitch_a);
  static size_t attr = 0;
  if (smem > attr) {
    DEI2I_CUDA(cudaFuncSetAttribute(wgrad_c8_kernel<NT, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = smem;
  }
  wgrad_c8_kernel<NT, P><<<grid, 256, smem, st>>>(p);
  count_launch();
  trace_kernel("wgrad_c8<%d,%d>", NT, P);
  return check_cuda(cudaGetLastError(), "wgrad_c8_kernel launch");
}

int run_wgrad_c8(const Gather& g, const void* dy, int Cout, int Cout_ld, int Kf, float* partial, cudaStream_t st) {
  WgC8Params p;
  memset(&p, 0, sizeof(p));
  int P, grid;
  wgrad_c8_plan(g, Cout_ld, p.warps_m, p.tpw, P, grid, p.chunks_per_cta, p.num_chunks);
  p.g = g; p.dy = (const __nv_bfloat16*)dy; p.partial = partial; p.Cout = Cout; p.Cout_ld = Cout_ld; p.Kf = Kf;
  if (p.tpw <= 9) return launch_wgrad_c8<9, 64>(p, grid, st);
  return launch_wgrad_c8<25, 32>(p, grid, st);
}


// ======================================================================================================================
// wgrad of the generator heads (3x3, stride 1, Cin % 64 == 0, <= 4 output channels; generator.py:224-241):
//   dW[co, (r, s), ci] = sum_pixels dY[pix, co] * x[pix + (r, s), ci]
//   A warp owns 64 input channels (lane = channel pair) and walks output rows with a 3 x 3 register window of x: per pixel
//   3 coalesced 128-byte loads + one broadcast 8-byte dY load feed 72 FMAs per lane (9 taps x 2 channels x 4 outputs).
//   Warps of a CTA are summed through shared memory in a fixed order; one FP32 partial tile per CTA.
constexpr int kHeadsWarps = 16;

__global__ void __launch_bounds__(kHeadsWarps * 32, 1)
wgrad_heads_kernel(const Gather g, const __nv_bfloat16* __restrict__ dy, int Cout, int Cout_ld, int Kf, float* __restrict__ partial) {
  __shared__ float red[72 * 32 * 2];  // two 64-channel slabs at most share a CTA
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slabs = g.C / 64;
  const int total_warps = gridDim.x * kHeadsWarps;
  const int gw = blockIdx.x * kHeadsWarps + warp;
  const int slab = gw % slabs;  // kHeadsWarps % slabs == 0: the warps of one slab inside a CTA are warp % slabs == slab
  const int ci = slab * 64 + 2 * lane;
  const int rows = g.N * g.Ho;
  float acc[3][3][4][2];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[r][s][c][0] = acc[r][s][c][1] = 0.f;

  for (int row = gw / slabs; row < rows; row += total_warps / slabs) {
    const int img = row / g.Ho, oy = row - img * g.Ho;
    const __nv_bfloat16* xr[3];
    bool yok[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int iy = oy - g.pad + r;
      yok[r] = iy >= 0 && iy < g.Hl;
      xr[r] = g.src + ((long long)img * g.Hs + (yok[r] ? iy : 0)) * g.Ws * g.C + ci;
    }
    auto ldx = [&](int r, int ix) -> uint32_t {
      return (yok[r] && ix >= 0 && ix < g.Wl) ? __ldg(reinterpret_cast<const uint32_t*>(xr[r] + (long long)ix * g.C)) : 0u;
    };
    uint32_t w[3][3], nxt[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      w[r][0] = 0;
      w[r][1] = ldx(r, -g.pad);
      w[r][2] = ldx(r, -g.pad + 1);
      nxt[r] = ldx(r, -g.pad + 2);
    }
    const __nv_bfloat16* dyp = dy + (long long)row * g.Wo * Cout_ld;
    uint2 dcur = __ldg(reinterpret_cast<const uint2*>(dyp));
    for (int ox = 0; ox < g.Wo; ++ox) {
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        w[r][0] = w[r][1];
        w[r][1] = w[r][2];
        w[r][2] = nxt[r];
      }
      const uint2 dv = dcur;
      if (ox + 1 < g.Wo) {
#pragma unroll
        for (int r = 0; r < 3; ++r) nxt[r] = ldx(r, ox + 1 - g.pad + 2);
        dcur = __ldg(reinterpret_cast<const uint2*>(dyp + (long long)(ox + 1) * Cout_ld));
      }
      const float d[4] = {bf16lo(dv.x), bf16hi(dv.x), bf16lo(dv.y), bf16hi(dv.y)};
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const float x0 = bf16lo(w[r][s]), x1 = bf16hi(w[r][s]);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            acc[r][s][c][0] = fmaf(x0, d[c], acc[r][s][c][0]);
            acc[r][s][c][1] = fmaf(x1, d[c], acc[r][s][c][1]);
          }
        }
    }
  }
  // CTA reduction in warp order (deterministic), one slab at a time
  for (int sl = 0; sl < slabs; ++sl) {
    float* buf = red + (sl & 1) * (72 * 32);
    for (int wv = 0; wv < kHeadsWarps; ++wv) {
      if (warp == wv && slab == sl) {
        const bool first = wv < slabs;  // first warp of this slab in the CTA
        int i = 0;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s = 0; s < 3; ++s)
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
              for (int e2 = 0; e2 < 2; ++e2, ++i) {
                float* q = buf + i * 32 + lane;
                *q = first ? acc[r][s][c][e2] : *q + acc[r][s][c][e2];
              }
      }
      __syncthreads();
    }
    // buf[((r*3+s)*4 + c)*2 + e][lane] -> partial[cta, c, (r*3+s)*C + sl*64 + 2*lane + e]
    for (int i = threadIdx.x; i < 72 * 32; i += blockDim.x) {
      const int ln = i & 31, idx = i >> 5;
      const int e2 = idx & 1, c = (idx >> 1) & 3, tap = idx >> 3;
      if (c < Cout) partial[((long long)blockIdx.x * Cout + c) * Kf + tap * g.C + sl * 64 + 2 * ln + e2] = buf[i];
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------------------------------------
// The same weight gradient on the tensor pipe (mma.sync m16n8k16; M = 16 input channels, N = the 8 stored output channels,
// K = 16 pixels), for Cin = 64:  a CTA walks a strip of output rows with a ring of four x rows ((Wo32 + 2) pixels x 128 B,
// 16-byte chunks XOR-swizzled by pixel so that ldmatrix.trans is conflict-free) and a double-buffered dY row in shared
// memory (cp.async, zero fill for the zero frame and for columns past Wo); every x row is read from HBM once per strip.
// Warp w: channel group w % 4, 32-pixel chunks (w / 4) mod 4; 9 taps x 4 FP32 accumulators per lane; the four pixel quarters
// are summed through shared memory in a fixed order.  One FP32 partial tile per CTA (same layout as above).
constexpr int kHmWarps = 16;
constexpr int kHmRing = 4;

__global__ void __launch_bounds__(kHmWarps * 32, 1)
wgrad_heads_mma_kernel(const Gather g, const __nv_bfloat16* __restrict__ dy, int Cout, int Kf, float* __restrict__ partial,
                       int rows_per_cta) {
  extern __shared__ __align__(128) uint8_t hm_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tq = lane & 3;
  const int cig = warp & 3, quarter = warp >> 2;
  const int Wo32 = (g.Wo + 31) & ~31;
  const int XW = Wo32 + 2;                 // pixels per staged x row
  const int x_row_bytes = XW * 128;
  const uint32_t s_x = smem_u32(hm_smem);
  const uint32_t s_dy = s_x + kHmRing * x_row_bytes;  // 2 x Wo32 x 16 B
  const int rows = g.N * g.Ho;
  const int rb = blockIdx.x * rows_per_cta;
  const int re = min(rows, rb + rows_per_cta);

  auto issue_x = [&](int img, int iy) {  // stored row iy of image img (zero row when outside the frame) -> slot (iy + 4) & 3
    const bool rok = iy >= 0 && iy < g.Hl;
    const __nv_bfloat16* src = g.src + ((long long)img * g.Hs + (rok ? iy : 0)) * g.Ws * 64;
    const uint32_t dst = s_x + ((iy + 4) & 3) * x_row_bytes;
    for (int i = threadIdx.x; i < XW * 8; i += blockDim.x) {
      const int j = i >> 3, c = i & 7;
      const int ix = j - g.pad;
      const bool ok = rok && ix >= 0 && ix < g.Wl;
      cp_async_16(dst + j * 128 + ((c ^ (j & 7)) << 4), src + (ok ? ix * 64 + c * 8 : 0), ok);
    }
  };
  auto issue_dy = [&](int row) {
    const __nv_bfloat16* src = dy + (long long)row * g.Wo * 8;
    const uint32_t dst = s_dy + (row & 1) * Wo32 * 16;
    for (int i = threadIdx.x; i < Wo32; i += blockDim.x) cp_async_16(dst + i * 16, src + (i < g.Wo ? i * 8 : 0), i < g.Wo);
  };

  float acc[9][4];
#pragma unroll
  for (int t = 0; t < 9; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
  const int mi = lane >> 3, mr = lane & 7;

  for (int row = rb; row < re; ++row) {
    const int img = row / g.Ho, oy = row - img * g.Ho;
    const int iy0 = oy - g.pad;
    if (row == rb || oy == 0) {  // (re)start of an image: nothing of it is staged yet
      __syncthreads();
      issue_x(img, iy0);
      issue_x(img, iy0 + 1);
      issue_x(img, iy0 + 2);
      issue_dy(row);
    }
    cp_async_commit();
    if (row + 1 < re && oy + 1 < g.Ho) {  // next output row of the same image: one new x row, the next dY row
      issue_x(img, iy0 + 3);
      issue_dy(row + 1);
    }
    cp_async_commit();
    cp_async_wait_group<1>();
    __syncthreads();
    const uint32_t dyr = s_dy + (row & 1) * Wo32 * 16;
    for (int q = quarter; q * 32 < g.Wo; q += 4) {
      const int ox0 = q * 32;
      uint32_t b[4];
      ldsm_x4_trans(b, dyr + (ox0 + mi * 8 + mr) * 16);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const uint32_t xr = s_x + ((iy0 + r + 4) & 3) * x_row_bytes;
#pragma unroll
        for (int sx = 0; sx < 3; ++sx) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int pix = ox0 + h * 16 + (mi >> 1) * 8 + mr + sx;
            const int chunk = cig * 2 + (mi & 1);
            uint32_t a[4];
            ldsm_x4_trans(a, xr + pix * 128 + ((chunk ^ (pix & 7)) << 4));
            mma_bf16_16816(acc[r * 3 + sx], a, b[2 * h], b[2 * h + 1]);
          }
        }
      }
    }
    __syncthreads();  // everyone is done with the slots the next iteration refills
  }
  cp_async_wait_group<0>();
  __syncthreads();
  // quarters 1..3 park their accumulators in shared memory, quarter 0 adds them in order and stores
  float* red = reinterpret_cast<float*>(hm_smem);
  if (quarter > 0) {
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
      for (int e = 0; e < 4; ++e) red[(((quarter - 1) * 4 + cig) * 36 + t * 4 + e) * 32 + lane] = acc[t][e];
  }
  __syncthreads();
  if (quarter == 0) {
#pragma unroll
    for (int t = 0; t < 9; ++t) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = acc[t][e];
        for (int qq = 0; qq < 3; ++qq) v += red[((qq * 4 + cig) * 36 + t * 4 + e) * 32 + lane];
        const int ci = cig * 16 + gid + 8 * (e >> 1);
        const int co = 2 * tq + (e & 1);
        if (co < Cout) partial[((long long)blockIdx.x * Cout + co) * Kf + t * 64 + ci] = v;
      }
    }
  }
}

static size_t wgrad_heads_mma_smem(int Wo) {
  const int Wo32 = (Wo + 31) & ~31;
  return (size_t)kHmRing * (Wo32 + 2) * 128 + 2 * (size_t)Wo32 * 16;
}
static bool wgrad_heads_mma_applicable(const Gather& g) {
  const size_t sm = wgrad_heads_mma_smem(g.Wo);
  return g.C == 64 && g.pad <= 1 && sm <= 200 * 1024 && sm >= (size_t)12 * 36 * 32 * 4 && g.Ho >= 1;
}

// ----------------------------------------------------------------------------------------------------------------------
// Forward of the heads (64 -> <= 8 stored output channels, 3x3, stride 1; generator.py:224-241) on the same row ring:
// mma.sync m16n8k16 with M = 16 pixels, N = 8 output channels, K = 16 input channels; the 9 x 4 B fragments (all weights)
// stay in registers, A fragments come from the swizzled x rows by ldmatrix.  A warp owns 16-pixel chunks of the row.
// HBM traffic = x once + y once (the tcgen05 path would pad N to 64 and run 16x the FLOPs).
__global__ void __launch_bounds__(kHmWarps * 32, 1)
conv_heads_fwd_kernel(const Gather g, const Epilogue e, const __nv_bfloat16* __restrict__ w, int rows_w, int Kf, int rows_per_cta) {
  extern __shared__ __align__(128) uint8_t hm_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tq = lane & 3;
  const int Wo16 = (g.Wo + 15) & ~15;
  const int XW = Wo16 + 2;
  const int x_row_bytes = XW * 128;
  const uint32_t s_x = smem_u32(hm_smem);
  const int rows = g.N * g.Ho;
  const int rb = blockIdx.x * rows_per_cta;
  const int re = min(rows, rb + rows_per_cta);

  auto issue_x = [&](int img, int iy) {
    const bool rok = iy >= 0 && iy < g.Hl;
    const __nv_bfloat16* src = g.src + ((long long)img * g.Hs + (rok ? iy : 0)) * g.Ws * 64;
    const uint32_t dst = s_x + ((iy + 4) & 3) * x_row_bytes;
    for (int i = threadIdx.x; i < XW * 8; i += blockDim.x) {
      const int j = i >> 3, c = i & 7;
      const int ix = j - g.pad;
      const bool ok = rok && ix >= 0 && ix < g.Wl;
      cp_async_16(dst + j * 128 + ((c ^ (j & 7)) << 4), src + (ok ? ix * 64 + c * 8 : 0), ok);
    }
  };

  // B fragments: b0 = W[co = gid][tap*64 + ks*16 + 2*tq .. +1], b1 = ... + 8
  uint32_t bw[36][2];
#pragma unroll
  for (int i = 0; i < 36; ++i) {
    bw[i][0] = bw[i][1] = 0u;
    if (gid < rows_w) {
      const __nv_bfloat16* q = w + (long long)gid * Kf + i * 16 + 2 * tq;
      bw[i][0] = __ldg(reinterpret_cast<const uint32_t*>(q));
      bw[i][1] = __ldg(reinterpret_cast<const uint32_t*>(q + 8));
    }
  }
  const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
  const float nzw = e.noise != nullptr ? __ldg(e.noise_weight) : 0.f;
  float bias[2] = {0.f, 0.f};
  if (e.bias != nullptr) {
    if (2 * tq < e.ncols) bias[0] = __ldg(e.bias + 2 * tq);
    if (2 * tq + 1 < e.ncols) bias[1] = __ldg(e.bias + 2 * tq + 1);
  }
  const int mi = lane >> 3, mr = lane & 7;

  for (int row = rb; row < re; ++row) {
    const int img = row / g.Ho, oy = row - img * g.Ho;
    const int iy0 = oy - g.pad;
    if (row == rb || oy == 0) {
      __syncthreads();
      issue_x(img, iy0);
      issue_x(img, iy0 + 1);
      issue_x(img, iy0 + 2);
    }
    cp_async_commit();
    if (row + 1 < re && oy + 1 < g.Ho) issue_x(img, iy0 + 3);
    cp_async_commit();
    cp_async_wait_group<1>();
    __syncthreads();
    for (int ox0 = warp * 16; ox0 < g.Wo; ox0 += kHmWarps * 16) {
      // one accumulator per filter row: three independent mma chains of 12 instead of one of 36 (the kernel is bound by the
      // ldmatrix -> mma.sync latency chain, profiles/r01l_ncu_full_conv_heads_fwd.txt)
      float accr[3][4];
#pragma unroll
      for (int r = 0; r < 3; ++r) accr[r][0] = accr[r][1] = accr[r][2] = accr[r][3] = 0.f;
#pragma unroll
      for (int sx = 0; sx < 3; ++sx) {
        const int pix = ox0 + (mi & 1) * 8 + mr + sx;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t coff = pix * 128 + (((ks * 2 + (mi >> 1)) ^ (pix & 7)) << 4);
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            uint32_t a[4];
            ldsm_x4(a, s_x + ((iy0 + r + 4) & 3) * x_row_bytes + coff);
            mma_bf16_16816(accr[r], a, bw[(r * 3 + sx) * 4 + ks][0], bw[(r * 3 + sx) * 4 + ks][1]);
          }
        }
      }
      float acc[4];
#pragma unroll
      for (int e2 = 0; e2 < 4; ++e2) acc[e2] = accr[0][e2] + accr[1][e2] + accr[2][e2];
      // epilogue: rows (pixels) gid / gid + 8, columns 2*tq, 2*tq + 1
      const int col = 2 * tq;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int ox = ox0 + gid + 8 * h;
        if (ox >= g.Wo) continue;
        const long long m = (long long)row * g.Wo + ox;
        const float nz = e.noise != nullptr ? __ldg(e.noise + m) * nzw : 0.f;
        float v0 = apply_act(fmaf(acc[2 * h], osc, bias[0] + nz), e.act);
        float v1 = apply_act(fmaf(acc[2 * h + 1], osc, bias[1] + nz), e.act);
        if (col < e.ld) {
          if (e.residual != nullptr) {
            const uint32_t rv = __ldg(reinterpret_cast<const uint32_t*>(e.residual + m * e.ld + col));
            v0 += bf16lo(rv);
            v1 += bf16hi(rv);
          }
          if (e.out_fp32) *reinterpret_cast<float2*>(reinterpret_cast<float*>(e.out) + m * e.ld + col) = make_float2(v0, v1);
          else *reinterpret_cast<uint32_t*>(reinterpret_cast<__nv_bfloat16*>(e.out) + m * e.ld + col) = pack_bf16(v0, v1);
        }
      }
    }
    __syncthreads();
  }
  cp_async_wait_group<0>();
}

bool conv_heads_fwd_applicable(const Gather& g, int rows_ld) {
  const size_t sm = (size_t)kHmRing * (((g.Wo + 15) & ~15) + 2) * 128;
  return g_tune_conv_path == 0 && g.mode == 0 && g.kh == 3 && g.kw == 3 && g.stride == 1 && g.C == 64 && rows_ld == 8 && g.pad <= 1 &&
         g.Hs == g.Hl && g.Ws == g.Wl && !(g.pad_mode == DEI2I_PAD_REFLECT && g.pad > 0) && sm <= 200 * 1024 &&
         g.M < (1ll << 31);
}
int run_conv_heads_fwd(const Gather& g, const Epilogue& e, const void* wmat, int rows, int Kf, cudaStream_t st) {
  const size_t sm = (size_t)kHmRing * (((g.Wo + 15) & ~15) + 2) * 128;
  static size_t attr = 0;
  if (sm > attr) {
    DEI2I_CUDA(cudaFuncSetAttribute(conv_heads_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    attr = sm;
  }
  const int nrows = g.N * g.Ho;
  const int rpc = (nrows + 147) / 148;
  const int grid = (nrows + rpc - 1) / rpc;
  conv_heads_fwd_kernel<<<grid, kHmWarps * 32, sm, st>>>(g, e, (const __nv_bfloat16*)wmat, rows, Kf, rpc);
  count_launch();
  trace_kernel("conv_heads_fwd");
  return check_cuda(cudaGetLastError(), "conv_heads_fwd_kernel launch");
}

// ----------------------------------------------------------------------------------------------------------------------
// Forward of the discriminator's two output convolutions (2048 -> 1, 3x3 on the 4x4 map; 2048 -> label_nc, 4x4 'valid':
// discriminator.py:60-75): <= 8 output channels, a handful of output pixels and K = taps * Cin of 18-32 k.  On the UMMA path
// that is ONE CTA walking 300-500 K blocks (145 us); here one CTA per output pixel streams its taps (contiguous 16-byte
// chunks of x against the <= 8 weight rows, which stay in L1/L2) and reduces in a fixed order.
__global__ void __launch_bounds__(256)
conv_narrow_fwd_kernel(const Gather g, const Epilogue e, const __nv_bfloat16* __restrict__ w, int rows_w, int Kf) {
  __shared__ float red[8][8];
  const int m = blockIdx.x;
  const int HoWo = g.Ho * g.Wo;
  const int img = m / HoWo, rem = m - img * HoWo;
  const int oy = rem / g.Wo, ox = rem - oy * g.Wo;
  const int c8n = g.C >> 3;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int tap = 0; tap < g.ntaps; ++tap) {
    const int r = tap / g.kw, s_ = tap - r * g.kw;
    const int iy = oy * g.stride + r - g.pad, ix = ox * g.stride + s_ - g.pad;
    if (iy < 0 || iy >= g.Hl || ix < 0 || ix >= g.Wl) continue;  // zero frame (block-uniform)
    const uint4* xp = reinterpret_cast<const uint4*>(g.src + (((long long)img * g.Hs + iy) * g.Ws + ix) * g.C);
    const uint4* wp = reinterpret_cast<const uint4*>(w + (long long)tap * g.C);
    for (int c8 = threadIdx.x; c8 < c8n; c8 += blockDim.x) {
      float fx[8];
      unpack8_bf16(__ldg(xp + c8), fx);
#pragma unroll
      for (int co = 0; co < 8; ++co) {
        if (co < rows_w) {
          float fw[8];
          unpack8_bf16(__ldg(wp + (long long)co * (Kf >> 3) + c8), fw);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[co] = fmaf(fx[k], fw[k], acc[co]);
        }
      }
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int co = 0; co < 8; ++co) {
    float v = acc[co];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][co] = v;
  }
  __syncthreads();
  if (threadIdx.x < 8 && (int)threadIdx.x < e.ld) {
    const int co = threadIdx.x;
    float v = 0.f;
    for (int wv = 0; wv < 8; ++wv) v += red[wv][co];
    const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
    v = v * osc + ((e.bias != nullptr && co < e.ncols) ? __ldg(e.bias + co) : 0.f);
    v = apply_act(v, e.act);
    if (e.out_fp32) reinterpret_cast<float*>(e.out)[(long long)m * e.ld + co] = v;
    else reinterpret_cast<__nv_bfloat16*>(e.out)[(long long)m * e.ld + co] = __float2bfloat16(v);
  }
}

bool conv_narrow_fwd_applicable(const Gather& g, int rows_ld) {
  return g_tune_conv_path == 0 && g.mode == 0 && rows_ld == 8 && g.C % 8 == 0 && g.Hs == g.Hl && g.Ws == g.Wl &&
         !(g.pad_mode == DEI2I_PAD_REFLECT && g.pad > 0) && (long long)g.ntaps * g.C >= 4096 && g.M <= 8192;
}
int run_conv_narrow_fwd(const Gather& g, const Epilogue& e, const void* wmat, int rows, int Kf, cudaStream_t st) {
  conv_narrow_fwd_kernel<<<(unsigned)g.M, 256, 0, st>>>(g, e, (const __nv_bfloat16*)wmat, rows < 8 ? rows : 8, Kf);
  count_launch();
  trace_kernel("conv_narrow_fwd");
  return check_cuda(cudaGetLastError(), "conv_narrow_fwd_kernel launch");
}

bool wgrad_heads_applicable(const Gather& g, int Cout, int Cout_ld) {
  return g.mode == 0 && g.kh == 3 && g.kw == 3 && g.stride == 1 && g.C % 64 == 0 && kHeadsWarps % (g.C / 64) == 0 && g.C <= 128 &&
         Cout <= 4 && Cout_ld == 8 && g.Hs == g.Hl && g.Ws == g.Wl && !(g.pad_mode == DEI2I_PAD_REFLECT && g.pad > 0);
}
int wgrad_heads_splits() { return 148; }
int run_wgrad_heads(const Gather& g, const void* dy, int Cout, int Cout_ld, int Kf, float* partial, cudaStream_t st) {
  if (g_tune_conv_path != 2 && wgrad_heads_mma_applicable(g)) {
    const size_t sm = wgrad_heads_mma_smem(g.Wo);
    static size_t attr = 0;
    if (sm > attr) {
      DEI2I_CUDA(cudaFuncSetAttribute(wgrad_heads_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      attr = sm;
    }
    const int rows = g.N * g.Ho;
    const int rpc = (rows + 147) / 148;
    // CTAs past the last strip still write their (zero) partial tile: the second pass sums all 148
    wgrad_heads_mma_kernel<<<148, kHmWarps * 32, sm, st>>>(g, (const __nv_bfloat16*)dy, Cout, Kf, partial, rpc);
    count_launch();
  trace_kernel("wgrad_heads_mma");
    return check_cuda(cudaGetLastError(), "wgrad_heads_mma_kernel launch");
  }
  wgrad_heads_kernel<<<148, kHeadsWarps * 32, 0, st>>>(g, (const __nv_bfloat16*)dy, Cout, Cout_ld, Kf, partial);
  count_launch();
  trace_kernel("wgrad_heads");
  return check_cuda(cudaGetLastError(), "wgrad_heads_kernel launch");
}

}  // namespace dei2i
