// Implicit-GEMM convolution kernels for sm_100a: tcgen05.mma (BF16 in, FP32 accumulate in TMEM).
//
//   forward / dgrad :  D[m, n] = sum_k A[m, k] * B[n, k]
//        m = output pixel (n_img, oy, ox), n = output channel, k = (tap, input channel)
//        A : gathered on the fly from the NHWC activation (im2col never materialised) by 4 producer warps with
//            16-byte cp.async into the 128B-swizzled K-major UMMA layout (reflect / zero padding, stride,
//            nearest resize and the transposed-conv index map of dgrad are all just address arithmetic);
//        B : packed weight matrix [n, k], TMA 2D tiled loads (SWIZZLE_128B), one elected thread;
//        MMA: one thread issues tcgen05.mma 128 x BLOCK_N x 16, accumulator = BLOCK_N TMEM columns;
//        epilogue: the 4 producer warps read TMEM (tcgen05.ld 32x32b), fuse bias / noise / activation /
//            residual and store NHWC rows.  Two CTAs are resident per SM (<= 100 KB smem, 2 x 256 TMEM
//            columns), so one CTA's epilogue overlaps the other's main loop.
//   wgrad :  dW[co, k] = sum_m dY[m, co] * A[m, k]   (reduction over pixels, split across CTAs)
//        both operands are "MN-major" (the reduction index m is the slow one in memory): dY tiles come by TMA,
//        im2col tiles by the same cp.async gather; FP32 partial tiles go to a workspace, a second pass sums the
//        splits in a fixed order (deterministic) and applies the spectral-norm backward.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"
#include "conv_common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {

extern int g_tune_conv_stages;  // 0 = auto
extern int g_tune_conv_bn;      // 0 = auto
extern int g_tune_wgrad_bn;     // 0 = auto
extern int g_tune_wgrad_waves;
extern int g_tune_wgrad_swap;
extern int g_tune_conv_path;
void count_launch(int n = 1);
void trace_kernel(const char* fmt, ...);
void trace_begin();

struct ConvKParams {
  Gather g;
  Epilogue e;
  int num_kb;  // K / 64
};

constexpr int kProducerThreads = 128;
constexpr int kThreads = 192;
constexpr int kATileBytes = 128 * 128;  // 128 rows x 64 bf16

template <int BLOCK_N, int STAGES>
struct ConvSmem {
  static constexpr int kBTileBytes = BLOCK_N * 128;
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  static constexpr int kBarrierOff = STAGES * kStageBytes;
  static constexpr int kTotal = kBarrierOff + (2 * STAGES + 1) * 8 + 16 + 1024 /*alignment slack*/;
};

template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(kThreads, (ConvSmem<BLOCK_N, STAGES>::kTotal <= 112 * 1024) ? 2 : 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_b, const ConvKParams p) {
  using S = ConvSmem<BLOCK_N, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarrierOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const long long m0 = (long long)blockIdx.x * 128;
  const int n0 = blockIdx.y * BLOCK_N;
  constexpr uint32_t kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], kProducerThreads + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4 && lane == 0) tma_prefetch_desc(&tmap_b);
  if (warp == 5) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp < 4) {
    // =========================== A producers (im2col gather) ===========================
    const Gather& g = p.g;
    const int tid = threadIdx.x;
    const int j = tid & 7;    // 16-byte chunk inside the 128-byte K row
    const int rb = tid >> 3;  // rows rb + 16*i
    int r_img[8], r_oy[8], r_ox[8];
    const int HoWo = g.Ho * g.Wo;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      long long m = m0 + rb + 16 * i;
      if (m < g.M) {
        int img = (int)(m / HoWo);
        int rem = (int)(m - (long long)img * HoWo);
        r_img[i] = img;
        r_oy[i] = rem / g.Wo;
        r_ox[i] = rem - r_oy[i] * g.Wo;
      } else {
        r_img[i] = -1;
        r_oy[i] = 0;
        r_ox[i] = 0;
      }
    }
    const uint32_t a_row_off = (uint32_t)(rb * 128 + ((j ^ (rb & 7)) << 4));
    for (int kb = 0; kb < p.num_kb; ++kb) {
      const int stage = kb % STAGES;
      const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
      mbar_wait(&empty_bar[stage], ph ^ 1u);
      const int q8 = kb * 8 + j;
      const int tap = q8 / g.cin8;
      const int c8 = q8 - tap * g.cin8;
      const bool tap_ok = tap < g.ntaps;
      const int r = tap / g.kw;
      const int s = tap - r * g.kw;
      const uint32_t a_stage = smem_u32(smem + stage * S::kStageBytes) + a_row_off;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const __nv_bfloat16* src = nullptr;
        if (tap_ok && r_img[i] >= 0) src = gather_ptr(g, r_img[i], r_oy[i], r_ox[i], r, s, c8);
        cp_async_16(a_stage + i * (16 * 128), src ? (const void*)src : (const void*)g.src, src != nullptr);
      }
      cp_async_mbar_arrive_noinc(&full_bar[stage]);
    }

    // =========================== epilogue ===========================
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const Epilogue& e = p.e;
    const long long m = m0 + warp * 32 + lane;
    const bool row_ok = m < g.M;
    float nz = 0.f;
    if (e.noise != nullptr && row_ok) nz = __ldg(e.noise + m) * __ldg(e.noise_weight);
    const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
    const float slope = e.act == ACT_LRELU ? 0.2f : (e.act == ACT_RELU ? 0.f : 1.f);  // none / leaky / relu = max(x, slope * x)
    const bool slow_act = e.act == ACT_TANH || e.act == ACT_SIGMOID;
    constexpr int CH = BLOCK_N >= 32 ? 32 : 16;
#pragma unroll 1
    for (int c = 0; c < BLOCK_N; c += CH) {
      uint32_t v[CH];
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c;
      if constexpr (CH == 32) tmem_ld_32x32b_x32(taddr, v);
      else tmem_ld_32x32b_x16(taddr, v);
      tc_wait_ld();
      const int nbase = n0 + c;
      if (row_ok && nbase < e.ld) {
        float f[CH];
#pragma unroll
        for (int t = 0; t < CH; ++t) {
          float x = __uint_as_float(v[t]) * osc;
          const int n = nbase + t;
          if (e.bias != nullptr && n < e.ncols) x += __ldg(e.bias + n);
          x += nz;
          f[t] = slow_act ? apply_act(x, e.act) : fmaxf(x, x * slope);
        }
        if (e.residual != nullptr) {
          const __nv_bfloat16* rp = e.residual + m * e.ld + nbase;
#pragma unroll
          for (int t8 = 0; t8 < CH; t8 += 8) {
            if (nbase + t8 < e.ld) {
              uint4 rv = __ldg(reinterpret_cast<const uint4*>(rp + t8));
              f[t8 + 0] += bf16lo(rv.x); f[t8 + 1] += bf16hi(rv.x);
              f[t8 + 2] += bf16lo(rv.y); f[t8 + 3] += bf16hi(rv.y);
              f[t8 + 4] += bf16lo(rv.z); f[t8 + 5] += bf16hi(rv.z);
              f[t8 + 6] += bf16lo(rv.w); f[t8 + 7] += bf16hi(rv.w);
            }
          }
        }
        if (e.out_fp32) {
          float* op = reinterpret_cast<float*>(e.out) + m * e.ld + nbase;
#pragma unroll
          for (int t4 = 0; t4 < CH; t4 += 4)
            if (nbase + t4 < e.ld) *reinterpret_cast<float4*>(op + t4) = make_float4(f[t4], f[t4 + 1], f[t4 + 2], f[t4 + 3]);
        } else {
          __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(e.out) + m * e.ld + nbase;
#pragma unroll
          for (int t8 = 0; t8 < CH; t8 += 8)
            if (nbase + t8 < e.ld)
              *reinterpret_cast<uint4*>(op + t8) = make_uint4(pack_bf16(f[t8], f[t8 + 1]), pack_bf16(f[t8 + 2], f[t8 + 3]),
                                                              pack_bf16(f[t8 + 4], f[t8 + 5]), pack_bf16(f[t8 + 6], f[t8 + 7]));
        }
      }
    }
  } else if (warp == 4) {
    // =========================== B producer (TMA) ===========================
    if (lane == 0) {
      for (int kb = 0; kb < p.num_kb; ++kb) {
        const int stage = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(&empty_bar[stage], ph ^ 1u);
        mbar_arrive_expect_tx(&full_bar[stage], S::kBTileBytes);
        tma_load_2d(smem + stage * S::kStageBytes + kATileBytes, &tmap_b, &full_bar[stage], kb * 64, n0);
      }
    }
  } else {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BLOCK_N, 0, 0);
      for (int kb = 0; kb < p.num_kb; ++kb) {
        const int stage = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(&full_bar[stage], ph);
        fence_proxy_async_smem();  // cp.async (generic proxy) data -> tensor-core (async proxy) reads
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * S::kStageBytes);
        const uint32_t b_addr = a_addr + kATileBytes;
        const uint64_t a_desc = make_smem_desc(a_addr, 16, 1024, 2);
        const uint64_t b_desc = make_smem_desc(b_addr, 16, 1024, 2);
#pragma unroll
        for (int k = 0; k < 4; ++k)  // 4 x (K = 16): +32 bytes inside the 128-byte swizzle span
          tc_mma_bf16(tmem_base, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
        tc_commit(&empty_bar[stage]);
      }
      tc_commit(tmem_full_bar);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ----------------------------------------------------------------------------------------------------------
// wgrad:  P[co, k] = sum_{m in split} dY[m, co] * A[m, k]       (M' = 128 co, N' = BLOCK_N k, K' = 64 pixels/step)
struct WgradKParams {
  Gather g;        // forward gather geometry of x
  float* partial;  // [splits, Cout_rows, Kf] fp32
  int cout_rows;   // rows stored per split (= Cout real)
  int Kf;          // packed K (multiple of 64)
  int pb_per_split;  // pixel blocks (64 pixels) per split
  int num_pb;        // total pixel blocks
};

template <int BLOCK_N, int STAGES>
struct WgradSmem {
  static constexpr int kATileBytes = 2 * 8192;              // 64 pixels x 128 co
  static constexpr int kBTileBytes = (BLOCK_N / 64) * 8192; // 64 pixels x BLOCK_N k
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  static constexpr int kBarrierOff = STAGES * kStageBytes;
  static constexpr int kTotal = kBarrierOff + (2 * STAGES + 1) * 8 + 16 + 1024;
};

template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(kThreads, (WgradSmem<BLOCK_N, STAGES>::kTotal <= 112 * 1024) ? 2 : 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const WgradKParams p) {
  using S = WgradSmem<BLOCK_N, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarrierOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * BLOCK_N;  // K offset (elements)
  const int co0 = blockIdx.y * 128;
  const int split = blockIdx.z;
  const int pb0 = split * p.pb_per_split;
  const int pb1 = min(p.num_pb, pb0 + p.pb_per_split);
  const int nsteps = pb1 - pb0;
  constexpr uint32_t kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], kProducerThreads + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4 && lane == 0) tma_prefetch_desc(&tmap_dy);
  if (warp == 5) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp < 4) {
    // ---------------- im2col producer: 64 pixels x BLOCK_N k-elements, MN-major 128B-swizzled atoms
    const Gather& g = p.g;
    const int tid = threadIdx.x;
    const int cj = tid & 7;
    const int pr = tid >> 3;  // pixels pr + 16*i
    constexpr int NG = BLOCK_N / 64;
    int t_r[NG], t_s[NG], t_c8[NG];
    bool t_ok[NG];
#pragma unroll
    for (int gi = 0; gi < NG; ++gi) {
      const int q8 = (k0 / 64 + gi) * 8 + cj;
      const int tap = q8 / g.cin8;
      t_c8[gi] = q8 - tap * g.cin8;
      t_ok[gi] = tap < g.ntaps;
      t_r[gi] = tap / g.kw;
      t_s[gi] = tap - t_r[gi] * g.kw;
    }
    const int HoWo = g.Ho * g.Wo;
    const uint32_t row_off = (uint32_t)((pr >> 3) * 1024 + (pr & 7) * 128 + ((cj ^ (pr & 7)) << 4));
    for (int it = 0; it < nsteps; ++it) {
      const int stage = it % STAGES;
      const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
      mbar_wait(&empty_bar[stage], ph ^ 1u);
      const uint32_t b_stage = smem_u32(smem + stage * S::kStageBytes + S::kATileBytes) + row_off;
      const long long mb = (long long)(pb0 + it) * 64;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const long long m = mb + pr + 16 * i;
        int img = -1, oy = 0, ox = 0;
        if (m < g.M) {
          img = (int)(m / HoWo);
          int rem = (int)(m - (long long)img * HoWo);
          oy = rem / g.Wo;
          ox = rem - oy * g.Wo;
        }
#pragma unroll
        for (int gi = 0; gi < NG; ++gi) {
          const __nv_bfloat16* src = nullptr;
          if (img >= 0 && t_ok[gi]) src = gather_ptr(g, img, oy, ox, t_r[gi], t_s[gi], t_c8[gi]);
          cp_async_16(b_stage + gi * 8192 + i * 2048, src ? (const void*)src : (const void*)g.src, src != nullptr);
        }
      }
      cp_async_mbar_arrive_noinc(&full_bar[stage]);
    }
    // ---------------- epilogue: fp32 partial tile -> workspace
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const int co = co0 + warp * 32 + lane;
    float* orow = p.partial + ((long long)split * p.cout_rows + co) * p.Kf + k0;
    constexpr int CH = 32;
#pragma unroll 1
    for (int c = 0; c < BLOCK_N; c += CH) {
      uint32_t v[CH];
      tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
      tc_wait_ld();
      if (co < p.cout_rows && k0 + c < p.Kf) {
#pragma unroll
        for (int t4 = 0; t4 < CH; t4 += 4)
          *reinterpret_cast<float4*>(orow + c + t4) =
              make_float4(__uint_as_float(v[t4]), __uint_as_float(v[t4 + 1]), __uint_as_float(v[t4 + 2]), __uint_as_float(v[t4 + 3]));
      }
    }
  } else if (warp == 4) {
    if (lane == 0) {
      for (int it = 0; it < nsteps; ++it) {
        const int stage = it % STAGES;
        const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
        mbar_wait(&empty_bar[stage], ph ^ 1u);
        mbar_arrive_expect_tx(&full_bar[stage], S::kATileBytes);
        uint8_t* a_dst = smem + stage * S::kStageBytes;
        tma_load_2d(a_dst, &tmap_dy, &full_bar[stage], co0, (pb0 + it) * 64);
        tma_load_2d(a_dst + 8192, &tmap_dy, &full_bar[stage], co0 + 64, (pb0 + it) * 64);
      }
    }
  } else {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BLOCK_N, 1, 1);
      for (int it = 0; it < nsteps; ++it) {
        const int stage = it % STAGES;
        const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
        mbar_wait(&full_bar[stage], ph);
        fence_proxy_async_smem();
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * S::kStageBytes);
        const uint32_t b_addr = a_addr + S::kATileBytes;
        // MN-major, 128B swizzle: LBO = stride between 64-element MN atoms (8192 B), SBO = between 8-row K atoms (1024 B)
        const uint64_t a_desc = make_smem_desc(a_addr, 8192, 1024, 2);
        const uint64_t b_desc = make_smem_desc(b_addr, 8192, 1024, 2);
#pragma unroll
        for (int k = 0; k < 4; ++k)  // 16 pixels = 2 K atoms = 2048 B per MMA
          tc_mma_bf16(tmem_base, a_desc + (uint64_t)(128 * k), b_desc + (uint64_t)(128 * k), idesc, (it | k) != 0);
        tc_commit(&empty_bar[stage]);
      }
      tc_commit(tmem_full_bar);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// second pass over the FP32 partial tiles P[split, co, tap*Cin + ci]: the weight gradient lives in weight memory order
// [co][ci][tap], i.e. (tap, ci) transposed.  One work tile = (co, CT input channels, all taps), staged through shared
// memory so that both the reads of P (CT consecutive ci per tap) and the accesses in weight order (CT*taps consecutive
// floats) are coalesced 16-byte accesses; 256 threads.  CT = 256 for the large discriminator weights (16 KB+ in flight per
// block), 64 / 16 for small problems so that enough blocks exist.  Split sums run in a fixed order (deterministic).
// Requires Cin_real % 4 == 0 for the vector path (VEC); the scalar path handles 3- and 6-channel inputs.
template <bool VEC>
__device__ __forceinline__ void wgrad_tile_load(const float* __restrict__ partial, int splits, long long split_stride, int Kf, int Cin,
                                                int Cin_real, int taps, int co, int ci0, int sh, int pitch, float* __restrict__ T) {
  const float* base = partial + (long long)co * Kf + ci0;
  if (VEC) {
    const int sh4 = sh - 2;
    for (int idx = threadIdx.x; idx < (taps << sh4); idx += blockDim.x) {
      const int tap = idx >> sh4, j = (idx & ((1 << sh4) - 1)) * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ci0 + j < Cin_real) {
        const float* q = base + tap * Cin + j;
        for (int s = 0; s < splits; ++s) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(q + (long long)s * split_stride));
          acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
      }
      *reinterpret_cast<float4*>(T + tap * pitch + j) = acc;
    }
  } else {
    for (int idx = threadIdx.x; idx < (taps << sh); idx += blockDim.x) {
      const int tap = idx >> sh, j = idx & ((1 << sh) - 1);
      float acc = 0.f;
      if (ci0 + j < Cin_real) {
        const float* q = base + tap * Cin + j;
        for (int s = 0; s < splits; ++s) acc += q[(long long)s * split_stride];
      }
      T[tap * pitch + j] = acc;
    }
  }
}

//   dw[co, ci, r, s] = beta*dw + scale * (sum_splits P - corr * u[co] * v[col])
template <bool VEC>
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ partial, int splits, int Cout, int Kf, int Cin, int Cin_real, int taps,
                    float* __restrict__ dw, float beta, const float* __restrict__ dot, const float* __restrict__ u,
                    const float* __restrict__ v, const float* __restrict__ inv_sigma, int sh) {
  extern __shared__ float4 T4[];
  float* T = reinterpret_cast<float*>(T4);
  const int ct = 1 << sh, pitch = ct + 4;
  const int ci_tiles = (Cin_real + ct - 1) >> sh;
  const int co = blockIdx.x / ci_tiles;
  const int ci0 = (blockIdx.x - co * ci_tiles) << sh;
  wgrad_tile_load<VEC>(partial, splits, (long long)Cout * Kf, Kf, Cin, Cin_real, taps, co, ci0, sh, pitch, T);
  __syncthreads();
  const int n = min(ct, Cin_real - ci0) * taps;
  const int cols = Cin_real * taps;
  const int col0 = ci0 * taps;
  float* out = dw + (long long)co * cols + col0;
  float corr = 0.f, scale = 1.f;
  if (u != nullptr) {
    corr = dot[0] * u[co];
    scale = inv_sigma[0];
  }
  if (VEC) {
    for (int o = threadIdx.x * 4; o < n; o += blockDim.x * 4) {
      float a[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cl = (o + k) / taps, tap = (o + k) - cl * taps;
        a[k] = T[tap * pitch + cl];
      }
      if (u != nullptr) {
        const float4 vv = __ldg(reinterpret_cast<const float4*>(v + col0 + o));
        a[0] = (a[0] - corr * vv.x) * scale; a[1] = (a[1] - corr * vv.y) * scale;
        a[2] = (a[2] - corr * vv.z) * scale; a[3] = (a[3] - corr * vv.w) * scale;
      }
      float4* q = reinterpret_cast<float4*>(out + o);
      if (beta != 0.f) {
        const float4 old = *q;
        a[0] += beta * old.x; a[1] += beta * old.y; a[2] += beta * old.z; a[3] += beta * old.w;
      }
      *q = make_float4(a[0], a[1], a[2], a[3]);
    }
  } else {
    for (int o = threadIdx.x; o < n; o += blockDim.x) {
      const int cl = o / taps, tap = o - cl * taps;
      float acc = T[tap * pitch + cl];
      if (u != nullptr) acc = (acc - corr * v[col0 + o]) * scale;
      out[o] = beta != 0.f ? beta * out[o] + acc : acc;
    }
  }
}

static __device__ unsigned int g_dot_ticket = 0;  // blocks of the wgrad_dot_kernel in flight that have finished (see there)
// <dW, W/sigma> over the real elements (spectral-norm backward); deterministic: fixed tile -> block assignment, fixed order
template <bool VEC>
__global__ void __launch_bounds__(256)
wgrad_dot_kernel(float* __restrict__ partial, int splits, int Cout, int Kf, int Cin, int Cin_real, int taps,
                 const float* __restrict__ w_orig, const float* __restrict__ inv_sigma, float* __restrict__ blocksums, int sh,
                 int store_sum, float* __restrict__ total_out) {
  extern __shared__ float4 T4[];
  float* T = reinterpret_cast<float*>(T4);
  const int ct = 1 << sh, pitch = ct + 4;
  const int ci_tiles = (Cin_real + ct - 1) >> sh;
  const int tiles = Cout * ci_tiles;
  const int cols = Cin_real * taps;
  float acc = 0.f;
  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int co = t / ci_tiles;
    const int ci0 = (t - co * ci_tiles) << sh;
    const int n = min(ct, Cin_real - ci0) * taps;
    const float* w = w_orig + (long long)co * cols + ci0 * taps;
    // the weight values of this tile are fetched before the barrier so that both streams are in flight together
    float4 wv[4];
    if (VEC) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int o = (threadIdx.x + r * 256) * 4;
        wv[r] = o < n ? __ldg(reinterpret_cast<const float4*>(w + o)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    __syncthreads();
    wgrad_tile_load<VEC>(partial, splits, (long long)Cout * Kf, Kf, Cin, Cin_real, taps, co, ci0, sh, pitch, T);
    __syncthreads();
    if (store_sum && splits > 1) {
      // the split sums of this tile replace split 0 in place (tiles are disjoint): wgrad_reduce_kernel then reads one split
      // instead of summing all of them a second time
      float* dst = partial + (long long)co * Kf + ci0;
      if (VEC) {
        const int sh4 = sh - 2;
        for (int idx = threadIdx.x; idx < (taps << sh4); idx += blockDim.x) {
          const int tap = idx >> sh4, j = (idx & ((1 << sh4) - 1)) * 4;
          if (ci0 + j < Cin_real) *reinterpret_cast<float4*>(dst + tap * Cin + j) = *reinterpret_cast<const float4*>(T + tap * pitch + j);
        }
      } else {
        for (int idx = threadIdx.x; idx < (taps << sh); idx += blockDim.x) {
          const int tap = idx >> sh, j = idx & ((1 << sh) - 1);
          if (ci0 + j < Cin_real) dst[tap * Cin + j] = T[tap * pitch + j];
        }
      }
    }
    if (VEC) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int o = (threadIdx.x + r * 256) * 4;
        if (o < n) {
          const float wk[4] = {wv[r].x, wv[r].y, wv[r].z, wv[r].w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int cl = (o + k) / taps, tap = (o + k) - cl * taps;
            acc += T[tap * pitch + cl] * wk[k];
          }
        }
      }
      for (int o = (threadIdx.x + 1024) * 4; o < n; o += blockDim.x * 4) {  // tiles beyond 4096 elements (taps > 16)
        const float4 t4 = __ldg(reinterpret_cast<const float4*>(w + o));
        const float wk[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int cl = (o + k) / taps, tap = (o + k) - cl * taps;
          acc += T[tap * pitch + cl] * wk[k];
        }
      }
    } else {
      for (int o = threadIdx.x; o < n; o += blockDim.x) {
        const int cl = o / taps, tap = o - cl * taps;
        acc += T[tap * pitch + cl] * w[o];
      }
    }
  }
  __shared__ float wsum[32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? wsum[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) blocksums[blockIdx.x] = t * inv_sigma[0];
  }
  if (total_out == nullptr) return;
  // the block that finishes last adds the block sums up in a fixed order (thread-strided partial sums, warp shuffles, one final warp - what the separate
  // one-block kernel did before, bit for bit), which saves that launch; the ticket lives in a file-scope counter that the last block resets (one wgrad in flight per device at a time:
  // they all run on the trainer's stream)
  __shared__ bool last;
  __threadfence();
  if (threadIdx.x == 0) last = atomicAdd(&g_dot_ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  float a2 = 0.f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) a2 += __ldcg(blocksums + i);
  a2 = warp_sum(a2);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = a2;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? wsum[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) {
      total_out[0] = t;
      g_dot_ticket = 0;
    }
  }
}

// ----------------------------------------------------------------------------------------------------------
// host side
static int conv_out_dim(int L, int k, int stride, int pad) { return (L + 2 * pad - k) / stride + 1; }
static int round_up(int a, int b) { return (a + b - 1) / b * b; }

template <int BN, int ST>
static int launch_conv(const CUtensorMap& tm, const ConvKParams& p, int n_tiles, cudaStream_t st) {
  using S = ConvSmem<BN, ST>;
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<BN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_done = true;
  }
  long long m_tiles = (p.g.M + 127) / 128;
  DEI2I_REQUIRE(m_tiles < (1ll << 31), "conv: too many M tiles");
  dim3 grid((unsigned)m_tiles, (unsigned)n_tiles, 1);
  conv_igemm_kernel<BN, ST><<<grid, kThreads, S::kTotal, st>>>(tm, p);
  count_launch();
  trace_kernel("conv_igemm<%d,%d>", BN, ST);
  return check_cuda(cudaGetLastError(), "conv_igemm_kernel launch");
}

static int run_igemm(const ConvKParams& p, const void* wmat, int rows, int K, cudaStream_t st) {
  // choose the N tile: largest of {256,128,64,32,16} that does not exceed round_up(rows,16)
  int bn = 256;
  int want = round_up(rows, 16);
  while (bn > 16 && bn / 2 >= want) bn /= 2;
  if (g_tune_conv_bn > 0 && g_tune_conv_bn <= bn) bn = g_tune_conv_bn;
  CUtensorMap tm;
  int rc = encode_tmap_2d_bf16(&tm, wmat, (uint64_t)K, (uint64_t)rows, (uint64_t)K * 2, 64, (uint32_t)bn, 1);
  if (rc) return rc;
  const int n_tiles = (rows + bn - 1) / bn;
  const int deep = g_tune_conv_stages;  // 0: two CTAs per SM; 1: one CTA per SM with a deep pipeline
  switch (bn) {
    case 256: return deep ? launch_conv<256, 4>(tm, p, n_tiles, st) : launch_conv<256, 2>(tm, p, n_tiles, st);
    case 128: return deep ? launch_conv<128, 6>(tm, p, n_tiles, st) : launch_conv<128, 3>(tm, p, n_tiles, st);
    case 64: return launch_conv<64, 4>(tm, p, n_tiles, st);
    case 32: return launch_conv<32, 4>(tm, p, n_tiles, st);
    default: return launch_conv<16, 4>(tm, p, n_tiles, st);
  }
}

static int check_geom(const dei2i_conv_geom* g) {
  DEI2I_REQUIRE(g != nullptr, "conv: null geometry");
  DEI2I_REQUIRE(g->N > 0 && g->H > 0 && g->W > 0 && g->Hs > 0 && g->Ws > 0, "conv: bad sizes");
  DEI2I_REQUIRE(g->Cin > 0 && g->Cin % 8 == 0, "conv: Cin (%d) must be a positive multiple of 8", g->Cin);
  DEI2I_REQUIRE(g->Cout > 0 && g->Cout_ld % 8 == 0 && g->Cout_ld >= g->Cout, "conv: bad Cout/Cout_ld (%d/%d)", g->Cout, g->Cout_ld);
  DEI2I_REQUIRE(g->kh > 0 && g->kw > 0 && g->stride > 0 && g->pad >= 0, "conv: bad kernel geometry");
  DEI2I_REQUIRE(g->pad_mode == DEI2I_PAD_ZEROS || g->pad_mode == DEI2I_PAD_REFLECT, "conv: bad pad_mode");
  if (g->pad_mode == DEI2I_PAD_REFLECT)
    DEI2I_REQUIRE(g->pad < g->H && g->pad < g->W, "conv: reflect padding (%d) must be smaller than the input (%dx%d)", g->pad, g->H, g->W);
  DEI2I_REQUIRE(conv_out_dim(g->H, g->kh, g->stride, g->pad) > 0 && conv_out_dim(g->W, g->kw, g->stride, g->pad) > 0,
                "conv: empty output");
  return 0;
}

// The TMA-im2col kernels take: zero / no padding, no on-the-fly resize, square filters, a K-operand channel count that is
// a multiple of 64 and whole 128-row tiles.  Everything else (tiny channel counts, resized label maps, ragged sizes,
// reflect padding that was not pre-applied) runs on the cp.async-gather kernels above.
static bool tma_applicable(const dei2i_conv_geom* g, int k_channels, long long M) {
  if (g_tune_conv_path == 1) return false;
  if (g->pad_mode == DEI2I_PAD_REFLECT && g->pad > 0) return false;
  if (g->Hs != g->H || g->Ws != g->W) return false;
  if (g->kh > 16 || g->kw > 16 || g->pad > 64 || g->stride > 8) return false;
  if (g->kh != g->kw && g->pad != 0) return false;  // non-square filters: 'valid' convolutions only (pre-padded inputs)
  (void)M;  // a ragged last tile is fine: im2col TMA zero-fills past the end of the tensor, the epilogue masks rows >= M
  if (k_channels % 64 != 0) return false;
  return true;
}

// Second pass of every weight gradient (shared by the BF16 tcgen05 kernels and the FP32 path): sums the split partials
// [splits, Cout, Kf] in a fixed order, re-lays [co][tap][ci] -> [co][ci][tap], and applies the spectral-norm backward
// dW_orig = (dW - <dW, W/sigma> u v^T) / sigma when the four sn pointers are given.  partial needs 1024 floats of slack at its end.
int wgrad_second_pass(float* partial, int splits, const dei2i_conv_geom* g, int Cin_real, int Kf, float* dw, float beta,
                      const float* sn_w_orig, const float* sn_u, const float* sn_v, const float* sn_inv_sigma, cudaStream_t st) {
  const int taps = g->kh * g->kw;
  float* dot = partial + (size_t)splits * g->Cout * Kf;  // [0] = <dW, W/sigma>, [1..] block sums
  const long long total = (long long)g->Cout * Cin_real * taps;
  // reduction tile width (input channels): 256 when that still leaves >= 4 blocks per SM and the tile fits 48 KB, else 64 / 16
  const bool red_vec = Cin_real % 4 == 0 && (((uintptr_t)dw | (uintptr_t)sn_v | (uintptr_t)sn_w_orig) & 15) == 0;
  int red_sh = 4;
  if (red_vec && Cin_real % 256 == 0 && (long long)g->Cout * (Cin_real / 256) >= 4 * 148 && taps * 260 * 4 <= 48 * 1024) red_sh = 8;
  else if ((long long)g->Cout * ((Cin_real + 63) / 64) >= 2 * 148) red_sh = 6;
  const int red_ci = 1 << red_sh;
  const int red_tiles = g->Cout * ((Cin_real + red_ci - 1) / red_ci);
  const size_t red_smem = (size_t)taps * (red_ci + 4) * sizeof(float);
  DEI2I_REQUIRE(red_smem <= 48 * 1024, "conv2d_wgrad: filter too large for the reduction tile");
  int red_splits = splits;
  if (sn_u != nullptr) {
    const int nb = red_tiles < 1023 ? red_tiles : 1023;
    if (red_vec)
      wgrad_dot_kernel<true><<<nb, 256, red_smem, st>>>(partial, splits, g->Cout, Kf, g->Cin, Cin_real, taps, sn_w_orig, sn_inv_sigma,
                                                        dot + 1, red_sh, 1, dot);
    else
      wgrad_dot_kernel<false><<<nb, 256, red_smem, st>>>(partial, splits, g->Cout, Kf, g->Cin, Cin_real, taps, sn_w_orig, sn_inv_sigma,
                                                         dot + 1, red_sh, 1, dot);
    red_splits = 1;  // the dot pass left the split sums in split 0
    count_launch(1);
    DEI2I_CUDA(cudaGetLastError());
  }
  if (red_vec)
    wgrad_reduce_kernel<true><<<(unsigned)red_tiles, 256, red_smem, st>>>(partial, red_splits, g->Cout, Kf, g->Cin, Cin_real, taps, dw, beta,
                                                                          dot, sn_u, sn_v, sn_inv_sigma, red_sh);
  else
    wgrad_reduce_kernel<false><<<(unsigned)red_tiles, 256, red_smem, st>>>(partial, red_splits, g->Cout, Kf, g->Cin, Cin_real, taps, dw, beta,
                                                                           dot, sn_u, sn_v, sn_inv_sigma, red_sh);
  count_launch();
  return check_cuda(cudaGetLastError(), "wgrad_reduce_kernel launch");
}

// ---- FP32-storage TF32 tensor-core variants (FP32-parity path with DEI2I_KERNELS=tf32): the same TMA-im2col geometry rules with 32
// channels per 128-byte row.  Return 1 when the geometry is not eligible (the caller then runs the exact FP32 FMA kernel).
extern int g_tune_f32_tf32;
static bool tf32_applicable(const dei2i_conv_geom* g, int k_channels) {
  if (!g_tune_f32_tf32) return false;
  if (g->pad_mode == DEI2I_PAD_REFLECT && g->pad > 0) return false;
  if (g->Hs != g->H || g->Ws != g->W) return false;
  if (g->kh > 16 || g->kw > 16 || g->pad > 64 || g->stride > 8) return false;
  if (g->kh != g->kw && g->pad != 0) return false;
  return k_channels % 32 == 0 && k_channels >= 32;
}
int conv_fwd_tf32(const float* x, const float* w_fwd, float* y, const dei2i_conv_geom* g, const dei2i_conv_epilogue* e, cudaStream_t st) {
  if (!tf32_applicable(g, g->Cin) || g->Cout_ld < 32) return 1;
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  TmaConvParams t;
  memset(&t, 0, sizeof(t));
  t.Ho = Ho; t.Wo = Wo; t.M = (long long)g->N * Ho * Wo; t.stride = g->stride; t.lower = -g->pad; t.lower_w = -g->pad;
  t.kh = g->kh; t.kw = g->kw; t.flip = 0; t.cin_blocks = g->Cin / 32; t.num_kb = g->kh * g->kw * t.cin_blocks;
  t.b_sub = 1; t.b_kw = g->kw;
  t.om.linear = 1;
  t.e.scale = e->scale; t.e.bias = e->bias; t.e.noise = e->noise; t.e.noise_weight = e->noise_weight;
  t.e.residual = (const __nv_bfloat16*)e->residual; t.e.res_fp32 = 1; t.e.out = y; t.e.act = e->act; t.e.out_fp32 = 1;
  t.e.ld = g->Cout_ld; t.e.ncols = g->Cout;
  // the packed FP32 weight rows are Kf = round_up(K, 64) long (K itself need only be a multiple of 32 here)
  return run_conv_tma(x, g->N, g->H, g->W, g->Cin, w_fwd, g->Cout, round_up(g->kh * g->kw * g->Cin, 64), t, g->pad - (g->kh - 1),
                      g->pad - (g->kw - 1), st, true);
}
int conv_dgrad_tf32(const float* dy, const float* w_dgrad, float* dx, const dei2i_conv_geom* g, const float* scale, cudaStream_t st) {
  if (!tf32_applicable(g, g->Cout_ld) || g->Cin < 32) return 1;
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  if (g->stride > 1) {
    if (!(g->pad == 0 && g->kh == g->kw && g->kh % g->stride == 0 && (g->H - g->kh) % g->stride == 0 && (g->W - g->kw) % g->stride == 0)) return 1;
    const int sd = g->stride, kt = g->kh / sd;
    TmaConvParams t;
    memset(&t, 0, sizeof(t));
    t.Ho = g->H / sd; t.Wo = g->W / sd; t.M = (long long)g->N * t.Ho * t.Wo; t.stride = 1; t.lower = -(kt - 1); t.lower_w = -(kt - 1);
    t.kh = kt; t.kw = kt; t.flip = 1; t.cin_blocks = g->Cout_ld / 32; t.num_kb = kt * kt * t.cin_blocks;
    t.b_sub = sd; t.b_py = 0; t.b_px = 0; t.b_kw = g->kw; t.npar = sd * sd;
    t.om.linear = 0; t.om.Ho = t.Ho; t.om.Wo = t.Wo; t.om.OH = g->H; t.om.OW = g->W; t.om.osy = sd; t.om.osx = sd;
    t.om.oy0 = 0; t.om.ox0 = 0;
    t.e.scale = scale; t.e.out = dx; t.e.act = 0; t.e.out_fp32 = 1; t.e.ld = g->Cin; t.e.ncols = g->Cin;
    return run_conv_tma(dy, g->N, Ho, Wo, g->Cout_ld, w_dgrad, g->Cin, round_up(g->kh * g->kw * g->Cout_ld, 64), t, 0, 0, st, true);
  }
  TmaConvParams t;
  memset(&t, 0, sizeof(t));
  t.Ho = g->H; t.Wo = g->W; t.M = (long long)g->N * g->H * g->W; t.stride = 1; t.lower = g->pad - (g->kh - 1); t.lower_w = g->pad - (g->kw - 1);
  t.kh = g->kh; t.kw = g->kw; t.flip = 1; t.cin_blocks = g->Cout_ld / 32; t.num_kb = g->kh * g->kw * t.cin_blocks;
  t.b_sub = 1; t.b_kw = g->kw;
  t.om.linear = 1;
  t.e.scale = scale; t.e.out = dx; t.e.act = 0; t.e.out_fp32 = 1; t.e.ld = g->Cin; t.e.ncols = g->Cin;
  return run_conv_tma(dy, g->N, Ho, Wo, g->Cout_ld, w_dgrad, g->Cin, round_up(g->kh * g->kw * g->Cout_ld, 64), t, -g->pad, -g->pad, st, true);
}

}  // namespace dei2i

using namespace dei2i;

// 1: the TMA-im2col tcgen05 kernel runs this forward geometry (the dispatch order of dei2i_conv2d_fwd)
static int fwd_path(const dei2i_conv_geom* g, bool plain_epilogue) {
  Gather gg;
  memset(&gg, 0, sizeof(gg));
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  gg.N = g->N; gg.Hl = g->H; gg.Wl = g->W; gg.Hs = g->Hs; gg.Ws = g->Ws; gg.C = g->Cin; gg.Ho = Ho; gg.Wo = Wo; gg.kh = g->kh;
  gg.kw = g->kw; gg.stride = g->stride; gg.pad = g->pad; gg.pad_mode = g->pad_mode; gg.mode = 0; gg.cin8 = g->Cin / 8;
  gg.ntaps = g->kh * g->kw; gg.M = (long long)g->N * Ho * Wo;
  if (plain_epilogue && conv_narrow_fwd_applicable(gg, g->Cout_ld)) return 0;
  if (conv_heads_fwd_applicable(gg, g->Cout_ld)) return 0;
  return tma_applicable(g, g->Cin, gg.M) ? 1 : 0;
}

extern "C" int dei2i_conv2d_fwd_path(const dei2i_conv_geom* g) {
  if (g == nullptr || check_geom(g)) return 0;
  return fwd_path(g, false);
}

extern "C" int dei2i_conv2d_fwd(const void* x, const void* w_fwd, void* y, const dei2i_conv_geom* g,
                                const dei2i_conv_epilogue* e, dei2i_stream stream) {
  trace_begin();
  if (int rc = check_geom(g)) return rc;
  DEI2I_REQUIRE(x && w_fwd && y && e, "conv2d_fwd: null pointer");
  DEI2I_REQUIRE(e->noise == nullptr || e->noise_weight != nullptr, "conv2d_fwd: noise without noise_weight");
  if (e->stats_partial != nullptr || e->out_pad > 0)
    DEI2I_REQUIRE(fwd_path(g, false) == 1 && !e->out_fp32 && e->out_pad >= 0,
                  "conv2d_fwd: stats_partial / out_pad need the TMA kernel (dei2i_conv2d_fwd_path) and a bf16 output");
  ConvKParams p;
  memset(&p, 0, sizeof(p));
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  p.g.src = (const __nv_bfloat16*)x;
  p.g.N = g->N; p.g.Hl = g->H; p.g.Wl = g->W; p.g.Hs = g->Hs; p.g.Ws = g->Ws; p.g.C = g->Cin;
  p.g.Ho = Ho; p.g.Wo = Wo; p.g.kh = g->kh; p.g.kw = g->kw; p.g.stride = g->stride; p.g.pad = g->pad;
  p.g.pad_mode = g->pad_mode; p.g.mode = 0; p.g.off = 0; p.g.cin8 = g->Cin / 8; p.g.ntaps = g->kh * g->kw;
  p.g.M = (long long)g->N * Ho * Wo;
  p.e.scale = e->scale; p.e.bias = e->bias; p.e.noise = e->noise; p.e.noise_weight = e->noise_weight;
  p.e.residual = (const __nv_bfloat16*)e->residual; p.e.out = y; p.e.act = e->act; p.e.out_fp32 = e->out_fp32;
  p.e.ld = g->Cout_ld; p.e.ncols = g->Cout;
  if (e->noise == nullptr && e->residual == nullptr && e->stats_partial == nullptr && e->out_pad == 0 &&
      conv_narrow_fwd_applicable(p.g, g->Cout_ld))
    return run_conv_narrow_fwd(p.g, p.e, w_fwd, g->Cout, round_up(g->kh * g->kw * g->Cin, 64), (cudaStream_t)stream);
  if (conv_heads_fwd_applicable(p.g, g->Cout_ld))
    return run_conv_heads_fwd(p.g, p.e, w_fwd, g->Cout, round_up(g->kh * g->kw * g->Cin, 64), (cudaStream_t)stream);
  if (tma_applicable(g, g->Cin, (long long)g->N * Ho * Wo)) {
    TmaConvParams t;
    memset(&t, 0, sizeof(t));
    t.Ho = Ho; t.Wo = Wo; t.M = (long long)g->N * Ho * Wo; t.stride = g->stride; t.lower = -g->pad; t.lower_w = -g->pad;
    t.kh = g->kh; t.kw = g->kw; t.flip = 0; t.cin_blocks = g->Cin / 64; t.num_kb = g->kh * g->kw * t.cin_blocks;
    t.b_sub = 1; t.b_kw = g->kw;
    t.om.linear = 1;
    if (e->out_pad > 0) {  // interior of the consumer-padded output; the frame is mirrored below
      t.om.linear = 0; t.om.Ho = Ho; t.om.Wo = Wo; t.om.OH = Ho + 2 * e->out_pad; t.om.OW = Wo + 2 * e->out_pad;
      t.om.osy = 1; t.om.osx = 1; t.om.oy0 = e->out_pad; t.om.ox0 = e->out_pad;
    }
    t.e.scale = e->scale; t.e.bias = e->bias; t.e.noise = e->noise; t.e.noise_weight = e->noise_weight;
    t.e.residual = (const __nv_bfloat16*)e->residual; t.e.out = y; t.e.act = e->act; t.e.out_fp32 = e->out_fp32;
    t.e.ld = g->Cout_ld; t.e.ncols = g->Cout; t.e.stats = e->stats_partial;
    if (int rc = run_conv_tma(x, g->N, g->H, g->W, g->Cin, w_fwd, g->Cout, g->kh * g->kw * g->Cin, t, g->pad - (g->kh - 1),
                              g->pad - (g->kw - 1), (cudaStream_t)stream))
      return rc;
    if (e->out_pad > 0) return dei2i_reflect_border(y, g->N, Ho, Wo, g->Cout_ld, e->out_pad, stream);
    return 0;
  }
  p.g.src = (const __nv_bfloat16*)x;
  p.g.N = g->N; p.g.Hl = g->H; p.g.Wl = g->W; p.g.Hs = g->Hs; p.g.Ws = g->Ws; p.g.C = g->Cin;
  p.g.Ho = Ho; p.g.Wo = Wo; p.g.kh = g->kh; p.g.kw = g->kw; p.g.stride = g->stride; p.g.pad = g->pad;
  p.g.pad_mode = g->pad_mode; p.g.mode = 0; p.g.off = 0; p.g.cin8 = g->Cin / 8; p.g.ntaps = g->kh * g->kw;
  p.g.M = (long long)g->N * Ho * Wo;
  const int K = round_up(g->kh * g->kw * g->Cin, 64);
  p.num_kb = K / 64;
  p.e.scale = e->scale; p.e.bias = e->bias; p.e.noise = e->noise; p.e.noise_weight = e->noise_weight;
  p.e.residual = (const __nv_bfloat16*)e->residual; p.e.out = y; p.e.act = e->act; p.e.out_fp32 = e->out_fp32;
  p.e.ld = g->Cout_ld; p.e.ncols = g->Cout;
  if (g_tune_conv_path != 1 && conv_c8_applicable(p.g, g->Cout_ld)) return run_conv_c8(p.g, p.e, w_fwd, g->Cout, K, (cudaStream_t)stream);
  return run_igemm(p, w_fwd, g->Cout, K, (cudaStream_t)stream);
}

extern "C" int dei2i_conv2d_dgrad(const void* dy, const void* w_dgrad, void* dx, const dei2i_conv_geom* g, const float* scale,
                                  int out_fp32, dei2i_stream stream) {
  trace_begin();
  if (int rc = check_geom(g)) return rc;
  DEI2I_REQUIRE(dy && w_dgrad && dx, "conv2d_dgrad: null pointer");
  DEI2I_REQUIRE(g->Hs == g->H && g->Ws == g->W, "conv2d_dgrad: resized inputs are not supported");
  ConvKParams p;
  memset(&p, 0, sizeof(p));
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  if (g->stride > 1 && g->pad == 0 && g->kh == g->kw && g->kh % g->stride == 0 && (g->H - g->kh) % g->stride == 0 &&
      (g->W - g->kw) % g->stride == 0 && tma_applicable(g, g->Cout_ld, 0)) {
    // fractionally-strided ("transposed") convolution as stride^2 dense stride-1 sub-convolutions over dY:
    //   dX[s*a + py, s*b + px] = sum_{t,u} dY[a - t, b - u] * W[py + s*t, px + s*u]      (no multiply-by-zero work)
    // All stride^2 parities in ONE launch (TmaConvParams::npar): the deep discriminator layers have 7 .. 16 work tiles per parity.
    const int sd = g->stride, kt = g->kh / sd;
    TmaConvParams t;
    memset(&t, 0, sizeof(t));
    t.Ho = g->H / sd; t.Wo = g->W / sd; t.M = (long long)g->N * t.Ho * t.Wo; t.stride = 1; t.lower = -(kt - 1); t.lower_w = -(kt - 1);
    t.kh = kt; t.kw = kt; t.flip = 1; t.cin_blocks = g->Cout_ld / 64; t.num_kb = kt * kt * t.cin_blocks;
    t.b_sub = sd; t.b_py = 0; t.b_px = 0; t.b_kw = g->kw; t.npar = sd * sd;
    t.om.linear = 0; t.om.Ho = t.Ho; t.om.Wo = t.Wo; t.om.OH = g->H; t.om.OW = g->W; t.om.osy = sd; t.om.osx = sd;
    t.om.oy0 = 0; t.om.ox0 = 0;
    t.e.scale = scale; t.e.out = dx; t.e.act = 0; t.e.out_fp32 = out_fp32; t.e.ld = g->Cin; t.e.ncols = g->Cin;
    return run_conv_tma(dy, g->N, Ho, Wo, g->Cout_ld, w_dgrad, g->Cin, g->kh * g->kw * g->Cout_ld, t, 0, 0, (cudaStream_t)stream);
  }
  const bool full = g->pad_mode == DEI2I_PAD_REFLECT;
  const int Hd = full ? g->H + 2 * g->pad : g->H, Wd = full ? g->W + 2 * g->pad : g->W;
  if (g->stride == 1 && tma_applicable(g, g->Cout_ld, (long long)g->N * Hd * Wd)) {
    // dX = conv(dY, W) with filter offsets flipped: traversal over dY [N, Ho, Wo, Cout_ld], one position per dX pixel
    TmaConvParams t;
    memset(&t, 0, sizeof(t));
    t.Ho = Hd; t.Wo = Wd; t.M = (long long)g->N * Hd * Wd; t.stride = 1; t.lower = g->pad - (g->kh - 1); t.lower_w = g->pad - (g->kw - 1);
    t.kh = g->kh; t.kw = g->kw; t.flip = 1; t.cin_blocks = g->Cout_ld / 64; t.num_kb = g->kh * g->kw * t.cin_blocks;
    t.b_sub = 1; t.b_kw = g->kw;
    t.om.linear = 1;
    t.e.scale = scale; t.e.out = dx; t.e.act = 0; t.e.out_fp32 = out_fp32; t.e.ld = g->Cin; t.e.ncols = g->Cin;
    return run_conv_tma(dy, g->N, Ho, Wo, g->Cout_ld, w_dgrad, g->Cin, g->kh * g->kw * g->Cout_ld, t, -g->pad, -g->pad,
                        (cudaStream_t)stream);
  }
  p.g.src = (const __nv_bfloat16*)dy;
  p.g.N = g->N; p.g.Hl = Ho; p.g.Wl = Wo; p.g.Hs = Ho; p.g.Ws = Wo; p.g.C = g->Cout_ld;
  p.g.Ho = Hd; p.g.Wo = Wd; p.g.kh = g->kh; p.g.kw = g->kw; p.g.stride = g->stride; p.g.pad = g->pad;
  p.g.pad_mode = g->pad_mode; p.g.mode = 1; p.g.off = full ? 0 : g->pad; p.g.cin8 = g->Cout_ld / 8;
  p.g.ntaps = g->kh * g->kw;
  p.g.M = (long long)g->N * Hd * Wd;
  const int K = round_up(g->kh * g->kw * g->Cout_ld, 64);
  p.num_kb = K / 64;
  p.e.scale = scale; p.e.out = dx; p.e.act = 0; p.e.out_fp32 = out_fp32; p.e.ld = g->Cin; p.e.ncols = g->Cin;
  if (g_tune_conv_path != 1 && conv_c8_applicable(p.g, g->Cin)) return run_conv_c8(p.g, p.e, w_dgrad, g->Cin, K, (cudaStream_t)stream);
  return run_igemm(p, w_dgrad, g->Cin, K, (cudaStream_t)stream);
}

static bool wgrad_tma_applicable(const dei2i_conv_geom* g, long long M) {
  if (g_tune_conv_path == 1) return false;
  if (g->pad_mode == DEI2I_PAD_REFLECT && g->pad > 0) return false;
  if (g->Hs != g->H || g->Ws != g->W) return false;
  if (g->kh > 16 || g->kw > 16 || g->pad > 64 || g->stride > 8) return false;
  if (g->kh != g->kw && g->pad != 0) return false;
  // M need not be a multiple of the 64-pixel block: rows of dY past M and im2col pixels past the last image are zero-filled
  // by TMA (the SPADE class maps are N x 5 x 5 = 800 pixels)
  (void)M;
  return g->Cin % 64 == 0 && g->Cout_ld >= 64;
}
static void fill_wgrad_tma(const dei2i_conv_geom* g, TmaWgradParams& t) {
  memset(&t, 0, sizeof(t));
  t.Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad);
  t.Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  t.M = (long long)g->N * t.Ho * t.Wo;
  t.stride = g->stride; t.lower = -g->pad; t.lower_w = -g->pad; t.kw = g->kw; t.cin_blocks = g->Cin / 64;
  t.num_boxes = g->kh * g->kw * t.cin_blocks;
  t.cout = g->Cout; t.Kf = g->kh * g->kw * g->Cin;
  wgrad_tma_plan(t);
}

static void wgrad_plan(const dei2i_conv_geom* g, int& bn, int& splits, int& pb_per_split, int& num_pb, int& Kf) {
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  const long long M = (long long)g->N * Ho * Wo;
  Kf = round_up(g->kh * g->kw * g->Cin, 64);
  bn = 256;
  while (bn > 64 && (Kf % bn) != 0) bn /= 2;
  if (g_tune_wgrad_bn > 0 && Kf % g_tune_wgrad_bn == 0) bn = g_tune_wgrad_bn;
  num_pb = (int)((M + 63) / 64);
  const int tiles = (Kf / bn) * ((g->Cout + 127) / 128);
  const int target = 148 * 2 * (g_tune_wgrad_waves > 0 ? g_tune_wgrad_waves : 1);
  splits = (target + tiles - 1) / tiles;
  if (splits > num_pb) splits = num_pb;
  if (splits < 1) splits = 1;
  // keep at least 8 pixel blocks per split so the prologue/epilogue amortise
  while (splits > 1 && num_pb / splits < 8) --splits;
  pb_per_split = (num_pb + splits - 1) / splits;
  splits = (num_pb + pb_per_split - 1) / pb_per_split;
}

// Operand-swapped wgrad for few output channels on many input channels (the 128 -> 64 decoder convolution at 256^2):
//   P'[ci, (tap', co)] = sum over INPUT pixels m' of  x[m', ci] * dY[m' - tap, co]        tap' = taps - 1 - tap
// x is read ONCE per pixel block as the plain [pixels, Cin] matrix (MMA M = Cin: no zero rows), the small dY tile is what gets
// re-read per tap (im2col over dY exactly as in dgrad).  Per 64 pixels: 3 x 16 KB of x + 9 x 8 KB of dY = 120 KB through
// L2 -> shared memory and three 128 x 192 accumulators, instead of 5 x 8 KB of dY + 18 x 8 KB of im2col'ed x = 184 KB and five
// 128 x 256 accumulators of which half the rows are padding - the kernel is bound by exactly that fill (9 TB/s measured).
static bool wgrad_swap_applicable(const dei2i_conv_geom* g) {
  if (g_tune_wgrad_swap == 0) return false;
  return g->stride == 1 && g->pad == 0 && g->Cout_ld == 64 && g->Cin % 128 == 0 && g->Cin <= 256 && g->kh * g->kw <= 16 &&
         g->Hs == g->H && g->Ws == g->W;
}
static void fill_wgrad_swap(const dei2i_conv_geom* g, TmaWgradParams& t) {
  memset(&t, 0, sizeof(t));
  t.Ho = g->H; t.Wo = g->W;  // traversal over input pixels
  t.M = (long long)g->N * g->H * g->W;
  t.stride = 1; t.lower = -(g->kh - 1); t.lower_w = -(g->kw - 1); t.kw = g->kw; t.cin_blocks = 1;
  t.num_boxes = g->kh * g->kw;
  t.cout = g->Cin; t.Kf = g->kh * g->kw * 64;
  wgrad_tma_plan(t);
}
// sum of the swapped partials over the splits (fixed order) into the ordinary layout P[co][tap * Cin + ci] with one split
__global__ void __launch_bounds__(256)
wgrad_swap_fold_kernel(const float* __restrict__ part, int splits, int Cin, int Cout, int taps, float* __restrict__ out) {
  const int Kfp = taps * 64;
  const int total = Cin * Kfp;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int ci = i / Kfp, k = i - ci * Kfp;
    const int tp = k >> 6, co = k & 63;
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) acc += part[(size_t)s * total + i];
    if (co < Cout) out[(size_t)co * taps * Cin + (taps - 1 - tp) * Cin + ci] = acc;
  }
}

extern "C" size_t dei2i_conv2d_wgrad_workspace(const dei2i_conv_geom* g, int Cin_real) {
  (void)Cin_real;
  if (g == nullptr || check_geom(g)) return 0;
  int bn, splits, pbs, npb, Kf;
  wgrad_plan(g, bn, splits, pbs, npb, Kf);
  {
    const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
    Gather gg;
    memset(&gg, 0, sizeof(gg));
    gg.C = g->Cin; gg.ntaps = g->kh * g->kw; gg.M = (long long)g->N * Ho * Wo;
    int s2 = 0;
    if (wgrad_c8_applicable(gg, g->Cout_ld)) s2 = wgrad_c8_splits(gg, g->Cout_ld);
    else if (g->Cout_ld < 64) s2 = wgrad_narrow_splits(g->kh * g->kw * g->Cin, g->Cout_ld, gg.M);
    if (g->Cout_ld == 8 && s2 < wgrad_heads_splits()) s2 = wgrad_heads_splits();
    if (s2 > splits) splits = s2;
  }
  {
    const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
    if (wgrad_tma_applicable(g, (long long)g->N * Ho * Wo)) {
      TmaWgradParams t;
      fill_wgrad_tma(g, t);
      if (t.splits > splits) splits = t.splits;
    }
  }
  if (wgrad_tma_applicable(g, 0)) {
    TmaWgradParams t;
    fill_wgrad_tma(g, t);
    if (wgrad_ring_applicable(t, g->kh, g->Cin, g->W, g->Cout_ld) && wgrad_ring_splits(t) > splits) splits = wgrad_ring_splits(t);
  }
  size_t bytes = (size_t)splits * g->Cout * Kf * sizeof(float) + 1024 * sizeof(float) + 256;
  if (wgrad_tma_applicable(g, 0) && wgrad_swap_applicable(g)) {
    TmaWgradParams t;
    fill_wgrad_swap(g, t);
    const size_t need = ((size_t)t.splits * g->Cin * t.Kf + (size_t)g->Cout * Kf + 1024) * sizeof(float) + 256;
    if (need > bytes) bytes = need;
  }
  return bytes;
}

template <int BN, int ST>
static int launch_wgrad(const CUtensorMap& tm, const WgradKParams& p, dim3 grid, cudaStream_t st) {
  using S = WgradSmem<BN, ST>;
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<BN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_done = true;
  }
  conv_wgrad_kernel<BN, ST><<<grid, kThreads, S::kTotal, st>>>(tm, p);
  count_launch();
  trace_kernel("conv_wgrad<%d,%d>", BN, ST);
  return check_cuda(cudaGetLastError(), "conv_wgrad_kernel launch");
}

extern "C" int dei2i_conv2d_wgrad(const void* x, const void* dy, float* dw, int Cin_real, float beta,
                                  const dei2i_conv_geom* g, const float* sn_w_orig, const float* sn_u, const float* sn_v,
                                  const float* sn_inv_sigma, void* workspace, size_t workspace_bytes, dei2i_stream stream) {
  trace_begin();
  if (int rc = check_geom(g)) return rc;
  DEI2I_REQUIRE(x && dy && dw && workspace, "conv2d_wgrad: null pointer");
  DEI2I_REQUIRE(Cin_real > 0 && Cin_real <= g->Cin, "conv2d_wgrad: bad Cin_real");
  DEI2I_REQUIRE(workspace_bytes >= dei2i_conv2d_wgrad_workspace(g, Cin_real), "conv2d_wgrad: workspace too small");
  DEI2I_REQUIRE((sn_u == nullptr) == (sn_v == nullptr) && (sn_u == nullptr) == (sn_inv_sigma == nullptr) &&
                    (sn_u == nullptr) == (sn_w_orig == nullptr),
                "conv2d_wgrad: spectral-norm pointers must be all set or all NULL");
  cudaStream_t st = (cudaStream_t)stream;
  const int Ho = conv_out_dim(g->H, g->kh, g->stride, g->pad), Wo = conv_out_dim(g->W, g->kw, g->stride, g->pad);
  int bn, splits, pbs, npb, Kf;
  wgrad_plan(g, bn, splits, pbs, npb, Kf);
  WgradKParams p;
  memset(&p, 0, sizeof(p));
  p.g.src = (const __nv_bfloat16*)x;
  p.g.N = g->N; p.g.Hl = g->H; p.g.Wl = g->W; p.g.Hs = g->Hs; p.g.Ws = g->Ws; p.g.C = g->Cin;
  p.g.Ho = Ho; p.g.Wo = Wo; p.g.kh = g->kh; p.g.kw = g->kw; p.g.stride = g->stride; p.g.pad = g->pad;
  p.g.pad_mode = g->pad_mode; p.g.mode = 0; p.g.off = 0; p.g.cin8 = g->Cin / 8; p.g.ntaps = g->kh * g->kw;
  p.g.M = (long long)g->N * Ho * Wo;
  float* partial = (float*)workspace;
  const int taps = g->kh * g->kw;
  if (wgrad_tma_applicable(g, p.g.M)) {
    TmaWgradParams t;
    fill_wgrad_tma(g, t);
    if (wgrad_ring_applicable(t, g->kh, g->Cin, g->W, g->Cout_ld)) {
      t.partial = partial;
      if (int rc = run_wgrad_ring(x, g->N, g->H, g->W, g->Cin, dy, g->Cout_ld, t, g->kh, g->pad - (g->kh - 1), g->pad - (g->kw - 1), st)) return rc;
      return wgrad_second_pass(partial, t.splits, g, Cin_real, Kf, dw, beta, sn_w_orig, sn_u, sn_v, sn_inv_sigma, st);
    }
  }
  if (wgrad_tma_applicable(g, p.g.M) && wgrad_swap_applicable(g)) {
    TmaWgradParams t;
    fill_wgrad_swap(g, t);
    t.partial = partial;
    if (int rc = run_wgrad_tma(dy, g->N, Ho, Wo, g->Cout_ld, x, g->Cin, t, 0, 0, st)) return rc;
    trace_kernel("+swap");
    float* folded = partial + (size_t)t.splits * g->Cin * t.Kf;
    wgrad_swap_fold_kernel<<<(g->Cin * t.Kf + 255) / 256, 256, 0, st>>>(partial, t.splits, g->Cin, g->Cout, taps, folded);
    count_launch();
    DEI2I_CUDA(cudaGetLastError());
    return wgrad_second_pass(folded, 1, g, Cin_real, Kf, dw, beta, sn_w_orig, sn_u, sn_v, sn_inv_sigma, st);
  } else if (wgrad_tma_applicable(g, p.g.M)) {
    TmaWgradParams t;
    fill_wgrad_tma(g, t);
    t.partial = partial;
    splits = t.splits;
    if (int rc = run_wgrad_tma(x, g->N, g->H, g->W, g->Cin, dy, g->Cout_ld, t, g->pad - (g->kh - 1), g->pad - (g->kw - 1), st)) return rc;
  } else if (g_tune_conv_path != 1 && wgrad_c8_applicable(p.g, g->Cout_ld)) {
    splits = wgrad_c8_splits(p.g, g->Cout_ld);
    if (int rc = run_wgrad_c8(p.g, dy, g->Cout, g->Cout_ld, Kf, partial, st)) return rc;
  } else if (g_tune_conv_path != 1 && wgrad_heads_applicable(p.g, g->Cout, g->Cout_ld)) {
    splits = wgrad_heads_splits();
    if (int rc = run_wgrad_heads(p.g, dy, g->Cout, g->Cout_ld, Kf, partial, st)) return rc;
  } else if (g->Cout_ld < 64) {
    splits = wgrad_narrow_splits(taps * g->Cin, g->Cout_ld, p.g.M);
    if (int rc = run_wgrad_narrow(p.g, dy, g->Cout, g->Cout_ld, Kf, partial, splits, st)) return rc;
  } else {
    p.partial = partial; p.cout_rows = g->Cout; p.Kf = Kf; p.pb_per_split = pbs; p.num_pb = npb;
    CUtensorMap tm;
    int rc = encode_tmap_2d_bf16(&tm, dy, (uint64_t)g->Cout_ld, (uint64_t)p.g.M, (uint64_t)g->Cout_ld * 2, 64, 64, 1);
    if (rc) return rc;
    dim3 grid(Kf / bn, (g->Cout + 127) / 128, splits);
    switch (bn) {
      case 256: rc = launch_wgrad<256, 2>(tm, p, grid, st); break;
      case 128: rc = launch_wgrad<128, 3>(tm, p, grid, st); break;
      default: rc = launch_wgrad<64, 4>(tm, p, grid, st); break;
    }
    if (rc) return rc;
  }
  return wgrad_second_pass(partial, splits, g, Cin_real, Kf, dw, beta, sn_w_orig, sn_u, sn_v, sn_inv_sigma, st);
}
