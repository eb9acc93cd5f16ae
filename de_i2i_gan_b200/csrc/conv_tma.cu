// TMA-fed persistent implicit-GEMM convolution for sm_100a (forward, and dgrad expressed as a convolution over dY).
//
//   D[m, n] = sum_k A[m, k] * B[n, k],   m = output pixel (img, p, q), n = output channel, k = (tap, input channel)
//
//   A : the im2col matrix is never materialised: one `cp.async.bulk.tensor.4d...im2col` per (filter tap, 64-channel block)
//       lands a 128-pixel x 64-channel K-major tile, 128B-swizzled, straight into the UMMA shared-memory layout; zero
//       padding comes from TMA's out-of-bounds fill, the stride from the descriptor's traversal stride (reflect padding is
//       applied by the producer of the activation, so here every convolution is 'zeros' / 'valid');
//   B : packed weights [n, k], tiled TMA (SWIZZLE_128B);
//   one CTA per SM, persistent over a static tile schedule; warp 4 = TMA producer, warp 5 = tcgen05.mma issuer,
//   warps 0-3 = epilogue.  STAGES-deep smem ring (full/empty mbarriers) and TWO accumulators in TMEM (2 x BLOCK_N
//   columns, tfull/tempty mbarriers) so that the epilogue of tile i runs under the main loop of tile i+1.
#include <string.h>

#include "common.cuh"
#include "conv_common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {

void count_launch(int n = 1);
void trace_kernel(const char* fmt, ...);

constexpr int kTmaThreads = 320;      // warps 0-7: two epilogue groups of 4, warp 8: TMA producer, warp 9: MMA issuer
constexpr int kWgTmaThreads = 192;    // wgrad kernel: warps 0-3 epilogue, warp 4 TMA, warp 5 MMA
constexpr int kTmaATile = 128 * 128;  // 128 pixels x 64 bf16

template <int BLOCK_N, int STAGES, int MS>
struct TmaSmem {
  static constexpr int kBTile = BLOCK_N * 128;
  static constexpr int kStage = MS * kTmaATile + kBTile;  // MS 128-pixel A tiles share one B tile
  static constexpr int kStageOut = STAGES * kStage;     // 8 epilogue warps x 2 KB output staging (32 rows x 64 B)
  static constexpr int kBarOff = kStageOut + 8 * 2048;
  static constexpr int kTotal = kBarOff + (2 * STAGES + 4) * 8 + 16 + 1024 /*alignment slack*/;
};

DEVINL void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// Chan et al. merge of two (count, mean, M2) summaries; fixed evaluation order
DEVINL void chan_merge(float& n, float& mean, float& m2, float nb, float mb, float m2b) {
  if (nb <= 0.f) return;
  const float nt = n + nb;
  const float d = mb - mean;
  mean += d * (nb / nt);
  m2 += m2b + d * d * (n * nb / nt);
  n = nt;
}

// TF32 = true: FP32 activations / weights in shared memory (32 channels per 128-byte row), kind::tf32 MMA (K = 8 per instruction),
// FP32 output - the tensor-core variant of the FP32-parity path.
// PAIR = true: launched as 2-CTA clusters; the pair runs ONE M = 256 (cta_group::2) MMA per k step - CTA r owns pixel tile 2 v + r of
// work item v and loads its own A tile plus HALF of the weight tile (rows r * BLOCK_N / 2 ..), so every MMA reads half the B bytes per
// SM from shared memory (the single-CTA SS-mode MMA is bound by its operand reads: 4 KB of A + 8 KB of B per 128-cycle N = 256 MMA)
// and the L2 -> shared-memory fill drops from 48 KB to 32 KB per k block.  Only the even CTA issues MMAs; full barriers live in the
// even CTA (both producers credit their bytes there), empty / accumulator-ready barriers are signalled in both by multicast commits.
template <int BLOCK_N, int STAGES, int MS, bool TF32 = false, bool PAIR = false>
__global__ void __launch_bounds__(kTmaThreads, 1)
conv_tma_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const TmaConvParams p) {
  constexpr int KC = TF32 ? 32 : 64;  // channels per 128-byte shared-memory row
  using S = TmaSmem<PAIR ? BLOCK_N / 2 : BLOCK_N, STAGES, MS>;
  static_assert(!PAIR || !TF32, "pair kernel: BF16 only");
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;
  const int tile0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;       // first work item of this CTA (pair)
  const int tstep = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;  // [2] accumulator ready
  uint64_t* tempty_bar = tfull_bar + 2;      // [2] accumulator drained
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 2 * MS * BLOCK_N < 32 ? 32 : 2 * MS * BLOCK_N;
  const int npar = p.npar > 1 ? p.npar : 1;
  const int total_tiles = (PAIR ? (p.m_tiles + 1) / 2 : p.m_tiles) * p.n_tiles * npar;
  __shared__ __align__(16) float bias_s[1024];
  __shared__ float2 wstat[2][2][4][32];  // [epilogue group][parity][row quarter][column]: (mean, M2) of 32 rows
  const bool bias_in_smem = p.e.bias != nullptr && p.n_tiles * BLOCK_N <= 1024;
  if (bias_in_smem)
    for (int i = threadIdx.x; i < p.n_tiles * BLOCK_N; i += kTmaThreads) bias_s[i] = i < p.e.ncols ? __ldg(p.e.bias + i) : 0.f;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull_bar[b], 1);
      mbar_init(&tempty_bar[b], PAIR ? 256 : 128);  // pair: the epilogue threads of both CTAs release the leader's accumulator
    }
    fence_barrier_init();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 9) {
    if (PAIR) tmem_alloc2(tmem_ptr_smem, kTmemCols);
    else tmem_alloc(tmem_ptr_smem, kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // both CTAs' barriers are initialised before either touches the other's
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 8) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      const int HoWo = p.Ho * p.Wo;
      uint32_t it = 0;
      for (int tile = tile0; tile < total_tiles; tile += tstep) {
        const int par = tile % npar, t2 = tile / npar;
        const int b_py = npar > 1 ? par / p.b_sub : p.b_py, b_px = npar > 1 ? par % p.b_sub : p.b_px;
        const int m_tile = PAIR ? (t2 / p.n_tiles) * 2 + (int)cta_rank : t2 / p.n_tiles;
        const int n_tile = t2 - (t2 / p.n_tiles) * p.n_tiles;
        int img[MS], cw[MS], ch[MS];
#pragma unroll
        for (int ms = 0; ms < MS; ++ms) {
          const long long m0 = ((long long)m_tile * MS + ms) * 128;
          img[ms] = (int)(m0 / HoWo);
          const int rem = (int)(m0 - (long long)img[ms] * HoWo);
          const int pr = rem / p.Wo;
          const int qc = rem - pr * p.Wo;
          cw[ms] = qc * p.stride + p.lower_w;
          ch[ms] = pr * p.stride + p.lower;
        }
        int cb = 0, r = 0, s = 0;
        for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
          const uint32_t stage = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1u;
          mbar_wait(&empty_bar[stage], ph ^ 1u);
          uint8_t* a_dst = smem + stage * S::kStage;
          const int os = p.flip ? p.kw - 1 - s : s;
          const int orr = p.flip ? p.kh - 1 - r : r;
          const int bcol = (((b_py + p.b_sub * r) * p.b_kw + (b_px + p.b_sub * s)) * p.cin_blocks + cb) * KC;
          if (PAIR) {
            if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * S::kStage);  // this CTA's bytes and the peer's
#pragma unroll
            for (int ms = 0; ms < MS; ++ms)
              tma2_load_im2col_4d(a_dst + ms * kTmaATile, &tmap_a, &full_bar[stage], cb * KC, cw[ms], ch[ms], img[ms], (uint16_t)os, (uint16_t)orr);
            tma2_load_2d(a_dst + MS * kTmaATile, &tmap_b, &full_bar[stage], bcol, n_tile * BLOCK_N + (int)cta_rank * (BLOCK_N / 2));
          } else {
            mbar_arrive_expect_tx(&full_bar[stage], S::kStage);
#pragma unroll
            for (int ms = 0; ms < MS; ++ms)
              tma_load_im2col_4d(a_dst + ms * kTmaATile, &tmap_a, &full_bar[stage], cb * KC, cw[ms], ch[ms], img[ms], (uint16_t)os, (uint16_t)orr);
            tma_load_2d(a_dst + MS * kTmaATile, &tmap_b, &full_bar[stage], bcol, n_tile * BLOCK_N);
          }
          if (++cb == p.cin_blocks) {
            cb = 0;
            if (++s == p.kw) { s = 0; ++r; }
          }
        }
      }
    }
  } else if (warp == 9) {
    // =========================== MMA issuer ===========================
    // The whole warp walks the loop, one elected lane issues (see conv_ring_kernel): stage addresses and descriptors stay in uniform
    // registers and the MMAs of a k block go out back to back.  With `if (lane == 0)` around the loop every UTCHMMA was preceded by
    // ~18 dependent instructions (~110 cycles) - hidden under a 128-cycle N = 256 MMA, but THE bound for N = 64 / 128.
    if (PAIR && cta_rank != 0) {
      // the odd CTA of a pair issues nothing: its tile is computed by the leader's M = 256 MMAs
    } else if (p.issue_mode == 0) {  // A/B switch (dei2i_set_tuning("mma_issue", 0)): the per-thread loop this kernel had before
      if (lane == 0) {
        constexpr uint32_t idesc = TF32 ? make_idesc_tf32(128, BLOCK_N, 0, 0) : make_idesc_bf16(PAIR ? 256 : 128, BLOCK_N, 0, 0);
        uint32_t it = 0, tcount = 0;
        for (int tile = tile0; tile < total_tiles; tile += tstep, ++tcount) {
          const uint32_t buf = tcount & 1u;
          mbar_wait(&tempty_bar[buf], ((tcount >> 1) & 1u) ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + buf * (MS * BLOCK_N);
          for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
            const uint32_t stage = it % STAGES;
            mbar_wait(&full_bar[stage], (it / STAGES) & 1u);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + stage * S::kStage);
            const uint64_t b_desc = make_smem_desc(a_addr + MS * kTmaATile, 16, 1024, 2);
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
              const uint64_t a_desc = make_smem_desc(a_addr + ms * kTmaATile, 16, 1024, 2);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (TF32) tc_mma_tf32(d_tmem + ms * BLOCK_N, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
                else if (PAIR) tc_mma2_bf16(d_tmem + ms * BLOCK_N, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
                else tc_mma_bf16(d_tmem + ms * BLOCK_N, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
              }
            }
            if (PAIR) tc_commit2(&empty_bar[stage]);
            else tc_commit(&empty_bar[stage]);
          }
          if (PAIR) tc_commit2(&tfull_bar[buf]);
          else tc_commit(&tfull_bar[buf]);
        }
      }
    } else {
      constexpr uint32_t idesc = TF32 ? make_idesc_tf32(128, BLOCK_N, 0, 0) : make_idesc_bf16(PAIR ? 256 : 128, BLOCK_N, 0, 0);
      const uint64_t desc_hi = make_smem_desc(0, 16, 1024, 2);
      const uint32_t smem_base = smem_u32(smem);
      uint32_t it = 0, tcount = 0;
      for (int tile = tile0; tile < total_tiles; tile += tstep, ++tcount) {
        const uint32_t buf = tcount & 1u;
        const uint32_t bph = (tcount >> 1) & 1u;
        mbar_wait(&tempty_bar[buf], bph ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * (MS * BLOCK_N);
        for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
          const uint32_t stage = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1u;
          mbar_wait(&full_bar[stage], ph);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * S::kStage;
          const uint64_t b_desc = desc_hi | (uint64_t)(((a_addr + MS * kTmaATile) & 0x3FFFFu) >> 4);
          if (elect_one()) {
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
              const uint64_t a_desc = desc_hi | (uint64_t)(((a_addr + ms * kTmaATile) & 0x3FFFFu) >> 4);
              if (TF32) {
                if (kb == 0) tc_mma_tf32(d_tmem + ms * BLOCK_N, a_desc, b_desc, idesc, 0);
                else tc_mma_tf32(d_tmem + ms * BLOCK_N, a_desc, b_desc, idesc, 1);
#pragma unroll
                for (int k = 1; k < 4; ++k) tc_mma_tf32(d_tmem + ms * BLOCK_N, a_desc + 2 * k, b_desc + 2 * k, idesc, 1);
              } else if (PAIR) {
                if (kb == 0) tc_mma2_bf16(d_tmem + ms * BLOCK_N, a_desc, b_desc, idesc, 0);
                else tc_mma2_bf16(d_tmem + ms * BLOCK_N, a_desc, b_desc, idesc, 1);
#pragma unroll
                for (int k = 1; k < 4; ++k) tc_mma2_bf16(d_tmem + ms * BLOCK_N, a_desc + 2 * k, b_desc + 2 * k, idesc, 1);
              } else {
                if (kb == 0) tc_mma_bf16(d_tmem + ms * BLOCK_N, a_desc, b_desc, idesc, 0);
                else tc_mma_bf16(d_tmem + ms * BLOCK_N, a_desc, b_desc, idesc, 1);
#pragma unroll
                for (int k = 1; k < 4; ++k) tc_mma_bf16(d_tmem + ms * BLOCK_N, a_desc + 2 * k, b_desc + 2 * k, idesc, 1);
              }
            }
            if (PAIR) tc_commit2(&empty_bar[stage]);
            else tc_commit(&empty_bar[stage]);
          }
          __syncwarp();
        }
        if (elect_one()) {
          if (PAIR) tc_commit2(&tfull_bar[buf]);
          else tc_commit(&tfull_bar[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    // =========================== epilogue ===========================
    // Two groups of four warps; group g drains accumulator buffer g, i.e. every other tile of this CTA, so each group has two
    // main loops of time per tile and two warps per scheduler hide each other's TMEM / shared-memory latencies (with one group the
    // N = 64 layers were bound by this loop: 32 % tensor-pipe activity, profiles/r02c_ncu_full_fwd64.txt).  Warp (g, q) owns TMEM
    // lanes 32q .. 32q + 31 = rows 32q .. of the 128-row tile.
    const int eg = warp >> 2, qw = warp & 3;
    const Epilogue& e = p.e;
    const OutMap& om = p.om;
    const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
    const float nzw = e.noise != nullptr ? __ldg(e.noise_weight) : 0.f;
    const float slope = e.act == ACT_LRELU ? 0.2f : (e.act == ACT_RELU ? 0.f : 1.f);
    const bool slow_act = e.act == ACT_TANH || e.act == ACT_SIGMOID;
    // 64 bytes of output per row and pass: 32 bf16 columns or 16 fp32 columns.  Each lane finalises its own row, parks it in
    // the warp's 2 KB staging tile (16-byte chunks XOR-swizzled by row pair: conflict-free both ways) and the warp then writes 8
    // rows x 64 B per store instruction - coalesced whatever the row -> address map is.  The residual comes in the same way.
    uint8_t* stage = smem + S::kStageOut + warp * 2048;
    const int CW = e.out_fp32 ? 16 : 32;
    const int esz = e.out_fp32 ? 4 : 2;
    const int sub_r = lane >> 2, sub_j = lane & 3;
    const int my_sw = (lane >> 1) & 3;  // swizzle of this lane's own row
    uint32_t ti = 0, spar = 0;
    for (int tile = tile0 + eg * tstep; tile < total_tiles; tile += 2 * tstep, ++ti) {
      const int par = tile % npar, t2 = tile / npar;
      const int oy0 = om.oy0 + (npar > 1 ? par / p.b_sub : 0), ox0 = om.ox0 + (npar > 1 ? par % p.b_sub : 0);
      const int m_tile = PAIR ? (t2 / p.n_tiles) * 2 + (int)cta_rank : t2 / p.n_tiles;
      const int n0 = (t2 - (t2 / p.n_tiles) * p.n_tiles) * BLOCK_N;
      mbar_wait(&tfull_bar[eg], ti & 1u);
      tc_fence_after();
#pragma unroll 1
      for (int ms = 0; ms < MS; ++ms) {
        const long long m_base = ((long long)m_tile * MS + ms) * 128;
        if (m_base >= p.M) break;
        const uint32_t tcol0 = eg * (MS * BLOCK_N) + ms * BLOCK_N;  // first TMEM column of this sub-tile's accumulator
        const long long m = m_base + qw * 32 + lane;
        const bool row_ok = m < p.M;
        long long orow = m;  // output row index (in units of ld elements)
        if (!om.linear && row_ok) {
          const int HoWo = om.Ho * om.Wo;
          const int img = (int)(m / HoWo);
          const int rem = (int)(m - (long long)img * HoWo);
          const int pr = rem / om.Wo;
          const int qc = rem - pr * om.Wo;
          orow = ((long long)img * om.OH + (pr * om.osy + oy0)) * om.OW + (qc * om.osx + ox0);
        }
        float nz = 0.f;
        if (e.noise != nullptr && row_ok) nz = __ldg(e.noise + m) * nzw;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += CW) {
          const int nbase = n0 + c;
          if (nbase >= e.ld) break;
          float f[32];
          {
            const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16) + tcol0 + (uint32_t)c;
            if (CW == 32) {
              uint32_t v[32];
              tmem_ld_32x32b_x32(taddr, v);
              tc_wait_ld();
#pragma unroll
              for (int t = 0; t < 32; ++t) f[t] = __uint_as_float(v[t]) * osc;
            } else {
              uint32_t v[16];
              tmem_ld_32x32b_x16(taddr, v);
              tc_wait_ld();
#pragma unroll
              for (int t = 0; t < 16; ++t) f[t] = __uint_as_float(v[t]) * osc;
#pragma unroll
              for (int t = 16; t < 32; ++t) f[t] = 0.f;
            }
          }
          if (bias_in_smem) {  // zero-padded copy of the bias vector: no bounds logic, broadcast LDS.128
#pragma unroll
            for (int t4 = 0; t4 < 32; t4 += 4) {
              if (t4 < CW) {
                const float4 bv = *reinterpret_cast<const float4*>(bias_s + nbase + t4);
                f[t4] += bv.x; f[t4 + 1] += bv.y; f[t4 + 2] += bv.z; f[t4 + 3] += bv.w;
              }
            }
          } else if (e.bias != nullptr) {
#pragma unroll
            for (int t = 0; t < 32; ++t)
              if (t < CW && nbase + t < e.ncols) f[t] += __ldg(e.bias + nbase + t);
          }
          if (slow_act) {  // tanh / sigmoid: rare on this kernel (the heads go through the compose kernel)
#pragma unroll
            for (int t = 0; t < 32; ++t)
              if (t < CW) f[t] = apply_act(f[t] + nz, e.act);
          } else if (e.noise != nullptr || e.act != ACT_NONE) {  // none / leaky-relu / relu = max(v, slope * v): branch-free
#pragma unroll
            for (int t = 0; t < 32; ++t) {
              const float v = f[t] + nz;
              f[t] = fmaxf(v, v * slope);
            }
          }
          if (e.residual != nullptr && e.res_fp32) {  // fp32 [M, ld] residual with fp32 output (CW == 16): 4 x 16-byte chunks per row
#pragma unroll
            for (int itr = 0; itr < 4; ++itr) {
              const int r = itr * 8 + sub_r;
              const long long mr = m_base + qw * 32 + r;
              float4 rv = make_float4(0.f, 0.f, 0.f, 0.f);
              if (mr < p.M && nbase + sub_j * 4 < e.ld)
                rv = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(e.residual) + mr * e.ld + nbase + sub_j * 4));
              *reinterpret_cast<float4*>(stage + r * 64 + ((sub_j ^ ((r >> 1) & 3)) << 4)) = rv;
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float4 rr = *reinterpret_cast<const float4*>(stage + lane * 64 + ((j ^ my_sw) << 4));
              f[4 * j] += rr.x; f[4 * j + 1] += rr.y; f[4 * j + 2] += rr.z; f[4 * j + 3] += rr.w;
            }
            __syncwarp();
          } else if (e.residual != nullptr) {  // bf16 [M, ld], linear rows; only with bf16 output (CW == 32)
#pragma unroll
            for (int itr = 0; itr < 4; ++itr) {
              const int r = itr * 8 + sub_r;
              const long long mr = m_base + qw * 32 + r;
              uint4 rv = make_uint4(0, 0, 0, 0);
              if (mr < p.M && nbase + sub_j * 8 < e.ld) rv = __ldg(reinterpret_cast<const uint4*>(e.residual + mr * e.ld + nbase + sub_j * 8));
              *reinterpret_cast<uint4*>(stage + r * 64 + ((sub_j ^ ((r >> 1) & 3)) << 4)) = rv;
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float rr[8];
              unpack8_bf16(*reinterpret_cast<const uint4*>(stage + lane * 64 + ((j ^ my_sw) << 4)), rr);
#pragma unroll
              for (int k = 0; k < 8; ++k) f[j * 8 + k] += rr[k];
            }
            __syncwarp();
          }
          if (e.out_fp32) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<float4*>(stage + lane * 64 + ((j ^ my_sw) << 4)) = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(stage + lane * 64 + ((j ^ my_sw) << 4)) =
                  make_uint4(pack_bf16(f[8 * j], f[8 * j + 1]), pack_bf16(f[8 * j + 2], f[8 * j + 3]), pack_bf16(f[8 * j + 4], f[8 * j + 5]),
                             pack_bf16(f[8 * j + 6], f[8 * j + 7]));
          }
          __syncwarp();
          if (e.stats != nullptr) {
            // per-channel (mean, M2) of this 128-row sub-tile over the STORED (BF16-rounded) values - what the consuming
            // BatchNorm / InstanceNorm would otherwise re-read the tensor for.  Lane = column: two exact passes over the warp's
            // 32 rows, then the four row quarters are merged in a fixed order (Chan) -> stats[tile128][2][ld]
            const int cnt_w = (int)max(0ll, min(32ll, p.M - (m_base + qw * 32)));
            const int cchunk = lane >> 3, coff = (lane & 7) * 2;
            float v[32], sum = 0.f;
#pragma unroll
            for (int r = 0; r < 32; ++r) {
              const uint16_t h = *reinterpret_cast<const uint16_t*>(stage + r * 64 + ((cchunk ^ ((r >> 1) & 3)) << 4) + coff);
              v[r] = __uint_as_float((uint32_t)h << 16);
              if (r < cnt_w) sum += v[r];
            }
            const float mean_w = cnt_w > 0 ? sum / (float)cnt_w : 0.f;
            float m2_w = 0.f;
#pragma unroll
            for (int r = 0; r < 32; ++r)
              if (r < cnt_w) m2_w = fmaf(v[r] - mean_w, v[r] - mean_w, m2_w);
            wstat[eg][spar][qw][lane] = make_float2(mean_w, m2_w);
            named_bar_sync(1 + eg, 128);
            if (qw == 0 && nbase + lane < e.ld) {
              float nn = 0.f, mean = 0.f, m2 = 0.f;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float cq = (float)max(0ll, min(32ll, p.M - (m_base + q * 32)));
                const float2 t = wstat[eg][spar][q][lane];
                chan_merge(nn, mean, m2, cq, t.x, t.y);
              }
              float* dst = e.stats + (size_t)(m_base >> 7) * 2 * e.ld + nbase + lane;
              dst[0] = mean;
              dst[e.ld] = m2;
            }
            spar ^= 1u;
          }
          const int col = nbase + sub_j * (16 / esz);  // first column of this lane's 16-byte chunk
#pragma unroll
          for (int itr = 0; itr < 4; ++itr) {
            const int r = itr * 8 + sub_r;
            const long long orow_r = __shfl_sync(0xffffffffu, orow, r);
            const int ok_r = __shfl_sync(0xffffffffu, (int)row_ok, r);
            const uint4 val = *reinterpret_cast<const uint4*>(stage + r * 64 + ((sub_j ^ ((r >> 1) & 3)) << 4));
            if (ok_r && col < e.ld)
              *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(e.out) + (orow_r * e.ld + col) * esz) = val;
          }
          __syncwarp();
        }
      }  // ms
      tc_fence_before();
      if (PAIR) mbar_arrive_leader(&tempty_bar[eg]);
      else mbar_arrive(&tempty_bar[eg]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // neither CTA leaves (or frees TMEM) while the other may still signal it
  if (warp == 9) {
    tc_fence_after();
    if (PAIR) tmem_dealloc2(tmem_base, kTmemCols);
    else tmem_dealloc(tmem_base, kTmemCols);
  }
}

static int g_num_sms = 0;
static int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

// ======================================================================================================================
// Row-ring variant for kh x 1 filters (the row-packed generator stem, generator.py:76-86: 7 x 1 over 64 packed channels, forward and
// dgrad).  The generic kernel fetches a 128-pixel A tile per filter tap: for a kh x 1 filter the tap r tile of output row y IS the
// tap r - 1 tile of output row y + 1, so every input tile travels L2 -> shared memory kh times and the kernel sits at the fill
// bandwidth (168 KB per 128 output pixels, 196 TFLOP/s).  Here a CTA walks DOWN a 128-pixel column block: input-row tiles live in a
// ring of kRingSlots slots, each fetched once and used by kh consecutive output rows; the weights of all taps stay resident.
// Per output row: ONE 16 KB tile in, kh x 4 MMAs (128 x 64 x 16), one accumulator out.  Epilogue: scale only, BF16, linear rows.
constexpr int kRingSlots = 9;
constexpr int kRingKh = 7;  // resident weight blocks (8 KB each)
struct RingSmem {
  static constexpr int kB = kRingKh * 64 * 128;
  static constexpr int kRing = kRingSlots * kTmaATile;
  static constexpr int kStageOut = kB + kRing;
  static constexpr int kBarOff = kStageOut + 8 * 2048;
  static constexpr int kTotal = kBarOff + (2 * kRingSlots + 5) * 8 + 16 + 1024;
};

__global__ void __launch_bounds__(kTmaThreads, 1)
conv_ring_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const TmaConvParams p,
                 long long tiles_per_cta, long long total_tiles) {
  using S = RingSmem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* b_smem = smem;
  uint8_t* ring = smem + S::kB;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + kRingSlots;
  uint64_t* tfull_bar = empty_bar + kRingSlots;  // [2]
  uint64_t* tempty_bar = tfull_bar + 2;          // [2]
  uint64_t* w_bar = tempty_bar + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(w_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 128;  // two 64-column accumulators
  const int kh = p.kh;
  const int CB = p.Wo / 128;
  const long long t0 = (long long)blockIdx.x * tiles_per_cta;
  const long long t1 = t0 + tiles_per_cta < total_tiles ? t0 + tiles_per_cta : total_tiles;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kRingSlots; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull_bar[b], 1);
      mbar_init(&tempty_bar[b], 128);
    }
    mbar_init(w_bar, 1);
    fence_barrier_init();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 9) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 8) {
    // =========================== TMA producer ===========================
    if (lane == 0 && t0 < t1) {
      mbar_arrive_expect_tx(w_bar, (uint32_t)kh * 8192u);
      for (int r = 0; r < kh; ++r) tma_load_2d(b_smem + r * 8192, &tmap_b, w_bar, r * p.b_kw * 64, 0);
      uint32_t L = 0;  // row-tile loads issued so far
      for (long long t = t0; t < t1;) {
        const long long strip = t / p.Ho;
        const int ya = (int)(t - strip * p.Ho);
        const int yb = (int)((long long)p.Ho < ya + (t1 - t) ? (long long)p.Ho : ya + (t1 - t));
        const int img = (int)(strip / CB), cb = (int)(strip - (long long)img * CB);
        for (int j = ya; j < yb + kh - 1; ++j, ++L) {
          const uint32_t slot = L % kRingSlots;
          mbar_wait(&empty_bar[slot], ((L / kRingSlots) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&full_bar[slot], (uint32_t)kTmaATile);
          const int jb = j < p.Ho - 1 ? j : p.Ho - 1;  // base rows are confined to the traversal box; the rest is a filter offset
          tma_load_im2col_4d(ring + slot * kTmaATile, &tmap_a, &full_bar[slot], 0, cb * 128 + p.lower_w, jb + p.lower, img, (uint16_t)0,
                             (uint16_t)(j - jb));
        }
        t += yb - ya;
      }
    }
  } else if (warp == 9) {
    // =========================== MMA issuer ===========================
    // The whole warp walks the loop and one elected lane issues: descriptors and addresses are then warp-uniform values that live
    // in uniform registers.  With `if (lane == 0)` around the loop the compiler computes them per thread and moves them to uniform
    // registers before every UTCHMMA - 18 dependent instructions, ~120 cycles per MMA, which is what bounds 128 x 64 x 16 MMAs
    // (32 cycles of tensor work each; measured: 205 us with the fill solved, 105 us without the MMAs).
    if (t0 < t1) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
      mbar_wait(w_bar, 0);
      tc_fence_after();
      const uint32_t b_addr = smem_u32(b_smem), ring_addr = smem_u32(ring);
      const uint64_t desc_hi = make_smem_desc(0, 16, 1024, 2);  // everything but the 14-bit start address
      uint32_t L0 = 0, tcount = 0;  // first load of the current segment, accumulator counter
      for (long long t = t0; t < t1;) {
        const long long strip = t / p.Ho;
        const int ya = (int)(t - strip * p.Ho);
        const int nrows = (int)((long long)p.Ho - ya < t1 - t ? (long long)p.Ho - ya : t1 - t);
        for (int y = 0; y < nrows; ++y, ++tcount) {
          const uint32_t buf = tcount & 1u;
          mbar_wait(&tempty_bar[buf], ((tcount >> 1) & 1u) ^ 1u);
          for (int r = (y == 0 ? 0 : kh - 1); r < kh; ++r) {  // rows y .. y + kh - 2 were waited for by earlier output rows
            const uint32_t L = L0 + y + r;
            mbar_wait(&full_bar[L % kRingSlots], (L / kRingSlots) & 1u);
          }
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + buf * 64;
          uint32_t slot = (L0 + y) % kRingSlots;
          for (int r = 0; r < kh; ++r) {
            const int wr = p.flip ? kh - 1 - r : r;
            const uint64_t a_desc = desc_hi | (uint64_t)(((ring_addr + slot * kTmaATile) & 0x3FFFFu) >> 4);
            const uint64_t b_desc = desc_hi | (uint64_t)(((b_addr + wr * 8192) & 0x3FFFFu) >> 4);
            if (elect_one()) {
              if (r == 0) tc_mma_bf16(d_tmem, a_desc, b_desc, idesc, 0);
              else tc_mma_bf16(d_tmem, a_desc, b_desc, idesc, 1);
              tc_mma_bf16(d_tmem, a_desc + 2, b_desc + 2, idesc, 1);
              tc_mma_bf16(d_tmem, a_desc + 4, b_desc + 4, idesc, 1);
              tc_mma_bf16(d_tmem, a_desc + 6, b_desc + 6, idesc, 1);
            }
            __syncwarp();
            if (++slot == kRingSlots) slot = 0;
          }
          if (elect_one()) {
            tc_commit(&empty_bar[(L0 + y) % kRingSlots]);  // input row y is not needed by later output rows
            if (y == nrows - 1)
              for (int r = 1; r < kh; ++r) tc_commit(&empty_bar[(L0 + y + r) % kRingSlots]);  // end of the segment: the tail rows too
            tc_commit(&tfull_bar[buf]);
          }
          __syncwarp();
        }
        L0 += nrows + kh - 1;
        t += nrows;
      }
    }
  } else if (t0 < t1) {
    // =========================== epilogue (two groups of four warps, one per accumulator buffer) ===========================
    const int eg = warp >> 2, qw = warp & 3;
    const Epilogue& e = p.e;
    const float osc = e.scale != nullptr ? __ldg(e.scale) : 1.f;
    uint8_t* stage = smem + S::kStageOut + warp * 2048;
    const int sub_r = lane >> 2, sub_j = lane & 3;
    const int my_sw = (lane >> 1) & 3;
    uint32_t tcount = 0;
    for (long long t = t0; t < t1; ++t, ++tcount) {
      if ((int)(tcount & 1u) != eg) continue;
      const long long strip = t / p.Ho;
      const int y = (int)(t - strip * p.Ho);
      const int img = (int)(strip / CB), cb = (int)(strip - (long long)img * CB);
      const long long m_base = ((long long)img * p.Ho + y) * p.Wo + cb * 128 + qw * 32;  // first output row of this warp
      mbar_wait(&tfull_bar[eg], (tcount >> 1) & 1u);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 64; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(qw * 32) << 16) + eg * 64 + (uint32_t)c, v);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float f[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) f[k] = __uint_as_float(v[8 * j + k]) * osc;
          *reinterpret_cast<uint4*>(stage + lane * 64 + ((j ^ my_sw) << 4)) =
              make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
        }
        __syncwarp();
#pragma unroll
        for (int itr = 0; itr < 4; ++itr) {
          const int r = itr * 8 + sub_r;
          const uint4 val = *reinterpret_cast<const uint4*>(stage + r * 64 + ((sub_j ^ ((r >> 1) & 3)) << 4));
          *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(e.out) + ((m_base + r) * e.ld + c + sub_j * 8) * 2) = val;
        }
        __syncwarp();
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[eg]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

extern int g_tune_conv_ring;  // 1 (default): kh x 1 convolutions that qualify run on conv_ring_kernel
static bool conv_ring_applicable(const TmaConvParams& p, int rows, int C, int W, int bn) {
  const Epilogue& e = p.e;
  return g_tune_conv_ring != 0 && bn == 64 && rows <= 64 && C == 64 && p.cin_blocks == 1 && p.kw == 1 && p.kh >= 2 && p.kh <= kRingKh &&
         p.stride == 1 && p.lower_w == 0 && p.Wo == W && p.Wo % 128 == 0 && p.om.linear && p.npar <= 1 && p.b_sub == 1 && e.bias == nullptr &&
         e.noise == nullptr && e.residual == nullptr && e.act == 0 && !e.out_fp32 && e.stats == nullptr && e.ld == 64 &&
         p.M == (p.M / ((long long)p.Ho * p.Wo)) * p.Ho * p.Wo && p.Ho >= p.kh;
}
static int launch_conv_ring(const CUtensorMap& ta, const CUtensorMap& tb, const TmaConvParams& p, cudaStream_t st) {
  static_assert(RingSmem::kTotal <= 227 * 1024 && RingSmem::kTotal > 114 * 1024, "shared memory budget / one CTA per SM");
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(conv_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RingSmem::kTotal));
    attr_done = true;
  }
  const long long total = (p.M / ((long long)p.Ho * p.Wo)) * (p.Wo / 128) * p.Ho;  // (image, column block, output row)
  const int grid = total < num_sms() ? (int)total : num_sms();
  const long long per = (total + grid - 1) / grid;
  conv_ring_kernel<<<grid, kTmaThreads, RingSmem::kTotal, st>>>(ta, tb, p, per, total);
  count_launch();
  trace_kernel("conv_ring<%d>", p.kh);
  return check_cuda(cudaGetLastError(), "conv_ring_kernel launch");
}

template <int BN, int ST, int MS, bool TF32 = false>
static int launch_tma(const CUtensorMap& ta, const CUtensorMap& tb, const TmaConvParams& p, cudaStream_t st) {
  using S = TmaSmem<BN, ST, MS>;
  static_assert(S::kTotal + 4096 + 4096 <= 227 * 1024, "shared memory budget (dynamic + the static bias copy + the statistics scratch)");
  static_assert(S::kTotal > 114 * 1024, "one CTA per SM is required (the CTA owns up to all 512 TMEM columns)");
  static_assert(2 * MS * BN <= 512, "TMEM columns");
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(conv_tma_kernel<BN, ST, MS, TF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_done = true;
  }
  const int total = p.m_tiles * p.n_tiles * (p.npar > 1 ? p.npar : 1);
  const int grid = total < num_sms() ? total : num_sms();
  conv_tma_kernel<BN, ST, MS, TF32><<<grid, kTmaThreads, S::kTotal, st>>>(ta, tb, p);
  count_launch();
  trace_kernel(TF32 ? "conv_tma_tf32<%d,%d,%d>" : "conv_tma<%d,%d,%d>", BN, ST, MS);
  return check_cuda(cudaGetLastError(), "conv_tma_kernel launch");
}

// cta_group::2 launch: 2-CTA clusters, one pair per two SMs
template <int BN, int ST, int MS>
static int launch_tma_pair(const CUtensorMap& ta, const CUtensorMap& tb, const TmaConvParams& p, cudaStream_t st) {
  using S = TmaSmem<BN / 2, ST, MS>;
  static_assert(S::kTotal + 4096 + 4096 <= 227 * 1024 && S::kTotal > 114 * 1024, "shared memory budget / one CTA per SM");
  auto kern = conv_tma_kernel<BN, ST, MS, false, true>;
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_done = true;
  }
  const int items = ((p.m_tiles + 1) / 2) * p.n_tiles;
  const int pairs = items < num_sms() / 2 ? items : num_sms() / 2;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * pairs, 1, 1);
  cfg.blockDim = dim3(kTmaThreads, 1, 1);
  cfg.dynamicSmemBytes = S::kTotal;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  DEI2I_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
  count_launch();
  trace_kernel("conv_tma_pair<%d,%d,%d>", BN, ST, MS);
  return check_cuda(cudaGetLastError(), "conv_tma_kernel (pair) launch");
}

extern int g_tune_tma_pair;  // 1: N = 256 tiles on CTA pairs (cta_group::2) where applicable, 0: single-CTA kernel
extern int g_tune_tma_ms;  // 0 = auto, 1 / 2 = force the number of 128-pixel sub-tiles per work tile
extern int g_tune_mma_issue;  // -1 (default): warp-uniform MMA issue loop for N <= 128, single-thread loop for N = 256; 0 / 1 force one
int run_conv_tma(const void* x, int N, int H, int W, int C, const void* wmat, int rows, int K, const TmaConvParams& p_in,
                 int upper, int upper_w, cudaStream_t st, bool tf32) {
  TmaConvParams p = p_in;
  int bn = 256;
  while (bn > 64 && bn / 2 >= rows) bn /= 2;
  // same-box A/B: N = 64 454 -> 392 us (128 -> 64 at 256^2), N = 128 339 -> 337 us, N = 256 114.0 -> 115.5 us (the elect / syncwarp per k
  // block costs where a 128-cycle MMA already hid the issue overhead)
  p.issue_mode = g_tune_mma_issue >= 0 ? g_tune_mma_issue : (bn <= 128 ? 1 : 0);
  if (tf32) {  // FP32 operands: 32 channels per 128-byte row, one 128-pixel tile per work tile
    CUtensorMap ta, tb;
    if (int rc = encode_tmap_im2col_f32(&ta, x, (uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N, p.lower_w, upper_w, p.lower, upper,
                                        (uint32_t)p.stride, 32, 128))
      return rc;
    if (int rc = encode_tmap_2d_f32(&tb, wmat, (uint64_t)K, (uint64_t)rows, (uint64_t)K * 4, 32, (uint32_t)bn, 1)) return rc;
    p.n_tiles = (rows + bn - 1) / bn;
    p.m_tiles = (int)((p.M + 127) / 128);
    switch (bn) {
      case 256: return launch_tma<256, 4, 1, true>(ta, tb, p, st);
      case 128: return launch_tma<128, 6, 1, true>(ta, tb, p, st);
      default: return launch_tma<64, 8, 1, true>(ta, tb, p, st);
    }
  }
  // (Narrowing BLOCK_N to put more CTAs on the few-pixel / many-channel discriminator tail was measured and dropped: 1024 -> 2048
  // at 4 x 4 went 101 -> 90 us but 512 -> 1024 at 8 x 8 went 50 -> 89 us - every extra N tile re-loads the same A tiles.)
  CUtensorMap ta, tb;
  if (int rc = encode_tmap_im2col_bf16(&ta, x, (uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N, p.lower_w, upper_w, p.lower, upper, (uint32_t)p.stride,
                                       64, 128))
    return rc;
  if (int rc = encode_tmap_2d_bf16(&tb, wmat, (uint64_t)K, (uint64_t)rows, (uint64_t)K * 2, 64, (uint32_t)bn, 1)) return rc;
  p.n_tiles = (rows + bn - 1) / bn;
  if (conv_ring_applicable(p, rows, C, W, bn)) return launch_conv_ring(ta, tb, p, st);
  if (g_tune_tma_pair != 0 && bn == 256 && rows % 256 == 0 && p.npar <= 1 && (p.M + 127) / 128 >= 2) {
    CUtensorMap tb2;  // each CTA of a pair loads 128 of the 256 weight rows of an N tile
    if (int rc = encode_tmap_2d_bf16(&tb2, wmat, (uint64_t)K, (uint64_t)rows, (uint64_t)K * 2, 64, 128, 1)) return rc;
    p.m_tiles = (int)((p.M + 127) / 128);
    p.issue_mode = 0;
    return launch_tma_pair<256, 6, 1>(ta, tb2, p, st);
  }
  if (g_tune_tma_pair >= 2 && bn == 128 && rows % 128 == 0 && p.npar <= 1 && (p.M + 255) / 256 * p.n_tiles >= 4ll * num_sms()) {
    CUtensorMap tb2;  // N = 128 tiles on pairs: two 128-pixel tiles per CTA, 64 of the 128 weight rows each
    if (int rc = encode_tmap_2d_bf16(&tb2, wmat, (uint64_t)K, (uint64_t)rows, (uint64_t)K * 2, 64, 64, 1)) return rc;
    p.m_tiles = (int)((p.M + 255) / 256);
    return launch_tma_pair<128, 5, 2>(ta, tb2, p, st);
  }
  if (g_tune_tma_pair >= 3 && bn == 64 && rows % 64 == 0 && p.npar <= 1 && (p.M + 255) / 256 * p.n_tiles >= 4ll * num_sms()) {
    CUtensorMap tb2;  // N = 64 tiles on pairs: 32 of the 64 weight rows per CTA
    if (int rc = encode_tmap_2d_bf16(&tb2, wmat, (uint64_t)K, (uint64_t)rows, (uint64_t)K * 2, 64, 32, 1)) return rc;
    p.m_tiles = (int)((p.M + 255) / 256);
    return launch_tma_pair<64, 5, 2>(ta, tb2, p, st);
  }
  // narrow N: two 128-pixel A tiles share each B tile (halves the weight traffic per FLOP through L2 / shared memory) as long
  // as that still leaves every SM a couple of work tiles
  const long long mt2 = (p.M + 255) / 256;
  int ms = (bn <= 128 && mt2 * p.n_tiles * (p.npar > 1 ? p.npar : 1) >= 2ll * num_sms()) ? 2 : 1;
  if (g_tune_tma_ms == 1 || (g_tune_tma_ms == 2 && bn <= 128)) ms = g_tune_tma_ms;
  p.m_tiles = (int)((p.M + 128 * ms - 1) / (128 * ms));
  if (ms == 2) return bn == 128 ? launch_tma<128, 4, 2>(ta, tb, p, st) : launch_tma<64, 5, 2>(ta, tb, p, st);
  switch (bn) {
    case 256: return launch_tma<256, 4, 1>(ta, tb, p, st);
    case 128: return launch_tma<128, 6, 1>(ta, tb, p, st);
    default: return launch_tma<64, 8, 1>(ta, tb, p, st);
  }
}



// ======================================================================================================================
// wgrad:  P[co, k] = sum_{m in split} dY[m, co] * A[m, k]      (k = (tap, ci); reduction over output pixels m)
//   per stage: 64 pixels.  MMA "A" = dY tile [64 pix x 128*CO_T co] (tiled TMA, MN-major), MMA "B" = up to four im2col
//   boxes [64 pix x 64 ci] of x, one per (tap, 64-channel block) (im2col TMA, MN-major): N' = 64 * boxes.
//   CO_T = 2 keeps TWO accumulators (co 0-127 and 128-255) on the same B boxes: 64 KB of operands per 2 x 512 MMA cycles.
//   One work tile per CTA: (co group, k tile, pixel split); FP32 partial tiles go to the workspace through the
//   coalescing staging tile; wgrad_reduce_kernel (conv_igemm.cu) sums the splits in a fixed order.
template <int CO_T, int STAGES>
struct WgTmaSmem {
  static constexpr int kATile = CO_T * 16384;  // 64 pix x 128*CO_T co
  static constexpr int kBTile = 4 * 8192;      // 4 boxes of 64 pix x 64 ci
  static constexpr int kStage = kATile + kBTile;
  static constexpr int kStageOut = STAGES * kStage;
  static constexpr int kBarOff = kStageOut + 4 * 4096;
  static constexpr int kTotal = kBarOff + (2 * STAGES + 1) * 8 + 16 + 1024;
};

template <int CO_T, int STAGES>
__global__ void __launch_bounds__(kWgTmaThreads, 1)
wgrad_tma_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x, const TmaWgradParams p) {
  using S = WgTmaSmem<CO_T, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tfull_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = CO_T * 256;

  // work tile
  int t = blockIdx.x;
  const int split = t % p.splits;
  t /= p.splits;
  const int k_tile = t % p.k_tiles;
  const int co_group = t / p.k_tiles;
  const int box0 = k_tile * p.boxes_per_tile;
  const int nb = min(p.boxes_per_tile, p.num_boxes - box0);  // boxes in this tile
  const int co0 = co_group * 128 * CO_T;
  const int pb0 = split * p.pb_per_split;
  const int pb1 = min(p.num_pb, pb0 + p.pb_per_split);
  const int nsteps = pb1 - pb0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
  }
  if (warp == 5) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 4) {
    if (lane == 0 && nsteps > 0) {
      const int HoWo = p.Ho * p.Wo;
      // filter offsets of this tile's boxes
      int b_cb[4], b_s[4], b_r[4];
      for (int g = 0; g < 4; ++g) {
        const int bx = box0 + (g < nb ? g : 0);
        const int tap = bx / p.cin_blocks;
        b_cb[g] = bx - tap * p.cin_blocks;
        b_r[g] = tap / p.kw;
        b_s[g] = tap - b_r[g] * p.kw;
      }
      const uint32_t stage_bytes = (uint32_t)(S::kATile + nb * 8192);
      for (int it = 0; it < nsteps; ++it) {
        const uint32_t stage = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        mbar_wait(&empty_bar[stage], ph ^ 1u);
        uint8_t* a_dst = smem + stage * S::kStage;
        uint8_t* b_dst = a_dst + S::kATile;
        mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
        const long long m0 = (long long)(pb0 + it) * 64;
#pragma unroll
        for (int h = 0; h < 2 * CO_T; ++h) tma_load_2d(a_dst + h * 8192, &tmap_dy, &full_bar[stage], co0 + h * 64, (int)m0);
        const int img = (int)(m0 / HoWo);
        const int rem = (int)(m0 - (long long)img * HoWo);
        const int pr = rem / p.Wo;
        const int qc = rem - pr * p.Wo;
        const int cw = qc * p.stride + p.lower_w, ch = pr * p.stride + p.lower;
        for (int g = 0; g < nb; ++g)
          tma_load_im2col_4d(b_dst + g * 8192, &tmap_x, &full_bar[stage], b_cb[g] * 64, cw, ch, img, (uint16_t)b_s[g], (uint16_t)b_r[g]);
      }
    }
  } else if (warp == 5) {
    if (lane == 0 && nsteps > 0) {
      const uint32_t idesc = make_idesc_bf16(128, nb * 64, 1, 1);
      for (int it = 0; it < nsteps; ++it) {
        const uint32_t stage = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        mbar_wait(&full_bar[stage], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * S::kStage);
        const uint32_t b_addr = a_addr + S::kATile;
        // MN-major, 128B swizzle: LBO = stride between 64-element MN atoms (8192 B), SBO = between 8-row K groups (1024 B)
        const uint64_t b_desc = make_smem_desc(b_addr, 8192, 1024, 2);
#pragma unroll
        for (int ct = 0; ct < CO_T; ++ct) {
          const uint64_t a_desc = make_smem_desc(a_addr + ct * 16384, 8192, 1024, 2);
#pragma unroll
          for (int k = 0; k < 4; ++k)  // 16 pixels = 2 K atoms = 2048 B per MMA
            tc_mma_bf16(tmem_base + ct * 256, a_desc + (uint64_t)(128 * k), b_desc + (uint64_t)(128 * k), idesc, (it | k) != 0);
        }
        tc_commit(&empty_bar[stage]);
      }
      tc_commit(tfull_bar);
    }
  } else if (nsteps > 0) {
    // epilogue: fp32 partial tile -> workspace [split, co, Kf], rows = co, 32 columns (128 B) per pass, coalesced
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    uint8_t* stage = smem + S::kStageOut + warp * 4096;
    const int sub_r = lane >> 3, sub_j = lane & 7;
    const int k0 = box0 * 64;
#pragma unroll 1
    for (int ct = 0; ct < CO_T; ++ct) {
      const int co_w = co0 + ct * 128 + warp * 32;  // first row of this warp
      if (co_w >= p.cout) break;
#pragma unroll 1
      for (int c = 0; c < nb * 64; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + ct * 256 + (uint32_t)c, v);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stage + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
#pragma unroll
        for (int itr = 0; itr < 8; ++itr) {
          const int r = itr * 4 + sub_r;
          const uint4 val = *reinterpret_cast<const uint4*>(stage + r * 128 + ((sub_j ^ (r & 7)) << 4));
          if (co_w + r < p.cout)
            *reinterpret_cast<uint4*>(p.partial + ((long long)split * p.cout + co_w + r) * p.Kf + k0 + c + sub_j * 4) = val;
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <int CO_T, int ST>
static int launch_wgrad_tma(const CUtensorMap& tdy, const CUtensorMap& tx, const TmaWgradParams& p, cudaStream_t st) {
  using S = WgTmaSmem<CO_T, ST>;
  static_assert(S::kTotal <= 227 * 1024 && S::kTotal > 114 * 1024, "shared memory budget / one CTA per SM");
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(wgrad_tma_kernel<CO_T, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_done = true;
  }
  const int grid = p.co_groups * p.k_tiles * p.splits;
  wgrad_tma_kernel<CO_T, ST><<<grid, kWgTmaThreads, S::kTotal, st>>>(tdy, tx, p);
  count_launch();
  trace_kernel("wgrad_tma<%d,%d>", CO_T, ST);
  return check_cuda(cudaGetLastError(), "wgrad_tma_kernel launch");
}


// ======================================================================================================================
// CTA-pair wgrad (Cout >= 256): the single-CTA kernel keeps a [256 co x 256 k] tile per CTA and needs 64 KB of operands per 1024 MMA
// cycles - more than the L2 -> shared-memory path delivers (it runs at 77 % of what the forward reaches).  A pair keeps
// [256 co x 512 k]: CTA r loads the dY tile of ITS 128 output channels (16 KB) and half of the up-to-eight x boxes (32 KB) per 64
// pixels, the leader issues two M = 256 (cta_group::2) MMAs per 16 pixels - 48 KB per 1024 cycles per CTA.  Accumulators: 128 lanes
// x 512 columns in each CTA (its co rows x all k columns of the tile).
// MEASURED (same box): 256 -> 256 3x3 137.0 vs 135.4 us, 512 -> 1024 58.7 vs 60.2 us - no gain: the first pass of these layers already runs
// at the power-limited tensor rate (~105 of the 135 us; the rest is the second pass over the split partials).  Bit-exact, kept behind
// dei2i_set_tuning("wgrad_pair", 1), off by default.
constexpr int kWPairStages = 4;
struct WPairSmem {
  static constexpr int kATile = 16384;     // 64 pix x 128 co (this CTA's half of the 256)
  static constexpr int kBTile = 4 * 8192;  // this CTA's half of up to 8 boxes
  static constexpr int kStage = kATile + kBTile;
  static constexpr int kStageOut = kWPairStages * kStage;
  static constexpr int kBarOff = kStageOut + 4 * 4096;
  static constexpr int kTotal = kBarOff + (2 * kWPairStages + 1) * 8 + 16 + 1024;
};

__global__ void __launch_bounds__(kWgTmaThreads, 1)
wgrad_pair_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x, const TmaWgradParams p) {
  using S = WPairSmem;
  constexpr int STAGES = kWPairStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tfull_bar + 1);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 512;
  const uint32_t rank = cluster_ctarank();

  // work tile of the pair
  int t = blockIdx.x >> 1;
  const int split = t % p.splits;
  t /= p.splits;
  const int k_tile = t % p.k_tiles;
  const int co_group = t / p.k_tiles;
  const int box0 = k_tile * p.boxes_per_tile;
  const int rest = p.num_boxes - box0;
  const int nb = rest < p.boxes_per_tile ? rest : p.boxes_per_tile;  // even, <= 8
  const int n1 = nb < 4 ? nb : 4, n2 = nb - n1;                       // boxes of the two MMAs (N = 64 n1, 64 n2)
  const int h1 = n1 / 2, h2 = n2 / 2;                                 // ... of which this CTA holds half
  const int co0 = co_group * 256 + (int)rank * 128;
  const int pb0 = split * p.pb_per_split;
  const int pb1 = p.num_pb < pb0 + p.pb_per_split ? p.num_pb : pb0 + p.pb_per_split;
  const int nsteps = pb1 - pb0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
  }
  if (warp == 5) tmem_alloc2(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 4) {
    if (lane == 0 && nsteps > 0) {
      const int HoWo = p.Ho * p.Wo;
      int b_cb[4], b_s[4], b_r[4];  // this CTA's boxes: h1 of MMA 1, then h2 of MMA 2
      for (int g = 0; g < 4; ++g) {
        int bx = box0;
        if (g < h1) bx = box0 + (int)rank * h1 + g;
        else if (g < h1 + h2) bx = box0 + n1 + (int)rank * h2 + (g - h1);
        const int tap = bx / p.cin_blocks;
        b_cb[g] = bx - tap * p.cin_blocks;
        b_r[g] = tap / p.kw;
        b_s[g] = tap - b_r[g] * p.kw;
      }
      const uint32_t stage_bytes = 2u * (uint32_t)(S::kATile + (h1 + h2) * 8192);  // both CTAs credit the leader's barrier
      for (int it = 0; it < nsteps; ++it) {
        const uint32_t stage = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        mbar_wait(&empty_bar[stage], ph ^ 1u);
        uint8_t* a_dst = smem + stage * S::kStage;
        uint8_t* b_dst = a_dst + S::kATile;
        if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
        const long long m0 = (long long)(pb0 + it) * 64;
        tma2_load_2d(a_dst, &tmap_dy, &full_bar[stage], co0, (int)m0);
        tma2_load_2d(a_dst + 8192, &tmap_dy, &full_bar[stage], co0 + 64, (int)m0);
        const int img = (int)(m0 / HoWo);
        const int rem = (int)(m0 - (long long)img * HoWo);
        const int pr = rem / p.Wo;
        const int qc = rem - pr * p.Wo;
        const int cw = qc * p.stride + p.lower_w, ch = pr * p.stride + p.lower;
        for (int g = 0; g < h1 + h2; ++g)
          tma2_load_im2col_4d(b_dst + g * 8192, &tmap_x, &full_bar[stage], b_cb[g] * 64, cw, ch, img, (uint16_t)b_s[g], (uint16_t)b_r[g]);
      }
    }
  } else if (warp == 5) {
    if (lane == 0 && nsteps > 0 && rank == 0) {
      const uint32_t idesc1 = make_idesc_bf16(256, n1 * 64, 1, 1);
      const uint32_t idesc2 = make_idesc_bf16(256, (n2 > 0 ? n2 : 2) * 64, 1, 1);
      for (int it = 0; it < nsteps; ++it) {
        const uint32_t stage = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        mbar_wait(&full_bar[stage], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * S::kStage);
        const uint32_t b_addr = a_addr + S::kATile;
        const uint64_t a_desc = make_smem_desc(a_addr, 8192, 1024, 2);
        const uint64_t b1 = make_smem_desc(b_addr, 8192, 1024, 2);
        const uint64_t b2 = make_smem_desc(b_addr + h1 * 8192, 8192, 1024, 2);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          tc_mma2_bf16(tmem_base, a_desc + (uint64_t)(128 * k), b1 + (uint64_t)(128 * k), idesc1, (it | k) != 0);
          if (n2 > 0) tc_mma2_bf16(tmem_base + 256, a_desc + (uint64_t)(128 * k), b2 + (uint64_t)(128 * k), idesc2, (it | k) != 0);
        }
        tc_commit2(&empty_bar[stage]);
      }
      tc_commit2(tfull_bar);
    }
  } else if (nsteps > 0) {
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    uint8_t* stage = smem + S::kStageOut + warp * 4096;
    const int sub_r = lane >> 3, sub_j = lane & 7;
    const int k0 = box0 * 64;
    const int co_w = co0 + warp * 32;
    if (co_w < p.cout) {
#pragma unroll 1
      for (int c = 0; c < nb * 64; c += 32) {  // n1 = 4 whenever a second MMA exists: TMEM column == c
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stage + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
#pragma unroll
        for (int itr = 0; itr < 8; ++itr) {
          const int r = itr * 4 + sub_r;
          const uint4 val = *reinterpret_cast<const uint4*>(stage + r * 128 + ((sub_j ^ (r & 7)) << 4));
          if (co_w + r < p.cout)
            *reinterpret_cast<uint4*>(p.partial + ((long long)split * p.cout + co_w + r) * p.Kf + k0 + c + sub_j * 4) = val;
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, kTmemCols);
  }
}

extern int g_tune_wgrad_pair;
static bool wgrad_pair_ok(const TmaWgradParams& p) {
  return g_tune_wgrad_pair != 0 && p.cout >= 256 && p.cout % 128 == 0 && p.num_boxes % 2 == 0 && p.num_boxes >= 8;
}
static int launch_wgrad_pair(const CUtensorMap& tdy, const CUtensorMap& tx, const TmaWgradParams& p, cudaStream_t st) {
  static_assert(WPairSmem::kTotal <= 227 * 1024 && WPairSmem::kTotal > 114 * 1024, "shared memory budget / one CTA per SM");
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(wgrad_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WPairSmem::kTotal));
    attr_done = true;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * p.co_groups * p.k_tiles * p.splits, 1, 1);
  cfg.blockDim = dim3(kWgTmaThreads, 1, 1);
  cfg.dynamicSmemBytes = WPairSmem::kTotal;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  DEI2I_CUDA(cudaLaunchKernelEx(&cfg, wgrad_pair_kernel, tdy, tx, p));
  count_launch();
  trace_kernel("wgrad_pair<%d>", p.boxes_per_tile);
  return check_cuda(cudaGetLastError(), "wgrad_pair_kernel launch");
}

// plan shared with the workspace-size query: fills co_groups / k_tiles / splits / pb_per_split / num_pb
void wgrad_tma_plan(TmaWgradParams& p) {
  p.num_pb = (int)((p.M + 63) / 64);
  p.pair = 0;
  if (wgrad_pair_ok(p)) {  // CTA pairs: [256 co x up to 8 boxes] per pair
    p.co_groups = (p.cout + 255) / 256;
    p.k_tiles = (p.num_boxes + 7) / 8;
    p.boxes_per_tile = ((p.num_boxes + p.k_tiles - 1) / p.k_tiles + 1) & ~1;
    p.k_tiles = (p.num_boxes + p.boxes_per_tile - 1) / p.boxes_per_tile;
    const int tiles = p.co_groups * p.k_tiles;
    int splits = (num_sms() / 2) / tiles;
    if (splits < 1) splits = 1;
    while (splits > 1 && p.num_pb / splits < 4) --splits;
    p.pb_per_split = (p.num_pb + splits - 1) / splits;
    p.splits = (p.num_pb + p.pb_per_split - 1) / p.pb_per_split;
    p.pair = 1;
    return;
  }
  const int co_t = p.cout > 128 ? 2 : 1;
  p.co_groups = (p.cout + 128 * co_t - 1) / (128 * co_t);
  p.k_tiles = (p.num_boxes + 3) / 4;
  p.boxes_per_tile = (p.num_boxes + p.k_tiles - 1) / p.k_tiles;
  p.k_tiles = (p.num_boxes + p.boxes_per_tile - 1) / p.boxes_per_tile;
  p.num_pb = (int)((p.M + 63) / 64);
  const int tiles = p.co_groups * p.k_tiles;
  int splits = num_sms() / tiles;
  if (splits < 1) splits = 1;
  while (splits > 1 && p.num_pb / splits < 4) --splits;
  p.pb_per_split = (p.num_pb + splits - 1) / splits;
  p.splits = (p.num_pb + p.pb_per_split - 1) / p.pb_per_split;
}

int run_wgrad_tma(const void* x, int N, int H, int W, int C, const void* dy, int cout_ld, TmaWgradParams& p, int upper,
                  int upper_w, cudaStream_t st) {
  CUtensorMap tdy, tx;
  if (int rc = encode_tmap_2d_bf16(&tdy, dy, (uint64_t)cout_ld, (uint64_t)p.M, (uint64_t)cout_ld * 2, 64, 64, 1)) return rc;
  if (int rc = encode_tmap_im2col_bf16(&tx, x, (uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N, p.lower_w, upper_w, p.lower, upper, (uint32_t)p.stride, 64, 64))
    return rc;
  if (p.pair) return launch_wgrad_pair(tdy, tx, p, st);
  if (p.cout > 128) return launch_wgrad_tma<2, 3>(tdy, tx, p, st);
  return launch_wgrad_tma<1, 4>(tdy, tx, p, st);
}


// ======================================================================================================================
// Row-ring wgrad for kh x 1 filters over 64 input channels (the row-packed stem).  The generic kernel fetches one im2col box of x per
// filter tap and 64-pixel block: 7 x 8 KB + 16 KB of dY per 64 pixels, and is bound by that fill (337 us).  Here a CTA walks down a
// 64-pixel column block; the x tile of input row j is fetched once and serves as tap r of output row j - r.  The ring is MIRRORED
// (row j lands in slots j % R and j % R + R) so that the kh tiles of any output row are contiguous in shared memory and can be fed
// to the tensor core as ONE wide B operand (N = 256 for taps 0-3, N = 64 (kh - 4) for the rest) exactly like the generic kernel's
// four-box stages: 16 + 16 KB per 64 pixels instead of 72.  One accumulator [128 co x kh * 64] per CTA for its whole pixel range.
constexpr int kWRingR = 10;      // ring rows (kh in use + lead)
constexpr int kWRingDy = 3;      // dY stages
struct WRingSmem {
  static constexpr int kX = 2 * kWRingR * 8192;  // mirrored ring of 64 pix x 64 ci tiles
  static constexpr int kDy = kWRingDy * 16384;   // 64 pix x 128 co
  static constexpr int kStageOut = kX + kDy;
  static constexpr int kBarOff = kStageOut + 4 * 4096;
  static constexpr int kTotal = kBarOff + (2 * kWRingR + 2 * kWRingDy + 1) * 8 + 16 + 1024;
};

__global__ void __launch_bounds__(kWgTmaThreads, 1)
wgrad_ring_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x, const TmaWgradParams p,
                  long long tiles_per_cta, long long total_tiles, int kh) {
  using S = WRingSmem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* xring = smem;
  uint8_t* dys = smem + S::kX;
  uint64_t* xfull = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* xempty = xfull + kWRingR;
  uint64_t* dfull = xempty + kWRingR;
  uint64_t* dempty = dfull + kWRingDy;
  uint64_t* tfull_bar = dempty + kWRingDy;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tfull_bar + 1);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 512;
  const int CB = p.Wo / 64;
  const long long t0 = (long long)blockIdx.x * tiles_per_cta;
  const long long t1 = t0 + tiles_per_cta < total_tiles ? t0 + tiles_per_cta : total_tiles;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kWRingR; ++s) {
      mbar_init(&xfull[s], 1);
      mbar_init(&xempty[s], 1);
    }
    for (int s = 0; s < kWRingDy; ++s) {
      mbar_init(&dfull[s], 1);
      mbar_init(&dempty[s], 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
  }
  if (warp == 5) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 4) {
    // =========================== TMA producer: per output row one dY tile, per input row one x tile (twice) ===========================
    if (lane == 0 && t0 < t1) {
      uint32_t L = 0, D = 0;  // x rows / dY tiles issued
      for (long long t = t0; t < t1;) {
        const long long strip = t / p.Ho;
        const int ya = (int)(t - strip * p.Ho);
        const int yb = (int)((long long)p.Ho < ya + (t1 - t) ? (long long)p.Ho : ya + (t1 - t));
        const int img = (int)(strip / CB), cb = (int)(strip - (long long)img * CB);
        // input rows ya .. yb + kh - 2 and output rows ya .. yb - 1, interleaved in the order the MMA warp consumes them
        int jx = ya;
        for (int y = ya; y < yb; ++y) {
          for (; jx <= y + kh - 1; ++jx, ++L) {
            const uint32_t slot = L % kWRingR;
            mbar_wait(&xempty[slot], ((L / kWRingR) & 1u) ^ 1u);
            mbar_arrive_expect_tx(&xfull[slot], 2 * 8192);
            const int jb = jx < p.Ho - 1 ? jx : p.Ho - 1;
            tma_load_im2col_4d(xring + slot * 8192, &tmap_x, &xfull[slot], 0, cb * 64 + p.lower_w, jb + p.lower, img, (uint16_t)0, (uint16_t)(jx - jb));
            tma_load_im2col_4d(xring + (slot + kWRingR) * 8192, &tmap_x, &xfull[slot], 0, cb * 64 + p.lower_w, jb + p.lower, img, (uint16_t)0,
                               (uint16_t)(jx - jb));
          }
          const uint32_t ds = D % kWRingDy;
          mbar_wait(&dempty[ds], ((D / kWRingDy) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&dfull[ds], 16384);
          const long long m0 = ((long long)img * p.Ho + y) * p.Wo + cb * 64;
          tma_load_2d(dys + ds * 16384, &tmap_dy, &dfull[ds], 0, (int)m0);
          tma_load_2d(dys + ds * 16384 + 8192, &tmap_dy, &dfull[ds], 64, (int)m0);
          ++D;
        }
        t += yb - ya;
      }
    }
  } else if (warp == 5) {
    // =========================== MMA issuer ===========================
    if (lane == 0 && t0 < t1) {
      const int n1 = kh < 4 ? kh : 4, n2 = kh - n1;  // taps per MMA group
      const uint32_t idesc1 = make_idesc_bf16(128, n1 * 64, 1, 1);
      const uint32_t idesc2 = make_idesc_bf16(128, (n2 > 0 ? n2 : 1) * 64, 1, 1);
      uint32_t L0 = 0, D = 0;
      bool first = true;
      for (long long t = t0; t < t1;) {
        const long long strip = t / p.Ho;
        const int ya = (int)(t - strip * p.Ho);
        const int nrows = (int)((long long)p.Ho - ya < t1 - t ? (long long)p.Ho - ya : t1 - t);
        for (int y = 0; y < nrows; ++y, ++D) {
          for (int r = (y == 0 ? 0 : kh - 1); r < kh; ++r) {
            const uint32_t L = L0 + y + r;
            mbar_wait(&xfull[L % kWRingR], (L / kWRingR) & 1u);
          }
          const uint32_t ds = D % kWRingDy;
          mbar_wait(&dfull[ds], (D / kWRingDy) & 1u);
          tc_fence_after();
          const uint32_t slot = (L0 + y) % kWRingR;  // tap r sits at slot + r (mirrored ring: no wrap)
          const uint64_t a_desc = make_smem_desc(smem_u32(dys + ds * 16384), 8192, 1024, 2);
          const uint64_t b1 = make_smem_desc(smem_u32(xring + slot * 8192), 8192, 1024, 2);
          const uint64_t b2 = make_smem_desc(smem_u32(xring + (slot + 4) * 8192), 8192, 1024, 2);
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // 16 pixels per MMA
            tc_mma_bf16(tmem_base, a_desc + (uint64_t)(128 * k), b1 + (uint64_t)(128 * k), idesc1, !(first && k == 0));
            if (n2 > 0) tc_mma_bf16(tmem_base + 256, a_desc + (uint64_t)(128 * k), b2 + (uint64_t)(128 * k), idesc2, !(first && k == 0));
          }
          first = false;
          tc_commit(&dempty[ds]);
          tc_commit(&xempty[slot]);  // input row y is not needed by later output rows
          if (y == nrows - 1)
            for (int r = 1; r < kh; ++r) tc_commit(&xempty[(L0 + y + r) % kWRingR]);
        }
        L0 += nrows + kh - 1;
        t += nrows;
      }
      tc_commit(tfull_bar);
    }
  } else if (t0 < t1) {
    // =========================== epilogue: fp32 partial [split = CTA][co][kh * 64] ===========================
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    uint8_t* stage = smem + S::kStageOut + warp * 4096;
    const int sub_r = lane >> 3, sub_j = lane & 7;
    const int co_w = warp * 32;
    if (co_w < p.cout) {
#pragma unroll 1
      for (int c = 0; c < kh * 64; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stage + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
#pragma unroll
        for (int itr = 0; itr < 8; ++itr) {
          const int r = itr * 4 + sub_r;
          const uint4 val = *reinterpret_cast<const uint4*>(stage + r * 128 + ((sub_j ^ (r & 7)) << 4));
          if (co_w + r < p.cout)
            *reinterpret_cast<uint4*>(p.partial + ((long long)blockIdx.x * p.cout + co_w + r) * p.Kf + c + sub_j * 4) = val;
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

extern int g_tune_wgrad_ring;
bool wgrad_ring_applicable(const TmaWgradParams& p, int kh, int C, int W, int cout_ld) {
  return g_tune_wgrad_ring != 0 && p.kw == 1 && kh >= 2 && kh < kWRingR && C == 64 && cout_ld == 64 && p.cout <= 64 && p.stride == 1 &&
         p.lower_w == 0 && p.Wo == W && p.Wo % 64 == 0 && p.Ho >= kh && p.Kf == kh * 64 &&
         p.M == (p.M / ((long long)p.Ho * p.Wo)) * p.Ho * p.Wo;
}
int wgrad_ring_splits(const TmaWgradParams& p) {
  const long long total = (p.M / ((long long)p.Ho * p.Wo)) * (p.Wo / 64) * p.Ho;
  const int grid = total < num_sms() ? (int)total : num_sms();
  const long long per = (total + grid - 1) / grid;
  return (int)((total + per - 1) / per);
}
int run_wgrad_ring(const void* x, int N, int H, int W, int C, const void* dy, int cout_ld, TmaWgradParams& p, int kh, int upper, int upper_w,
                   cudaStream_t st) {
  CUtensorMap tdy, tx;
  if (int rc = encode_tmap_2d_bf16(&tdy, dy, (uint64_t)cout_ld, (uint64_t)p.M, (uint64_t)cout_ld * 2, 64, 64, 1)) return rc;
  if (int rc = encode_tmap_im2col_bf16(&tx, x, (uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N, p.lower_w, upper_w, p.lower, upper, 1u, 64, 64))
    return rc;
  static_assert(WRingSmem::kTotal <= 227 * 1024 && WRingSmem::kTotal > 114 * 1024, "shared memory budget / one CTA per SM");
  static bool attr_done = false;
  if (!attr_done) {
    DEI2I_CUDA(cudaFuncSetAttribute(wgrad_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WRingSmem::kTotal));
    attr_done = true;
  }
  const long long total = (p.M / ((long long)p.Ho * p.Wo)) * (p.Wo / 64) * p.Ho;
  const int grid = wgrad_ring_splits(p);
  const long long per = (total + (total < num_sms() ? total : num_sms()) - 1) / (total < num_sms() ? total : num_sms());
  p.splits = grid;
  wgrad_ring_kernel<<<grid, kWgTmaThreads, WRingSmem::kTotal, st>>>(tdy, tx, p, per, total, kh);
  count_launch();
  trace_kernel("wgrad_ring<%d>", kh);
  return check_cuda(cudaGetLastError(), "wgrad_ring_kernel launch");
}

}  // namespace dei2i
